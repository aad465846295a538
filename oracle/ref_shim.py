"""TEST INFRASTRUCTURE ONLY — loads the *real* reference modules from /root/reference/src.

Only usable in the build container (the GPU box has no /root/reference); used by
oracle/make_golden.py to generate the committed fixtures and by CPU tests that are
skipped when the reference tree is absent. Shims: a stub
``nptyping.Array`` and the removed numpy aliases ``np.int`` / ``np.float``.
Reference files are never edited.
"""
import os
import sys
import types
import contextlib
import io

REF_SRC = '/root/reference/src'
REF_SCENE = '/root/reference/configs/dummy_scene.json'


def available():
    return os.path.isdir(os.path.join(REF_SRC, 'lib'))


def load():
    """Returns (calib, utils, sba) reference modules."""
    import numpy as np
    if not hasattr(np, 'int'):
        np.int = int
    if not hasattr(np, 'float'):
        np.float = float
    if 'nptyping' not in sys.modules:
        m = types.ModuleType('nptyping')

        class Array:
            def __class_getitem__(cls, _k):
                return cls
        m.Array = Array
        sys.modules['nptyping'] = m
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)
    from lib import calib, utils, sba  # noqa
    return calib, utils, sba


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield
