"""N1 host logic (no GPU): DeepLabCut table ingest against the reference's load_dlc_points_as_df
(src/lib/utils.py:77-151; golden made by oracle/make_golden_n3.py from the real function) and the direct dense loader.
PyTables is absent from the image, so pd.read_hdf is pointed at in-memory tables, exactly as the golden generator did."""
import os
import numpy as np
import pandas as pd
import pytest

from acinoset_b200 import pipeline, synth, utils

MARKERS = ['nose', 'r_eye', 'l_eye', 'neck_base', 'spine', 'tail_base']


@pytest.fixture
def dlc_files(monkeypatch):
    tabs = synth.make_dlc_tables(12, MARKERS, 4, seed=7, first_frame=3)
    paths = [f'/data/cam{i + 1}DLC.h5' for i in range(4)]
    store = dict(zip(paths, tabs))
    monkeypatch.setattr(pd, 'read_hdf', lambda p, *a, **k: store[p].copy())
    return paths


@pytest.mark.parametrize('name,shifts', [('noshift', None), ('shift', [0, 2, -1, 0])])
def test_load_dlc_points_as_df_matches_reference(dlc_files, golden_dir, name, shifts):
    g = np.load(os.path.join(golden_dir, 'golden_dlc_load.npz'))
    df = utils.load_dlc_points_as_df(dlc_files, frame_shifts=shifts)
    assert list(df.columns) == ['frame', 'camera', 'marker', 'x', 'y', 'likelihood']
    assert np.array_equal(df['frame'].to_numpy(), g[f'{name}_frame'])
    assert np.array_equal(df['camera'].to_numpy(), g[f'{name}_camera'])
    assert np.array_equal(df['marker'].to_numpy().astype(str), g[f'{name}_marker'])
    for col in ('x', 'y', 'likelihood'):
        assert np.array_equal(df[col].to_numpy(), g[f'{name}_{col}'], equal_nan=True), col
    assert not df['likelihood'].isna().any()          # fillna(0), utils.py:138


def test_dense_loader_equals_table_path(dlc_files):
    """The direct array ingest gives the same slots as load_dlc_points_as_df -> dlc_df_to_dense, window and marker order
    included, without building the long table."""
    markers = ['spine', 'nose', 'l_eye', 'lure']            # a marker no file holds stays empty
    shifts = [0, 2, -1, 0]
    df = utils.load_dlc_points_as_df(dlc_files, frame_shifts=shifts)
    uv_a, lik_a = pipeline.dlc_df_to_dense(df, 4, markers, 5, 12)
    uv_b, lik_b = pipeline.load_dlc_points_dense(dlc_files, markers, 5, 12, frame_shifts=shifts)
    assert np.array_equal(lik_a, lik_b)
    seen = lik_a >= 0
    assert np.array_equal(uv_a[seen], uv_b[seen], equal_nan=True)
    assert (lik_b.reshape(8, 4, 4)[:, 3, :] == -1).all()
    # thresholding drops NaN coordinates' companions only through the likelihood, as in the reference
    assert (lik_b[np.isnan(uv_b[..., 0]) & seen] >= 0).all()


def test_frame_shift_assert(dlc_files):
    with pytest.raises(AssertionError):
        utils.load_dlc_points_as_df(dlc_files, frame_shifts=[0, 1])


@pytest.mark.parametrize('name', ['nolik', 'head'])
def test_tables_without_likelihood_match_reference(monkeypatch, golden_dir, name):
    """tables without a likelihood column take it from the y column (src/lib/utils.py:104-118); the 'dlc_head'
    tables take it from x and keep nose / r_eye / l_eye only (utils.py:84-103). The synthetic tables hold cells that are
    NaN in x only and in y only, so the two rules give different tables."""
    g = np.load(os.path.join(golden_dir, 'golden_dlc_load.npz'))
    tabs, paths = (synth.make_dlc_xy_tables(9, ['nose', 'r_eye', 'spine'], 3, seed=11) if name == 'nolik'
                   else synth.make_dlc_head_tables(7, 2, seed=12))
    store = dict(zip(paths, tabs))
    monkeypatch.setattr(pd, 'read_hdf', lambda p, *a, **k: store[p].copy())
    df = utils.load_dlc_points_as_df(paths)
    assert np.array_equal(df['frame'].to_numpy(), g[f'{name}_frame'])
    assert np.array_equal(df['camera'].to_numpy(), g[f'{name}_camera'])
    assert np.array_equal(df['marker'].to_numpy().astype(str), g[f'{name}_marker'])
    for col in ('x', 'y', 'likelihood'):
        assert np.array_equal(df[col].to_numpy(), g[f'{name}_{col}'], equal_nan=True), col
    x_only = np.isnan(df['x'].to_numpy()) & ~np.isnan(df['y'].to_numpy())
    y_only = ~np.isnan(df['x'].to_numpy()) & np.isnan(df['y'].to_numpy())
    assert x_only.any() and y_only.any()
