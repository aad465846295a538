"""Config C2: extrinsic calibration with N checkerboard poses (default 500 boards = 27,000 points) through
the reference-shaped call chain: board preparation (host + K2) -> bundle_adjust_points_and_extrinsics. Needs a GPU."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, sba, calib, dense, _capi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 500
b = synth.make_board_scene(n, seed=3)
t0 = time.perf_counter()
p2, p3, pi, ci = sba.prepare_calib_board_data_for_bundle_adjustment(b['img_pts_arr'], b['fnames_arr'], (9, 6), b['k'], b['d'], b['r0'], b['t0'],
                                                                    calib.triangulate_points_fisheye)
t1 = time.perf_counter()
for rep in range(3):
    t2 = time.perf_counter()
    obj, r1, tt1, res = sba.bundle_adjust_points_and_extrinsics(p2, p3, pi, ci, b['k'], b['d'], b['r0'].copy(), b['t0'].copy(), calib.project_points_fisheye)
    t3 = time.perf_counter()
    st = sba.bundle_adjust_points_and_extrinsics.last_stats
    print(f"{n} boards: {len(p3)} points, {len(p2)} observations; prep {t1 - t0:.3f} s; solve call {t3 - t2:.4f} s, device {st['solve_ms']:.2f} ms, "
          f"{st['iterations']} iterations ({st['solve_ms'] / max(st['iterations'], 1) * 1e3:.1f} us each), status {st['status']}, "
          f"RMSE {np.sqrt(np.mean(res['before'] ** 2)):.4f} -> {np.sqrt(np.mean(res['after'] ** 2)):.4f} px")
