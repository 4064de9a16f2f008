"""Regenerate tests/golden/egomotion_stereo.json: outputs of the COMPILED REFERENCE's VisualOdometryStereo::estimateMotion
(/root/reference/src/viso_stereo.cpp:54, through oracle/_ref/libvoref_native.so) on synthetic quad match lists of a
rectified rig (hls_final_visual_odometry_b200.synth.make_scene_matches with base > 0).  Run where /root/reference exists:
    python tests/golden/make_golden_egomotion_stereo.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from hls_final_visual_odometry_b200.synth import KITTI_CALIB, make_scene_matches  # noqa: E402
from oracle.ref_lib import RefLib, stereo_motion_seq  # noqa: E402

BASE = 0.54  # KITTI rig
CASES = [dict(seed=s, n_lists=5, n_points=320, outlier_frac=0.05) for s in range(3)] + \
        [dict(seed=20, n_lists=3, n_points=120, outlier_frac=0.3), dict(seed=21, n_lists=3, n_points=400, outlier_frac=0.0),
         dict(seed=22, n_lists=2, n_points=320, outlier_frac=0.05, step=0.0), dict(seed=23, n_lists=2, n_points=8, outlier_frac=0.0)]


def main():
    ref = RefLib("native")
    out = []
    for case in CASES:
        kw = {k: v for k, v in case.items() if k not in ("seed", "n_lists")}
        lists, _ = make_scene_matches(case["n_lists"], seed=case["seed"], base=BASE, **kw)
        tr, ok, ninl = stereo_motion_seq(ref, lists, base=BASE, **KITTI_CALIB)
        out.append(dict(case=case, n_matches=[len(m) for m in lists], ok=ok.tolist(), n_inliers=ninl.tolist(),
                        tr=[[float.hex(float(x)) for x in row] for row in tr]))
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "egomotion_stereo.json"), "w") as f:
        json.dump(dict(source="VisualOdometryStereo::estimateMotion of the compiled reference (oracle/_ref/libvoref_native.so)",
                       calib=KITTI_CALIB, base=BASE, cases=out), f, indent=1)
    print("wrote egomotion_stereo.json:", [(c["ok"], c["n_inliers"]) for c in out])


if __name__ == "__main__":
    main()
