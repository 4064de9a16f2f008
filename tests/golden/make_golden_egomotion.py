"""Regenerate tests/golden/egomotion.json: outputs of the COMPILED REFERENCE's VisualOdometryMono::estimateMotion
(/root/reference/src/viso_mono.cpp:41, through oracle/_ref/libvoref_native.so) on the synthetic 3-D-scene match
lists of hls_final_visual_odometry_b200.synth.make_scene_matches.  Run where /root/reference exists:
    python tests/golden/make_golden_egomotion.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from hls_final_visual_odometry_b200.synth import KITTI_CALIB, KITTI_HEIGHT, KITTI_PITCH, make_scene_matches  # noqa: E402
from oracle.ref_lib import RefLib, mono_motion_seq  # noqa: E402

CASES = [dict(seed=s, n_lists=5, n_points=320, outlier_frac=0.05) for s in range(4)] + \
        [dict(seed=10, n_lists=3, n_points=120, outlier_frac=0.2), dict(seed=11, n_lists=3, n_points=400, outlier_frac=0.0),
         dict(seed=12, n_lists=2, n_points=320, outlier_frac=0.05, step=0.0)]


def main():
    ref = RefLib("native")
    out = []
    for case in CASES:
        kw = {k: v for k, v in case.items() if k not in ("seed", "n_lists")}
        lists, _ = make_scene_matches(case["n_lists"], seed=case["seed"], **kw)
        tr, ok, ninl = mono_motion_seq(ref, lists, height=KITTI_HEIGHT, pitch=KITTI_PITCH, **KITTI_CALIB)
        out.append(dict(case=case, n_matches=[len(m) for m in lists], ok=ok.tolist(), n_inliers=ninl.tolist(),
                        tr=[[float.hex(float(x)) for x in row] for row in tr]))
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "egomotion.json"), "w") as f:
        json.dump(dict(source="VisualOdometryMono::estimateMotion of the compiled reference (oracle/_ref/libvoref_native.so)",
                       calib=KITTI_CALIB, height=KITTI_HEIGHT, pitch=KITTI_PITCH, cases=out), f, indent=1)
    print("wrote egomotion.json:", [(c["ok"], c["n_inliers"]) for c in out])


if __name__ == "__main__":
    main()
