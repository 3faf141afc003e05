"""Feature-extraction figures alone (what bench.py reports under "features"); optional --batch."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--variant", type=int, default=0, help="response kernel: 0 auto, 1 warp-specialised, 2 column walk")
    a = ap.parse_args()
    from visual_slam_b200 import _cabi
    _cabi.check(_cabi.lib().vslam_features_set_variant(a.variant))
    print(json.dumps(bench.features_figures(torch, a.batch, a.steps)))
