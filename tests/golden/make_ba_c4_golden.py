"""Generates tests/golden/ba_c4_restatement.npz: the LM trajectory of the CPU restatement (oracle/ba.py, "restatement,
not g2o": PARITY UNPINNED for the arithmetic) on BASELINE.json configs[3] -- 200 key frames x 100k points x ~1M
observations, seed 1, the same generator call bench.py and tests/test_ba_gpu.py make.  Per outer iteration: cost,
lambda; totals: iterations, trials; final cost, RMS reprojection error, and a digest of the final state.  Takes a
few minutes of CPU; run from the repo root:  python tests/golden/make_ba_c4_golden.py
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ba as oba  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402

CFG = dict(n_cams=200, n_pts=100000, obs_per_pt=10.0, seed=1)


def main():
    p = synthetic_ba(**CFG)
    t0 = time.time()
    h = oba.optimize(p, 10, verbose=True)
    out = {"cfg": np.array(repr(CFG)), "n_cams": p.n_cams, "n_pts": p.n_pts, "n_obs": p.n_obs,
           "costs": np.array(h["cost"]), "lambdas": np.array(h["lambda"]), "iterations": h["iterations"],
           "trials": h["trials"], "final_cost": h["final_cost"], "rms_px": h["rms_px"],
           "cams": p.cams, "pts_sample_idx": np.arange(0, p.n_pts, 997), "pts_sample": p.pts[::997]}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ba_c4_restatement.npz")
    np.savez_compressed(path, **out)
    print("written", path, os.path.getsize(path), "bytes; %.1f s; iterations %d trials %d final cost %.9g rms %.6f"
          % (time.time() - t0, h["iterations"], h["trials"], h["final_cost"], h["rms_px"]))


if __name__ == "__main__":
    main()
