"""Generates tests/golden/essential_g1.npz by RUNNING THE REFERENCE in this container.

Golden vector G1: the notebook src/testing.py records, at :510-518, R and t for
ICL-NUIM frames 1 -> 30.  This script executes the notebook's own definitions cell (cell 1, with the
absent `import g2o` lines dropped) and its driver cell (cell 2, data directory re-pointed to
/root/reference/data and the loop cut after frame 30), with recording wrappers around
cv2.undistortPoints / cv2.findEssentialMat / cv2.recoverPose, and stores the arguments and results of
those calls together with the digits the notebook printed.  Nothing else is computed here.

Run from the repo root:  python tests/golden/make_essential_golden.py
Needs /root/reference (read-only); the resulting .npz travels with the repo.
"""
import json
import os
import sys

import cv2
import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "essential_g1.npz")

# digits printed by the notebook (src/testing.py:510-518)
G1_R = np.array([[9.99504922e-01, -1.67506360e-02, -2.66331853e-02],
                 [1.67488668e-02, 9.99859686e-01, -2.89518124e-04],
                 [2.66342979e-02, -1.56700885e-04, 9.99645232e-01]])
G1_T = np.array([0.31807891, 0.10384802, 0.94235948])


def main():
    nb = json.load(open(os.path.join(REF, "src", "testing.py")))
    cell1 = "".join(nb["cells"][1]["source"])
    cell2 = "".join(nb["cells"][2]["source"])
    cell1 = "\n".join(l for l in cell1.split("\n") if l.strip() != "import g2o")
    cell2 = cell2.replace('cur_dir = "/home/juuso"', 'cur_dir = "/root/reference/.."')
    cell2 = cell2.replace('cur_dir + "/visual_slam/data/ICL_NUIM/', '"/root/reference/data/ICL_NUIM/')
    cell2 = cell2.replace("range(2,1000)", "range(2,31)")
    # the display tail of the loop (imshow/waitKey) cannot run headless: cut it
    cut = cell2.find("# Display")
    if cut > 0:
        cell2 = cell2[:cut] + "\n        prev_frame = cur_frame\n"
    rec = {}
    real_undist, real_fe, real_rp = cv2.undistortPoints, cv2.findEssentialMat, cv2.recoverPose

    def undist(src, *a, **kw):
        out = real_undist(src, *a, **kw)
        rec.setdefault("undist_in", []).append(np.array(src).copy())
        rec.setdefault("undist_out", []).append(np.array(out).copy())
        return out

    def fe(p1, p2, *a, **kw):
        out = real_fe(p1, p2, *a, **kw)
        rec["fe_p1"], rec["fe_p2"] = np.array(p1).copy(), np.array(p2).copy()
        rec["fe_args"] = (a, kw)
        rec["fe_E"], rec["fe_mask"] = out[0].copy(), out[1].copy()
        return out

    def rp(E, p1, p2, *a, **kw):
        out = real_rp(E, p1, p2, *a, **kw)
        rec["rp_E"], rec["rp_p1"], rec["rp_p2"] = np.array(E).copy(), np.array(p1).copy(), np.array(p2).copy()
        rec["rp_args"] = (a, kw)
        rec["rp_out"] = out
        return out

    cv2.undistortPoints, cv2.findEssentialMat, cv2.recoverPose = undist, fe, rp
    g = {"cv2": cv2, "np": np}
    try:
        exec(compile("import cv2\n" + cell1, "testing.py:cell1", "exec"), g)
        exec(compile(cell2, "testing.py:cell2", "exec"), g)
    finally:
        cv2.undistortPoints, cv2.findEssentialMat, cv2.recoverPose = real_undist, real_fe, real_rp
    print("findEssentialMat args:", rec["fe_args"])
    print("recoverPose args:", rec["rp_args"])
    out = rec["rp_out"]
    R, t = out[1], out[2]
    print("R\n", R, "\nt", t.ravel())
    assert np.allclose(R, G1_R, atol=5e-9) and np.allclose(t.ravel(), G1_T, atol=5e-9), "G1 not reproduced"
    K = np.asarray(g["K"], dtype=np.float64)
    np.savez_compressed(
        OUT, K=K, fe_p1=rec["fe_p1"], fe_p2=rec["fe_p2"], fe_E=rec["fe_E"], fe_mask=rec["fe_mask"],
        fe_threshold=np.float64(rec["fe_args"][1].get("threshold", np.nan)),
        rp_E=rec["rp_E"], rp_p1=rec["rp_p1"], rp_p2=rec["rp_p2"], rp_n=np.int64(out[0]), rp_R=R, rp_t=t,
        rp_mask=(out[3] if len(out) > 3 and out[3] is not None else np.zeros(0)),
        undist_in=rec["undist_in"][0], undist_out=rec["undist_out"][0],
        g1_R=G1_R, g1_t=G1_T, cv2_version=np.array(cv2.__version__))
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    sys.exit(main())
