"""Converts the reference's only BA artefact, /root/reference/debug.txt (g2o's "Cholesky failure"
dump of a 90x90 damped reduced camera system, Octave sparse text, SURVEY.md section 4 fixture F1),
into tests/golden/ba_debug_reduced_system.npz.  Runs only in the build container."""
import os

import numpy as np


def main():
    rows, cols, vals = [], [], []
    n = None
    with open("/root/reference/debug.txt") as f:
        for line in f:
            if line.startswith("# rows:"):
                n = int(line.split(":")[1])
            if line.startswith("#") or not line.strip():
                continue
            r, c, v = line.split()
            rows.append(int(r) - 1); cols.append(int(c) - 1); vals.append(float(v))
    S = np.zeros((n, n))
    S[rows, cols] = vals
    path = os.path.join(os.path.dirname(__file__), "ba_debug_reduced_system.npz")
    np.savez_compressed(path, S=S)
    w = np.linalg.eigvalsh(S)
    print(path, S.shape, "sym", np.array_equal(S, S.T), "eig min/max", w[0], w[-1])


if __name__ == "__main__":
    main()
