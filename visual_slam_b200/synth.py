"""Seeded synthetic workloads of the shapes BASELINE.json names."""
import numpy as np


def descriptors_pair(nq=2048, nt=2048, seed=0, planted=0.7, flip=0.10):
    """C2: train = uniform bits; `planted` of the queries are distinct train rows with each
    bit flipped w.p. `flip`; the rest are fresh uniform rows.  Returns (q, t, truth[nq])."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    truth = np.full(nq, -1, dtype=np.int32)
    n_pl = min(int(round(planted * nq)), nt)
    if n_pl > 0:
        src = rng.permutation(nt)[:n_pl]
        rows = rng.permutation(nq)[:n_pl]
        noise = np.packbits(rng.random((n_pl, 256)) < flip, axis=1)
        q[rows] = t[src] ^ noise
        truth[rows] = src
    return q, t, truth


def descriptors_batch(batch, nq=2048, nt=2048, seed=0, **kw):
    qs, ts = [], []
    for b in range(batch):
        q, t, _ = descriptors_pair(nq, nt, seed=seed * 1000003 + b, **kw)
        qs.append(q); ts.append(t)
    return np.stack(qs), np.stack(ts)


def low_entropy_pair(nq, nt, seed=0):
    """Adversarial tie set: every byte is 0 or 1, so distances collide massively."""
    rng = np.random.default_rng(seed)
    return (rng.integers(0, 2, size=(nq, 32), dtype=np.uint8),
            rng.integers(0, 2, size=(nt, 32), dtype=np.uint8))
