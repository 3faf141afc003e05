"""CPU experiment behind the decision to keep P3P + 1 as the RANSAC minimal solver (deviation i).

cv2.solvePnPRansac's default kernel is EPnP on 5-point samples.  With 5 points the 10 x 12 matrix M of EPnP has an
EXACT two-dimensional null space, and EPnP's "approximate betas + five Gauss-Newton steps" depend on which basis of
that null space the SVD happens to return (rounding decides).  This script restates OpenCV's epnp.cpp in NumPy and
measures, on 5-point samples of a synthetic tracking pair, (a) how often the restatement lands within a tolerance of
cv2.solvePnP(flags=SOLVEPNP_EPNP), and (b) how often the restatement agrees WITH ITSELF when only the null-space
basis is changed (LAPACK's eigenvectors vs a canonical basis).  (b) bounds (a): an independent implementation
cannot reproduce cv2's 5-point EPnP hypotheses, so parity on a "given hypothesis set" (north_star) cannot be
defined through that solver.  (c) Control on 50-point samples, where the null space is not degenerate: the
restatement equals cv2 to 1e-14 on noise-free data (both recover the true pose) but differs from it by ~4e-4 rad /
3e-3 m at 0.5 px noise -- the size of the noise-induced error itself -- so OpenCV 4.13's EPnP also differs from
the published epnp.cpp in a step that cannot be identified without its source (absent from this image).
Measured here (OpenCV 4.13, 300 samples): restatement vs cv2 within 1e-6: 2 %; restatement vs itself under a change
of null-space basis within 1e-6: 27 %, within 1e-3: 53 %.
Run: python tools/probes/epnp5_basis_dependence.py
"""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))

def qr_solve(A, b):
    return np.linalg.lstsq(A, b, rcond=None)[0]

def epnp(X, uv, K, canon=True):
    X = np.asarray(X, np.float64); uv = np.asarray(uv, np.float64); n = len(X)
    fu, fv, uc, vc = K
    # control points
    c0 = X.mean(0)
    P0 = X - c0
    w, V = np.linalg.eigh(P0.T @ P0)          # ascending
    order = np.argsort(-w); w = w[order]; V = V[:, order]
    cws = np.vstack([c0] + [c0 + np.sqrt(w[i] / n) * V[:, i] for i in range(3)])
    CC = (cws[1:] - cws[0]).T                  # columns
    al = np.linalg.solve(CC, (X - cws[0]).T).T
    alphas = np.hstack([1 - al.sum(1, keepdims=True), al])
    M = np.zeros((2 * n, 12))
    for i in range(n):
        for j in range(4):
            M[2*i, 3*j] = alphas[i, j] * fu; M[2*i, 3*j+2] = alphas[i, j] * (uc - uv[i, 0])
            M[2*i+1, 3*j+1] = alphas[i, j] * fv; M[2*i+1, 3*j+2] = alphas[i, j] * (vc - uv[i, 1])
    MtM = M.T @ M
    ew, ev = np.linalg.eigh(MtM)               # ascending: ev[:,0] smallest
    v = [ev[:, k].copy() for k in range(4)]    # v[0] = smallest (cv: ut[11])
    if canon and n == 5:
        # canonical basis of the exact 2-D null space: v0 = normalised projection of a fixed vector, v1 = complement
        N = np.stack([v[0], v[1]], 1)
        ref = np.ones(12) / np.sqrt(12)
        p = N @ (N.T @ ref)
        v0 = p / np.linalg.norm(p)
        q = N[:, 0] - v0 * (v0 @ N[:, 0])
        if np.linalg.norm(q) < 0.5:
            q = N[:, 1] - v0 * (v0 @ N[:, 1])
        v1 = q / np.linalg.norm(q)
        v[0], v[1] = v0, v1
    for k in range(4):
        i = np.argmax(np.abs(v[k]))
        if v[k][i] < 0: v[k] = -v[k]
    pairs = [(0,1),(0,2),(0,3),(1,2),(1,3),(2,3)]
    L = np.zeros((6, 10)); rho = np.zeros(6)
    dv = np.zeros((4, 6, 3))
    for k in range(4):
        for r,(a,b) in enumerate(pairs):
            dv[k, r] = v[k][3*a:3*a+3] - v[k][3*b:3*b+3]
    for r,(a,b) in enumerate(pairs):
        d = dv[:, r]
        L[r] = [d[0]@d[0], 2*d[0]@d[1], d[1]@d[1], 2*d[0]@d[2], 2*d[1]@d[2], d[2]@d[2], 2*d[0]@d[3], 2*d[1]@d[3], 2*d[2]@d[3], d[3]@d[3]]
        rho[r] = np.sum((cws[a]-cws[b])**2)
    def gauss_newton(b):
        b = b.copy()
        for _ in range(5):
            A = np.zeros((6,4)); rhs = np.zeros(6)
            for i in range(6):
                l = L[i]
                A[i] = [2*l[0]*b[0]+l[1]*b[1]+l[3]*b[2]+l[6]*b[3], l[1]*b[0]+2*l[2]*b[1]+l[4]*b[2]+l[7]*b[3],
                        l[3]*b[0]+l[4]*b[1]+2*l[5]*b[2]+l[8]*b[3], l[6]*b[0]+l[7]*b[1]+l[8]*b[2]+2*l[9]*b[3]]
                rhs[i] = rho[i] - (l[0]*b[0]*b[0]+l[1]*b[0]*b[1]+l[2]*b[1]*b[1]+l[3]*b[0]*b[2]+l[4]*b[1]*b[2]+l[5]*b[2]*b[2]+l[6]*b[0]*b[3]+l[7]*b[1]*b[3]+l[8]*b[2]*b[3]+l[9]*b[3]*b[3])
            b += qr_solve(A, rhs)
        return b
    def R_and_t(b):
        ccs = sum(b[k] * v[k] for k in range(4)).reshape(4,3)
        pcs = alphas @ ccs
        if pcs[0,2] < 0: ccs = -ccs; pcs = -pcs
        pc0 = pcs.mean(0); pw0 = X.mean(0)
        ABt = (pcs - pc0).T @ (X - pw0)
        U,S,Vt = np.linalg.svd(ABt)
        R = U @ Vt
        if np.linalg.det(R) < 0: R[2] = -R[2]
        t = pc0 - R @ pw0
        pc = X @ R.T + t
        ue = uc + fu*pc[:,0]/pc[:,2]; ve = vc + fv*pc[:,1]/pc[:,2]
        err = np.mean(np.sqrt((uv[:,0]-ue)**2 + (uv[:,1]-ve)**2))
        return R, t, err
    cands = []
    # approx 1
    b4 = np.linalg.lstsq(L[:, [0,1,3,6]], rho, rcond=None)[0]
    b = np.zeros(4)
    if b4[0] < 0: b[0] = np.sqrt(-b4[0]); b[1:] = -b4[1:]/b[0]
    else: b[0] = np.sqrt(b4[0]); b[1:] = b4[1:]/b[0]
    cands.append(R_and_t(gauss_newton(b)))
    b3 = np.linalg.lstsq(L[:, [0,1,2]], rho, rcond=None)[0]
    b = np.zeros(4)
    if b3[0] < 0: b[0] = np.sqrt(-b3[0]); b[1] = np.sqrt(-b3[2]) if b3[2] < 0 else 0.0
    else: b[0] = np.sqrt(b3[0]); b[1] = np.sqrt(b3[2]) if b3[2] > 0 else 0.0
    if b3[1] < 0: b[0] = -b[0]
    cands.append(R_and_t(gauss_newton(b)))
    b5 = np.linalg.lstsq(L[:, [0,1,2,3,4]], rho, rcond=None)[0]
    b = np.zeros(4)
    if b5[0] < 0: b[0] = np.sqrt(-b5[0]); b[1] = np.sqrt(-b5[2]) if b5[2] < 0 else 0.0
    else: b[0] = np.sqrt(b5[0]); b[1] = np.sqrt(b5[2]) if b5[2] > 0 else 0.0
    if b5[1] < 0: b[0] = -b[0]
    b[2] = b5[3]/b[0]
    cands.append(R_and_t(gauss_newton(b)))
    N = int(np.argmin([c[2] for c in cands]))
    return cands[N][0], cands[N][1], N, [c[2] for c in cands]

from visual_slam_b200 import synth
K = synth.ICL_K
Kmat = np.array([[K[0],0,K[2]],[0,K[1],K[3]],[0,0,1.0]])
rng = np.random.default_rng(0)
X, uv, rvec, tvec, _ = synth.tracking_pair(5, 400)
X = X.astype(np.float64); uv = uv.astype(np.float64)
dR=[]; dt=[]; Ns=[]
for trial in range(300):
    idx = rng.choice(400, 5, replace=False)
    ok, rv, tv = cv2.solvePnP(X[idx].reshape(-1,1,3), uv[idx].reshape(-1,1,2), Kmat, None, flags=cv2.SOLVEPNP_EPNP)
    Rc = cv2.Rodrigues(rv)[0]
    for canon in (False, True):
        R, t, N, errs = epnp(X[idx], uv[idx], K, canon)
        if canon:
            dR.append(np.abs(R-Rc).max()); dt.append(np.abs(t-tv.ravel()).max()); Ns.append(N)
        else:
            d0 = np.abs(R-Rc).max()
dR=np.array(dR); dt=np.array(dt)
for thr in (1e-9,1e-6,1e-3,1e-1):
    print(thr, (dR<thr).mean(), (dt<thr).mean())
print(np.bincount(Ns))
print("---- normalised coordinates like cv2 (K = I), and basis dependence")
dA=[]; dB=[]; dC=[]
for trial in range(300):
    idx = rng.choice(400, 5, replace=False)
    ok, rv, tv = cv2.solvePnP(X[idx].reshape(-1,1,3), uv[idx].reshape(-1,1,2), Kmat, None, flags=cv2.SOLVEPNP_EPNP)
    Rc = cv2.Rodrigues(rv)[0]
    un = np.stack([(uv[idx,0]-K[2])/K[0], (uv[idx,1]-K[3])/K[1]],1)
    R1,t1,N1,_ = epnp(X[idx], un, (1,1,0,0), False)
    R2,t2,N2,_ = epnp(X[idx], un, (1,1,0,0), True)
    dA.append(np.abs(R1-Rc).max()); dB.append(np.abs(R2-Rc).max()); dC.append(np.abs(R1-R2).max())
dA=np.array(dA); dB=np.array(dB); dC=np.array(dC)
for thr in (1e-9,1e-6,1e-3):
    print(thr, 'eigh-basis vs cv2', (dA<thr).mean(), 'canon vs cv2', (dB<thr).mean(), 'eigh vs canon', (dC<thr).mean())
print("---- control: 50-point samples (no degenerate null space), restatement vs cv2")
dA = []
for trial in range(100):
    idx = rng.choice(400, 50, replace=False)
    ok, rv, tv = cv2.solvePnP(X[idx].reshape(-1, 1, 3), uv[idx].reshape(-1, 1, 2), Kmat, None, flags=cv2.SOLVEPNP_EPNP)
    Rc = cv2.Rodrigues(rv)[0]
    un = np.stack([(uv[idx, 0] - K[2]) / K[0], (uv[idx, 1] - K[3]) / K[1]], 1)
    R1, t1, N1, _ = epnp(X[idx], un, (1, 1, 0, 0), False)
    dA.append(max(np.abs(R1 - Rc).max(), np.abs(t1 - tv.ravel()).max()))
dA = np.array(dA)
for thr in (1e-9, 1e-6, 1e-3):
    print(thr, 'restatement vs cv2', (dA < thr).mean())
