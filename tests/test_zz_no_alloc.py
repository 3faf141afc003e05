"""GPU: the no-allocation entry point vslam_pcg_solve_static (no
counterpart in the reference) against the existing vslam_pcg_solve, through the C ABI, on two calls in a row.

The file name sorts last on purpose: a failure here must not hide other tests.
Every buffer the entry point touches is allocated by the TEST before the calls; the entry point itself allocates
nothing (its scratch is a file-scope device array in csrc/ba.cu) and does not synchronise.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _system(n, seed, spd=True):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(n, n))
    S = A @ A.T + n * np.eye(n)
    if not spd:
        S[n // 2, n // 2] = -abs(S[n // 2, n // 2])          # an indefinite diagonal block: status 1, x = 0
    return S, rng.normal(size=n)


def _both(lib, torch, S, b, tol=1e-10, max_iter=5000):
    """(x, status, iters) of the existing entry point and of the static one, same inputs."""
    from visual_slam_b200 import _cabi
    St, bt = torch.from_numpy(S).cuda(), torch.from_numpy(b).cuda()
    n = len(b)
    x_ref = torch.zeros(n, dtype=torch.float64, device="cuda")
    host = (C.c_int32 * 2)()
    _cabi.check(lib.vslam_pcg_solve(_cabi.ptr(St), _cabi.ptr(bt), _cabi.ptr(x_ref), n, tol, max_iter, host,
                                    _cabi.stream_ptr()))
    x_new = torch.full((n,), 7.0, dtype=torch.float64, device="cuda")
    out = torch.full((2,), -5, dtype=torch.int32, device="cuda")
    _cabi.check(lib.vslam_pcg_solve_static(_cabi.ptr(St), _cabi.ptr(bt), _cabi.ptr(x_new), n, tol, max_iter,
                                           _cabi.ptr(out), _cabi.stream_ptr()))
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    return (x_ref.cpu().numpy(), int(host[0]), int(host[1])), (x_new.cpu().numpy(), int(o[0]), int(o[1]))


def test_static_pcg_equals_the_allocating_entry_point_on_consecutive_calls():
    import torch
    from visual_slam_b200 import _cabi
    lib = _cabi.lib()
    assert lib.vslam_pcg_solve_static_max_n() == 1536
    # two calls in a row (and more): a large system first, so that the shared scratch holds stale vectors, flags and
    # barrier counts when the smaller ones run; the small-kernel path (n <= 64), the one-CTA path with the matrix in
    # shared memory (n = 96) and without (n = 300); an indefinite system between two good ones
    for n, seed, spd in ((300, 1, True), (54, 2, True), (96, 3, True), (90, 4, False), (54, 5, True), (6, 6, True)):
        S, b = _system(n, seed, spd)
        (x_ref, st_ref, it_ref), (x_new, st_new, it_new) = _both(lib, torch, S, b)
        assert (st_new, it_new) == (st_ref, it_ref), (n, st_new, it_new, st_ref, it_ref)
        assert st_ref == (0 if spd else 1)
        # same kernels, same order of operations: identical; the tolerance of tests/test_ba_gpu.py::test_pcg_solves_spd_
        # systems (1e-9 of |b| on the residual) is the fallback statement of "equal"
        assert np.array_equal(x_new, x_ref), (n, np.abs(x_new - x_ref).max())
        if spd:
            assert np.linalg.norm(S @ x_new - b) <= 1e-9 * np.linalg.norm(b)
        else:
            assert not x_new.any()
    # the same call twice in a row gives the same bits
    S, b = _system(120, 9)
    a1 = _both(lib, torch, S, b)[1]
    a2 = _both(lib, torch, S, b)[1]
    assert np.array_equal(a1[0], a2[0]) and a1[1:] == a2[1:]


def test_static_pcg_rejects_sizes_beyond_its_scratch_and_bad_arguments():
    import torch
    from visual_slam_b200 import _cabi
    lib = _cabi.lib()
    n = lib.vslam_pcg_solve_static_max_n() + 6
    d = torch.zeros(8, dtype=torch.float64, device="cuda")        # never dereferenced: the size check comes first
    out = torch.zeros(2, dtype=torch.int32, device="cuda")
    rc = lib.vslam_pcg_solve_static(_cabi.ptr(d), _cabi.ptr(d), _cabi.ptr(d), n, 1e-10, 100, _cabi.ptr(out), None)
    assert rc == 6 and lib.vslam_status_string(rc) == b"size exceeds the static scratch capacity of this entry point"
    assert b"1542" in lib.vslam_last_cuda_error() and b"1536" in lib.vslam_last_cuda_error()
    assert lib.vslam_pcg_solve_static(_cabi.ptr(d), _cabi.ptr(d), _cabi.ptr(d), 7, 1e-10, 100, _cabi.ptr(out), None) == 1
    assert lib.vslam_pcg_solve_static(None, _cabi.ptr(d), _cabi.ptr(d), 6, 1e-10, 100, _cabi.ptr(out), None) == 1
    # the largest size it accepts runs (a diagonally dominant system: a handful of iterations)
    n = lib.vslam_pcg_solve_static_max_n()
    S = np.eye(n) * 4.0 + np.diag(np.ones(n - 1), 1) + np.diag(np.ones(n - 1), -1)
    b = np.arange(n, dtype=np.float64) / n
    (x_ref, st_ref, it_ref), (x_new, st_new, it_new) = _both(lib, torch, S, b)
    assert st_new == st_ref == 0 and it_new == it_ref and np.array_equal(x_new, x_ref)
    assert np.linalg.norm(S @ x_new - b) <= 1e-9 * np.linalg.norm(b)     # the solver's criterion: |r|_2 <= 1e-10 |b|_2
