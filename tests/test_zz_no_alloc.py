"""Allocation-free entry point `cpe_eom_slack_noalloc_f64` (include/cpe.h; no counterpart in the reference):
its scratch is a static device array inside the library, so it can be called where
cudaMalloc cannot (CUDA graph capture, tight serving loops).  Called here through the C ABI (ctypes) and compared with the
existing `cpe_eom_slack_f64` -- same kernels, so bit-identical -- on two calls in a row of DIFFERENT shape (the shared
array keeps the previous call's contents), and the size limit is checked.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_FRAMES = 65536          # CPE_NOALLOC_MAX_FRAMES
E_SCRATCH = -4              # CPE_E_SCRATCH


def _batch(_lib, S, N, nf):
    b = _lib.CpeBatch()
    b.S, b.N, b.C, b.W = S, N, 0, 1
    b.n_frames = nf.data_ptr()
    return b


def _both(lib, _lib, torch, rng, S, N, n_frames):
    dev = "cuda:0"
    f64 = dict(dtype=torch.float64, device=dev)
    eom = torch.as_tensor(rng.normal(size=(S, N, 6)), device=dev)
    Bm = torch.as_tensor(rng.normal(size=(S, N, 6, 20)), device=dev)
    nf = torch.as_tensor(np.asarray(n_frames, np.int32), device=dev)
    b = _batch(_lib, S, N, nf)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    work = torch.full((S, N), float("nan"), **f64)
    cost_ref, grad_ref = torch.zeros(S, **f64), torch.zeros(S, N, 20, **f64)
    assert lib.cpe_eom_slack_f64(C.byref(b), p(eom), p(Bm), p(work), p(cost_ref), p(grad_ref), st) == 0
    cost, grad = torch.full((S,), -1.0, **f64), torch.zeros(S, N, 20, **f64)       # outputs are the caller's; no scratch passed
    assert lib.cpe_eom_slack_noalloc_f64(C.byref(b), p(eom), p(Bm), p(cost), p(grad), st) == 0
    torch.cuda.synchronize()
    # the reference value itself, so that "both wrong in the same way" cannot pass
    e = eom.cpu().numpy()
    want = np.array([np.sum(e[s, :n_frames[s]] ** 2) for s in range(S)])
    assert np.allclose(cost_ref.cpu().numpy(), want, rtol=1e-12, atol=0)
    return cost.cpu().numpy(), cost_ref.cpu().numpy(), grad.cpu().numpy(), grad_ref.cpu().numpy()


def test_noalloc_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    import torch
    from cheetah_pose_estimation_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(17)
    # first call: large, every frame used -> fills the shared scratch; second call: smaller and ragged -> must not see it
    for S, N, n_frames in ((7, 300, [300] * 7), (4, 50, [50, 17, 1, 33])):
        cost, cost_ref, grad, grad_ref = _both(lib, _lib, torch, rng, S, N, n_frames)
        assert np.array_equal(cost, cost_ref), (S, N)                  # tolerance of tests/test_kinetic_costs.py: 1e-12;
        assert np.abs(cost - cost_ref).max() <= 1e-12 * np.abs(cost_ref).max()      # same kernels, so in fact identical
        assert np.array_equal(grad, grad_ref)


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    import torch
    from cheetah_pose_estimation_b200 import _lib
    lib = _lib.load()
    S, N = 2, MAX_FRAMES // 2 + 1                                      # S * N = MAX_FRAMES + 2
    nf = torch.full((S,), N, dtype=torch.int32, device="cuda:0")
    b = _batch(_lib, S, N, nf)
    eom = torch.zeros(S, N, 6, dtype=torch.float64, device="cuda:0")
    cost = torch.zeros(S, dtype=torch.float64, device="cuda:0")
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.cpe_eom_slack_noalloc_f64(C.byref(b), p(eom), None, p(cost), None, None)
    assert rc == E_SCRATCH
    with pytest.raises(_lib.CpeError, match="static scratch"):
        _lib.check(rc, "cpe_eom_slack_noalloc_f64")
    # exactly at the limit it runs
    S, N = 2, MAX_FRAMES // 2
    nf = torch.full((S,), N, dtype=torch.int32, device="cuda:0")
    b = _batch(_lib, S, N, nf)
    eom = torch.ones(S, N, 6, dtype=torch.float64, device="cuda:0")
    assert lib.cpe_eom_slack_noalloc_f64(C.byref(b), p(eom), None, p(cost), None, None) == 0
    torch.cuda.synchronize()
    assert cost.tolist() == [6.0 * N, 6.0 * N]
