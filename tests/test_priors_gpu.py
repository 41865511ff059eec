"""GPU parity of the monocular data-driven priors through the C ABI: cpe_prior_cost_f64,
cpe_prior_assemble_f64, cpe_band_solve_general_f64 and the monocular LM loop, against oracle.priors / oracle.lm with
the frozen fitted parameters of tests/golden/priors_ref.npz.  FP64 tolerance 1e-9 relative (north star)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def ctx(golden_dir):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    from cheetah_pose_estimation_b200 import engine
    from cheetah_pose_estimation_b200.priors import Priors
    g = np.load(os.path.join(golden_dir, "priors_ref.npz"))
    pr = {k: g[k] for k in ("gmm_w", "gmm_mu", "gmm_cov", "ar_W", "ar_b", "ar_var")}
    return engine.Model("cuda:0"), Priors(pr, "cuda:0"), pr, g


def _problem(gold, S=2, N=12, C=1, seed=31):
    """Synthetic monocular problem whose trajectory is the fixture's (so that the GMM has mass there): the synthetic
    camera ring is moved onto the fixture track and the detections are the oracle's projections of the (perturbed)
    fixture trajectory plus pixel noise, all gates on."""
    from cheetah_pose_estimation_b200 import synthetic
    from oracle import adapters
    prob = synthetic.make_problem(S=S, N=N, C=C, seed=seed, projector=adapters.projector)
    rng = np.random.default_rng(seed)
    q = np.stack([gold["q"][:N] + rng.normal(0, 0.01, (N, 54)) * (s > 0) for s in range(S)])
    for s in range(S):
        shift = q[s, :, 0:3].mean(axis=0) - prob.q_true[s, :, 0:3].mean(axis=0)
        for c in range(C):
            R = prob.cams[s, c, 8:17].reshape(3, 3)
            prob.cams[s, c, 17:20] -= R @ shift
    uv = adapters.projector(q, prob.cams, prob.chain_vec)                 # [S,N,C,24,2]
    prob.meas[..., 0, 0:2] = (uv + rng.normal(0, 2.0, uv.shape)).astype(np.float32)
    prob.meas[..., 0, 2] = 1.0
    prob.q_true = q.copy()
    return prob, q


def test_prior_costs_match_oracle(ctx):
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import priors
    model, P, pr, gold = ctx
    prob, q = _problem(gold, S=3)
    batch = prob.to_batch("cuda:0")
    nf = np.array([12, 9, 3], np.int32)                                   # ragged; 3 < window: no motion rows
    batch.n_frames = torch.as_tensor(nf, device="cuda:0")
    sv = engine.Solver(model, batch, with_solver=False, priors=P)
    out = sv.evaluate(torch.as_tensor(q, device="cuda:0"))
    got = out["prior_cost"].cpu().numpy()
    assert np.allclose(got[0], gold["costs"], rtol=RTOL)                  # the fixture's known answer
    for s in range(3):
        ref = priors.prior_costs(pr, q[s, :nf[s]])
        assert np.allclose(got[s], ref, rtol=RTOL, atol=1e-9), (s, got[s], ref)


def test_prior_assemble_adds_gradient_and_majoriser(ctx):
    import torch
    from cheetah_pose_estimation_b200 import engine, skeleton
    from oracle import priors
    model, P, pr, gold = ctx
    prob, q = _problem(gold, S=2)
    ind = skeleton.independent_coords()
    qd = torch.as_tensor(q, device="cuda:0")
    damp = torch.tensor([1e-3, 5e-2], dtype=torch.float64, device="cuda:0")
    base = engine.Solver(model, prob.to_batch("cuda:0"))
    base.assemble(qd, damp)
    sv = engine.Solver(model, prob.to_batch("cuda:0"), priors=P)
    sv.assemble(qd, damp)
    torch.cuda.synchronize()
    dS = (sv.Sblk - base.Sblk).cpu().numpy()
    dgh = (sv.ghat - base.ghat).cpu().numpy()
    dg = (sv.g - base.g).cpu().numpy()
    dD = (sv.diagH - base.diagH).cpu().numpy()
    for s in range(2):
        gp, Hp = priors.pose_grad_hess(pr, q[s])
        gm, _ = priors.motion_grad_hess(pr, q[s])
        gref = gp + gm.reshape(-1, 54)
        assert not np.any(np.delete(gref, ind, axis=1))
        sc = np.abs(gref).max()
        assert np.abs(dgh[s] - gref[:, ind]).max() <= 1e-8 * sc          # difference of two large sums
        assert np.abs(dg[s] - gref).max() <= 1e-8 * sc
        Href = Hp[:, ind][:, :, ind]
        dref = np.einsum("kaa->ka", Href)
        want = Href + float(damp[s]) * np.einsum("ka,ab->kab", dref, np.eye(28))
        assert np.abs(dS[s] - want).max() <= 1e-8 * np.abs(want).max()
        assert np.abs(dD[s][:, ind] - dref).max() <= 1e-8 * np.abs(dref).max()


@pytest.mark.parametrize("nf,N", [(12, 12), (7, 9), (3, 4), (1, 3), (2, 2), (5, 6), (6, 6), (0, 3), (40, 40)])
def test_general_band_solve_matches_dense(ctx, nf, N):
    import ctypes as C
    import torch
    from cheetah_pose_estimation_b200 import _lib, engine
    model, P, pr, gold = ctx
    rng = np.random.default_rng(nf)
    S_ = 2
    A = rng.normal(size=(S_, N, 28, 28))
    Sblk = np.einsum("snab,sncb->snac", A, A) + 5.0 * np.eye(28)          # SPD blocks
    ghat = rng.normal(size=(S_, N, 28))
    h = 1.0 / 120.0
    band = P.mband(nf, h, N) * 1e-9                                       # keep the blocks well conditioned
    dev = "cuda:0"
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
    Sd, gd, Md = t(Sblk), t(ghat), t(band)
    Lb = torch.empty(S_, N * 28, P.bw + 1, dtype=torch.float64, device=dev)
    delta = torch.full((S_, N, 28), 7.0, dtype=torch.float64, device=dev)
    status = torch.zeros(S_, dtype=torch.int32, device=dev)
    batch = engine.Batch(n_frames=torch.full((S_,), nf, dtype=torch.int32, device=dev),
                         h=torch.full((S_,), h, dtype=torch.float64, device=dev),
                         meas=torch.zeros(S_, N, 1, 24, 1, 4, device=dev), cams=torch.zeros(S_, 1, 24, dtype=torch.float64, device=dev),
                         chain_vec=torch.zeros(S_, 24, 5, 3, dtype=torch.float64, device=dev))
    bs = batch.struct()
    rc = model.lib.cpe_band_solve_general_f64(C.byref(bs), Sd.data_ptr(), gd.data_ptr(), Md.data_ptr(), C.c_int64(0),
                                              C.c_int32(P.bw), Lb.data_ptr(), delta.data_ptr(), status.data_ptr(), None)
    _lib.check(rc, "cpe_band_solve_general_f64")
    torch.cuda.synchronize()
    assert int(status.sum()) == 0
    n = nf * 28
    M = np.zeros((n, n))
    for i in range(n):
        for d in range(min(i, P.bw) + 1):
            M[i, i - d] = band[i, d]
    M = M + np.tril(M, -1).T
    got = delta.cpu().numpy()
    for s in range(S_):
        Afull = M.copy()
        for k in range(nf):
            Afull[k * 28:(k + 1) * 28, k * 28:(k + 1) * 28] += Sblk[s, k]
        if nf > 0:
            ref = np.linalg.solve(Afull, -ghat[s, :nf].reshape(-1)).reshape(nf, 28)
            assert np.abs(got[s, :nf] - ref).max() <= 1e-9 * np.abs(ref).max()
        assert not np.any(got[s, nf:])


def test_monocular_lm_step_matches_oracle(ctx):
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, lm
    model, P, pr, gold = ctx
    prob, q = _problem(gold, S=2, N=10)
    sv = engine.Solver(model, prob.to_batch("cuda:0"), priors=P)
    damp = torch.tensor([1e-3, 1e-1], dtype=torch.float64, device="cuda:0")
    delta, q_try, frame_pred = sv.lm_step(torch.as_tensor(q, device="cuda:0"), damp)
    torch.cuda.synchronize()
    assert int(sv.status.sum()) == 0
    for s in range(2):
        pb = adapters.from_synthetic(prob, s)
        d_ref, pred_ref, _ = lm.lm_step(pb, q[s], np.zeros((10, 26)), np.zeros((10, 24)), float(damp[s]), pr=pr)
        got = delta[s].cpu().numpy()
        assert np.abs(got - d_ref).max() <= 1e-6 * max(1.0, np.abs(d_ref).max())
        assert abs(float(frame_pred[s].sum()) - pred_ref) <= 1e-6 * abs(pred_ref)


def test_monocular_solve_converges_like_oracle(ctx):
    """Config 3 in miniature: one camera, priors on; final merit within 0.5 % of the oracle LM on the same inputs."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, lm
    model, P, pr, gold = ctx
    prob, q = _problem(gold, S=2, N=12)
    q0 = q + np.random.default_rng(3).normal(0, 0.03, q.shape)
    sv = engine.Solver(model, prob.to_batch("cuda:0"), priors=P)
    res = sv.solve(torch.as_tensor(q0, device="cuda:0"), max_iter=25, n_outer=2)
    torch.cuda.synchronize()
    for s in range(2):
        pb = adapters.from_synthetic(prob, s)
        ref = lm.solve(pb, q0[s], max_iter=25, n_outer=2, pr=pr)
        f_ref = ref["cost_terms"][:2].sum() + ref["prior_costs"].sum()
        f_gpu = float(res["cost"][s, :2].sum() + res["prior_cost"][s].sum())
        assert abs(f_gpu - f_ref) <= 5e-3 * abs(f_ref), (f_gpu, f_ref)
        assert np.allclose(res["prior_cost"][s].cpu().numpy(), __import__("oracle.priors", fromlist=["x"]).prior_costs(
            pr, res["q"][s].cpu().numpy()), rtol=1e-8, atol=1e-8)
