"""Edge cases of the hot path on the GPU: empty and very short sequences, frames and sequences without a single usable
detection, and -- at the sweep's frame count -- properties that need no oracle (world-translation invariance, batch
permutation)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-9


@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    from cheetah_pose_estimation_b200 import engine
    return engine.Model("cuda:0")


def _problem(S, N, C, seed):
    import torch
    from cheetah_pose_estimation_b200 import synthetic
    prob = synthetic.make_problem(S=S, N=N, C=C, seed=seed)
    rng = np.random.default_rng(seed + 100)
    q = prob.q_true + rng.normal(0, 0.05, prob.q_true.shape)
    return prob, q, torch.as_tensor(q, device="cuda:0")


def test_empty_and_short_sequences(gpu):
    """n_frames = 0, 1, 2, 3 (no third difference exists below 4 frames) next to a full sequence."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, cost, lm
    prob, q, qd = _problem(5, 6, 6, 41)
    nf = np.array([0, 1, 2, 3, 6], np.int32)
    prob.n_frames = nf
    batch = prob.to_batch("cuda:0")
    batch.n_frames = torch.as_tensor(nf, device="cuda:0")
    sv = engine.Solver(gpu, batch)
    c = sv.evaluate(qd)["cost"].cpu().numpy().copy()
    assert not np.any(c[0])
    damp = torch.full((5,), 1e-2, dtype=torch.float64, device="cuda:0")
    delta, q_try, _ = sv.lm_step(qd, damp)
    d = delta.cpu().numpy().copy()
    assert np.isfinite(d).all() and not np.any(d[0]) and int(sv.status.sum()) == 0
    for s in range(1, 5):
        pb = adapters.from_synthetic(prob, s)
        assert np.allclose(c[s], cost.cost_terms(pb, q[s, :nf[s]]), rtol=RTOL, atol=1e-9)
        d_ref, _, _ = lm.lm_step(pb, q[s, :nf[s]], np.zeros((nf[s], 26)), np.zeros((nf[s], 24)), 1e-2)
        assert np.abs(d[s, :nf[s]] - d_ref).max() <= 1e-7 * max(1.0, np.abs(d_ref).max())
        assert not np.any(d[s, nf[s]:])


def test_frames_and_sequences_without_detections(gpu):
    """All gates off in one frame / in a whole sequence: every entry there contributes the constant rho(0) = -0.213 of the
    reference's redescending loss (acinoset_misc.py:2001-2015; the slack itself stays proj - meas) and nothing to the
    step, which stays finite (motion term, joint rows and damping keep the system positive definite); both as the oracle."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, cost, lm
    prob, q, qd = _problem(2, 8, 6, 43)
    prob.meas[0, 2, ..., 2] = 0.0
    prob.meas[1, ..., 2] = 0.0
    sv = engine.Solver(gpu, prob.to_batch("cuda:0"))
    c = sv.evaluate(qd)["cost"].cpu().numpy().copy()
    rho0 = c[1, 0] / (8 * 6 * 24 * 2)
    assert abs(rho0 + 0.2132) < 1e-3, rho0
    damp = torch.full((2,), 1e-2, dtype=torch.float64, device="cuda:0")
    d = sv.lm_step(qd, damp)[0].cpu().numpy().copy()
    assert np.isfinite(d).all() and int(sv.status.sum()) == 0
    for s in range(2):
        pb = adapters.from_synthetic(prob, s)
        assert np.allclose(c[s], cost.cost_terms(pb, q[s]), rtol=RTOL, atol=1e-9)
        d_ref, _, _ = lm.lm_step(pb, q[s], np.zeros((8, 26)), np.zeros((8, 24)), 1e-2)
        assert np.abs(d[s] - d_ref).max() <= 1e-7 * max(1.0, np.abs(d_ref).max())


def test_sweep_size_properties(gpu):
    """500-frame sequences (the sweep's length), 6 cameras: (i) moving the world origin (base positions and camera
    translations together) leaves every cost term unchanged up to rounding, in both flavours; (ii) a sequence's costs and LM
    step do not depend on its position in the batch (bit-identical under permutation)."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    S, N, C = 48, 500, 6
    batch, q_true, q_init = synthetic.make_device_problem(S, N, C, seed=5, device="cuda:0", dlc_native=True)
    g = torch.Generator(device="cuda:0")
    g.manual_seed(1)
    q = (q_true + 0.03 * torch.randn(q_true.shape, generator=g, device="cuda:0", dtype=torch.float64)).contiguous()
    sv = engine.Solver(gpu, batch)
    c0 = sv.evaluate(q)["cost"].clone()
    c0f = sv.evaluate(q, f32=True)["cost"].clone()
    # (i) world translation d: X' = X + d, t' = t - R d
    d = torch.tensor([3.0, -2.0, 0.25], dtype=torch.float64, device="cuda:0")
    q2 = q.clone()
    q2[..., 0:3] += d
    b2 = batch.window(0, S)
    cams = batch.cams.clone()
    R = cams[..., 8:17].reshape(S, C, 3, 3)
    cams[..., 17:20] -= torch.einsum("scij,j->sci", R, d)
    b2.cams = cams.contiguous()
    s2 = engine.Solver(gpu, b2)
    c2 = s2.evaluate(q2)["cost"].clone()
    c2f = s2.evaluate(q2, f32=True)["cost"].clone()
    rel = ((c2 - c0).abs() / c0.abs().clamp_min(1.0)).max(dim=0).values.tolist()
    relf = ((c2f - c0f).abs() / c0f.abs().clamp_min(1.0)).max(dim=0).values.tolist()
    # the constant-acceleration term is a third difference of positions divided by h^3: rounding of x + d is amplified
    assert rel[0] < 1e-12 and rel[1] < 1e-9 and rel[2] < 1e-12 and rel[3] < 1e-12, rel
    assert relf[0] < 1e-6 and relf[1] < 1e-9 and relf[2] < 1e-12 and relf[3] < 1e-12, relf
    # (ii) permutation of the batch
    perm = torch.randperm(S, generator=torch.Generator().manual_seed(3)).to("cuda:0")
    bp = engine.Batch(n_frames=batch.n_frames[perm].contiguous(), h=batch.h[perm].contiguous(),
                      meas=batch.meas[perm].contiguous(), cams=batch.cams[perm].contiguous(),
                      chain_vec=batch.chain_vec[perm].contiguous(), dlc_thresh=batch.dlc_thresh)
    sp = engine.Solver(gpu, bp)
    qp = q[perm].contiguous()
    assert torch.equal(sp.evaluate(qp)["cost"], c0[perm])
    damp = torch.full((S,), 1e-2, dtype=torch.float64, device="cuda:0")
    dp = sp.lm_step(qp, damp)[0].clone()
    d0 = sv.lm_step(q, damp)[0].clone()
    assert torch.equal(dp, d0[perm])


def test_full_sweep_size_wave_split_and_checksum(gpu):
    """The whole C4 configuration of BASELINE.json -- 4096 sequences x 500 frames x 6 cameras on one GPU (12.3 M
    frame.views) -- with properties that need no oracle: (i) one launch over all sequences and eight launches over
    waves of 512 give bit-identical per-sequence costs (a sequence's result does not depend on its batch), (ii) the sum
    of the wave checksums equals the checksum of the whole, (iii) tiled sequences (the generator repeats 32 base
    trajectories with fresh noise) have pairwise different measurement costs but identical joint / inequality terms
    at the shared q, (iv) the FP32 flavour agrees with FP64 to 1e-6 on the measurement cost of every sequence."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    S, N, C, WAVE = 4096, 500, 6, 512
    batch, q_true, _ = synthetic.make_device_problem(S, N, C, seed=0, device="cuda:0", dlc_native=True)
    q = q_true.contiguous()
    sv = engine.Solver(gpu, batch, with_solver=False)
    whole = sv.evaluate(q)["cost"].clone()
    torch.cuda.synchronize()
    assert whole.shape == (S, 4) and bool(torch.isfinite(whole).all())
    parts = []
    for s0 in range(0, S, WAVE):
        w = engine.Solver(gpu, batch.window(s0, s0 + WAVE), with_solver=False)
        parts.append(w.evaluate(q[s0:s0 + WAVE])["cost"].clone())
    waves = torch.cat(parts)
    assert torch.equal(waves, whole)
    chk = torch.stack([p.sum(dim=0) for p in parts]).sum(dim=0)
    assert bool(((chk - whole.sum(dim=0)).abs() <= 1e-12 * whole.sum(dim=0).abs()).all())
    # sequences s and s + 32 share trajectory and cameras (q_true) but not the detection noise
    assert bool((whole[:32, 0] != whole[32:64, 0]).all())
    assert torch.equal(whole[:32, 2:], whole[32:64, 2:]) and torch.equal(whole[:32, 1], whole[32:64, 1])
    w32 = sv.evaluate(q, f32=True)["cost"]
    assert float(((w32[:, 0] - whole[:, 0]).abs() / whole[:, 0].abs()).max()) < 1e-6


def test_shutter_delay_estimation_on_ragged_and_tiny_sequences(gpu):
    """solve_shutter next to sequences that are too short to carry a velocity or an acceleration (0, 1, 2 frames) and a
    ragged tail: everything stays finite, the delays stay inside +-1/fps with camera 1 at 0, frames beyond n_frames are
    left untouched, and the full-length sequence still lowers its cost against tau = 0."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    prob, q, qd = _problem(5, 24, 6, 47)
    nf = np.array([0, 1, 2, 17, 24], np.int32)
    batch = prob.to_batch("cuda:0")
    batch.n_frames = torch.as_tensor(nf, device="cuda:0")
    sv = engine.Solver(gpu, batch)
    c0 = sv.evaluate(qd)["cost"].sum(dim=1).clone()
    res = sv.solve_shutter(qd, max_iter=15, n_outer=2)
    torch.cuda.synchronize()
    tau, qs = res["tau"].cpu().numpy(), res["q"].cpu().numpy()
    h = np.asarray(prob.h)
    assert np.isfinite(tau).all() and np.isfinite(qs).all() and np.isfinite(res["cost"].cpu().numpy()).all()
    assert np.all(tau[:, 0] == 0.0) and np.all(np.abs(tau) <= h[:, None])
    assert not np.any(tau[0]) and not np.any(tau[1])          # no frames / one frame: no rates, nothing to estimate from
    for s in range(5):
        assert np.array_equal(qs[s, nf[s]:], q[s, nf[s]:])
    assert int((res["status"] & 5).sum()) == 0              # no failed pivot, nothing non-finite (15 iterations may run out)
    assert float(res["cost"][4, 0:2].sum()) < float(c0[4])


def test_exit_status_flags_and_nan_gating(gpu):
    """cpe_solve status beyond "failed pivot" (SURVEY.md section 5): 0 = converged, CPE_STATUS_MAXITER when the last round
    ran out of iterations, CPE_STATUS_NONFINITE for a NaN in q; a NaN detection in the DLC-native float3 layout is gated on
    the device (weight 0) and leaves every result finite."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=3, N=10, C=4, seed=31)
    q0 = torch.as_tensor(prob.q_init, device="cuda:0")
    sv = engine.Solver(gpu, prob.to_batch("cuda:0"))
    res = sv.solve(q0, max_iter=40, n_outer=3)
    assert res["status"].tolist() == [0, 0, 0]
    res = sv.solve(q0, max_iter=2, n_outer=1)
    assert res["status"].tolist() == [engine.STATUS_MAXITER] * 3
    qn = q0.clone()
    qn[1, 4, 7] = float("nan")
    res = sv.solve(qn, max_iter=5, n_outer=1)
    st = res["status"].tolist()
    assert st[1] & engine.STATUS_NONFINITE and not (st[0] & engine.STATUS_NONFINITE) and not (st[2] & engine.STATUS_NONFINITE)
    # NaN in DLC-native detections: same result as the same detection with likelihood 0
    m3 = np.concatenate([prob.meas[..., 0:2], np.where(prob.meas[..., 2:3] > 0, 0.9, 0.1)], -1).astype(np.float32)
    m_nan, m_off = m3.copy(), m3.copy()
    m_nan[0, 3, 1, 5, 0, 0] = np.nan
    m_off[0, 3, 1, 5, 0, 2] = 0.0
    outs = []
    for m in (m_nan, m_off):
        b = engine.Batch.from_numpy(m, prob.cams, prob.chain_vec, prob.h, dlc_native=True)
        s2 = engine.Solver(gpu, b)
        o = s2.evaluate(q0, resid=True)
        s2.assemble(q0, torch.full((3,), 1e-3, dtype=torch.float64, device="cuda:0"))
        outs.append((o["cost"].clone(), s2.Sblk.clone(), s2.ghat.clone()))
    assert torch.isfinite(outs[0][0]).all() and torch.isfinite(outs[0][1]).all()
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)
