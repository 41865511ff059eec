"""GPU parity tests proper: the CUDA path (through the C ABI of libcpe_b200.so) against the CPU oracle on the same
seeded inputs and against the golden fixtures.  Tolerances: FP64 per-residual relative error <= 1e-9 (north star)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def _rel(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    from cheetah_pose_estimation_b200 import engine
    return engine.Model("cuda:0")


def _setup(model, S=3, N=12, C=6, seed=1):
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=S, N=N, C=C, seed=seed)
    batch = prob.to_batch("cuda:0")
    solver = engine.Solver(model, batch)
    rng = np.random.default_rng(seed + 100)
    q = prob.q_true + rng.normal(0, 0.05, prob.q_true.shape)
    return prob, batch, solver, q, torch.as_tensor(q, device="cuda:0")


def test_eval_matches_oracle(gpu):
    import torch
    from oracle import adapters, cost, kinematics
    prob, batch, solver, q, qd = _setup(gpu)
    rng = np.random.default_rng(5)
    lam_eq = rng.normal(0, 3, (batch.S, batch.N, 26))
    lam_in = rng.normal(0, 3, (batch.S, batch.N, gpu.n_ineq))
    batch.lam_eq = torch.as_tensor(lam_eq, device="cuda:0")
    batch.lam_in = torch.as_tensor(lam_in, device="cuda:0")
    out = solver.evaluate(qd, pose=True, resid=True, constraints=True)
    torch.cuda.synchronize()
    for s in range(batch.S):
        pb = adapters.from_synthetic(prob, s)
        assert _rel(out["pose"][s].cpu().numpy(), kinematics.fk(q[s], pb.chain_link, pb.chain_vec)) < 1e-13
        r = cost.residuals(pb, q[s])
        got = out["resid"][s].cpu().numpy()
        # per-residual relative error (floor 1 px on the denominator)
        assert (np.abs(got - r) / np.maximum(np.abs(r), 1.0)).max() < RTOL
        ref = cost.cost_terms(pb, q[s], lam_eq[s], lam_in[s])
        assert np.allclose(out["cost"][s].cpu().numpy(), ref, rtol=RTOL, atol=1e-9)
        _, _, c, v = cost.al_terms(pb, q[s], lam_eq[s], lam_in[s])
        assert np.abs(out["c_eq"][s].cpu().numpy() - c).max() < 1e-13
        assert np.abs(out["v_in"][s].cpu().numpy() - v).max() < 1e-12


@pytest.mark.parametrize("fname,animal", [("clip_jules_run1_1.npz", "jules"), ("clip_phantom_run.npz", "phantom")])
def test_eval_real_clip_known_answers(gpu, golden_dir, fname, animal):
    """Shipped optimum of a real clip (cameras/detections recovered): residuals equal the saved
    `meas_err`, the measurement cost equals the reference-function value, FK equals the saved `positions`."""
    import torch
    from cheetah_pose_estimation_b200 import engine, skeleton
    g = np.load(os.path.join(golden_dir, fname))
    N, C = g["detections"].shape[:2]
    lengths, radii, _ = skeleton.animal(animal)
    _, cvec = skeleton.marker_chains(lengths, radii)
    meas = np.zeros((1, N, C, 24, 1, 4), np.float32)
    meas[0, :, :, :, 0, 0:2] = g["detections"]
    meas[..., 2] = 1.0
    cams = np.zeros((1, C, 24))
    cams[0, :, :20] = g["cams"]
    cams[0, :, 20] = 1.0
    batch = engine.Batch.from_numpy(meas, cams, cvec[None], [1.0 / float(g["fps"])])
    solver = engine.Solver(gpu, batch, with_solver=False)
    out = solver.evaluate(torch.as_tensor(g["q"][None].copy(), device="cuda:0"), pose=True, resid=True)
    torch.cuda.synchronize()
    assert np.abs(out["pose"][0].cpu().numpy() - g["positions"]).max() < 1e-13
    assert np.abs(out["resid"][0, :, :, :, 0, :].cpu().numpy() - g["meas_err"][..., 0]).max() < 2e-4
    cost = out["cost"][0].cpu().numpy()
    assert abs(cost[0] - float(g["known_meas_cost_all_on"])) / float(g["known_meas_cost_all_on"]) < 1e-6
    dd = g["ddq"][3:] - g["ddq"][2:-1]
    from cheetah_pose_estimation_b200.skeleton import motion_weights
    assert abs(cost[1] - float(np.sum(motion_weights() * dd * dd))) / cost[1] < 1e-8


def test_jacobian_matches_oracle(gpu):
    """Packed chain-sparse Jacobian d(u,v)/dq against the oracle's dense one (entries outside the pattern are
    structurally zero in the oracle)."""
    import torch
    from cheetah_pose_estimation_b200 import skeleton
    from oracle import adapters, cameras, kinematics
    for (S, N, C) in ((2, 5, 4), (2, 7, 6), (1, 3, 1)):          # partial tail chunk, 6 cams, monocular
        prob, batch, solver, q, qd = _setup(gpu, S=S, N=N, C=C)
        out = solver.evaluate(qd, jac=True, resid=True)
        torch.cuda.synchronize()
        assert out["jac"].shape == (S, N, C, 540)
        dense = skeleton.jacobian_to_dense(out["jac"].cpu().numpy())                      # [S,N,C,24,2,54]
        for s in range(batch.S):
            pb = adapters.from_synthetic(prob, s)
            pose = kinematics.fk(q[s], pb.chain_link, pb.chain_vec)
            G = kinematics.fk_jacobian(q[s], pb.chain_link, pb.chain_vec)                 # [N,24,3,54]
            P = cameras.project_jacobian(pose[:, None], pb.cams[None, :, None, :])        # [N,C,24,2,3]
            J = np.einsum("ncldi,nlia->nclda", P, G)                                      # dense [N,C,24,2,54]
            scale = np.abs(J).max(axis=-1, keepdims=True)
            assert (np.abs(dense[s] - J) / np.maximum(scale, 1.0)).max() < RTOL


@pytest.mark.parametrize("with_mult", [False, True])
def test_normal_equations_match_oracle(gpu, with_mult):
    import torch
    from oracle import adapters, cost
    prob, batch, solver, q, qd = _setup(gpu, S=2, N=9, C=6, seed=7)
    rng = np.random.default_rng(11)
    if with_mult:
        # push a few coordinates outside the inequality box so that PHR rows are active
        q[:, :, 19] += 2.5
        qd = torch.as_tensor(q, device="cuda:0")
        lam_eq = rng.normal(0, 5, (batch.S, batch.N, 26))
        lam_in = rng.normal(0, 5, (batch.S, batch.N, gpu.n_ineq))
        batch.lam_eq = torch.as_tensor(lam_eq, device="cuda:0")
        batch.lam_in = torch.as_tensor(lam_in, device="cuda:0")
    else:
        lam_eq = np.zeros((batch.S, batch.N, 26))
        lam_in = np.zeros((batch.S, batch.N, gpu.n_ineq))
    damp = torch.full((batch.S,), 1e-3, dtype=torch.float64, device="cuda:0")
    Hraw = solver.assemble(qd, damp, raw=True)
    torch.cuda.synchronize()
    for s in range(batch.S):
        pb = adapters.from_synthetic(prob, s)
        H, g = cost.frame_normal_equations(pb, q[s], lam_eq[s], lam_in[s])
        g = g + cost.motion_gradient(pb, q[s])
        assert _rel(Hraw[s].cpu().numpy(), H) < RTOL
        assert _rel(solver.g[s].cpu().numpy(), g) < RTOL


def test_lm_step_matches_oracle(gpu):
    import torch
    from oracle import adapters, lm
    prob, batch, solver, q, qd = _setup(gpu, S=2, N=14, C=6, seed=9)
    damp = torch.tensor([1e-3, 5e-2], dtype=torch.float64, device="cuda:0")
    delta, q_try, frame_pred = solver.lm_step(qd, damp)
    torch.cuda.synchronize()
    assert int(solver.status.sum()) == 0
    for s in range(batch.S):
        pb = adapters.from_synthetic(prob, s)
        d_ref, pred_ref, _ = lm.lm_step(pb, q[s], np.zeros((batch.N, 26)), np.zeros((batch.N, 24)), float(damp[s]))
        assert _rel(delta[s].cpu().numpy(), d_ref) < 1e-7
        assert abs(float(frame_pred[s].sum()) - pred_ref) <= 1e-7 * abs(pred_ref)
        assert _rel(q_try[s].cpu().numpy(), q[s] + d_ref) < 1e-7


def test_ragged_frames_and_determinism(gpu):
    """Padded frames are ignored, and a sequence's result does not depend on its batch neighbours."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    prob, batch, solver, q, qd = _setup(gpu, S=3, N=10, C=6, seed=13)
    out_full = solver.evaluate(qd)["cost"].cpu().numpy().copy()
    nf = np.array([10, 7, 4], np.int32)
    batch2 = prob.to_batch("cuda:0")
    batch2.n_frames = torch.as_tensor(nf, device="cuda:0")
    s2 = engine.Solver(gpu, batch2)
    c2 = s2.evaluate(qd)["cost"].cpu().numpy().copy()
    assert np.array_equal(c2[0], out_full[0])
    from oracle import adapters, cost, lm
    prob.n_frames = nf
    for s in (1, 2):
        pb = adapters.from_synthetic(prob, s)
        assert np.allclose(c2[s], cost.cost_terms(pb, q[s, :nf[s]]), rtol=RTOL, atol=1e-9)
    damp = torch.full((3,), 1e-2, dtype=torch.float64, device="cuda:0")
    delta, _, _ = s2.lm_step(qd, damp)
    d = delta.cpu().numpy().copy()
    for s in (1, 2):
        pb = adapters.from_synthetic(prob, s)
        d_ref, _, _ = lm.lm_step(pb, q[s, :nf[s]], np.zeros((nf[s], 26)), np.zeros((nf[s], 24)), 1e-2)
        assert _rel(d[s, :nf[s]], d_ref) < 1e-7
    # single-sequence batch gives bit-identical output for that sequence
    one = engine.Batch(n_frames=batch2.n_frames[1:2].contiguous(), h=batch2.h[1:2].contiguous(), meas=batch2.meas[1:2].contiguous(),
                       cams=batch2.cams[1:2].contiguous(), chain_vec=batch2.chain_vec[1:2].contiguous())
    s1 = engine.Solver(gpu, one)
    d1, _, _ = s1.lm_step(qd[1:2].contiguous(), damp[:1].contiguous())
    assert np.array_equal(d1[0].cpu().numpy()[:nf[1]], d[1, :nf[1]])


def test_pinhole_and_pairwise_layout(gpu):
    """cam_model = pinhole (kinetic dataset) and W = 3 hypotheses per marker."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    from oracle import adapters, cameras, cost
    prob = synthetic.make_problem(S=1, N=6, C=4, seed=21)
    rng = np.random.default_rng(3)
    meas3 = np.repeat(prob.meas, 3, axis=4).copy()
    meas3[..., 1, 0:2] += rng.normal(0, 3, meas3[..., 1, 0:2].shape).astype(np.float32)
    meas3[..., 2, 0:2] += rng.normal(0, 5, meas3[..., 2, 0:2].shape).astype(np.float32)
    cams = prob.cams.copy()
    cams[:, :, 4:8] *= 0.2
    cams[:, 2:, 20] = 0.6
    batch = engine.Batch.from_numpy(meas3, cams, prob.chain_vec, prob.h, cam_model=engine.PINHOLE)
    solver = engine.Solver(gpu, batch, with_solver=False)
    q = prob.q_true + rng.normal(0, 0.03, prob.q_true.shape)
    out = solver.evaluate(torch.as_tensor(q, device="cuda:0"), resid=True)
    pb = adapters.problem_from_arrays(meas3[0], cams[0], prob.chain_link, prob.chain_vec[0], prob.h[0],
                                      cam_model=cameras.PINHOLE)
    r = cost.residuals(pb, q[0])
    got = out["resid"][0].cpu().numpy()
    assert (np.abs(got - r) / np.maximum(np.abs(r), 1.0)).max() < RTOL
    assert np.allclose(out["cost"][0].cpu().numpy(), cost.cost_terms(pb, q[0]), rtol=RTOL, atol=1e-9)


def test_solver_converges_like_oracle(gpu):
    """Converged trajectories: final cost within 0.5 % of the oracle solver's and MPJPE within 1 mm of it."""
    import torch
    from oracle import adapters, kinematics, lm
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=2, N=30, C=6, seed=4)
    batch = prob.to_batch("cuda:0")
    solver = engine.Solver(gpu, batch)
    res = solver.solve(torch.as_tensor(prob.q_init, device="cuda:0"), max_iter=30, n_outer=3)
    torch.cuda.synchronize()
    qg = res["q"].cpu().numpy()
    for s in range(2):
        pb = adapters.from_synthetic(prob, s)
        ref = lm.solve(pb, prob.q_init[s], max_iter=30, n_outer=3)
        f_ref = ref["cost_terms"][:2].sum()
        f_gpu = res["cost"][s, :2].sum().item()
        assert abs(f_gpu - f_ref) / f_ref < 5e-3, (f_gpu, f_ref)
        pg = kinematics.fk(qg[s], pb.chain_link, pb.chain_vec)
        pr = kinematics.fk(ref["q"], pb.chain_link, pb.chain_vec)
        pt = kinematics.fk(prob.q_true[s], pb.chain_link, pb.chain_vec)

        def mpjpe(a, b):
            a = a - a.mean(axis=1, keepdims=True)
            b = b - b.mean(axis=1, keepdims=True)
            return 1000 * np.linalg.norm(a - b, axis=-1).mean()
        assert abs(mpjpe(pg, pt) - mpjpe(pr, pt)) < 1.0, (mpjpe(pg, pt), mpjpe(pr, pt))
        assert np.abs(res["c_eq"][s].cpu().numpy()).max() < 1e-4


# FP32 flavour (north star: "<= 1e-4 in FP32").  Tolerances, written out:
#   residual  |r32 - r64| <= 1e-4 * max(|r64|, 1 px)          (the FP32 floor at ~1000 px offsets is ~6e-5 px)
#   Jacobian  |J32 - J64| <= 1e-4 * max(row max |J64|, 1)
#   cost      relative 1e-4 on the measurement term; the FP64 terms (model, joints, inequalities) stay at 1e-9
RTOL32 = 1e-4


@pytest.mark.parametrize("S,N,C", [(2, 7, 6), (2, 5, 4), (1, 9, 1)])
def test_f32_flavour_matches_oracle(gpu, S, N, C):
    import torch
    from cheetah_pose_estimation_b200 import skeleton
    from oracle import adapters, cameras, cost, kinematics
    prob, batch, solver, q, qd = _setup(gpu, S=S, N=N, C=C, seed=17)
    out = solver.evaluate(qd, pose=True, resid=True, jac=True, constraints=True, f32=True)
    torch.cuda.synchronize()
    assert out["resid"].dtype == torch.float32 and out["jac"].dtype == torch.float32
    dense = skeleton.jacobian_to_dense(out["jac"].cpu().numpy().astype(np.float64))
    for s in range(S):
        pb = adapters.from_synthetic(prob, s)
        pose = kinematics.fk(q[s], pb.chain_link, pb.chain_vec)
        assert _rel(out["pose"][s].cpu().numpy(), pose) < 1e-13              # FK front end is FP64 in both flavours
        r = cost.residuals(pb, q[s])
        got = out["resid"][s].cpu().numpy().astype(np.float64)
        assert (np.abs(got - r) / np.maximum(np.abs(r), 1.0)).max() < RTOL32
        ref = cost.cost_terms(pb, q[s])
        c = out["cost"][s].cpu().numpy()
        assert abs(c[0] - ref[0]) <= RTOL32 * abs(ref[0])
        assert np.allclose(c[1:], ref[1:], rtol=RTOL, atol=1e-9)
        G = kinematics.fk_jacobian(q[s], pb.chain_link, pb.chain_vec)
        P = cameras.project_jacobian(pose[:, None], pb.cams[None, :, None, :])
        J = np.einsum("ncldi,nlia->nclda", P, G)
        scale = np.abs(J).max(axis=-1, keepdims=True)
        assert (np.abs(dense[s] - J) / np.maximum(scale, 1.0)).max() < RTOL32


def test_f32_flavour_far_from_origin(gpu):
    """The scaled path must not lose accuracy when the animal is 40 m from the world origin (cameras moved along)."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, cost
    prob, batch, solver, q, qd = _setup(gpu, S=1, N=6, C=6, seed=23)
    shift = np.array([40.0, -25.0, 3.0])
    q2 = q.copy()
    q2[..., 0:3] += shift
    cams = prob.cams.copy()
    for c in range(cams.shape[1]):
        R = cams[0, c, 8:17].reshape(3, 3)
        cams[0, c, 17:20] -= R @ shift
    batch2 = engine.Batch.from_numpy(prob.meas, cams, prob.chain_vec, prob.h)
    s2 = engine.Solver(gpu, batch2, with_solver=False)
    out = s2.evaluate(torch.as_tensor(q2, device="cuda:0"), resid=True, f32=True)
    pb = adapters.from_synthetic(prob, 0)
    r = cost.residuals(pb, q[0])                                             # same residuals as the unshifted problem
    got = out["resid"][0].cpu().numpy().astype(np.float64)
    assert (np.abs(got - r) / np.maximum(np.abs(r), 1.0)).max() < RTOL32


def test_dlc_native_layout_equals_pregated(gpu):
    """CPE_MEAS_F32X3 (x, y, likelihood; strict `> dlc_thresh` on the device) gives bit-identical costs,
    residuals and LM steps to the host-gated float4 records."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    prob, batch, solver, q, qd = _setup(gpu, S=2, N=8, C=6, seed=29)
    rng = np.random.default_rng(2)
    lik = rng.random(prob.meas.shape[:-1]).astype(np.float32)
    lik[0, 0, 0, :4, 0] = 0.5                                                # exactly at the threshold: gate off
    m3 = np.concatenate([prob.meas[..., 0:2], lik[..., None]], -1).astype(np.float32)
    m4 = prob.meas.copy()
    m4[..., 2] = (lik > 0.5).astype(np.float32)
    b3 = engine.Batch.from_numpy(m3, prob.cams, prob.chain_vec, prob.h, dlc_native=True)
    b4 = engine.Batch.from_numpy(m4, prob.cams, prob.chain_vec, prob.h)
    assert b3.meas_layout == engine.MEAS_F32X3 and b4.meas_layout == engine.MEAS_F32X4
    s3, s4 = engine.Solver(gpu, b3), engine.Solver(gpu, b4)
    o3 = s3.evaluate(qd, resid=True, jac=True)
    o4 = s4.evaluate(qd, resid=True, jac=True)
    for k in ("cost", "resid", "jac"):
        assert torch.equal(o3[k], o4[k]), k
    damp = torch.full((2,), 1e-2, dtype=torch.float64, device="cuda:0")
    d3 = s3.lm_step(qd, damp)[0].clone()
    d4 = s4.lm_step(qd, damp)[0].clone()
    assert torch.equal(d3, d4)


def test_compact_layout_equals_dlc_native_and_pregated(gpu):
    """CPE_MEAS_F32X2G (float2 u, v + one gate byte, 9 B per detection; the likelihood / NaN test done once by
    Batch.compact) against the DLC-native float3 and the host-gated float4 records: costs, residuals, Jacobians, the
    normal equations' LM step and a whole solve are bit-identical; NaN detections and the exact-threshold case included;
    W = 3 (pairwise layout) as well."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    prob, batch, solver, q, qd = _setup(gpu, S=3, N=8, C=6, seed=31)
    rng = np.random.default_rng(4)
    for W in (1, 3):
        meas = np.repeat(prob.meas, W, axis=4) if W > 1 else prob.meas
        lik = rng.random(meas.shape[:-1]).astype(np.float32)
        lik[0, 0, 0, :4, 0] = 0.5                                            # exactly at the threshold: gate off
        m3 = np.concatenate([meas[..., 0:2], lik[..., None]], -1).astype(np.float32)
        m3[1, 2, 1, 5, 0, 0] = np.nan                                        # NaN coordinate: gate off, finite stand-in
        m3[2, 3, 4, 7, 0, 2] = np.nan                                        # NaN likelihood: gate off
        b3 = engine.Batch.from_numpy(m3, prob.cams, prob.chain_vec, prob.h, dlc_native=True)
        b2 = b3.compact()
        assert b2.meas_layout == engine.MEAS_F32X2G and b2.meas.shape[-1] == 2 and b2.meas_gate.dtype == torch.uint8
        assert b2.compact() is b2
        m4 = np.zeros(meas.shape[:-1] + (4,), np.float32)
        m4[..., 0:2] = np.where(np.isnan(m3[..., 0:2]).any(-1, keepdims=True), 0.0, m3[..., 0:2])
        with np.errstate(invalid="ignore"):
            m4[..., 2] = ((m3[..., 2].astype(np.float64) > 0.5) & ~np.isnan(m3[..., 0:2]).any(-1)).astype(np.float32)
        b4 = engine.Batch.from_numpy(m4, prob.cams, prob.chain_vec, prob.h)
        assert np.array_equal(b2.meas_gate.cpu().numpy(), m4[..., 2].astype(np.uint8))
        assert int(b2.meas_gate[1, 2, 1, 5, 0]) == 0 and int(b2.meas_gate[2, 3, 4, 7, 0]) == 0
        # gates packed 8 per byte (CPE_MEAS_F32X2B): numpy's little-endian packbits of the byte gates, from either layout
        bb = b3.compact(bits=True)
        assert bb.meas_layout == engine.MEAS_F32X2B and tuple(bb.meas_gate.shape) == (3, 8 * 6 * 24 * W // 8)
        want_bits = np.packbits(b2.meas_gate.cpu().numpy().reshape(3, -1), axis=1, bitorder="little")
        assert np.array_equal(bb.meas_gate.cpu().numpy(), want_bits)
        assert torch.equal(b2.compact(bits=True).meas_gate, bb.meas_gate) and bb.compact(bits=True) is bb
        sb = engine.Solver(gpu, bb)
        ob = sb.evaluate(qd, resid=True, jac=True)
        s2, s3, s4 = (engine.Solver(gpu, b) for b in (b2, b3, b4))
        o2, o3, o4 = (s.evaluate(qd, resid=True, jac=True) for s in (s2, s3, s4))
        for k in ("cost", "resid", "jac"):
            assert torch.equal(ob[k], o3[k]), (W, k, "bit-packed vs DLC-native")
        for k in ("cost", "resid", "jac"):
            assert torch.equal(o2[k], o3[k]), (W, k, "compact vs DLC-native")
            assert torch.equal(o2[k], o4[k]), (W, k, "compact vs host-gated float4")
        assert bool(torch.isfinite(o2["cost"]).all())
        damp = torch.full((3,), 1e-2, dtype=torch.float64, device="cuda:0")
        d2, d3 = s2.lm_step(qd, damp)[0].clone(), s3.lm_step(qd, damp)[0].clone()
        assert torch.equal(d2, d3) and torch.equal(sb.lm_step(qd, damp)[0], d3), W
        q0 = torch.as_tensor(prob.q_init, device="cuda:0")
        r2, r3 = s2.solve(q0, max_iter=8, n_outer=2), s3.solve(q0, max_iter=8, n_outer=2)
        for k in ("q", "cost", "counters", "status"):
            assert torch.equal(r2[k], r3[k]), (W, k)
    # the f32 flavour reads the same records
    o2, o3 = s2.evaluate(qd, resid=True, jac=True, f32=True), s3.evaluate(qd, resid=True, jac=True, f32=True)
    for k in ("cost", "resid", "jac"):
        assert torch.equal(o2[k], o3[k]), k
    # double-precision detections (kinetic dataset) are not silently narrowed
    b64 = engine.Batch.from_numpy(m4, prob.cams, prob.chain_vec, prob.h, meas_dtype=torch.float64)
    with pytest.raises(ValueError, match="float32"):
        b64.compact()
    # a compact batch without its gate array is refused at the boundary
    import ctypes as C
    from cheetah_pose_estimation_b200 import _lib
    bs = b2.struct()
    bs.meas_gate = None
    eo = _lib.CpeEvalOut()
    fc = torch.zeros(3, 8, 4, dtype=torch.float64, device="cuda:0")
    eo.frame_cost = fc.data_ptr()
    rc = gpu.lib.cpe_eval_f64(C.byref(gpu.struct), C.byref(bs), C.c_void_p(qd.data_ptr()), C.byref(eo), None)
    assert rc == -3                                                          # CPE_E_NULL


def test_skipping_converged_sequences_changes_nothing(gpu):
    """cpe_batch.active: sequences that converged are skipped by every kernel; results must be bit-identical."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=4, N=16, C=6, seed=41)
    q0 = torch.as_tensor(prob.q_init, device="cuda:0")
    res = []
    for skip in (True, False):
        sv = engine.Solver(gpu, prob.to_batch("cuda:0"))
        res.append(sv.solve(q0, max_iter=12, n_outer=2, skip_converged=skip))
    torch.cuda.synchronize()
    for k in ("q", "cost", "counters"):
        assert torch.equal(res[0][k], res[1][k]), k


def test_solve_split_is_bit_identical(gpu):
    """Two sub-batches on two streams (K3 of one overlapping K2 of the other) == one batch."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=5, N=16, C=6, seed=43)
    q0 = torch.as_tensor(prob.q_init, device="cuda:0")
    one = engine.Solver(gpu, prob.to_batch("cuda:0")).solve(q0, max_iter=10, n_outer=2)
    two = engine.solve_split(gpu, prob.to_batch("cuda:0"), q0, n_split=2, max_iter=10, n_outer=2)
    torch.cuda.synchronize()
    for k in ("q", "cost", "counters", "c_eq"):
        assert torch.equal(one[k], two[k]), k


def test_square_loss_mode_matches_oracle(gpu):
    """CPE_LOSS_SQUARE = measurement_cost(hand_labeled_data=True): plain (w slack)^2 without the camera multiplier
    (acinoset_misc.py:471-474); cost and normal equations against the oracle."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    from oracle import adapters, cost
    prob = synthetic.make_problem(S=2, N=7, C=4, seed=47)
    cams = prob.cams.copy()
    cams[:, 2:, 20] = 0.6                                                  # must be ignored in this mode
    batch = engine.Batch.from_numpy(prob.meas, cams, prob.chain_vec, prob.h, loss_mode=engine.LOSS_SQUARE)
    sv = engine.Solver(gpu, batch)
    rng = np.random.default_rng(1)
    q = prob.q_true + rng.normal(0, 0.02, prob.q_true.shape)
    qd = torch.as_tensor(q, device="cuda:0")
    out = sv.evaluate(qd)
    out32 = engine.Solver(gpu, batch, with_solver=False).evaluate(qd, f32=True)["cost"].cpu().numpy()
    damp = torch.full((2,), 1e-3, dtype=torch.float64, device="cuda:0")
    Hraw = sv.assemble(qd, damp, raw=True)
    torch.cuda.synchronize()
    for s in range(2):
        pb = adapters.problem_from_arrays(prob.meas[s], cams[s], prob.chain_link, prob.chain_vec[s], prob.h[s],
                                          square_loss=True)
        ref = cost.cost_terms(pb, q[s])
        assert np.allclose(out["cost"][s].cpu().numpy(), ref, rtol=RTOL, atol=1e-9)
        assert abs(out32[s, 0] - ref[0]) <= RTOL32 * ref[0]
        H, g = cost.frame_normal_equations(pb, q[s])
        g = g + cost.motion_gradient(pb, q[s])
        assert _rel(Hraw[s].cpu().numpy(), H) < RTOL
        assert _rel(sv.g[s].cpu().numpy(), g) < RTOL


def test_shutter_delay_evaluation(gpu, golden_dir):
    """Shutter-delay term of the measurement model (acinoset_misc.py:273-288): K1's residuals, cost and Jacobian with
    per-camera delays against the oracle and against the reference-function golden; d slack / d tau from the packed
    Jacobian (the estimation of the delays is tested below)."""
    import torch
    from cheetah_pose_estimation_b200 import engine, skeleton
    from oracle import cameras, cost, kinematics, tables
    g = np.load(os.path.join(golden_dir, "shutter_ref.npz"))
    clip = np.load(os.path.join(golden_dir, "clip_jules_run1_1.npz"))
    N, Cn = clip["detections"].shape[:2]
    lengths, radii, _ = skeleton.animal("jules")
    cl, cvec = skeleton.marker_chains(lengths, radii)
    meas = np.zeros((2, N, Cn, 24, 1, 4), np.float32)
    meas[:, :, :, :, 0, 0:2] = clip["detections"]
    meas[..., 2] = 1.0
    cams = np.zeros((2, Cn, 24))
    cams[:, :, :20] = clip["cams"]
    cams[:, :, 20] = 1.0
    h = 1.0 / float(clip["fps"])
    batch = engine.Batch.from_numpy(meas, cams, np.stack([cvec, cvec]), [h, h])
    tau = np.stack([g["tau"], -0.5 * g["tau"]])                     # second sequence: different delays, same clip
    dev = dict(device="cuda:0", dtype=torch.float64)
    batch.shutter_tau = torch.as_tensor(tau, **dev).contiguous()
    batch.base_dq = torch.as_tensor(np.stack([clip["dq"][:, 0:3]] * 2), **dev).contiguous()
    batch.base_ddq = torch.as_tensor(np.stack([clip["ddq"][:, 0:3]] * 2), **dev).contiguous()
    solver = engine.Solver(gpu, batch, with_solver=False)
    qd = torch.as_tensor(np.stack([clip["q"]] * 2), **dev).contiguous()
    out = solver.evaluate(qd, pose=True, resid=True, jac=True)
    torch.cuda.synchronize()
    got = out["resid"][:, :, :, :, 0, :].cpu().numpy()
    # golden from the reference's own functions (10 frames of sequence 0)
    ref = g["slack"]
    assert (np.abs(got[0][g["frames"]] - ref) / np.maximum(np.abs(ref), 1.0)).max() < RTOL
    # the saved positions stay the undelayed FK (acinoset_opt.py:354: positions_arr[0], camera 1 has tau = 0)
    assert np.abs(out["pose"][0].cpu().numpy() - clip["positions"]).max() < 1e-13
    r_pw, _ = tables.uncertainty_models()
    wl = np.broadcast_to(1.0 / r_pw[0][None, None, :, None], (N, Cn, 24, 1))      # all-on likelihood mask
    om = np.concatenate([clip["detections"].astype(float), wl], -1)[:, :, :, None, :]
    pb = cost.Problem(meas=om, cams=clip["cams"], chain_link=cl, chain_vec=cvec, h=h)
    dense = skeleton.jacobian_to_dense(out["jac"].cpu().numpy())
    Jt = engine.shutter_jacobian(out["jac"], batch).cpu().numpy()
    for s in range(2):
        args = (tau[s], clip["dq"][:, 0:3], clip["ddq"][:, 0:3])
        r = cost.residuals(pb, clip["q"], *args)[:, :, :, 0, :]
        assert (np.abs(got[s] - r) / np.maximum(np.abs(r), 1.0)).max() < RTOL
        mc = cost.measurement_cost(pb, clip["q"], *args)
        assert abs(float(out["cost"][s, 0]) - mc) <= RTOL * mc
        pts = cost.delayed_pose(pb, clip["q"], *args)                                   # [N,C,24,3]
        G = kinematics.fk_jacobian(clip["q"], cl, cvec)                                 # [N,24,3,54]
        P = cameras.project_jacobian(pts, pb.cams[None, :, None, :])                    # [N,C,24,2,3]
        J = np.einsum("ncldi,nlia->nclda", P, G)
        scale = np.abs(J).max(axis=-1, keepdims=True)
        assert (np.abs(dense[s] - J) / np.maximum(scale, 1.0)).max() < RTOL
        Jo = cost.shutter_jacobian(pb, clip["q"], *args)
        assert np.abs(Jt[s] - Jo).max() <= RTOL * np.abs(Jo).max()
    # FP32 flavour honours the delay as well
    out32 = solver.evaluate(qd, resid=True, f32=True)
    torch.cuda.synchronize()
    r32 = out32["resid"][:, :, :, :, 0, :].cpu().numpy()
    # 1e-4 relative; floor 4 px: the FP32 image coordinates of this clip reach ~1700 px (ulp 1.2e-4 px), so residuals
    # below a few px carry an absolute rounding error of a few 1e-4 px whatever the delay is
    assert (np.abs(r32 - got) / np.maximum(np.abs(got), 4.0)).max() < 1e-4
    out32_0 = engine.Solver(gpu, engine.Batch.from_numpy(meas, cams, np.stack([cvec, cvec]), [h, h]),
                            with_solver=False).evaluate(qd, resid=True, f32=True)
    torch.cuda.synchronize()
    assert np.abs(out32_0["resid"][:, :, 1:, :, 0, :].cpu().numpy() - r32[:, :, 1:]).max() > 1.0   # the delay did act


def _delayed_problem(S, N, C, seed, tau_scale, noise=1.0):
    """Synthetic sequences whose detections follow the delayed measurement model exactly (oracle.delayed_pose with
    the q-only rates) plus Gaussian pixel noise; returns (prob, tau_true[S,C])."""
    from cheetah_pose_estimation_b200 import synthetic
    from oracle import adapters, cameras, cost
    prob = synthetic.make_problem(S=S, N=N, C=C, seed=seed)
    rng = np.random.default_rng(seed + 7)
    tau = rng.uniform(-1.0, 1.0, (S, C)) * tau_scale * np.asarray(prob.h)[:, None]
    tau[:, 0] = 0.0
    for s in range(S):
        pb = adapters.from_synthetic(prob, s)
        v, a = cost.base_rates(prob.q_true[s], pb.h)
        pts = cost.delayed_pose(pb, prob.q_true[s], tau[s], v, a)
        uv = cameras.project(pts, pb.cams[None, :, None, :], pb.cam_model)
        prob.meas[s, :, :, :, 0, 0:2] = (uv + rng.normal(0, noise, uv.shape)).astype(np.float32)
        prob.meas[s, :, :, :, 0, 2] = 1.0
    return prob, tau


def test_shutter_delay_gradient_is_exact(gpu):
    """d cost / d q (incl. the terms through dq_base, ddq_base of frames n, n-1, n-2) and d cost / d tau from K2's
    per-camera blocks against central differences of the oracle's delayed cost."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import adapters, cost
    S, N, C = 4, 8, 4
    prob, tau_true = _delayed_problem(S, N, C, 61, 0.5, noise=3.0)
    nf = np.array([8, 8, 3, 2], np.int32)        # the first-node rules differ for 3- and 2-frame sequences
    prob.n_frames = nf
    batch = prob.to_batch("cuda:0")
    batch.n_frames = torch.as_tensor(nf, device="cuda:0")
    sv = engine.Solver(gpu, batch)
    rng = np.random.default_rng(3)
    q = prob.q_true + rng.normal(0, 0.03, prob.q_true.shape)
    tau = tau_true + rng.uniform(-0.2, 0.2, tau_true.shape) * np.asarray(prob.h)[:, None]
    lam_eq = rng.normal(0, 1, (S, N, 26))
    lam_in = rng.normal(0, 1, (S, N, gpu.n_ineq))
    batch.lam_eq = torch.as_tensor(lam_eq, device="cuda:0")
    batch.lam_in = torch.as_tensor(lam_in, device="cuda:0")
    qd, td = torch.as_tensor(q, device="cuda:0"), torch.as_tensor(tau, device="cuda:0")
    damp = torch.full((S,), 1e-3, dtype=torch.float64, device="cuda:0")
    g_tau, curv = sv.assemble_delayed(qd, td, damp, reduced=False)
    torch.cuda.synchronize()
    g = sv.g.cpu().numpy()
    g_tau, curv = g_tau.cpu().numpy(), curv.cpu().numpy()
    c_gpu = sv.evaluate(qd)["cost"].cpu().numpy()
    for s in range(S):
        pb = adapters.from_synthetic(prob, s)
        pb.mu_eq, pb.mu_in = batch.mu_eq, batch.mu_in
        n = int(nf[s])
        qs, le, li = q[s, :n], lam_eq[s, :n], lam_in[s, :n]
        f = lambda qq, tt: cost.delayed_cost_terms(pb, qq, tt, le, li).sum()
        assert np.allclose(c_gpu[s], cost.delayed_cost_terms(pb, qs, tau[s], le, li), rtol=RTOL, atol=1e-9)
        for (k, p) in [(0, 0), (0, 2), (1, 1), (2, 0), (3, 2), (5, 1), (7, 0), (7, 2), (4, 5), (2, 30), (6, 4), (1, 0)]:
            if k >= n:
                continue
            d = np.zeros_like(qs)
            eps = 1e-6
            d[k, p] = eps
            fd = (f(qs + d, tau[s]) - f(qs - d, tau[s])) / (2 * eps)
            assert abs(fd - g[s, k, p]) <= 2e-5 * max(1.0, abs(fd)), (s, k, p, fd, g[s, k, p])
        for c in range(C):
            d = np.zeros(C)
            eps = 2e-7                     # seconds; tau is O(1e-3) and the cost O(1e5): a smaller step drowns in rounding
            d[c] = eps
            fd = (f(qs, tau[s] + d) - f(qs, tau[s] - d)) / (2 * eps)
            assert abs(fd - g_tau[s, c]) <= 2e-4 * max(1.0, abs(fd)), (s, c, fd, g_tau[s, c])
            assert curv[s, c] > 0


def test_shutter_delay_estimation_recovers_delays(gpu):
    """solve_shutter on sequences generated with known per-camera delays: the delays come back, camera 1 stays at 0, and
    the final cost is below that of the plain solve (tau = 0) of the same data."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    S, N, C = 3, 40, 6
    prob, tau_true = _delayed_problem(S, N, C, 71, 0.4, noise=1.0)
    h = np.asarray(prob.h)
    rng = np.random.default_rng(5)
    q0 = torch.as_tensor(prob.q_true + rng.normal(0, 0.02, prob.q_true.shape), device="cuda:0")
    plain = engine.Solver(gpu, prob.to_batch("cuda:0")).solve(q0, max_iter=30, n_outer=3)
    res = engine.Solver(gpu, prob.to_batch("cuda:0")).solve_shutter(q0, max_iter=30, n_outer=3)
    torch.cuda.synchronize()
    tau = res["tau"].cpu().numpy()
    assert np.all(tau[:, 0] == 0.0) and np.all(np.abs(tau) <= h[:, None])
    err = np.abs(tau - tau_true) / h[:, None]
    assert err.max() < 0.1, (err.max(), tau / h[:, None], tau_true / h[:, None])
    c_plain = plain["cost"][:, 0:2].sum(dim=1).cpu().numpy()
    c_delay = res["cost"][:, 0:2].sum(dim=1).cpu().numpy()
    assert np.all(c_delay < c_plain), (c_delay, c_plain)
    assert float(res["c_eq"].abs().max()) < 1e-3


def test_host_evaluator_modes_match_the_resident_evaluation(gpu):
    """engine.HostEvaluator (cpe_eval_host_f64, the end-to-end call of bench.py) in its three modes -- all inputs in /
    detections resident, q only / residuals copied back -- against the device-resident evaluation, over two waves."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=5, N=9, C=3, seed=23)
    batch = prob.to_batch("cuda:0")
    q = torch.as_tensor(prob.q_init, device="cuda:0")
    ref = engine.Solver(gpu, batch, with_solver=False).evaluate(q, resid=True, jac=True)     # same kernel flavour (JAC)
    h_q = q.cpu().pin_memory()
    h_meas = batch.meas.cpu().pin_memory()
    for resident, resid_out in ((False, False), (True, False), (False, True)):
        hev = engine.HostEvaluator(gpu, batch, 5, 3, resident_meas=resident, resid_out=resid_out)
        cost = hev(h_q, None if resident else h_meas)
        assert torch.equal(cost, ref["cost"].cpu())
        h2d, d2h = hev.bytes_per_call()
        assert h2d == 5 * 9 * 54 * 8 + (0 if resident else h_meas.numel() * h_meas.element_size())
        if resid_out:
            assert torch.equal(hev.h_resid, ref["resid"].cpu()) and d2h == 5 * 4 * 8 + ref["resid"].numel() * 8
    # the compact layout through its own entry point (cpe_eval_host_uvg_f64): float2 records + gate bytes cross the bus
    for bits in (False, True):
        cb = batch.compact(bits=bits)
        h_uv, h_gate = cb.meas.cpu().pin_memory(), cb.meas_gate.cpu().pin_memory()
        n_det = 5 * 9 * 3 * 24
        for resident, resid_out in ((False, False), (True, False), (False, True)):
            hev = engine.HostEvaluator(gpu, cb, 5, 3, resident_meas=resident, resid_out=resid_out)
            cost = hev(h_q, None if resident else h_uv, None if resident else h_gate)
            assert torch.equal(cost, ref["cost"].cpu())
            h2d, _ = hev.bytes_per_call()
            assert h2d == 5 * 9 * 54 * 8 + (0 if resident else n_det * 8 + (n_det // 8 if bits else n_det))
            if resid_out:
                assert torch.equal(hev.h_resid, ref["resid"].cpu())


def test_graph_replay_of_the_lm_iteration_is_bit_identical(gpu):
    """`Solver.solve(graph=True)`: the launches of one LM iteration captured per multiplier round into a CUDA graph and
    replayed -- same kernels, order and arguments, so q, cost, counters and status equal the eager loop bit for bit
    (ragged batch; a second solve on the same solver re-captures against its new buffers)."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=5, N=14, C=4, seed=52)
    prob.n_frames[:] = (14, 9, 14, 4, 11)
    q0 = torch.as_tensor(prob.q_init, device="cuda:0")
    ref = engine.Solver(gpu, prob.to_batch("cuda:0")).solve(q0, max_iter=12, n_outer=3)
    sv = engine.Solver(gpu, prob.to_batch("cuda:0"))
    for _ in range(2):
        res = sv.solve(q0, max_iter=12, n_outer=3, graph=True)
        torch.cuda.synchronize()
        for k in ("q", "cost", "counters", "status", "c_eq", "v_in"):
            assert torch.equal(res[k], ref[k]), k
        assert res["lockstep_iterations"] == ref["lockstep_iterations"]


def test_extra_rounds_only_touch_sequences_that_are_not_yet_feasible(gpu):
    """`solve(extra_outer=k, feas_tol=...)`: after the scheduled multiplier rounds, further rounds run for the sequences
    whose constraints still exceed feas_tol and for nobody else.  With a schedule that is cut short (2 rounds) and a
    tolerance between the violations of the batch, the sequences already below it come out bit-identical to the plain
    solve and the others end below the tolerance (or at least strictly better)."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob = synthetic.make_problem(S=6, N=14, C=4, seed=61)
    prob.n_frames[:] = (14, 9, 14, 6, 11, 14)
    q0 = torch.as_tensor(prob.q_init, device="cuda:0")
    kw = dict(max_iter=12, n_outer=2, mu_schedule=(1e4, 1e5))             # the extra rounds stay at the last penalty
    base = engine.Solver(gpu, prob.to_batch("cuda:0")).solve(q0, **kw)
    viol = base["c_eq"].abs().amax(dim=(1, 2))
    tol = float(viol.sort().values[2:4].mean())                          # three sequences below, three above
    sv = engine.Solver(gpu, prob.to_batch("cuda:0"))
    res = sv.solve(q0, extra_outer=6, feas_tol=(tol, 1.0), **kw)
    torch.cuda.synchronize()
    ok = viol < tol
    assert int(ok.sum()) == 3 and 1 <= sv.extra_rounds <= 6
    for k in ("q", "cost", "c_eq", "v_in", "counters", "status"):
        assert torch.equal(res[k][ok], base[k][ok]), k
    after = res["c_eq"].abs().amax(dim=(1, 2))
    assert bool((after[~ok] < viol[~ok]).all()) and bool((res["counters"][~ok, 0] > base["counters"][~ok, 0]).all())
    assert bool((after[~ok] < tol).all()) or sv.extra_rounds == 6
    # nothing to do: no extra round runs and the result is the plain one
    sv2 = engine.Solver(gpu, prob.to_batch("cuda:0"))
    res2 = sv2.solve(q0, extra_outer=6, feas_tol=(1.0, 1.0), **kw)
    assert sv2.extra_rounds == 0 and all(torch.equal(res2[k], base[k]) for k in ("q", "cost", "status"))


def test_redamp_after_a_rejected_step_equals_a_full_assembly(gpu):
    """cpe_assemble_out.ent / redamp: K2 keeps the undamped entries of every frame; for sequences flagged as rejected it
    reads them back and only repeats the damped elimination and the Schur gather.  (i) at the kernel level the re-damped
    S, ghat, Y, y equal a full assembly at the new damping bit for bit, flagged and unflagged sequences alike; (ii) a
    whole solve with the shortcut equals the solve without it bit for bit, on a problem whose LM loop does reject steps."""
    import torch
    from cheetah_pose_estimation_b200 import engine, synthetic
    prob, batch, solver, q, qd = _setup(gpu, S=4, N=10, C=6, seed=71)
    rng = np.random.default_rng(9)
    batch.lam_eq = torch.as_tensor(rng.normal(0, 3, (4, 10, 26)), device="cuda:0")
    batch.lam_in = torch.as_tensor(rng.normal(0, 3, (4, 10, gpu.n_ineq)), device="cuda:0")
    f64 = dict(dtype=torch.float64, device="cuda:0")
    d0 = torch.full((4,), 1e-3, **f64)
    d1 = torch.tensor([1e-3, 7e-2, 1e-3, 3.0], **f64)                      # sequences 1 and 3 were "rejected": new damping
    solver.assemble(qd, d0)                                                # full: fills ent
    assert float(solver.ent.abs().max()) > 0
    solver.state[3].copy_(torch.tensor([0.5, -1.0, 0.0, -1.0], **f64))
    solver._redamp_on = True
    solver.assemble(qd, d1)                                                # 1 and 3 re-damp, 0 and 2 assemble in full
    solver._redamp_on = False
    got = {k: getattr(solver, k).clone() for k in ("Sblk", "ghat", "Y", "y", "g", "diagH")}
    ref = engine.Solver(gpu, batch, redamp=False)
    ref.assemble(qd, d1)
    for k in got:
        assert torch.equal(got[k], getattr(ref, k)), k
    assert not torch.equal(got["Sblk"][1], engine.Solver(gpu, batch, redamp=False).Sblk[1])      # (and it is not all zeros)
    # (ii) whole solves
    prob2 = synthetic.make_problem(S=5, N=14, C=4, seed=52)
    prob2.n_frames[:] = (14, 9, 14, 4, 11)
    q0 = torch.as_tensor(prob2.q_init, device="cuda:0")
    a = engine.Solver(gpu, prob2.to_batch("cuda:0"), redamp=False).solve(q0, max_iter=15, n_outer=3)
    sb = engine.Solver(gpu, prob2.to_batch("cuda:0"))
    assert sb._redamp and sb.ent is not None
    b = sb.solve(q0, max_iter=15, n_outer=3)
    assert bool((a["counters"][:, 0] > a["counters"][:, 1]).any()), "no rejected step: the test does not exercise the path"
    for k in ("q", "cost", "counters", "status", "c_eq", "v_in"):
        assert torch.equal(a[k], b[k]), k
    # with the iteration captured in a CUDA graph as well
    c = engine.Solver(gpu, prob2.to_batch("cuda:0")).solve(q0, max_iter=15, n_outer=3, graph=True)
    assert torch.equal(a["q"], c["q"]) and torch.equal(a["counters"], c["counters"])
