"""Kinetic path, first slice: rows 0-5 of the equations of motion.  The golden values are the
reference's OWN symbolic model evaluated through sympy.lambdify (oracle/derive_eom.py, run in the build container);
the CUDA code is generated from the same expressions, so this checks the generator, the input order and the kernel."""
import os

import numpy as np
import pytest


def test_fixture_and_parameter_order(golden_dir):
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    assert tuple(g["feet"]) == kinetics.FEET
    assert np.allclose(g["par"][0], kinetics.link_params("jules"), rtol=0, atol=1e-15)
    z, xy = np.arange(4.0), np.arange(16.0).reshape(4, 4) + 10
    assert np.array_equal(kinetics.pack_grf(z, xy), np.concatenate([z, np.arange(16.0) + 10]))
    # free fall sanity of the golden rows: no velocity, no acceleration, no forces -> row 2 = total weight
    # (checked on the reference expressions when the fixture was made: gravity enters only through dEp/dz)
    assert g["eom"].shape == (g["q"].shape[0], 6)


@pytest.mark.gpu
def test_eom_rows_match_reference_expressions(golden_dir):
    import torch
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    n = g["q"].shape[0]
    # every golden row as its own sequence (per-sequence link parameters), plus one batch with N > 1 and ragged frames
    out = kinetics.eom_rows(g["q"][:, None], g["dq"][:, None], g["ddq"][:, None], g["F"][:, None], g["par"])
    got = out[:, 0].cpu().numpy()
    scale = np.maximum(np.abs(g["eom"]), 1.0)
    assert (np.abs(got - g["eom"]) / scale).max() < 1e-9, (np.abs(got - g["eom"]) / scale).max()
    h = n // 2
    out2 = kinetics.eom_rows(g["q"][None, :h], g["dq"][None, :h], g["ddq"][None, :h], g["F"][None, :h], g["par"][:1],
                             n_frames=[h - 3])
    got2 = out2[0].cpu().numpy()
    assert (np.abs(got2[:h - 3] - g["eom"][:h - 3]) / scale[:h - 3]).max() < 1e-9
    assert not np.any(got2[h - 3:])
    # gravity check: at rest with zero forces row 2 (base z) equals the total weight
    z = np.zeros((1, 1, 54))
    q0 = z.copy()
    q0[0, 0, 2] = 0.7
    r = kinetics.eom_rows(q0, z, z, np.zeros((1, 1, 20)), g["par"][:1])[0, 0].cpu().numpy()
    assert abs(r[2] - g["par"][0, 17:34].sum() * 9.81) < 1e-9 and abs(r[0]) < 1e-9 and abs(r[1]) < 1e-9


CONTACTS = ([1, 1, 1, 1], [1, 0, 0, 0], [1, 0, 1, 1], [0, 0, 1, 1], [0, 1, 1, 0], [0, 0, 0, 0], [0, 0, 0, 1])


def test_force_matrix_fixture_is_affine_split(golden_dir):
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    # eom(F) = a - Bm F on the reference expressions themselves
    assert np.abs(g["a"] - np.einsum("nij,nj->ni", g["Bm"], g["F"]) - g["eom"]).max() < 1e-9 * np.abs(g["eom"]).max()


def test_grf_oracle_is_feasible_and_not_worse_than_zero(golden_dir):
    from oracle import grf as ogrf
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    for i, ct in enumerate(CONTACTS):
        F, r = ogrf.solve_frame(g["a"][i], g["Bm"][i], ct)
        assert F.min() >= 0 and F.max() <= ogrf.FMAX and r <= np.linalg.norm(g["a"][i]) + 1e-12
        for f in range(4):
            assert ogrf.MU * F[f] - F[4 + 4 * f:8 + 4 * f].sum() >= -1e-9
            if not ct[f]:
                assert F[f] == 0 and not F[4 + 4 * f:8 + 4 * f].any()


META = {"start_frame": 53, "end_frame": 83,
        "contacts": {"HFL_foot": [[58, 64, 0, "leading"]], "HFR_foot": [[54, 60, 1, "trailing"]],
                     "HBL_foot": [[69, 75, 2, "trailing"], [80, 90, 2, "x"]], "HBR_foot": None}}


def test_contact_mask_follows_the_reference_ranges():
    from cheetah_pose_estimation_b200 import kinetics
    m = kinetics.contact_mask(META)
    assert m.shape == (30, 4)
    for fe in range(30):                                   # literal restatement of acinoset_opt.py:221-229
        for f, name in enumerate(kinetics.FEET):
            on = False
            seqs = META["contacts"].get(name)
            if seqs is not None:
                idx = []
                for s in seqs:
                    idx = idx + list(range(s[0], s[1]))
                on = (53 + fe) in idx
            assert bool(m[fe, f]) == on, (fe, name)


@pytest.mark.gpu
def test_estimate_grf_states_reference_return_shape(golden_dir):
    from cheetah_pose_estimation_b200 import kinetics
    from oracle import grf as ogrf
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    n = 30
    rep = lambda a: np.concatenate([a] * (n // a.shape[0] + 1))[:n + 2]
    st = {"q": rep(g["q"]), "dq": rep(g["dq"]), "ddq": rep(g["ddq"])}
    gz, gxy = kinetics.estimate_grf_states(st, META, "jules")
    assert set(gz) == set(kinetics.FEET) and all(len(v) == n for v in gz.values())
    assert all(len(v) == n and len(v[0]) == 4 for v in gxy.values())
    m = kinetics.contact_mask(META)
    assert not any(gz["HBR_foot"]) and all(gz["HFL_foot"][fe] == 0 for fe in range(n) if not m[fe, 0])
    assert not any(any(gz[f][fe] for f in kinetics.FEET) for fe in range(n) if not m[fe].any())   # :231-236
    fe = 6                                                  # frame 59: both front feet down
    F = np.array([gz[f][fe] for f in kinetics.FEET] + [v for f in kinetics.FEET for v in gxy[f][fe]])
    i = fe % g["q"].shape[0]
    _, ro = ogrf.solve_frame(g["a"][i], g["Bm"][i], m[fe])
    assert abs(np.linalg.norm(g["a"][i] - g["Bm"][i] @ F) - ro) <= 1e-9 * np.linalg.norm(g["a"][i])


@pytest.mark.gpu
def test_estimator_estimate_grf_reads_saved_trajectory(golden_dir, tmp_path):
    """The mirror of CheetahEstimator.estimate_grf: files in the reference's places, same return value."""
    import json
    import pickle
    from types import SimpleNamespace
    from cheetah_pose_estimation_b200 import acinoset_opt, kinetics
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    n = 30
    rep = lambda a: np.concatenate([a] * (n // a.shape[0] + 1))[:n]
    st = {"q": rep(g["q"]), "dq": rep(g["dq"]), "ddq": rep(g["ddq"])}
    os.makedirs(tmp_path / "fte_kinematic")
    os.makedirs(tmp_path / "grf")
    with open(tmp_path / "fte_kinematic" / "fte.pickle", "wb") as f:
        pickle.dump(st, f)
    with open(tmp_path / "grf" / "autogen-contact.json", "w", encoding="utf-8") as f:
        json.dump(META, f)
    est = acinoset_opt.CheetahEstimator("jules", "x", None, SimpleNamespace(data_dir=str(tmp_path)),
                                        SimpleNamespace(cam_idx=None), 1.0, None, True, True)
    gz, gxy = est.estimate_grf()
    gz2, gxy2 = kinetics.estimate_grf_states(st, META, "jules")
    assert gz == gz2 and gxy == gxy2 and len(gz["HFL_foot"]) == n


@pytest.mark.gpu
def test_grf_estimate_matches_oracle_value(golden_dir):
    """Optimal VALUE parity (the minimiser is not unique): | r_gpu - r_oracle | <= 1e-9 ||a|| and never below the oracle's
    by more than that; the GPU forces are feasible to 1e-12 and reproduce their own residual through cpe_eom_f64."""
    from cheetah_pose_estimation_b200 import kinetics
    from oracle import grf as ogrf
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    n = min(g["q"].shape[0], 2 * len(CONTACTS))
    ct = np.array([CONTACTS[i % len(CONTACTS)] for i in range(n)])
    grf, resid = kinetics.estimate_grf(g["q"][:n, None], g["dq"][:n, None], g["ddq"][:n, None], g["par"][:n], ct[:, None])
    F, r = grf[:, 0].cpu().numpy(), resid[:, 0].cpu().numpy()
    for i in range(n):
        Fo, ro = ogrf.solve_frame(g["a"][i], g["Bm"][i], ct[i])
        tol = 1e-9 * np.linalg.norm(g["a"][i])
        assert abs(r[i] - ro) <= tol, (i, r[i], ro, tol)
        assert F[i].min() >= 0 and F[i].max() <= ogrf.FMAX
        for f in range(4):
            assert ogrf.MU * F[i, f] - F[i, 4 + 4 * f:8 + 4 * f].sum() >= -1e-12
            if not ct[i, f]:
                assert F[i, f] == 0 and not F[i, 4 + 4 * f:8 + 4 * f].any()
    rows = kinetics.eom_rows(g["q"][:n, None], g["dq"][:n, None], g["ddq"][:n, None], F[:, None], g["par"][:n])
    rr = np.linalg.norm(rows[:, 0].cpu().numpy(), axis=1)
    assert np.abs(rr - r).max() <= 1e-9 * np.abs(g["a"][:n]).max()


# grf/autogen-contact.json of data/test_set/2017_08_29/top/jules/run1_1 (what the reference's estimate_grf reads next to
# fte_kinematic/fte.pickle); the trajectory itself is clip 0 of tests/golden/shipped_clips.npz
CLIP_META = {"start_frame": 53, "end_frame": 83,
             "contacts": {"HFL_foot": [[58, 64, 0, "leading"]], "HFR_foot": [[54, 60, 1, "trailing"]],
                          "HBL_foot": [[69, 75, 2, "trailing"]], "HBR_foot": [[72, 78, 3, "leading"]]}}


@pytest.mark.gpu
def test_grf_of_a_real_shipped_trajectory(golden_dir):
    """estimate_grf on the reference's own kinematic solution of jules/run1_1 (30 frames) with its shipped contact
    ranges: a and Bm against the reference expressions (1e-9), optimal value per frame against the SLSQP oracle."""
    from cheetah_pose_estimation_b200 import kinetics
    from oracle import grf as ogrf
    g = np.load(os.path.join(golden_dir, "eom_clip_ref.npz"))
    n = g["q"].shape[0]
    assert np.allclose(g["par"], kinetics.link_params("jules"), rtol=0, atol=1e-15)
    ct = kinetics.contact_mask(CLIP_META)
    assert ct.shape == (n, 4) and ct.sum(axis=0).tolist() == [6, 6, 6, 6]
    st = {k: g[k] for k in ("q", "dq", "ddq")}
    gz, gxy = kinetics.estimate_grf_states(st, CLIP_META, "jules")
    zero = np.zeros((1, n, 20))
    rows0 = kinetics.eom_rows(g["q"][None], g["dq"][None], g["ddq"][None], zero, g["par"][None])[0].cpu().numpy()
    assert np.abs(rows0 - g["a"]).max() <= 1e-9 * np.abs(g["a"]).max()
    for fe in range(n):
        F = np.array([gz[f][fe] for f in kinetics.FEET] + [v for f in kinetics.FEET for v in gxy[f][fe]])
        r_gpu = np.linalg.norm(g["a"][fe] - g["Bm"][fe] @ F)
        _, r_ref = ogrf.solve_frame(g["a"][fe], g["Bm"][fe], ct[fe])
        assert abs(r_gpu - r_ref) <= 1e-8 * np.linalg.norm(g["a"][fe]), (fe, r_gpu, r_ref)
        if not ct[fe].any():
            assert not F.any()
    # stance forces point up: vertical force of every foot in contact is positive somewhere in its stance
    assert all(max(gz[f]) > 0.0 for f in kinetics.FEET)


def test_rates_adjoint_matches_autograd():
    """kinetics.states_gradient_to_q is the adjoint of kinetics.rates_from_q (the mirrored API's dq / ddq from q, incl. the
    free boundary values): checked against torch.autograd on the CPU."""
    import torch
    from cheetah_pose_estimation_b200 import kinetics
    g = torch.Generator().manual_seed(3)
    q = torch.randn(2, 9, 54, dtype=torch.float64, generator=g, requires_grad=True)
    a, b, c = (torch.randn(2, 9, 54, dtype=torch.float64, generator=g) for _ in range(3))
    h = 1.0 / 120.0
    dq, ddq = kinetics.rates_from_q(q, h)
    ((a * q).sum() + (b * dq).sum() + (c * ddq).sum()).backward()
    got = kinetics.states_gradient_to_q(a, b, c, h)
    assert torch.allclose(got, q.grad, rtol=1e-12, atol=1e-9 * float(q.grad.abs().max()))


@pytest.mark.gpu
def test_eom_vjp_matches_complex_step_jacobian(golden_dir):
    """cpe_eom_vjp_f64 (adjoint sweep of the generated EOM program) against d rows / d (q, dq, ddq) obtained by complex-step
    differentiation of sympy.lambdify of the reference-derived expressions (tests/golden/eom_grad_ref.npz,
    oracle/derive_eom_grad.py): random weights, the slack-cost default w = 2 rows, rows returned alongside, ragged batch."""
    import torch
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "eom_grad_ref.npz"))
    n = g["q"].shape[0]
    jac = g["jac"]                                                   # [n, 6, 162]
    rng = np.random.default_rng(5)
    w = rng.normal(size=(n, 6))
    sh = lambda x: x[:, None]                                        # n sequences of one frame: par differs per sample
    eom, gq, gdq, gddq = kinetics.eom_vjp(sh(g["q"]), sh(g["dq"]), sh(g["ddq"]), sh(g["F"]), g["par"], w=sh(w))
    got = torch.cat([gq, gdq, gddq], dim=-1)[:, 0].cpu().numpy()
    want = np.einsum("ni,nij->nj", w, jac)
    scale = np.abs(want).max(axis=1, keepdims=True)
    assert np.abs(got - want).max() / scale.max() < 1e-10 and (np.abs(got - want) / scale).max() < 1e-9
    assert np.abs(eom[:, 0].cpu().numpy() - g["eom"]).max() < 1e-9 * np.abs(g["eom"]).max()
    # default weights: gradient of sum_i row_i^2
    _, gq, gdq, gddq = kinetics.eom_vjp(sh(g["q"]), sh(g["dq"]), sh(g["ddq"]), sh(g["F"]), g["par"])
    got = torch.cat([gq, gdq, gddq], dim=-1)[:, 0].cpu().numpy()
    want = 2.0 * np.einsum("ni,nij->nj", g["eom"], jac)
    assert (np.abs(got - want) / np.abs(want).max(axis=1, keepdims=True)).max() < 1e-9
    # one sequence of n frames, the last two beyond n_frames: zeros there, the others unchanged
    par0 = np.tile(g["par"][0], (n, 1))
    one = lambda x: x[None]
    e1, a1, b1, c1 = kinetics.eom_vjp(one(g["q"]), one(g["dq"]), one(g["ddq"]), one(g["F"]), par0[:1], n_frames=[n - 2])
    e2, a2, b2, c2 = kinetics.eom_vjp(one(g["q"]), one(g["dq"]), one(g["ddq"]), one(g["F"]), par0[:1])
    for x1, x2 in ((e1, e2), (a1, a2), (b1, b2), (c1, c2)):
        assert torch.equal(x1[0, :n - 2], x2[0, :n - 2]) and float(x1[0, n - 2:].abs().max()) == 0.0
    # folded onto q through the mirrored API's difference operator: a directional finite difference of the total cost
    q = torch.as_tensor(g["q"][None], device="cuda")
    h = 1.0 / 120.0
    F, par = torch.as_tensor(g["F"][None], device="cuda"), torch.as_tensor(par0[:1], device="cuda")

    def total(qq):
        dq_, ddq_ = kinetics.rates_from_q(qq, h)
        e = kinetics.eom_rows(qq, dq_, ddq_, F, par)
        return float((e * e).sum())

    dq_, ddq_ = kinetics.rates_from_q(q, h)
    _, a, b, c = kinetics.eom_vjp(q, dq_, ddq_, F, par)
    grad = kinetics.states_gradient_to_q(a, b, c, h)
    d = torch.as_tensor(rng.normal(size=tuple(q.shape)), device="cuda")
    eps = 1e-7
    fd = (total(q + eps * d) - total(q - eps * d)) / (2 * eps)
    an = float((grad * d).sum())
    assert abs(fd - an) <= 1e-5 * abs(an), (fd, an)
