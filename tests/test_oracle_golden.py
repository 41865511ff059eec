"""Pins the CPU oracle (oracle/) to the reference: fixtures in tests/golden/ were produced by
oracle/make_golden.py from the reference's OWN functions and from its shipped data/test_set outputs."""
import os

import numpy as np
import pytest

from oracle import cameras, cost, kinematics, tables


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


def test_tables_match_reference(golden_dir):
    g = _load(golden_dir, "tables_ref.npz")
    r_pw, q = tables.uncertainty_models()
    assert np.array_equal(r_pw, g["R_pw"])
    assert np.array_equal(q, g["Q"])
    assert list(g["markers"]) == tables.MARKERS
    assert [tables.DLC_INDEX[m] for m in tables.MARKERS] == list(g["dlc_index"])
    assert [tables.PAIRWISE_GRAPH[m] for m in tables.MARKERS] == g["pairwise"].tolist()
    rp = _load(golden_dir, "reduced_pose_ref.npz")
    assert np.array_equal(tables.relative_angle_mask(), rp["mask"])


@pytest.mark.parametrize("name", tables.ANIMAL_NAMES)
def test_fk_matches_reference_lambdified(golden_dir, name):
    g = _load(golden_dir, "fk_ref.npz")
    lengths, radii, masses = tables.animal(name)
    assert np.array_equal(lengths, g[f"{name}_lengths"])
    assert np.allclose(radii, g[f"{name}_radii"], rtol=0, atol=1e-16)
    assert np.allclose(masses, g[f"{name}_masses"], rtol=1e-15, atol=0)
    cl, cv = kinematics.marker_chains(lengths, radii)
    pos = kinematics.fk(g[f"{name}_q"], cl, cv)
    assert np.abs(pos - g[f"{name}_positions"]).max() < 1e-13


def test_fk_matches_all_shipped_clips(golden_dir):
    g = _load(golden_dir, "shipped_clips.npz")
    names = list(g["names"])
    assert len(names) == 55
    worst = 0.0
    for i, nm in enumerate(names):
        animal = nm.split("|")[1]
        lengths, radii, _ = tables.animal(animal)
        cl, cv = kinematics.marker_chains(lengths, radii)
        pos = kinematics.fk(g[f"{i}_q"], cl, cv)
        worst = max(worst, np.abs(pos - g[f"{i}_positions"]).max())
        x = kinematics.reduced_pose(g[f"{i}_q"])
        assert np.abs(x - g[f"{i}_x"]).max() < 1e-12
    assert worst < 1e-13


def test_fk_jacobian_matches_sympy_of_reference(golden_dir):
    g = _load(golden_dir, "fk_jac_ref.npz")
    cl, cv = kinematics.marker_chains(g["lengths"], g["radii"])
    J = kinematics.fk_jacobian(g["q"], cl, cv)
    assert np.abs(J - g["jac"]).max() < 1e-13
    # chain sparsity of SURVEY 8(a4): sum of active q inputs = 270
    nz = (np.abs(g["jac"]).max(axis=(0, 2)) > 0).sum()
    assert 240 <= nz <= 270   # d/dphi of R @ [x,0,0] vanishes identically, hence < 270


def test_joint_constraints_match_reference(golden_dir):
    g = _load(golden_dir, "constraints_ref.npz")
    c = kinematics.joint_constraints(g["q"])
    assert np.abs(c[:, g["oracle_index_of_ref_row"]] - g["c_ref"]).max() < 1e-13
    # Jacobian by central differences
    q = g["q"][:2]
    J = kinematics.joint_constraints_jacobian(q)
    eps = 1e-6
    for p in range(54):
        dq = np.zeros(54)
        dq[p] = eps
        fd = (kinematics.joint_constraints(q + dq) - kinematics.joint_constraints(q - dq)) / (2 * eps)
        assert np.abs(fd - J[:, :, p]).max() < 1e-8


def test_joint_constraints_hold_at_shipped_optima(golden_dir):
    g = _load(golden_dir, "shipped_clips.npz")
    for i, nm in enumerate(g["names"]):
        if not nm.split("|")[0].endswith("fte_kinematic"):
            continue
        c = kinematics.joint_constraints(g[f"{i}_q"])
        assert np.abs(c).max() < 1e-4, (nm, np.abs(c).max())
        d = kinematics.inequality_values(g[f"{i}_q"], kinematics.inequality_rows("kinetic_dataset" in nm))
        rows = kinematics.inequality_rows("kinetic_dataset" in nm)
        lo = np.array([r[2] for r in rows])
        up = np.array([r[3] for r in rows])
        assert (d > lo - 1e-3).all() and (d < up + 1e-3).all(), nm


def test_projection_matches_reference(golden_dir):
    g = _load(golden_dir, "projection_ref.npz")
    for c in range(g["cams"].shape[0]):
        uv = cameras.project(g["X"], g["cams"][c], cameras.FISHEYE)
        assert np.abs(uv - g["uv_fisheye"][c]).max() < 1e-9
        uv = cameras.project(g["X"], g["cams"][c], cameras.PINHOLE)
        ref = g["uv_pinhole"][c]
        assert np.abs(uv - ref).max() <= 1e-12 * max(1.0, np.abs(ref).max())
        for model in (cameras.FISHEYE, cameras.PINHOLE):
            J = cameras.project_jacobian(g["X"], g["cams"][c], model)
            eps = 1e-6
            for k in range(3):
                d = np.zeros(3)
                d[k] = eps
                fd = (cameras.project(g["X"] + d, g["cams"][c], model) - cameras.project(g["X"] - d, g["cams"][c], model)) / (2 * eps)
                assert np.abs(fd - J[..., k]).max() <= 2e-6 * max(1.0, np.abs(J).max())


def test_redescending_loss_matches_reference(golden_dir):
    g = _load(golden_dir, "loss_ref.npz")
    assert np.abs(cost.redescending_loss(g["err"]) - g["loss"]).max() < 1e-12
    e = np.linspace(0.01, 35, 500)
    fd = (cost.redescending_loss(e + 1e-6) - cost.redescending_loss(e - 1e-6)) / 2e-6
    assert np.abs(fd - cost.redescending_dloss(e)).max() < 1e-7


def test_reduced_pose_matches_reference(golden_dir):
    g = _load(golden_dir, "reduced_pose_ref.npz")
    assert np.abs(kinematics.reduced_pose(g["q"]) - g["x"]).max() < 1e-14


def _problem_from_clip(g, animal, weight_on=True):
    lengths, radii, _ = tables.animal(animal)
    cl, cv = kinematics.marker_chains(lengths, radii)
    det = g["detections"].astype(float)                     # [N,C,24,2]
    r_pw, _ = tables.uncertainty_models()
    w = np.broadcast_to(1.0 / r_pw[0], det.shape[:3])
    meas = np.concatenate([det, w[..., None]], -1)[:, :, :, None, :]
    return cost.Problem(meas=meas, cams=g["cams"], chain_link=cl, chain_vec=cv, h=1.0 / float(g["fps"]))


@pytest.mark.parametrize("fname,animal", [("clip_jules_run1_1.npz", "jules"), ("clip_phantom_run.npz", "phantom")])
def test_real_clip_known_answers(golden_dir, fname, animal):
    g = _load(golden_dir, fname)
    pb = _problem_from_clip(g, animal)
    r = cost.residuals(pb, g["q"])
    # recovered cameras/detections reproduce the saved residuals (float32 rounding of the detections only)
    assert np.abs(r[:, :, :, 0, :] - g["meas_err"][..., 0]).max() < 2e-4
    mc = cost.measurement_cost(pb, g["q"])
    assert abs(mc - float(g["known_meas_cost_all_on"])) / float(g["known_meas_cost_all_on"]) < 1e-6
    # motion term in q-only form equals the ddq-based reference expression
    model_q = cost.motion_cost(pb, g["q"])
    dd = g["ddq"][3:] - g["ddq"][2:-1]
    model_ddq = float(np.sum(pb.wq * dd * dd))
    assert abs(model_q - model_ddq) / model_ddq < 1e-8
    # and the boundary terms the reference leaves free are small at its optimum
    assert model_ddq <= float(g["model_cost_from_ddq"]) * (1 + 1e-12)
    assert model_ddq > 0.95 * float(g["model_cost_from_ddq"])


def test_cost_assembly_matches_reference_functions(golden_dir):
    g = _load(golden_dir, "cost_ref.npz")
    cl, cv = kinematics.marker_chains(g["lengths"], g["radii"])
    r_pw, _ = tables.uncertainty_models()
    w = np.where(g["likelihood"] > 0.5, 1.0 / r_pw[0][None, None, :], 0.0)
    meas = np.concatenate([g["meas"], w[..., None]], -1)[:, :, :, None, :]
    pb = cost.Problem(meas=meas, cams=g["cams"], chain_link=cl, chain_vec=cv, h=1 / 90)
    r = cost.residuals(pb, g["q"])[:, :, :, 0, :]
    assert np.abs(r - g["resid"]).max() <= 1e-9 * np.abs(g["resid"]).max()
    assert abs(cost.measurement_cost(pb, g["q"]) - float(g["meas_cost"])) < 1e-9 * float(g["meas_cost"])


def test_gradient_and_gn_blocks_are_consistent(golden_dir):
    g = _load(golden_dir, "cost_ref.npz")
    cl, cv = kinematics.marker_chains(g["lengths"], g["radii"])
    r_pw, _ = tables.uncertainty_models()
    w = np.where(g["likelihood"] > 0.5, 1.0 / r_pw[0][None, None, :], 0.0)
    meas = np.concatenate([g["meas"], w[..., None]], -1)[:, :, :, None, :]
    pb = cost.Problem(meas=meas, cams=g["cams"], chain_link=cl, chain_vec=cv, h=1 / 90)
    rng = np.random.default_rng(1)
    q = g["q"] + rng.normal(0, 0.02, g["q"].shape)
    lam_eq = rng.normal(0, 1, (q.shape[0], 26))
    lam_in = rng.normal(0, 1, (q.shape[0], len(pb.ineq_rows)))
    grad = cost.full_gradient(pb, q, lam_eq, lam_in)
    f = lambda qq: cost.cost_terms(pb, qq, lam_eq, lam_in).sum()
    for (k, p) in [(0, 0), (1, 4), (2, 7), (3, 19), (4, 52), (2, 37), (0, 23)]:
        d = np.zeros_like(q)
        eps = 1e-6
        d[k, p] = eps
        fd = (f(q + d) - f(q - d)) / (2 * eps)
        assert abs(fd - grad[k, p]) <= 1e-5 * max(1.0, abs(fd)), (k, p, fd, grad[k, p])
    A, gvec = cost.dense_hessian(pb, q, lam_eq, lam_in)
    assert np.allclose(A, A.T)
    assert np.linalg.eigvalsh(A).min() > -1e-6 * np.abs(A).max()


def test_shutter_delay_residuals_match_reference_expression(golden_dir):
    """slack = proj(pose + dq_base tau_c + ddq_base tau_c^2) - meas, evaluated by the reference's own lambdified FK
    and pt3d_to_2d_fisheye on the shipped solution of jules/run1_1 (oracle/make_golden_shutter.py)."""
    g = _load(golden_dir, "shutter_ref.npz")
    clip = _load(golden_dir, "clip_jules_run1_1.npz")
    cl, cv = kinematics.marker_chains(g["lengths"], g["radii"])
    N, C = clip["detections"].shape[:2]
    meas = np.concatenate([clip["detections"].astype(float), np.ones((N, C, 24, 1))], -1)[:, :, :, None, :]
    pb = cost.Problem(meas=meas, cams=clip["cams"], chain_link=cl, chain_vec=cv, h=1 / float(clip["fps"]))
    r = cost.residuals(pb, clip["q"], g["tau"], clip["dq"][:, 0:3], clip["ddq"][:, 0:3])[g["frames"], :, :, 0, :]
    assert (np.abs(r - g["slack"]) / np.maximum(np.abs(g["slack"]), 1.0)).max() < 1e-9
    # tau = 0 is the undelayed model, and camera 1 (tau fixed at 0, acinoset_misc.py:274-275) is untouched
    r0 = cost.residuals(pb, clip["q"])[g["frames"], :, :, 0, :]
    assert np.array_equal(r[:, 0], r0[:, 0]) and np.abs(r[:, 1:] - r0[:, 1:]).max() > 1.0
    # d slack / d tau against a central difference
    J = cost.shutter_jacobian(pb, clip["q"], g["tau"], clip["dq"][:, 0:3], clip["ddq"][:, 0:3])
    eps = 1e-7
    for c in (1, 4):
        d = np.zeros(C)
        d[c] = eps
        fd = (cost.residuals(pb, clip["q"], g["tau"] + d, clip["dq"][:, 0:3], clip["ddq"][:, 0:3])
              - cost.residuals(pb, clip["q"], g["tau"] - d, clip["dq"][:, 0:3], clip["ddq"][:, 0:3]))[:, :, :, 0, :] / (2 * eps)
        assert np.abs(fd[:, c] - J[:, c]).max() <= 1e-6 * np.abs(J[:, c]).max()
        assert np.abs(np.delete(fd, c, axis=1)).max() == 0.0


def test_oracle_residuals_match_every_shipped_multi_view_clip(golden_dir):
    """All 15 multi-view clips of the reference's data/test_set (10 AcinoSet fisheye, 5 kinetic-dataset pinhole),
    cameras / detections rebuilt by oracle/make_golden_clips.py with the reference's own pt3d_to_2d[_fisheye]: the
    oracle's FK reproduces the saved `positions` and its residuals the saved `meas_err` at the shipped optimum
    (AcinoSet detections are stored as float32 -> half a float32 ulp of a pixel coordinate; kinetic ones in double)."""
    g = _load(golden_dir, "test_set_clips.npz")
    names = list(g["names"])
    assert len(names) == 15
    for i, nm in enumerate(names):
        kin = bool(g[f"{i}_kinetic"])
        lengths, radii, _ = tables.animal(str(g[f"{i}_animal"]))
        cl, cv = kinematics.marker_chains(lengths, radii)
        q, det, cams = g[f"{i}_q"], g[f"{i}_detections"], g[f"{i}_cams"]
        assert float(g[f"{i}_rms"].max()) < 1e-8, nm                      # camera recovery converged
        pos = kinematics.fk(q, cl, cv)
        assert np.abs(pos - g[f"{i}_positions"]).max() < 1e-13, nm
        N, C = det.shape[:2]
        meas = np.concatenate([det.astype(np.float64)[:, :, :, None, :], np.ones((N, C, 24, 1, 1))], -1)
        pb = cost.Problem(meas=meas, cams=cams, chain_link=cl, chain_vec=cv, h=1.0 / float(g[f"{i}_fps"]),
                          cam_model=cameras.PINHOLE if kin else cameras.FISHEYE)
        r = cost.residuals(pb, q)[:, :, :, 0, :]
        tol = 1e-9 if kin else 2e-4
        assert np.abs(r - g[f"{i}_meas_err"][..., 0]).max() < tol, (nm, np.abs(r - g[f"{i}_meas_err"][..., 0]).max())
