"""Trajectory-level terms of the reference's kinetic objective (acinoset_opt.py:905-921): motion smoothing,
EOM slack, torque squared, change in torque.  tests/golden/kinetic_costs_ref.npz holds the values of the reference's OWN
functions (oracle/make_golden_kinetic.py); the oracle restates them with gradients; the CUDA kernels are compared to both."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "kinetic_costs_ref.npz"))


def test_oracle_matches_the_reference_functions(gold):
    from oracle import kinetic_costs as kc
    assert abs(kc.motion_smoothing_cost(gold["pose"], float(gold["fps"])) - float(gold["smooth_ref"])) <= 1e-12 * float(gold["smooth_ref"])
    assert abs(kc.eom_slack(gold["eom"])[0] - float(gold["slack_ref"])) <= 1e-12 * float(gold["slack_ref"])
    c1, _, c2, _ = kc.torque_costs(gold["Tc"], float(gold["weight"]), float(gold["hm0"]))
    assert abs(c1 - float(gold["torque_ref"])) <= 1e-12 * c1 and abs(c2 - float(gold["dtau_ref"])) <= 1e-12 * c2


def test_oracle_gradients_by_finite_differences(gold):
    from cheetah_pose_estimation_b200 import skeleton as sk
    from oracle import kinetic_costs as kc
    lengths, radii, _ = sk.animal("jules")
    cl, cv = sk.marker_chains(lengths, radii)
    q, fps = gold["q"][:9].copy(), float(gold["fps"])
    c0, g = kc.motion_smoothing(q, cl, cv, fps)
    rng = np.random.default_rng(1)
    for _ in range(8):
        n, p = int(rng.integers(0, 9)), int(rng.integers(0, 54))
        qp, qm = q.copy(), q.copy()
        qp[n, p] += 1e-6
        qm[n, p] -= 1e-6
        fd = (kc.motion_smoothing(qp, cl, cv, fps)[0] - kc.motion_smoothing(qm, cl, cv, fps)[0]) / 2e-6
        assert abs(fd - g[n, p]) <= 1e-5 * max(1.0, abs(g[n, p]), abs(c0) * 1e-6), (n, p, fd, g[n, p])
    Tc = gold["Tc"][:7]
    c1, g1, c2, g2 = kc.torque_costs(Tc, 1.3, 0.01)
    Tp = Tc.copy()
    Tp[3, 5] += 1e-6
    a1, _, a2, _ = kc.torque_costs(Tp, 1.3, 0.01)
    assert abs((a1 - c1) / 1e-6 - g1[3, 5]) < 1e-4 * abs(g1[3, 5]) and abs((a2 - c2) / 1e-6 - g2[3, 5]) < 1e-4 * abs(g2[3, 5])


@pytest.mark.gpu
def test_kernels_match_the_oracle_and_the_reference_values(gold):
    import torch
    from cheetah_pose_estimation_b200 import engine, kinetics, skeleton as sk
    from oracle import kinetic_costs as kc
    model = engine.Model("cuda:0")
    lengths, radii, _ = sk.animal("jules")
    cl, cv = sk.marker_chains(lengths, radii)
    fps = float(gold["fps"])
    N = gold["q"].shape[0]
    # ragged batch: the real clip and two truncated copies (incl. one too short for any second difference)
    nf = np.array([N, 11, 3], np.int32)
    q = np.stack([gold["q"], gold["q"], gold["q"]])
    meas = np.zeros((3, N, 1, 24, 1, 4), np.float32)
    cams = np.zeros((3, 1, sk.CAM_STRIDE))
    cams[:, 0, [0, 1]] = 900.0
    cams[:, 0, [8, 12, 16]] = 1.0
    cams[:, 0, 19] = 10.0
    batch = engine.Batch.from_numpy(meas, cams, np.stack([cv] * 3), np.full(3, 1.0 / fps), n_frames=nf)
    sv = engine.Solver(model, batch, with_solver=False)
    cost, grad = kinetics.motion_smoothing(sv, torch.as_tensor(q, device="cuda:0"))
    cost, grad = cost.cpu().numpy(), grad.cpu().numpy()
    assert abs(cost[0] - float(gold["smooth_ref"])) <= 1e-9 * float(gold["smooth_ref"])        # the reference's own value
    for s in range(3):
        c_o, g_o = kc.motion_smoothing(q[s, :nf[s]], cl, cv, fps)
        assert abs(cost[s] - c_o) <= 1e-9 * max(1.0, c_o)
        assert np.abs(grad[s, :nf[s]] - g_o).max() <= 1e-9 * max(1.0, np.abs(g_o).max())
        assert not np.any(grad[s, nf[s]:])
    assert cost[2] == 0.0
    # EOM slack (+ gradient with respect to the forces)
    rng = np.random.default_rng(4)
    eom = np.stack([gold["eom"], rng.normal(size=gold["eom"].shape)])
    Bm = rng.normal(size=(2, N, 6, 20))
    c, g = kinetics.eom_slack(eom, Bm, n_frames=[N, 7])
    assert abs(float(c[0]) - float(gold["slack_ref"])) <= 1e-12 * float(gold["slack_ref"])
    c_o, g_o = kc.eom_slack(eom[1, :7], Bm[1, :7])
    assert abs(float(c[1]) - c_o) <= 1e-12 * c_o and np.abs(g[1, :7].cpu().numpy() - g_o).max() < 1e-12 * np.abs(g_o).max()
    # torque terms
    Tc = np.stack([gold["Tc"], rng.normal(size=gold["Tc"].shape)])
    w, hm0 = float(gold["weight"]), float(gold["hm0"])
    c, g1, g2 = kinetics.torque_costs(Tc, w, hm0, n_frames=[N, 9])
    assert abs(float(c[0, 0]) - float(gold["torque_ref"])) <= 1e-12 * float(gold["torque_ref"])
    assert abs(float(c[0, 1]) - float(gold["dtau_ref"])) <= 1e-12 * float(gold["dtau_ref"])
    a1, b1, a2, b2 = kc.torque_costs(Tc[1, :9], w, hm0)
    assert abs(float(c[1, 0]) - a1) <= 1e-12 * a1 and abs(float(c[1, 1]) - a2) <= 1e-12 * a2
    assert np.abs(g1[1, :9].cpu().numpy() - b1).max() <= 1e-12 * np.abs(b1).max()
    assert np.abs(g2[1, :9].cpu().numpy() - b2).max() <= 1e-12 * np.abs(b2).max()
    obj = kinetics.kinetic_objective(100.0, 2.0, float(c[0, 0]), float(cost[0]), float(gold["slack_ref"]), fps)
    assert abs(obj - 1e-3 * (102.0 + float(c[0, 0]) + 0.1 * fps ** -2 * float(cost[0]) + 10e3 * float(gold["slack_ref"]))) < 1e-9
