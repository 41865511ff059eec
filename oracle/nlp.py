"""Independent witness for the converged solution (oracle; TEST INFRASTRUCTURE -- never imported by the product).

The reference hands a constrained NLP to IPOPT + MA97 (acinoset_opt.py:601-618, shared/physical_education/utils.py:
459-548); IPOPT is not installable here.  This file states the SAME NLP -- not the augmented-Lagrangian / LM algorithm
of the CUDA path (oracle/lm.py restates that one) -- and gives it to a general-purpose constrained solver from scipy:

    minimise   1e-3 * ( sum rho(w * (proj(FK(q)) - meas))  +  sum_{k>=3} wq ((q_k - 3 q_{k-1} + 3 q_{k-2} - q_{k-3}) / h^2)^2 )
    subject to the 26 joint equalities per frame  (R_a e_i) . (R_b e_j) = 0          (system.py:574-586, links.py:552-586)
               the two-sided inequality rows       lo <= q[+] - q[-] <= up           (cheetah.py:306-352)
               the bounds  q in [-100, 100], z_base >= 0                              (links.py:602-605, 636-639)

with the constraints as CONSTRAINTS (SLSQP's active-set SQP or trust-constr's interior point), not penalties.  The
slack / dq / ddq variables of the Pyomo model are defined by equalities and are eliminated analytically, which does
not change the feasible set or the optimum (the free first-node rates make the first three motion terms
vanish).  Gradient and constraint Jacobians are analytic, written here from the pinned kinematics / camera pieces and
checked against finite differences in tests/test_nlp_witness.py.

Use: small problems only (N <= 12 frames: 648 unknowns, dense SQP).
"""
import time

import numpy as np
from scipy import optimize

from . import cameras, cost, kinematics

OBJ_SCALE = 1e-3                      # acinoset_opt.py:602


def objective(pb, x):
    """(f, grad) of the reference objective as a function of q.flatten()."""
    N = pb.n_frames
    q = x.reshape(N, 54)
    pose = kinematics.fk(q, pb.chain_link, pb.chain_vec)                                    # [N,24,3]
    G = kinematics.fk_jacobian(q, pb.chain_link, pb.chain_vec)                              # [N,24,3,54]
    uv = cameras.project(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)            # [N,C,24,2]
    P = cameras.project_jacobian(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)    # [N,C,24,2,3]
    w = pb.meas[..., 2:3]                                                                   # [N,C,24,W,1]
    e = w * (uv[:, :, :, None, :] - pb.meas[..., 0:2])                                      # [N,C,24,W,2]
    if pb.square_loss:
        f_meas = float(np.sum(e * e))
        de = 2.0 * e
    else:
        f_meas = float(np.sum(cost.redescending_loss(e)))
        de = cost.redescending_dloss(e) * np.sign(e)
    b = np.einsum("nclwd,ncldi->nli", de * w, P)                                            # d f / d pose
    g = np.einsum("nli,nlia->na", b, G)
    d3 = cost.third_difference(q, pb.h)
    f_mod = float(np.sum(pb.wq * d3 * d3))
    g = g + cost.motion_gradient(pb, q)
    return OBJ_SCALE * (f_meas + f_mod), OBJ_SCALE * g.reshape(-1)


def equalities(pb, x):
    q = x.reshape(pb.n_frames, 54)
    return kinematics.joint_constraints(q).reshape(-1)


def equalities_jacobian(pb, x):
    N = pb.n_frames
    q = x.reshape(N, 54)
    Jc = kinematics.joint_constraints_jacobian(q)                                           # [N,26,54]
    J = np.zeros((N * 26, N * 54))
    for k in range(N):
        J[k * 26:(k + 1) * 26, k * 54:(k + 1) * 54] = Jc[k]
    return J


def inequality_matrix(pb):
    """(A, lo, up) with lo <= A q.flatten() <= up for all frames."""
    N, rows = pb.n_frames, pb.ineq_rows
    A = np.zeros((N * len(rows), N * 54))
    lo = np.zeros(N * len(rows))
    up = np.zeros(N * len(rows))
    for k in range(N):
        for j, (ip, im, l, u) in enumerate(rows):
            r = k * len(rows) + j
            A[r, k * 54 + ip] = 1.0
            if im >= 0:
                A[r, k * 54 + im] = -1.0
            lo[r], up[r] = l, u
    return A, lo, up


def solve(pb, q0, method="SLSQP", maxiter=400, ftol=1e-12, verbose=False):
    """Constrained minimum from q0[N,54].  Returns dict(q, obj, cost[2] = (measurement, model) unscaled, c_eq_max,
    ineq_violation, nit, success, message, seconds)."""
    N = pb.n_frames
    x0 = np.asarray(q0, float).reshape(-1).copy()
    A, lo, up = inequality_matrix(pb)
    bounds_lo = np.full(N * 54, -100.0)
    bounds_up = np.full(N * 54, 100.0)
    bounds_lo[2::54] = 0.0                                    # z_base >= 0
    t0 = time.time()
    if method == "SLSQP":
        cons = [{"type": "eq", "fun": lambda x: equalities(pb, x), "jac": lambda x: equalities_jacobian(pb, x)},
                {"type": "ineq", "fun": lambda x: np.concatenate([A @ x - lo, up - A @ x]),
                 "jac": lambda x: np.concatenate([A, -A])}]
        res = optimize.minimize(lambda x: objective(pb, x), x0, jac=True, method="SLSQP", constraints=cons,
                                bounds=list(zip(bounds_lo, bounds_up)),
                                options=dict(maxiter=maxiter, ftol=ftol, disp=verbose))
    else:
        cons = [optimize.NonlinearConstraint(lambda x: equalities(pb, x), 0.0, 0.0,
                                             jac=lambda x: equalities_jacobian(pb, x), hess=optimize.BFGS()),
                optimize.LinearConstraint(A, lo, up)]
        res = optimize.minimize(lambda x: objective(pb, x), x0, jac=True, hess=optimize.BFGS(), method="trust-constr",
                                constraints=cons, bounds=optimize.Bounds(bounds_lo, bounds_up),
                                options=dict(maxiter=maxiter, gtol=1e-8, xtol=1e-12, verbose=2 if verbose else 0))
    q = res.x.reshape(N, 54)
    d = A @ res.x
    return dict(q=q, obj=float(res.fun), cost=np.array([cost.measurement_cost(pb, q), cost.motion_cost(pb, q)]),
                c_eq_max=float(np.abs(equalities(pb, res.x)).max()),
                ineq_violation=float(max(0.0, (lo - d).max(), (d - up).max())),
                nit=int(res.get("nit", -1)), success=bool(res.success), message=str(res.message),
                seconds=time.time() - t0)
