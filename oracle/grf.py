"""ORACLE (test infrastructure only): CPU restatement of CheetahEstimator.estimate_grf's per-frame problem
(acinoset_opt.py:176-249): minimise sqrt(sum_i eom_i^2) over GRFz[f] in [0, 5], GRFxy[f, 0..3] in [0, 5] of the feet in
contact, subject to the friction polyhedron 1.3 GRFz[f] >= sum_i GRFxy[f, i] (:187-190); forces of the other feet are
fixed to 0 (:221-223); frames without any contact return zeros (:231-236).  eom = a - Bm F is affine in the forces.
IPOPT is replaced by scipy SLSQP on the equivalent squared objective (same minimiser); parity is on the optimal
VALUE -- the minimiser itself is not unique (opposite friction directions cancel)."""
import numpy as np
from scipy.optimize import minimize

MU, FMAX = 1.3, 5.0


def solve_frame(a, Bm, contact):
    """Returns (F[20], residual norm).  contact[4] booleans in foot order HFL, HFR, HBL, HBR."""
    contact = np.asarray(contact, bool)
    if not contact.any():
        return np.zeros(20), float(np.linalg.norm(a))
    free = np.concatenate([contact, np.repeat(contact, 4)])
    idx = np.nonzero(free)[0]
    B = Bm[:, idx]
    sc = np.abs(B).max()
    cons = []
    for f in np.nonzero(contact)[0]:
        z = int(np.nonzero(idx == f)[0][0])
        xs = [int(np.nonzero(idx == 4 + 4 * f + i)[0][0]) for i in range(4)]
        cons.append({"type": "ineq", "fun": (lambda x, z=z, xs=xs: MU * x[z] - x[xs].sum()),
                     "jac": (lambda x, z=z, xs=xs: np.bincount([z] + xs, weights=[MU] + [-1.0] * 4, minlength=len(idx)))})
    best = None
    for x0 in (np.full(len(idx), 0.1), np.zeros(len(idx)), np.full(len(idx), 1.0)):
        res = minimize(lambda x: 0.5 * np.sum(((a - B @ x) / sc) ** 2), x0, jac=lambda x: -(B.T @ (a - B @ x)) / sc ** 2,
                       bounds=[(0.0, FMAX)] * len(idx), constraints=cons, method="SLSQP",
                       options={"ftol": 1e-16, "maxiter": 500})
        if best is None or res.fun < best.fun:
            best = res
    F = np.zeros(20)
    F[idx] = np.clip(best.x, 0.0, FMAX)
    return F, float(np.linalg.norm(a - Bm @ F))
