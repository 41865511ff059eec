"""ORACLE (test infrastructure only): the cost terms of the reference's kinetic objective that sit on top of the
trajectory, restated in numpy with their gradients.

  motion smoothing     acinoset_misc.py:601-613   sum_{fe < nfe-2} sum_l sum_d (fps^2 (pose[fe+2] - 2 pose[fe+1] + pose[fe]))^2
  EOM slack            acinoset_misc.py:631-636   sum_fe sum_i slack_eom[fe, i]^2 ; the slack IS the EOM row (system.py:558-567)
  torque squared       motor.py:412-417, 496-503  sum over motors, frames and axes of (weight * Tc)^2
  change in torque     acinoset_misc.py:616-628   sum_{fe < nfe-1} ((Tc[fe+1] - Tc[fe]) / hm0)^2
  kinetic objective    acinoset_opt.py:905-921    1e-3 (kinematic + pose + (torque + 0.1 fps^-2 smoothing) + 10e3 slack)

The generator oracle/make_golden_kinetic.py evaluates the reference's OWN functions (on duck-typed stand-ins for the Pyomo
model) to pin these restatements.
"""
import numpy as np

from . import kinematics


def second_differences(pose):
    """r[fe] = pose[fe+2] - 2 pose[fe+1] + pose[fe] for the reference's frames: 1-based fe with `fe < nfe - 2`
    (acinoset_misc.py:607), i.e. 0-based fe = 0 .. N-4 -- ONE TERM FEWER than the data allows (the last second difference,
    which ends at the last frame, is left out).  Follow the code."""
    pose = np.asarray(pose, float)
    if pose.shape[0] < 4:
        return np.zeros((0,) + pose.shape[1:])
    return pose[2:-1] - 2.0 * pose[1:-2] + pose[:-3]


def motion_smoothing_cost(pose, fps):
    r = second_differences(pose)
    return float(np.sum((fps ** 2 * r) ** 2))


def motion_smoothing_grad_pose(pose, fps):
    """d cost / d pose[N,24,3]."""
    pose = np.asarray(pose, float)
    g = np.zeros_like(pose)
    r = second_differences(pose)
    if r.shape[0]:
        w = 2.0 * fps ** 4 * r
        g[2:-1] += w
        g[1:-2] -= 2.0 * w
        g[:-3] += w
    return g


def motion_smoothing(q, chain_link, chain_vec, fps):
    """(cost, d cost / d q [N,54]) of the smoothing term as a function of the generalised coordinates."""
    pose = kinematics.fk(q, chain_link, chain_vec)
    G = kinematics.fk_jacobian(q, chain_link, chain_vec)                    # [N,24,3,54]
    return motion_smoothing_cost(pose, fps), np.einsum("nli,nlia->na", motion_smoothing_grad_pose(pose, fps), G)


def eom_slack(eom, Bm=None):
    """(cost, d cost / d F [N,20]) with eom[N,6] = a - Bm F (rows 0-5; acinoset_opt.py:120-174)."""
    eom = np.asarray(eom, float)
    cost = float(np.sum(eom * eom))
    g = None if Bm is None else -2.0 * np.einsum("nif,ni->nf", np.asarray(Bm, float), eom)
    return cost, g


def torque_costs(Tc, weight, hm0):
    """Tc[N,M] torques (all motors' axes flattened), weight[M].  Returns (torque^2, d/dTc, change^2, d/dTc)."""
    Tc = np.asarray(Tc, float)
    w = np.broadcast_to(np.asarray(weight, float), Tc.shape[1:])
    c1 = float(np.sum((w * Tc) ** 2))
    g1 = 2.0 * w * w * Tc
    # 1-based fe with `fe < nfe - 1` (acinoset_misc.py:623): 0-based fe = 0 .. N-3, the last difference is left out
    d = (Tc[1:-1] - Tc[:-2]) / hm0
    c2 = float(np.sum(d * d))
    g2 = np.zeros_like(Tc)
    g2[1:-1] += 2.0 * d / hm0
    g2[:-2] -= 2.0 * d / hm0
    return c1, g1, c2, g2
