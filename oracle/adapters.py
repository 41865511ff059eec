"""Glue between the product's array layouts and the oracle's `cost.Problem` (test infrastructure)."""
import numpy as np

from . import cameras, cost, kinematics, tables


def oracle_weights(gate, cams, W=1, kinetic_dataset=False):
    """weight[N,C,L,W] = gate * (1/R_pw[w][l]) * cam_mult[c]  (acinoset_misc.py:231, 462-482)."""
    r_pw, _ = tables.uncertainty_models(kinetic_dataset)
    w = (1.0 / r_pw[:W]).T[None, None, :, :]                       # [1,1,L,W]
    return gate * w * cams[None, :, None, None, 20]


def problem_from_arrays(meas, cams, chain_link, chain_vec, h, cam_model=cameras.FISHEYE, kinetic_dataset=False,
                        mu=1e4, square_loss=False):
    """meas[N,C,24,W,4] (u,v,gate,_), cams[C,24] -> cost.Problem for ONE sequence."""
    meas = np.asarray(meas, float)
    wgt = oracle_weights(meas[..., 2], cams, meas.shape[3], kinetic_dataset)
    if square_loss:                                   # the camera multiplier is not applied in this branch (:471-474)
        wgt = wgt / np.asarray(cams, float)[None, :, None, None, 20]
    m3 = np.concatenate([meas[..., 0:2], wgt[..., None]], -1)
    return cost.Problem(meas=m3, cams=np.asarray(cams, float)[:, :20], chain_link=chain_link, chain_vec=chain_vec,
                        h=float(h), cam_model=cam_model, ineq_rows=kinematics.inequality_rows(kinetic_dataset),
                        mu_eq=mu, mu_in=mu, square_loss=square_loss)


def from_synthetic(prob, s, mu=1e4):
    n = int(prob.n_frames[s])
    return problem_from_arrays(prob.meas[s, :n], prob.cams[s], prob.chain_link, prob.chain_vec[s], prob.h[s], mu=mu)


def projector(q, cams, chain_vec):
    """Oracle stand-in for the CUDA projector of cheetah_pose_estimation_b200.synthetic (CPU tests)."""
    S = q.shape[0]
    out = []
    for s in range(S):
        lengths, radii, _ = tables.animal("jules")
        cl, _ = kinematics.marker_chains(lengths, radii)
        pose = kinematics.fk(q[s], cl, chain_vec[s])
        out.append(cameras.project(pose[:, None], cams[s][None, :, None, :20], cameras.FISHEYE))
    return np.stack(out)
