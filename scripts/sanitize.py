"""Every product kernel once at CI shapes, for `compute-sanitizer` (race / memory checking):

    compute-sanitizer --tool memcheck  --error-exitcode 1 python scripts/sanitize.py
    compute-sanitizer --tool racecheck --error-exitcode 1 python scripts/sanitize.py
    compute-sanitizer --tool synccheck --error-exitcode 1 python scripts/sanitize.py
    compute-sanitizer --tool initcheck --error-exitcode 1 python scripts/sanitize.py

Shapes: S = 3 ragged sequences of up to 12 frames, C in {1, 6}, W in {1, 3}; multi-view LM / AL loop (K1 both flavours
with Jacobian, K2, K3, K4, K5, multiplier update), padded-camera batch (cpe_batch.n_cams), monocular priors (prior
kernels + blocked band Cholesky), shutter-delay estimation (K6), equations of motion + GRF solve.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cheetah_pose_estimation_b200 import engine, kinetics, synthetic  # noqa: E402
from cheetah_pose_estimation_b200.priors import Priors  # noqa: E402


def main():
    dev = "cuda:0"
    model = engine.Model(dev)
    ran = []
    for C, W in ((6, 1), (1, 1), (6, 3)):
        prob = synthetic.make_problem(S=3, N=12, C=C, seed=5 + C)
        prob.n_frames[:] = (12, 9, 5)
        meas = prob.meas
        if W == 3:
            meas = np.repeat(meas, 3, axis=4)
        batch = engine.Batch.from_numpy(meas, prob.cams, prob.chain_vec, prob.h, prob.n_frames, device=dev)
        if C == 6:
            batch.n_cams = torch.as_tensor(np.array([6, 4, 5], np.int32), device=dev)
        sv = engine.Solver(model, batch)
        q0 = torch.as_tensor(prob.q_init, device=dev)
        sv.evaluate(q0, pose=True, resid=True, constraints=True, jac=True)
        sv.evaluate(q0, resid=True, jac=True, f32=True)
        res = sv.solve(q0, max_iter=6, n_outer=2)
        ran.append(f"multi-view C={C} W={W}: {int(res['counters'][:, 0].sum())} LM iterations")
        if C == 6 and W == 1:
            r2 = sv.solve_shutter(res["q"], max_iter=4, n_outer=1)
            ran.append(f"shutter: tau {r2['tau'][0].cpu().numpy().round(5)}")
    # monocular priors: synthetic GMM / AR parameters of the right shapes
    rng = np.random.default_rng(3)
    K = 3
    A = rng.normal(size=(K, 22, 22))
    pr = dict(gmm_w=np.full(K, 1.0 / K), gmm_mu=rng.normal(0, 0.2, (K, 22)),
              gmm_cov=np.einsum("kij,klj->kil", A, A) * 0.05 + np.eye(22)[None] * 0.2,
              ar_W=np.concatenate([np.zeros((28, 84)), np.eye(28)], 1) + rng.normal(0, 0.01, (28, 112)),
              ar_b=rng.normal(0, 0.01, 28), ar_var=np.full(28, 0.05))
    prob = synthetic.make_problem(S=3, N=12, C=1, seed=11)
    prob.n_frames[:] = (12, 10, 7)
    batch = engine.Batch.from_numpy(prob.meas, prob.cams, prob.chain_vec, prob.h, prob.n_frames, device=dev)
    sv = engine.Solver(model, batch, priors=Priors(pr, dev))
    res = sv.solve(torch.as_tensor(prob.q_init, device=dev), max_iter=5, n_outer=2)
    ran.append(f"monocular priors: {int(res['counters'][:, 0].sum())} LM iterations")
    # equations of motion + GRF solve
    q = torch.as_tensor(prob.q_true, device=dev)
    h = float(prob.h[0])
    dq = torch.zeros_like(q)
    dq[:, 1:] = (q[:, 1:] - q[:, :-1]) / h
    ddq = torch.zeros_like(q)
    ddq[:, 2:] = (dq[:, 2:] - dq[:, 1:-1]) / h
    par = np.tile(kinetics.link_params("jules"), (3, 1))
    contact = np.ones((3, 12, 4), np.int32)
    grf, resid = kinetics.estimate_grf(q, dq, ddq, par, contact, n_frames=prob.n_frames, iters=50, device=dev)
    eom = kinetics.eom_rows(q, dq, ddq, grf, par, n_frames=prob.n_frames, device=dev)
    ran.append(f"eom + grf: residual {float(torch.as_tensor(resid).abs().max()):.3e}, eom {tuple(eom.shape)}")
    torch.cuda.synchronize()
    for r in ran:
        print(r)
    print("sanitize: all kernels ran")


if __name__ == "__main__":
    main()
