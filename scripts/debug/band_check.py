"""Debug: K3 (band solve) alone against a numpy solve of the SAME reduced system, across LM iterations."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from cheetah_pose_estimation_b200 import engine, synthetic, skeleton
from oracle import cost as ocost, lm as olm, tables

S, N = 4, int(sys.argv[1]) if len(sys.argv) > 1 else 60
prob = synthetic.make_problem(S=S, N=N, C=6, seed=4)
model = engine.Model("cuda:0")
batch = prob.to_batch("cuda:0")
sv = engine.Solver(model, batch)
q = torch.as_tensor(prob.q_init, device="cuda:0")
ind = tables.relative_angle_mask()
wq = skeleton.motion_weights()
m = ocost.motion_stencil(N)
for damp_v in (1e-3, 1e-6, 1e-9, 1e-12):
    damp = torch.full((S,), damp_v, dtype=torch.float64, device="cuda:0")
    sv.assemble(q, damp)
    sv.band_solve()
    torch.cuda.synchronize()
    Sb, gh, d = sv.Sblk.cpu().numpy(), sv.ghat.cpu().numpy(), sv.delta_ind.cpu().numpy()
    for s in range(S):
        wd = 2.0 * wq / prob.h[s] ** 4
        ref = olm.band_solve(Sb[s], gh[s], m, wd[ind])
        err = np.abs(d[s] - ref).max() / max(1.0, np.abs(ref).max())
        print(f"damp {damp_v:g} seq {s} status {int(sv.status[s])} rel err {err:.3e} |ref| {np.abs(ref).max():.3e}")
    # move q a bit along the step so that later systems differ
    sv.step_apply(q, damp)
    q = sv.q_try.clone()
