"""Compare LM/AL schedules on the bench's LM problem: final merit, joint violation, iterations, time."""
import sys
import time

import torch

sys.path.insert(0, ".")
from cheetah_pose_estimation_b200 import engine, synthetic

S = int(sys.argv[1]) if len(sys.argv) > 1 else 296
dev = torch.device("cuda:0")
model = engine.Model(dev)
batch, q_true, q_init = synthetic.make_device_problem(S, 500, 6, seed=0, device=dev, dlc_native=True)
q0 = q_init.contiguous()
cfgs = {
    "default": dict(max_iter=20, n_outer=3),
    "ftol1e-5": dict(max_iter=20, n_outer=3, ftol=1e-5),
    "ftol1e-4": dict(max_iter=20, n_outer=3, ftol=1e-4),
    "mu1e5,1e6": dict(max_iter=20, n_outer=2, mu_schedule=(1e5, 1e6)),
    "mu1e6x2": dict(max_iter=20, n_outer=2, mu_schedule=(1e6, 1e6)),
    "mu1e5,1e6,1e6": dict(max_iter=20, n_outer=3, mu_schedule=(1e5, 1e6, 1e6)),
    "ftol 1e-3,1e-5,1e-6": dict(max_iter=20, n_outer=3, ftol=(1e-3, 1e-5, 1e-6)),
    "ftol 1e-4,1e-5,1e-6": dict(max_iter=20, n_outer=3, ftol=(1e-4, 1e-5, 1e-6)),
    "ftol 1e-2,1e-4,1e-6": dict(max_iter=20, n_outer=3, ftol=(1e-2, 1e-4, 1e-6)),
    "ftol 1e-4,1e-4,1e-6": dict(max_iter=20, n_outer=3, ftol=(1e-4, 1e-4, 1e-6)),
    "it 12,20,20": dict(max_iter=12, n_outer=3),
}
sv = engine.Solver(model, batch)
sv.solve(q0, max_iter=2, n_outer=1)
for name, kw in cfgs.items():
    torch.cuda.synchronize()
    t0 = time.time()
    r = sv.solve(q0, **kw)
    torch.cuda.synchronize()
    dt = time.time() - t0
    cost = r["cost"][:, :2].sum(dim=1)
    err = (r["q"][:, :, :3] - q_true[:, :, :3]).norm(dim=-1).mean()
    print(f"{name:20s} {dt*1e3:8.1f} ms  lockstep {r['lockstep_iterations']:3d}  mean it {r['counters'][:,0].double().mean():5.1f}"
          f"  cost mean {cost.mean():.4f} max {cost.max():.2f}  c_eq {r['c_eq'].abs().max():.2e}  v_in {r['v_in'].abs().max():.2e}"
          f"  base err {err*1e3:.3f} mm  fail {int(r['status'].sum())}")
