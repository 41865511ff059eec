"""The LM leg of bench.py on its own: S sequences x 500 frames x 6 cameras, 3 multiplier rounds of at most 20 iterations,
`n_split` sub-batches on their own streams.  usage: time_lm.py [S] [n_split] [n_outer]   (env CPE_K2_TEAMS, CPE_LIB_PATH)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from cheetah_pose_estimation_b200 import engine, synthetic  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 1776
n_split = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n_outer = int(sys.argv[3]) if len(sys.argv) > 3 else 3
model = engine.Model("cuda:0")
batch, q_true, q_init = synthetic.make_device_problem(S, 500, 6, seed=0, device="cuda:0", dlc_native=True, n_unique=S)
engine.solve_split(model, batch, q_init, n_split=n_split, max_iter=2, n_outer=1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
res = engine.solve_split(model, batch, q_init, n_split=n_split, max_iter=20, n_outer=n_outer)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"S {S} split {n_split} teams {os.environ.get('CPE_K2_TEAMS', '5')} lib {os.path.basename(os.environ.get('CPE_LIB_PATH', 'default'))}: "
      f"{ms:8.1f} ms  {S / ms * 1e3:7.1f} sequences/s  lockstep {res['lockstep_iterations']}  "
      f"cost {float(res['cost'][:, :2].sum()):.9e}  max|c_eq| {float(res['c_eq'].abs().max()):.2e}")
