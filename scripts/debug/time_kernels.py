"""CUDA-event timing of each product kernel at bench scale. usage: time_kernels.py [S] [N]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from cheetah_pose_estimation_b200 import engine, synthetic

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 500
model = engine.Model("cuda:0")
batch, q_true, q_init = synthetic.make_device_problem(S, N, 6, seed=0, device="cuda:0")
sv = engine.Solver(model, batch)
q = (q_true + 0.05 * torch.randn_like(q_true)).contiguous()
damp = torch.full((S,), 1e-3, dtype=torch.float64, device="cuda:0")

def t(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

frames = S * N
for name, fn in (("eval (cost only)", lambda: sv.evaluate(q)), ("assemble", lambda: sv.assemble(q, damp)),
                 ("band_solve", lambda: sv.band_solve()), ("step_apply", lambda: sv.step_apply(q, damp))):
    ms = t(fn)
    print(f"{name:18s} {ms:9.3f} ms   {ms*1e6/frames:7.2f} ns/frame")
