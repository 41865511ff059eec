"""CUDA-event timing of K1 (residual + Jacobian) in both flavours. usage: time_eval.py [S] [N] [C]
Knob: CPE_EVAL_CFG (launch-bounds variant)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from cheetah_pose_estimation_b200 import engine, synthetic

S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
N = int(sys.argv[2]) if len(sys.argv) > 2 else 500
C = int(sys.argv[3]) if len(sys.argv) > 3 else 6
dev = "cuda:0"
model = engine.Model(dev)
batch, q_true, _ = synthetic.make_device_problem(S, N, C, seed=0, device=dev, dlc_native=True)
sv = engine.Solver(model, batch, with_solver=False)
q = (q_true + 0.05 * torch.randn_like(q_true)).contiguous()
cost = torch.zeros(S, 4, dtype=torch.float64, device=dev)
for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
    for jac in (True, False):
        bufs = {"resid": torch.empty(S, N, C, 24, 1, 2, dtype=dt, device=dev)}
        if jac:
            bufs["jac"] = torch.empty(S, N, C, model.jac_width, dtype=dt, device=dev)
        for _ in range(3):
            sv.evaluate_launch(q, bufs, cost)
        torch.cuda.synchronize()
        ms = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sv.evaluate_launch(q, bufs, cost, record_mid=e1)
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        m = sorted(ms)[len(ms) // 2]
        print(f"cfg={os.environ.get('CPE_EVAL_CFG','0')} {name} jac={int(jac)}: {m:8.3f} ms  {S*N*C/m/1e3:8.1f} M frame.views/s  cost0={float(cost[:,0].sum()):.6e}")
        del bufs
