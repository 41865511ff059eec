"""Eager loop against CUDA-graph replay of the LM iteration: one real 30-frame clip (S = 1) and the config-2 batch."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from cheetah_pose_estimation_b200 import engine, skeleton as sk

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
g = np.load(os.path.join(ROOT, "tests", "golden", "test_set_clips.npz"))
idx = [i for i in range(len(g["names"])) if not bool(g[f"{i}_kinetic"])]
model = engine.Model("cuda:0")


def pack(sel, copies):
    n_pad = max(g[f"{i}_q"].shape[0] for i in sel); c_pad = max(g[f"{i}_cams"].shape[0] for i in sel)
    S0 = len(sel)
    meas = np.zeros((S0, n_pad, c_pad, 24, 1, 4), np.float32); cams = np.zeros((S0, c_pad, sk.CAM_STRIDE))
    cvec = np.zeros((S0, 24, 5, 3)); q0 = np.zeros((S0, n_pad, 54)); nf = np.zeros(S0, np.int32); nc = np.zeros(S0, np.int32); h = np.zeros(S0)
    for j, i in enumerate(sel):
        det, cam, q = g[f"{i}_detections"], g[f"{i}_cams"], g[f"{i}_q"]
        n, c = det.shape[:2]
        meas[j, :n, :c, :, 0, 0:2] = det; meas[j, :n, :c, :, 0, 2] = 1.0
        cams[j, :c, :20] = cam; cams[j, :c, 20] = 1.0; cams[j, c:] = cams[j, 0]
        L, R, _ = sk.animal(str(g[f"{i}_animal"])); cvec[j] = sk.marker_chains(L, R)[1]
        q0[j, :n, 0:3] = q[:, 0:3]; q0[j, :n, 5::3] = q[:, 5:6]; q0[j, n:] = q0[j, n - 1]
        nf[j], nc[j], h[j] = n, c, 1.0 / float(g[f"{i}_fps"])
    rep = lambda x: np.tile(x, (copies,) + (1,) * (x.ndim - 1))
    b = engine.Batch.from_numpy(rep(meas), rep(cams), rep(cvec), rep(h), n_frames=rep(nf), n_cams=rep(nc))
    return b, torch.as_tensor(rep(q0), device="cuda:0")


for name, sel, copies in (("one clip (S = 1)", idx[6:7], 1), ("10 clips", idx, 1), ("c2: 10 clips x 45", idx, 45)):
    b, q0 = pack(sel, copies)
    sv = engine.Solver(model, b)
    out = {}
    for graph in (False, True):
        sv.solve(q0, max_iter=40, n_outer=8, graph=graph)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(3):
            res = sv.solve(q0, max_iter=40, n_outer=8, graph=graph)
        torch.cuda.synchronize(); out[graph] = ((time.perf_counter() - t0) / 3, res)
    same = torch.equal(out[False][1]["q"], out[True][1]["q"])
    print(f"{name:22s} eager {1e3 * out[False][0]:8.2f} ms   graph {1e3 * out[True][0]:8.2f} ms   "
          f"lockstep iterations {out[True][1]['lockstep_iterations']}   identical {same}")
