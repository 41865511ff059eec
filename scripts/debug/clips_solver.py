"""Solver experiments on the 15 rebuilt data/test_set clips (tests/golden/test_set_clips.npz): per clip the objective,
feasibility and distance to the shipped IPOPT trajectory for a few LM / AL settings.  GPU box only.

    python scripts/debug/clips_solver.py [variant ...]
"""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt, engine  # noqa: E402
from tests._dataset import build_clip_dir, inferred_likelihood, load_test_set_clips  # noqa: E402


def main():
    golden = os.path.join(ROOT, "tests", "golden")
    variants = sys.argv[1:] or ["base"]
    for kinetic in (False, True):
        clips = load_test_set_clips(golden, kinetic=kinetic)
        tmp = tempfile.mkdtemp()
        root = os.path.join(tmp, "cheetah_videos")
        ests = []
        for c in clips:
            build_clip_dir(root, c["name"], c, monocular_cam=max(int(c["monocular_cam"]), 0),
                           resolution=[1280, 720] if kinetic else None,
                           likelihood=inferred_likelihood(c) if kinetic else None)
            ests.append(opt.init_trajectory(root, c["name"], c["animal"], kinetic_dataset=kinetic, kinematic_model=True))
        q0s = [opt.initial_guess(e) for e in ests]
        batch, q0 = opt.pack_ragged(ests, q0s, "cuda")
        S, N = q0.shape[:2]
        qs = torch.zeros_like(q0)
        for i, c in enumerate(clips):
            n = c["q"].shape[0]
            qs[i, :n] = torch.as_tensor(c["q"], device="cuda")
            qs[i, n:] = qs[i, n - 1]
        sv = engine.Solver(ests[0].model, batch)
        o = sv.evaluate(qs, constraints=True)
        ship = o["cost"].cpu().numpy()
        ship_ceq = o["c_eq"].abs().amax(dim=(1, 2)).cpu().numpy()
        for v in variants:
            kw = dict(max_iter=40, n_outer=8)
            start = q0
            if v == "from_shipped":
                start = qs
            elif v.startswith("outer"):
                kw["n_outer"] = int(v[5:])
                kw["mu_schedule"] = (1e4, 1e5, 1e6, 1e6, 1e7, 1e7, 1e7, 1e7)
            elif v.startswith("gnc"):
                # graduated non-convexity: first a wide (nearly quadratic) loss, then the reference's break points
                k = float(v[3:] or 4)
                batch.loss = (3.0 * k, 10.0 * k, 20.0 * k)
                r0 = sv.solve(q0, max_iter=40, n_outer=2)
                batch.loss = (3.0, 10.0, 20.0)
                start = r0["q"]
            elif v == "ftol8":
                kw["ftol"] = 1e-8
                kw["max_iter"] = 80
            torch.cuda.synchronize()
            import time
            t0 = time.time()
            res = sv.solve(start, **kw)
            torch.cuda.synchronize()
            dt = time.time() - t0
            out = sv.evaluate(res["q"], pose=True)
            cost = res["cost"].cpu().numpy()
            print(f"--- variant {v} ({'kinetic' if kinetic else 'acinoset'}): {dt:.3f} s, lockstep iterations {res['lockstep_iterations']}")
            for i, c in enumerate(clips):
                n = c["q"].shape[0]
                pos = out["pose"][i, :n].cpu().numpy()
                mpjpe = misc.traj_error(pos.copy(), c["positions"].copy(), centered=True).mean()
                mpe = misc.traj_error(pos.copy(), c["positions"].copy()).mean()
                ceq = float(res["c_eq"][i].abs().max())
                vin = float(res["v_in"][i].abs().max())
                it = res["counters"][i].cpu().numpy()
                print(f"{c['name']:42s} obj {1e-3 * (cost[i, 0] + cost[i, 1]):9.4f} (meas {cost[i, 0]:9.1f} model {cost[i, 1]:8.1f}) "
                      f"shipped q {1e-3 * (ship[i, 0] + ship[i, 1]):9.4f} (meas {ship[i, 0]:9.1f} model {ship[i, 1]:8.1f}) "
                      f"MPJPE {mpjpe:6.1f} MPE {mpe:6.1f} |c_eq| {ceq:.1e} (shipped {ship_ceq[i]:.1e}) |v_in| {vin:.1e} it {it[0]}/{it[1]} "
                      f"status {int(res['status'][i])}")


if __name__ == "__main__":
    main()
