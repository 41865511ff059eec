"""Why a single-view solve of the test set is (not) accepted: per clip status bits, feasibility and iteration counts of
the "default" (no priors) and "data-driven" monocular batches of run_dataset.run_monocular.  GPU box only.

    python scripts/debug/mono_status.py [n_outer] [max_iter]
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt, run_dataset  # noqa: E402
from tests._dataset import build_clip_dir, load_test_set_clips, write_data_driven_dataset  # noqa: E402


def main():
    n_outer = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    max_iter = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    golden = os.path.join(ROOT, "tests", "golden")
    clips = load_test_set_clips(golden)
    tmp = tempfile.mkdtemp()
    root = os.path.join(tmp, "cheetah_videos")
    for c in clips:
        build_clip_dir(root, c["name"], c, monocular_cam=max(int(c["monocular_cam"]), 0))
    models_dir = os.path.join(tmp, "models", "data-driven")
    write_data_driven_dataset(golden, models_dir)
    for mono in (False, True):
        ests = [run_dataset._init(root, t, "cuda", monocular_enable=True) for t in run_dataset.DATASET]
        oks = opt.estimate_kinematics_batch(ests, monocular_constraints=mono, solver_output=False, save=False,
                                            models_dir=models_dir, n_outer=n_outer, max_iter=max_iter)
        print(f"--- {'data-driven' if mono else 'default'} (n_outer {n_outer}, max_iter {max_iter}): accepted {sum(oks)}/{len(oks)}")
        for c, e, ok in zip(clips, ests, oks):
            r = e.result
            mpjpe = float(misc.traj_error(c["positions"].copy(), e.positions.copy(), centered=True).mean())
            mpe = float(misc.traj_error(c["positions"].copy(), e.positions.copy()).mean())
            ce = np.abs(r["c_eq"])
            vi = np.abs(r["v_in"])
            print(f"{c['name']:34s} ok {int(ok)} status {int(r['status'])} it {int(r['counters'][0])}/{int(r['counters'][1])} "
                  f"|c_eq| {ce.max():.2e} (frame {int(ce.max(1).argmax())}) |v_in| {vi.max():.2e} (frame {int(vi.max(1).argmax())}, "
                  f"row {int(vi.max(0).argmax())}) cost {r['cost'].round(1)} MPJPE {mpjpe:.1f} MPE {mpe:.1f}")


if __name__ == "__main__":
    main()
