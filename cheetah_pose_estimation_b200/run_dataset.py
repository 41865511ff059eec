"""Mirror of the reference's batch driver `run_dataset.py` (flags :1334-1383, clip list :1387-1398, `run_acinoset`
:1234-1331, `run_monocular` :1143-1231, `run_monocular_all` :918-1015, `run_data_driven_ablation_study` :635-717):
same command line, same clip list, same output directories and files.

What changes is the execution shape.  The reference reconstructs one clip after the other (a Pyomo model + one IPOPT
run each, `gc.collect()` in between); here every loop over clips / camera choices / test types becomes ONE ragged
batch per test type (`acinoset_opt.estimate_kinematics_batch`): clips x hypotheses are the sequences of one LM solve,
dealt to the GPUs of the box by frames x cameras when launched under torchrun (one process per GPU, no inter-GPU
traffic inside the solve, one NCCL gather of the results), and rank 0 writes each clip's `fte.pickle` /
`cam*_fte.csv`.

    python -m cheetah_pose_estimation_b200.run_dataset --root_dir <...>/cheetah_videos --out_dir_prefix <out> \\
        --run_monocular --clean
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m \\
        cheetah_pose_estimation_b200.run_dataset ... (same flags)

Out of scope here: the "physics-based" test type (`estimate_kinetics`, full 54-row dynamics --
its `.robot` blobs are missing from the reference checkout), `--run_kinetic`, `--run_grid_search` (AR windows above
the kernel's CPE_PRIOR_MAX_WINDOW) and the matplotlib analysis (`distance_vs_error`, `kinetic_analysis`); those flags
are accepted and report that they are not part of this path.  `dataset_post_process` keeps the numbers (MPE / MPJPE
table, `dataset_results.csv`) and drops the plots.
"""
import json
import os
import pickle
from argparse import ArgumentParser
from time import time
from typing import List, Optional, Tuple

import numpy as np

from . import acinoset_misc as misc
from . import acinoset_opt as opt

# run_dataset.py:1387-1398
DATASET = (
    ("jules", "2017_12_09/bottom", "flick2"),
    ("jules", "2019_03_09", "flick1"),
    ("phantom", "2019_03_03", "run"),
    ("phantom", "2017_09_02/top", "run1_2"),
    ("jules", "2017_08_29/top", "run1_2"),
    ("phantom", "2017_08_29/top", "run1_1"),
    ("jules", "2017_08_29/top", "run1_1"),
    ("jules", "2017_09_02/top", "run1"),
    ("phantom", "2019_03_07", "run"),
    ("jules", "2017_09_02/bottom", "run2"),
)
MONOCULAR_CAM_OVERRIDES = [0, 0, 0, 3, 3, 3, 5, 0, 3, 0]           # run_dataset.py:1401
# run_dataset.py:921-932
CAM_SPACE = ((0, 2, 3, 4), (0, 1, 2, 3, 4, 5), (0, 1, 2, 3), (0, 1, 2, 3, 4, 5), (0, 1, 2, 3, 4, 5), (0, 1, 2, 3, 4, 5),
             (0, 1, 2, 3, 4, 5), (0, 1, 2, 3, 4, 5), (0, 1, 2, 3, 4, 5), (0, 1, 2, 3, 4, 5))
# run_dataset.py:1245-1298 (the entries that are not commented out)
MANUALLY_SELECTED_FRAMES = {
    "2019_03_03/phantom/run": (100, 220), "2019_03_09/lily/run": (80, 170), "2017_08_29/top/phantom/run1_1": (20, 160),
    "2017_12_21/top/lily/run1": (10, 105), "2017_12_21/bottom/jules/flick2_2": (5, 150),
    "2017_12_10/top/zorro/flick1": (115, 210), "2017_12_10/bottom/zorro/flick2": (5, 140),
    "2017_09_03/bottom/zorro/run2_1": (130, 270), "2017_12_09/bottom/phantom/run2": (20, 115),
    "2017_09_03/bottom/zorro/run2_3": (5, 150), "2017_08_29/top/jules/run1_1": (10, 110),
    "2017_09_02/top/jules/run1": (10, 110), "2019_03_07/menya/run": (60, 160), "2017_09_02/top/phantom/run1_2": (20, 160),
    "2019_03_07/phantom/run": (100, 200), "2019_02_27/romeo/run": (40, 150), "2019_02_27/romeo/flick": (10, 150),
    "2017_08_29/top/jules/run1_2": (30, 130), "2017_12_16/top/cetane/run1": (110, 210), "2019_02_27/kiara/run": (20, 100),
    "2017_09_02/bottom/jules/run2": (50, 160), "2017_09_03/bottom/zorro/run2_2": (32, 141),
    "2019_03_09/jules/flick1": (40, 160), "2017_09_03/bottom/zorro/flick2": (10, 100),
    "2017_08_29/bottom/zorro/flick2": (75, 135), "2017_12_09/bottom/jules/flick2": (5, 75),
    "2017_12_17/bottom/zorro/flick2": (5, 145),
}


def _data_path(test_run):
    return f"{test_run[1]}/{test_run[0]}/{test_run[2]}"


def _rank():
    import torch.distributed as dist
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def _init(root_dir, test_run, device, **kw):
    return opt.init_trajectory(root_dir, _data_path(test_run), test_run[0], solver_path=None, kinetic_dataset=False,
                               kinematic_model=True, include_camera_constraints=True, device=device, **kw)


def run_monocular(root_dir: str, dir_prefix: str, test_set: Tuple, cam_overrides: Optional[List[int]] = None,
                  device: str = "cuda", models_dir: Optional[str] = None, test_types=None):
    """run_dataset.py:1143-1231.  Test types "ground-truth" (all cameras), "default" (one camera) and "data-driven"
    (one camera + GMM pose prior + AR(4) motion model): one ragged batch each over all clips of `test_set`.
    Returns {test type: [bool per clip]}."""
    time0 = time()
    test_types = test_types or ("ground-truth", "default", "data-driven", "physics-based")
    out = {}
    for t in test_types:
        if t == "physics-based":
            if _rank() == 0:
                print("physics-based (estimate_kinetics): not part of this path -- skipped")
            continue
        ests = []
        for idx, test_run in enumerate(test_set):
            cam = None if cam_overrides is None else cam_overrides[idx]
            if t == "ground-truth":
                ests.append(_init(root_dir, test_run, device))
            else:
                ests.append(_init(root_dir, test_run, device, monocular_enable=True, override_monocular_cam=cam))
        out[t] = opt.estimate_kinematics_batch(ests, solver_output=False, monocular_constraints=(t == "data-driven"),
                                               out_dir_prefix=dir_prefix, models_dir=models_dir)
    if _rank() == 0:
        print(f"Run through all videos took {time() - time0:.2f}s")
    return out


def run_monocular_all(root_dir: str, dir_prefix: str, test_set: Tuple, device: str = "cuda",
                      models_dir: Optional[str] = None, cam_space=None):
    """run_dataset.py:918-1015: every clip x every camera of `cam_space` x ("default", "data-driven") -- 56 camera
    hypotheses per test type for the shipped list, one ragged batch each."""
    time0 = time()
    cam_space = cam_space or CAM_SPACE
    out = {}
    for t in ("default", "data-driven"):
        ests = [_init(root_dir, test_run, device, monocular_enable=True, override_monocular_cam=cam)
                for i, test_run in enumerate(test_set) for cam in cam_space[i]]
        out[t] = opt.estimate_kinematics_batch(ests, solver_output=False, monocular_constraints=(t == "data-driven"),
                                               out_dir_prefix=dir_prefix, models_dir=models_dir)
    if _rank() == 0:
        print("physics-based (estimate_kinetics): not part of this path -- skipped")
        print(f"Run through all videos took {time() - time0:.2f}s")
    return out


def run_acinoset(root_dir: str, out_dir_prefix: str, device: str = "cuda", tests: Optional[List[str]] = None):
    """run_dataset.py:1234-1331: every directory of `<root>/../test_set.pickle` ("test_dirs") that has manually selected
    frames, skeleton "acinoset", pairwise pseudo-measurements for the "flick" clips.  Two batches: W = 1 (runs) and
    W = 3 (flicks).  Returns {data_path: bool}."""
    if tests is None:
        path = os.path.join(root_dir.split("/cheetah_videos")[0], "test_set.pickle")
        try:
            import dill
            with open(path, "rb") as f:
                tests = dill.load(f)["test_dirs"]
        except ImportError:
            with open(path, "rb") as f:
                tests = pickle.load(f)["test_dirs"]
    time0 = time()
    if _rank() == 0:
        print("Run reconstruction on all videos...")
    groups = {False: [], True: []}
    for test_vid in tests:
        current_dir = test_vid.split("/cheetah_videos/")[1] if "/cheetah_videos/" in test_vid else test_vid
        if current_dir not in MANUALLY_SELECTED_FRAMES:
            continue
        start, end = MANUALLY_SELECTED_FRAMES[current_dir]
        ppm = "flick" in current_dir
        est = opt.init_trajectory(root_dir, current_dir, "acinoset", solver_path=None, kinetic_dataset=False,
                                  start_frame=start, end_frame=end, shutter_delay_estimation=False, enable_ppm=ppm,
                                  kinematic_model=True, device=device)
        groups[ppm].append((current_dir, est))
    result = {}
    for ppm, items in groups.items():
        if not items:
            continue
        oks = opt.estimate_kinematics_batch([e for _, e in items], solver_output=False, out_dir_prefix=out_dir_prefix)
        for (name, _), ok in zip(items, oks):
            result[name] = ok
            if not ok and _rank() == 0:
                print(f"Failed to find optimal solution for test: {name}")
    if _rank() == 0:
        print(f"Run through all videos took {time() - time0:.2f}s")
    return result


def run_data_driven_ablation_study(root_dir: str, dir_prefix: str, test_set: Tuple, device: str = "cuda",
                                   models_dir: Optional[str] = None):
    """run_dataset.py:635-717: the four (disable_pose_prior, disable_motion_prior) scenarios, one batch each over all
    clips; MPJPE / MPE against the multi-view reconstruction in `<dir_prefix>/<clip>/fte_kinematic` (mean and std over
    the clips), written to `data_driven_ablation.pickle` by rank 0."""
    results = {"mpjpe": [], "mpjpe_std": [], "mpe": [], "mpe_std": [], "optimisation_time": [],
               "optimisation_time_std": []}
    for scenario in ((True, True), (True, False), (False, True), (False, False)):
        ests = [_init(root_dir, test_run, device, monocular_enable=True) for test_run in test_set]
        opt.estimate_kinematics_batch(ests, monocular_constraints=True, solver_output=False,
                                      disable_pose_prior=scenario[0], disable_motion_prior=scenario[1],
                                      out_dir_prefix=dir_prefix, models_dir=models_dir, save=False)
        mpe, mpjpe, tt = [], [], []
        for test_run, e in zip(test_set, ests):
            with open(os.path.join(dir_prefix, _data_path(test_run), "fte_kinematic", "fte.pickle"), "rb") as f:
                gt = pickle.load(f)["positions"]
            mpjpe.append(float(misc.traj_error(gt.copy(), e.positions.copy(), centered=True).mean()))
            mpe.append(float(misc.traj_error(gt.copy(), e.positions.copy()).mean()))
            tt.append(e.opt_time_s)
        for k, v in (("mpjpe", mpjpe), ("mpe", mpe), ("optimisation_time", tt)):
            results[k].append(float(np.mean(v)))
            results[k + "_std"].append(float(np.std(v)))
    if _rank() == 0:
        with open(os.path.join(dir_prefix, "data_driven_ablation.pickle"), "wb") as f:
            pickle.dump(results, f)
        print(results)
    return results


def dataset_post_process(root_dir: str, dir_prefix: str, test_set: Tuple, cam_overrides: Optional[List[int]] = None):
    """The table of run_dataset.py:365-633 without the plots and without the physics-based column: mean position error
    (MPE) and centred error (MPJPE), mm, of the single-view ("default") and data-driven reconstructions against the
    multi-view one, per clip; written as `<dir_prefix>/dataset_results.csv` (clip / test-type column header, rows mpe,
    mpjpe) and returned as a DataFrame."""
    import pandas as pd
    cols, mpe_row, mpjpe_row = [], [], []
    for idx, test_run in enumerate(test_set):
        data_path = _data_path(test_run)
        with open(os.path.join(root_dir, data_path, "metadata.json"), "r", encoding="utf-8") as f:
            metadata = json.load(f)
        cam_idx = metadata["monocular_cam"] if cam_overrides is None else cam_overrides[idx]
        with open(os.path.join(dir_prefix, data_path, "fte_kinematic", "fte.pickle"), "rb") as f:
            gt = pickle.load(f)["positions"]
        for name, sub in (("default", f"fte_kinematic_orig_{cam_idx}"), ("data-driven", f"fte_kinematic_{cam_idx}")):
            p = os.path.join(dir_prefix, data_path, sub, "fte.pickle")
            cols.append((data_path, name))
            if not os.path.exists(p):
                mpe_row.append(np.nan)
                mpjpe_row.append(np.nan)
                continue
            with open(p, "rb") as f:
                pos = pickle.load(f)["positions"]
            mpe_row.append(round(float(misc.traj_error(gt.copy(), pos.copy()).mean()), 1))
            mpjpe_row.append(round(float(misc.traj_error(gt.copy(), pos.copy(), centered=True).mean()), 1))
    df = pd.DataFrame([mpe_row, mpjpe_row], index=["mpe", "mpjpe"], columns=pd.MultiIndex.from_tuples(cols))
    if _rank() == 0:
        df.to_csv(os.path.join(dir_prefix, "dataset_results.csv"))
        print(df.T)
    return df


def _maybe_init_distributed():
    """One process per GPU under torchrun (RANK / LOCAL_RANK / WORLD_SIZE in the environment); a plain `python -m`
    run stays single-process."""
    import torch
    import torch.distributed as dist
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and not dist.is_initialized():
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        return f"cuda:{local}"
    return "cuda"


def build_parser():
    """The reference's flags, verbatim (run_dataset.py:1335-1383)."""
    hlp = ("A flag to run a subset of AcinoSet that forms the dataset used for train/test procedures in monocular 3D "
           "reconstruction.")
    parser = ArgumentParser(description="Monocular 3D Reconstruction of Cheetahs in the Wild")
    parser.add_argument("--root_dir", type=str, help="The data directory path to all cheetah videos.")
    parser.add_argument("--out_dir_prefix", type=str, help="The directory to place the outputs.")
    parser.add_argument("--override_default_cam", action="store_true",
                        help="A flag to override the default cameras used for monocular reconstruction.")
    for flag in ("--run_acinoset", "--run_monocular", "--run_kinetic", "--run_analysis", "--run_grid_search",
                 "--run_physics_based_ablation_study", "--run_data_driven_ablation_study"):
        parser.add_argument(flag, action="store_true", help=hlp)
    parser.add_argument("--clean", action="store_true",
                        help="A flag that enables the re-generation of the results prior to post-processing for output "
                             "metrics. Note you will need to have the original AcinoSet data and the optimisation "
                             "framework setup to do this.")
    return parser


def main(argv=None):
    args = build_parser().parse_args(argv)
    device = _maybe_init_distributed()
    root_dir = os.path.normpath(args.root_dir)
    out_dir = os.path.normpath(args.out_dir_prefix)
    overrides = MONOCULAR_CAM_OVERRIDES if args.override_default_cam else None
    if args.run_acinoset:
        print("Running AcinoSet subset")
        if args.clean:
            run_acinoset(root_dir, out_dir, device)
    if args.run_monocular:
        print("Running monocular procedure")
        if args.clean:
            run_monocular(root_dir, out_dir, DATASET, overrides, device)
        dataset_post_process(root_dir, out_dir, DATASET, overrides)
    if args.run_analysis:
        print("Running monocular analysis")
        if args.clean:
            run_monocular_all(root_dir, out_dir, DATASET, device)
    if args.run_data_driven_ablation_study:
        print("Running data driven ablation study")
        run_data_driven_ablation_study(root_dir, out_dir, DATASET, device)
    for flag in ("run_kinetic", "run_grid_search", "run_physics_based_ablation_study"):
        if getattr(args, flag):
            print(f"--{flag}: not part of this path (SURVEY.md section 8: out of scope)")
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
