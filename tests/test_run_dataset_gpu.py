"""BASELINE config 2 through the mirrored `run_dataset` entry points: every multi-view clip of the reference's
data/test_set (cameras / detections rebuilt from the shipped outputs, tests/golden/test_set_clips.npz) reconstructed
as ONE ragged batch per test type, each clip checked against its own single-clip run bit for bit and against the
shipped IPOPT trajectory."""
import os
import pickle

import numpy as np
import pytest

from tests._dataset import (build_clip_dir, inferred_likelihood, load_test_set_clips, write_data_driven_dataset,
                            write_pairwise_files)

pytestmark = pytest.mark.gpu


def _build_root(tmp_path, clips):
    root = os.path.join(str(tmp_path), "cheetah_videos")
    for c in clips:
        # AcinoSet clips: every detection on (the all-on mask of the known-answer tests); kinetic dataset: a fifth of its
        # points are unlabelled markers parked hundreds of pixels away -- gate those (tests/_dataset.inferred_likelihood)
        build_clip_dir(root, c["name"], c, monocular_cam=max(int(c["monocular_cam"]), 0),
                       resolution=[1280, 720] if bool(c["kinetic"]) else None,
                       likelihood=inferred_likelihood(c) if bool(c["kinetic"]) else None)
    return root


def test_run_monocular_batches_the_whole_test_set(golden_dir, tmp_path):
    """`run_dataset.run_monocular` (run_dataset.py:1143-1231) on the 10 clips of the reference's `dataset` tuple:
    "ground-truth" (4-6 cameras, ragged 30-57 frames), "default" and "data-driven" (one camera) -- three batches.
    (i) every output file of every clip exists with the reference's keys; (ii) each clip's multi-view trajectory is
    bit-identical to a single-clip solve, padded like the batch or not padded at all;
    (iii) distance and cost relative to the shipped IPOPT optimum, per clip, inside the measured envelope."""
    import torch
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt, engine, run_dataset
    clips = load_test_set_clips(golden_dir)
    assert [c["name"] for c in clips] == [run_dataset._data_path(t) for t in run_dataset.DATASET]
    root = _build_root(tmp_path, clips)
    out = os.path.join(str(tmp_path), "out")
    models_dir = os.path.join(str(tmp_path), "models", "data-driven")
    write_data_driven_dataset(golden_dir, models_dir)
    oks = run_dataset.run_monocular(root, out, run_dataset.DATASET, None, models_dir=models_dir)
    assert set(oks) == {"ground-truth", "default", "data-driven"}
    assert all(oks["ground-truth"]), oks
    ref_keys = sorted(np.load(os.path.join(golden_dir, "save_ref.npz"))["jules_keys"])
    n_pad = max(c["detections"].shape[0] for c in clips)
    c_pad = max(c["detections"].shape[1] for c in clips)
    rows = []
    for c, test_run in zip(clips, run_dataset.DATASET):
        name, (N, C) = c["name"], c["detections"].shape[:2]
        d = pickle.load(open(os.path.join(out, name, "fte_kinematic", "fte.pickle"), "rb"))
        assert sorted(d.keys()) == ref_keys
        assert d["q"].shape == (N, 54) and d["meas_err"].shape == (N, C, 24, 2, 1) and d["positions"].shape == (N, 24, 3)
        for i in range(C):
            assert os.path.exists(os.path.join(out, name, "fte_kinematic", f"cam{i + 1}_fte.csv"))
        # (ii) the same clip on its own
        est = run_dataset._init(root, test_run, "cuda")
        assert opt.estimate_kinematics_batch([est], solver_output=False, save=False, pad=(n_pad, c_pad))[0]
        assert np.array_equal(est.q, d["q"]), name
        assert np.array_equal(est.meas_err, d["meas_err"]) and est.get_objective_cost() == d["obj_cost"]
        est1 = run_dataset._init(root, test_run, "cuda")
        assert opt.estimate_kinematics(est1, solver_output=False)
        # unpadded (C cameras, N frames): cpe_batch.n_cams / n_frames keep every sum in the order of the unpadded clip
        assert np.array_equal(est1.q, d["q"]) and est1.get_objective_cost() == d["obj_cost"], name
        # (iii) against the shipped IPOPT optimum: same objective (all-on mask) evaluated at the shipped q
        batch, _ = opt.pack_ragged([est], [c["q"]], "cuda")
        sv = engine.Solver(est.model, batch, with_solver=False)
        o = sv.evaluate(torch.as_tensor(c["q"][None].copy(), device="cuda"), constraints=True)
        c_ship = o["cost"][0].cpu().numpy()
        obj_ship = 1e-3 * (c_ship[0] + c_ship[1])
        mpjpe = float(misc.traj_error(d["positions"].copy(), c["positions"].copy(), centered=True).mean())
        mpe = float(misc.traj_error(d["positions"].copy(), c["positions"].copy()).mean())
        rows.append((name, N, C, d["obj_cost"], obj_ship, mpjpe, mpe, float(np.abs(est.result["c_eq"]).max()),
                     float(o["c_eq"].abs().max())))
    print("\nclip | N | C | obj (ours) | obj (shipped q, same objective) | MPJPE mm | MPE mm | max|c_eq| ours | shipped")
    for r in rows:
        print(f"{r[0]:32s} {r[1]:3d} {r[2]} {r[3]:9.4f} {r[4]:9.4f} {r[5]:6.1f} {r[6]:6.1f} {r[7]:.1e} {r[8]:.1e}")
    # Measured envelope (profiles/r2b_run_dataset.log): under the all-on mask our optimum costs
    # 1-5 % LESS than the shipped trajectory and holds the joint equalities as tightly (<= 3.2e-5 against 5e-8 .. 8.5e-5);
    # it sits 13-30 mm (centred) from the shipped one for the runs and 28 / 61 mm for the two tail-flick clips -- the
    # shipped optimum belongs to a different objective (its likelihood mask is lost), so the north star's 1 mm bar is
    # checked against the independent NLP witness on identical inputs (tests/test_nlp_witness.py), not here.
    for r in rows:
        assert 0.93 * r[4] <= r[3] <= r[4] * 1.005, r          # two-sided: no worse, and not a different problem
        assert r[7] < 5e-5, r                                  # joint equalities at the level of the shipped optima
        flick = "flick" in r[0]
        assert r[5] < (70.0 if flick else 35.0) and r[6] < (65.0 if flick else 40.0), r
    # single-view outputs: directory names of the reference, per clip's monocular camera
    for c, ok_def, ok_dd in zip(clips, oks["default"], oks["data-driven"]):
        cam = max(int(c["monocular_cam"]), 0)
        if ok_def:
            assert os.path.exists(os.path.join(out, c["name"], f"fte_kinematic_orig_{cam}", "fte.pickle"))
        if ok_dd:
            assert os.path.exists(os.path.join(out, c["name"], f"fte_kinematic_{cam}", "fte.pickle"))
    # every single-view solve is accepted (the two tail-flick clips need the extra multiplier rounds of
    # engine.Solver.solve_steps: profiles/r2q_mono_status.log before, r2q_mono_status2.log after)
    assert all(oks["default"]) and all(oks["data-driven"]), oks
    df = run_dataset.dataset_post_process(root, out, run_dataset.DATASET)
    assert os.path.exists(os.path.join(out, "dataset_results.csv")) and df.shape == (2, 20)
    assert bool(np.isfinite(df.to_numpy()).all())
    dd = df.loc["mpjpe", [c for c in df.columns if c[1] == "data-driven"]].to_numpy()
    de = df.loc["mpjpe", [c for c in df.columns if c[1] == "default"]].to_numpy()
    # the reference's own table (data/test_set/dataset_results.csv): data-driven MPJPE 49.7 .. 115.2 mm, mean 78.8;
    # default 77.8 .. 176.7, mean 118.6.  Measured here: 50.1 .. 96.6, mean 69.8; default mean 135 (priors fitted on the
    # shipped trajectories of the test set, the only pose data in the checkout).
    assert dd.max() < 115.2 and dd.mean() < 78.8 and dd.mean() < de.mean(), (dd, de)


def test_kinetic_dataset_clips_as_one_pinhole_batch(golden_dir, tmp_path):
    """The five kinetic-dataset clips (4 pinhole cameras, sigma = 7 px, camera multipliers [1, 1, .6, .6], 200 fps,
    detections kept in double): one batch through `estimate_kinematics_batch`, checked like the AcinoSet ones."""
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt
    clips = load_test_set_clips(golden_dir, kinetic=True)
    assert len(clips) == 5
    root = _build_root(tmp_path, clips)
    ests = [opt.init_trajectory(root, c["name"], c["animal"], kinetic_dataset=True, kinematic_model=True) for c in clips]
    assert all(e.meas.dtype == np.float64 for e in ests)
    oks = opt.estimate_kinematics_batch(ests, solver_output=False)
    assert all(oks), oks
    for c, e in zip(clips, ests):
        d = pickle.load(open(os.path.join(root, c["name"], "fte_kinematic", "fte.pickle"), "rb"))
        assert np.array_equal(d["q"], e.q)
        mpjpe = float(misc.traj_error(d["positions"].copy(), c["positions"].copy(), centered=True).mean())
        mpe = float(misc.traj_error(d["positions"].copy(), c["positions"].copy()).mean())
        print(f"{c['name']}: obj {d['obj_cost']:.4f} (shipped {float(c['obj_cost']):.4f}), MPJPE {mpjpe:.1f} mm, MPE {mpe:.1f} mm")
        # 4 pinhole cameras at sigma = 7 px leave the pose loosely determined: 32-70 mm centred / 59-124 mm absolute from
        # the shipped optimum (whose mask is lost), with a 4-8 % LOWER objective at the shipped level of feasibility
        assert mpjpe < 80.0 and mpe < 140.0
        assert np.abs(e.result["c_eq"]).max() < 5e-5


def test_run_acinoset_groups_runs_and_flicks(golden_dir, tmp_path, monkeypatch):
    """`run_dataset.run_acinoset` (run_dataset.py:1234-1331): explicit start / end frames, the generic "acinoset"
    skeleton, pairwise pseudo-measurements (W = 3) for the "flick" clips -- runs and flicks form two batches."""
    from cheetah_pose_estimation_b200 import run_dataset
    clips = [c for c in load_test_set_clips(golden_dir) if c["name"] in
             ("2019_03_09/jules/flick1", "2017_12_09/bottom/jules/flick2", "2017_08_29/top/jules/run1_1",
              "2019_03_07/phantom/run")]
    root = _build_root(tmp_path, clips)
    frames = {}
    for c in clips:
        s0 = int(c["start_frame"])
        frames[c["name"]] = (s0, s0 + c["detections"].shape[0])
        if "flick" in c["name"]:
            write_pairwise_files(root, c["name"], c)
    monkeypatch.setattr(run_dataset, "MANUALLY_SELECTED_FRAMES", frames)
    out = os.path.join(str(tmp_path), "out")
    res = run_dataset.run_acinoset(root, out, tests=[f"/data/cheetah_videos/{c['name']}" for c in clips] + ["x/cheetah_videos/unknown/clip"])
    assert set(res) == set(frames) and all(res.values()), res
    for c in clips:
        d = pickle.load(open(os.path.join(out, c["name"], "fte_kinematic", "fte.pickle"), "rb"))
        W = 3 if "flick" in c["name"] else 1
        assert d["meas_err"].shape == c["detections"].shape[:2] + (24, 2, W)


def test_run_dataset_cli_on_two_gpus_matches_one(golden_dir, tmp_path):
    """The command line of the reference (`--root_dir --out_dir_prefix --run_monocular --clean`) under torchrun with two
    ranks: clips are dealt to the GPUs by frames x cameras, solved without inter-GPU traffic, gathered over NCCL and
    written by rank 0 -- every clip's trajectory is bit-identical to the single-process run."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    from cheetah_pose_estimation_b200 import run_dataset
    clips = load_test_set_clips(golden_dir)
    root = _build_root(tmp_path, clips)
    write_data_driven_dataset(golden_dir, os.path.join(str(tmp_path), "models", "data-driven"))
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PYTHONPATH=repo + os.pathsep + os.environ.get("PYTHONPATH", ""))
    outs = {}
    for n in (1, 2):
        outs[n] = os.path.join(str(tmp_path), f"out{n}")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
               "127.0.0.1", "--master-port", str(29500 + n), "-m", "cheetah_pose_estimation_b200.run_dataset",
               "--root_dir", root, "--out_dir_prefix", outs[n], "--run_monocular", "--clean"]
        r = subprocess.run(cmd, cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for c in clips:
        cam = max(int(c["monocular_cam"]), 0)
        for sub in ("fte_kinematic", f"fte_kinematic_orig_{cam}", f"fte_kinematic_{cam}"):
            p1 = os.path.join(outs[1], c["name"], sub, "fte.pickle")
            p2 = os.path.join(outs[2], c["name"], sub, "fte.pickle")
            assert os.path.exists(p1) == os.path.exists(p2), (c["name"], sub)
            if os.path.exists(p1):
                d1, d2 = pickle.load(open(p1, "rb")), pickle.load(open(p2, "rb"))
                assert np.array_equal(d1["q"], d2["q"]) and d1["obj_cost"] == d2["obj_cost"], (c["name"], sub)
    assert os.path.exists(os.path.join(outs[2], "dataset_results.csv"))


def test_run_monocular_all_and_ablation_on_a_subset(golden_dir, tmp_path):
    """`run_monocular_all` (run_dataset.py:918-1015: every clip x every camera of its camera space x two test types) and
    `run_data_driven_ablation_study` (:635-717) on two clips: one ragged batch of camera hypotheses per test type, output
    directories per camera, the ablation's four prior scenarios with errors against the multi-view reconstruction."""
    from cheetah_pose_estimation_b200 import run_dataset
    clips = [c for c in load_test_set_clips(golden_dir) if c["name"] in
             ("2017_08_29/top/jules/run1_1", "2019_03_03/phantom/run")]
    root = _build_root(tmp_path, clips)
    subset = tuple(t for t in run_dataset.DATASET if run_dataset._data_path(t) in [c["name"] for c in clips])
    assert len(subset) == 2
    out = os.path.join(str(tmp_path), "out")
    models_dir = os.path.join(str(tmp_path), "models", "data-driven")
    write_data_driven_dataset(golden_dir, models_dir)
    cam_space = {"2019_03_03/phantom/run": (0, 1, 2, 3), "2017_08_29/top/jules/run1_1": (0, 2, 5)}
    spaces = tuple(cam_space[run_dataset._data_path(t)] for t in subset)
    oks = run_dataset.run_monocular_all(root, out, subset, models_dir=models_dir, cam_space=spaces)
    assert len(oks["default"]) == 7 and len(oks["data-driven"]) == 7
    i = 0
    for t, space in zip(subset, spaces):
        for cam in space:
            if oks["default"][i]:
                assert os.path.exists(os.path.join(out, run_dataset._data_path(t), f"fte_kinematic_orig_{cam}", "fte.pickle"))
            if oks["data-driven"][i]:
                d = pickle.load(open(os.path.join(out, run_dataset._data_path(t), f"fte_kinematic_{cam}", "fte.pickle"), "rb"))
                assert d["meas_err"].shape[1] == 1
            i += 1
    assert sum(oks["data-driven"]) >= 5, oks
    # the ablation study needs the multi-view reconstruction as its reference
    run_dataset.run_monocular(root, out, subset, None, models_dir=models_dir, test_types=("ground-truth",))
    res = run_dataset.run_data_driven_ablation_study(root, out, subset, models_dir=models_dir)
    assert len(res["mpjpe"]) == 4 and all(np.isfinite(res["mpjpe"])) and os.path.exists(os.path.join(out, "data_driven_ablation.pickle"))
    print("ablation (pose off+motion off, pose off, motion off, both on): MPJPE", np.round(res["mpjpe"], 1), "MPE", np.round(res["mpe"], 1))
