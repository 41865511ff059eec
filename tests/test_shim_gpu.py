"""End-to-end drop-in test on real clips: init_trajectory -> estimate_kinematics -> save through the mirrored API,
from an AcinoSet-style directory rebuilt out of the reference's shipped outputs (cameras / detections recovered,
SURVEY.md G8).  The DLC likelihood mask of the original run is lost, so optimiser parity against IPOPT is UNPINNED:
the checks are (i) file format and keys equal the shipped fte.pickle, (ii) our optimum costs no more than the shipped
trajectory under the SAME (all-on) objective, (iii) it stays within centimetres of the shipped 3D trajectory."""
import os
import pickle

import numpy as np
import pytest

from tests._dataset import build_clip_dir, write_data_driven_dataset as _write_data_driven_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("animal,fname,path", [("jules", "clip_jules_run1_1.npz", "2017_08_29/top/jules/run1_1"),
                                               ("phantom", "clip_phantom_run.npz", "2019_03_07/phantom/run")])
def test_kinematic_reconstruction_of_a_real_clip(golden_dir, tmp_path, animal, fname, path):
    import torch
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt, engine, skeleton as sk
    g = np.load(os.path.join(golden_dir, fname))
    ref = np.load(os.path.join(golden_dir, "save_ref.npz"))
    root = str(tmp_path)
    build_clip_dir(root, path, g)
    est = opt.init_trajectory(root, path, animal, kinetic_dataset=False, solver_path="ignored", kinematic_model=True)
    assert est.meas.shape == g["detections"].shape[:2] + (24, 1, 4)
    ok = opt.estimate_kinematics(est, solver_output=True)
    assert ok
    out_dir = os.path.join(root, path, "fte_kinematic")
    d = pickle.load(open(os.path.join(out_dir, "fte.pickle"), "rb"))
    assert sorted(d.keys()) == list(ref[f"{animal}_keys"])
    N, C = g["detections"].shape[:2]
    assert d["q"].shape == (N, 54) and d["x"].shape == (N, 28) and d["positions"].shape == (N, 24, 3)
    assert d["meas_err"].shape == (N, C, 24, 2, 1) and d["com_vel"].shape == (N - 1, 3)
    assert d["start_frame"] == int(g["start_frame"]) and d["processing_time_s"] > 0
    for c in range(C):
        assert os.path.exists(os.path.join(out_dir, f"cam{c + 1}_fte.csv"))
    # implicit-Euler relations of the saved derivatives
    h = 1.0 / float(g["fps"])
    assert np.abs(d["q"][1:] - d["q"][:-1] - h * d["dq"][1:]).max() < 1e-12
    assert np.abs(d["dq"][1:] - d["dq"][:-1] - h * d["ddq"][1:]).max() < 1e-9
    # (ii) same objective evaluated at the shipped optimum
    _, cvec = sk.marker_chains(est.lengths, est.radii)
    batch = engine.Batch.from_numpy(est.meas[None], est.packed_cameras()[None], cvec[None], [h])
    sv = engine.Solver(est.model, batch, with_solver=False)
    c_ship = sv.evaluate(torch.as_tensor(g["q"][None].copy(), device="cuda"))["cost"][0].cpu().numpy()
    obj_ship = 1e-3 * (c_ship[0] + c_ship[1])
    # two-sided (measured, profiles/r2b_run_dataset.log: 38.209 / 39.358 and 83.397 / 84.801, i.e. 1.7-2.9 % below the
    # shipped trajectory's cost under the same all-on objective): not worse, and not a different problem
    assert 0.95 * obj_ship <= d["obj_cost"] <= obj_ship * 1.005, (d["obj_cost"], obj_ship)
    assert np.abs(est.result["c_eq"]).max() < 5e-5            # feasibility at the level of the shipped IPOPT optima
    # (iii) distance to the shipped trajectory (different likelihood mask => not identical)
    mpjpe = misc.traj_error(d["positions"].copy(), g["positions"].copy(), centered=True).mean()
    mpe = misc.traj_error(d["positions"].copy(), g["positions"].copy()).mean()
    print(f"{path}: obj {d['obj_cost']:.3f} (shipped trajectory under the same objective {obj_ship:.3f}; shipped "
          f"obj_cost {float(g['obj_cost']):.3f} with the lost mask), MPJPE {mpjpe:.1f} mm, MPE {mpe:.1f} mm, "
          f"{d['processing_time_s']:.3f} s vs IPOPT {float(g['processing_time_s']):.1f} s")
    # measured 18.8 / 21.6 mm (jules run1_1) and 19.7 / 21.7 mm (phantom run) from the shipped trajectory, whose
    # likelihood mask is lost (the 1 mm bar is checked against the NLP witness on identical inputs)
    assert mpjpe < 25.0 and mpe < 28.0


def test_monocular_reconstruction_with_data_driven_priors(golden_dir, tmp_path):
    """BASELINE config 3 through the mirrored API: single camera, GMM pose prior + AR(4) motion model fitted on the
    CPU (scikit-learn, as the reference), evaluated on the GPU inside the LM loop; output goes to fte_kinematic_<cam>."""
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt
    path = "2017_08_29/top/jules/run1_1"
    g = np.load(os.path.join(golden_dir, "clip_jules_run1_1.npz"))
    root = str(tmp_path)
    build_clip_dir(root, path, g)
    models_dir = os.path.join(root, "models", "data-driven")
    _write_data_driven_dataset(golden_dir, models_dir)
    est = opt.init_trajectory(root, path, "jules", kinetic_dataset=False, solver_path="ignored", kinematic_model=True,
                              monocular_enable=True)
    assert est.scene.cam_idx == 0 and est.meas.shape[1] == 1
    ok = opt.estimate_kinematics(est, solver_output=True, monocular_constraints=True, models_dir=models_dir)
    assert ok
    assert est.pose_model is not None and est.motion_model is not None
    out_dir = os.path.join(root, path, "fte_kinematic_0")
    d = pickle.load(open(os.path.join(out_dir, "fte.pickle"), "rb"))
    N = g["detections"].shape[0]
    assert d["q"].shape == (N, 54) and d["meas_err"].shape == (N, 1, 24, 2, 1)
    assert np.isfinite(d["obj_cost"]) and est.costs["motion"] > 0 and np.isfinite(est.costs["pose"])
    assert abs(d["obj_cost"] - 1e-3 * sum(est.costs.values())) < 1e-9
    mpjpe = misc.traj_error(d["positions"].copy(), g["positions"].copy(), centered=True).mean()
    print(f"monocular cam 0: obj {d['obj_cost']:.3f}, costs {est.costs}, MPJPE vs the shipped multi-view trajectory "
          f"{mpjpe:.1f} mm, {d['processing_time_s']:.3f} s")
    # the reference reports 64.7 mm (MPJPE, data-driven) for this clip (its own camera and its own
    # training set); bound = 1.5 x that, the prior here being fitted on the 55 shipped trajectories only
    assert mpjpe < 1.5 * 64.7


def test_shutter_delay_estimation_through_the_mirrored_api(golden_dir, tmp_path):
    """`init_trajectory(..., shutter_delay_estimation=True)` (acinoset_opt.py:428, acinoset_misc.py:182-183, 273-286) on
    a real 6-camera clip: delays inside the reference's bounds with camera 1 fixed at 0, an objective no worse than the
    plain reconstruction of the same data, and per-camera 2-D outputs written from the delayed positions
    (acinoset_opt.py:342-352)."""
    import pandas as pd
    from cheetah_pose_estimation_b200 import acinoset_opt as opt
    path = "2017_08_29/top/jules/run1_1"
    g = np.load(os.path.join(golden_dir, "clip_jules_run1_1.npz"))
    h = 1.0 / float(g["fps"])
    objs = {}
    for flag in (False, True):
        root = os.path.join(str(tmp_path), "delay" if flag else "plain")
        build_clip_dir(root, path, g)
        est = opt.init_trajectory(root, path, "jules", kinetic_dataset=False, solver_path="ignored",
                                  kinematic_model=True, shutter_delay_estimation=flag)
        assert opt.estimate_kinematics(est, solver_output=True)
        objs[flag] = (est.get_objective_cost(), est, root)
    obj_plain, obj_delay = objs[False][0], objs[True][0]
    est, root = objs[True][1], objs[True][2]
    tau = est.shutter_delay
    assert objs[False][1].shutter_delay is None
    assert tau.shape == (6,) and tau[0] == 0.0 and np.all(np.abs(tau) <= h)
    print(f"shutter delays / h: {np.round(tau / h, 4)}, objective {obj_delay:.4f} vs plain {obj_plain:.4f}")
    assert obj_delay <= obj_plain * 1.002
    out_dir = os.path.join(root, path, "fte_kinematic")
    d = pickle.load(open(os.path.join(out_dir, "fte.pickle"), "rb"))
    # camera c's CSV holds the projection of pose + dq tau_c + ddq tau_c^2; camera 1 (tau = 0) that of the pose itself
    from cheetah_pose_estimation_b200 import acinoset_misc as misc
    from oracle import cameras
    for c in (0, 3):
        df = pd.read_csv(os.path.join(out_dir, f"cam{c + 1}_fte.csv"), header=[0, 1], index_col=0)
        pts = d["positions"] + d["dq"][:, None, 0:3] * tau[c] + d["ddq"][:, None, 0:3] * tau[c] ** 2
        uv = cameras.project(pts, g["cams"][c][None, None, :])                     # [N,24,2]
        got = df.to_numpy().reshape(len(df), 24, 3)[:, :, 0:2]                    # x, y, likelihood (NaN) per bodypart
        m = np.isfinite(got)
        assert m.mean() > 0.5 and np.abs(got[m] - uv[m]).max() < 1e-3


def test_hand_labelled_reconstruction_through_the_mirrored_api(golden_dir, tmp_path):
    """`init_trajectory(hand_labeled_data=True)` + `estimate_kinematics` (acinoset_opt.py:413-430, 539-635;
    acinoset_misc.py:216-221, 425-439): (x, y) label tables named by image file, likelihood 1, NaN labels gated, the
    weighted-spline initial guess, output into `fte_kinematic_gt`.  With labels equal to the clip's detections the result
    is the plain reconstruction of that clip."""
    from tests._dataset import write_hand_labels
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, acinoset_opt as opt
    path = "2017_08_29/top/jules/run1_1"
    g = np.load(os.path.join(golden_dir, "clip_jules_run1_1.npz"))
    root = str(tmp_path)
    build_clip_dir(root, path, g)
    write_hand_labels(root, path, g, missing=[(3, 1, 7), (4, 1, 7), (10, 0, 2)])
    est = opt.init_trajectory(root, path, "jules", kinetic_dataset=False, kinematic_model=True, hand_labeled_data=True)
    assert est.meas.shape == g["detections"].shape[:2] + (24, 1, 4)
    assert est.meas[3, 1, 7, 0, 2] == 0.0 and est.meas[3, 1, 8, 0, 2] == 1.0 and est.base_data.first_frame == [0] * 6
    assert opt.estimate_kinematics(est, solver_output=False)
    out = os.path.join(root, path, "fte_kinematic_gt", "fte.pickle")
    assert os.path.exists(out)
    d = pickle.load(open(out, "rb"))
    mpjpe = misc.traj_error(d["positions"].copy(), g["positions"].copy(), centered=True).mean()
    print(f"hand-labelled run: obj {d['obj_cost']:.3f}, MPJPE vs the shipped trajectory {mpjpe:.1f} mm")
    assert mpjpe < 35.0
