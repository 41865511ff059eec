"""Drop-in mirror of the reference's estimation API for the kinematic reconstruction (acinoset_opt.py:21-38, 278-373,
413-536, 539-635): same function names, keyword arguments, output directories and file formats; the Pyomo model and
the `ipopt` call are replaced by the CUDA path (engine.Solver).  `solver_path` is accepted and ignored.

Mirrored beyond the multi-view reconstruction: the monocular data-driven priors (`monocular_constraints=True`,
SURVEY.md section 8f.1), pairwise pseudo-measurements (`enable_ppm`), hand-labelled detections, shutter-delay
estimation (`shutter_delay_estimation=True`, section 8f.4: engine.Solver.solve_shutter) and
`CheetahEstimator.estimate_grf` (section 8f.3, first stage of the kinetic mode).  Not mirrored (out of scope,
SURVEY.md section 8): `determine_contacts`, `estimate_kinetics` (full 54-row dynamics).
"""
import json
import os
from dataclasses import dataclass, field
from time import time
from typing import Dict, Optional, Tuple

import numpy as np

from . import acinoset_misc as misc
from . import skeleton as sk


@dataclass
class CheetahEstimator:
    """Fields follow the reference dataclass (acinoset_opt.py:21-38); `model` is the device model (engine.Model)
    instead of a `System3D`, and the solution lives in `q`, `dq`, `ddq` instead of Pyomo variables."""
    name: str
    data_path: str
    model: object
    params: misc.TrajectoryParams
    scene: misc.Scene
    scale_forces_by: float
    position_funcs: object
    kinematic_model: bool
    enable_eom_slack: bool
    costs: Optional[Dict] = None
    synthesised_grf: Optional[Dict] = None
    com_pos: Optional[np.ndarray] = None
    com_vel: Optional[np.ndarray] = None
    opt_time_s: Optional[float] = None
    pose_model: object = None
    motion_model: object = None
    # ---- state of the replacement path ----
    lengths: np.ndarray = None
    radii: np.ndarray = None
    masses: np.ndarray = None
    meas: np.ndarray = None             # [N,C,24,W,4] float32 (u, v, gate, 0)
    cams_used: list = field(default_factory=list)
    base_data: list = None
    q: Optional[np.ndarray] = None
    dq: Optional[np.ndarray] = None
    ddq: Optional[np.ndarray] = None
    positions: Optional[np.ndarray] = None
    meas_err: Optional[np.ndarray] = None
    result: Optional[Dict] = None

    # -- helpers -------------------------------------------------------------------------------------------------
    def packed_cameras(self):
        p, s = self.params, self.scene
        mult = [1.0, 1.0, 0.6, 0.6] if p.kinetic_dataset else [1.0] * s.n_cams       # acinoset_misc.py:462-464
        return np.stack([sk.pack_camera(s.k_arr[c], s.d_arr[c], s.r_arr[c], s.t_arr[c], mult[c])
                         for c in self.cams_used])

    def get_objective_cost(self) -> float:
        return float(1e-3 * (self.costs["measurement"] + self.costs["model"] + self.costs["pose"] + self.costs["motion"]))

    def solution_details(self, print_infeasible_contraints: bool = False, tolerance: float = 1e-1):
        print("Total cost:", self.get_objective_cost())
        if getattr(self, "shutter_delay", None) is not None:
            print("Shutter delay estimation:", [float(t) for t in self.shutter_delay])       # acinoset_opt.py:397-398
        for k, v in (self.costs or {}).items():
            print(f"-- {k}: {v}")
        if print_infeasible_contraints and self.result is not None:
            c = np.abs(self.result["c_eq"])
            print("Infeasible constraints:", int((c > tolerance).sum()), "joint rows above", tolerance)

    def is_solution_acceptable(self, results, base_error: float, relative_error: float):
        """Reference: solver status ok and termination optimal (acinoset_opt.py:400-406).  Here: no failed pivot,
        finite cost, joint equalities (dot products of unit axes) satisfied to 2e-3 (~0.1 degree) and inequality
        violations below 2e-3 rad."""
        if results is None:
            return False
        ok = int(results["status"]) == 0 and np.isfinite(results["cost"]).all()
        return bool(ok and np.abs(results["c_eq"]).max() < 2e-3 and np.abs(results["v_in"]).max() < 2e-3)

    def estimate_grf(self, plot: bool = False, monocular: bool = False, out_dir_prefix: Optional[str] = None,
                     device: str = "cuda"):
        """Reference: CheetahEstimator.estimate_grf (acinoset_opt.py:176-268).  Reads the kinematic trajectory saved by
        `estimate_kinematics` (`fte_kinematic[_<cam>]/fte.pickle`) and `grf/autogen-contact.json`, solves every frame's
        bounded least-squares problem on the GPU (kinetics.estimate_grf) and returns `(grfz_est, grfxy_est)`, dicts
        keyed by foot name with one entry per frame.  `plot` is accepted and ignored (no matplotlib on this path)."""
        import pickle
        from . import kinetics
        data_dir = self.params.data_dir if out_dir_prefix is None else os.path.join(
            out_dir_prefix, self.params.data_dir.split("cheetah_videos")[1][1:])
        sub = "fte_kinematic" if not monocular else f"fte_kinematic_{self.scene.cam_idx}"
        with open(os.path.join(data_dir, sub, "fte.pickle"), "rb") as f:
            fte_states = pickle.load(f)
        with open(os.path.join(data_dir, "grf/autogen-contact.json"), "r", encoding="utf-8") as f:
            metadata = json.load(f)
        _ = plot
        return kinetics.estimate_grf_states(fte_states, metadata, self.name, device=device)

    def joint_error(self, q: np.ndarray) -> Tuple[float, float]:
        return misc_rmse(q[:, :6], self.q[:, :6]), misc_rmse(q[:, 6:], self.q[:, 6:])

    def save(self, out_dir: str, fname: str = "fte", out_dir_prefix: Optional[str] = None):
        """Writes `<out_dir>/<fname>.pickle` and `cam<i>_<fname>.{csv,h5}` with the reference's keys and layouts
        (acinoset_opt.py:278-373)."""
        params, scene = self.params, self.scene
        markers = misc.get_markers()
        if out_dir_prefix:
            out_dir = os.path.join(out_dir_prefix, self.data_path, "fte" if out_dir is None else out_dir)
        else:
            out_dir = os.path.join(params.data_dir, "fte" if out_dir is None else out_dir)
        os.makedirs(out_dir, exist_ok=True)
        T = sk.relative_matrix()[sk.independent_coords()]
        output = dict(x=self.q @ T.T, dx=self.dq @ T.T, ddx=self.ddq @ T.T, q=self.q, dq=self.dq, ddq=self.ddq,
                      com_pos=self.com_pos, com_vel=self.com_vel, tau={}, meas_err=self.meas_err,
                      obj_cost=self.get_objective_cost(), processing_time_s=self.opt_time_s)
        sync_offset_arr = [0] * scene.n_cams
        if params.sync_offset is not None:
            for offset in params.sync_offset:
                sync_offset_arr[offset["cam"]] = offset["frame"]
        out_fpath = os.path.join(out_dir, f"{fname}.pickle")
        misc.save_optimised_cheetah(self.positions, out_fpath, extra_data=dict(**output, start_frame=params.start_frame))
        positions_arr = [self.positions] * scene.n_cams
        if getattr(self, "shutter_delay", None) is not None:
            # what camera c sees: pose + dq_base tau_c + ddq_base tau_c^2 (acinoset_opt.py:342-352)
            assert len(self.shutter_delay) == scene.n_cams
            positions_arr = [self.positions + self.dq[:, None, 0:3] * t + self.ddq[:, None, 0:3] * (t ** 2)
                             for t in self.shutter_delay]
        misc.save_3d_cheetah_as_2d(positions_arr, out_dir, scene.scene_fpath, markers,
                                   misc.project_points if params.kinetic_dataset else misc.project_points_fisheye,
                                   params.start_frame, sync_offset_arr, vid_dir=params.data_dir, out_fname=fname)
        return out_fpath


def misc_rmse(predictions, targets):
    return float(np.sqrt(np.nanmean(((predictions - targets) ** 2).flatten())))


def init_trajectory(root_dir: str,
                    data_path: str,
                    cheetah_name: str,
                    kinetic_dataset: bool,
                    solver_path: str = None,
                    start_frame: int = -1,
                    end_frame: int = -1,
                    dlc_thresh: float = 0.5,
                    include_camera_constraints: bool = True,
                    enable_eom_slack: bool = True,
                    kinematic_model: bool = False,
                    monocular_enable: bool = False,
                    override_monocular_cam: Optional[int] = None,
                    disable_contact_lcp: bool = True,
                    bound_eom_error: Optional[Tuple[float, float]] = None,
                    shutter_delay_estimation: bool = False,
                    enable_ppm: bool = False,
                    hand_labeled_data: bool = False,
                    device: str = "cuda") -> CheetahEstimator:
    """Same arguments as the reference (acinoset_opt.py:413-430) plus `device`.  Reads `metadata.json`, the scene
    file and the DLC predictions, and packs the measurements for the CUDA path."""
    from . import engine
    if not kinematic_model:
        raise NotImplementedError("kinetic (physics-based) model: out of scope of this path (SURVEY.md section 8f.3)")
    if cheetah_name not in ("jules", "phantom", "shiraz", "arabia"):
        cheetah_name = "acinoset"
    ground_plane_height = 0.0
    cam_idx = None
    if start_frame < 0 or end_frame < 0:
        with open(os.path.join(root_dir, data_path, "metadata.json"), "r", encoding="utf-8") as f:
            metadata = json.load(f)
        start_frame = metadata["start_frame"]
        end_frame = metadata["end_frame"]
        sync_offset = metadata["cam_sync"]
        ground_plane_height = metadata["ground_plane_height"]
        cam_idx = metadata["monocular_cam"] if monocular_enable else None
        cam_idx = cam_idx if override_monocular_cam is None else override_monocular_cam
        total_length = metadata["end_frame"] - start_frame
    else:
        total_length = end_frame - start_frame
        sync_offset = None
    data_dir = os.path.join(root_dir, data_path)
    assert os.path.exists(data_dir)
    dlc_dir = os.path.join(data_dir, "dlc" if not hand_labeled_data else "dlc_hand_labeled")
    assert os.path.exists(dlc_dir)
    k_arr, d_arr, r_arr, t_arr, cam_res, n_cams, scene_fpath = misc.find_scene_file(data_dir)
    fps = 200
    if not kinetic_dataset and "2019" in data_path:
        fps = 120
    elif not kinetic_dataset and "2017" in data_path:
        fps = 90
    d_arr = d_arr.reshape((-1, 4))
    params = misc.TrajectoryParams(data_dir, start_frame, end_frame, total_length, dlc_thresh, sync_offset,
                                   hand_labeled_data, kinetic_dataset, shutter_delay_estimation, enable_ppm)
    scene = misc.Scene(scene_fpath, k_arr, d_arr, r_arr, t_arr, cam_res, fps, n_cams, cam_idx)
    lengths, radii, masses = sk.animal(cheetah_name)
    est = CheetahEstimator(cheetah_name, data_path, engine.Model(device, kinetic_dataset), params, scene,
                           float(masses.sum() * 9.81), None, kinematic_model, enable_eom_slack,
                           lengths=lengths, radii=radii, masses=masses)
    print(f"Total cheetah mass: {masses.sum():.2f} Kg")
    if include_camera_constraints:
        print("Create camera constraints based on DLC predictions")
        est.base_data = misc.load_dlc_arrays(dlc_dir, scene.n_cams)
        pw_data = misc.load_pairwise(dlc_dir, scene.n_cams) if enable_ppm else None
        est.meas, est.cams_used = misc.pack_measurements(params, scene, est.base_data, pw_data)
    _ = (ground_plane_height, disable_contact_lcp, bound_eom_error, solver_path)     # kinetic-mode / IPOPT settings
    return est


def estimate_kinematics(estimator: CheetahEstimator,
                        solver_output: bool = True,
                        monocular_constraints: bool = False,
                        disable_pose_prior: bool = False,
                        disable_motion_prior: bool = False,
                        pose_model_num_components: int = 5,
                        motion_model_window_size: int = 4,
                        motion_model_sparse_solution: bool = True,
                        out_dir_prefix: Optional[str] = None,
                        max_iter: int = 40,
                        n_outer: int = 4,
                        models_dir: Optional[str] = None) -> bool:
    """Same arguments and side effects as the reference (acinoset_opt.py:539-635): initial guess from the spline
    track, optimisation, COM, acceptance test, `save` into `fte_kinematic[_orig][_<cam>]`.  With
    `monocular_constraints` on a single-camera scene the GMM pose prior and the linear motion model are fitted on the
    CPU from `<models_dir or ./models/data-driven>/dataset_full_pose.h5` (its `.csv` twin when PyTables is absent), as
    the reference does (acinoset_misc.py:294-302, 684-689), and evaluated on the GPU inside the LM loop."""
    import torch
    from . import acinoset_models, engine
    from .priors import Priors
    params, scene = estimator.params, estimator.scene
    priors = None
    estimator.pose_model = estimator.motion_model = None
    if monocular_constraints and scene.cam_idx is not None and not (disable_pose_prior and disable_motion_prior):
        ds = os.path.join(models_dir or os.path.join(".", "models", "data-driven"), "dataset_full_pose.h5")
        if not disable_pose_prior:
            estimator.pose_model = acinoset_models.PoseModelGMM(ds, num_vars=28, ext_dim=6,
                                                                n_comps=pose_model_num_components)
        if not disable_motion_prior:
            estimator.motion_model = acinoset_models.MotionModel(ds, 28, window_size=motion_model_window_size,
                                                                 window_time=1, lasso=motion_model_sparse_solution)
        priors = Priors(acinoset_models.pack_priors(estimator.pose_model, estimator.motion_model), estimator.model.device)
    N = params.end_frame - params.start_frame
    x_est, y_est, z_est, psi_est = misc.create_trajectory_estimate(estimator.lengths[sk.BASE], params, scene,
                                                                   params.kinetic_dataset, estimator.base_data)
    q0 = np.zeros((N, 54))
    fr = params.start_frame + np.arange(N)
    q0[:, 5::3] = psi_est[fr][:, None]                       # every link's psi = heading (acinoset_opt.py:574-583)
    q0[:, 0], q0[:, 1], q0[:, 2] = x_est[fr], y_est[fr], z_est[fr]
    dev = estimator.model.device
    _, chain_vec = sk.marker_chains(estimator.lengths, estimator.radii)
    batch = engine.Batch.from_numpy(estimator.meas[None], estimator.packed_cameras()[None], chain_vec[None],
                                    [1.0 / scene.fps], device=dev,
                                    cam_model=engine.PINHOLE if params.kinetic_dataset else engine.FISHEYE)
    solver = engine.Solver(estimator.model, batch, priors=priors)
    time0 = time()
    estimator.shutter_delay = None
    if scene.cam_idx is None and params.enable_shutter_delay_estimation and priors is None:
        # per-camera shutter delays next to q (acinoset_misc.py:182-183, 273-286): first the plain solve, then the
        # joint (q, tau) iterations from its optimum
        res = solver.solve(torch.as_tensor(q0[None], device=dev), max_iter=max_iter, n_outer=n_outer)
        res = solver.solve_shutter(res["q"], max_iter=max_iter, n_outer=n_outer)
        estimator.shutter_delay = res["tau"][0].cpu().numpy()
    else:
        # the priors pull against the joint equalities: two more multiplier updates keep the violation below the
        # acceptance threshold of is_solution_acceptable
        res = solver.solve(torch.as_tensor(q0[None], device=dev), max_iter=max_iter,
                           n_outer=n_outer + (2 if priors is not None else 0))
    torch.cuda.synchronize()
    estimator.opt_time_s = time() - time0
    q = res["q"][0].cpu().numpy()
    out = solver.evaluate(res["q"], pose=True, resid=True)
    estimator.q = q
    h = 1.0 / scene.fps
    dq = np.zeros_like(q)
    ddq = np.zeros_like(q)
    dq[1:] = (q[1:] - q[:-1]) / h                            # implicit Euler (collocation.py:58-68)
    ddq[2:] = (dq[2:] - dq[1:-1]) / h
    if N > 2:                                                # free boundary values: ddq_0 = ddq_1 = ddq_2
        ddq[1] = ddq[2]
        ddq[0] = ddq[2]
        dq[0] = dq[1] - h * ddq[1]
    estimator.dq, estimator.ddq = dq, ddq
    estimator.positions = out["pose"][0].cpu().numpy()
    estimator.meas_err = np.transpose(out["resid"][0].cpu().numpy(), (0, 1, 2, 4, 3))      # [N,C,24,2,W]
    cost = res["cost"][0].cpu().numpy()
    pc = res["prior_cost"][0].cpu().numpy() if priors is not None else np.zeros(2)
    estimator.costs = {"measurement": float(cost[0]), "model": float(cost[1]), "pose": float(pc[0]),
                       "motion": float(pc[1])}
    estimator.result = dict(status=int(res["status"][0]), cost=cost, c_eq=res["c_eq"][0].cpu().numpy(),
                            v_in=res["v_in"][0].cpu().numpy(), counters=res["counters"][0].cpu().numpy())
    if solver_output:
        it, acc = estimator.result["counters"]
        print(f"LM/AL: {it} iterations ({acc} accepted), max |joint eq| {np.abs(estimator.result['c_eq']).max():.2e}, "
              f"{estimator.opt_time_s:.3f} s")
    estimator.solution_details()
    estimator.com_pos, estimator.com_vel = misc.get_com(q, estimator.lengths, estimator.radii, estimator.masses, scene.fps)
    ret = estimator.is_solution_acceptable(estimator.result, 0.1, 0.2)
    if ret:
        fname = f"fte_kinematic{'_gt' if params.hand_labeled_data else ''}"
        fname = fname if scene.cam_idx is None or monocular_constraints else "fte_kinematic_orig"
        fname = fname if scene.cam_idx is None else f"{fname}_{scene.cam_idx}"
        estimator.save(fname, out_dir_prefix=out_dir_prefix)
    return ret
