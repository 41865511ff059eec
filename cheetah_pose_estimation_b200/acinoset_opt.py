"""Drop-in mirror of the reference's estimation API for the kinematic reconstruction (acinoset_opt.py:21-38, 278-373,
413-536, 539-635): same function names, keyword arguments, output directories and file formats; the Pyomo model and
the `ipopt` call are replaced by the CUDA path (engine.Solver).  `solver_path` is accepted and ignored.

Mirrored beyond the multi-view reconstruction: the monocular data-driven priors (`monocular_constraints=True`,
SURVEY.md section 8f.1), pairwise pseudo-measurements (`enable_ppm`), hand-labelled detections (`hand_labeled_data`:
label tables named by image file, weighted-spline initial guess, `fte_kinematic_gt` output), shutter-delay
estimation (`shutter_delay_estimation=True`, section 8f.4: engine.Solver.solve_shutter) and
`CheetahEstimator.estimate_grf` and `determine_contacts` (section 8f.3, first stages of the kinetic mode).  Not mirrored
(out of scope): `estimate_kinetics` (full 54-row dynamics).
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
        finite cost, joint equalities (dot products of unit axes) satisfied to 1e-4 -- the shipped IPOPT optima hold
        them to 5e-8 .. 9e-5 (tests/test_run_dataset_gpu.py prints both) -- and inequality violations below 5e-4 rad."""
        if results is None:
            return False
        # CPE_STATUS_PIVOT | CPE_STATUS_NONFINITE disqualify; CPE_STATUS_MAXITER alone does not (the feasibility test below
        # decides, as IPOPT's "acceptable" termination does)
        ok = (int(results["status"]) & 5) == 0 and np.isfinite(results["cost"]).all()
        return bool(ok and np.abs(results["c_eq"]).max() < CEQ_TOL and np.abs(results["v_in"]).max() < VIN_TOL)

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


_MODELS = {}
EXTRA_OUTER = 8          # further multiplier rounds for sequences that are not yet feasible (engine.Solver.solve_steps)
CEQ_TOL = 1e-4          # acceptance: max |joint equality| (unit-axis dot products)
VIN_TOL = 5e-4          # acceptance: max inequality violation, rad


def device_model(device, kinetic_dataset):
    """One `engine.Model` (device copy of the static skeleton tables) per (device, dataset kind), shared by every
    estimator of a batch."""
    import torch
    from . import engine
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    key = (str(dev), bool(kinetic_dataset))
    if key not in _MODELS:
        _MODELS[key] = engine.Model(dev, kinetic_dataset)
    return _MODELS[key]


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
    est = CheetahEstimator(cheetah_name, data_path, device_model(device, kinetic_dataset), params, scene,
                           float(masses.sum() * 9.81), None, kinematic_model, enable_eom_slack,
                           lengths=lengths, radii=radii, masses=masses)
    print(f"Total cheetah mass: {masses.sum():.2f} Kg")
    if include_camera_constraints:
        print("Create camera constraints based on DLC predictions")
        est.base_data = misc.load_dlc_arrays(dlc_dir, scene.n_cams, hand_labeled=hand_labeled_data)
        pw_data = misc.load_pairwise(dlc_dir, scene.n_cams) if enable_ppm else None
        est.meas, est.cams_used = misc.pack_measurements(params, scene, est.base_data, pw_data)
    est.ground_plane_height = float(ground_plane_height)      # Foot3D.ground_plane_height (acinoset_opt.py:500)
    _ = (disable_contact_lcp, bound_eom_error, solver_path)                           # kinetic-mode / IPOPT settings
    return est


def determine_contacts(estimator: CheetahEstimator, monocular: bool = False, verbose: bool = True,
                       out_dir_prefix: Optional[str] = None):
    """Reference: determine_contacts (acinoset_opt.py:638-690).  Reads the kinematic trajectory saved by
    `estimate_kinematics` (`fte_kinematic[_<cam>]/fte.pickle`), runs the contact heuristic on the CPU as the reference does
    (kinetics.contact_detection = misc.contact_detection, acinoset_misc.py:745-856) and writes `grf/autogen-contact.json`
    and `grf/autogen-contact-02.json`, the files `estimate_grf` reads.  The synthetic GRF tables of `misc.synth_grf_data`
    (initial guesses of the kinetic NLP) are not produced: that NLP is out of scope here.  Returns (contacts, contacts_tmp)."""
    import pickle
    from . import kinetics
    data_dir = estimator.params.data_dir if out_dir_prefix is None else os.path.join(
        out_dir_prefix, estimator.params.data_dir.split("cheetah_videos")[1][1:])
    sub = "fte_kinematic" if not monocular else f"fte_kinematic_{estimator.scene.cam_idx}"
    with open(os.path.join(data_dir, sub, "fte.pickle"), "rb") as f:
        fte_states = pickle.load(f)
    estimator.com_vel, estimator.com_pos = fte_states["com_vel"], fte_states["com_pos"]
    contacts, contacts_tmp = kinetics.determine_contacts_from_states(
        fte_states, estimator.name, estimator.params.start_frame, estimator.scene.fps,
        getattr(estimator, "ground_plane_height", 0.0), data_dir)
    if verbose:
        print("Height, velocity, stance time heuristic:")
        print(contacts)
        print("Height:")
        print(contacts_tmp)
    return contacts, contacts_tmp


def initial_guess(estimator: CheetahEstimator) -> np.ndarray:
    """q0[N,54] of the reference (acinoset_opt.py:568-583): every angle 0 except psi = heading of the spline track for
    every link, base x, y, z on the track."""
    params, scene = estimator.params, estimator.scene
    N = params.end_frame - params.start_frame
    x_est, y_est, z_est, psi_est = misc.create_trajectory_estimate(estimator.lengths[sk.BASE], params, scene,
                                                                   params.kinetic_dataset, estimator.base_data)
    q0 = np.zeros((N, 54))
    fr = params.start_frame + np.arange(N)
    q0[:, 5::3] = psi_est[fr][:, None]
    q0[:, 0], q0[:, 1], q0[:, 2] = x_est[fr], y_est[fr], z_est[fr]
    return q0


def pack_ragged(estimators, q0s, device, n_pad=None, c_pad=None):
    """One `engine.Batch` (and q0[S,N,54]) for clips of different length and camera count: frames are padded to the
    longest clip (`cpe_batch.n_frames` tells the kernels where each one ends); cameras are padded to the largest rig
    with copies of the clip's first camera and `cpe_batch.n_cams` tells the kernels how many are real: the padding adds
    nothing to cost, gradient or curvature and every sum keeps the order of the unpadded clip, so a clip's result is
    bit-identical whatever it is batched with.  `n_pad` / `c_pad` fix the padded sizes (the distributed driver
    passes the maxima over ALL ranks so that a clip's arithmetic does not depend on which rank it lands on)."""
    from . import engine
    import torch
    S = len(estimators)
    kin = estimators[0].params.kinetic_dataset
    W = estimators[0].meas.shape[3]
    assert all(e.params.kinetic_dataset == kin and e.meas.shape[3] == W for e in estimators), \
        "one batch = one camera model and one hypothesis count"
    N = max(e.meas.shape[0] for e in estimators) if n_pad is None else n_pad
    Cn = max(e.meas.shape[1] for e in estimators) if c_pad is None else c_pad
    # the kinetic dataset's 2-D points are not float32 DLC output: keep them in double (CPE_MEAS_F64X4)
    mdt = np.float64 if kin else np.float32
    meas = np.zeros((S, N, Cn, 24, W, 4), mdt)
    cams = np.zeros((S, Cn, sk.CAM_STRIDE))
    cvec = np.zeros((S, 24, 5, 3))
    q0 = np.zeros((S, N, 54))
    nf = np.zeros(S, np.int32)
    nc = np.zeros(S, np.int32)
    h = np.zeros(S)
    for i, (e, q) in enumerate(zip(estimators, q0s)):
        n, c = e.meas.shape[:2]
        meas[i, :n, :c] = e.meas
        pc = e.packed_cameras()
        cams[i, :c] = pc
        cams[i, c:] = pc[0]
        cvec[i] = sk.marker_chains(e.lengths, e.radii)[1]
        q0[i, :n] = q
        q0[i, n:] = q[-1]                    # frames beyond n_frames are never read; keep them finite
        nf[i], nc[i], h[i] = n, c, 1.0 / e.scene.fps
    batch = engine.Batch.from_numpy(meas, cams, cvec, h, n_frames=nf, device=device,
                                    cam_model=engine.PINHOLE if kin else engine.FISHEYE,
                                    meas_dtype=torch.float64 if kin else torch.float32, n_cams=nc)
    return batch, torch.as_tensor(q0, device=device)


def estimate_kinematics_batch(estimators,
                              solver_output: bool = True,
                              monocular_constraints: bool = False,
                              disable_pose_prior: bool = False,
                              disable_motion_prior: bool = False,
                              pose_model_num_components: int = 5,
                              motion_model_window_size: int = 4,
                              motion_model_sparse_solution: bool = True,
                              out_dir_prefix: Optional[str] = None,
                              max_iter: int = 40,
                              n_outer: int = 8,
                              models_dir: Optional[str] = None,
                              distributed: Optional[bool] = None,
                              save: bool = True,
                              pad: Optional[Tuple[int, int]] = None):
    """`estimate_kinematics` for a LIST of estimators in one ragged batch: every clip (or monocular camera hypothesis)
    is one sequence of one LM / augmented-Lagrangian solve (the reference runs them one after the other,
    run_dataset.py:1143-1231, 1234-1331, 918-1015).  All estimators must share the camera model (kinetic_dataset), the
    hypothesis count (enable_ppm) and the single-/multi-view kind.  With torch.distributed initialised (one process per
    GPU) the sequences are dealt to the ranks by frames x cameras (sharding.partition_balanced), solved without any
    inter-GPU traffic, gathered on every rank (one all_gather per result tensor) and written by rank 0.
    Per-sequence results do not depend on the batch neighbours or the rank count (`pad` = (frames, cameras) fixes the
    padded sizes; default = the maxima over the list).  Returns [bool] per estimator."""
    import torch
    import torch.distributed as dist
    from . import acinoset_models, engine, sharding
    from .priors import Priors
    ests = list(estimators)
    assert ests, "no estimators"
    dev = ests[0].model.device
    mono = [e.scene.cam_idx is not None for e in ests]
    assert all(m == mono[0] for m in mono), "one batch = all multi-view or all single-view"
    shutter = [bool(e.params.enable_shutter_delay_estimation) for e in ests]
    assert all(x == shutter[0] for x in shutter)
    priors = None
    pose_model = motion_model = None
    if monocular_constraints and mono[0] and not (disable_pose_prior and disable_motion_prior):
        ds = os.path.join(models_dir or os.path.join(".", "models", "data-driven"), "dataset_full_pose.h5")
        if not disable_pose_prior:
            pose_model = acinoset_models.PoseModelGMM(ds, num_vars=28, ext_dim=6, n_comps=pose_model_num_components)
        if not disable_motion_prior:
            motion_model = acinoset_models.MotionModel(ds, 28, window_size=motion_model_window_size, window_time=1,
                                                       lasso=motion_model_sparse_solution)
        priors = Priors(acinoset_models.pack_priors(pose_model, motion_model), dev)
    for e in ests:
        e.pose_model, e.motion_model = pose_model, motion_model
    q0s = [initial_guess(e) for e in ests]
    if distributed is None:
        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    work = [e.meas.shape[0] * e.meas.shape[1] for e in ests]
    shard = sharding.my_shard(len(ests), work) if distributed else sharding.ShardInfo(0, 1, np.arange(len(ests)))
    n_pad = max(e.meas.shape[0] for e in ests)
    c_pad = max(e.meas.shape[1] for e in ests)
    if pad is not None:
        assert pad[0] >= n_pad and pad[1] >= c_pad
        n_pad, c_pad = pad
    mine = [int(i) for i in shard.index]
    n_ineq = ests[0].model.n_ineq
    f64 = dict(dtype=torch.float64, device=dev)
    time0 = time()
    tau = None
    if mine:
        batch, q0 = pack_ragged([ests[i] for i in mine], [q0s[i] for i in mine], dev, n_pad, c_pad)
        solver = engine.Solver(ests[0].model, batch, priors=priors)
        if not mono[0] and shutter[0] and priors is None:
            # per-camera shutter delays next to q (acinoset_misc.py:182-183, 273-286): first the plain solve, then the
            # joint (q, tau) iterations from its optimum
            res = solver.solve(q0, max_iter=max_iter, n_outer=n_outer)
            res = solver.solve_shutter(res["q"], max_iter=max_iter, n_outer=n_outer)
            tau = res["tau"]
        else:
            # the priors pull against the joint equalities: two more multiplier updates keep the violation below the
            # acceptance threshold of is_solution_acceptable
            # Sequences that still violate the acceptance thresholds after the scheduled rounds (the ill-conditioned
            # single-view tail flicks) get up to EXTRA_OUTER more rounds on their own; the others are untouched.
            res = solver.solve(q0, max_iter=max_iter, n_outer=n_outer + (2 if priors is not None else 0),
                               extra_outer=EXTRA_OUTER, feas_tol=(0.5 * CEQ_TOL, 0.5 * VIN_TOL))
        out = solver.evaluate(res["q"], pose=True, resid=True)
        loc = dict(q=res["q"], cost=res["cost"], c_eq=res["c_eq"], v_in=res["v_in"],
                   status=res["status"].to(torch.int32), counters=res["counters"].to(torch.int32),
                   pose=out["pose"], resid=out["resid"],
                   prior_cost=res["prior_cost"] if priors is not None else torch.zeros(len(mine), 2, **f64),
                   tau=tau if tau is not None else torch.zeros(len(mine), c_pad, **f64))
    else:                                    # more ranks than sequences: an empty shard still joins the gather
        W = ests[0].meas.shape[3]
        loc = dict(q=torch.zeros(0, n_pad, 54, **f64), cost=torch.zeros(0, 4, **f64),
                   c_eq=torch.zeros(0, n_pad, 26, **f64), v_in=torch.zeros(0, n_pad, n_ineq, **f64),
                   status=torch.zeros(0, dtype=torch.int32, device=dev),
                   counters=torch.zeros(0, 2, dtype=torch.int32, device=dev),
                   pose=torch.zeros(0, n_pad, 24, 3, **f64), resid=torch.zeros(0, n_pad, c_pad, 24, W, 2, **f64),
                   prior_cost=torch.zeros(0, 2, **f64), tau=torch.zeros(0, c_pad, **f64))
    loc = {k: v.contiguous() for k, v in loc.items()}
    full = sharding.gather_results(shard, len(ests), loc, dst=None)
    torch.cuda.synchronize()
    elapsed = time() - time0
    full = {k: v.cpu().numpy() for k, v in full.items()}
    rets = []
    for i, e in enumerate(ests):
        n, c = e.meas.shape[:2]
        e.shutter_delay = full["tau"][i, :c].copy() if (not mono[0] and shutter[0] and priors is None) else None
        e.opt_time_s = elapsed / len(ests) if len(ests) > 1 else elapsed
        rets.append(_finish_estimate(e, full["q"][i, :n], full["cost"][i], full["prior_cost"][i], full["c_eq"][i, :n],
                                     full["v_in"][i, :n], int(full["status"][i]), full["counters"][i],
                                     full["pose"][i, :n], full["resid"][i, :n, :c], priors is not None,
                                     monocular_constraints, solver_output, out_dir_prefix,
                                     save and shard.rank == 0))
    return rets


def _finish_estimate(estimator, q, cost, pc, c_eq, v_in, status, counters, pose, resid, has_priors,
                     monocular_constraints, solver_output, out_dir_prefix, save):
    """Host-side epilogue of one sequence: dq / ddq, COM, costs, acceptance test, `save` (acinoset_opt.py:619-635)."""
    params, scene = estimator.params, estimator.scene
    N = q.shape[0]
    estimator.q = q = np.ascontiguousarray(q)
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
    estimator.positions = np.ascontiguousarray(pose)
    estimator.meas_err = np.ascontiguousarray(np.transpose(resid, (0, 1, 2, 4, 3)))      # [N,C,24,2,W]
    estimator.costs = {"measurement": float(cost[0]), "model": float(cost[1]),
                       "pose": float(pc[0]) if has_priors else 0.0, "motion": float(pc[1]) if has_priors else 0.0}
    estimator.result = dict(status=status, cost=cost, c_eq=c_eq, v_in=v_in, counters=counters)
    if solver_output:
        it, acc = counters
        print(f"LM/AL: {it} iterations ({acc} accepted), max |joint eq| {np.abs(c_eq).max():.2e}, "
              f"{estimator.opt_time_s:.3f} s")
        estimator.solution_details()
    estimator.com_pos, estimator.com_vel = misc.get_com(q, estimator.lengths, estimator.radii, estimator.masses, scene.fps)
    ret = estimator.is_solution_acceptable(estimator.result, 0.1, 0.2)
    if ret and save:
        fname = f"fte_kinematic{'_gt' if params.hand_labeled_data else ''}"
        fname = fname if scene.cam_idx is None or monocular_constraints else "fte_kinematic_orig"
        fname = fname if scene.cam_idx is None else f"{fname}_{scene.cam_idx}"
        estimator.save(fname, out_dir_prefix=out_dir_prefix)
        if scene.cam_idx is not None and monocular_constraints:
            misc.compare_traj_error(params, scene, include_kinetic=False, out_dir_prefix=out_dir_prefix)
    return ret


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
                        n_outer: int = 8,
                        models_dir: Optional[str] = None) -> bool:
    """Same arguments and side effects as the reference (acinoset_opt.py:539-635): initial guess from the spline
    track, optimisation, COM, acceptance test, `save` into `fte_kinematic[_orig][_<cam>]`.  With
    `monocular_constraints` on a single-camera scene the GMM pose prior and the linear motion model are fitted on the
    CPU from `<models_dir or ./models/data-driven>/dataset_full_pose.h5` (its `.csv` twin when PyTables is absent), as
    the reference does (acinoset_misc.py:294-302, 684-689), and evaluated on the GPU inside the LM loop.  A batch of
    one through `estimate_kinematics_batch`."""
    return estimate_kinematics_batch([estimator], solver_output, monocular_constraints, disable_pose_prior,
                                     disable_motion_prior, pose_model_num_components, motion_model_window_size,
                                     motion_model_sparse_solution, out_dir_prefix, max_iter, n_outer, models_dir,
                                     distributed=False)[0]
