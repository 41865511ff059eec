"""Host-side mirror of the reference's `acinoset_misc` functions that sit on either side of the hot path: scene and
DLC loading, measurement packing, trajectory initialisation, COM, metrics and the on-disk output formats.
Same names, argument meaning and file formats as the reference (paths below are in the reference repo); none of
this is timed arithmetic of the path -- it runs once per clip on the CPU.

  TrajectoryParams / Scene      acinoset_misc.py:40-65
  load_scene / find_scene_file  acinoset_misc.py:1496-1542
  DLC loading                   acinoset_misc.py:1545-1572 (+ :195-208)
  measurement packing           acinoset_misc.py:211-256 (weights/likelihood gate :231, frame offsets :224, :246)
  create_trajectory_estimate    acinoset_misc.py:381-456, triangulation :1422-1493
  get_com                       acinoset_misc.py:722-742
  save_optimised_cheetah        acinoset_misc.py:1402-1411 ; save_3d_cheetah_as_2d :1346-1399
  traj_error                    acinoset_misc.py:1179-1199
"""
import json
import os
import pickle
from dataclasses import dataclass
from errno import ENOENT
from glob import glob
from typing import Dict, List, Optional

import numpy as np

from . import skeleton as sk


@dataclass
class TrajectoryParams:
    data_dir: str
    start_frame: int
    end_frame: int
    total_length: int
    dlc_thresh: float
    sync_offset: Optional[List[Dict]]
    hand_labeled_data: bool
    kinetic_dataset: bool
    enable_shutter_delay_estimation: bool
    enable_ppms: bool


@dataclass
class Scene:
    scene_fpath: str
    k_arr: np.ndarray
    d_arr: np.ndarray
    r_arr: np.ndarray
    t_arr: np.ndarray
    cam_res: tuple
    fps: float
    n_cams: int
    cam_idx: Optional[int] = None


def get_markers():
    return list(sk.MARKERS)


def get_dlc_marker_indices():
    return {m: i for m, i in zip(sk.MARKERS, sk.DLC_INDEX)}


def load_scene(fpath, verbose=True):
    with open(fpath, "r", encoding="utf-8") as f:
        data = json.load(f)
    cam_res = tuple(data["camera_resolution"])
    k_arr = np.array([c["k"] for c in data["cameras"]], dtype=np.float64)
    d_arr = np.array([c["d"] for c in data["cameras"]], dtype=np.float64)
    r_arr = np.array([c["r"] for c in data["cameras"]], dtype=np.float64)
    t_arr = np.array([c["t"] for c in data["cameras"]], dtype=np.float64)
    if verbose:
        print(f"Loaded extrinsics from {fpath}\n")
    return k_arr, d_arr, r_arr, t_arr, cam_res


def find_scene_file(dir_path, scene_fname=None, verbose=True):
    if scene_fname is None:
        n_cams = len(glob(os.path.join(dir_path, "cam[1-9].mp4")))
        scene_fname = f"{n_cams}_cam_scene_sba.json" if n_cams else "[1-9]_cam_scene*.json"
    if dir_path and dir_path != os.path.sep and dir_path != os.path.join("..", "data"):
        scene_fpath = os.path.join(dir_path, "extrinsic_calib", scene_fname)
        scene_files = sorted(f for f in glob(scene_fpath) if ("before_corrections" not in f) or (f == scene_fpath))
        if scene_files:
            k_arr, d_arr, r_arr, t_arr, cam_res = load_scene(scene_files[-1], verbose)
            scene_fname = os.path.basename(scene_files[-1])
            return k_arr, d_arr, r_arr, t_arr, cam_res, int(scene_fname[0]), scene_files[-1]
        return find_scene_file(os.path.dirname(dir_path), scene_fname, verbose)
    raise FileNotFoundError(ENOENT, os.strerror(ENOENT), os.path.join("extrinsic_calib", scene_fname))


class DlcData(list):
    """Per-camera DLC tables, `self[c]` = float array [rows, 75] (or [rows, 50] hand-labelled) addressed by ROW POSITION
    as the reference's `base_data[cam] = list(df.to_numpy())` (acinoset_misc.py:207); `first_frame[c]` = frame number
    of row 0 (the DataFrame index the reference's `load_dlc_points_as_df` works with, :1545-1572)."""

    def __init__(self, arrays=(), first_frame=None):
        super().__init__(arrays)
        self.first_frame = list(first_frame) if first_frame is not None else [0] * len(self)


def _hand_label_frame(ix):
    """Frame number of a hand-labelled row: characters 3..5 of the image name, `img053.png` -> 53
    (`int(dlc_df.index[0][-1][3:6])`, acinoset_misc.py:1550-1551); the index entry is a tuple of path parts (h5) or
    one path string (DeepLabCut's CSV twin)."""
    name = ix[-1] if isinstance(ix, tuple) else os.path.basename(str(ix).replace("\\", "/"))
    return int(name[3:6])


def _read_dlc_file(path, hand_labeled=False):
    """One camera's DLC predictions as (float array [rows, 75], frame number of row 0).  `.h5` is the reference's
    format (pandas HDF, needs PyTables); `.csv` is DeepLabCut's own CSV twin of the same table (3 header rows) and
    `.npy` a raw dump starting at frame 0 -- both accepted because PyTables may be absent.  Hand-labelled tables hold
    (x, y) pairs per part, [rows, 50], and name their rows by image file."""
    ext = os.path.splitext(path)[1]
    if ext == ".npy":
        return np.load(path).astype(np.float64), 0
    import pandas as pd
    if ext == ".csv":
        # the three header rows (scorer / bodyparts / coords) are skipped rather than parsed: with header=[0, 1, 2] pandas
        # takes an all-NaN first data row (an unlabelled frame) for the index-name row and drops it
        df = pd.read_csv(path, header=None, skiprows=3, index_col=0)
    else:
        df = pd.read_hdf(path)          # raises ImportError without PyTables -- no silent substitute
    if not len(df.index):
        return df.to_numpy(dtype=np.float64), 0
    first = _hand_label_frame(df.index[0]) if hand_labeled else int(df.index[0])
    return df.to_numpy(dtype=np.float64), first


def dlc_files(dlc_dir):
    """Sorted per-camera files; `.h5` wins when present (reference behaviour, acinoset_misc.py:197)."""
    for ext in ("h5", "csv", "npy"):
        files = sorted(glob(os.path.join(dlc_dir, f"*.{ext}")))
        if files:
            return files
    return []


def load_dlc_arrays(dlc_dir, n_cams=None, hand_labeled=False):
    files = dlc_files(dlc_dir)
    if n_cams is not None:
        assert n_cams == len(files), f"# of dlc files != # of cams in {n_cams}_cam_scene_sba.json"
    tabs = [_read_dlc_file(p, hand_labeled) for p in files]
    return DlcData([t[0] for t in tabs], [t[1] for t in tabs])


PW_SUFFIX = "DLC_resnet152_CheetahOct14shuffle4_650000.pickle"      # acinoset_misc.py:203-205


def load_pairwise(dlc_dir, n_cams=None):
    """Pairwise pseudo-measurement files of `enable_ppm` (acinoset_misc.py:200-205): `<dlc_dir>_pw/<first 4 characters
    of the camera's DLC file name>DLC_resnet152_..._650000.pickle`, one per camera; each is indexable by frame row and
    yields {"pose": [75] (x, y, likelihood per DLC part), "pws": [k, 25, 25, 2] pairwise offsets}."""
    import pickle
    out = []
    for path in dlc_files(dlc_dir):
        with open(os.path.join(dlc_dir + "_pw", os.path.basename(path)[:4] + PW_SUFFIX), "rb") as f:
            out.append(pickle.load(f))
    if n_cams is not None:
        assert n_cams == len(out)
    return out


def pack_measurements(params: TrajectoryParams, scene: Scene, base_data, pw_data=None):
    """meas[N,C,24,W,4] float32 (float64 for the kinetic dataset) = (u, v, gate, 0) in marker order, C = all cameras or the single monocular one.
    Row of camera c for frame n: (n-1) + start_frame - sync_offset[c]  (acinoset_misc.py:224, 246), a ROW POSITION of
    the table as in the reference (not a frame label); gate = likelihood > dlc_thresh (strict, :231).  NaN detections get gate 0 (documented divergence).
    W = 1, or 3 with `params.enable_ppms`: hypotheses w = 2, 3 are pairwise pseudo-measurements
    `pose[base] + pws[0, base, dlc_index[marker]]` with `base = pairwise_graph[marker][w-2]`, gated by the
    likelihood of the base part (acinoset_misc.py:225-228, 249-254)."""
    N = params.end_frame - params.start_frame
    W = 3 if params.enable_ppms else 1
    sync = [0] * scene.n_cams
    if params.sync_offset is not None:
        for off in params.sync_offset:
            sync[off["cam"]] = off["frame"]
    cams = list(range(scene.n_cams)) if scene.cam_idx is None else [scene.cam_idx]
    idx = np.array(sk.DLC_INDEX)
    # AcinoSet detections are float32 DLC output; the kinetic dataset's 2-D points are not: keep those in double
    meas = np.zeros((N, len(cams), 24, W, 4), np.float64 if params.kinetic_dataset else np.float32)
    for ci, c in enumerate(cams):
        if params.hand_labeled_data:
            # hand labels: (x, y) pairs per part, rows NOT shifted by the sync offset, likelihood 1
            # (acinoset_misc.py:216-221, 241-242)
            rows = np.arange(N) + params.start_frame
            d = np.asarray(base_data[c])[rows]
            x, y = d[:, 0::2][:, idx], d[:, 1::2][:, idx]
            lik = np.ones_like(x)
        else:
            rows = np.arange(N) + params.start_frame - sync[c]
            d = np.asarray(base_data[c])[rows]
            x, y, lik = d[:, 0::3][:, idx], d[:, 1::3][:, idx], d[:, 2::3][:, idx]
        bad = ~(np.isfinite(x) & np.isfinite(y) & np.isfinite(lik))
        gate = (lik > params.dlc_thresh) & ~bad
        meas[:, ci, :, 0, 0] = np.where(bad, 0.0, x)
        meas[:, ci, :, 0, 1] = np.where(bad, 0.0, y)
        meas[:, ci, :, 0, 2] = gate
        if W == 3:
            assert pw_data is not None, "enable_ppm needs the dlc_pw/*.pickle files (load_pairwise)"
            pair = np.array(sk.PAIRWISE)                                   # [24, 2] DLC part indices
            for n in range(N):
                values = pw_data[c][int(rows[n])]
                pose = np.asarray(values["pose"], dtype=np.float64)
                pws = np.asarray(values["pws"], dtype=np.float64)
                for w in (1, 2):
                    base = pair[:, w - 1]
                    xw = pose[0::3][base] + pws[0, base, idx, 0]
                    yw = pose[1::3][base] + pws[0, base, idx, 1]
                    lw = pose[2::3][base]
                    badw = ~(np.isfinite(xw) & np.isfinite(yw) & np.isfinite(lw))
                    meas[n, ci, :, w, 0] = np.where(badw, 0.0, xw)
                    meas[n, ci, :, w, 1] = np.where(badw, 0.0, yw)
                    meas[n, ci, :, w, 2] = (lw > params.dlc_thresh) & ~badw
    return meas, cams


# ---- initialisation (acinoset_misc.py:381-456) -------------------------------------------------------------------
def triangulate_points_fisheye(img_pts_1, img_pts_2, k1, d1, r1, t1, k2, d2, r2, t2):
    import cv2 as cv
    pts_1 = cv.fisheye.undistortPoints(img_pts_1.reshape((-1, 1, 2)), k1, d1)
    pts_2 = cv.fisheye.undistortPoints(img_pts_2.reshape((-1, 1, 2)), k2, d2)
    pts_4d = cv.triangulatePoints(np.hstack((r1, t1)), np.hstack((r2, t2)), pts_1, pts_2)
    return (pts_4d[:3] / pts_4d[3]).T


def triangulate_points(img_pts_1, img_pts_2, k1, d1, r1, t1, k2, d2, r2, t2):
    import cv2 as cv
    pts_1 = cv.undistortPoints(img_pts_1.reshape((-1, 1, 2)), k1, d1)
    pts_2 = cv.undistortPoints(img_pts_2.reshape((-1, 1, 2)), k2, d2)
    pts_4d = cv.triangulatePoints(np.hstack((r1, t1)), np.hstack((r2, t2)), pts_1, pts_2)
    return (pts_4d[:3] / pts_4d[3]).T


def triangulate_points_single_img(img_pts, dist_to_plane, K, D, R, t):
    import cv2 as cv
    norm_pts = cv.fisheye.undistortPoints(img_pts.reshape((-1, 1, 2)), K, D).reshape((-1, 2))
    norm_pts = np.append(norm_pts, np.ones((len(norm_pts), 1)), 1).T
    X_c = dist_to_plane * norm_pts
    return np.dot(R.T, X_c) - np.dot(R.T, t)


def create_trajectory_estimate(base_length, params: TrajectoryParams, scene: Scene, kinetic_dataset: bool,
                               base_data=None, robust=True):
    """Initial base track and heading (acinoset_misc.py:381-456): pairwise triangulation of the `spine` detections
    above the likelihood threshold, a smoothing spline per axis over absolute frame numbers, yaw from finite
    differences (+ pi: "assumes running in +x").  Returns x_est, y_est, z_est, psi_est indexed by absolute frame.
    `base_length` replaces `robot["base"].length`.  Rows are addressed by FRAME NUMBER here (row position +
    `base_data.first_frame[c]`), as the reference's DataFrame path does.  Hand-labelled tables (:425-439): (x, y) pairs
    per part, no likelihood filter, unlabelled frames stay NaN and enter the spline with weight 0.
    `robust=True` (default) adds two guards that do nothing on clean input and are the only deviations from the
    reference's initial guess: one re-fit of the splines without gross triangulation outliers, and unwrapping of the
    heading across atan2's branch cut; `robust=False` is the reference's literal procedure."""
    from scipy.interpolate import UnivariateSpline
    hand = params.hand_labeled_data
    if base_data is None:
        dlc_dir = os.path.join(params.data_dir, "dlc" if not hand else "dlc_hand_labeled")
        base_data = load_dlc_arrays(dlc_dir, hand_labeled=hand)
    first = getattr(base_data, "first_frame", [0] * len(base_data))
    spine = sk.DLC_INDEX[sk.MARKERS.index("spine")]
    sync = {}
    if params.sync_offset is not None:
        for off in params.sync_offset:
            sync[off["cam"]] = off["frame"]
    per_cam = {}
    for c, d in enumerate(base_data):
        d = np.asarray(d)
        if hand:
            x, y = d[:, 2 * spine], d[:, 2 * spine + 1]
            ok = np.ones(len(x), bool)                       # NaN labels are kept (weight 0 in the spline)
        else:
            x, y, lik = d[:, 3 * spine], d[:, 3 * spine + 1], d[:, 3 * spine + 2]
            ok = np.isfinite(x) & np.isfinite(y) & (lik > params.dlc_thresh)
        frames = np.nonzero(ok)[0] + first[c] + sync.get(c, 0)
        per_cam[c] = (frames, np.stack([x[ok], y[ok]], 1).astype(np.float32))
    k = 1 if kinetic_dataset else 3
    if scene.cam_idx is None:
        n_cams = len(scene.k_arr)
        use = [c for c in per_cam if (c < 2 or not kinetic_dataset)]
        tri = triangulate_points if kinetic_dataset else triangulate_points_fisheye
        acc = {}
        for i in range(n_cams):
            a, b = i % n_cams, (i + 1) % n_cams
            if a not in use or b not in use:
                continue
            fa, pa = per_cam[a]
            fb, pb = per_cam[b]
            common, ia, ib = np.intersect1d(fa, fb, return_indices=True)
            if len(common) == 0:
                continue
            p3 = tri(pa[ia], pb[ib], scene.k_arr[a], scene.d_arr[a], scene.r_arr[a], scene.t_arr[a],
                     scene.k_arr[b], scene.d_arr[b], scene.r_arr[b], scene.t_arr[b])
            for f, p in zip(common, p3):
                acc.setdefault(int(f), []).append(p)
        frames = np.array(sorted(acc))
        with np.errstate(invalid="ignore"):
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore", RuntimeWarning)           # all-NaN frames of hand labels
                spine_pts = np.array([[f, *np.nanmean(np.asarray(acc[f], dtype=np.float64), axis=0)] for f in frames],
                                     dtype=np.float64)
    else:
        c = scene.cam_idx
        fa, pa = per_cam[c]
        xw = triangulate_points_single_img(pa, 3, scene.k_arr[c], scene.d_arr[c], scene.r_arr[c], scene.t_arr[c]).T
        spine_pts = np.c_[fa.astype(np.float64), xw]
    spine_pts[:, 1] = spine_pts[:, 1] + float(base_length) / 2.0
    if hand:
        # one (possibly NaN) sample per frame 0..max, NaN samples weighted 0 (acinoset_misc.py:425-439); frames before
        # the first label are absent in the reference's table -- here they are NaN samples of weight 0 as well
        num_frames = int(spine_pts[:, 0].max()) + 1
        frame_est = np.arange(num_frames)
        vals = np.full((num_frames, 3), np.nan)
        vals[spine_pts[:, 0].astype(int)] = spine_pts[:, 1:4]
        w = np.isnan(vals).any(axis=1)
        vals[w] = 0.0
        x_est, y_est, z_est = (np.array(UnivariateSpline(frame_est, vals[:, i], w=(~w).astype(float), k=k)(frame_est))
                               for i in range(3))
    else:
        frame_est = np.arange(params.end_frame)
        keep = np.ones(len(spine_pts), bool)
        for _ in range(2 if robust else 1):
            fits = [UnivariateSpline(spine_pts[keep, 0], spine_pts[keep, i], k=k) for i in (1, 2, 3)]
            if not robust:
                break
            # one re-fit without gross outliers of the triangulation (a wrong spine detection that passed the likelihood
            # gate): points further than max(5 MAD, 0.25 m) from the first fit.  Nothing is dropped on clean tracks, where
            # this is exactly the reference's single fit.
            dev = np.sqrt(sum((f(spine_pts[:, 0]) - spine_pts[:, i]) ** 2 for f, i in zip(fits, (1, 2, 3))))
            thr = max(5.0 * 1.4826 * float(np.median(dev)), 0.25)
            new_keep = dev <= thr
            if new_keep.all() or new_keep.sum() <= k + 1:
                break
            keep = new_keep
        x_est, y_est, z_est = (np.array(f(frame_est)) for f in fits)
    dx_est = np.diff(x_est) * scene.fps
    dy_est = np.diff(y_est) * scene.fps
    psi_est = np.arctan2(dy_est, dx_est)
    psi_est = np.pi + np.append(psi_est, [psi_est[-1]])
    if robust:
        # atan2's branch cut: a heading that crosses +-pi would jump by 2 pi between frames, which the motion model
        # penalises although it is the same orientation; identical to the reference whenever there is no crossing
        psi_est = np.unwrap(psi_est)
    return x_est, y_est, z_est, psi_est


# ---- outputs -------------------------------------------------------------------------------------------------------
def link_centres(q, lengths, radii):
    """Pb_I of the 17 links, [N,17,3] (links.py:496-519 with the topology of cheetah.py:109-200)."""
    q = np.asarray(q, float)
    L, rad = np.asarray(lengths, float), np.asarray(radii, float)
    ang = q[:, 3:].reshape(-1, 17, 3)
    sf, cf = np.sin(ang[..., 0]), np.cos(ang[..., 0])
    st, ct = np.sin(ang[..., 1]), np.cos(ang[..., 1])
    sp, cp = np.sin(ang[..., 2]), np.cos(ang[..., 2])
    R = np.empty(ang.shape[:2] + (3, 3))
    R[..., 0, 0] = cp * ct; R[..., 0, 1] = cp * st * sf - sp * cf; R[..., 0, 2] = cp * st * cf + sp * sf
    R[..., 1, 0] = sp * ct; R[..., 1, 1] = sp * st * sf + cp * cf; R[..., 1, 2] = sp * st * cf - cp * sf
    R[..., 2, 0] = -st;     R[..., 2, 1] = ct * sf;                R[..., 2, 2] = ct * cf

    def rv(link, v):
        return R[:, link] @ np.asarray(v, float)

    Pb = np.zeros((q.shape[0], 17, 3))
    Pb[:, sk.BASE] = q[:, 0:3]
    base_bottom = Pb[:, sk.BASE] + rv(sk.BASE, [-L[sk.BASE] / 2, 0, 0])
    base_top = Pb[:, sk.BASE] + rv(sk.BASE, [L[sk.BASE] / 2, 0, 0])
    Pb[:, sk.BODYF] = base_bottom + rv(sk.BODYF, [-L[sk.BODYF] / 2, 0, 0])
    bodyf_bottom = base_bottom + rv(sk.BODYF, [-L[sk.BODYF], 0, 0])
    Pb[:, sk.NECK] = bodyf_bottom + rv(sk.NECK, [-L[sk.NECK] / 2, 0, 0])
    Pb[:, sk.TAIL0] = base_top + rv(sk.TAIL0, [L[sk.TAIL0] / 2, 0, 0])
    tail0_bottom = base_top + rv(sk.TAIL0, [L[sk.TAIL0], 0, 0])
    Pb[:, sk.TAIL1] = tail0_bottom + rv(sk.TAIL1, [L[sk.TAIL1] / 2, 0, 0])
    for front, right, U, Lo, H in ((True, True, sk.UFR, sk.LFR, sk.HFR), (True, False, sk.UFL, sk.LFL, sk.HFL),
                                   (False, True, sk.UBR, sk.LBR, sk.HBR), (False, False, sk.UBL, sk.LBL, sk.HBL)):
        sy = 1.0 if right else -1.0
        if front:
            top = Pb[:, sk.BODYF] + rv(sk.BODYF, [-L[sk.BODYF] / 2, sy * rad[sk.BODYF], 0])
        else:
            top = Pb[:, sk.BASE] + rv(sk.BASE, [L[sk.BASE] / 2, sy * rad[sk.BASE], 0])
        for link in (U, Lo, H):
            Pb[:, link] = top + rv(link, [0, 0, -L[link] / 2])
            top = top + rv(link, [0, 0, -L[link]])
    return Pb


def get_com(q, lengths, radii, masses, fps):
    """(com_pos[N,3], com_vel[N-1,3]) as acinoset_misc.py:722-742."""
    Pb = link_centres(q, lengths, radii)
    m = np.asarray(masses, float)
    com = np.einsum("l,nlk->nk", m, Pb) * (1.0 / m.sum())
    return com, (com[1:, :] - com[:-1, :]) * fps


def save_optimised_cheetah(positions, out_fpath, extra_data=None):
    file_data = dict(positions=positions)
    if extra_data is not None:
        assert type(extra_data) is dict
        file_data.update(extra_data)
    with open(out_fpath, "wb") as f:
        pickle.dump(file_data, f)
    print("Saved", out_fpath)


def project_points_fisheye(obj_pts, k, d, r, t):
    import cv2 as cv
    r_vec = cv.Rodrigues(r)[0]
    return cv.fisheye.projectPoints(obj_pts.reshape((-1, 1, 3)), r_vec, t, k, d)[0].reshape((-1, 2))


def project_points(obj_pts, k, d, r, t):
    import cv2 as cv
    r_vec = cv.Rodrigues(r)[0]
    return cv.projectPoints(obj_pts.reshape((-1, 1, 3)), r_vec, t, k, d)[0].reshape((-1, 2))


def save_3d_cheetah_as_2d(positions_3d_arr, out_dir, scene_fpath, bodyparts, project_func, start_frame,
                          sync_offset_arr, vid_dir=None, save_as_csv=True, out_fname=None):
    """Per-camera 2D projections in DLC layout (acinoset_misc.py:1346-1399): `cam<i>_<fname>.csv` always, `.h5`
    (key `<fname>_df`, table format) when PyTables is importable.  As in the reference nothing is written when no
    `cam[1-9].mp4` is found next to the data."""
    import pandas as pd
    if vid_dir:
        video_fpaths = sorted(glob(os.path.join(vid_dir, "cam[1-9].mp4")))
    else:
        video_fpaths = sorted(glob(os.path.join(out_dir, "cam[1-9].mp4")))
        if not video_fpaths:
            video_fpaths = sorted(glob(os.path.join(os.path.dirname(out_dir), "cam[1-9].mp4")))
    if not video_fpaths:
        print("Could not save 3D cheetah to 2D - No videos were found in", out_dir, "or", os.path.dirname(out_dir))
        return []
    k_arr, d_arr, r_arr, t_arr, cam_res = load_scene(scene_fpath, verbose=False)
    assert len(k_arr) == len(video_fpaths)
    pdindex = pd.MultiIndex.from_product([bodyparts, ["x", "y", "likelihood"]], names=["bodyparts", "coords"])
    out_fname = os.path.basename(out_dir) if out_fname is None else out_fname
    written = []
    for i in range(len(video_fpaths)):
        position_3d = np.array(positions_3d_arr[i])
        n_frames = len(position_3d)
        projections = project_func(position_3d, k_arr[i], d_arr[i], r_arr[i], t_arr[i])
        out_of_range = np.where((projections > cam_res) | (projections < [0] * 2))[0]
        projections[out_of_range] = np.nan
        data = np.full(position_3d.shape, np.nan)
        data[:, :, 0:2] = projections.reshape((n_frames, -1, 2))
        cam_name = os.path.splitext(os.path.basename(video_fpaths[i]))[0]
        fpath = os.path.join(out_dir, cam_name + "_" + out_fname + ".h5")
        df = pd.DataFrame(data.reshape((n_frames, -1)), columns=pdindex,
                          index=range(start_frame - sync_offset_arr[i], start_frame + n_frames - sync_offset_arr[i]))
        if save_as_csv:
            df.to_csv(os.path.splitext(fpath)[0] + ".csv")
            written.append(os.path.splitext(fpath)[0] + ".csv")
        try:
            import tables  # noqa: F401
            df.to_hdf(fpath, key=f"{out_fname}_df", format="table", mode="w")
            written.append(fpath)
        except ImportError:
            pass                      # PyTables absent on this box: CSV twin only
    return written


def traj_error(X, Y, centered=False):
    """Per-marker position error in mm (acinoset_misc.py:1179-1199): mean over frames of ||X - Y||, optionally after
    removing each frame's marker centroid (MPJPE)."""
    X, Y = np.asarray(X, float), np.asarray(Y, float)
    if centered:
        X = X - np.expand_dims(np.mean(X, axis=1), axis=1)
        Y = Y - np.expand_dims(np.mean(Y, axis=1), axis=1)
    distances = np.sqrt(np.sum((X - Y) ** 2, axis=2))
    return 1000.0 * np.mean(distances, axis=0)


def compare_traj_error(params: TrajectoryParams, scene: Scene, include_kinetic=False, kinetic_out_fname: str = "fte",
                       out_dir_prefix: Optional[str] = None):
    """The numbers of the reference's `compare_traj_error` (acinoset_misc.py:1202-1264) without its plots: position
    error (mm) of the single-view and the data-driven single-view reconstructions against the multi-view one.
    Returns {"single_view": (mpe, mpjpe), "data_driven": (mpe, mpjpe)}; entries whose file is missing are skipped."""
    data_dir = params.data_dir if out_dir_prefix is None else os.path.join(
        out_dir_prefix, params.data_dir.split("cheetah_videos")[1][1:])
    _ = (include_kinetic, kinetic_out_fname)
    paths = dict(multi=os.path.join(data_dir, "fte_kinematic", "fte.pickle"),
                 single_view=os.path.join(data_dir, f"fte_kinematic_orig_{scene.cam_idx}", "fte.pickle"),
                 data_driven=os.path.join(data_dir, f"fte_kinematic_{scene.cam_idx}", "fte.pickle"))
    if not os.path.exists(paths["multi"]):
        return {}
    with open(paths["multi"], "rb") as f:
        gt = pickle.load(f)["positions"]
    out = {}
    for key in ("single_view", "data_driven"):
        if not os.path.exists(paths[key]):
            continue
        with open(paths[key], "rb") as f:
            pos = pickle.load(f)["positions"]
        out[key] = (float(traj_error(gt.copy(), pos.copy()).mean()), float(traj_error(gt.copy(), pos.copy(), True).mean()))
        print(f"{key}: MPE {out[key][0]:.1f} mm, MPJPE {out[key][1]:.1f} mm")
    return out
