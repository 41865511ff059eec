"""Rebuilds an AcinoSet-style clip directory (inputs of init_trajectory) from a golden fixture whose cameras and
DLC detections were recovered from the reference's shipped outputs."""
import json
import os

import numpy as np

from cheetah_pose_estimation_b200 import skeleton as sk


class Clip(dict):
    """Clip `i` of tests/golden/test_set_clips.npz (oracle/make_golden_clips.py) with the key prefix stripped."""

    def __init__(self, g, i):
        pre = f"{i}_"
        super().__init__({k[len(pre):]: g[k] for k in g.keys() if k.startswith(pre)})
        self["name"] = str(g["names"][i])
        self["animal"] = str(self["animal"])


def load_test_set_clips(golden_dir, kinetic=False):
    """[Clip] of the 10 AcinoSet clips in the order of the reference's `dataset` tuple (run_dataset.py:1387-1398), or of
    the 5 kinetic-dataset clips."""
    import os
    g = np.load(os.path.join(golden_dir, "test_set_clips.npz"))
    clips = [Clip(g, i) for i in range(len(g["names"]))]
    return [c for c in clips if bool(c["kinetic"]) == kinetic]


def inferred_likelihood(clip):
    """A stand-in for the DLC likelihoods the shipped outputs do not carry: 0.1 (below the 0.5 gate) for detections whose
    weighted residual AT THE SHIPPED OPTIMUM lies beyond the redescending loss's last break point (|w r| > c = 20,
    acinoset_misc.py:476-482) in either coordinate, 0.9 otherwise.  Beyond c the loss is constant, so such a detection
    exerts no force on the optimum whether it was gated or not: the gate cannot change the shipped solution's
    stationarity, it only keeps the garbage (unlabelled points parked far away: 20-24 % of the kinetic dataset's
    detections sit > 150 px from the model) out of the initial spine track."""
    from cheetah_pose_estimation_b200 import skeleton as sk2
    kin = bool(clip["kinetic"])
    w = np.asarray(sk2.build_tables(kin)["meas_w"], float).reshape(3, 24)[0]          # 1 / R_pw[0][l]
    C = clip["detections"].shape[1]
    mult = np.array([1.0, 1.0, 0.6, 0.6])[:C] if kin else np.ones(C)
    e = np.abs(clip["meas_err"][..., 0]) * w[None, None, :, None] * mult[None, :, None, None]
    return np.where((e > 20.0).any(-1), 0.1, 0.9)


def build_clip_dir(root, data_path, g, likelihood=None, sync_offset=None, pad_frames=5, resolution=None,
                   monocular_cam=0):
    """root/data_path gets metadata.json, extrinsic_calib/<C>_cam_scene_sba.json, cam<i>.mp4 (empty), dlc/cam<i>.csv
    (DeepLabCut CSV layout: scorer/bodyparts/coords header rows, 25 body parts x (x, y, likelihood))."""
    import pandas as pd
    N, C = g["detections"].shape[:2]
    start = int(g["start_frame"])
    d = os.path.join(root, data_path)
    os.makedirs(os.path.join(d, "dlc"), exist_ok=True)
    os.makedirs(os.path.join(d, "extrinsic_calib"), exist_ok=True)
    cams = g["cams"]
    if resolution is None:                       # AcinoSet: 2017 clips 1920x1080, 2019 clips 2704x1520
        resolution = [2704, 1520] if "2019" in data_path else [1920, 1080]
    scene = {"camera_resolution": list(resolution), "cameras": []}
    for c in range(C):
        cam = cams[c]
        scene["cameras"].append({"k": [[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], "d": [[x] for x in cam[4:8]],
                                 "r": cam[8:17].reshape(3, 3).tolist(), "t": [[x] for x in cam[17:20]]})
        open(os.path.join(d, f"cam{c + 1}.mp4"), "wb").close()
    with open(os.path.join(d, "extrinsic_calib", f"{C}_cam_scene_sba.json"), "w") as f:
        json.dump(scene, f)
    sync = [0] * C
    for off in (sync_offset or []):
        sync[off["cam"]] = off["frame"]
    with open(os.path.join(d, "metadata.json"), "w") as f:
        json.dump({"start_frame": start, "end_frame": start + N, "cam_sync": sync_offset, "ground_plane_height": 0.0,
                   "monocular_cam": int(monocular_cam)}, f)
    parts = [f"part{i}" for i in range(25)]
    for m, i in zip(sk.MARKERS, sk.DLC_INDEX):
        parts[i] = m
    cols = pd.MultiIndex.from_product([["DLC_resnet"], parts, ["x", "y", "likelihood"]],
                                      names=["scorer", "bodyparts", "coords"])
    lik = np.ones((N, C, 24)) if likelihood is None else likelihood
    for c in range(C):
        n_rows = start + N + pad_frames
        arr = np.zeros((n_rows, 75))
        arr[:, 2::3] = 0.01                                            # frames outside the clip: low likelihood
        rows = np.arange(N) + start - sync[c]
        for l, i in enumerate(sk.DLC_INDEX):
            arr[rows, 3 * i] = g["detections"][:, c, l, 0]
            arr[rows, 3 * i + 1] = g["detections"][:, c, l, 1]
            arr[rows, 3 * i + 2] = lik[:, c, l]
        pd.DataFrame(arr, columns=cols).to_csv(os.path.join(d, "dlc", f"cam{c + 1}DLC_resnet.csv"),
                                               float_format="%.9g" if g["detections"].dtype == np.float32 else "%.17g")
    return d


def write_data_driven_dataset(golden_dir, models_dir):
    """A stand-in for the reference's models/data-driven/dataset_full_pose.csv (which does not travel to the GPU box):
    the reduced poses `x` of the reference's shipped trajectories (tests/golden/shipped_clips.npz), one segment per
    clip, index restarting at 0 -- the same table layout (28 pose columns, 28 derivative columns)."""
    import pandas as pd
    g = np.load(os.path.join(golden_dir, "shipped_clips.npz"))
    rows, idx = [], []
    for key in sorted((k for k in g.keys() if k.endswith("_x")), key=lambda k: int(k.split("_")[0])):
        x = g[key]
        if x.shape[0] < 9:
            continue
        rows.append(np.concatenate([x, np.gradient(x, axis=0)], axis=1))
        idx.append(np.arange(x.shape[0]))
    os.makedirs(models_dir, exist_ok=True)
    pd.DataFrame(np.concatenate(rows), index=np.concatenate(idx)).to_csv(os.path.join(models_dir, "dataset_full_pose.csv"))


def write_pairwise_files(root, data_path, g, sync_offset=None, pad_frames=5):
    """`dlc_pw/cam<i>DLC_resnet152_..._650000.pickle` next to a clip built by `build_clip_dir` (enable_ppm,
    acinoset_misc.py:200-205, 249-254): per row {"pose": the row's 75 DLC values, "pws": [1,25,25,2] offsets}, with
    pws[0, base, part] = detection(part) - detection(base), i.e. pseudo-measurements consistent with the detections."""
    import pickle
    from cheetah_pose_estimation_b200 import acinoset_misc as misc
    N, C = g["detections"].shape[:2]
    start = int(g["start_frame"])
    sync = [0] * C
    for off in (sync_offset or []):
        sync[off["cam"]] = off["frame"]
    d = os.path.join(root, data_path, "dlc_pw")
    os.makedirs(d, exist_ok=True)
    for c in range(C):
        n_rows = start + N + pad_frames
        rows = []
        for r in range(n_rows):
            pose = np.zeros(75)
            pose[2::3] = 0.01
            pws = np.zeros((1, 25, 25, 2))
            n = r - start + sync[c]
            if 0 <= n < N:
                det = np.zeros((25, 2))
                for l, i in enumerate(sk.DLC_INDEX):
                    det[i] = g["detections"][n, c, l]
                    pose[3 * i:3 * i + 3] = (det[i, 0], det[i, 1], 1.0)
                pws[0] = det[None, :, :] - det[:, None, :]
            rows.append({"pose": pose, "pws": pws})
        with open(os.path.join(d, f"cam{c + 1}{misc.PW_SUFFIX}"), "wb") as f:
            pickle.dump(rows, f)


def write_hand_labels(root, data_path, g, missing=()):
    """`dlc_hand_labeled/cam<i>.csv` next to a clip built by `build_clip_dir`: DeepLabCut's CollectedData layout (3 header
    rows, rows named by image file `img<frame>.png`, (x, y) per part, NaN = not labelled), one row per frame from 0 to the
    end of the clip (the reference addresses these tables by row position = frame, acinoset_misc.py:216-221)."""
    import pandas as pd
    N, C = g["detections"].shape[:2]
    start = int(g["start_frame"])
    d = os.path.join(root, data_path, "dlc_hand_labeled")
    os.makedirs(d, exist_ok=True)
    parts = [f"part{i}" for i in range(25)]
    for m, i in zip(sk.MARKERS, sk.DLC_INDEX):
        parts[i] = m
    cols = pd.MultiIndex.from_product([["labeller"], parts, ["x", "y"]], names=["scorer", "bodyparts", "coords"])
    for c in range(C):
        arr = np.full((start + N, 50), np.nan)
        for l, i in enumerate(sk.DLC_INDEX):
            arr[start:, 2 * i] = g["detections"][:, c, l, 0]
            arr[start:, 2 * i + 1] = g["detections"][:, c, l, 1]
        for (n, cc, l) in missing:
            if cc == c:
                arr[start + n, 2 * sk.DLC_INDEX[l]:2 * sk.DLC_INDEX[l] + 2] = np.nan
        idx = [f"labeled-data/cam{c + 1}/img{f:03d}.png" for f in range(start + N)]
        pd.DataFrame(arr, columns=cols, index=idx).to_csv(os.path.join(d, f"cam{c + 1}.csv"))
