"""Rebuilds an AcinoSet-style clip directory (inputs of init_trajectory) from a golden fixture whose cameras and
DLC detections were recovered from the reference's shipped outputs."""
import json
import os

import numpy as np

from cheetah_pose_estimation_b200 import skeleton as sk


def build_clip_dir(root, data_path, g, likelihood=None, sync_offset=None, pad_frames=5, resolution=None):
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
                   "monocular_cam": 0}, f)
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
                                               float_format="%.9g")
    return d
