"""CPU tests of the host-side mirror of the reference API (acinoset_misc / acinoset_opt look-alikes): output
formats and derived quantities against the reference's shipped files, input packing and initialisation."""
import os

import numpy as np
import pytest

from cheetah_pose_estimation_b200 import acinoset_misc as misc, skeleton as sk
from tests._dataset import build_clip_dir

CLIPS = [("jules", "clip_jules_run1_1.npz", "2017_08_29/top/jules/run1_1"),
         ("phantom", "clip_phantom_run.npz", "2019_03_07/phantom/run")]


@pytest.mark.parametrize("animal,fname,path", CLIPS)
def test_com_and_reduced_pose_match_shipped_outputs(golden_dir, animal, fname, path):
    g = np.load(os.path.join(golden_dir, fname))
    ref = np.load(os.path.join(golden_dir, "save_ref.npz"))
    lengths, radii, masses = sk.animal(animal)
    com, vel = misc.get_com(g["q"], lengths, radii, masses, float(g["fps"]))
    assert np.abs(com - ref[f"{animal}_com_pos"]).max() < 1e-12
    assert np.abs(vel - ref[f"{animal}_com_vel"]).max() < 1e-9
    T = sk.relative_matrix()[sk.independent_coords()]
    for k, a in (("x", g["q"]), ("dx", g["dq"]), ("ddx", g["ddq"])):
        assert np.abs(a @ T.T - ref[f"{animal}_{k}"]).max() <= 1e-12 * max(1.0, np.abs(ref[f"{animal}_{k}"]).max())


@pytest.mark.parametrize("animal,fname,path", CLIPS)
def test_2d_projection_csv_matches_shipped_file(golden_dir, tmp_path, animal, fname, path):
    g = np.load(os.path.join(golden_dir, fname))
    ref = np.load(os.path.join(golden_dir, "save_ref.npz"))
    d = build_clip_dir(str(tmp_path), path, g)
    out_dir = os.path.join(d, "fte_kinematic")
    os.makedirs(out_dir)
    C = g["cams"].shape[0]
    scene_fpath = os.path.join(d, "extrinsic_calib", f"{C}_cam_scene_sba.json")
    written = misc.save_3d_cheetah_as_2d([g["positions"]] * C, out_dir, scene_fpath, misc.get_markers(),
                                         misc.project_points_fisheye, int(g["start_frame"]), [0] * C, vid_dir=d,
                                         out_fname="fte")
    assert os.path.join(out_dir, "cam1_fte.csv") in written
    with open(os.path.join(out_dir, "cam1_fte.csv")) as f:
        head = [next(f) for _ in range(2)]
    assert head == list(ref[f"{animal}_cam1_head"])[:2]          # identical 2-row header (bodyparts / coords)
    import pandas as pd
    df = pd.read_csv(os.path.join(out_dir, "cam1_fte.csv"), header=[0, 1], index_col=0)
    assert np.array_equal(df.index.to_numpy(), ref[f"{animal}_cam1_index"])
    got, want = df.to_numpy(), ref[f"{animal}_cam1_csv"]
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.nanmax(np.abs(got - want)) < 1e-6                  # recovered camera reproduces the file to < 1e-6 px


def test_measurement_packing_and_sync_offsets(golden_dir, tmp_path):
    g = np.load(os.path.join(golden_dir, "clip_jules_run1_1.npz"))
    N, C = g["detections"].shape[:2]
    rng = np.random.default_rng(0)
    lik = rng.uniform(0, 1, (N, C, 24))
    lik[0, 0, 0] = 0.5                                            # strict '>' : exactly at the threshold is off
    sync = [{"cam": 2, "frame": 3}, {"cam": 4, "frame": -2}]
    d = build_clip_dir(str(tmp_path), "2017_08_29/top/jules/run1_1", g, likelihood=lik, sync_offset=sync)
    k_arr, d_arr, r_arr, t_arr, cam_res, n_cams, scene_fpath = misc.find_scene_file(d)
    assert n_cams == C and cam_res == (1920, 1080)
    start = int(g["start_frame"])
    params = misc.TrajectoryParams(d, start, start + N, N, 0.5, sync, False, False, False, False)
    scene = misc.Scene(scene_fpath, k_arr, d_arr.reshape(-1, 4), r_arr, t_arr, cam_res, 90.0, n_cams, None)
    base = misc.load_dlc_arrays(os.path.join(d, "dlc"), n_cams)
    meas, cams = misc.pack_measurements(params, scene, base)
    assert meas.shape == (N, C, 24, 1, 4) and cams == list(range(C))
    assert np.array_equal(meas[..., 0, 0:2], g["detections"])     # float32-exact detections, offsets undone
    assert np.array_equal(meas[..., 0, 2] > 0, lik > 0.5) and meas[0, 0, 0, 0, 2] == 0.0
    scene.cam_idx = 3
    meas1, cams1 = misc.pack_measurements(params, scene, base)
    assert cams1 == [3] and np.array_equal(meas1[:, 0], meas[:, 3])


def test_trajectory_estimate_tracks_the_spine(golden_dir, tmp_path):
    g = np.load(os.path.join(golden_dir, "clip_phantom_run.npz"))
    N, C = g["detections"].shape[:2]
    d = build_clip_dir(str(tmp_path), "2019_03_07/phantom/run", g)
    k_arr, d_arr, r_arr, t_arr, cam_res, n_cams, scene_fpath = misc.find_scene_file(d)
    start = int(g["start_frame"])
    params = misc.TrajectoryParams(d, start, start + N, N, 0.5, None, False, False, False, False)
    scene = misc.Scene(scene_fpath, k_arr, d_arr.reshape(-1, 4), r_arr, t_arr, cam_res, 120.0, n_cams, None)
    lengths, _, _ = sk.animal("phantom")
    x, y, z, psi = misc.create_trajectory_estimate(lengths[sk.BASE], params, scene, False)
    fr = start + np.arange(N)
    track = np.stack([x[fr], y[fr], z[fr]], 1)
    # the reference shifts x by L_base/2 (acinoset_misc.py:424); the shipped optimum's base stays within ~0.5 m
    assert np.abs(track - g["q"][:, 0:3]).max() < 0.6
    heading = g["q"][:, 5]
    dpsi = np.angle(np.exp(1j * (psi[fr] - heading)))
    assert np.abs(dpsi).max() < 0.5


def test_pairwise_pseudo_measurements_follow_the_reference_indexing(tmp_path):
    """enable_ppm (W = 3): literal per-element restatement of acinoset_misc.py:211-256 against pack_measurements."""
    import pickle
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, skeleton as sk
    rng = np.random.default_rng(5)
    n_cams, n_rows, start, N = 2, 14, 3, 6
    dlc_dir = tmp_path / "dlc"
    (tmp_path / "dlc_pw").mkdir()
    dlc_dir.mkdir()
    base_data, pw_files = [], []
    for c in range(n_cams):
        arr = rng.uniform(0, 1000, (n_rows, 75))
        arr[:, 2::3] = rng.uniform(0, 1, (n_rows, 25))
        np.save(dlc_dir / f"cam{c + 1}DLC_x.npy", arr)
        base_data.append(arr)
        pw = [{"pose": np.concatenate([rng.uniform(0, 1000, 50), rng.uniform(0, 1, 25)])[np.argsort(np.tile(np.arange(25), 3), kind="stable")],
               "pws": rng.normal(0, 20, (2, 25, 25, 2))} for _ in range(n_rows)]
        for v in pw:                                                        # (x, y, lik) interleaved per part
            p = np.empty(75)
            p[0::3], p[1::3], p[2::3] = rng.uniform(0, 1000, 25), rng.uniform(0, 1000, 25), rng.uniform(0, 1, 25)
            v["pose"] = p
        with open(tmp_path / "dlc_pw" / f"cam{c + 1}{misc.PW_SUFFIX}", "wb") as f:
            pickle.dump(pw, f)
        pw_files.append(pw)
    sync = [{"cam": 1, "frame": 2}]
    params = misc.TrajectoryParams(str(tmp_path), start, start + N, N, 0.5, sync, False, False, False, True)
    scene = misc.Scene("", None, None, None, None, None, 90, n_cams, None)
    pw_data = misc.load_pairwise(str(dlc_dir), n_cams)
    meas, cams = misc.pack_measurements(params, scene, misc.load_dlc_arrays(str(dlc_dir), n_cams), pw_data)
    assert meas.shape == (N, n_cams, 24, 3, 4) and cams == [0, 1]
    # the reference's dictionaries (acinoset_misc.py:1943-1998), from the pinned tables
    index_dict = dict(zip(sk.MARKERS, sk.DLC_INDEX))
    pair_dict = dict(zip(sk.MARKERS, sk.PAIRWISE))
    sync_arr = [0, 2]
    for n in range(1, N + 1):
        for c in range(1, n_cams + 1):
            row = (n - 1) + start - sync_arr[c - 1]
            for l, marker in enumerate(sk.MARKERS):
                for w in (1, 2, 3):
                    if w < 2:
                        base = index_dict[marker]
                        lik = base_data[c - 1][row][2::3][base]
                        val = [base_data[c - 1][row][d::3][base] for d in (0, 1)]
                    else:
                        values = pw_files[c - 1][row]
                        base = pair_dict[marker][w - 2]
                        lik = values["pose"][2::3][base]
                        val = [values["pose"][d::3][base] + values["pws"][:, :, :, d][0, base, index_dict[marker]]
                               for d in (0, 1)]
                    got = meas[n - 1, c - 1, l, w - 1]
                    assert got[0] == np.float32(val[0]) and got[1] == np.float32(val[1])
                    assert got[2] == (1.0 if lik > 0.5 else 0.0)


def test_hand_labelled_packing(tmp_path):
    """dlc_hand_labeled arrays hold (x, y) pairs, are not shifted by the sync offset and have likelihood 1
    (acinoset_misc.py:216-221, 241-242); NaN labels get gate 0 (SURVEY A5 divergence)."""
    rng = np.random.default_rng(8)
    arr = [rng.uniform(0, 1000, (12, 50)) for _ in range(2)]
    arr[1][5, 2 * sk.DLC_INDEX[3]] = np.nan
    params = misc.TrajectoryParams(str(tmp_path), 2, 9, 7, 0.5, [{"cam": 1, "frame": 1}], True, False, False, False)
    scene = misc.Scene("", None, None, None, None, None, 90, 2, None)
    meas, cams = misc.pack_measurements(params, scene, arr)
    assert meas.shape == (7, 2, 24, 1, 4)
    for c in range(2):
        for l, base in enumerate(sk.DLC_INDEX):
            for n in range(7):
                x, y = arr[c][n + 2][0::2][base], arr[c][n + 2][1::2][base]
                if np.isnan(x):
                    assert meas[n, c, l, 0, 2] == 0.0
                else:
                    assert meas[n, c, l, 0, 0] == np.float32(x) and meas[n, c, l, 0, 1] == np.float32(y)
                    assert meas[n, c, l, 0, 2] == 1.0


def _hand_label_csv(path, arr, first_frame, parts):
    """DeepLabCut `CollectedData_<scorer>.csv` layout: 3 header rows, rows named by image path, (x, y) per part."""
    import pandas as pd
    cols = pd.MultiIndex.from_product([["scorer"], parts, ["x", "y"]], names=["scorer", "bodyparts", "coords"])
    idx = [f"labeled-data/cam/img{first_frame + i:03d}.png" for i in range(arr.shape[0])]
    pd.DataFrame(arr, columns=cols, index=idx).to_csv(path)


def test_hand_labelled_initialisation(tmp_path):
    """`hand_labeled_data=True` through the loaders and `create_trajectory_estimate` (acinoset_misc.py:381-456,
    1545-1572): rows named by image file (frame = characters 3..5), (x, y) pairs per part, no likelihood filter, NaN
    labels kept and weighted 0 in the spline.  A point moving on a straight line seen by two cameras must come back
    as that line (spine + L_base / 2) with heading pi + atan2(dy, dx)."""
    import cv2
    rng = np.random.default_rng(11)
    first, n_rows = 4, 16
    parts = [f"part{i}" for i in range(25)]
    for m, i in zip(sk.MARKERS, sk.DLC_INDEX):
        parts[i] = m
    spine = sk.DLC_INDEX[sk.MARKERS.index("spine")]
    K = np.array([[900.0, 0, 960], [0, 900, 540], [0, 0, 1]])
    D = np.zeros(4)
    cams = []
    for ang in (0.0, 0.5):
        R = cv2.Rodrigues(np.array([0.0, ang, 0.0]))[0] @ cv2.Rodrigues(np.array([np.pi / 2, 0.0, 0.0]))[0]
        t = -R @ np.array([[2.0 + 4 * ang], [-10.0], [1.0]])
        cams.append((R, t))
    frames = first + np.arange(n_rows)
    track = np.stack([1.0 + 0.1 * frames, 0.3 + 0.02 * frames, 0.7 + 0 * frames], 1)
    os.makedirs(tmp_path / "dlc_hand_labeled")
    for c, (R, t) in enumerate(cams):
        arr = rng.uniform(100, 900, (n_rows, 50))
        uv = cv2.fisheye.projectPoints(track.reshape(-1, 1, 3), cv2.Rodrigues(R)[0], t, K, D)[0].reshape(-1, 2)
        arr[:, 2 * spine], arr[:, 2 * spine + 1] = uv[:, 0], uv[:, 1]
        if c == 1:
            arr[6, 2 * spine] = arr[6, 2 * spine + 1] = np.nan              # an unlabelled frame
        _hand_label_csv(tmp_path / "dlc_hand_labeled" / f"cam{c + 1}.csv", arr, first, parts)
    data = misc.load_dlc_arrays(str(tmp_path / "dlc_hand_labeled"), 2, hand_labeled=True)
    assert data.first_frame == [first, first] and data[0].shape == (n_rows, 50)
    params = misc.TrajectoryParams(str(tmp_path), first, first + n_rows, n_rows, 0.5, None, True, False, False, False)
    scene = misc.Scene("", np.stack([K, K]), np.stack([D, D]), np.stack([c[0] for c in cams]),
                       np.stack([c[1] for c in cams]), (1920, 1080), 90, 2, None)
    base_len = 0.4
    x, y, z, psi = misc.create_trajectory_estimate(base_len, params, scene, False, data)
    assert len(x) == first + n_rows
    assert np.abs(x[frames] - (track[:, 0] + base_len / 2)).max() < 2e-3
    assert np.abs(y[frames] - track[:, 1]).max() < 2e-3 and np.abs(z[frames] - track[:, 2]).max() < 2e-3
    assert np.abs(psi[frames[:-1]] - (np.pi + np.arctan2(0.02, 0.1))).max() < 2e-2
    # packing: hand-labelled rows are addressed by row position start_frame + n, not shifted by the sync offset
    meas, _ = misc.pack_measurements(params, scene, DlcPad(data))
    assert meas.shape == (n_rows, 2, 24, 1, 4)


class DlcPad(list):
    """Hand-labelled tables as the reference indexes them: row position = frame number (the shipped label files start at
    frame 0); pads the fixture's rows to that layout."""

    def __init__(self, data):
        super().__init__(np.concatenate([np.full((f, a.shape[1]), np.nan), a]) for a, f in zip(data, data.first_frame))
        self.first_frame = [0] * len(data)


def test_dlc_rows_are_addressed_by_position_like_the_reference(tmp_path):
    """`init_measurements` indexes `list(df.to_numpy())` by (n-1) + start_frame - sync (acinoset_misc.py:207, 224, 246): a
    table whose index starts above 0 is NOT re-based for the packing, only for the trajectory initialisation."""
    import pandas as pd
    rng = np.random.default_rng(3)
    arr = rng.uniform(0, 1000, (10, 75))
    arr[:, 2::3] = 0.9
    cols = pd.MultiIndex.from_product([["s"], [f"p{i}" for i in range(25)], ["x", "y", "likelihood"]])
    os.makedirs(tmp_path / "dlc")
    pd.DataFrame(arr, columns=cols, index=np.arange(5, 15)).to_csv(tmp_path / "dlc" / "cam1.csv")
    data = misc.load_dlc_arrays(str(tmp_path / "dlc"), 1)
    assert data.first_frame == [5] and data[0].shape == (10, 75)
    params = misc.TrajectoryParams(str(tmp_path), 2, 6, 4, 0.5, None, False, False, False, False)
    scene = misc.Scene("", None, None, None, None, None, 90, 1, None)
    meas, _ = misc.pack_measurements(params, scene, data)
    l = 4
    assert meas[0, 0, l, 0, 0] == np.float32(arr[2, 3 * sk.DLC_INDEX[l]])


def test_base_rates_match_the_oracle_and_the_collocation_rows():
    """engine.base_rates (host-side torch plumbing of the shutter-delay driver) against the oracle's restatement, on a
    ragged batch, and against the implicit-Euler rows it eliminates (collocation.py:58-68)."""
    import torch
    from cheetah_pose_estimation_b200 import engine
    from oracle import cost
    rng = np.random.default_rng(4)
    q = rng.normal(size=(4, 9, 54))
    h = np.array([1 / 90, 1 / 120, 1 / 200, 1 / 90])
    nf = np.array([9, 3, 2, 1], np.int32)
    v, a = engine.base_rates(torch.as_tensor(q[..., 0:3]), torch.as_tensor(h), torch.as_tensor(nf))
    for s in range(4):
        vo, ao = cost.base_rates(q[s, :nf[s]], h[s])
        assert np.array_equal(v[s, :nf[s]].numpy(), vo) and np.array_equal(a[s, :nf[s]].numpy(), ao)
    b, vv, aa = q[0, :, 0:3], v[0].numpy(), a[0].numpy()
    assert np.abs(b[1:] - b[:-1] - h[0] * vv[1:]).max() < 1e-14          # q_n = q_{n-1} + h dq_n
    assert np.abs(vv[1:] - vv[:-1] - h[0] * aa[1:]).max() < 1e-10        # dq_n = dq_{n-1} + h ddq_n, incl. the first node
    assert np.array_equal(aa[0], aa[1]) and np.array_equal(aa[1], aa[2])


def test_contact_detection_reproduces_the_shipped_contact_files(golden_dir):
    """kinetics.contact_detection (= misc.contact_detection, acinoset_misc.py:745-856, driven as opt.determine_contacts
    does, acinoset_opt.py:638-690) on the shipped monocular trajectories of all ten clips: both shipped contact files
    (`autogen-contact.json`: height + velocity + stance-time heuristic with leading / trailing labels;
    `autogen-contact-02.json`: height threshold only) come back exactly.  The clip's ground-plane height is the one input
    the shipped files do not carry; the fixture holds the value the second file pins (oracle/make_golden_contacts.py)."""
    import json
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "contacts_ref.npz"))
    assert len(g["names"]) == 10
    for i, name in enumerate(g["names"]):
        st = {"q": g[f"{i}_q"], "dq": g[f"{i}_dq"], "com_vel": g[f"{i}_com_vel"]}
        c, c2 = kinetics.determine_contacts_from_states(st, str(g[f"{i}_animal"]), int(g[f"{i}_start_frame"]),
                                                        float(g[f"{i}_fps"]), float(g[f"{i}_ground"]))
        assert c == json.loads(str(g[f"{i}_contacts"])), name
        assert c2 == json.loads(str(g[f"{i}_contacts_02"])), name
        lo, hi = g[f"{i}_ground_interval"]
        assert 0.002 < hi - lo < 0.02                      # the height-only file pins the ground plane to millimetres


def test_foot_velocity_is_the_time_derivative_of_the_foot_position(golden_dir):
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "contacts_ref.npz"))
    lengths, radii, _ = sk.animal("jules")
    q, dq = g["0_q"][:5], g["0_dq"][:5]
    pos, vel = kinetics.foot_height_and_velocity(q, dq, lengths, radii)
    eps = 1e-6
    pp, _ = kinetics.foot_height_and_velocity(q + eps * dq, dq, lengths, radii)
    pm, _ = kinetics.foot_height_and_velocity(q - eps * dq, dq, lengths, radii)
    assert np.abs((pp - pm) / (2 * eps) - vel).max() < 1e-6 * max(1.0, np.abs(vel).max())
    # the foot point is the paw marker of the pose function (acinoset_misc.py:1640-1659)
    from oracle import kinematics
    cl, cv = kinematics.marker_chains(lengths, radii)
    pose = kinematics.fk(q, cl, cv)
    for f, name in enumerate(kinetics.FEET):
        assert np.abs(pose[:, sk.MARKERS.index(kinetics.PAW_MARKERS[name])] - pos[:, f]).max() < 1e-13


def test_contact_prescription_as_bound_arrays(golden_dir):
    """kinetics.contact_times / prescribe_contact_order (acinoset_opt.py:795-812, acinoset_misc.py:1140-1167) on a shipped
    contact file: 1-based INCLUSIVE stance indices, GRF fixed to zero in flight, GRFz >= 0.01 and the foot height within
    the uncertainty band in stance; the shipped trajectory itself keeps its feet within 0.1 m of the ground plane in the
    stance frames the shipped file prescribes (the reference's band for AcinoSet clips, acinoset_opt.py:782)."""
    import json
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "contacts_ref.npz"))
    q, dq = g["0_q"], g["0_dq"]
    n = q.shape[0]
    start = int(g["0_start_frame"])
    cj = {"start_frame": start, "end_frame": start + n, "contacts": json.loads(str(g["0_contacts"]))}
    times = kinetics.contact_times(cj, n_frames=n)
    assert len(times) == 4 and any(times)
    for f, name in enumerate(kinetics.FEET):
        want = []
        for seq in (cj["contacts"].get(name) or []):
            want += [fe for fe in range(seq[0] - start + 1, seq[1] - start + 2) if 1 <= fe <= n]      # inclusive, 1-based
        assert times[f] == want
    b = kinetics.prescribe_contact_order(times, n, foot_height_uncertainty=0.1)
    st = b["stance"]
    assert st.shape == (n, 4) and [int(x) for x in st.sum(0)] == [len(set(t)) for t in times]
    assert np.all(b["grfz_ub"][~st] == 0.0) and np.all(b["grfxy_ub"][~st] == 0.0) and np.all(b["grfz_lb"][~st] == 0.0)
    assert np.all(b["grfz_lb"][st] == 0.01) and np.all(np.isinf(b["grfz_ub"][st]))
    assert np.all(b["height_ub"][st] == 0.1) and np.all(b["height_lb"][st] == -0.1) and np.all(np.isinf(b["height_ub"][~st]))
    # the half-open mask estimate_grf uses covers the same ranges minus their last frame
    mask = kinetics.contact_mask(cj).astype(bool)
    assert np.all(st[:-1][mask[:-1]]) and st.sum() >= mask.sum()
    lengths, radii, _ = sk.animal(str(g["0_animal"]))
    grf = np.zeros((n, 20))
    grf[:, 0:4] = np.where(st, 0.5, 0.0)
    grf[0, 0] = 0.25 if not st[0, 0] else 0.0                           # a force in flight (if the foot is in flight there)
    v = kinetics.contact_violations(q, dq, lengths, radii, b, grf=grf, ground_plane_height=float(g["0_ground"]))
    assert v["height"].shape == (4,) and v["height"].max() < 0.05, v
    assert v["stance_shortfall"].max() == 0.0 and v["flight_force"][0] == (0.25 if not st[0, 0] else 0.0)
    unbounded = kinetics.prescribe_contact_order(times, n, foot_height_uncertainty=None)
    assert np.all(np.isinf(unbounded["height_ub"]))
