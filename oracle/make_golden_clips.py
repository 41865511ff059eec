"""Rebuild cameras + DLC detections of EVERY multi-view clip shipped under the reference's data/test_set (SURVEY.md
section 8c G8, section 8d config C2) and write them as ONE fixture, tests/golden/test_set_clips.npz.

    python oracle/make_golden_clips.py          # build container only: needs /root/reference

TEST INFRASTRUCTURE.  Nothing here is on the product path, and nothing in tests/, bench.py or smoke() reads
/root/reference at run time: this script is run once where the reference exists and its output is committed.

Per clip the reference ships `fte_kinematic/fte.pickle` (positions[N,24,3], meas_err[N,C,24,2,1], q, ...) and
`cam<i>_fte.csv` (OpenCV projections of `positions`, NaN outside the image).  Fitting the 14 camera parameters to the
(positions -> csv) pairs reproduces the calibration to ~1e-10 px, and then  DLC = proj_ref(positions) - meas_err
with `proj_ref` the function the reference's cost uses (misc.pt3d_to_2d_fisheye for AcinoSet, misc.pt3d_to_2d for the
kinetic dataset, acinoset_misc.py:1663-1696) lands on the float32 grid of the original detections.

  AcinoSet clips (10, the `dataset` tuple of run_dataset.py:1387-1398): csv = cv.fisheye.projectPoints, same model
     as the cost (Kannala-Brandt) -> fit with oracle.cameras.project(FISHEYE).
  kinetic_dataset clips (5): csv = cv.projectPoints with D = (k1, k2, p1, p2) (acinoset_misc.py:1332-1336) while the
     cost treats D[2] as the r^6 coefficient -> fit the OpenCV pinhole model, keep (K, D, R, t) and
     evaluate the reference's own pt3d_to_2d with them for the detections.

Keys of the fixture: `names` and, per clip index i, `<i>_cams [C,20]`, `<i>_rms [C]`, `<i>_detections [N,C,24,2]` (f32 for AcinoSet: they land on the float32 grid; f64 for the kinetic dataset),
`<i>_q`, `<i>_positions`, `<i>_meas_err`, `<i>_obj_cost`, `<i>_fps`, `<i>_start_frame`, `<i>_processing_time_s`,
`<i>_grid_err` (distance of the recovered detections from the float32 grid), `<i>_kinetic`, `<i>_animal`,
`<i>_monocular_cam` (from the shipped `fte_kinematic_<cam>` directory name).
"""
import glob
import os
import pickle
import sys
from multiprocessing import Pool

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, os.path.join(HERE, "ref_shim"))
sys.path.insert(1, REF)
sys.path.insert(2, REPO)

import numpy as np  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")

# run_dataset.py:1387-1398 (`dataset`), in the reference's order, then the five kinetic-dataset clips
ACINOSET = (("jules", "2017_12_09/bottom", "flick2"), ("jules", "2019_03_09", "flick1"), ("phantom", "2019_03_03", "run"),
            ("phantom", "2017_09_02/top", "run1_2"), ("jules", "2017_08_29/top", "run1_2"),
            ("phantom", "2017_08_29/top", "run1_1"), ("jules", "2017_08_29/top", "run1_1"),
            ("jules", "2017_09_02/top", "run1"), ("phantom", "2019_03_07", "run"), ("jules", "2017_09_02/bottom", "run2"))
KINETIC = (("arabia", "kinetic_dataset/2009_09_07", "trial06"), ("shiraz", "kinetic_dataset/2009_09_07", "trial04"),
           ("shiraz", "kinetic_dataset/2009_09_08", "trial04"), ("shiraz", "kinetic_dataset/2009_09_11", "trial01"),
           ("shiraz", "kinetic_dataset/2009_09_11", "trial02"))


def opencv_pinhole(X, p, R):
    """cv.projectPoints with D = (k1, k2, p1, p2): radial r^2, r^4 and tangential terms."""
    Xc = X @ R.T + p[11:14]
    a, b = Xc[:, 0] / Xc[:, 2], Xc[:, 1] / Xc[:, 2]
    r2 = a * a + b * b
    rad = 1 + p[4] * r2 + p[5] * r2 * r2
    xp = a * rad + 2 * p[6] * a * b + p[7] * (r2 + 2 * a * a)
    yp = b * rad + p[6] * (r2 + 2 * b * b) + 2 * p[7] * a * b
    return np.stack([p[0] * xp + p[2], p[1] * yp + p[3]], 1)


def recover_clip(args):
    animal, date, trial = args
    import cv2
    import pandas as pd
    from scipy.optimize import least_squares
    from oracle import cameras as ocam
    os.chdir(REF)
    import acinoset_misc as misc
    kinetic = date.startswith("kinetic_dataset")
    clip = f"{date}/{animal}/{trial}"
    base = f"data/test_set/{clip}"
    d = pickle.load(open(f"{base}/fte_kinematic/fte.pickle", "rb"))
    pos = d["positions"]
    N = pos.shape[0]
    csvs = sorted(glob.glob(f"{base}/fte_kinematic/cam*_fte.csv"))
    assert len(csvs) == d["meas_err"].shape[1]
    cam_list, dets, rms = [], [], []
    good = None

    def unpack(p):
        R, _ = cv2.Rodrigues(p[8:11])
        return np.concatenate([p[:8], R.reshape(-1), p[11:14]])

    for ci, path in enumerate(csvs):
        # header=None: an all-NaN first frame would otherwise be taken for the index-name row of the 2-row header
        arr = pd.read_csv(path, header=None, skiprows=2, index_col=0).to_numpy(dtype=np.float64).reshape(N, 24, 3)[:, :, :2]
        ok = np.isfinite(arr).all(-1)
        obj, img = pos[ok], arr[ok]

        def fun(p):
            pp = p.copy()
            pp[0], pp[1] = abs(pp[0]), abs(pp[1])
            if kinetic:
                R, _ = cv2.Rodrigues(pp[8:11])
                return (opencv_pinhole(obj, pp, R) - img).reshape(-1)
            return (ocam.project(obj, unpack(pp), ocam.FISHEYE) - img).reshape(-1)

        starts = []
        for f0, c0 in ((1000.0, (960.0, 540.0)), (860.0, (960.0, 540.0)), (1500.0, (1352.0, 760.0)),
                       (1200.0, (640.0, 360.0)), (2500.0, (640.0, 360.0)), (600.0, (960.0, 540.0))):
            K0 = np.array([[f0, 0, c0[0]], [0, f0, c0[1]], [0, 0, 1]])
            for flag in (cv2.SOLVEPNP_EPNP, cv2.SOLVEPNP_ITERATIVE):
                try:
                    okp, rv, tv = cv2.solvePnP(obj, img, K0, None, flags=flag)
                except cv2.error:
                    continue
                if okp:
                    starts.append(np.concatenate([[f0, f0, c0[0], c0[1], 0, 0, 0, 0], rv.reshape(-1), tv.reshape(-1)]))
        if good is not None:                      # intrinsics of a converged camera of the same rig
            Kg = np.array([[good[0], 0, good[2]], [0, good[1], good[3]], [0, 0, 1]])
            und = (cv2.undistortPoints(img.reshape(-1, 1, 2), Kg, good[4:8]) if kinetic else
                   cv2.fisheye.undistortPoints(img.reshape(-1, 1, 2), Kg, good[4:8])).reshape(-1, 2)
            okp, rv, tv = cv2.solvePnP(obj, und, np.eye(3), None, flags=cv2.SOLVEPNP_ITERATIVE)
            starts.insert(0, np.concatenate([good[:8], rv.reshape(-1), tv.reshape(-1)]))
        best, best_rr = None, np.inf
        for p0 in starts:
            res = least_squares(fun, p0, method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=6000)
            rr = float(np.sqrt(np.mean(res.fun ** 2)))
            if rr < best_rr:
                best, best_rr = res.x.copy(), rr
            if best_rr < 1e-8:
                break
        best[0], best[1] = abs(best[0]), abs(best[1])
        cam = unpack(best)
        if best_rr < 1e-8 and good is None:
            good = cam[:8].copy()
        rms.append(best_rr)
        cam_list.append(cam)
        # detections = what the reference's COST projects - the saved residual
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]])
        D, R, t = cam[4:8], cam[8:17].reshape(3, 3), cam[17:20].reshape(3, 1)
        fproj = misc.pt3d_to_2d if kinetic else misc.pt3d_to_2d_fisheye
        proj = np.array([[fproj(*pos[n, l], K, D, R, t) for l in range(24)] for n in range(N)], dtype=np.float64)
        dets.append(proj - d["meas_err"][:, ci, :, :, 0])
        print(f"{clip} cam {ci}: rms {best_rr:.2e} fx {cam[0]:.2f} fy {cam[1]:.2f} c ({cam[2]:.1f}, {cam[3]:.1f}) "
              f"D {np.round(cam[4:8], 4)}", flush=True)
    det = np.stack(dets, 1)
    det32 = det.astype(np.float32)
    grid = float(np.abs(det - det32).max())
    if kinetic:                 # the kinetic dataset's 2-D points are not float32 DLC output: keep them in double
        det32 = det
    mono = sorted(glob.glob(f"{base}/fte_kinematic_[0-9]"))
    mono_cam = int(mono[0].rsplit("_", 1)[1]) if mono else -1
    fps = 200.0 if kinetic else (120.0 if "2019" in clip else 90.0)
    print(f"{clip}: N {N} C {len(csvs)} grid error {grid:.2e} monocular cam {mono_cam}", flush=True)
    return dict(name=clip, animal=animal, cams=np.array(cam_list), rms=np.array(rms), detections=det32, q=d["q"],
                positions=pos, meas_err=d["meas_err"], obj_cost=np.float64(d["obj_cost"]), fps=np.float64(fps),
                start_frame=np.int64(d["start_frame"]), processing_time_s=np.float64(d["processing_time_s"]),
                grid_err=np.float64(grid), kinetic=np.bool_(kinetic), monocular_cam=np.int64(mono_cam))


def main():
    clips = list(ACINOSET) + list(KINETIC)
    with Pool(min(8, len(clips))) as pool:
        out = pool.map(recover_clip, clips)
    save = {"names": np.array([o["name"] for o in out])}
    for i, o in enumerate(out):
        for k, v in o.items():
            if k != "name":
                save[f"{i}_{k}"] = np.array(v) if k == "animal" else v
    np.savez_compressed(os.path.join(OUT, "test_set_clips.npz"), **save)
    for i, o in enumerate(out):
        print(i, o["name"], "max cam rms", o["rms"].max(), "grid", o["grid_err"])


if __name__ == "__main__":
    main()
