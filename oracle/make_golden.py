"""Generate tests/golden/*.npz from the REAL reference (run in the build container only).

    python oracle/make_golden.py            # needs /root/reference; writes tests/golden/

TEST INFRASTRUCTURE.  The reference is Python; it is imported here (with the pyomo/matplotlib
shims in oracle/ref_shim, because neither is installed) and its OWN functions are evaluated on seeded
inputs.  The outputs are committed as small fixtures because /root/reference does not exist on
the GPU box.  Nothing in tests/, bench.py or smoke() reads /root/reference at run time.

Fixtures written (reference function -> file):
  cheetah.model + misc.get_pose_state (lambdified FK)          -> fk_ref.npz
  sympy.diff of the same symbolic marker positions               -> fk_jac_ref.npz
  robot link angle_constraints (links.py:552-586), lambdified    -> constraints_ref.npz
  misc.pt3d_to_2d_fisheye / misc.pt3d_to_2d                      -> projection_ref.npz
  misc.redescending_loss                                         -> loss_ref.npz
  misc.get_relative_angles + get_relative_angle_mask             -> reduced_pose_ref.npz
  misc.get_uncertainty_models / get_markers / get_dlc_marker_indices / get_pairwise_graph -> tables_ref.npz
  shipped data/test_set fte.pickle files (q, positions, x, ...)  -> shipped_clips.npz
  camera + detection recovery from a shipped clip     -> clip_jules_run1_1.npz
  reference-function cost on a small synthetic sequence           -> cost_ref.npz
"""
import glob
import os
import pickle
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, os.path.join(HERE, "ref_shim"))
sys.path.insert(1, REF)
sys.path.insert(2, REPO)

import numpy as np  # noqa: E402
import sympy as sp  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")


def build_robot(name):
    import cheetah
    import cheetah_params
    import acinoset_misc as misc
    robot, _ = cheetah.model(cheetah_params.parameters[name], include_dynamics=False, relative_orientation=False)
    sp_vars, position_funcs = misc.get_pose_state(robot)
    return robot, sp_vars, position_funcs


def random_q(rng, n, heading=True):
    q = rng.uniform(-1.4, 1.4, size=(n, 54))
    q[:, 0] = rng.uniform(-3, 12, n)
    q[:, 1] = rng.uniform(-3, 3, n)
    q[:, 2] = rng.uniform(0.3, 1.0, n)
    return q


def eval_positions(position_funcs, q, lengths):
    states = np.concatenate([q, np.tile(lengths, (q.shape[0], 1))], axis=1)
    return np.array([[[position_funcs[l][d](states[f]) for d in range(3)] for l in range(24)]
                     for f in range(q.shape[0])])


def main():
    os.makedirs(OUT, exist_ok=True)
    os.chdir(REF)
    import acinoset_misc as misc
    import cheetah_params
    rng = np.random.default_rng(20261104)

    # ---- FK from the reference's lambdified functions, all animals --------------------------------
    fk = {}
    robots = {}
    for name in ("jules", "phantom", "arabia", "shiraz", "acinoset"):
        robot, sp_vars, pfs = build_robot(name)
        robots[name] = (robot, sp_vars, pfs)
        lengths = np.array([l.length for l in robot.links], float)
        radii = np.array([l.radius for l in robot.links], float)
        masses = np.array([l.mass for l in robot.links], float)
        q = random_q(rng, 6)
        fk[f"{name}_q"] = q
        fk[f"{name}_lengths"] = lengths
        fk[f"{name}_radii"] = radii
        fk[f"{name}_masses"] = masses
        fk[f"{name}_positions"] = eval_positions(pfs, q, lengths)
        print("fk", name)
    np.savez_compressed(os.path.join(OUT, "fk_ref.npz"), **fk)

    # ---- FK Jacobian by sympy differentiation of the reference's symbolic expressions -------------
    robot, sp_vars, pfs = robots["jules"]
    # rebuild the symbolic marker positions exactly as get_pose_state does (acinoset_misc.py:1585-1592)
    r = robot
    exprs = {
        "nose": r["neck"].bottom_I + r["neck"].Rb_I @ sp.Matrix([-0.055, 0, -0.055]),
        "r_eye": r["neck"].bottom_I + r["neck"].Rb_I @ sp.Matrix([0, 0.045, 0]),
        "l_eye": r["neck"].bottom_I + r["neck"].Rb_I @ sp.Matrix([0, -0.045, 0]),
        "neck_base": r["neck"].top_I, "spine": r["base"].bottom_I, "tail_base": r["base"].top_I,
        "tail1": r["tail1"].top_I, "tail2": r["tail1"].bottom_I,
        "r_shoulder": r["bodyF"].bottom_I + r["bodyF"].Rb_I @ sp.Matrix([0.06, 0.075, -0.15]),
        "r_front_knee": r["UFR"].bottom_I, "r_front_ankle": r["HFR"].top_I, "r_front_paw": r["HFR"].bottom_I,
        "l_shoulder": r["bodyF"].bottom_I + r["bodyF"].Rb_I @ sp.Matrix([0.06, -0.075, -0.15]),
        "l_front_knee": r["UFL"].bottom_I, "l_front_ankle": r["HFL"].top_I, "l_front_paw": r["HFL"].bottom_I,
        "r_hip": r["base"].top_I + r["base"].Rb_I @ sp.Matrix([-0.06, 0.06, -0.1]),
        "r_back_knee": r["UBR"].bottom_I, "r_back_ankle": r["HBR"].top_I, "r_back_paw": r["HBR"].bottom_I,
        "l_hip": r["base"].top_I + r["base"].Rb_I @ sp.Matrix([-0.06, -0.06, -0.1]),
        "l_back_knee": r["UBL"].bottom_I, "l_back_ankle": r["HBL"].top_I, "l_back_paw": r["HBL"].bottom_I,
    }
    markers = misc.get_markers()
    qsyms = sp_vars[:54]
    P = sp.Matrix([exprs[m][d] for m in markers for d in range(3)])
    Jsym = P.jacobian(qsyms)
    fJ = sp.lambdify([sp_vars], Jsym, "numpy")
    fP = sp.lambdify([sp_vars], P, "numpy")
    lengths = fk["jules_lengths"]
    qj = random_q(rng, 3)
    Jref = np.array([np.asarray(fJ(np.concatenate([qq, lengths])), float).reshape(24, 3, 54) for qq in qj])
    Pref = np.array([np.asarray(fP(np.concatenate([qq, lengths])), float).reshape(24, 3) for qq in qj])
    assert np.abs(Pref - eval_positions(pfs, qj, lengths)).max() < 1e-13
    np.savez_compressed(os.path.join(OUT, "fk_jac_ref.npz"), q=qj, lengths=lengths, radii=fk["jules_radii"],
                        jac=Jref, positions=Pref)
    print("fk jac")

    # ---- joint equality constraints (links.py:552-586; system.py:574-586 order) ------------------
    cons = [c for link in robot.links for c in link.angle_constraints]
    assert len(cons) == 26
    fC = sp.lambdify([qsyms], sp.Matrix(cons), "numpy")
    qc = random_q(rng, 5)
    cref = np.array([np.asarray(fC(qq), float).reshape(26) for qq in qc])
    from oracle import kinematics as okin
    cmine = okin.joint_constraints(qc)
    perm = []
    for j in range(26):
        dif = np.abs(cmine - cref[:, j:j + 1]).max(axis=0)
        k = int(np.argmin(dif))
        assert dif[k] < 1e-12, (j, dif[k])
        perm.append(k)
    assert sorted(perm) == list(range(26))
    np.savez_compressed(os.path.join(OUT, "constraints_ref.npz"), q=qc, c_ref=cref, oracle_index_of_ref_row=np.array(perm))
    print("constraints")

    # ---- projections ---------------------------------------------------------------------------
    n = 64
    X = np.stack([rng.uniform(-4, 14, n), rng.uniform(-4, 4, n), rng.uniform(0, 1.2, n)], -1)
    cams = []
    uvf, uvp = [], []
    import cv2
    for c in range(4):
        K = np.array([[rng.uniform(800, 1200), 0, rng.uniform(900, 1000)], [0, rng.uniform(800, 1200), rng.uniform(500, 580)], [0, 0, 1]])
        D = np.array([0.12, -0.054, 0.015, -0.001]) * rng.uniform(0.5, 1.5, 4)
        rvec = np.array([np.pi / 2, 0, 0]) + rng.normal(0, 0.3, 3)
        R, _ = cv2.Rodrigues(rvec)
        t = np.array([[rng.uniform(-6, -2)], [rng.uniform(0.5, 1.5)], [rng.uniform(6, 12)]])
        cams.append(np.concatenate([[K[0, 0], K[1, 1], K[0, 2], K[1, 2]], D, R.reshape(-1), t.reshape(-1)]))
        uvf.append([misc.pt3d_to_2d_fisheye(x, y, z, K, D, R, t) for x, y, z in X])
        uvp.append([misc.pt3d_to_2d(x, y, z, K, D, R, t) for x, y, z in X])
    np.savez_compressed(os.path.join(OUT, "projection_ref.npz"), X=X, cams=np.array(cams), uv_fisheye=np.array(uvf, float),
                        uv_pinhole=np.array(uvp, float))
    print("projection")

    # ---- robust loss ---------------------------------------------------------------------------
    err = np.concatenate([np.linspace(-40, 40, 801), rng.normal(0, 3, 200), [0.0, 1e-9, -1e-9, 3, 10, 20]])
    loss = np.array([misc.redescending_loss(e, 3, 10, 20) for e in err], float)
    np.savez_compressed(os.path.join(OUT, "loss_ref.npz"), err=err, loss=loss)

    # ---- reduced pose ---------------------------------------------------------------------------
    qr = random_q(rng, 7)
    mask = misc.get_relative_angle_mask()
    x = np.array([np.array([v for part in misc.get_relative_angles(qr, fe) for v in part])[mask] for fe in range(qr.shape[0])])
    np.savez_compressed(os.path.join(OUT, "reduced_pose_ref.npz"), q=qr, x=x, mask=mask[0])

    # ---- tables ---------------------------------------------------------------------------------
    R_pw, Q = misc.get_uncertainty_models()
    idx = misc.get_dlc_marker_indices()
    pw = misc.get_pairwise_graph()
    np.savez_compressed(os.path.join(OUT, "tables_ref.npz"), R_pw=R_pw, Q=Q, markers=np.array(markers),
                        dlc_index=np.array([idx[m] for m in markers]),
                        pairwise=np.array([pw[m] for m in markers]))

    # ---- shipped clips (every 2nd frame of each fte*.pickle, first-level content only) ------------
    ship = {}
    files = sorted(glob.glob("data/test_set/**/fte*.pickle", recursive=True))
    names = []
    for f in files:
        d = pickle.load(open(f, "rb"))
        parts = f.split("/")
        animal = [p for p in parts if p in ("jules", "phantom", "arabia", "shiraz")][0]
        key = "/".join(parts[2:-1])
        names.append(key + "|" + animal)
        i = len(names) - 1
        full = "fte_kinematic/" in f + "/" and parts[-2] == "fte_kinematic"
        sl = slice(None) if full else slice(None, None, 4)
        ship[f"{i}_q"] = d["q"][sl]
        ship[f"{i}_positions"] = d["positions"][sl]
        ship[f"{i}_x"] = d["x"][sl]
        if full:
            ship[f"{i}_dq"] = d["dq"]
            ship[f"{i}_ddq"] = d["ddq"]
            ship[f"{i}_obj_cost"] = np.float64(d["obj_cost"])
            ship[f"{i}_time"] = np.float64(d["processing_time_s"])
            ship[f"{i}_meas_err_shape"] = np.array(d["meas_err"].shape)
    ship["names"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "shipped_clips.npz"), **ship)
    print("shipped", len(names))

    # ---- G8: camera + detection recovery for real clips --------------------------------------------
    import pandas as pd
    from scipy.optimize import least_squares
    from oracle import cameras as ocam

    def recover(clip, fname):
        d = pickle.load(open(f"data/test_set/{clip}/fte_kinematic/fte.pickle", "rb"))
        pos = d["positions"]
        N = pos.shape[0]
        csvs = sorted(glob.glob(f"data/test_set/{clip}/fte_kinematic/cam*_fte.csv"))
        cam_list, dets, rms = [], [], []
        good_intr = None
        for ci, path in enumerate(csvs):
            df = pd.read_csv(path, header=[0, 1], index_col=0)
            arr = df.to_numpy().reshape(N, 24, 3)[:, :, :2]
            ok = np.isfinite(arr).all(-1)
            obj = pos[ok]
            img = arr[ok]
            K0 = np.array([[1000.0, 0, 960], [0, 1000, 540], [0, 0, 1]])
            _, rvec, tvec = cv2.solvePnP(obj, img, K0, None, flags=cv2.SOLVEPNP_EPNP)

            def unpack(p):
                R, _ = cv2.Rodrigues(p[8:11])
                return np.concatenate([p[:8], R.reshape(-1), p[11:14]])

            def fun(p):
                return (ocam.project(obj, unpack(p), ocam.FISHEYE) - img).reshape(-1)

            best = None
            for fy_sign in (1.0,):
                p0 = np.concatenate([[1000, 1000 * fy_sign, 960, 540, 0, 0, 0, 0], rvec.reshape(-1), tvec.reshape(-1)])
                res = least_squares(fun, p0, method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=4000)
                if best is None or res.cost < best.cost:
                    best = res
            p = best.x
            if p[1] < 0 or p[0] < 0:
                # mirrored twin: restart from a perturbed pose with positive focal lengths enforced by |.|
                def fun2(p):
                    pp = p.copy()
                    pp[0], pp[1] = abs(pp[0]), abs(pp[1])
                    return (ocam.project(obj, unpack(pp), ocam.FISHEYE) - img).reshape(-1)
                _, rvec2, tvec2 = cv2.solvePnP(obj, img, K0, None, flags=cv2.SOLVEPNP_ITERATIVE)
                p0 = np.concatenate([[860, 860, 960, 540, 0.1, -0.05, 0.01, 0], rvec2.reshape(-1), tvec2.reshape(-1)])
                res = least_squares(fun2, p0, method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=8000)
                p = res.x
                p[0], p[1] = abs(p[0]), abs(p[1])
            cam = unpack(p)
            rr = np.sqrt(np.mean((ocam.project(obj, cam, ocam.FISHEYE) - img) ** 2))
            if rr > 1e-6 and good_intr is not None:
                # local minimum: restart from the intrinsics of an already converged camera of the same rig
                Kg = np.array([[good_intr[0], 0, good_intr[2]], [0, good_intr[1], good_intr[3]], [0, 0, 1]])
                und = cv2.fisheye.undistortPoints(img.reshape(-1, 1, 2), Kg, good_intr[4:8]).reshape(-1, 2)
                _, rv, tv = cv2.solvePnP(obj, und, np.eye(3), None, flags=cv2.SOLVEPNP_ITERATIVE)
                p0 = np.concatenate([good_intr, rv.reshape(-1), tv.reshape(-1)])
                res = least_squares(fun, p0, method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=8000)
                cam2 = unpack(res.x)
                rr2 = np.sqrt(np.mean((ocam.project(obj, cam2, ocam.FISHEYE) - img) ** 2))
                if rr2 < rr:
                    cam, rr = cam2, rr2
            if rr < 1e-8 and good_intr is None:
                good_intr = cam[:8].copy()
            rms.append(rr)
            cam_list.append(cam)
            # detections = projection - meas_err ; use our projection for NaN (out of image) entries
            proj = ocam.project(pos, cam, ocam.FISHEYE)
            det = proj - d["meas_err"][:, ci, :, :, 0]
            dets.append(det)
            print(clip, "cam", ci, "rms", rr, "fx fy", cam[0], cam[1])
        cams = np.array(cam_list)
        det = np.stack(dets, 1)                               # [N,C,24,2]
        det32 = det.astype(np.float32)
        print(clip, "float32 grid error", np.abs(det - det32).max())
        # known answer: reference's own loss on the saved residuals, all weights on
        R_pw, _ = misc.get_uncertainty_models()
        w = 1.0 / R_pw[0]
        known = 0.0
        me = d["meas_err"][..., 0]
        for n_ in range(N):
            for c_ in range(me.shape[1]):
                for l_ in range(24):
                    for d2 in range(2):
                        known += misc.redescending_loss(w[l_] * me[n_, c_, l_, d2], 3, 10, 20)
        fps = 120.0 if "2019" in clip else 90.0
        _, Q = misc.get_uncertainty_models()
        wq = np.where(Q != 0, 1.0 / np.where(Q != 0, Q, 1), 0.0)
        dd = d["ddq"][1:] - d["ddq"][:-1]
        model_cost = float(np.sum(wq * dd * dd))
        np.savez_compressed(os.path.join(OUT, fname), cams=cams, rms=np.array(rms), detections=det32,
                            q=d["q"], dq=d["dq"], ddq=d["ddq"], positions=pos, meas_err=d["meas_err"],
                            obj_cost=np.float64(d["obj_cost"]), known_meas_cost_all_on=np.float64(known),
                            model_cost_from_ddq=np.float64(model_cost), fps=np.float64(fps),
                            processing_time_s=np.float64(d["processing_time_s"]), start_frame=np.int64(d["start_frame"]))
        print(clip, "known cost", known, "model", model_cost, "obj", d["obj_cost"])

    recover("2017_08_29/top/jules/run1_1", "clip_jules_run1_1.npz")
    recover("2019_03_07/phantom/run", "clip_phantom_run.npz")

    # ---- small synthetic sequence: cost by the reference's own functions (pure python loops) -------
    robot, sp_vars, pfs = robots["jules"]
    lengths = fk["jules_lengths"]
    N, C = 5, 2
    qs = np.zeros((N, 54))
    qs[:, 0] = np.linspace(0, 0.4, N) + 3.0
    qs[:, 2] = 0.7
    qs[:, 3:] = rng.normal(0, 0.25, (N, 51))
    pos = eval_positions(pfs, qs, lengths)
    g8 = np.load(os.path.join(OUT, "clip_jules_run1_1.npz"))
    cams = g8["cams"][:C]
    meas = np.zeros((N, C, 24, 2))
    lik = rng.beta(8, 2, (N, C, 24))
    resid = np.zeros((N, C, 24, 2))
    total = 0.0
    R_pw, Q = misc.get_uncertainty_models()
    for n_ in range(N):
        for c_ in range(C):
            cam = cams[c_]
            K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]])
            D = cam[4:8]
            R = cam[8:17].reshape(3, 3)
            t = cam[17:20].reshape(3, 1)
            for l_ in range(24):
                # place the point in front of the first cams by shifting to the real clip's region
                x, y, z = pos[n_, l_] + g8["positions"][0, 4] - np.array([3.0, 0, 0.7])
                u, v = misc.pt3d_to_2d_fisheye(x, y, z, K, D, R, t)
                noise = rng.normal(0, 4.0, 2) + (rng.uniform(-60, 60, 2) if rng.uniform() < 0.15 else 0)
                meas[n_, c_, l_] = np.float32([u, v] + noise)
                wgt = 1 / R_pw[0][l_] if lik[n_, c_, l_] > 0.5 else 0.0
                for d2, val in enumerate((u, v)):
                    resid[n_, c_, l_, d2] = val - meas[n_, c_, l_, d2]
                    total += misc.redescending_loss(wgt * resid[n_, c_, l_, d2], 3, 10, 20)
    shift = g8["positions"][0, 4] - np.array([3.0, 0, 0.7])
    qs2 = qs.copy()
    qs2[:, 0:3] += shift
    np.savez_compressed(os.path.join(OUT, "cost_ref.npz"), q=qs2, cams=cams, meas=meas, likelihood=lik, resid=resid,
                        meas_cost=np.float64(total), lengths=lengths, radii=fk["jules_radii"])
    print("cost_ref", total)


if __name__ == "__main__":
    main()
