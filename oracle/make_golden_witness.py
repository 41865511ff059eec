"""Generates tests/golden/nlp_witness.npz: converged solutions of the reference's constrained NLP from an INDEPENDENT
solver (oracle/nlp.py: scipy SLSQP on the objective with the 26 joint equalities and the inequality rows per frame as
constraints), for small problems the CUDA LM / augmented-Lagrangian path is then compared with (tests/
test_nlp_witness.py).  TEST INFRASTRUCTURE; runs anywhere (needs no /root/reference: the real-clip problems come from
the committed fixture tests/golden/test_set_clips.npz), but takes 10-20 CPU-minutes, so its output is committed.

    python -m oracle.make_golden_witness

Problems:
    syn_c6   synthetic, 8 frames x 6 cameras (seed 7)          syn_c4   synthetic, 10 frames x 4 cameras (seed 11)
    syn_c1   synthetic, 8 frames x 1 camera (seed 13; monocular without priors is ill-conditioned: cost only)
    jules    first 12 frames of 2017_08_29/top/jules/run1_1, 6 cameras, all-on likelihood mask
    phantom  first 12 frames of 2019_03_07/phantom/run, 6 cameras
Start point of every problem = the initial guess the product would use (synthetic.q_init / acinoset_opt.initial_guess).
"""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def synthetic_problem(N, C, seed):
    from cheetah_pose_estimation_b200 import synthetic
    from oracle import adapters
    prob = synthetic.make_problem(S=1, N=N, C=C, seed=seed, projector=adapters.projector)
    return dict(meas=prob.meas[0], cams=prob.cams[0], chain_vec=prob.chain_vec[0], h=float(prob.h[0]),
                q0=prob.q_init[0], q_true=prob.q_true[0])


def clip_problem(name, n_frames):
    """First `n_frames` frames of a rebuilt real clip; q0 from the mirror's own initial guess (CPU part only)."""
    from cheetah_pose_estimation_b200 import acinoset_misc as misc, skeleton as sk
    from tests._dataset import build_clip_dir, load_test_set_clips
    c = next(c for c in load_test_set_clips(os.path.join(ROOT, "tests", "golden")) if c["name"] == name)
    root = os.path.join(tempfile.mkdtemp(), "cheetah_videos")
    d = build_clip_dir(root, name, c)
    md = json.load(open(os.path.join(d, "metadata.json")))
    k_arr, d_arr, r_arr, t_arr, cam_res, n_cams, scene_fpath = misc.find_scene_file(d, verbose=False)
    fps = float(c["fps"])
    params = misc.TrajectoryParams(d, md["start_frame"], md["end_frame"], md["end_frame"] - md["start_frame"], 0.5,
                                   md["cam_sync"], False, False, False, False)
    scene = misc.Scene(scene_fpath, k_arr, d_arr.reshape(-1, 4), r_arr, t_arr, cam_res, fps, n_cams, None)
    base = misc.load_dlc_arrays(os.path.join(d, "dlc"), n_cams)
    lengths, radii, _ = sk.animal(c["animal"])
    x, y, z, psi = misc.create_trajectory_estimate(lengths[sk.BASE], params, scene, False, base)
    N = c["q"].shape[0]
    fr = params.start_frame + np.arange(N)
    q0 = np.zeros((N, 54))
    q0[:, 5::3] = psi[fr][:, None]
    q0[:, 0], q0[:, 1], q0[:, 2] = x[fr], y[fr], z[fr]
    meas, _ = misc.pack_measurements(params, scene, base)
    cams = np.stack([sk.pack_camera(k_arr[i], d_arr.reshape(-1, 4)[i], r_arr[i], t_arr[i], 1.0) for i in range(n_cams)])
    _, cvec = sk.marker_chains(lengths, radii)
    return dict(meas=meas[:n_frames], cams=cams, chain_vec=cvec, h=1.0 / fps, q0=q0[:n_frames],
                q_true=c["q"][:n_frames])          # "truth" = the shipped IPOPT trajectory (of the lost-mask objective)


def main():
    from cheetah_pose_estimation_b200 import skeleton as sk
    from oracle import adapters, nlp
    lengths, radii, _ = sk.animal("jules")
    chain_link, _ = sk.marker_chains(lengths, radii)
    problems = {"syn_c6": synthetic_problem(8, 6, 7), "syn_c4": synthetic_problem(10, 4, 11),
                "syn_c1": synthetic_problem(8, 1, 13),
                "jules": clip_problem("2017_08_29/top/jules/run1_1", 12),
                "phantom": clip_problem("2019_03_07/phantom/run", 12)}
    out = {"names": np.array(list(problems))}
    for name, pr in problems.items():
        pb = adapters.problem_from_arrays(pr["meas"], pr["cams"], chain_link, pr["chain_vec"], pr["h"])
        res = nlp.solve(pb, pr["q0"], maxiter=1500)
        print(name, {k: v for k, v in res.items() if k != "q"}, flush=True)
        for k, v in pr.items():
            out[f"{name}_{k}"] = np.asarray(v)
        out[f"{name}_q"] = res["q"]
        for k in ("obj", "c_eq_max", "ineq_violation", "nit", "success", "seconds"):
            out[f"{name}_{k}"] = np.asarray(res[k])
        out[f"{name}_cost"] = res["cost"]
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", "nlp_witness.npz"), **out)


if __name__ == "__main__":
    main()
