"""Host side of the kinetic path's first slice: rows 0-5 of the equations of motion on the GPU.

Reference: CheetahEstimator.calc_grf_eom / estimate_grf (acinoset_opt.py:120-270): `eom[i](q, dq, ddq, GRFz, GRFxy)`.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, skeleton

FEET = ("HFL_foot", "HFR_foot", "HBL_foot", "HBR_foot")        # order of pe.foot.feet(robot) (link order)
N_FORCES, N_PAR = 20, 52


def link_params(animal):
    """par[52] = lengths[17], masses[17], radii[17] in link order, F_scale = total_mass * 9.81
    (acinoset_opt.py:151-161)."""
    lengths, radii, masses = skeleton.animal(animal)
    return np.concatenate([lengths, masses, radii, [float(np.sum(masses)) * 9.81]])


def pack_grf(grfz, grfxy):
    """grf[..., 20] from GRFz[..., 4] and GRFxy[..., 4, 4] (foot-major), the order of eom_vars
    (acinoset_opt.py:170-171: all Lz first, then the flattened Lx)."""
    grfz, grfxy = np.asarray(grfz, np.float64), np.asarray(grfxy, np.float64)
    return np.concatenate([grfz, grfxy.reshape(grfxy.shape[:-2] + (16,))], axis=-1)


def eom_rows(q, dq, ddq, grf, par, n_frames=None, device="cuda"):
    """eom[S,N,6] on the GPU.  q, dq, ddq [S,N,54], grf[S,N,20], par[S,52] (numpy or torch)."""
    if not torch.cuda.is_available():
        raise _lib.CpeError("CUDA device required: the hot path has no CPU fallback")
    lib = _lib.load()
    dev = torch.device(device)
    t = [torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a, dtype=torch.float64,
                         device=dev).contiguous() for a in (q, dq, ddq, grf, par)]
    S, N = t[0].shape[:2]
    assert t[0].shape == (S, N, 54) and t[1].shape == (S, N, 54) and t[2].shape == (S, N, 54)
    assert t[3].shape == (S, N, N_FORCES) and t[4].shape == (S, N_PAR)
    nf = torch.full((S,), N, dtype=torch.int32, device=dev) if n_frames is None else \
        torch.as_tensor(n_frames, dtype=torch.int32, device=dev)
    b = _lib.CpeBatch()
    b.S, b.N, b.C, b.W = S, N, 0, 1
    b.n_frames = nf.data_ptr()
    out = torch.zeros(S, N, 6, dtype=torch.float64, device=dev)
    _lib.check(lib.cpe_eom_f64(C.byref(b), *[C.c_void_p(x.data_ptr()) for x in t], C.c_void_p(out.data_ptr()),
                               C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cpe_eom_f64")
    return out


def eom_vjp(q, dq, ddq, grf, par, w=None, n_frames=None, device="cuda"):
    """Vector-Jacobian product of the EOM rows with respect to the states (`cpe_eom_vjp_f64`): returns
    (eom[S,N,6], g_q, g_dq, g_ddq [S,N,54]) with g = sum_i w[..., i] d row_i / d (q, dq, ddq).  w = None takes
    w = 2 * rows: the gradient of misc.eom_slack_cost (acinoset_misc.py:631-636) -- in the reference's kinetic NLP q, dq
    and ddq are separate variables tied by the collocation constraints, so these three partials are what ASL evaluates
    for that term (acinoset_opt.py:905-921).  `states_gradient_to_q` folds them onto q alone."""
    if not torch.cuda.is_available():
        raise _lib.CpeError("CUDA device required: the hot path has no CPU fallback")
    lib = _lib.load()
    dev = torch.device(device)
    t = [_t(a, dev) for a in (q, dq, ddq, grf, par)]
    S, N = t[0].shape[:2]
    assert all(x.shape == (S, N, 54) for x in t[:3]) and t[3].shape == (S, N, N_FORCES) and t[4].shape == (S, N_PAR)
    wt = None if w is None else _t(w, dev)
    assert wt is None or wt.shape == (S, N, 6)
    b, keep = _dev_batch(S, N, 1.0, n_frames, dev)
    f64 = dict(dtype=torch.float64, device=dev)
    eom = torch.zeros(S, N, 6, **f64)
    g = [torch.zeros(S, N, 54, **f64) for _ in range(3)]
    p = lambda x: None if x is None else C.c_void_p(x.data_ptr())
    _lib.check(lib.cpe_eom_vjp_f64(C.byref(b), *[p(x) for x in t], p(wt), p(eom), p(g[0]), p(g[1]), p(g[2]),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cpe_eom_vjp_f64")
    return eom, g[0], g[1], g[2]


def rates_from_q(q, h):
    """dq, ddq of the mirrored API from q[..., N, 54] (acinoset_opt._finish_estimate: implicit Euler, collocation.py:58-68,
    free boundary values ddq_0 = ddq_1 = ddq_2 and dq_0 = dq_1 - h ddq_1).  torch, any device."""
    dq = torch.zeros_like(q)
    ddq = torch.zeros_like(q)
    dq[..., 1:, :] = (q[..., 1:, :] - q[..., :-1, :]) / h
    ddq[..., 2:, :] = (dq[..., 2:, :] - dq[..., 1:-1, :]) / h
    if q.shape[-2] > 2:
        ddq[..., 1, :] = ddq[..., 2, :]
        ddq[..., 0, :] = ddq[..., 2, :]
        dq[..., 0, :] = dq[..., 1, :] - h * ddq[..., 1, :]
    return dq, ddq


def states_gradient_to_q(g_q, g_dq, g_ddq, h):
    """Adjoint of `rates_from_q`: d cost / d q [..., N, 54] of a cost whose partials with respect to q, dq, ddq are given
    and whose dq, ddq are the mirrored API's functions of q.  N > 2."""
    N = g_q.shape[-2]
    assert N > 2
    g_dq, g_ddq = g_dq.clone(), g_ddq.clone()
    out = g_q.clone()
    # dq_0 = dq_1 - h ddq_1 ; ddq_0 = ddq_1 = ddq_2
    g_dq[..., 1, :] += g_dq[..., 0, :]
    g_ddq[..., 1, :] -= h * g_dq[..., 0, :]
    g_dq[..., 0, :] = 0.0
    g_ddq[..., 2, :] += g_ddq[..., 0, :] + g_ddq[..., 1, :]
    g_ddq[..., 0:2, :] = 0.0
    # ddq_n = (dq_n - dq_{n-1}) / h, n >= 2
    g_dq[..., 2:, :] += g_ddq[..., 2:, :] / h
    g_dq[..., 1:-1, :] -= g_ddq[..., 2:, :] / h
    # dq_n = (q_n - q_{n-1}) / h, n >= 1
    out[..., 1:, :] += g_dq[..., 1:, :] / h
    out[..., :-1, :] -= g_dq[..., 1:, :] / h
    return out


def estimate_grf(q, dq, ddq, par, contact, n_frames=None, iters=600, device="cuda"):
    """Per-frame GRF estimate of CheetahEstimator.estimate_grf (acinoset_opt.py:176-249) for S sequences at once.

    q, dq, ddq [S,N,54]; par[S,52]; contact[S,N,4] (non-zero = foot on the ground, feet in `FEET` order).
    Returns (grf[S,N,20] in units of F_scale, resid[S,N] = the minimal ||eom||).  The reference runs one IPOPT solve per
    frame (:204-249); here `cpe_eom_affine_f64` splits the rows into a - Bm F and `cpe_grf_solve_f64` solves every
    frame's bounded least-squares problem in one launch."""
    if not torch.cuda.is_available():
        raise _lib.CpeError("CUDA device required: the hot path has no CPU fallback")
    lib = _lib.load()
    dev = torch.device(device)
    t = [torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a, dtype=torch.float64,
                         device=dev).contiguous() for a in (q, dq, ddq, par)]
    S, N = t[0].shape[:2]
    assert all(x.shape == (S, N, 54) for x in t[:3]) and t[3].shape == (S, N_PAR)
    ct = torch.as_tensor(np.asarray(contact) != 0 if not torch.is_tensor(contact) else contact != 0, device=dev)
    ct = ct.to(torch.int32).contiguous()
    assert ct.shape == (S, N, 4)
    nf = torch.full((S,), N, dtype=torch.int32, device=dev) if n_frames is None else \
        torch.as_tensor(n_frames, dtype=torch.int32, device=dev)
    b = _lib.CpeBatch()
    b.S, b.N, b.C, b.W = S, N, 0, 1
    b.n_frames = nf.data_ptr()
    a = torch.zeros(S, N, 6, dtype=torch.float64, device=dev)
    Bm = torch.zeros(S, N, 6, N_FORCES, dtype=torch.float64, device=dev)
    grf = torch.zeros(S, N, N_FORCES, dtype=torch.float64, device=dev)
    resid = torch.zeros(S, N, dtype=torch.float64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda x: C.c_void_p(x.data_ptr())
    _lib.check(lib.cpe_eom_affine_f64(C.byref(b), p(t[0]), p(t[1]), p(t[2]), p(t[3]), p(a), p(Bm), st),
               "cpe_eom_affine_f64")
    _lib.check(lib.cpe_grf_solve_f64(C.byref(b), p(a), p(Bm), p(ct), int(iters), p(grf), p(resid), st),
               "cpe_grf_solve_f64")
    return grf, resid


def contact_mask(metadata):
    """contact[N,4] from an `autogen-contact.json` dict: N = end_frame - start_frame, foot f is on the ground at row fe
    when start_frame + fe lies in one of its [first, last) ranges (acinoset_opt.py:203-229; entries of a foot may be
    null, extra fields after the two frame numbers are ignored as the reference does)."""
    start, end = int(metadata["start_frame"]), int(metadata["end_frame"])
    order = metadata["contacts"]
    out = np.zeros((max(end - start, 0), 4), dtype=np.int32)
    for f, name in enumerate(FEET):
        for seq in (order.get(name) or []):
            lo, hi = max(int(seq[0]) - start, 0), min(int(seq[1]) - start, out.shape[0])
            if hi > lo:
                out[lo:hi, f] = 1
    return out


def estimate_grf_states(fte_states, metadata, animal, iters=600, device="cuda"):
    """CheetahEstimator.estimate_grf (acinoset_opt.py:176-268) for one trajectory: `fte_states` is the dict of a saved
    fte.pickle (q, dq, ddq [>=N,54]), `metadata` the contact JSON.  Returns (grfz_est, grfxy_est): per foot name a list
    over the N frames of GRFz and of the 4 GRFxy values -- the reference's return value."""
    ct = contact_mask(metadata)
    N = ct.shape[0]
    q, dq, ddq = (np.asarray(fte_states[k], np.float64)[None, :N] for k in ("q", "dq", "ddq"))
    grf, _ = estimate_grf(q, dq, ddq, link_params(animal)[None], ct[None], iters=iters, device=device)
    F = grf[0].cpu().numpy()
    grfz_est = {name: [float(v) for v in F[:, f]] for f, name in enumerate(FEET)}
    grfxy_est = {name: [[float(v) for v in row] for row in F[:, 4 + 4 * f:8 + 4 * f]] for f, name in enumerate(FEET)}
    return grfz_est, grfxy_est


# ---- trajectory-level cost terms of the kinetic objective (acinoset_opt.py:905-921), cost + gradient on the GPU ------------
def _dev_batch(S, N, h, n_frames, dev, chain_vec=None):
    nf = torch.full((S,), N, dtype=torch.int32, device=dev) if n_frames is None else \
        torch.as_tensor(n_frames, dtype=torch.int32, device=dev)
    ht = torch.as_tensor(np.broadcast_to(np.asarray(h, np.float64), (S,)).copy(), device=dev)
    b = _lib.CpeBatch()
    b.S, b.N, b.C, b.W = S, N, 0, 1
    b.n_frames, b.h = nf.data_ptr(), ht.data_ptr()
    if chain_vec is not None:
        b.chain_vec = chain_vec.data_ptr()
    return b, (nf, ht)


def _t(x, dev):
    return torch.as_tensor(np.ascontiguousarray(x) if isinstance(x, np.ndarray) else x, dtype=torch.float64,
                           device=dev).contiguous()


def motion_smoothing(solver, q):
    """misc.motion_smoothing_cost (acinoset_misc.py:601-613) of q[S,N,54] on the batch of `solver` (an engine.Solver):
    returns (cost[S], grad[S,N,54]) as device tensors; fps = 1 / batch.h."""
    lib, b = solver.lib, solver.b
    pose = solver.evaluate(q, pose=True)["pose"]
    S, N = q.shape[:2]
    work = torch.zeros(S, N, dtype=torch.float64, device=q.device)
    cost = torch.zeros(S, dtype=torch.float64, device=q.device)
    grad = torch.zeros(S, N, 54, dtype=torch.float64, device=q.device)
    bs = b.struct()
    p = lambda x: C.c_void_p(x.data_ptr())
    _lib.check(lib.cpe_motion_smoothing_f64(C.byref(solver.m.struct), C.byref(bs), p(q), p(pose), p(work), p(cost), p(grad),
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream)),
               "cpe_motion_smoothing_f64")
    return cost, grad


def eom_slack(eom, Bm=None, n_frames=None, device="cuda"):
    """misc.eom_slack_cost (acinoset_misc.py:631-636) on eom[S,N,6] (rows 0-5, `eom_rows`); with Bm[S,N,6,20] also the
    gradient with respect to the forces.  Returns (cost[S], grad_F[S,N,20] or None)."""
    lib = _lib.load()
    dev = torch.device(device)
    e = _t(eom, dev)
    S, N = e.shape[:2]
    B = None if Bm is None else _t(Bm, dev)
    b, keep = _dev_batch(S, N, 1.0, n_frames, dev)
    work = torch.zeros(S, N, dtype=torch.float64, device=dev)
    cost = torch.zeros(S, dtype=torch.float64, device=dev)
    g = None if B is None else torch.zeros(S, N, N_FORCES, dtype=torch.float64, device=dev)
    p = lambda x: None if x is None else C.c_void_p(x.data_ptr())
    _lib.check(lib.cpe_eom_slack_f64(C.byref(b), p(e), p(B), p(work), p(cost), p(g),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cpe_eom_slack_f64")
    return cost, g


def torque_costs(Tc, weight, h, n_frames=None, device="cuda"):
    """Torque terms on Tc[S,N,M] (all motors' axes flattened), weight[M], hm0 = h[S]: returns (cost[S,2], grad_sq, grad_ch)
    with cost[:,0] = pe.motor.torque_squared_penalty (motor.py:412-417, 496-503) and cost[:,1] =
    misc.change_in_torque_squared_cost (acinoset_misc.py:616-628)."""
    lib = _lib.load()
    dev = torch.device(device)
    T = _t(Tc, dev)
    S, N, M = T.shape
    w = _t(np.broadcast_to(np.asarray(weight, np.float64), (M,)).copy(), dev)
    b, keep = _dev_batch(S, N, h, n_frames, dev)
    work = torch.zeros(2, S, N, dtype=torch.float64, device=dev)
    cost = torch.zeros(S, 2, dtype=torch.float64, device=dev)
    g1, g2 = torch.zeros_like(T), torch.zeros_like(T)
    p = lambda x: C.c_void_p(x.data_ptr())
    _lib.check(lib.cpe_torque_costs_f64(C.byref(b), p(T), C.c_int32(M), p(w), p(work), p(cost), p(g1), p(g2),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cpe_torque_costs_f64")
    return cost, g1, g2


def kinetic_objective(kinematic_cost, pose_cost, torque_cost, smoothing_cost, slack_cost, fps, enable_eom_slack=True,
                      disable_motion_prior=False):
    """The reference's kinetic objective from its terms (acinoset_opt.py:905-921, use_2d_reprojections branch):
    1e-3 (kinematic + pose + (torque + 0.1 fps^-2 smoothing) + 10e3 slack)."""
    motion_prior = 0.0 if disable_motion_prior else torque_cost + 0.1 * fps ** -2 * smoothing_cost
    return 1e-3 * (kinematic_cost + pose_cost + motion_prior + 10e3 * (slack_cost if enable_eom_slack else 0.0))


# ---- contact detection (CPU, as in the reference: acinoset_opt.py:638-690 -> acinoset_misc.py:745-856) --------------------
PAW_MARKERS = {"HFL_foot": "l_front_paw", "HFR_foot": "r_front_paw", "HBL_foot": "l_back_paw", "HBR_foot": "r_back_paw"}


def foot_height_and_velocity(q, dq, lengths, radii):
    """Foot positions and velocities of the four feet, [N,4,3] each, in FEET order: the foot point of the reference's
    `Foot3D` is the bottom of the H link = the paw marker (cheetah.py:96-101, acinoset_misc.py:1640-1659), and its velocity
    `Pb_I_vel = J dq` (foot.py:394-395) is the time derivative of that position along the saved dq."""
    q, dq = np.asarray(q, np.float64), np.asarray(dq, np.float64)
    chain_link, chain_vec = skeleton.marker_chains(lengths, radii)
    idx = [skeleton.MARKERS.index(PAW_MARKERS[f]) for f in FEET]
    ang, dang = q[:, 3:].reshape(-1, 17, 3), dq[:, 3:].reshape(-1, 17, 3)
    sf, cf, st, ct, sp, cp = (np.sin(ang[..., 0]), np.cos(ang[..., 0]), np.sin(ang[..., 1]), np.cos(ang[..., 1]),
                              np.sin(ang[..., 2]), np.cos(ang[..., 2]))
    z = np.zeros_like(sf)
    R = np.stack([np.stack([cp * ct, cp * st * sf - sp * cf, cp * st * cf + sp * sf], -1),
                  np.stack([sp * ct, sp * st * sf + cp * cf, sp * st * cf - cp * sf], -1),
                  np.stack([-st, ct * sf, ct * cf], -1)], -2)                               # [N,17,3,3]
    dRf = np.stack([np.stack([z, cp * st * cf + sp * sf, -cp * st * sf + sp * cf], -1),
                    np.stack([z, sp * st * cf - cp * sf, -sp * st * sf - cp * cf], -1),
                    np.stack([z, ct * cf, -ct * sf], -1)], -2)
    dRt = np.stack([np.stack([-cp * st, cp * ct * sf, cp * ct * cf], -1),
                    np.stack([-sp * st, sp * ct * sf, sp * ct * cf], -1),
                    np.stack([-ct, -st * sf, -st * cf], -1)], -2)
    dRp = np.stack([np.stack([-sp * ct, -sp * st * sf - cp * cf, -sp * st * cf + cp * sf], -1),
                    np.stack([cp * ct, cp * st * sf - sp * cf, cp * st * cf + sp * sf], -1),
                    np.stack([z, z, z], -1)], -2)
    Rdot = dRf * dang[..., 0, None, None] + dRt * dang[..., 1, None, None] + dRp * dang[..., 2, None, None]
    pos = np.zeros((q.shape[0], 4, 3))
    vel = np.zeros((q.shape[0], 4, 3))
    for f, l in enumerate(idx):
        pos[:, f], vel[:, f] = q[:, 0:3], dq[:, 0:3]
        for j in range(skeleton.MAX_CHAIN):
            link = int(chain_link[l, j])
            if link >= 0:
                pos[:, f] += R[:, link] @ chain_vec[l, j]
                vel[:, f] += Rdot[:, link] @ chain_vec[l, j]
    return pos, vel


def _positive_zero_crossings(x):
    """acinoset_misc.py:2033-2046 (indices refer to the array with exact zeros removed, as in the reference)."""
    x = x[np.nonzero(x)]
    out = []
    for i in range(1, len(x)):
        if x[i - 1] < 0 and x[i] > 0:
            out += [i + 2, i + 1, i, i - 1, i - 2]
    return out


def _group_by_consecutive_values(x):
    spl = [0] + [i for i in range(1, len(x)) if x[i] - x[i - 1] > 1] + [None]
    return [x[b:e] for (b, e) in [(spl[i - 1], spl[i]) for i in range(1, len(spl))]]


def contact_detection(foot_height, foot_vel_z, start_frame, speed, fps, ground_plane_height=0.0, data_dir=None):
    """misc.contact_detection (acinoset_misc.py:745-856) on foot_height[N,4] and the vertical foot velocity [N,4] (FEET
    order): stance length from the speed (linear model through (9 m/s, 0.09 s), (14 m/s, 0.06 s)), a 5 cm height
    threshold above the ground plane, and a positive zero crossing of the vertical velocity near the height minimum.
    Returns (contacts, contacts_tmp) and, with `data_dir`, writes `grf/autogen-contact.json` / `autogen-contact-02.json`."""
    import json
    import os
    n = foot_height.shape[0]
    m_, c_ = np.linalg.lstsq(np.vstack([[9.0, 14.0], [1.0, 1.0]]).T, [0.09, 0.06], rcond=None)[0]
    stance_time_fe = round((m_ * speed + c_) * fps)
    mid_way = stance_time_fe // 2
    is_even = (stance_time_fe % 2) == 0
    contacts, contacts_tmp = {}, {}
    for i, name in enumerate(FEET):
        h = foot_height[:, i]
        under = _group_by_consecutive_values(np.where(h < ground_plane_height + 0.05)[0])
        crossings = _positive_zero_crossings(foot_vel_z[:, i])
        contacts[name], contacts_tmp[name] = [], []
        arg_min = -1
        for j, grp in enumerate(under):
            if len(grp) == 0:
                continue
            lo = int(arg_min + 1)
            hi = under[j + 1][0] if j + 1 < len(under) else -1
            arg_min = lo + int(np.argmin(h[lo:hi]))
            possible = np.intersect1d(grp, crossings)
            if all(arg_min + k not in possible for k in (-2, -1, 0, 1, 2)):
                arg_min = grp[-1]
                continue
            start_idx, end_idx = int(arg_min - mid_way), int(arg_min + mid_way)
            arg_min = grp[-1]
            if is_even:
                start_idx += 1
            if start_idx < 0:
                end_idx -= start_idx
                start_idx = 0
            if end_idx >= n:
                start_idx -= end_idx - n - 1
                end_idx = n - 1
            contacts[name].append([int(start_frame + start_idx), int(start_frame + end_idx), i, "TBD"])
            contacts_tmp[name].append([int(start_frame + grp[0]), int(start_frame + grp[-1]), i, "TBD"])
        if not contacts[name]:
            contacts[name] = contacts_tmp[name] = None
    for a, b in (("HFL_foot", "HFR_foot"), ("HBL_foot", "HBR_foot")):          # rotary gallop: first contact = trailing limb
        if contacts[a] is not None and contacts[b] is not None:
            first_a = contacts[a][0][0] > contacts[b][0][0]
            contacts[a][0][3] = "leading" if first_a else "trailing"
            contacts[b][0][3] = "trailing" if first_a else "leading"
    if data_dir is not None:
        grf_dir = os.path.join(data_dir, "grf")
        os.makedirs(grf_dir, exist_ok=True)
        for fname, c in (("autogen-contact.json", contacts), ("autogen-contact-02.json", contacts_tmp)):
            with open(os.path.join(grf_dir, fname), "w", encoding="utf-8") as f:
                json.dump({"start_frame": int(start_frame), "end_frame": int(start_frame + n), "contacts": c}, f)
    return contacts, contacts_tmp


def determine_contacts_from_states(fte_states, animal, start_frame, fps, ground_plane_height=0.0, data_dir=None):
    """opt.determine_contacts (acinoset_opt.py:638-690) from the dict of a saved fte.pickle: foot heights / velocities from
    q, dq, speed = mean norm of `com_vel`."""
    lengths, radii, _ = skeleton.animal(animal)
    pos, vel = foot_height_and_velocity(fte_states["q"], fte_states["dq"], lengths, radii)
    speed = float(np.mean(np.linalg.norm(np.asarray(fte_states["com_vel"]), axis=1)))
    return contact_detection(pos[:, :, 2], vel[:, :, 2], start_frame, speed, fps, ground_plane_height, data_dir)


# ---- contact prescription (acinoset_opt.py:783-812, acinoset_misc.py:1140-1167) as bound arrays ---------------------------
def contact_times(contact_json, n_frames=None):
    """Per foot (FEET order) the sorted 1-based frame indices `fe` at which it is prescribed to be on the ground, from an
    `autogen-contact.json` / metadata dict: every [first, last] range INCLUSIVE of both ends, shifted by the file's
    start_frame, plus one for the Pyomo indexing (acinoset_opt.py:795-803; `estimate_grf` reads the same ranges half-open,
    see contact_mask).  Feet whose entry is null get an empty list.  `n_frames` clips to the trajectory."""
    start = int(contact_json["start_frame"])
    out = []
    for name in FEET:
        times = []
        for seq in (contact_json["contacts"].get(name) or []):
            times += list(range(int(seq[0]) - start + 1, int(seq[1]) - start + 2))
        if n_frames is not None:
            times = [t for t in times if 1 <= t <= n_frames]
        out.append(times)
    return out


def prescribe_contact_order(ground_timings, n_frames, foot_height_uncertainty=0.05, min_GRFz=0.01):
    """misc.prescribe_contact_order (acinoset_misc.py:1140-1167) without the Pyomo model: what it fixes and bounds, as
    arrays [n_frames, 4] in FEET order (row fe - 1 for the 1-based index fe of `ground_timings`, one list per foot):
      stance           bool: the foot is prescribed on the ground
      grfz_lb/grfz_ub  flight: both 0 (GRFz and GRFxy are fixed to 0); stance: [min_GRFz, inf)
      grfxy_ub         flight: 0; stance: inf
      height_lb/ub     stance: -/+ foot_height_uncertainty (None: unbounded), flight: unbounded
    `contact_violations` evaluates a trajectory against them."""
    stance = np.zeros((n_frames, 4), dtype=bool)
    for f, times in enumerate(ground_timings):
        for fe in times:
            assert 1 <= fe <= n_frames, (fe, n_frames)
            stance[fe - 1, f] = True
    inf = np.inf
    hu = inf if foot_height_uncertainty is None else float(foot_height_uncertainty)
    return dict(stance=stance, grfz_lb=np.where(stance, float(min_GRFz), 0.0), grfz_ub=np.where(stance, inf, 0.0),
                grfxy_ub=np.where(stance, inf, 0.0), height_lb=np.where(stance, -hu, -inf), height_ub=np.where(stance, hu, inf))


def contact_violations(q, dq, lengths, radii, bounds, grf=None, ground_plane_height=0.0):
    """How far a trajectory is from a contact prescription: per foot the largest distance of the foot height (above the
    ground plane) outside [height_lb, height_ub] over the prescribed stance frames, the largest horizontal foot speed
    there (the reference's no-slip rows, acinoset_opt.py:805-806), and -- with grf[N,20] (`pack_grf` layout) -- the
    largest force applied in flight and the largest shortfall below min_GRFz in stance."""
    pos, vel = foot_height_and_velocity(q, dq, lengths, radii)
    n = bounds["stance"].shape[0]
    hgt = pos[:n, :, 2] - ground_plane_height
    st = bounds["stance"]
    out_h = np.maximum(np.maximum(bounds["height_lb"] - hgt, hgt - bounds["height_ub"]), 0.0)
    slip = np.where(st, np.linalg.norm(vel[:n, :, 0:2], axis=-1), 0.0)
    res = dict(height=out_h.max(axis=0), slip_speed=slip.max(axis=0))
    if grf is not None:
        g = np.asarray(grf, np.float64)[:n]
        fz, fxy = g[:, 0:4], g[:, 4:20].reshape(n, 4, 4)
        res["flight_force"] = np.where(st, 0.0, np.abs(fz) + np.abs(fxy).sum(-1)).max(axis=0)
        res["stance_shortfall"] = np.where(st, np.maximum(bounds["grfz_lb"] - fz, 0.0), 0.0).max(axis=0)
    return res
