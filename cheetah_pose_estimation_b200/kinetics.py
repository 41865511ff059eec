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
