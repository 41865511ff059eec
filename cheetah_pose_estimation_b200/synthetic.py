"""Seeded synthetic sequences of the named shape: trajectories that satisfy the joint
equalities exactly, a camera ring, and DLC-like detections (projection + noise, float32, likelihood gate).

Host-side data generation only.  Detections are produced by projecting the ground truth with the CUDA kernel
(`projector=None`) or with any callable `(q[S,N,54], cams[S,C,24], chain_vec[S,24,5,3]) -> uv[S,N,C,24,2]`
(CPU tests pass the oracle here).
"""
from dataclasses import dataclass

import numpy as np

from . import skeleton as sk


def _rz(a):
    c, s = np.cos(a), np.sin(a)
    z, o = np.zeros_like(a), np.ones_like(a)
    return np.stack([np.stack([c, -s, z], -1), np.stack([s, c, z], -1), np.stack([z, z, o], -1)], -2)


def _ry(a):
    c, s = np.cos(a), np.sin(a)
    z, o = np.zeros_like(a), np.ones_like(a)
    return np.stack([np.stack([c, z, s], -1), np.stack([z, o, z], -1), np.stack([-s, z, c], -1)], -2)


def _rx(a):
    c, s = np.cos(a), np.sin(a)
    z, o = np.zeros_like(a), np.ones_like(a)
    return np.stack([np.stack([o, z, z], -1), np.stack([z, c, -s], -1), np.stack([z, s, c], -1)], -2)


def euler_to_rot(phi, theta, psi):
    """Rb_I = Rz(psi) Ry(theta) Rx(phi)  (= euler_321(phi,theta,psi).T of the reference, links.py:483)."""
    return _rz(psi) @ _ry(theta) @ _rx(phi)


def rot_to_euler(R):
    """Inverse of euler_to_rot on the |theta| < pi/2 branch."""
    theta = -np.arcsin(np.clip(R[..., 2, 0], -1.0, 1.0))
    phi = np.arctan2(R[..., 2, 1], R[..., 2, 2])
    psi = np.arctan2(R[..., 1, 0], R[..., 0, 0])
    return phi, theta, psi


def make_trajectories(S, N, fps, rng, max_travel=16.0):
    """q_true[S,N,54]: galloping-like motion; children composed as R_parent Ry(d) (revolute) and
    R_parent Ry(b) Rz(a) (tail hooke) so that the 26 joint equalities hold exactly."""
    t = np.arange(N)[None, :] / fps
    T = N / fps

    def U(lo, hi):
        return rng.uniform(lo, hi, (S, 1))

    v = np.minimum(U(8.0, 14.0), max_travel / T)
    f = U(2.5, 3.5)
    w = 2 * np.pi * f
    ph = U(0, 2 * np.pi)
    heading = np.pi + U(-0.2, 0.2)
    x0 = -0.5 * v * T
    q = np.zeros((S, N, 54))
    s_along = x0 + v * t
    lateral = rng.normal(0, 0.1, (S, 1)) + 0.02 * np.sin(w * t / 2 + ph)
    # the animal runs along its facing direction, passing the origin mid-clip
    # (links point along body -x, so an animal with yaw psi faces the world direction -(cos psi, sin psi))
    q[:, :, 0] = -np.cos(heading) * s_along - np.sin(heading) * lateral
    q[:, :, 1] = -np.sin(heading) * s_along + np.cos(heading) * lateral
    q[:, :, 2] = 0.7 + 0.03 * np.sin(w * t + ph)
    q[:, :, 3] = 0.05 * np.sin(w * t + U(0, 6.28))                 # base roll
    q[:, :, 4] = 0.10 * np.sin(w * t + ph + 0.5)                   # base pitch
    q[:, :, 5] = heading + 0.05 * np.sin(w * t / 2 + U(0, 6.28))
    q[:, :, 6] = q[:, :, 3] + 0.04 * np.sin(w * t + U(0, 6.28))    # bodyF (free joint w.r.t. base)
    q[:, :, 7] = q[:, :, 4] + 0.15 * np.sin(w * t + ph + 2.0)
    q[:, :, 8] = q[:, :, 5] + 0.05 * np.sin(w * t + U(0, 6.28))
    q[:, :, 9] = q[:, :, 6] + 0.05 * np.sin(w * t + U(0, 6.28))    # neck
    q[:, :, 10] = q[:, :, 7] - 0.3 + 0.1 * np.sin(w * t + U(0, 6.28))
    q[:, :, 11] = q[:, :, 8] + 0.08 * np.sin(w * t / 2 + U(0, 6.28))
    R = {sk.BASE: euler_to_rot(q[:, :, 3], q[:, :, 4], q[:, :, 5]),
         sk.BODYF: euler_to_rot(q[:, :, 6], q[:, :, 7], q[:, :, 8])}

    def put(link, Rl):
        R[link] = Rl
        a, b, c = rot_to_euler(Rl)
        q[:, :, 3 + 3 * link], q[:, :, 4 + 3 * link], q[:, :, 5 + 3 * link] = a, b, c

    # tail: hooke joints
    put(sk.TAIL0, R[sk.BASE] @ _ry(0.3 + 0.25 * np.sin(w * t + U(0, 6.28))) @ _rz(0.2 * np.sin(w * t + U(0, 6.28))))
    put(sk.TAIL1, R[sk.TAIL0] @ _ry(0.25 * np.sin(w * t + U(0, 6.28))) @ _rz(0.25 * np.sin(w * t + U(0, 6.28))))
    # legs: revolute joints about the parent's y axis
    for body, Ul, Ll, Hl, back, lag in ((sk.BODYF, sk.UFL, sk.LFL, sk.HFL, False, 0.0), (sk.BODYF, sk.UFR, sk.LFR, sk.HFR, False, 0.6),
                                        (sk.BASE, sk.UBL, sk.LBL, sk.HBL, True, 3.0), (sk.BASE, sk.UBR, sk.LBR, sk.HBR, True, 3.6)):
        p = ph + lag
        d_u = 0.4 * np.sin(w * t + p)
        if back:
            d_l = -(0.5 + 0.3 * np.sin(w * t + p + 1.0))        # theta_U - theta_L in [0, pi]
            d_h = 0.5 + 0.3 * np.sin(w * t + p + 2.0)           # theta_L - theta_H in [-0.75 pi, 0]
        else:
            d_l = 0.45 + 0.25 * np.sin(w * t + p + 1.0)         # theta_U - theta_L in [-pi, 0]
            d_h = -(0.3 + 0.3 * np.sin(w * t + p + 2.0))        # theta_L - theta_H in [-pi/4, 0.75 pi]
        put(Ul, R[body] @ _ry(d_u))
        put(Ll, R[Ul] @ _ry(d_l))
        put(Hl, R[Ll] @ _ry(d_h))
    # unwrap yaw-like angles along time so that the trajectories are smooth (psi ~ pi sits on the atan2 cut)
    for c in range(3, 54):
        q[:, :, c] = np.unwrap(q[:, :, c], axis=1)
        if c % 3 == 2:   # psi columns: keep the same 2*pi sheet as the base heading
            q[:, :, c] += 2 * np.pi * np.round((q[:, :1, 5] - q[:, :1, c]) / (2 * np.pi))
    return q


def make_cameras(S, C, rng, radius=12.0):
    """cams[S,C,24] on a ring looking at the track centre; intrinsics as recovered from the shipped clips (G8)."""
    cams = np.zeros((S, C, sk.CAM_STRIDE))
    for s in range(S):
        a0 = rng.uniform(0, 2 * np.pi)
        for c in range(C):
            ang = a0 + 2 * np.pi * c / C + rng.uniform(-0.15, 0.15)
            pos = np.array([radius * np.cos(ang), radius * np.sin(ang), rng.uniform(0.5, 1.5)])
            fwd = np.array([0.0, 0.0, 0.5]) - pos
            fwd /= np.linalg.norm(fwd)
            right = np.cross(fwd, [0.0, 0.0, 1.0])
            right /= np.linalg.norm(right)
            down = np.cross(fwd, right)
            Rm = np.stack([right, down, fwd])
            K = np.array([[860.0, 0, 960.0], [0, 860.0, 540.0], [0, 0, 1]])
            cams[s, c] = sk.pack_camera(K, [0.12, -0.054, 0.015, -0.001], Rm, -Rm @ pos)
    return cams


@dataclass
class SyntheticProblem:
    q_true: np.ndarray       # [S,N,54]
    q_init: np.ndarray       # [S,N,54]
    cams: np.ndarray         # [S,C,24]
    chain_link: np.ndarray   # [24,5]
    chain_vec: np.ndarray    # [S,24,5,3]
    h: np.ndarray            # [S]
    meas: np.ndarray         # [S,N,C,24,W,4] float32: u, v, gate, 0
    n_frames: np.ndarray     # [S]
    fps: float

    def to_batch(self, device="cuda"):
        from . import engine
        return engine.Batch.from_numpy(self.meas, self.cams, self.chain_vec, self.h, self.n_frames, device=device)


def cuda_projector(device="cuda"):
    """uv = projection of the FK of q through the CUDA eval kernel (residual against zero detections)."""
    import torch
    from . import engine

    def project(q, cams, chain_vec):
        S, N = q.shape[:2]
        C = cams.shape[1]
        model = engine.Model(device)
        meas = np.zeros((S, N, C, 24, 1, 4), np.float32)
        batch = engine.Batch.from_numpy(meas, cams, chain_vec, np.ones(S), device=device)
        solver = engine.Solver(model, batch, with_solver=False)
        out = solver.evaluate(torch.as_tensor(np.ascontiguousarray(q), device=device), resid=True)
        return out["resid"][:, :, :, :, 0, :].cpu().numpy()

    return project


def make_problem(S=1, N=100, C=6, seed=0, fps=120.0, animal="jules", projector=None, outlier_frac=0.10,
                 n_unique=None, image_size=(1920, 1080)):
    """Synthetic batch C1(b)/C4 of SURVEY 8d.  `n_unique` < S tiles that many distinct trajectories (detection
    noise still differs per sequence) to bound host-side generation time for the large sweep."""
    rng = np.random.default_rng(seed)
    nu = S if n_unique is None else min(S, n_unique)
    lengths, radii, _ = sk.animal(animal)
    chain_link, cvec = sk.marker_chains(lengths, radii)
    q_u = make_trajectories(nu, N, fps, rng)
    cams_u = make_cameras(nu, C, rng)
    reps = (S + nu - 1) // nu
    q_true = np.tile(q_u, (reps, 1, 1))[:S]
    cams = np.tile(cams_u, (reps, 1, 1))[:S]
    chain_vec = np.tile(cvec[None], (S, 1, 1, 1))
    if projector is None:
        projector = cuda_projector()
    uv_u = projector(q_u, cams_u, chain_vec[:nu])
    uv = np.tile(uv_u, (reps, 1, 1, 1, 1))[:S]                              # [S,N,C,24,2]
    sigma = 1.0 / sk.measurement_weights()[0] / 2.0                          # the un-doubled sigma table
    det = uv + rng.normal(0.0, 1.0, uv.shape) * sigma[None, None, None, :, None]
    lik = rng.beta(8, 2, uv.shape[:-1])
    out = rng.uniform(size=uv.shape[:-1]) < outlier_frac
    det[out] = np.stack([rng.uniform(0, image_size[0], out.sum()), rng.uniform(0, image_size[1], out.sum())], -1)
    low = out & (rng.uniform(size=out.shape) < 0.5)
    lik[low] = rng.uniform(0.0, 0.5, low.sum())
    outside = (uv[..., 0] < 0) | (uv[..., 0] >= image_size[0]) | (uv[..., 1] < 0) | (uv[..., 1] >= image_size[1])
    outside |= ~np.isfinite(uv).all(-1)
    lik[outside] = 0.0
    det[outside] = 0.0
    meas = np.zeros(uv.shape[:-1] + (1, 4), np.float32)
    meas[..., 0, 0:2] = det.astype(np.float32)
    meas[..., 0, 2] = (lik > 0.5).astype(np.float32)                        # strict > dlc_thresh (acinoset_misc.py:231)
    # reference-style initial guess (acinoset_opt.py:574-583): angles 0, psi = heading, base on a rough track
    q_init = np.zeros_like(q_true)
    q_init[:, :, 0:3] = q_true[:, :, 0:3] + rng.normal(0, 0.03, (S, N, 3))
    head = np.pi + np.arctan2(q_true[:, -1, 1] - q_true[:, 0, 1], q_true[:, -1, 0] - q_true[:, 0, 0])
    head = head + 2 * np.pi * np.round((q_true[:, 0, 5] - head) / (2 * np.pi))
    q_init[:, :, 5::3] = head[:, None, None]
    return SyntheticProblem(q_true=q_true, q_init=q_init, cams=cams, chain_link=chain_link, chain_vec=chain_vec,
                            h=np.full(S, 1.0 / fps), meas=meas, n_frames=np.full(S, N, np.int32), fps=fps)


def make_device_problem(S, N=500, C=6, seed=0, fps=120.0, animal="jules", device="cuda", n_unique=32, chunk=256,
                        outlier_frac=0.10, image_size=(1920, 1080), dlc_native=False):
    """The large sweep built directly in HBM: `n_unique` distinct trajectories / camera rings are
    generated on the host exactly as in make_problem, projected once by the CUDA kernel, tiled to S sequences, and
    the per-sequence detection noise, likelihoods and outliers are drawn on the device (torch generator).
    With dlc_native=True the detections are kept as DLC triples (x, y, likelihood) -- CPE_MEAS_F32X3, gated on the
    device -- instead of the pre-gated float4 records; both describe the same problem.
    Returns (Batch, q_true[S,N,54], q_init[S,N,54]) as device tensors."""
    import torch
    from . import engine
    dev = torch.device(device)
    rng = np.random.default_rng(seed)
    nu = min(S, n_unique)
    lengths, radii, _ = sk.animal(animal)
    _, cvec = sk.marker_chains(lengths, radii)
    q_u = make_trajectories(nu, N, fps, rng)
    cams_u = make_cameras(nu, C, rng)
    cvec_u = np.tile(cvec[None], (nu, 1, 1, 1))
    uv_u = torch.as_tensor(cuda_projector(device)(q_u, cams_u, cvec_u), device=dev)           # [nu,N,C,24,2]
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(seed) + 12345)
    sigma = torch.as_tensor(1.0 / sk.measurement_weights()[0] / 2.0, device=dev)               # [24]
    meas = torch.empty(S, N, C, 24, 1, 3 if dlc_native else 4, dtype=torch.float32, device=dev)
    idx_all = torch.arange(S, device=dev) % nu
    W_, H_ = image_size
    for s0 in range(0, S, chunk):
        s1 = min(S, s0 + chunk)
        uv = uv_u[idx_all[s0:s1]]
        shp = uv.shape[:-1]
        det = uv + torch.randn(uv.shape, generator=gen, device=dev, dtype=torch.float64) * sigma[None, None, None, :, None]
        # likelihood ~ Beta(8,2) as the ratio of two gamma draws
        ga = torch._standard_gamma(torch.full(shp, 8.0, device=dev, dtype=torch.float32), generator=gen)
        gb = torch._standard_gamma(torch.full(shp, 2.0, device=dev, dtype=torch.float32), generator=gen)
        lik = ga / (ga + gb)
        out = torch.rand(shp, generator=gen, device=dev) < outlier_frac
        ou = torch.rand(shp, generator=gen, device=dev, dtype=torch.float64) * W_
        ov = torch.rand(shp, generator=gen, device=dev, dtype=torch.float64) * H_
        det = torch.where(out[..., None], torch.stack([ou, ov], -1), det)
        low = out & (torch.rand(shp, generator=gen, device=dev) < 0.5)
        lik = torch.where(low, torch.rand(shp, generator=gen, device=dev) * 0.5, lik)
        outside = (uv[..., 0] < 0) | (uv[..., 0] >= W_) | (uv[..., 1] < 0) | (uv[..., 1] >= H_)
        outside |= ~torch.isfinite(uv).all(-1)
        lik = torch.where(outside, torch.zeros_like(lik), lik)
        det = torch.where(outside[..., None], torch.zeros_like(det), det)
        m = meas[s0:s1, :, :, :, 0]
        m[..., 0:2] = det.to(torch.float32)
        if dlc_native:
            m[..., 2] = lik.to(torch.float32)
        else:
            m[..., 2] = (lik > 0.5).to(torch.float32)
            m[..., 3] = 0.0
    q_true_u = torch.as_tensor(q_u, device=dev)
    q_true = q_true_u[idx_all].contiguous()
    q_init = torch.zeros_like(q_true)
    q_init[:, :, 0:3] = q_true[:, :, 0:3] + 0.03 * torch.randn(S, N, 3, generator=gen, device=dev, dtype=torch.float64)
    head = np.pi + np.arctan2(q_u[:, -1, 1] - q_u[:, 0, 1], q_u[:, -1, 0] - q_u[:, 0, 0])
    head = head + 2 * np.pi * np.round((q_u[:, 0, 5] - head) / (2 * np.pi))
    q_init[:, :, 5::3] = torch.as_tensor(head, device=dev)[idx_all][:, None, None]
    batch = engine.Batch(
        n_frames=torch.full((S,), N, dtype=torch.int32, device=dev),
        h=torch.full((S,), 1.0 / fps, dtype=torch.float64, device=dev),
        meas=meas,
        cams=torch.as_tensor(cams_u, device=dev)[idx_all].contiguous(),
        chain_vec=torch.as_tensor(cvec, device=dev)[None].expand(S, -1, -1, -1).contiguous())
    return batch, q_true, q_init
