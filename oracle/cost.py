"""Cost, residuals, gradient and Gauss-Newton normal equations of one sequence (oracle; test infra).

Reference being restated (paths under /root/reference):
  residual definition   acinoset_misc.py:278-288   slack = proj(pose) - meas
  weights               acinoset_misc.py:211-231   1/R_pw[w][l] if likelihood > thresh else 0
  measurement cost      acinoset_misc.py:459-484   sum rho(mult_c * w * slack; 3, 10, 20)
  robust loss           acinoset_misc.py:2001-2015 redescending_loss
  motion cost           acinoset_misc.py:639-677 + collocation.py:58-68 (implicit Euler), weights :234
  objective             acinoset_opt.py:602        1e-3 * (meas + model + pose + motion)

The reference hands these to IPOPT as a constrained NLP; here the slack / dq / ddq variables are
eliminated analytically (SURVEY.md section 7 "hard parts"), which leaves q[N,54] with
  * measurement term   sum rho(w * (proj(FK(q)) - meas))
  * motion term        sum_{k>=3} sum_p wq_p ((q_k - 3 q_{k-1} + 3 q_{k-2} - q_{k-3}) / h^2)^2
  * 26 joint equalities / frame and 24 two-sided linear inequalities / frame, carried as
    augmented-Lagrangian terms (mu/2 (c + lam/mu)^2 ; PHR form for the inequalities).
"""
from dataclasses import dataclass, field

import numpy as np

from . import cameras, kinematics, tables

LOSS_A, LOSS_B, LOSS_C = 3.0, 10.0, 20.0


def _sig(x):
    return 1.0 / (1.0 + np.exp(-x))


def redescending_loss(err, a=LOSS_A, b=LOSS_B, c=LOSS_C):
    """acinoset_misc.py:2001-2015, vectorised (np.e ** x == exp(x))."""
    e = np.abs(err)
    sa, sb, sc = _sig(e - a), _sig(e - b), _sig(e - c)
    cost = (1 - sa) / 2 * e ** 2
    cost = cost + (sa - sb) * (a * e - (a ** 2) / 2)
    cost = cost + (sb - sc) * (a * b - (a ** 2) / 2 + (a * (c - b) / 2) * (1 - ((c - e) / (c - b)) ** 2))
    cost = cost + sc * (a * b - (a ** 2) / 2 + (a * (c - b) / 2))
    return cost


def redescending_dloss(err, a=LOSS_A, b=LOSS_B, c=LOSS_C):
    """d rho / d|err| (derivative w.r.t. the magnitude E = |err|); d rho/d err = sign(err) * this."""
    e = np.abs(err)
    sa, sb, sc = _sig(e - a), _sig(e - b), _sig(e - c)
    da, db, dc = sa * (1 - sa), sb * (1 - sb), sc * (1 - sc)
    p2 = a * e - a * a / 2
    p3 = a * b - a * a / 2 + (a * (c - b) / 2) * (1 - ((c - e) / (c - b)) ** 2)
    k4 = a * b - a * a / 2 + a * (c - b) / 2
    d = -da * e * e / 2 + (1 - sa) * e
    d = d + (da - db) * p2 + (sa - sb) * a
    d = d + (db - dc) * p3 + (sb - sc) * a * (c - e) / (c - b)
    d = d + dc * k4
    return d


def irls_weight(err):
    """Gauss-Newton curvature weight  max(rho'(E)/E, 0)  (majoriser of the robust loss)."""
    e = np.abs(err)
    d = redescending_dloss(err)
    with np.errstate(divide="ignore", invalid="ignore"):
        w = np.where(e > 0, d / e, 0.0)
    return np.maximum(w, 0.0)


@dataclass
class Problem:
    """One sequence.  meas[N,C,L,W,3] = (u, v, weight); weight already includes the camera
    multiplier (acinoset_misc.py:462-464) and the likelihood gate (:231)."""
    meas: np.ndarray
    cams: np.ndarray                      # [C,20]
    chain_link: np.ndarray                # [24,5]
    chain_vec: np.ndarray                 # [24,5,3]
    h: float                              # 1/fps
    cam_model: int = cameras.FISHEYE
    wq: np.ndarray = field(default_factory=tables.model_err_weight)      # [54]
    ineq_rows: list = field(default_factory=kinematics.inequality_rows)
    mu_eq: float = 1e4
    mu_in: float = 1e4
    square_loss: bool = False             # hand-labelled data: plain (w slack)^2 (acinoset_misc.py:471-474)

    @property
    def n_frames(self):
        return self.meas.shape[0]


def delayed_pose(pb: Problem, q, tau=None, dq_base=None, ddq_base=None):
    """Marker positions as camera c sees them, [N,C,24,3]:  pose + dq_base tau_c + ddq_base tau_c^2
    (acinoset_misc.py:280-286; the same expression in acinoset_opt.py:342-352 for the saved 2-D outputs).
    tau[C] seconds (the reference fixes tau[0] = 0, :274-275), dq_base / ddq_base [N,3] = the base x, y, z entries of
    the reference's dq / ddq variables.  tau None = no delay."""
    pose = kinematics.fk(q, pb.chain_link, pb.chain_vec)                     # [N,24,3]
    C = pb.cams.shape[0]
    if tau is None:
        return np.broadcast_to(pose[:, None], (pose.shape[0], C, 24, 3))
    tau = np.asarray(tau, float)
    off = (np.asarray(dq_base, float)[:, None, :] * tau[None, :, None]
           + np.asarray(ddq_base, float)[:, None, :] * (tau ** 2)[None, :, None])          # [N,C,3]
    return pose[:, None, :, :] + off[:, :, None, :]


def base_rates(q, h):
    """dq_base, ddq_base [N,3] once the collocation rows (collocation.py:58-68) are eliminated: backward differences
    of the base x, y, z.  The reference keeps dq and ddq of the first node (hence ddq of the second) as free
    variables; in the q-only formulation they take the values that zero their motion cost (the same as in
    the saved dq / ddq): ddq_0 = ddq_1 = ddq_2, dq_0 = dq_1 - h ddq_1 (0-based; stated deviation, first two frames)."""
    b = np.asarray(q, float)[:, 0:3]
    N = b.shape[0]
    v, a = np.zeros_like(b), np.zeros_like(b)
    if N >= 2:
        v[1:] = (b[1:] - b[:-1]) / h
        v[0] = v[1]
    if N >= 3:
        a[2:] = (v[2:] - v[1:-1]) / h
        a[0] = a[1] = a[2]
        v[0] = v[1] - h * a[1]
    return v, a


def delayed_cost_terms(pb: Problem, q, tau, lam_eq=None, lam_in=None):
    """cost_terms of the problem WITH shutter-delay estimation: the measurement term sees
    pose + dq_base(q) tau_c + ddq_base(q) tau_c^2 with the rates of base_rates."""
    v, a = base_rates(q, pb.h)
    t = cost_terms(pb, q, lam_eq, lam_in)
    t[0] = measurement_cost(pb, q, tau, v, a)
    return t


def residuals(pb: Problem, q, tau=None, dq_base=None, ddq_base=None):
    """slack[N,C,L,W,2] = proj(pose [+ shutter-delay displacement]) - meas  (what the reference saves as `meas_err`)."""
    pts = delayed_pose(pb, q, tau, dq_base, ddq_base)                        # [N,C,24,3]
    uv = cameras.project(pts, pb.cams[None, :, None, :], pb.cam_model)       # [N,C,L,2]
    return uv[:, :, :, None, :] - pb.meas[..., 0:2]


def shutter_jacobian(pb: Problem, q, tau, dq_base, ddq_base):
    """d slack / d tau_c, [N,C,L,2] = dproj/dpoint . (dq_base + 2 tau_c ddq_base)  (derivative of :283-286)."""
    pts = delayed_pose(pb, q, tau, dq_base, ddq_base)
    P = cameras.project_jacobian(pts, pb.cams[None, :, None, :], pb.cam_model)             # [N,C,L,2,3]
    tau = np.asarray(tau, float)
    d = np.asarray(dq_base, float)[:, None, :] + 2.0 * tau[None, :, None] * np.asarray(ddq_base, float)[:, None, :]
    return np.einsum("ncldi,nci->ncld", P, d)


def measurement_cost(pb: Problem, q, tau=None, dq_base=None, ddq_base=None):
    r = residuals(pb, q, tau, dq_base, ddq_base)
    e = pb.meas[..., 2:3] * r
    if pb.square_loss:
        return float(np.sum(e * e))
    return float(np.sum(redescending_loss(e)))


def third_difference(q, h):
    """(ddq_k - ddq_{k-1}) for k = 3..N-1 in q-only form, [N-3,54]."""
    q = np.asarray(q, float)
    if q.shape[0] < 4:
        return np.zeros((0, q.shape[1]))
    return (q[3:] - 3 * q[2:-1] + 3 * q[1:-2] - q[:-3]) / (h * h)


def motion_cost(pb: Problem, q):
    d = third_difference(q, pb.h)
    return float(np.sum(pb.wq * d * d))


def al_terms(pb: Problem, q, lam_eq, lam_in):
    """(eq_term, in_term, c[N,26], v[N,24]) ; v = t - clamp(t, lo, up), t = d + lam/mu."""
    c = kinematics.joint_constraints(q)
    eq = 0.5 * pb.mu_eq * np.sum((c + lam_eq / pb.mu_eq) ** 2)
    d = kinematics.inequality_values(q, pb.ineq_rows)
    lo = np.array([r[2] for r in pb.ineq_rows])
    up = np.array([r[3] for r in pb.ineq_rows])
    t = d + lam_in / pb.mu_in
    v = t - np.clip(t, lo, up)
    ine = 0.5 * pb.mu_in * np.sum(v * v)
    return float(eq), float(ine), c, v


def cost_terms(pb: Problem, q, lam_eq=None, lam_in=None):
    """[measurement, model, al_eq, al_in]; the reference objective is 1e-3 * (t[0] + t[1])."""
    N = pb.n_frames
    lam_eq = np.zeros((N, 26)) if lam_eq is None else lam_eq
    lam_in = np.zeros((N, len(pb.ineq_rows))) if lam_in is None else lam_in
    eq, ine, _, _ = al_terms(pb, q, lam_eq, lam_in)
    return np.array([measurement_cost(pb, q), motion_cost(pb, q), eq, ine])


def motion_stencil(N):
    """Band of D3^T D3 : m[k, j] = coefficient coupling frame k with frame k-j (j = 0..3)."""
    D = np.zeros((max(N - 3, 0), N))
    for r in range(max(N - 3, 0)):
        D[r, r:r + 4] = [-1.0, 3.0, -3.0, 1.0]
    G = D.T @ D
    m = np.zeros((N, 4))
    for k in range(N):
        for j in range(4):
            if k - j >= 0:
                m[k, j] = G[k, k - j]
    return m


def frame_normal_equations(pb: Problem, q, lam_eq=None, lam_in=None):
    """Frame-local Gauss-Newton pieces, WITHOUT the motion term:
       H[N,54,54], g[N,54]  with  H = sum w^2 omega Jr^T Jr + mu Jc^T Jc + mu (active ineq),
       g = sum rho'(e) w Jr^T sign + mu (c + lam/mu) Jc + mu v Jd."""
    N = pb.n_frames
    lam_eq = np.zeros((N, 26)) if lam_eq is None else lam_eq
    lam_in = np.zeros((N, len(pb.ineq_rows))) if lam_in is None else lam_in
    pose = kinematics.fk(q, pb.chain_link, pb.chain_vec)
    G = kinematics.fk_jacobian(q, pb.chain_link, pb.chain_vec)               # [N,24,3,54]
    P = cameras.project_jacobian(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)   # [N,C,L,2,3]
    uv = cameras.project(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)
    r = uv[:, :, :, None, :] - pb.meas[..., 0:2]                              # [N,C,L,W,2]
    w = pb.meas[..., 2:3]
    e = w * r
    if pb.square_loss:
        gsc = 2.0 * e * w
        hsc = 2.0 * w * w * np.ones_like(e)
    else:
        gsc = redescending_dloss(e) * np.sign(e) * w                          # d rho / d slack
        hsc = irls_weight(e) * w * w
    # per marker 3x3 / 3 accumulations over cameras and hypotheses
    A = np.einsum("nclwd,ncldi,ncldj->nlij", hsc, P, P)
    b = np.einsum("nclwd,ncldi->nli", gsc, P)
    H = np.einsum("nlia,nlij,nljb->nab", G, A, G)
    g = np.einsum("nlia,nli->na", G, b)
    # equality AL
    c = kinematics.joint_constraints(q)
    Jc = kinematics.joint_constraints_jacobian(q)
    H += pb.mu_eq * np.einsum("nka,nkb->nab", Jc, Jc)
    g += pb.mu_eq * np.einsum("nk,nka->na", c + lam_eq / pb.mu_eq, Jc)
    # inequality AL (PHR)
    _, _, _, v = al_terms(pb, q, lam_eq, lam_in)
    for j, (ip, im, lo, up) in enumerate(pb.ineq_rows):
        act = (v[:, j] != 0).astype(float)
        g[:, ip] += pb.mu_in * v[:, j]
        H[:, ip, ip] += pb.mu_in * act
        if im >= 0:
            g[:, im] -= pb.mu_in * v[:, j]
            H[:, im, im] += pb.mu_in * act
            H[:, ip, im] -= pb.mu_in * act
            H[:, im, ip] -= pb.mu_in * act
    return H, g


def motion_gradient(pb: Problem, q):
    """d/dq of the motion term, [N,54]."""
    N = q.shape[0]
    d = third_difference(q, pb.h)
    g = np.zeros_like(q)
    if N < 4:
        return g
    wd = 2.0 * pb.wq * d / (pb.h * pb.h)
    g[3:] += wd
    g[2:-1] -= 3 * wd
    g[1:-2] += 3 * wd
    g[:-3] -= wd
    return g


def full_gradient(pb: Problem, q, lam_eq=None, lam_in=None):
    _, g = frame_normal_equations(pb, q, lam_eq, lam_in)
    return g + motion_gradient(pb, q)


def dense_hessian(pb: Problem, q, lam_eq=None, lam_in=None):
    """Full GN Hessian [N*54, N*54] incl. motion term (small N only) and gradient."""
    N = q.shape[0]
    H, g = frame_normal_equations(pb, q, lam_eq, lam_in)
    g = g + motion_gradient(pb, q)
    A = np.zeros((N * 54, N * 54))
    m = motion_stencil(N)
    wd = 2.0 * pb.wq / pb.h ** 4
    for k in range(N):
        A[k * 54:(k + 1) * 54, k * 54:(k + 1) * 54] = H[k]
        for j in range(4):
            if k - j >= 0 and m[k, j] != 0:
                blk = np.diag(m[k, j] * wd)
                A[k * 54:(k + 1) * 54, (k - j) * 54:(k - j + 1) * 54] += blk
                if j > 0:
                    A[(k - j) * 54:(k - j + 1) * 54, k * 54:(k + 1) * 54] += blk
    return A, g.reshape(-1)
