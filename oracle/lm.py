"""Levenberg-Marquardt / augmented-Lagrangian reference solver (oracle; test infrastructure).

Stands in for the reference's solver seam  `default_solver(...).solve(robot.m)`  (acinoset_opt.py:611-617;
shared/physical_education/utils.py:459-548), which hands the NLP to IPOPT + MA97 -- neither is installed on
this box, so optimiser parity is UNPINNED against IPOPT; this file is the FP64 CPU statement of the algorithm
the CUDA path implements, used to check it step by step; its optimum is cross-checked against an independent
scipy SLSQP solve of the constrained NLP (oracle/nlp.py, tests/test_nlp_witness.py).

One LM step, per sequence:
  1. frame-local Gauss-Newton blocks H_k[54,54], g_k[54]            (cost.frame_normal_equations)
  2. Marquardt damping  H_k += damp * diag(H_k + motion diagonal)
  3. Schur-eliminate the 26 coordinates that carry no motion prior (Q_p = 0) inside each frame
        S_k = H_ii - H_id H_dd^-1 H_di ,  ghat_k = g_i - H_id H_dd^-1 g_d
  4. solve the block-banded SPD system  (blockdiag(S_k) + M) d_i = -ghat  with M the constant motion
     Hessian (block half-bandwidth 3, blocks are diagonal: 2 wq/h^4 * stencil) by block Cholesky
  5. back-substitute d_d = -H_dd^-1 (g_d + H_di d_i)
  6. accept if the merit decreased (Nielsen damping update), else raise the damping.
"""
import numpy as np

from . import cost, priors as opriors, tables


def split_indices():
    ind = tables.relative_angle_mask()
    dep = np.array([p for p in range(54) if p not in set(ind.tolist())])
    return ind, dep


def merit(pb, q, lam_eq, lam_in, pr=None):
    """Sum of the cost terms; with `pr` (monocular priors, oracle.priors) the pose and motion priors are added."""
    f = float(cost.cost_terms(pb, q, lam_eq, lam_in).sum())
    if pr is not None:
        f += float(opriors.prior_costs(pr, q).sum())
    return f


def schur_frames(H, g, diagH, damp):
    """Step 2+3.  Returns S[N,28,28], ghat[N,28], Y[N,26,28] (= Hdd^-1 Hdi), y[N,26] (= Hdd^-1 g_d)."""
    ind, dep = split_indices()
    N = H.shape[0]
    Hd = H.copy()
    idx = np.arange(54)
    Hd[:, idx, idx] += damp * diagH
    S = np.empty((N, 28, 28))
    ghat = np.empty((N, 28))
    Y = np.empty((N, 26, 28))
    y = np.empty((N, 26))
    for k in range(N):
        Hdd = Hd[k][np.ix_(dep, dep)]
        Hdi = Hd[k][np.ix_(dep, ind)]
        Hii = Hd[k][np.ix_(ind, ind)]
        L = np.linalg.cholesky(Hdd)
        Y[k] = np.linalg.solve(L.T, np.linalg.solve(L, Hdi))
        y[k] = np.linalg.solve(L.T, np.linalg.solve(L, g[k][dep]))
        S[k] = Hii - Hdi.T @ Y[k]
        ghat[k] = g[k][ind] - Hdi.T @ y[k]
    return S, ghat, Y, y


def band_solve(S, ghat, m, wd_i):
    """Step 4: block Cholesky of A (A_kk = S_k + m[k,0] diag(wd_i), A_{k,k-j} = m[k,j] diag(wd_i)),
    returns d_i[N,28] with A d = -ghat."""
    N = S.shape[0]
    n = S.shape[1]
    B = 3
    # L blocks: Ld[k] (lower-tri), Lo[k][j-1] = L_{k,k-j}
    Ld = np.zeros((N, n, n))
    Lo = np.zeros((N, B, n, n))
    for k in range(N):
        # off-diagonal blocks of row k: L_{k,k-j} for j = B..1 (left to right)
        for j in range(min(B, k), 0, -1):
            c = k - j
            Akc = m[k, j] * np.diag(wd_i)
            # subtract sum_{p<c} L_{k,p} L_{c,p}^T  over p within both bands
            for p in range(max(k - B, c - B, 0), c):
                Akc = Akc - Lo[k, k - p - 1] @ Lo[c, c - p - 1].T
            Lo[k, j - 1] = np.linalg.solve(Ld[c], Akc.T).T        # L_kc = A_kc L_cc^-T
        Akk = S[k] + m[k, 0] * np.diag(wd_i)
        for j in range(1, min(B, k) + 1):
            Akk = Akk - Lo[k, j - 1] @ Lo[k, j - 1].T
        Ld[k] = np.linalg.cholesky(Akk)
    # forward  L z = -ghat
    z = np.zeros((N, n))
    for k in range(N):
        r = -ghat[k].copy()
        for j in range(1, min(B, k) + 1):
            r -= Lo[k, j - 1] @ z[k - j]
        z[k] = np.linalg.solve(Ld[k], r)
    # backward L^T d = z
    d = np.zeros((N, n))
    for k in range(N - 1, -1, -1):
        r = z[k].copy()
        for j in range(1, B + 1):
            if k + j < N:
                r -= Lo[k + j, j - 1].T @ d[k + j]
        d[k] = np.linalg.solve(Ld[k].T, r)
    return d


def dense_band_solve(S, ghat, m, wd_i, H_extra):
    """(blockdiag(S) + const-acc band + H_extra) d = -ghat by a dense solve (monocular path: H_extra = AR Hessian)."""
    N, n = S.shape[0], S.shape[1]
    A = np.array(H_extra, dtype=float).copy()
    for k in range(N):
        A[k * n:(k + 1) * n, k * n:(k + 1) * n] += S[k]
        for j in range(0, 4):
            if k - j >= 0:
                blk = m[k, j] * np.diag(wd_i)
                A[k * n:(k + 1) * n, (k - j) * n:(k - j + 1) * n] += blk
                if j > 0:
                    A[(k - j) * n:(k - j + 1) * n, k * n:(k + 1) * n] += blk
    return np.linalg.solve(A, -ghat.reshape(-1)).reshape(N, n)


def lm_step(pb, q, lam_eq, lam_in, damp, pr=None):
    """Returns (delta[N,54], predicted_reduction, g[N,54]).  With `pr` the monocular priors enter: gradients of both,
    the PSD majoriser of the pose prior per frame (damped like the rest), the exact AR Hessian in the band."""
    ind, dep = split_indices()
    N = q.shape[0]
    H, g = cost.frame_normal_equations(pb, q, lam_eq, lam_in)
    g = g + cost.motion_gradient(pb, q)
    H_ar = None
    if pr is not None:
        if "gmm_w" in pr:
            gp, Hp = opriors.pose_grad_hess(pr, q)
            H = H + Hp
            g = g + gp
        if "ar_W" in pr:
            gm, Hm = opriors.motion_grad_hess(pr, q)
            g = g + gm.reshape(N, 54)
            sel = (np.arange(N)[:, None] * 54 + ind[None, :]).reshape(-1)
            H_ar = Hm[np.ix_(sel, sel)]
    m = cost.motion_stencil(N)
    wd = 2.0 * pb.wq / pb.h ** 4
    idx = np.arange(54)
    diagH = H[:, idx, idx] + m[:, 0:1] * wd[None, :]
    S, ghat, Y, y = schur_frames(H, g, diagH, damp)
    if pr is not None:
        d_i = dense_band_solve(S, ghat, m, wd[ind], H_ar if H_ar is not None else np.zeros((N * 28, N * 28)))
    else:
        d_i = band_solve(S, ghat, m, wd[ind])
    delta = np.zeros((N, 54))
    delta[:, ind] = d_i
    delta[:, dep] = -(y + np.einsum("nab,nb->na", Y, d_i))
    pred = -0.5 * np.sum(g * delta) + 0.5 * damp * np.sum(diagH * delta * delta)
    return delta, float(pred), g


def update_multipliers(pb, q, lam_eq, lam_in):
    _, _, c, v = cost.al_terms(pb, q, lam_eq, lam_in)
    return lam_eq + pb.mu_eq * c, pb.mu_in * v


def solve(pb, q0, max_iter=60, n_outer=4, damp0=1e-3, ftol=1e-6, mu_schedule=(1e4, 1e5, 1e6, 1e6), verbose=False,
          pr=None):
    """Returns dict(q, cost_terms, iters, accepted, history).  `mu_schedule[o]` is the AL penalty of outer
    iteration o (continuation: a soft penalty far from the optimum, stiff once the multipliers are good)."""
    N = q0.shape[0]
    q = q0.copy()
    lam_eq = np.zeros((N, 26))
    lam_in = np.zeros((N, len(pb.ineq_rows)))
    damp, nu = damp0, 2.0
    hist = []
    iters = accepted = 0
    for outer in range(n_outer):
        pb.mu_eq = pb.mu_in = float(mu_schedule[min(outer, len(mu_schedule) - 1)])
        F = merit(pb, q, lam_eq, lam_in, pr)
        stall = 0
        for it in range(max_iter):
            delta, pred, _ = lm_step(pb, q, lam_eq, lam_in, damp, pr)
            q_try = q + delta
            F_try = merit(pb, q_try, lam_eq, lam_in, pr)
            iters += 1
            if np.isfinite(F_try) and F_try < F:
                rho = (F - F_try) / max(pred, 1e-300)
                rel = (F - F_try) / max(abs(F), 1e-300)
                q, F = q_try, F_try
                damp = damp * max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3)
                damp = max(damp, 1e-12)
                nu = 2.0
                accepted += 1
                hist.append(F)
                if verbose:
                    print(f"outer {outer} it {it} F {F:.6f} rho {rho:.3f} damp {damp:.2e}")
                if rel < ftol:
                    break
            else:
                damp = min(damp * nu, 1e12)
                nu *= 2.0
                stall += 1
                if verbose:
                    print(f"outer {outer} it {it} reject F_try {F_try:.6f} damp {damp:.2e}")
                if damp >= 1e12:
                    break
        lam_eq, lam_in = update_multipliers(pb, q, lam_eq, lam_in)
    out = dict(q=q, cost_terms=cost.cost_terms(pb, q, lam_eq * 0, lam_in * 0), lam_eq=lam_eq, lam_in=lam_in,
               iters=iters, accepted=accepted, history=hist)
    if pr is not None:
        out["prior_costs"] = opriors.prior_costs(pr, q)
    return out
