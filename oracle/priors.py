"""ORACLE (test infrastructure only -- never imported by the product package): CPU restatement of the reference's
data-driven monocular priors, evaluation side.

  reduced pose      x = get_relative_angles(q)[mask]                    acinoset_misc.py:487-528, 1699-1757
  pose prior        sum_frames -log( sum_c w_c N(x[6:]; mu_c, Sigma_c) + 1e-12 )   acinoset_misc.py:680-714
                    (norm_pdf_multivariate with np.linalg.det / np.linalg.inv, :691-694)
  motion prior      for fe > window: r = x[fe] - (W [x(fe-w); ...; x(fe-1)] + b),  cost = sum r^2 / var_p
                    acinoset_misc.py:291-336, acinoset_models.py:242-247 (predict), data_ops.py:108-142 (lag layout)

Parity status: the lag layout is pinned against the reference's own `series_to_supervised` and the reduced pose
against `get_relative_angles` (tests/golden/priors_ref.npz, reduced_pose_ref.npz).  The fitted parameters are NOT
pinned (the reference's fit needs PyTables and a removed sklearn argument -- SURVEY.md 8c iii): they are frozen into
the fixture and only the evaluation is compared ("prior-fit parity unpinned").
"""
import numpy as np

from . import kinematics, tables

WINDOW = 4
EXT_DIM = 6


def reduced_matrix():
    """Tm[28,54]: x = Tm @ q."""
    T = kinematics.relative_matrix()
    return T[np.asarray(tables.relative_angle_mask())]


def series_to_supervised(data, n_in, n_step=1):
    """Literal restatement of data_ops.py:108-142 (n_out = 1, dropnan): shift(n_in*n_step) ... shift(n_step), shift(0)."""
    data = np.asarray(data, dtype=float)
    n, lag = data.shape[0], n_in * n_step
    rows = []
    for t in range(lag, n):
        rows.append(np.concatenate([data[t - i] for i in range(lag, 0, -n_step)] + [data[t]]))
    return np.array(rows).reshape(len(rows), data.shape[1] * (n_in + 1))


def norm_pdf_multivariate(x, mu, cov):
    """acinoset_misc.py:691-694, with math.exp in place of the Pyomo exp."""
    part1 = 1 / (((2 * np.pi) ** (len(x) / 2)) * (np.linalg.det(cov) ** (1 / 2)))
    part2 = (-1 / 2) * ((x - mu).T.dot(np.linalg.inv(cov))).dot((x - mu))
    return part1 * np.exp(part2)


def pose_cost(pr, q):
    """sum over frames of -log(sum_c w_c N(x[6:]) + 1e-12)  (acinoset_misc.py:696-707)."""
    X = q @ reduced_matrix().T
    total = 0.0
    for x in X:
        xs = x[EXT_DIM:]
        total += -np.log(sum(w * norm_pdf_multivariate(xs, mu, cov)
                             for w, mu, cov in zip(pr["gmm_w"], pr["gmm_mu"], pr["gmm_cov"])) + 1e-12)
    return float(total)


def motion_residuals(pr, q, window=WINDOW):
    """r[N-window, 28] for frames window..N-1 (acinoset_misc.py:310-323)."""
    X = q @ reduced_matrix().T
    df = series_to_supervised(X, n_in=window)
    if df.shape[0] == 0:
        return np.zeros((0, X.shape[1]))
    X_in = df[:, :X.shape[1] * window]
    y_pred = np.dot(pr["ar_W"], X_in.T).T + np.tile(pr["ar_b"], (X_in.shape[0], 1))
    return X[window:] - y_pred


def motion_weights(pr):
    v = np.asarray(pr["ar_var"], dtype=float)
    return np.where(v != 0, 1.0 / np.where(v != 0, v, 1.0), 0.0)          # acinoset_misc.py:308


def motion_prior_cost(pr, q, window=WINDOW):
    r = motion_residuals(pr, q, window)
    return float(np.sum(motion_weights(pr)[None, :] * r * r))


def prior_costs(pr, q):
    return np.array([pose_cost(pr, q) if "gmm_w" in pr else 0.0, motion_prior_cost(pr, q) if "ar_W" in pr else 0.0])


# ---- derivatives used by the LM formulation (not in the reference: IPOPT gets them from ASL) --------------------------
def pose_grad_hess(pr, q):
    """Per frame: gradient of the pose term w.r.t. q (54) and the PSD majoriser sum_c r_c Sigma_c^-1 of its Hessian,
    both mapped through x = Tm q.  Returns g[N,54], H[N,54,54]."""
    Tm = reduced_matrix()
    Ts = Tm[EXT_DIM:]
    N = q.shape[0]
    g = np.zeros((N, 54))
    H = np.zeros((N, 54, 54))
    P = [np.linalg.inv(c) for c in pr["gmm_cov"]]
    for k in range(N):
        xs = Ts @ q[k]
        pdf = np.array([w * norm_pdf_multivariate(xs, mu, cov)
                        for w, mu, cov in zip(pr["gmm_w"], pr["gmm_mu"], pr["gmm_cov"])])
        r = pdf / (pdf.sum() + 1e-12)
        gx = sum(rc * Pc @ (xs - mu) for rc, Pc, mu in zip(r, P, pr["gmm_mu"]))
        Hx = sum(rc * Pc for rc, Pc in zip(r, P))
        g[k] = Ts.T @ gx
        H[k] = Ts.T @ Hx @ Ts
    return g, H


def motion_jacobian(pr, N, window=WINDOW):
    """Dense J[(N-window)*28, N*54] of the (linear) motion-prior residuals."""
    Tm = reduced_matrix()
    n = Tm.shape[0]
    J = np.zeros(((N - window) * n, N * 54))
    for k in range(window, N):
        row = slice((k - window) * n, (k - window + 1) * n)
        J[row, k * 54:(k + 1) * 54] += Tm
        for j in range(1, window + 1):                                    # lag j sits at columns (window-j)*28
            Wj = pr["ar_W"][:, (window - j) * n:(window - j + 1) * n]
            J[row, (k - j) * 54:(k - j + 1) * 54] -= Wj @ Tm
    return J


def motion_grad_hess(pr, q, window=WINDOW):
    """g[N*54] = 2 J^T (w r), H[N*54, N*54] = 2 J^T W J (exact: the term is a linear least-squares)."""
    N = q.shape[0]
    if N <= window:
        return np.zeros(N * 54), np.zeros((N * 54, N * 54))
    J = motion_jacobian(pr, N, window)
    w = np.tile(motion_weights(pr), N - window)
    r = motion_residuals(pr, q, window).reshape(-1)
    return 2.0 * J.T @ (w * r), 2.0 * (J.T * w) @ J
