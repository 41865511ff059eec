"""Mirror of the reference's data-driven priors (acinoset_models.py:173-300): the AR(window) linear motion model on the
reduced pose and the GMM pose prior.  As in the reference, FITTING is scikit-learn on the CPU (it happens once per
run); what moves to the GPU is the evaluation of the two cost terms, their gradients and
Gauss-Newton blocks inside the LM loop (csrc/cpe_prior.cu).

The reference reads `models/data-driven/*.h5` through PyTables; that is not available here, so the `.csv` twins the
reference ships beside them are read instead (same rows, same 56 columns; segment starts where the index is 0).
"""
import os

import numpy as np
import pandas as pd

NUM_VARS = 28        # reduced pose (acinoset_misc.py:1699-1757)
EXT_DIM = 6          # base x, y, z, phi, theta, psi are excluded from the pose prior (acinoset_misc.py:698)


def read_dataset(fname):
    """DataFrame of a data-driven dataset; accepts the reference's `.h5` name and falls back to its `.csv` twin."""
    base, ext = os.path.splitext(fname)
    if ext == ".h5":
        try:
            return pd.read_hdf(fname)
        except (ImportError, FileNotFoundError):                                 # PyTables absent: the shipped CSV holds the same table
            return pd.read_csv(base + ".csv", index_col=0)
    return pd.read_csv(fname, index_col=0)


def series_to_supervised(data, n_in=1, n_step=1):
    """Rows [x(t-n_in*n_step), ..., x(t-n_step) ; x(t)], oldest lag first, first n_in*n_step rows dropped
    (common/py_utils/data_ops.py:108-142 with n_out = 1, dropnan = True), as a plain array."""
    data = np.asarray(data)
    lag = n_in * n_step
    n = data.shape[0] - lag
    if n <= 0:
        return np.zeros((0, data.shape[1] * (n_in + 1)))
    cols = [data[lag - i:lag - i + n] for i in range(lag, 0, -n_step)]
    cols.append(data[lag:])
    return np.concatenate(cols, axis=1)


class MotionModel:
    """acinoset_models.py:173-271 (pose_model = None branch): MultiTaskLasso(alpha=1e-2, random_state=42) or plain
    LinearRegression on windows of the reduced pose; `error_variance` = training residual variance."""

    def __init__(self, dataset_fname, num_params=NUM_VARS, start_idx=0, window_size=4, window_time=1, lasso=True):
        from sklearn.linear_model import LinearRegression, MultiTaskLasso
        self.window_size, self.window_time = window_size, window_time
        X, y = self._read_dataset(dataset_fname, start_idx, num_params)
        self.lr_model = MultiTaskLasso(alpha=1e-2, random_state=42, max_iter=20000) if lasso else LinearRegression()
        self.lr_model.fit(X, y)
        y_pred = self.lr_model.predict(X)
        self.model_non_zeros = int(np.count_nonzero(self.lr_model.coef_))
        residuals = y - y_pred
        self.error_variance = np.var(residuals, axis=0)
        self.train_rmse = float(np.sqrt(np.mean(residuals ** 2)))

    @property
    def coef_(self):
        return np.asarray(self.lr_model.coef_, dtype=np.float64)           # [28, 28*window]

    @property
    def intercept_(self):
        return np.asarray(self.lr_model.intercept_, dtype=np.float64)      # [28]

    def predict(self, X, matrix=False):
        X = np.asarray(X)
        if matrix:
            return np.dot(self.lr_model.coef_, X.T).T + np.tile(self.lr_model.intercept_, (X.shape[0], 1))
        return np.dot(self.lr_model.coef_, X.flatten()) + self.lr_model.intercept_

    def _read_dataset(self, dataset_fname, start_idx, num_params):
        df = read_dataset(dataset_fname)
        idx = np.where(df.index.values == 0)[0]
        data_in = df.iloc[:, start_idx:start_idx + num_params].to_numpy()
        parts, end_segment = [], 0
        for begin_segment, end_segment in zip(idx, idx[1:]):
            parts.append(series_to_supervised(data_in[begin_segment:end_segment], self.window_size, self.window_time))
        parts.append(series_to_supervised(data_in[end_segment:], self.window_size, self.window_time))
        xy = np.concatenate(parts, axis=0)
        n = num_params * self.window_size
        return xy[:, :n], xy[:, n:]


class PoseModelGMM:
    """acinoset_models.py:277-300 (pose_model = None): GaussianMixture(n_comps, random_state=42, max_iter=20000) on the
    22 joint-angle entries of the reduced pose."""

    def __init__(self, dataset_fname, num_vars=NUM_VARS, ext_dim=EXT_DIM, n_comps=5):
        from sklearn.mixture import GaussianMixture
        self.n_comps, self.num_vars, self.ext_dim = n_comps, num_vars, ext_dim
        df = read_dataset(dataset_fname)
        self.X = df.iloc[:, ext_dim:num_vars].to_numpy()
        self.gmm = GaussianMixture(n_components=n_comps, random_state=42, max_iter=20000).fit(self.X)
        self.log_likelihood_train_set = float(self.gmm.score(self.X))


def pack_priors(pose_model=None, motion_model=None):
    """Frozen parameters in the layout of `cpe_prior` (include/cpe.h): dict of float64 arrays."""
    out = {}
    if pose_model is not None:
        g = pose_model.gmm if hasattr(pose_model, "gmm") else pose_model
        w, mu, cov = np.asarray(g.weights_), np.asarray(g.means_), np.asarray(g.covariances_)
        out.update(gmm_w=w, gmm_mu=mu, gmm_cov=cov)
    if motion_model is not None:
        out.update(ar_W=np.asarray(motion_model.coef_, np.float64), ar_b=np.asarray(motion_model.intercept_, np.float64),
                   ar_var=np.asarray(motion_model.error_variance, np.float64))
    return out
