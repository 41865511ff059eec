"""CPU tests of the monocular data-driven priors: the oracle against the reference-pinned fixture, the product's host
logic (lag layout, reduced matrix, band builder) against the oracle's independent dense construction."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "priors_ref.npz"))


def _pr(g):
    return {k: g[k] for k in ("gmm_w", "gmm_mu", "gmm_cov", "ar_W", "ar_b", "ar_var")}


def test_lag_layout_matches_reference_series_to_supervised(gold):
    """`s2s_out` was produced by the reference's own data_ops.series_to_supervised (oracle/make_golden_priors.py)."""
    from cheetah_pose_estimation_b200 import acinoset_models as am
    from oracle import priors
    assert np.array_equal(priors.series_to_supervised(gold["s2s_in"], 4), gold["s2s_out"])
    assert np.array_equal(am.series_to_supervised(gold["s2s_in"], 4, 1), gold["s2s_out"])
    assert am.series_to_supervised(gold["s2s_in"][:3], 4, 1).shape == (0, 15)          # ragged: shorter than the window


def test_oracle_reproduces_fixture_costs(gold):
    from oracle import priors
    pr = _pr(gold)
    assert np.allclose(priors.prior_costs(pr, gold["q"]), gold["costs"], rtol=1e-12)
    assert np.allclose(priors.motion_residuals(pr, gold["q"]), gold["motion_resid"], atol=1e-12)


def test_gmm_pose_term_is_pinned_by_three_independent_evaluations(gold):
    """Row a11: the oracle's per-frame pose cost against (i) the reference's OWN `norm_pdf_multivariate` (its source
    cut out of misc.gmm_pose_cost and executed at fixture-generation time), (ii) scikit-learn's mixture density
    (`-log(exp(score_samples) + 1e-12)`) and (iii) scipy's multivariate normal -- oracle/make_golden_priors.py."""
    from oracle import priors
    pr = _pr(gold)
    q = gold["q"]
    frame = np.array([priors.pose_cost(pr, q[k:k + 1]) for k in range(q.shape[0])])
    for key in ("pose_frame_reference", "pose_frame_sklearn", "pose_frame_scipy"):
        assert np.abs(frame - gold[key]).max() < 1e-10, key
    assert abs(frame.sum() - gold["costs"][0]) < 1e-10


def test_reduced_matrix_agrees_with_oracle_and_only_uses_prior_coordinates():
    from cheetah_pose_estimation_b200 import priors as pp, skeleton
    from oracle import priors
    ind = skeleton.independent_coords()
    Tm = priors.reduced_matrix()
    assert np.array_equal(pp.reduced_matrix(), Tm[:, ind])
    assert not np.any(np.delete(Tm, ind, axis=1))


def test_oracle_derivatives_by_finite_differences(gold):
    from oracle import priors
    pr = _pr(gold)
    q = gold["q"][:7].copy()
    gp, Hp = priors.pose_grad_hess(pr, q)
    gm, Hm = priors.motion_grad_hess(pr, q)
    rng = np.random.default_rng(0)
    d = rng.normal(size=q.shape) * 1e-6
    f = lambda z: priors.prior_costs(pr, z)
    num = (f(q + d) - f(q - d)) / 2
    assert abs(num[0] - np.sum(gp * d)) <= 1e-6 * abs(num[0]) + 1e-12
    assert abs(num[1] - gm @ d.reshape(-1)) <= 1e-6 * abs(num[1]) + 1e-12
    # motion prior is quadratic: the Hessian is exact
    d2 = rng.normal(size=q.shape) * 1e-2
    quad = f(q + d2)[1] - f(q)[1] - gm @ d2.reshape(-1)
    assert abs(quad - 0.5 * d2.reshape(-1) @ Hm @ d2.reshape(-1)) <= 1e-9 * abs(quad)
    # the pose majoriser is PSD
    assert min(np.linalg.eigvalsh(Hp[k]).min() for k in range(q.shape[0])) > -1e-9


@pytest.mark.parametrize("nf,N", [(9, 9), (6, 8), (4, 4), (3, 5)])
def test_host_band_equals_oracle_dense_hessian(gold, nf, N):
    """Priors.mband (scalar lower band of const-acc + AR terms on the 28 prior coordinates) against
    2 J^T W J (oracle.priors) + the oracle's constant-acceleration Hessian."""
    from cheetah_pose_estimation_b200 import skeleton
    from cheetah_pose_estimation_b200.priors import Priors
    from oracle import cost, priors
    import torch
    pr = _pr(gold)
    P = Priors(pr, device="cpu")
    h = 1.0 / 90.0
    band = P.mband(nf, h, N)
    assert band.shape == (N * 28, P.bw + 1) and P.bw == 139
    ind = skeleton.independent_coords()
    _, Hm = priors.motion_grad_hess(pr, np.zeros((nf, 54)))
    idx = (np.arange(nf)[:, None] * 54 + ind[None, :]).reshape(-1)
    dense = Hm[np.ix_(idx, idx)]
    # constant-acceleration part: 2 D3^T W D3 / h^4
    wq = skeleton.motion_weights()[ind]
    D3 = np.zeros((max(nf - 3, 0), nf))
    for r in range(nf - 3):
        D3[r, r:r + 4] = [-1, 3, -3, 1]
    dense = dense + np.kron(D3.T @ D3, np.diag(2 * wq / h ** 4))
    rec = np.zeros((N * 28, N * 28))
    for i in range(N * 28):
        for d in range(min(i, P.bw) + 1):
            rec[i, i - d] = band[i, d]
    rec = rec + np.tril(rec, -1).T
    n = nf * 28
    scale = np.abs(dense).max()
    assert np.abs(rec[:n, :n] - dense).max() <= 1e-12 * scale
    assert not np.any(rec[n:]) and not np.any(rec[:, n:])


def test_models_mirror_fits_and_packs(tmp_path):
    """MotionModel / PoseModelGMM mirrors on a small synthetic dataset written in the reference's CSV format."""
    import pandas as pd
    from cheetah_pose_estimation_b200 import acinoset_models as am
    rng = np.random.default_rng(1)
    segs = [np.cumsum(rng.normal(0, 0.05, (40, 56)), axis=0) for _ in range(3)]
    idx = np.concatenate([np.arange(40)] * 3)
    df = pd.DataFrame(np.concatenate(segs), index=idx)
    f = tmp_path / "dataset_full_pose.csv"
    df.to_csv(f)
    mm = am.MotionModel(str(tmp_path / "dataset_full_pose.h5"), window_size=4, lasso=False)   # .h5 name -> .csv twin
    assert mm.coef_.shape == (28, 112) and mm.error_variance.shape == (28,)
    X = np.concatenate(segs)[:, :28]
    row = np.concatenate([X[0], X[1], X[2], X[3]])
    assert np.allclose(mm.predict(row), mm.predict(row[None], matrix=True)[0])
    pm = am.PoseModelGMM(str(f), n_comps=2)
    pk = am.pack_priors(pm, mm)
    assert pk["gmm_cov"].shape == (2, 22, 22) and pk["ar_W"].shape == (28, 112)
