"""Generates tests/golden/priors_ref.npz (run in the BUILD container, where /root/reference exists):

  * `s2s_in`, `s2s_out`: the reference's own `data_ops.series_to_supervised` (common/py_utils/data_ops.py:108-142) on a
    small seeded array with n_in = 4 -> pins the lag layout of oracle.priors / the product mirror;
  * frozen fitted parameters of the default monocular configuration (5-component GMM on the 22 joint-angle entries,
    MultiTaskLasso AR(4) on the 28 reduced-pose entries), fitted HERE with the installed scikit-learn from the
    reference's shipped `models/data-driven/dataset_full_pose.csv` by the product mirror
    `cheetah_pose_estimation_b200.acinoset_models` (the reference's own classes need PyTables and a removed sklearn
    argument: SURVEY.md 8c) -> "prior-fit parity unpinned", evaluation parity only;
  * known answers of the pose prior on seeded trajectories from THREE sources that do not share code with oracle.priors:
    `pose_frame_reference` -- the reference's OWN nested function `norm_pdf_multivariate` (acinoset_misc.py:691-694),
    whose source text is cut out of `misc.gmm_pose_cost` at generation time with `inspect` and executed with
    `exp = math.exp` (it cannot be imported: it is local to a function that builds a Pyomo model), summed as :696-707;
    `pose_frame_sklearn` -- `-log(exp(GaussianMixture.score_samples(x)) + 1e-12)` on the frozen parameters;
    `pose_frame_scipy` -- `scipy.stats.multivariate_normal.pdf`.  All three must agree with oracle.priors to 1e-10
    before the file is written.  `costs` / `motion_resid` stay the oracle's values (the AR term is pinned through the
    lag layout above; its prediction is one matrix product).

usage: python -m oracle.make_golden_priors
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def main():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "ref_shim"))
    sys.path.insert(1, REF)
    sys.path.insert(2, ROOT)
    from common.py_utils import data_ops                               # the reference's own module
    from cheetah_pose_estimation_b200 import acinoset_models as am
    from oracle import priors

    rng = np.random.default_rng(7)
    s2s_in = rng.normal(size=(9, 3))
    s2s_out = data_ops.series_to_supervised(s2s_in, n_in=4, n_step=1).to_numpy()
    assert np.array_equal(s2s_out, priors.series_to_supervised(s2s_in, 4))
    assert np.array_equal(s2s_out, am.series_to_supervised(s2s_in, 4, 1))

    ds = os.path.join(REF, "models", "data-driven", "dataset_full_pose.h5")
    fixture = os.path.join(ROOT, "tests", "golden", "priors_ref.npz")
    if os.path.exists(fixture) and "--refit" not in sys.argv:
        old = np.load(fixture)                                         # keep the frozen parameters of the first run
        pr = {k: old[k] for k in old.keys() if k.startswith(("gmm_", "ar_")) or k in ("window", "Tr")}
        print("reusing the frozen parameters of", fixture)
    else:
        pose = am.PoseModelGMM(ds, n_comps=5)
        motion = am.MotionModel(ds, window_size=4, lasso=True)
        pr = am.pack_priors(pose, motion)
        print("GMM train log-likelihood", pose.log_likelihood_train_set, "LR non-zeros", motion.model_non_zeros,
              "train rmse", motion.train_rmse)

    # seeded trajectories around dataset rows (so that the GMM is evaluated where it has mass)
    df = am.read_dataset(ds)
    X = df.iloc[:, :28].to_numpy()
    Tm = priors.reduced_matrix()
    from oracle import tables
    N = 12
    x = X[100:100 + N] + rng.normal(0, 0.02, (N, 28))
    # q with the given reduced pose: least-norm solution of Tm q = x
    q = (np.linalg.pinv(Tm) @ x.T).T
    assert np.abs(q @ Tm.T - x).max() < 1e-12
    # ---- independent pins of the GMM term --------------------------------------------------------------------------
    import inspect
    import math
    import textwrap
    import acinoset_misc as misc                                       # the reference's own module (Pyomo stubbed)
    src = inspect.getsource(misc.gmm_pose_cost).splitlines()
    i0 = next(i for i, ln in enumerate(src) if "def norm_pdf_multivariate" in ln)
    i1 = next(i for i in range(i0 + 1, len(src)) if src[i].strip().startswith("return")) + 1
    ns = {"np": np, "exp": math.exp}
    exec(textwrap.dedent("\n".join(src[i0:i1])), ns)                    # the reference's text, executed, not copied
    ref_pdf = ns["norm_pdf_multivariate"]
    x = q @ Tm.T
    frame_ref = np.array([-math.log(sum(w * ref_pdf(xx[6:], mu, cov) for w, mu, cov in
                                        zip(pr["gmm_w"], pr["gmm_mu"], pr["gmm_cov"])) + 1e-12) for xx in x])
    from sklearn.mixture import GaussianMixture
    from scipy.stats import multivariate_normal
    gm = GaussianMixture(n_components=len(pr["gmm_w"]), covariance_type="full")
    gm.weights_, gm.means_, gm.covariances_ = pr["gmm_w"], pr["gmm_mu"], pr["gmm_cov"]
    gm.precisions_cholesky_ = np.array([np.linalg.cholesky(np.linalg.inv(c)) for c in pr["gmm_cov"]])
    frame_skl = -np.log(np.exp(gm.score_samples(x[:, 6:])) + 1e-12)
    frame_sci = -np.log(sum(w * multivariate_normal(mu, cov).pdf(x[:, 6:]) for w, mu, cov in
                            zip(pr["gmm_w"], pr["gmm_mu"], pr["gmm_cov"])) + 1e-12)
    frame_orc = np.array([priors.pose_cost(pr, q[k:k + 1]) for k in range(N)])
    for name, v in (("reference fn", frame_ref), ("sklearn", frame_skl), ("scipy", frame_sci)):
        err = np.abs(v - frame_orc).max()
        print(f"pose prior per frame: oracle vs {name}: max abs diff {err:.2e}")
        assert err < 1e-10 * max(1.0, np.abs(frame_orc).max())
    out = dict(s2s_in=s2s_in, s2s_out=s2s_out, q=q, costs=priors.prior_costs(pr, q),
               motion_resid=priors.motion_residuals(pr, q), pose_frame_reference=frame_ref,
               pose_frame_sklearn=frame_skl, pose_frame_scipy=frame_sci, **pr)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "priors_ref.npz"), **out)
    print("costs", out["costs"])


if __name__ == "__main__":
    main()
