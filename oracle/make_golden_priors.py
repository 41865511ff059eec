"""Generates tests/golden/priors_ref.npz (run in the BUILD container, where /root/reference exists):

  * `s2s_in`, `s2s_out`: the reference's own `data_ops.series_to_supervised` (common/py_utils/data_ops.py:108-142) on a
    small seeded array with n_in = 4 -> pins the lag layout of oracle.priors / the product mirror;
  * frozen fitted parameters of the default monocular configuration (5-component GMM on the 22 joint-angle entries,
    MultiTaskLasso AR(4) on the 28 reduced-pose entries), fitted HERE with the installed scikit-learn from the
    reference's shipped `models/data-driven/dataset_full_pose.csv` by the product mirror
    `cheetah_pose_estimation_b200.acinoset_models` (the reference's own classes need PyTables and a removed sklearn
    argument: SURVEY.md 8c) -> "prior-fit parity unpinned", evaluation parity only;
  * known answers of the reference formulas (oracle.priors, a literal restatement) on seeded trajectories.

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
    out = dict(s2s_in=s2s_in, s2s_out=s2s_out, q=q, costs=priors.prior_costs(pr, q),
               motion_resid=priors.motion_residuals(pr, q), **pr)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "priors_ref.npz"), **out)
    print("costs", out["costs"])


if __name__ == "__main__":
    main()
