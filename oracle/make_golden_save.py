"""Golden vectors for the output side of the path (CheetahEstimator.save, acinoset_opt.py:278-373): COM, reduced
pose and the per-camera 2D projection CSV of a shipped clip.  Run in the BUILD container only (reads /root/reference):
    python oracle/make_golden_save.py
"""
import os
import pickle

import numpy as np
import pandas as pd

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    out = {}
    for tag, clip in (("jules", "2017_08_29/top/jules/run1_1"), ("phantom", "2019_03_07/phantom/run")):
        d = pickle.load(open(f"{REF}/data/test_set/{clip}/fte_kinematic/fte.pickle", "rb"))
        for k in ("com_pos", "com_vel", "x", "dx", "ddx"):
            out[f"{tag}_{k}"] = np.asarray(d[k], dtype=np.float64)
        out[f"{tag}_keys"] = np.array(sorted(d.keys()))
        df = pd.read_csv(f"{REF}/data/test_set/{clip}/fte_kinematic/cam1_fte.csv", header=[0, 1], index_col=0)
        out[f"{tag}_cam1_csv"] = df.to_numpy()
        out[f"{tag}_cam1_index"] = df.index.to_numpy()
        out[f"{tag}_cam1_columns"] = np.array(["|".join(c) for c in df.columns])
        with open(f"{REF}/data/test_set/{clip}/fte_kinematic/cam1_fte.csv") as f:
            out[f"{tag}_cam1_head"] = np.array([next(f) for _ in range(3)])
    np.savez_compressed(os.path.join(OUT, "save_ref.npz"), **out)
    for k, v in out.items():
        print(k, v.shape)


if __name__ == "__main__":
    main()
