"""Generates tests/golden/kinetic_costs_ref.npz (run in the BUILD container, where /root/reference exists): the cost
terms of the reference's kinetic objective evaluated by the reference's OWN functions

    misc.motion_smoothing_cost (acinoset_misc.py:601-613), misc.change_in_torque_squared_cost (:616-628),
    misc.eom_slack_cost (:631-636), Motor3D.torque_squared_cost (shared/physical_education/motor.py:412-417)

on duck-typed stand-ins for the Pyomo model (they only index `m.pose[fe, l, d]`, `m.slack_eom[fe, cp, i]`,
`Tc[fe, idx]` ...), with `pose` = the shipped `positions` of a real clip.  TEST INFRASTRUCTURE.

usage: python -m oracle.make_golden_kinetic
"""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


class _Idx(dict):
    """`m.pose[fe, l, d]` style access on a dense array with 1-based indices."""

    def __init__(self, arr):
        super().__init__()
        self.arr = np.asarray(arr, float)

    def __getitem__(self, key):
        if isinstance(key, tuple) and all(isinstance(k, slice) for k in key):          # Tc[:, :]
            return [float(v) for v in self.arr.reshape(-1)]
        return float(self.arr[tuple(int(k) - 1 for k in key)])


class _M:
    pass


class _Robot:
    def __init__(self, m, n, eom_rows=6):
        self.m = m
        self.n = n
        self.eom = [None] * eom_rows

    def indices(self, one_based=True):
        return [(fe, 1) for fe in range(1, self.n + 1)]


def main():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "ref_shim"))
    sys.path.insert(1, REF)
    sys.path.insert(2, os.path.join(REF, "shared"))
    sys.path.insert(3, ROOT)
    os.chdir(REF)
    import acinoset_misc as misc
    import shared.physical_education as pe
    from oracle import kinetic_costs as kc
    rng = np.random.default_rng(11)
    d = pickle.load(open("data/test_set/2017_08_29/top/jules/run1_1/fte_kinematic/fte.pickle", "rb"))
    pose, q = d["positions"], d["q"]
    N = pose.shape[0]
    fps = 90.0
    m = _M()
    m.fe, m.L, m.D3 = range(1, N + 1), range(1, 25), range(1, 4)
    m.pose = _Idx(pose)
    smooth_ref = float(misc.motion_smoothing_cost(_Robot(m, N), fps))
    eom = rng.normal(0, 0.3, (N, 6))
    m.slack_eom = _Idx(eom[:, None, :])
    slack_ref = float(misc.eom_slack_cost(_Robot(m, N)))
    M = 12
    Tc = rng.normal(0, 2.0, (N, M))
    wt = rng.uniform(0.5, 2.0)

    class _Motor:
        pyomo_vars = {"Tc": _Idx(Tc)}
        pyomo_sets = {"Tc_set": range(1, M + 1)}
    torque_ref = float(pe.motor.Motor3D.torque_squared_cost(_Motor, wt))
    hm0 = 1.0 / fps
    m.hm0 = hm0
    orig = pe.motor.torques
    pe.motor.torques = lambda robot: [_Motor]
    try:
        dtau_ref = float(misc.change_in_torque_squared_cost(_Robot(m, N)))
    finally:
        pe.motor.torques = orig
    # the restatement must agree before the fixture is written
    assert abs(kc.motion_smoothing_cost(pose, fps) - smooth_ref) <= 1e-12 * smooth_ref
    assert abs(kc.eom_slack(eom)[0] - slack_ref) <= 1e-12 * slack_ref
    c1, _, c2, _ = kc.torque_costs(Tc, wt, hm0)
    assert abs(c1 - torque_ref) <= 1e-12 * torque_ref and abs(c2 - dtau_ref) <= 1e-12 * dtau_ref
    print("motion smoothing", smooth_ref, "eom slack", slack_ref, "torque^2", torque_ref, "change in torque^2", dtau_ref)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "kinetic_costs_ref.npz"), q=q, pose=pose, fps=fps,
                        smooth_ref=smooth_ref, eom=eom, slack_ref=slack_ref, Tc=Tc, weight=wt, hm0=hm0,
                        torque_ref=torque_ref, dtau_ref=dtau_ref)


if __name__ == "__main__":
    main()
