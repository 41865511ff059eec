"""Kinetic path, first slice: rows 0-5 of the equations of motion.  The golden values are the
reference's OWN symbolic model evaluated through sympy.lambdify (oracle/derive_eom.py, run in the build container);
the CUDA code is generated from the same expressions, so this checks the generator, the input order and the kernel."""
import os

import numpy as np
import pytest


def test_fixture_and_parameter_order(golden_dir):
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    assert tuple(g["feet"]) == kinetics.FEET
    assert np.allclose(g["par"][0], kinetics.link_params("jules"), rtol=0, atol=1e-15)
    z, xy = np.arange(4.0), np.arange(16.0).reshape(4, 4) + 10
    assert np.array_equal(kinetics.pack_grf(z, xy), np.concatenate([z, np.arange(16.0) + 10]))
    # free fall sanity of the golden rows: no velocity, no acceleration, no forces -> row 2 = total weight
    # (checked on the reference expressions when the fixture was made: gravity enters only through dEp/dz)
    assert g["eom"].shape == (g["q"].shape[0], 6)


@pytest.mark.gpu
def test_eom_rows_match_reference_expressions(golden_dir):
    import torch
    from cheetah_pose_estimation_b200 import kinetics
    g = np.load(os.path.join(golden_dir, "eom_ref.npz"))
    n = g["q"].shape[0]
    # every golden row as its own sequence (per-sequence link parameters), plus one batch with N > 1 and ragged frames
    out = kinetics.eom_rows(g["q"][:, None], g["dq"][:, None], g["ddq"][:, None], g["F"][:, None], g["par"])
    got = out[:, 0].cpu().numpy()
    scale = np.maximum(np.abs(g["eom"]), 1.0)
    assert (np.abs(got - g["eom"]) / scale).max() < 1e-9, (np.abs(got - g["eom"]) / scale).max()
    h = n // 2
    out2 = kinetics.eom_rows(g["q"][None, :h], g["dq"][None, :h], g["ddq"][None, :h], g["F"][None, :h], g["par"][:1],
                             n_frames=[h - 3])
    got2 = out2[0].cpu().numpy()
    assert (np.abs(got2[:h - 3] - g["eom"][:h - 3]) / scale[:h - 3]).max() < 1e-9
    assert not np.any(got2[h - 3:])
    # gravity check: at rest with zero forces row 2 (base z) equals the total weight
    z = np.zeros((1, 1, 54))
    q0 = z.copy()
    q0[0, 0, 2] = 0.7
    r = kinetics.eom_rows(q0, z, z, np.zeros((1, 1, 20)), g["par"][:1])[0, 0].cpu().numpy()
    assert abs(r[2] - g["par"][0, 17:34].sum() * 9.81) < 1e-9 and abs(r[0]) < 1e-9 and abs(r[1]) < 1e-9
