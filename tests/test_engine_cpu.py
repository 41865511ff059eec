"""Host-side logic of engine.py that needs no GPU: the compact detection layouts (Batch.compact / _pack_bits) and the
balanced block lists.  The kernels that read these layouts are covered by the GPU parity tests."""
import numpy as np
import pytest
import torch

from cheetah_pose_estimation_b200 import engine


def _batch(meas):
    S, N, C = meas.shape[:3]
    return engine.Batch(n_frames=torch.full((S,), N, dtype=torch.int32), h=torch.full((S,), 1 / 120.0, dtype=torch.float64),
                        meas=torch.as_tensor(meas), cams=torch.zeros(S, C, 24, dtype=torch.float64),
                        chain_vec=torch.zeros(S, 24, 5, 3, dtype=torch.float64), dlc_thresh=0.5)


def test_compact_layouts_on_the_host():
    rng = np.random.default_rng(0)
    m3 = rng.random((2, 5, 3, 24, 1, 3)).astype(np.float32)
    m3[..., 0:2] *= 1000.0
    m3[0, 0, 0, 0, 0, 2] = 0.5                      # exactly the threshold: off (strict >, acinoset_misc.py:231)
    m3[0, 1, 1, 3, 0, 0] = np.nan                   # NaN coordinate: off, finite stand-in
    m3[1, 2, 2, 5, 0, 2] = np.nan                   # NaN likelihood: off
    b = _batch(m3)
    assert b.meas_layout == engine.MEAS_F32X3
    c = b.compact()
    assert c.meas_layout == engine.MEAS_F32X2G and c.meas.dtype == torch.float32 and c.meas_gate.dtype == torch.uint8
    with np.errstate(invalid="ignore"):
        want = (m3[..., 2].astype(np.float64) > 0.5) & ~np.isnan(m3[..., 0:2]).any(-1)
    assert np.array_equal(c.meas_gate.numpy().astype(bool), want)
    assert not want[0, 0, 0, 0, 0] and not want[0, 1, 1, 3, 0] and not want[1, 2, 2, 5, 0]
    uv = np.where(np.isnan(m3[..., 0:2]).any(-1, keepdims=True), 0.0, m3[..., 0:2])
    assert np.array_equal(c.meas.numpy(), uv) and np.isfinite(c.meas.numpy()).all()
    bits = b.compact(bits=True)
    assert bits.meas_layout == engine.MEAS_F32X2B and tuple(bits.meas_gate.shape) == (2, 5 * 3 * 24 // 8)
    assert np.array_equal(bits.meas_gate.numpy(), np.packbits(want.reshape(2, -1), axis=1, bitorder="little"))
    assert torch.equal(c.compact(bits=True).meas_gate, bits.meas_gate) and bits.compact(bits=True) is bits
    # windows slice both gate shapes along the sequence axis
    w = bits.window(1, 2)
    assert w.S == 1 and torch.equal(w.meas_gate, bits.meas_gate[1:2]) and w.meas_layout == engine.MEAS_F32X2B
    # host-gated float4 records: the gate column is taken as it is
    m4 = np.zeros((2, 5, 3, 24, 1, 4), np.float32)
    m4[..., 0:2] = uv
    m4[..., 2] = want
    c4 = _batch(m4).compact()
    assert np.array_equal(c4.meas_gate.numpy().astype(bool), want) and torch.equal(c4.meas, c.meas)
    # double-precision detections are refused, not narrowed
    with pytest.raises(ValueError, match="float32"):
        _batch(m4.astype(np.float64)).compact()
