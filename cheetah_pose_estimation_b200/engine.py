"""Host-side driver of the CUDA hot path: device tables, batches, and the LM / augmented-Lagrangian loop.

Replaces the reference's solver seam `default_solver(...).solve(robot.m)` (acinoset_opt.py:611-617) for the
kinematic reconstruction: the arithmetic lives in libcpe_b200.so (include/cpe.h); this file only owns device
memory (torch tensors), launches, and the few scalar decisions of the outer loop.
"""
import ctypes as C
import os
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib, skeleton

FISHEYE, PINHOLE = 0, 1
MEAS_F32X4, MEAS_F64X4, MEAS_F32X3, MEAS_F32X2G, MEAS_F32X2B = 0, 1, 2, 3, 4      # include/cpe.h CPE_MEAS_*
LOSS_REDESCENDING, LOSS_SQUARE = 0, 1             # include/cpe.h CPE_LOSS_*
STATUS_PIVOT, STATUS_MAXITER, STATUS_NONFINITE = 1, 2, 4      # include/cpe.h CPE_STATUS_* (bit flags; 0 = converged)
_I32_KEYS = ("chain_link", "eq_pairs", "ineq_idx", "ind_idx", "ind_pos", "grp_ndep", "grp_nnb", "grp_dep", "grp_nb",
             "u_off", "u_item", "col_item", "ent_pq", "a_ptr", "a_items", "e_ptr", "e_items", "i_ptr", "i_items", "diag_ent",
             "grad_ent", "grp_src", "s_src", "s_ptr", "s_items", "ent_order", "s_order")
_F64_KEYS = ("ineq_bounds", "wq", "meas_w")
# penalty continuation of the augmented-Lagrangian outer loop: rounds 1-4 as in round 1 (the bench's LM leg runs 4);
# the mirrored API runs 8, whose 1e7 rounds bring the joint equalities to ~1e-6, the level of the shipped IPOPT optima
MU_SCHEDULE = (1e4, 1e5, 1e6, 1e6, 1e7, 1e7, 1e7, 1e7)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Model:
    """Device copy of the static skeleton tables (`cpe_model`)."""

    def __init__(self, device="cuda", kinetic_dataset=False):
        if not torch.cuda.is_available():
            raise _lib.CpeError("CUDA device required: the hot path has no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.kinetic_dataset = kinetic_dataset
        self.tables = skeleton.build_tables(kinetic_dataset)
        self.tables.update(skeleton.assemble_tables(self.tables))
        self.t = {}
        for k in _I32_KEYS:
            self.t[k] = torch.as_tensor(np.ascontiguousarray(self.tables[k], dtype=np.int32), device=self.device)
        for k in _F64_KEYS:
            self.t[k] = torch.as_tensor(np.ascontiguousarray(self.tables[k], dtype=np.float64), device=self.device)
        self.n_ineq = int(self.tables["n_ineq"])
        self.n_groups = int(self.tables["n_groups"])
        s = _lib.CpeModel()
        s.n_ineq, s.n_groups = self.n_ineq, self.n_groups
        s.n_u, s.n_ent = int(self.tables["n_u"]), int(self.tables["n_ent"])
        self.jac_width = 2 * (72 + s.n_u)
        for k in _I32_KEYS + _F64_KEYS:
            setattr(s, k, self.t[k].data_ptr())
        self.t["k2_tab"] = torch.as_tensor(np.ascontiguousarray(self.tables["k2_tab"], dtype=np.uint8), device=self.device)
        s.k2_tab = self.t["k2_tab"].data_ptr()
        for i, v in enumerate(self.tables["k2_off"]):
            s.k2_off[i] = int(v)
        self.struct = s


def _pack_bits(gate):
    """[S, ...] uint8 0 / 1 -> [S, n / 8] uint8, bit (i & 7) of byte i >> 3 = gate i (numpy.packbits, bitorder little)."""
    S = gate.shape[0]
    g = gate.reshape(S, -1, 8).to(torch.int32)
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.int32, device=gate.device)
    return (g * w).sum(dim=2).to(torch.uint8).contiguous()


@dataclass
class Batch:
    """Device-resident inputs of S sequences (`cpe_batch`)."""
    n_frames: torch.Tensor       # [S] int32
    h: torch.Tensor              # [S] f64
    meas: torch.Tensor           # [S,N,C,24,W,4] f32 (or f64): u, v, gate, unused -- or [S,N,C,24,W,3] f32:
                                 # DLC-native x, y, likelihood (gated on the device by likelihood > dlc_thresh)
    cams: torch.Tensor           # [S,C,24] f64
    chain_vec: torch.Tensor      # [S,24,5,3] f64
    cam_model: int = FISHEYE
    lam_eq: torch.Tensor = None  # [S,N,26]
    lam_in: torch.Tensor = None  # [S,N,n_ineq]
    mu_eq: float = 1e4
    mu_in: float = 1e4
    loss: tuple = (3.0, 10.0, 20.0)
    dlc_thresh: float = 0.5
    loss_mode: int = LOSS_REDESCENDING
    active: torch.Tensor = None  # [S] int32, optional: sequences with 0 are skipped by every kernel
    # shutter delay of the measurement model (acinoset_misc.py:273-286); evaluation only, see include/cpe.h
    shutter_tau: torch.Tensor = None   # [S,C] f64 seconds
    base_dq: torch.Tensor = None       # [S,N,3] f64
    base_ddq: torch.Tensor = None      # [S,N,3] f64
    n_cams: torch.Tensor = None        # [S] int32, optional: cameras c >= n_cams[s] are padding (ragged batch)
    meas_gate: torch.Tensor = None     # [S,N,C,24,W] uint8: gates of the compact layout (meas = [..., 2] float32 u, v), or
                                       # [S, N*C*24*W/8] uint8: the same packed 8 per byte (little-endian bit order)

    @property
    def meas_layout(self):
        if self.meas.shape[5] == 2:
            return MEAS_F32X2B if self.meas_gate.dim() == 2 else MEAS_F32X2G
        if self.meas.shape[5] == 3:
            return MEAS_F32X3
        return MEAS_F64X4 if self.meas.dtype == torch.float64 else MEAS_F32X4

    @property
    def S(self):
        return self.meas.shape[0]

    @property
    def N(self):
        return self.meas.shape[1]

    @property
    def C(self):
        return self.meas.shape[2]

    @property
    def W(self):
        return self.meas.shape[4]

    def struct(self):
        for t in (self.n_frames, self.h, self.meas, self.cams, self.chain_vec):
            assert t.is_cuda and t.is_contiguous()
        assert self.n_frames.dtype == torch.int32 and self.h.dtype == torch.float64
        assert self.meas.dtype in (torch.float32, torch.float64) and self.meas.shape[3] == 24
        assert self.meas.shape[5] == 4 or (self.meas.shape[5] in (2, 3) and self.meas.dtype == torch.float32)
        if self.meas.shape[5] == 2:
            g = self.meas_gate
            assert g is not None and g.is_cuda and g.is_contiguous() and g.dtype == torch.uint8
            assert tuple(g.shape) in (tuple(self.meas.shape[:5]), (self.S, self.N * self.C * 24 * self.W // 8))
        assert self.cams.shape == (self.S, self.C, skeleton.CAM_STRIDE) and self.cams.dtype == torch.float64
        assert self.chain_vec.shape == (self.S, 24, 5, 3) and self.chain_vec.dtype == torch.float64
        b = _lib.CpeBatch()
        b.S, b.N, b.C, b.W = self.S, self.N, self.C, self.W
        b.cam_model = int(self.cam_model)
        b.meas_layout = self.meas_layout
        b.loss_mode = int(self.loss_mode)
        b.dlc_thresh = float(self.dlc_thresh)
        b.mu_eq, b.mu_in = float(self.mu_eq), float(self.mu_in)
        b.loss_a, b.loss_b, b.loss_c = self.loss
        b.n_frames, b.h = self.n_frames.data_ptr(), self.h.data_ptr()
        b.meas, b.cams, b.chain_vec = self.meas.data_ptr(), self.cams.data_ptr(), self.chain_vec.data_ptr()
        b.lam_eq = self.lam_eq.data_ptr() if self.lam_eq is not None else None
        b.lam_in = self.lam_in.data_ptr() if self.lam_in is not None else None
        b.active = self.active.data_ptr() if self.active is not None else None
        if self.meas_gate is not None and self.meas.shape[5] == 2:
            b.meas_gate = self.meas_gate.data_ptr()
        if self.n_cams is not None:
            assert self.n_cams.is_cuda and self.n_cams.dtype == torch.int32 and self.n_cams.shape == (self.S,)
            b.n_cams = self.n_cams.data_ptr()
        if self.shutter_tau is not None:
            for t, shp in ((self.shutter_tau, (self.S, self.C)), (self.base_dq, (self.S, self.N, 3)),
                           (self.base_ddq, (self.S, self.N, 3))):
                assert t is not None and t.is_cuda and t.is_contiguous() and t.dtype == torch.float64
                assert tuple(t.shape) == shp
            b.shutter_tau, b.base_dq, b.base_ddq = (self.shutter_tau.data_ptr(), self.base_dq.data_ptr(),
                                                    self.base_ddq.data_ptr())
        return b

    def window(self, s0, s1):
        """Sequences [s0, s1) of this batch as a Batch of views (no copy): the unit the wave loops launch on."""
        return Batch(n_frames=self.n_frames[s0:s1], h=self.h[s0:s1], meas=self.meas[s0:s1], cams=self.cams[s0:s1],
                     chain_vec=self.chain_vec[s0:s1], cam_model=self.cam_model,
                     lam_eq=None if self.lam_eq is None else self.lam_eq[s0:s1],
                     lam_in=None if self.lam_in is None else self.lam_in[s0:s1],
                     mu_eq=self.mu_eq, mu_in=self.mu_in, loss=self.loss, dlc_thresh=self.dlc_thresh,
                     loss_mode=self.loss_mode,
                     active=None if self.active is None else self.active[s0:s1],
                     shutter_tau=None if self.shutter_tau is None else self.shutter_tau[s0:s1],
                     base_dq=None if self.base_dq is None else self.base_dq[s0:s1],
                     base_ddq=None if self.base_ddq is None else self.base_ddq[s0:s1],
                     n_cams=None if self.n_cams is None else self.n_cams[s0:s1],
                     meas_gate=None if self.meas_gate is None else self.meas_gate[s0:s1])

    def compact(self, bits=False):
        """The same batch in the compact layout CPE_MEAS_F32X2G: float2 (u, v) records plus one gate byte each, the
        likelihood test (strict >, acinoset_misc.py:231; NaN -> 0) done here once.  9 bytes per detection; with
        `bits` the gates are packed 8 per byte (CPE_MEAS_F32X2B, 8.125 bytes per detection)."""
        m = self.meas
        if m.shape[5] == 2:
            if bits and self.meas_gate.dim() != 2:
                out = self.window(0, self.S)
                out.meas_gate = _pack_bits(self.meas_gate)
                return out
            assert bits == (self.meas_gate.dim() == 2)
            return self
        if m.dtype != torch.float32:
            raise ValueError("Batch.compact: the compact layouts hold float32 coordinates; this batch keeps its detections in "
                             "double (kinetic dataset) and would lose bits")
        uv = m[..., 0:2]
        fin = ~torch.isnan(uv).any(dim=-1)                     # the device gate of the DLC-native layout (meas_unpack)
        gate = (m[..., 2].double() > float(self.dlc_thresh)) if m.shape[5] == 3 else (m[..., 2] != 0)
        gate = (gate & fin).to(torch.uint8).contiguous()
        uv = torch.where(fin[..., None], uv, torch.zeros_like(uv)).contiguous()
        out = self.window(0, self.S)
        out.meas, out.meas_gate = uv, (_pack_bits(gate) if bits else gate)
        return out

    @staticmethod
    def from_numpy(meas, cams, chain_vec, h, n_frames=None, device="cuda", cam_model=FISHEYE, meas_dtype=torch.float32,
                   dlc_native=False, dlc_thresh=0.5, loss_mode=LOSS_REDESCENDING, n_cams=None):
        """meas[S,N,C,24,W,3|4] (u, v, gate), cams[S,C,24], chain_vec[S,24,5,3], h[S].  With dlc_native=True the
        last axis is the DLC triple (x, y, likelihood) and is kept as float3 records (gated on the device)."""
        meas = np.asarray(meas)
        S, N = meas.shape[:2]
        if dlc_native:
            assert meas.shape[-1] == 3
            meas_dtype = torch.float32
        elif meas.shape[-1] == 3:
            meas = np.concatenate([meas, np.zeros(meas.shape[:-1] + (1,), meas.dtype)], -1)
        if n_frames is None:
            n_frames = np.full(S, N)
        dev = torch.device(device)
        return Batch(
            n_frames=torch.as_tensor(np.asarray(n_frames, np.int32), device=dev),
            h=torch.as_tensor(np.broadcast_to(np.asarray(h, np.float64), (S,)).copy(), device=dev),
            meas=torch.as_tensor(np.ascontiguousarray(meas), device=dev).to(meas_dtype).contiguous(),
            cams=torch.as_tensor(np.ascontiguousarray(cams, dtype=np.float64), device=dev),
            chain_vec=torch.as_tensor(np.ascontiguousarray(chain_vec, dtype=np.float64), device=dev),
            cam_model=cam_model, dlc_thresh=dlc_thresh, loss_mode=loss_mode,
            n_cams=None if n_cams is None else torch.as_tensor(np.asarray(n_cams, np.int32), device=dev))


class Solver:
    """Workspaces + launches for one (model, batch)."""

    def __init__(self, model: Model, batch: Batch, with_solver=True, priors=None, redamp=True):
        """priors: a `priors.Priors` (monocular data-driven terms) or None.
        redamp: after a REJECTED LM step q is unchanged, so K2 only re-damps the stored undamped entries
        (cpe_assemble_out.ent / redamp) instead of assembling the frame again -- bit-identical results; plain multi-view
        solves only (the prior and shutter-delay paths add to g / diagH in place after K2 and keep assembling in full).
        Environment CPE_REDAMP=0 switches it off for A/B timing (profiles/r2w_redamp_ab.txt)."""
        self.m, self.b = model, batch
        self._redamp = bool(redamp) and with_solver and priors is None and os.environ.get("CPE_REDAMP", "1") != "0"
        self._redamp_on = False
        self.lib = model.lib
        self.priors = priors
        S, N, dev = batch.S, batch.N, model.device
        f64 = dict(dtype=torch.float64, device=dev)
        self.frame_cost = torch.zeros(S, N, 4, **f64)
        self.cost = torch.zeros(S, 4, **f64)
        self.launches = 0
        if priors is not None:
            self.prior_frame = torch.zeros(S, N, 2, **f64)
            self.prior_cost = torch.zeros(S, 2, **f64)
            if with_solver:
                # constant band of the motion terms, one per sequence (depends on n_frames and h)
                nfs, hs = batch.n_frames.cpu().numpy(), batch.h.cpu().numpy()
                cache, rows = {}, []
                for nf, h in zip(nfs, hs):
                    key = (int(nf), float(h))
                    if key not in cache:
                        cache[key] = priors.mband(int(nf), float(h), N)
                    rows.append(cache[key])
                self.Mband = torch.as_tensor(np.stack(rows), device=dev)
                self.Lband = torch.empty(S, N * 28, priors.bw + 1, **f64)
        if with_solver:
            G = model.n_groups
            self.Sblk = torch.zeros(S, N, 28, 28, **f64)
            self.ghat = torch.zeros(S, N, 28, **f64)
            self.Y = torch.zeros(S, N, G, 6, 12, **f64)
            self.y = torch.zeros(S, N, G, 6, **f64)
            self.g = torch.zeros(S, N, 54, **f64)
            self.diagH = torch.zeros(S, N, 54, **f64)
            self.Lfac = torch.empty(S, N, 3, 28, 28, **f64)
            self.zbuf = torch.zeros(S, N, 28, **f64)
            self.delta_ind = torch.zeros(S, N, 28, **f64)
            self.status = torch.zeros(S, dtype=torch.int32, device=dev)
            self.delta = torch.zeros(S, N, 54, **f64)
            self.q_try = torch.zeros(S, N, 54, **f64)
            self.frame_pred = torch.zeros(S, N, **f64)
            self.state = torch.zeros(4, S, **f64)
            self.ent = torch.zeros(S, N, 704, **f64) if self._redamp else None       # CPE_K2_ENT_STORE
            self.counters = torch.zeros(S, 2, dtype=torch.int32, device=dev)
            self.active = torch.ones(S, dtype=torch.int32, device=dev)
            self.c_eq = torch.zeros(S, N, 26, **f64)
            self.v_in = torch.zeros(S, N, model.n_ineq, **f64)

    # ---- K1 --------------------------------------------------------------------------------------------------
    def evaluate(self, q, pose=False, resid=False, constraints=False, jac=False, cost_out=None, buffers=None, f32=False,
                 prior_out=None):
        """cost[S,4] = (measurement, model, al_eq, al_in) and optional per-frame outputs.  `buffers` may hold
        preallocated output tensors ("pose", "resid", "jac") to write into instead of allocating.  f32=True runs the
        scaled FP32 flavour (cpe_eval_f32): resid / jac come back as float32."""
        b = self.b
        assert q.is_cuda and q.dtype == torch.float64 and q.is_contiguous() and q.shape == (b.S, b.N, 54)
        out = {}
        buffers = buffers or {}
        eo = _lib.CpeEvalOutF32() if f32 else _lib.CpeEvalOut()
        eo.frame_cost = self.frame_cost.data_ptr()
        f64 = dict(dtype=torch.float64, device=q.device)
        fT = dict(dtype=torch.float32 if f32 else torch.float64, device=q.device)
        if pose:
            out["pose"] = buffers["pose"] if "pose" in buffers else torch.zeros(b.S, b.N, 24, 3, **f64)
            eo.pose = out["pose"].data_ptr()
        if resid:
            out["resid"] = buffers["resid"] if "resid" in buffers else torch.zeros(b.S, b.N, b.C, 24, b.W, 2, **fT)
            assert out["resid"].dtype == fT["dtype"]
            eo.resid = out["resid"].data_ptr()
        if constraints:
            out["c_eq"] = self.c_eq if hasattr(self, "c_eq") else torch.zeros(b.S, b.N, 26, **f64)
            out["v_in"] = self.v_in if hasattr(self, "v_in") else torch.zeros(b.S, b.N, self.m.n_ineq, **f64)
            eo.c_eq, eo.v_in = out["c_eq"].data_ptr(), out["v_in"].data_ptr()
        if jac:
            out["jac"] = buffers["jac"] if "jac" in buffers else torch.zeros(b.S, b.N, b.C, self.m.jac_width, **fT)
            assert out["jac"].dtype == fT["dtype"]
            eo.jac = out["jac"].data_ptr()
        bs = b.struct()
        st = _stream()
        fn, name = (self.lib.cpe_eval_f32, "cpe_eval_f32") if f32 else (self.lib.cpe_eval_f64, "cpe_eval_f64")
        _lib.check(fn(C.byref(self.m.struct), C.byref(bs), _ptr(q), C.byref(eo), st), name)
        cost = self.cost if cost_out is None else cost_out
        _lib.check(self.lib.cpe_reduce_cost_f64(C.byref(bs), _ptr(self.frame_cost), _ptr(cost), st), "cpe_reduce_cost_f64")
        self.launches += 2
        out["cost"] = cost
        if self.priors is not None:
            out["prior_cost"] = self.evaluate_priors(q, prior_out)
        return out

    def evaluate_priors(self, q, prior_out=None):
        """prior_cost[S,2] = (pose, motion) of the monocular priors at q."""
        pc = self.prior_cost if prior_out is None else prior_out
        bs = self.b.struct()
        _lib.check(self.lib.cpe_prior_cost_f64(C.byref(self.m.struct), C.byref(self.priors.struct), C.byref(bs), _ptr(q),
                                               _ptr(self.prior_frame), _ptr(pc), _stream()), "cpe_prior_cost_f64")
        self.launches += 2
        return pc

    def evaluate_launch(self, q, buffers, cost_out, record_mid=None):
        """Lean K1 launch for wave loops: residuals + chain-sparse Jacobian into preallocated `buffers`
        ("resid", "jac"), then the per-sequence cost reduction into cost_out[S,4].  `record_mid` (a CUDA event)
        is recorded between the two kernels so that the K1 launch can be timed on its own."""
        f32 = any(t.dtype == torch.float32 for t in buffers.values())
        assert all((t.dtype == torch.float32) == f32 for t in buffers.values())
        eo = _lib.CpeEvalOutF32() if f32 else _lib.CpeEvalOut()
        eo.frame_cost = self.frame_cost.data_ptr()
        if "resid" in buffers:
            eo.resid = buffers["resid"].data_ptr()
        if "jac" in buffers:
            eo.jac = buffers["jac"].data_ptr()
        bs = self.b.struct()
        st = _stream()
        fn, name = (self.lib.cpe_eval_f32, "cpe_eval_f32") if f32 else (self.lib.cpe_eval_f64, "cpe_eval_f64")
        _lib.check(fn(C.byref(self.m.struct), C.byref(bs), _ptr(q), C.byref(eo), st), name)
        if record_mid is not None:
            record_mid.record()
        _lib.check(self.lib.cpe_reduce_cost_f64(C.byref(bs), _ptr(self.frame_cost), _ptr(cost_out), st),
                   "cpe_reduce_cost_f64")
        self.launches += 2

    # ---- K2 --------------------------------------------------------------------------------------------------
    def _asm_struct(self, raw=None, reduced=True):
        ao = _lib.CpeAssembleOut()
        if reduced:
            ao.S, ao.ghat, ao.Y, ao.y = (self.Sblk.data_ptr(), self.ghat.data_ptr(), self.Y.data_ptr(), self.y.data_ptr())
        ao.g, ao.diagH = self.g.data_ptr(), self.diagH.data_ptr()
        if hasattr(self, "cam_blocks"):
            ao.cam_blocks = self.cam_blocks.data_ptr()
        if raw is not None:
            ao.Hraw = raw.data_ptr()
        if reduced and self._redamp:
            ao.ent = self.ent.data_ptr()
            if self._redamp_on:                      # inside solve_steps: row 3 of `state` is -1 where the last step was rejected
                ao.redamp = self.state[3].data_ptr()
        return ao

    def assemble(self, q, damp, raw=False, reduced=True):
        b = self.b
        assert q.is_contiguous() and damp.is_contiguous() and damp.dtype == torch.float64
        Hraw = torch.zeros(b.S, b.N, 54, 54, dtype=torch.float64, device=q.device) if raw else None
        ao = self._asm_struct(Hraw, reduced)
        bs = b.struct()
        _lib.check(self.lib.cpe_assemble_f64(C.byref(self.m.struct), C.byref(bs), _ptr(q), _ptr(damp), C.byref(ao), _stream()),
                   "cpe_assemble_f64")
        self.launches += 1
        if self.priors is not None and reduced:
            _lib.check(self.lib.cpe_prior_assemble_f64(C.byref(self.m.struct), C.byref(self.priors.struct), C.byref(bs),
                                                       _ptr(q), _ptr(damp), C.byref(ao), _stream()),
                       "cpe_prior_assemble_f64")
            self.launches += 1
        return Hraw

    # ---- K3 --------------------------------------------------------------------------------------------------
    def band_solve(self):
        hp = getattr(self, "k3_stream", None)
        if hp is None:
            return self._band_solve()
        # K3 on a high-priority stream: its few, long-lived CTAs get SM slots ahead of the queued K2 CTAs of another
        # sub-batch (solve_split), which is what lets the latency-bound sweep overlap the throughput-bound assembly
        cur = torch.cuda.current_stream()
        hp.wait_stream(cur)
        with torch.cuda.stream(hp):
            out = self._band_solve()
        cur.wait_stream(hp)
        return out

    def _band_solve(self):
        if self.priors is not None:
            bs = self.b.struct()
            _lib.check(self.lib.cpe_band_solve_general_f64(
                C.byref(bs), _ptr(self.Sblk), _ptr(self.ghat), _ptr(self.Mband), C.c_int64(self.Mband.stride(0)),
                C.c_int32(self.priors.bw), _ptr(self.Lband), _ptr(self.delta_ind), _ptr(self.status), _stream()),
                "cpe_band_solve_general_f64")
            self.launches += 1
            return self.delta_ind
        io = _lib.CpeSolveIO()
        io.Lfac, io.zbuf, io.delta_ind, io.status = (self.Lfac.data_ptr(), self.zbuf.data_ptr(), self.delta_ind.data_ptr(),
                                                     self.status.data_ptr())
        bs = self.b.struct()
        _lib.check(self.lib.cpe_band_solve_f64(C.byref(self.m.struct), C.byref(bs), _ptr(self.Sblk), _ptr(self.ghat),
                                               C.byref(io), _stream()), "cpe_band_solve_f64")
        self.launches += 1
        return self.delta_ind

    # ---- K4 --------------------------------------------------------------------------------------------------
    def step_apply(self, q, damp):
        ao = self._asm_struct()
        bs = self.b.struct()
        _lib.check(self.lib.cpe_step_apply_f64(C.byref(self.m.struct), C.byref(bs), _ptr(q), _ptr(damp), C.byref(ao),
                                               _ptr(self.delta_ind), _ptr(self.delta), _ptr(self.q_try),
                                               _ptr(self.frame_pred), _stream()), "cpe_step_apply_f64")
        self.launches += 1

    def lm_step(self, q, damp):
        """One trial step at fixed damping: returns (delta[S,N,54], q_try, frame_pred[S,N])."""
        self.assemble(q, damp)
        self.band_solve()
        self.step_apply(q, damp)
        return self.delta, self.q_try, self.frame_pred

    # ---- full LM / AL loop -----------------------------------------------------------------------------------------
    def solve(self, q0, **kw):
        """Minimise the reference objective over q starting from q0[S,N,54]; returns a dict of device tensors.

        Outer loop = augmented-Lagrangian multiplier updates with penalty continuation; inner loop = LM with
        per-sequence damping (all sequences advance in lock step; converged ones are masked and skipped).
        Keywords: max_iter=40, n_outer=4, damp0=1e-3, ftol=1e-6, mu_schedule, check_every=4, skip_converged=True,
        graph=False (True: one CUDA graph per multiplier round replays the iteration's launches -- for small batches,
        where the loop is launch-bound; bit-identical results).
        The host looks at the convergence flags (one device synchronisation) every `check_every` iterations only; the
        iterations queued after every sequence has stopped are no-ops on the device, so results do not depend on it."""
        for _ in self.solve_steps(q0, **kw):
            pass
        return self.result

    def solve_steps(self, q0, max_iter=40, n_outer=4, damp0=1e-3, ftol=1e-6, mu_schedule=MU_SCHEDULE,
                    check_every=4, skip_converged=True, graph=False, extra_outer=0, feas_tol=(5e-5, 2.5e-4)):
        """Generator form of `solve`: yields after the launches of every LM iteration have been queued (before the
        host looks at the convergence flags), so that a driver can interleave several solvers on different streams
        (`solve_split`).  The result dict is left in `self.result`.

        extra_outer > 0: after the n_outer scheduled rounds, up to that many further rounds at the last penalty for the
        sequences whose joint equalities / inequality rows still exceed feas_tol = (|c_eq|, |v_in|) -- IPOPT keeps going
        until its constraint tolerance is met; a fixed round count does not.  Every other sequence is inactive in those
        rounds (no step, no multiplier update), so its result is bit-identical to a solve without them."""
        b, lib = self.b, self.lib
        S, N, dev = b.S, b.N, q0.device
        if not hasattr(self, "_q_work"):
            # work buffers of the loop, allocated once per solver: their addresses stay the same from solve to solve, which
            # is what lets a captured iteration (graph=True) be replayed by later solves
            f64 = dict(dtype=torch.float64, device=dev)
            self._q_work = torch.empty(S, N, 54, **f64)
            self._lam_eq = torch.empty(S, N, 26, **f64)
            self._lam_in = torch.empty(S, N, self.m.n_ineq, **f64)
            self._cost_try = torch.zeros(S, 4, **f64)
            self._prior_try = torch.zeros(S, 2, **f64) if self.priors is not None else None
            self._graphs = {}
        q = self._q_work
        q.copy_(q0)
        b.lam_eq, b.lam_in = self._lam_eq.zero_(), self._lam_in.zero_()
        self.state.zero_()
        self.state[1].fill_(damp0)
        self.state[2].fill_(2.0)
        self.counters.zero_()
        cost_try, prior_try = self._cost_try, self._prior_try
        damp = self.state[1]
        total_inner = 0
        ftols = tuple(ftol) if isinstance(ftol, (tuple, list)) else (ftol,)
        need = held = None
        extra_rounds = 0
        for outer in range(n_outer + extra_outer):
            if outer >= n_outer:
                # self.c_eq / self.v_in hold the constraints at q (evaluated when the previous round ended)
                need = ((self.c_eq.abs().amax(dim=(1, 2)) >= feas_tol[0])
                        | (self.v_in.abs().amax(dim=(1, 2)) >= feas_tol[1])).to(torch.int32)
                if not bool(need.any()):
                    need = None
                    break
                if held is None:
                    held = self.active.clone()                  # MAXITER flags of the last scheduled round
                    ever = torch.zeros_like(need)
                ever |= need
                extra_rounds += 1
            b.mu_eq = b.mu_in = float(mu_schedule[min(outer, len(mu_schedule) - 1)])
            ftol = float(ftols[min(outer, len(ftols) - 1)])      # inexact inner solves while the multipliers are rough
            self.evaluate(q)
            self.state[0].copy_(self.cost.sum(dim=1))
            self.state[3].zero_()                    # new penalty / multipliers: the first assembly of a round is a full one
            self._redamp_on = self._redamp
            if self.priors is not None:
                self.state[0].add_(self.prior_cost.sum(dim=1))
            if need is None:
                self.active.fill_(1)
            else:
                self.active.copy_(need)
            masked = skip_converged or need is not None
            def iteration():
                b.active = self.active if masked else None      # converged sequences cost nothing from here on
                self.lm_step(q, damp)
                self.evaluate(self.q_try, cost_out=cost_try, prior_out=prior_try)
                b.active = None
                bs = b.struct()
                if self.priors is not None:
                    _lib.check(lib.cpe_step_accept_prior_f64(C.byref(bs), _ptr(cost_try), _ptr(prior_try),
                                                             _ptr(self.frame_pred), _ptr(q), _ptr(self.q_try),
                                                             _ptr(self.state), _ptr(self.counters), _ptr(self.active),
                                                             C.c_double(ftol), _stream()), "cpe_step_accept_prior_f64")
                else:
                    _lib.check(lib.cpe_step_accept_f64(C.byref(bs), _ptr(cost_try), _ptr(self.frame_pred), _ptr(q),
                                                       _ptr(self.q_try), _ptr(self.state), _ptr(self.counters),
                                                       _ptr(self.active), C.c_double(ftol), _stream()),
                               "cpe_step_accept_f64")
                self.launches += 1

            # graph=True: the launches of ONE iteration (K2, K3, K4, K1, reduce, K5 -- six to ten kernels whose arguments
            # do not change while the penalty stays the same: every step size, damping and convergence flag lives in
            # device memory) are captured into a CUDA graph and replayed; the very first iteration of a solver runs
            # eagerly so that no module is loaded under capture.  Same kernels, same order, same arguments: results are
            # bit-identical to the eager loop.
            use_graph = graph and getattr(self, "k3_stream", None) is None
            # every scalar and pointer the captured launches bake in: the batch struct itself (penalties, loss, pointers)
            key = (bytes(b.struct()), ftol, bool(masked)) if use_graph else None
            for it in range(max_iter):
                if not use_graph or not getattr(self, "_graph_warm", False):
                    iteration()
                    self._graph_warm = True
                else:
                    if key not in self._graphs:
                        # captured once per (penalty, tolerance) of this solver and kept: later rounds with the same
                        # penalty and later solves replay it.  Plain capture_begin / capture_end on a side stream
                        # (torch.cuda.graph() would add a garbage collection and a device synchronisation per capture)
                        l0 = self.launches
                        g = torch.cuda.CUDAGraph()
                        cur, side = torch.cuda.current_stream(), self._capture_stream()
                        side.wait_stream(cur)
                        with torch.cuda.stream(side):
                            g.capture_begin()
                            try:
                                iteration()
                            finally:
                                g.capture_end()
                        cur.wait_stream(side)
                        self._graphs[key] = (g, self.launches - l0)
                        self.launches = l0
                    g, per_iter = self._graphs[key]
                    g.replay()
                    self.launches += per_iter
                total_inner += 1
                yield total_inner
                if (it + 1) % check_every == 0 and not bool(self.active.any()):
                    break
            self.evaluate(q, constraints=True)
            b.active = need                                     # extra rounds: only the sequences they ran for
            bs = b.struct()
            b.active = None
            _lib.check(lib.cpe_update_multipliers_f64(C.byref(self.m.struct), C.byref(bs), _ptr(self.c_eq), _ptr(self.v_in),
                                                      _ptr(b.lam_eq), _ptr(b.lam_in), _stream()),
                       "cpe_update_multipliers_f64")
            self.launches += 1
        self._redamp_on = False
        if held is not None:
            # sequences that sat out the extra rounds keep the activity (= MAXITER) flag of their own last round
            self.active.copy_(torch.where(ever != 0, self.active, held))
        self.extra_rounds = extra_rounds
        out = self.evaluate(q, constraints=True)
        res = dict(q=q.clone(), cost=out["cost"].clone(), c_eq=out["c_eq"].clone(), v_in=out["v_in"].clone(),
                   counters=self.counters.clone(), status=self.exit_status(out), lockstep_iterations=total_inner)
        if self.priors is not None:
            res["prior_cost"] = out["prior_cost"].clone()
        self.result = res

    def _capture_stream(self):
        if not hasattr(self, "_cap_stream"):
            self._cap_stream = torch.cuda.Stream(device=self.m.device)
        return self._cap_stream

    def exit_status(self, out):
        """status[S] of CPE_STATUS_* flags from the final cost, the activity flags of the last round and K3's pivots."""
        st = torch.zeros(self.b.S, dtype=torch.int32, device=self.cost.device)
        bs = self.b.struct()
        _lib.check(self.lib.cpe_finalize_status_f64(C.byref(bs), _ptr(out["cost"]), _ptr(out.get("prior_cost")),
                                                    _ptr(self.active), _ptr(self.status), _ptr(st), _stream()),
                   "cpe_finalize_status_f64")
        self.launches += 1
        return st


    # ---- shutter-delay estimation (acinoset_misc.py:182-183, 273-286) ------------------------------------
    def set_delay(self, q, tau):
        """Point the batch at the delay tau[S,C] and at the base velocity / acceleration of q (cpe_base_rates_f64)."""
        b = self.b
        if not hasattr(self, "_dqb"):
            self._dqb = torch.zeros(b.S, b.N, 3, dtype=torch.float64, device=q.device)
            self._ddqb = torch.zeros_like(self._dqb)
        b.shutter_tau, b.base_dq, b.base_ddq = tau.contiguous(), self._dqb, self._ddqb
        bs = b.struct()
        _lib.check(self.lib.cpe_base_rates_f64(C.byref(bs), _ptr(q), _ptr(self._dqb), _ptr(self._ddqb), _stream()),
                   "cpe_base_rates_f64")
        self.launches += 1

    def delay_terms(self, q, tau):
        """From the per-camera blocks of the last `assemble` at (q, tau): the part of d cost / d q that goes through
        dq_base and ddq_base (g_extra[S,N,3], on the base x, y, z of frames n, n-1, n-2), d cost / d tau [S,C], and the
        Gauss-Newton curvature of tau [S,C] (cameras do not share residuals, so it is diagonal)."""
        b = self.b
        assert b.shutter_tau is not None and q.is_contiguous()
        if not hasattr(self, "_g_extra"):
            self._g_extra = torch.zeros(b.S, b.N, 3, dtype=torch.float64, device=q.device)
            self._g_tau = torch.zeros(b.S, b.C, dtype=torch.float64, device=q.device)
            self._curv_tau = torch.zeros_like(self._g_tau)
        bs = b.struct()
        _lib.check(self.lib.cpe_shutter_terms_f64(C.byref(bs), _ptr(self.cam_blocks), _ptr(self._g_extra),
                                                  _ptr(self._g_tau), _ptr(self._curv_tau), _stream()),
                   "cpe_shutter_terms_f64")
        self.launches += 2
        return self._g_extra, self._g_tau, self._curv_tau

    def assemble_delayed(self, q, tau, damp, reduced=True):
        """`assemble` at (q, tau) with the exact gradient: K2 gives d cost / d q_n at fixed dq, ddq; the terms through
        dq_base / ddq_base are added to the gradient (they only touch base x, y, z, which are not eliminated, so the
        reduced gradient takes them unchanged).  Returns (g_tau, curv_tau)."""
        self.set_delay(q, tau)
        if not hasattr(self, "cam_blocks"):
            self.cam_blocks = torch.zeros(self.b.S, self.b.N, self.b.C, 9, dtype=torch.float64, device=q.device)
            self._xyz_ind = torch.as_tensor(np.asarray(self.m.tables["ind_pos"][0:3], np.int64), device=q.device)
            assert int(self._xyz_ind.min()) >= 0
        self.assemble(q, damp, reduced=reduced)
        g_extra, g_tau, curv = self.delay_terms(q, tau)
        self.g[..., 0:3] += g_extra
        if reduced:
            self.ghat[..., self._xyz_ind] += g_extra
        return g_tau, curv

    def solve_shutter(self, q0, tau0=None, max_iter=40, n_outer=4, damp0=1e-3, ftol=1e-7,
                      mu_schedule=MU_SCHEDULE, fix_first=True):
        """Multi-view reconstruction WITH shutter-delay estimation (`shutter_delay_estimation=True`): minimises the same
        objective as `solve` over q and the per-camera delays tau[S,C] (camera 1 fixed at 0, |tau| <= 1/fps, as
        acinoset_misc.py:183, 274-275).  Each LM iteration takes ONE joint step from one assembly at (q, tau): the
        band solve for q with the exact gradient (assemble_delayed) and a damped scalar Newton step per camera for tau,
        accepted or rejected together on the true cost.  The Hessian is block diagonal in (q, tau) and frame local in q
        (the coupling of frame n to n-1, n-2 through dq, ddq is in the gradient only): the stationary point is exact,
        the convergence rate degrades as |tau| approaches 1/fps.  Returns the dict of `solve` plus `tau`."""
        b, lib = self.b, self.lib
        assert self.priors is None
        S, N, Cn, dev = b.S, b.N, b.C, q0.device
        q = q0.clone().contiguous()
        tau = torch.zeros(S, Cn, dtype=torch.float64, device=dev) if tau0 is None else tau0.clone()
        lim = b.h[:, None].expand(S, Cn)
        b.lam_eq = torch.zeros(S, N, 26, dtype=torch.float64, device=dev)
        b.lam_in = torch.zeros(S, N, self.m.n_ineq, dtype=torch.float64, device=dev)
        b.active = None
        self.state.zero_()
        self.state[1].fill_(damp0)
        self.state[2].fill_(2.0)
        self.counters.zero_()
        cost_try = torch.zeros(S, 4, dtype=torch.float64, device=dev)
        damp = self.state[1]
        total_inner = 0
        for outer in range(n_outer):
            b.mu_eq = b.mu_in = float(mu_schedule[min(outer, len(mu_schedule) - 1)])
            self.set_delay(q, tau)
            self.evaluate(q)
            self.state[0].copy_(self.cost.sum(dim=1))
            self.active.fill_(1)
            for it in range(max_iter):
                g_tau, curv = self.assemble_delayed(q, tau, damp)
                self.band_solve()
                self.step_apply(q, damp)
                step = -g_tau / (curv * (1.0 + damp[:, None]) + 1e-300)
                step = torch.where(curv > 0, step, torch.zeros_like(step))
                tau_try = torch.minimum(torch.maximum(tau + step, -lim), lim)
                if fix_first:
                    tau_try[:, 0] = tau[:, 0]
                self.set_delay(self.q_try, tau_try)
                self.evaluate(self.q_try, cost_out=cost_try)
                n_acc = self.counters[:, 1].clone()
                bs = b.struct()
                _lib.check(lib.cpe_step_accept_f64(C.byref(bs), _ptr(cost_try), _ptr(self.frame_pred), _ptr(q),
                                                   _ptr(self.q_try), _ptr(self.state), _ptr(self.counters),
                                                   _ptr(self.active), C.c_double(ftol), _stream()), "cpe_step_accept_f64")
                self.launches += 1
                accepted = self.counters[:, 1] > n_acc
                tau = torch.where(accepted[:, None], tau_try, tau)
                total_inner += 1
                if not bool(self.active.any()):
                    break
            self.set_delay(q, tau)
            self.evaluate(q, constraints=True)
            bs = b.struct()
            _lib.check(lib.cpe_update_multipliers_f64(C.byref(self.m.struct), C.byref(bs), _ptr(self.c_eq), _ptr(self.v_in),
                                                      _ptr(b.lam_eq), _ptr(b.lam_in), _stream()),
                       "cpe_update_multipliers_f64")
            self.launches += 1
        self.set_delay(q, tau)
        out = self.evaluate(q, constraints=True)
        self.result = dict(q=q, tau=tau, cost=out["cost"].clone(), c_eq=out["c_eq"], v_in=out["v_in"],
                           counters=self.counters.clone(), status=self.exit_status(out), lockstep_iterations=total_inner)
        return self.result


def base_rates(bxyz, h, n_frames=None):
    """dq_base, ddq_base [S,N,3] of the q-only formulation: the reference's implicit-Euler collocation
    (collocation.py:58-68: q_n = q_{n-1} + h dq_n, dq_n = dq_{n-1} + h ddq_n) eliminated, i.e. backward differences of
    the base x, y, z.  The reference leaves dq and ddq of the first node free (hence ddq of the second); with them
    eliminated nothing is left to define them, so they take the values that zero their motion cost, as in the saved
    dq / ddq: ddq_0 = ddq_1 = ddq_2 and dq_0 = dq_1 - h ddq_1 (0-based).  A deviation from the reference
    NLP confined to the first two frames, stated in DESIGN.md.  Plain differentiable torch ops on [S,N,3]."""
    S, N, _ = bxyz.shape
    hh = h[:, None, None]
    if N < 2:
        z = torch.zeros_like(bxyz)
        return z, z
    v1 = (bxyz[:, 1:] - bxyz[:, :-1]) / hh                                   # frames 1..N-1
    if N < 3:
        return torch.cat([v1[:, :1], v1], dim=1), torch.zeros_like(bxyz)
    a2 = (v1[:, 1:] - v1[:, :-1]) / hh                                       # frames 2..N-1
    a = torch.cat([a2[:, :1], a2[:, :1], a2], dim=1)
    if n_frames is not None:                                                 # ragged batch: sequences shorter than 3 / 2
        a = a * (n_frames >= 3).to(a.dtype)[:, None, None]                   # frames have no acceleration / velocity
        v1 = v1 * (n_frames >= 2).to(a.dtype)[:, None, None]
    v = torch.cat([v1[:, :1] - hh * a[:, 1:2], v1], dim=1)
    return v, a


def shutter_jacobian(jac, batch):
    """d slack / d tau_c, [S,N,C,24,2], from the packed Jacobian of `Solver.evaluate(jac=True)` on a batch with a
    shutter delay: the base x, y, z columns of a marker ARE d(u,v)/d(point), so
    d slack / d tau_c = d(u,v)/d(point) . (dq_base + 2 tau_c ddq_base)   (derivative of acinoset_misc.py:283-286).
    Likewise d slack / d dq_base = tau_c d(u,v)/d(point) and d slack / d ddq_base = tau_c^2 d(u,v)/d(point).
    Device-side torch indexing of the kernel's output; no kernel of its own."""
    u_off, _, _ = skeleton.jacobian_pattern()
    start = torch.as_tensor([2 * (3 * l + int(u_off[l])) for l in range(24)], device=jac.device)
    idx = start[:, None] + torch.arange(6, device=jac.device)[None, :]                     # [24,6]
    P = jac[..., idx].reshape(*jac.shape[:-1], 24, 3, 2).to(torch.float64)                 # [S,N,C,24,xyz,uv]
    tau = batch.shutter_tau[:, None, :, None]                                              # [S,1,C,1]
    d = batch.base_dq[:, :, None, :] + 2.0 * tau * batch.base_ddq[:, :, None, :]           # [S,N,C,3]
    return torch.einsum("snclid,snci->sncld", P, d)


_SPLIT_STREAMS = {}


def solve_split(model, batch, q0, n_split=2, priors=None, **kw):
    """LM-solve `batch` as `n_split` independent sub-batches on their own CUDA streams, launches interleaved from one
    host thread.  The block-banded solve K3 is a latency-bound sweep along time (one CTA per sequence, ~10 % issue
    utilisation) while the assembly K2 is throughput-bound, so K3 of one sub-batch overlaps K2 of the other.  Per-sequence
    results are bit-identical to `Solver.solve` (no cross-sequence arithmetic anywhere).  Returns the same dict."""
    S = batch.S
    n_split = max(1, min(n_split, S))
    cuts = [S * i // n_split for i in range(n_split + 1)]
    cur = torch.cuda.current_stream()
    # the same stream objects from call to call: torch's caching allocator keeps its free blocks PER STREAM, so fresh streams
    # would make every call cudaMalloc its ~40 GB of workspaces anew (measured: 23 cudaMalloc calls / 32 GB inside a solve)
    key = (str(model.device), n_split)
    if key not in _SPLIT_STREAMS:
        _SPLIT_STREAMS[key] = ([torch.cuda.Stream(device=model.device) for _ in range(n_split)],
                               [torch.cuda.Stream(device=model.device, priority=-1) for _ in range(n_split)])
    streams, k3_streams = _SPLIT_STREAMS[key]
    solvers, gens = [], []
    for i, st in enumerate(streams):
        st.wait_stream(cur)
        with torch.cuda.stream(st):
            sv = Solver(model, batch.window(cuts[i], cuts[i + 1]), priors=priors)
            if n_split > 1:
                sv.k3_stream = k3_streams[i]
            solvers.append(sv)
            gens.append(sv.solve_steps(q0[cuts[i]:cuts[i + 1]].contiguous(), **kw))
    live = list(range(n_split))
    while live:
        for i in list(live):
            with torch.cuda.stream(streams[i]):
                try:
                    next(gens[i])
                except StopIteration:
                    live.remove(i)
    for st in streams:
        cur.wait_stream(st)
    res = {}
    for k in solvers[0].result:
        vals = [sv.result[k] for sv in solvers]
        res[k] = max(vals) if k == "lockstep_iterations" else torch.cat(vals, dim=0)
    res["launches"] = sum(sv.launches for sv in solvers)
    return res


class HostEvaluator:
    """Residual + Jacobian evaluation for callers that keep their data in HOST memory (the reference-facing call:
    what the L4 shim does once per LM iteration when q lives in numpy).  Sequences are streamed through the GPU
    in waves of `wave` sequences on `n_streams` (default 4) streams with as many staging sets, so the host->device copies
    of the next waves overlap the kernel of wave i (measured, detections resident: 471 / 557 / 613 M frame.views/s with
    2 / 3 / 4 sets; the cold call is PCIe-bound and does not care).  Per wave it calls `cpe_eval_host_f64` (include/cpe.h): H2D of q and the
    detections, K1 with the chain-sparse Jacobian written to a device buffer, the per-sequence cost reduction,
    and D2H of cost[S,4].  Modes:
      resident_meas=True  the detections stay on the device (they are constant across the iterations of a solve) and
                          only q crosses the bus -- what one LM iteration of a host-side caller moves;
      resid_out=True      the residuals (`meas_err`, the array the reference saves: acinoset_opt.py:318) come back too."""

    def __init__(self, model: Model, proto: Batch, S, wave, jac=True, resident_meas=False, resid_out=False, n_streams=None):
        self.m, self.lib, self.proto = model, model.lib, proto
        self.S, self.wave, self.N, self.C, self.W = S, wave, proto.N, proto.C, proto.W
        self.resident_meas, self.resid_out = resident_meas, resid_out
        self.compact = proto.meas.shape[5] == 2          # CPE_MEAS_F32X2G / B: float2 (u, v) + gate byte / bit (cpe_eval_host_uvg_f64)
        self.bits = self.compact and proto.meas_gate.dim() == 2
        self.gate_bytes = (self.N * self.C * 24 * self.W // (8 if self.bits else 1)) if self.compact else 0   # per sequence
        dev = model.device
        f64 = dict(dtype=torch.float64, device=dev)
        # staging sets = streams: with two, the copy of wave i + 2 waits for the kernel of wave i (same stream); a third set
        # lets the copy engine run ahead, which matters when the kernel, not the copy, is the longer leg (resident_meas)
        self.n_streams = int(n_streams or os.environ.get("CPE_HOST_STREAMS", 4))
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(self.n_streams)]
        self.stage = []
        for _ in range(self.n_streams):
            st = dict(q=torch.empty(wave, self.N, 54, **f64),
                      meas=None if resident_meas else torch.empty(wave, self.N, self.C, 24, self.W, proto.meas.shape[5],
                                                                  dtype=proto.meas.dtype, device=dev),
                      gate=torch.empty(wave, self.gate_bytes, dtype=torch.uint8, device=dev)
                      if (self.compact and not resident_meas) else None,
                      frame_cost=torch.zeros(wave, self.N, 4, **f64), cost=torch.zeros(wave, 4, **f64),
                      jac=torch.empty(wave, self.N, self.C, model.jac_width, **f64) if jac else None,
                      resid=torch.empty(wave, self.N, self.C, 24, self.W, 2, **f64) if resid_out else None)
            self.stage.append(st)
        self.h_cost = torch.zeros(S, 4, dtype=torch.float64).pin_memory()
        self.h_resid = torch.zeros(S, self.N, self.C, 24, self.W, 2, dtype=torch.float64).pin_memory() if resid_out else None
        self.launches = 0

    def bytes_per_call(self):
        mb = 0 if self.resident_meas else self.C * 24 * self.W * self.proto.meas.shape[5] * self.proto.meas.element_size()
        h2d = self.S * self.N * (54 * 8 + mb) + (0 if self.resident_meas else self.S * self.gate_bytes)
        d2h = self.S * 4 * 8 + (self.S * self.N * self.C * 24 * self.W * 2 * 8 if self.resid_out else 0)
        return h2d, d2h

    def __call__(self, h_q, h_meas=None, h_gate=None):
        """h_q[S,N,54] f64 (and, unless resident_meas, h_meas[S,N,C,24,W,4|3|2] -- with h_gate[S,N,C,24,W] uint8, or its
        bits [S, N*C*24*W/8], for the compact layouts) pinned host tensors -> pinned cost[S,4] (valid after return; `self.h_resid` holds the residuals
        with resid_out)."""
        assert not h_q.is_cuda and h_q.is_pinned()
        assert self.resident_meas == (h_meas is None)
        if h_meas is not None:
            assert not h_meas.is_cuda and h_meas.is_pinned()
            assert (h_gate is not None and h_gate.is_pinned() and h_gate.dtype == torch.uint8) == self.compact
            assert not self.compact or h_gate[0].numel() == self.gate_bytes
        cur = torch.cuda.current_stream()
        for st in self.streams:
            st.wait_stream(cur)
        for i, s0 in enumerate(range(0, self.S, self.wave)):
            s1 = min(self.S, s0 + self.wave)
            stg, stream = self.stage[i % self.n_streams], self.streams[i % self.n_streams]
            bw = self.proto.window(s0, s1)
            bs = bw.struct()
            eo = _lib.CpeEvalOut()
            eo.frame_cost = stg["frame_cost"].data_ptr()
            if stg["jac"] is not None:
                eo.jac = stg["jac"].data_ptr()
            if stg["resid"] is not None:
                eo.resid = stg["resid"].data_ptr()
            hr = None if self.h_resid is None else C.c_void_p(self.h_resid[s0:s1].data_ptr())
            if self.compact:
                _lib.check(self.lib.cpe_eval_host_uvg_f64(
                    C.byref(self.m.struct), C.byref(bs), C.c_void_p(h_q[s0:s1].data_ptr()),
                    None if h_meas is None else C.c_void_p(h_meas[s0:s1].data_ptr()),
                    None if h_gate is None else C.c_void_p(h_gate[s0:s1].data_ptr()), _ptr(stg["q"]), _ptr(stg["meas"]),
                    _ptr(stg["gate"]), C.byref(eo), _ptr(stg["cost"]), C.c_void_p(self.h_cost[s0:s1].data_ptr()), hr,
                    C.c_void_p(stream.cuda_stream)), "cpe_eval_host_uvg_f64")
            else:
                _lib.check(self.lib.cpe_eval_host_f64(
                    C.byref(self.m.struct), C.byref(bs), C.c_void_p(h_q[s0:s1].data_ptr()),
                    None if h_meas is None else C.c_void_p(h_meas[s0:s1].data_ptr()), _ptr(stg["q"]), _ptr(stg["meas"]),
                    C.byref(eo), _ptr(stg["cost"]), C.c_void_p(self.h_cost[s0:s1].data_ptr()), hr,
                    C.c_void_p(stream.cuda_stream)), "cpe_eval_host_f64")
            self.launches += 2
        for st in self.streams:
            cur.wait_stream(st)
        cur.synchronize()
        return self.h_cost
