"""Host-side driver of the CUDA hot path: device tables, batches, and the LM / augmented-Lagrangian loop.

Replaces the reference's solver seam `default_solver(...).solve(robot.m)` (acinoset_opt.py:611-617) for the
kinematic reconstruction: the arithmetic lives in libcpe_b200.so (include/cpe.h); this file only owns device
memory (torch tensors), launches, and the few scalar decisions of the outer loop.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib, skeleton

FISHEYE, PINHOLE = 0, 1
_I32_KEYS = ("chain_link", "slot_of", "eq_pairs", "ineq_idx", "ind_idx", "ind_pos", "blk_nodes", "blk_id",
             "blk_item_ptr", "blk_items", "blk_eq_ptr", "blk_eq_items", "grp_ndep", "grp_nnb", "grp_dep", "grp_nb",
             "grp_nbpos")
_F64_KEYS = ("ineq_bounds", "wq", "meas_w")


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
        self.t = {}
        for k in _I32_KEYS:
            self.t[k] = torch.as_tensor(np.ascontiguousarray(self.tables[k], dtype=np.int32), device=self.device)
        for k in _F64_KEYS:
            self.t[k] = torch.as_tensor(np.ascontiguousarray(self.tables[k], dtype=np.float64), device=self.device)
        self.n_ineq = int(self.tables["n_ineq"])
        self.n_groups = int(self.tables["n_groups"])
        s = _lib.CpeModel()
        s.n_ineq, s.n_blocks = self.n_ineq, int(self.tables["n_blocks"])
        s.n_groups, s.n_blk_items = self.n_groups, int(self.tables["n_blk_items"])
        for k in _I32_KEYS + _F64_KEYS:
            setattr(s, k, self.t[k].data_ptr())
        self.struct = s


@dataclass
class Batch:
    """Device-resident inputs of S sequences (`cpe_batch`)."""
    n_frames: torch.Tensor       # [S] int32
    h: torch.Tensor              # [S] f64
    meas: torch.Tensor           # [S,N,C,24,W,4] f32 (or f64): u, v, gate, unused
    cams: torch.Tensor           # [S,C,24] f64
    chain_vec: torch.Tensor      # [S,24,5,3] f64
    cam_model: int = FISHEYE
    lam_eq: torch.Tensor = None  # [S,N,26]
    lam_in: torch.Tensor = None  # [S,N,n_ineq]
    mu_eq: float = 1e4
    mu_in: float = 1e4
    loss: tuple = (3.0, 10.0, 20.0)

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
        assert self.meas.dtype in (torch.float32, torch.float64) and self.meas.shape[3] == 24 and self.meas.shape[5] == 4
        assert self.cams.shape == (self.S, self.C, skeleton.CAM_STRIDE) and self.cams.dtype == torch.float64
        assert self.chain_vec.shape == (self.S, 24, 5, 3) and self.chain_vec.dtype == torch.float64
        b = _lib.CpeBatch()
        b.S, b.N, b.C, b.W = self.S, self.N, self.C, self.W
        b.cam_model = int(self.cam_model)
        b.meas_is_f64 = 1 if self.meas.dtype == torch.float64 else 0
        b.mu_eq, b.mu_in = float(self.mu_eq), float(self.mu_in)
        b.loss_a, b.loss_b, b.loss_c = self.loss
        b.n_frames, b.h = self.n_frames.data_ptr(), self.h.data_ptr()
        b.meas, b.cams, b.chain_vec = self.meas.data_ptr(), self.cams.data_ptr(), self.chain_vec.data_ptr()
        b.lam_eq = self.lam_eq.data_ptr() if self.lam_eq is not None else None
        b.lam_in = self.lam_in.data_ptr() if self.lam_in is not None else None
        return b

    @staticmethod
    def from_numpy(meas, cams, chain_vec, h, n_frames=None, device="cuda", cam_model=FISHEYE, meas_dtype=torch.float32):
        """meas[S,N,C,24,W,3|4] (u, v, gate), cams[S,C,24], chain_vec[S,24,5,3], h[S]."""
        meas = np.asarray(meas)
        S, N = meas.shape[:2]
        if meas.shape[-1] == 3:
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
            cam_model=cam_model)


class Solver:
    """Workspaces + launches for one (model, batch)."""

    def __init__(self, model: Model, batch: Batch, with_solver=True):
        self.m, self.b = model, batch
        self.lib = model.lib
        S, N, dev = batch.S, batch.N, model.device
        f64 = dict(dtype=torch.float64, device=dev)
        self.frame_cost = torch.zeros(S, N, 4, **f64)
        self.cost = torch.zeros(S, 4, **f64)
        self.launches = 0
        if with_solver:
            G = model.n_groups
            self.Sblk = torch.zeros(S, N, 28, 28, **f64)
            self.ghat = torch.zeros(S, N, 28, **f64)
            self.Y = torch.zeros(S, N, G, 6, 12, **f64)
            self.y = torch.zeros(S, N, G, 6, **f64)
            self.g = torch.zeros(S, N, 54, **f64)
            self.diagH = torch.zeros(S, N, 54, **f64)
            self.Lfac = torch.empty(S, N, 4, 28, 28, **f64)
            self.zbuf = torch.zeros(S, N, 28, **f64)
            self.delta_ind = torch.zeros(S, N, 28, **f64)
            self.status = torch.zeros(S, dtype=torch.int32, device=dev)
            self.delta = torch.zeros(S, N, 54, **f64)
            self.q_try = torch.zeros(S, N, 54, **f64)
            self.frame_pred = torch.zeros(S, N, **f64)
            self.state = torch.zeros(4, S, **f64)
            self.counters = torch.zeros(S, 2, dtype=torch.int32, device=dev)
            self.active = torch.ones(S, dtype=torch.int32, device=dev)
            self.c_eq = torch.zeros(S, N, 26, **f64)
            self.v_in = torch.zeros(S, N, model.n_ineq, **f64)

    # ---- K1 --------------------------------------------------------------------------------------------------
    def evaluate(self, q, pose=False, resid=False, constraints=False, jac=False, cost_out=None):
        """cost[S,4] = (measurement, model, al_eq, al_in) and optional per-frame outputs."""
        b = self.b
        assert q.is_cuda and q.dtype == torch.float64 and q.is_contiguous() and q.shape == (b.S, b.N, 54)
        out = {}
        eo = _lib.CpeEvalOut()
        eo.frame_cost = self.frame_cost.data_ptr()
        f64 = dict(dtype=torch.float64, device=q.device)
        if pose:
            out["pose"] = torch.zeros(b.S, b.N, 24, 3, **f64)
            eo.pose = out["pose"].data_ptr()
        if resid:
            out["resid"] = torch.zeros(b.S, b.N, b.C, 24, b.W, 2, **f64)
            eo.resid = out["resid"].data_ptr()
        if constraints:
            out["c_eq"] = self.c_eq if hasattr(self, "c_eq") else torch.zeros(b.S, b.N, 26, **f64)
            out["v_in"] = self.v_in if hasattr(self, "v_in") else torch.zeros(b.S, b.N, self.m.n_ineq, **f64)
            eo.c_eq, eo.v_in = out["c_eq"].data_ptr(), out["v_in"].data_ptr()
        if jac:
            out["jac"] = torch.zeros(b.S, b.N, b.C, 24, b.W, 2, 18, **f64)
            eo.jac = out["jac"].data_ptr()
        bs = b.struct()
        st = _stream()
        _lib.check(self.lib.cpe_eval_f64(C.byref(self.m.struct), C.byref(bs), _ptr(q), C.byref(eo), st), "cpe_eval_f64")
        cost = self.cost if cost_out is None else cost_out
        _lib.check(self.lib.cpe_reduce_cost_f64(C.byref(bs), _ptr(self.frame_cost), _ptr(cost), st), "cpe_reduce_cost_f64")
        self.launches += 2
        out["cost"] = cost
        return out

    # ---- K2 --------------------------------------------------------------------------------------------------
    def _asm_struct(self, raw=None, reduced=True):
        ao = _lib.CpeAssembleOut()
        if reduced:
            ao.S, ao.ghat, ao.Y, ao.y = (self.Sblk.data_ptr(), self.ghat.data_ptr(), self.Y.data_ptr(), self.y.data_ptr())
        ao.g, ao.diagH = self.g.data_ptr(), self.diagH.data_ptr()
        if raw is not None:
            ao.Hraw = raw.data_ptr()
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
        return Hraw

    # ---- K3 --------------------------------------------------------------------------------------------------
    def band_solve(self):
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
    def solve(self, q0, max_iter=40, n_outer=4, damp0=1e-3, ftol=1e-6, mu_schedule=(1e4, 1e5, 1e6, 1e6), check_every=1):
        """Minimise the reference objective over q starting from q0[S,N,54]; returns a dict of device tensors.

        Outer loop = augmented-Lagrangian multiplier updates with penalty continuation; inner loop = LM with
        per-sequence damping (all sequences advance in lock step; converged ones are masked)."""
        b, lib = self.b, self.lib
        S, N, dev = b.S, b.N, q0.device
        q = q0.clone().contiguous()
        b.lam_eq = torch.zeros(S, N, 26, dtype=torch.float64, device=dev)
        b.lam_in = torch.zeros(S, N, self.m.n_ineq, dtype=torch.float64, device=dev)
        self.state.zero_()
        self.state[1].fill_(damp0)
        self.state[2].fill_(2.0)
        self.counters.zero_()
        cost_try = torch.zeros(S, 4, dtype=torch.float64, device=dev)
        damp = self.state[1]
        total_inner = 0
        for outer in range(n_outer):
            b.mu_eq = b.mu_in = float(mu_schedule[min(outer, len(mu_schedule) - 1)])
            self.evaluate(q)
            self.state[0].copy_(self.cost.sum(dim=1))
            self.active.fill_(1)
            for it in range(max_iter):
                self.lm_step(q, damp)
                self.evaluate(self.q_try, cost_out=cost_try)
                bs = b.struct()
                _lib.check(lib.cpe_step_accept_f64(C.byref(bs), _ptr(cost_try), _ptr(self.frame_pred), _ptr(q),
                                                   _ptr(self.q_try), _ptr(self.state), _ptr(self.counters),
                                                   _ptr(self.active), C.c_double(ftol), _stream()), "cpe_step_accept_f64")
                self.launches += 1
                total_inner += 1
                if (it + 1) % check_every == 0 and not bool(self.active.any()):
                    break
            self.evaluate(q, constraints=True)
            bs = b.struct()
            _lib.check(lib.cpe_update_multipliers_f64(C.byref(self.m.struct), C.byref(bs), _ptr(self.c_eq), _ptr(self.v_in),
                                                      _ptr(b.lam_eq), _ptr(b.lam_in), _stream()),
                       "cpe_update_multipliers_f64")
            self.launches += 1
        out = self.evaluate(q, constraints=True)
        return dict(q=q, cost=out["cost"].clone(), c_eq=out["c_eq"], v_in=out["v_in"], counters=self.counters.clone(),
                    status=self.status.clone(), lockstep_iterations=total_inner)
