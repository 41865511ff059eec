"""ctypes binding of libcpe_b200.so (the C ABI declared in include/cpe.h).

There is NO fallback: if the CUDA library is missing or a call fails, this raises.  torch is used only for
device memory and streams (`tensor.data_ptr()`, `torch.cuda.current_stream().cuda_stream`).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CPE_LIB_PATH: another build of the same ABI (A/B timing of kernel variants); never a fallback
LIB_PATH = os.environ.get("CPE_LIB_PATH") or os.path.join(_HERE, "libcpe_b200.so")

c_i32p = C.c_void_p
c_f64p = C.c_void_p


class CpeModel(C.Structure):
    _fields_ = [
        ("n_ineq", C.c_int32), ("n_groups", C.c_int32), ("n_u", C.c_int32), ("n_ent", C.c_int32),
        ("chain_link", c_i32p), ("eq_pairs", c_i32p), ("ineq_idx", c_i32p),
        ("ineq_bounds", c_f64p), ("wq", c_f64p), ("meas_w", c_f64p), ("ind_idx", c_i32p), ("ind_pos", c_i32p),
        ("grp_ndep", c_i32p), ("grp_nnb", c_i32p), ("grp_dep", c_i32p), ("grp_nb", c_i32p),
        ("u_off", c_i32p), ("u_item", c_i32p), ("col_item", c_i32p),
        ("ent_pq", c_i32p), ("a_ptr", c_i32p), ("a_items", c_i32p), ("e_ptr", c_i32p), ("e_items", c_i32p),
        ("i_ptr", c_i32p), ("i_items", c_i32p), ("diag_ent", c_i32p), ("grad_ent", c_i32p), ("grp_src", c_i32p),
        ("s_src", c_i32p), ("s_ptr", c_i32p), ("s_items", c_i32p), ("ent_order", c_i32p), ("s_order", c_i32p),
        ("k2_tab", C.c_void_p), ("k2_off", C.c_int32 * 48),
    ]


class CpeBatch(C.Structure):
    _fields_ = [
        ("S", C.c_int32), ("N", C.c_int32), ("C", C.c_int32), ("W", C.c_int32),
        ("cam_model", C.c_int32), ("meas_layout", C.c_int32), ("loss_mode", C.c_int32), ("reserved0", C.c_int32),
        ("dlc_thresh", C.c_double),
        ("mu_eq", C.c_double), ("mu_in", C.c_double),
        ("loss_a", C.c_double), ("loss_b", C.c_double), ("loss_c", C.c_double),
        ("n_frames", c_i32p), ("h", c_f64p), ("meas", C.c_void_p), ("cams", c_f64p), ("chain_vec", c_f64p),
        ("lam_eq", c_f64p), ("lam_in", c_f64p), ("active", c_i32p),
        ("shutter_tau", c_f64p), ("base_dq", c_f64p), ("base_ddq", c_f64p),
        ("n_cams", c_i32p), ("meas_gate", C.c_void_p),
    ]


class CpeEvalOut(C.Structure):
    _fields_ = [("frame_cost", c_f64p), ("pose", c_f64p), ("resid", c_f64p), ("c_eq", c_f64p), ("v_in", c_f64p),
                ("jac", c_f64p)]


class CpeEvalOutF32(C.Structure):
    _fields_ = [("frame_cost", c_f64p), ("pose", c_f64p), ("resid", C.c_void_p), ("c_eq", c_f64p), ("v_in", c_f64p),
                ("jac", C.c_void_p)]


class CpeAssembleOut(C.Structure):
    _fields_ = [("S", c_f64p), ("ghat", c_f64p), ("Y", c_f64p), ("y", c_f64p), ("g", c_f64p), ("diagH", c_f64p),
                ("Hraw", c_f64p), ("cam_blocks", c_f64p), ("ent", c_f64p), ("redamp", c_f64p)]


class CpePrior(C.Structure):
    _fields_ = [("n_comp", C.c_int32), ("window", C.c_int32), ("Tr", c_f64p), ("gmm_logc", c_f64p), ("gmm_mu", c_f64p),
                ("gmm_prec", c_f64p), ("ar_W", c_f64p), ("ar_b", c_f64p), ("ar_wt", c_f64p)]


class CpeSolveIO(C.Structure):
    _fields_ = [("Lfac", c_f64p), ("zbuf", c_f64p), ("delta_ind", c_f64p), ("status", c_i32p)]


EXPORTS = {
    "cpe_version": (C.c_int, []),
    "cpe_eval_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, C.POINTER(CpeEvalOut), C.c_void_p]),
    "cpe_eval_f32": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, C.POINTER(CpeEvalOutF32), C.c_void_p]),
    "cpe_reduce_cost_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, C.c_void_p]),
    "cpe_assemble_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, c_f64p,
                                   C.POINTER(CpeAssembleOut), C.c_void_p]),
    "cpe_band_solve_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, c_f64p,
                                     C.POINTER(CpeSolveIO), C.c_void_p]),
    "cpe_step_apply_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, c_f64p,
                                     C.POINTER(CpeAssembleOut), c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_step_accept_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_i32p, c_i32p,
                                      C.c_double, C.c_void_p]),
    "cpe_step_accept_prior_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_i32p,
                                            c_i32p, C.c_double, C.c_void_p]),
    "cpe_prior_cost_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpePrior), C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p,
                                     C.c_void_p]),
    "cpe_prior_assemble_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpePrior), C.POINTER(CpeBatch), c_f64p, c_f64p,
                                         C.POINTER(CpeAssembleOut), C.c_void_p]),
    "cpe_band_solve_general_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, C.c_int64, C.c_int32, c_f64p,
                                             c_f64p, c_i32p, C.c_void_p]),
    "cpe_eom_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_eom_affine_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_grf_solve_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_i32p, C.c_int32, c_f64p, c_f64p, C.c_void_p]),
    "cpe_base_rates_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_shutter_terms_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_update_multipliers_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p,
                                             c_f64p, C.c_void_p]),
    "cpe_motion_smoothing_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p,
                                           C.c_void_p]),
    "cpe_eom_slack_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_eom_slack_noalloc_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, C.c_void_p]),
    "cpe_torque_costs_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, C.c_int32, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p,
                                       C.c_void_p]),
    "cpe_eom_vjp_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p,
                                  c_f64p, C.c_void_p]),
    "cpe_eval_host_uvg_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), C.c_void_p, C.c_void_p, C.c_void_p, c_f64p,
                                        C.c_void_p, C.c_void_p, C.POINTER(CpeEvalOut), c_f64p, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
    "cpe_finalize_status_f64": (C.c_int, [C.POINTER(CpeBatch), c_f64p, c_f64p, c_i32p, c_i32p, c_i32p, C.c_void_p]),
    "cpe_eval_host_f64": (C.c_int, [C.POINTER(CpeModel), C.POINTER(CpeBatch), C.c_void_p, C.c_void_p, c_f64p,
                                    C.c_void_p, C.POINTER(CpeEvalOut), c_f64p, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


class CpeError(RuntimeError):
    pass


ABI_VERSION = 121          # CPE_VERSION of include/cpe.h the ctypes mirrors were written for (tests/test_abi.py)


def load():
    """Load libcpe_b200.so and type its entry points.  Raises if the library is absent (no CPU fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CpeError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if lib.cpe_version() != ABI_VERSION:
        # the struct mirrors above are laid out for exactly this header version: a stale .so would read fields at wrong offsets
        raise CpeError(f"{LIB_PATH} is ABI version {lib.cpe_version()}, this package mirrors {ABI_VERSION} (include/cpe.h): "
                       "rebuild with `python -c 'import __graft_entry__ as g; g.build()'`")
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        kind = ("problem exceeds the static scratch of a *_noalloc entry point" if rc == -4 else
                "argument error" if rc < 0 else "CUDA error")
        raise CpeError(f"{what} failed: {kind} {rc}")
