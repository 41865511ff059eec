/* cpe.h -- C ABI of the B200-native cheetah trajectory-optimisation hot path (libcpe_b200.so).
 *
 * The reference (zicodasilva/cheetah_pose_estimation) has NO FFI: its hot path is the Pyomo model it hands to the
 * external `ipopt` executable.  This ABI is what a maintainer binds (ctypes) to replace that seam:
 *
 *   reference seam being replaced                                         entry point here
 *   -------------------------------------------------------------------  ---------------------------------------
 *   misc.get_pose_state position_funcs          acinoset_misc.py:1581-1659  cpe_eval_f64  (pose output)
 *   misc.pt3d_to_2d_fisheye / pt3d_to_2d        acinoset_misc.py:1663-1696  cpe_eval_f64  (resid output)
 *   create_camera_contraints `measurement` rows acinoset_misc.py:278-288    cpe_eval_f64  (resid = slack_meas)
 *   measurement_cost + redescending_loss        acinoset_misc.py:459-484,2001-2015  cpe_eval_f64 (cost[0])
 *   constant_acc_cost + implicit-Euler colloc.  acinoset_misc.py:639-677; collocation.py:58-68  cpe_eval_f64 (cost[1])
 *   angle_constr equalities                     system.py:574-586; links.py:552-586  cpe_eval_f64 (c_eq, cost[2])
 *   add_pyomo_constraints inequalities          cheetah.py:306-352; links.py:602-605,636-639  cpe_eval_f64 (v_in, cost[3])
 *   ASL Jacobian / Hessian evaluation inside ipopt (utils.py:522-535)     cpe_eval_f64 (jac), cpe_assemble_f64
 *   MA97 KKT factorisation inside ipopt (acinoset_opt.py:611-617)         cpe_band_solve_f64
 *   ipopt step acceptance                                                 cpe_step_apply_f64 / cpe_step_accept_f64
 *
 * Conventions: plain `extern "C"`; every pointer inside the structs is a DEVICE pointer owned by the caller
 * (contiguous, 16-byte aligned) unless its comment says "host"; nothing is allocated or freed inside; every call is
 * asynchronous on `stream`; no global state, thread-safe per stream.  Return value: 0 = OK, <0 = argument error
 * (CPE_E_*), >0 = cudaError_t of the launch.  Per-sequence solver status is written to device arrays.
 *
 * Index conventions: S sequences, N = frame stride (max frames), C cameras, L = 24 markers, W hypotheses per marker
 * (1, or 3 with pairwise pseudo-measurements), P = 54 generalised coordinates in the reference's order
 * (cheetah.py:197-198): base x,y,z,phi,theta,psi then 16 x (phi,theta,psi).
 */
#ifndef CPE_H_
#define CPE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CPE_VERSION 121          /* 0.1.21 */

#define CPE_N_MARKERS 24
#define CPE_N_LINKS 17
#define CPE_N_COORDS 54
#define CPE_MAX_CHAIN 5          /* links between the base origin and a marker */
#define CPE_N_NODES 18           /* node 0 = base xyz, node 1+i = link i: coordinate p lives in node p/3 */
#define CPE_N_IND 28             /* coordinates with a motion prior (Q_p != 0) */
#define CPE_N_DEP 26             /* coordinates without (eliminated per frame)  */
#define CPE_N_EQ 26              /* joint equalities per frame */
#define CPE_MAX_INEQ 32
#define CPE_CAM_STRIDE 24        /* doubles per camera: fx fy cx cy D0..D3 R[9] t[3] mult pad[3] */
#define CPE_MAX_CAMS 8
#define CPE_GRP_MAX_DEP 6        /* elimination group: <= 6 dependent coords ... */
#define CPE_GRP_MAX_NB 12        /* ... coupled to <= 12 independent coords */
#define CPE_MAX_GROUPS 8
#define CPE_COST_TERMS 4         /* measurement, model, al_eq, al_in */
#define CPE_K2_ENT_PAD 640       /* work items of K2's entry gather, padded (>= n_ent) */
#define CPE_K2_S_PAD 832         /* work items of K2's Schur gather, padded (>= 28 * 29) */

#define CPE_CAM_FISHEYE 0        /* acinoset_misc.py:1663-1679 */
#define CPE_CAM_PINHOLE 1        /* acinoset_misc.py:1682-1696 */

/* layouts of cpe_batch.meas, one record per (sequence, frame, camera, marker, hypothesis) */
#define CPE_MEAS_F32X4 0         /* float4  {u, v, gate, unused}, gate = likelihood > dlc_thresh ? 1 : 0 packed on the host */
#define CPE_MEAS_F64X4 1         /* double4 {u, v, gate, unused} */
#define CPE_MEAS_F32X3 2         /* float3  {x, y, likelihood}: the DLC-native triple (acinoset_misc.py:237-256); the
                                    strict test likelihood > dlc_thresh (acinoset_misc.py:231) runs on the device */

#define CPE_MEAS_F32X2G 3        /* float2 {u, v} in cpe_batch.meas plus one byte per record in cpe_batch.meas_gate (0 / 1,
                                    the likelihood test done once on the host): 9 bytes per detection instead of 12 / 16 --
                                    the layout of choice when the detections cross PCIe (cpe_eval_host_uvg_f64) */
#define CPE_MEAS_F32X2B 4        /* the same with the gates packed 8 per byte: bit (i & 7) of meas_gate[i >> 3] for the flat
                                    record index i (numpy.packbits(..., bitorder="little")); 8.125 bytes per detection.  A
                                    batch holds S * N * C * 24 * W records, always a multiple of 8 */

#define CPE_LOSS_REDESCENDING 0  /* acinoset_misc.py:476-482, 2001-2015 */
#define CPE_LOSS_SQUARE 1        /* hand-labelled data: (w slack)^2, camera multiplier not applied (acinoset_misc.py:471-474) */

/* per-sequence exit status of a solve (bit flags; 0 = converged: the last multiplier round stopped on ftol or at a point
 * no damping could improve) */
#define CPE_STATUS_PIVOT 1       /* a block of the banded factorisation had a non-positive pivot */
#define CPE_STATUS_MAXITER 2     /* the last multiplier round ran out of LM iterations */
#define CPE_STATUS_NONFINITE 4   /* the final cost is NaN or infinite (NaN in q, the cameras or pre-gated detections) */

#define CPE_E_ARG (-1)
#define CPE_E_DIM (-2)
#define CPE_E_NULL (-3)
#define CPE_E_SCRATCH (-4)       /* the problem exceeds the static scratch of a *_noalloc entry point */

/* Static skeleton tables (device pointers; built once on the host from the chain topology). */
typedef struct cpe_model {
    int32_t n_ineq;                 /* two-sided linear inequality rows per frame (<= CPE_MAX_INEQ) */
    int32_t n_groups;               /* elimination groups of dependent coordinates */
    int32_t n_u;                    /* packed chain columns over all markers: sum_l 3*chainlen(l) (198) */
    int32_t n_ent;                  /* Hessian lower-triangle nonzeros + 54 gradient entries (639) */
    const int32_t *chain_link;      /* [24*5] link index or -1 */
    const int32_t *eq_pairs;        /* [26*4] link_a, link_b, axis_a, axis_b : (R_a e_axa).(R_b e_axb) = 0 */
    const int32_t *ineq_idx;        /* [n_ineq*2] coord_plus, coord_minus (-1 = none): lo <= q[+] - q[-] <= up */
    const double *ineq_bounds;      /* [n_ineq*2] lo, up */
    const double *wq;               /* [54] motion weights 1/Q_p^2 (0 where Q_p = 0), acinoset_misc.py:234 */
    const double *meas_w;           /* [3*24] 1/R_pw[w][l]  (acinoset_misc.py:231) */
    const int32_t *ind_idx;         /* [28] coordinates with wq != 0 */
    const int32_t *ind_pos;         /* [54] position of coord in ind_idx or -1 */
    const int32_t *grp_ndep;        /* [n_groups] */
    const int32_t *grp_nnb;         /* [n_groups] */
    const int32_t *grp_dep;         /* [n_groups*6] coordinate indices */
    const int32_t *grp_nb;          /* [n_groups*12] coordinate indices (independent) */
    const int32_t *u_off;           /* [25] first packed chain column of marker l (prefix sum of 3*chainlen) */
    const int32_t *u_item;          /* [n_u] (marker << 8) | (chain slot << 4) | angle, per packed chain column */
    const int32_t *col_item;        /* [24+n_u/3] K1 Jacobian items of one frame.view in packed order, one per 3 adjacent
                                       columns (base x,y,z of a marker, or phi,theta,psi of a chain link):
                                       (marker << 16) | 10 * block, block 0 = identity, 1 + k = (marker, link) pair k */
    /* static gather lists of K2 (skeleton.assemble_tables); entry e = Hessian nonzero (p >= q) or gradient p */
    const int32_t *ent_pq;          /* [n_ent] p*64 + q ; q = 63 marks a gradient entry */
    const int32_t *a_ptr;           /* [n_ent+1] marker items: (offset in G space << 16) | offset in B space */
    const int32_t *a_items;
    const int32_t *e_ptr;           /* [n_ent+1] joint-equality items: (eqg index << 8) | (eqg index | eq row) */
    const int32_t *e_items;
    const int32_t *i_ptr;           /* [n_ent+1] inequality items: (row << 1) | negative */
    const int32_t *i_items;
    const int32_t *diag_ent;        /* [54] entry of H(p,p) */
    const int32_t *grad_ent;        /* [54] entry of g(p) */
    const int32_t *grp_src;         /* [n_groups*6*19] tableau sources: entry, -1 = zero, -2 = identity padding */
    const int32_t *s_src;           /* [28*28] entry of H(ind_r, ind_c) or -1 */
    const int32_t *s_ptr;           /* [28*29+1] Schur items of S(r,c) (c = 28: ghat): (group << 8) | (a << 4) | b */
    const int32_t *s_items;
    const int32_t *ent_order;       /* [n_ent] processing order of the entries, longest gather list first */
    const int32_t *s_order;         /* [28*29] processing order of the Schur items, longest list first */
    /* The same lists as ONE blob in the layout K2 copies to shared memory once per CTA (skeleton.k2_tables): work items in
       processing order, list items slot-major, each slot truncated to the prefix of work items that reach it.  The CSR
       tables above remain the definition (tests/test_assemble_tables.py checks the blob against them).
       k2_off[] = byte offsets of the sections: [0] meta uint2[CPE_K2_ENT_PAD], [1] smeta uint32[CPE_K2_S_PAD],
       [2] grp int16, [3..26] a slots int32, [27..31] e slots uint16, [32..35] i slots uint8, [36..40] s slots uint16,
       [41..44] uint8 block lists: which blocks of 32 entries / Schur items each warp of a team gathers (41 and 42 for
       5-warp teams, 43 and 44 for 6-warp teams, 8 slots per warp, 255 = none), [45] total bytes (multiple of 16),
       [46] number of Schur work items (lower triangle of S + reduced gradient). */
    const void *k2_tab;
    int32_t k2_off[48];
} cpe_model;

/* One batch of sequences. */
typedef struct cpe_batch {
    int32_t S, N, C, W;
    int32_t cam_model;              /* CPE_CAM_FISHEYE / CPE_CAM_PINHOLE */
    int32_t meas_layout;            /* CPE_MEAS_* */
    int32_t loss_mode;              /* CPE_LOSS_* */
    int32_t reserved0;
    double dlc_thresh;              /* likelihood threshold of CPE_MEAS_F32X3 (reference default 0.5) */
    double mu_eq, mu_in;            /* augmented-Lagrangian penalties */
    double loss_a, loss_b, loss_c;  /* redescending loss break points (3, 10, 20) */
    const int32_t *n_frames;        /* [S] frames used of each sequence (<= N) */
    const double *h;                /* [S] time step 1/fps (acinoset_opt.py:483-487, 508) */
    const void *meas;               /* [S,N,C,24,W] records in meas_layout */
    const double *cams;             /* [S,C,24]  see CPE_CAM_STRIDE */
    const double *chain_vec;        /* [S,24,5,3] body-frame vectors of the marker chains */
    const double *lam_eq;           /* [S,N,26] multipliers (may be NULL = 0) */
    const double *lam_in;           /* [S,N,n_ineq] (may be NULL = 0) */
    const int32_t *active;          /* [S] optional (NULL = all): sequences with active == 0 are skipped by every kernel
                                       (their outputs keep their previous contents) -- converged sequences of an LM loop */
    /* Shutter delay of the measurement model (acinoset_misc.py:182-183, 273-286; acinoset_opt.py:342-352): camera c sees
       every marker of frame n displaced by  dq_base[n] tau_c + ddq_base[n] tau_c^2  (no 1/2, as in the reference).
       All three NULL = no delay.  cpe_eval_* and cpe_assemble_f64 honour them: residuals, costs, and derivatives with
       respect to q_n at fixed dq, ddq, tau -- the reference's own partial derivatives, dq and ddq being separate NLP
       variables there; eo.pose stays the undelayed FK, the reference's `positions`.  The derivatives with respect to
       tau, dq_base and ddq_base come from cpe_assemble_out.cam_blocks (engine.Solver.solve_shutter is the driver that
       estimates tau with them). */
    const double *shutter_tau;      /* [S,C] seconds; the reference fixes camera 1 at 0 and bounds the rest by +-1/fps */
    const double *base_dq;          /* [S,N,3] velocity of base x, y, z     (robot["base"]["dq"][n, 1, .]) */
    const double *base_ddq;         /* [S,N,3] acceleration of base x, y, z (robot["base"]["ddq"][n, 1, .]) */
    /* Ragged camera count (the clips of data/test_set have 4, 5 or 6 cameras; acinoset_misc.py:171-174 makes C
       per-clip data): cameras c >= n_cams[s] of sequence s are PADDING.  Their residuals are still written (layout
       stays [.., C, ..]) but they add nothing to the cost, the gradient or the curvature, and because they sit behind
       the real cameras every sum keeps the order of the unpadded problem: results are bit-identical to a batch with
       C = n_cams[s].  The padded cams[] rows must still hold a valid camera.  NULL = C cameras for every sequence. */
    const int32_t *n_cams;          /* [S] optional */
    const uint8_t *meas_gate;       /* [S,N,C,24,W] gates of CPE_MEAS_F32X2G, or their bits for CPE_MEAS_F32X2B (NULL for the
                                       other layouts) */
} cpe_batch;

/* Optional outputs of cpe_eval (NULL = skip). */
typedef struct cpe_eval_out {
    double *frame_cost;             /* [S,N,4] per-frame partial sums (measurement, model, al_eq, al_in) */
    double *pose;                   /* [S,N,24,3] marker positions */
    double *resid;                  /* [S,N,C,24,W,2] slack = proj - meas (what the reference saves as meas_err) */
    double *c_eq;                   /* [S,N,26] */
    double *v_in;                   /* [S,N,n_ineq]  t - clamp(t, lo, up), t = d + lam/mu */
    double *jac;                    /* [S,N,C,2*(72+n_u)] packed chain-sparse d(u,v)/dq (540 per frame.view): marker l owns
                                       2*ncol(l) entries at 2*(3l+u_off[l]), ncol = 3 + u_off[l+1]-u_off[l]: one
                                       (du/dq_j, dv/dq_j) pair per column j, columns = base x,y,z then
                                       (phi,theta,psi) of each chain link in chain order.  Independent of w. */
} cpe_eval_out;

/* Outputs of the FP32 flavour: residuals and Jacobian in float, everything else as above. */
typedef struct cpe_eval_out_f32 {
    double *frame_cost;             /* [S,N,4] (per-thread loss in float, summed in double) */
    double *pose;                   /* [S,N,24,3] (FP64: the FK front end is FP64 in both flavours) */
    float *resid;                   /* [S,N,C,24,W,2] */
    double *c_eq;                   /* [S,N,26] */
    double *v_in;                   /* [S,N,n_ineq] */
    float *jac;                     /* [S,N,C,2*(72+n_u)] same packing as cpe_eval_out.jac */
} cpe_eval_out_f32;

/* Outputs of cpe_assemble: the per-frame reduced normal equations after Marquardt damping and elimination. */
typedef struct cpe_assemble_out {
    double *S;                      /* [S,N,28,28] Schur complement on the independent coordinates (no motion term) */
    double *ghat;                   /* [S,N,28] reduced gradient (incl. motion gradient) */
    double *Y;                      /* [S,N,n_groups,6,12] H_dd^-1 H_di per group */
    double *y;                      /* [S,N,n_groups,6]    H_dd^-1 g_d */
    double *g;                      /* [S,N,54] full gradient (incl. motion gradient) */
    double *diagH;                  /* [S,N,54] diag of the undamped GN Hessian incl. motion diagonal */
    double *Hraw;                   /* optional [S,N,54,54] dense undamped frame Hessian (tests / cpe_normal_eq) */
    double *cam_blocks;             /* optional [S,N,C,9] per camera, summed over markers and hypotheses, in point space:
                                       IRLS-weighted P^T P (xx, xy, xz, yy, yz, zz) and P^T rho' (x, y, z), P = d(u,v)/d(point).
                                       The shutter-delay terms are built from these: d cost / d tau_c =
                                       sum_n b_nc . (dq_base + 2 tau_c ddq_base), and the gradient through dq_base, ddq_base */
    double *ent;                    /* optional [S,N,CPE_K2_ENT_STORE]: the UNDAMPED Hessian / gradient entries of the frame (the
                                       gather results, cpe_model.n_ent <= 640 values) followed at 640 by diagH[54]; written by
                                       every full assembly, read back by a re-damp (next field) */
    const double *redamp;           /* optional [S], needs `ent`: sequences with redamp[s] < 0 only RE-DAMP -- their q has not
                                       changed since the last call (a rejected LM step: cpe_step_accept leaves -1 in row 3 of
                                       its `state`), so the entries are read back from `ent` and only the damped elimination
                                       and the Schur gather run again.  S, ghat, Y, y come out bit-identical to a full
                                       assembly; g, diagH, Hraw, cam_blocks are left as the last full assembly wrote them */
} cpe_assemble_out;

/* Data-driven monocular priors on the reduced pose x = Tr q_ind (28 = the coordinates with a motion prior; the
 * reduced pose only involves those: acinoset_misc.py:487-528, 1699-1757).  Parameters are fitted on the CPU as in the
 * reference (acinoset_models.py:173-300) and frozen here; the kernels evaluate
 *   pose   sum_frames -log( sum_c exp(logc_c - 1/2 d_c^T P_c d_c) + 1e-12 ),  d_c = x[6:28] - mu_c   (acinoset_misc.py:680-714)
 *   motion sum_{k >= window} sum_p wt_p ( x_k - W [x_{k-window}; ...; x_{k-1}] - b )_p^2              (acinoset_misc.py:291-336)
 * with logc_c = log w_c - 1/2 (22 log 2 pi + log det Sigma_c), P_c = Sigma_c^-1. */
#define CPE_PRIOR_MAX_COMP 8
#define CPE_PRIOR_MAX_WINDOW 4
#define CPE_PRIOR_DIM 22          /* entries 6..27 of the reduced pose */
typedef struct cpe_prior {
    int32_t n_comp;                 /* GMM components (0 = pose prior off) */
    int32_t window;                 /* AR window (0 = motion prior off; <= CPE_PRIOR_MAX_WINDOW) */
    const double *Tr;               /* [28*28] x = Tr q_ind */
    const double *gmm_logc;         /* [n_comp] */
    const double *gmm_mu;           /* [n_comp*22] */
    const double *gmm_prec;         /* [n_comp*22*22] */
    const double *ar_W;             /* [28*(28*window)] oldest lag first (data_ops.py:108-142) */
    const double *ar_b;             /* [28] */
    const double *ar_wt;            /* [28] 1/var_p, 0 where var_p = 0 (acinoset_misc.py:308) */
} cpe_prior;

/* Workspace + outputs of the block-banded solve. */
typedef struct cpe_solve_io {
    double *Lfac;                   /* [S,N,3,28,28] workspace: L1, L2, L3 (unit-lower block factors of each block row) */
    double *zbuf;                   /* [S,N,28] workspace */
    double *delta_ind;              /* [S,N,28] out */
    int32_t *status;                /* [S] out: 0 ok, 1 non-positive pivot encountered (CPE_STATUS_PIVOT) */
} cpe_solve_io;

int cpe_version(void);

/* K1: FK + projection + robust loss (+ optional chain-sparse Jacobian). q[S,N,54]. */
int cpe_eval_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const cpe_eval_out *out,
                 void *stream);

/* K1, scaled FP32 flavour: marker offsets relative to the base origin, camera translation and principal-point
 * minus detection folded in FP64, then projection / residual / loss / Jacobian in FP32 (per-residual error <= 1e-4
 * relative to max(|r|, 1 px); see tests/test_gpu_parity.py). */
int cpe_eval_f32(const cpe_model *model, const cpe_batch *batch, const double *q, const cpe_eval_out_f32 *out,
                 void *stream);

/* Deterministic per-sequence reduction of frame_cost[S,N,4] -> cost[S,4] (fixed order). */
int cpe_reduce_cost_f64(const cpe_batch *batch, const double *frame_cost, double *cost, void *stream);

/* K2: fused residual + Jacobian + J^T J / J^T r accumulation + damping + per-frame elimination. damp[S]. */
int cpe_assemble_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const double *damp,
                     const cpe_assemble_out *out, void *stream);

/* K3: block-banded (block half-bandwidth 3, n = 28) LDL^T factor + solve of (blockdiag(S) + M) d = -ghat. */
int cpe_band_solve_f64(const cpe_model *model, const cpe_batch *batch, const double *Sblk, const double *ghat,
                       const cpe_solve_io *io, void *stream);

/* K4: back-substitute the eliminated coordinates, form q_try = q + delta and the predicted reduction.
 * frame_pred[S,N] = -1/2 g.delta + 1/2 damp sum diagH delta^2 (per frame partials). */
int cpe_step_apply_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const double *damp,
                       const cpe_assemble_out *asm_out, const double *delta_ind, double *delta, double *q_try,
                       double *frame_pred, void *stream);

/* K5: per-sequence accept / reject with Nielsen damping update.
 * state[4,S] (structure of arrays) = {merit F, damp, nu, last relative decrease (-1 = rejected)};
 * counters[S,2] = {iterations, accepted}.  cost_try[S,4] from cpe_reduce_cost of q_try; q is overwritten by q_try
 * where accepted.  active[S] (may be NULL): sequences with active == 0 are skipped; a sequence is switched off when
 * an accepted step decreased the merit by less than ftol (relative) or its damping hit the ceiling. */
int cpe_step_accept_f64(const cpe_batch *batch, const double *cost_try, const double *frame_pred, double *q,
                        const double *q_try, double *state, int32_t *counters, int32_t *active, double ftol,
                        void *stream);
/* Same, with the prior costs of the trial point added to the merit: cost_extra[S,2] = (pose, motion). */
int cpe_step_accept_prior_f64(const cpe_batch *batch, const double *cost_try, const double *cost_extra,
                              const double *frame_pred, double *q, const double *q_try, double *state,
                              int32_t *counters, int32_t *active, double ftol, void *stream);

/* Monocular priors, cost only: prior_cost[S,2] = (pose, motion) of q[S,N,54]; frame_work[S,N,2] is scratch. */
int cpe_prior_cost_f64(const cpe_model *model, const cpe_prior *prior, const cpe_batch *batch, const double *q,
                       double *frame_work, double *prior_cost, void *stream);

/* Monocular priors, normal-equation side: call AFTER cpe_assemble_f64 on the same q / damp.  Adds, per frame, the
 * gradient of both priors to out->ghat and out->g, and the PSD majoriser sum_c r_c Tr^T P_c Tr of the pose-prior
 * Hessian to out->S (plus damp * its diagonal) and its diagonal to out->diagH.  The motion prior's exact (constant)
 * Hessian 2 J^T W J is not added here: it is part of the band `Mband` given to cpe_band_solve_general_f64. */
int cpe_prior_assemble_f64(const cpe_model *model, const cpe_prior *prior, const cpe_batch *batch, const double *q,
                           const double *damp, const cpe_assemble_out *out, void *stream);

/* General banded SPD solve of (blockdiag(S) + M) d = -ghat per sequence, n = 28 N unknowns, scalar half-bandwidth
 * bw (<= 139 = 28 * 5 - 1: AR window 4).  Mband[n][bw+1] holds the lower band of the constant part M (motion terms):
 * Mband[i][d] = M(i, i-d); sequence s reads Mband + s * mband_stride (0 = shared by all sequences).  Lband[S][n][bw+1]
 * is workspace (the Cholesky factor, column-major band).  Used by the monocular path, whose AR(4) prior couples five
 * consecutive frames with dense 28x28 blocks; the multi-view path keeps cpe_band_solve_f64. */
int cpe_band_solve_general_f64(const cpe_batch *batch, const double *Sblk, const double *ghat, const double *Mband,
                               int64_t mband_stride, int32_t bw, double *Lband, double *delta_ind, int32_t *status,
                               void *stream);

/* Exit status of a solve, status[S] = CPE_STATUS_* flags: `pivot` = cpe_solve_io.status of the last band solve (may be
 * NULL), `active` = the LM activity flags when the last round ended (NULL = none), cost[S,4] (+ cost_extra[S,2], may be
 * NULL) = the final cost terms. */
int cpe_finalize_status_f64(const cpe_batch *batch, const double *cost, const double *cost_extra, const int32_t *active,
                            const int32_t *pivot, int32_t *status, void *stream);

/* Augmented-Lagrangian multiplier update: lam_eq += mu_eq * c_eq ; lam_in = mu_in * v_in.  Sequences with
 * batch->active[s] == 0 keep their multipliers (NULL: all are updated). */
int cpe_update_multipliers_f64(const cpe_model *model, const cpe_batch *batch, const double *c_eq,
                               const double *v_in, double *lam_eq, double *lam_in, void *stream);

/* Host-buffer convenience wrapper used for the end-to-end metric: copies q and meas from (pinned) host memory,
 * evaluates, and copies cost[S,4] (and resid = the reference's `meas_err`, if h_resid and out_dev->resid are non-NULL)
 * back.  d_* are caller-owned device staging buffers.  h_meas == NULL: the detections are already resident in
 * batch_dev->meas (they are constant across the iterations of a solve) and only q crosses the bus. */
int cpe_eval_host_f64(const cpe_model *model, const cpe_batch *batch_dev, const double *h_q, const void *h_meas,
                      double *d_q, void *d_meas, const cpe_eval_out *out_dev, double *d_cost, double *h_cost,
                      double *h_resid, void *stream);

/* Kinetic path, first slice: rows 0-5 of the equations of motion M ddq + C + G - B F, the functions
 * eom[i](q, dq, ddq, GRFz, GRFxy) of CheetahEstimator.calc_grf_eom / estimate_grf (acinoset_opt.py:120-270).
 * q, dq, ddq [S,N,54]; grf[S,N,20] = GRFz of the feet HFL, HFR, HBL, HBR, then GRFxy[foot][4], in units of
 * F_scale (acinoset_opt.py:142-149); link_par[S,52] = lengths[17], masses[17], radii[17] in link order, then
 * F_scale = total_mass * 9.81; eom[S,N,6].  The generated code carries the trunk dimensions of `jules` / `phantom` in
 * its hip offsets, exactly as the reference bakes them (cheetah.py:32-38). */
#define CPE_EOM_N_FORCES 20
#define CPE_EOM_N_PAR 52
int cpe_eom_f64(const cpe_batch *batch, const double *q, const double *dq, const double *ddq, const double *grf,
                const double *link_par, double *eom, void *stream);
/* The rows are affine in the forces: eom = a - Bm grf.  a[S,N,6], Bm[S,N,6,20]. */
int cpe_eom_affine_f64(const cpe_batch *batch, const double *q, const double *dq, const double *ddq,
                       const double *link_par, double *a, double *Bm, void *stream);
/* Per-frame GRF estimate of CheetahEstimator.estimate_grf (acinoset_opt.py:176-249): minimise || a - Bm F || over
 * 0 <= F <= 5 with the friction polyhedron 1.3 GRFz[f] >= sum_i GRFxy[f][i] (:187-190); contact[S,N,4] != 0 marks the
 * feet on the ground, the forces of the others are fixed to 0 (:221-229).  Accelerated projected gradient with adaptive
 * restart, at most `iters` iterations (a frame stops once its iterate is a fixed point to 1e-13; 600 reaches the
 * optimum value to ~1e-14 ||a|| on the fixture).  grf[S,N,20], resid[S,N] = the optimal || eom ||. */
int cpe_grf_solve_f64(const cpe_batch *batch, const double *a, const double *Bm, const int32_t *contact, int32_t iters,
                      double *grf, double *resid, void *stream);

/* The same for the compact layout CPE_MEAS_F32X2G: h_uv / d_uv hold float2 {u, v} records, h_gate / d_gate one byte each
 * (9 bytes per detection over PCIe instead of 12), or one BIT each for CPE_MEAS_F32X2B (8.125 bytes).
 * batch_dev->meas_layout must be one of the two; h_uv == NULL means the detections are already resident in
 * batch_dev->meas / meas_gate. */
int cpe_eval_host_uvg_f64(const cpe_model *model, const cpe_batch *batch_dev, const double *h_q, const void *h_uv,
                          const uint8_t *h_gate, double *d_q, void *d_uv, uint8_t *d_gate, const cpe_eval_out *out_dev,
                          double *d_cost, double *h_cost, double *h_resid, void *stream);

/* Trajectory-level cost terms of the kinetic objective (acinoset_opt.py:905-921), each with its gradient.  Frame ranges
 * follow the reference's code (1-based `fe < nfe - 2` / `fe < nfe - 1`: one term fewer than the data allows).
 * cpe_motion_smoothing_f64: misc.motion_smoothing_cost (acinoset_misc.py:601-613) with fps = 1 / batch->h[s]; pose[S,N,24,3]
 *   = cpe_eval_out.pose of the same q; cost[S]; grad[S,N,54] = d cost / d q (may be NULL); frame_work[S,N] scratch.
 * cpe_eom_slack_f64: misc.eom_slack_cost (:631-636) on eom[S,N,6] of cpe_eom_f64; with Bm[S,N,6,20] of cpe_eom_affine_f64
 *   also grad_f[S,N,20] = d cost / d F = -2 Bm^T eom (both may be NULL).
 * cpe_torque_costs_f64: Tc[S,N,n_axes] joint torques; cost[S,2] = { sum (weight Tc)^2 (motor.py:412-417, 496-503),
 *   sum ((Tc[fe+1] - Tc[fe]) / hm0)^2 with hm0 = batch->h[s] (acinoset_misc.py:616-628) }; grad_sq, grad_ch [S,N,n_axes]
 *   (may be NULL); frame_work[2,S,N] scratch. */
int cpe_motion_smoothing_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const double *pose,
                             double *frame_work, double *cost, double *grad, void *stream);
int cpe_eom_slack_f64(const cpe_batch *batch, const double *eom, const double *Bm, double *frame_work, double *cost,
                      double *grad_f, void *stream);
int cpe_torque_costs_f64(const cpe_batch *batch, const double *Tc, int32_t n_axes, const double *weight, double *frame_work,
                         double *cost, double *grad_sq, double *grad_ch, void *stream);

/* Vector-Jacobian product of rows 0-5 of the equations of motion (cpe_eom_f64, same inputs) with respect to the states:
 *   g_q, g_dq, g_ddq [S,N,54] = sum_i w[s,n,i] * d row_i / d (q, dq, ddq)      (frames beyond n_frames[s]: zeros)
 * w[S,N,6] == NULL takes w = 2 * rows: the gradient of the EOM slack cost sum_i row_i^2 (misc.eom_slack_cost,
 * acinoset_misc.py:631-636) -- the derivative ASL hands IPOPT for that term of the kinetic objective (acinoset_opt.py:905-921).
 * eom[S,N,6] (may be NULL) receives the rows.  Adjoint sweep of the generated program (oracle/derive_eom_grad.py). */
int cpe_eom_vjp_f64(const cpe_batch *batch, const double *q, const double *dq, const double *ddq, const double *grf,
                    const double *par, const double *w, double *eom, double *g_q, double *g_dq, double *g_ddq, void *stream);

/* Allocation-free variant of cpe_eom_slack_f64.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs and outputs as cpe_eom_slack_f64 minus `frame_work`: the per-frame scratch is a static device
 * array inside the library, CPE_NOALLOC_MAX_FRAMES doubles (512 KB per GPU that loads the library, held until exit).
 * Returns CPE_E_SCRATCH when S * N exceeds CPE_NOALLOC_MAX_FRAMES (use cpe_eom_slack_f64 with caller scratch then).
 * Calls no cudaMalloc / cudaFree / allocating routine; results are bit-identical to cpe_eom_slack_f64.
 * EVERY call on a device shares that one array (the part a call uses is cleared first, on `stream`): calls that overlap
 * on the GPU -- different streams, or graphs replayed concurrently -- RACE; serialise them. */
#define CPE_K2_ENT_STORE 704       /* doubles per frame of cpe_assemble_out.ent: 640 entries + diagH[54], padded */
#define CPE_NOALLOC_MAX_FRAMES 65536
int cpe_eom_slack_noalloc_f64(const cpe_batch *batch, const double *eom, const double *Bm, double *cost, double *grad_f,
                              void *stream);

/* Shutter-delay estimation (acinoset_misc.py:182-183, 273-286): the pieces that are not per-residual work.
 * cpe_base_rates_f64: dq_base, ddq_base [S,N,3] of q in the q-only formulation -- backward differences of the base
 * x, y, z (collocation.py:58-68 eliminated), first-node values as in the saved dq / ddq (ddq_0 = ddq_1 = ddq_2,
 * dq_0 = dq_1 - h ddq_1); zero for frames beyond n_frames and for sequences too short to have them.
 * cpe_shutter_terms_f64: from cpe_assemble_out.cam_blocks of an assembly at (q, tau) -- batch carries shutter_tau,
 * base_dq, base_ddq -- g_extra[S,N,3] = the part of d cost / d (base x, y, z) that goes through dq_base and ddq_base
 * (frame n feeds frames n, n-1, n-2), g_tau[S,C] = d cost / d tau_c, curv_tau[S,C] = its Gauss-Newton curvature. */
int cpe_base_rates_f64(const cpe_batch *batch, const double *q, double *dq_base, double *ddq_base, void *stream);
int cpe_shutter_terms_f64(const cpe_batch *batch, const double *cam_blocks, double *g_extra, double *g_tau,
                          double *curv_tau, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CPE_H_ */
