// cpe_eval.cu -- K1: skeleton FK + camera projection + likelihood-weighted redescending reprojection loss, residuals,
// constraint values, motion cost and (optionally) the chain-sparse Jacobian d(u,v)/dq, for batches of sequences.
//
// Reference arithmetic being replaced (paths under the reference repo; seam table in include/cpe.h):
//   rotations   shared/physical_education/utils.py:39-64, links.py:483         R = Rz(psi) Ry(theta) Rx(phi)
//   FK chains   links.py:496-519, cheetah.py:19-200, acinoset_misc.py:1581-1659
//   projection  acinoset_misc.py:1663-1696
//   residual    acinoset_misc.py:278-288 ; weights :211-231 ; cost :459-484 ; loss :2001-2015
//   motion      acinoset_misc.py:639-677 + collocation.py:58-68 (q-only third difference)
//   joints      links.py:552-586 ; inequalities cheetah.py:306-352
//
// Two flavours of ONE kernel template: T = double (cpe_eval_f64, <= 1e-9 of the reference) and T = float
// (cpe_eval_f32, the "scaled FP32 residual path" of the north star, <= 1e-4).  In both, everything that is shared by
// the cameras of a frame and cheap -- sincos, link rotations, marker offsets, joint / inequality / motion terms --
// is FP64.  The FP32 flavour is *scaled*: markers are kept as offsets d_l from the base origin (|d| < 2 m) and the
// base enters through the per-(frame, camera) translation tc = R_c base + t_c formed in FP64, so no FP32 value
// ever carries the ~10 m world coordinates; the principal point is folded with the detection in FP64
// (cx - u_meas) before rounding, so the residual does not lose the ~1000 px offset either.  Per (camera, marker):
// projection, residual, robust loss and the Jacobian columns J = dproj(2x3) . [I | U] run in T.
//
// Mapping (sm_100a): ONE WARP PER FRAME.  A CTA is WPC independent warps working on WPC consecutive frames of one
// sequence; after the cameras are staged (the only CTA barrier) a warp never waits for another warp.  Per frame:
//   A  q -> the warp's shared slice; first 32 detections prefetched into registers (coalesced)
//   B  51 sincos (2 rounds over the lanes), 17 rotation matrices, FP64
//   C  rounds of warp-uniform items: chain derivative blocks U = dR/dangle . v (one item per (marker, link)),
//      marker offsets, camera translations, joint equalities, inequalities, motion terms; the cost terms stay in
//      lane registers
//   D  rounds of 32 (camera, marker) pairs.  Pair pass (lane <-> pair): projection, residuals, loss, and the 2x3
//      projection derivative of each pair into shared memory; the next round's detections are in flight meanwhile.
//      Item pass (lane <-> 3 adjacent Jacobian columns: base x,y,z of a marker or phi,theta,psi of a chain link):
//      J = dproj(2x3) . block(3x3).  The 32 pairs of a round own ONE contiguous span of the packed Jacobian and
//      consecutive lanes own consecutive 48-byte (24-byte in FP32) pieces of it, so the three stores of a round
//      fill the span without gaps (per-pair stores from registers were LSU-wavefront bound: 32 lines per store)
//   E  fixed-order shuffle reductions -> frame_cost
// No atomics, no cross-warp reductions: a sequence's results are bit-identical whatever its batch neighbours are.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "cpe.h"
#include "cpe_internal.cuh"

namespace cpe {

struct EvalParams {
    cpe_model m;
    cpe_batch b;
    const double *q;
    cpe_eval_out eo;        // resid / jac are T* (float for the f32 flavour)
    int TPF;                // (camera, marker) pairs per frame = C * 24
    int chunks;             // ceil(N / WPC)
    int n_u;                // packed chain columns per marker set (198 for the cheetah)
    int n_pair;             // n_u / 3 (marker, link) pairs
    int nitem;              // 24 + n_pair Jacobian items (3 columns each) per frame.view
    unsigned item_magic;    // ceil(2^24 / nitem): g / nitem == (g * item_magic) >> 24 for g < C * nitem
    int wbytes;             // bytes of one warp's shared slice
    int off_warp;           // byte offset of the first warp slice (after the camera rows)
    double ea, eb, ec;      // exp(loss_a), exp(loss_b), exp(loss_c)
    double icb;             // 1 / (loss_c - loss_b)
};

// per-warp FP64 slice (doubles): q, sin/cos, rotations
constexpr int QO = 0, SCO = 56, RO = 160, WD = 314;
// compact FP64 layout (WPC = 7, 4 CTAs / SM = 28 warps): the marker offsets and camera translations, which are only
// written at the very end of phase C, move into doubles that are dead by then -- pos behind the phase-D projection
// derivatives ([248, 320): the tail of the rotations), tc over q ([0, 24)) -- and the T slice holds U alone
constexpr int WD_C = 320, POS_C = 248;
// per-warp T slice (elements): marker offsets, camera translations, U = 3x3 blocks padded to 10 (block 0 = identity,
// block 1 + k = dR/d(phi,theta,psi) . v of (marker, link) pair k); the f32 flavour also keeps a float copy of sin/cos
// in front.  The d(u,v)/d(point) rows of the 32 pairs of a round (6 T each) reuse the sin/cos + rotation doubles,
// which are dead after phase C.
constexpr int T_POS = 0, T_TC = 72, T_U = 96;

template <typename T> struct Vec;
template <> struct Vec<double> { using t2 = double2; };
template <> struct Vec<float> { using t2 = float2; };

__device__ __forceinline__ double2 mk2(double a, double b) { return make_double2(a, b); }
__device__ __forceinline__ float2 mk2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ double r_exp(double x) { return exp(x); }
__device__ __forceinline__ float r_exp(float x) { return expf(x); }

// Redescending loss (acinoset_misc.py:2001-2015), value only: sigma(E - s) = 1 / (1 + exp(s) exp(-E)) needs one exp per
// residual.  Both (u, v) residuals of one detection at once: ONE division serves both (1 / (P_u P_v), P = da db dc <= 2.2e14
// each, so the product stays far inside the range of either type) and (c - E) / (c - b) is a multiplication by the
// host-side reciprocal -- 2 divisions and 2 exps per detection instead of 6 and 2.
template <typename T>
__device__ __forceinline__ T redescending_pair(T Eu, T Ev, T a, T b, T c, T icb, T ea, T eb, T ec) {
    const T tu = r_exp(-Eu), tv = r_exp(-Ev);
    const T dau = T(1) + tu * ea, dbu = T(1) + tu * eb, dcu = T(1) + tu * ec;
    const T dav = T(1) + tv * ea, dbv = T(1) + tv * eb, dcv = T(1) + tv * ec;
    const T dabu = dau * dbu, dabv = dav * dbv;
    const T Pu = dabu * dcu, Pv = dabv * dcv;
    const T inv = T(1) / (Pu * Pv);
    const T iu = inv * Pv, iv = inv * Pu;
    const T sau = iu * (dbu * dcu), sbu = iu * (dau * dcu), scu = iu * dabu;
    const T sav = iv * (dbv * dcv), sbv = iv * (dav * dcv), scv = iv * dabv;
    const T cb = c - b;
    const T k2 = T(0.5) * a * a, k3 = a * b - k2, hk = T(0.5) * a * cb, k4 = k3 + hk;
    const T qu = (c - Eu) * icb, qv = (c - Ev) * icb;
    const T p3u = k3 + hk * (T(1) - qu * qu), p3v = k3 + hk * (T(1) - qv * qv);
    const T ru = (T(1) - sau) * T(0.5) * Eu * Eu + (sau - sbu) * (a * Eu - k2) + (sbu - scu) * p3u + scu * k4;
    const T rv = (T(1) - sav) * T(0.5) * Ev * Ev + (sav - sbv) * (a * Ev - k2) + (sbv - scv) * p3v + scv * k4;
    return ru + rv;
}

template <typename T, bool JAC, int WPC, int MINB>
__global__ void __launch_bounds__(WPC * 32, MINB) eval_kernel(const __grid_constant__ EvalParams p) {
    using T2 = typename Vec<T>::t2;
    constexpr bool F64 = sizeof(T) == 8;
    constexpr int TS0 = F64 ? 0 : 104;                     // f32: float copy of sin/cos in front of the T slice
    constexpr bool CMP = F64 && WPC == 7;                  // compact layout (see WD_C)
    extern __shared__ __align__(16) unsigned char smraw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.b.N, C = p.b.C, W = p.b.W, TPF = p.TPF;
    const int s = blockIdx.x / p.chunks;
    const int k = (blockIdx.x - s * p.chunks) * WPC + warp;
    if (p.b.active && !p.b.active[s]) return;              // converged sequence of an LM loop (uniform per CTA)
    double *camd = reinterpret_cast<double *>(smraw);      // [C*24] FP64 camera rows
    T *camt = F64 ? reinterpret_cast<T *>(camd) : reinterpret_cast<T *>(smraw + sizeof(double) * C * CPE_CAM_STRIDE);
    int32_t *coltab = reinterpret_cast<int32_t *>(smraw + p.off_warp) - ((p.nitem + 3) & ~3);   // [nitem] Jacobian items
    for (int i = tid; i < C * CPE_CAM_STRIDE; i += WPC * 32) {
        const double v = p.b.cams[(size_t)s * C * CPE_CAM_STRIDE + i];
        camd[i] = v;
        if (!F64) camt[i] = (T)v;
    }
    int32_t *goff = coltab - 28;                           // [25] first Jacobian item of marker l: l + u_off[l] / 3
    if (JAC) {
        for (int i = tid; i < p.nitem; i += WPC * 32) coltab[i] = p.m.col_item[i];
        if (tid < 25) goff[tid] = tid + p.m.u_off[tid] / 3;
    }
    __syncthreads();                                       // the only CTA-wide barrier
    const int nf = p.b.n_frames[s];
    if (k >= nf) return;
    const int TPFs = p.b.n_cams ? min(p.b.n_cams[s], C) * 24 : TPF;     // pairs of the real cameras (ragged batch)
    double *wd = reinterpret_cast<double *>(smraw + p.off_warp + warp * p.wbytes);
    T *wt = reinterpret_cast<T *>(wd + (CMP ? WD_C : WD));
    const T *sct = F64 ? reinterpret_cast<const T *>(wd + SCO) : wt;
    T *pos = CMP ? reinterpret_cast<T *>(wd + POS_C) : wt + TS0 + T_POS;
    T *tcv = CMP ? reinterpret_cast<T *>(wd + QO) : wt + TS0 + T_TC;
    T *U = CMP ? wt : wt + TS0 + T_U;
    T *dp = reinterpret_cast<T *>(wd + SCO);               // [32][6] per-round projection derivatives (phase D)
    const int fi = s * N + k;                              // global frame index
    const int n_ineq = p.m.n_ineq;
    const size_t m0 = (size_t)fi * TPF;                    // detection record of (camera 0, marker 0), in units of W

    // ---- A: q -> shared; first detections in flight ----------------------------------------------------------------
    MeasRaw cur;
    cur.d = make_double4(0.0, 0.0, 0.0, 0.0);
    if (lane < TPF) meas_fetch(p.b, (m0 + lane) * W, cur);
    const double *qf = p.q + (size_t)fi * 54;
    for (int c = lane; c < 54; c += 32) wd[QO + c] = qf[c];
    __syncwarp();
    // ---- B: sin/cos of the 51 angles, then the 17 link rotations (FP64) --------------------------------------------
    for (int a = lane; a < 51; a += 32) {
        double sn, cs;
        sincos(wd[QO + 3 + a], &sn, &cs);
        wd[SCO + 2 * a] = sn; wd[SCO + 2 * a + 1] = cs;
        if (!F64) { wt[2 * a] = (T)sn; wt[2 * a + 1] = (T)cs; }
    }
    __syncwarp();
    if (lane < 17) rot_from_sc(wd + SCO + 6 * lane, wd + RO + 9 * lane);
    __syncwarp();

    // ---- C: camera-independent per-frame items, one warp-uniform round per kind ---------------------------------------
    double a_eq = 0.0, a_in = 0.0, a_mo = 0.0;
    {
        const double *cvec = p.b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
        const int32_t *chain_link = p.m.chain_link;
        if (JAC) {
            // U: identity columns, then one block per (marker, chain link): 3 angles x 3 components
            if (lane < 10) U[lane] = (lane == 0 || lane == 4 || lane == 8) ? T(1) : T(0);
            for (int it = lane; it < p.n_pair; it += 32) {
                const int code = p.m.u_item[3 * it];
                const int l = code >> 8, j = (code >> 4) & 15;
                const int link = chain_link[l * CPE_MAX_CHAIN + j];
                const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                T o[9];
                drot_all<T>(sct + 6 * link, (T)v[0], (T)v[1], (T)v[2], o);
                T *u = U + 10 + 10 * it;                 // rows = angle, columns = xyz
#pragma unroll
                for (int i = 0; i < 9; ++i) u[i] = o[i];
            }
        }
        double px = 0.0, py = 0.0, pz = 0.0, tcr = 0.0;   // written to shared memory after the last read of q and R
        if (lane < 24) {
            // marker offset from the base origin (FP64 accumulate; the FP32 flavour never sees world metres)
            const int l = lane;
            double x = 0.0, y = 0.0, z = 0.0;
#pragma unroll
            for (int j = 0; j < CPE_MAX_CHAIN; ++j) {
                const int link = chain_link[l * CPE_MAX_CHAIN + j];
                if (link >= 0) {
                    const double *Rl = wd + RO + link * 9;
                    const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                    const double v0 = v[0], v1 = v[1], v2 = v[2];
                    x += Rl[0] * v0 + Rl[1] * v1 + Rl[2] * v2;
                    y += Rl[3] * v0 + Rl[4] * v1 + Rl[5] * v2;
                    z += Rl[6] * v0 + Rl[7] * v1 + Rl[8] * v2;
                }
            }
            px = x; py = y; pz = z;
            if (p.eo.pose) {
                double *o = p.eo.pose + ((size_t)fi * 24 + l) * 3;
                o[0] = wd[QO + 0] + x; o[1] = wd[QO + 1] + y; o[2] = wd[QO + 2] + z;
            }
        }
        if (lane < 3 * C) {
            // base origin in the camera frame: tc = R_c base + t_c (FP64)
            const int c = lane / 3, r = lane - 3 * c;
            const double *cm = camd + c * CPE_CAM_STRIDE;
            double bx = wd[QO + 0], by = wd[QO + 1], bz = wd[QO + 2];
            if (p.b.shutter_tau) {
                // shutter delay: camera c sees the frame displaced by dq tau_c + ddq tau_c^2 (acinoset_misc.py:280-286)
                const double tau = p.b.shutter_tau[(size_t)s * C + c];
                const double *vb = p.b.base_dq + (size_t)fi * 3, *ab = p.b.base_ddq + (size_t)fi * 3;
                bx = bx + vb[0] * tau + ab[0] * (tau * tau);
                by = by + vb[1] * tau + ab[1] * (tau * tau);
                bz = bz + vb[2] * tau + ab[2] * (tau * tau);
            }
            tcr = cm[8 + 3 * r] * bx + cm[9 + 3 * r] * by + cm[10 + 3 * r] * bz + cm[17 + r];
        }
        if (lane < 26) {
            const int e = lane;
            const int4 pr4 = reinterpret_cast<const int4 *>(p.m.eq_pairs)[e];
            const double *Ra = wd + RO + pr4.x * 9, *Rb = wd + RO + pr4.y * 9;
            const int ia = pr4.z, ib = pr4.w;
            const double c = Ra[ia] * Rb[ib] + Ra[3 + ia] * Rb[3 + ib] + Ra[6 + ia] * Rb[6 + ib];
            const double lam = p.b.lam_eq ? p.b.lam_eq[(size_t)fi * 26 + e] : 0.0;
            const double t = c + lam / p.b.mu_eq;
            a_eq = 0.5 * p.b.mu_eq * t * t;
            if (p.eo.c_eq) p.eo.c_eq[(size_t)fi * 26 + e] = c;
        }
        if (lane < n_ineq) {
            const int j = lane;
            const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
            const double lo = p.m.ineq_bounds[2 * j], up = p.m.ineq_bounds[2 * j + 1];
            const double d = wd[QO + ip] - (im >= 0 ? wd[QO + im] : 0.0);
            const double lam = p.b.lam_in ? p.b.lam_in[(size_t)fi * n_ineq + j] : 0.0;
            const double t = d + lam / p.b.mu_in;
            const double v = t - fmin(fmax(t, lo), up);
            a_in = 0.5 * p.b.mu_in * v * v;
            if (p.eo.v_in) p.eo.v_in[(size_t)fi * n_ineq + j] = v;
        }
        if (k >= 3) {
            const double h = p.b.h[s];
            const double ih2 = 1.0 / (h * h);
            for (int c = lane; c < 54; c += 32) {
                const double w = p.m.wq[c];
                if (w != 0.0) {
                    const double d3 = (wd[QO + c] - 3.0 * qf[c - 54] + 3.0 * qf[c - 108] - qf[c - 162]) * ih2;
                    a_mo += w * d3 * d3;
                }
            }
        }
        if (CMP) __syncwarp();                             // pos / tc overwrite the rotations / q
        if (lane < 24) { pos[lane * 3 + 0] = (T)px; pos[lane * 3 + 1] = (T)py; pos[lane * 3 + 2] = (T)pz; }
        if (lane < 3 * C) tcv[lane] = (T)tcr;
    }
    __syncwarp();

    // ---- D: rounds of 32 (camera, marker) pairs: projection, residuals, robust loss, Jacobian columns -----------------
    T rho = T(0);
    {
        const T la_ = (T)p.b.loss_a, lb_ = (T)p.b.loss_b, lc_ = (T)p.b.loss_c;
        const T ea = (T)p.ea, eb = (T)p.eb, ec = (T)p.ec, icb = (T)p.icb;
        for (int p0 = 0; p0 < TPF; p0 += 32) {
            const int pi = p0 + lane;
            MeasRaw nxt;
            nxt.d = make_double4(0.0, 0.0, 0.0, 0.0);
            if (pi + 32 < TPF) meas_fetch(p.b, (m0 + pi + 32) * W, nxt);        // next round's detections in flight
            if (pi < TPF) {
                const int mc = pi / 24, ml = pi - mc * 24;
                const T *cm = camt + mc * CPE_CAM_STRIDE;
                const double *cmd = camd + mc * CPE_CAM_STRIDE;
                const T dx = pos[3 * ml], dy = pos[3 * ml + 1], dz = pos[3 * ml + 2];
                const T *tc = tcv + 3 * mc;
                const T xc = cm[8] * dx + cm[9] * dy + cm[10] * dz + tc[0];
                const T yc = cm[11] * dx + cm[12] * dy + cm[13] * dz + tc[1];
                const T zc = cm[14] * dx + cm[15] * dy + cm[16] * dz + tc[2];
                ProjT<T> pr;
                project_cam<T, JAC>(xc, yc, zc, cm, p.b.cam_model, pr);
                const double cmult = cmd[20], cx = cmd[2], cy = cmd[3];
                const size_t mi = (m0 + pi) * W;
                for (int w = 0; w < W; ++w) {
                    if (w > 0) meas_fetch(p.b, mi + w, cur);
                    double um, vm, gate;
                    meas_unpack(p.b, cur, um, vm, gate);
                    const bool sq = p.b.loss_mode == CPE_LOSS_SQUARE;
                    const T wgt = (T)(gate * p.m.meas_w[w * 24 + ml] * (sq ? 1.0 : cmult));
                    // residual = (fx a s) + (cx - u_meas): the ~1000 px offsets cancel in FP64 before rounding to T
                    const T ru = pr.pu + (T)(cx - um), rv = pr.pv + (T)(cy - vm);
                    if (p.eo.resid) reinterpret_cast<T2 *>(p.eo.resid)[mi + w] = mk2(ru, rv);
                    if (pi >= TPFs) continue;                   // padding camera: rho(0) != 0 must not be counted
                    if (sq) {
                        rho += (wgt * ru) * (wgt * ru) + (wgt * rv) * (wgt * rv);
                    } else {
                        rho += redescending_pair<T>(fabs(wgt * ru), fabs(wgt * rv), la_, lb_, lc_, icb, ea, eb, ec);
                    }
                }
                if (JAC) {
                    T2 *d2 = reinterpret_cast<T2 *>(dp + 6 * lane);
                    d2[0] = mk2(pr.du[0], pr.du[1]);
                    d2[1] = mk2(pr.du[2], pr.dv[0]);
                    d2[2] = mk2(pr.dv[1], pr.dv[2]);
                }
            }
            if (JAC) {
                // item pass: the 32 pairs of this round own ONE contiguous span of packed columns; lane <-> item of 3
                // adjacent columns, so the three stores of a round fill the span without gaps
                __syncwarp();
                const int nitem = p.nitem;
                const int pe = min(p0 + 32, TPF);
                const int c0 = p0 / 24, l0 = p0 - 24 * c0;
                const int g0 = c0 * nitem + goff[l0];
                int g1 = C * nitem;
                if (pe < TPF) {
                    const int c1 = pe / 24, l1 = pe - 24 * c1;
                    g1 = c1 * nitem + goff[l1];
                }
                T2 *jb = reinterpret_cast<T2 *>(p.eo.jac) + (size_t)fi * C * 3 * nitem;
                for (int g = g0 + lane; g < g1; g += 32) {
                    const int c = (int)(((unsigned)g * p.item_magic) >> 24);
                    const int e = coltab[g - c * nitem];
                    const int slot = c * 24 + (e >> 16) - p0;
                    const T2 *d2 = reinterpret_cast<const T2 *>(dp + 6 * slot);
                    const T2 da = d2[0], db = d2[1], dc = d2[2];       // du0 du1 | du2 dv0 | dv1 dv2
                    const T2 *u2 = reinterpret_cast<const T2 *>(U + (e & 0xffff));
                    const T2 ua = u2[0], ub = u2[1], uc = u2[2], ud = u2[3], ue = u2[4];   // u0 u1|u2 u3|u4 u5|u6 u7|u8 -
                    T2 *o = jb + 3 * (size_t)g;
                    o[0] = mk2(da.x * ua.x + da.y * ua.y + db.x * ub.x, db.y * ua.x + dc.x * ua.y + dc.y * ub.x);
                    o[1] = mk2(da.x * ub.y + da.y * uc.x + db.x * uc.y, db.y * ub.y + dc.x * uc.x + dc.y * uc.y);
                    o[2] = mk2(da.x * ud.x + da.y * ud.y + db.x * ue.x, db.y * ud.x + dc.x * ud.y + dc.y * ue.x);
                }
                __syncwarp();
            }
            cur = nxt;
        }
    }

    // ---- E: fixed-order warp reductions of the four cost terms ---------------------------------------------------------
    if (p.eo.frame_cost) {
        const double a0 = warp_sum((double)rho), a1 = warp_sum(a_mo), a2 = warp_sum(a_eq), a3 = warp_sum(a_in);
        if (lane == 0) {
            double *o = p.eo.frame_cost + (size_t)fi * 4;
            o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3;
        }
    }
}

template <typename T, bool JAC, int WPC, int MINB>
static int launch_cfg(const EvalParams &p, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(eval_kernel<T, JAC, WPC, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
    eval_kernel<T, JAC, WPC, MINB><<<grid, WPC * 32, smem, st>>>(p);
    return (int)cudaGetLastError();
}

static int eval_cfg() {
    static int cfg = -1;
    if (cfg < 0) {
        const char *e = getenv("CPE_EVAL_CFG");      // tuning knob for experiments; default chosen by measurement
        cfg = e ? atoi(e) : 1;      // 1: 6 CTAs/SM (80 regs) for FP64, 8 CTAs/SM (64 regs) for FP32
                                    // 3: FP64 compact layout, 7 warps / CTA x 4 CTAs / SM (72 regs); FP32 as 1
    }
    return cfg;
}

template <typename T>
static int launch_eval(const cpe_model *model, const cpe_batch *b, const double *q, const cpe_eval_out *eo,
                       cudaStream_t st) {
    if (!model || !b || !q || !eo) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS || (b->W != 1 && b->W != 3)) return CPE_E_DIM;
    if (model->n_ineq > CPE_MAX_INEQ || model->n_u <= 0 || model->n_u > 240 || model->n_u % 3) return CPE_E_DIM;
    if (b->meas_layout < 0 || b->meas_layout > CPE_MEAS_F32X2B) return CPE_E_ARG;
    if (b->meas_layout >= CPE_MEAS_F32X2G && !b->meas_gate) return CPE_E_NULL;
    if (!b->meas || !b->cams || !b->chain_vec || !b->n_frames || !b->h) return CPE_E_NULL;
    if (!model->u_off || !model->u_item) return CPE_E_NULL;
    if ((b->shutter_tau || b->base_dq || b->base_ddq) && !(b->shutter_tau && b->base_dq && b->base_ddq)) return CPE_E_NULL;
    constexpr bool f64 = sizeof(T) == 8;
    const int cfg = eval_cfg();
    const bool cmp = f64 && cfg == 3;                    // compact layout: 7 warps / CTA
    const int WPC = cmp ? 7 : 4;
    const bool jac = eo->jac != nullptr;
    EvalParams p;
    p.m = *model; p.b = *b; p.q = q; p.eo = *eo;
    p.ea = exp(b->loss_a); p.eb = exp(b->loss_b); p.ec = exp(b->loss_c);
    p.icb = 1.0 / (b->loss_c - b->loss_b);
    p.TPF = b->C * 24;
    p.chunks = (b->N + WPC - 1) / WPC;
    p.n_u = model->n_u;
    p.n_pair = model->n_u / 3;
    p.nitem = 24 + p.n_pair;
    p.item_magic = ((1u << 24) + p.nitem - 1) / p.nitem;
    if (jac && !model->col_item) return CPE_E_NULL;
    const int t_elems = cmp ? 10 + 10 * p.n_pair : (f64 ? 0 : 104) + T_U + 10 + 10 * p.n_pair;
    p.wbytes = (int)(((cmp ? WD_C : WD) * sizeof(double) + t_elems * sizeof(T) + 15) & ~(size_t)15);
    p.off_warp = (int)(b->C * CPE_CAM_STRIDE * (sizeof(double) + (f64 ? 0 : sizeof(T))) +
                       sizeof(int32_t) * (28 + ((p.nitem + 3) & ~3)));
    const size_t smem = (size_t)p.off_warp + (size_t)WPC * p.wbytes;
    const long long grid_ll = (long long)b->S * p.chunks;
    if (grid_ll > 0x7fffffffLL) return CPE_E_DIM;
    const int grid = (int)grid_ll;
    if constexpr (f64) {
        if (cfg == 3) return jac ? launch_cfg<T, true, 7, 4>(p, grid, smem, st) : launch_cfg<T, false, 7, 4>(p, grid, smem, st);
        if (cfg == 0) return jac ? launch_cfg<T, true, 4, 7>(p, grid, smem, st) : launch_cfg<T, false, 4, 7>(p, grid, smem, st);
        if (cfg == 2) return jac ? launch_cfg<T, true, 4, 5>(p, grid, smem, st) : launch_cfg<T, false, 4, 5>(p, grid, smem, st);
        return jac ? launch_cfg<T, true, 4, 6>(p, grid, smem, st) : launch_cfg<T, false, 4, 6>(p, grid, smem, st);
    } else {
        if (cfg == 0) return jac ? launch_cfg<T, true, 4, 10>(p, grid, smem, st) : launch_cfg<T, false, 4, 10>(p, grid, smem, st);
        if (cfg == 2) return jac ? launch_cfg<T, true, 4, 12>(p, grid, smem, st) : launch_cfg<T, false, 4, 12>(p, grid, smem, st);
        return jac ? launch_cfg<T, true, 4, 8>(p, grid, smem, st) : launch_cfg<T, false, 4, 8>(p, grid, smem, st);
    }
}

}  // namespace cpe

extern "C" int cpe_eval_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const cpe_eval_out *out,
                            void *stream) {
    return cpe::launch_eval<double>(model, batch, q, out, (cudaStream_t)stream);
}

extern "C" int cpe_eval_f32(const cpe_model *model, const cpe_batch *batch, const double *q,
                            const cpe_eval_out_f32 *out, void *stream) {
    if (!out) return CPE_E_NULL;
    cpe_eval_out eo;                                    // same slots; resid / jac carry float data
    eo.frame_cost = out->frame_cost; eo.pose = out->pose; eo.c_eq = out->c_eq; eo.v_in = out->v_in;
    eo.resid = reinterpret_cast<double *>(out->resid);
    eo.jac = reinterpret_cast<double *>(out->jac);
    return cpe::launch_eval<float>(model, batch, q, &eo, (cudaStream_t)stream);
}
