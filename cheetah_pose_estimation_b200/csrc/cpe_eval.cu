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
// Mapping (sm_100a): one CTA owns F consecutive frames of ONE sequence (F*C*24 <= 160 threads).  Phases:
//   A  q, cameras -> shared; detections prefetched into registers (coalesced)
//   B  one thread per link: 3 sincos + rotation matrix (FP64)
//   C  warp-uniform item lists: chain derivative blocks U = dR/dangle . v (one item per (marker, link): 3 angles),
//      marker offsets, camera translations, joint equalities, inequalities, motion terms
//   D  one thread per (frame, camera, marker): projection, residuals, loss, Jacobian columns into a shared staging
//      tile laid out exactly like the packed global Jacobian
//   E  the tile leaves the SM as ONE bulk async copy (cp.async.bulk, the TMA engine) while warps reduce the costs.
// No atomics: every reduction has a fixed order, so a sequence's results are bit-identical whatever its batch
// neighbours are.
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
    int F;                  // frames per CTA
    int TPF;                // threads per frame = C * 24
    int chunks;             // ceil(N / F)
    int n_u;                // packed chain columns per marker set (198 for the cheetah)
    int n_pair;             // n_u / 3 (marker, link) pairs
    int i_mk, i_tc, i_eq, i_in, i_mo, i_end;   // warp-aligned item list boundaries of phase C
    int fd;                 // per-frame FP64 shared stride (doubles)
    int ft;                 // per-frame T shared stride (elements)
    int off_t;              // byte offset of the T region
    int off_stage;          // byte offset of the Jacobian staging tile
    double ea, eb, ec;      // exp(loss_a), exp(loss_b), exp(loss_c)
};

// per-frame FP64 shared layout (doubles): q, rotations, cost contributions eq[26] in[32] mo[54] me[TPF]
constexpr int QO = 0, RO = 56, CTR = 210;
constexpr int EQC = 0, INC = 26, MOC = 58, MEC = 112;
// per-frame T shared layout (elements): sin/cos, marker offsets, camera translations, U blocks
constexpr int T_SC = 0, T_POS = 104, T_TC = 176, T_U = 200;

template <typename T> struct Vec;
template <> struct Vec<double> { using t2 = double2; static constexpr int UPAD = 8; };
template <> struct Vec<float> { using t2 = float2; static constexpr int UPAD = 8; };

__device__ __forceinline__ double2 mk2(double a, double b) { return make_double2(a, b); }
__device__ __forceinline__ float2 mk2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ double r_exp(double x) { return exp(x); }
__device__ __forceinline__ float r_exp(float x) { return expf(x); }

// Redescending loss value only (acinoset_misc.py:2001-2015); one exp and one division per residual.
template <typename T>
__device__ __forceinline__ T redescending_value(T E, T a, T b, T c, T ea, T eb, T ec) {
    const T t = r_exp(-E);
    const T da = T(1) + t * ea, db = T(1) + t * eb, dc = T(1) + t * ec;
    const T dab = da * db;
    const T inv = T(1) / (dab * dc);
    const T sa = inv * (db * dc), sb = inv * (da * dc), sc = inv * dab;
    const T cb = c - b;
    const T p2 = a * E - T(0.5) * a * a;
    const T qq = (c - E) / cb;
    const T p3 = a * b - T(0.5) * a * a + (T(0.5) * a * cb) * (T(1) - qq * qq);
    const T k4 = a * b - T(0.5) * a * a + T(0.5) * a * cb;
    return (T(1) - sa) * T(0.5) * E * E + (sa - sb) * p2 + (sb - sc) * p3 + sc * k4;
}

template <typename T, bool JAC, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) eval_kernel(const __grid_constant__ EvalParams p) {
    using T2 = typename Vec<T>::t2;
    constexpr int UPAD = Vec<T>::UPAD;
    constexpr bool F64 = sizeof(T) == 8;
    extern __shared__ __align__(16) unsigned char smraw[];
    const int tid = threadIdx.x;
    const int N = p.b.N, C = p.b.C, W = p.b.W, F = p.F, TPF = p.TPF;
    const int s = blockIdx.x / p.chunks;
    const int k0 = (blockIdx.x - s * p.chunks) * F;
    const int nf = p.b.n_frames[s];
    if (k0 >= nf) return;
    const int nv = min(F, nf - k0);                       // valid frames of this CTA
    const int n_ineq = p.m.n_ineq, n_u = p.n_u;
    double *camd = reinterpret_cast<double *>(smraw);      // [C*24] FP64 camera rows
    double *frd0 = camd + C * CPE_CAM_STRIDE;              // per-frame FP64 regions
    T *camt = F64 ? reinterpret_cast<T *>(camd) : reinterpret_cast<T *>(smraw + p.off_t);   // [C*24] in T
    T *frt0 = reinterpret_cast<T *>(smraw + p.off_t) + (F64 ? 0 : C * CPE_CAM_STRIDE);
    T *stage = reinterpret_cast<T *>(smraw + p.off_stage);  // [F][C][540]
    const int f0 = s * N + k0;                            // global index of the first frame
    const int jw = 2 * (72 + n_u);                        // packed Jacobian elements per frame.view

    // Every thread belongs to ONE frame of the CTA in phases A and D: frame mf, local index mt in [0, TPF).
    int mf = 0, mt = tid;
    if (F > 1) { mf = tid / TPF; mt = tid - mf * TPF; }
    const bool mvalid = mf < nv && mt < TPF;
    const int fidx = f0 + mf;
    const int mc = mt / 24, ml = mt - mc * 24;
    const size_t midx0 = ((size_t)(fidx * C + mc) * 24 + ml) * W;
    // ---- A: issue the detection load before the FK phases; q and cameras to shared --------------------------------
    MeasRaw mraw;
    mraw.d = make_double4(0.0, 0.0, 0.0, 0.0);
    if (mvalid) meas_fetch(p.b, midx0, mraw);
    for (int i = tid; i < nv * 54; i += blockDim.x) {
        const int f = i / 54, c = i - f * 54;
        frd0[f * p.fd + QO + c] = p.q[(size_t)(f0 + f) * 54 + c];
    }
    for (int i = tid; i < C * CPE_CAM_STRIDE; i += blockDim.x) {
        const double v = p.b.cams[(size_t)s * C * CPE_CAM_STRIDE + i];
        camd[i] = v;
        if (!F64) camt[i] = (T)v;
    }
    __syncthreads();
    // ---- B: one thread per (frame, link): sincos of its three angles + rotation matrix, FP64 ------------------------
    if (tid < nv * 17) {
        const int f = tid / 17, link = tid - f * 17;
        double *frd = frd0 + f * p.fd;
        T *frt = frt0 + f * p.ft;
        double sc[6];
        sincos(frd[QO + 3 + 3 * link + 0], &sc[0], &sc[1]);
        sincos(frd[QO + 3 + 3 * link + 1], &sc[2], &sc[3]);
        sincos(frd[QO + 3 + 3 * link + 2], &sc[4], &sc[5]);
        double M[9];
        rot_from_sc(sc, M);
#pragma unroll
        for (int i = 0; i < 9; ++i) frd[RO + 9 * link + i] = M[i];
#pragma unroll
        for (int i = 0; i < 6; ++i) frt[T_SC + 6 * link + i] = (T)sc[i];
    }
    __syncthreads();

    // ---- C: camera-independent per-frame items, in warp-uniform lists ------------------------------------------------
    {
        const double *cvec = p.b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
        const int32_t *chain_link = p.m.chain_link;
        const int i_begin = JAC ? 0 : p.i_mk;
        const int per = p.i_end - i_begin;
        for (int w0 = tid; w0 < nv * per; w0 += blockDim.x) {
            int f = 0, it = w0;
            if (F > 1) { f = w0 / per; it = w0 - f * per; }
            it += i_begin;
            double *frd = frd0 + f * p.fd;
            T *frt = frt0 + f * p.ft;
            double *ctr = frd + CTR;
            const int fi = f0 + f;
            if (it < p.i_mk) {
                // U block of one (marker, chain link): 3 angles x 3 components (8 stored)
                if (it < p.n_pair) {
                    const int code = p.m.u_item[3 * it];
                    const int l = code >> 8, j = (code >> 4) & 15;
                    const int link = chain_link[l * CPE_MAX_CHAIN + j];
                    const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                    T o[9];
                    drot_all<T>(frt + T_SC + 6 * link, (T)v[0], (T)v[1], (T)v[2], o);
                    T *u = frt + T_U + UPAD * it;                 // o[8] = dz/dpsi = 0 is not stored
#pragma unroll
                    for (int i = 0; i < 8; ++i) u[i] = o[i];
                }
            } else if (it < p.i_tc) {
                const int l = it - p.i_mk;
                if (l < 24) {
                    // marker offset from the base origin (FP64 accumulate; the FP32 flavour never sees world metres)
                    double x = 0.0, y = 0.0, z = 0.0;
#pragma unroll
                    for (int j = 0; j < CPE_MAX_CHAIN; ++j) {
                        const int link = chain_link[l * CPE_MAX_CHAIN + j];
                        if (link >= 0) {
                            const double *Rl = frd + RO + link * 9;
                            const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                            const double v0 = v[0], v1 = v[1], v2 = v[2];
                            x += Rl[0] * v0 + Rl[1] * v1 + Rl[2] * v2;
                            y += Rl[3] * v0 + Rl[4] * v1 + Rl[5] * v2;
                            z += Rl[6] * v0 + Rl[7] * v1 + Rl[8] * v2;
                        }
                    }
                    frt[T_POS + l * 3 + 0] = (T)x; frt[T_POS + l * 3 + 1] = (T)y; frt[T_POS + l * 3 + 2] = (T)z;
                    if (p.eo.pose) {
                        double *o = p.eo.pose + ((size_t)fi * 24 + l) * 3;
                        o[0] = frd[QO + 0] + x; o[1] = frd[QO + 1] + y; o[2] = frd[QO + 2] + z;
                    }
                }
            } else if (it < p.i_eq) {
                const int e = it - p.i_tc;
                if (e < 3 * C) {
                    // base origin in the camera frame: tc = R_c base + t_c (FP64)
                    const int c = e / 3, r = e - 3 * c;
                    const double *cm = camd + c * CPE_CAM_STRIDE;
                    frt[T_TC + e] = (T)(cm[8 + 3 * r] * frd[QO + 0] + cm[9 + 3 * r] * frd[QO + 1] +
                                        cm[10 + 3 * r] * frd[QO + 2] + cm[17 + r]);
                }
            } else if (it < p.i_in) {
                const int e = it - p.i_eq;
                if (e < 26) {
                    const int4 pr4 = reinterpret_cast<const int4 *>(p.m.eq_pairs)[e];
                    const double *Ra = frd + RO + pr4.x * 9, *Rb = frd + RO + pr4.y * 9;
                    const int ia = pr4.z, ib = pr4.w;
                    const double c = Ra[ia] * Rb[ib] + Ra[3 + ia] * Rb[3 + ib] + Ra[6 + ia] * Rb[6 + ib];
                    const double lam = p.b.lam_eq ? p.b.lam_eq[(size_t)fi * 26 + e] : 0.0;
                    const double t = c + lam / p.b.mu_eq;
                    ctr[EQC + e] = 0.5 * p.b.mu_eq * t * t;
                    if (p.eo.c_eq) p.eo.c_eq[(size_t)fi * 26 + e] = c;
                }
            } else if (it < p.i_mo) {
                const int j = it - p.i_in;
                if (j < n_ineq) {
                    const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
                    const double lo = p.m.ineq_bounds[2 * j], up = p.m.ineq_bounds[2 * j + 1];
                    const double d = frd[QO + ip] - (im >= 0 ? frd[QO + im] : 0.0);
                    const double lam = p.b.lam_in ? p.b.lam_in[(size_t)fi * n_ineq + j] : 0.0;
                    const double t = d + lam / p.b.mu_in;
                    const double v = t - fmin(fmax(t, lo), up);
                    ctr[INC + j] = 0.5 * p.b.mu_in * v * v;
                    if (p.eo.v_in) p.eo.v_in[(size_t)fi * n_ineq + j] = v;
                }
            } else {
                const int c = it - p.i_mo;
                if (c < 54) {
                    const double w = p.m.wq[c];
                    double val = 0.0;
                    if (w != 0.0 && k0 + f >= 3) {
                        const double h = p.b.h[s];
                        const double *qf = p.q + (size_t)fi * 54;
                        const double d3 = (frd[QO + c] - 3.0 * qf[c - 54] + 3.0 * qf[c - 108] - qf[c - 162]) / (h * h);
                        val = w * d3 * d3;
                    }
                    ctr[MOC + c] = val;
                }
            }
        }
    }
    __syncthreads();

    // ---- D: one thread per (frame, camera, marker): projection, residuals, robust loss, Jacobian columns --------------
    if (mvalid) {
        double *frd = frd0 + mf * p.fd;
        const T *frt = frt0 + mf * p.ft;
        const T la_ = (T)p.b.loss_a, lb_ = (T)p.b.loss_b, lc_ = (T)p.b.loss_c;
        const T ea = (T)p.ea, eb = (T)p.eb, ec = (T)p.ec;
        const T *cm = camt + mc * CPE_CAM_STRIDE;
        const double *cmd = camd + mc * CPE_CAM_STRIDE;
        const T dx = frt[T_POS + 3 * ml], dy = frt[T_POS + 3 * ml + 1], dz = frt[T_POS + 3 * ml + 2];
        const T *tc = frt + T_TC + 3 * mc;
        const T xc = cm[8] * dx + cm[9] * dy + cm[10] * dz + tc[0];
        const T yc = cm[11] * dx + cm[12] * dy + cm[13] * dz + tc[1];
        const T zc = cm[14] * dx + cm[15] * dy + cm[16] * dz + tc[2];
        ProjT<T> pr;
        project_cam<T, JAC>(xc, yc, zc, cm, p.b.cam_model, pr);
        T rho = T(0);
        const double cmult = cmd[20], cx = cmd[2], cy = cmd[3];
        for (int w = 0; w < W; ++w) {
            if (w > 0) meas_fetch(p.b, midx0 + w, mraw);
            double um, vm, gate;
            meas_unpack(p.b, mraw, um, vm, gate);
            const T wgt = (T)(gate * p.m.meas_w[w * 24 + ml] * cmult);
            // residual = (fx a s) + (cx - u_meas): the ~1000 px offsets cancel in FP64 before rounding to T
            const T ru = pr.pu + (T)(cx - um), rv = pr.pv + (T)(cy - vm);
            if (p.eo.resid) reinterpret_cast<T2 *>(p.eo.resid)[midx0 + w] = mk2(ru, rv);
            rho += redescending_value<T>(fabs(wgt * ru), la_, lb_, lc_, ea, eb, ec);
            rho += redescending_value<T>(fabs(wgt * rv), la_, lb_, lc_, ea, eb, ec);
        }
        frd[CTR + MEC + mt] = (double)rho;
        if (JAC) {
            const int uo0 = p.m.u_off[ml], nlk = (p.m.u_off[ml + 1] - uo0) / 3;      // chain links of this marker
            // marker block of the packed Jacobian: (3 + 3 nlk) columns, each a (du, dv) pair
            T2 *jc = reinterpret_cast<T2 *>(stage + (size_t)(mf * C + mc) * jw + 2 * (3 * ml + uo0));
            jc[0] = mk2(pr.du[0], pr.dv[0]);
            jc[1] = mk2(pr.du[1], pr.dv[1]);
            jc[2] = mk2(pr.du[2], pr.dv[2]);
            const T *U = frt + T_U + UPAD * (uo0 / 3);
            for (int j = 0; j < nlk; ++j) {
                T u[9];
                if (F64) {
                    const double2 *U2 = reinterpret_cast<const double2 *>(U + UPAD * j);
                    const double2 a0 = U2[0], a1 = U2[1], a2 = U2[2], a3 = U2[3];
                    u[0] = a0.x; u[1] = a0.y; u[2] = a1.x; u[3] = a1.y; u[4] = a2.x; u[5] = a2.y; u[6] = a3.x; u[7] = a3.y;
                    u[8] = (T)0;
                } else {
                    const float4 *U4 = reinterpret_cast<const float4 *>(U + UPAD * j);
                    const float4 a0 = U4[0], a1 = U4[1];
                    u[0] = a0.x; u[1] = a0.y; u[2] = a0.z; u[3] = a0.w; u[4] = a1.x; u[5] = a1.y; u[6] = a1.z; u[7] = a1.w;
                    u[8] = (T)0;
                }
                T2 *o = jc + 3 + 3 * j;
                o[0] = mk2(pr.du[0] * u[0] + pr.du[1] * u[1] + pr.du[2] * u[2],
                           pr.dv[0] * u[0] + pr.dv[1] * u[1] + pr.dv[2] * u[2]);
                o[1] = mk2(pr.du[0] * u[3] + pr.du[1] * u[4] + pr.du[2] * u[5],
                           pr.dv[0] * u[3] + pr.dv[1] * u[4] + pr.dv[2] * u[5]);
                o[2] = mk2(pr.du[0] * u[6] + pr.du[1] * u[7], pr.dv[0] * u[6] + pr.dv[1] * u[7]);   // dz/dpsi = 0
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
    }
    __syncthreads();

    // ---- E: the Jacobian tile leaves as one bulk async copy; meanwhile warps reduce the cost terms ----------------
    if (JAC && tid == 0) {
        const uint32_t bytes = (uint32_t)(nv * C * jw * sizeof(T));
        T *g = reinterpret_cast<T *>(p.eo.jac) + (size_t)f0 * C * jw;
        const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(saddr), "r"(bytes)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    if (p.eo.frame_cost) {
        const int wid = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
        for (int f = wid; f < nv; f += nw) {
            const double *ctr = frd0 + f * p.fd + CTR;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            for (int i = lane; i < TPF; i += 32) a0 += ctr[MEC + i];
            for (int i = lane; i < 54; i += 32) a1 += ctr[MOC + i];
            if (lane < 26) a2 = ctr[EQC + lane];
            if (lane < n_ineq) a3 = ctr[INC + lane];
            a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); a3 = warp_sum(a3);
            if (lane == 0) {
                double *o = p.eo.frame_cost + (size_t)(f0 + f) * 4;
                o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3;
            }
        }
    }
    if (JAC && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

template <typename T, bool JAC, int MAXT, int MINB>
static int launch_cfg(const EvalParams &p, int grid, int threads, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(eval_kernel<T, JAC, MAXT, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
    eval_kernel<T, JAC, MAXT, MINB><<<grid, threads, smem, st>>>(p);
    return (int)cudaGetLastError();
}

static inline int up32(int x) { return (x + 31) & ~31; }

template <typename T>
static int launch_eval(const cpe_model *model, const cpe_batch *b, const double *q, const cpe_eval_out *eo,
                       cudaStream_t st) {
    if (!model || !b || !q || !eo) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS || (b->W != 1 && b->W != 3)) return CPE_E_DIM;
    if (model->n_ineq > CPE_MAX_INEQ || model->n_u <= 0 || model->n_u > 240 || model->n_u % 3) return CPE_E_DIM;
    if (b->meas_layout < 0 || b->meas_layout > CPE_MEAS_F32X3) return CPE_E_ARG;
    if (!b->meas || !b->cams || !b->chain_vec || !b->n_frames || !b->h) return CPE_E_NULL;
    const bool jac = eo->jac != nullptr;
    if (!model->u_off || !model->u_item) return CPE_E_NULL;
    EvalParams p;
    p.m = *model; p.b = *b; p.q = q; p.eo = *eo;
    p.ea = exp(b->loss_a); p.eb = exp(b->loss_b); p.ec = exp(b->loss_c);
    p.TPF = b->C * 24;
    int F = 160 / p.TPF;                                // frames packed into one 160-thread CTA
    if (F < 1) F = 1;
    if (F > 6) F = 6;
    if (F > b->N) F = b->N;
    p.F = F;
    p.chunks = (b->N + F - 1) / F;
    p.n_u = model->n_u;
    p.n_pair = model->n_u / 3;
    p.i_mk = up32(p.n_pair);
    p.i_tc = p.i_mk + 32;
    p.i_eq = p.i_tc + up32(3 * b->C);
    p.i_in = p.i_eq + 32;
    p.i_mo = p.i_in + up32(model->n_ineq > 0 ? model->n_ineq : 1);
    p.i_end = p.i_mo + 64;
    p.fd = (CTR + MEC + p.TPF + 1) & ~1;
    p.ft = (T_U + 8 * p.n_pair + 3) & ~3;
    const bool f64 = sizeof(T) == 8;
    p.off_t = (int)sizeof(double) * (b->C * CPE_CAM_STRIDE + F * p.fd);
    const size_t t_elems = (size_t)(f64 ? 0 : b->C * CPE_CAM_STRIDE) + (size_t)F * p.ft;
    p.off_stage = (int)((p.off_t + t_elems * sizeof(T) + 15) & ~(size_t)15);
    const size_t smem = (size_t)p.off_stage + (jac ? (size_t)F * b->C * 2 * (72 + p.n_u) * sizeof(T) : 0);
    int threads = up32(F * p.TPF);
    if (threads < 128) threads = 128;
    const long long grid_ll = (long long)b->S * p.chunks;
    if (grid_ll > 0x7fffffffLL) return CPE_E_DIM;
    const int grid = (int)grid_ll;
    if (threads > 160)
        return jac ? launch_cfg<T, true, 192, 4>(p, grid, threads, smem, st)
                   : launch_cfg<T, false, 192, 4>(p, grid, threads, smem, st);
    constexpr int MINB = sizeof(T) == 8 ? 6 : 8;
    return jac ? launch_cfg<T, true, 160, MINB>(p, grid, threads, smem, st)
               : launch_cfg<T, false, 160, MINB>(p, grid, threads, smem, st);
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
