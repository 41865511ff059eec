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
// Mapping (sm_100a, FP64): one CTA owns F consecutive frames of ONE sequence (F*C*24 ~ 288 threads).  The camera
// independent part (sincos, link rotations and their angle derivatives, marker positions, chain derivative vectors
// U = dR/dangle . v) is computed once per frame into shared memory; then one thread per (frame, camera, marker)
// projects, forms the residuals / robust loss, and writes its Jacobian rows J = dproj(2x3) . [I | U] into a shared
// staging tile that is laid out exactly like the packed global Jacobian, so the whole tile leaves the SM as ONE
// bulk async copy (cp.async.bulk, the TMA engine) while the warps reduce the cost terms.  Detections are fetched
// with coalesced float4 loads issued before the FK phase.  No atomics: every reduction has a fixed order, so a
// sequence's results are bit-identical whatever its batch neighbours are.
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
    cpe_eval_out eo;
    int F;          // frames per CTA
    int TPF;        // threads per frame = C * 24
    int chunks;     // ceil(N / F)
    int n_u;        // packed chain columns per marker set (198 for the cheetah)
    int fs;         // per-frame shared stride (doubles)
    int off_stage;  // offset (doubles) of the Jacobian staging tile
    double ea, eb, ec;  // exp(loss_a), exp(loss_b), exp(loss_c)
};

// per-frame shared layout (doubles)
constexpr int QO = 0, SCO = 56, RO = 160, POSO = 314, UO = 386;
__host__ __device__ inline int contrib_off(int n_u) { return UO + 10 * (n_u / 3); }   // U: 10 doubles per (marker, link)
// contributions: eq[26] in[32] mo[54] me[TPF]
constexpr int EQC = 0, INC = 26, MOC = 58, MEC = 112;

// Redescending loss value only (acinoset_misc.py:2001-2015); one exp and one division per residual.
__device__ __forceinline__ double redescending_value(double E, double a, double b, double c, double ea, double eb,
                                                      double ec) {
    const double t = exp(-E);
    const double da = 1.0 + t * ea, db = 1.0 + t * eb, dc = 1.0 + t * ec;
    const double dab = da * db;
    const double inv = 1.0 / (dab * dc);
    const double sa = inv * (db * dc), sb = inv * (da * dc), sc = inv * dab;
    const double cb = c - b;
    const double p2 = a * E - 0.5 * a * a;
    const double qq = (c - E) / cb;
    const double p3 = a * b - 0.5 * a * a + (0.5 * a * cb) * (1.0 - qq * qq);
    const double k4 = a * b - 0.5 * a * a + 0.5 * a * cb;
    return (1.0 - sa) * 0.5 * E * E + (sa - sb) * p2 + (sb - sc) * p3 + sc * k4;
}

template <bool JAC, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) eval_kernel(const __grid_constant__ EvalParams p) {
    extern __shared__ __align__(16) double sm[];
    const int tid = threadIdx.x;
    const int N = p.b.N, C = p.b.C, W = p.b.W, F = p.F, TPF = p.TPF;
    const int s = blockIdx.x / p.chunks;
    const int k0 = (blockIdx.x - s * p.chunks) * F;
    const int nf = p.b.n_frames[s];
    if (k0 >= nf) return;
    const int nv = min(F, nf - k0);                       // valid frames of this CTA
    const int n_ineq = p.m.n_ineq, n_u = p.n_u, n_ul = n_u / 3;
    const int FS = p.fs, CO = contrib_off(n_u);
    double *cam = sm;                                      // [C*24]
    double *stage = sm + p.off_stage;                      // [F][C][540]
    const size_t f0 = (size_t)s * N + k0;                 // global index of the first frame

    // Every thread belongs to ONE frame of the CTA for all phases: frame mf, local index mt in [0, TPF).
    const int mf = tid / TPF, mt = tid - mf * TPF;
    const bool mvalid = mf < nv;
    double *fr = sm + C * CPE_CAM_STRIDE + mf * FS;
    const size_t fidx = f0 + mf;
    const int mc = mt / 24, ml = mt - mc * 24;
    const size_t midx0 = ((fidx * C + mc) * 24 + ml) * W;
    // issue the detection load before the FK phases
    float4 m4 = make_float4(0.f, 0.f, 0.f, 0.f);
    double4 m8 = make_double4(0.0, 0.0, 0.0, 0.0);
    const bool m64 = p.b.meas_is_f64 != 0;
    if (mvalid) {
        if (m64) m8 = reinterpret_cast<const double4 *>(p.b.meas)[midx0];
        else m4 = __ldg(reinterpret_cast<const float4 *>(p.b.meas) + midx0);
    }
    const double *qf = p.q + fidx * 54;
    if (mvalid) {
        for (int c = mt; c < 54; c += TPF) fr[QO + c] = qf[c];
    }
    for (int i = tid; i < C * CPE_CAM_STRIDE; i += blockDim.x) cam[i] = p.b.cams[(size_t)s * C * CPE_CAM_STRIDE + i];
    __syncthreads();
    if (mvalid) {
        for (int a = mt; a < 51; a += TPF) sincos(fr[QO + 3 + a], &fr[SCO + 2 * a], &fr[SCO + 2 * a + 1]);
    }
    __syncthreads();
    if (mvalid) {
        for (int link = mt; link < 17; link += TPF) rot_from_sc(fr + SCO + 6 * link, fr + RO + 9 * link);
    }
    __syncthreads();

    // ---- camera-independent per-frame items -------------------------------------------------------------------
    if (mvalid) {
        const double *cvec = p.b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
        const int32_t *chain_link = p.m.chain_link;
        const int n_uit = JAC ? n_u : 0;
        const int n_items = 24 + n_uit + 26 + n_ineq + 54;
        double *ctr = fr + CO;
        for (int it0 = mt; it0 < n_items; it0 += TPF) {
            int it = it0;
            if (it < 24) {
                const int l = it;
                double x = fr[QO + 0], y = fr[QO + 1], z = fr[QO + 2];
#pragma unroll
                for (int j = 0; j < CPE_MAX_CHAIN; ++j) {
                    const int link = chain_link[l * CPE_MAX_CHAIN + j];
                    if (link >= 0) {
                        const double *Rl = fr + RO + link * 9;
                        const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                        const double v0 = v[0], v1 = v[1], v2 = v[2];
                        x += Rl[0] * v0 + Rl[1] * v1 + Rl[2] * v2;
                        y += Rl[3] * v0 + Rl[4] * v1 + Rl[5] * v2;
                        z += Rl[6] * v0 + Rl[7] * v1 + Rl[8] * v2;
                    }
                }
                fr[POSO + l * 3 + 0] = x; fr[POSO + l * 3 + 1] = y; fr[POSO + l * 3 + 2] = z;
                if (p.eo.pose) {
                    double *o = p.eo.pose + (fidx * 24 + l) * 3;
                    o[0] = x; o[1] = y; o[2] = z;
                }
                continue;
            }
            it -= 24;
            if (JAC) {
                if (it < n_uit) {
                    const int code = p.m.u_item[it];
                    const int l = code >> 8, j = (code >> 4) & 15, a = code & 15;
                    const int link = chain_link[l * CPE_MAX_CHAIN + j];
                    const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                    // U block of one (marker, chain link): 3 angles x 3 components, padded to 10 doubles
                    drot_times(fr + SCO + 6 * link, a, v[0], v[1], v[2], fr + UO + 10 * (it / 3) + 3 * a);
                    continue;
                }
                it -= n_uit;
            }
            if (it < 26) {
                const int e = it;
                const int4 pr4 = reinterpret_cast<const int4 *>(p.m.eq_pairs)[e];
                const double *Ra = fr + RO + pr4.x * 9, *Rb = fr + RO + pr4.y * 9;
                const int ia = pr4.z, ib = pr4.w;
                const double c = Ra[ia] * Rb[ib] + Ra[3 + ia] * Rb[3 + ib] + Ra[6 + ia] * Rb[6 + ib];
                const double lam = p.b.lam_eq ? p.b.lam_eq[fidx * 26 + e] : 0.0;
                const double t = c + lam / p.b.mu_eq;
                ctr[EQC + e] = 0.5 * p.b.mu_eq * t * t;
                if (p.eo.c_eq) p.eo.c_eq[fidx * 26 + e] = c;
                continue;
            }
            it -= 26;
            if (it < n_ineq) {
                const int j = it;
                const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
                const double lo = p.m.ineq_bounds[2 * j], up = p.m.ineq_bounds[2 * j + 1];
                const double d = fr[QO + ip] - (im >= 0 ? fr[QO + im] : 0.0);
                const double lam = p.b.lam_in ? p.b.lam_in[fidx * n_ineq + j] : 0.0;
                const double t = d + lam / p.b.mu_in;
                const double v = t - fmin(fmax(t, lo), up);
                ctr[INC + j] = 0.5 * p.b.mu_in * v * v;
                if (p.eo.v_in) p.eo.v_in[fidx * n_ineq + j] = v;
                continue;
            }
            it -= n_ineq;
            {
                const int c = it;
                const double w = p.m.wq[c];
                double val = 0.0;
                if (w != 0.0 && k0 + mf >= 3) {
                    const double h = p.b.h[s];
                    const double d3 = (fr[QO + c] - 3.0 * qf[c - 54] + 3.0 * qf[c - 108] - qf[c - 162]) / (h * h);
                    val = w * d3 * d3;
                }
                ctr[MOC + c] = val;
            }
        }
    }
    __syncthreads();

    // ---- one thread per (frame, camera, marker): projection, residuals, robust loss, Jacobian columns -------------
    if (mvalid) {
        const double la_ = p.b.loss_a, lb_ = p.b.loss_b, lc_ = p.b.loss_c;
        const double ea = p.ea, eb = p.eb, ec = p.ec;
        const double *cm = cam + mc * CPE_CAM_STRIDE;
        Proj pr;
        project_point<JAC>(fr + POSO + ml * 3, cm, p.b.cam_model, pr);
        double rho = 0.0;
        const double cmult = cm[20];
        for (int w = 0; w < W; ++w) {
            double um, vm, gate;
            if (w > 0) {
                if (m64) m8 = reinterpret_cast<const double4 *>(p.b.meas)[midx0 + w];
                else m4 = __ldg(reinterpret_cast<const float4 *>(p.b.meas) + midx0 + w);
            }
            if (m64) { um = m8.x; vm = m8.y; gate = m8.z; }
            else { um = (double)m4.x; vm = (double)m4.y; gate = (double)m4.z; }
            const double wgt = gate * p.m.meas_w[w * 24 + ml] * cmult;
            const double ru = pr.u - um, rv = pr.v - vm;
            if (p.eo.resid) reinterpret_cast<double2 *>(p.eo.resid)[midx0 + w] = make_double2(ru, rv);
            rho += redescending_value(fabs(wgt * ru), la_, lb_, lc_, ea, eb, ec);
            rho += redescending_value(fabs(wgt * rv), la_, lb_, lc_, ea, eb, ec);
        }
        fr[CO + MEC + mt] = rho;
        if (JAC) {
            const int uo0 = p.m.u_off[ml], nlk = (p.m.u_off[ml + 1] - uo0) / 3;      // chain links of this marker
            // marker block of the packed Jacobian: (3 + 3 nlk) columns, each a (du, dv) pair
            double2 *jc = reinterpret_cast<double2 *>(stage + ((size_t)mf * C + mc) * (2 * (72 + n_u)) + 2 * (3 * ml + uo0));
            jc[0] = make_double2(pr.du[0], pr.dv[0]);
            jc[1] = make_double2(pr.du[1], pr.dv[1]);
            jc[2] = make_double2(pr.du[2], pr.dv[2]);
            const double2 *U2 = reinterpret_cast<const double2 *>(fr + UO + 10 * (uo0 / 3));
            for (int j = 0; j < nlk; ++j) {
                const double2 a0 = U2[5 * j], a1 = U2[5 * j + 1], a2 = U2[5 * j + 2], a3 = U2[5 * j + 3],
                              a4 = U2[5 * j + 4];
                double2 *o = jc + 3 + 3 * j;
                o[0] = make_double2(pr.du[0] * a0.x + pr.du[1] * a0.y + pr.du[2] * a1.x,
                                    pr.dv[0] * a0.x + pr.dv[1] * a0.y + pr.dv[2] * a1.x);
                o[1] = make_double2(pr.du[0] * a1.y + pr.du[1] * a2.x + pr.du[2] * a2.y,
                                    pr.dv[0] * a1.y + pr.dv[1] * a2.x + pr.dv[2] * a2.y);
                o[2] = make_double2(pr.du[0] * a3.x + pr.du[1] * a3.y + pr.du[2] * a4.x,
                                    pr.dv[0] * a3.x + pr.dv[1] * a3.y + pr.dv[2] * a4.x);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
    }
    __syncthreads();

    // ---- the Jacobian tile leaves as one bulk async copy; meanwhile warps reduce the cost terms -------------------
    if (JAC && tid == 0) {
        const uint32_t bytes = (uint32_t)(nv * C * 2 * (72 + n_u) * sizeof(double));
        double *g = p.eo.jac + f0 * C * 2 * (72 + n_u);
        const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(saddr), "r"(bytes)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    if (p.eo.frame_cost) {
        const int wid = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
        for (int f = wid; f < nv; f += nw) {
            const double *ctr = sm + C * CPE_CAM_STRIDE + f * FS + CO;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            for (int i = lane; i < TPF; i += 32) a0 += ctr[MEC + i];
            for (int i = lane; i < 54; i += 32) a1 += ctr[MOC + i];
            if (lane < 26) a2 = ctr[EQC + lane];
            if (lane < n_ineq) a3 = ctr[INC + lane];
            a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); a3 = warp_sum(a3);
            if (lane == 0) {
                double *o = p.eo.frame_cost + (f0 + f) * 4;
                o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3;
            }
        }
    }
    if (JAC && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    (void)n_ul;
}

template <bool JAC, int MAXT, int MINB>
static int launch_cfg(const EvalParams &p, int grid, int threads, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(eval_kernel<JAC, MAXT, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
    eval_kernel<JAC, MAXT, MINB><<<grid, threads, smem, st>>>(p);
    return (int)cudaGetLastError();
}

static int eval_cfg() {
    static int cfg = -1;
    if (cfg < 0) {
        const char *e = getenv("CPE_EVAL_CFG");      // tuning knob for experiments; default chosen by measurement
        cfg = e ? atoi(e) : 3;
    }
    return cfg;
}

static int launch_eval(const cpe_model *model, const cpe_batch *b, const double *q, const cpe_eval_out *eo,
                       cudaStream_t st) {
    if (!model || !b || !q || !eo) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS || (b->W != 1 && b->W != 3)) return CPE_E_DIM;
    if (model->n_ineq > CPE_MAX_INEQ || model->n_u <= 0 || model->n_u > 240) return CPE_E_DIM;
    const bool jac = eo->jac != nullptr;
    if (jac && (!model->u_off || !model->u_item)) return CPE_E_NULL;
    const int cfg = eval_cfg();
    EvalParams p;
    p.m = *model; p.b = *b; p.q = q; p.eo = *eo;
    p.ea = exp(b->loss_a); p.eb = exp(b->loss_b); p.ec = exp(b->loss_c);
    p.TPF = b->C * 24;
    const int budget = (cfg >= 2) ? 160 : 288;          // threads per CTA the frames are packed into
    int F = budget / p.TPF;
    if (F < 1) F = 1;
    if (F > 4) F = 4;
    if (F > b->N) F = b->N;
    p.F = F;
    p.chunks = (b->N + F - 1) / F;
    p.n_u = model->n_u;
    p.fs = (contrib_off(p.n_u) + MEC + p.TPF + 1) & ~1;
    p.off_stage = (b->C * CPE_CAM_STRIDE + F * p.fs + 1) & ~1;
    const size_t doubles = (size_t)p.off_stage + (jac ? (size_t)F * b->C * 2 * (72 + p.n_u) : 0);
    const size_t smem = doubles * sizeof(double);
    int threads = ((F * p.TPF + 31) / 32) * 32;
    if (threads < 128) threads = 128;
    const int grid = b->S * p.chunks;
    if (threads > 160 || cfg < 2) {
        if (cfg == 1) return jac ? launch_cfg<true, 288, 3>(p, grid, threads, smem, st)
                                 : launch_cfg<false, 288, 3>(p, grid, threads, smem, st);
        return jac ? launch_cfg<true, 288, 2>(p, grid, threads, smem, st)
                   : launch_cfg<false, 288, 2>(p, grid, threads, smem, st);
    }
    if (cfg == 3) return jac ? launch_cfg<true, 160, 6>(p, grid, threads, smem, st)
                             : launch_cfg<false, 160, 6>(p, grid, threads, smem, st);
    return jac ? launch_cfg<true, 160, 5>(p, grid, threads, smem, st)
               : launch_cfg<false, 160, 5>(p, grid, threads, smem, st);
}

}  // namespace cpe

extern "C" int cpe_eval_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const cpe_eval_out *out,
                            void *stream) {
    return cpe::launch_eval(model, batch, q, out, (cudaStream_t)stream);
}
