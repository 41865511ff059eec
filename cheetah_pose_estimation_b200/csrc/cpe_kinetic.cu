// cpe_kinetic.cu -- cost + gradient kernels of the trajectory-level terms of the reference's kinetic objective
// (acinoset_opt.py:905-921):
//   motion smoothing   acinoset_misc.py:601-613  sum_{fe < nfe-2} sum_l sum_d (fps^2 (pose[fe+2] - 2 pose[fe+1] + pose[fe]))^2
//   EOM slack          acinoset_misc.py:631-636  sum_fe sum_i slack_eom[fe,i]^2 on the rows of cpe_eom_f64 (eom = a - Bm F)
//   torque squared     motor.py:412-417, 496-503 sum (weight Tc)^2
//   change in torque   acinoset_misc.py:616-628  sum_{fe < nfe-1} ((Tc[fe+1] - Tc[fe]) / hm0)^2
// The frame ranges follow the reference's code (1-based `fe < nfe - 2` / `fe < nfe - 1`): one term fewer than the data
// allows.  Every sum is taken in a fixed order (no atomics): results do not depend on the batch or the launch.
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"
#include "cpe_internal.cuh"

namespace cpe {

// ---- motion smoothing: one 96-thread CTA per frame ------------------------------------------------------------------------
// a[l,d] = d cost / d pose[k,l,d] from the five-frame stencil of second differences; gradient through the marker chains:
// base x, y, z collect sum_l a_l, angle (link, a) collects (dR_link/da . v_{l,slot}) . a_l over the chain slots on that link.
__global__ void __launch_bounds__(96) smoothing_kernel(cpe_model m, cpe_batch b, const double *q, const double *pose,
                                                       double *frame_cost, double *grad) {
    __shared__ double a_sm[72], sc[102], part[3];
    const int s = blockIdx.x / b.N, k = blockIdx.x - s * b.N;
    const int nf = b.n_frames[s];
    if (k >= nf) return;
    const int tid = threadIdx.x;
    const size_t f = (size_t)s * b.N + k;
    const double fps = 1.0 / b.h[s];
    const double f2 = fps * fps, f4 = f2 * f2;
    const int last = nf - 4;                              // second differences r_0 .. r_last (0-based), see the header
    double c_loc = 0.0;
    if (tid < 72) {
        const double *p = pose + f * 72 + tid;            // pose[s,k,l,d], stride 72 per frame
        double acc = 0.0;
        // r_j = pose[j+2] - 2 pose[j+1] + pose[j]; pose[k] enters r_k (+1), r_{k-1} (-2), r_{k-2} (+1)
        if (k <= last) {
            const double r = p[144] - 2.0 * p[72] + p[0];
            acc += r;
            c_loc = (f2 * r) * (f2 * r);
        }
        if (k - 1 >= 0 && k - 1 <= last) acc -= 2.0 * (p[72] - 2.0 * p[0] + p[-72]);
        if (k - 2 >= 0 && k - 2 <= last) acc += p[0] - 2.0 * p[-72] + p[-144];
        a_sm[tid] = 2.0 * f4 * acc;
    }
    if (tid < 51) sincos(q[f * 54 + 3 + tid], &sc[2 * tid], &sc[2 * tid + 1]);
    const double cw = warp_sum(c_loc);
    if ((tid & 31) == 0) part[tid >> 5] = cw;
    __syncthreads();
    if (tid == 0 && frame_cost) frame_cost[f] = part[0] + part[1] + part[2];
    if (!grad || tid >= 54) return;
    double g = 0.0;
    if (tid < 3) {
        for (int l = 0; l < 24; ++l) g += a_sm[3 * l + tid];
    } else {
        const int link = (tid - 3) / 3, ang = (tid - 3) - 3 * link;
        const double *cvec = b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
        for (int it = 0; it < m.n_u; ++it) {
            const int code = m.u_item[it];
            const int l = code >> 8, j = (code >> 4) & 15, aa = code & 15;
            if (aa != ang || m.chain_link[l * CPE_MAX_CHAIN + j] != link) continue;
            const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
            double o[3];
            drot_times(sc + 6 * link, ang, v[0], v[1], v[2], o);
            g += o[0] * a_sm[3 * l] + o[1] * a_sm[3 * l + 1] + o[2] * a_sm[3 * l + 2];
        }
    }
    grad[f * 54 + tid] = g;
}

// cost[s * stride + col] = sum over the frames of sequence s of frame_cost (fixed order)
__global__ void __launch_bounds__(128) reduce_frames_kernel(int S, int N, const int32_t *n_frames, const double *frame_cost,
                                                            double *cost, int stride, int col) {
    const int s = blockIdx.x;
    __shared__ double part[4];
    double acc = 0.0;
    for (int k = threadIdx.x; k < n_frames[s]; k += 128) acc += frame_cost[(size_t)s * N + k];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) cost[(size_t)s * stride + col] = part[0] + part[1] + part[2] + part[3];
}

// ---- EOM slack: one warp per frame ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) eom_slack_kernel(int S, int N, const int32_t *n_frames, const double *eom,
                                                        const double *Bm, double *frame_cost, double *grad_f) {
    const int lane = threadIdx.x & 31;
    const long long f = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (f >= (long long)S * N) return;
    const int s = (int)(f / N), k = (int)(f - (long long)s * N);
    if (k >= n_frames[s]) return;
    const double *e = eom + f * 6;
    if (lane == 0) {
        double c = 0.0;
        for (int i = 0; i < 6; ++i) c += e[i] * e[i];
        frame_cost[f] = c;
    }
    if (grad_f && Bm && lane < CPE_EOM_N_FORCES) {        // eom = a - Bm F  =>  d cost / d F = -2 Bm^T eom
        const double *B = Bm + f * 6 * CPE_EOM_N_FORCES + lane;
        double g = 0.0;
        for (int i = 0; i < 6; ++i) g += B[i * CPE_EOM_N_FORCES] * e[i];
        grad_f[f * CPE_EOM_N_FORCES + lane] = -2.0 * g;
    }
}

// ---- torque terms: one thread per (frame, motor axis) ----------------------------------------------------------------------------
__global__ void __launch_bounds__(128) torque_kernel(int S, int N, int M, const int32_t *n_frames, const double *h,
                                                     const double *Tc, const double *weight, double *fc_sq, double *fc_ch,
                                                     double *grad_sq, double *grad_ch) {
    const long long f = blockIdx.x;
    const int s = (int)(f / N), k = (int)(f - (long long)s * N);
    const int nf = n_frames[s];
    if (k >= nf) return;
    __shared__ double p1[4], p2[4];
    const double ih = 1.0 / h[s];
    double c1 = 0.0, c2 = 0.0;
    for (int j = threadIdx.x; j < M; j += 128) {
        const double *t = Tc + f * M + j;
        const double w = weight[j], wt = w * t[0];
        c1 += wt * wt;
        if (grad_sq) grad_sq[f * M + j] = 2.0 * w * wt;
        // differences d_fe = (Tc[fe+1] - Tc[fe]) / hm0 exist for 0-based fe = 0 .. nf-3
        double g = 0.0;
        if (k <= nf - 3) {
            const double d = (t[M] - t[0]) * ih;
            c2 += d * d;
            g -= 2.0 * d * ih;
        }
        if (k >= 1 && k - 1 <= nf - 3) g += 2.0 * (t[0] - t[-M]) * ih * ih;
        if (grad_ch) grad_ch[f * M + j] = g;
    }
    c1 = warp_sum(c1);
    c2 = warp_sum(c2);
    if ((threadIdx.x & 31) == 0) { p1[threadIdx.x >> 5] = c1; p2[threadIdx.x >> 5] = c2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        fc_sq[f] = p1[0] + p1[1] + p1[2] + p1[3];
        fc_ch[f] = p2[0] + p2[1] + p2[2] + p2[3];
    }
}

}  // namespace cpe

extern "C" int cpe_motion_smoothing_f64(const cpe_model *model, const cpe_batch *b, const double *q, const double *pose,
                                        double *frame_work, double *cost, double *grad, void *stream) {
    if (!model || !b || !q || !pose || !frame_work || !cost) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || (long long)b->S * b->N > 2147483647LL || model->n_u > 256) return CPE_E_DIM;
    if (!b->n_frames || !b->h || !b->chain_vec || !model->u_item || !model->chain_link) return CPE_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    cpe::smoothing_kernel<<<(unsigned)(b->S * b->N), 96, 0, st>>>(*model, *b, q, pose, frame_work, grad);
    cpe::reduce_frames_kernel<<<b->S, 128, 0, st>>>(b->S, b->N, b->n_frames, frame_work, cost, 1, 0);
    return (int)cudaGetLastError();
}

extern "C" int cpe_eom_slack_f64(const cpe_batch *b, const double *eom, const double *Bm, double *frame_work, double *cost,
                                 double *grad_f, void *stream) {
    if (!b || !eom || !frame_work || !cost || !b->n_frames) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0) return CPE_E_DIM;
    cudaStream_t st = (cudaStream_t)stream;
    const long long frames = (long long)b->S * b->N;
    cpe::eom_slack_kernel<<<(unsigned)((frames + 3) / 4), 128, 0, st>>>(b->S, b->N, b->n_frames, eom, Bm, frame_work, grad_f);
    cpe::reduce_frames_kernel<<<b->S, 128, 0, st>>>(b->S, b->N, b->n_frames, frame_work, cost, 1, 0);
    return (int)cudaGetLastError();
}

// ---- allocation-free variant of cpe_eom_slack_f64 ---------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-frame scratch `frame_work` lives in a
// file-scope device array of CPE_NOALLOC_MAX_FRAMES doubles (512 KB, held by every GPU that loads the library until exit).
// In HOST code the array's name is a host-side stand-in, not a device address: the address is obtained once per device, with
// that device current, through cudaGetSymbolAddress and cached; the kernels get that pointer.
static __device__ __align__(256) double g_noalloc_frame_work[CPE_NOALLOC_MAX_FRAMES];

namespace cpe {
constexpr int NOALLOC_MAX_DEVICES = 64;
static double *g_noalloc_ptr[NOALLOC_MAX_DEVICES];      // device address of the scratch, per device ordinal (0 = not set up)

static int noalloc_scratch(double **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return CPE_E_ARG;
    double *p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        // set-up, once per device, with that device current (two threads racing here obtain the same address)
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_frame_work);
        if (e != cudaSuccess) return (int)e;
        if (!p) return CPE_E_NULL;
        __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}
}  // namespace cpe

extern "C" int cpe_eom_slack_noalloc_f64(const cpe_batch *b, const double *eom, const double *Bm, double *cost,
                                         double *grad_f, void *stream) {
    if (!b || !eom || !cost || !b->n_frames) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0) return CPE_E_DIM;
    const long long frames = (long long)b->S * b->N;
    if (frames > CPE_NOALLOC_MAX_FRAMES) return CPE_E_SCRATCH;          // beyond the static scratch: use cpe_eom_slack_f64
    double *work = nullptr;
    const int rc = cpe::noalloc_scratch(&work);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the previous call's contents: the part this call uses starts at zero
    cudaError_t e = cudaMemsetAsync(work, 0, (size_t)frames * sizeof(double), st);
    if (e != cudaSuccess) return (int)e;
    cpe::eom_slack_kernel<<<(unsigned)((frames + 3) / 4), 128, 0, st>>>(b->S, b->N, b->n_frames, eom, Bm, work, grad_f);
    cpe::reduce_frames_kernel<<<b->S, 128, 0, st>>>(b->S, b->N, b->n_frames, work, cost, 1, 0);
    return (int)cudaGetLastError();
}

extern "C" int cpe_torque_costs_f64(const cpe_batch *b, const double *Tc, int32_t n_axes, const double *weight,
                                    double *frame_work, double *cost, double *grad_sq, double *grad_ch, void *stream) {
    if (!b || !Tc || !weight || !frame_work || !cost || !b->n_frames || !b->h) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || n_axes <= 0 || (long long)b->S * b->N > 2147483647LL) return CPE_E_DIM;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t frames = (size_t)b->S * b->N;
    cpe::torque_kernel<<<(unsigned)frames, 128, 0, st>>>(b->S, b->N, n_axes, b->n_frames, b->h, Tc, weight, frame_work,
                                                         frame_work + frames, grad_sq, grad_ch);
    cpe::reduce_frames_kernel<<<b->S, 128, 0, st>>>(b->S, b->N, b->n_frames, frame_work, cost, 2, 0);
    cpe::reduce_frames_kernel<<<b->S, 128, 0, st>>>(b->S, b->N, b->n_frames, frame_work + frames, cost, 2, 1);
    return (int)cudaGetLastError();
}
