// cpe_prior.cu -- data-driven monocular priors on the reduced pose: GMM pose prior and AR(window)
// linear motion prior.  Reference arithmetic being replaced:
//   reduced pose   acinoset_misc.py:487-528 (get_relative_angles) + :1699-1757 (mask)        x = Tr q_ind
//   pose prior     acinoset_misc.py:680-714 (gmm_pose_cost, norm_pdf_multivariate :691-694)
//   motion prior   acinoset_misc.py:291-336 (add_linear_motion_model), acinoset_models.py:242-247 (predict),
//                  common/py_utils/data_ops.py:108-142 (lag layout: oldest lag first)
// and, for the LM formulation, what ASL differentiates inside ipopt: the gradient of both terms and a PSD
// majoriser of the pose-prior Hessian (sum_c r_c P_c, r = responsibilities); the motion prior is a linear least
// squares whose exact Hessian is constant and lives in the band handed to cpe_band_solve_general_f64.
//
// Mapping: one CTA (128 threads) per frame; everything is 28- or 22-dimensional dense algebra in shared memory.  The
// monocular path runs tens of hypotheses x ~100 frames, so this kernel is launch-latency sized, not throughput sized.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "cpe.h"

namespace cpe {

constexpr int NX = CPE_N_IND;            // 28
constexpr int ND = CPE_PRIOR_DIM;        // 22
constexpr int NE = NX - ND;              // 6 leading entries (base pose) outside the pose prior
constexpr int MAXW = CPE_PRIOR_MAX_WINDOW;

struct PriorParams {
    cpe_prior pr;
    cpe_batch b;
    const int32_t *ind_idx;
    const double *q;
    const double *damp;
    cpe_assemble_out ao;
    double *frame_cost;      // [S,N,2] (cost mode)
};

template <bool ASSEMBLE>
__global__ void __launch_bounds__(128) prior_kernel(const __grid_constant__ PriorParams p) {
    __shared__ double x[2 * MAXW + 1][NX];          // reduced pose of frames k-window .. k+window
    __shared__ double qi[2 * MAXW + 1][NX];
    __shared__ double wr[MAXW + 1][NX];             // wt .* residual of frames k .. k+window
    __shared__ double gx[NX];
    __shared__ double vc[CPE_PRIOR_MAX_COMP][ND];   // P_c (x - mu_c)
    __shared__ double resp[CPE_PRIOR_MAX_COMP];
    __shared__ double Hx[ND][ND];
    __shared__ double M1[ND][NX];
    __shared__ double cost2[2];
    const int tid = threadIdx.x, nt = blockDim.x;
    const int N = p.b.N;
    const int s = blockIdx.y, k = blockIdx.x;
    const int nf = p.b.n_frames[s];
    if (k >= nf || (p.b.active && !p.b.active[s])) return;
    const int win = p.pr.window, ncomp = p.pr.n_comp;
    const size_t fidx = (size_t)s * N + k;
    const int nfr = 2 * win + 1;

    // independent coordinates and reduced pose of the neighbouring frames
    for (int i = tid; i < nfr * NX; i += nt) {
        const int f = i / NX, a = i - f * NX;
        const int kk = k - win + f;
        qi[f][a] = (kk >= 0 && kk < nf) ? p.q[((size_t)s * N + kk) * 54 + p.ind_idx[a]] : 0.0;
    }
    if (tid < 2) cost2[tid] = 0.0;
    __syncthreads();
    for (int i = tid; i < nfr * NX; i += nt) {
        const int f = i / NX, a = i - f * NX;
        const double *tr = p.pr.Tr + a * NX;
        double acc = 0.0;
#pragma unroll 4
        for (int j = 0; j < NX; ++j) acc += tr[j] * qi[f][j];
        x[f][a] = acc;
    }
    __syncthreads();

    // ---- motion prior: residuals of frames k .. k+window (those that exist), weighted ---------------------------------
    if (win > 0) {
        const int nres = ASSEMBLE ? win + 1 : 1;
        for (int i = tid; i < nres * NX; i += nt) {
            const int j = i / NX, a = i - j * NX;           // residual row of frame k + j
            const int kk = k + j;
            double r = 0.0;
            if (kk >= win && kk < nf) {
                const double *w = p.pr.ar_W + (size_t)a * NX * win;
                r = x[win + j][a] - p.pr.ar_b[a];
                for (int lag = win; lag >= 1; --lag) {      // lag `lag` sits at columns (win - lag) * 28
                    const double *wl = w + (win - lag) * NX;
                    const double *xl = x[win + j - lag];
                    double acc = 0.0;
#pragma unroll 4
                    for (int c = 0; c < NX; ++c) acc += wl[c] * xl[c];
                    r -= acc;
                }
            }
            wr[j][a] = p.pr.ar_wt[a] * r;
            if (j == 0 && !ASSEMBLE) gx[a] = p.pr.ar_wt[a] * r * r;        // frame cost contributions
        }
    }
    __syncthreads();
    if (!ASSEMBLE && win > 0 && tid == 0) {
        double acc = 0.0;
        for (int a = 0; a < NX; ++a) acc += gx[a];
        cost2[1] = acc;
    }
    if (ASSEMBLE) {
        // gradient w.r.t. x_k: 2 wt r_k - sum_j 2 W_j^T (wt r_{k+j})
        for (int a = tid; a < NX; a += nt) {
            double acc = win > 0 ? 2.0 * wr[0][a] : 0.0;
            for (int j = 1; j <= win; ++j) {
                const double *wj = p.pr.ar_W + (win - j) * NX + a;          // column a of lag-j block
                double t = 0.0;
                for (int r = 0; r < NX; ++r) t += wj[(size_t)r * NX * win] * wr[j][r];
                acc -= 2.0 * t;
            }
            gx[a] = acc;
        }
    }
    __syncthreads();

    // ---- pose prior ---------------------------------------------------------------------------------------------------------
    if (ncomp > 0) {
        const double *xs = x[win] + NE;
        for (int i = tid; i < ncomp * ND; i += nt) {
            const int c = i / ND, r = i - c * ND;
            const double *P = p.pr.gmm_prec + ((size_t)c * ND + r) * ND;
            const double *mu = p.pr.gmm_mu + c * ND;
            double acc = 0.0;
            for (int j = 0; j < ND; ++j) acc += P[j] * (xs[j] - mu[j]);
            vc[c][r] = acc;
        }
        __syncthreads();
        if (tid < ncomp) {
            const double *mu = p.pr.gmm_mu + tid * ND;
            double m = 0.0;
            for (int j = 0; j < ND; ++j) m += (xs[j] - mu[j]) * vc[tid][j];
            resp[tid] = exp(p.pr.gmm_logc[tid] - 0.5 * m);                   // w_c N(x; mu_c, Sigma_c)
        }
        __syncthreads();
        if (tid == 0) {
            double pdf = 0.0;
            for (int c = 0; c < ncomp; ++c) pdf += resp[c];
            cost2[0] = -log(pdf + 1e-12);
            const double inv = 1.0 / (pdf + 1e-12);
            for (int c = 0; c < ncomp; ++c) resp[c] *= inv;
        }
        __syncthreads();
        if (ASSEMBLE) {
            for (int r = tid; r < ND; r += nt) {
                double acc = 0.0;
                for (int c = 0; c < ncomp; ++c) acc += resp[c] * vc[c][r];
                gx[NE + r] += acc;
            }
            for (int i = tid; i < ND * ND; i += nt) {
                double acc = 0.0;
                for (int c = 0; c < ncomp; ++c) acc += resp[c] * p.pr.gmm_prec[(size_t)c * ND * ND + i];
                Hx[i / ND][i % ND] = acc;
            }
        }
        __syncthreads();
    }
    if constexpr (!ASSEMBLE) {
        if (tid < 2) p.frame_cost[fidx * 2 + tid] = cost2[tid];
    } else {
    // ---- back to the independent coordinates: g_q = Tr^T gx, H_q = Trs^T Hx Trs (Trs = rows 6..27 of Tr) ------------------
    for (int a = tid; a < NX; a += nt) {
        double acc = 0.0;
        for (int i = 0; i < NX; ++i) acc += p.pr.Tr[i * NX + a] * gx[i];
        p.ao.ghat[fidx * NX + a] += acc;
        if (p.ao.g) p.ao.g[fidx * 54 + p.ind_idx[a]] += acc;
    }
    if (ncomp > 0) {
        for (int i = tid; i < ND * NX; i += nt) {
            const int r = i / NX, b = i - r * NX;
            double acc = 0.0;
            for (int j = 0; j < ND; ++j) acc += Hx[r][j] * p.pr.Tr[(NE + j) * NX + b];
            M1[r][b] = acc;
        }
        __syncthreads();
        const double damp = p.damp ? p.damp[s] : 0.0;
        double *So = p.ao.S + fidx * NX * NX;
        for (int i = tid; i < NX * NX; i += nt) {
            const int a = i / NX, b = i - a * NX;
            double acc = 0.0;
            for (int r = 0; r < ND; ++r) acc += p.pr.Tr[(NE + r) * NX + a] * M1[r][b];
            if (a == b) {
                if (p.ao.diagH) p.ao.diagH[fidx * 54 + p.ind_idx[a]] += acc;
                So[i] += acc + damp * acc;
            } else {
                So[i] += acc;
            }
        }
    }
    }
}

__global__ void reduce_prior_cost_kernel(int S, int N, const int32_t *n_frames, const int32_t *active,
                                         const double *frame_cost, double *cost) {
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (s >= S || (active && !active[s])) return;
    const int nf = n_frames[s];
    double a0 = 0.0, a1 = 0.0;
    for (int k = lane; k < nf; k += 32) {
        a0 += frame_cost[((size_t)s * N + k) * 2];
        a1 += frame_cost[((size_t)s * N + k) * 2 + 1];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a0 += __shfl_down_sync(0xffffffffu, a0, o);
        a1 += __shfl_down_sync(0xffffffffu, a1, o);
    }
    if (lane == 0) { cost[s * 2] = a0; cost[s * 2 + 1] = a1; }
}

static int check_prior(const cpe_model *model, const cpe_prior *pr, const cpe_batch *b) {
    if (!model || !pr || !b) return CPE_E_NULL;
    if (pr->n_comp < 0 || pr->n_comp > CPE_PRIOR_MAX_COMP || pr->window < 0 || pr->window > CPE_PRIOR_MAX_WINDOW)
        return CPE_E_DIM;
    if (b->S <= 0 || b->N <= 0 || b->S > 65535) return CPE_E_DIM;
    if (!pr->Tr || !model->ind_idx || !b->n_frames) return CPE_E_NULL;
    if (pr->n_comp > 0 && (!pr->gmm_logc || !pr->gmm_mu || !pr->gmm_prec)) return CPE_E_NULL;
    if (pr->window > 0 && (!pr->ar_W || !pr->ar_b || !pr->ar_wt)) return CPE_E_NULL;
    return 0;
}

}  // namespace cpe

extern "C" int cpe_prior_cost_f64(const cpe_model *model, const cpe_prior *prior, const cpe_batch *b, const double *q,
                                  double *frame_work, double *prior_cost, void *stream) {
    using namespace cpe;
    const int rc = check_prior(model, prior, b);
    if (rc) return rc;
    if (!q || !frame_work || !prior_cost) return CPE_E_NULL;
    PriorParams p;
    p.pr = *prior; p.b = *b; p.ind_idx = model->ind_idx; p.q = q; p.damp = nullptr; p.frame_cost = frame_work;
    p.ao = cpe_assemble_out{};
    prior_kernel<false><<<dim3(b->N, b->S), 128, 0, (cudaStream_t)stream>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    reduce_prior_cost_kernel<<<(b->S + 3) / 4, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, b->active, frame_work, prior_cost);
    return (int)cudaGetLastError();
}

extern "C" int cpe_prior_assemble_f64(const cpe_model *model, const cpe_prior *prior, const cpe_batch *b, const double *q,
                                      const double *damp, const cpe_assemble_out *out, void *stream) {
    using namespace cpe;
    const int rc = check_prior(model, prior, b);
    if (rc) return rc;
    if (!q || !out || !out->S || !out->ghat) return CPE_E_NULL;
    PriorParams p;
    p.pr = *prior; p.b = *b; p.ind_idx = model->ind_idx; p.q = q; p.damp = damp; p.frame_cost = nullptr; p.ao = *out;
    prior_kernel<true><<<dim3(b->N, b->S), 128, 0, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}

// ---- general banded SPD solve (monocular path) ----------------------------------------------------------------------------
// Right-looking Cholesky on a sliding (bw+1) x (bw+1) window of the band kept in shared memory (circular indexing);
// the right-hand side rides along, so the forward substitution is fused.  One CTA (16 x 16 threads) per sequence.
namespace cpe {

struct BandParams {
    cpe_batch b;
    const double *S;
    const double *ghat;
    const double *Mband;
    int64_t mstride;
    int bw;
    double *Lband;
    double *delta;
    int32_t *status;
};

__global__ void __launch_bounds__(256) band_general_kernel(const __grid_constant__ BandParams p) {
    extern __shared__ __align__(16) double smg[];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4, tid = threadIdx.x;
    const int s = blockIdx.x;
    const int N = p.b.N, bw = p.bw, WN = bw + 1;
    const int nf = p.b.n_frames[s];
    const int n = nf * NX;
    if (p.b.active && !p.b.active[s]) return;
    double *Wn = smg;                    // [WN][WN]
    double *lcol = Wn + WN * WN;         // [WN]
    double *bvec = lcol + WN;            // [WN] running right-hand side
    double *dc = bvec + WN;              // [WN] backward-substitution ring
    __shared__ int bad;
    if (tid == 0) bad = 0;
    const double *Sg = p.S + (size_t)s * N * NX * NX;
    const double *gh = p.ghat + (size_t)s * N * NX;
    const double *Mb = p.Mband + (size_t)s * p.mstride;
    double *Lb = p.Lband + (size_t)s * N * NX * WN;
    double *dl = p.delta + (size_t)s * N * NX;
    if (n == 0) {
        if (tid == 0 && p.status) p.status[s] = 0;
        return;
    }
    // A(i, c), c <= i <= c + bw
    auto Aval = [&](int i, int c) -> double {
        double v = Mb[(size_t)i * WN + (i - c)];
        const int fi = i / NX, fc = c / NX;
        if (fi == fc) v += Sg[(size_t)fi * NX * NX + (i - fi * NX) * NX + (c - fc * NX)];
        return v;
    };
    // initial window: rows 0 .. min(bw, n-1)
    const int m0 = min(bw, n - 1);
    for (int r = ty; r <= m0; r += 16)
        for (int c = tx; c <= r; c += 16) Wn[r * WN + c] = Aval(r, c);
    for (int i = tid; i <= m0; i += 256) bvec[i] = -gh[i];
    __syncthreads();

    for (int j = 0; j < n; ++j) {
        const int sj = j % WN;
        const int m = min(bw, n - 1 - j);            // rows j+1 .. j+m exist
        const double piv = Wn[sj * WN + sj];
        const double d = sqrt(piv > 0.0 ? piv : 1.0);
        if (tid == 0 && !(piv > 0.0)) bad = 1;
        const double id = 1.0 / d;
        if (tid <= m) {
            const int si = (j + tid) % WN;
            if (tid == 0) {
                const double z = bvec[sj] * id;
                lcol[0] = z;                          // z_j, read back after the barrier
                Lb[(size_t)j * WN] = d;
                dl[j] = z;
            } else {
                const double l = Wn[si * WN + sj] * id;
                lcol[tid] = l;
                Lb[(size_t)j * WN + tid] = l;
            }
        }
        __syncthreads();
        const double zj = lcol[0];
        if (tid >= 1 && tid <= m) bvec[(j + tid) % WN] -= lcol[tid] * zj;
        for (int r = ty + 1; r <= m; r += 16) {
            const double lr = lcol[r];
            const int sr = ((j + r) % WN) * WN;
            for (int c = tx + 1; c <= r; c += 16) Wn[sr + (j + c) % WN] -= lr * lcol[c];
        }
        // the freed slot takes row j + bw + 1 of A
        const int inew = j + bw + 1;
        if (inew < n) {
            if (tid <= bw) {
                const int c = inew - tid;             // columns inew .. inew - bw = j + 1
                Wn[sj * WN + (c % WN)] = Aval(inew, c);
            }
            if (tid == 255) bvec[sj] = -gh[inew];
        }
        __syncthreads();
    }
    // ---- backward substitution L^T delta = z (warp 0; ring of the last bw+1 solutions in shared memory) ---------------------
    if (tid < 32) {
        for (int j = n - 1; j >= 0; --j) {
            const int m = min(bw, n - 1 - j);
            const double *Lj = Lb + (size_t)j * WN;
            double acc = 0.0;
            for (int t = 1 + tid; t <= m; t += 32) acc += Lj[t] * dc[(j + t) % WN];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            const double v = (dl[j] - acc) / Lj[0];
            __syncwarp();
            if (tid == 0) { dc[j % WN] = v; dl[j] = v; }
            __syncwarp();
        }
    }
    __syncthreads();
    // padded frames: zero step
    for (int i = n + tid; i < N * NX; i += 256) dl[i] = 0.0;
    if (tid == 0 && p.status) p.status[s] = bad;
}

}  // namespace cpe

extern "C" int cpe_band_solve_general_f64(const cpe_batch *b, const double *Sblk, const double *ghat, const double *Mband,
                                          int64_t mband_stride, int32_t bw, double *Lband, double *delta_ind,
                                          int32_t *status, void *stream) {
    if (!b || !Sblk || !ghat || !Mband || !Lband || !delta_ind) return CPE_E_NULL;
    if (bw < cpe::NX - 1 || bw > 255 || b->S <= 0 || b->N <= 0) return CPE_E_DIM;
    cpe::BandParams p;
    p.b = *b; p.S = Sblk; p.ghat = ghat; p.Mband = Mband; p.mstride = mband_stride; p.bw = bw; p.Lband = Lband;
    p.delta = delta_ind; p.status = status;
    const int WN = bw + 1;
    const size_t smem = (size_t)(WN * WN + 3 * WN) * sizeof(double);
    if (smem > 227 * 1024) return CPE_E_DIM;
    cudaError_t e = cudaFuncSetAttribute(cpe::band_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cpe::band_general_kernel<<<b->S, 256, smem, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}
