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

// ---- blocked variant: the band is a whole number of 28-wide frame blocks (always the case for the monocular path: ----
// half-bandwidth = max(window, 3) frames).  One frame = one panel, two barriers per frame:
//   B      one thread per panel row solves X L_kk^T = A_panel (right-looking, L_kk broadcast from shared memory); the
//          right-hand side is one more row, so the forward substitution rides along;
//   A | C  warp 0 applies the panel to the NEXT diagonal block in registers (lane = row) and factorises it with shuffle
//          broadcasts of the pivot column, while warps 1-7 apply the rank-28 update to the rest of the lower window
//          in 7x7 register tiles and move the rows of frame k + R (fetched into registers before B) into the ring
//          slot frame k frees.
// band_general_kernel needs two barriers per scalar column (56 per frame).  The backward substitution streams the
// column panels of L back through a double buffer in shared memory.
constexpr int LDP = 29;

// Column J of the in-register Cholesky (lane i holds row i).  A template recursion instead of `#pragma unroll`: nvcc
// leaves the triangular double loop partly rolled, which puts the row into local memory.
// The finished column goes straight to its place in LdT (LdT[J * 28 + i] = L[i][J]) and is read back by every lane
// with 128-bit broadcast loads: 14 loads per column instead of 54 shuffles.
template <int J>
__device__ __forceinline__ void chol_column(double (&a)[NX], int i, double *LdT, double *idg, int &bad) {
    const unsigned FULL = 0xffffffffu;
    double ajj = __shfl_sync(FULL, a[J], J);
    const bool pos = ajj > 0.0;
    bad |= !pos;
    ajj = pos ? ajj : 1.0;
    const double id = rsqrt(ajj), d = ajj * id;
    const double l = (i == J) ? d : a[J] * id;               // lanes i < J carry stale values that are never read
    a[J] = l;
    if (i == J) idg[J] = id;
    if (i < NX) LdT[J * NX + i] = l;
    __syncwarp();
    constexpr int C0 = (J + 1) + ((J + 1) & 1);              // first even row > J
    if constexpr (((J + 1) & 1) && J + 1 < NX) {
        const double lc = LdT[J * NX + J + 1];
        if (i >= J + 1) a[J + 1] -= l * lc;
    }
#pragma unroll
    for (int c = C0; c + 1 < NX; c += 2) {
        const double2 v = *reinterpret_cast<const double2 *>(LdT + J * NX + c);
        if (i >= c) a[c] -= l * v.x;
        if (i >= c + 1) a[c + 1] -= l * v.y;
    }
    if constexpr (J + 1 < NX) chol_column<J + 1>(a, i, LdT, idg, bad);
}

// Row-wise solve x L^T = a against LdT[t * 28 + c] = L[c][t] (right-looking), same reason for the recursion.  LdT is
// 16-byte aligned: aligned pairs of a row are fetched with one 128-bit broadcast load.
template <int T>
__device__ __forceinline__ void panel_column(double (&a)[NX], const double *LdT, const double *idg) {
    const double x = a[T] * idg[T];
    a[T] = x;
    constexpr int C0 = (T + 1) + ((T + 1) & 1);              // first even column > T
    if constexpr (((T + 1) & 1) && T + 1 < NX) a[T + 1] -= x * LdT[T * NX + T + 1];
#pragma unroll
    for (int c = C0; c + 1 < NX; c += 2) {
        const double2 v = *reinterpret_cast<const double2 *>(LdT + T * NX + c);
        a[c] -= x * v.x;
        a[c + 1] -= x * v.y;
    }
    if constexpr (T + 1 < NX) panel_column<T + 1>(a, LdT, idg);
}

template <int R>
__global__ void __launch_bounds__(256) band_block_kernel(const __grid_constant__ BandParams p) {
    extern __shared__ __align__(16) double smg[];
    constexpr int WN = R * NX, nbk = R - 1, LD = WN + 1;
    constexpr int NPRE = (NX * WN + 223) / 224;             // entries of one frame row per thread of warps 1-7
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned FULL = 0xffffffffu;
    const int s = blockIdx.x;
    const int N = p.b.N;
    const int nf = p.b.n_frames[s];
    if (p.b.active && !p.b.active[s]) return;
    double *Wn = smg;                       // [WN][LD] ring window: frame f lives in slot f % R
    double *Lp = Wn + WN * LD;              // [112][LDP] panel of the current frame
    double *LdT = Lp + 112 * LDP;           // [28][28]  L_kk transposed: LdT[t * 28 + c] = L_kk[c][t]
    double *idg = LdT + NX * NX;            // [32] 1 / diag(L_kk)
    double *bv = idg + 32;                  // [160] right-hand-side ring / backward solution ring
    double *zk = bv + 160;                  // [2][32]
    double *Sst = zk + 64;                  // [2][28][LDP] frame-local blocks S_f, f & 1 (staged one frame ahead)
    __shared__ int bad;
    if (tid == 0) bad = 0;
    const double *Sg = p.S + (size_t)s * N * NX * NX;
    const double *gh = p.ghat + (size_t)s * N * NX;
    const double *Mb = p.Mband + (size_t)s * p.mstride;
    double *Lb = p.Lband + (size_t)s * N * NX * WN;
    double *dl = p.delta + (size_t)s * N * NX;
    // entry e of the rows of frame f (lower part of the constant band: column frames f - nbk .. f): offset of the value in
    // Mband (0 where there is none, so that the load needs no branch), ring position, and whether it is stored
    auto entry = [&](int f, int e, size_t &src, int &pos) -> bool {
        const int r = e / WN, cc = e - r * WN;
        const int g = f - nbk + cc / NX, c = cc % NX;
        pos = ((f % R) * NX + r) * LD + ((g + R) % R) * NX + c;
        const bool ok = g >= 0 && !(g == f && c > r);        // above the diagonal: never read
        src = ok ? (size_t)(f * NX + r) * WN + ((f - g) * NX + r - c) : 0;
        return ok;
    };
    for (int f = 0; f < min(R, nf); ++f) {
        for (int e = tid; e < NX * WN; e += 256) {
            size_t src; int pos;
            if (entry(f, e, src, pos)) Wn[pos] = Mb[src];
        }
        if (tid < NX) bv[(f % R) * NX + tid] = -gh[(size_t)f * NX + tid];
        if (f < 2)
            for (int e = tid; e < NX * NX; e += 256) Sst[f * NX * LDP + (e / NX) * LDP + e % NX] = Sg[(size_t)f * NX * NX + e];
    }
    // tile of this thread in the trailing update (warps 1-7): lower tiles except the 10 of the next diagonal block
    int ty = -1, tx = 0;
    if (tid >= 32) {
        int rem = tid - 32;
        for (int y = 4; y < 4 * nbk; ++y) {
            if (rem <= y) { ty = y; tx = rem; break; }
            rem -= y + 1;
        }
    }
    // Cholesky of the diagonal block of frame f (the panel of frame f - 1 has been applied by warps 5-7); warp 0 only
    auto diag_factor = [&](int f) {
        const int i = lane, ic = lane < NX ? lane : NX - 1;
        const int sf = (f % R) * NX;
        double a[NX];
        {
            // the part above the diagonal is stale shared memory: it is loaded but never feeds a lower entry
            const double *row = Wn + (sf + ic) * LD + sf, *srow = Sst + (f & 1) * NX * LDP + ic * LDP;
#pragma unroll
            for (int c = 0; c < NX; ++c) a[c] = row[c] + srow[c];
        }
        int flag = 0;
        chol_column<0>(a, i, LdT, idg, flag);                  // leaves L_kk in LdT (entries with t > i are never read)
        if (flag && lane == 0) bad = 1;
    };
    __syncthreads();
    if (tid < 32 && nf > 0) diag_factor(0);
    __syncthreads();

    for (int k = 0; k < nf; ++k) {
        const int m = min(nbk, nf - 1 - k);
        const int s0 = (k % R) * NX;
        // rows of frame k + R: loads issued now, parked in the ring after the panel has been consumed
        double pre[NPRE], preS[4];
        const bool fetch = tid >= 32 && k + R < nf, fetchS = tid >= 32 && k + 2 < nf;
        if (fetch) {
#pragma unroll
            for (int j = 0; j < NPRE; ++j) {
                const int e = min(tid - 32 + 224 * j, NX * WN - 1);
                size_t src; int pos;
                entry(k + R, e, src, pos);
                pre[j] = Mb[src];
            }
        }
        if (fetchS) {
#pragma unroll
            for (int j = 0; j < 4; ++j) preS[j] = Sg[(size_t)(k + 2) * NX * NX + min(tid - 32 + 224 * j, NX * NX - 1)];
        }
        // ---- B: panel rows (and the right-hand side as row 112); in-block columns of L to HBM ----------------------------
        if (tid < NX * m || tid == 112) {
            const bool rhs = tid == 112;
            const double *src = rhs ? bv + s0 : Wn + ((((k + 1 + tid / NX) % R) * NX + tid % NX) * LD + s0);
            double a[NX];
#pragma unroll
            for (int t = 0; t < NX; ++t) a[t] = src[t];
            panel_column<0>(a, LdT, idg);
            if (rhs) {
#pragma unroll
                for (int t = 0; t < NX; ++t) { zk[t] = a[t]; dl[(size_t)k * NX + t] = a[t]; }
            } else {
#pragma unroll
                for (int t = 0; t < NX; ++t) {
                    Lp[tid * LDP + t] = a[t];
                    Lb[(size_t)(k * NX + t) * WN + (NX - t + tid)] = a[t];
                }
            }
        } else if (tid >= 128) {
            for (int e = tid - 128; e < NX * NX; e += 128) {
                const int t = e / NX, i = e - t * NX;
                if (i >= t) Lb[(size_t)(k * NX + t) * WN + (i - t)] = LdT[e];
            }
        }
        __syncthreads();
        if (tid < 32) {
            // ---- A: next diagonal block ------------------------------------------------------------------------------
            if (k + 1 < nf) {
                asm volatile("bar.sync 1, 128;" ::: "memory");     // the 70 strips of warps 5-7 are in place
                diag_factor(k + 1);
            }
        } else {
            if (tid >= 160 && k + 1 < nf) {
                // warps 5-7 own no trailing tile: they apply the panel to the NEXT diagonal block first, as 1x7 strips
                // of its ten lower 7x7 tiles, and hand it to warp 0 through named barrier 1
                const int v = tid - 160;
                if (v < 70) {
                    const int tl = v / 7, ri = v - 7 * tl;
                    const int ty2 = tl < 1 ? 0 : (tl < 3 ? 1 : (tl < 6 ? 2 : 3));
                    const int tx2 = tl - (ty2 * (ty2 + 1)) / 2;
                    const int r = 7 * ty2 + ri, c0 = 7 * tx2;
                    double d[7];
#pragma unroll
                    for (int j = 0; j < 7; ++j) d[j] = 0.0;
#pragma unroll 4
                    for (int t = 0; t < NX; ++t) {
                        const double x = Lp[r * LDP + t];
#pragma unroll
                        for (int j = 0; j < 7; ++j) d[j] += x * Lp[(c0 + j) * LDP + t];
                    }
                    const int s1 = ((k + 1) % R) * NX;
                    double *w = Wn + (s1 + r) * LD + s1 + c0;
#pragma unroll
                    for (int j = 0; j < 7; ++j) w[j] -= d[j];
                }
                __threadfence_block();
                __syncwarp();                                      // bar.* is warp-aligned: reconverge first
                asm volatile("bar.arrive 1, 128;" ::: "memory");
            }
            // ---- C: trailing update (lower 7x7 tiles) + right-hand side ----------------------------------------------------
            if (ty >= 0 && 7 * ty < NX * m) {
                double acc[7][7];
#pragma unroll
                for (int i = 0; i < 7; ++i)
#pragma unroll
                    for (int j = 0; j < 7; ++j) acc[i][j] = 0.0;
                const double *lr = Lp + 7 * ty * LDP, *lc = Lp + 7 * tx * LDP;
#pragma unroll 4
                for (int t = 0; t < NX; ++t) {
                    double xr[7], xc[7];
#pragma unroll
                    for (int i = 0; i < 7; ++i) { xr[i] = lr[i * LDP + t]; xc[i] = lc[i * LDP + t]; }
#pragma unroll
                    for (int i = 0; i < 7; ++i)
#pragma unroll
                        for (int j = 0; j < 7; ++j) acc[i][j] += xr[i] * xc[j];
                }
                const int prow = ((k + 1 + (7 * ty) / NX) % R) * NX + (7 * ty) % NX;
                const int pcol = ((k + 1 + (7 * tx) / NX) % R) * NX + (7 * tx) % NX;
#pragma unroll
                for (int i = 0; i < 7; ++i)
#pragma unroll
                    for (int j = 0; j < 7; ++j) Wn[(prow + i) * LD + pcol + j] -= acc[i][j];
            }
            const int u = tid - 32;
            if (u < NX * m) {
                double acc = 0.0;
#pragma unroll
                for (int t = 0; t < NX; ++t) acc += Lp[u * LDP + t] * zk[t];
                bv[((k + 1 + u / NX) % R) * NX + u % NX] -= acc;
            }
            // ---- D: the freed slot takes frame k + R -----------------------------------------------------------------
            if (fetch) {
#pragma unroll
                for (int j = 0; j < NPRE; ++j) {
                    const int e = u + 224 * j;
                    size_t src; int pos;
                    if (e < NX * WN && entry(k + R, e, src, pos)) Wn[pos] = pre[j];
                }
                if (u < NX) bv[s0 + u] = -gh[(size_t)(k + R) * NX + u];
            }
            if (fetchS) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int e = u + 224 * j;
                    if (e < NX * NX) Sst[(k & 1) * NX * LDP + (e / NX) * LDP + e % NX] = preS[j];
                }
            }
        }
        __syncthreads();
    }
    // ---- backward substitution L^T delta = z, one frame per step ---------------------------------------------------------
    // The raw column panel of frame k (28 columns x WN entries, contiguous in Lband) is staged in shared memory; the panel
    // of frame k - 1 is fetched into registers meanwhile.  bv becomes the ring of the last R solutions.
    double *P0 = Wn, *P1 = Wn + NX * WN, *accv = LdT;
    const int np = NX * WN;
    if (nf > 0) {
        const double *src = Lb + (size_t)(nf - 1) * np;
        double *dst = ((nf - 1) & 1) ? P1 : P0;
        for (int e = tid; e < np; e += 256) dst[e] = src[e];
        if (tid < NX) zk[((nf - 1) & 1) * 32 + tid] = dl[(size_t)(nf - 1) * NX + tid];
    }
    __syncthreads();
    for (int k = nf - 1; k >= 0; --k) {
        const double *cur = (k & 1) ? P1 : P0;
        double *nxt = (k & 1) ? P0 : P1;
        const int m = min(nbk, nf - 1 - k);
        double pre[16], zpre = 0.0;
        if (k > 0) {
            const double *src = Lb + (size_t)(k - 1) * np;
#pragma unroll
            for (int e = 0; e < 16; ++e) pre[e] = (tid + 256 * e < np) ? src[tid + 256 * e] : 0.0;
            if (tid < NX) zpre = dl[(size_t)(k - 1) * NX + tid];
        }
        // step 1: contributions of the later frames, 8 threads per column
        {
            const int c = tid >> 3, g = tid & 7;
            double part = 0.0;
            if (c < NX) {
                const double *col = cur + c * WN - c;                   // col[u] = L(frame row u, column c), u >= c
                for (int u = NX + g; u < NX * (m + 1); u += 8)
                    part += col[u] * bv[((k + u / NX) % R) * NX + u % NX];
            }
            part += __shfl_xor_sync(FULL, part, 1);
            part += __shfl_xor_sync(FULL, part, 2);
            part += __shfl_xor_sync(FULL, part, 4);
            if (c < NX && g == 0) accv[c] = zk[(k & 1) * 32 + c] - part;
        }
        __syncthreads();
        // step 2: triangular solve with L_kk^T (warp 0) while the other warps park the prefetched panel
        if (tid < 32) {
            const int c1 = lane < NX ? lane : 0;
            double v = lane < NX ? accv[lane] : 0.0;
            const double inv = 1.0 / cur[c1 * WN];
            for (int c = NX - 1; c >= 0; --c) {
                if (lane == c) v *= inv;
                const double dcur = __shfl_sync(FULL, v, c);
                if (lane < c) v -= cur[lane * WN + (c - lane)] * dcur;
            }
            if (lane < NX) {
                bv[(k % R) * NX + lane] = v;
                dl[(size_t)k * NX + lane] = v;
            }
        }
        if (k > 0) {
#pragma unroll
            for (int e = 0; e < 16; ++e)
                if (tid + 256 * e < np) nxt[tid + 256 * e] = pre[e];
            if (tid < NX) zk[((k - 1) & 1) * 32 + tid] = zpre;
        }
        __syncthreads();
    }
    for (int i = nf * NX + tid; i < N * NX; i += 256) dl[i] = 0.0;
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
    if (WN % cpe::NX == 0 && WN / cpe::NX >= 2 && WN / cpe::NX <= 5) {
        // whole frame blocks: blocked factorisation (3 barriers per frame)
        const size_t smb = (size_t)(WN * (WN + 1) + 112 * cpe::LDP + cpe::NX * cpe::NX + 32 + 160 + 64 + 2 * cpe::NX * cpe::LDP) *
                           sizeof(double);
        auto launch = [&](auto kern) -> int {
            cudaError_t eb = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smb);
            if (eb != cudaSuccess) return (int)eb;
            kern<<<b->S, 256, smb, (cudaStream_t)stream>>>(p);
            return (int)cudaGetLastError();
        };
        switch (WN / cpe::NX) {
            case 2: return launch(cpe::band_block_kernel<2>);
            case 3: return launch(cpe::band_block_kernel<3>);
            case 4: return launch(cpe::band_block_kernel<4>);
            default: return launch(cpe::band_block_kernel<5>);
        }
    }
    const size_t smem = (size_t)(WN * WN + 3 * WN) * sizeof(double);
    if (smem > 227 * 1024) return CPE_E_DIM;
    cudaError_t e = cudaFuncSetAttribute(cpe::band_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cpe::band_general_kernel<<<b->S, 256, smem, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}
