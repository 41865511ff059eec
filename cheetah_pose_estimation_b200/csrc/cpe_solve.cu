// cpe_solve.cu -- K3: batched block-banded LDL^T solve of the reduced normal equations, K4: step application,
// K5: step acceptance, plus the small reductions.  Replaces what IPOPT + MA97 do inside the reference's solver seam
// (acinoset_opt.py:611-617; shared/physical_education/utils.py:459-548) for the Gauss-Newton/LM formulation.
//
// System per sequence (n = 28 independent coordinates per frame, block half-bandwidth 3 frames):
//     A_kk = S_k + m(k,0) Wd ,  A_{k,k-j} = m(k,j) Wd  (j = 1..3) ,  Wd = diag(2 wq / h^4)
// S_k comes from cpe_assemble (frame-local Schur complement, already damped); m() is the D3^T D3 stencil of the
// constant-acceleration term (acinoset_misc.py:639-677 in q-only form).  Grouping 3 frames per super-block makes this
// the block-tridiagonal system of the north star; it is factorised directly in its natural 28x28 blocks:
//     A = Lt D Lt^T,  Lt unit-lower block-banded, D block-diagonal;  P_k = D_k^-1 by in-place Gauss-Jordan.
// One CTA per sequence, FP64, window of 10 blocks in shared memory.
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"

namespace cpe {

constexpr int NB = CPE_N_IND;      // 28
constexpr int LD = 29;             // padded leading dimension in shared memory (bank spread for column access)
constexpr int BLK = NB * LD;
constexpr int SOLVE_THREADS = 224; // 196 = 28 rows x 7 column-quads do the block algebra; warp 6 does the vectors

__device__ __forceinline__ double stencil(int k, int j, int nf) {
    const double c[4] = {-1.0, 3.0, -3.0, 1.0};
    double s = 0.0;
    for (int r = k - 3; r <= k - j; ++r) {
        if (r < 0 || r > nf - 4) continue;
        s += c[k - r] * c[k - j - r];
    }
    return s;
}

// out[r][c0..c0+3] -= sum_x X[r][x] * Y[c][x]   (out may be SUBTRACT = false -> overwrite with +sum)
template <bool SUBTRACT>
__device__ __forceinline__ void gemm_nt(double *out, const double *X, const double *Y, int r, int c0) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const double *xr = X + r * LD;
    const double *y0 = Y + c0 * LD, *y1 = y0 + LD, *y2 = y1 + LD, *y3 = y2 + LD;
#pragma unroll 7
    for (int x = 0; x < NB; ++x) {
        const double xv = xr[x];
        a0 += xv * y0[x];
        a1 += xv * y1[x];
        a2 += xv * y2[x];
        a3 += xv * y3[x];
    }
    double *o = out + r * LD + c0;
    if (SUBTRACT) {
        o[0] -= a0; o[1] -= a1; o[2] -= a2; o[3] -= a3;
    } else {
        o[0] = a0; o[1] = a1; o[2] = a2; o[3] = a3;
    }
}

struct SolveParams {
    cpe_batch b;
    const double *wq;
    const int32_t *ind_idx;
    const double *S;
    const double *ghat;
    cpe_solve_io io;
};

__global__ void __launch_bounds__(SOLVE_THREADS) band_solve_kernel(const __grid_constant__ SolveParams p) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x;
    const int s = blockIdx.x;
    const int N = p.b.N;
    const int nf = p.b.n_frames[s];
    double *buf[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) buf[i] = sm + i * BLK;
    double *wd = sm + 10 * BLK;          // [28]
    double *wv = wd + 28;                // w_{k}, ring of 4 : [4][28]
    double *uv = wv + 4 * 28;            // u_{k}, ring of 4 : [4][28]
    double *dv = uv + 4 * 28;            // backward ring [4][28]
    double *part = dv + 4 * 28;          // [3][28]
    __shared__ int bad;
    if (tid == 0) bad = 0;
    const double h = p.b.h[s];
    const double h4 = h * h * h * h;
    if (tid < 28) wd[tid] = 2.0 * p.wq[p.ind_idx[tid]] / h4;
    for (int i = tid; i < 12 * 28; i += SOLVE_THREADS) wv[i] = 0.0;
    __syncthreads();

    const bool alg = tid < 196;
    const int r = tid / 7, c0 = (tid % 7) * 4;
    // roles of the ten buffers (indices rotate each row)
    int iW1a = 0, iW1b = 1, iW2a = 2, iPa = 3, iPb = 4, iPc = 5, iT2 = 6, iT1 = 7, iL2 = 8, iL1 = 9;
    double *Lf = p.io.Lfac + (size_t)s * N * 4 * 784;
    const double *Sg = p.S + (size_t)s * N * 784;
    const double *gh = p.ghat + (size_t)s * N * 28;
    double *zb = p.io.zbuf + (size_t)s * N * 28;

    for (int k = 0; k < nf; ++k) {
        double *W1a = buf[iW1a], *W1b = buf[iW1b], *W2a = buf[iW2a];
        double *Pa = buf[iPa], *Pb = buf[iPb], *Pc = buf[iPc];
        double *T2 = buf[iT2], *T1 = buf[iT1], *L2 = buf[iL2], *L1 = buf[iL1];
        const double m0 = stencil(k, 0, nf), m1 = stencil(k, 1, nf), m2 = stencil(k, 2, nf), m3 = stencil(k, 3, nf);
        double *Lk = Lf + (size_t)k * 4 * 784;
        // ---- stage 1: L3 (to global), D init (into Pc), T2 ------------------------------------------------------
        double dreg[4];
        if (alg) {
            const double wr = wd[r];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = c0 + i;
                double d = Sg[(size_t)k * 784 + r * 28 + c] + (r == c ? m0 * wr : 0.0);
                double l3 = 0.0;
                if (k >= 3) {
                    const double pc = Pc[r * LD + c];
                    l3 = m3 * wr * pc;
                    d -= m3 * m3 * wr * wd[c] * pc;       // T3 L3^T, P symmetric
                }
                Lk[2 * 784 + r * 28 + c] = l3;
                dreg[i] = d;
                double t2 = (r == c ? m2 * wr : 0.0);
                if (k >= 3) t2 -= m3 * wr * W1b[c * LD + r];
                T2[r * LD + c] = t2;
            }
        }
        __syncthreads();
        if (alg) {
#pragma unroll
            for (int i = 0; i < 4; ++i) Pc[r * LD + c0 + i] = dreg[i];     // D_k now lives in Pc
        }
        // ---- stage 2: L2 = T2 Pb ; T1 init ------------------------------------------------------------------------
        if (alg) {
            if (k >= 2) gemm_nt<false>(L2, T2, Pb, r, c0);
            const double wr = wd[r];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = c0 + i;
                double t1 = (r == c ? m1 * wr : 0.0);
                if (k >= 3) t1 -= m3 * wr * W2a[c * LD + r];
                T1[r * LD + c] = t1;
                if (k < 2) L2[r * LD + c] = 0.0;
            }
        }
        __syncthreads();
        // ---- stage 3: T1 -= T2 W1a^T ; D -= T2 L2^T ----------------------------------------------------------------
        if (alg && k >= 2) {
            gemm_nt<true>(T1, T2, W1a, r, c0);
            gemm_nt<true>(Pc, T2, L2, r, c0);
        }
        __syncthreads();
        // ---- stage 4: L1 = T1 Pa -------------------------------------------------------------------------------------
        if (alg) {
            if (k >= 1) {
                gemm_nt<false>(L1, T1, Pa, r, c0);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) L1[r * LD + c0 + i] = 0.0;
            }
        }
        __syncthreads();
        // ---- stage 5: D -= T1 L1^T ; store L1, L2 ; forward substitution on warp 6 ----------------------------------
        if (alg) {
            if (k >= 1) gemm_nt<true>(Pc, T1, L1, r, c0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Lk[0 * 784 + r * 28 + c0 + i] = L1[r * LD + c0 + i];
                Lk[1 * 784 + r * 28 + c0 + i] = L2[r * LD + c0 + i];
            }
        } else {
            const int rr = tid - 196;
            // w_k = -ghat_k - L1 w_{k-1} - L2 w_{k-2} - m3 Wd u_{k-3}
            double acc = -gh[(size_t)k * 28 + rr];
            const double *w1 = wv + ((k + 3) & 3) * 28, *w2 = wv + ((k + 2) & 3) * 28;
            const double *u3 = uv + ((k + 1) & 3) * 28;
            if (k >= 1)
                for (int x = 0; x < 28; ++x) acc -= L1[rr * LD + x] * w1[x];
            if (k >= 2)
                for (int x = 0; x < 28; ++x) acc -= L2[rr * LD + x] * w2[x];
            if (k >= 3) acc -= m3 * wd[rr] * u3[rr];
            wv[(k & 3) * 28 + rr] = acc;
        }
        __syncthreads();
        // ---- stage 6: P_k = D_k^-1 by in-place Gauss-Jordan --------------------------------------------------------
        double *D = Pc;
        for (int pv = 0; pv < NB; ++pv) {
            double f = 0.0, piv = 1.0, rowp[4] = {0, 0, 0, 0};
            if (alg) {
                f = D[r * LD + pv];
                piv = D[pv * LD + pv];
#pragma unroll
                for (int i = 0; i < 4; ++i) rowp[i] = D[pv * LD + c0 + i];
            }
            __syncthreads();
            if (alg) {
                if (!(piv > 0.0)) {
                    if (r == pv && c0 == 0) bad = 1;
                    piv = (piv == 0.0 || piv != piv) ? 1e-300 : piv;
                }
                const double ip = 1.0 / piv;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = c0 + i;
                    double v;
                    if (r == pv) v = (c == pv) ? ip : rowp[i] * ip;
                    else v = (c == pv) ? -f * ip : D[r * LD + c] - f * rowp[i] * ip;
                    D[r * LD + c] = v;
                }
            }
            __syncthreads();
        }
        // symmetrise, store P_k, u_k = P_k w_k
        double sym[4];
        if (alg) {
#pragma unroll
            for (int i = 0; i < 4; ++i) sym[i] = 0.5 * (D[r * LD + c0 + i] + D[(c0 + i) * LD + r]);
        }
        __syncthreads();
        if (alg) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                D[r * LD + c0 + i] = sym[i];
                Lk[3 * 784 + r * 28 + c0 + i] = sym[i];
            }
        }
        __syncthreads();
        if (!alg) {
            const int rr = tid - 196;
            const double *wk = wv + (k & 3) * 28;
            double acc = 0.0;
            for (int x = 0; x < 28; ++x) acc += D[rr * LD + x] * wk[x];
            uv[(k & 3) * 28 + rr] = acc;
            zb[(size_t)k * 28 + rr] = acc;
        }
        __syncthreads();
        // rotate roles: new W1a = L1, W1b = old W1a, W2a = L2, Pa = D(Pc), Pb = old Pa, Pc = old Pb; the rest is scratch
        const int nW1a = iL1, nW1b = iW1a, nW2a = iL2, nPa = iPc, nPb = iPa, nPc = iPb;
        const int f0 = iW1b, f1 = iW2a, f2 = iT2, f3 = iT1;
        iW1a = nW1a; iW1b = nW1b; iW2a = nW2a; iPa = nPa; iPb = nPb; iPc = nPc;
        iT2 = f0; iT1 = f1; iL2 = f2; iL1 = f3;
    }
    // ---- backward: d_k = u_k - L1(k+1)^T d_{k+1} - L2(k+2)^T d_{k+2} - L3(k+3)^T d_{k+3} ----------------------------
    for (int i = tid; i < 4 * 28; i += SOLVE_THREADS) dv[i] = 0.0;
    __syncthreads();
    double *dout = p.io.delta_ind + (size_t)s * N * 28;
    for (int k = nf - 1; k >= 0; --k) {
        if (tid < 84) {
            const int j = tid / 28, c = tid - j * 28;     // block offset j+1
            const int kk = k + j + 1;
            double acc = 0.0;
            if (kk < nf) {
                const double *Lb = Lf + ((size_t)kk * 4 + j) * 784 + c;
                const double *dk = dv + (kk & 3) * 28;
#pragma unroll 7
                for (int x = 0; x < 28; ++x) acc += Lb[x * 28] * dk[x];
            }
            part[tid] = acc;
        }
        __syncthreads();
        if (tid < 28) {
            const double d = zb[(size_t)k * 28 + tid] - part[tid] - part[28 + tid] - part[56 + tid];
            dv[(k & 3) * 28 + tid] = d;
            dout[(size_t)k * 28 + tid] = d;
        }
        __syncthreads();
    }
    if (tid == 0 && p.io.status) p.io.status[s] = bad;
}

// ---- K4: back-substitute eliminated coords, q_try = q + delta, predicted reduction -----------------------------
struct ApplyParams {
    cpe_model m;
    cpe_batch b;
    const double *q;
    const double *damp;
    cpe_assemble_out ao;
    const double *delta_ind;
    double *delta;
    double *q_try;
    double *frame_pred;
};

__global__ void __launch_bounds__(128) step_apply_kernel(const __grid_constant__ ApplyParams p) {
    __shared__ double dl[4][54];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int fidx = blockIdx.x * 4 + wid;
    const int N = p.b.N;
    if (fidx >= p.b.S * N) return;
    const int s = fidx / N, k = fidx - s * N;
    if (k >= p.b.n_frames[s]) return;
    double *d = dl[wid];
    const double *di = p.delta_ind + (size_t)fidx * 28;
    for (int c = lane; c < 54; c += 32) {
        const int ip = p.m.ind_pos[c];
        d[c] = ip >= 0 ? di[ip] : 0.0;
    }
    __syncwarp();
    constexpr int GD = CPE_GRP_MAX_DEP, GN = CPE_GRP_MAX_NB;
    for (int e = lane; e < p.m.n_groups * GD; e += 32) {
        const int g = e / GD, i = e - g * GD;
        if (i < p.m.grp_ndep[g]) {
            const double *Y = p.ao.Y + (((size_t)fidx * p.m.n_groups + g) * GD + i) * GN;
            double acc = p.ao.y[((size_t)fidx * p.m.n_groups + g) * GD + i];
            const int nn = p.m.grp_nnb[g];
            for (int j = 0; j < nn; ++j) acc += Y[j] * d[p.m.grp_nb[g * GN + j]];
            d[p.m.grp_dep[g * GD + i]] = -acc;
        }
    }
    __syncwarp();
    const double damp = p.damp[s];
    double pred = 0.0;
    for (int c = lane; c < 54; c += 32) {
        const double dc = d[c];
        const size_t o = (size_t)fidx * 54 + c;
        if (p.delta) p.delta[o] = dc;
        p.q_try[o] = p.q[o] + dc;
        pred += -0.5 * p.ao.g[o] * dc + 0.5 * damp * p.ao.diagH[o] * dc * dc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) pred += __shfl_down_sync(0xffffffffu, pred, o);
    if (lane == 0) p.frame_pred[fidx] = pred;
}

// ---- deterministic per-sequence reduction ------------------------------------------------------------------------
__global__ void reduce_cost_kernel(int S, int N, const int32_t *n_frames, const double *frame_cost, double *cost) {
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (s >= S) return;
    const int nf = n_frames[s];
    double a[4] = {0, 0, 0, 0};
    for (int k = lane; k < nf; k += 32) {
        const double *f = frame_cost + ((size_t)s * N + k) * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] += f[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a[i] += __shfl_down_sync(0xffffffffu, a[i], o);
        if (lane == 0) cost[(size_t)s * 4 + i] = a[i];
    }
}

// ---- K5: accept / reject ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) step_accept_kernel(int S, int N, const int32_t *n_frames, const double *cost_try,
                                                          const double *frame_pred, double *q, const double *q_try,
                                                          double *state, int32_t *counters, int32_t *active, double ftol) {
    const int s = blockIdx.x;
    __shared__ int accept;
    if (active && !active[s]) return;
    const int nf = n_frames[s];
    if (threadIdx.x < 32) {
        double pred = 0.0;
        for (int k = threadIdx.x; k < nf; k += 32) pred += frame_pred[(size_t)s * N + k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pred += __shfl_down_sync(0xffffffffu, pred, o);
        if (threadIdx.x == 0) {
            const double *ct = cost_try + (size_t)s * 4;
            const double Ft = ct[0] + ct[1] + ct[2] + ct[3];
            double F = state[s], damp = state[S + s], nu = state[2 * S + s];
            int acc = 0;
            double rel = -1.0;
            if (Ft == Ft && Ft < F) {           // finite and decreased
                const double rho = (F - Ft) / fmax(pred, 1e-300);
                const double t = 2.0 * rho - 1.0;
                rel = (F - Ft) / fmax(fabs(F), 1e-300);
                F = Ft;
                damp = fmax(damp * fmax(1.0 / 3.0, 1.0 - t * t * t), 1e-12);
                nu = 2.0;
                acc = 1;
                counters[s * 2 + 1] += 1;
            } else {
                damp = fmin(damp * nu, 1e12);
                nu *= 2.0;
            }
            counters[s * 2 + 0] += 1;
            state[s] = F; state[S + s] = damp; state[2 * S + s] = nu; state[3 * S + s] = rel;
            if (active && ((acc && rel < ftol) || damp >= 1e12)) active[s] = 0;
            accept = acc;
        }
    }
    __syncthreads();
    if (accept) {
        const size_t base = (size_t)s * N * 54;
        for (int i = threadIdx.x; i < nf * 54; i += blockDim.x) q[base + i] = q_try[base + i];
    }
}

__global__ void update_multipliers_kernel(int S, int N, int n_ineq, const int32_t *n_frames, double mu_eq, double mu_in,
                                          const double *c_eq, const double *v_in, double *lam_eq, double *lam_in) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t per = 26 + n_ineq;
    if (i >= (size_t)S * N * per) return;
    const size_t fidx = i / per;
    const int e = (int)(i - fidx * per);
    const int s = (int)(fidx / N), k = (int)(fidx - (size_t)s * N);
    if (k >= n_frames[s]) return;
    if (e < 26) lam_eq[fidx * 26 + e] += mu_eq * c_eq[fidx * 26 + e];
    else lam_in[fidx * n_ineq + e - 26] = mu_in * v_in[fidx * n_ineq + e - 26];
}

}  // namespace cpe

extern "C" int cpe_version(void) { return CPE_VERSION; }

extern "C" int cpe_reduce_cost_f64(const cpe_batch *b, const double *frame_cost, double *cost, void *stream) {
    if (!b || !frame_cost || !cost) return CPE_E_NULL;
    cpe::reduce_cost_kernel<<<(b->S + 3) / 4, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, frame_cost, cost);
    return (int)cudaGetLastError();
}

extern "C" int cpe_band_solve_f64(const cpe_model *model, const cpe_batch *b, const double *Sblk, const double *ghat,
                                  const cpe_solve_io *io, void *stream) {
    if (!model || !b || !Sblk || !ghat || !io || !io->Lfac || !io->zbuf || !io->delta_ind) return CPE_E_NULL;
    cpe::SolveParams p;
    p.b = *b;
    p.wq = model->wq;
    p.ind_idx = model->ind_idx;
    p.S = Sblk;
    p.ghat = ghat;
    p.io = *io;
    const size_t smem = (size_t)(10 * cpe::BLK + 28 + 12 * 28 + 3 * 28) * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(cpe::band_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cpe::band_solve_kernel<<<b->S, cpe::SOLVE_THREADS, smem, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}

extern "C" int cpe_step_apply_f64(const cpe_model *model, const cpe_batch *b, const double *q, const double *damp,
                                  const cpe_assemble_out *ao, const double *delta_ind, double *delta, double *q_try,
                                  double *frame_pred, void *stream) {
    if (!model || !b || !q || !damp || !ao || !delta_ind || !q_try || !frame_pred) return CPE_E_NULL;
    if (!ao->Y || !ao->y || !ao->g || !ao->diagH) return CPE_E_NULL;
    cpe::ApplyParams p;
    p.m = *model; p.b = *b; p.q = q; p.damp = damp; p.ao = *ao; p.delta_ind = delta_ind; p.delta = delta;
    p.q_try = q_try; p.frame_pred = frame_pred;
    const int frames = b->S * b->N;
    cpe::step_apply_kernel<<<(frames + 3) / 4, 128, 0, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}

extern "C" int cpe_step_accept_f64(const cpe_batch *b, const double *cost_try, const double *frame_pred, double *q,
                                   const double *q_try, double *state, int32_t *counters, int32_t *active,
                                   double ftol, void *stream) {
    if (!b || !cost_try || !frame_pred || !q || !q_try || !state || !counters) return CPE_E_NULL;
    cpe::step_accept_kernel<<<b->S, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, cost_try, frame_pred, q, q_try,
                                                                    state, counters, active, ftol);
    return (int)cudaGetLastError();
}

extern "C" int cpe_update_multipliers_f64(const cpe_model *model, const cpe_batch *b, const double *c_eq, const double *v_in,
                                          double *lam_eq, double *lam_in, void *stream) {
    if (!model || !b || !c_eq || !v_in || !lam_eq || !lam_in) return CPE_E_NULL;
    const size_t n = (size_t)b->S * b->N * (26 + model->n_ineq);
    cpe::update_multipliers_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        b->S, b->N, model->n_ineq, b->n_frames, b->mu_eq, b->mu_in, c_eq, v_in, lam_eq, lam_in);
    return (int)cudaGetLastError();
}

extern "C" int cpe_eval_host_f64(const cpe_model *model, const cpe_batch *bd, const double *h_q, const void *h_meas,
                                 double *d_q, void *d_meas, const cpe_eval_out *out_dev, double *d_cost, double *h_cost,
                                 double *h_resid, void *stream) {
    if (!model || !bd || !h_q || !h_meas || !d_q || !d_meas || !out_dev || !d_cost || !h_cost) return CPE_E_NULL;
    if (!out_dev->frame_cost) return CPE_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t frames = (size_t)bd->S * bd->N;
    const size_t mbytes = frames * bd->C * 24 * bd->W * 4 * (bd->meas_is_f64 ? 8 : 4);
    cudaError_t e = cudaMemcpyAsync(d_q, h_q, frames * 54 * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    e = cudaMemcpyAsync(d_meas, h_meas, mbytes, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    cpe_batch b = *bd;
    b.meas = d_meas;
    int rc = cpe_eval_f64(model, &b, d_q, out_dev, stream);
    if (rc) return rc;
    rc = cpe_reduce_cost_f64(&b, out_dev->frame_cost, d_cost, stream);
    if (rc) return rc;
    e = cudaMemcpyAsync(h_cost, d_cost, (size_t)bd->S * 4 * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return (int)e;
    if (h_resid && out_dev->resid) {
        e = cudaMemcpyAsync(h_resid, out_dev->resid, frames * bd->C * 24 * bd->W * 2 * sizeof(double),
                            cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return (int)e;
    }
    return 0;
}
