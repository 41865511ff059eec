// cpe_bandsolve.cu -- K3: batched block-banded LDL^T factorisation + solve of the reduced normal equations.
// Replaces what IPOPT + MA97 do inside the reference's solver seam (acinoset_opt.py:611-617;
// shared/physical_education/utils.py:459-548) for the Gauss-Newton/LM formulation.
//
// System per sequence (n = 28 independent coordinates per frame, block half-bandwidth 3 frames):
//     A_kk = S_k + m(k,0) Wd ,  A_{k,k-j} = m(k,j) Wd  (j = 1..3) ,  Wd = diag(2 wq / h^4)
// S_k comes from cpe_assemble (frame-local Schur complement, already damped); m() is the D3^T D3 stencil of the
// constant-acceleration term (acinoset_misc.py:639-677 in q-only form).  Grouping 3 frames per super-block makes this
// the block-tridiagonal system of the north star; it is factorised directly in its natural 28x28 blocks:
//     A = L D L^T,  L unit-lower block-banded (L1, L2, L3 per row), D block-diagonal, P_k = D_k^-1.
//     T3 = A_{k,k-3} (diagonal)              L3 = T3 P_{k-3}
//     T2 = A_{k,k-2} - T3 L1_{k-2}^T         L2 = T2 P_{k-2}
//     T1 = A_{k,k-1} - T3 L2_{k-1}^T - T2 L1_{k-1}^T        L1 = T1 P_{k-1}
//     D_k = A_kk - T3 L3^T - T2 L2^T - T1 L1^T
//
// Mapping (sm_100a, FP64): one CTA per sequence, 4 math warps + 1 vector warp.  The five 28x28x28 products of a
// block row run on the FP64 tensor path (mma.sync m8n8k4 f64 -> DMMA): warp w owns rows 8w..8w+7 of every result
// as accumulator fragments in registers; operands are read from a rotating window of ten 28x28 blocks in shared
// memory (leading dimension 28 is bank-conflict-free for the 8x4 fragment loads).  P_k = D_k^-1 is formed in the
// accumulator registers by block Gauss-Jordan with 4x4 pivots: each of the 7 pivot steps is a rank-4 update = one
// DMMA per 8x8 tile; P_k is symmetrised before use.  12 barriers per block row.  The fifth warp carries the forward substitution one row
// behind the factorisation.  The backward substitution streams L1..L3 back from HBM.
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"

namespace cpe {

constexpr int NB = CPE_N_IND;          // 28
constexpr int BLK = NB * NB;           // 784
constexpr int SOLVE_THREADS = 160;

__device__ __forceinline__ double stencil(int k, int j, int nf) {
    const double c[4] = {-1.0, 3.0, -3.0, 1.0};
    double s = 0.0;
    for (int r = k - 3; r <= k - j; ++r) {
        if (r < 0 || r > nf - 4) continue;
        s += c[k - r] * c[k - j - r];
    }
    return s;
}

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__device__ __forceinline__ void bar_math() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// ---- mbarrier + 1-D bulk async copy (global -> shared), used by the backward substitution ----------------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *b) {
    asm volatile("{ .reg .b64 t; mbarrier.arrive.shared::cta.b64 t, [%0]; }" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, unsigned bytes) {
    asm volatile("{ .reg .b64 t; mbarrier.arrive.expect_tx.shared::cta.b64 t, [%0], %1; }" ::"r"(smem_u32(b)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity) {
    asm volatile(
        "{ .reg .pred p;\n"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra W; }" ::"r"(smem_u32(b)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}

// acc (rows 8w..8w+7, all 32 padded columns) += sgn * X[rows] . Y^T ; X, Y row-major 28x28 in shared memory.
// Lane (gr, gc): A fragment X[r][4ks+gc], B fragment Y[8t+gr][4ks+gc]; padded rows/columns contribute zeros.
template <bool NEG>
__device__ __forceinline__ void mma_nt(double (&acc)[4][2], const double *X, const double *Y, int r, int gr, int gc) {
    const bool rv = r < NB;
    const double *xr = X + (rv ? r : 0) * NB + gc;
    const double *y0 = Y + gr * NB + gc, *y1 = y0 + 8 * NB, *y2 = y0 + 16 * NB;
    const bool v3 = gr < 4;                                     // rows 24..27 of the last tile
    const double *y3 = Y + (24 + (v3 ? gr : 0)) * NB + gc;
#pragma unroll
    for (int ks = 0; ks < 7; ++ks) {
        double a = rv ? xr[4 * ks] : 0.0;
        if (NEG) a = -a;
        dmma(acc[0][0], acc[0][1], a, y0[4 * ks]);
        dmma(acc[1][0], acc[1][1], a, y1[4 * ks]);
        dmma(acc[2][0], acc[2][1], a, y2[4 * ks]);
        dmma(acc[3][0], acc[3][1], a, v3 ? y3[4 * ks] : 0.0);
    }
}

// store an accumulator set to a 28x28 row-major block (shared or global)
__device__ __forceinline__ void store_frag(double *M, const double (&acc)[4][2], int r, int gc) {
    if (r >= NB) return;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int c = 8 * t + 2 * gc;
        if (c < NB) *reinterpret_cast<double2 *>(M + r * NB + c) = make_double2(acc[t][0], acc[t][1]);
    }
}

// inverse of a symmetric positive definite 4x4 (upper entries m[i][j], i <= j) via 2x2 blocks; ok = all pivots > 0
__device__ __forceinline__ bool inv4_spd(const double (&m)[4][4], double (&o)[4][4]) {
    const double a00 = m[0][0], a01 = m[0][1], a11 = m[1][1];
    const double b00 = m[0][2], b01 = m[0][3], b10 = m[1][2], b11 = m[1][3];
    const double c00 = m[2][2], c01 = m[2][3], c11 = m[3][3];
    const double detA = a00 * a11 - a01 * a01;
    const double idA = 1.0 / detA;
    const double i00 = a11 * idA, i01 = -a01 * idA, i11 = a00 * idA;
    // X = A^-1 B
    const double x00 = i00 * b00 + i01 * b10, x01 = i00 * b01 + i01 * b11;
    const double x10 = i01 * b00 + i11 * b10, x11 = i01 * b01 + i11 * b11;
    // S = C - B^T X
    const double s00 = c00 - (b00 * x00 + b10 * x10);
    const double s01 = c01 - (b00 * x01 + b10 * x11);
    const double s11 = c11 - (b01 * x01 + b11 * x11);
    const double detS = s00 * s11 - s01 * s01;
    const double idS = 1.0 / detS;
    const double t00 = s11 * idS, t01 = -s01 * idS, t11 = s00 * idS;
    // Z = X S^-1
    const double z00 = x00 * t00 + x01 * t01, z01 = x00 * t01 + x01 * t11;
    const double z10 = x10 * t00 + x11 * t01, z11 = x10 * t01 + x11 * t11;
    o[0][0] = i00 + z00 * x00 + z01 * x01;
    o[0][1] = i01 + z00 * x10 + z01 * x11;
    o[1][1] = i11 + z10 * x10 + z11 * x11;
    o[1][0] = o[0][1];
    o[0][2] = -z00; o[0][3] = -z01; o[1][2] = -z10; o[1][3] = -z11;
    o[2][0] = -z00; o[3][0] = -z01; o[2][1] = -z10; o[3][1] = -z11;
    o[2][2] = t00; o[2][3] = t01; o[3][2] = t01; o[3][3] = t11;
    return a00 > 0.0 && detA > 0.0 && s00 > 0.0 && detS > 0.0;
}

struct SolveParams {
    cpe_batch b;
    const double *wq;
    const int32_t *ind_idx;
    const double *S;
    const double *ghat;
    cpe_solve_io io;
};

__global__ void __launch_bounds__(SOLVE_THREADS, 3) band_solve_kernel(const __grid_constant__ SolveParams p) {
    extern __shared__ __align__(16) double sm[];
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int gr = lane >> 2, gc = lane & 3;
    const int s = blockIdx.x;
    const int N = p.b.N;
    const int nf = p.b.n_frames[s];
    if (p.b.active && !p.b.active[s]) return;
    double *colp = sm + 10 * BLK;        // [32][4]  pivot column panel (rows private to each warp)
    double *rowp = colp + 128;           // [4][32]  pivot row panel (owner warp only)
    double *Rp = rowp + 128;             // [2][4][32] scaled pivot rows, ping-pong
    double *wd = Rp + 256;               // [28] (padded to 32)
    double *wv = wd + 32;                // w ring [4][28]
    double *uv = wv + 112;               // u ring [4][28]
    double *dv = uv + 112;               // backward ring [4][28]
    double *part = dv + 112;             // [3][28]
    __shared__ int bad;
    if (tid == 0) bad = 0;
    const double h = p.b.h[s];
    const double h4 = h * h * h * h;
    for (int i = tid; i < 10 * BLK + 128 + 128 + 256; i += SOLVE_THREADS) sm[i] = 0.0;
    if (tid < 32) wd[tid] = tid < NB ? 2.0 * p.wq[p.ind_idx[tid]] / h4 : 0.0;
    for (int i = tid; i < 3 * 112 + 84; i += SOLVE_THREADS) wv[i] = 0.0;
    __syncthreads();

    // roles of the ten buffers (indices rotate each row)
    int iW1a = 0, iW1b = 1, iW2a = 2, iPa = 3, iPb = 4, iPc = 5, iT2 = 6, iT1 = 7, iL2 = 8, iL1 = 9;
    double *Lf = p.io.Lfac + (size_t)s * N * 3 * BLK;
    const double *Sg = p.S + (size_t)s * N * BLK;
    const double *gh = p.ghat + (size_t)s * N * NB;
    double *zb = p.io.zbuf + (size_t)s * N * NB;

    const int r = 8 * warp + gr;                      // this lane's row (math warps)
    const bool rvalid = warp < 4 && r < NB;
    double snext[4][2];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int c = 8 * t + 2 * gc;
        snext[t][0] = snext[t][1] = 0.0;
        if (rvalid && c < NB && nf > 0) {
            const double2 v = *reinterpret_cast<const double2 *>(Sg + r * NB + c);
            snext[t][0] = v.x; snext[t][1] = v.y;
        }
    }

    for (int k = 0; k <= nf; ++k) {
        if (warp < 4) {
            if (k < nf) {
                double *W1a = sm + iW1a * BLK, *W1b = sm + iW1b * BLK, *W2a = sm + iW2a * BLK;
                double *Pa = sm + iPa * BLK, *Pb = sm + iPb * BLK, *Pc = sm + iPc * BLK;
                double *T2 = sm + iT2 * BLK, *T1 = sm + iT1 * BLK, *L2 = sm + iL2 * BLK, *L1 = sm + iL1 * BLK;
                const double m0 = stencil(k, 0, nf), m1 = stencil(k, 1, nf), m2 = stencil(k, 2, nf), m3 = stencil(k, 3, nf);
                double *Lk = Lf + (size_t)k * 3 * BLK;
                double dacc[4][2], t1acc[4][2], l2acc[4][2];
                // ---- E: elementwise set-up in fragment coordinates --------------------------------------------------
                {
                    const double wr = rvalid ? wd[r] : 0.0;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        double l3[2], t2v[2];
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int c = 8 * t + 2 * gc + e;
                            const bool v = rvalid && c < NB;
                            const double pc = v ? Pc[r * NB + c] : 0.0;
                            const double dg = (r == c) ? wr : 0.0;
                            l3[e] = m3 * wr * pc;
                            dacc[t][e] = v ? snext[t][e] + m0 * dg - m3 * m3 * wr * wd[c & 31] * pc : 0.0;
                            t2v[e] = v ? m2 * dg - m3 * wr * W1b[c * NB + r] : 0.0;
                            t1acc[t][e] = v ? m1 * dg - m3 * wr * W2a[c * NB + r] : 0.0;
                            l2acc[t][e] = 0.0;
                        }
                        const int c0 = 8 * t + 2 * gc;
                        if (rvalid && c0 < NB) {
                            *reinterpret_cast<double2 *>(T2 + r * NB + c0) = make_double2(t2v[0], t2v[1]);
                            *reinterpret_cast<double2 *>(Lk + 2 * BLK + r * NB + c0) = make_double2(l3[0], l3[1]);
                        }
                    }
                    // prefetch S_{k+1}
                    if (k + 1 < nf) {
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const int c = 8 * t + 2 * gc;
                            if (rvalid && c < NB) {
                                const double2 v = *reinterpret_cast<const double2 *>(Sg + (size_t)(k + 1) * BLK + r * NB + c);
                                snext[t][0] = v.x; snext[t][1] = v.y;
                            }
                        }
                    }
                }
                bar_math();
                // ---- G1: L2 = T2 Pb ; G2: T1 -= T2 W1a^T ---------------------------------------------------------------
                mma_nt<false>(l2acc, T2, Pb, r, gr, gc);
                mma_nt<true>(t1acc, T2, W1a, r, gr, gc);
                store_frag(L2, l2acc, r, gc);
                store_frag(T1, t1acc, r, gc);
                store_frag(Lk + BLK, l2acc, r, gc);
                bar_math();
                // ---- G3: D -= T2 L2^T ; G4: L1 = T1 Pa -----------------------------------------------------------------
                {
                    double l1acc[4][2];
#pragma unroll
                    for (int t = 0; t < 4; ++t) l1acc[t][0] = l1acc[t][1] = 0.0;
                    mma_nt<true>(dacc, T2, L2, r, gr, gc);
                    mma_nt<false>(l1acc, T1, Pa, r, gr, gc);
                    store_frag(L1, l1acc, r, gc);
                    store_frag(Lk, l1acc, r, gc);
                }
                bar_math();
                // ---- G5: D -= T1 L1^T --------------------------------------------------------------------------------------
                mma_nt<true>(dacc, T1, L1, r, gr, gc);
                // ---- P = D^-1: block Gauss-Jordan with 4x4 pivots on the accumulator fragments ------------------------------
#pragma unroll
                for (int pv = 0; pv < 7; ++pv) {
                    const int tp = pv >> 1, half = pv & 1, wo = pv >> 1;
                    double *R = Rp + (pv & 1) * 128;
                    if ((gc >> 1) == half) {
                        colp[(8 * warp + gr) * 4 + 2 * (gc & 1) + 0] = dacc[tp][0];
                        colp[(8 * warp + gr) * 4 + 2 * (gc & 1) + 1] = dacc[tp][1];
                    }
                    if (warp == wo) {
                        if ((gr >> 2) == half) {
                            const int m = gr & 3;
#pragma unroll
                            for (int t = 0; t < 4; ++t)
                                *reinterpret_cast<double2 *>(rowp + m * 32 + 8 * t + 2 * gc) = make_double2(dacc[t][0], dacc[t][1]);
                        }
                        __syncwarp();
                        double mm[4][4], iv[4][4];
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = i; j < 4; ++j) mm[i][j] = rowp[i * 32 + 4 * pv + j];
                        if (!inv4_spd(mm, iv) && lane == 0) bad = 1;
                        // lane = column j of the scaled pivot rows R[kk][j]
                        const int j = lane;
                        const double r0 = rowp[j], r1 = rowp[32 + j], r2 = rowp[64 + j], r3 = rowp[96 + j];
                        const bool inp = (j >> 2) == pv;
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            double v = iv[kk][0] * r0 + iv[kk][1] * r1 + iv[kk][2] * r2 + iv[kk][3] * r3;
                            if (inp) v = iv[kk][j & 3];
                            R[kk * 32 + j] = v;
                        }
                    }
                    bar_math();
                    {
                        const double a = -colp[(8 * warp + gr) * 4 + gc];
                        if ((gc >> 1) == half) dacc[tp][0] = dacc[tp][1] = 0.0;
#pragma unroll
                        for (int t = 0; t < 4; ++t) dmma(dacc[t][0], dacc[t][1], a, R[gc * 32 + 8 * t + gr]);
                        if (warp == wo && (gr >> 2) == half) {
                            const int m = gr & 3;
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                const double2 v = *reinterpret_cast<const double2 *>(R + m * 32 + 8 * t + 2 * gc);
                                dacc[t][0] = v.x; dacc[t][1] = v.y;
                            }
                        }
                    }
                }
                // symmetrise P_k (asymmetric rounding would otherwise compound along the recursion): raw copy to the
                // dead T2 block, then P = (raw + raw^T) / 2 replaces P_{k-3}
                store_frag(T2, dacc, r, gc);
                bar_math();
                if (rvalid) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int c = 8 * t + 2 * gc;
                        if (c < NB) {
                            dacc[t][0] = 0.5 * (dacc[t][0] + T2[c * NB + r]);
                            dacc[t][1] = 0.5 * (dacc[t][1] + T2[(c + 1) * NB + r]);
                        }
                    }
                }
                store_frag(Pc, dacc, r, gc);
            }
        } else if (k >= 1) {
            // ---- vector warp: forward substitution of row kv = k-1 (its L1, L2, P are W1a, W2a, Pa during row k) ------
            const int kv = k - 1;
            const double *L1v = sm + iW1a * BLK, *L2v = sm + iW2a * BLK, *Pv = sm + iPa * BLK;
            const int rr = lane;
            double acc = 0.0;
            if (rr < NB) {
                acc = -gh[(size_t)kv * NB + rr];
                const double *w1 = wv + ((kv + 3) & 3) * NB, *w2 = wv + ((kv + 2) & 3) * NB;
                const double *u3 = uv + ((kv + 1) & 3) * NB;
                double a1 = 0.0, a2 = 0.0;
#pragma unroll 7
                for (int x = 0; x < NB; ++x) {
                    a1 += L1v[rr * NB + x] * w1[x];
                    a2 += L2v[rr * NB + x] * w2[x];
                }
                acc -= a1 + a2;
                acc -= stencil(kv, 3, nf) * wd[rr] * u3[rr];
                wv[(kv & 3) * NB + rr] = acc;
            }
            __syncwarp();
            if (rr < NB) {
                const double *wk = wv + (kv & 3) * NB;
                double u = 0.0;
#pragma unroll 7
                for (int x = 0; x < NB; ++x) u += Pv[rr * NB + x] * wk[x];
                uv[(kv & 3) * NB + rr] = u;
                zb[(size_t)kv * NB + rr] = u;
            }
        }
        __syncthreads();
        if (k < nf) {
            // rotate roles: new W1a = L1, W1b = old W1a, W2a = L2, Pa = P_k (in Pc), Pb = old Pa, Pc = old Pb
            const int nW1a = iL1, nW1b = iW1a, nW2a = iL2, nPa = iPc, nPb = iPa, nPc = iPb;
            const int f0 = iW1b, f1 = iW2a, f2 = iT2, f3 = iT1;
            iW1a = nW1a; iW1b = nW1b; iW2a = nW2a; iPa = nPa; iPb = nPb; iPc = nPc;
            iT2 = f0; iT1 = f1; iL2 = f2; iL1 = f3;
        }
    }
    // ---- backward: d_k = u_k - L1(k+1)^T d_{k+1} - L2(k+2)^T d_{k+2} - L3(k+3)^T d_{k+3} ----------------------------
    // The factor blocks come back from HBM through the (now dead) operand window: three stages of three 28x28 blocks,
    // filled two rows ahead by bulk async copies (cp.async.bulk -> mbarrier complete_tx), so the 500 dependent steps of
    // a sequence no longer wait for a DRAM round trip each.
    double *dout = p.io.delta_ind + (size_t)s * N * NB;
    __shared__ __align__(8) unsigned long long mbar[3];
    // every thread's factor stores (global) and window stores (shared) went through the generic proxy: order them before
    // the async-proxy copies that follow
    asm volatile("fence.proxy.async;" ::: "memory");
    if (tid == 0) {
        for (int i = 0; i < 3; ++i) mbar_init(&mbar[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int kq) {                            // thread 0: blocks needed by step kq into stage kq % 3
        const int st = kq % 3;
        int nv = 0;
        for (int j = 0; j < 3; ++j) nv += (kq + j + 1 < nf) ? 1 : 0;
        if (nv == 0) { mbar_arrive(&mbar[st]); return; }
        mbar_expect_tx(&mbar[st], (unsigned)(nv * BLK * sizeof(double)));
        for (int j = 0; j < 3; ++j) {
            const int kk = kq + j + 1;
            if (kk < nf) bulk_g2s(sm + (st * 3 + j) * BLK, Lf + ((size_t)kk * 3 + j) * BLK, (unsigned)(BLK * sizeof(double)), &mbar[st]);
        }
    };
    if (tid == 0) {
        if (nf >= 1) issue(nf - 1);
        if (nf >= 2) issue(nf - 2);
    }
    unsigned phase = 0;                                   // bit st = parity the next wait on stage st expects
    double z_next = (tid < NB && nf >= 1) ? zb[(size_t)(nf - 1) * NB + tid] : 0.0;
    for (int k = nf - 1; k >= 0; --k) {
        const int st = k % 3;
        const double z_cur = z_next;
        if (tid < NB && k >= 1) z_next = zb[(size_t)(k - 1) * NB + tid];
        if (tid == 0 && k >= 2) issue(k - 2);             // stage (k + 1) % 3: its readers passed the barrier ending step k + 1
        mbar_wait(&mbar[st], (phase >> st) & 1u);
        phase ^= 1u << st;
        if (tid < 84) {
            const int j = tid / NB, c = tid - j * NB;     // block offset j+1
            const int kk = k + j + 1;
            double acc = 0.0;
            if (kk < nf) {
                const double *Lb = sm + (st * 3 + j) * BLK + c;
                const double *dk = dv + (kk & 3) * NB;
                double a0 = 0.0, a1 = 0.0;
#pragma unroll
                for (int x = 0; x < NB; x += 2) {
                    a0 += Lb[x * NB] * dk[x];
                    a1 += Lb[(x + 1) * NB] * dk[x + 1];
                }
                acc = a0 + a1;
            }
            part[tid] = acc;
        }
        __syncthreads();
        if (tid < NB) {
            const double d = z_cur - part[tid] - part[NB + tid] - part[2 * NB + tid];
            dv[(k & 3) * NB + tid] = d;
            dout[(size_t)k * NB + tid] = d;
        }
        __syncthreads();
    }
    if (tid == 0 && p.io.status) p.io.status[s] = bad;
}

}  // namespace cpe

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
    const size_t smem = (size_t)(10 * cpe::BLK + 128 + 128 + 256 + 32 + 3 * 112 + 84) * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(cpe::band_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cpe::band_solve_kernel<<<b->S, cpe::SOLVE_THREADS, smem, (cudaStream_t)stream>>>(p);
    return (int)cudaGetLastError();
}
