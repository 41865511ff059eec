// cpe_frame.cu -- K1/K2: per-frame skeleton FK + camera projection + redescending reprojection loss, and (assemble
// mode) forward-mode Jacobians, J^T J / J^T r accumulation in the tree-sparse 3x3 block pattern, Marquardt damping
// and the per-frame Schur elimination of the coordinates that carry no motion prior.
//
// Reference arithmetic being replaced (paths under the reference repo; see include/cpe.h for the seam table):
//   rotations   shared/physical_education/utils.py:39-64, links.py:483         R = Rz(psi) Ry(theta) Rx(phi)
//   FK chains   links.py:496-519, cheetah.py:19-200, acinoset_misc.py:1581-1659
//   projection  acinoset_misc.py:1663-1696
//   residual    acinoset_misc.py:278-288 ; weights :211-231 ; cost :459-484 ; loss :2001-2015
//   motion      acinoset_misc.py:639-677 + collocation.py:58-68 (q-only third difference)
//   joints      links.py:552-586 ; inequalities cheetah.py:306-352
//
// One CTA per frame.  sm_100a, FP64 throughout (the parity bar is 1e-9 relative per residual); no tensor cores:
// this is small per-element algebra dominated by sincos/atan/exp and 3x3 products.
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"
#include "cpe_internal.cuh"

namespace cpe {

enum { MODE_EVAL = 0, MODE_EVAL_JAC = 1, MODE_ASSEMBLE = 2 };

struct FrameParams {
    cpe_model m;
    cpe_batch b;
    const double *q;
    const double *damp;
    cpe_eval_out eo;
    cpe_assemble_out ao;
    int ntask;       // C * W * 24
    int off_asm;     // shared-memory offset (doubles) of the assemble region
};

// ---- shared memory layout (in doubles) -------------------------------------------------------------------------
// common   : qs[54] sc[102] R[153] pos[72] cam[C*24] red[64] vin[32] eqr[26] pad
// deriv    : dR[459]
// assemble : A[216] eqg[156] gvec[54] diagH[54] Hb[n_blocks*9] U[1296] B[max(1296, ntask*9)]
//            group scratch aliases U
__host__ __device__ inline int common_doubles(int C) { return 54 + 102 + 153 + 72 + C * CPE_CAM_STRIDE + 64 + 32 + 26 + 1; }

__device__ __forceinline__ void build_rot(const double *sc, int link, int part, double *M) {
    const double sf = sc[6 * link + 0], cf = sc[6 * link + 1];
    const double st = sc[6 * link + 2], ct = sc[6 * link + 3];
    const double sp = sc[6 * link + 4], cp = sc[6 * link + 5];
    if (part == 0) {           // R
        M[0] = cp * ct; M[1] = cp * st * sf - sp * cf; M[2] = cp * st * cf + sp * sf;
        M[3] = sp * ct; M[4] = sp * st * sf + cp * cf; M[5] = sp * st * cf - cp * sf;
        M[6] = -st;     M[7] = ct * sf;                M[8] = ct * cf;
    } else if (part == 1) {    // dR/dphi
        M[0] = 0.0; M[1] = cp * st * cf + sp * sf;  M[2] = -cp * st * sf + sp * cf;
        M[3] = 0.0; M[4] = sp * st * cf - cp * sf;  M[5] = -sp * st * sf - cp * cf;
        M[6] = 0.0; M[7] = ct * cf;                 M[8] = -ct * sf;
    } else if (part == 2) {    // dR/dtheta
        M[0] = -cp * st; M[1] = cp * ct * sf; M[2] = cp * ct * cf;
        M[3] = -sp * st; M[4] = sp * ct * sf; M[5] = sp * ct * cf;
        M[6] = -ct;      M[7] = -st * sf;     M[8] = -st * cf;
    } else {                   // dR/dpsi
        M[0] = -sp * ct; M[1] = -sp * st * sf - cp * cf; M[2] = -sp * st * cf + cp * sf;
        M[3] = cp * ct;  M[4] = cp * st * sf - sp * cf;  M[5] = cp * st * cf + sp * sf;
        M[6] = 0.0;      M[7] = 0.0;                     M[8] = 0.0;
    }
}

// coefficient of the constant motion Hessian D3^T D3 coupling frame k with frame k-j (j = 0..3), nf frames.
__device__ __forceinline__ double motion_stencil(int k, int j, int nf) {
    const double c[4] = {-1.0, 3.0, -3.0, 1.0};
    double s = 0.0;
    // third-difference row r covers frames r..r+3, r = 0..nf-4
    for (int r = k - 3; r <= k - j; ++r) {
        if (r < 0 || r > nf - 4) continue;
        s += c[k - r] * c[k - j - r];
    }
    return s;
}

template <int MODE>
__global__ void __launch_bounds__(192) frame_kernel(const __grid_constant__ FrameParams p) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, nt = blockDim.x;
    const int fidx = blockIdx.x;
    const int N = p.b.N, C = p.b.C, W = p.b.W;
    const int s = fidx / N, k = fidx - s * N;
    const int nf = p.b.n_frames[s];
    if (k >= nf) return;
    constexpr bool DERIV = (MODE != MODE_EVAL);

    double *qs = sm;
    double *sc = qs + 54;
    double *R = sc + 102;
    double *pos = R + 153;
    double *cam = pos + 72;
    double *red = cam + C * CPE_CAM_STRIDE;
    double *vin = red + 64;
    double *eqr = vin + 32;
    double *dR = sm + common_doubles(C);                 // [17][3][9]
    double *A = sm + p.off_asm;                          // [24][9] : 6 sym + 3 grad
    double *eqg = A + 216;                               // [26][6]
    double *gvec = eqg + 156;                            // [54]
    double *diagH = gvec + 54;                           // [54]
    double *Hb = diagH + 54;                             // [n_blocks][9]
    double *U = Hb + p.m.n_blocks * 9;                   // [24][6][3][3]
    double *Bm = U + 1296;                               // [24][6][3][3], aliased by Ab[ntask][9]

    const double h = p.b.h[s];
    const double mu_eq = p.b.mu_eq, mu_in = p.b.mu_in;
    const double *qf = p.q + (size_t)fidx * 54;

    for (int i = tid; i < 54; i += nt) qs[i] = qf[i];
    for (int i = tid; i < C * CPE_CAM_STRIDE; i += nt) cam[i] = p.b.cams[(size_t)s * C * CPE_CAM_STRIDE + i];
    __syncthreads();
    for (int i = tid; i < 51; i += nt) sincos(qs[3 + i], &sc[2 * i], &sc[2 * i + 1]);
    __syncthreads();
    for (int it = tid; it < 17 * (DERIV ? 4 : 1); it += nt) {
        const int link = it % 17, part = it / 17;
        build_rot(sc, link, part, part == 0 ? (R + link * 9) : (dR + (link * 3 + part - 1) * 9));
    }
    __syncthreads();

    double cost4[4] = {0.0, 0.0, 0.0, 0.0};
    const double *cvec = p.b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
    const int n_ineq = p.m.n_ineq;
    // ---- marker positions, joint equalities, inequalities, motion cost ------------------------------------------
    for (int it = tid; it < 24 + 26 + n_ineq + 54; it += nt) {
        if (it < 24) {
            const int l = it;
            double x = qs[0], y = qs[1], z = qs[2];
#pragma unroll
            for (int j = 0; j < CPE_MAX_CHAIN; ++j) {
                const int link = p.m.chain_link[l * CPE_MAX_CHAIN + j];
                if (link < 0) break;
                const double *Rl = R + link * 9;
                const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                const double v0 = v[0], v1 = v[1], v2 = v[2];
                x += Rl[0] * v0 + Rl[1] * v1 + Rl[2] * v2;
                y += Rl[3] * v0 + Rl[4] * v1 + Rl[5] * v2;
                z += Rl[6] * v0 + Rl[7] * v1 + Rl[8] * v2;
            }
            pos[l * 3 + 0] = x; pos[l * 3 + 1] = y; pos[l * 3 + 2] = z;
            if (MODE != MODE_ASSEMBLE && p.eo.pose) {
                double *o = p.eo.pose + ((size_t)fidx * 24 + l) * 3;
                o[0] = x; o[1] = y; o[2] = z;
            }
        } else if (it < 24 + 26) {
            const int e = it - 24;
            const int la = p.m.eq_pairs[e * 4 + 0], lb = p.m.eq_pairs[e * 4 + 1];
            const int ia = p.m.eq_pairs[e * 4 + 2], ib = p.m.eq_pairs[e * 4 + 3];
            const double *Ra = R + la * 9, *Rb = R + lb * 9;
            const double c = Ra[ia] * Rb[ib] + Ra[3 + ia] * Rb[3 + ib] + Ra[6 + ia] * Rb[6 + ib];
            const double lam = p.b.lam_eq ? p.b.lam_eq[(size_t)fidx * 26 + e] : 0.0;
            const double t = c + lam / mu_eq;
            cost4[2] += 0.5 * mu_eq * t * t;
            if (MODE != MODE_ASSEMBLE && p.eo.c_eq) p.eo.c_eq[(size_t)fidx * 26 + e] = c;
            if (MODE == MODE_ASSEMBLE) {
                eqr[e] = mu_eq * t;
#pragma unroll
                for (int x = 0; x < 3; ++x) {
                    const double *dA = dR + (la * 3 + x) * 9, *dB = dR + (lb * 3 + x) * 9;
                    eqg[e * 6 + x] = dA[ia] * Rb[ib] + dA[3 + ia] * Rb[3 + ib] + dA[6 + ia] * Rb[6 + ib];
                    eqg[e * 6 + 3 + x] = Ra[ia] * dB[ib] + Ra[3 + ia] * dB[3 + ib] + Ra[6 + ia] * dB[6 + ib];
                }
            }
        } else if (it < 24 + 26 + n_ineq) {
            const int j = it - 50;
            const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
            const double lo = p.m.ineq_bounds[2 * j], up = p.m.ineq_bounds[2 * j + 1];
            const double d = qs[ip] - (im >= 0 ? qs[im] : 0.0);
            const double lam = p.b.lam_in ? p.b.lam_in[(size_t)fidx * n_ineq + j] : 0.0;
            const double t = d + lam / mu_in;
            const double v = t - fmin(fmax(t, lo), up);
            cost4[3] += 0.5 * mu_in * v * v;
            vin[j] = v;
            if (MODE != MODE_ASSEMBLE && p.eo.v_in) p.eo.v_in[(size_t)fidx * n_ineq + j] = v;
        } else {
            const int c = it - 50 - n_ineq;
            const double w = p.m.wq[c];
            if (w != 0.0 && k >= 3) {
                const double d3 = (qs[c] - 3.0 * qf[c - 54] + 3.0 * qf[c - 108] - qf[c - 162]) / (h * h);
                cost4[1] += w * d3 * d3;
            }
        }
    }
    __syncthreads();

    // ---- (camera, hypothesis, marker) tasks: projection + robust loss -------------------------------------------
    const double la_ = p.b.loss_a, lb_ = p.b.loss_b, lc_ = p.b.loss_c;
    const double ea = exp(la_), eb = exp(lb_), ec = exp(lc_);
    for (int task = tid; task < p.ntask; task += nt) {
        const int l = task % 24;
        const int cw = task / 24;
        const int c = cw / W, w = cw - c * W;
        const size_t midx = (((size_t)fidx * C + c) * 24 + l) * W + w;
        double um, vm, gate;
        if (p.b.meas_is_f64) {
            const double4 mv = reinterpret_cast<const double4 *>(p.b.meas)[midx];
            um = mv.x; vm = mv.y; gate = mv.z;
        } else {
            const float4 mv = reinterpret_cast<const float4 *>(p.b.meas)[midx];
            um = (double)mv.x; vm = (double)mv.y; gate = (double)mv.z;
        }
        const double *cm = cam + c * CPE_CAM_STRIDE;
        const double wgt = gate * p.m.meas_w[w * 24 + l] * cm[20];
        Proj pr;
        project_point<DERIV>(pos + l * 3, cm, p.b.cam_model, pr);
        const double ru = pr.u - um, rv = pr.v - vm;
        if (MODE != MODE_ASSEMBLE && p.eo.resid) {
            double *o = p.eo.resid + midx * 2;
            o[0] = ru; o[1] = rv;
        }
        double rho_u, drho_u, rho_v, drho_v;
        const double eu = wgt * ru, ev = wgt * rv;
        redescending(fabs(eu), la_, lb_, lc_, ea, eb, ec, rho_u, drho_u);
        redescending(fabs(ev), la_, lb_, lc_, ea, eb, ec, rho_v, drho_v);
        cost4[0] += rho_u + rho_v;
        if (MODE == MODE_ASSEMBLE) {
            const double Eu = fabs(eu), Ev = fabs(ev);
            const double gu = drho_u * (eu > 0.0 ? 1.0 : (eu < 0.0 ? -1.0 : 0.0)) * wgt;
            const double gv = drho_v * (ev > 0.0 ? 1.0 : (ev < 0.0 ? -1.0 : 0.0)) * wgt;
            const double hu = (Eu > 0.0 ? fmax(drho_u / Eu, 0.0) : 0.0) * wgt * wgt;
            const double hv = (Ev > 0.0 ? fmax(drho_v / Ev, 0.0) : 0.0) * wgt * wgt;
            double *ab = Bm + task * 9;
            ab[0] = hu * pr.du[0] * pr.du[0] + hv * pr.dv[0] * pr.dv[0];
            ab[1] = hu * pr.du[0] * pr.du[1] + hv * pr.dv[0] * pr.dv[1];
            ab[2] = hu * pr.du[0] * pr.du[2] + hv * pr.dv[0] * pr.dv[2];
            ab[3] = hu * pr.du[1] * pr.du[1] + hv * pr.dv[1] * pr.dv[1];
            ab[4] = hu * pr.du[1] * pr.du[2] + hv * pr.dv[1] * pr.dv[2];
            ab[5] = hu * pr.du[2] * pr.du[2] + hv * pr.dv[2] * pr.dv[2];
            ab[6] = gu * pr.du[0] + gv * pr.dv[0];
            ab[7] = gu * pr.du[1] + gv * pr.dv[1];
            ab[8] = gu * pr.du[2] + gv * pr.dv[2];
        }
        if (MODE == MODE_EVAL_JAC && p.eo.jac) {
            double *J = p.eo.jac + midx * 36;
            for (int slot = 0; slot < 6; ++slot) {
                const int link = slot == 0 ? -1 : p.m.chain_link[l * CPE_MAX_CHAIN + slot - 1];
                for (int a = 0; a < 3; ++a) {
                    double c0 = 0.0, c1 = 0.0, c2 = 0.0;
                    if (slot == 0) {
                        c0 = a == 0; c1 = a == 1; c2 = a == 2;
                    } else if (link >= 0) {
                        const double *D = dR + (link * 3 + a) * 9;
                        const double *v = cvec + (l * CPE_MAX_CHAIN + slot - 1) * 3;
                        c0 = D[0] * v[0] + D[1] * v[1] + D[2] * v[2];
                        c1 = D[3] * v[0] + D[4] * v[1] + D[5] * v[2];
                        c2 = D[6] * v[0] + D[7] * v[1] + D[8] * v[2];
                    }
                    J[slot * 3 + a] = pr.du[0] * c0 + pr.du[1] * c1 + pr.du[2] * c2;
                    J[18 + slot * 3 + a] = pr.dv[0] * c0 + pr.dv[1] * c1 + pr.dv[2] * c2;
                }
            }
        }
    }

    // ---- deterministic block reduction of the four cost partials -------------------------------------------------
    if (MODE != MODE_ASSEMBLE && p.eo.frame_cost) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            double v = cost4[i];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if ((tid & 31) == 0) red[(tid >> 5) * 4 + i] = v;
        }
        __syncthreads();
        if (tid < 4) {
            double v = 0.0;
            for (int wi = 0; wi < (nt + 31) / 32; ++wi) v += red[wi * 4 + tid];
            p.eo.frame_cost[(size_t)fidx * 4 + tid] = v;
        }
    }
    if (MODE != MODE_ASSEMBLE) return;

    // =============================== assemble ====================================================================
    __syncthreads();
    const int ncw = C * W;
    for (int i = tid; i < 216; i += nt) {
        const int l = i / 9, comp = i - l * 9;
        double v = 0.0;
        for (int cw = 0; cw < ncw; ++cw) v += Bm[(cw * 24 + l) * 9 + comp];
        A[i] = v;
    }
    __syncthreads();
    // U = d pos_l / d (node coords), B = A_l U   (columns)
    for (int it = tid; it < 24 * 6 * 3; it += nt) {
        const int a = it % 3, slot = (it / 3) % 6, l = it / 18;
        double c0, c1, c2;
        if (slot == 0) {
            c0 = a == 0; c1 = a == 1; c2 = a == 2;
        } else {
            const int link = p.m.chain_link[l * CPE_MAX_CHAIN + slot - 1];
            if (link < 0) continue;
            const double *D = dR + (link * 3 + a) * 9;
            const double *v = cvec + (l * CPE_MAX_CHAIN + slot - 1) * 3;
            const double v0 = v[0], v1 = v[1], v2 = v[2];
            c0 = D[0] * v0 + D[1] * v1 + D[2] * v2;
            c1 = D[3] * v0 + D[4] * v1 + D[5] * v2;
            c2 = D[6] * v0 + D[7] * v1 + D[8] * v2;
        }
        const double *Al = A + l * 9;
        double *u = U + it * 3, *bb = Bm + it * 3;
        u[0] = c0; u[1] = c1; u[2] = c2;
        bb[0] = Al[0] * c0 + Al[1] * c1 + Al[2] * c2;
        bb[1] = Al[1] * c0 + Al[3] * c1 + Al[4] * c2;
        bb[2] = Al[2] * c0 + Al[4] * c1 + Al[5] * c2;
    }
    __syncthreads();
    // tree-sparse block Hessian entries and gradient entries
    const int nH = p.m.n_blocks * 9;
    for (int e = tid; e < nH + 54; e += nt) {
        if (e < nH) {
            const int blk = e / 9, a = (e % 9) / 3, b = e % 3;
            const int ni = p.m.blk_nodes[2 * blk], nj = p.m.blk_nodes[2 * blk + 1];
            double sum = 0.0;
            for (int it = p.m.blk_item_ptr[blk]; it < p.m.blk_item_ptr[blk + 1]; ++it) {
                const int l = p.m.blk_items[it];
                const int si = p.m.slot_of[l * 18 + ni], sj = p.m.slot_of[l * 18 + nj];
                const double *u = U + ((l * 6 + si) * 3 + a) * 3;
                const double *bb = Bm + ((l * 6 + sj) * 3 + b) * 3;
                sum += u[0] * bb[0] + u[1] * bb[1] + u[2] * bb[2];
            }
            for (int it = p.m.blk_eq_ptr[blk]; it < p.m.blk_eq_ptr[blk + 1]; ++it) {
                const int code = p.m.blk_eq_items[it];
                const int ke = code >> 2, rs = (code >> 1) & 1, cs = code & 1;
                sum += mu_eq * eqg[ke * 6 + rs * 3 + a] * eqg[ke * 6 + cs * 3 + b];
            }
            Hb[e] = sum;
        } else {
            const int c = e - nH;
            const int node = c / 3, a = c - node * 3;
            const int blk = p.m.blk_id[node * 18 + node];
            double sum = 0.0;
            for (int it = p.m.blk_item_ptr[blk]; it < p.m.blk_item_ptr[blk + 1]; ++it) {
                const int l = p.m.blk_items[it];
                const int si = p.m.slot_of[l * 18 + node];
                const double *u = U + ((l * 6 + si) * 3 + a) * 3;
                const double *bl = A + l * 9 + 6;
                sum += u[0] * bl[0] + u[1] * bl[1] + u[2] * bl[2];
            }
            if (node >= 1) {
                const int link = node - 1;
                for (int ke = 0; ke < 26; ++ke) {
                    if (p.m.eq_pairs[ke * 4] == link) sum += eqr[ke] * eqg[ke * 6 + a];
                    if (p.m.eq_pairs[ke * 4 + 1] == link) sum += eqr[ke] * eqg[ke * 6 + 3 + a];
                }
            }
            const double w = p.m.wq[c];
            if (w != 0.0) {
                // motion gradient: 2 w / h^2 * sum_r coef[k-r] d3_r, rows r = k-3..k within [0, nf-4]
                const double cf[4] = {-1.0, 3.0, -3.0, 1.0};
                double acc = 0.0;
                for (int r = k - 3; r <= k; ++r) {
                    if (r < 0 || r > nf - 4) continue;
                    const double *qr = p.q + ((size_t)s * N + r) * 54 + c;
                    const double d3 = (qr[162] - 3.0 * qr[108] + 3.0 * qr[54] - qr[0]) / (h * h);
                    acc += cf[k - r] * d3;
                }
                sum += 2.0 * w / (h * h) * acc;
            }
            gvec[c] = sum;
        }
    }
    __syncthreads();
    if (tid == 0) {   // PHR inequality rows: mu v grad d, mu grad d grad d^T where active (serial => deterministic)
        for (int j = 0; j < n_ineq; ++j) {
            const double v = vin[j];
            if (v == 0.0) continue;
            const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
            gvec[ip] += mu_in * v;
            const int np_ = ip / 3, ap = ip % 3;
            Hb[p.m.blk_id[np_ * 18 + np_] * 9 + ap * 4] += mu_in;
            if (im >= 0) {
                gvec[im] -= mu_in * v;
                const int nm = im / 3, am = im % 3;
                Hb[p.m.blk_id[nm * 18 + nm] * 9 + am * 4] += mu_in;
                if (np_ == nm) {
                    Hb[p.m.blk_id[np_ * 18 + np_] * 9 + ap * 3 + am] -= mu_in;
                    Hb[p.m.blk_id[np_ * 18 + np_] * 9 + am * 3 + ap] -= mu_in;
                } else if (np_ > nm) {
                    Hb[p.m.blk_id[np_ * 18 + nm] * 9 + ap * 3 + am] -= mu_in;
                } else {
                    Hb[p.m.blk_id[nm * 18 + np_] * 9 + am * 3 + ap] -= mu_in;
                }
            }
        }
    }
    __syncthreads();
    const double h4 = h * h * h * h;
    const double m0 = motion_stencil(k, 0, nf);
    for (int c = tid; c < 54; c += nt) {
        const int node = c / 3, a = c - node * 3;
        const double d = Hb[p.m.blk_id[node * 18 + node] * 9 + a * 4] + m0 * 2.0 * p.m.wq[c] / h4;
        diagH[c] = d;
        if (p.ao.diagH) p.ao.diagH[(size_t)fidx * 54 + c] = d;
        if (p.ao.g) p.ao.g[(size_t)fidx * 54 + c] = gvec[c];
    }
    auto Hget = [&](int pr, int pc) -> double {
        const int ni = pr / 3, nj = pc / 3;
        if (ni >= nj) {
            const int id = p.m.blk_id[ni * 18 + nj];
            return id < 0 ? 0.0 : Hb[id * 9 + (pr - ni * 3) * 3 + (pc - nj * 3)];
        }
        const int id = p.m.blk_id[nj * 18 + ni];
        return id < 0 ? 0.0 : Hb[id * 9 + (pc - nj * 3) * 3 + (pr - ni * 3)];
    };
    if (p.ao.Hraw) {
        double *o = p.ao.Hraw + (size_t)fidx * 54 * 54;
        for (int e = tid; e < 54 * 54; e += nt) o[e] = Hget(e / 54, e % 54);
    }
    __syncthreads();
    if (!p.ao.S) return;

    // ---- per-frame elimination of the dependent coordinates, one warp per group ---------------------------------
    const double damp = p.damp[s];
    constexpr int GD = CPE_GRP_MAX_DEP, GN = CPE_GRP_MAX_NB, TW = GD + GN + 1;
    constexpr int GSCR = GD * TW + GD * GN + GN * GN + GN;     // doubles of scratch per group
    double *scr = U;                                           // U/B no longer needed
    const int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int g = wid; g < p.m.n_groups; g += nw) {
        double *T = scr + g * GSCR;
        double *Ddi = T + GD * TW;
        double *dS = Ddi + GD * GN;
        double *dg = dS + GN * GN;
        const int nd = p.m.grp_ndep[g], nn = p.m.grp_nnb[g];
        const int *dep = p.m.grp_dep + g * GD, *nb = p.m.grp_nb + g * GN;
        for (int e = lane; e < GD * TW; e += 32) {
            const int i = e / TW, j = e - i * TW;
            double v = 0.0;
            if (i < nd) {
                if (j < GD) {
                    if (j < nd) v = Hget(dep[i], dep[j]) + (i == j ? damp * diagH[dep[i]] : 0.0);
                } else if (j < GD + GN) {
                    if (j - GD < nn) v = Hget(dep[i], nb[j - GD]);
                } else {
                    v = gvec[dep[i]];
                }
            } else if (i == j) {
                v = 1.0;
            }
            T[e] = v;
            if (j >= GD && j < GD + GN) Ddi[i * GN + (j - GD)] = v;
        }
        __syncwarp();
        // Gauss-Jordan on the SPD pivot block (no pivoting needed)
        for (int pc = 0; pc < nd; ++pc) {
            const double ipiv = 1.0 / T[pc * TW + pc];
            double nv[(GD * TW + 31) / 32];
#pragma unroll
            for (int r = 0; r < (GD * TW + 31) / 32; ++r) {
                const int e = lane + 32 * r;
                if (e < GD * TW) {
                    const int i = e / TW, j = e - i * TW;
                    const double prow = T[pc * TW + j] * ipiv;
                    nv[r] = (i == pc) ? prow : T[e] - T[i * TW + pc] * prow;
                }
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < (GD * TW + 31) / 32; ++r) {
                const int e = lane + 32 * r;
                if (e < GD * TW) T[e] = nv[r];
            }
            __syncwarp();
        }
        // dS = Ddi^T Y , dg = Ddi^T y ; export Y, y
        for (int e = lane; e < GN * GN + GN; e += 32) {
            if (e < GN * GN) {
                const int a = e / GN, b = e - a * GN;
                double v = 0.0;
                for (int i = 0; i < nd; ++i) v += Ddi[i * GN + a] * T[i * TW + GD + b];
                dS[e] = v;
            } else {
                const int a = e - GN * GN;
                double v = 0.0;
                for (int i = 0; i < nd; ++i) v += Ddi[i * GN + a] * T[i * TW + GD + GN];
                dg[a] = v;
            }
        }
        double *Yo = p.ao.Y + ((size_t)fidx * p.m.n_groups + g) * GD * GN;
        double *yo = p.ao.y + ((size_t)fidx * p.m.n_groups + g) * GD;
        for (int e = lane; e < GD * GN; e += 32) {
            const int i = e / GN, j = e - i * GN;
            Yo[e] = (i < nd && j < nn) ? T[i * TW + GD + j] : 0.0;
        }
        if (lane < GD) yo[lane] = lane < nd ? T[lane * TW + GD + GN] : 0.0;
    }
    __syncthreads();
    double *So = p.ao.S + (size_t)fidx * 784;
    for (int e = tid; e < 784 + 28; e += nt) {
        if (e < 784) {
            const int r = e / 28, c = e - r * 28;
            const int pr = p.m.ind_idx[r], pc = p.m.ind_idx[c];
            double v = Hget(pr, pc) + (r == c ? damp * diagH[pr] : 0.0);
            for (int g = 0; g < p.m.n_groups; ++g) {
                const int a = p.m.grp_nbpos[g * 28 + r], b = p.m.grp_nbpos[g * 28 + c];
                if (a >= 0 && b >= 0) v -= scr[g * GSCR + GD * TW + GD * GN + a * GN + b];
            }
            So[e] = v;
        } else {
            const int r = e - 784;
            double v = gvec[p.m.ind_idx[r]];
            for (int g = 0; g < p.m.n_groups; ++g) {
                const int a = p.m.grp_nbpos[g * 28 + r];
                if (a >= 0) v -= scr[g * GSCR + GD * TW + GD * GN + GN * GN + a];
            }
            p.ao.ghat[(size_t)fidx * 28 + r] = v;
        }
    }
}

static int frame_threads(int ntask) {
    int t = ((ntask + 31) / 32) * 32;
    if (t < 160) t = 160;      // >= 5 warps: one per elimination group, and enough lanes for the aux items
    if (t > 192) t = 192;
    return t;
}

template <int MODE>
static int launch_frame(const cpe_model *model, const cpe_batch *b, const double *q, const double *damp,
                        const cpe_eval_out *eo, const cpe_assemble_out *ao, cudaStream_t st) {
    if (!model || !b || !q) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS || (b->W != 1 && b->W != 3)) return CPE_E_DIM;
    if (model->n_ineq > CPE_MAX_INEQ || model->n_groups > CPE_MAX_GROUPS) return CPE_E_DIM;
    FrameParams p;
    p.m = *model;
    p.b = *b;
    p.q = q;
    p.damp = damp;
    p.eo = eo ? *eo : cpe_eval_out{};
    p.ao = ao ? *ao : cpe_assemble_out{};
    p.ntask = b->C * b->W * 24;
    int doubles = common_doubles(b->C);
    if (MODE != MODE_EVAL) doubles += 459;
    p.off_asm = doubles;
    if (MODE == MODE_ASSEMBLE) {
        if (p.ao.S && (!damp || !p.ao.ghat || !p.ao.Y || !p.ao.y)) return CPE_E_NULL;
        const int bsz = p.ntask * 9 > 1296 ? p.ntask * 9 : 1296;
        doubles += 216 + 156 + 54 + 54 + model->n_blocks * 9 + 1296 + bsz;
        const int gscr = CPE_GRP_MAX_DEP * (CPE_GRP_MAX_DEP + CPE_GRP_MAX_NB + 1) + CPE_GRP_MAX_DEP * CPE_GRP_MAX_NB +
                         CPE_GRP_MAX_NB * CPE_GRP_MAX_NB + CPE_GRP_MAX_NB;
        if (model->n_groups * gscr > 1296 + bsz) return CPE_E_DIM;
    }
    const size_t smem = (size_t)doubles * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(frame_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    const int threads = frame_threads(p.ntask);
    frame_kernel<MODE><<<b->S * b->N, threads, smem, st>>>(p);
    e = cudaGetLastError();
    return (int)e;
}

}  // namespace cpe

extern "C" int cpe_assemble_f64(const cpe_model *model, const cpe_batch *batch, const double *q, const double *damp,
                                const cpe_assemble_out *out, void *stream) {
    if (!out) return CPE_E_NULL;
    return cpe::launch_frame<cpe::MODE_ASSEMBLE>(model, batch, q, damp, nullptr, out, (cudaStream_t)stream);
}
