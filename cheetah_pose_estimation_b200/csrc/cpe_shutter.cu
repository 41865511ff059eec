// cpe_shutter.cu -- K6: the pieces of the shutter-delay estimation that are not per-residual work
// (acinoset_misc.py:182-183, 273-286).  Camera c sees frame n displaced by dq_base[n] tau_c +
// ddq_base[n] tau_c^2.  In the q-only formulation the rates are backward differences of the base x, y, z
// (collocation.py:58-68 eliminated); the reference's free first-node values take what zeroes their motion cost, as in
// the saved dq / ddq:
//     v_n = (b_n - b_{n-1}) / h            (n >= 1)        v_0 = v_1 - h a_1 = (-2 b_0 + 3 b_1 - b_2) / h
//     a_n = (b_n - 2 b_{n-1} + b_{n-2}) / h^2   (n >= 2)   a_0 = a_1 = a_2
// (sequences of fewer than 3 / 2 frames have no acceleration / velocity).
//   cpe_base_rates_f64     q -> dq_base, ddq_base                                   one thread per (sequence, frame, axis)
//   cpe_shutter_terms_f64  K2's per-camera point-space blocks -> the gradient through the rates (the ADJOINT of the
//                          operator above applied to sum_c tau_c b_nc and sum_c tau_c^2 b_nc), d cost / d tau_c and
//                          the Gauss-Newton curvature of tau_c (one warp per (sequence, camera), fixed-order sums)
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"

namespace cpe {

__global__ void __launch_bounds__(128) base_rates_kernel(cpe_batch b, const double *q, double *dqb, double *ddqb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int N = b.N;
    if (i >= b.S * N * 3) return;
    const int j = i % 3, n = (i / 3) % N, s = i / (3 * N);
    const int nf = b.n_frames[s];
    double v = 0.0, a = 0.0;
    if (n < nf) {
        const double h = b.h[s];
        const double *bq = q + (size_t)s * N * 54 + j;            // b(k) = bq[k * 54]
        if (nf >= 3) {
            const int k = n < 2 ? 2 : n;
            a = ((bq[k * 54] - bq[(k - 1) * 54]) / h - (bq[(k - 1) * 54] - bq[(k - 2) * 54]) / h) / h;
        }
        if (nf >= 2) {
            const int k = n < 1 ? 1 : n;
            v = (bq[k * 54] - bq[(k - 1) * 54]) / h;
            if (n == 0) v -= h * a;
        }
    }
    dqb[i] = v;
    ddqb[i] = a;
}

// g_extra[s, m, j]: d/d b_m of  sum_n ( P_n . v_n + Q_n . a_n ),  P_n = sum_c tau_c b_nc,  Q_n = sum_c tau_c^2 b_nc
__global__ void __launch_bounds__(128) shutter_grad_kernel(cpe_batch b, const double *blk, double *g_extra) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int N = b.N, C = b.C;
    if (i >= b.S * N * 3) return;
    const int j = i % 3, m = (i / 3) % N, s = i / (3 * N);
    const int nf = b.n_frames[s];
    double g = 0.0;
    if (m < nf && nf >= 2) {
        const double h = b.h[s];
        const double *tau = b.shutter_tau + (size_t)s * C;
        auto PQ = [&](int n, double &P, double &Q) {
            P = 0.0; Q = 0.0;
            if (n >= nf) return;
            const double *bb = blk + ((size_t)(s * N + n) * C) * 9 + 6 + j;
            for (int c = 0; c < C; ++c) {
                const double t = tau[c], v = bb[c * 9];
                P += t * v;
                Q += t * t * v;
            }
        };
        double P0, Q0, P1, Q1, P2, Q2;
        PQ(m, P0, Q0); PQ(m + 1, P1, Q1); PQ(m + 2, P2, Q2);
        // velocity rows n >= 1: + P_n / h at m = n, - P_n / h at m = n - 1
        if (m >= 1) g += P0 / h;
        g -= P1 / h;
        // velocity row 0
        double Pf, Qf, Pa, Qa;
        PQ(0, Pf, Qf);
        if (nf >= 3) {
            if (m == 0) g += -2.0 * Pf / h;
            else if (m == 1) g += 3.0 * Pf / h;
            else if (m == 2) g += -Pf / h;
        } else {
            if (m == 0) g += -Pf / h;
            else if (m == 1) g += Pf / h;
        }
        if (nf >= 3) {
            // acceleration rows n >= 2 with weights Q'_n (Q'_2 collects the copies a_0 = a_1 = a_2)
            PQ(1, Pa, Qa);
            double Q2f, Pd;
            PQ(2, Pd, Q2f);
            const double Qp2 = Qf + Qa + Q2f;
            auto Qp = [&](int n, double Qn) { return n < 2 ? 0.0 : (n == 2 ? Qp2 : Qn); };
            g += (Qp(m, Q0) - 2.0 * Qp(m + 1, Q1) + Qp(m + 2, Q2)) / (h * h);
        }
    }
    g_extra[i] = g;
}

__global__ void __launch_bounds__(128) shutter_tau_kernel(cpe_batch b, const double *blk, double *g_tau, double *curv) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int N = b.N, C = b.C;
    if (w >= b.S * C) return;
    const int s = w / C, c = w - s * C;
    const int nf = b.n_frames[s];
    const double tau = b.shutter_tau[w];
    double g = 0.0, k = 0.0;
    for (int n = lane; n < nf; n += 32) {
        const double *v = b.base_dq + (size_t)(s * N + n) * 3, *a = b.base_ddq + (size_t)(s * N + n) * 3;
        const double *A = blk + ((size_t)(s * N + n) * C + c) * 9;
        const double w0 = v[0] + 2.0 * tau * a[0], w1 = v[1] + 2.0 * tau * a[1], w2 = v[2] + 2.0 * tau * a[2];
        g += A[6] * w0 + A[7] * w1 + A[8] * w2;
        k += A[0] * w0 * w0 + A[3] * w1 * w1 + A[5] * w2 * w2 + 2.0 * (A[1] * w0 * w1 + A[2] * w0 * w2 + A[4] * w1 * w2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        g += __shfl_down_sync(0xffffffffu, g, o);
        k += __shfl_down_sync(0xffffffffu, k, o);
    }
    if (lane == 0) { g_tau[w] = g; curv[w] = k; }
}

}  // namespace cpe

extern "C" int cpe_base_rates_f64(const cpe_batch *b, const double *q, double *dq_base, double *ddq_base, void *stream) {
    if (!b || !q || !dq_base || !ddq_base || !b->n_frames || !b->h) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0) return CPE_E_DIM;
    const long long n = (long long)b->S * b->N * 3;
    cpe::base_rates_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(*b, q, dq_base, ddq_base);
    return (int)cudaGetLastError();
}

extern "C" int cpe_shutter_terms_f64(const cpe_batch *b, const double *cam_blocks, double *g_extra, double *g_tau,
                                     double *curv_tau, void *stream) {
    if (!b || !cam_blocks || !g_extra || !g_tau || !curv_tau) return CPE_E_NULL;
    if (!b->shutter_tau || !b->base_dq || !b->base_ddq || !b->n_frames || !b->h) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS) return CPE_E_DIM;
    const long long n = (long long)b->S * b->N * 3;
    cpe::shutter_grad_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(*b, cam_blocks, g_extra);
    const long long nw = (long long)b->S * b->C;
    cpe::shutter_tau_kernel<<<(unsigned)((nw * 32 + 127) / 128), 128, 0, (cudaStream_t)stream>>>(*b, cam_blocks, g_tau,
                                                                                               curv_tau);
    return (int)cudaGetLastError();
}
