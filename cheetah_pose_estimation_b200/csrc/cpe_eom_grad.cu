// cpe_eom_grad.cu -- vector-Jacobian products of rows 0-5 of the equations of motion with respect to q, dq, ddq.
//
// Reference being replaced: the derivatives ASL evaluates for the EOM-slack term of the kinetic objective
// (misc.eom_slack_cost, acinoset_misc.py:631-636, on the rows of calc_grf_eom, acinoset_opt.py:120-174; objective
// acinoset_opt.py:905-921).  The code is the adjoint sweep of the generated straight-line program
// (cpe_eom_grad_gen.cuh, written by oracle/derive_eom_grad.py from the reference's own symbolic model).
//
// Mapping: one thread per frame, as the forward rows (cpe_eom.cu); ~2 k adjoint temporaries live in local memory, so the
// kernel is bound by FP64 issue and L1 traffic of its own spills, not by HBM (486 B in, 1.3 KB out per frame).
#include <cuda_runtime.h>
#include <stdint.h>

#include "cpe.h"
#include "cpe_eom_grad_gen.cuh"

namespace cpe {

__global__ void __launch_bounds__(64) eom_vjp_kernel(int S, int N, const int32_t *n_frames, const double *q, const double *dq,
                                                     const double *ddq, const double *grf, const double *par, const double *w,
                                                     double *eom, double *g_q, double *g_dq, double *g_ddq) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)S * N) return;
    const int s = (int)(idx / N), k = (int)(idx - (long long)s * N);
    double r[6], gq[54], gdq[54], gddq[54];
#pragma unroll
    for (int i = 0; i < 54; ++i) gq[i] = gdq[i] = gddq[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) r[i] = 0.0;
    if (k < n_frames[s])
        eom_rows_0_5_vjp(q + idx * 54, dq + idx * 54, ddq + idx * 54, grf + idx * CPE_EOM_N_FORCES, par + (size_t)s * CPE_EOM_N_PAR,
                         w ? w + idx * 6 : nullptr, r, gq, gdq, gddq);
    if (eom)
        for (int i = 0; i < 6; ++i) eom[idx * 6 + i] = r[i];
    for (int i = 0; i < 54; ++i) {                     // frames beyond n_frames[s] get zeros
        g_q[idx * 54 + i] = gq[i];
        g_dq[idx * 54 + i] = gdq[i];
        g_ddq[idx * 54 + i] = gddq[i];
    }
}

}  // namespace cpe

extern "C" int cpe_eom_vjp_f64(const cpe_batch *b, const double *q, const double *dq, const double *ddq, const double *grf,
                               const double *par, const double *w, double *eom, double *g_q, double *g_dq, double *g_ddq,
                               void *stream) {
    if (!b || !q || !dq || !ddq || !grf || !par || !g_q || !g_dq || !g_ddq || !b->n_frames) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0) return CPE_E_DIM;
    const long long n = (long long)b->S * b->N;
    cpe::eom_vjp_kernel<<<(unsigned)((n + 63) / 64), 64, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, q, dq, ddq, grf, par,
                                                                                  w, eom, g_q, g_dq, g_ddq);
    return (int)cudaGetLastError();
}
