// cpe_solve.cu -- K4: step application, K5: step acceptance, plus the small reductions (K3, the block-banded
// solve, lives in cpe_bandsolve.cu).  Replaces what IPOPT + MA97 do inside the reference's solver seam
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
    if (k >= p.b.n_frames[s] || (p.b.active && !p.b.active[s])) return;
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
__global__ void reduce_cost_kernel(int S, int N, const int32_t *n_frames, const int32_t *active, const double *frame_cost,
                                   double *cost) {
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (s >= S || (active && !active[s])) return;
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
                                                          const double *cost_extra, const double *frame_pred, double *q, const double *q_try,
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
            double Ft = ct[0] + ct[1] + ct[2] + ct[3];
            if (cost_extra) Ft += cost_extra[(size_t)s * 2] + cost_extra[(size_t)s * 2 + 1];   // monocular priors
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

// per-sequence exit status of an LM / AL solve (cpe.h CPE_STATUS_*)
__global__ void finalize_status_kernel(int S, const double *cost, const double *cost_extra, const int32_t *active,
                                       const int32_t *pivot, int32_t *status) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    double f = cost[s * 4] + cost[s * 4 + 1] + cost[s * 4 + 2] + cost[s * 4 + 3];
    if (cost_extra) f += cost_extra[s * 2] + cost_extra[s * 2 + 1];
    int st = 0;
    if (pivot && pivot[s]) st |= CPE_STATUS_PIVOT;
    if (active && active[s]) st |= CPE_STATUS_MAXITER;
    if (!(fabs(f) <= 1.7976931348623157e308)) st |= CPE_STATUS_NONFINITE;      // NaN or infinity
    status[s] = st;
}

__global__ void update_multipliers_kernel(int S, int N, int n_ineq, const int32_t *n_frames, const int32_t *active,
                                          double mu_eq, double mu_in, const double *c_eq, const double *v_in,
                                          double *lam_eq, double *lam_in) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t per = 26 + n_ineq;
    if (i >= (size_t)S * N * per) return;
    const size_t fidx = i / per;
    const int e = (int)(i - fidx * per);
    const int s = (int)(fidx / N), k = (int)(fidx - (size_t)s * N);
    if (k >= n_frames[s] || (active && !active[s])) return;
    if (e < 26) lam_eq[fidx * 26 + e] += mu_eq * c_eq[fidx * 26 + e];
    else lam_in[fidx * n_ineq + e - 26] = mu_in * v_in[fidx * n_ineq + e - 26];
}

}  // namespace cpe

extern "C" int cpe_version(void) { return CPE_VERSION; }

extern "C" int cpe_reduce_cost_f64(const cpe_batch *b, const double *frame_cost, double *cost, void *stream) {
    if (!b || !frame_cost || !cost) return CPE_E_NULL;
    cpe::reduce_cost_kernel<<<(b->S + 3) / 4, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, b->active, frame_cost, cost);
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
    cpe::step_accept_kernel<<<b->S, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, cost_try, nullptr, frame_pred,
                                                                    q, q_try, state, counters, active, ftol);
    return (int)cudaGetLastError();
}

extern "C" int cpe_step_accept_prior_f64(const cpe_batch *b, const double *cost_try, const double *cost_extra,
                                         const double *frame_pred, double *q, const double *q_try, double *state,
                                         int32_t *counters, int32_t *active, double ftol, void *stream) {
    if (!b || !cost_try || !cost_extra || !frame_pred || !q || !q_try || !state || !counters) return CPE_E_NULL;
    cpe::step_accept_kernel<<<b->S, 128, 0, (cudaStream_t)stream>>>(b->S, b->N, b->n_frames, cost_try, cost_extra,
                                                                    frame_pred, q, q_try, state, counters, active, ftol);
    return (int)cudaGetLastError();
}

extern "C" int cpe_finalize_status_f64(const cpe_batch *b, const double *cost, const double *cost_extra,
                                       const int32_t *active, const int32_t *pivot, int32_t *status, void *stream) {
    if (!b || !cost || !status) return CPE_E_NULL;
    cpe::finalize_status_kernel<<<(b->S + 127) / 128, 128, 0, (cudaStream_t)stream>>>(b->S, cost, cost_extra, active, pivot,
                                                                                      status);
    return (int)cudaGetLastError();
}

extern "C" int cpe_update_multipliers_f64(const cpe_model *model, const cpe_batch *b, const double *c_eq, const double *v_in,
                                          double *lam_eq, double *lam_in, void *stream) {
    if (!model || !b || !c_eq || !v_in || !lam_eq || !lam_in) return CPE_E_NULL;
    const size_t n = (size_t)b->S * b->N * (26 + model->n_ineq);
    cpe::update_multipliers_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        b->S, b->N, model->n_ineq, b->n_frames, b->active, b->mu_eq, b->mu_in, c_eq, v_in, lam_eq, lam_in);
    return (int)cudaGetLastError();
}

extern "C" int cpe_eval_host_f64(const cpe_model *model, const cpe_batch *bd, const double *h_q, const void *h_meas,
                                 double *d_q, void *d_meas, const cpe_eval_out *out_dev, double *d_cost, double *h_cost,
                                 double *h_resid, void *stream) {
    if (!model || !bd || !h_q || !d_q || !out_dev || !d_cost || !h_cost) return CPE_E_NULL;
    if (!out_dev->frame_cost || (h_meas && !d_meas) || (!h_meas && !bd->meas)) return CPE_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t frames = (size_t)bd->S * bd->N;
    if (bd->meas_layout >= CPE_MEAS_F32X2G) return CPE_E_ARG;       // two arrays cross the bus: cpe_eval_host_uvg_f64
    const size_t rec = bd->meas_layout == CPE_MEAS_F64X4 ? 32 : (bd->meas_layout == CPE_MEAS_F32X3 ? 12 : 16);
    const size_t mbytes = frames * bd->C * 24 * bd->W * rec;
    cudaError_t e = cudaMemcpyAsync(d_q, h_q, frames * 54 * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    cpe_batch b = *bd;
    if (h_meas) {                  // NULL: the detections are already resident in bd->meas (they do not change between
        e = cudaMemcpyAsync(d_meas, h_meas, mbytes, cudaMemcpyHostToDevice, st);      // the iterations of a solve)
        if (e != cudaSuccess) return (int)e;
        b.meas = d_meas;
    }
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

extern "C" int cpe_eval_host_uvg_f64(const cpe_model *model, const cpe_batch *bd, const double *h_q, const void *h_uv,
                                     const uint8_t *h_gate, double *d_q, void *d_uv, uint8_t *d_gate,
                                     const cpe_eval_out *out_dev, double *d_cost, double *h_cost, double *h_resid,
                                     void *stream) {
    if (!model || !bd || !h_q || !d_q || !out_dev || !d_cost || !h_cost || !out_dev->frame_cost) return CPE_E_NULL;
    if (bd->meas_layout != CPE_MEAS_F32X2G && bd->meas_layout != CPE_MEAS_F32X2B) return CPE_E_ARG;
    if (h_uv ? (!h_gate || !d_uv || !d_gate) : (!bd->meas || !bd->meas_gate)) return CPE_E_NULL;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t frames = (size_t)bd->S * bd->N;
    const size_t recs = frames * bd->C * 24 * bd->W;
    cudaError_t e = cudaMemcpyAsync(d_q, h_q, frames * 54 * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return (int)e;
    cpe_batch b = *bd;
    if (h_uv) {
        e = cudaMemcpyAsync(d_uv, h_uv, recs * 2 * sizeof(float), cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return (int)e;
        e = cudaMemcpyAsync(d_gate, h_gate, bd->meas_layout == CPE_MEAS_F32X2B ? (recs + 7) / 8 : recs, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return (int)e;
        b.meas = d_uv;
        b.meas_gate = d_gate;
    }
    int rc = cpe_eval_f64(model, &b, d_q, out_dev, stream);
    if (rc) return rc;
    rc = cpe_reduce_cost_f64(&b, out_dev->frame_cost, d_cost, stream);
    if (rc) return rc;
    e = cudaMemcpyAsync(h_cost, d_cost, (size_t)bd->S * 4 * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return (int)e;
    if (h_resid && out_dev->resid) {
        e = cudaMemcpyAsync(h_resid, out_dev->resid, recs * 2 * sizeof(double), cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return (int)e;
    }
    return 0;
}
