// cpe_assemble.cu -- K2: fused residual + forward-mode Jacobian + J^T J / J^T r accumulation in the tree-sparse
// pattern of the skeleton, augmented-Lagrangian rows, Marquardt damping and the per-frame Schur elimination of the
// 26 coordinates that carry no motion prior.  The Jacobian is never materialised.
//
// Reference arithmetic being replaced: what ASL evaluates inside `ipopt` (gradient / Hessian of the objective and
// constraints of acinoset_opt.py:601-618) for the terms of
//   acinoset_misc.py:278-288, 459-484, 2001-2015 (measurement), :639-677 (motion), system.py:574-586 (joints),
//   cheetah.py:306-352 (inequalities).
//
// Mapping (sm_100a, FP64): PERSISTENT CTAs, one per SM, of NTEAM "frame teams" (5 x 160 threads, or 4 x 192 when a frame has
// more than 160 (camera, marker) pairs).  The static gather tables (~24 KB, cpe_model.k2_tab) are copied into shared
// memory ONCE per CTA and shared by its teams; each team walks over frames (team-strided), synchronises on its own named
// barrier, and fetches the next frame's coordinates and detections into registers while it works on the current one.
// Per frame: front end as K1 (sincos -> R -> marker positions, chain derivative
// vectors U).  One thread per (camera, marker) projects, weights by the IRLS curvature max(rho'/E, 0) and leaves a
// 3x3 + 3 block; blocks are summed over cameras (fixed order) into A_l, b_l.  Every nonzero of the frame Hessian
// H = sum_l G_l^T A_l G_l (+ mu Jc^T Jc + active inequality rows) and of the gradient is then ONE gather over a
// static item list built on the host (skeleton.assemble_tables): no atomics, no index arithmetic in the kernel.
// The five elimination groups are reduced by one warp each, Gauss-Jordan in registers with warp shuffles
// (lane = tableau column), and the reduced 28x28 block S and ghat are gathered the same way.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "cpe.h"
#include "cpe_internal.cuh"

namespace cpe {

struct AsmParams {
    cpe_model m;
    cpe_batch b;
    const double *q;
    const double *damp;
    cpe_assemble_out ao;
    int TPF;        // C * 24
    int at_size;    // doubles of the aliased region (per-task blocks, then entries + group scratch)
    int tab_doubles;    // shared-memory doubles taken by the table blob (k2_off[K2_OFF_END] / 8, rounded up to 2)
    int team_doubles;   // shared-memory doubles of one team: cam[C*24] + A_AT + at_size
    double ea, eb, ec, icb;     // exp(loss_a), exp(loss_b), exp(loss_c), 1 / (loss_c - loss_b)
};

// per-frame shared layout (doubles), after cam[C*24]
// sin/cos and the link rotations are dead once the per-frame items are done, i.e. before the first write to the aliased
// region AT (per-pair blocks, then entries + elimination scratch): they live at its start.
// EQX = mu_eq * EQG sits right in front of EQR, and ACT (1 where an inequality row is violated) right in front of VIN, so
// that the second factor of a joint / inequality item is ONE indexed load from a base chosen once per entry (Hessian
// entry: EQX / ACT, gradient entry: EQR / VIN = the same index 156 / 32 doubles further)
constexpr int A_QO = 0, A_POSO = 56, A_GV = 128, A_BV = 798, A_EQG = 1746, A_EQX = 1902, A_EQR = 2058, A_ACT = 2084,
              A_VIN = 2116, A_MOG = 2148, A_DIAG = 2202, A_AT = 2256, A_SCO = A_AT, A_RO = A_AT + 104;
static_assert(A_EQR - A_EQX == 156 && A_VIN - A_ACT == 32, "gradient entries read 156 / 32 doubles behind the Hessian ones");
constexpr int K2_OFF_BLK = 41, K2_OFF_END = 45, K2_BLK_SLOTS = 8;      // must match skeleton.py
constexpr int G_U0 = 10, B_U0 = 216, B_G0 = 876, GRAD_Q = 63;       // must match skeleton.py
static_assert(CPE_K2_ENT_STORE >= 640 + 54, "ent record holds the entries and the diagonal");
constexpr int ENT_PAD = 640, DDI_SZ = CPE_MAX_GROUPS * 72, YS_SZ = CPE_MAX_GROUPS * 78;

__device__ __forceinline__ double motion_m0(int k, int nf) {
    const double c[4] = {-1.0, 3.0, -3.0, 1.0};
    double s = 0.0;
    for (int r = k - 3; r <= k; ++r) {
        if (r < 0 || r > nf - 4) continue;
        s += c[k - r] * c[k - r];
    }
    return s;
}

template <int TEAM_T>
__device__ __forceinline__ void team_sync(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(TEAM_T) : "memory");
}

template <int TEAM_T, int NTEAM>
__global__ void __launch_bounds__(TEAM_T *NTEAM, 1) assemble_kernel(const __grid_constant__ AsmParams p) {
    extern __shared__ __align__(16) double sm[];
    // ---- once per CTA: the gather tables ----------------------------------------------------------------------------------
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(p.m.k2_tab);
        uint4 *dst = reinterpret_cast<uint4 *>(sm);
        for (int i = threadIdx.x; i < p.m.k2_off[K2_OFF_END] / 16; i += TEAM_T * NTEAM) dst[i] = src[i];
    }
    __syncthreads();
    const unsigned char *tab = reinterpret_cast<const unsigned char *>(sm);
    const uint2 *t_meta = reinterpret_cast<const uint2 *>(tab + p.m.k2_off[0]);
    const uint32_t *t_smeta = reinterpret_cast<const uint32_t *>(tab + p.m.k2_off[1]);
    const int16_t *t_grp = reinterpret_cast<const int16_t *>(tab + p.m.k2_off[2]);
    const int team = threadIdx.x / TEAM_T, tid = threadIdx.x - team * TEAM_T;
    constexpr int nt = TEAM_T;
    const int lane = tid & 31, wid = tid >> 5;
    // blocks of 32 gather work items this warp owns (balanced on the host: skeleton._balanced_blocks)
    const unsigned char *h_blk = tab + p.m.k2_off[K2_OFF_BLK + (TEAM_T == 160 ? 0 : 2)] + wid * K2_BLK_SLOTS;
    const unsigned char *s_blk = tab + p.m.k2_off[K2_OFF_BLK + (TEAM_T == 160 ? 1 : 3)] + wid * K2_BLK_SLOTS;
    const int N = p.b.N, C = p.b.C, W = p.b.W, TPF = p.TPF;
    const int n_ineq = p.m.n_ineq, n_u = p.m.n_u, n_ent = p.m.n_ent;
    double *cam = sm + p.tab_doubles + (size_t)team * p.team_doubles;
    double *fr = cam + C * CPE_CAM_STRIDE;
    double *GV = fr + A_GV, *BV = fr + A_BV, *AT = fr + A_AT;
    double *ENT = AT, *DDI = AT + ENT_PAD, *YS = DDI + DDI_SZ;
    const bool mvalid = tid < TPF;
    const int mc = tid / 24, ml = tid - mc * 24;
    const long long total = (long long)p.b.S * N, stride = (long long)gridDim.x * NTEAM;
    long long f = (long long)blockIdx.x * NTEAM + team;
    // (sequence, frame) of f, advanced with the stride in 32-bit arithmetic (S * N < 2^31 is checked by the host)
    const int st_s = (int)(stride / N), st_k = (int)(stride - (long long)st_s * N);
    int s = (int)(f / N), k = (int)(f - (long long)s * N);
    s -= st_s; k -= st_k;
    if (k < 0) { k += N; --s; }
    // registers of the frame in flight: one coordinate and one detection per thread, fetched one frame ahead
    double q_pre = 0.0;
    MeasRaw m_pre;
    m_pre.d = make_double4(0.0, 0.0, 0.0, 0.0);
    if (f < total) {
        if (tid < 54) q_pre = p.q[f * 54 + tid];
        if (mvalid) meas_fetch(p.b, ((f * C + mc) * 24 + ml) * W, m_pre);
    }
    int s_cam = -1;                                      // sequence whose cameras sit in `cam`
    for (; f < total; f += stride) {
    s += st_s; k += st_k;
    if (k >= N) { k -= N; ++s; }
    const int nf = p.b.n_frames[s];
    const double q_cur = q_pre;
    MeasRaw mraw = m_pre;
    {
        const long long f2 = f + stride;                 // next frame of this team: loads in flight during this one
        if (f2 < total) {
            if (tid < 54) q_pre = p.q[f2 * 54 + tid];
            if (mvalid) meas_fetch(p.b, ((f2 * C + mc) * 24 + ml) * W, m_pre);
        }
    }
    if (k >= nf || (p.b.active && !p.b.active[s])) continue;
    const size_t fidx = (size_t)f;
    const int ncam_s = p.b.n_cams ? p.b.n_cams[s] : C;      // cameras beyond it are padding of a ragged batch (weight 0)
    const size_t midx0 = ((fidx * C + mc) * 24 + ml) * W;
    // A rejected LM step leaves q where it was: every entry of the frame Hessian and gradient is what the last full assembly
    // computed.  Only the damping changed, and it enters behind the gather (elimination pivots, diagonal of S).
    const bool redamp = p.ao.ent && p.ao.redamp && p.ao.redamp[s] < 0.0;
    if (redamp) {
        const double *src = p.ao.ent + fidx * CPE_K2_ENT_STORE;
        for (int i = tid; i < ENT_PAD; i += nt) ENT[i] = src[i];
        if (tid < 54) fr[A_DIAG + tid] = src[ENT_PAD + tid];
    } else {
    if (tid < 54) fr[A_QO + tid] = q_cur;
    if (s != s_cam) {                                    // consecutive frames of a team mostly share the sequence
        for (int i = tid; i < C * CPE_CAM_STRIDE; i += nt) cam[i] = p.b.cams[(size_t)s * C * CPE_CAM_STRIDE + i];
        s_cam = s;
    }
    if (tid < 10) GV[tid] = (tid == 0 || tid == 4 || tid == 8) ? 1.0 : 0.0;      // identity columns e_x, e_y, e_z
    team_sync<TEAM_T>(team);
    for (int a = tid; a < 51; a += nt) sincos(fr[A_QO + 3 + a], &fr[A_SCO + 2 * a], &fr[A_SCO + 2 * a + 1]);
    team_sync<TEAM_T>(team);
    for (int link = tid; link < 17; link += nt) rot_from_sc(fr + A_SCO + 6 * link, fr + A_RO + 9 * link);
    team_sync<TEAM_T>(team);

    // ---- camera-independent items: positions, U, joint rows, inequality rows, motion gradient --------------------
    const double h = p.b.h[s];
    const double mu_eq = p.b.mu_eq, mu_in = p.b.mu_in;
    {
        const double *cvec = p.b.chain_vec + (size_t)s * 24 * CPE_MAX_CHAIN * 3;
        const int32_t *chain_link = p.m.chain_link;
        const int n_items = 24 + n_u + 26 + n_ineq + 54;
        for (int it0 = tid; it0 < n_items; it0 += nt) {
            int it = it0;
            if (it < 24) {
                const int l = it;
                double x = fr[A_QO + 0], y = fr[A_QO + 1], z = fr[A_QO + 2];
#pragma unroll
                for (int j = 0; j < CPE_MAX_CHAIN; ++j) {
                    const int link = chain_link[l * CPE_MAX_CHAIN + j];
                    if (link >= 0) {
                        const double *Rl = fr + A_RO + link * 9;
                        const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                        const double v0 = v[0], v1 = v[1], v2 = v[2];
                        x += Rl[0] * v0 + Rl[1] * v1 + Rl[2] * v2;
                        y += Rl[3] * v0 + Rl[4] * v1 + Rl[5] * v2;
                        z += Rl[6] * v0 + Rl[7] * v1 + Rl[8] * v2;
                    }
                }
                fr[A_POSO + l * 3 + 0] = x; fr[A_POSO + l * 3 + 1] = y; fr[A_POSO + l * 3 + 2] = z;
                continue;
            }
            it -= 24;
            if (it < n_u) {
                const int code = p.m.u_item[it];
                const int l = code >> 8, j = (code >> 4) & 15, a = code & 15;
                const int link = chain_link[l * CPE_MAX_CHAIN + j];
                const double *v = cvec + (l * CPE_MAX_CHAIN + j) * 3;
                drot_times(fr + A_SCO + 6 * link, a, v[0], v[1], v[2], GV + G_U0 + 10 * (it / 3) + 3 * a);
                continue;
            }
            it -= n_u;
            if (it < 26) {
                const int e = it;
                const int4 pr4 = reinterpret_cast<const int4 *>(p.m.eq_pairs)[e];
                const int la = pr4.x, lb = pr4.y, ia = pr4.z, ib = pr4.w;
                const double *Ra = fr + A_RO + la * 9, *Rb = fr + A_RO + lb * 9;
                const double a0 = Ra[ia], a1 = Ra[3 + ia], a2 = Ra[6 + ia];
                const double b0 = Rb[ib], b1 = Rb[3 + ib], b2 = Rb[6 + ib];
                const double c = a0 * b0 + a1 * b1 + a2 * b2;
                const double lam = p.b.lam_eq ? p.b.lam_eq[fidx * 26 + e] : 0.0;
                fr[A_EQR + e] = mu_eq * (c + lam / mu_eq);
                // d c / d angles: (dR_a/dangle e_ia) . (R_b e_ib) and (R_a e_ia) . (dR_b/dangle e_ib)
                const double ea0 = ia == 0, ea1 = ia == 1, ea2 = ia == 2;
                const double eb0 = ib == 0, eb1 = ib == 1, eb2 = ib == 2;
#pragma unroll
                for (int x = 0; x < 3; ++x) {
                    double d[3];
                    drot_times(fr + A_SCO + 6 * la, x, ea0, ea1, ea2, d);
                    const double ga = d[0] * b0 + d[1] * b1 + d[2] * b2;
                    drot_times(fr + A_SCO + 6 * lb, x, eb0, eb1, eb2, d);
                    const double gb = a0 * d[0] + a1 * d[1] + a2 * d[2];
                    fr[A_EQG + e * 6 + x] = ga; fr[A_EQX + e * 6 + x] = mu_eq * ga;
                    fr[A_EQG + e * 6 + 3 + x] = gb; fr[A_EQX + e * 6 + 3 + x] = mu_eq * gb;
                }
                continue;
            }
            it -= 26;
            if (it < n_ineq) {
                const int j = it;
                const int ip = p.m.ineq_idx[2 * j], im = p.m.ineq_idx[2 * j + 1];
                const double lo = p.m.ineq_bounds[2 * j], up = p.m.ineq_bounds[2 * j + 1];
                const double d = fr[A_QO + ip] - (im >= 0 ? fr[A_QO + im] : 0.0);
                const double lam = p.b.lam_in ? p.b.lam_in[fidx * n_ineq + j] : 0.0;
                const double t = d + lam / mu_in;
                const double v = t - fmin(fmax(t, lo), up);
                fr[A_VIN + j] = v;
                fr[A_ACT + j] = v != 0.0 ? 1.0 : 0.0;
                continue;
            }
            it -= n_ineq;
            {
                const int c = it;
                const double w = p.m.wq[c];
                double val = 0.0;
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
                    val = 2.0 * w / (h * h) * acc;
                }
                fr[A_MOG + c] = val;
            }
        }
    }
    team_sync<TEAM_T>(team);

    // ---- one thread per (camera, marker): projection + IRLS-weighted 3x3 block and gradient 3-vector -------------
    if (mvalid) {
        const double la_ = p.b.loss_a, lb_ = p.b.loss_b, lc_ = p.b.loss_c;
        const double *cm = cam + mc * CPE_CAM_STRIDE;
        Proj pr;
        double pt[3] = {fr[A_POSO + ml * 3], fr[A_POSO + ml * 3 + 1], fr[A_POSO + ml * 3 + 2]};
        if (p.b.shutter_tau) {
            // shutter delay: camera mc sees the frame displaced by dq tau + ddq tau^2 (acinoset_misc.py:280-286)
            const double tau = p.b.shutter_tau[(size_t)s * C + mc];
            const double *vb = p.b.base_dq + fidx * 3, *ab = p.b.base_ddq + fidx * 3;
            pt[0] += vb[0] * tau + ab[0] * (tau * tau);
            pt[1] += vb[1] * tau + ab[1] * (tau * tau);
            pt[2] += vb[2] * tau + ab[2] * (tau * tau);
        }
        project_point<true>(pt, cm, p.b.cam_model, pr);
        const double cmult = cm[20];
        double hu = 0.0, hv = 0.0, gu = 0.0, gv = 0.0;
        for (int w = 0; w < W; ++w) {
            double um, vm, gate;
            if (w > 0) meas_fetch(p.b, midx0 + w, mraw);
            meas_unpack(p.b, mraw, um, vm, gate);
            const bool sq = p.b.loss_mode == CPE_LOSS_SQUARE;
            const double wgt = (mc < ncam_s ? gate : 0.0) * p.m.meas_w[w * 24 + ml] * (sq ? 1.0 : cmult);
            const double eu = wgt * (pr.u - um), ev = wgt * (pr.v - vm);
            const double Eu = fabs(eu), Ev = fabs(ev);
            double du_, dv_, wu_, wv_;
            if (sq) {
                du_ = 2.0 * Eu; dv_ = 2.0 * Ev;
                wu_ = Eu > 0.0 ? 2.0 : 0.0; wv_ = Ev > 0.0 ? 2.0 : 0.0;
            } else {
                redescending_pair_d(Eu, Ev, la_, lb_, lc_, p.icb, p.ea, p.eb, p.ec, du_, dv_, wu_, wv_);
            }
            gu += du_ * (eu > 0.0 ? 1.0 : (eu < 0.0 ? -1.0 : 0.0)) * wgt;
            gv += dv_ * (ev > 0.0 ? 1.0 : (ev < 0.0 ? -1.0 : 0.0)) * wgt;
            hu += wu_ * wgt * wgt;
            hv += wv_ * wgt * wgt;
        }
        double *ab = AT + tid * 9;
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
    team_sync<TEAM_T>(team);
    if (p.ao.cam_blocks) {
        // per camera: sum over markers of the point-space curvature (6) and gradient (3) -- what the shutter-delay
        // unknowns see (d slack / d tau_c and the cross-frame base terms are d(u,v)/d(point) times a 3-vector)
        for (int i = tid; i < C * 9; i += nt) {
            const int c = i / 9, comp = i - c * 9;
            double v = 0.0;
            for (int l = 0; l < 24; ++l) v += AT[(c * 24 + l) * 9 + comp];
            p.ao.cam_blocks[(fidx * C + c) * 9 + comp] = v;
        }
    }
    // ---- A_l = sum over cameras (fixed order), expanded to full 3x3 columns in B space; b_l ------------------------
    for (int i = tid; i < 216; i += nt) {
        const int l = i / 9, comp = i - l * 9;
        double v = 0.0;
        for (int c = 0; c < C; ++c) v += AT[(c * 24 + l) * 9 + comp];
        if (comp >= 6) {
            BV[B_G0 + 3 * l + comp - 6] = v;
        } else {
            const int ri = comp < 3 ? 0 : (comp < 5 ? 1 : 2);
            const int ci = comp < 3 ? comp : (comp < 5 ? comp - 2 : 2);
            BV[9 * l + 3 * ci + ri] = v;
            BV[9 * l + 3 * ri + ci] = v;
        }
    }
    team_sync<TEAM_T>(team);
    // ---- B = A_l U for every chain column ------------------------------------------------------------------------------
    for (int it = tid; it < n_u; it += nt) {
        const int l = p.m.u_item[it] >> 8;
        const int o = 10 * (it / 3) + 3 * (it % 3);
        const double *Al = BV + 9 * l;
        const double u0 = GV[G_U0 + o], u1 = GV[G_U0 + o + 1], u2 = GV[G_U0 + o + 2];
        BV[B_U0 + o + 0] = Al[0] * u0 + Al[3] * u1 + Al[6] * u2;
        BV[B_U0 + o + 1] = Al[1] * u0 + Al[4] * u1 + Al[7] * u2;
        BV[B_U0 + o + 2] = Al[2] * u0 + Al[5] * u1 + Al[8] * u2;
    }
    team_sync<TEAM_T>(team);
    // ---- every Hessian / gradient nonzero = one gather over its static item list -------------------------------------
    // slot-major tables in shared memory (skeleton.k2_tables): lane <-> consecutive eo, no bank conflicts, no address
    // that depends on a loaded index
    {
        for (int bi = 0; bi < K2_BLK_SLOTS; ++bi) {
            const int blk = h_blk[bi];
            if (blk == 255) break;
            const int eo = blk * 32 + lane;
            if (eo >= n_ent) continue;
            const uint2 m = t_meta[eo];
            const int e = m.x & 1023, pp = (m.x >> 10) & 63, qq = (m.x >> 16) & 63;
            const int na = m.y & 255, ne = (m.y >> 8) & 255, ni = m.y >> 16;
            const bool grad = qq == GRAD_Q;
            double acc = 0.0;
#pragma unroll 4
            for (int i = 0; i < na; ++i) {
                const int it = reinterpret_cast<const int32_t *>(tab + p.m.k2_off[3 + i])[eo];
                const double *g = GV + (it >> 16), *b = BV + (it & 0xFFFF);
                acc += g[0] * b[0] + g[1] * b[1] + g[2] * b[2];
            }
            double acce = 0.0;
            const double *ex = fr + (grad ? A_EQR : A_EQX);        // second factor: mu_eq * eqg[j], or the row's residual
            for (int i = 0; i < ne; ++i) {
                const int it = reinterpret_cast<const uint16_t *>(tab + p.m.k2_off[27 + i])[eo];
                acce += fr[A_EQG + (it >> 8)] * ex[it & 255];
            }
            double acci = 0.0;
            const double *vx = fr + (grad ? A_VIN : A_ACT);
            for (int i = 0; i < ni; ++i) {
                const int it = (tab + p.m.k2_off[32 + i])[eo];
                const double val = vx[it >> 1];
                acci += (it & 1) ? -val : val;
            }
            acc += acce + mu_in * acci;
            if (grad) {
                acc += fr[A_MOG + pp];
                if (p.ao.g) p.ao.g[fidx * 54 + pp] = acc;
            } else if (p.ao.Hraw) {
                double *o = p.ao.Hraw + fidx * 54 * 54;
                o[pp * 54 + qq] = acc;
                o[qq * 54 + pp] = acc;
            }
            ENT[e] = acc;
        }
    }
    team_sync<TEAM_T>(team);
    const double h4 = h * h * h * h;
    if (tid < 54) {
        const double d = ENT[p.m.diag_ent[tid]] + motion_m0(k, nf) * 2.0 * p.m.wq[tid] / h4;
        fr[A_DIAG + tid] = d;
        if (p.ao.diagH) p.ao.diagH[fidx * 54 + tid] = d;
        if (p.ao.ent) p.ao.ent[fidx * CPE_K2_ENT_STORE + ENT_PAD + tid] = d;
    }
    if (p.ao.ent) {
        double *dst = p.ao.ent + fidx * CPE_K2_ENT_STORE;
        for (int i = tid; i < ENT_PAD; i += nt) dst[i] = ENT[i];
    }
    }   // full assembly
    if (!p.ao.S) { team_sync<TEAM_T>(team); continue; }
    team_sync<TEAM_T>(team);

    // ---- per-frame elimination of the dependent coordinates: one warp per group, tableau column per lane ----------
    const double damp = p.damp[s];
    constexpr int GD = CPE_GRP_MAX_DEP, GN = CPE_GRP_MAX_NB, TW = GD + GN + 1;
    constexpr int nw = nt >> 5;
    const int n_groups = p.m.n_groups;
    for (int g = wid; g < n_groups; g += nw) {
        double t[GD];
        const int16_t *src = t_grp + g * GD * TW;
#pragma unroll
        for (int i = 0; i < GD; ++i) {
            double v = 0.0;
            if (lane < TW) {
                const int si = src[i * TW + lane];
                v = si == -2 ? 1.0 : (si < 0 ? 0.0 : ENT[si]);
                if (lane == i && si >= 0) v += damp * fr[A_DIAG + p.m.grp_dep[g * GD + i]];
            }
            t[i] = v;
            if (lane >= GD && lane < GD + GN) DDI[(g * GD + i) * GN + lane - GD] = v;
        }
#pragma unroll
        for (int pc = 0; pc < GD; ++pc) {
            const double piv = __shfl_sync(0xffffffffu, t[pc], pc);
            const double prow = t[pc] * (1.0 / piv);
#pragma unroll
            for (int i = 0; i < GD; ++i) {
                if (i == pc) continue;
                const double f = __shfl_sync(0xffffffffu, t[i], pc);
                t[i] -= f * prow;
            }
            t[pc] = prow;
        }
        // lanes GD..GD+GN-1 hold Y = Hdd^-1 Hdi (column lane-GD), lane GD+GN holds y = Hdd^-1 g_d
        if (lane >= GD && lane < TW) {
            const int j = lane - GD;
#pragma unroll
            for (int i = 0; i < GD; ++i) {
                YS[(g * GD + i) * (GN + 1) + j] = t[i];
                if (j < GN) p.ao.Y[((fidx * n_groups + g) * GD + i) * GN + j] = t[i];
                else p.ao.y[(fidx * n_groups + g) * GD + i] = t[i];
            }
        }
    }
    team_sync<TEAM_T>(team);
    // ---- reduced block S = H_ii + damping - sum_g Hdi^T Y and ghat = g_i - sum_g Hdi^T y (static gather) ----------------
    double *So = p.ao.S + fidx * 784;
    const int n_swork = p.m.k2_off[K2_OFF_END + 1];   // lower triangle of S + the reduced gradient
    for (int bi = 0; bi < K2_BLK_SLOTS; ++bi) {
        const int blk = s_blk[bi];
        if (blk == 255) break;
        const int eo = blk * 32 + lane;
        if (eo >= n_swork) continue;
        const uint32_t m = t_smeta[eo];
        const int r = m & 31, c = (m >> 5) & 31, ns = (m >> 10) & 7, si = m >> 13;
        double v = si != 1023 ? ENT[si] : 0.0;
        if (r == c) v += damp * fr[A_DIAG + p.m.ind_idx[r]];
        for (int i = 0; i < ns; ++i) {
            const int it = reinterpret_cast<const uint16_t *>(tab + p.m.k2_off[36 + i])[eo];
            const int g = it >> 8, a = (it >> 4) & 15, b = it & 15;
            const double *D = DDI + g * GD * GN + a;
            const double *Yc = YS + g * GD * (GN + 1) + (b == 15 ? GN : b);
            double sacc = 0.0;
#pragma unroll
            for (int ii = 0; ii < GD; ++ii) sacc += D[ii * GN] * Yc[ii * (GN + 1)];
            v -= sacc;
        }
        if (c == 28) {
            p.ao.ghat[fidx * 28 + r] = v;
        } else {
            So[r * 28 + c] = v;
            if (r != c) So[c * 28 + r] = v;               // S is symmetric: one evaluation, two stores
        }
    }
    team_sync<TEAM_T>(team);       // the next frame overwrites this team's scratch
    }   // frames of this team
}

template <int TEAM_T, int NTEAM>
static int launch_asm(const AsmParams &p, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(assemble_kernel<TEAM_T, NTEAM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long frames = (long long)p.b.S * p.b.N;
    long long grid = (frames + NTEAM - 1) / NTEAM;
    if (grid > sms) grid = sms;                      // persistent: one CTA per SM, its teams stride over the frames
    assemble_kernel<TEAM_T, NTEAM><<<(unsigned)grid, TEAM_T * NTEAM, smem, st>>>(p);
    return (int)cudaGetLastError();
}

}  // namespace cpe

extern "C" int cpe_assemble_f64(const cpe_model *model, const cpe_batch *b, const double *q, const double *damp,
                                const cpe_assemble_out *out, void *stream) {
    using namespace cpe;
    if (!model || !b || !q || !out) return CPE_E_NULL;
    if (b->S <= 0 || b->N <= 0 || b->C <= 0 || b->C > CPE_MAX_CAMS || (b->W != 1 && b->W != 3)) return CPE_E_DIM;
    if ((long long)b->S * b->N > 2147483647LL || model->n_ineq > CPE_MAX_INEQ || model->n_groups > CPE_MAX_GROUPS)
        return CPE_E_DIM;
    if (model->n_u != 198 || model->n_ent <= 0 || model->n_ent > ENT_PAD) return CPE_E_DIM;
    if (!model->k2_tab || !model->diag_ent || !model->grp_dep || !model->ind_idx) return CPE_E_NULL;
    if (out->S && (!damp || !out->ghat || !out->Y || !out->y)) return CPE_E_NULL;
    if (out->redamp && (!out->ent || !out->S)) return CPE_E_NULL;
    if ((b->shutter_tau || b->base_dq || b->base_ddq) && !(b->shutter_tau && b->base_dq && b->base_ddq)) return CPE_E_NULL;
    AsmParams p;
    p.m = *model; p.b = *b; p.q = q; p.damp = damp; p.ao = *out;
    p.TPF = b->C * 24;
    p.ea = exp(b->loss_a); p.eb = exp(b->loss_b); p.ec = exp(b->loss_c);
    p.icb = 1.0 / (b->loss_c - b->loss_b);
    const int need = ENT_PAD + DDI_SZ + YS_SZ;
    p.at_size = p.TPF * 9 > need ? p.TPF * 9 : need;
    p.team_doubles = (b->C * CPE_CAM_STRIDE + A_AT + p.at_size + 1) & ~1;
    p.tab_doubles = ((model->k2_off[K2_OFF_END] + 15) / 16) * 2;
    if (model->k2_off[K2_OFF_END] <= 0 || model->k2_off[K2_OFF_END] % 16) return CPE_E_DIM;
    if (p.TPF <= 160) {
        // five teams: 25 warps at 72 registers.  Measured at 1776 x 500 (profiles/r2i_k2_teams.txt): 6 teams / 64 registers
        // 31.2 ms, 5 / 72: 29.3 ms, 4 / 96: 30.0 ms, 3 / 128: 32.9 ms
        const size_t smem = (size_t)(p.tab_doubles + 5 * p.team_doubles) * sizeof(double);
        return launch_asm<160, 5>(p, smem, (cudaStream_t)stream);
    }
    const size_t smem = (size_t)(p.tab_doubles + 4 * p.team_doubles) * sizeof(double);
    return launch_asm<192, 4>(p, smem, (cudaStream_t)stream);
}
