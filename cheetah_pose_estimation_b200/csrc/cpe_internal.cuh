// cpe_internal.cuh -- device helpers shared by the kernels (projection, robust loss).
#pragma once
#include <cuda_runtime.h>

#include "cpe.h"

namespace cpe {

struct Proj {
    double u, v;
    double du[3], dv[3];   // d(u,v)/d(world point)
};

// Camera projection of one world point.  cam = fx fy cx cy D0..D3 R[9] t[3] mult.
//   fisheye: reference acinoset_misc.py:1663-1679 (Kannala-Brandt, `r + 1e-12`, no skew)
//   pinhole: reference acinoset_misc.py:1682-1696 (radial r^2, r^4, r^6 with D[0..2])
template <bool DERIV>
__device__ __forceinline__ void project_point(const double *X, const double *cam, int model, Proj &o) {
    const double *R = cam + 8;
    const double x = X[0] * R[0] + X[1] * R[1] + X[2] * R[2] + cam[17];
    const double y = X[0] * R[3] + X[1] * R[4] + X[2] * R[5] + cam[18];
    const double z = X[0] * R[6] + X[1] * R[7] + X[2] * R[8] + cam[19];
    // two divisions per point: 1/z, then the reciprocals of r + 1e-12, 1 + r^2 and r from a single 1/(product)
    double iz, s, g = 0.0, a, b;
    if (model == CPE_CAM_FISHEYE) {
        const double izp = 1.0 / z;                         // a, b first (r depends on them)
        a = x * izp; b = y * izp; iz = izp;
        const double r2 = a * a + b * b;
        const double r = sqrt(r2);
        const double th = atan(r);
        const double t2 = th * th;
        const double poly = 1.0 + t2 * (cam[4] + t2 * (cam[5] + t2 * (cam[6] + t2 * cam[7])));
        const double thd = th * poly;
        const double re = r + 1e-12;
        if (DERIV) {
            const double rr = r > 0.0 ? r : 1.0;
            const double p1 = re * (1.0 + r2);
            const double inv = 1.0 / (p1 * rr);
            const double ire = inv * (1.0 + r2) * rr, i1r2 = inv * re * rr, ir = inv * p1;
            s = thd * ire;
            const double dpoly = 1.0 + t2 * (3.0 * cam[4] + t2 * (5.0 * cam[5] + t2 * (7.0 * cam[6] + t2 * 9.0 * cam[7])));
            const double ds_dr = (dpoly * i1r2 * re - thd) * (ire * ire);
            g = r > 0.0 ? ds_dr * ir : 0.0;
        } else {
            s = thd / re;
        }
    } else {
        iz = 1.0 / z;
        a = x * iz; b = y * iz;
        const double r2 = a * a + b * b;
        s = 1.0 + r2 * (cam[4] + r2 * (cam[5] + r2 * cam[6]));
        if (DERIV) g = 2.0 * cam[4] + r2 * (4.0 * cam[5] + 6.0 * r2 * cam[6]);
    }
    o.u = cam[0] * (a * s) + cam[2];
    o.v = cam[1] * (b * s) + cam[3];
    if (DERIV) {
        const double xa = s + a * a * g, xb = a * b * g, yb = s + b * b * g;
        // d(u)/d(Xc) and d(v)/d(Xc)
        const double ux = cam[0] * xa * iz, uy = cam[0] * xb * iz, uz = -cam[0] * (xa * a + xb * b) * iz;
        const double vx = cam[1] * xb * iz, vy = cam[1] * yb * iz, vz = -cam[1] * (xb * a + yb * b) * iz;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            o.du[j] = ux * R[j] + uy * R[3 + j] + uz * R[6 + j];
            o.dv[j] = vx * R[j] + vy * R[3 + j] + vz * R[6 + j];
        }
    }
}

// ---- detections in the three accepted layouts (cpe_batch.meas_layout) -----------------------------------------------
// Raw registers of one detection: fetched early (before the FK phases), unpacked where it is used.
union MeasRaw {
    float4 f;
    double4 d;
};

__device__ __forceinline__ void meas_fetch(const cpe_batch &b, size_t idx, MeasRaw &m) {
    if (b.meas_layout == CPE_MEAS_F64X4) {
        m.d = reinterpret_cast<const double4 *>(b.meas)[idx];
    } else if (b.meas_layout == CPE_MEAS_F32X4) {
        m.f = __ldg(reinterpret_cast<const float4 *>(b.meas) + idx);
    } else if (b.meas_layout == CPE_MEAS_F32X2G) {         // float2 {u, v} + gate byte
        const float2 uv = __ldg(reinterpret_cast<const float2 *>(b.meas) + idx);
        m.f.x = uv.x; m.f.y = uv.y; m.f.z = (float)__ldg(b.meas_gate + idx);
    } else if (b.meas_layout == CPE_MEAS_F32X2B) {         // float2 {u, v} + gate bit
        const float2 uv = __ldg(reinterpret_cast<const float2 *>(b.meas) + idx);
        m.f.x = uv.x; m.f.y = uv.y; m.f.z = (float)((__ldg(b.meas_gate + (idx >> 3)) >> (idx & 7)) & 1);
    } else {                                   // CPE_MEAS_F32X3: DLC-native (x, y, likelihood)
        const float *f = reinterpret_cast<const float *>(b.meas) + 3 * idx;
        m.f.x = __ldg(f); m.f.y = __ldg(f + 1); m.f.z = __ldg(f + 2);
    }
}

// (u, v, gate); for the DLC-native layout the gate is the reference's strict test likelihood > dlc_thresh
// (acinoset_misc.py:231).
__device__ __forceinline__ void meas_unpack(const cpe_batch &b, const MeasRaw &m, double &u, double &v, double &gate) {
    if (b.meas_layout == CPE_MEAS_F64X4) {
        u = m.d.x; v = m.d.y; gate = m.d.z;
    } else {
        u = (double)m.f.x; v = (double)m.f.y;
        if (b.meas_layout != CPE_MEAS_F32X3) {
            gate = (double)m.f.z;
        } else {
            // DLC-native record, gated here: a NaN coordinate or likelihood gets gate 0 and a finite stand-in, so that it
            // cannot poison the sums (the reference's own NaN guards are no-ops; the host packer of the
            // pre-gated layouts does the same: acinoset_misc.pack_measurements)
            const bool fin = (u == u) && (v == v);
            gate = (fin && (double)m.f.z > b.dlc_thresh) ? 1.0 : 0.0;
            if (!fin) { u = 0.0; v = 0.0; }
        }
    }
}

// ---- projection from CAMERA-frame coordinates, templated on the scalar type (K1) --------------------------------------
// cam (type T) = fx fy cx cy D0..D3 R[9] ...; returns pu = fx a s, pv = fy b s (principal point NOT added: the caller
// folds cx - u_meas in FP64) and d(pu,pv)/d(world point) through R.
template <typename T>
struct ProjT {
    T pu, pv;
    T du[3], dv[3];
};

__device__ __forceinline__ double r_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float r_sqrt(float x) { return sqrtf(x); }
__device__ __forceinline__ double r_atan(double x) { return atan(x); }
__device__ __forceinline__ float r_atan(float x) { return atanf(x); }

template <typename T, bool DERIV>
__device__ __forceinline__ void project_cam(T x, T y, T z, const T *cam, int model, ProjT<T> &o) {
    const T *R = cam + 8;
    const T one = T(1);
    T iz, s, g = T(0), a, b;
    if (model == CPE_CAM_FISHEYE) {
        iz = one / z;
        a = x * iz; b = y * iz;
        const T r2 = a * a + b * b;
        const T r = r_sqrt(r2);
        const T th = r_atan(r);
        const T t2 = th * th;
        const T poly = one + t2 * (cam[4] + t2 * (cam[5] + t2 * (cam[6] + t2 * cam[7])));
        const T thd = th * poly;
        const T re = r + T(1e-12);
        if (DERIV) {
            const T rr = r > T(0) ? r : one;
            const T p1 = re * (one + r2);
            const T inv = one / (p1 * rr);
            const T ire = inv * (one + r2) * rr, i1r2 = inv * re * rr, ir = inv * p1;
            s = thd * ire;
            const T dpoly = one + t2 * (T(3) * cam[4] + t2 * (T(5) * cam[5] + t2 * (T(7) * cam[6] + t2 * T(9) * cam[7])));
            const T ds_dr = (dpoly * i1r2 * re - thd) * (ire * ire);
            g = r > T(0) ? ds_dr * ir : T(0);
        } else {
            s = thd / re;
        }
    } else {
        iz = one / z;
        a = x * iz; b = y * iz;
        const T r2 = a * a + b * b;
        s = one + r2 * (cam[4] + r2 * (cam[5] + r2 * cam[6]));
        if (DERIV) g = T(2) * cam[4] + r2 * (T(4) * cam[5] + T(6) * r2 * cam[6]);
    }
    o.pu = cam[0] * (a * s);
    o.pv = cam[1] * (b * s);
    if (DERIV) {
        const T xa = s + a * a * g, xb = a * b * g, yb = s + b * b * g;
        const T ux = cam[0] * xa * iz, uy = cam[0] * xb * iz, uz = -cam[0] * (xa * a + xb * b) * iz;
        const T vx = cam[1] * xb * iz, vy = cam[1] * yb * iz, vz = -cam[1] * (xb * a + yb * b) * iz;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            o.du[j] = ux * R[j] + uy * R[3 + j] + uz * R[6 + j];
            o.dv[j] = vx * R[j] + vy * R[3 + j] + vz * R[6 + j];
        }
    }
}

// all three angle derivatives (dR/dphi, dR/dtheta, dR/dpsi) . v of one link: o[0..8] = {phi xyz, theta xyz, psi xyz}
template <typename T>
__device__ __forceinline__ void drot_all(const T *sc, T v0, T v1, T v2, T *o) {
    const T sf = sc[0], cf = sc[1], st = sc[2], ct = sc[3], sp = sc[4], cp = sc[5];
    const T stsf = st * sf, stcf = st * cf;
    const T m01 = cp * stsf - sp * cf, m02 = cp * stcf + sp * sf;      // R[0][1], R[0][2]
    const T m11 = sp * stsf + cp * cf, m12 = sp * stcf - cp * sf;      // R[1][1], R[1][2]
    // d/dphi: columns rotate (col1 -> col2, col2 -> -col1)
    o[0] = m02 * v1 - m01 * v2;
    o[1] = m12 * v1 - m11 * v2;
    o[2] = (ct * cf) * v1 - (ct * sf) * v2;
    // d/dtheta
    const T w = -st * v0 + ct * (sf * v1 + cf * v2);
    o[3] = cp * w;
    o[4] = sp * w;
    o[5] = -ct * v0 - st * (sf * v1 + cf * v2);
    // d/dpsi: rows rotate (row0 -> -row1, row1 -> row0)
    const T x0 = cp * ct * v0 + m01 * v1 + m02 * v2;                   // (R v)[0]
    const T x1 = sp * ct * v0 + m11 * v1 + m12 * v2;                   // (R v)[1]
    o[6] = -x1;
    o[7] = x0;
    o[8] = T(0);
}

// Redescending loss of the reference (acinoset_misc.py:2001-2015) at magnitude E = |err| and its derivative
// d rho / dE.  sigma(E - s) = 1 / (1 + exp(s) exp(-E)) needs a single exp per residual.
__device__ __forceinline__ void redescending(double E, double a, double b, double c, double ea, double eb, double ec,
                                             double &rho, double &drho) {
    const double t = exp(-E);
    const double sa = 1.0 / (1.0 + t * ea), sb = 1.0 / (1.0 + t * eb), sc = 1.0 / (1.0 + t * ec);
    const double da = sa * (1.0 - sa), db = sb * (1.0 - sb), dc = sc * (1.0 - sc);
    const double cb = c - b;
    const double p2 = a * E - 0.5 * a * a;
    const double q = (c - E) / cb;
    const double p3 = a * b - 0.5 * a * a + (0.5 * a * cb) * (1.0 - q * q);
    const double k4 = a * b - 0.5 * a * a + 0.5 * a * cb;
    rho = (1.0 - sa) * 0.5 * E * E + (sa - sb) * p2 + (sb - sc) * p3 + sc * k4;
    drho = -da * 0.5 * E * E + (1.0 - sa) * E + (da - db) * p2 + (sa - sb) * a + (db - dc) * p3 +
           (sb - sc) * a * (c - E) / cb + dc * k4;
}

// d rho / dE and the Gauss-Newton curvature weight max(rho'(E) / E, 0) of BOTH residuals of one detection (K2).  All
// eight divisions of the two redescending() calls and the two rho'/E share ONE reciprocal, 1 / (P_u P_v E_u E_v) with
// P = (1 + t ea)(1 + t eb)(1 + t ec) <= 2.2e14 (E = 0 is replaced by 1 in the product and its weight set to 0), and
// (c - E) / (c - b) is a multiplication by icb = 1 / (c - b).
__device__ __forceinline__ void redescending_pair_d(double Eu, double Ev, double a, double b, double c, double icb,
                                                    double ea, double eb, double ec, double &du, double &dv,
                                                    double &wu, double &wv) {
    const double tu = exp(-Eu), tv = exp(-Ev);
    const double dau = 1.0 + tu * ea, dbu = 1.0 + tu * eb, dcu = 1.0 + tu * ec;
    const double dav = 1.0 + tv * ea, dbv = 1.0 + tv * eb, dcv = 1.0 + tv * ec;
    const double dabu = dau * dbu, dabv = dav * dbv;
    const double Pu = dabu * dcu, Pv = dabv * dcv;
    const double Eus = Eu > 0.0 ? Eu : 1.0, Evs = Ev > 0.0 ? Ev : 1.0;
    const double PP = Pu * Pv, EE = Eus * Evs;
    const double inv = 1.0 / (PP * EE);
    const double iP = inv * EE, iE = inv * PP;
    const double iPu = iP * Pv, iPv = iP * Pu, iEu = iE * Evs, iEv = iE * Eus;
    const double cb = c - b;
    const double k2 = 0.5 * a * a, k3 = a * b - k2, hk = 0.5 * a * cb, k4 = k3 + hk;
    {
        const double sa = iPu * (dbu * dcu), sb = iPu * (dau * dcu), sc = iPu * dabu;
        const double da = sa * (1.0 - sa), db = sb * (1.0 - sb), dc = sc * (1.0 - sc);
        const double q = (c - Eu) * icb;
        du = -da * 0.5 * Eu * Eu + (1.0 - sa) * Eu + (da - db) * (a * Eu - k2) + (sa - sb) * a +
             (db - dc) * (k3 + hk * (1.0 - q * q)) + (sb - sc) * a * q + dc * k4;
        wu = Eu > 0.0 ? fmax(du * iEu, 0.0) : 0.0;
    }
    {
        const double sa = iPv * (dbv * dcv), sb = iPv * (dav * dcv), sc = iPv * dabv;
        const double da = sa * (1.0 - sa), db = sb * (1.0 - sb), dc = sc * (1.0 - sc);
        const double q = (c - Ev) * icb;
        dv = -da * 0.5 * Ev * Ev + (1.0 - sa) * Ev + (da - db) * (a * Ev - k2) + (sa - sb) * a +
             (db - dc) * (k3 + hk * (1.0 - q * q)) + (sb - sc) * a * q + dc * k4;
        wv = Ev > 0.0 ? fmax(dv * iEv, 0.0) : 0.0;
    }
}

// R = Rz(psi) Ry(theta) Rx(phi) from sc = {sin phi, cos phi, sin theta, cos theta, sin psi, cos psi}
// (reference shared/physical_education/utils.py:39-64, links.py:483)
__device__ __forceinline__ void rot_from_sc(const double *sc, double *M) {
    const double sf = sc[0], cf = sc[1], st = sc[2], ct = sc[3], sp = sc[4], cp = sc[5];
    M[0] = cp * ct; M[1] = cp * st * sf - sp * cf; M[2] = cp * st * cf + sp * sf;
    M[3] = sp * ct; M[4] = sp * st * sf + cp * cf; M[5] = sp * st * cf - cp * sf;
    M[6] = -st;     M[7] = ct * sf;                M[8] = ct * cf;
}

// (dR/dangle_a) . v  for a = 0 (phi), 1 (theta), 2 (psi)
__device__ __forceinline__ void drot_times(const double *sc, int a, double v0, double v1, double v2, double *o) {
    const double sf = sc[0], cf = sc[1], st = sc[2], ct = sc[3], sp = sc[4], cp = sc[5];
    if (a == 0) {
        o[0] = (cp * st * cf + sp * sf) * v1 + (-cp * st * sf + sp * cf) * v2;
        o[1] = (sp * st * cf - cp * sf) * v1 + (-sp * st * sf - cp * cf) * v2;
        o[2] = (ct * cf) * v1 + (-ct * sf) * v2;
    } else if (a == 1) {
        o[0] = (-cp * st) * v0 + (cp * ct * sf) * v1 + (cp * ct * cf) * v2;
        o[1] = (-sp * st) * v0 + (sp * ct * sf) * v1 + (sp * ct * cf) * v2;
        o[2] = (-ct) * v0 + (-st * sf) * v1 + (-st * cf) * v2;
    } else {
        o[0] = (-sp * ct) * v0 + (-sp * st * sf - cp * cf) * v1 + (-sp * st * cf + cp * sf) * v2;
        o[1] = (cp * ct) * v0 + (cp * st * sf - sp * cf) * v1 + (cp * st * cf + sp * sf) * v2;
        o[2] = 0.0;
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace cpe
