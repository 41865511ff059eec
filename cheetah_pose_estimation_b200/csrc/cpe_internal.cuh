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
