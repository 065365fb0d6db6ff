// Device-side building blocks of the explicit step (3D tets, FP64).
//
// Every function restates one piece of reference arithmetic in the SAME operation order
// (compiled with -fmad=false the results are bit-identical to the reference's g++ -O2
// x86-64 build wherever no libm transcendental is involved).  Citations are
// /root/reference paths.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <float.h>
#include "../../include/des_gpu.h"

#define DES_MAX_MEMO (4 * DES_MAX_MAT)

// Scalars and per-material tables the kernels read, passed BY VALUE as a
// __grid_constant__ kernel parameter (constant-bank loads, no pointer chase).
struct DevParams {
    double gravity, damping_factor, surf_diff, surf_base_level;
    double winkler_delta_rho, elastic_foundation_constant, sea_water_density;
    double surface_temperature, zlength;
    double pseudo_speed;            // max_vbc_val * inertial_scaling (geometry.cxx:1754)
    double visc_min, visc_max, tension_max, therm_diff_max;
    double vbc_values[DES_NBDRY];
    double vbc_val_lateral[4];
    double vbc_val_z1_loading_period;
    double bnormals[DES_NBDRY][3];
    int vbc_types[DES_NBDRY];
    int damping_option, is_quasi_static, has_thermal_diffusion, is_using_mixed_stress;
    int surface_process_option, ref_pressure_option, has_winkler_foundation, has_water_loading;
    int has_elastic_foundation, rheol_type, nmat, mattype_ref;
    unsigned sbc_mask;              // bcflag bits of boundaries with a live stress bc
    unsigned nbc_mask;              // bcflag bits of boundaries with a prescribed traction (stress_bc_types)
    int stress_bc_types[6];         // 1 / 2 / 3: traction along x / y / z (bc.cxx:862-873)
    double stress_bc_values[6];
    int nmemo;
    double rho0[DES_MAX_MAT], alpha[DES_MAX_MAT];
    double bulk_modulus[DES_MAX_MAT], shear_modulus[DES_MAX_MAT];
    // Chen & Morgan creep law, material-only terms (matprops.cxx:236-252), built on the
    // host with the host libm so they carry the reference's bits.
    double visc_pow_edot[DES_MAX_MAT], visc_coef_term[DES_MAX_MAT], visc_nR[DES_MAX_MAT];
    double visc_activation_energy[DES_MAX_MAT], visc_activation_volume[DES_MAX_MAT];
    double heat_capacity[DES_MAX_MAT], therm_cond[DES_MAX_MAT];
    double pls0[DES_MAX_MAT], pls1[DES_MAX_MAT];
    double cohesion0[DES_MAX_MAT], cohesion1[DES_MAX_MAT];
    double friction_angle0[DES_MAX_MAT], friction_angle1[DES_MAX_MAT];
    double dilation_angle0[DES_MAX_MAT], dilation_angle1[DES_MAX_MAT];
    double porosity[DES_MAX_MAT];
    // Host-libm memo of sin/tan at the configured friction/dilation angles: an element
    // whose weakened angle equals one of them exactly (constant-friction materials: every
    // element, every step) gets the reference's own libm bits for sin/tan
    // (matprops.cxx:589-606) instead of CUDA's, which may differ in the last ulp.
    double memo_angle[DES_MAX_MEMO], memo_sin[DES_MAX_MEMO], memo_tan[DES_MAX_MEMO];
};

// 256-bit global loads/stores (LDG.E.ENL2.256 / STG.E.ENL2.256 on sm_100a): a 32-byte node or
// contribution record is ONE L1TEX wavefront per lane instead of two 128-bit halves.
__device__ __forceinline__ double4 ld256_nc(const void *p)      // read-only in this kernel
{
    double4 r;
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ double ld64_v(const double *p)       // ordered among the volatile loads
{
    double r;
    asm volatile("ld.global.f64 %0, [%1];" : "=d"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ unsigned ld8_v(const unsigned char *p)
{
    unsigned r;
    asm volatile("ld.global.u8 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
// L2 prefetch hint (no register, no scoreboard entry): used for the rows a later block streams
__device__ __forceinline__ void pf_l2(const void *p)
{
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}
#ifndef DES_PF_DEFAULT
#define DES_PF_DEFAULT 65536
#endif
__device__ __forceinline__ double4 ld256(const void *p)
{
    double4 r;
    asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st256(void *p, const double4 &r)
{
    asm volatile("st.global.v4.f64 [%4], {%0,%1,%2,%3};"
                 :: "d"(r.x), "d"(r.y), "d"(r.z), "d"(r.w), "l"(p) : "memory");
}

// constants.hpp:75-79 (3D)
__constant__ int c_node_of_facet[4][3] = {{1,2,3},{0,3,2},{0,1,3},{0,2,1}};

#define BOUNDX0 1u
#define BOUNDX1 2u
#define BOUNDY0 4u
#define BOUNDY1 8u
#define BOUNDZ0 16u
#define BOUNDZ1 32u
#define BOUND_ANY 1023u

// ---------------------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------------------

// tetrahedron_volume, geometry.cxx:36-56
__device__ __forceinline__ double tet_volume(const double4 &d0, const double4 &d1,
                                             const double4 &d2, const double4 &d3)
{
    double x01 = d0.x - d1.x, x12 = d1.x - d2.x, x23 = d2.x - d3.x;
    double y01 = d0.y - d1.y, y12 = d1.y - d2.y, y23 = d2.y - d3.y;
    double z01 = d0.z - d1.z, z12 = d1.z - d2.z, z23 = d2.z - d3.z;
    return (x01*(y23*z12 - y12*z23) +
            x12*(y01*z23 - y23*z01) +
            x23*(y12*z01 - y01*z12)) / 6;
}

// get_local_shape_fn (3D), fields.cxx:11-38
__device__ __forceinline__ void shape_fn(const double4 &d0, const double4 &d1,
                                         const double4 &d2, const double4 &d3, double vol,
                                         double sx[4], double sy[4], double sz[4])
{
    double iv = 1.0 / (6.0 * vol);
    double x01 = d0.x - d1.x, x02 = d0.x - d2.x, x03 = d0.x - d3.x;
    double x12 = d1.x - d2.x, x13 = d1.x - d3.x, x23 = d2.x - d3.x;
    double y01 = d0.y - d1.y, y02 = d0.y - d2.y, y03 = d0.y - d3.y;
    double y12 = d1.y - d2.y, y13 = d1.y - d3.y, y23 = d2.y - d3.y;
    double z01 = d0.z - d1.z, z02 = d0.z - d2.z, z03 = d0.z - d3.z;
    double z12 = d1.z - d2.z, z13 = d1.z - d3.z, z23 = d2.z - d3.z;

    sx[0] = iv * (y13*z12 - y12*z13);
    sx[1] = iv * (y02*z23 - y23*z02);
    sx[2] = iv * (y13*z03 - y03*z13);
    sx[3] = iv * (y01*z02 - y02*z01);

    sy[0] = iv * (z13*x12 - z12*x13);
    sy[1] = iv * (z02*x23 - z23*x02);
    sy[2] = iv * (z13*x03 - z03*x13);
    sy[3] = iv * (z01*x02 - z02*x01);

    sz[0] = iv * (x13*y12 - x12*y13);
    sz[1] = iv * (x02*y23 - x23*y02);
    sz[2] = iv * (x13*y03 - x03*y13);
    sz[3] = iv * (x01*y02 - x02*y01);
}

// triangle_area (3D), geometry.cxx:76-106
__device__ __forceinline__ double tri_area(const double4 &a, const double4 &b, const double4 &c)
{
    double ab0 = b.x - a.x, ab1 = b.y - a.y, ac0 = c.x - a.x, ac1 = c.y - a.y;
    double ab2 = b.z - a.z, ac2 = c.z - a.z;
    double d0 = ab1*ac2 - ab2*ac1;
    double d1 = ab2*ac0 - ab0*ac2;
    double d2 = ab0*ac1 - ab1*ac0;
    return sqrt(d0*d0 + d1*d1 + d2*d2) / 2;
}

// triangle_area2d, geometry.cxx:58-73 (what compute_area_facet uses in 3D, :108-120)
__device__ __forceinline__ double tri_area2d(const double4 &a, const double4 &b, const double4 &c)
{
    double ab0 = b.x - a.x, ab1 = b.y - a.y, ac0 = c.x - a.x, ac1 = c.y - a.y;
    return fabs(ab0*ac1 - ab1*ac0) / 2;
}

// ---------------------------------------------------------------------------------------
// tensors, utils.hpp:210-245
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double trace3(const double *s) { return s[0] + s[1] + s[2]; }

__device__ __forceinline__ double second_invariant2(const double *t)
{
    double a = (t[0] + t[1] + t[2]) / 3;
    return ( 0.5 * ((t[0]-a)*(t[0]-a) + (t[1]-a)*(t[1]-a) + (t[2]-a)*(t[2]-a))
             + t[3]*t[3] + t[4]*t[4] + t[5]*t[5] );
}

// ---------------------------------------------------------------------------------------
// material properties, matprops.cxx
// ---------------------------------------------------------------------------------------

// arithmetic_mean, matprops.cxx:115-130
__device__ __forceinline__ double arith_mean(const double *s, const int *cnt, int nmat)
{
    if (nmat == 1) return s[0];
    double result = 0;
    int m = 0;
    for (int i = 0; i < nmat; ++i) {
        if (cnt[i] == 0) continue;
        result += cnt[i] * s[i];
        m += cnt[i];
    }
    return result / m;
}

// harmonic_mean, matprops.cxx:133-150
__device__ __forceinline__ double harm_mean(const double *s, const int *cnt, int nmat)
{
    if (nmat == 1) return s[0];
    double result = 0;
    int m = 0;
    for (int i = 0; i < nmat; ++i) {
        if (cnt[i] == 0) continue;
        result += cnt[i] / s[i];
        m += cnt[i];
    }
    return m / result;
}

// Marker counts of one element, in registers.
struct Markers {
    int cnt[DES_MAX_MAT];
};

__device__ __forceinline__ void load_markers(const int *__restrict__ elemmarkers, int e, int nmat,
                                             Markers &mk)
{
    if (nmat == 1) { mk.cnt[0] = elemmarkers[e]; return; }
    const int *p = elemmarkers + size_t(e) * nmat;
    #pragma unroll 1
    for (int m = 0; m < nmat; ++m) mk.cnt[m] = p[m];
}

// MatProps::rho, matprops.cxx:642-664 (does NOT skip absent materials)
__device__ __forceinline__ double mat_rho(const DevParams &P, int nmat, const Markers &mk,
                                          double T0, double T1, double T2, double T3)
{
    const double celsius0 = 273;
    double T = 0;
    T += T0; T += T1; T += T2; T += T3;
    T /= 4;
    double TinCelsius = T - celsius0;
    double result = 0;
    int n = 0;
    for (int m = 0; m < nmat; ++m) {
        result += P.rho0[m] * (1 - P.alpha[m] * TinCelsius) * mk.cnt[m];
        n += mk.cnt[m];
    }
    return result / n;
}

// MatProps::visc, matprops.cxx:333-377.  `s` is the PRE-update stress, `edot` the
// anti-locked strain rate of this step (rheology.cxx:786-799 runs before :877).
__device__ __forceinline__ double mat_visc(const DevParams &P, int nmat, const Markers &mk,
                                           double T0, double T1, double T2, double T3,
                                           const double *s, const double *edot_t)
{
    const double min_strain_rate = 1e-30;
    double T = 0;
    T += T0; T += T1; T += T2; T += T3;
    T /= 4;
    double s0 = trace3(s) / 3;
    double edot = sqrt(second_invariant2(edot_t));
    edot = fmax(edot, min_strain_rate);
    double result = 0;
    int n = 0;
    for (int m = 0; m < nmat; ++m) {
        const int marker_count = mk.cnt[m];
        if (marker_count == 0) continue;
        double visc0 = 0.25 * pow(edot, P.visc_pow_edot[m]) * P.visc_coef_term[m]
            * exp((P.visc_activation_energy[m] + P.visc_activation_volume[m] * s0)
                  / (P.visc_nR[m] * T)) * 1e6;
        result += marker_count / visc0;
        n += marker_count;
    }
    double visc = n / result;
    visc = fmin(fmax(visc, P.visc_min), P.visc_max);
    return visc;
}

__device__ __forceinline__ double memo_sin(const DevParams &P, double angle_deg)
{
    for (int i = 0; i < P.nmemo; ++i)
        if (angle_deg == P.memo_angle[i]) return P.memo_sin[i];
    return sin(angle_deg * (M_PI / 180));
}

__device__ __forceinline__ double memo_tan(const DevParams &P, double angle_deg)
{
    for (int i = 0; i < P.nmemo; ++i)
        if (angle_deg == P.memo_angle[i]) return P.memo_tan[i];
    return tan(angle_deg * (M_PI / 180));
}

// MatProps::plastic_weakening + plastic_props, matprops.cxx:380-418,589-606
__device__ __forceinline__ void plastic_props(const DevParams &P, int nmat, const Markers &mk, double pls,
                                              double &amc, double &anphi, double &anpsi,
                                              double &hardn, double &ten_max)
{
    double c = 0, f = 0, d = 0, h = 0;
    int n = 0;
    for (int m = 0; m < nmat; ++m) {
        int k = mk.cnt[m];
        if (k == 0) continue;
        n += k;
        if (pls < P.pls0[m]) {
            c += P.cohesion0[m] * k;
            f += P.friction_angle0[m] * k;
            d += P.dilation_angle0[m] * k;
            h += 0;
        }
        else if (pls < P.pls1[m]) {
            double p = (pls - P.pls0[m]) / (P.pls1[m] - P.pls0[m]);
            c += (P.cohesion0[m] + p * (P.cohesion1[m] - P.cohesion0[m])) * k;
            f += (P.friction_angle0[m] + p * (P.friction_angle1[m] - P.friction_angle0[m])) * k;
            d += (P.dilation_angle0[m] + p * (P.dilation_angle1[m] - P.dilation_angle0[m])) * k;
            h += (P.cohesion1[m] - P.cohesion0[m]) / (P.pls1[m] - P.pls0[m]) * k;
        }
        else {
            c += P.cohesion1[m] * k;
            f += P.friction_angle1[m] * k;
            d += P.dilation_angle1[m] * k;
            h += 0;
        }
    }
    double cohesion = c / n, phi = f / n, psi = d / n;
    hardn = h / n;

    double sphi = memo_sin(P, phi);
    double spsi = memo_sin(P, psi);
    anphi = (1 + sphi) / (1 - sphi);
    anpsi = (1 + spsi) / (1 - spsi);
    amc = 2 * cohesion * sqrt(anphi);
    ten_max = (phi == 0) ? P.tension_max : fmin(P.tension_max, cohesion / memo_tan(P, phi));
}

// PREM tables, matprops.cxx:12-100
__device__ __forceinline__ double prem_pressure(double depth, bool modified)
{
    const double ref_depth[46] = { 0e3, 3e3, 15e3, 24.4e3, 40e3,
        60e3, 80e3, 115e3, 150e3, 185e3, 220e3, 265e3, 310e3, 355e3, 400e3,
        450e3, 500e3, 550e3, 600e3, 635e3, 670e3, 721e3, 771e3, 871e3, 971e3,
        1071e3, 1171e3, 1271e3, 1371e3, 1471e3, 1571e3, 1671e3, 1771e3, 1871e3, 1971e3,
        2071e3, 2171e3, 2271e3, 2371e3, 2471e3, 2571e3, 2671e3, 2741e3, 2771e3, 2871e3,
        2891e3 };
    const double ref_p[46] = { 0e8, 0.3e8, 3.3e8, 6.0e8, 11.2e8,
        17.8e8, 24.5e8, 36.1e8, 47.8e8, 59.4e8, 71.1e8, 86.4e8, 102.0e8, 117.7e8, 133.5e8,
        152.2e8, 171.3e8, 190.7e8, 210.4e8, 224.3e8, 238.3e8, 260.7e8, 282.9e8, 327.6e8, 372.8e8,
        418.6e8, 464.8e8, 511.6e8, 558.9e8, 606.8e8, 655.2e8, 704.1e8, 753.5e8, 803.6e8, 854.3e8,
        905.6e8, 957.6e8, 1010.3e8, 1063.8e8, 1118.2e8, 1173.4e8, 1229.7e8, 1269.7e8, 1287.0e8,
        1345.6e8, 1357.5e8 };
    if (depth <= 0) return 0;
    int n;
    for (n = 1; n < 46; n++)
        if (depth <= ref_depth[n]) break;
    double p0 = ref_p[n-1], p1 = ref_p[n];
    if (modified) {     // the continental-crust variant differs in entries 1..3 only
        if (n-1 == 1) p0 = 0.82e8; else if (n-1 == 2) p0 = 4.1e8; else if (n-1 == 3) p0 = 6.7e8;
        if (n == 1) p1 = 0.82e8; else if (n == 2) p1 = 4.1e8; else if (n == 3) p1 = 6.7e8;
    }
    return p0 + (p1 - p0) * (depth - ref_depth[n-1]) / (ref_depth[n] - ref_depth[n-1]);
}

// ref_pressure, matprops.cxx:153-174 (no hydraulic diffusion)
__device__ __forceinline__ double ref_pressure(const DevParams &P, double z)
{
    double depth = -z;
    double p = 0;
    if (P.ref_pressure_option == 0)
        p = P.rho0[P.mattype_ref] * P.gravity * depth;
    else if (P.ref_pressure_option == 1)
        p = prem_pressure(depth, false);
    else if (P.ref_pressure_option == 2)
        p = prem_pressure(depth, true);
    return p;
}

// ---------------------------------------------------------------------------------------
// 3x3 symmetric eigen-solvers: a device restatement of J. Kopp's algorithms (arXiv
// physics/0610206) as the reference vendors them in 3x3-C/ (dsyevc3.c, dsytrd3.c, dsyevq3.c,
// dsyevh3.c).  The branch decisions and the operation order must match that library for the
// yield decisions to match the reference, so eig3_cardano / eig3_tridiag / eig3_ql /
// eig3_hybrid below follow it statement by statement; they are a derived work of it:
//
//   Numerical diagonalization of 3x3 matrcies
//   Copyright (C) 2006  Joachim Kopp
//
//   This library is free software; you can redistribute it and/or modify it under the terms
//   of the GNU Lesser General Public License as published by the Free Software Foundation;
//   either version 2.1 of the License, or (at your option) any later version.
//   This library is distributed in the hope that it will be useful, but WITHOUT ANY WARRANTY;
//   without even the implied warranty of MERCHANTABILITY or FITNESS FOR A PARTICULAR PURPOSE.
//   See the GNU Lesser General Public License for more details.
//
// A is the upper triangle a00 a01 a02 / a11 a12 / a22.
// ---------------------------------------------------------------------------------------
struct Sym3 { double a00, a11, a22, a01, a02, a12; };

// Cardano eigenvalues, 3x3-C/dsyevc3.c:33-81
__device__ __forceinline__ void eig3_cardano(const Sym3 &A, double w[3])
{
    const double sqrt3 = 1.73205080756887729352744634151;
    double de = A.a01 * A.a12;
    double dd = A.a01 * A.a01;
    double ee = A.a12 * A.a12;
    double ff = A.a02 * A.a02;
    double m  = A.a00 + A.a11 + A.a22;
    double c1 = (A.a00*A.a11 + A.a00*A.a22 + A.a11*A.a22) - (dd + ee + ff);
    double c0 = A.a22*dd + A.a00*ee + A.a11*ff - A.a00*A.a11*A.a22 - 2.0 * A.a02*de;

    double p = m*m - 3.0*c1;
    double q = m*(p - (3.0/2.0)*c1) - (27.0/2.0)*c0;
    double sqrt_p = sqrt(fabs(p));

    double phi = 27.0 * ( 0.25*(c1*c1)*(p - c1) + c0*(q + 27.0/4.0*c0));
    phi = (1.0/3.0) * atan2(sqrt(fabs(phi)), q);

    double c = sqrt_p*cos(phi);
    double s = (1.0/sqrt3)*sqrt_p*sin(phi);

    w[1]  = (1.0/3.0)*(m - c);
    w[2]  = w[1] + s;
    w[0]  = w[1] + c;
    w[1] -= s;
}

// Householder tridiagonalisation, 3x3-C/dsytrd3.c:39-105
__device__ __noinline__ void eig3_tridiag(const Sym3 &A, double Q[3][3], double d[3], double e[2])
{
    double u1, u2, q1, q2, omega, f, K, h, g;
    Q[0][0] = Q[1][1] = Q[2][2] = 1.0;
    Q[1][0] = Q[0][1] = Q[2][0] = Q[0][2] = Q[2][1] = Q[1][2] = 0.0;

    h = A.a01*A.a01 + A.a02*A.a02;
    g = (A.a01 > 0) ? -sqrt(h) : sqrt(h);
    e[0] = g;
    f    = g * A.a01;
    u1 = A.a01 - g;
    u2 = A.a02;

    omega = h - f;
    if (omega > 0.0) {
        omega = 1.0 / omega;
        K = 0.0;
        f  = A.a11 * u1 + A.a12 * u2;  q1 = omega * f;  K += u1 * f;
        f  = A.a12 * u1 + A.a22 * u2;  q2 = omega * f;  K += u2 * f;
        K *= 0.5 * (omega*omega);
        q1 = q1 - K * u1;
        q2 = q2 - K * u2;

        d[0] = A.a00;
        d[1] = A.a11 - 2.0*q1*u1;
        d[2] = A.a22 - 2.0*q2*u2;

        f = omega * u1;
        Q[1][1] = Q[1][1] - f*u1;
        Q[2][1] = Q[2][1] - f*u2;
        f = omega * u2;
        Q[1][2] = Q[1][2] - f*u1;
        Q[2][2] = Q[2][2] - f*u2;

        e[1] = A.a12 - q1*u2 - u1*q2;
    }
    else {
        d[0] = A.a00; d[1] = A.a11; d[2] = A.a22;
        e[1] = A.a12;
    }
}

// QL with implicit shifts, 3x3-C/dsyevq3.c:49-135
__device__ __noinline__ int eig3_ql(const Sym3 &A, double Q[3][3], double w[3])
{
    const int n = 3;
    double e[3];
    double g, r, p, f, b, s, c, t;
    int m;

    eig3_tridiag(A, Q, w, e);

    for (int l = 0; l < n-1; l++) {
        int nIter = 0;
        while (1) {
            for (m = l; m <= n-2; m++) {
                g = fabs(w[m]) + fabs(w[m+1]);
                if (fabs(e[m]) + g == g) break;
            }
            if (m == l) break;
            if (nIter++ >= 30) return -1;

            g = (w[l+1] - w[l]) / (e[l] + e[l]);
            r = sqrt(g*g + 1.0);
            if (g > 0) g = w[m] - w[l] + e[l]/(g + r);
            else       g = w[m] - w[l] + e[l]/(g - r);

            s = c = 1.0;
            p = 0.0;
            for (int i = m-1; i >= l; i--) {
                f = s * e[i];
                b = c * e[i];
                if (fabs(f) > fabs(g)) {
                    c      = g / f;
                    r      = sqrt(c*c + 1.0);
                    e[i+1] = f * r;
                    c     *= (s = 1.0/r);
                }
                else {
                    s      = f / g;
                    r      = sqrt(s*s + 1.0);
                    e[i+1] = g * r;
                    s     *= (c = 1.0/r);
                }
                g = w[i+1] - p;
                r = (w[i] - g)*s + 2.0*c*b;
                p = s * r;
                w[i+1] = g + p;
                g = c*r - b;
                for (int k = 0; k < n; k++) {
                    t = Q[k][i+1];
                    Q[k][i+1] = s*Q[k][i] + c*t;
                    Q[k][i]   = c*Q[k][i] - s*t;
                }
            }
            w[l] -= p;
            e[l]  = g;
            e[m]  = 0.0;
        }
    }
    return 0;
}

// Cardano + cross-product eigenvectors with QL fallback, 3x3-C/dsyevh3.c:57-135
__device__ __noinline__ void eig3_hybrid(const Sym3 &A, double Q[3][3], double w[3])
{
    eig3_cardano(A, w);

    double t = fabs(w[0]), u;
    if ((u = fabs(w[1])) > t) t = u;
    if ((u = fabs(w[2])) > t) t = u;
    if (t < 1.0) u = t; else u = t*t;
    double error = 256.0 * DBL_EPSILON * (u*u);

    Q[0][1] = A.a01*A.a12 - A.a02*A.a11;
    Q[1][1] = A.a02*A.a01 - A.a12*A.a00;
    Q[2][1] = A.a01*A.a01;

    Q[0][0] = Q[0][1] + A.a02*w[0];
    Q[1][0] = Q[1][1] + A.a12*w[0];
    Q[2][0] = (A.a00 - w[0]) * (A.a11 - w[0]) - Q[2][1];
    double norm = Q[0][0]*Q[0][0] + Q[1][0]*Q[1][0] + Q[2][0]*Q[2][0];
    if (norm <= error) { eig3_ql(A, Q, w); return; }
    norm = sqrt(1.0 / norm);
    for (int j = 0; j < 3; j++) Q[j][0] = Q[j][0] * norm;

    Q[0][1] = Q[0][1] + A.a02*w[1];
    Q[1][1] = Q[1][1] + A.a12*w[1];
    Q[2][1] = (A.a00 - w[1]) * (A.a11 - w[1]) - Q[2][1];
    norm = Q[0][1]*Q[0][1] + Q[1][1]*Q[1][1] + Q[2][1]*Q[2][1];
    if (norm <= error) { eig3_ql(A, Q, w); return; }
    norm = sqrt(1.0 / norm);
    for (int j = 0; j < 3; j++) Q[j][1] = Q[j][1] * norm;

    Q[0][2] = Q[1][0]*Q[2][1] - Q[2][0]*Q[1][1];
    Q[1][2] = Q[2][0]*Q[0][1] - Q[0][0]*Q[2][1];
    Q[2][2] = Q[0][0]*Q[1][1] - Q[1][0]*Q[0][1];
}

// sort_principal3, rheology.cxx:24-45: (0,1),(1,2),(0,1) compare-swap network
__device__ __forceinline__ void sort3(double p[3], double (*v)[3])
{
    #define DES_CSWAP(i, j) \
        if (p[i] > p[j]) { double tmp = p[i]; p[i] = p[j]; p[j] = tmp; \
            if (v) { for (int r = 0; r < 3; ++r) { double b = v[r][i]; v[r][i] = v[r][j]; v[r][j] = b; } } }
    DES_CSWAP(0, 1) DES_CSWAP(1, 2) DES_CSWAP(0, 1)
    #undef DES_CSWAP
}

// ---------------------------------------------------------------------------------------
// rheology, rheology.cxx
// ---------------------------------------------------------------------------------------

// elastic, rheology.cxx:250-260
__device__ __forceinline__ void rh_elastic(double bulkm, double shearm, const double *de, double *s)
{
    double lambda = bulkm - 2. / 3 * shearm;
    double dev = trace3(de);
    for (int i = 0; i < 3; ++i) s[i] += 2 * shearm * de[i] + lambda * dev;
    for (int i = 3; i < 6; ++i) s[i] += 2 * shearm * de[i];
}

// maxwell, rheology.cxx:279-295
__device__ __forceinline__ void rh_maxwell(double bulkm, double shearm, double viscosity, double dt,
                                           double dv, const double *de, double *s)
{
    double tmp = 0.5 * dt * shearm / viscosity;
    double f1 = 1 - tmp;
    double f2 = 1 / (1 + tmp);
    double dev = trace3(de) / 3;
    double s0 = trace3(s) / 3;
    for (int i = 0; i < 3; ++i)
        s[i] = ((s[i] - s0) * f1 + 2 * shearm * (de[i] - dev)) * f2 + s0 + bulkm * dv;
    for (int i = 3; i < 6; ++i)
        s[i] = (s[i] * f1 + 2 * shearm * de[i]) * f2;
}

// The part of elasto_plastic() after the prefilter: accurate eigen-decomposition, yield
// test, return mapping, rotation back (rheology.cxx:363-475).  Rare (~0.2 % of elements),
// kept out of line so the common path stays lean in registers.
__device__ __noinline__ void rh_plastic_return(double bulkm, double shearm, double amc, double anphi,
                                               double anpsi, double hardn, double ten_max,
                                               double &depls, double *s)
{
    double p[3], v[3][3];
    Sym3 A = { s[0], s[1], s[2], s[3], s[4], s[5] };
    eig3_hybrid(A, v, p);
    sort3(p, v);

    double fs = p[0] - p[2] * anphi + amc;
    double ft = p[2] - ten_max;
    if (fs > 0 && ft < 0) return;

    double pa = sqrt(1 + anphi*anphi) + anphi;
    double ps = ten_max * anphi - amc;
    double h = p[2] - ten_max + pa * (p[0] - ps);
    double a1 = bulkm + 4. / 3 * shearm;
    double a2 = bulkm - 2. / 3 * shearm;

    double alam;
    if (h < 0) {    // shear failure
        alam = fs / (a1 - a2*anpsi + a1*anphi*anpsi - a2*anphi + 2*sqrt(anphi)*hardn);
        p[0] -= alam * (a1 - a2 * anpsi);
        p[1] -= alam * (a2 - a2 * anpsi);
        p[2] -= alam * (a2 - a1 * anpsi);
        depls = fabs(alam) * sqrt((7 + 4*anpsi + 7*anpsi*anpsi) / 18);
    }
    else {          // tensile failure
        alam = ft / a1;
        p[0] -= alam * a2;
        p[1] -= alam * a2;
        p[2] -= alam * a1;
        depls = fabs(alam) * sqrt(7. / 18);
    }

    double ss[3][3] = {{0,0,0},{0,0,0},{0,0,0}};
    for (int m = 0; m < 3; m++)
        for (int n = m; n < 3; n++)
            for (int k = 0; k < 3; k++)
                ss[m][n] += v[m][k] * v[n][k] * p[k];
    s[0] = ss[0][0]; s[1] = ss[1][1]; s[2] = ss[2][2];
    s[3] = ss[0][1]; s[4] = ss[0][2]; s[5] = ss[1][2];
}

// elasto_plastic (3D Mohr-Coulomb with tension cut-off), rheology.cxx:314-484
//
// The reference runs a Cardano eigenvalue PREFILTER (rheology.cxx:351-361) whose only effect
// is to skip the accurate eigen-decomposition for elements that are clearly elastic; an
// element the prefilter lets through and the accurate test (fs > 0 && ft < 0, :383) then
// finds elastic comes out exactly like a skipped one (s = elastic trial, depls = 0).  The
// outcome therefore only requires that no truly yielding element is skipped, and any
// conservative test may run first.  QUICK test: with m = tr(s)/3 and a = 2 sqrt(J2/3) every
// eigenvalue lies in [m - a, m + a], so fs >= (m - a) - (m + a) anphi + amc and
// ft <= m + a - ten_max; elements beyond the reference's own 1e-2 band (widened to the
// bounds' magnitudes) on both are skipped without any transcendental.  Everything else runs
// the reference's Cardano prefilter and accurate path unchanged.
__device__ __forceinline__ void rh_elasto_plastic(double bulkm, double shearm, double amc, double anphi,
                                                  double anpsi, double hardn, double ten_max,
                                                  const double *de, double &depls, double *s)
{
    rh_elastic(bulkm, shearm, de, s);
    depls = 0;
    {
        const double m = (s[0] + s[1] + s[2]) / 3;
        const double a = 2 * sqrt(second_invariant2(s) / 3);
        const double big = fabs(m) + a;
        const double bandq = 1e-2 * (big + anphi * big + fabs(amc));
        if ((m - a) - (m + a) * anphi + amc > bandq && (m + a) - ten_max < -bandq) return;
    }
    {   // eigenvalue-only prefilter, rheology.cxx:351-361 (band 1e-2)
        double pf[3];
        Sym3 A = { s[0], s[1], s[2], s[3], s[4], s[5] };
        eig3_cardano(A, pf);
        sort3(pf, nullptr);
        const double band = 1e-2 * (fabs(pf[0]) + anphi * fabs(pf[2]) + fabs(amc));
        if (pf[0] - pf[2] * anphi + amc > band && pf[2] - ten_max < -band) return;
    }
    rh_plastic_return(bulkm, shearm, amc, anphi, anpsi, hardn, ten_max, depls, s);
}

// jaumann_rate_3d, fields.cxx:791-805
__device__ __forceinline__ void jaumann3(double *s, double dt, double w3, double w4, double w5)
{
    double inc[6];
    inc[0] = -2.0 * s[3] * w3 - 2.0 * s[4] * w4;
    inc[1] =  2.0 * s[3] * w3 - 2.0 * s[5] * w5;
    inc[2] =  2.0 * s[4] * w4 + 2.0 * s[5] * w5;
    inc[3] = s[0] * w3 - s[1] * w3 - s[4] * w5 - s[5] * w4;
    inc[4] = s[0] * w4 - s[2] * w4 + s[3] * w5 - s[5] * w3;
    inc[5] = s[1] * w5 - s[2] * w5 + s[3] * w4 + s[4] * w3;
    for (int i = 0; i < 6; ++i) s[i] += dt * inc[i];
}

// ---------------------------------------------------------------------------------------
// apply_vbcs for one boundary node (3D), bc.cxx:396-651
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void vbc_node(const DevParams &P, unsigned flag, double v[3], int bc_z1,
                                         const int *__restrict__ edge_slot,
                                         const double *__restrict__ edge_vec)
{
    // lateral faces x0 x1 y0 y1: {mask, normal comp, lateral comp} (bc.cxx:359-365,412-424)
    #pragma unroll
    for (int fi = 0; fi < 4; ++fi) {
        if (!(flag & (1u << fi))) continue;
        const int ni = (fi < 2) ? 0 : 1, li = (fi < 2) ? 1 : 0;
        const double val = P.vbc_values[fi], val_l = P.vbc_val_lateral[fi];
        switch (P.vbc_types[fi]) {
        case 0: break;
        case 1: v[ni] = val; break;
        case 2: v[li] = 0; v[2] = 0; break;
        case 3: v[ni] = val; v[li] = 0; v[2] = 0; break;
        case 4: v[li] = val; v[2] = 0; break;
        case 5: v[ni] = 0; v[li] = val; v[2] = 0; break;
        case 6: v[ni] = val; v[li] = val_l; break;
        case 7: v[ni] = val; v[li] = 0; break;
        }
    }

    // slant boundaries n0..n3 (bc.cxx:491-585)
    if (flag & (15u << 6)) {
        for (int ib = 6; ib <= 9; ib++) {
            if (!(flag & (1u << ib))) continue;
            const double *n = P.bnormals[ib];
            double fac = 0;
            const int type = P.vbc_types[ib];
            if (type == 1 || type == 11) {
                const int nd = (type == 1) ? 3 : 2;
                double target = P.vbc_values[ib];
                if (type == 11) {
                    fac = 1 / sqrt(1 - n[2]*n[2]);
                    target = P.vbc_values[ib] * fac;
                }
                if (flag == (1u << ib)) {
                    double vn = 0;
                    for (int d = 0; d < nd; d++) vn += v[d] * n[d];
                    for (int d = 0; d < nd; d++) v[d] += (target - vn) * n[d];
                }
                else {
                    for (int ic = 0; ic < ib; ic++) {
                        if (!(flag & (1u << ic))) continue;
                        if (P.vbc_types[ic] == 0) {
                            double vn = 0;
                            for (int d = 0; d < nd; d++) vn += v[d] * n[d];
                            for (int d = 0; d < nd; d++) v[d] += (target - vn) * n[d];
                        }
                        else if (P.vbc_types[ic] == 1) {
                            const int slot = edge_slot[ic*DES_NBDRY + ib];
                            if (slot < 0) continue;
                            const double *edge = &edge_vec[slot*3];
                            double ve = 0;
                            for (int d = 0; d < 3; d++) ve += v[d] * edge[d];
                            for (int d = 0; d < 3; d++) v[d] = ve * edge[d];
                        }
                    }
                }
            }
            else if (type == 3) {
                for (int d = 0; d < 3; d++) v[d] = P.vbc_values[ib] * n[d];
            }
            else if (type == 13) {
                fac = 1 / sqrt(1 - n[2]*n[2]);
                for (int d = 0; d < 2; d++) v[d] = P.vbc_values[ib] * fac * n[d];
                v[2] = 0;
            }
        }
    }

    // z faces, dealt last (bc.cxx:588-650)
    const int bc_z0 = P.vbc_types[4];
    if (bc_z0 == 0 && bc_z1 == 0) return;
    if (flag & BOUNDZ0) {
        switch (bc_z0) {
        case 1: v[2] = P.vbc_values[4]; break;
        case 2: v[0] = 0; v[1] = 0; break;
        case 3: v[0] = 0; v[1] = 0; v[2] = P.vbc_values[4]; break;
        }
    }
    if (flag & BOUNDZ1) {
        switch (bc_z1) {
        case 1: v[2] = P.vbc_values[5]; break;
        case 2: v[0] = 0; v[1] = 0; break;
        case 3: v[0] = 0.0; v[1] = 0; v[2] = P.vbc_values[5]; break;
        case 4: v[0] = P.vbc_values[5]; v[1] = 0; v[2] = 0; break;
        }
    }
}

// apply_damping for one component, fields.cxx:483-579 (options 0..4)
// crit = 2 sqrt(mass * ymass) of the node (option 4 only)
__device__ __forceinline__ double damp(const DevParams &P, double f, double v, double crit = 0)
{
    const double small_vel = 1e-13;
    switch (P.damping_option) {
    case 4:     // Rayleigh, fields.cxx:547-570
        if (fabs(v) > small_vel) {
            const double f_C = P.damping_factor * copysign(f, v);
            const double f_V = crit * v;
            const double f_damping = (fabs(f_C) < fabs(f_V)) ? f_V : f_C;
            f -= f_damping;
        }
        break;
    case 1:
        if (fabs(v) > small_vel) f -= P.damping_factor * copysign(f, v);
        break;
    case 2:
        f -= P.damping_factor * f;
        break;
    case 3:
        // fields.cxx:538 has a comma operator: the velocity factor is a no-op
        if ((f < 0) == (v < 0)) f -= P.damping_factor * f;
        else f += (1 - P.damping_factor) * f;
        break;
    default: break;
    }
    return f;
}
