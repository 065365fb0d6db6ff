// Schedule v2 of the fused explicit step (sm_100a, FP64) -- see DESIGN.md section 4.
//
//   E-A  k2_elem_a<TAIL>      tail(n-1): volume, mass contributions
//                             head(n)  : T-diffusion contributions, tr(edot)*vol
//   N-A  k3_node_a<MASS>      CSR gather: volume_n/mass/tmass, T update, dvoldt
//   E-B  k2_elem_b<ROT>       tail(n-1): Jaumann rotation of stress and strain (dt_prev)
//                             head(n)  : strain rate, edvoldt, anti-locking, strain, rheology,
//                                        plstrain, dpressure*vol
//   N-B  k_node_b             (des_kernels.cuh) CSR gather: dp_nd
//   E-C  k_elem_c<MIXED>      (des_kernels.cuh) NMD correction, element forces
//   N-C  k3_node_c            CSR gather of forces, stress bcs, damping, v, vbc, x
//
// Differences from v1 (des_kernels.cuh, kept as the un-fused twins and the flush path):
//  * strain rate and spin are computed where they are consumed (E-B), so strain_rate is
//    written once per step and stress/strain cross HBM once per direction in E-B instead
//    of once more for the rotation in E-A;
//  * the node gathers read each support entry's record with one 256-bit load (+ one double)
//    and keep four entries in flight per trip; they sum in ascending element order (the
//    reference's order, fields.cxx:659-676).  Three alternatives were built, measured and
//    dropped (profiles/experiments/): staging a block's entries through
//    shared memory, a warp-interleaved copy of the support list, and staging an element
//    tile's distinct nodes in shared memory for the element kernels.
#pragma once
#include "des_kernels.cuh"

// ---------------------------------------------------------------------------------------
// E-A (v2)
// ---------------------------------------------------------------------------------------
#define DES_EA_BLOCK 128      // measured: 128-thread blocks are 7 % faster than 256 here (72 registers)
// Measured on the 50M box (profiles/experiments/): requesting the velocities only after the
// shape functions are done (their registers reuse the coordinates') and asking for 6 resident
// blocks (80 registers, 8 bytes spilled) takes E-A from 1.63 to 1.50 ms; 7 blocks (72
// registers, 56 bytes spilled) is slower (1.84 ms), and so is the bound without the late loads.
#ifndef DES_EA_LATE
#define DES_EA_LATE 1
#endif
#ifndef DES_EA_MINB
#define DES_EA_MINB 6
#endif
#if DES_EA_MINB > 0
#define DES_EA_BOUNDS __launch_bounds__(DES_EA_BLOCK, DES_EA_MINB)
#else
#define DES_EA_BOUNDS __launch_bounds__(DES_EA_BLOCK)
#endif
// DT: also the element loop and reductions of compute_dt (geometry.cxx:1511-1593), which reads
// the very coordinates, velocities, volume and density this launch already holds -- on the
// steps where the host asks for a new dt (every 10th) E-A of the NEXT step runs early with DT
// set (des_gpu_compute_dt) and writes one DtPartial per block; k_dt_final reduces them.
// block reduction of the five compute_dt quantities -> partials[blockIdx.x] (min and max are
// exact in any order)
__device__ __forceinline__ void k2_dt_reduce(double minl, double dtm, double dtd, double gmin, double vem,
                                             DtPartial *__restrict__ partials)
{
    __shared__ DtPartial sh[DES_EA_BLOCK / 32];
    minl = warp_min(minl); dtm = warp_min(dtm); dtd = warp_min(dtd); gmin = warp_min(gmin);
    vem = warp_max(vem);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) sh[w] = DtPartial{minl, dtm, dtd, gmin, vem};
    __syncthreads();
    if (threadIdx.x == 0) {
        DtPartial r = sh[0];
        for (int i = 1; i < DES_EA_BLOCK / 32; ++i) {
            r.minl = fmin(r.minl, sh[i].minl); r.dt_maxwell = fmin(r.dt_maxwell, sh[i].dt_maxwell);
            r.dt_diffusion = fmin(r.dt_diffusion, sh[i].dt_diffusion);
            r.global_dt_min = fmin(r.global_dt_min, sh[i].global_dt_min);
            r.max_vem = fmax(r.max_vem, sh[i].max_vem);
        }
        partials[blockIdx.x] = r;
    }
}

template <bool SINGLE, bool TAIL, bool DT = false>
__global__ void __launch_bounds__(DES_EA_BLOCK, DT ? 5 : (DES_EA_MINB > 0 ? DES_EA_MINB : 1))
k2_elem_a(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, int tail_flags, int head_flags,
          DtPartial *__restrict__ dt_partials = nullptr)
{
    // tail_flags: 1 volume, 2 mass contributions;  head_flags: 1 T-diffusion, 4 tr(edot)*vol
    int e = D.e0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (DT) e = min(e, D.e1 - 1);   // every thread reaches the block reduction at the end: the
                                    // surplus threads of the last block redo its last element
    else if (e >= D.e1) return;
    const int nmat = SINGLE ? 1 : P.nmat;
    // every load of the kernel is issued here, before the first use of any of them: the
    // mid-kernel store of volume[e] (which may alias, as far as the compiler knows) and the
    // flag branches otherwise serialise them into five dependent memory round trips
    // (conn -> xT -> markers -> radiogenic -> vv in the SASS of the previous version)
    const int4 c = D.conn[e];
    Markers mk;
    load_markers(D.elemmarkers, e, nmat, mk);
    double radiogenic = 0, vol = 0;
    if (head_flags & 1) radiogenic = D.radiogenic[e];
    if (!(TAIL && (tail_flags & 1))) vol = D.volume[e];
    const double4 a0 = ld256_nc(D.xT + c.x), a1 = ld256_nc(D.xT + c.y), a2 = ld256_nc(D.xT + c.z), a3 = ld256_nc(D.xT + c.w);
    double4 v0 = make_double4(0, 0, 0, 0), v1 = v0, v2 = v0, v3 = v0;
#if !DES_EA_LATE
    if (head_flags & 4) {
        v0 = ld256_nc(D.vv + c.x); v1 = ld256_nc(D.vv + c.y); v2 = ld256_nc(D.vv + c.z); v3 = ld256_nc(D.vv + c.w);
    }
#endif

    if (D.pf) {
        const int ep = e + D.pf;
        if (ep < D.nelem) {
            pf_l2(D.conn + ep);
            if (SINGLE) pf_l2(D.elemmarkers + ep);
            if (head_flags & 1) pf_l2(D.radiogenic + ep);
        }
    }

    if (TAIL && (tail_flags & 1)) {
        vol = tet_volume(a0, a1, a2, a3);       // compute_volume, geometry.cxx:180-197
        D.volume[e] = vol;
        if (!(vol > 0.0)) atomicOr(D.status, 1);    // inverted / NaN element: DES_E_NAN at the next sync point
    }

    double sx[4], sy[4], sz[4];
    shape_fn(a0, a1, a2, a3, vol, sx, sy, sz);

#if DES_EA_LATE
    // the velocities are requested when the coordinates' registers are free again
    if (head_flags & 4) {
        v0 = ld256_nc(D.vv + c.x); v1 = ld256_nc(D.vv + c.y); v2 = ld256_nc(D.vv + c.z); v3 = ld256_nc(D.vv + c.w);
    }
#endif
    const double rho = mat_rho(P, nmat, mk, a0.w, a1.w, a2.w, a3.w);

    double rec[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    rec[7] = vol;
    if (TAIL && (tail_flags & 2)) {
        // compute_mass element loop, geometry.cxx:1794-1840
        const double bulkm = harm_mean(P.bulk_modulus, mk.cnt, nmat);
        double rho_m = P.is_quasi_static ? bulkm / (P.pseudo_speed * P.pseudo_speed) : rho;
        rec[5] = rho_m * vol / 4;
        const double cp = arith_mean(P.heat_capacity, mk.cnt, nmat);
        rec[6] = rho * cp * vol / 4;
    }
    if (head_flags & 1) {
        // update_temperature element loop, fields.cxx:210-239
        const double k = arith_mean(P.therm_cond, mk.cnt, nmat);
        const double kv = k * vol;
        const double rh = radiogenic * vol * rho / 4;
        const double T[4] = {a0.w, a1.w, a2.w, a3.w};
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            double diffusion = 0.;
            #pragma unroll
            for (int j = 0; j < 4; ++j)
                diffusion += (sx[i] * sx[j] + sy[i] * sy[j] + sz[i] * sz[j]) * T[j];
            rec[i] = diffusion * kv - rh;
        }
    }
    if (head_flags & 4) {
        // diagonal of update_strain_rate (fields.cxx:414-476) -> compute_dvoldt element
        // loop (geometry.cxx:217-225)
        double s[3];
        s[0] = 0; s[0] += v0.x * sx[0]; s[0] += v1.x * sx[1]; s[0] += v2.x * sx[2]; s[0] += v3.x * sx[3];
        s[1] = 0; s[1] += v0.y * sy[0]; s[1] += v1.y * sy[1]; s[1] += v2.y * sy[2]; s[1] += v3.y * sy[3];
        s[2] = 0; s[2] += v0.z * sz[0]; s[2] += v1.z * sz[1]; s[2] += v2.z * sz[2]; s[2] += v3.z * sz[3];
        rec[4] = trace3(s) * vol;
    }
    // two 256-bit stores per element, each contiguous across the warp (8 L1TEX wavefronts per
    // store instead of the 16 per 128-bit store of an interleaved 64-byte record)
    if (head_flags & 1) st256(D.ca + size_t(e) * 4, make_double4(rec[0], rec[1], rec[2], rec[3]));
    if (TAIL) st256(D.ca_hi + size_t(e) * 4, make_double4(rec[4], rec[5], rec[6], rec[7]));
    else if (head_flags & 4) D.ca_hi[size_t(e) * 4] = rec[4];

    if (DT) {
        // compute_dt's element loop, statement for statement as k_dt (des_kernels.cuh)
        const double weight = 1.0 / 4;
        double vxe = 0.0, vye = 0.0, vze = 0.0;
        vxe += v0.x * weight; vye += v0.y * weight; vze += v0.z * weight;
        vxe += v1.x * weight; vye += v1.y * weight; vze += v1.z * weight;
        vxe += v2.x * weight; vye += v2.y * weight; vze += v2.z * weight;
        vxe += v3.x * weight; vye += v3.y * weight; vze += v3.z * weight;
        const double vem = sqrt(vxe*vxe + vye*vye + vze*vze);
        const double maxa = fmax(fmax(tri_area(a0, a1, a2), tri_area(a0, a1, a3)),
                                 fmax(tri_area(a2, a3, a0), tri_area(a2, a3, a1)));
        const double minh = 3 * vol / maxa;
        const double shearm = harm_mean(P.shear_modulus, mk.cnt, nmat);
        const double dtm = 0.5 * P.visc_min / (1e-40 + shearm);
        const double dtd = P.has_thermal_diffusion ? 0.5 * minh * minh / P.therm_diff_max : DBL_MAX;
        const double gmin = minh / sqrt(shearm / rho) / 5.0;
        k2_dt_reduce(minh, dtm, dtd, gmin, vem, dt_partials);
    }
}

// ---------------------------------------------------------------------------------------
// E-B (v2): rotate_stress of the step that ended (fields.cxx:851-898) + update_strain_rate
// (fields.cxx:414-476) + compute_edvoldt (geometry.cxx:263-272) + update_stress
// (rheology.cxx:727-1026) + first loop of NMD_stress (geometry.cxx:293-295).
// The coordinates, velocities and volume the rotation needs (end of step n-1) are the ones
// the strain rate of step n needs, so one shape-function evaluation serves both.
// ---------------------------------------------------------------------------------------
// Measured on the 50M box: stress and strain requested after the strain rate and spin are
// computed (116 -> 96 registers) and 6 resident blocks (80 registers, 8 bytes spilled): E-B
// 3.69 -> 3.44 ms; 7 and 8 blocks spill (3.98, 4.78 ms).
#ifndef DES_EB_LATE
#define DES_EB_LATE 1
#endif
#ifndef DES_EB_MINB
#define DES_EB_MINB 6
#endif
#if DES_EB_MINB > 0
#define DES_EB_BOUNDS __launch_bounds__(128, DES_EB_MINB)
#else
#define DES_EB_BOUNDS __launch_bounds__(128)
#endif
// AVG (only with ROT): Output::average_fields (output.cxx:312-352) of the step that just ended,
// whose call the host deferred to this launch -- the rotated stress it must see is in
// registers here.  avg_first: the interval's first step (snapshot) instead of a running sum.
// TWIN (never with ROT / AVG): the un-fused reference call update_stress (rheology.cxx:703) --
// the strain rate and edvoldt are the STORED ones (update_strain_rate and compute_edvoldt ran
// as their own calls), everything after that is the same code as the fused step.
template <bool SINGLE, bool ROT, int RHEOL, bool AVG = false, bool TWIN = false>
__global__ void DES_EB_BOUNDS
k2_elem_b(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt, double dt_rot,
          int avg_first = 0, int pre = 0)
{
    if (D.sc) { dt = D.sc[0]; dt_rot = D.sc[1]; }
    // pre: what the previous step left for this launch to apply to the stored stress first,
    // in the reference's order -- 1: NMD_stress's diagonal correction ddp[e] (geometry.cxx:
    // 315-331, computed by E-C), 2: correct_surface_element's divisions by rdv for the top
    // elements (bc.cxx:1669-1686); then (ROT) rotate_stress
    const int e = D.e0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.e1) return;
    const int nmat = SINGLE ? 1 : P.nmat;
    const int E = D.nelem;
    const int4 c = D.conn[e];
    const double vol = D.volume[e];
    // the streaming (coalesced) loads are issued before the dependent gathers -- all of them,
    // including the ones the rheology needs after the strain stores (the compiler cannot
    // move a load above a store that may alias)
    Markers mk;
    load_markers(D.elemmarkers, e, nmat, mk);
    double pls = 0, vol_old = 1;
    if (RHEOL == DES_RH_EP || RHEOL == DES_RH_EVP) pls = D.plstrain[e];
    if (RHEOL == DES_RH_MAXWELL || RHEOL == DES_RH_EVP) vol_old = D.volume_old[e];
    double s[6], es[6];
#if !DES_EB_LATE
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
        s[i] = D.stress[size_t(i) * E + e];
        es[i] = D.strain[size_t(i) * E + e];
    }
#endif

    double edot[6], w3 = 0, w4 = 0, w5 = 0;
    double T0, T1, T2, T3;
    double edvoldt;
    if (TWIN) {
        edvoldt = D.edvoldt[e];
        #pragma unroll
        for (int i = 0; i < 6; ++i) edot[i] = D.strain_rate[size_t(i) * E + e];
        T0 = D.xT[c.x].w; T1 = D.xT[c.y].w; T2 = D.xT[c.z].w; T3 = D.xT[c.w].w;
    }
    else {
        const double4 a0 = ld256_nc(D.xT + c.x), a1 = ld256_nc(D.xT + c.y), a2 = ld256_nc(D.xT + c.z), a3 = ld256_nc(D.xT + c.w);
        const double4 v0 = ld256_nc(D.vv + c.x), v1 = ld256_nc(D.vv + c.y), v2 = ld256_nc(D.vv + c.z), v3 = ld256_nc(D.vv + c.w);
        {
            double dj = 0;
            dj += D.ntmp[c.x]; dj += D.ntmp[c.y]; dj += D.ntmp[c.z]; dj += D.ntmp[c.w];
            edvoldt = dj / 4;
        }
        if (D.pf) {
            const int ep = e + D.pf;
            if (ep < E) {
                pf_l2(D.conn + ep);
                pf_l2(D.volume + ep);
                if (SINGLE) pf_l2(D.elemmarkers + ep);
                #pragma unroll
                for (int i = 0; i < 6; ++i) {
                    pf_l2(D.stress + size_t(i) * E + ep);
                    pf_l2(D.strain + size_t(i) * E + ep);
                }
                if (RHEOL == DES_RH_EP || RHEOL == DES_RH_EVP) pf_l2(D.plstrain + ep);
                if (RHEOL == DES_RH_MAXWELL || RHEOL == DES_RH_EVP) pf_l2(D.volume_old + ep);
                if (pre & 1) pf_l2(D.ddp + ep);
                if (pre & 2) pf_l2(D.eflag + ep);
            }
        }
        T0 = a0.w; T1 = a1.w; T2 = a2.w; T3 = a3.w;
        double sx[4], sy[4], sz[4];
        shape_fn(a0, a1, a2, a3, vol, sx, sy, sz);
        const double vx[4] = {v0.x, v1.x, v2.x, v3.x};
        const double vy[4] = {v0.y, v1.y, v2.y, v3.y};
        const double vz[4] = {v0.z, v1.z, v2.z, v3.z};
        if (ROT) {
            #pragma unroll
            for (int i = 0; i < 4; ++i) w3 += 0.5 * (vx[i] * sy[i] - vy[i] * sx[i]);
            #pragma unroll
            for (int i = 0; i < 4; ++i) w4 += 0.5 * (vx[i] * sz[i] - vz[i] * sx[i]);
            #pragma unroll
            for (int i = 0; i < 4; ++i) w5 += 0.5 * (vy[i] * sz[i] - vz[i] * sy[i]);
        }
        #pragma unroll
        for (int k = 0; k < 6; ++k) edot[k] = 0;
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[0] += vx[i] * sx[i];
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[1] += vy[i] * sy[i];
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[2] += vz[i] * sz[i];
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[3] += 0.5 * (vx[i] * sy[i] + vy[i] * sx[i]);
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[4] += 0.5 * (vx[i] * sz[i] + vz[i] * sx[i]);
        #pragma unroll
        for (int i = 0; i < 4; ++i) edot[5] += 0.5 * (vy[i] * sz[i] + vz[i] * sy[i]);
    }

    // what the previous step left for this launch: requested in the same batch as the stress
    // rows (one more dependent round trip otherwise: measured +0.45 ms at 51 M tets)
    double ddp = 0;
    unsigned ef = 0;
    if (pre & 1) ddp = ld64_v(D.ddp + e);
    if (pre & 2) ef = ld8_v(D.eflag + e);
#if DES_EB_LATE
    // stress and strain are requested only now (volatile: the order is kept), when the 64
    // registers of the node records are free again; the rows were prefetched into L2
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
        s[i] = ld64_v(D.stress + size_t(i) * E + e);
        es[i] = ld64_v(D.strain + size_t(i) * E + e);
    }
#endif
    if (pre & 1) {
        #pragma unroll
        for (int i = 0; i < 3; ++i) s[i] += ddp;
    }
    double rdv = 0;
    if (ef) rdv = D.escale[e];          // top elements only
    const bool scaled = rdv >= 1.0;
    bool pls_stored = false;
    if (scaled) {
        if (!(RHEOL == DES_RH_EP || RHEOL == DES_RH_EVP)) pls = D.plstrain[e];
        pls = pls / rdv;
        #pragma unroll
        for (int i = 0; i < 6; ++i) { s[i] = s[i] / rdv; es[i] = es[i] / rdv; }
    }
    if (ROT) {
        jaumann3(s, dt_rot, w3, w4, w5);
        jaumann3(es, dt_rot, w3, w4, w5);
    }
    if (AVG) {
        // s, es: stress and strain as the previous step left them (after rotate_stress);
        // delta_plstrain[e] still holds that step's value (it is rewritten below)
        const double dpl_prev = D.delta_plstrain[e];
        if (avg_first) {
            #pragma unroll
            for (int i = 0; i < 6; ++i) {
                D.avg_stress[size_t(i) * E + e] = s[i];
                D.avg_strain0[size_t(i) * E + e] = es[i];
            }
            D.avg_dpl[e] = dpl_prev;
        } else {
            double a[6];
            #pragma unroll
            for (int i = 0; i < 6; ++i) a[i] = D.avg_stress[size_t(i) * E + e];
            const double ad = D.avg_dpl[e];
            #pragma unroll
            for (int i = 0; i < 6; ++i) D.avg_stress[size_t(i) * E + e] = a[i] + s[i];
            D.avg_dpl[e] = ad + dpl_prev;
        }
    }
    const double old_s = trace3(s);

    // anti-mesh-locking correction, rheology.cxx:786-793
    {
        const double div = trace3(edot);
        for (int i = 0; i < 3; ++i) edot[i] += (edvoldt - div) / 3;
    }
    double de[6];
    for (int i = 0; i < 6; ++i) es[i] += edot[i] * dt;
    for (int i = 0; i < 6; ++i) de[i] = edot[i] * dt;
    #pragma unroll
    for (int i = 0; i < 6; ++i) D.strain_rate[size_t(i) * E + e] = edot[i];
    #pragma unroll
    for (int i = 0; i < 6; ++i) D.strain[size_t(i) * E + e] = es[i];

    const double bulkm = harm_mean(P.bulk_modulus, mk.cnt, nmat);
    const double shearm = harm_mean(P.shear_modulus, mk.cnt, nmat);
    double dpl = 0.;

    switch (RHEOL) {
    case DES_RH_ELASTIC:
        rh_elastic(bulkm, shearm, de, s);
        break;
    case DES_RH_MAXWELL: {
        const double visc = mat_visc(P, nmat, mk, T0, T1, T2, T3, s, edot);
        D.viscosity[e] = visc;
        const double dv = vol / vol_old - 1;
        rh_maxwell(bulkm, shearm, visc, dt, dv, de, s);
        break;
    }
    case DES_RH_EP: {
        double depls = 0, amc, anphi, anpsi, hardn, ten_max;
        plastic_props(P, nmat, mk, pls, amc, anphi, anpsi, hardn, ten_max);
        rh_elasto_plastic(bulkm, shearm, amc, anphi, anpsi, hardn, ten_max, de, depls, s);
        D.plstrain[e] = pls + depls;
        pls_stored = true;
        dpl = depls;
        break;
    }
    case DES_RH_EVP: {
        double depls = 0;
        const double visc = mat_visc(P, nmat, mk, T0, T1, T2, T3, s, edot);
        D.viscosity[e] = visc;
        const double dv = vol / vol_old - 1;
        double sv[6];
        for (int i = 0; i < 6; ++i) sv[i] = s[i];
        rh_maxwell(bulkm, shearm, visc, dt, dv, de, sv);
        const double svII = second_invariant2(sv);

        double amc, anphi, anpsi, hardn, ten_max;
        plastic_props(P, nmat, mk, pls, amc, anphi, anpsi, hardn, ten_max);
        double sp[6];
        for (int i = 0; i < 6; ++i) sp[i] = s[i];
        rh_elasto_plastic(bulkm, shearm, amc, anphi, anpsi, hardn, ten_max, de, depls, sp);
        const double spII = second_invariant2(sp);
        if (svII < spII) {
            for (int i = 0; i < 6; ++i) s[i] = sv[i];
        }
        else {
            for (int i = 0; i < 6; ++i) s[i] = sp[i];
            D.plstrain[e] = pls + depls;
            pls_stored = true;
            dpl = depls;
        }
        break;
    }
    default: break;
    }
    if (scaled && !pls_stored) D.plstrain[e] = pls;     // the division alone (top elements)
    D.delta_plstrain[e] = dpl;
    #pragma unroll
    for (int i = 0; i < 6; ++i) D.stress[size_t(i) * E + e] = s[i];
    if (P.is_using_mixed_stress) {
        const double dp = trace3(s) - old_s;
        D.dpressure[e] = dp;
        D.etmp[e] = dp * vol;
    }
}

// ---------------------------------------------------------------------------------------
// direct node gathers with wide loads (one thread per node, entries walked in ascending
// element order, 4 entries' loads in flight per trip).  Per entry: N-A reads one 32-byte
// sector {tr(edot)*vol, m, tm, vol} + one double; N-C reads one 32-byte sector {fx,fy,fz,-}
// when DES_FC_STRIDE == 16.  The L1TEX wavefront count per entry -- what bounded the v1
// gathers (l1tex 94 % busy, DRAM 28 %) -- drops from 5 to 2 (N-A) and from 3 to 1 (N-C).
// ---------------------------------------------------------------------------------------
// (min-blocks launch bounds of 5/6 for the node gathers were measured: 48/40 registers with
// spills, N-A 0.16 -> 0.38 ms at 6 M tets -- dropped)
template <bool MASS>
__global__ void __launch_bounds__(256)
k3_node_a(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt, int head_flags)
{
    if (D.sc) dt = D.sc[0];
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int beg = D.sup_idx[n], end = D.sup_idx[n + 1];
    double vn = 0, m = 0, tm = 0, tdot = 0, dvol = 0;
    // the support list is padded to a multiple of four entries per node (sentinel -1), so a
    // trip reads its four entries with ONE 128-bit load
    const int4 *ents = reinterpret_cast<const int4 *>(D.sup_ent);
    for (int k = beg >> 2; k < (end >> 2); ++k) {
        const int4 q = ents[k];
        const int ent[4] = {q.x, q.y, q.z, q.w};
        double lo[4]; double4 hi[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (ent[j] >= 0) {
                lo[j] = D.ca[ent[j]];
                hi[j] = ld256_nc(D.ca_hi + size_t(ent[j] >> 2) * 4);
            }
        }
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (ent[j] >= 0) {
                tdot += lo[j]; dvol += hi[j].x;
                if (MASS) { m += hi[j].y; tm += hi[j].z; vn += hi[j].w; }
            }
        }
    }
    if (MASS) {
        D.volume_n[n] = vn;
        D.mass[n] = m;
        if (!P.has_thermal_diffusion) tm = 0;   // geometry.cxx:1836-1837,1859-1860
        D.tmass[n] = tm;
    }
    else {
        vn = D.volume_n[n];
        tm = D.tmass[n];
    }
    if (head_flags & 1) {
        double T;
        if (D.bcflag[n] & BOUNDZ1) T = P.surface_temperature;
        else T = D.xT[n].w - dt * tdot / tm;
        D.xT[n].w = T;
    }
    if (head_flags & 4) D.ntmp[n] = dvol / vn;
}

__global__ void __launch_bounds__(256)
k3_node_c(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt, int bc_z1, int mode)
{
    if (D.sc) dt = D.sc[0];
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int N = D.nnode;
    const int beg = D.sup_idx[n], end = D.sup_idx[n + 1];
    double f[3] = {0, 0, 0};
    const int4 *ents = reinterpret_cast<const int4 *>(D.sup_ent);
    for (int k = beg >> 2; k < (end >> 2); ++k) {
        const int4 q = ents[k];
        const int ent[4] = {q.x, q.y, q.z, q.w};
#if DES_FC_STRIDE == 16
        double4 t[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) if (ent[j] >= 0) t[j] = ld256_nc(D.fc + DES_FC_OFF(ent[j]));
        #pragma unroll
        for (int j = 0; j < 4; ++j) if (ent[j] >= 0) { f[0] -= t[j].x; f[1] -= t[j].y; f[2] -= t[j].z; }
#else
        // a 24-byte entry {fx,fy,fz} of the 96-byte element record always lies in exactly two
        // aligned 16-byte chunks: two 128-bit loads per entry instead of three 64-bit ones
        double2 A[4], B[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (ent[j] >= 0) {
                const int l = ent[j] & 3;
                const double2 *c2 = reinterpret_cast<const double2 *>(D.fc + size_t(ent[j] >> 2) * 12) + ((3 * l) >> 1);
                A[j] = __ldg(c2); B[j] = __ldg(c2 + 1);
            }
        }
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (ent[j] >= 0) {
                const bool odd = ent[j] & 1;
                f[0] -= odd ? A[j].y : A[j].x;
                f[1] -= odd ? B[j].x : A[j].y;
                f[2] -= odd ? B[j].y : B[j].x;
            }
        }
#endif
    }
    const unsigned flag = D.bcflag[n];
    double4 x = ld256(D.xT + n);
    const double4 v4 = ld256(D.vv + n);
    double v[3] = {v4.x, v4.y, v4.z};
    if (flag & P.sbc_mask) {
        const int row = D.sbc_row[n];
        if (row >= 0) {
            for (int q = D.sbc_idx[row]; q < D.sbc_idx[row + 1]; ++q) {
                const int se = D.sbc_ent[q];
                const double *tr = D.sbc_ftr + size_t(se / 3) * 9 + (se % 3) * 3;
                f[0] -= tr[0]; f[1] -= tr[1]; f[2] -= tr[2];
            }
        }
    }
    if (P.has_elastic_foundation && (flag & BOUNDZ0))
        f[2] -= P.elastic_foundation_constant * (x.z - D.coord0[n].z);
    if (flag & P.nbc_mask) nbc_node(D, n, f);
    const double crit = (P.damping_option == 4) ? 2.0 * sqrt(D.mass[n] * D.ymass[n]) : 0.0;
    #pragma unroll
    for (int j = 0; j < 3; ++j) {
        f[j] = damp(P, f[j], v[j], crit);
        D.force[size_t(j) * N + n] = f[j];
    }
    if (mode & 2) {
        const double mass = D.mass[n];
        #pragma unroll
        for (int j = 0; j < 3; ++j) v[j] += dt * f[j] / mass;
    }
    if ((mode & 4) && (flag & BOUND_ANY)) vbc_node(P, flag, v, bc_z1, D.edge_slot, D.edge_vec);
    if (mode & 6) st256(D.vv + n, make_double4(v[0], v[1], v[2], 0.0));
    if (mode & 8) {
        x.x += v[0] * dt; x.y += v[1] * dt; x.z += v[2] * dt;
        st256(D.xT + n, x);
    }
}

// ---------------------------------------------------------------------------------------
// ghost-node exchange buffers (multi-GPU): entry j of a list <-> node idx[j]; AoS per entry
// so that each neighbour's entries are one contiguous message.
//   phase 0 (after N-A): {temperature, dvoldt}            2 doubles
//   phase 1 (after N-B): {dp_nd}                          1 double
//   phase 2 (after N-C): {x, y, z, vx, vy, vz}            6 doubles
//   phase 3 (after surface diffusion): {z, dh}            2 doubles
// ---------------------------------------------------------------------------------------
__host__ __device__ inline int halo_width(int phase) { return phase == 0 ? 2 : (phase == 1 ? 1 : (phase == 2 ? 6 : 2)); }

__global__ void k_top_of_node(int ntop, const int *top_nodes, int *top_of_node)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ntop) top_of_node[top_nodes[i]] = i;
}

__global__ void k_halo_pack(int phase, int n, const int *__restrict__ idx, const __grid_constant__ Dev D,
                            const int *__restrict__ top_of_node, double *__restrict__ buf)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int node = idx[j];
    if (phase == 0) { buf[2 * j] = D.xT[node].w; buf[2 * j + 1] = D.ntmp[node]; }
    else if (phase == 1) buf[j] = D.ntmp[node];
    else if (phase == 2) {
        const double4 x = D.xT[node], v = D.vv[node];
        double *b = buf + size_t(j) * 6;
        b[0] = x.x; b[1] = x.y; b[2] = x.z; b[3] = v.x; b[4] = v.y; b[5] = v.z;
    }
    else {
        const int t = top_of_node[node];
        buf[2 * j] = D.xT[node].z;
        buf[2 * j + 1] = (t >= 0) ? D.dh[t] : 0.0;
    }
}

__global__ void k_halo_unpack(int phase, int n, const int *__restrict__ idx, const __grid_constant__ Dev D,
                              const int *__restrict__ top_of_node, const double *__restrict__ buf)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int node = idx[j];
    if (phase == 0) { D.xT[node].w = buf[2 * j]; D.ntmp[node] = buf[2 * j + 1]; }
    else if (phase == 1) D.ntmp[node] = buf[j];
    else if (phase == 2) {
        const double *b = buf + size_t(j) * 6;
        double4 x = D.xT[node];
        x.x = b[0]; x.y = b[1]; x.z = b[2];
        D.xT[node] = x;
        D.vv[node] = make_double4(b[3], b[4], b[5], 0.0);
    }
    else {
        D.xT[node].z = buf[2 * j];
        const int t = top_of_node[node];
        if (t >= 0) { D.dh[t] = buf[2 * j + 1]; D.dhacc[node] += buf[2 * j + 1]; }     // the owner's dhacc += dh
    }
}
