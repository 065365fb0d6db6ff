// Kernels of the fused explicit step (sm_100a, FP64).  See DESIGN.md for the schedule:
//
//   E-A  k_elem_a<TAIL,HEAD>   tail(n-1): volume, mass contributions, Jaumann rotation
//                              head(n)  : T-diffusion contributions, strain rate, tr(edot)*vol
//   N-A  k_node_a<MASS,HEAD>   CSR gather: volume_n/mass/tmass, T update, dvoldt
//   E-B  (des_kernels2.cuh: k2_elem_b, also the update_stress twin)
//   N-B  k_node_b              CSR gather: dp_nd
//   E-C  k_elem_c<MIXED>       NMD correction of the stress diagonal, element forces
//   N-C  k_node_c              CSR gather of forces, stress bcs, damping, v, vbc, x
//   S-*  surface kernels       simple_diffusion, edvacc_surf, correct_surface_element
//
// One thread per element or node.  Node data is gathered as 32-byte records
// (double4 {x,y,z,T}, {vx,vy,vz,-}); element fields are SoA so warps load 256-byte rows;
// element->node contributions are AoS per element so one CSR entry touches one or two sectors.
#pragma once
#include "des_device.cuh"

struct Dev {
    int nnode, nelem;
    // topology
    const int4 *conn;           // [E]
    const unsigned *bcflag;     // [N]
    const int *sup_idx;         // [N+1] offsets into sup_ent, each a multiple of 4
    const int *sup_ent;         // packed (elem << 2 | local node), ascending elem per node; each node's list
                                // is padded with -1 to a multiple of 4 entries (one 128-bit load per 4)
    const int *elemmarkers;     // [E*nmat]
    // node fields
    double4 *xT;                // {x, y, z, temperature}
    double4 *vv;                // {vx, vy, vz, unused}
    double *force;              // [3N] SoA (derived, host-visible)
    double *mass, *tmass, *volume_n, *ntmp;
    double *ymass;              // nodal sum of Young's moduli, only for damping option 4 (else NULL)
    double4 *coord0;            // only for the elastic foundation
    // element fields (SoA, component-major)
    double *stress, *strain, *strain_rate;      // [6E]
    double *plstrain, *delta_plstrain;
    double *volume, *volume_old;
    double *dpressure, *edvoldt, *viscosity;
    const double *radiogenic;
    // element -> node contribution records
    double *ca;                 // [E][4] {tdiff0..3}: T-diffusion contribution per local node
    double *ca_hi;              // [E][4] {tr(edot)*vol, m, tm, vol}  (both halves: one 256-bit store per element)
    double *etmp;               // [E]    dpressure*vol
    double *fc;                 // [E][4][DES_FC_STRIDE/4] element forces, local node major
    // stress boundary conditions (apply_stress_bcs)
    int nsbc_facets;
    const int *sbc_felem, *sbc_ffacet, *sbc_fbdry;  // [nsbc_facets]
    double *sbc_ftr;            // [nsbc_facets][9]
    const int *sbc_row;         // [N] row in sbc_idx or -1
    const int *sbc_idx, *sbc_ent;   // CSR over boundary nodes: packed facet*3 + l
    // prescribed tractions (apply_stress_bcs_neumann, bc.cxx:829-912): the facets of the boundaries
    // with stress_bc_types != 0 and, per node, the facets it belongs to in the reference's serial
    // order (boundary, then facet)
    int nnbc_facets;
    const int *nbc_felem, *nbc_ffacet, *nbc_fbdry;  // [nnbc_facets]
    double *nbc_ftr;            // [nnbc_facets][3]
    const int *nbc_row;         // [N] row in nbc_idx or -1
    const int *nbc_idx, *nbc_ent;
    // slant-boundary tables for apply_vbcs
    const int *edge_slot;       // [100]
    const double *edge_vec;
    // surface processes
    int ntop, etop, ntop_elems;
    const int *top_nodes, *top_felem, *top_ffacet, *ssup_idx, *ssup_arr;
    const int *elem_and_nodes;  // [3*etop] SoA
    const int *conn_surf;       // [4*etop] SoA (first 3 columns used)
    const int *top_elems;
    double *dh, *dhacc, *edvacc_surf, *s_area, *s_slope;   // [ntop],[N],[etop],[etop],[etop][3]
    unsigned long long *maxdh_bits;
    // multi-GPU: 1 for ghost nodes (values received from the owner), NULL on a single GPU
    const unsigned char *ghost;
    // Output::average_fields accumulators (des_aux.cuh), NULL until des_gpu_average_fields is used
    double *avg_stress, *avg_strain0, *avg_dpl;     // [6E], [6E], [E]
    // element range [e0, e1) the v2 element kernels (k2_elem_a, k2_elem_b, k_elem_c) work on:
    // the whole mesh, or -- multi-GPU overlap -- the elements that touch no ghost node / the rest
    int e0, e1;
    // L2 prefetch distance of the element kernels, in elements (0 = off): thread e asks L2 for
    // the streamed rows of element e + pf, so that block finds them in L2 instead of HBM
    int pf;
    // work of step n that the fused schedule leaves to the next E-B, which streams the stress
    // rows anyway (des_api.cu: resolve_pending applies it on its own before anything else
    // looks at the stress):
    //  ddp[e]     NMD_stress's diagonal correction (geometry.cxx:315-331); E-C uses it for the
    //             forces but stores the scalar instead of rewriting three stress rows
    //  eflag[e]   1 = top element (var.top_elems); escale[e] = its rdv of correct_surface_element
    //             (bc.cxx:1669-1686), whose division of stress / strain / plstrain waits for E-B
    double *ddp;
    const unsigned char *eflag;
    double *escale;
    // bit 0: compute_volume produced a non-positive or NaN volume (the next des_gpu_compute_dt,
    // the step loop's sync point, returns DES_E_NAN; downloads still work so that the host can
    // write its error frame)
    int *status;
    // When the step runs as a CUDA Graph (small meshes, des_api.cu: step_graphed) the scalars
    // that change from step to step are read from here instead of from the baked-in launch
    // arguments: sc[0] = dt, sc[1] = dt of the step whose rotation is pending.  NULL otherwise.
    const double *sc;
};

#define DES_GRID(n, b) (((n) + (b) - 1) / (b))

// Element-force record: DES_FC_STRIDE doubles per element, DES_FC_STRIDE/4 per local node.
// 16 pads each node's {fx,fy,fz} to one aligned 32-byte sector, so the node gather is a
// single 256-bit load (one L1TEX wavefront) at the price of 32 B more per element each way.
#ifndef DES_FC_STRIDE
#define DES_FC_STRIDE 12
#endif
#define DES_FC_OFF(ent) (size_t((ent) >> 2) * DES_FC_STRIDE + ((ent) & 3) * (DES_FC_STRIDE / 4))

// ---------------------------------------------------------------------------------------
// E-A
// ---------------------------------------------------------------------------------------
// TAIL == update_mesh's compute_volume + compute_mass element loop (geometry.cxx:180-197,
// 1794-1840) and rotate_stress (fields.cxx:851-898) of the step that just ended;
// HEAD == update_temperature's element loop (fields.cxx:210-239), update_strain_rate
// (fields.cxx:414-476) and compute_dvoldt's element loop (geometry.cxx:217-225) of the
// step that begins.  Both halves see the same coordinates, velocities and volume, so the
// shape-function derivatives are computed once.
template <bool TAIL, bool HEAD>
__global__ void __launch_bounds__(256)
k_elem_a(const __grid_constant__ DevParams P, const __grid_constant__ Dev D,
         double dt_tail, int tail_flags, int head_flags)
{
    // tail_flags: 1 volume, 2 mass contributions, 4 rotation
    // head_flags: 1 T-diffusion contributions, 2 strain rate, 4 tr(edot)*vol
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const int nmat = P.nmat;
    const int E = D.nelem;
    const int4 c = D.conn[e];
    const double4 a0 = D.xT[c.x], a1 = D.xT[c.y], a2 = D.xT[c.z], a3 = D.xT[c.w];

    double vol;
    if (TAIL && (tail_flags & 1)) {
        vol = tet_volume(a0, a1, a2, a3);
        D.volume[e] = vol;
        if (!(vol > 0.0)) atomicOr(D.status, 1);
    }
    else
        vol = D.volume[e];

    double sx[4], sy[4], sz[4];
    shape_fn(a0, a1, a2, a3, vol, sx, sy, sz);
    const double4 v0 = D.vv[c.x], v1 = D.vv[c.y], v2 = D.vv[c.z], v3 = D.vv[c.w];
    const double vx[4] = {v0.x, v1.x, v2.x, v3.x};
    const double vy[4] = {v0.y, v1.y, v2.y, v3.y};
    const double vz[4] = {v0.z, v1.z, v2.z, v3.z};

    Markers mk;
    load_markers(D.elemmarkers, e, nmat, mk);
    const double rho = mat_rho(P, nmat, mk, a0.w, a1.w, a2.w, a3.w);

    double rec[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    rec[7] = vol;

    if (TAIL) {
        if (tail_flags & 2) {
            // compute_mass element loop, geometry.cxx:1794-1840
            const double bulkm = harm_mean(P.bulk_modulus, mk.cnt, nmat);
            double rho_m = P.is_quasi_static ? bulkm / (P.pseudo_speed * P.pseudo_speed) : rho;
            rec[5] = rho_m * vol / 4;
            const double cp = arith_mean(P.heat_capacity, mk.cnt, nmat);
            rec[6] = rho * cp * vol / 4;
        }
        if (tail_flags & 4) {
            // rotate_stress, fields.cxx:851-898
            double w3 = 0, w4 = 0, w5 = 0;
            for (int i = 0; i < 4; ++i) w3 += 0.5 * (vx[i] * sy[i] - vy[i] * sx[i]);
            for (int i = 0; i < 4; ++i) w4 += 0.5 * (vx[i] * sz[i] - vz[i] * sx[i]);
            for (int i = 0; i < 4; ++i) w5 += 0.5 * (vy[i] * sz[i] - vz[i] * sy[i]);
            double s[6];
            #pragma unroll
            for (int i = 0; i < 6; ++i) s[i] = D.stress[size_t(i) * E + e];
            jaumann3(s, dt_tail, w3, w4, w5);
            #pragma unroll
            for (int i = 0; i < 6; ++i) D.stress[size_t(i) * E + e] = s[i];
            #pragma unroll
            for (int i = 0; i < 6; ++i) s[i] = D.strain[size_t(i) * E + e];
            jaumann3(s, dt_tail, w3, w4, w5);
            #pragma unroll
            for (int i = 0; i < 6; ++i) D.strain[size_t(i) * E + e] = s[i];
        }
    }

    if (HEAD) {
        if (head_flags & 1) {
            // update_temperature element loop, fields.cxx:210-239
            const double k = arith_mean(P.therm_cond, mk.cnt, nmat);
            const double kv = k * vol;
            const double rh = D.radiogenic[e] * vol * rho / 4;
            const double T[4] = {a0.w, a1.w, a2.w, a3.w};
            for (int i = 0; i < 4; ++i) {
                double diffusion = 0.;
                for (int j = 0; j < 4; ++j)
                    diffusion += (sx[i] * sx[j] + sy[i] * sy[j] + sz[i] * sz[j]) * T[j];
                rec[i] = diffusion * kv - rh;
            }
        }
        if (head_flags & 6) {
        // update_strain_rate, fields.cxx:414-476
        double s[6];
        s[0] = 0; for (int i = 0; i < 4; ++i) s[0] += vx[i] * sx[i];
        s[1] = 0; for (int i = 0; i < 4; ++i) s[1] += vy[i] * sy[i];
        s[2] = 0; for (int i = 0; i < 4; ++i) s[2] += vz[i] * sz[i];
        s[3] = 0; for (int i = 0; i < 4; ++i) s[3] += 0.5 * (vx[i] * sy[i] + vy[i] * sx[i]);
        s[4] = 0; for (int i = 0; i < 4; ++i) s[4] += 0.5 * (vx[i] * sz[i] + vz[i] * sx[i]);
        s[5] = 0; for (int i = 0; i < 4; ++i) s[5] += 0.5 * (vy[i] * sz[i] + vz[i] * sy[i]);
        if (head_flags & 2) {
            #pragma unroll
            for (int i = 0; i < 6; ++i) D.strain_rate[size_t(i) * E + e] = s[i];
        }
        // compute_dvoldt element loop, geometry.cxx:217-225
        rec[4] = trace3(s) * vol;
        }
    }

    double *lo = D.ca + size_t(e) * 4, *hi = D.ca_hi + size_t(e) * 4;
    if (HEAD && (head_flags & 1)) st256(lo, make_double4(rec[0], rec[1], rec[2], rec[3]));
    if (HEAD && TAIL) st256(hi, make_double4(rec[4], rec[5], rec[6], rec[7]));
    else if (HEAD) { if (head_flags & 4) hi[0] = rec[4]; }
    else if (tail_flags & 2) { hi[1] = rec[5]; hi[2] = rec[6]; hi[3] = rec[7]; }
}

// ---------------------------------------------------------------------------------------
// N-A
// ---------------------------------------------------------------------------------------
// MASS == compute_mass node loop (geometry.cxx:1843-1864); HEAD == update_temperature
// node loop (fields.cxx:244-262) and compute_dvoldt node loop (geometry.cxx:231-238).
// Sums run over the node's support in ascending element order, as the reference does.
template <bool MASS, bool HEAD>
__global__ void __launch_bounds__(256)
k_node_a(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt, int head_flags)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int beg = D.sup_idx[n], end = D.sup_idx[n + 1];
    double vn = 0, m = 0, tm = 0, tdot = 0, dvol = 0;
    for (int k = beg; k < end; ++k) {
        const int ent = D.sup_ent[k];
        if (ent < 0) break;     // padding
        const double *lo = D.ca + size_t(ent >> 2) * 4, *r = D.ca_hi + size_t(ent >> 2) * 4;
        if (MASS) { vn += r[3]; m += r[1]; tm += r[2]; }
        if (HEAD) { tdot += lo[ent & 3]; dvol += r[0]; }
    }
    if (MASS) {
        D.volume_n[n] = vn;
        D.mass[n] = m;
        if (!P.has_thermal_diffusion) tm = 0;   // geometry.cxx:1836-1837,1859-1860
        D.tmass[n] = tm;
    }
    else if (HEAD) {
        vn = D.volume_n[n];
        tm = D.tmass[n];
    }
    if (HEAD) {
        if (head_flags & 1) {
            double T;
            if (D.bcflag[n] & BOUNDZ1) T = P.surface_temperature;
            else T = D.xT[n].w - dt * tdot / tm;
            D.xT[n].w = T;
        }
        if (head_flags & 4) D.ntmp[n] = dvol / vn;
    }
}

// ---------------------------------------------------------------------------------------
// N-B: NMD_stress node loop, geometry.cxx:301-309
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_node_b(const __grid_constant__ Dev D)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int beg = D.sup_idx[n], end = D.sup_idx[n + 1];
    const double vn = D.volume_n[n];
    double acc = 0;
    const int4 *ents = reinterpret_cast<const int4 *>(D.sup_ent);
    for (int k = beg >> 2; k < (end >> 2); ++k) {
        const int4 q = ents[k];
        const int ent[4] = {q.x, q.y, q.z, q.w};
        double v[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) if (ent[j] >= 0) v[j] = D.etmp[ent[j] >> 2];
        #pragma unroll
        for (int j = 0; j < 4; ++j) if (ent[j] >= 0) acc += v[j];
    }
    D.ntmp[n] = acc / vn;
}

// ---------------------------------------------------------------------------------------
// E-C: third loop of NMD_stress (geometry.cxx:315-331) + update_force element loop
// (fields.cxx:622-653)
// ---------------------------------------------------------------------------------------
// SINGLE: one material (nmat == 1) -- the marker counts live in registers, not local memory
// defer (MIXED only): keep the corrected diagonal in registers for the forces and store the
// scalar ddp[e] (8 B) instead of rewriting three stress rows (24 B); the next E-B, or
// resolve_pending, adds it to the stored diagonal as its first operation -- the same sum
template <bool SINGLE, bool MIXED>
__global__ void __launch_bounds__(256)
k_elem_c(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, int defer)
{
    const int e = D.e0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.e1) return;
    const int nmat = SINGLE ? 1 : P.nmat;
    const int E = D.nelem;
    // all loads first: the stores of the corrected stress diagonal below would otherwise
    // fence the coordinate gathers behind the dp_nd gathers (one more dependent round trip)
    const int4 c = D.conn[e];
    double s[6];
    #pragma unroll
    for (int i = 0; i < 6; ++i) s[i] = D.stress[size_t(i) * E + e];
    const double vol = D.volume[e];
    Markers mk;
    if (P.gravity != 0) load_markers(D.elemmarkers, e, nmat, mk);
    double dp_orig = 0, dpn0 = 0, dpn1 = 0, dpn2 = 0, dpn3 = 0;
    if (MIXED) {
        dp_orig = D.dpressure[e];
        dpn0 = D.ntmp[c.x]; dpn1 = D.ntmp[c.y]; dpn2 = D.ntmp[c.z]; dpn3 = D.ntmp[c.w];
    }
    const double4 a0 = ld256_nc(D.xT + c.x), a1 = ld256_nc(D.xT + c.y), a2 = ld256_nc(D.xT + c.z), a3 = ld256_nc(D.xT + c.w);
    if (D.pf) {
        const int ep = e + D.pf;
        if (ep < E) {
            pf_l2(D.conn + ep);
            pf_l2(D.volume + ep);
            if (SINGLE) pf_l2(D.elemmarkers + ep);
            #pragma unroll
            for (int i = 0; i < 6; ++i) pf_l2(D.stress + size_t(i) * E + ep);
            if (MIXED) pf_l2(D.dpressure + ep);
        }
    }
    if (MIXED) {
        double dp = 0;
        dp += dpn0; dp += dpn1; dp += dpn2; dp += dpn3;
        const double dp_el = dp / 4;
        const double ddp = (-dp_orig + dp_el) / 3;
        #pragma unroll
        for (int i = 0; i < 3; ++i) s[i] += ddp;
        if (defer) D.ddp[e] = ddp;
        else {
            #pragma unroll
            for (int i = 0; i < 3; ++i) D.stress[size_t(i) * E + e] = s[i];
        }
    }
    double sx[4], sy[4], sz[4];
    shape_fn(a0, a1, a2, a3, vol, sx, sy, sz);

    double buoy = 0;
    if (P.gravity != 0) {
        const double rho = mat_rho(P, nmat, mk, a0.w, a1.w, a2.w, a3.w);
        const double phi = arith_mean(P.porosity, mk.cnt, nmat);
        buoy = (rho * (1 - phi) + 1000.0 * phi) * P.gravity / 4;
    }
    double f[12];
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[i*3 + 0] = (s[0]*sx[i] + s[3]*sy[i] + s[4]*sz[i]) * vol;
        f[i*3 + 1] = (s[3]*sx[i] + s[1]*sy[i] + s[5]*sz[i]) * vol;
        f[i*3 + 2] = (s[4]*sx[i] + s[5]*sy[i] + s[2]*sz[i] + buoy) * vol;
    }
#if DES_FC_STRIDE == 16
    double *out = D.fc + size_t(e) * 16;
    #pragma unroll
    for (int i = 0; i < 4; ++i) st256(out + 4 * i, make_double4(f[i*3], f[i*3 + 1], f[i*3 + 2], 0.0));
#else
    // three 256-bit stores (the 96-byte record is 32-byte aligned): half the L1TEX wavefronts
    // of six 128-bit stores
    double *out = D.fc + size_t(e) * 12;
    #pragma unroll
    for (int i = 0; i < 3; ++i) st256(out + 4 * i, make_double4(f[4*i], f[4*i + 1], f[4*i + 2], f[4*i + 3]));
#endif
}

// ---------------------------------------------------------------------------------------
// stress-bc facet tractions: first loop of apply_stress_bcs, bc.cxx:700-780
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_sbc_facets(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double compensation_pressure)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nsbc_facets) return;
    const int nmat = P.nmat;
    const int e = D.sbc_felem[n], f = D.sbc_ffacet[n], ib = D.sbc_fbdry[n];
    const int4 c4 = D.conn[e];
    const int cn[4] = {c4.x, c4.y, c4.z, c4.w};
    const double4 p0 = D.xT[cn[c_node_of_facet[f][0]]];
    const double4 p1 = D.xT[cn[c_node_of_facet[f][1]]];
    const double4 p2 = D.xT[cn[c_node_of_facet[f][2]]];
    // normal_vector_of_facet, bc.cxx:24-40
    const double v01[3] = {p1.x - p0.x, p1.y - p0.y, p1.z - p0.z};
    const double v02[3] = {p2.x - p0.x, p2.y - p0.y, p2.z - p0.z};
    double normal[3];
    normal[0] = (v01[1] * v02[2] - v01[2] * v02[1]) / 2;
    normal[1] = (v01[2] * v02[0] - v01[0] * v02[2]) / 2;
    normal[2] = (v01[0] * v02[1] - v01[1] * v02[0]) / 2;
    const double zcenter = (p0.z + p1.z + p2.z) / 3;

    double p;
    if (ib == 4 && P.has_winkler_foundation) {
        Markers mk;
        load_markers(D.elemmarkers, e, nmat, mk);
        const double rho = mat_rho(P, nmat, mk, D.xT[cn[0]].w, D.xT[cn[1]].w, D.xT[cn[2]].w, D.xT[cn[3]].w);
        p = compensation_pressure - (rho + P.winkler_delta_rho) * P.gravity * (zcenter + P.zlength);
    }
    else if (ib == 5 && P.has_water_loading) {
        p = 0;
        if (zcenter < P.surf_base_level)
            p = P.sea_water_density * P.gravity * (P.surf_base_level - zcenter);
    }
    else {
        p = ref_pressure(P, zcenter);
        if (p < 0.0) p = 0.0;
    }
    double *tmp = D.sbc_ftr + size_t(n) * 9;
    for (int j = 0; j < 3; ++j)
        for (int d = 0; d < 3; ++d)
            tmp[j*3 + d] = p * normal[d] / 3;
}

// apply_stress_bcs_neumann's facet part, bc.cxx:844-885: what every node of facet n receives
__global__ void __launch_bounds__(128)
k_nbc_facets(const __grid_constant__ DevParams P, const __grid_constant__ Dev D)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnbc_facets) return;
    const int e = D.nbc_felem[n], f = D.nbc_ffacet[n], ib = D.nbc_fbdry[n];
    const int4 c4 = D.conn[e];
    const int cn[4] = {c4.x, c4.y, c4.z, c4.w};
    const double4 p0 = D.xT[cn[c_node_of_facet[f][0]]];
    const double4 p1 = D.xT[cn[c_node_of_facet[f][1]]];
    const double4 p2 = D.xT[cn[c_node_of_facet[f][2]]];
    // normal_vector_of_facet, bc.cxx:24-40
    const double v01[3] = {p1.x - p0.x, p1.y - p0.y, p1.z - p0.z};
    const double v02[3] = {p2.x - p0.x, p2.y - p0.y, p2.z - p0.z};
    double normal[3];
    normal[0] = (v01[1] * v02[2] - v01[2] * v02[1]) / 2;
    normal[1] = (v01[2] * v02[0] - v01[0] * v02[2]) / 2;
    normal[2] = (v01[0] * v02[1] - v01[1] * v02[0]) / 2;
    double traction[3] = {0.0, 0.0, 0.0};
    traction[P.stress_bc_types[ib] - 1] = P.stress_bc_values[ib];
    for (int d = 0; d < 3; ++d) D.nbc_ftr[size_t(n) * 3 + d] = traction[d] * normal[d] / 3;
}

// ---------------------------------------------------------------------------------------
// N-C: update_force node loop (fields.cxx:658-676), apply_stress_bcs node loops
// (bc.cxx:783-823), apply_damping (fields.cxx:483-579), update_velocity (:734-737),
// apply_vbcs (bc.cxx:396-651), update_coordinate (fields.cxx:774-778)
// ---------------------------------------------------------------------------------------
// apply_stress_bcs_neumann's node part (bc.cxx:888-900) for one node: its facets in the
// reference's order, `force += traction * normal / 3` component by component
__device__ __forceinline__ void nbc_node(const Dev &D, int n, double f[3])
{
    const int row = D.nbc_row[n];
    if (row < 0) return;
    for (int q = D.nbc_idx[row]; q < D.nbc_idx[row + 1]; ++q) {
        const double *tr = D.nbc_ftr + size_t(D.nbc_ent[q]) * 3;
        f[0] += tr[0]; f[1] += tr[1]; f[2] += tr[2];
    }
}

__global__ void __launch_bounds__(256)
k_node_c(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt, int bc_z1,
         int mode)
{
    // mode bits: 1 update_force (+stress bcs, damping), 2 update_velocity, 4 apply_vbcs,
    //            8 update_coordinate.  The fused step passes all that apply at once.
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int N = D.nnode;
    const unsigned flag = D.bcflag[n];
    double4 x = D.xT[n];
    double4 v4 = D.vv[n];
    double v[3] = {v4.x, v4.y, v4.z};
    double f[3] = {0, 0, 0};
    if (mode & 1) {
        const int beg = D.sup_idx[n], end = D.sup_idx[n + 1];
        for (int k = beg; k < end; ++k) {
            const int ent = D.sup_ent[k];
            if (ent < 0) break;     // padding
            const double *tr = D.fc + DES_FC_OFF(ent);
            f[0] -= tr[0]; f[1] -= tr[1]; f[2] -= tr[2];
        }
        if (flag & P.sbc_mask) {
            const int row = D.sbc_row[n];
            if (row >= 0) {
                for (int k = D.sbc_idx[row]; k < D.sbc_idx[row + 1]; ++k) {
                    const int ent = D.sbc_ent[k];
                    const double *tr = D.sbc_ftr + size_t(ent / 3) * 9 + (ent % 3) * 3;
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
    }
    else if (mode & 2) {
        #pragma unroll
        for (int j = 0; j < 3; ++j) f[j] = D.force[size_t(j) * N + n];
    }
    if (mode & 2) {
        const double mass = D.mass[n];
        #pragma unroll
        for (int j = 0; j < 3; ++j) v[j] += dt * f[j] / mass;
    }
    if ((mode & 4) && (flag & BOUND_ANY)) vbc_node(P, flag, v, bc_z1, D.edge_slot, D.edge_vec);
    if (mode & 6) D.vv[n] = make_double4(v[0], v[1], v[2], 0.0);
    if (mode & 8) {
        x.x += v[0] * dt; x.y += v[1] * dt; x.z += v[2] * dt;
        D.xT[n] = x;
    }
}

// ---------------------------------------------------------------------------------------
// surface processes (3D), bc.cxx:916-1112, 1655-1872
// ---------------------------------------------------------------------------------------
// simple_diffusion facet loop, bc.cxx:958-1030
__global__ void k_surf_facets(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.etop) return;
    const int e = D.top_felem[i], f = D.top_ffacet[i];
    const int4 c4 = D.conn[e];
    const int cn[4] = {c4.x, c4.y, c4.z, c4.w};
    const double4 q0 = D.xT[cn[c_node_of_facet[f][0]]];
    const double4 q1 = D.xT[cn[c_node_of_facet[f][1]]];
    const double4 q2 = D.xT[cn[c_node_of_facet[f][2]]];
    const double x01 = q1.x - q0.x, y01 = q1.y - q0.y;
    const double x02 = q2.x - q0.x, y02 = q2.y - q0.y;
    const double normal2 = x01*y02 - y01*x02;
    const double projected_area = 0.5 * normal2;
    D.s_area[i] = projected_area;
    const double iv = 1 / (2 * projected_area);
    double sdx[3], sdy[3];
    sdx[0] = iv * (q1.y - q2.y); sdx[1] = iv * (q2.y - q0.y); sdx[2] = iv * (q0.y - q1.y);
    sdy[0] = iv * (q2.x - q1.x); sdy[1] = iv * (q0.x - q2.x); sdy[2] = iv * (q1.x - q0.x);
    const double z[3] = {q0.z, q1.z, q2.z};
    for (int j = 0; j < 3; j++) {
        double slope = 0;
        for (int k = 0; k < 3; k++)
            slope += (sdx[j] * sdx[k] + sdy[j] * sdy[k]) * z[k];
        D.s_slope[size_t(i) * 3 + j] = slope * projected_area;
    }
}

// simple_diffusion node loops (bc.cxx:1048-1107) + surface_processes coordinate loop
// (bc.cxx:1772-1781) + max|dh| (bc.cxx:1811-1821)
__global__ void k_surf_nodes(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double dt)
{
    if (D.sc) dt = D.sc[0];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.ntop) return;
    const int n = D.top_nodes[i];
    double total_dx = 0., total_slope = 0.;
    for (int j = D.ssup_idx[i]; j < D.ssup_idx[i + 1]; ++j) {
        const int k = D.ssup_arr[j];
        total_dx += D.s_area[k];
        const int e = D.top_felem[k], f = D.top_ffacet[k];
        const int4 c4 = D.conn[e];
        const int cn[4] = {c4.x, c4.y, c4.z, c4.w};
        for (int m = 0; m < 3; ++m) {
            if (cn[c_node_of_facet[f][m]] == n) {
                total_slope += D.s_slope[size_t(k) * 3 + m];
                break;
            }
        }
    }
    const double conv = P.surf_diff * dt * total_slope / total_dx;
    double dh = 0.;
    dh -= conv;
    D.dh[i] = dh;
    D.xT[n].z += dh;
    // a ghost's dh is a partial sum (it is overwritten by the owner's, and dhacc then takes the
    // owner's value in k_halo_unpack): keep it out of dhacc and of max|dh|
    if (!D.ghost || !D.ghost[n]) {
        D.dhacc[n] += dh;
        atomicMax(D.maxdh_bits, (unsigned long long)__double_as_longlong(fabs(dh)));
    }
}

// edvacc_surf loop, bc.cxx:1787-1797 (after the coordinate update)
__global__ void k_surf_edvacc(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.etop) return;
    double dh_e = 0.;
    for (int j = 0; j < 3; j++) dh_e += D.dh[D.elem_and_nodes[size_t(j) * D.etop + i]];
    const double4 a = D.xT[D.conn_surf[i]];
    const double4 b = D.xT[D.conn_surf[size_t(D.etop) + i]];
    const double4 c = D.xT[D.conn_surf[size_t(2) * D.etop + i]];
    const double base = tri_area2d(a, b, c);
    D.edvacc_surf[i] += dh_e * base / 3;
}

// correct_surface_element element loop, bc.cxx:1669-1686
__global__ void k_surf_correct_elems(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.ntop_elems) return;
    const int e = D.top_elems[i];
    const int E = D.nelem;
    const int4 c = D.conn[e];
    const double nv = tet_volume(D.xT[c.x], D.xT[c.y], D.xT[c.z], D.xT[c.w]);
    const double rdv = nv / D.volume[e];
    D.volume[e] = nv;
    if (rdv < 1.0) return;
    // all 19 loads before the first store: the rows may alias as far as the compiler knows, and
    // a load/divide/store chain per row was 19 dependent DRAM round trips (ncu: 890 cycles of
    // long-scoreboard stall per issue; the top elements are isolated in every row)
    double pl = D.plstrain[e], s[6], es[6], er[6];
    #pragma unroll
    for (int j = 0; j < 6; j++) {
        s[j] = D.stress[size_t(j) * E + e];
        es[j] = D.strain[size_t(j) * E + e];
        er[j] = D.strain_rate[size_t(j) * E + e];
    }
    D.plstrain[e] = pl / rdv;
    #pragma unroll
    for (int j = 0; j < 6; j++) {
        D.stress[size_t(j) * E + e] = s[j] / rdv;
        D.strain[size_t(j) * E + e] = es[j] / rdv;
        D.strain_rate[size_t(j) * E + e] = er[j] / rdv;
    }
}

// The same loop as the fused step runs it: the new volume and rdv now, the divisions later --
// stress / strain / plstrain in the next E-B (which streams those rows; here every top element
// is alone in its 32-byte sector of each of the 19 rows: 2.8 kB of DRAM traffic per element,
// profiles/r01_v8_ncu_surface_kernels_50M.csv), strain_rate only if something reads it before
// E-B overwrites it (k_surf_apply_scale below).
__global__ void k_surf_correct_vol(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.ntop_elems) return;
    const int e = D.top_elems[i];
    const int4 c = D.conn[e];
    const double nv = tet_volume(D.xT[c.x], D.xT[c.y], D.xT[c.z], D.xT[c.w]);
    const double rdv = nv / D.volume[e];
    D.volume[e] = nv;
    D.escale[e] = rdv;
}

// the deferred divisions of k_surf_correct_vol, on their own (flush path)
__global__ void k_surf_apply_scale(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.ntop_elems) return;
    const int e = D.top_elems[i];
    const int E = D.nelem;
    const double rdv = D.escale[e];
    if (rdv < 1.0) return;
    double pl = D.plstrain[e], s[6], es[6], er[6];
    #pragma unroll
    for (int j = 0; j < 6; j++) {
        s[j] = D.stress[size_t(j) * E + e];
        es[j] = D.strain[size_t(j) * E + e];
        er[j] = D.strain_rate[size_t(j) * E + e];
    }
    D.plstrain[e] = pl / rdv;
    #pragma unroll
    for (int j = 0; j < 6; j++) {
        D.stress[size_t(j) * E + e] = s[j] / rdv;
        D.strain[size_t(j) * E + e] = es[j] / rdv;
        D.strain_rate[size_t(j) * E + e] = er[j] / rdv;
    }
}

// the deferred diagonal correction of E-C, on its own (flush path)
__global__ void k_apply_ddp(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const size_t E = D.nelem;
    const double ddp = D.ddp[e];
    #pragma unroll
    for (int i = 0; i < 3; ++i) D.stress[i * E + e] += ddp;
}

__global__ void k_set_eflag(int n, const int *top_elems, unsigned char *eflag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) eflag[top_elems[i]] = 1;
}

// correct_surface_element node loop, bc.cxx:1692-1701
__global__ void k_surf_correct_nodes(const __grid_constant__ Dev D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D.ntop) return;
    const int nt = D.top_nodes[i];
    double acc = 0.;
    for (int k = D.sup_idx[nt]; k < D.sup_idx[nt + 1]; ++k) {
        const int ent = D.sup_ent[k];
        if (ent < 0) break;     // padding
        acc += D.volume[ent >> 2];
    }
    D.volume_n[nt] = acc;
}

// ---------------------------------------------------------------------------------------
// compute_dt element loop + reductions, geometry.cxx:1511-1593
// ---------------------------------------------------------------------------------------
struct DtPartial { double minl, dt_maxwell, dt_diffusion, global_dt_min, max_vem; };

__device__ __forceinline__ double warp_min(double v)
{
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v)
{
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <bool SINGLE>
__global__ void __launch_bounds__(256)
k_dt(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, DtPartial *partials, int fresh_volume)
{
    const int nmat = SINGLE ? 1 : P.nmat;
    double minl = DBL_MAX, dtm = DBL_MAX, dtd = DBL_MAX, gmin = DBL_MAX, vem = 0.0;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < D.nelem; e += gridDim.x * blockDim.x) {
        const int4 c = D.conn[e];
        const double4 a = ld256_nc(D.xT + c.x), b = ld256_nc(D.xT + c.y), cc = ld256_nc(D.xT + c.z), d = ld256_nc(D.xT + c.w);
        const double4 v0 = ld256_nc(D.vv + c.x), v1 = ld256_nc(D.vv + c.y), v2 = ld256_nc(D.vv + c.z), v3 = ld256_nc(D.vv + c.w);
        const double weight = 1.0 / 4;
        double vxe = 0.0, vye = 0.0, vze = 0.0;
        vxe += v0.x * weight; vye += v0.y * weight; vze += v0.z * weight;
        vxe += v1.x * weight; vye += v1.y * weight; vze += v1.z * weight;
        vxe += v2.x * weight; vye += v2.y * weight; vze += v2.z * weight;
        vxe += v3.x * weight; vye += v3.y * weight; vze += v3.z * weight;
        const double max_vem = sqrt(vxe*vxe + vye*vye + vze*vze);
        vem = fmax(vem, max_vem);
        const double maxa = fmax(fmax(tri_area(a, b, cc), tri_area(a, b, d)),
                                 fmax(tri_area(cc, d, a), tri_area(cc, d, b)));
        const double vol = fresh_volume ? tet_volume(a, b, cc, d) : D.volume[e];
        const double minh = 3 * vol / maxa;
        Markers mk;
        load_markers(D.elemmarkers, e, nmat, mk);
        const double shearm = harm_mean(P.shear_modulus, mk.cnt, nmat);
        dtm = fmin(dtm, 0.5 * P.visc_min / (1e-40 + shearm));
        if (P.has_thermal_diffusion) dtd = fmin(dtd, 0.5 * minh * minh / P.therm_diff_max);
        minl = fmin(minl, minh);
        const double rho = mat_rho(P, nmat, mk, a.w, b.w, cc.w, d.w);
        gmin = fmin(gmin, minh / sqrt(shearm / rho) / 5.0);
    }
    __shared__ DtPartial sh[8];
    minl = warp_min(minl); dtm = warp_min(dtm); dtd = warp_min(dtd); gmin = warp_min(gmin);
    vem = warp_max(vem);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) sh[w] = DtPartial{minl, dtm, dtd, gmin, vem};
    __syncthreads();
    if (threadIdx.x == 0) {
        DtPartial r = sh[0];
        for (int i = 1; i < (blockDim.x >> 5); ++i) {
            r.minl = fmin(r.minl, sh[i].minl); r.dt_maxwell = fmin(r.dt_maxwell, sh[i].dt_maxwell);
            r.dt_diffusion = fmin(r.dt_diffusion, sh[i].dt_diffusion);
            r.global_dt_min = fmin(r.global_dt_min, sh[i].global_dt_min);
            r.max_vem = fmax(r.max_vem, sh[i].max_vem);
        }
        partials[blockIdx.x] = r;
    }
}

// partials[0..n) -> out[0]: the reduction finishes on the device, 40 bytes cross PCIe
__global__ void __launch_bounds__(1024)
k_dt_final(int n, const DtPartial *__restrict__ partials, DtPartial *__restrict__ out)
{
    double minl = DBL_MAX, dtm = DBL_MAX, dtd = DBL_MAX, gmin = DBL_MAX, vem = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const DtPartial p = partials[i];
        minl = fmin(minl, p.minl); dtm = fmin(dtm, p.dt_maxwell); dtd = fmin(dtd, p.dt_diffusion);
        gmin = fmin(gmin, p.global_dt_min); vem = fmax(vem, p.max_vem);
    }
    __shared__ DtPartial sh[32];
    minl = warp_min(minl); dtm = warp_min(dtm); dtd = warp_min(dtd); gmin = warp_min(gmin);
    vem = warp_max(vem);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) sh[w] = DtPartial{minl, dtm, dtd, gmin, vem};
    __syncthreads();
    if (threadIdx.x == 0) {
        DtPartial r = sh[0];
        for (int i = 1; i < (blockDim.x >> 5); ++i) {
            r.minl = fmin(r.minl, sh[i].minl); r.dt_maxwell = fmin(r.dt_maxwell, sh[i].dt_maxwell);
            r.dt_diffusion = fmin(r.dt_diffusion, sh[i].dt_diffusion);
            r.global_dt_min = fmin(r.global_dt_min, sh[i].global_dt_min);
            r.max_vem = fmax(r.max_vem, sh[i].max_vem);
        }
        out[0] = r;
    }
}

// ---------------------------------------------------------------------------------------
// layout conversion between the reference's host SoA and the device records
// ---------------------------------------------------------------------------------------
// Internal element order: elements are stored sorted by their smallest
// node index, perm[internal] = caller's element id, einv = its inverse (both NULL when the
// caller's order is kept).  Everything element-indexed crosses the C-ABI in the caller's order.
// ninv (caller's node id -> internal) is NULL when the caller's node order is kept
__global__ void k_pack_conn(int n, const int *soa, const int *perm, const int *ninv, int4 *out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const size_t o = perm ? perm[e] : e;
    int4 c = make_int4(soa[o], soa[size_t(n) + o], soa[size_t(2) * n + o], soa[size_t(3) * n + o]);
    if (ninv) c = make_int4(ninv[c.x], ninv[c.y], ninv[c.z], ninv[c.w]);
    out[e] = c;
}
// one thread per node: its support entries -> packed, into the padded list (out pre-filled with -1)
// (internal node n holds the list of the caller's node nperm[n], entries in the caller's order:
// the summation order never changes)
__global__ void k_pack_support(int nnode, const int *idx, const int *pidx, const int *arr, const int *lidx,
                               const int *einv, const int *nperm, int *out)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= nnode) return;
    const int cn = nperm ? nperm[n] : n;
    const int beg = idx[cn], end = idx[cn + 1], o = pidx[n] - beg;
    for (int k = beg; k < end; ++k) {
        const int e = einv ? einv[arr[k]] : arr[k];
        out[o + k] = (e << 2) | lidx[k];
    }
}
// rows of an element field, caller's order -> internal order: out[r][i] = in[r][perm[i]]
template <class T>
__global__ void k_rows_to_internal(int n, int rows, const T *in, const int *perm, T *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = perm[i];
    for (int r = 0; r < rows; ++r) out[size_t(r) * n + i] = in[size_t(r) * n + o];
}
// internal order -> caller's order: out[r][o] = in[r][einv[o]]
template <class T>
__global__ void k_rows_to_caller(int n, int rows, const T *in, const int *einv, T *out)
{
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    const size_t i = einv[o];
    for (int r = 0; r < rows; ++r) out[size_t(r) * n + o] = in[size_t(r) * n + i];
}
// marker counts are [E][nmat] (element-major): out[i][m] = in[perm[i]][m]
__global__ void k_markers_to_internal(int n, int nmat, const int *in, const int *perm, int *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = perm[i];
    for (int m = 0; m < nmat; ++m) out[size_t(i) * nmat + m] = in[o * nmat + m];
}
// node fields between the caller's SoA and the device records; perm[internal] = caller's node
// id (NULL: same order)
__global__ void k_pack_vec3(int n, const double *soa, const int *perm, double4 *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t c = perm ? perm[i] : i;
    double4 r = out[i];
    r.x = soa[c]; r.y = soa[size_t(n) + c]; r.z = soa[size_t(2) * n + c];
    out[i] = r;
}
__global__ void k_pack_w(int n, const double *src, const int *perm, double4 *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i].w = src[perm ? perm[i] : i];
}
__global__ void k_unpack_vec3(int n, const double4 *in, const int *perm, double *soa)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t c = perm ? perm[i] : i;
    const double4 r = in[i];
    soa[c] = r.x; soa[size_t(n) + c] = r.y; soa[size_t(2) * n + c] = r.z;
}
__global__ void k_unpack_w(int n, const double4 *in, const int *perm, double *dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dst[perm ? perm[i] : i] = in[i].w;
}
