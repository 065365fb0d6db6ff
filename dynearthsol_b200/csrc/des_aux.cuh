// Host touch points of the step loop moved onto the device:
// the reductions of bad_mesh_quality (remeshing.cxx:2752-2866, geometry.cxx:1873-1927) and
// the accumulators of Output::average_fields (output.cxx:312-352).  Both used to cost a
// PCIe download of coord/volume (every quality_check_step_interval steps) or of stress and
// delta_plstrain (EVERY step when sim.is_outputting_averaged_fields, the default).
#pragma once

// ---------------------------------------------------------------------------------------
// bad_mesh_quality
// ---------------------------------------------------------------------------------------
// One partial per block: the worst element (quality, caller's element id) in the reference's
// scan order -- `quality < q` strictly, so among equal minima the LOWEST id wins
// (geometry.cxx:1919-1925) -- and the lowest id with volume < smallest_vol
// (remeshing.cxx:2766-2775, first hit returns).
struct QPartial { double q; int worst; int small; };

__device__ __forceinline__ bool q_better(double q, int e, double bq, int be)
{
    return q < bq || (q == bq && e < be);
}

__global__ void __launch_bounds__(256)
k_quality_elems(const __grid_constant__ Dev D, const int *perm, double smallest_vol, double normf,
                QPartial *partials)
{
    double bq = 1.0;        // worst_elem_quality starts from q = 1, worst_elem = 0
    int be = 0, small = INT_MAX;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < D.nelem; e += gridDim.x * blockDim.x) {
        const int ce = perm ? perm[e] : e;
        const int4 c = D.conn[e];
        const double4 a = D.xT[c.x], b = D.xT[c.y], cc = D.xT[c.z], d = D.xT[c.w];
        const double vol = D.volume[e];
        if (vol < smallest_vol) small = min(small, ce);
        // elem_quality, geometry.cxx:1893-1899
        const double area_sum = tri_area(a, b, cc) + tri_area(a, b, d) + tri_area(cc, d, a) + tri_area(cc, d, b);
        const double quality = normf * vol * vol / (area_sum * area_sum * area_sum);
        if (q_better(quality, ce, bq, be)) { bq = quality; be = ce; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double oq = __shfl_xor_sync(0xffffffffu, bq, o);
        const int oe = __shfl_xor_sync(0xffffffffu, be, o);
        const int os = __shfl_xor_sync(0xffffffffu, small, o);
        if (q_better(oq, oe, bq, be)) { bq = oq; be = oe; }
        small = min(small, os);
    }
    __shared__ QPartial sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = QPartial{bq, be, small};
    __syncthreads();
    if (threadIdx.x == 0) {
        QPartial r = sh[0];
        for (int i = 1; i < (blockDim.x >> 5); ++i) {
            if (q_better(sh[i].q, sh[i].worst, r.q, r.worst)) { r.q = sh[i].q; r.worst = sh[i].worst; }
            r.small = min(r.small, sh[i].small);
        }
        partials[blockIdx.x] = r;
    }
}

// Boundary-node distortion checks, remeshing.cxx:2778-2838: out[0] = lowest bottom node with
// |z - bottom| > dist, out[1] = lowest side node (remeshing_option 13; the reference's
// x0 / x1 / y0 / y1 else-if chain) off its plane by more than dist.  out[] preset to INT_MAX.
// (ids reported are the caller's: nperm[internal] when the nodes are stored in another order)
__global__ void k_quality_nodes(const __grid_constant__ Dev D, double bottom, double dist, double xlength,
                                double ylength, int check_bottom, int check_sides, const int *nperm, int *out)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    const int cn = nperm ? nperm[n] : n;
    const unsigned flag = D.bcflag[n];
    if (!(flag & (BOUNDX0 | BOUNDX1 | BOUNDY0 | BOUNDY1 | BOUNDZ0))) return;
    const double4 x = D.xT[n];
    if (check_bottom && (flag & BOUNDZ0) && fabs(x.z - bottom) > dist) atomicMin(out, cn);
    if (check_sides) {
        bool bad = false;
        if (flag & BOUNDX0) bad = fabs(x.x) > dist;
        else if (flag & BOUNDX1) bad = fabs(x.x - xlength) > dist;
        else if (flag & BOUNDY0) bad = fabs(x.y) > dist;
        else if (flag & BOUNDY1) bad = fabs(x.y - ylength) > dist;
        if (bad) atomicMin(out + 1, cn);
    }
}

// ---------------------------------------------------------------------------------------
// Output::average_fields
// ---------------------------------------------------------------------------------------
// acc[i] += src[i], the reference's plain running sum (output.cxx:343-350): same operation
// order per entry, so the accumulators match the host loop bit for bit.
__global__ void k_avg_accumulate(size_t n, const double *__restrict__ src, double *__restrict__ acc)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) acc[i] += src[i];
}

// ---------------------------------------------------------------------------------------
// isostasy_adjustment's velocity constraint, dynearthsol.cxx:522-533: displacement is vertical
// only; without a Winkler foundation the bottom is held
// ---------------------------------------------------------------------------------------
__global__ void k_isostasy_velocity(const __grid_constant__ Dev D, int hold_bottom)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    double4 v = D.vv[n];
    v.x = 0; v.y = 0;
    if (hold_bottom && (D.bcflag[n] & BOUNDZ0)) v.z = 0;
    D.vv[n] = v;
}

// ---------------------------------------------------------------------------------------
// ymass for Rayleigh damping (control.damping_option 4): compute_mass's fifth column,
// geometry.cxx:1833 (element: Young's modulus / 4, material only) and :1850-1866 (nodal sum in
// support order).  Two small kernels run wherever the mass is recomputed; off the hot path.
// ---------------------------------------------------------------------------------------
__global__ void k_ymass_elem(const __grid_constant__ DevParams P, const __grid_constant__ Dev D, double *ym)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    Markers mk;
    load_markers(D.elemmarkers, e, P.nmat, mk);
    const double bulkm = harm_mean(P.bulk_modulus, mk.cnt, P.nmat);
    const double shearm = harm_mean(P.shear_modulus, mk.cnt, P.nmat);
    ym[e] = 9 * bulkm * shearm / (3 * bulkm + shearm) / 4;
}
__global__ void k_ymass_node(const __grid_constant__ Dev D, const double *__restrict__ ym)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= D.nnode) return;
    double acc = 0;
    for (int k = D.sup_idx[n]; k < D.sup_idx[n + 1]; ++k) {
        const int ent = D.sup_ent[k];
        if (ent < 0) break;     // padding
        acc += ym[ent >> 2];
    }
    D.ymass[n] = acc;
}

// ---------------------------------------------------------------------------------------
// check_nan, utils.hpp:323-385: number of NaN entries in the arrays the reference scans
// (volume, dpressure, viscosity, stress; temperature, tmass, force, vel, coord)
// ---------------------------------------------------------------------------------------
__global__ void k_check_nan(const __grid_constant__ Dev D, int *count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const size_t E = D.nelem, N = D.nnode;
    int c = 0;
    if (i < D.nelem) {
        c += isnan(D.volume[i]) + isnan(D.dpressure[i]) + isnan(D.viscosity[i]);
        #pragma unroll
        for (int j = 0; j < 6; ++j) c += isnan(D.stress[size_t(j) * E + i]);
    }
    if (i < D.nnode) {
        const double4 x = D.xT[i], v = D.vv[i];
        c += isnan(x.w) + isnan(D.tmass[i]);
        c += isnan(x.x) + isnan(x.y) + isnan(x.z) + isnan(v.x) + isnan(v.y) + isnan(v.z);
        #pragma unroll
        for (int j = 0; j < 3; ++j) c += isnan(D.force[size_t(j) * N + i]);
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}
