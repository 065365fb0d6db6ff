// Topology builders: the integer tables the step gathers through,
// bit-identical to the reference's (mesh.cxx:2824-3329), without its O(nseg * nelem) facet
// search -- the reason the reference cannot initialise >= 10 M-tet meshes.
//
//   device: create_support (mesh.cxx:3287-3329)            des_gpu_build_support
//           create_boundary_facets' search (:3161-3225)    des_gpu_find_boundary_facets
//   host:   create_boundary_flags2 (:2824-2835), create_boundary_nodes (:2854-2880), the
//           per-boundary sort of create_boundary_facets (:3240-3247), connectivity_surface
//           (:3249-3262), create_top_elems (:2882-2931), create_surface_info (:3031-3098) and
//           create_support_surf (:2937-2964)               des_host_* below
// The host parts are O(nnode) or O(surface) integer loops; they use the same std::sort calls
// with the same comparators on the same input order as the reference, so ties (top nodes
// sharing an x coordinate) fall the same way.
#pragma once
#include <algorithm>
#include <set>
#include <cub/device/device_scan.cuh>

// ---- create_support on the device -------------------------------------------------------
__global__ void k_topo_count(int E, int N, const int *__restrict__ conn, int *__restrict__ cnt, int *bad)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = conn[size_t(i) * E + e];
        if (n < 0 || n >= N) { atomicExch(bad, 1); return; }
        atomicAdd(cnt + n, 1);
    }
}
// entries land in arbitrary order inside a node's range ...
__global__ void k_topo_fill(int E, const int *__restrict__ conn, const int *__restrict__ idx,
                            int *__restrict__ cursor, int *__restrict__ packed)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = conn[size_t(i) * E + e];
        packed[idx[n] + atomicAdd(cursor + n, 1)] = (e << 2) | i;
    }
}
// ... and are put in the reference's order (ascending element id; an element lists a node once)
// by one thread per node: insertion sort of a patch of ~8-40 packed entries
__global__ void k_topo_sort(int N, const int *__restrict__ idx, int *__restrict__ packed,
                            int *__restrict__ arr, int *__restrict__ lidx)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const int b = idx[n], e = idx[n + 1];
    for (int i = b + 1; i < e; ++i) {
        const int v = packed[i];
        int j = i - 1;
        while (j >= b && packed[j] > v) { packed[j + 1] = packed[j]; --j; }
        packed[j + 1] = v;
    }
    for (int i = b; i < e; ++i) { arr[i] = packed[i] >> 2; lidx[i] = packed[i] & 3; }
}

// ---- create_boundary_facets' search on the device -----------------------------------------
// One thread per segment.  The reference scans ALL elements in ascending order for the first
// (element, facet) whose sorted node triple equals the segment's and whose nodes share a flag
// with it; every candidate contains the segment's smallest node, so scanning that node's
// support (ascending element ids) visits the same candidates in the same order.
// out_elem: element id, -1 = not a boundary segment, -2 = on no element (the reference dies).
__device__ __forceinline__ void topo_sort3(int &a, int &b, int &c)
{
    if (a > b) { const int t = a; a = b; b = t; }
    if (b > c) { const int t = b; b = c; c = t; }
    if (a > b) { const int t = a; a = b; b = t; }
}
__global__ void k_topo_facets(int nseg, int E, const int *__restrict__ segment, const int *__restrict__ segflag,
                              const int *__restrict__ conn, const unsigned *__restrict__ bcflag,
                              const int *__restrict__ idx, const int *__restrict__ arr,
                              int *__restrict__ out_elem, int *__restrict__ out_facet)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    const unsigned flag = unsigned(segflag[s]);
    out_facet[s] = -1;
    if ((flag & BOUND_ANY) == 0) { out_elem[s] = -1; return; }
    int a = segment[s], b = segment[size_t(nseg) + s], c = segment[size_t(2) * nseg + s];
    topo_sort3(a, b, c);
    // `flag == (1U << k)` for some k < nbdrytypes, mesh.cxx:3205-3211
    const bool single = (flag & (flag - 1)) == 0 && flag < (1u << DES_NBDRY);
    int found = -2, ff = -1;
    for (int k = idx[a]; k < idx[a + 1] && found < 0; ++k) {
        const int e = arr[k];
        const int cn[4] = {conn[e], conn[size_t(E) + e], conn[size_t(2) * E + e], conn[size_t(3) * E + e]};
        for (int f = 0; f < 4; ++f) {
            int x = cn[c_node_of_facet[f][0]], y = cn[c_node_of_facet[f][1]], z = cn[c_node_of_facet[f][2]];
            if ((flag & bcflag[x] & bcflag[y] & bcflag[z]) == 0u) continue;
            topo_sort3(x, y, z);
            if (x == a && y == b && z == c && single) { found = e; ff = f; break; }
        }
    }
    out_elem[s] = found;
    out_facet[s] = ff;
}
