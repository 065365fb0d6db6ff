// C-ABI of the topology builders (include/des_gpu.h, "topology builders" block).
#pragma once
#include "des_topology.cuh"

namespace {
struct TopoBuf {    // device scratch freed on scope exit
    std::vector<void *> p;
    ~TopoBuf() { for (void *q : p) cudaFree(q); }
    template <class T> cudaError_t get(T **out, size_t n)
    {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) p.push_back(q);
        *out = static_cast<T *>(q);
        return e;
    }
};
#define TCK(call) do { if ((call) != cudaSuccess) { cudaGetLastError(); return DES_E_CUDA; } } while (0)
}

extern "C" {

// create_support (mesh.cxx:3287-3329) on device `device`: same outputs as des_host_build_support
int des_gpu_build_support(int device, int nnode, int nelem, const int *connectivity,
                          int *idx, int *arr, int *lidx)
{
    if (!connectivity || !idx || !arr || !lidx || nnode <= 0 || nelem <= 0) return DES_E_ARG;
    if (size_t(nelem) * 4 >= (size_t(1) << 31)) return DES_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return DES_E_NODEVICE; }
    const size_t E = nelem, N = nnode;
    TopoBuf buf;
    int *d_conn, *d_cnt, *d_idx, *d_packed, *d_arr, *d_lidx, *d_bad;
    TCK(buf.get(&d_conn, E * 4)); TCK(buf.get(&d_cnt, N)); TCK(buf.get(&d_idx, N + 1));
    TCK(buf.get(&d_packed, E * 4)); TCK(buf.get(&d_arr, E * 4)); TCK(buf.get(&d_lidx, E * 4));
    TCK(buf.get(&d_bad, 1));
    TCK(cudaMemcpy(d_conn, connectivity, E * 4 * sizeof(int), cudaMemcpyHostToDevice));
    TCK(cudaMemset(d_cnt, 0, N * sizeof(int)));
    TCK(cudaMemset(d_idx, 0, sizeof(int)));
    TCK(cudaMemset(d_bad, 0, sizeof(int)));
    k_topo_count<<<DES_GRID(nelem, 256), 256>>>(nelem, nnode, d_conn, d_cnt, d_bad);
    int bad = 0;
    TCK(cudaMemcpy(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) return DES_E_ARG;      // connectivity entry out of range
    size_t tmp_bytes = 0;
    TCK(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, d_cnt, d_idx + 1, nnode));
    char *d_tmp; TCK(buf.get(&d_tmp, tmp_bytes));
    TCK(cub::DeviceScan::InclusiveSum(d_tmp, tmp_bytes, d_cnt, d_idx + 1, nnode));
    TCK(cudaMemset(d_cnt, 0, N * sizeof(int)));     // reused as the fill cursor
    k_topo_fill<<<DES_GRID(nelem, 256), 256>>>(nelem, d_conn, d_idx, d_cnt, d_packed);
    k_topo_sort<<<DES_GRID(nnode, 128), 128>>>(nnode, d_idx, d_packed, d_arr, d_lidx);
    TCK(cudaGetLastError());
    TCK(cudaMemcpy(idx, d_idx, (N + 1) * sizeof(int), cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(arr, d_arr, E * 4 * sizeof(int), cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(lidx, d_lidx, E * 4 * sizeof(int), cudaMemcpyDeviceToHost));
    return DES_OK;
}

// the search of create_boundary_facets (mesh.cxx:3176-3225) through the support lists
int des_gpu_find_boundary_facets(int device, int nnode, int nelem, int nseg, const int *connectivity,
                                 const int *segment, const int *segflag, const unsigned *bcflag,
                                 const int *support_idx, const int *support_arr,
                                 int *seg_elem, int *seg_facet)
{
    if (!connectivity || !segment || !segflag || !bcflag || !support_idx || !support_arr || !seg_elem || !seg_facet)
        return DES_E_ARG;
    if (nnode <= 0 || nelem <= 0 || nseg < 0) return DES_E_ARG;
    if (nseg == 0) return DES_OK;
    for (int s = 0; s < 3 * nseg; ++s)
        if (segment[s] < 0 || segment[s] >= nnode) return DES_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return DES_E_NODEVICE; }
    const size_t E = nelem, N = nnode, S = nseg;
    TopoBuf buf;
    int *d_conn, *d_seg, *d_flag, *d_idx, *d_arr, *d_e, *d_f; unsigned *d_bc;
    TCK(buf.get(&d_conn, E * 4)); TCK(buf.get(&d_seg, S * 3)); TCK(buf.get(&d_flag, S));
    TCK(buf.get(&d_bc, N)); TCK(buf.get(&d_idx, N + 1)); TCK(buf.get(&d_arr, E * 4));
    TCK(buf.get(&d_e, S)); TCK(buf.get(&d_f, S));
    TCK(cudaMemcpy(d_conn, connectivity, E * 4 * sizeof(int), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(d_seg, segment, S * 3 * sizeof(int), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(d_flag, segflag, S * sizeof(int), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(d_bc, bcflag, N * sizeof(unsigned), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(d_idx, support_idx, (N + 1) * sizeof(int), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(d_arr, support_arr, E * 4 * sizeof(int), cudaMemcpyHostToDevice));
    k_topo_facets<<<DES_GRID(nseg, 128), 128>>>(nseg, nelem, d_seg, d_flag, d_conn, d_bc, d_idx, d_arr, d_e, d_f);
    TCK(cudaGetLastError());
    TCK(cudaMemcpy(seg_elem, d_e, S * sizeof(int), cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(seg_facet, d_f, S * sizeof(int), cudaMemcpyDeviceToHost));
    return DES_OK;
}

// create_boundary_flags2 (mesh.cxx:2824-2835): bcflag must be zeroed by the caller, as
// create_boundary_flags does (:2844)
int des_host_boundary_flags(int nnode, int nseg, const int *segment, const int *segflag, unsigned *bcflag)
{
    if (!segment || !segflag || !bcflag) return DES_E_ARG;
    for (int i = 0; i < nseg; ++i)
        for (int j = 0; j < 3; ++j) {
            const int n = segment[size_t(j) * nseg + i];
            if (n < 0 || n >= nnode) return DES_E_ARG;
            bcflag[n] |= unsigned(segflag[i]);
        }
    return DES_OK;
}

// create_boundary_nodes (mesh.cxx:2854-2880): call with bnodes == NULL for the counts, then
// with bnodes[j] pointing at nbnodes[j] ints
int des_host_boundary_nodes(int nnode, const unsigned *bcflag, int nbnodes[DES_NBDRY], int *const bnodes[DES_NBDRY])
{
    if (!bcflag || !nbnodes) return DES_E_ARG;
    int cnt[DES_NBDRY] = {0};
    for (int i = 0; i < nnode; ++i)
        for (int j = 0; j < DES_NBDRY; ++j)
            if (bcflag[i] & (1u << j)) {
                if (bnodes && bnodes[j]) bnodes[j][cnt[j]] = i;
                cnt[j]++;
            }
    for (int j = 0; j < DES_NBDRY; ++j) nbnodes[j] = cnt[j];
    return DES_OK;
}

// Second half of create_boundary_facets (mesh.cxx:3205-3262): bucket the per-segment hits by
// boundary in segment order (the reference's single-thread order), std::sort each bucket by
// element id with the reference's comparator, gather connectivity_surface from the z1 bucket.
// pairs: [2*nseg] ints, bucket k starts at pairs + 2*offset[k] as (elem, facet) pairs;
// conn_surface: conn_t(etop) SoA [4*etop], 3 columns written.  Returns DES_E_STATE if a boundary
// segment was on no element (the reference dies with EXIT_INTERNAL_UNREACHABLE there).
int des_host_bucket_facets(int nelem, int nseg, const int *segflag, const int *seg_elem, const int *seg_facet,
                           const int *connectivity, int nbfacets[DES_NBDRY], int offset[DES_NBDRY + 1],
                           int *pairs, int *conn_surface)
{
    if (!segflag || !seg_elem || !seg_facet || !nbfacets || !offset || !pairs) return DES_E_ARG;
    typedef std::pair<int, int> int_pair;
    std::vector<int_pair> b[DES_NBDRY];
    for (int i = 0; i < nseg; ++i) {
        const unsigned flag = unsigned(segflag[i]);
        if ((flag & BOUND_ANY) == 0) continue;
        if (seg_elem[i] < 0) return DES_E_STATE;
        for (int k = 0; k < DES_NBDRY; ++k)
            if (flag == (1u << k)) { b[k].emplace_back(seg_elem[i], seg_facet[i]); break; }
    }
    int o = 0;
    for (int k = 0; k < DES_NBDRY; ++k) {
        std::sort(b[k].begin(), b[k].end(), [](const int_pair &x, const int_pair &y) { return x.first < y.first; });
        offset[k] = o; nbfacets[k] = int(b[k].size());
        for (const int_pair &q : b[k]) { pairs[2 * o] = q.first; pairs[2 * o + 1] = q.second; ++o; }
    }
    offset[DES_NBDRY] = o;
    if (conn_surface && connectivity) {
        static const int nof[4][3] = {{1, 2, 3}, {0, 3, 2}, {0, 1, 3}, {0, 2, 1}};      // constants.hpp:75-79
        const int etop = nbfacets[5];
        for (int i = 0; i < etop; ++i) {
            const int e = pairs[2 * (offset[5] + i)], f = pairs[2 * (offset[5] + i) + 1];
            for (int j = 0; j < 3; ++j)
                conn_surface[size_t(j) * etop + i] = connectivity[size_t(nof[f][j]) * nelem + e];
        }
    }
    return DES_OK;
}

// create_top_elems (mesh.cxx:2882-2931) and create_surface_info's integer tables (:3031-3098,
// 2937-2964).  coord_x = first row of coord ([nnode]).  ntop = nbnodes[z1], etop = nbfacets[z1].
// Outputs: top_nodes [ntop] (z1 nodes sorted by x with the reference's std::sort), top_elems
// (room for max_top_elems ints; *ntop_elems is set; DES_E_ARG if it does not fit),
// elem_and_nodes [3*etop] SoA, support_surf idx [ntop+1] / arr [3*etop].
int des_host_surface_topology(int nnode, int ntop, const int *bnodes_z1, const double *coord_x,
                              const int *support_idx, const int *support_arr,
                              int etop, const int *conn_surface,
                              int *top_nodes, int max_top_elems, int *ntop_elems, int *top_elems,
                              int *elem_and_nodes, int *ssup_idx, int *ssup_arr)
{
    if (!bnodes_z1 || !coord_x || !support_idx || !support_arr || !top_nodes || !ntop_elems || !top_elems)
        return DES_E_ARG;
    // mesh.cxx:3037-3056: sort an index vector by the x of the z1 nodes
    std::vector<int> top_ind(ntop);
    std::vector<double> top_x(ntop);
    for (int i = 0; i < ntop; ++i) { top_x[i] = coord_x[bnodes_z1[i]]; top_ind[i] = i; }
    std::sort(top_ind.begin(), top_ind.end(), [&](const int &a, const int &b) { return top_x[a] < top_x[b]; });
    for (int i = 0; i < ntop; ++i) top_nodes[i] = bnodes_z1[top_ind[i]];
    // mesh.cxx:2908-2925: the set of elements around the top nodes, ascending
    std::set<int> elem_set;
    for (int i = 0; i < ntop; ++i) {
        const int n = top_nodes[i];
        for (int k = support_idx[n]; k < support_idx[n + 1]; ++k) elem_set.insert(support_arr[k]);
    }
    if (int(elem_set.size()) > max_top_elems) return DES_E_ARG;
    *ntop_elems = int(elem_set.size());
    { int i = 0; for (int e : elem_set) top_elems[i++] = e; }
    if (etop > 0 && conn_surface && elem_and_nodes && ssup_idx && ssup_arr) {
        std::vector<int> arctop(nnode, -1);
        for (int i = 0; i < ntop; ++i) arctop[top_nodes[i]] = i;
        for (int i = 0; i < etop; ++i)
            for (int k = 0; k < 3; ++k)
                elem_and_nodes[size_t(k) * etop + i] = arctop[conn_surface[size_t(k) * etop + i]];
        for (int n = 0; n <= ntop; ++n) ssup_idx[n] = 0;
        for (int i = 0; i < etop; ++i)
            for (int k = 0; k < 3; ++k) ssup_idx[elem_and_nodes[size_t(k) * etop + i] + 1]++;
        for (int n = 1; n <= ntop; ++n) ssup_idx[n] += ssup_idx[n - 1];
        std::vector<int> cursor(ssup_idx, ssup_idx + ntop);
        for (int i = 0; i < etop; ++i)
            for (int k = 0; k < 3; ++k) ssup_arr[cursor[elem_and_nodes[size_t(k) * etop + i]]++] = i;
    }
    return DES_OK;
}

} // extern "C"
#undef TCK
