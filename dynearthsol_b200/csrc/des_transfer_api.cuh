// Kernels and C-ABI of the remesh field transfer (include/des_gpu.h, "remesh field transfer"
// block).  The per-item routines live in des_transfer.cuh.
#pragma once
#include <chrono>
#include "des_transfer.cuh"

#define TRK(call) do { if ((call) != cudaSuccess) { cudaGetLastError(); return DES_E_CUDA; } } while (0)

namespace {

__global__ void k_tr_centroids(int ncell, int npc, const int *__restrict__ conn, int nnode,
                               const double *__restrict__ coord, double *__restrict__ out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < ncell) tr::centroid_one(e, npc, ncell, conn, nnode, coord, out);
}
__global__ void k_tr_cell_count(tr::Grid g, int *__restrict__ cnt)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < g.npts) atomicAdd(cnt + tr::cell_of_point(g, i), 1);
}
// the order of the points inside one cell is whatever the atomics give: the searches sort by
// (distance, index), so no result depends on it
__global__ void k_tr_cell_fill(tr::Grid g, const int *__restrict__ start, int *__restrict__ cursor,
                               int *__restrict__ pts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.npts) return;
    const size_t c = tr::cell_of_point(g, i);
    pts[start[c] + atomicAdd(cursor + c, 1)] = i;
}
__global__ void k_tr_coeff(int ncell, int dim, const int *__restrict__ conn, int nnode,
                           const double *__restrict__ coord, const double *__restrict__ size,
                           double *__restrict__ coeff)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < ncell) tr::coeff_one(e, dim, ncell, conn, nnode, coord, size, coeff);
}
__global__ void k_tr_nearest(tr::Grid g, int nq, const double *__restrict__ q, int *__restrict__ idx,
                             int *__restrict__ is_changed, int *__restrict__ flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    tr::nearest_one(i, g, nq, q, idx, is_changed);
    flag[i] = is_changed[i] > 0;
}
// compaction in ascending order: idx_changed[i] = rank among the flagged, list[rank] = i
__global__ void k_tr_scatter(int n, const int *__restrict__ flag, const int *__restrict__ pos,
                             int *__restrict__ list)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flag[i]) list[pos[i]] = i;
}
__global__ void k_tr_acm_samples(int start, int count, int nsample, const double *__restrict__ eta, int npc,
                                 const int *__restrict__ changed, int new_ncell, const int *__restrict__ new_conn,
                                 int new_nnode, const double *__restrict__ new_coord, tr::Grid g,
                                 const double *__restrict__ coeff, int max_el, int *__restrict__ found)
{
    const size_t t = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= size_t(count) * nsample) return;
    const int i = int(t / nsample), j = int(t % nsample);
    found[t] = tr::acm_sample_one(changed[start + i], eta + size_t(j) * npc, npc, new_ncell, new_conn,
                                  new_nnode, new_coord, g, coeff, max_el);
}
__global__ void k_tr_acm_merge(int start, int count, int nsample, const int *__restrict__ found,
                               const int *__restrict__ changed, int *__restrict__ is_changed,
                               int *__restrict__ elems, double *__restrict__ ratios, double *__restrict__ empty)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    tr::acm_merge_one(start + i, changed[start + i], nsample, found + size_t(i) * nsample, is_changed,
                      elems, ratios, empty);
}
__global__ void k_tr_inject(int n_new, int n_old, int ncomp, const double *__restrict__ src,
                            double *__restrict__ dst, const int *__restrict__ idx,
                            const int *__restrict__ is_changed, const int *__restrict__ idx_changed,
                            const int *__restrict__ elems, const double *__restrict__ ratios)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_new) return;
    for (int d = 0; d < ncomp; ++d)
        dst[size_t(d) * n_new + i] = tr::inject_one(i, src + size_t(d) * n_old, idx, is_changed, idx_changed,
                                                    elems, ratios);
}
__global__ void k_tr_node_locate(int nnew, const double *__restrict__ new_coord, tr::Grid g, tr::OldMesh m,
                                 int *__restrict__ nn, int *__restrict__ el, double *__restrict__ brc,
                                 int *__restrict__ flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nnew) return;
    flag[i] = tr::node_locate_one(i, nnew, new_coord, g, m, nn, el, brc) ? 0 : 1;
}
// The capped breadth-first search of prepare_interpolation (brc-interpolation.cxx:226-329), one
// WARP per node.  tr::node_bfs_one (des_transfer.cuh) is the same search written serially -- the
// statement of what this kernel must produce, and what the host-loop check runs; one thread
// per node spent 22 ms on 354 hull nodes in chains of dependent global-memory trips.  Here the
// visited set lives in shared memory and the elements around one node of a frontier element (one
// patch, <= ~40 entries) are handled by the 32 lanes at once, in lane order = patch order:
//   * new = not in the visited set and the lowest lane among equal ids of the batch, i.e. exactly
//     the entries the serial walk would insert, in the same order (they are appended to the next
//     frontier at rank popc(lower lanes));
//   * "is inside" is decided per lane and the LOWEST hit lane wins -- the first hit of the serial walk;
//   * the running "least outside" element takes the batch maximum only if it is strictly larger,
//     ties inside the batch going to the lowest lane -- the serial walk's first maximum.
__device__ __forceinline__ unsigned tr_hash(int key) { return (unsigned(key) * 2654435761u) & unsigned(tr::BFS_HCAP - 1); }
__device__ __forceinline__ bool tr_probe(const int *table, int key)
{
    unsigned h = tr_hash(key);
    for (;;) {
        const int v = table[h];
        if (v == key) return true;
        if (v < 0) return false;
        h = (h + 1) & unsigned(tr::BFS_HCAP - 1);
    }
}
__device__ __forceinline__ void tr_insert(int *table, int key)
{
    unsigned h = tr_hash(key);
    for (;;) {
        const int old = atomicCAS(table + h, -1, key);
        if (old == -1 || old == key) return;
        h = (h + 1) & unsigned(tr::BFS_HCAP - 1);
    }
}
__global__ void __launch_bounds__(32) k_tr_node_bfs(int start, int count, const int *__restrict__ list, int nnew,
                                                    const double *__restrict__ new_coord, tr::OldMesh m,
                                                    const int *__restrict__ nn_of, int *__restrict__ frontier_all,
                                                    int *__restrict__ el, double *__restrict__ brc,
                                                    int *__restrict__ outcome)
{
    extern __shared__ int table[];                  // tr::BFS_HCAP slots, -1 = free
    const unsigned FULL = 0xffffffffu;
    const int t = blockIdx.x, lane = threadIdx.x;
    if (t >= count) return;
    for (int s = lane; s < tr::BFS_HCAP; s += 32) table[s] = -1;
    __syncwarp();
    const int i = list[start + t];
    const int nn = nn_of[i];
    const double q[3] = {new_coord[i], new_coord[size_t(nnew) + i], new_coord[size_t(2) * nnew + i]};
    const int *patch = m.sup_arr + m.sup_idx[nn];
    const int np = m.sup_idx[nn + 1] - m.sup_idx[nn];
    int *frontier = frontier_all + size_t(t) * 2 * tr::BFS_FCAP, *next = frontier + tr::BFS_FCAP;
    if (np > tr::BFS_FCAP) { if (lane == 0) outcome[start + t] = -1; return; }
    for (int j = lane; j < np; j += 32) { tr_insert(table, patch[j]); frontier[j] = patch[j]; }
    __syncwarp();
    int used = np, nf = np;
    int best_e = (np == 0) ? 0 : patch[0];
    double best_min = -1e30;
    const int MAX_LAYERS = 3;
    for (int layer = 0; layer < MAX_LAYERS && nf > 0; ++layer) {
        int nnext = 0;
        for (int f = 0; f < nf; ++f) {
            const int ee = frontier[f];
            for (int mm = 0; mm < 4; ++mm) {
                const int npn = m.conn[size_t(mm) * m.nelem + ee];
                const int base = m.sup_idx[npn], npe = m.sup_idx[npn + 1] - base;
                for (int j0 = 0; j0 < npe; j0 += 32) {
                    const int jj = j0 + lane;
                    const bool valid = jj < npe;
                    const int cand = valid ? m.sup_arr[base + jj] : -1 - lane;
                    const bool seen = valid ? tr_probe(table, cand) : true;
                    const unsigned same = __match_any_sync(FULL, cand);
                    const bool is_new = valid && !seen && (__ffs(same) - 1 == lane);
                    const unsigned newmask = __ballot_sync(FULL, is_new);
                    if (!newmask) continue;
                    const int nnew_here = __popc(newmask);
                    if (used + nnew_here > tr::BFS_HCAP / 2) { if (lane == 0) outcome[start + t] = -1; return; }
                    __syncwarp();
                    if (is_new) tr_insert(table, cand);
                    used += nnew_here;
                    double r[3] = {0, 0, 0}, min_bary = -DBL_MAX;
                    bool inside = false;
                    if (is_new) {
                        tr::transform(m.coeff + size_t(cand) * 12, 3, q, r);
                        min_bary = r[0];
                        for (int d = 1; d < 3; ++d) if (r[d] < min_bary) min_bary = r[d];
                        double last = 1.0;
                        for (int d = 0; d < 3; ++d) last -= r[d];
                        if (last < min_bary) min_bary = last;
                        inside = tr::is_inside(3, r);
                    }
                    __syncwarp();
                    const unsigned hit = __ballot_sync(FULL, inside);
                    if (hit) {
                        if (lane == __ffs(hit) - 1) { tr::brc_store(i, cand, r, el, brc); outcome[start + t] = 0; }
                        return;
                    }
                    double bm = min_bary; int bl = lane;
                    for (int off = 16; off; off >>= 1) {
                        const double om = __shfl_down_sync(FULL, bm, off);
                        const int ol = __shfl_down_sync(FULL, bl, off);
                        if (om > bm || (om == bm && ol < bl)) { bm = om; bl = ol; }
                    }
                    bm = __shfl_sync(FULL, bm, 0); bl = __shfl_sync(FULL, bl, 0);
                    const int be = __shfl_sync(FULL, cand, bl);
                    if (bm > best_min) { best_min = bm; best_e = be; }
                    if (layer + 1 < MAX_LAYERS) {           // the last layer's list is never walked
                        if (nnext + nnew_here > tr::BFS_FCAP) { if (lane == 0) outcome[start + t] = -1; return; }
                        if (is_new) next[nnext + __popc(newmask & ((1u << lane) - 1u))] = cand;
                        nnext += nnew_here;
                    }
                }
            }
        }
        __syncwarp();
        int *tmp = frontier; frontier = next; next = tmp;
        nf = nnext;
    }
    if (lane != 0) return;
    double r[3];
    if (best_min > -1e-10) {                        // numerically on a face of best_e
        tr::transform(m.coeff + size_t(best_e) * 12, 3, q, r);
        tr::brc_store(i, best_e, r, el, brc);
        outcome[start + t] = 1;
        return;
    }
    // outside the old domain: the nearest old node stands in for q
    const int e = patch[0];
    const double qn[3] = {m.coord[nn], m.coord[size_t(m.nnode) + nn], m.coord[size_t(2) * m.nnode + nn]};
    tr::transform(m.coeff + size_t(e) * 12, 3, qn, r);
    tr::brc_store(i, e, r, el, brc);
    outcome[start + t] = 2;
}
__global__ void k_tr_interp(int n_new, int n_old, int ncomp, const double *__restrict__ src,
                            double *__restrict__ dst, int old_nelem, const int *__restrict__ old_conn,
                            const int *__restrict__ el, const double *__restrict__ brc)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_new) return;
    for (int d = 0; d < ncomp; ++d)
        dst[size_t(d) * n_new + i] = tr::interp_one(i, src + size_t(d) * n_old, old_nelem, old_conn, el, brc);
}

// the cell grid of a device-resident point set [3][npts] inside [lo, hi]
int tr_build_grid(TopoBuf &buf, int npts, const double *d_pts, const double lo[3], const double hi[3], tr::Grid &g)
{
    tr::grid_shape(npts, lo, hi, g);
    g.p = d_pts;
    const size_t ncell = size_t(g.n[0]) * g.n[1] * g.n[2];
    int *d_cnt, *d_start, *d_pts_of;
    TRK(buf.get(&d_cnt, ncell)); TRK(buf.get(&d_start, ncell + 1)); TRK(buf.get(&d_pts_of, size_t(npts)));
    TRK(cudaMemset(d_cnt, 0, ncell * sizeof(int)));
    k_tr_cell_count<<<DES_GRID(npts, 256), 256>>>(g, d_cnt);
    size_t tmp_bytes = 0;
    TRK(cudaMemset(d_start, 0, sizeof(int)));
    TRK(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, d_cnt, d_start + 1, int(ncell)));
    char *d_tmp; TRK(buf.get(&d_tmp, tmp_bytes));
    TRK(cub::DeviceScan::InclusiveSum(d_tmp, tmp_bytes, d_cnt, d_start + 1, int(ncell)));
    TRK(cudaMemset(d_cnt, 0, ncell * sizeof(int)));
    k_tr_cell_fill<<<DES_GRID(npts, 256), 256>>>(g, d_start, d_cnt, d_pts_of);
    TRK(cudaGetLastError());
    g.cell_start = d_start; g.cell_pts = d_pts_of;
    return DES_OK;
}

// flags -> (rank of each flagged entry, ascending list of the flagged, how many)
int tr_compact(TopoBuf &buf, int n, const int *d_flag, int **d_pos, int **d_list, int *count)
{
    TRK(buf.get(d_pos, size_t(n))); TRK(buf.get(d_list, size_t(n)));
    size_t tmp_bytes = 0;
    TRK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_flag, *d_pos, n));
    char *d_tmp; TRK(buf.get(&d_tmp, tmp_bytes));
    TRK(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_flag, *d_pos, n));
    k_tr_scatter<<<DES_GRID(n, 256), 256>>>(n, d_flag, *d_pos, *d_list);
    TRK(cudaGetLastError());
    int last_pos = 0, last_flag = 0;
    TRK(cudaMemcpy(&last_pos, *d_pos + (n - 1), sizeof(int), cudaMemcpyDeviceToHost));
    TRK(cudaMemcpy(&last_flag, d_flag + (n - 1), sizeof(int), cudaMemcpyDeviceToHost));
    *count = last_pos + last_flag;
    return DES_OK;
}

bool tr_check_mesh(const DesTransferMeshes *m, int npc_old, int npc_new, bool need_new_conn)
{
    if (!m || !m->old_coord || !m->old_conn || !m->old_size || !m->new_coord) return false;
    if (m->old_nnode <= 0 || m->old_ncell <= 0 || m->new_nnode <= 0) return false;
    if (size_t(m->old_ncell) * 12 >= (size_t(1) << 31)) return false;
    for (int j = 0; j < npc_old; ++j)
        for (int e = 0; e < m->old_ncell; ++e) {
            const int n = m->old_conn[size_t(j) * m->old_ncell + e];
            if (n < 0 || n >= m->old_nnode) return false;
        }
    if (need_new_conn) {
        if (!m->new_conn || m->new_ncell <= 0) return false;
        for (int j = 0; j < npc_new; ++j)
            for (int e = 0; e < m->new_ncell; ++e) {
                const int n = m->new_conn[size_t(j) * m->new_ncell + e];
                if (n < 0 || n >= m->new_nnode) return false;
            }
    }
    return true;
}

void tr_bbox(int n, const double *coord, bool flat_z, double lo[3], double hi[3])
{
    for (int d = 0; d < 3; ++d) {
        lo[d] = DBL_MAX; hi[d] = -DBL_MAX;
        if (d == 2 && flat_z) { lo[d] = hi[d] = 0; continue; }
        for (int i = 0; i < n; ++i) {
            const double x = coord[size_t(d) * n + i];
            if (x < lo[d]) lo[d] = x;
            if (x > hi[d]) hi[d] = x;
        }
    }
}

double tr_now()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
} // namespace

extern "C" {

// nearest_neighbor_interpolation (nn-interpolation.cxx:661-702): elements (is_surface = 0) or
// the facets of the top surface (is_surface = 1)
int des_remesh_transfer_cells(int device, const DesTransferMeshes *m, int is_surface, int nfields,
                              const DesTransferField *fields, int *is_changed, DesTransferStats *stats)
{
    const int npc = is_surface ? 3 : 4, dim = npc - 1;
    if (!tr_check_mesh(m, npc, npc, true) || nfields < 0 || (nfields && !fields)) return DES_E_ARG;
    for (int f = 0; f < nfields; ++f)
        if (!fields[f].src || !fields[f].dst || fields[f].ncomp < 1 || fields[f].ncomp > 6) return DES_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return DES_E_NODEVICE; }
    const double t0 = tr_now();
    const size_t ON = m->old_nnode, OC = m->old_ncell, NN = m->new_nnode, NC = m->new_ncell;
    TopoBuf buf;
    double *d_ocoord, *d_osize, *d_ncoord, *d_pts, *d_q, *d_coeff;
    int *d_oconn, *d_nconn, *d_idx, *d_changed, *d_flag;
    TRK(buf.get(&d_ocoord, ON * 3)); TRK(buf.get(&d_oconn, OC * 4)); TRK(buf.get(&d_osize, OC));
    TRK(buf.get(&d_ncoord, NN * 3)); TRK(buf.get(&d_nconn, NC * 4));
    TRK(buf.get(&d_pts, OC * 3)); TRK(buf.get(&d_q, NC * 3)); TRK(buf.get(&d_coeff, OC * 12));
    TRK(buf.get(&d_idx, NC)); TRK(buf.get(&d_changed, NC)); TRK(buf.get(&d_flag, NC));
    TRK(cudaMemcpy(d_ocoord, m->old_coord, ON * 3 * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_oconn, m->old_conn, OC * 4 * sizeof(int), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_osize, m->old_size, OC * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_ncoord, m->new_coord, NN * 3 * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_nconn, m->new_conn, NC * 4 * sizeof(int), cudaMemcpyHostToDevice));

    // the old cells' centroids in a cell grid, their barycentric coefficients
    k_tr_centroids<<<DES_GRID(m->old_ncell, 256), 256>>>(m->old_ncell, npc, d_oconn, m->old_nnode, d_ocoord, d_pts);
    double lo[3], hi[3];
    tr_bbox(m->old_nnode, m->old_coord, is_surface != 0, lo, hi);
    tr::Grid g;
    { const int rc = tr_build_grid(buf, m->old_ncell, d_pts, lo, hi, g); if (rc) return rc; }
    k_tr_coeff<<<DES_GRID(m->old_ncell, 256), 256>>>(m->old_ncell, dim, d_oconn, m->old_nnode, d_ocoord, d_osize, d_coeff);

    // find_nearest_neighbor
    k_tr_centroids<<<DES_GRID(m->new_ncell, 256), 256>>>(m->new_ncell, npc, d_nconn, m->new_nnode, d_ncoord, d_q);
    k_tr_nearest<<<DES_GRID(m->new_ncell, 128), 128>>>(g, m->new_ncell, d_q, d_idx, d_changed, d_flag);
    TRK(cudaGetLastError());
    int *d_idx_changed, *d_list, nchanged = 0;
    { const int rc = tr_compact(buf, m->new_ncell, d_flag, &d_idx_changed, &d_list, &nchanged); if (rc) return rc; }
    const double t1 = tr_now();

    // find_acm_elem_ratios
    int *d_elems = nullptr; double *d_ratios = nullptr, *d_empty = nullptr;
    const std::vector<double> eta = tr::acm_samples(is_surface != 0);
    const int nsample = int(eta.size() / npc);
    TRK(buf.get(&d_elems, size_t(nchanged) * tr::KMAX)); TRK(buf.get(&d_ratios, size_t(nchanged) * tr::KMAX));
    TRK(buf.get(&d_empty, size_t(nchanged)));
    if (nchanged) {
        const int max_el = std::min(tr::KMAX, m->old_ncell);
        double *d_eta; TRK(buf.get(&d_eta, eta.size()));
        TRK(cudaMemcpy(d_eta, eta.data(), eta.size() * sizeof(double), cudaMemcpyHostToDevice));
        // cells per pass: 64 Mi sample points (256 MB of scratch); DES_TRANSFER_BLOCK overrides (tests)
        int per_block = std::max(1, std::min(nchanged, (64 << 20) / nsample));
        if (const char *e = std::getenv("DES_TRANSFER_BLOCK")) per_block = std::max(1, std::min(per_block, std::atoi(e)));
        int *d_found; TRK(buf.get(&d_found, size_t(per_block) * nsample));
        for (int start = 0; start < nchanged; start += per_block) {
            const int count = std::min(per_block, nchanged - start);
            const size_t nt = size_t(count) * nsample;
            k_tr_acm_samples<<<unsigned((nt + 127) / 128), 128>>>(start, count, nsample, d_eta, npc, d_list,
                m->new_ncell, d_nconn, m->new_nnode, d_ncoord, g, d_coeff, max_el, d_found);
            k_tr_acm_merge<<<DES_GRID(count, 128), 128>>>(start, count, nsample, d_found, d_list, d_changed,
                d_elems, d_ratios, d_empty);
        }
        TRK(cudaGetLastError());
    }
    TRK(cudaDeviceSynchronize());
    const double t2 = tr_now();

    // nn_interpolate_elem_fields
    size_t wmax = 1;
    for (int f = 0; f < nfields; ++f) wmax = std::max(wmax, size_t(fields[f].ncomp));
    double *d_src, *d_dst;
    TRK(buf.get(&d_src, OC * wmax)); TRK(buf.get(&d_dst, NC * wmax));
    for (int f = 0; f < nfields; ++f) {
        const size_t w = fields[f].ncomp;
        TRK(cudaMemcpy(d_src, fields[f].src, OC * w * sizeof(double), cudaMemcpyHostToDevice));
        k_tr_inject<<<DES_GRID(m->new_ncell, 256), 256>>>(m->new_ncell, m->old_ncell, int(w), d_src, d_dst, d_idx,
                                                         d_changed, d_idx_changed, d_elems, d_ratios);
        TRK(cudaGetLastError());
        TRK(cudaMemcpy(fields[f].dst, d_dst, NC * w * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (is_changed) TRK(cudaMemcpy(is_changed, d_changed, NC * sizeof(int), cudaMemcpyDeviceToHost));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->nchanged = nchanged; stats->nsample = nsample;
        stats->search_seconds = t1 - t0; stats->acm_seconds = t2 - t1; stats->field_seconds = tr_now() - t2;
    }
    return DES_OK;
}

// barycentric_node_interpolation (brc-interpolation.cxx:345-423); old_support_* are the OLD
// mesh's node -> element lists (create_support's idx / arr)
int des_remesh_transfer_nodes(int device, const DesTransferMeshes *m, const int *old_support_idx,
                              const int *old_support_arr, int nfields, const DesTransferField *fields,
                              int *el, double *brc, DesTransferStats *stats)
{
    if (!tr_check_mesh(m, 4, 4, false) || !old_support_idx || !old_support_arr || nfields < 0 || (nfields && !fields))
        return DES_E_ARG;
    for (int f = 0; f < nfields; ++f)
        if (!fields[f].src || !fields[f].dst || fields[f].ncomp < 1 || fields[f].ncomp > 6) return DES_E_ARG;
    const size_t ON = m->old_nnode, OC = m->old_ncell, NN = m->new_nnode;
    if (old_support_idx[0] != 0 || size_t(old_support_idx[ON]) != OC * 4) return DES_E_ARG;
    for (size_t n = 0; n < ON; ++n)
        if (old_support_idx[n + 1] <= old_support_idx[n]) return DES_E_ARG;      // a node on no element
    for (size_t s = 0; s < OC * 4; ++s)
        if (old_support_arr[s] < 0 || size_t(old_support_arr[s]) >= OC) return DES_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return DES_E_NODEVICE; }
    const double t0 = tr_now();
    TopoBuf buf;
    double *d_ocoord, *d_osize, *d_ncoord, *d_coeff, *d_brc;
    int *d_oconn, *d_sidx, *d_sarr, *d_nn, *d_el, *d_flag;
    TRK(buf.get(&d_ocoord, ON * 3)); TRK(buf.get(&d_oconn, OC * 4)); TRK(buf.get(&d_osize, OC));
    TRK(buf.get(&d_ncoord, NN * 3)); TRK(buf.get(&d_coeff, OC * 12)); TRK(buf.get(&d_brc, NN * 4));
    TRK(buf.get(&d_sidx, ON + 1)); TRK(buf.get(&d_sarr, OC * 4));
    TRK(buf.get(&d_nn, NN)); TRK(buf.get(&d_el, NN)); TRK(buf.get(&d_flag, NN));
    TRK(cudaMemcpy(d_ocoord, m->old_coord, ON * 3 * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_oconn, m->old_conn, OC * 4 * sizeof(int), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_osize, m->old_size, OC * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_ncoord, m->new_coord, NN * 3 * sizeof(double), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_sidx, old_support_idx, (ON + 1) * sizeof(int), cudaMemcpyHostToDevice));
    TRK(cudaMemcpy(d_sarr, old_support_arr, OC * 4 * sizeof(int), cudaMemcpyHostToDevice));

    double lo[3], hi[3];
    tr_bbox(m->old_nnode, m->old_coord, false, lo, hi);
    tr::Grid g;
    { const int rc = tr_build_grid(buf, m->old_nnode, d_ocoord, lo, hi, g); if (rc) return rc; }
    k_tr_coeff<<<DES_GRID(m->old_ncell, 256), 256>>>(m->old_ncell, 3, d_oconn, m->old_nnode, d_ocoord, d_osize, d_coeff);
    tr::OldMesh om = { m->old_nnode, m->old_ncell, d_ocoord, d_oconn, d_sidx, d_sarr, d_coeff };
    k_tr_node_locate<<<DES_GRID(m->new_nnode, 128), 128>>>(m->new_nnode, d_ncoord, g, om, d_nn, d_el, d_brc, d_flag);
    TRK(cudaGetLastError());
    int *d_pos, *d_list, nflag = 0;
    { const int rc = tr_compact(buf, m->new_nnode, d_flag, &d_pos, &d_list, &nflag); if (rc) return rc; }
    const double t1 = tr_now();

    int n_face = 0, n_outside = 0;
    if (nflag) {
        int per_batch = std::min(nflag, 4096);      // nodes per pass (64 KB of frontier scratch each)
        if (const char *e = std::getenv("DES_TRANSFER_BLOCK")) per_batch = std::max(1, std::min(per_batch, std::atoi(e)));
        int *d_frontier, *d_outcome;
        TRK(buf.get(&d_frontier, size_t(per_batch) * 2 * tr::BFS_FCAP));
        TRK(buf.get(&d_outcome, size_t(nflag)));
        const int smem = tr::BFS_HCAP * int(sizeof(int));
        TRK(cudaFuncSetAttribute(k_tr_node_bfs, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        for (int start = 0; start < nflag; start += per_batch) {
            const int count = std::min(per_batch, nflag - start);
            k_tr_node_bfs<<<count, 32, smem>>>(start, count, d_list, m->new_nnode, d_ncoord, om, d_nn,
                                               d_frontier, d_el, d_brc, d_outcome);
        }
        TRK(cudaGetLastError());
        std::vector<int> outcome(nflag);
        TRK(cudaMemcpy(outcome.data(), d_outcome, size_t(nflag) * sizeof(int), cudaMemcpyDeviceToHost));
        for (int v : outcome) {
            // the search around a new node ran out of its scratch: more than 16 384 old elements
            // within three rings of the nearest old node (see the header)
            if (v < 0) return DES_E_UNSUPPORTED;
            n_face += (v == 1); n_outside += (v == 2);
        }
    }
    TRK(cudaDeviceSynchronize());
    const double t2 = tr_now();

    size_t wmax = 1;
    for (int f = 0; f < nfields; ++f) wmax = std::max(wmax, size_t(fields[f].ncomp));
    double *d_src, *d_dst;
    TRK(buf.get(&d_src, ON * wmax)); TRK(buf.get(&d_dst, NN * wmax));
    for (int f = 0; f < nfields; ++f) {
        const size_t w = fields[f].ncomp;
        TRK(cudaMemcpy(d_src, fields[f].src, ON * w * sizeof(double), cudaMemcpyHostToDevice));
        k_tr_interp<<<DES_GRID(m->new_nnode, 256), 256>>>(m->new_nnode, m->old_nnode, int(w), d_src, d_dst,
                                                         m->old_ncell, d_oconn, d_el, d_brc);
        TRK(cudaGetLastError());
        TRK(cudaMemcpy(fields[f].dst, d_dst, NN * w * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (el) TRK(cudaMemcpy(el, d_el, NN * sizeof(int), cudaMemcpyDeviceToHost));
    if (brc) TRK(cudaMemcpy(brc, d_brc, NN * 4 * sizeof(double), cudaMemcpyDeviceToHost));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->nchanged = nflag; stats->n_on_face = n_face; stats->n_outside = n_outside;
        stats->search_seconds = t1 - t0; stats->acm_seconds = t2 - t1; stats->field_seconds = tr_now() - t2;
    }
    return DES_OK;
}

} // extern "C"
#undef TRK
