// Remesh field transfer on the device (include/des_gpu.h, "remesh
// field transfer" block): what remesh() (remeshing.cxx:3058-3069) asks of
//   nearest_neighbor_interpolation   nn-interpolation.cxx:661-702  (elements, top-surface facets)
//   barycentric_node_interpolation   brc-interpolation.cxx:345-423 (nodes)
// with Barycentric_transformation (barycentric-fn.cxx) and the k-nearest-neighbour searches the
// reference gives to nanoflann (knn.cxx:139-180).
//
// The searches are EXACT k-NN in double precision over a uniform cell grid (counting sort of the
// points into cells, expanding shells of cells around the query until the k-th distance is
// proven), results ascending by (distance^2, index) -- nanoflann's KNNResultSet order
// (nanoflann.hpp:247-282); equal distances, which nanoflann resolves by its tree traversal
// order, are resolved here by the lower index.  Distances are evaluated as L2_Simple_Adaptor
// does, ((dx*dx + dy*dy) + dz*dz), so the `dist2 < eps` / `dist2 == 0` decisions fall the same
// way.  Every arithmetic statement below keeps the reference's operation order (the build has
// FMA contraction off), so the transferred fields are bit-identical.
//
// Every per-item routine is a host/device inline function and the kernels are one-line loops
// over them -- except the breadth-first search of the node transfer, which the device runs one
// warp per node (des_transfer_api.cuh: k_tr_node_bfs) in the visiting order node_bfs_one below
// states serially.  tests/hostcheck/transfer_hostcheck.cpp runs the same functions in host loops
// so that the logic can be checked where there is no GPU (test infrastructure; the library's
// entry points only ever launch the kernels).
#pragma once
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <vector>

#ifdef __CUDACC__
#define TR_HD __host__ __device__ inline
#else
#define TR_HD inline
#endif

namespace tr {

constexpr int KMAX = 32;            // nn_t / ratio_t width (nn-interpolation.hpp:4-5)
#ifndef DES_ACM_K1
#define DES_ACM_K1 6
#endif
constexpr int ACM_K1 = DES_ACM_K1;  // first, short search of find_acm_elem_ratios' sample points
constexpr int BFS_HCAP = 32768;     // visited-set slots per node in the breadth-first search
constexpr int BFS_FCAP = 8192;      // frontier entries per node

// ---- uniform cell grid over a point set ----------------------------------------------------
struct Grid {
    double o[3];            // lower corner
    double h;               // cell edge
    int n[3];               // cells per axis
    int npts;
    const double *p;        // [3][npts] point coordinates
    const int *cell_start;  // [ncell + 1]
    const int *cell_pts;    // [npts] point ids grouped by cell
};

TR_HD int cell_axis(const Grid &g, int d, double x)
{
    const double t = (x - g.o[d]) / g.h;
    if (!(t > 0)) return 0;
    if (t >= g.n[d]) return g.n[d] - 1;
    return int(t);
}
TR_HD size_t cell_id(const Grid &g, int cx, int cy, int cz)
{
    return (size_t(cz) * g.n[1] + cy) * g.n[0] + cx;
}
TR_HD size_t cell_of_point(const Grid &g, int i)
{
    return cell_id(g, cell_axis(g, 0, g.p[i]), cell_axis(g, 1, g.p[size_t(g.npts) + i]),
                   cell_axis(g, 2, g.p[size_t(2) * g.npts + i]));
}

// Result lists: KNNResultSet (nanoflann.hpp:213-296) -- ascending by distance^2, a new entry goes
// behind the entries that are not farther -- with equal distances ordered by the lower index.
// ListRT holds a run-time k (<= KMAX) in caller-provided arrays; ListFixed<K> holds a compile-time
// K in members that the compiler keeps in registers (every index is static after unrolling): the
// nearest-neighbour searches use K = 1.
TR_HD bool knn_before(double da, int ia, double db, int ib)        // (da, ia) sorts before (db, ib)
{
    return da < db || (da == db && ia < ib);
}
struct ListRT {
    int k, cnt; int *idx; double *d2;
    TR_HD void insert(int i, double d)              // KNNResultSet::addPoint (nanoflann.hpp:247-282)
    {
        if (cnt == k && !knn_before(d, i, d2[k - 1], idx[k - 1])) return;
        int j = cnt < k ? cnt : k - 1;
        while (j > 0 && knn_before(d, i, d2[j - 1], idx[j - 1])) { d2[j] = d2[j - 1]; idx[j] = idx[j - 1]; --j; }
        d2[j] = d; idx[j] = i;
        if (cnt < k) ++cnt;
    }
    TR_HD bool full() const { return cnt == k; }
    TR_HD double worst() const { return d2[k - 1]; }
};
template <int K> struct ListFixed {
    int cnt; int idx[K]; double d2[K];
    TR_HD ListFixed() : cnt(0)
    {
        for (int j = 0; j < K; ++j) { idx[j] = 0x7fffffff; d2[j] = DBL_MAX; }       // free slots sort last
    }
    TR_HD void insert(int i, double d)
    {
        if (!knn_before(d, i, d2[K - 1], idx[K - 1])) return;
        bool placed = false;
#ifdef __CUDA_ARCH__
        #pragma unroll
#endif
        for (int j = K - 1; j > 0; --j) {
            if (placed) continue;
            if (knn_before(d, i, d2[j - 1], idx[j - 1])) { d2[j] = d2[j - 1]; idx[j] = idx[j - 1]; }
            else { d2[j] = d; idx[j] = i; placed = true; }
        }
        if (!placed) { d2[0] = d; idx[0] = i; }
        if (cnt < K) ++cnt;
    }
    TR_HD bool full() const { return cnt == K; }
    TR_HD double worst() const { return d2[K - 1]; }
};

template <class List>
TR_HD void knn_cell(const Grid &g, size_t c, const double *q, List &L)
{
    const size_t N = g.npts;
    for (int s = g.cell_start[c]; s < g.cell_start[c + 1]; ++s) {
        const int i = g.cell_pts[s];
        // L2_Simple_Adaptor::evalMetric: result += diff * diff over the dimensions, from 0
        const double dx = q[0] - g.p[i], dy = q[1] - g.p[N + i], dz = q[2] - g.p[2 * N + i];
        double d = 0;
        d += dx * dx; d += dy * dy; d += dz * dz;
        L.insert(i, d);
    }
}

// the nearest points of q into L, shell by shell of cells around q's cell, until the list is full
// and its last entry is closer than anything an unvisited cell can hold
template <class List>
TR_HD void knn_walk(const Grid &g, const double *q, List &L)
{
    int c[3], maxr = 0;
    for (int d = 0; d < 3; ++d) {
        c[d] = cell_axis(g, d, q[d]);
        const int a = c[d] > g.n[d] - 1 - c[d] ? c[d] : g.n[d] - 1 - c[d];
        if (a > maxr) maxr = a;
    }
    for (int r = 0; ; ++r) {
        const int lo0 = c[0] - r, hi0 = c[0] + r, lo1 = c[1] - r, hi1 = c[1] + r, lo2 = c[2] - r, hi2 = c[2] + r;
        const int x0 = lo0 < 0 ? 0 : lo0, x1 = hi0 > g.n[0] - 1 ? g.n[0] - 1 : hi0;
        const int y0 = lo1 < 0 ? 0 : lo1, y1 = hi1 > g.n[1] - 1 ? g.n[1] - 1 : hi1;
        const int z0 = lo2 < 0 ? 0 : lo2, z1 = hi2 > g.n[2] - 1 ? g.n[2] - 1 : hi2;
        for (int z = z0; z <= z1; ++z) {
            const bool zedge = (z == lo2 || z == hi2);
            for (int y = y0; y <= y1; ++y) {
                if (zedge || y == lo1 || y == hi1) {
                    for (int x = x0; x <= x1; ++x) knn_cell(g, cell_id(g, x, y, z), q, L);
                } else {
                    if (lo0 >= 0) knn_cell(g, cell_id(g, lo0, y, z), q, L);
                    if (hi0 <= g.n[0] - 1 && hi0 != lo0) knn_cell(g, cell_id(g, hi0, y, z), q, L);
                }
            }
        }
        if (r >= maxr) break;                       // every cell has been visited
        if (L.full()) {
            // no unvisited point is closer than the nearest face of the visited block that
            // has cells beyond it
            double b = DBL_MAX;
            const int lo[3] = {lo0, lo1, lo2}, hi[3] = {hi0, hi1, hi2};
            for (int d = 0; d < 3; ++d) {
                if (lo[d] > 0) { const double t = q[d] - (g.o[d] + lo[d] * g.h); if (t < b) b = t; }
                if (hi[d] < g.n[d] - 1) { const double t = (g.o[d] + (hi[d] + 1) * g.h) - q[d]; if (t < b) b = t; }
            }
            b -= 1e-9 * g.h;                        // rounding of the cell assignment
            if (b > 0 && L.worst() < b * b) break;
        }
    }
}

// the k nearest points of q (k <= KMAX), ascending; returns how many there are (min(k, npts))
TR_HD int knn(const Grid &g, const double *q, int k, int *idx, double *d2)
{
    ListRT L = { k, 0, idx, d2 };
    knn_walk(g, q, L);
    return L.cnt;
}
// kept for the brute-force comparison of the host-loop check
TR_HD void knn_insert(int k, int &cnt, int *idx, double *d2, int i, double d)
{
    ListRT L = { k, cnt, idx, d2 };
    L.insert(i, d);
    cnt = L.cnt;
}

// ---- Barycentric_transformation (barycentric-fn.cxx) ----------------------------------------
// coefficients of one cell, index3d(node, dim) = node*3 + dim / index2d = node*2 + dim; 12 per cell
TR_HD void coeff3d(const double *a, const double *b, const double *c, const double *d, double volume, double *cf)
{
    const double det = 6 * volume;                  // barycentric-fn.cxx:252-293
    cf[0] = (b[0] * (c[1]*d[2] - d[1]*c[2]) + c[0] * (d[1]*b[2] - b[1]*d[2]) + d[0] * (b[1]*c[2] - c[1]*b[2])) / det;
    cf[1] = (a[0] * (d[1]*c[2] - c[1]*d[2]) + c[0] * (a[1]*d[2] - d[1]*a[2]) + d[0] * (c[1]*a[2] - a[1]*c[2])) / det;
    cf[2] = (a[0] * (b[1]*d[2] - d[1]*b[2]) + b[0] * (d[1]*a[2] - a[1]*d[2]) + d[0] * (a[1]*b[2] - b[1]*a[2])) / det;
    cf[3] = ((d[1] - b[1]) * (c[2] - b[2]) - (c[1] - b[1]) * (d[2] - b[2])) / det;
    cf[4] = ((c[1] - a[1]) * (d[2] - c[2]) - (d[1] - c[1]) * (c[2] - a[2])) / det;
    cf[5] = ((b[1] - d[1]) * (a[2] - d[2]) - (a[1] - d[1]) * (b[2] - d[2])) / det;
    cf[6] = ((d[2] - b[2]) * (c[0] - b[0]) - (c[2] - b[2]) * (d[0] - b[0])) / det;
    cf[7] = ((c[2] - a[2]) * (d[0] - c[0]) - (d[2] - c[2]) * (c[0] - a[0])) / det;
    cf[8] = ((b[2] - d[2]) * (a[0] - d[0]) - (a[2] - d[2]) * (b[0] - d[0])) / det;
    cf[9] = ((d[0] - b[0]) * (c[1] - b[1]) - (c[0] - b[0]) * (d[1] - b[1])) / det;
    cf[10] = ((c[0] - a[0]) * (d[1] - c[1]) - (d[0] - c[0]) * (c[1] - a[1])) / det;
    cf[11] = ((b[0] - d[0]) * (a[1] - d[1]) - (a[0] - d[0]) * (b[1] - d[1])) / det;
}
TR_HD void coeff2d(const double *a, const double *b, const double *c, double area, double *cf)
{
    const double det = 2 * area;                    // barycentric-fn.cxx:234-241
    cf[0] = (b[0]*c[1] - b[1]*c[0]) / det;
    cf[1] = (c[0]*a[1] - c[1]*a[0]) / det;
    cf[2] = (b[1] - c[1]) / det;
    cf[3] = (c[1] - a[1]) / det;
    cf[4] = (c[0] - b[0]) / det;
    cf[5] = (a[0] - c[0]) / det;
}
// one old cell's coefficients from the SoA mesh (the constructors, barycentric-fn.cxx:4-111)
TR_HD void coeff_one(int e, int dim, int ncell, const int *conn, int nnode, const double *coord,
                     const double *size, double *coeff)
{
    double x[4][3];
    for (int j = 0; j <= dim; ++j) {
        const int n = conn[size_t(j) * ncell + e];
        for (int d = 0; d < 3; ++d) x[j][d] = coord[size_t(d) * nnode + n];
    }
    if (dim == 3) coeff3d(x[0], x[1], x[2], x[3], size[e], coeff + size_t(e) * 12);
    else coeff2d(x[0], x[1], x[2], size[e], coeff + size_t(e) * 12);
}
// transform (barycentric-fn.cxx:134-165) and is_inside (:186-220)
TR_HD void transform(const double *cf, int dim, const double *pt, double *r)
{
    for (int d = 0; d < dim; ++d) {
        r[d] = cf[d];
        for (int i = 0; i < dim; ++i) r[d] += cf[(i + 1) * dim + d] * pt[i];
    }
}
TR_HD bool is_inside(int dim, const double *r)
{
    if (dim == 3) {
        const double tolerance = 5e-11;
        return r[0] >= -tolerance && r[1] >= -tolerance && r[2] >= -tolerance &&
               (r[0] + r[1] + r[2]) <= 1 + tolerance;
    }
    const double tolerance = 1e-12;
    return r[0] >= -tolerance && r[1] >= -tolerance && (r[0] + r[1]) <= 1 + tolerance;
}

// ---- nearest_neighbor_interpolation ---------------------------------------------------------
// average_nodal_to_elem (mesh.cxx:3509-3533) of the coordinates; the surface pass keeps x, y
// and zeroes z (nn-interpolation.cxx:32-35, 355-359)
TR_HD void centroid_one(int e, int npc, int ncell, const int *conn, int nnode, const double *coord, double *out)
{
    const int ncomp = npc == 4 ? 3 : 2;
    for (int d = 0; d < ncomp; ++d) {
        double avg = 0.0;
        for (int k = 0; k < npc; ++k) avg += coord[size_t(d) * nnode + conn[size_t(k) * ncell + e]];
        out[size_t(d) * ncell + e] = avg / npc;
    }
    if (ncomp == 2) out[size_t(2) * ncell + e] = 0;
}
// find_nearest_neighbor (nn-interpolation.cxx:53-62)
TR_HD void nearest_one(int i, const Grid &g, int nq, const double *queries, int *idx, int *is_changed)
{
    const double q[3] = {queries[i], queries[size_t(nq) + i], queries[size_t(2) * nq + i]};
    ListFixed<1> L;
    knn_walk(g, q, L);
    idx[i] = L.idx[0];
    is_changed[i] = (L.d2[0] < 1e-15) ? 0 : 1;
}
// find_acm_elem_ratios, one sample point of one changed cell (nn-interpolation.cxx:213-253):
// the old cell that encloses it among the max_el nearest old centroids, or -1
TR_HD int acm_sample_one(int e, const double *eta, int npc, int new_ncell, const int *new_conn,
                         int new_nnode, const double *new_coord, const Grid &g, const double *coeff,
                         int max_el)
{
    double x[3];
    for (int d = 0; d < 3; ++d) {
        x[d] = 0;
        for (int n = 0; n < npc; ++n)
            x[d] += new_coord[size_t(d) * new_nnode + new_conn[size_t(n) * new_ncell + e]] * eta[n];
    }
    const int dim = npc - 1;
    // The enclosing cell is nearly always among the few nearest centroids.  The k nearest in
    // ascending (distance, index) order are the first k of the max_el nearest, so a hit among the
    // first ACM_K1 is the hit the full list would give; only a miss pays for the full search.
    // (A register-resident ListFixed<ACM_K1> for this first search was measured on B200: 122
    // registers instead of 56 halve the occupancy and the kernel takes 13.4 ms instead of 7.9 ms
    // at 18 M sample points -- the run-time list in local memory stays.)
    int id[KMAX]; double d2[KMAX];
    const int k1 = max_el < ACM_K1 ? max_el : ACM_K1;
    int cnt = knn(g, x, k1, id, d2);
    for (int jj = 0; jj < cnt; ++jj) {
        double r[3];
        transform(coeff + size_t(id[jj]) * 12, dim, x, r);
        if (is_inside(dim, r)) return id[jj];
    }
    if (k1 == max_el) return -1;
    cnt = knn(g, x, max_el, id, d2);
    for (int jj = k1; jj < cnt; ++jj) {
        double r[3];
        transform(coeff + size_t(id[jj]) * 12, dim, x, r);
        if (is_inside(dim, r)) return id[jj];
    }
    return -1;
}
// the tally over the samples of one changed cell, in sample order (nn-interpolation.cxx:240-312)
TR_HD void acm_merge_one(int i, int e, int nsample, const int *found, int *is_changed,
                         int *elems, double *ratios, double *empty)
{
    int count[KMAX], keys[KMAX], size = 0;
    for (int j = 0; j < nsample; ++j) {
        const int old_e = found[j];
        if (old_e < 0) continue;
        bool seen = false;
        for (int k = 0; k < size; ++k)
            if (keys[k] == old_e) { count[k]++; seen = true; break; }
        if (!seen && size < KMAX) { keys[size] = old_e; count[size] = 1; size++; }
    }
    int total = 0;
    for (int k = 0; k < size; ++k) total += count[k];
    empty[i] = double(nsample - total) / nsample;
    for (int k = 0; k < KMAX; ++k) elems[size_t(i) * KMAX + k] = -1;
    if (total == 0 || size == 1) { is_changed[e] = -1; return; }
    const double inv = 1.0 / total;
    for (int k = 0; k < size; ++k) {
        elems[size_t(i) * KMAX + k] = keys[k];
        ratios[size_t(i) * KMAX + k] = count[k] * inv;
    }
}
// inject_field (nn-interpolation.cxx:433-523), component d of cell i
TR_HD double inject_one(int i, const double *src, const int *idx, const int *is_changed,
                        const int *idx_changed, const int *elems, const double *ratios)
{
    if (is_changed[i] > 0) {
        const size_t n = idx_changed[i];
        double t = 0;
        for (int j = 0; j < KMAX; ++j) {
            const int e = elems[n * KMAX + j];
            if (e < 0) break;
            t += ratios[n * KMAX + j] * src[e];
        }
        return t;
    }
    return src[idx[i]];
}

// ---- barycentric_node_interpolation ---------------------------------------------------------
struct OldMesh {
    int nnode, nelem;
    const double *coord;    // [3][nnode]
    const int *conn;        // [4][nelem]
    const int *sup_idx;     // [nnode + 1]
    const int *sup_arr;     // [4 * nelem]
    const double *coeff;    // [nelem][12]
};
TR_HD void brc_store(int i, int e, const double *r, int *el, double *brc)
{
    el[i] = e;                                      // brc-interpolation.cxx:331-337
    double sum = 0;
    for (int d = 0; d < 3; ++d) { brc[size_t(i) * 4 + d] = r[d]; sum += r[d]; }
    brc[size_t(i) * 4 + 3] = 1 - sum;
}
// prepare_interpolation up to the one-ring search (brc-interpolation.cxx:183-222); returns
// false when the breadth-first search has to take over (nn[i] then holds the nearest old node)
TR_HD bool node_locate_one(int i, int nnew, const double *new_coord, const Grid &g, const OldMesh &m,
                           int *nn_out, int *el, double *brc)
{
    const double q[3] = {new_coord[i], new_coord[size_t(nnew) + i], new_coord[size_t(2) * nnew + i]};
    ListFixed<1> L;
    knn_walk(g, q, L);
    const int nn = L.idx[0];
    const double d2[1] = { L.d2[0] };
    nn_out[i] = nn;
    const int *patch = m.sup_arr + m.sup_idx[nn];
    const int np = m.sup_idx[nn + 1] - m.sup_idx[nn];
    double r[3];
    if (d2[0] == 0) {
        const int e = patch[0];
        transform(m.coeff + size_t(e) * 12, 3, q, r);
        for (int d = 0; d < 3; ++d) r[d] = (r[d] > 0.9) ? 1 : 0;
        brc_store(i, e, r, el, brc);
        return true;
    }
    for (int j = 0; j < np; ++j) {
        const int e = patch[j];
        transform(m.coeff + size_t(e) * 12, 3, q, r);
        if (is_inside(3, r)) { brc_store(i, e, r, el, brc); return true; }
    }
    return false;
}
// std::unordered_set<int>::insert(...).second on an open-addressing table (-1 = free)
TR_HD int bfs_insert(int *table, int key, int &used)
{
    unsigned h = (unsigned(key) * 2654435761u) & unsigned(BFS_HCAP - 1);
    for (;;) {
        const int v = table[h];
        if (v == key) return 0;
        if (v < 0) { table[h] = key; ++used; return 1; }
        h = (h + 1) & unsigned(BFS_HCAP - 1);
    }
}
// the breadth-first search and the two fallbacks (brc-interpolation.cxx:226-329); `table`
// (BFS_HCAP ints, all -1), `fa`, `fb` (BFS_FCAP ints each) are this node's scratch.
// returns 0 found in the search, 1 taken as lying on a face, 2 outside the old domain,
// -1 scratch exhausted
TR_HD int node_bfs_one(int i, int nnew, const double *new_coord, const OldMesh &m, int nn,
                       int *table, int *fa, int *fb, int *el, double *brc)
{
    const double q[3] = {new_coord[i], new_coord[size_t(nnew) + i], new_coord[size_t(2) * nnew + i]};
    const int *patch = m.sup_arr + m.sup_idx[nn];
    const int np = m.sup_idx[nn + 1] - m.sup_idx[nn];
    int used = 0, nf = 0;
    if (np > BFS_FCAP) return -1;
    for (int j = 0; j < np; ++j) { bfs_insert(table, patch[j], used); fa[nf++] = patch[j]; }
    int best_e = (np == 0) ? 0 : patch[0];
    double best_min = -1e30;
    double r[3];
    const int MAX_LAYERS = 3;
    int *frontier = fa, *next = fb;
    for (int layer = 0; layer < MAX_LAYERS && nf > 0; ++layer) {
        int nnext = 0;
        for (int f = 0; f < nf; ++f) {
            const int ee = frontier[f];
            for (int mm = 0; mm < 4; ++mm) {
                const int npn = m.conn[size_t(mm) * m.nelem + ee];
                const int *pe = m.sup_arr + m.sup_idx[npn];
                const int npe = m.sup_idx[npn + 1] - m.sup_idx[npn];
                for (int jj = 0; jj < npe; ++jj) {
                    const int cand = pe[jj];
                    if (used >= BFS_HCAP / 2) return -1;
                    if (!bfs_insert(table, cand, used)) continue;
                    transform(m.coeff + size_t(cand) * 12, 3, q, r);
                    double min_bary = r[0];
                    for (int d = 1; d < 3; ++d) if (r[d] < min_bary) min_bary = r[d];
                    double last = 1.0;
                    for (int d = 0; d < 3; ++d) last -= r[d];
                    if (last < min_bary) min_bary = last;
                    if (min_bary > best_min) { best_min = min_bary; best_e = cand; }
                    if (is_inside(3, r)) { brc_store(i, cand, r, el, brc); return 0; }
                    if (layer + 1 < MAX_LAYERS) {   // the last layer's list is never walked
                        if (nnext >= BFS_FCAP) return -1;
                        next[nnext++] = cand;
                    }
                }
            }
        }
        int *t = frontier; frontier = next; next = t;
        nf = nnext;
    }
    if (best_min > -1e-10) {
        transform(m.coeff + size_t(best_e) * 12, 3, q, r);
        brc_store(i, best_e, r, el, brc);
        return 1;
    }
    // outside the old domain: the nearest old node stands in for q
    const int e = patch[0];
    const double qn[3] = {m.coord[nn], m.coord[size_t(m.nnode) + nn], m.coord[size_t(2) * m.nnode + nn]};
    transform(m.coeff + size_t(e) * 12, 3, qn, r);
    brc_store(i, e, r, el, brc);
    return 2;
}
// interpolate_field (brc-interpolation.cxx:18-96), component d of node i
TR_HD double interp_one(int i, const double *src, int old_nelem, const int *old_conn, const int *el,
                        const double *brc)
{
    const int e = el[i];
    double result = 0;
    for (int j = 0; j < 4; ++j) result += src[old_conn[size_t(j) * old_nelem + e]] * brc[size_t(i) * 4 + j];
    return result;
}

// sample points of find_acm_elem_ratios (nn-interpolation.cxx:89-153): barycentric weights of
// the new cell's nodes, [nsample][npc]
inline std::vector<double> acm_samples(bool is_surface)
{
    const int neta0 = is_surface ? 20 : 10;
    const int neta1 = neta0 + 1;
    const int neta2 = neta0;
    const double spacing0 = 1.0 / neta0, spacing1 = 1.0 / neta1, spacing2 = 1.0 / neta2;
    const double eps = 1e-15;
    std::vector<double> out;
    for (int i = 0; i < neta0; i++)
        for (int j = 0; j < neta1; j++) {
            if (is_surface) {
                const double eta[3] = {(i + 0.5) * spacing0, (j + 0.5) * spacing1,
                                       1 - (i + 0.5) * spacing0 - (j + 0.5) * spacing1};
                if (eta[2] < eps) continue;
                out.insert(out.end(), eta, eta + 3);
            } else {
                for (int k = 0; k < neta2; k++) {
                    const double eta[4] = {(i + 0.5) * spacing0, (j + 0.5) * spacing1, (k + 0.5) * spacing2,
                                           1 - (i + 0.5) * spacing0 - (j + 0.5) * spacing1 - (k + 0.5) * spacing2};
                    if (eta[3] < eps) continue;
                    out.insert(out.end(), eta, eta + 4);
                }
            }
        }
    return out;
}

// cell edge and counts for a point set inside the box [lo, hi]: about two points per cell
inline void grid_shape(int npts, const double lo[3], const double hi[3], Grid &g)
{
    double ext[3], vol = 1; int nd = 0;
    for (int d = 0; d < 3; ++d) {
        ext[d] = hi[d] - lo[d];
        if (ext[d] > 0) { vol *= ext[d]; ++nd; }
        g.o[d] = lo[d];
    }
    const double cells = npts > 2 ? npts / 2.0 : 1.0;
    double h = nd ? std::pow(vol / cells, 1.0 / nd) : 1.0;
    double emax = 0; for (int d = 0; d < 3; ++d) if (ext[d] > emax) emax = ext[d];
    if (!(h > emax / 1024)) h = emax / 1024;     // at most 1024 cells along an axis
    if (!(h > 0)) h = 1.0;
    for (;;) {
        double total = 1;
        for (int d = 0; d < 3; ++d) {
            int n = int(ext[d] / h) + 1;
            g.n[d] = n < 1 ? 1 : (n > 1025 ? 1025 : n);
            total *= g.n[d];
        }
        if (total <= 4.0 * npts + 4096) break;      // flat or elongated clouds: coarser cells
        h *= 1.25;
    }
    g.h = h;
    g.npts = npts;
}

} // namespace tr
