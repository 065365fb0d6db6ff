// TEST INFRASTRUCTURE ONLY -- never linked or loaded by the product path.
//
// The per-item routines of the remesh field transfer (dynearthsol_b200/csrc/des_transfer.cuh:
// the same source the CUDA kernels are one-line loops over) run in host loops, so that the
// search / tally / interpolation logic can be compared with the reference's own
// nearest_neighbor_interpolation and barycentric_node_interpolation where there is no GPU
// (tests/test_transfer.py, `-m "not gpu"`).  The library itself has no such path: its entry
// points des_remesh_transfer_cells / _nodes only launch kernels and fail without a device.
// Built by __graft_entry__.build() into tests/hostcheck/_build/ (g++ -O2 -ffp-contract=off).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>
#include "../../include/des_gpu.h"
#include "../../dynearthsol_b200/csrc/des_transfer.cuh"

namespace {
void bbox(int n, const double *coord, bool flat_z, double lo[3], double hi[3])
{
    for (int d = 0; d < 3; ++d) {
        lo[d] = DBL_MAX; hi[d] = -DBL_MAX;
        if (d == 2 && flat_z) { lo[d] = hi[d] = 0; continue; }
        for (int i = 0; i < n; ++i) {
            const double x = coord[size_t(d) * n + i];
            lo[d] = std::min(lo[d], x); hi[d] = std::max(hi[d], x);
        }
    }
}
struct HostGrid {
    tr::Grid g;
    std::vector<int> start, pts;
    void build(int npts, const double *p, const double lo[3], const double hi[3])
    {
        tr::grid_shape(npts, lo, hi, g);
        g.p = p;
        const size_t ncell = size_t(g.n[0]) * g.n[1] * g.n[2];
        start.assign(ncell + 1, 0); pts.resize(npts);
        for (int i = 0; i < npts; ++i) start[tr::cell_of_point(g, i) + 1]++;
        for (size_t c = 0; c < ncell; ++c) start[c + 1] += start[c];
        std::vector<int> cur(start.begin(), start.end() - 1);
        // filled backwards on purpose: no result may depend on the order inside a cell
        for (int i = npts - 1; i >= 0; --i) pts[cur[tr::cell_of_point(g, i)]++] = i;
        g.cell_start = start.data(); g.cell_pts = pts.data();
    }
};
}

extern "C" {

int hostcheck_transfer_cells(const DesTransferMeshes *m, int is_surface, int nfields,
                             const DesTransferField *fields, int *is_changed_out, DesTransferStats *stats,
                             int *idx_out)
{
    const int npc = is_surface ? 3 : 4, dim = npc - 1;
    const int OC = m->old_ncell, NC = m->new_ncell;
    std::vector<double> pts(size_t(OC) * 3), q(size_t(NC) * 3), coeff(size_t(OC) * 12);
    for (int e = 0; e < OC; ++e) tr::centroid_one(e, npc, OC, m->old_conn, m->old_nnode, m->old_coord, pts.data());
    double lo[3], hi[3];
    bbox(m->old_nnode, m->old_coord, is_surface != 0, lo, hi);
    HostGrid hg; hg.build(OC, pts.data(), lo, hi);
    for (int e = 0; e < OC; ++e) tr::coeff_one(e, dim, OC, m->old_conn, m->old_nnode, m->old_coord, m->old_size, coeff.data());
    for (int e = 0; e < NC; ++e) tr::centroid_one(e, npc, NC, m->new_conn, m->new_nnode, m->new_coord, q.data());
    std::vector<int> idx(NC), changed(NC), idx_changed(NC, 0), list;
    #pragma omp parallel for
    for (int i = 0; i < NC; ++i) tr::nearest_one(i, hg.g, NC, q.data(), idx.data(), changed.data());
    for (int i = 0; i < NC; ++i) if (changed[i] > 0) { idx_changed[i] = int(list.size()); list.push_back(i); }
    const int nchanged = int(list.size());
    const std::vector<double> eta = tr::acm_samples(is_surface != 0);
    const int nsample = int(eta.size() / npc);
    std::vector<int> elems(size_t(nchanged) * tr::KMAX + 1);
    std::vector<double> ratios(size_t(nchanged) * tr::KMAX + 1), empty(nchanged + 1);
    const int max_el = std::min(tr::KMAX, OC);
    #pragma omp parallel for schedule(dynamic, 16)
    for (int i = 0; i < nchanged; ++i) {
        std::vector<int> found(nsample);
        for (int j = 0; j < nsample; ++j)
            found[j] = tr::acm_sample_one(list[i], eta.data() + size_t(j) * npc, npc, NC, m->new_conn,
                                          m->new_nnode, m->new_coord, hg.g, coeff.data(), max_el);
        tr::acm_merge_one(i, list[i], nsample, found.data(), changed.data(), elems.data(), ratios.data(), empty.data());
    }
    for (int f = 0; f < nfields; ++f)
        for (int d = 0; d < fields[f].ncomp; ++d)
            for (int i = 0; i < NC; ++i)
                fields[f].dst[size_t(d) * NC + i] = tr::inject_one(i, fields[f].src + size_t(d) * OC, idx.data(),
                                                                   changed.data(), idx_changed.data(), elems.data(), ratios.data());
    if (is_changed_out) std::memcpy(is_changed_out, changed.data(), sizeof(int) * NC);
    if (idx_out) std::memcpy(idx_out, idx.data(), sizeof(int) * NC);
    if (stats) { std::memset(stats, 0, sizeof(*stats)); stats->nchanged = nchanged; stats->nsample = nsample; }
    return 0;
}

int hostcheck_transfer_nodes(const DesTransferMeshes *m, const int *sup_idx, const int *sup_arr, int nfields,
                             const DesTransferField *fields, int *el_out, double *brc_out, DesTransferStats *stats)
{
    const int ON = m->old_nnode, OC = m->old_ncell, NN = m->new_nnode;
    std::vector<double> coeff(size_t(OC) * 12), brc(size_t(NN) * 4);
    std::vector<int> nn(NN), el(NN), flagged;
    double lo[3], hi[3];
    bbox(ON, m->old_coord, false, lo, hi);
    HostGrid hg; hg.build(ON, m->old_coord, lo, hi);
    for (int e = 0; e < OC; ++e) tr::coeff_one(e, 3, OC, m->old_conn, ON, m->old_coord, m->old_size, coeff.data());
    tr::OldMesh om = { ON, OC, m->old_coord, m->old_conn, sup_idx, sup_arr, coeff.data() };
    for (int i = 0; i < NN; ++i)
        if (!tr::node_locate_one(i, NN, m->new_coord, hg.g, om, nn.data(), el.data(), brc.data())) flagged.push_back(i);
    int n_face = 0, n_out = 0, bad = 0;
    std::vector<int> table(tr::BFS_HCAP), fa(tr::BFS_FCAP), fb(tr::BFS_FCAP);
    for (int i : flagged) {
        std::fill(table.begin(), table.end(), -1);
        const int v = tr::node_bfs_one(i, NN, m->new_coord, om, nn[i], table.data(), fa.data(), fb.data(), el.data(), brc.data());
        bad += v < 0; n_face += v == 1; n_out += v == 2;
    }
    if (bad) return 4;
    for (int f = 0; f < nfields; ++f)
        for (int d = 0; d < fields[f].ncomp; ++d)
            for (int i = 0; i < NN; ++i)
                fields[f].dst[size_t(d) * NN + i] = tr::interp_one(i, fields[f].src + size_t(d) * ON, OC, m->old_conn,
                                                                   el.data(), brc.data());
    if (el_out) std::memcpy(el_out, el.data(), sizeof(int) * NN);
    if (brc_out) std::memcpy(brc_out, brc.data(), sizeof(double) * NN * 4);
    if (stats) { std::memset(stats, 0, sizeof(*stats)); stats->nchanged = int(flagged.size()); stats->n_on_face = n_face; stats->n_outside = n_out; }
    return 0;
}

// brute-force k nearest (ascending by distance^2, then index) against the grid search
int hostcheck_knn(int npts, const double *pts, int nq, const double *q, int k, int *idx_grid, int *idx_brute)
{
    double lo[3], hi[3];
    bbox(npts, pts, false, lo, hi);
    HostGrid hg; hg.build(npts, pts, lo, hi);
    int bad = 0;
    #pragma omp parallel for
    for (int i = 0; i < nq; ++i) {
        const double qq[3] = {q[i], q[size_t(nq) + i], q[size_t(2) * nq + i]};
        int id[tr::KMAX]; double d2[tr::KMAX];
        const int cnt = tr::knn(hg.g, qq, k, id, d2);
        for (int j = 0; j < k; ++j) idx_grid[size_t(i) * k + j] = j < cnt ? id[j] : -1;
        // the fixed-size lists (the kernels use K = 1 for the nearest-neighbour searches) give the same entries
        if (k == 1) {
            tr::ListFixed<1> L; tr::knn_walk(hg.g, qq, L);
            if (L.idx[0] != id[0] || L.d2[0] != d2[0]) {
#pragma omp atomic
                bad++; }
        }
        if (k == tr::ACM_K1 && cnt == k) {
            tr::ListFixed<tr::ACM_K1> L; tr::knn_walk(hg.g, qq, L);
            for (int j = 0; j < k; ++j) if (L.idx[j] != id[j] || L.d2[j] != d2[j]) {
#pragma omp atomic
                bad++; }
        }
        int cb = 0; int ib[tr::KMAX]; double db[tr::KMAX];
        for (int p = 0; p < npts; ++p) {
            const double dx = qq[0] - pts[p], dy = qq[1] - pts[size_t(npts) + p], dz = qq[2] - pts[size_t(2) * npts + p];
            double d = 0; d += dx * dx; d += dy * dy; d += dz * dz;
            tr::knn_insert(k, cb, ib, db, p, d);
        }
        for (int j = 0; j < k; ++j) idx_brute[size_t(i) * k + j] = j < cb ? ib[j] : -1;
    }
    return bad;
}
}

#ifdef HOSTCHECK_AS_ABI
// oracle/Makefile's wrapcheck build only: the two entry points bridge/des_transfer_wrap.cxx calls,
// answered by the host loops above, so that the wrap file's handling of `Variables` can be run
// against the reference where there is no GPU.  The product library never contains this.
extern "C" {
int des_remesh_transfer_cells(int, const DesTransferMeshes *m, int is_surface, int nfields,
                              const DesTransferField *fields, int *is_changed, DesTransferStats *stats)
{
    return hostcheck_transfer_cells(m, is_surface, nfields, fields, is_changed, stats, nullptr);
}
int des_remesh_transfer_nodes(int, const DesTransferMeshes *m, const int *sup_idx, const int *sup_arr, int nfields,
                              const DesTransferField *fields, int *el, double *brc, DesTransferStats *stats)
{
    return hostcheck_transfer_nodes(m, sup_idx, sup_arr, nfields, fields, el, brc, stats);
}
}
#endif
