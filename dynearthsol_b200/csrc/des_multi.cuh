// One handle, several devices (include/des_gpu.h: des_gpu_create_multi).
//
// The reference's driver is one process with one `Variables`; so that it can use the GPUs of a
// box without becoming an MPI program, the partitioning, the ghost exchange and the gather /
// scatter of the host arrays live behind the same C-ABI calls it already makes.  The group
// handle owns one ordinary DesGpu per part; every entry point of des_api.cu that the bridge
// uses forwards to the functions below when the handle is a group.
//
// Scheme (the same as dynearthsol_b200/partition.py, which the NCCL
// one-process-per-GPU harness uses, and tested against it):
//   * elements are split by recursive coordinate bisection of their centroids;
//   * a node is owned by the lowest part among the elements around it;
//   * a part holds every element that touches a node it owns (its own elements plus a
//     one-element layer computed redundantly) and their nodes, both in ascending global order,
//     so an owned node sums its complete support in the reference's ascending element order:
//     results are bit-identical to one GPU and to the reference;
//   * after each nodal barrier the owner's values are copied into the other parts' ghost nodes
//     (cudaMemcpyPeerAsync between the parts' packed buffers, ordered by events) -- pure
//     copies, no reduction, deterministic;
//   * compute_dt's min / max are combined on the host (exact in any order).
// A group may place several parts on ONE device (devices = {0, 0}): that is how the single-GPU
// test box exercises every line of this file.
#pragma once
#include <chrono>
#include <numeric>

struct MultiPart {
    DesGpu *g = nullptr;
    int rank = 0, device = 0;
    std::vector<int> elem_gid, node_gid;        // local -> global, ascending
    std::vector<int> owned_elems, owned_nodes;  // local indices whose values this part reports back
    std::vector<int> facet_gid;                 // local z1 facet -> global z1 facet (edvacc_surf)
    std::vector<int> owned_facets;              // local z1 facets of owned elements
    std::vector<int> nbr;                       // neighbour parts, ascending
    std::vector<int> send_off, recv_off;        // per neighbour, in nodes (size nbr+1)
    std::vector<int> send_nodes, recv_nodes;    // local node indices, concatenated over neighbours
    cudaEvent_t ev_packed = nullptr, ev_unpacked = nullptr;
};

struct DesMulti {
    std::vector<MultiPart> parts;
    int N = 0, E = 0, etop = 0, ntop = 0;
    double partition_seconds = 0, upload_seconds = 0;
    std::vector<double> hbuf;       // host staging for gather / scatter
};

namespace {

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// recursive coordinate bisection of the element centroids: split the longest extent at the
// count-proportional rank (ties by element id), recurse.  part[e] in [0, nparts).
void rcb_split(const double *cx, const double *cy, const double *cz, std::vector<int> &ids, size_t b, size_t e,
               int p0, int np_, std::vector<int> &part)
{
    if (np_ == 1) { for (size_t i = b; i < e; ++i) part[ids[i]] = p0; return; }
    double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (size_t i = b; i < e; ++i) {
        const int k = ids[i];
        const double c[3] = {cx[k], cy[k], cz[k]};
        for (int d = 0; d < 3; ++d) { lo[d] = std::min(lo[d], c[d]); hi[d] = std::max(hi[d], c[d]); }
    }
    int ax = 0;
    for (int d = 1; d < 3; ++d) if (hi[d] - lo[d] > hi[ax] - lo[ax]) ax = d;
    const double *c = ax == 0 ? cx : (ax == 1 ? cy : cz);
    const int left = np_ / 2;
    const size_t k = b + size_t(std::llround(double(e - b) * left / np_));
    // only the split matters: nth_element under the same total order (coordinate, then id)
    if (k > b && k < e)
        std::nth_element(ids.begin() + b, ids.begin() + k, ids.begin() + e,
                         [c](int x, int y) { return c[x] < c[y] || (c[x] == c[y] && x < y); });
    rcb_split(cx, cy, cz, ids, b, k, p0, left, part);
    rcb_split(cx, cy, cz, ids, k, e, p0 + left, np_ - left, part);
}

// RCB of the element centroids into P parts; a node is owned by the lowest part among the
// elements around it (0 for a node on no element).  Pure host integer work (also behind
// des_host_partition, which the CPU tests call).
void partition_elements(int N, int E, const int *conn, const double *coord, int P,
                        std::vector<int> &part, std::vector<int> &owner)
{
    part.assign(E, 0);
    {
        std::vector<double> cx(E), cy(E), cz(E);
        #pragma omp parallel for
        for (int e = 0; e < E; ++e) {
            double s[3] = {0, 0, 0};
            for (int i = 0; i < 4; ++i) {
                const int n = conn[size_t(i) * E + e];
                for (int d = 0; d < 3; ++d) s[d] += coord[size_t(d) * N + n];
            }
            cx[e] = s[0] / 4; cy[e] = s[1] / 4; cz[e] = s[2] / 4;
        }
        std::vector<int> ids(E);
        std::iota(ids.begin(), ids.end(), 0);
        rcb_split(cx.data(), cy.data(), cz.data(), ids, 0, size_t(E), 0, P, part);
    }
    owner.assign(N, INT_MAX);
    for (int i = 0; i < 4; ++i)
        for (int e = 0; e < E; ++e) {
            int &o = owner[conn[size_t(i) * E + e]];
            o = std::min(o, part[e]);
        }
    for (int n = 0; n < N; ++n) if (owner[n] == INT_MAX) owner[n] = 0;
}

int multi_fail(DesGpu *h, int code, const std::string &msg) { return fail(h, code, "multi: " + msg); }

#define MCK(part, call) do { int rc_ = (call); if (rc_) return multi_fail(h, rc_, std::string(#call) + ": " + \
    des_gpu_last_error((part).g)); } while (0)

// owner -> ghosts copy after segment `phase` of the step (des_kernels2.cuh: halo_width)
int multi_exchange(DesGpu *h, int phase)
{
    DesMulti &M = *h->multi;
    const int w = halo_width(phase);
    for (MultiPart &p : M.parts) {
        if (p.send_nodes.empty()) continue;
        CK(cudaSetDevice(p.device));
        // the buffer may still be read by the receivers of the previous exchange
        for (int q : p.nbr) CK(cudaStreamWaitEvent(p.g->stream, M.parts[q].ev_unpacked, 0));
        MCK(p, des_gpu_halo_pack(p.g, phase));
        CK(cudaEventRecord(p.ev_packed, p.g->stream));
    }
    for (MultiPart &p : M.parts) {
        if (p.recv_nodes.empty()) continue;
        CK(cudaSetDevice(p.device));
        for (size_t j = 0; j < p.nbr.size(); ++j) {
            MultiPart &q = M.parts[p.nbr[j]];
            const int cnt = p.recv_off[j + 1] - p.recv_off[j];
            if (cnt == 0) continue;
            size_t k = std::lower_bound(q.nbr.begin(), q.nbr.end(), p.rank) - q.nbr.begin();
            CK(cudaStreamWaitEvent(p.g->stream, q.ev_packed, 0));
            CK(cudaMemcpyPeerAsync(p.g->recv_buf + size_t(p.recv_off[j]) * w, p.device,
                                   q.g->send_buf + size_t(q.send_off[k]) * w, q.device,
                                   size_t(cnt) * w * sizeof(double), p.g->stream));
        }
        MCK(p, des_gpu_halo_unpack(p.g, phase));
        CK(cudaEventRecord(p.ev_unpacked, p.g->stream));
    }
    return DES_OK;
}

void multi_free_parts(DesGpu *h)
{
    if (!h->multi) return;
    for (MultiPart &p : h->multi->parts) {
        cudaSetDevice(p.device);
        if (p.ev_packed) cudaEventDestroy(p.ev_packed);
        if (p.ev_unpacked) cudaEventDestroy(p.ev_unpacked);
        if (p.g) des_gpu_destroy(p.g);
    }
    h->multi->parts.clear();
}

// ---- forwarding bodies of the C-ABI calls (des_api.cu) --------------------------------------

int multi_upload_mesh(DesGpu *h, const DesMesh *m, const double *coord)
{
    DesMulti &M = *h->multi;
    const int P = int(h->devices.size());
    const int N = m->nnode, E = m->nelem;
    if (!coord) return multi_fail(h, DES_E_ARG, "des_gpu_upload_mesh_multi needs the coordinates (RCB of the element centroids)");
    int rc;
    if ((rc = validate_mesh(h, m))) return rc;
    const double t0 = now_s();
    multi_free_parts(h);
    M.N = N; M.E = E; M.etop = m->etop; M.ntop = m->ntop;
    const int *conn = m->connectivity;
    // 1. RCB of the element centroids, 2. node owner = lowest part around it
    std::vector<int> part, owner;
    partition_elements(N, E, conn, coord, P, part, owner);
    // global top tables
    std::vector<int> gtop(N, -1);
    for (int i = 0; i < m->ntop; ++i) gtop[m->top_nodes[i]] = i;
    M.parts.resize(P);
    std::vector<int> n2l(N), e2l(E), f2l(std::max(m->etop, 1)), t2l(std::max(m->ntop, 1));
    std::vector<std::vector<std::vector<int> > > recv_gid(P, std::vector<std::vector<int> >(P));
    struct Local {
        std::vector<int> conn, sidx, sarr, slidx, te, tn, ssidx, ssarr, ean, cs;
        std::vector<double> coord;
        std::vector<unsigned> bc;
        std::vector<int> bf[DES_NBDRY], bn[DES_NBDRY];
    };
    std::vector<Local> loc(P);
    for (int r = 0; r < P; ++r) {
        MultiPart &p = M.parts[r];
        p.rank = r; p.device = h->devices[r];
        Local &L = loc[r];
        // elements touching an owned node, ascending; their nodes, ascending
        std::fill(n2l.begin(), n2l.end(), -1);
        std::fill(e2l.begin(), e2l.end(), -1);
        for (int e = 0; e < E; ++e) {
            bool touch = false;
            for (int i = 0; i < 4; ++i) touch = touch || owner[conn[size_t(i) * E + e]] == r;
            if (!touch) continue;
            e2l[e] = int(p.elem_gid.size());
            p.elem_gid.push_back(e);
            for (int i = 0; i < 4; ++i) n2l[conn[size_t(i) * E + e]] = 0;
        }
        for (int n = 0; n < N; ++n) if (n2l[n] == 0) { n2l[n] = int(p.node_gid.size()); p.node_gid.push_back(n); }
        const int El = int(p.elem_gid.size()), Nl = int(p.node_gid.size());
        if (El == 0) return multi_fail(h, DES_E_ARG, "a part received no element (mesh smaller than the device count)");
        L.conn.resize(size_t(El) * 4);
        for (int le = 0; le < El; ++le) {
            const int e = p.elem_gid[le];
            int mo = INT_MAX;
            for (int i = 0; i < 4; ++i) {
                const int n = conn[size_t(i) * E + e];
                L.conn[size_t(i) * El + le] = n2l[n];
                mo = std::min(mo, owner[n]);
            }
            if (mo == r) p.owned_elems.push_back(le);       // each element reported by exactly one part
        }
        L.bc.resize(Nl);
        L.coord.resize(size_t(Nl) * 3);
        for (int ln = 0; ln < Nl; ++ln) {
            const int n = p.node_gid[ln];
            L.bc[ln] = m->bcflag[n];
            for (int d = 0; d < 3; ++d) L.coord[size_t(d) * Nl + ln] = coord[size_t(d) * N + n];
            if (owner[n] == r) p.owned_nodes.push_back(ln);
            else recv_gid[r][owner[n]].push_back(n);
        }
        L.sidx.resize(size_t(Nl) + 1); L.sarr.resize(size_t(El) * 4); L.slidx.resize(size_t(El) * 4);
        if ((rc = des_host_build_support(Nl, El, L.conn.data(), L.sidx.data(), L.sarr.data(), L.slidx.data())))
            return multi_fail(h, rc, "local support");
        // boundary lists keep the global relative order
        for (int i = 0; i < DES_NBDRY; ++i) {
            for (int k = 0; k < m->nbfacets[i]; ++k) {
                const int le = e2l[m->bfacets[i][2 * k]];
                if (le < 0) continue;
                if (i == 5) { f2l[k] = int(L.bf[i].size() / 2); p.facet_gid.push_back(k); }
                L.bf[i].push_back(le); L.bf[i].push_back(m->bfacets[i][2 * k + 1]);
            }
            for (int k = 0; k < m->nbnodes[i]; ++k) {
                const int ln = n2l[m->bnodes[i][k]];
                if (ln >= 0) L.bn[i].push_back(ln);
            }
        }
        for (int k = 0; k < m->ntop_elems; ++k) { const int le = e2l[m->top_elems[k]]; if (le >= 0) L.te.push_back(le); }
        // surface tables: local top nodes in the global list's order; a facet is local iff its element is
        std::fill(t2l.begin(), t2l.end(), -1);
        for (int i = 0; i < m->ntop; ++i) {
            const int ln = n2l[m->top_nodes[i]];
            if (ln >= 0) { t2l[i] = int(L.tn.size()); L.tn.push_back(ln); }
        }
        const int etl = int(p.facet_gid.size()), ntl = int(L.tn.size());
        std::vector<char> flocal(std::max(m->etop, 1), 0);
        for (int lf = 0; lf < etl; ++lf) flocal[p.facet_gid[lf]] = 1;
        L.ean.resize(size_t(etl) * 3); L.cs.assign(size_t(etl) * 4, 0);
        for (int lf = 0; lf < etl; ++lf) {
            const int k = p.facet_gid[lf];
            for (int j = 0; j < 3; ++j) {
                L.ean[size_t(j) * etl + lf] = t2l[m->elem_and_nodes[size_t(j) * m->etop + k]];
                L.cs[size_t(j) * etl + lf] = n2l[m->connectivity_surface[size_t(j) * m->etop + k]];
            }
            const int le = L.bf[5][2 * lf];
            if (std::binary_search(p.owned_elems.begin(), p.owned_elems.end(), le)) p.owned_facets.push_back(lf);
        }
        L.ssidx.assign(size_t(ntl) + 1, 0);
        for (int i = 0; i < m->ntop; ++i) {
            if (t2l[i] < 0) continue;
            for (int q = m->support_surf_idx[i]; q < m->support_surf_idx[i + 1]; ++q) {
                const int k = m->support_surf_arr[q];
                if (flocal[k]) L.ssarr.push_back(f2l[k]);
            }
            L.ssidx[size_t(t2l[i]) + 1] = int(L.ssarr.size());
        }
        // top nodes come in the global list's order, so the CSR rows were filled in row order
        for (int i = 1; i <= ntl; ++i) L.ssidx[i] = std::max(L.ssidx[i], L.ssidx[i - 1]);
    }
    // 3. halo plan: recv blocks per owner, ascending global id; send blocks are their transpose
    for (int r = 0; r < P; ++r) {
        MultiPart &p = M.parts[r];
        for (int q = 0; q < P; ++q)
            if (q != r && (!recv_gid[r][q].empty() || !recv_gid[q][r].empty())) p.nbr.push_back(q);
        p.send_off.assign(1, 0); p.recv_off.assign(1, 0);
        for (int q : p.nbr) {
            for (int n : recv_gid[r][q])
                p.recv_nodes.push_back(int(std::lower_bound(p.node_gid.begin(), p.node_gid.end(), n) - p.node_gid.begin()));
            for (int n : recv_gid[q][r])
                p.send_nodes.push_back(int(std::lower_bound(p.node_gid.begin(), p.node_gid.end(), n) - p.node_gid.begin()));
            p.recv_off.push_back(int(p.recv_nodes.size()));
            p.send_off.push_back(int(p.send_nodes.size()));
        }
    }
    M.partition_seconds = now_s() - t0;
    // 4. one ordinary handle per part
    const double t1 = now_s();
    for (int r = 0; r < P; ++r) {
        MultiPart &p = M.parts[r];
        Local &L = loc[r];
        if ((rc = des_gpu_create(&h->hp, p.device, &p.g))) {
            std::string msg = p.g ? des_gpu_last_error(p.g) : "no device";
            return multi_fail(h, rc, "des_gpu_create on device " + std::to_string(p.device) + ": " + msg);
        }
        DesMesh lm;
        memset(&lm, 0, sizeof(lm));
        lm.nnode = int(p.node_gid.size()); lm.nelem = int(p.elem_gid.size());
        lm.connectivity = L.conn.data(); lm.bcflag = L.bc.data();
        lm.support_idx = L.sidx.data(); lm.support_arr = L.sarr.data(); lm.support_lidx = L.slidx.data();
        for (int i = 0; i < DES_NBDRY; ++i) {
            lm.nbfacets[i] = int(L.bf[i].size() / 2); lm.bfacets[i] = L.bf[i].data();
            lm.nbnodes[i] = int(L.bn[i].size()); lm.bnodes[i] = L.bn[i].data();
        }
        lm.bnormals = m->bnormals; lm.n_edge_vec = m->n_edge_vec; lm.edge_vec = m->edge_vec; lm.edge_slot = m->edge_slot;
        lm.ntop_elems = int(L.te.size()); lm.top_elems = L.te.data();
        lm.ntop = int(L.tn.size()); lm.top_nodes = L.tn.data();
        lm.etop = int(p.facet_gid.size());
        lm.support_surf_idx = L.ssidx.data(); lm.support_surf_arr = L.ssarr.data();
        lm.elem_and_nodes = L.ean.data(); lm.connectivity_surface = L.cs.data();
        lm.coord = L.coord.data();
        p.g->reorder_mode = h->reorder_mode;    // each part orders its own nodes and elements
        MCK(p, des_gpu_upload_mesh(p.g, &lm));
        MCK(p, des_gpu_set_halo(p.g, int(p.send_nodes.size()), p.send_nodes.data(),
                                int(p.recv_nodes.size()), p.recv_nodes.data()));
        CK(cudaSetDevice(p.device));
        CK(cudaEventCreateWithFlags(&p.ev_packed, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&p.ev_unpacked, cudaEventDisableTiming));
        CK(cudaEventRecord(p.ev_unpacked, p.g->stream));
        // peer access where the hardware has it (cudaMemcpyPeerAsync works either way)
        for (int q = 0; q < P; ++q) {
            const int dq = h->devices[q];
            int can = 0;
            if (dq != p.device && cudaDeviceCanAccessPeer(&can, p.device, dq) == cudaSuccess && can)
                if (cudaDeviceEnablePeerAccess(dq, 0) != cudaSuccess) cudaGetLastError();   // already enabled
        }
    }
    M.upload_seconds = now_s() - t1;
    h->have_mesh = true;
    h->have_state = false;
    h->D.nnode = N; h->D.nelem = E;
    return DES_OK;
}

// rows of a global SoA array [rows][n] restricted to idx (local -> global)
void restrict_rows(const double *g, size_t n, int rows, const std::vector<int> &gid, double *out)
{
    const size_t nl = gid.size();
    for (int r = 0; r < rows; ++r) {
        const double *src = g + size_t(r) * n;
        double *dst = out + size_t(r) * nl;
        #pragma omp parallel for
        for (long long i = 0; i < (long long)nl; ++i) dst[i] = src[gid[i]];
    }
}
// the values a part reports back: local rows [rows][nl] at `own` -> global rows at gid[own]
void scatter_rows(const double *l, size_t nl, int rows, const std::vector<int> &gid, const std::vector<int> &own,
                  size_t n, double *g)
{
    for (int r = 0; r < rows; ++r) {
        const double *src = l + size_t(r) * nl;
        double *dst = g + size_t(r) * n;
        #pragma omp parallel for
        for (long long i = 0; i < (long long)own.size(); ++i) dst[gid[own[i]]] = src[own[i]];
    }
}

struct FieldDesc { double *DesState::*f; int rows; int kind; };    // kind 0 node, 1 element, 2 z1 facet
const FieldDesc kFields[] = {
    {&DesState::coord, 3, 0}, {&DesState::vel, 3, 0}, {&DesState::coord0, 3, 0}, {&DesState::temperature, 1, 0},
    {&DesState::stress, 6, 1}, {&DesState::strain, 6, 1}, {&DesState::strain_rate, 6, 1},
    {&DesState::plstrain, 1, 1}, {&DesState::delta_plstrain, 1, 1}, {&DesState::volume, 1, 1},
    {&DesState::volume_old, 1, 1}, {&DesState::viscosity, 1, 1}, {&DesState::dpressure, 1, 1},
    {&DesState::edvoldt, 1, 1}, {&DesState::radiogenic_source, 1, 1},
    {&DesState::force, 3, 0}, {&DesState::volume_n, 1, 0}, {&DesState::mass, 1, 0}, {&DesState::tmass, 1, 0},
    {&DesState::ntmp, 1, 0}, {&DesState::dhacc, 1, 0}, {&DesState::edvacc_surf, 1, 2},
};

int multi_move_state(DesGpu *h, const DesState *s_in, bool up)
{
    DesMulti &M = *h->multi;
    DesState want = *s_in;
    if (!up) want.coord0 = nullptr;     // the device never changes coord0: des_gpu_download_state leaves it alone
    const DesState *s = &want;
    for (MultiPart &p : M.parts) {
        const size_t nl[3] = {p.node_gid.size(), p.elem_gid.size(), p.facet_gid.size()};
        const size_t ng[3] = {size_t(M.N), size_t(M.E), size_t(M.etop)};
        const std::vector<int> *gid[3] = {&p.node_gid, &p.elem_gid, &p.facet_gid};
        const std::vector<int> *own[3] = {&p.owned_nodes, &p.owned_elems, &p.owned_facets};
        size_t need = 0;
        for (const FieldDesc &fd : kFields) if (s->*(fd.f)) need += size_t(fd.rows) * nl[fd.kind];
        if (M.hbuf.size() < need) M.hbuf.resize(need);
        DesState ls;
        memset(&ls, 0, sizeof(ls));
        size_t off = 0;
        for (const FieldDesc &fd : kFields) {
            if (!(s->*(fd.f))) continue;
            double *buf = M.hbuf.data() + off;
            off += size_t(fd.rows) * nl[fd.kind];
            ls.*(fd.f) = buf;
            if (up) restrict_rows(s->*(fd.f), ng[fd.kind], fd.rows, *gid[fd.kind], buf);
        }
        if (up) { MCK(p, des_gpu_upload_state(p.g, &ls)); continue; }
        MCK(p, des_gpu_download_state(p.g, &ls));
        off = 0;
        for (const FieldDesc &fd : kFields) {
            if (!(s->*(fd.f))) continue;
            const double *buf = M.hbuf.data() + off;
            off += size_t(fd.rows) * nl[fd.kind];
            scatter_rows(buf, nl[fd.kind], fd.rows, *gid[fd.kind], *own[fd.kind], ng[fd.kind], s->*(fd.f));
        }
    }
    if (up) h->have_state = true;
    return DES_OK;
}

int multi_upload_elemmarkers(DesGpu *h, const int *mk)
{
    const int nmat = h->hp.nmat;
    std::vector<int> lm;
    for (MultiPart &p : h->multi->parts) {
        lm.resize(p.elem_gid.size() * size_t(nmat));
        for (size_t le = 0; le < p.elem_gid.size(); ++le)
            for (int k = 0; k < nmat; ++k) lm[le * nmat + k] = mk[size_t(p.elem_gid[le]) * nmat + k];
        MCK(p, des_gpu_upload_elemmarkers(p.g, lm.data()));
    }
    return DES_OK;
}

// des_gpu_step_segment on every part, then the exchange that follows that segment
int multi_step_segment(DesGpu *h, int seg, double dt, double time)
{
    DesMulti &M = *h->multi;
    for (MultiPart &p : M.parts) MCK(p, des_gpu_step_segment(p.g, seg, dt, time));
    const bool mm = h->hp.has_moving_mesh;
    if (seg == 0) return multi_exchange(h, 0);
    if (seg == 1 && h->hp.is_using_mixed_stress) return multi_exchange(h, 1);
    if (seg == 2) return multi_exchange(h, 2);
    if (seg == 3 && mm && h->hp.surface_process_option == 1) return multi_exchange(h, 3);
    return DES_OK;
}

int multi_compute_dt(DesGpu *h, DesDtReductions *out)
{
    bool first = true;
    for (MultiPart &p : h->multi->parts) {
        DesDtReductions r;
        MCK(p, des_gpu_compute_dt(p.g, &r));
        if (first) { *out = r; first = false; continue; }
        out->minl = std::min(out->minl, r.minl); out->dt_maxwell = std::min(out->dt_maxwell, r.dt_maxwell);
        out->dt_diffusion = std::min(out->dt_diffusion, r.dt_diffusion);
        out->global_dt_min = std::min(out->global_dt_min, r.global_dt_min);
        out->global_max_vem = std::max(out->global_max_vem, r.global_max_vem);
        out->max_surf_vel = std::max(out->max_surf_vel, r.max_surf_vel);
    }
    return DES_OK;
}

// bad_mesh_quality's reductions: per part on local ids, combined on global ids (the lowest id
// among equal minima, as one device reports it)
int multi_mesh_quality(DesGpu *h, const DesQualityParam *qp, DesQuality *out)
{
    bool first = true;
    for (MultiPart &p : h->multi->parts) {
        DesQuality q;
        MCK(p, des_gpu_mesh_quality(p.g, qp, &q));
        if (q.first_small_elem >= 0) q.first_small_elem = p.elem_gid[q.first_small_elem];
        if (q.worst_elem >= 0) q.worst_elem = p.elem_gid[q.worst_elem];
        if (q.first_bottom_node >= 0) q.first_bottom_node = p.node_gid[q.first_bottom_node];
        if (q.first_side_node >= 0) q.first_side_node = p.node_gid[q.first_side_node];
        if (first) { *out = q; first = false; continue; }
        auto lower = [](int a, int b) { return a < 0 ? b : (b < 0 ? a : std::min(a, b)); };
        out->first_small_elem = lower(out->first_small_elem, q.first_small_elem);
        if (q.first_bottom_node >= 0 && (out->first_bottom_node < 0 || q.first_bottom_node < out->first_bottom_node)) {
            out->first_bottom_node = q.first_bottom_node; out->bottom_value = q.bottom_value;
        }
        if (q.first_side_node >= 0 && (out->first_side_node < 0 || q.first_side_node < out->first_side_node)) {
            out->first_side_node = q.first_side_node; out->side_value = q.side_value;
        }
        if (q.worst_quality < out->worst_quality || (q.worst_quality == out->worst_quality && q.worst_elem < out->worst_elem)) {
            out->worst_quality = q.worst_quality; out->worst_elem = q.worst_elem;
        }
    }
    return DES_OK;
}

int multi_download_averages(DesGpu *h, double *coord0, double *strain0, double *stress_avg, double *dpl_avg)
{
    DesMulti &M = *h->multi;
    for (MultiPart &p : M.parts) {
        const size_t Nl = p.node_gid.size(), El = p.elem_gid.size();
        if (M.hbuf.size() < 3 * Nl + 13 * El) M.hbuf.resize(3 * Nl + 13 * El);
        double *c = M.hbuf.data(), *s0 = c + 3 * Nl, *sa = s0 + 6 * El, *dp = sa + 6 * El;
        MCK(p, des_gpu_download_averages(p.g, coord0 ? c : nullptr, strain0 ? s0 : nullptr,
                                         stress_avg ? sa : nullptr, dpl_avg ? dp : nullptr));
        if (coord0) scatter_rows(c, Nl, 3, p.node_gid, p.owned_nodes, size_t(M.N), coord0);
        if (strain0) scatter_rows(s0, El, 6, p.elem_gid, p.owned_elems, size_t(M.E), strain0);
        if (stress_avg) scatter_rows(sa, El, 6, p.elem_gid, p.owned_elems, size_t(M.E), stress_avg);
        if (dpl_avg) scatter_rows(dp, El, 1, p.elem_gid, p.owned_elems, size_t(M.E), dpl_avg);
    }
    return DES_OK;
}

} // namespace
