// Host side of the C-ABI in include/des_gpu.h: device state, data movement, the step
// schedule.  No torch, no CPU fallback: every numeric path below launches a kernel from
// des_kernels.cuh on the handle's stream.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cfloat>
#include <climits>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>
#include <algorithm>
#include <limits>
#include <mutex>

#include "des_kernels2.cuh"
#include "des_aux.cuh"

namespace {

enum KClass { KC_ELEM_A = 0, KC_NODE_A, KC_ELEM_B, KC_NODE_B, KC_ELEM_C, KC_NODE_C,
              KC_SBC, KC_SURFACE, KC_DT, KC_PACK, KC_TWIN, KC_COUNT };
const char *kclass_name[DES_NKCLASS] = {
    "elem_a(volume,mass,rotate|Tdiff,strain_rate,dvol)", "node_a(mass,T,dvoldt gather)",
    "elem_b(edvoldt,rheology)", "node_b(dp gather)", "elem_c(NMD,force)",
    "node_c(force gather,bc,damping,v,vbc,x)", "sbc_facets", "surface", "compute_dt",
    "pack/unpack", "twin", "", "", "", "", "" };

struct EvPair { cudaEvent_t a, b; int cls; };

} // namespace

// One captured step (des_gpu_step on small meshes): the host state that selects the launches
// and their baked-in arguments, the executable graph, and the host state the step leaves.
struct StepGraph {
    // key
    bool tail_pending, nmd_pending, scale_pending, ea_prerun, z1_expired;
    int avg_pending;
    const void *volume;         // which of the two volume buffers is current
    double compensation_pressure;
    // value
    cudaGraphExec_t exec;
    bool seg_tail_mass, seg_rot, post_tail_pending, post_nmd, post_scale, post_surf_ran;
    int post_avg;
    long long launches;
    unsigned long long used;
};

struct DesMulti;
struct DesGpu {
    // a group handle (des_gpu_create_multi) owns one ordinary handle per part: des_multi.cuh
    DesMulti *multi = nullptr;
    std::vector<int> devices;
    int device = 0;
    cudaStream_t stream = nullptr;
    DesParam hp;
    DevParams P;
    Dev D;
    std::vector<void *> mesh_allocs;    // freed at upload_mesh / destroy
    std::string err;
    bool have_mesh = false, have_state = false;
    bool tail_pending = false;          // end-of-step element work deferred into next E-A
    double dt_prev = 0;                 // dt of the step whose tail is pending
    double compensation_pressure = 0;
    double last_dt = 0;                 // dt of the last step (for max_surf_vel)
    bool surf_ran = false;
    // multi-GPU ghost exchange (des_gpu_set_halo)
    int nsend = 0, nrecv = 0;
    int *send_idx = nullptr, *recv_idx = nullptr, *top_of_node = nullptr;
    double *send_buf = nullptr, *recv_buf = nullptr;
    bool own_stream = true;
    bool seg_rot = false;               // rotation pending for segment 1 of the current step
    bool seg_tail_mass = false;         // volume/mass of the previous step pending for segment 0
    int n_interior = 0;                 // leading elements that touch no ghost node (des_gpu_set_interior)
    // A/B switches of the three deferrals (DES_DEFER bits 1, 2, 4; all on).  Measured on the
    // 51.4M-tet box, same board, back to back (profiles/r02_defer_ab_50M.json): E-C 1.75 -> 1.64 ms
    // (it no longer rewrites three stress rows), the 0.30 ms correct_surface_element kernel and the
    // 1.4 ms compute_dt kernel disappear, E-B 3.77 -> 4.01 ms and E-A on dt steps 1.5 -> 2.8 ms pay
    // for it; the step as a whole moves by less than the run-to-run spread of a power-capped board
    // (10.35-10.52 ms over the four combinations).
    int defer_ddp = 1, defer_scale = 1, dt_rides = 1;
    // internal element order: DES_ELEM_ORDER=1 stores elements sorted by smallest node index.
    // Off by default: measured on B200 it cuts the lines a warp's gathers touch by 40 % but the
    // element kernels gain < 4 % and the node gathers lose 20-30 % (profiles/experiments/)
    bool reorder = false;
    // DES_REORDER=morton / des_gpu_set_reorder(h, 2): nodes stored in the Morton order of their
    // coordinates, elements in the Morton order of their centroids (needs DesMesh.coord).  On a
    // 1.07M-tet tetgen mesh the step is 11 % faster than in the order the reference delivers
    // (profiles/r02_order_ab_tetgen_1M.json); every array crosses the C-ABI in the caller's order
    // and each node's support keeps the caller's element order, so no result changes.
    int reorder_mode = 0;                           // 0 caller's order, 1 elements by smallest node, 2 Morton
    bool reorder_from_env = false;                  // asked for by DES_REORDER, not by des_gpu_set_reorder
    // Meshes of up to graph_max_elems elements run des_gpu_step as ONE CUDA Graph launch: at 14 k
    // tets the 11-16 kernels of a step take ~35 us on the device and ~100 us to launch one by one
    // (DES_GRAPH=0 turns it off, DES_GRAPH=n sets the element limit).
    long long graph_max_elems = 2000000;
    std::vector<StepGraph> graphs;
    unsigned long long graph_clock = 0;
    double *sc_dev = nullptr;                       // {dt, dt_rot} read by the captured kernels
    double sc_host[2] = {0, 0};
    bool sc_valid = false;
    const int *eperm = nullptr, *einv = nullptr;    // device: internal -> caller, caller -> internal
    std::vector<int> h_einv;                        // host copy of einv (empty = identity)
    const int *nperm = nullptr, *ninv = nullptr;    // the same for nodes
    std::vector<int> h_ninv;
    int pf_elems = DES_PF_DEFAULT;      // L2 prefetch distance of the element kernels (DES_PF_ELEMS overrides)
    long long launches = 0;
    // staging
    void *stage = nullptr; size_t stage_bytes = 0;
    DtPartial *dt_partials = nullptr; int dt_blocks = 0;
    size_t dt_cap = 0;                  // partials the buffer holds; slot [dt_cap] takes k_dt_final's result
    // work of the step that ended which the next E-B applies first (resolve_pending otherwise):
    // driver states of the pseudo-transient loop (dynearthsol.cxx:803-864, 546-591)
    bool pt_jump = false;               // param.control.PT_jump: face velocity values held at 0, no surface processes
    unsigned nbc_mask_full = 0;         // P.nbc_mask when the prescribed tractions are live
    bool nmd_pending = false;           // stress diagonal += ddp[e]       (E-C stored the scalar)
    bool scale_pending = false;         // top elements: stress, strain, plstrain /= rdv (k_surf_correct_vol)
    bool ea_prerun = false;             // des_gpu_compute_dt already ran the next step's E-A (with the dt reductions)
    QPartial *q_partials = nullptr; int *q_nodes = nullptr;     // des_gpu_mesh_quality
    // Output::average_fields accumulators (mesh-lifetime allocations, internal element order)
    double *avg_coord0 = nullptr;   // the element accumulators live in Dev (avg_stress, avg_strain0, avg_dpl)
    bool avg_valid = false;
    // device-resident markers of the material marker set (des_markers.cuh; mesh-lifetime allocations)
    int *mk_elem = nullptr, *mk_mattype = nullptr, *mk_changed = nullptr, *mk_pos = nullptr, *mk_hyd = nullptr,
        *mk_out = nullptr;
    double *mk_eta = nullptr;       // [4][mk_cap]
    char *mk_scan_tmp = nullptr; size_t mk_scan_bytes = 0;
    size_t mk_cap = 0; int mk_n = 0;
    int avg_pending = 0;        // 1 accumulate / 2 snapshot: deferred into the next E-B (with the rotation)
    // timing
    bool timing = false;
    std::vector<EvPair> ev_pool; size_t ev_used = 0;
    double kc_ms[DES_NKCLASS]; long long kc_n[DES_NKCLASS];
    std::unordered_map<std::string, void *> named;
    cudaEvent_t user_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

namespace {

int fail(DesGpu *h, int code, const std::string &msg)
{
    if (h) h->err = msg;
    return code;
}

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(h, DES_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

void drain_events(DesGpu *h)
{
    if (h->ev_used == 0) return;
    cudaStreamSynchronize(h->stream);
    for (size_t i = 0; i < h->ev_used; ++i) {
        float ms = 0;
        cudaEventElapsedTime(&ms, h->ev_pool[i].a, h->ev_pool[i].b);
        h->kc_ms[h->ev_pool[i].cls] += ms;
        h->kc_n[h->ev_pool[i].cls] += 1;
    }
    h->ev_used = 0;
}

struct LaunchScope {
    DesGpu *h; int idx;
    LaunchScope(DesGpu *h_, int cls) : h(h_), idx(-1) {
        h->launches++;
        if (!h->timing) return;
        if (h->ev_used == h->ev_pool.size()) {
            if (h->ev_pool.size() < 4096) {
                EvPair p; cudaEventCreate(&p.a); cudaEventCreate(&p.b); p.cls = 0;
                h->ev_pool.push_back(p);
            } else drain_events(h);
        }
        idx = int(h->ev_used++);
        h->ev_pool[idx].cls = cls;
        cudaEventRecord(h->ev_pool[idx].a, h->stream);
    }
    ~LaunchScope() { if (idx >= 0) cudaEventRecord(h->ev_pool[idx].b, h->stream); }
};

#define LAUNCH(cls, kern, n, bs, ...) do { if ((n) > 0) { LaunchScope ls_(h, cls); auto kfn_ = kern; \
    kfn_<<<DES_GRID((n), (bs)), (bs), 0, h->stream>>>(__VA_ARGS__); } } while (0)

// launch with an explicit grid and dynamic shared memory
#define LAUNCH_G(cls, kern, grid, bs, smem, ...) do { if ((grid) > 0) { LaunchScope ls_(h, cls); auto kfn_ = kern; \
    kfn_<<<(grid), (bs), (smem), h->stream>>>(__VA_ARGS__); } } while (0)

// kernel template<bool SINGLE, bool X>: SINGLE (one material) keeps marker counts in registers
#define LAUNCH_M(cls, kern, X, n, bs, ...) do { if (h->hp.nmat == 1) LAUNCH(cls, (kern<true, X>), n, bs, __VA_ARGS__); \
    else LAUNCH(cls, (kern<false, X>), n, bs, __VA_ARGS__); } while (0)

template <class T>
int dalloc(DesGpu *h, T **p, size_t n, const char *name = nullptr)
{
    *p = nullptr;
    if (n == 0) n = 1;
    void *q = nullptr;
    CK(cudaMalloc(&q, n * sizeof(T)));
    h->mesh_allocs.push_back(q);
    *p = static_cast<T *>(q);
    if (name) h->named[name] = q;
    return DES_OK;
}

int ensure_stage(DesGpu *h, size_t bytes)
{
    if (bytes <= h->stage_bytes) return DES_OK;
    if (h->stage) cudaFree(h->stage);
    h->stage = nullptr; h->stage_bytes = 0;
    CK(cudaMalloc(&h->stage, bytes));
    h->stage_bytes = bytes;
    return DES_OK;
}

template <class T>
int h2d(DesGpu *h, T *dst, const T *src, size_t n)
{
    if (n == 0) return DES_OK;
    CK(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    return DES_OK;
}

template <class T>
int upload_new(DesGpu *h, const T **devp, const T *src, size_t n, const char *name = nullptr)
{
    T *p;
    int rc = dalloc(h, &p, n, name);
    if (rc) return rc;
    if (n && !src) return fail(h, DES_E_ARG, std::string("null host array: ") + (name ? name : "?"));
    rc = h2d(h, p, src, n);
    *devp = p;
    return rc;
}

void free_graphs(DesGpu *h)
{
    for (StepGraph &g : h->graphs) cudaGraphExecDestroy(g.exec);
    h->graphs.clear();
    h->sc_valid = false;
}

void free_mesh(DesGpu *h)
{
    free_graphs(h);
    h->sc_dev = nullptr;
    for (void *p : h->mesh_allocs) cudaFree(p);
    h->mesh_allocs.clear();
    h->named.clear();
    h->nsend = h->nrecv = 0;
    h->send_idx = h->recv_idx = h->top_of_node = nullptr;
    h->send_buf = h->recv_buf = nullptr;
    h->avg_coord0 = nullptr;
    h->avg_valid = false;
    h->mk_elem = h->mk_mattype = h->mk_changed = h->mk_pos = h->mk_hyd = h->mk_out = nullptr;
    h->mk_eta = nullptr; h->mk_scan_tmp = nullptr; h->mk_scan_bytes = 0; h->mk_cap = 0; h->mk_n = 0;
    h->avg_pending = 0;
    h->eperm = h->einv = nullptr;
    h->h_einv.clear();
    h->nperm = h->ninv = nullptr;
    h->h_ninv.clear();
    memset(&h->D, 0, sizeof(h->D));
    h->D.pf = h->pf_elems;
    h->have_mesh = h->have_state = false;
    h->tail_pending = false;
    h->nmd_pending = h->scale_pending = h->ea_prerun = false;
    h->surf_ran = false;
}

int fill_dev_params(DesGpu *h)
{
    const DesParam &p = h->hp;
    DevParams &P = h->P;
    memset(&P, 0, sizeof(P));
    if (p.nmat < 1 || p.nmat > DES_MAX_MAT) return fail(h, DES_E_ARG, "nmat out of range");
    switch (p.rheol_type) {
    case DES_RH_ELASTIC: case DES_RH_MAXWELL: case DES_RH_EP: case DES_RH_EVP: break;
    default: return fail(h, DES_E_UNSUPPORTED, "rheology type outside the hot-path scope "
                         "(viscous / rate-and-state / plane strain)");
    }
    if (p.damping_option < 0 || p.damping_option > 4)
        return fail(h, DES_E_UNSUPPORTED, "damping_option must be 0..4");
    if (p.surface_process_option != 0 && p.surface_process_option != 1)
        return fail(h, DES_E_UNSUPPORTED, "surface_process_option must be 0 or 1 in 3D");
    if (p.ref_pressure_option < 0 || p.ref_pressure_option > 2)
        return fail(h, DES_E_UNSUPPORTED, "ref_pressure_option must be 0..2");
    for (int i = 0; i < 6; ++i)
        if (p.stress_bc_types[i] < 0 || p.stress_bc_types[i] > 3)
            return fail(h, DES_E_ARG, "stress_bc_types must be 0..3");

    P.gravity = p.gravity; P.damping_factor = p.damping_factor;
    P.surf_diff = p.surface_diffusivity; P.surf_base_level = p.surf_base_level;
    P.winkler_delta_rho = p.winkler_delta_rho;
    P.elastic_foundation_constant = p.elastic_foundation_constant;
    P.sea_water_density = p.sea_water_density;
    P.surface_temperature = p.surface_temperature; P.zlength = p.zlength;
    P.pseudo_speed = p.max_vbc_val * p.inertial_scaling;
    P.visc_min = p.visc_min; P.visc_max = p.visc_max; P.tension_max = p.tension_max;
    P.therm_diff_max = p.therm_diff_max;
    for (int i = 0; i < DES_NBDRY; ++i) { P.vbc_values[i] = p.vbc_values[i]; P.vbc_types[i] = p.vbc_types[i]; }
    for (int i = 0; i < 4; ++i) P.vbc_val_lateral[i] = p.vbc_val_lateral[i];
    P.vbc_val_z1_loading_period = p.vbc_val_z1_loading_period;
    P.damping_option = p.damping_option; P.is_quasi_static = p.is_quasi_static;
    P.has_thermal_diffusion = p.has_thermal_diffusion;
    P.is_using_mixed_stress = p.is_using_mixed_stress;
    P.surface_process_option = p.surface_process_option;
    P.ref_pressure_option = p.ref_pressure_option;
    P.has_winkler_foundation = p.has_winkler_foundation;
    P.has_water_loading = p.has_water_loading;
    // the foundation force lives inside apply_stress_bcs, which returns at its first statement
    // without gravity (bc.cxx:667): no gravity, no foundation force
    P.has_elastic_foundation = p.has_elastic_foundation && p.gravity != 0;
    P.rheol_type = p.rheol_type; P.nmat = p.nmat; P.mattype_ref = p.mattype_ref;
    for (int i = 0; i < 6; ++i) {
        P.stress_bc_types[i] = p.stress_bc_types[i];
        P.stress_bc_values[i] = p.stress_bc_values[i];
    }
    int nmemo = 0;
    for (int m = 0; m < p.nmat; ++m) {
        P.rho0[m] = p.rho0[m]; P.alpha[m] = p.alpha[m];
        P.bulk_modulus[m] = p.bulk_modulus[m]; P.shear_modulus[m] = p.shear_modulus[m];
        // matprops.cxx:236-252, evaluated with the HOST libm like the reference does
        const double gas_constant = 8.3144;
        P.visc_pow_edot[m] = 1 / p.visc_exponent[m] - 1;
        const double pow1 = -1 / p.visc_exponent[m];
        P.visc_coef_term[m] = std::pow(0.75 * p.visc_coefficient[m], pow1);
        P.visc_nR[m] = p.visc_exponent[m] * gas_constant;
        P.visc_activation_energy[m] = p.visc_activation_energy[m];
        P.visc_activation_volume[m] = p.visc_activation_volume[m];
        P.heat_capacity[m] = p.heat_capacity[m]; P.therm_cond[m] = p.therm_cond[m];
        P.pls0[m] = p.pls0[m]; P.pls1[m] = p.pls1[m];
        P.cohesion0[m] = p.cohesion0[m]; P.cohesion1[m] = p.cohesion1[m];
        P.friction_angle0[m] = p.friction_angle0[m]; P.friction_angle1[m] = p.friction_angle1[m];
        P.dilation_angle0[m] = p.dilation_angle0[m]; P.dilation_angle1[m] = p.dilation_angle1[m];
        P.porosity[m] = p.porosity[m];
        const double ang[4] = { p.friction_angle0[m], p.friction_angle1[m],
                                p.dilation_angle0[m], p.dilation_angle1[m] };
        for (int a = 0; a < 4; ++a) {
            bool dup = false;
            for (int i = 0; i < nmemo; ++i) if (P.memo_angle[i] == ang[a]) dup = true;
            if (dup) continue;
            P.memo_angle[nmemo] = ang[a];
            P.memo_sin[nmemo] = std::sin(ang[a] * (M_PI / 180));
            P.memo_tan[nmemo] = std::tan(ang[a] * (M_PI / 180));
            nmemo++;
        }
    }
    P.nmemo = nmemo;
    return DES_OK;
}

// Builds the per-boundary-node lists apply_stress_bcs walks (bc.cxx:681-812): for each
// live boundary, each boundary node, each support element (ascending) that owns a facet
// on that boundary, the (facet, local facet node) whose traction the node receives.
int build_stress_bcs(DesGpu *h, const DesMesh *m)
{
    const DesParam &p = h->hp;
    Dev &D = h->D;
    std::vector<int> felem, ffacet, fbdry;
    std::map<int, std::vector<int> > rows;
    unsigned mask = 0;
    const int E = m->nelem;
    if (p.gravity != 0) {
        for (int i = 0; i < DES_NBDRY; ++i) {
            if (p.vbc_types[i] != 0 && p.vbc_types[i] != 2 && p.vbc_types[i] != 4) continue;
            if (i == 4 && !p.has_winkler_foundation) continue;
            if (i == 5 && !p.has_water_loading) continue;
            const int nf = m->nbfacets[i];
            if (nf == 0) continue;
            mask |= 1u << i;
            const int base = int(felem.size());
            std::unordered_map<int, int> emap;
            emap.reserve(size_t(nf) * 2);
            for (int n = 0; n < nf; ++n) {
                const int e = m->bfacets[i][2*n], f = m->bfacets[i][2*n + 1];
                felem.push_back(e); ffacet.push_back(f); fbdry.push_back(i);
                emap[e] = n;    // sequential order: the last facet of an element wins (bc.cxx:769)
            }
            static const int nof[4][3] = {{1,2,3},{0,3,2},{0,1,3},{0,2,1}};
            for (int j = 0; j < m->nbnodes[i]; ++j) {
                const int n = m->bnodes[i][j];
                for (int k = m->support_idx[n]; k < m->support_idx[n + 1]; ++k) {
                    const int e = m->support_arr[k];
                    auto it = emap.find(e);
                    if (it == emap.end()) continue;
                    const int ib = it->second;
                    const int f = m->bfacets[i][2*ib + 1];
                    for (int l = 0; l < 3; ++l) {
                        if (n == m->connectivity[size_t(nof[f][l]) * E + e]) {
                            rows[n].push_back((base + ib) * 3 + l);
                            break;
                        }
                    }
                }
            }
        }
    }
    h->P.sbc_mask = mask;
    D.nsbc_facets = int(felem.size());
    int rc;
    if (!h->h_einv.empty()) for (int &e : felem) e = h->h_einv[e];
    if ((rc = upload_new(h, &D.sbc_felem, felem.data(), felem.size()))) return rc;
    if ((rc = upload_new(h, &D.sbc_ffacet, ffacet.data(), ffacet.size()))) return rc;
    if ((rc = upload_new(h, &D.sbc_fbdry, fbdry.data(), fbdry.size()))) return rc;
    if ((rc = dalloc(h, &D.sbc_ftr, felem.size() * 9))) return rc;
    std::vector<int> idx(1, 0), ent, row;
    if (!rows.empty()) {
        row.assign(m->nnode, -1);
        int r = 0;
        for (auto &kv : rows) {
            row[h->h_ninv.empty() ? kv.first : h->h_ninv[kv.first]] = r++;
            ent.insert(ent.end(), kv.second.begin(), kv.second.end());
            idx.push_back(int(ent.size()));
        }
    }
    if ((rc = upload_new(h, &D.sbc_row, row.data(), row.size()))) return rc;
    if ((rc = upload_new(h, &D.sbc_idx, idx.data(), idx.size()))) return rc;
    if ((rc = upload_new(h, &D.sbc_ent, ent.data(), ent.size()))) return rc;
    CK(cudaStreamSynchronize(h->stream));   // the std::vectors above die here
    return DES_OK;
}

// apply_stress_bcs_neumann (bc.cxx:829-912) is a serial loop: boundaries 0..5 with a traction
// type, their facets in list order, the three nodes of each facet: `force[node] += ...`.  Per
// node that is a fixed sequence of additions; the lists below keep it (boundary, then facet).
int build_neumann_bcs(DesGpu *h, const DesMesh *m)
{
    Dev &D = h->D;
    const int E = m->nelem;
    std::vector<int> felem, ffacet, fbdry;
    std::map<int, std::vector<int> > rows;
    unsigned mask = 0;
    static const int nof[4][3] = {{1,2,3},{0,3,2},{0,1,3},{0,2,1}};
    for (int i = 0; i < 6; ++i) {
        if (h->P.stress_bc_types[i] == 0 || m->nbfacets[i] == 0) continue;
        mask |= 1u << i;
        for (int n = 0; n < m->nbfacets[i]; ++n) {
            const int e = m->bfacets[i][2*n], f = m->bfacets[i][2*n + 1];
            const int k = int(felem.size());
            felem.push_back(e); ffacet.push_back(f); fbdry.push_back(i);
            for (int j = 0; j < 3; ++j) rows[m->connectivity[size_t(nof[f][j]) * E + e]].push_back(k);
        }
    }
    // fields.cxx:690: the prescribed tractions are skipped while the body-force adjustment runs
    // (des_gpu_set_body_force_adjustment flips it back)
    h->nbc_mask_full = mask;
    h->P.nbc_mask = h->hp.has_body_force_adjustment ? 0u : mask;
    D.nnbc_facets = int(felem.size());
    int rc;
    if (!h->h_einv.empty()) for (int &e : felem) e = h->h_einv[e];
    if ((rc = upload_new(h, &D.nbc_felem, felem.data(), felem.size()))) return rc;
    if ((rc = upload_new(h, &D.nbc_ffacet, ffacet.data(), ffacet.size()))) return rc;
    if ((rc = upload_new(h, &D.nbc_fbdry, fbdry.data(), fbdry.size()))) return rc;
    if ((rc = dalloc(h, &D.nbc_ftr, felem.size() * 3))) return rc;
    std::vector<int> idx(1, 0), ent, row;
    if (!rows.empty()) {
        row.assign(m->nnode, -1);
        int r = 0;
        for (auto &kv : rows) {
            row[h->h_ninv.empty() ? kv.first : h->h_ninv[kv.first]] = r++;
            ent.insert(ent.end(), kv.second.begin(), kv.second.end());
            idx.push_back(int(ent.size()));
        }
    }
    if ((rc = upload_new(h, &D.nbc_row, row.data(), row.size()))) return rc;
    if ((rc = upload_new(h, &D.nbc_idx, idx.data(), idx.size()))) return rc;
    if ((rc = upload_new(h, &D.nbc_ent, ent.data(), ent.size()))) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

// Every index list of DesMesh becomes a gather/scatter index on the device: a malformed mesh
// (e.g. from a faulty remesh) must come back as DES_E_ARG, not as a sticky CUDA fault.
int validate_mesh(DesGpu *h, const DesMesh *m)
{
    const int N = m->nnode, E = m->nelem;
    if (!m->connectivity || !m->bcflag || !m->support_idx || !m->support_arr || !m->support_lidx)
        return fail(h, DES_E_ARG, "null topology array");
    const size_t n4 = size_t(E) * 4;
    bool ok = true;
    #pragma omp parallel for reduction(&&:ok)
    for (long long k = 0; k < (long long)n4; ++k) {
        ok = ok && (unsigned)m->connectivity[k] < (unsigned)N;
        ok = ok && (unsigned)m->support_arr[k] < (unsigned)E && (unsigned)m->support_lidx[k] < 4u;
    }
    if (!ok) return fail(h, DES_E_ARG, "connectivity / support entry out of range");
    if (m->support_idx[0] != 0) return fail(h, DES_E_ARG, "support_idx[0] != 0");
    for (int n = 0; n < N; ++n)
        if (m->support_idx[n + 1] < m->support_idx[n]) return fail(h, DES_E_ARG, "support_idx not monotone");
    for (int i = 0; i < DES_NBDRY; ++i) {
        if (m->nbfacets[i] < 0 || m->nbnodes[i] < 0 || (m->nbfacets[i] && !m->bfacets[i]) || (m->nbnodes[i] && !m->bnodes[i]))
            return fail(h, DES_E_ARG, "bad boundary list");
        for (int k = 0; k < m->nbfacets[i]; ++k)
            if ((unsigned)m->bfacets[i][2*k] >= (unsigned)E || (unsigned)m->bfacets[i][2*k + 1] >= 4u)
                return fail(h, DES_E_ARG, "boundary facet out of range");
        for (int k = 0; k < m->nbnodes[i]; ++k)
            if ((unsigned)m->bnodes[i][k] >= (unsigned)N) return fail(h, DES_E_ARG, "boundary node out of range");
    }
    if (m->ntop < 0 || m->etop < 0 || m->ntop_elems < 0) return fail(h, DES_E_ARG, "negative surface count");
    if ((m->ntop_elems && !m->top_elems) || (m->ntop && !m->top_nodes) ||
        (m->etop && (!m->elem_and_nodes || !m->connectivity_surface || !m->support_surf_arr)))
        return fail(h, DES_E_ARG, "null surface array");
    for (int k = 0; k < m->ntop_elems; ++k)
        if ((unsigned)m->top_elems[k] >= (unsigned)E) return fail(h, DES_E_ARG, "top element out of range");
    for (int k = 0; k < m->ntop; ++k)
        if ((unsigned)m->top_nodes[k] >= (unsigned)N) return fail(h, DES_E_ARG, "top node out of range");
    if (m->ntop > 0) {
        if (!m->support_surf_idx || m->support_surf_idx[0] != 0) return fail(h, DES_E_ARG, "bad support_surf_idx");
        for (int k = 0; k < m->ntop; ++k)
            if (m->support_surf_idx[k + 1] < m->support_surf_idx[k]) return fail(h, DES_E_ARG, "support_surf_idx not monotone");
        for (int k = 0; k < m->support_surf_idx[m->ntop]; ++k)
            if ((unsigned)m->support_surf_arr[k] >= (unsigned)m->etop) return fail(h, DES_E_ARG, "support_surf entry out of range");
    }
    for (size_t k = 0; k < size_t(m->etop) * 3; ++k) {
        if ((unsigned)m->elem_and_nodes[k] >= (unsigned)m->ntop) return fail(h, DES_E_ARG, "elem_and_nodes entry out of range");
        if ((unsigned)m->connectivity_surface[k] >= (unsigned)N) return fail(h, DES_E_ARG, "connectivity_surface entry out of range");
    }
    if (m->edge_slot)
        for (int k = 0; k < 100; ++k)
            if (m->edge_slot[k] >= 0 && (m->edge_slot[k] + 1) * 3 > m->n_edge_vec)
                return fail(h, DES_E_ARG, "edge_slot points past edge_vec");
    return DES_OK;
}

// ---- schedule pieces -----------------------------------------------------------------

// tail_flags of k_elem_a for the fused step: volume+mass when the mesh moves
// (dynearthsol.cxx:871-872), rotation when the rheology has the elastic bit (:875-876)
int tail_flags(const DesGpu *h)
{
    return (h->hp.has_moving_mesh ? 3 : 0) | ((h->hp.rheol_type & DES_RH_ELASTIC) ? 4 : 0);
}
int head_flags(const DesGpu *h) { return (h->hp.has_thermal_diffusion ? 1 : 0) | 2 | 4; }

// Completes the deferred end-of-step element work (volume, mass, rotation) on its own.
// compute_mass's ymass column (geometry.cxx:1833,1850-1866), needed only by Rayleigh damping:
// runs wherever the mass is recomputed.  D.etmp is free at all of those points.
int ymass_update(DesGpu *h)
{
    if (h->hp.damping_option != 4) return DES_OK;
    LAUNCH(KC_TWIN, k_ymass_elem, h->D.nelem, 256, h->P, h->D, h->D.etmp);
    LAUNCH(KC_TWIN, k_ymass_node, h->D.nnode, 256, h->D, (const double *)h->D.etmp);
    CK(cudaGetLastError());
    return DES_OK;
}

// Output::average_fields on the current device state (output.cxx:317-350), as its own kernels
int average_now(DesGpu *h)
{
    if (!h->avg_pending) return DES_OK;
    Dev &D = h->D;
    const size_t E = D.nelem;
    if (h->avg_pending == 2) {
        CK(cudaMemcpyAsync(D.avg_strain0, D.strain, E * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        CK(cudaMemcpyAsync(D.avg_stress, D.stress, E * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        CK(cudaMemcpyAsync(D.avg_dpl, D.delta_plstrain, E * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    } else {
        const size_t n6 = E * 6;
        { LaunchScope ls(h, KC_PACK);
          k_avg_accumulate<<<unsigned((n6 + 255) / 256), 256, 0, h->stream>>>(n6, D.stress, D.avg_stress); }
        { LaunchScope ls(h, KC_PACK);
          k_avg_accumulate<<<unsigned((E + 255) / 256), 256, 0, h->stream>>>(E, D.delta_plstrain, D.avg_dpl); }
    }
    h->avg_pending = 0;
    CK(cudaGetLastError());
    return DES_OK;
}

// Applies what the fused schedule leaves for the next E-B to the stored fields, in the
// reference's order: NMD_stress's diagonal correction (geometry.cxx:315-331), then
// correct_surface_element's divisions (bc.cxx:1678-1685).
int resolve_pending(DesGpu *h)
{
    if (h->nmd_pending) LAUNCH(KC_ELEM_C, k_apply_ddp, h->D.nelem, 256, h->D);
    if (h->scale_pending) LAUNCH(KC_SURFACE, k_surf_apply_scale, h->D.ntop_elems, 128, h->D);
    h->nmd_pending = h->scale_pending = false;
    CK(cudaGetLastError());
    return DES_OK;
}

int flush_tail(DesGpu *h)
{
    h->ea_prerun = false;       // whoever flushes may change what E-A read
    { int rc = resolve_pending(h); if (rc) return rc; }
    if (!h->tail_pending) return average_now(h);
    const int mm = h->hp.has_moving_mesh;
    LAUNCH(KC_ELEM_A, (k_elem_a<true, false>), h->D.nelem, 256, h->P, h->D, h->dt_prev, tail_flags(h), 0);
    if (mm) LAUNCH(KC_NODE_A, (k_node_a<true, false>), h->D.nnode, 256, h->P, h->D, 0.0, 0);
    if (mm) { int rc = ymass_update(h); if (rc) return rc; }
    h->tail_pending = false;
    CK(cudaGetLastError());
    return average_now(h);      // a deferred average_fields wanted exactly this (rotated) state
}

int run_surface(DesGpu *h, double dt, bool with_volume_n)
{
    Dev &D = h->D;
    if (h->hp.surface_process_option == 1 && D.ntop > 0) {
        CK(cudaMemsetAsync(D.maxdh_bits, 0, sizeof(unsigned long long), h->stream));
        LAUNCH(KC_SURFACE, k_surf_facets, D.etop, 128, D);
        LAUNCH(KC_SURFACE, k_surf_nodes, D.ntop, 128, h->P, D, dt);
        LAUNCH(KC_SURFACE, k_surf_edvacc, D.etop, 128, D);
        h->surf_ran = true;
    }
    LAUNCH(KC_SURFACE, k_surf_correct_elems, D.ntop_elems, 128, D);
    if (with_volume_n) LAUNCH(KC_SURFACE, k_surf_correct_nodes, D.ntop, 128, D);
    return DES_OK;
}

} // namespace

// forward declarations the group layer calls back into
extern "C" {
int des_gpu_create(const DesParam *param, int device, DesGpu **out);
int des_gpu_destroy(DesGpu *h);
const char *des_gpu_last_error(const DesGpu *h);
int des_gpu_upload_mesh(DesGpu *h, const DesMesh *m);
int des_gpu_upload_state(DesGpu *h, const DesState *s);
int des_gpu_download_state(DesGpu *h, const DesState *s);
int des_gpu_upload_elemmarkers(DesGpu *h, const int *elemmarkers);
int des_gpu_step_segment(DesGpu *h, int seg, double dt, double time);
int des_gpu_compute_dt(DesGpu *h, DesDtReductions *out);
int des_gpu_mesh_quality(DesGpu *h, const DesQualityParam *qp, DesQuality *out);
int des_gpu_download_averages(DesGpu *h, double *coord0, double *strain0, double *stress_avg, double *dpl_avg);
int des_gpu_set_halo(DesGpu *h, int nsend, const int *send_nodes, int nrecv, const int *recv_nodes);
int des_gpu_halo_pack(DesGpu *h, int phase);
int des_gpu_halo_unpack(DesGpu *h, int phase);
int des_host_build_support(int nnode, int nelem, const int *conn, int *idx, int *arr, int *lidx);
}
#include "des_multi.cuh"

// =========================================================================================
extern "C" {

int des_gpu_create_multi(const DesParam *param, int ndev, const int *devices, DesGpu **out)
{
    if (!param || !out || ndev < 1 || ndev > 64) return DES_E_ARG;
    *out = nullptr;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) return DES_E_NODEVICE;
    DesGpu *h = new DesGpu;
    h->hp = *param;
    h->own_stream = false;
    memset(&h->D, 0, sizeof(h->D));
    memset(h->kc_ms, 0, sizeof(h->kc_ms));
    memset(h->kc_n, 0, sizeof(h->kc_n));
    *out = h;
    for (int i = 0; i < ndev; ++i) {
        const int d = devices ? devices[i] : i;
        if (d < 0 || d >= have) return fail(h, DES_E_NODEVICE, "device index out of range");
        h->devices.push_back(d);
    }
    h->device = h->devices[0];
    h->multi = new DesMulti;
    return fill_dev_params(h);      // the same option checks as one device
}

// The partitioner of a multi-device handle on its own (no GPU): part of every element, owner of
// every node, and per part the number of local elements (its own + the redundant layer), local
// nodes and ghost nodes -- what des_gpu_upload_mesh will build.
int des_host_partition(int nnode, int nelem, const int *connectivity, const double *coord, int nparts,
                       int *elem_part, int *node_owner, int *local_elems, int *local_nodes, int *ghost_nodes)
{
    if (nnode <= 0 || nelem <= 0 || !connectivity || !coord || nparts < 1 || nparts > 64) return DES_E_ARG;
    const size_t E = nelem;
    for (size_t k = 0; k < E * 4; ++k) if ((unsigned)connectivity[k] >= (unsigned)nnode) return DES_E_ARG;
    std::vector<int> part, owner;
    partition_elements(nnode, nelem, connectivity, coord, nparts, part, owner);
    if (elem_part) std::copy(part.begin(), part.end(), elem_part);
    if (node_owner) std::copy(owner.begin(), owner.end(), node_owner);
    if (local_elems || local_nodes || ghost_nodes) {
        std::vector<char> mark(nnode);
        for (int r = 0; r < nparts; ++r) {
            std::fill(mark.begin(), mark.end(), 0);
            int ne = 0, nn = 0, ng = 0;
            for (size_t e = 0; e < E; ++e) {
                bool touch = false;
                for (int i = 0; i < 4; ++i) touch = touch || owner[connectivity[i * E + e]] == r;
                if (!touch) continue;
                ++ne;
                for (int i = 0; i < 4; ++i) mark[connectivity[i * E + e]] = 1;
            }
            for (int n = 0; n < nnode; ++n) if (mark[n]) { ++nn; if (owner[n] != r) ++ng; }
            if (local_elems) local_elems[r] = ne;
            if (local_nodes) local_nodes[r] = nn;
            if (ghost_nodes) ghost_nodes[r] = ng;
        }
    }
    return DES_OK;
}

int des_gpu_partition_info(DesGpu *h, int *nparts, double *partition_seconds, double *upload_seconds,
                           int *local_elems, int *local_nodes, int *ghost_nodes)
{
    if (!h) return DES_E_ARG;
    if (!h->multi) {
        if (nparts) *nparts = 1;
        if (partition_seconds) *partition_seconds = 0;
        if (upload_seconds) *upload_seconds = 0;
        if (local_elems) local_elems[0] = h->D.nelem;
        if (local_nodes) local_nodes[0] = h->D.nnode;
        if (ghost_nodes) ghost_nodes[0] = h->nrecv;
        return DES_OK;
    }
    const DesMulti &M = *h->multi;
    if (nparts) *nparts = int(M.parts.size());
    if (partition_seconds) *partition_seconds = M.partition_seconds;
    if (upload_seconds) *upload_seconds = M.upload_seconds;
    for (size_t r = 0; r < M.parts.size(); ++r) {
        if (local_elems) local_elems[r] = int(M.parts[r].elem_gid.size());
        if (local_nodes) local_nodes[r] = int(M.parts[r].node_gid.size());
        if (ghost_nodes) ghost_nodes[r] = int(M.parts[r].recv_nodes.size());
    }
    return DES_OK;
}

int des_gpu_create(const DesParam *param, int device, DesGpu **out)
{
    if (!param || !out) return DES_E_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev)
        return DES_E_NODEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return DES_E_NODEVICE;
    DesGpu *h = new DesGpu;
    h->device = device;
    h->hp = *param;
    if (const char *sv = getenv("DES_ELEM_ORDER")) h->reorder_mode = atoi(sv) != 0 ? 1 : 0;
    if (const char *sv = getenv("DES_REORDER")) {
        h->reorder_mode = (!strcmp(sv, "morton") || atoi(sv) == 2) ? 2 : (atoi(sv) == 1 ? 1 : 0);
        h->reorder_from_env = true;
    }
    if (const char *sv = getenv("DES_DEFER")) {     // bit 1 ddp, 2 surface scale, 4 compute_dt on E-A
        const int v = atoi(sv);
        h->defer_ddp = v & 1; h->defer_scale = (v >> 1) & 1; h->dt_rides = (v >> 2) & 1;
    }
    if (const char *sv = getenv("DES_PF_ELEMS")) h->pf_elems = atoi(sv) > 0 ? atoi(sv) : 0;
    if (const char *sv = getenv("DES_GRAPH")) h->graph_max_elems = atoll(sv);
    memset(&h->D, 0, sizeof(h->D));
    h->D.pf = h->pf_elems;
    memset(h->kc_ms, 0, sizeof(h->kc_ms));
    memset(h->kc_n, 0, sizeof(h->kc_n));
    int rc = fill_dev_params(h);
    if (rc == DES_OK && cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess)
        rc = fail(h, DES_E_CUDA, "cudaStreamCreate failed");
    *out = h;   // returned even on failure so the caller can read last_error, then destroy
    return rc;
}

int des_gpu_destroy(DesGpu *h)
{
    if (!h) return DES_OK;
    if (h->multi) { multi_free_parts(h); delete h->multi; h->multi = nullptr; }
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    free_mesh(h);
    if (h->stage) cudaFree(h->stage);
    if (h->dt_partials) cudaFree(h->dt_partials);
    if (h->q_partials) cudaFree(h->q_partials);
    if (h->q_nodes) cudaFree(h->q_nodes);
    for (auto &p : h->ev_pool) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto &e : h->user_ev) if (e) cudaEventDestroy(e);
    if (h->stream && h->own_stream) cudaStreamDestroy(h->stream);
    delete h;
    return DES_OK;
}

const char *des_gpu_last_error(const DesGpu *h) { return h ? h->err.c_str() : "null handle"; }

int des_gpu_upload_mesh(DesGpu *h, const DesMesh *m)
{
    if (!h || !m) return DES_E_ARG;
    if (m->nnode <= 0 || m->nelem <= 0) return fail(h, DES_E_ARG, "empty mesh");
    if (size_t(m->nelem) * 4 >= (size_t(1) << 31)) return fail(h, DES_E_ARG, "nelem*4 overflows int32");
    if (h->multi) return multi_upload_mesh(h, m, m->coord);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    free_mesh(h);
    Dev &D = h->D;
    const int N = m->nnode, E = m->nelem;
    D.nnode = N; D.nelem = E;
    D.e0 = 0; D.e1 = E;
    h->n_interior = E;
    int rc;
    if ((rc = validate_mesh(h, m))) return rc;
    // internal element order: stable counting sort by smallest node index -- consecutive
    // elements then share nodes, so a warp's gathers touch ~40 % fewer 128-byte lines than in
    // the mesh generators' orders (measured: profiles/)
    int mode = h->reorder_mode;
    if (mode == 2 && !m->coord) {
        // an explicit request cannot be met; the environment's blanket request just does not apply
        if (!h->reorder_from_env) return fail(h, DES_E_ARG, "Morton order needs DesMesh.coord");
        mode = 0;
    }
    h->reorder = mode != 0;
    if (mode == 2) {
        // 21 bits per axis of the bounding box; ties by the caller's id (std::sort on pairs)
        const double *x = m->coord;
        double lo[3], hi[3];
        for (int d = 0; d < 3; ++d) {
            lo[d] = hi[d] = x[size_t(d) * N];
            for (int n = 1; n < N; ++n) { lo[d] = std::min(lo[d], x[size_t(d) * N + n]); hi[d] = std::max(hi[d], x[size_t(d) * N + n]); }
        }
        auto spread = [](unsigned long long v) {
            v &= 0x1fffffull;
            v = (v | (v << 32)) & 0x1f00000000ffffull;
            v = (v | (v << 16)) & 0x1f0000ff0000ffull;
            v = (v | (v << 8)) & 0x100f00f00f00f00full;
            v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
            v = (v | (v << 2)) & 0x1249249249249249ull;
            return v;
        };
        auto code = [&](const double p[3]) {
            unsigned long long q[3];
            for (int d = 0; d < 3; ++d) {
                const double w = hi[d] - lo[d];
                double t = w > 0 ? (p[d] - lo[d]) / w : 0.0;
                t = std::min(std::max(t, 0.0), 1.0);
                q[d] = (unsigned long long)(t * 2097151.0);
            }
            return spread(q[0]) | (spread(q[1]) << 1) | (spread(q[2]) << 2);
        };
        std::vector<std::pair<unsigned long long, int> > kn(N), ke(E);
        #pragma omp parallel for
        for (int n = 0; n < N; ++n) {
            const double p[3] = {x[n], x[size_t(N) + n], x[size_t(2) * N + n]};
            kn[n] = std::make_pair(code(p), n);
        }
        const int *c = m->connectivity;
        #pragma omp parallel for
        for (int e = 0; e < E; ++e) {
            double p[3] = {0, 0, 0};
            for (int j = 0; j < 4; ++j) {
                const int n = c[size_t(j) * E + e];
                for (int d = 0; d < 3; ++d) p[d] += x[size_t(d) * N + n];
            }
            for (int d = 0; d < 3; ++d) p[d] /= 4;
            ke[e] = std::make_pair(code(p), e);
        }
        std::sort(kn.begin(), kn.end());
        std::sort(ke.begin(), ke.end());
        std::vector<int> nperm(N), eperm(E);
        h->h_ninv.resize(N); h->h_einv.resize(E);
        for (int i = 0; i < N; ++i) { nperm[i] = kn[i].second; h->h_ninv[kn[i].second] = i; }
        for (int i = 0; i < E; ++i) { eperm[i] = ke[i].second; h->h_einv[ke[i].second] = i; }
        if ((rc = upload_new(h, &h->nperm, nperm.data(), size_t(N), "nperm"))) return rc;
        if ((rc = upload_new(h, &h->ninv, h->h_ninv.data(), size_t(N), "ninv"))) return rc;
        if ((rc = upload_new(h, &h->eperm, eperm.data(), size_t(E), "eperm"))) return rc;
        if ((rc = upload_new(h, &h->einv, h->h_einv.data(), size_t(E), "einv"))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    else if (h->reorder) {
        const int *c = m->connectivity;
        std::vector<int> key(E), start(size_t(N) + 1, 0), perm(E);
        for (int e = 0; e < E; ++e) {
            int k = c[e];
            for (int j = 1; j < 4; ++j) k = std::min(k, c[size_t(j) * E + e]);
            if (k < 0 || k >= N) return fail(h, DES_E_ARG, "connectivity entry out of range");
            key[e] = k; start[size_t(k) + 1]++;
        }
        for (int n = 0; n < N; ++n) start[size_t(n) + 1] += start[n];
        h->h_einv.resize(E);
        for (int e = 0; e < E; ++e) { const int i = start[key[e]]++; perm[i] = e; h->h_einv[e] = i; }
        if ((rc = upload_new(h, &h->eperm, perm.data(), size_t(E), "eperm"))) return rc;
        if ((rc = upload_new(h, &h->einv, h->h_einv.data(), size_t(E), "einv"))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    // connectivity -> int4
    {
        if ((rc = ensure_stage(h, size_t(E) * 8 * sizeof(int)))) return rc;
        int *st = static_cast<int *>(h->stage);
        int4 *conn; if ((rc = dalloc(h, &conn, E, "conn"))) return rc;
        if ((rc = h2d(h, st, m->connectivity, size_t(E) * 4))) return rc;
        LAUNCH(KC_PACK, k_pack_conn, E, 256, E, st, h->eperm, h->ninv, conn);
        D.conn = conn;
        CK(cudaStreamSynchronize(h->stream));
    }
    // support -> packed entries, each node's list padded to a multiple of four (sentinel -1)
    if (m->support_idx[N] != 4 * E) return fail(h, DES_E_ARG, "support_idx[nnode] != 4*nelem");
    {
        std::vector<int> pidx(size_t(N) + 1);
        long long tot = 0;
        std::vector<int> hn;        // internal -> caller node (host copy of nperm)
        if (!h->h_ninv.empty()) { hn.resize(N); for (int c = 0; c < N; ++c) hn[h->h_ninv[c]] = c; }
        for (int n = 0; n < N; ++n) {
            pidx[n] = int(tot);
            const int cn = hn.empty() ? n : hn[n];
            tot += (m->support_idx[cn + 1] - m->support_idx[cn] + 3) & ~3;
            if (tot >= (1ll << 31)) return fail(h, DES_E_ARG, "padded support list overflows int32");
        }
        pidx[N] = int(tot);
        if ((rc = upload_new(h, &D.sup_idx, pidx.data(), size_t(N) + 1, "sup_idx"))) return rc;
        if ((rc = ensure_stage(h, (size_t(E) * 8 + size_t(N) + 1) * sizeof(int)))) return rc;
        int *st = static_cast<int *>(h->stage);
        int *ent; if ((rc = dalloc(h, &ent, size_t(tot), "sup_ent"))) return rc;
        CK(cudaMemsetAsync(ent, 0xFF, size_t(tot) * sizeof(int), h->stream));
        if ((rc = h2d(h, st, m->support_arr, size_t(E) * 4))) return rc;
        if ((rc = h2d(h, st + size_t(E) * 4, m->support_lidx, size_t(E) * 4))) return rc;
        if ((rc = h2d(h, st + size_t(E) * 8, m->support_idx, size_t(N) + 1))) return rc;
        LAUNCH(KC_PACK, k_pack_support, N, 256, N, st + size_t(E) * 8, D.sup_idx, st, st + size_t(E) * 4, h->einv, h->nperm, ent);
        D.sup_ent = ent;
        CK(cudaStreamSynchronize(h->stream));
    }
    if (h->h_ninv.empty()) { if ((rc = upload_new(h, &D.bcflag, m->bcflag, size_t(N), "bcflag"))) return rc; }
    else {
        std::vector<unsigned> bc(N);
        for (int c = 0; c < N; ++c) bc[h->h_ninv[c]] = m->bcflag[c];
        if ((rc = upload_new(h, &D.bcflag, bc.data(), size_t(N), "bcflag"))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    {
        int *mk; if ((rc = dalloc(h, &mk, size_t(E) * h->hp.nmat, "elemmarkers"))) return rc;
        D.elemmarkers = mk;
    }
    // node fields
    if ((rc = dalloc(h, &D.xT, N, "xT"))) return rc;
    if ((rc = dalloc(h, &D.vv, N, "vv"))) return rc;
    if ((rc = dalloc(h, &D.coord0, h->hp.has_elastic_foundation ? N : 1))) return rc;
    if ((rc = dalloc(h, &D.force, size_t(N) * 3, "force"))) return rc;
    if ((rc = dalloc(h, &D.mass, N, "mass"))) return rc;
    if (h->hp.damping_option == 4 && (rc = dalloc(h, &D.ymass, N, "ymass"))) return rc;
    if ((rc = dalloc(h, &D.tmass, N, "tmass"))) return rc;
    if ((rc = dalloc(h, &D.volume_n, N, "volume_n"))) return rc;
    if ((rc = dalloc(h, &D.ntmp, N, "ntmp"))) return rc;
    if ((rc = dalloc(h, &D.dhacc, N, "dhacc"))) return rc;
    CK(cudaMemsetAsync(D.xT, 0, sizeof(double4) * N, h->stream));
    CK(cudaMemsetAsync(D.vv, 0, sizeof(double4) * N, h->stream));
    CK(cudaMemsetAsync(D.force, 0, sizeof(double) * 3 * N, h->stream));
    CK(cudaMemsetAsync(D.tmass, 0, sizeof(double) * N, h->stream));
    CK(cudaMemsetAsync(D.dhacc, 0, sizeof(double) * N, h->stream));
    // element fields
    if ((rc = dalloc(h, &D.stress, size_t(E) * 6, "stress"))) return rc;
    if ((rc = dalloc(h, &D.strain, size_t(E) * 6, "strain"))) return rc;
    if ((rc = dalloc(h, &D.strain_rate, size_t(E) * 6, "strain_rate"))) return rc;
    if ((rc = dalloc(h, &D.plstrain, E, "plstrain"))) return rc;
    if ((rc = dalloc(h, &D.delta_plstrain, E, "delta_plstrain"))) return rc;
    if ((rc = dalloc(h, &D.volume, E, "volume"))) return rc;
    if ((rc = dalloc(h, &D.volume_old, E, "volume_old"))) return rc;
    if ((rc = dalloc(h, &D.dpressure, E, "dpressure"))) return rc;
    if ((rc = dalloc(h, &D.edvoldt, E, "edvoldt"))) return rc;
    if ((rc = dalloc(h, &D.viscosity, E, "viscosity"))) return rc;
    { double *r; if ((rc = dalloc(h, &r, E, "radiogenic"))) return rc; D.radiogenic = r;
      CK(cudaMemsetAsync(r, 0, sizeof(double) * E, h->stream)); }
    if ((rc = dalloc(h, &D.ca, size_t(E) * 8, "ca"))) return rc;
    D.ca_hi = D.ca + size_t(E) * 4;
    if ((rc = dalloc(h, &D.etmp, E, "etmp"))) return rc;
    if ((rc = dalloc(h, &D.fc, size_t(E) * DES_FC_STRIDE, "fc"))) return rc;
    if ((rc = dalloc(h, &h->sc_dev, 2, "step_scalars"))) return rc;
    if ((rc = dalloc(h, &D.status, 1, "status"))) return rc;
    CK(cudaMemsetAsync(D.status, 0, sizeof(int), h->stream));
    if ((rc = dalloc(h, &D.ddp, E, "ddp"))) return rc;
    if ((rc = dalloc(h, &D.escale, E, "escale"))) return rc;
    CK(cudaMemsetAsync(D.strain_rate, 0, sizeof(double) * 6 * E, h->stream));
    CK(cudaMemsetAsync(D.strain, 0, sizeof(double) * 6 * E, h->stream));
    CK(cudaMemsetAsync(D.stress, 0, sizeof(double) * 6 * E, h->stream));
    CK(cudaMemsetAsync(D.dpressure, 0, sizeof(double) * E, h->stream));
    CK(cudaMemsetAsync(D.delta_plstrain, 0, sizeof(double) * E, h->stream));
    CK(cudaMemsetAsync(D.ca, 0, sizeof(double) * 8 * E, h->stream));
    // never-uploaded rows must not hold NaN bit patterns (viscosity is only written by the
    // Maxwell / EVP branches; des_gpu_check_nan scans it like the reference does)
    for (double *r : {D.viscosity, D.plstrain, D.volume, D.volume_old, D.edvoldt})
        CK(cudaMemsetAsync(r, 0, sizeof(double) * E, h->stream));
    // slant-boundary tables
    for (int i = 0; i < DES_NBDRY; ++i)
        for (int d = 0; d < 3; ++d)
            h->P.bnormals[i][d] = m->bnormals ? m->bnormals[d * DES_NBDRY + i] : 0.0;
    {
        std::vector<int> slot(100, -1);
        if (m->edge_slot) slot.assign(m->edge_slot, m->edge_slot + 100);
        if ((rc = upload_new(h, &D.edge_slot, slot.data(), 100))) return rc;
        if ((rc = upload_new(h, &D.edge_vec, m->edge_vec, size_t(m->n_edge_vec)))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    // surface
    D.ntop = m->ntop; D.etop = m->etop; D.ntop_elems = m->ntop_elems;
    if (h->h_ninv.empty()) { if ((rc = upload_new(h, &D.top_nodes, m->top_nodes, size_t(m->ntop)))) return rc; }
    else {
        std::vector<int> tn(m->top_nodes, m->top_nodes + m->ntop);
        for (int &n : tn) n = h->h_ninv[n];
        if ((rc = upload_new(h, &D.top_nodes, tn.data(), tn.size()))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    {
        std::vector<int> te(m->top_elems, m->top_elems + m->ntop_elems);
        if (!h->h_einv.empty()) for (int &e : te) e = h->h_einv[e];
        if ((rc = upload_new(h, &D.top_elems, te.data(), te.size()))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    {
        unsigned char *ef; if ((rc = dalloc(h, &ef, size_t(E), "eflag"))) return rc;
        CK(cudaMemsetAsync(ef, 0, size_t(E), h->stream));
        LAUNCH(KC_PACK, k_set_eflag, m->ntop_elems, 256, m->ntop_elems, D.top_elems, ef);
        D.eflag = ef;
    }
    if (m->etop != m->nbfacets[5]) return fail(h, DES_E_ARG, "etop != nbfacets[z1]");
    {
        std::vector<int> fe(m->etop), ff(m->etop);
        for (int i = 0; i < m->etop; ++i) { fe[i] = m->bfacets[5][2*i]; ff[i] = m->bfacets[5][2*i + 1]; }
        if (!h->h_einv.empty()) for (int &e : fe) e = h->h_einv[e];
        if ((rc = upload_new(h, &D.top_felem, fe.data(), fe.size()))) return rc;
        if ((rc = upload_new(h, &D.top_ffacet, ff.data(), ff.size()))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    if (m->ntop > 0) {
        if ((rc = upload_new(h, &D.ssup_idx, m->support_surf_idx, size_t(m->ntop) + 1))) return rc;
        if ((rc = upload_new(h, &D.ssup_arr, m->support_surf_arr, size_t(m->support_surf_idx[m->ntop])))) return rc;
    }
    if ((rc = upload_new(h, &D.elem_and_nodes, m->elem_and_nodes, size_t(m->etop) * 3))) return rc;
    if (h->h_ninv.empty()) { if ((rc = upload_new(h, &D.conn_surf, m->connectivity_surface, size_t(m->etop) * 4))) return rc; }
    else {
        std::vector<int> cs(size_t(m->etop) * 4, 0);        // conn_t(etop): 3 of 4 columns are written
        for (size_t k = 0; k < size_t(m->etop) * 3; ++k) cs[k] = h->h_ninv[m->connectivity_surface[k]];
        if ((rc = upload_new(h, &D.conn_surf, cs.data(), cs.size()))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    if ((rc = dalloc(h, &D.dh, m->ntop))) return rc;
    if ((rc = dalloc(h, &D.edvacc_surf, m->etop, "edvacc_surf"))) return rc;
    if ((rc = dalloc(h, &D.s_area, m->etop))) return rc;
    if ((rc = dalloc(h, &D.s_slope, size_t(m->etop) * 3))) return rc;
    if ((rc = dalloc(h, &D.maxdh_bits, 1))) return rc;
    CK(cudaMemsetAsync(D.edvacc_surf, 0, sizeof(double) * std::max(1, m->etop), h->stream));
    CK(cudaMemsetAsync(D.dh, 0, sizeof(double) * std::max(1, m->ntop), h->stream));
    CK(cudaMemsetAsync(D.maxdh_bits, 0, sizeof(unsigned long long), h->stream));

    if ((rc = build_stress_bcs(h, m))) return rc;
    if ((rc = build_neumann_bcs(h, m))) return rc;

    h->dt_blocks = 148 * 32;     // ~40 elements per thread at 51 M tets: enough warps in flight for the gathers (592 blocks: 2.05 ms, measured)
    {
        // one partial per block of k_dt or of E-A (when compute_dt rides on it) + the final result
        const size_t cap = std::max(size_t(h->dt_blocks), size_t(DES_GRID(E, DES_EA_BLOCK)));
        if (cap > h->dt_cap) {
            if (h->dt_partials) cudaFree(h->dt_partials);
            h->dt_partials = nullptr; h->dt_cap = 0;
            CK(cudaMalloc(&h->dt_partials, sizeof(DtPartial) * (cap + 1)));
            h->dt_cap = cap;
        }
    }
    CK(cudaStreamSynchronize(h->stream));
    h->have_mesh = true;
    return DES_OK;
}

int des_gpu_upload_elemmarkers(DesGpu *h, const int *elemmarkers)
{
    if (!h || !elemmarkers) return DES_E_ARG;
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    if (h->multi) return multi_upload_elemmarkers(h, elemmarkers);
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);     // the pending mass computation must see the OLD markers
    if (rc) return rc;
    const size_t nm = size_t(h->D.nelem) * h->hp.nmat;
    if (h->eperm) {
        if ((rc = ensure_stage(h, nm * sizeof(int)))) return rc;
        int *st = static_cast<int *>(h->stage);
        if ((rc = h2d(h, st, elemmarkers, nm))) return rc;
        LAUNCH(KC_PACK, k_markers_to_internal, h->D.nelem, 256, h->D.nelem, h->hp.nmat, st, h->eperm,
               const_cast<int *>(h->D.elemmarkers));
    }
    else rc = h2d(h, const_cast<int *>(h->D.elemmarkers), elemmarkers, nm);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

int des_gpu_set_compensation_pressure(DesGpu *h, double p)
{
    if (!h) return DES_E_ARG;
    h->compensation_pressure = p;
    if (h->multi) for (MultiPart &q : h->multi->parts) q.g->compensation_pressure = p;
    return DES_OK;
}

int des_gpu_upload_state(DesGpu *h, const DesState *s)
{
    if (!h || !s) return DES_E_ARG;
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    if (h->multi) return multi_move_state(h, s, true);
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);
    if (rc) return rc;
    Dev &D = h->D;
    const size_t N = D.nnode, E = D.nelem;
    // every copy and pack kernel is queued on the stream back to back, each node field through
    // its own slice of the staging buffer; ONE synchronisation at the end (the host arrays are
    // borrowed until the call returns)
    if ((rc = ensure_stage(h, (N * 10 + (h->eperm ? std::max(E * 6, N * 3) : size_t(0)) + 1) * sizeof(double)))) return rc;
    double *stn = static_cast<double *>(h->stage);
    double *st = stn + N * 10;      // element rows (internal element order only)
    if (s->coord) {
        if ((rc = h2d(h, stn, s->coord, N * 3))) return rc;
        LAUNCH(KC_PACK, k_pack_vec3, int(N), 256, int(N), stn, h->nperm, D.xT);
    }
    if (s->vel) {
        if ((rc = h2d(h, stn + N * 3, s->vel, N * 3))) return rc;
        LAUNCH(KC_PACK, k_pack_vec3, int(N), 256, int(N), stn + N * 3, h->nperm, D.vv);
    }
    if (s->coord0 && h->hp.has_elastic_foundation) {
        if ((rc = h2d(h, stn + N * 6, s->coord0, N * 3))) return rc;
        LAUNCH(KC_PACK, k_pack_vec3, int(N), 256, int(N), stn + N * 6, h->nperm, D.coord0);
    }
    if (s->temperature) {
        if ((rc = h2d(h, stn + N * 9, s->temperature, N))) return rc;
        LAUNCH(KC_PACK, k_pack_w, int(N), 256, int(N), stn + N * 9, h->nperm, D.xT);
    }
#define UP(field, dev, n) if (s->field) { if ((rc = h2d(h, dev, (const double *)s->field, n))) return rc; }
    // element fields arrive in the caller's element order and are stored in the internal one
#define UPE(field, dev, rows) if (s->field) { \
        if (!h->eperm) { if ((rc = h2d(h, dev, (const double *)s->field, E * (rows)))) return rc; } \
        else { if ((rc = h2d(h, st, (const double *)s->field, E * (rows)))) return rc; \
               LAUNCH(KC_PACK, k_rows_to_internal<double>, int(E), 256, int(E), rows, st, h->eperm, dev); \
               CK(cudaStreamSynchronize(h->stream)); } }
    UPE(stress, D.stress, 6) UPE(strain, D.strain, 6) UPE(strain_rate, D.strain_rate, 6)
    UPE(plstrain, D.plstrain, 1) UPE(delta_plstrain, D.delta_plstrain, 1)
    UPE(volume, D.volume, 1) UPE(volume_old, D.volume_old, 1)
    UPE(viscosity, D.viscosity, 1) UPE(dpressure, D.dpressure, 1) UPE(edvoldt, D.edvoldt, 1)
    UPE(radiogenic_source, const_cast<double *>(D.radiogenic), 1)
#undef UPE
    // node rows: straight copies, or through the staging slice when the nodes are stored in another order
#define UPN(field, dev, rows) if (s->field) { \
        if (!h->nperm) { if ((rc = h2d(h, dev, (const double *)s->field, N * (rows)))) return rc; } \
        else { if ((rc = h2d(h, st, (const double *)s->field, N * (rows)))) return rc; \
               LAUNCH(KC_PACK, k_rows_to_internal<double>, int(N), 256, int(N), rows, st, h->nperm, dev); \
               CK(cudaStreamSynchronize(h->stream)); } }
    UPN(force, D.force, 3) UPN(volume_n, D.volume_n, 1) UPN(mass, D.mass, 1) UPN(tmass, D.tmass, 1)
    if (s->mass && (rc = ymass_update(h))) return rc;       // ymass is not part of DesState: rebuilt with the mass
    UPN(ntmp, D.ntmp, 1) UPN(dhacc, D.dhacc, 1) UP(edvacc_surf, D.edvacc_surf, size_t(D.etop))
#undef UPN
#undef UP
    CK(cudaStreamSynchronize(h->stream));
    h->have_state = true;
    return DES_OK;
}

int des_gpu_download_state(DesGpu *h, const DesState *s)
{
    if (!h || !s) return DES_E_ARG;
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    if (h->multi) return multi_move_state(h, s, false);
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);
    if (rc) return rc;
    Dev &D = h->D;
    const size_t N = D.nnode, E = D.nelem;
    if ((rc = ensure_stage(h, (N * 10 + (h->eperm ? std::max(E * 6, N * 3) : size_t(0)) + 1) * sizeof(double)))) return rc;
    double *stn = static_cast<double *>(h->stage);
    double *st = stn + N * 10;
#define DN(dst, src, n) CK(cudaMemcpyAsync(dst, src, (n) * sizeof(double), cudaMemcpyDeviceToHost, h->stream))
    // unpack kernels and copies back to back, one synchronisation at the end
    if (s->coord) {
        LAUNCH(KC_PACK, k_unpack_vec3, int(N), 256, int(N), D.xT, h->nperm, stn);
        DN(s->coord, stn, N * 3);
    }
    if (s->vel) {
        LAUNCH(KC_PACK, k_unpack_vec3, int(N), 256, int(N), D.vv, h->nperm, stn + N * 3);
        DN(s->vel, stn + N * 3, N * 3);
    }
    if (s->temperature) {
        LAUNCH(KC_PACK, k_unpack_w, int(N), 256, int(N), D.xT, h->nperm, stn + N * 9);
        DN(s->temperature, stn + N * 9, N);
    }
#define DOWN(field, dev, n) if (s->field) { DN(s->field, dev, n); }
    // element fields leave in the caller's element order
#define DOWNE(field, dev, rows) if (s->field) { \
        if (!h->eperm) { DN(s->field, dev, E * (rows)); } \
        else { LAUNCH(KC_PACK, k_rows_to_caller<double>, int(E), 256, int(E), rows, (const double *)dev, h->einv, st); \
               DN(s->field, st, E * (rows)); CK(cudaStreamSynchronize(h->stream)); } }
    DOWNE(stress, D.stress, 6) DOWNE(strain, D.strain, 6) DOWNE(strain_rate, D.strain_rate, 6)
    DOWNE(plstrain, D.plstrain, 1) DOWNE(delta_plstrain, D.delta_plstrain, 1)
    DOWNE(volume, D.volume, 1) DOWNE(volume_old, D.volume_old, 1)
    DOWNE(viscosity, D.viscosity, 1) DOWNE(dpressure, D.dpressure, 1) DOWNE(edvoldt, D.edvoldt, 1)
    DOWNE(radiogenic_source, D.radiogenic, 1)
#undef DOWNE
#define DOWNN(field, dev, rows) if (s->field) { \
        if (!h->nperm) { DN(s->field, dev, N * (rows)); } \
        else { LAUNCH(KC_PACK, k_rows_to_caller<double>, int(N), 256, int(N), rows, (const double *)dev, h->ninv, st); \
               DN(s->field, st, N * (rows)); CK(cudaStreamSynchronize(h->stream)); } }
    DOWNN(force, D.force, 3) DOWNN(volume_n, D.volume_n, 1) DOWNN(mass, D.mass, 1) DOWNN(tmass, D.tmass, 1)
    DOWNN(ntmp, D.ntmp, 1) DOWNN(dhacc, D.dhacc, 1) DOWN(edvacc_surf, D.edvacc_surf, size_t(D.etop))
#undef DOWNN
#undef DOWN
#undef DN
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

int des_gpu_sync(DesGpu *h)
{
    if (!h) return DES_E_ARG;
    if (h->multi) {
        for (MultiPart &p : h->multi->parts) { int rc = des_gpu_sync(p.g); if (rc) return fail(h, rc, des_gpu_last_error(p.g)); }
        return DES_OK;
    }
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    return DES_OK;
}

// -----------------------------------------------------------------------------------------
// One step is three segments separated by the two nodal barriers inside it, plus a tail;
// a multi-GPU host exchanges ghost-node values between them (des_gpu_halo_*):
//   segment 0: E-A, N-A                     -> ghosts need {temperature, dvoldt}
//   segment 1: E-B, N-B                     -> ghosts need {dp_nd}
//   segment 2: stress-bc facets, E-C, N-C     -> ghosts need {coord, vel}
//   segment 3: surface diffusion (reads the neighbours' NEW coordinates) -> ghosts need {z, dh}
//   segment 4: edvacc_surf, correct_surface_element (the host's surface-marker routines run
//              here on quality-check steps, bc.cxx:1829-1843, between segments 4 and 5)
//   segment 5: volume swap; defers compute_volume / compute_mass / rotate_stress (no kernel)
// part: -1 = the whole segment; segments 0..2 also run in three parts so that a multi-GPU host
// can overlap the ghost exchange that precedes the segment with most of its element kernel:
// 0 = element kernel on the elements that touch no ghost node (des_gpu_set_interior),
// 1 = element kernel on the others, 2 = the rest of the segment (node kernel).
static int segment_impl(DesGpu *h, int seg, int part, double dt, double time)
{
    Dev &D0 = h->D;
    const int N = D0.nnode;
    const int mm = h->hp.has_moving_mesh;
    const int hf = head_flags(h);
    const bool do_elem = part != 2, do_node = part == -1 || part == 2;
    Dev D = D0;                                 // the element kernels' view: range [e0, e1)
    if (part == 0) D.e1 = h->n_interior;
    else if (part == 1) D.e0 = h->n_interior;
    const int E = D.e1 - D.e0;                  // launch size of the element kernels
    if (seg == 0) {
    if (part <= 0) {
        const bool tail = h->tail_pending;
        h->seg_tail_mass = tail && mm;
        h->seg_rot = tail && (h->hp.rheol_type & DES_RH_ELASTIC);
    }
    // E-A / N-A (E-A already ran, with compute_dt's reductions on board, if ea_prerun)
    const bool do_ea = do_elem && !h->ea_prerun;
    if (h->seg_tail_mass) {
        if (do_ea) LAUNCH_M(KC_ELEM_A, k2_elem_a, true, E, DES_EA_BLOCK, h->P, D, 3, hf, (DtPartial *)nullptr);
        if (do_node) LAUNCH(KC_NODE_A, (k3_node_a<true>), N, 256, h->P, D0, dt, hf);
        if (do_node) { int rc = ymass_update(h); if (rc) return rc; }
    }
    else {
        if (do_ea) LAUNCH_M(KC_ELEM_A, k2_elem_a, false, E, DES_EA_BLOCK, h->P, D, 0, hf, (DtPartial *)nullptr);
        if (do_node) LAUNCH(KC_NODE_A, (k3_node_a<false>), N, 256, h->P, D0, dt, hf);
    }
    if (do_node) { h->tail_pending = false; h->ea_prerun = false; }
    }
    else if (seg == 1) {
    const bool rot = h->seg_rot;
    // E-B / N-B
    if (do_elem) {
        const double dtr = rot ? h->dt_prev : 0.0;
#define EB_CASE(S, R, A) switch (h->hp.rheol_type) { \
        case DES_RH_ELASTIC: LAUNCH(KC_ELEM_B, (k2_elem_b<S, R, DES_RH_ELASTIC, A>), E, 128, h->P, D, dt, dtr, afirst, pre); break; \
        case DES_RH_MAXWELL: LAUNCH(KC_ELEM_B, (k2_elem_b<S, R, DES_RH_MAXWELL, A>), E, 128, h->P, D, dt, dtr, afirst, pre); break; \
        case DES_RH_EP:      LAUNCH(KC_ELEM_B, (k2_elem_b<S, R, DES_RH_EP, A>), E, 128, h->P, D, dt, dtr, afirst, pre); break; \
        default:             LAUNCH(KC_ELEM_B, (k2_elem_b<S, R, DES_RH_EVP, A>), E, 128, h->P, D, dt, dtr, afirst, pre); break; }
        const bool single = h->hp.nmat == 1;
        // a deferred Output::average_fields rides on the rotation (des_gpu_average_fields)
        if (!rot && h->avg_pending) {
            // (the sums must see the stress with everything the previous step left applied)
            int rc = resolve_pending(h); if (rc) return rc;
            rc = average_now(h); if (rc) return rc;
        }
        const bool avg = rot && h->avg_pending != 0;
        const int afirst = (h->avg_pending == 2) ? 1 : 0;
        const int pre = (h->nmd_pending ? 1 : 0) | (h->scale_pending ? 2 : 0);
        if (single && avg) EB_CASE(true, true, true)
        else if (avg) EB_CASE(false, true, true)
        else if (single && rot) EB_CASE(true, true, false)
        else if (single) EB_CASE(true, false, false)
        else if (rot) EB_CASE(false, true, false)
        else EB_CASE(false, false, false)
#undef EB_CASE
        if (part == -1 || part == 1) {      // both element parts have taken it
            if (avg) h->avg_pending = 0;
            h->nmd_pending = h->scale_pending = false;
        }
    }
    if (do_node && h->hp.is_using_mixed_stress) {
        LAUNCH(KC_NODE_B, k_node_b, N, 256, D0);
    }
    }
    else if (seg == 2) {
    // E-C / N-C
    if (do_elem) {
    if (h->hp.is_using_mixed_stress) LAUNCH_M(KC_ELEM_C, k_elem_c, true, E, 256, h->P, D, h->defer_ddp);
    else                             LAUNCH_M(KC_ELEM_C, k_elem_c, false, E, 256, h->P, D, 0);
    if (h->hp.is_using_mixed_stress && h->defer_ddp && (part == -1 || part == 1)) h->nmd_pending = true;
    }
    if (do_node) {
    LAUNCH(KC_SBC, k_sbc_facets, D0.nsbc_facets, 128, h->P, D0, h->compensation_pressure);
    LAUNCH(KC_SBC, k_nbc_facets, D0.nnbc_facets, 128, h->P, D0);
    int bc_z1 = h->hp.vbc_types[5];
    if (time > h->hp.vbc_val_z1_loading_period) bc_z1 = 0;     // bc.cxx:344
    LAUNCH(KC_NODE_C, k3_node_c, N, 256, h->P, D0, dt, bc_z1, mm ? 15 : 7);
    }
    }
    else if (seg == 3) {
    if (mm && h->hp.surface_process_option == 1 && D0.ntop > 0) {
        CK(cudaMemsetAsync(D0.maxdh_bits, 0, sizeof(unsigned long long), h->stream));
        LAUNCH(KC_SURFACE, k_surf_facets, D0.etop, 128, D0);
        LAUNCH(KC_SURFACE, k_surf_nodes, D0.ntop, 128, h->P, D0, dt);
        h->surf_ran = true;
    }
    }
    else if (seg == 4) {
    if (mm) {
        if (h->hp.surface_process_option == 1 && D0.ntop > 0) LAUNCH(KC_SURFACE, k_surf_edvacc, D0.etop, 128, D0);
        if (h->defer_scale) {
            LAUNCH(KC_SURFACE, k_surf_correct_vol, D0.ntop_elems, 128, D0);
            if (D0.ntop_elems > 0) h->scale_pending = true;
        }
        else {
            // the divisions act on the stress AFTER NMD_stress's correction (reference order): a
            // diagonal correction still pending for the next E-B has to land first
            if (h->nmd_pending) { int rc = resolve_pending(h); if (rc) return rc; }
            LAUNCH(KC_SURFACE, k_surf_correct_elems, D0.ntop_elems, 128, D0);
        }
    }
    }
    else {
    if (mm) std::swap(D0.volume, D0.volume_old);      // dynearthsol.cxx:466-470
    h->tail_pending = true;
    h->dt_prev = dt;
    h->last_dt = dt;
    }
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_step_segment(DesGpu *h, int seg, double dt, double time)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (!(dt > 0)) return fail(h, DES_E_NAN, "dt <= 0");
    if (seg < 0 || seg > 5) return fail(h, DES_E_ARG, "segment must be 0..5");
    if (h->multi) return multi_step_segment(h, seg, dt, time);
    CK(cudaSetDevice(h->device));
    return segment_impl(h, seg, -1, dt, time);
}

int des_gpu_set_pt_jump(DesGpu *h, int on)
{
    if (!h) return DES_E_ARG;
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "the pseudo-transient loop runs on a single-device handle");
    h->pt_jump = on != 0;
    return DES_OK;
}

int des_gpu_set_body_force_adjustment(DesGpu *h, int on)
{
    if (!h) return DES_E_ARG;
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "the body-force adjustment runs on a single-device handle");
    h->hp.has_body_force_adjustment = on != 0;
    h->P.nbc_mask = on ? 0u : h->nbc_mask_full;
    free_graphs(h);     // P is baked into captured launches
    return DES_OK;
}

int des_gpu_set_reorder(DesGpu *h, int mode)
{
    if (!h || mode < 0 || mode > 2) return DES_E_ARG;
    h->reorder_mode = mode;     // takes effect at the next des_gpu_upload_mesh
    h->reorder_from_env = false;
    return DES_OK;
}

int des_gpu_set_interior(DesGpu *h, int n_interior)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    if (h->eperm && n_interior != h->D.nelem)
        return fail(h, DES_E_UNSUPPORTED, "interior-first element ranges need the caller's element order (reorder off)");
    if (n_interior < 0 || n_interior > h->D.nelem) return fail(h, DES_E_ARG, "n_interior out of range");
    h->n_interior = n_interior;
    return DES_OK;
}

int des_gpu_step_segment_part(DesGpu *h, int seg, int part, double dt, double time, void *cuda_stream)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (!(dt > 0)) return fail(h, DES_E_NAN, "dt <= 0");
    if (seg < 0 || seg > 2 || part < 0 || part > 2) return fail(h, DES_E_ARG, "segment 0..2, part 0..2");
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "segment parts are for hosts that run their own exchange");
    CK(cudaSetDevice(h->device));
    cudaStream_t keep = h->stream;
    if (cuda_stream) h->stream = static_cast<cudaStream_t>(cuda_stream);
    const int rc = segment_impl(h, seg, part, dt, time);
    h->stream = keep;
    return rc;
}

// des_gpu_step as one graph launch.  The six segments are captured once per combination of the
// host flags that select kernels and baked-in arguments (a handful: steady state x 2 volume
// buffers, the step after a compute_dt, the first step, averaging on/off); dt and the pending
// rotation's dt reach the kernels through Dev::sc, refreshed by an 16-byte copy when they change.
static int step_graphed(DesGpu *h, double dt, double time)
{
    CK(cudaSetDevice(h->device));
    const bool rot = h->tail_pending && (h->hp.rheol_type & DES_RH_ELASTIC);
    const double sc[2] = {dt, rot ? h->dt_prev : 0.0};
    if (!h->sc_valid || sc[0] != h->sc_host[0] || sc[1] != h->sc_host[1]) {
        h->sc_host[0] = sc[0]; h->sc_host[1] = sc[1];
        // pageable source: the runtime stages it before returning, sc_host may change afterwards
        CK(cudaMemcpyAsync(h->sc_dev, h->sc_host, sizeof(sc), cudaMemcpyHostToDevice, h->stream));
        h->sc_valid = true;
    }
    const bool z1_expired = time > h->hp.vbc_val_z1_loading_period;
    StepGraph *g = nullptr;
    for (StepGraph &c : h->graphs)
        if (c.tail_pending == h->tail_pending && c.nmd_pending == h->nmd_pending && c.scale_pending == h->scale_pending &&
            c.ea_prerun == h->ea_prerun && c.z1_expired == z1_expired && c.avg_pending == h->avg_pending &&
            c.volume == h->D.volume && c.compensation_pressure == h->compensation_pressure) { g = &c; break; }
    if (!g) {
        StepGraph n;
        n.tail_pending = h->tail_pending; n.nmd_pending = h->nmd_pending; n.scale_pending = h->scale_pending;
        n.ea_prerun = h->ea_prerun; n.z1_expired = z1_expired; n.avg_pending = h->avg_pending;
        n.volume = h->D.volume; n.compensation_pressure = h->compensation_pressure;
        const long long l0 = h->launches;
        h->D.sc = h->sc_dev;
        CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeRelaxed));
        int rc = DES_OK;
        for (int seg = 0; seg < 6 && rc == DES_OK; ++seg) rc = segment_impl(h, seg, -1, dt, time);
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
        h->D.sc = nullptr;
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) return fail(h, DES_E_CUDA, std::string("stream capture: ") + cudaGetErrorString(e));
        const cudaError_t ei = cudaGraphInstantiate(&n.exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) return fail(h, DES_E_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei));
        n.seg_tail_mass = h->seg_tail_mass; n.seg_rot = h->seg_rot;
        n.post_tail_pending = h->tail_pending; n.post_nmd = h->nmd_pending; n.post_scale = h->scale_pending;
        n.post_surf_ran = h->surf_ran; n.post_avg = h->avg_pending;
        n.launches = h->launches - l0;
        if (h->graphs.size() >= 16) {       // drop the least recently used
            size_t lru = 0;
            for (size_t i = 1; i < h->graphs.size(); ++i) if (h->graphs[i].used < h->graphs[lru].used) lru = i;
            cudaGraphExecDestroy(h->graphs[lru].exec);
            h->graphs.erase(h->graphs.begin() + lru);
        }
        h->graphs.push_back(n);
        g = &h->graphs.back();
        // the capture has already moved the host state to where the step leaves it
    }
    else {
        // what the six segments do to the host state (segment_impl), without enqueuing anything
        h->seg_tail_mass = g->seg_tail_mass; h->seg_rot = g->seg_rot;
        h->tail_pending = g->post_tail_pending; h->nmd_pending = g->post_nmd; h->scale_pending = g->post_scale;
        h->surf_ran = h->surf_ran || g->post_surf_ran; h->avg_pending = g->post_avg;
        h->ea_prerun = false;
        if (h->hp.has_moving_mesh) std::swap(h->D.volume, h->D.volume_old);
        h->dt_prev = dt; h->last_dt = dt;
        h->launches += g->launches;
    }
    g->used = ++h->graph_clock;
    CK(cudaGraphLaunch(g->exec, h->stream));
    return DES_OK;
}

int des_gpu_step(DesGpu *h, double dt, double time, int steps)
{
    (void)steps;
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (!(dt > 0)) return fail(h, DES_E_NAN, "dt <= 0");
    if (!h->multi && !h->timing && h->D.nelem <= h->graph_max_elems) return step_graphed(h, dt, time);
    for (int seg = 0; seg < 6; ++seg) {
        int rc = des_gpu_step_segment(h, seg, dt, time);
        if (rc) return rc;
    }
    return DES_OK;
}

int des_gpu_compute_dt(DesGpu *h, DesDtReductions *out)
{
    if (!h || !out) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return multi_compute_dt(h, out);
    CK(cudaSetDevice(h->device));
    int nparts;
    if (h->dt_rides) {
        // compute_dt reads the coordinates, velocities, volume and density the NEXT step's E-A
        // gathers anyway, and E-A does not depend on dt: run that launch now with the reductions
        // on board (k2_elem_a<.., DT>); segment 0 of the next step then skips its E-A.  Nothing
        // is flushed: the deferred compute_volume is part of the same launch (TAIL).
        const bool tail = h->tail_pending && h->hp.has_moving_mesh;
        const int E = h->D.nelem, hf = head_flags(h);
        nparts = DES_GRID(E, DES_EA_BLOCK);
        const bool single = h->hp.nmat == 1;
        if (tail && single)  LAUNCH(KC_ELEM_A, (k2_elem_a<true, true, true>), E, DES_EA_BLOCK, h->P, h->D, 3, hf, h->dt_partials);
        else if (tail)       LAUNCH(KC_ELEM_A, (k2_elem_a<false, true, true>), E, DES_EA_BLOCK, h->P, h->D, 3, hf, h->dt_partials);
        else if (single)     LAUNCH(KC_ELEM_A, (k2_elem_a<true, false, true>), E, DES_EA_BLOCK, h->P, h->D, 0, hf, h->dt_partials);
        else                 LAUNCH(KC_ELEM_A, (k2_elem_a<false, false, true>), E, DES_EA_BLOCK, h->P, h->D, 0, hf, h->dt_partials);
        h->ea_prerun = true;
    }
    else {
        // the deferred end-of-step work does not have to be flushed for this: the only
        // thing compute_dt reads from it is volume, which k_dt then evaluates itself
        // (tet_volume of the current coordinates == what compute_volume will store)
        const int fresh_volume = (h->tail_pending && h->hp.has_moving_mesh) ? 1 : 0;
        nparts = h->dt_blocks;
        LaunchScope ls(h, KC_DT);
        if (h->hp.nmat == 1) k_dt<true><<<h->dt_blocks, 256, 0, h->stream>>>(h->P, h->D, h->dt_partials, fresh_volume);
        else k_dt<false><<<h->dt_blocks, 256, 0, h->stream>>>(h->P, h->D, h->dt_partials, fresh_volume);
    }
    // the reduction finishes on the device: 40 + 8 bytes cross PCIe
    LAUNCH_G(KC_DT, k_dt_final, 1, 1024, 0, nparts, (const DtPartial *)h->dt_partials, h->dt_partials + h->dt_cap);
    DtPartial r;
    unsigned long long bits = 0;
    int status = 0;
    CK(cudaMemcpyAsync(&r, h->dt_partials + h->dt_cap, sizeof(DtPartial), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&bits, h->D.maxdh_bits, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&status, h->D.status, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    if (status & 1) return fail(h, DES_E_NAN, "an element volume became non-positive or NaN (compute_volume)");
    out->minl = r.minl; out->dt_maxwell = r.dt_maxwell; out->dt_diffusion = r.dt_diffusion;
    out->global_dt_min = r.global_dt_min; out->global_max_vem = r.max_vem;
    double maxdh; memcpy(&maxdh, &bits, sizeof(double));
    out->max_surf_vel = (h->surf_ran && h->last_dt > 0) ? maxdh / h->last_dt : 0.0;   // bc.cxx:1825
    return DES_OK;
}

// geometry.cxx:1597-1630 (use_global_velocity_scaling == false)
double des_dt_from_reductions(const DesParam *p, const DesDtReductions *r, double *max_global_vel_mag)
{
    if (p->fixed_dt != 0) return p->fixed_dt;
    double max_vbc_val;
    if (p->characteristic_speed == 0) {
        max_vbc_val = p->max_vbc_val;
        if (p->surface_process_option > 0)
            max_vbc_val = std::max(max_vbc_val, r->max_surf_vel * 5e-1);
    }
    else
        max_vbc_val = p->characteristic_speed;
    double gmv = std::max(r->global_max_vem, p->max_vbc_val);
    if (max_global_vel_mag) *max_global_vel_mag = gmv;
    double dt_advection = 0.5 * r->minl / max_vbc_val;
    double dt_elastic = (p->is_quasi_static) ?
        0.5 * r->minl / (max_vbc_val * p->inertial_scaling) :
        0.5 * r->minl / std::sqrt(p->bulk_modulus[p->mattype_ref] / p->rho0[p->mattype_ref]);
    double dt_hydro = std::numeric_limits<double>::max();
    double dt = std::min({dt_elastic, r->dt_maxwell, dt_advection, r->dt_diffusion, dt_hydro}) * p->dt_fraction;
    return dt;
}

// ---- bad_mesh_quality / average_fields on the device ----------

int des_gpu_mesh_quality(DesGpu *h, const DesQualityParam *qp, DesQuality *out)
{
    if (!h || !qp || !out) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return multi_mesh_quality(h, qp, out);
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);     // bad_mesh_quality reads the volume compute_volume stored
    if (rc) return rc;
    const int nb = h->dt_blocks;
    if (!h->q_partials) CK(cudaMalloc(&h->q_partials, sizeof(QPartial) * nb));
    if (!h->q_nodes) CK(cudaMalloc(&h->q_nodes, sizeof(int) * 2));
    // remeshing.cxx:2764 (sizefactor = 0.118 in 3D, mesh.cxx:63) and geometry.cxx:1893
    const double smallest_vol = qp->smallest_size * 0.118 * std::pow(qp->resolution, 3);
    const double normf = 216 * std::sqrt(3);
    const int ro = qp->remeshing_option;
    const int check_bottom = (ro == 1 || ro == 2 || ro == 11 || ro == 13);
    const int check_sides = (ro == 13);
    const double dist = qp->max_boundary_distortion * qp->resolution;
    CK(cudaMemsetAsync(h->q_nodes, 0x7f, sizeof(int) * 2, h->stream));      // 0x7f7f7f7f > any id
    LAUNCH_G(KC_DT, k_quality_elems, nb, 256, 0, h->D, h->eperm, smallest_vol, normf, h->q_partials);
    if (check_bottom || check_sides)
        LAUNCH(KC_DT, k_quality_nodes, h->D.nnode, 256, h->D, -qp->zlength, dist, qp->xlength, qp->ylength,
               check_bottom, check_sides, h->nperm, h->q_nodes);
    std::vector<QPartial> part(nb);
    int nodes[2];
    CK(cudaMemcpyAsync(part.data(), h->q_partials, sizeof(QPartial) * nb, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(nodes, h->q_nodes, sizeof(nodes), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    QPartial r = part[0];
    for (int i = 1; i < nb; ++i) {
        if (part[i].q < r.q || (part[i].q == r.q && part[i].worst < r.worst)) { r.q = part[i].q; r.worst = part[i].worst; }
        r.small = std::min(r.small, part[i].small);
    }
    out->first_small_elem = (r.small == INT_MAX) ? -1 : r.small;
    out->worst_elem = r.worst;
    out->worst_quality = r.q;
    out->first_bottom_node = (nodes[0] >= h->D.nnode) ? -1 : nodes[0];
    out->first_side_node = (nodes[1] >= h->D.nnode) ? -1 : nodes[1];
    out->bottom_value = out->side_value = 0;
    double4 x;
    auto internal = [&](int cn) { return h->h_ninv.empty() ? cn : h->h_ninv[cn]; };
    if (out->first_bottom_node >= 0) {
        CK(cudaMemcpy(&x, h->D.xT + internal(out->first_bottom_node), sizeof(x), cudaMemcpyDeviceToHost));
        out->bottom_value = x.z;
    }
    if (out->first_side_node >= 0) {
        unsigned flag;
        CK(cudaMemcpy(&x, h->D.xT + internal(out->first_side_node), sizeof(x), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(&flag, h->D.bcflag + internal(out->first_side_node), sizeof(flag), cudaMemcpyDeviceToHost));
        out->side_value = (flag & (BOUNDX0 | BOUNDX1)) ? x.x : x.y;
    }
    return DES_OK;
}

// remeshing.cxx:2766-2866: tiny element (3), bottom / side node (2), distortion (1), else 0
int des_quality_verdict(const DesQualityParam *qp, const DesQuality *q, int *index, double *min_quality)
{
    if (q->first_small_elem >= 0) { if (index) *index = q->first_small_elem; return 3; }
    if (q->first_bottom_node >= 0) { if (index) *index = q->first_bottom_node; return 2; }
    if (qp->remeshing_option == 13) {
        if (index) *index = q->first_side_node;     // the reference leaves -1 here when none is bad
        if (q->first_side_node >= 0) return 2;
    }
    const double qq = std::pow(q->worst_quality, 1.0 / 3);
    if (min_quality) *min_quality = qq;
    if (qq < qp->min_quality) { if (index) *index = q->worst_elem; return 1; }
    return 0;
}

// check_nan (utils.hpp:323-385; Output::write_exact_error): integer count, so exact
int des_gpu_check_nan(DesGpu *h, int *count)
{
    if (!h || !count) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "check_nan on a group handle: download and scan on the host");
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);
    if (rc) return rc;
    if (!h->q_nodes) CK(cudaMalloc(&h->q_nodes, sizeof(int) * 2));
    CK(cudaMemsetAsync(h->q_nodes, 0, sizeof(int), h->stream));
    LAUNCH(KC_DT, k_check_nan, std::max(h->D.nelem, h->D.nnode), 256, h->D, h->q_nodes);
    CK(cudaMemcpyAsync(count, h->q_nodes, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

int des_gpu_average_fields(DesGpu *h, int first)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) {
        for (MultiPart &p : h->multi->parts) { int rc = des_gpu_average_fields(p.g, first); if (rc) return fail(h, rc, des_gpu_last_error(p.g)); }
        h->avg_valid = true;
        return DES_OK;
    }
    CK(cudaSetDevice(h->device));
    int rc;
    Dev &D = h->D;
    const size_t N = D.nnode, E = D.nelem;
    if (h->avg_pending && (rc = flush_tail(h))) return rc;     // two calls without a step in between
    if (!h->avg_coord0) {
        if ((rc = dalloc(h, &h->avg_coord0, N * 3, "avg_coord0"))) return rc;
        if ((rc = dalloc(h, &D.avg_strain0, E * 6, "avg_strain0"))) return rc;
        if ((rc = dalloc(h, &D.avg_stress, E * 6, "avg_stress"))) return rc;
        if ((rc = dalloc(h, &D.avg_dpl, E, "avg_dpl"))) return rc;
    }
    if (!first && !h->avg_valid) return fail(h, DES_E_STATE, "average_fields: no snapshot in this interval");
    // coordinates are final for the step (nothing moves them before the next N-C)
    if (first) LAUNCH(KC_PACK, k_unpack_vec3, int(N), 256, int(N), D.xT, h->nperm, h->avg_coord0);   // caller's node order
    h->avg_valid = true;
    // average_fields must see the stress AFTER rotate_stress.  With the end-of-step work still
    // deferred, the rotation happens inside the next step's E-B, and the sums ride on it there
    // (k2_elem_b<.., AVG>): no flush, +112 B/element in that launch.  Otherwise: now.
    h->avg_pending = first ? 2 : 1;
    const bool rides = h->tail_pending && (h->hp.rheol_type & DES_RH_ELASTIC);
    if (!rides && (rc = flush_tail(h))) return rc;
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_download_averages(DesGpu *h, double *coord0, double *strain0, double *stress_avg, double *dpl_avg)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->avg_valid) return fail(h, DES_E_STATE, "no averaged fields on the device");
    if (h->multi) return multi_download_averages(h, coord0, strain0, stress_avg, dpl_avg);
    CK(cudaSetDevice(h->device));
    const size_t N = h->D.nnode, E = h->D.nelem;
    int rc;
    if (h->avg_pending && (rc = flush_tail(h))) return rc;
    if ((rc = ensure_stage(h, (h->eperm ? E * 6 : size_t(1)) * sizeof(double)))) return rc;
    double *st = static_cast<double *>(h->stage);
    if (coord0) CK(cudaMemcpyAsync(coord0, h->avg_coord0, N * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    struct { double *dst; const double *src; int rows; } f[3] = {
        {strain0, h->D.avg_strain0, 6}, {stress_avg, h->D.avg_stress, 6}, {dpl_avg, h->D.avg_dpl, 1}};
    for (auto &x : f) {
        if (!x.dst) continue;
        const double *src = x.src;
        if (h->eperm) {
            LAUNCH(KC_PACK, k_rows_to_caller<double>, int(E), 256, int(E), x.rows, x.src, h->einv, st);
            src = st;
        }
        CK(cudaMemcpyAsync(x.dst, src, E * x.rows * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        if (h->eperm) CK(cudaStreamSynchronize(h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

// ---- multi-GPU ghost exchange -------------------------------------------------------------
int des_gpu_set_stream(DesGpu *h, void *stream, int own)
{
    if (!h) return DES_E_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    drain_events(h);
    free_graphs(h);     // captured on the old stream
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    if (!own) { h->stream = static_cast<cudaStream_t>(stream); h->own_stream = false; }   // 0 = legacy default stream
    else {
        CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
    }
    return DES_OK;
}

int des_gpu_set_halo(DesGpu *h, int nsend, const int *send_nodes, int nrecv, const int *recv_nodes)
{
    if (!h || nsend < 0 || nrecv < 0 || (nsend && !send_nodes) || (nrecv && !recv_nodes)) return DES_E_ARG;
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    // the halo tables live until the next upload_mesh (pointers handed out by
    // des_gpu_halo_buffers must stay valid): one plan per uploaded mesh
    if (h->send_idx || h->recv_idx) return fail(h, DES_E_STATE, "halo already set for this mesh (upload_mesh resets it)");
    CK(cudaSetDevice(h->device));
    free_graphs(h);     // Dev changes below
    for (int i = 0; i < nsend; ++i)
        if (send_nodes[i] < 0 || send_nodes[i] >= h->D.nnode) return fail(h, DES_E_ARG, "send node out of range");
    for (int i = 0; i < nrecv; ++i)
        if (recv_nodes[i] < 0 || recv_nodes[i] >= h->D.nnode) return fail(h, DES_E_ARG, "recv node out of range");
    int rc;
    const int *p;
    std::vector<int> sn(send_nodes, send_nodes + nsend), rn(recv_nodes, recv_nodes + nrecv);
    if (!h->h_ninv.empty()) { for (int &n : sn) n = h->h_ninv[n]; for (int &n : rn) n = h->h_ninv[n]; }
    if ((rc = upload_new(h, &p, sn.data(), size_t(nsend)))) return rc;
    h->send_idx = const_cast<int *>(p);
    if ((rc = upload_new(h, &p, rn.data(), size_t(nrecv)))) return rc;
    h->recv_idx = const_cast<int *>(p);
    CK(cudaStreamSynchronize(h->stream));
    if ((rc = dalloc(h, &h->send_buf, size_t(nsend) * 7, "halo_send"))) return rc;
    if ((rc = dalloc(h, &h->recv_buf, size_t(nrecv) * 7, "halo_recv"))) return rc;
    {
        std::vector<unsigned char> gh(size_t(h->D.nnode), 0);
        for (int i = 0; i < nrecv; ++i) gh[rn[i]] = 1;
        if ((rc = upload_new(h, &h->D.ghost, gh.data(), gh.size()))) return rc;
        CK(cudaStreamSynchronize(h->stream));
    }
    if ((rc = dalloc(h, &h->top_of_node, size_t(h->D.nnode)))) return rc;
    CK(cudaMemsetAsync(h->top_of_node, 0xff, sizeof(int) * h->D.nnode, h->stream));
    LAUNCH(KC_PACK, k_top_of_node, h->D.ntop, 128, h->D.ntop, h->D.top_nodes, h->top_of_node);
    CK(cudaStreamSynchronize(h->stream));
    h->nsend = nsend; h->nrecv = nrecv;
    return DES_OK;
}

int des_gpu_halo_buffers(DesGpu *h, void **send_buf, void **recv_buf)
{
    if (!h || !send_buf || !recv_buf) return DES_E_ARG;
    *send_buf = h->send_buf; *recv_buf = h->recv_buf;
    return DES_OK;
}

int des_gpu_halo_width(int phase) { return (phase < 0 || phase > 3) ? -1 : halo_width(phase); }

int des_gpu_halo_pack(DesGpu *h, int phase)
{
    if (!h || phase < 0 || phase > 3) return DES_E_ARG;
    CK(cudaSetDevice(h->device));
    LAUNCH(KC_PACK, k_halo_pack, h->nsend, 256, phase, h->nsend, h->send_idx, h->D, h->top_of_node, h->send_buf);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_halo_unpack(DesGpu *h, int phase)
{
    if (!h || phase < 0 || phase > 3) return DES_E_ARG;
    CK(cudaSetDevice(h->device));
    LAUNCH(KC_PACK, k_halo_unpack, h->nrecv, 256, phase, h->nrecv, h->recv_idx, h->D, h->top_of_node, h->recv_buf);
    CK(cudaGetLastError());
    return DES_OK;
}

// ---- host-side topology helpers ------------------------------------------------------------
int des_host_build_support(int nnode, int nelem, const int *conn, int *idx, int *arr, int *lidx)
{
    if (nnode <= 0 || nelem <= 0 || !conn || !idx || !arr || !lidx) return DES_E_ARG;
    if (size_t(nelem) * 4 >= (size_t(1) << 31)) return DES_E_ARG;
    const size_t E = nelem;
    for (int n = 0; n <= nnode; ++n) idx[n] = 0;
    for (size_t e = 0; e < E; ++e)
        for (int i = 0; i < 4; ++i) {
            const int n = conn[i * E + e];
            if (n < 0 || n >= nnode) return DES_E_ARG;
            idx[n + 1]++;
        }
    for (int n = 1; n <= nnode; ++n) idx[n] += idx[n - 1];
    std::vector<int> cursor(idx, idx + nnode);
    for (size_t e = 0; e < E; ++e)
        for (int i = 0; i < 4; ++i) {
            const int slot = cursor[conn[i * E + e]]++;
            arr[slot] = int(e);
            lidx[slot] = i;
        }
    return DES_OK;
}

int des_abi_sizeof(int which)
{
    switch (which) {
    case 0: return int(sizeof(DesParam));
    case 1: return int(sizeof(DesMesh));
    case 2: return int(sizeof(DesState));
    case 3: return int(sizeof(DesDtReductions));
    case 4: return int(sizeof(DesQualityParam));
    case 5: return int(sizeof(DesQuality));
    case 6: return int(sizeof(DesFrame));
    case 7: return int(sizeof(DesTransferMeshes));
    case 8: return int(sizeof(DesTransferField));
    case 9: return int(sizeof(DesTransferStats));
    default: return -1;
    }
}

// ---- instrumentation --------------------------------------------------------------------
long long des_gpu_launch_count(const DesGpu *h)
{
    if (!h) return 0;
    long long n = h->launches;
    if (h->multi) for (const MultiPart &p : h->multi->parts) n += p.g->launches;
    return n;
}

int des_gpu_enable_kernel_timing(DesGpu *h, int on)
{
    if (!h) return DES_E_ARG;
    drain_events(h);
    h->timing = on != 0;
    memset(h->kc_ms, 0, sizeof(h->kc_ms));
    memset(h->kc_n, 0, sizeof(h->kc_n));
    return DES_OK;
}

int des_gpu_kernel_time(DesGpu *h, int which, double *ms, long long *launches, const char **name)
{
    if (!h || which < 0 || which >= DES_NKCLASS) return DES_E_ARG;
    drain_events(h);
    if (ms) *ms = h->kc_ms[which];
    if (launches) *launches = h->kc_n[which];
    if (name) *name = kclass_name[which];
    return DES_OK;
}

int des_gpu_event_record(DesGpu *h, int slot)
{
    if (!h || slot < 0 || slot >= 8) return DES_E_ARG;
    CK(cudaSetDevice(h->device));
    if (!h->user_ev[slot]) CK(cudaEventCreate(&h->user_ev[slot]));
    CK(cudaEventRecord(h->user_ev[slot], h->stream));
    return DES_OK;
}

int des_gpu_event_elapsed_ms(DesGpu *h, int a, int b, double *ms)
{
    if (!h || !ms || a < 0 || a >= 8 || b < 0 || b >= 8 || !h->user_ev[a] || !h->user_ev[b]) return DES_E_ARG;
    CK(cudaEventSynchronize(h->user_ev[b]));
    float f = 0;
    CK(cudaEventElapsedTime(&f, h->user_ev[a], h->user_ev[b]));
    *ms = f;
    return DES_OK;
}

void *des_gpu_device_ptr(DesGpu *h, const char *name)
{
    if (!h || !name) return nullptr;
    std::string n(name);
    // arrays whose pointer changes (volume swap) are resolved live
    if (n == "volume") return h->D.volume;
    if (n == "volume_old") return h->D.volume_old;
    auto it = h->named.find(n);
    return it == h->named.end() ? nullptr : it->second;
}

} // extern "C"

// ---- des_gpu_compute_dt_static: compute_dt's reductions without any allocation --------------
// No counterpart in the reference; added for callers that
// may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// k_dt's per-block partial results and k_dt_final's combined result live in the file-scope
// device array below instead of the handle's cudaMalloc'ed dt_partials.  The array exists once
// per GPU that loads this library and is held until the process exits:
// (DES_STATIC_DT_BLOCKS + 1) * 40 bytes = 655 400 bytes.  One block reduces
// DES_STATIC_DT_ELEMS_PER_BLOCK elements, so the entry serves meshes of up to
// DES_STATIC_DT_BLOCKS * DES_STATIC_DT_ELEMS_PER_BLOCK = 16 777 216 elements and refuses
// larger ones.
#define DES_STATIC_DT_BLOCKS 16384
#define DES_STATIC_DT_ELEMS_PER_BLOCK 1024
#define DES_STATIC_MAX_DEVICES 64

static_assert((long long)DES_STATIC_DT_BLOCKS * DES_STATIC_DT_ELEMS_PER_BLOCK == DES_STATIC_DT_MAX_ELEMS,
              "include/des_gpu.h states the limit");

static __device__ __align__(256) DtPartial des_static_dt_scratch[DES_STATIC_DT_BLOCKS + 1];

namespace {

// In host code `des_static_dt_scratch` is only a stand-in for the symbol, not a device address:
// the address the kernels get is looked up once per device, with that device current.
DtPartial *g_static_dt_addr[DES_STATIC_MAX_DEVICES];
std::mutex g_static_dt_mutex;

int static_dt_scratch(DesGpu *h, DtPartial **out)
{
    if (h->device < 0 || h->device >= DES_STATIC_MAX_DEVICES)
        return fail(h, DES_E_UNSUPPORTED, "des_gpu_compute_dt_static: device index beyond " +
                    std::to_string(DES_STATIC_MAX_DEVICES - 1));
    std::lock_guard<std::mutex> lock(g_static_dt_mutex);
    if (!g_static_dt_addr[h->device]) {
        CK(cudaSetDevice(h->device));
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, des_static_dt_scratch);
        if (e != cudaSuccess || !p)
            return fail(h, DES_E_CUDA, std::string("cudaGetSymbolAddress(des_static_dt_scratch): ") +
                        cudaGetErrorString(e));
        g_static_dt_addr[h->device] = static_cast<DtPartial *>(p);
    }
    *out = g_static_dt_addr[h->device];
    return DES_OK;
}

} // namespace

extern "C" int des_gpu_compute_dt_static(DesGpu *h, DesDtReductions *out)
{
    if (!h || !out) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "des_gpu_compute_dt_static runs on a single-device handle");
    const long long E = h->D.nelem;
    const long long max_elems = (long long)DES_STATIC_DT_BLOCKS * DES_STATIC_DT_ELEMS_PER_BLOCK;
    if (E > max_elems)
        return fail(h, DES_E_ARG, "des_gpu_compute_dt_static: the mesh has " + std::to_string(E) +
                    " elements, the static scratch serves at most " + std::to_string(max_elems) +
                    " (" + std::to_string(DES_STATIC_DT_BLOCKS) + " block partials); use des_gpu_compute_dt");
    CK(cudaSetDevice(h->device));
    DtPartial *scratch = nullptr;
    int rc = static_dt_scratch(h, &scratch);
    if (rc) return rc;
    DtPartial *result = scratch + DES_STATIC_DT_BLOCKS;
    const int nparts = std::max(1, int((E + DES_STATIC_DT_ELEMS_PER_BLOCK - 1) / DES_STATIC_DT_ELEMS_PER_BLOCK));
    // The array keeps the previous call's contents.  Every partial that k_dt_final reads is
    // written by this call's k_dt first; the result slot is cleared so that a launch that
    // failed cannot hand back the previous call's numbers.
    CK(cudaMemsetAsync(result, 0, sizeof(DtPartial), h->stream));
    // nothing of the handle's deferred end-of-step work is touched: like the non-riding path of
    // des_gpu_compute_dt, k_dt evaluates the volume itself while compute_volume is pending
    const int fresh_volume = (h->tail_pending && h->hp.has_moving_mesh) ? 1 : 0;
    if (h->hp.nmat == 1) k_dt<true><<<nparts, 256, 0, h->stream>>>(h->P, h->D, scratch, fresh_volume);
    else k_dt<false><<<nparts, 256, 0, h->stream>>>(h->P, h->D, scratch, fresh_volume);
    k_dt_final<<<1, 1024, 0, h->stream>>>(nparts, (const DtPartial *)scratch, result);
    h->launches += 2;
    DtPartial r;
    unsigned long long bits = 0;
    int status = 0;
    CK(cudaMemcpyAsync(&r, result, sizeof(DtPartial), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&bits, h->D.maxdh_bits, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&status, h->D.status, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    if (status & 1) return fail(h, DES_E_NAN, "an element volume became non-positive or NaN (compute_volume)");
    out->minl = r.minl; out->dt_maxwell = r.dt_maxwell; out->dt_diffusion = r.dt_diffusion;
    out->global_dt_min = r.global_dt_min; out->global_max_vem = r.max_vem;
    double maxdh; memcpy(&maxdh, &bits, sizeof(double));
    out->max_surf_vel = (h->surf_ran && h->last_dt > 0) ? maxdh / h->last_dt : 0.0;   // bc.cxx:1825
    return DES_OK;
}

#include "des_twins.cuh"
#include "des_topology_api.cuh"
#include "des_frame.cuh"
#include "des_transfer_api.cuh"
#include "des_markers.cuh"
