// See des_bridge.hpp.  Compiled with the reference's flags (-DTHREED -DSOA) against its headers.
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <iostream>
#include <string>
#include <vector>

#include "constants.hpp"
#include "parameters.hpp"
#include "matprops.hpp"
#include "markerset.hpp"
#include "phasechanges.hpp"
#include "output.hpp"
#include "remeshing.hpp"
#include "utils.hpp"
#include "des_bridge.hpp"

#ifndef THREED
#error "the B200 explicit core is 3D only (SURVEY.md section 2: 2D is out of scope)"
#endif

struct DesBridge {
    DesGpu *h = nullptr;
    DesParam p;
    std::string err;
    std::vector<int> flat_markers;
    std::vector<std::vector<int> > flat_facets;
    double avg_time0 = 0;       // Output::time0 of the running averaging interval
    // device-resident markers (des_gpu_upload_markers): stale after every host change to the
    // material marker set -- init / remesh (des_bridge_upload), the surface-marker routines and the
    // host phase_changes (des_bridge_markers_changed)
    bool markers_stale = true;
    int ndev = 1;
    std::vector<int> mk_elem, mk_mattype, mk_changed, mk_new, mk_hyd, hyd_count;
    std::vector<double> mk_eta;
};

namespace {

template <class V> void copy_mat(double *dst, const V &src, int nmat)
{
    for (int m = 0; m < nmat && m < DES_MAX_MAT; ++m)
        dst[m] = src[src.size() == 1 ? 0 : m];
}

void fill_param(DesParam &q, const Param &param, const Variables &var)
{
    std::memset(&q, 0, sizeof(q));
    const Control &c = param.control;
    q.gravity = c.gravity;
    q.inertial_scaling = c.inertial_scaling;
    q.damping_factor = c.damping_factor;
    q.surface_diffusivity = var.surfinfo.surf_diff;
    q.surf_base_level = c.surf_base_level;
    q.dt_fraction = c.dt_fraction;
    q.fixed_dt = c.fixed_dt;
    q.characteristic_speed = c.characteristic_speed;
    q.damping_option = c.damping_option;
    q.is_quasi_static = c.is_quasi_static;
    q.has_thermal_diffusion = c.has_thermal_diffusion;
    q.is_using_mixed_stress = c.is_using_mixed_stress;
    q.surface_process_option = c.surface_process_option;
    q.ref_pressure_option = c.ref_pressure_option;
    q.has_moving_mesh = c.has_moving_mesh;
    q.has_body_force_adjustment = param.ic.has_body_force_adjustment;
    const BC &bc = param.bc;
    q.has_winkler_foundation = bc.has_winkler_foundation;
    q.has_water_loading = bc.has_water_loading;
    q.has_elastic_foundation = bc.has_elastic_foundation;
    q.winkler_delta_rho = bc.winkler_delta_rho;
    q.elastic_foundation_constant = bc.elastic_foundation_constant;
    q.sea_water_density = bc.sea_water_density;
    q.surface_temperature = bc.surface_temperature;
    for (int i = 0; i < DES_NBDRY; ++i) { q.vbc_types[i] = var.vbc_types[i]; q.vbc_values[i] = var.vbc_values[i]; }
    q.vbc_val_lateral[0] = bc.vbc_val_x0_l; q.vbc_val_lateral[1] = bc.vbc_val_x1_l;
    q.vbc_val_lateral[2] = bc.vbc_val_y0_l; q.vbc_val_lateral[3] = bc.vbc_val_y1_l;
    q.vbc_val_z1_loading_period = var.vbc_val_z1_loading_period;
    for (int i = 0; i < 6; ++i) { q.stress_bc_types[i] = var.stress_bc_types[i]; q.stress_bc_values[i] = var.stress_bc_values[i]; }
    q.max_vbc_val = var.max_vbc_val;
    q.zlength = param.mesh.zlength;
    const Mat &m = param.mat;
    q.rheol_type = m.rheol_type;
    q.nmat = m.nmat;
    q.mattype_ref = m.mattype_ref;
    q.visc_min = m.visc_min; q.visc_max = m.visc_max;
    q.tension_max = m.tension_max; q.therm_diff_max = m.therm_diff_max;
    copy_mat(q.rho0, m.rho0, m.nmat); copy_mat(q.alpha, m.alpha, m.nmat);
    copy_mat(q.bulk_modulus, m.bulk_modulus, m.nmat); copy_mat(q.shear_modulus, m.shear_modulus, m.nmat);
    copy_mat(q.visc_exponent, m.visc_exponent, m.nmat); copy_mat(q.visc_coefficient, m.visc_coefficient, m.nmat);
    copy_mat(q.visc_activation_energy, m.visc_activation_energy, m.nmat);
    copy_mat(q.visc_activation_volume, m.visc_activation_volume, m.nmat);
    copy_mat(q.heat_capacity, m.heat_capacity, m.nmat); copy_mat(q.therm_cond, m.therm_cond, m.nmat);
    copy_mat(q.pls0, m.pls0, m.nmat); copy_mat(q.pls1, m.pls1, m.nmat);
    copy_mat(q.cohesion0, m.cohesion0, m.nmat); copy_mat(q.cohesion1, m.cohesion1, m.nmat);
    copy_mat(q.friction_angle0, m.friction_angle0, m.nmat); copy_mat(q.friction_angle1, m.friction_angle1, m.nmat);
    copy_mat(q.dilation_angle0, m.dilation_angle0, m.nmat); copy_mat(q.dilation_angle1, m.dilation_angle1, m.nmat);
    copy_mat(q.porosity, m.porosity, m.nmat);
}

int fail(DesBridge *b, int rc, const char *what)
{
    b->err = std::string(what) + ": " + (b->h ? des_gpu_last_error(b->h) : "no device model");
    std::fprintf(stderr, "des_bridge: %s (code %d)\n", b->err.c_str(), rc);
    return rc;
}

void flatten_markers(DesBridge *b, const Param &param, const Variables &var)
{
    const int nmat = param.mat.nmat;
    b->flat_markers.resize(size_t(var.nelem) * nmat);
    for (int e = 0; e < var.nelem; ++e)
        for (int m = 0; m < nmat; ++m)
            b->flat_markers[size_t(e) * nmat + m] = (*var.elemmarkers)[e][m];
}

// every field array of Variables the device owns a copy of (SURVEY.md section 8b classes)
void state_pointers(DesState &s, const Variables &var, bool all)
{
    std::memset(&s, 0, sizeof(s));
    s.coord = var.coord->data(); s.vel = var.vel->data();
    s.temperature = var.temperature->data();
    s.stress = var.stress->data(); s.strain = var.strain->data(); s.strain_rate = var.strain_rate->data();
    s.plstrain = var.plstrain->data(); s.delta_plstrain = var.delta_plstrain->data();
    s.volume = var.volume->data();
    s.viscosity = var.viscosity->data();
    s.force = var.force->data();
    if (all) {
        s.coord0 = var.coord0->data();
        s.volume_old = var.volume_old->data();
        s.dpressure = var.dpressure->data(); s.edvoldt = var.edvoldt->data();
        s.radiogenic_source = var.radiogenic_source->data();
        s.volume_n = var.volume_n->data(); s.mass = var.mass->data(); s.tmass = var.tmass->data();
        s.dhacc = var.surfinfo.dhacc->data();
        s.edvacc_surf = var.surfinfo.edvacc_surf->data();
    }
}

} // namespace

int des_bridge_upload(DesBridge *b, const Param &param, const Variables &var)
{
    b->markers_stale = true;
    DesMesh m;
    std::memset(&m, 0, sizeof(m));
    m.nnode = var.nnode; m.nelem = var.nelem;
    m.connectivity = var.connectivity->data();
    m.bcflag = var.bcflag->data();
    m.support_idx = var.support.idx_data.data();
    m.support_arr = var.support.arr_data.data();
    m.support_lidx = var.support.lidx_data.data();
    b->flat_facets.assign(DES_NBDRY, std::vector<int>());
    for (int i = 0; i < DES_NBDRY; ++i) {
        const auto &bf = *var.bfacets[i];
        std::vector<int> &f = b->flat_facets[i];
        f.resize(bf.size() * 2);
        for (size_t k = 0; k < bf.size(); ++k) { f[2*k] = bf[k].first; f[2*k + 1] = bf[k].second; }
        m.nbfacets[i] = int(bf.size());
        m.bfacets[i] = f.data();
        m.nbnodes[i] = int(var.bnodes[i]->size());
        m.bnodes[i] = var.bnodes[i]->data();
    }
    m.bnormals = var.bnormals->data();
    m.n_edge_vec = int(var.edge_vec.size());
    m.edge_vec = var.edge_vec.data();
    m.edge_slot = var.edge_slot;
    m.ntop_elems = var.ntop_elems;
    m.top_elems = var.top_elems->data();
    m.ntop = var.surfinfo.ntop;
    m.top_nodes = var.surfinfo.top_nodes->data();
    m.etop = var.surfinfo.etop;
    m.support_surf_idx = var.surfinfo.support_surf.idx_data.data();
    m.support_surf_arr = var.surfinfo.support_surf.arr_data.data();
    m.elem_and_nodes = var.surfinfo.elem_and_nodes->data();
    m.connectivity_surface = var.connectivity_surface->data();
    m.coord = var.coord->data();        // a handle over several devices bisects the element centroids
    const int64_t t0 = get_nanoseconds();
    int rc = des_gpu_upload_mesh(b->h, &m);
    if (rc) return fail(b, rc, "upload_mesh");
    const int64_t t1 = get_nanoseconds();
    {
        int nparts = 1;
        double ps = 0, us = 0;
        std::vector<int> le(64), ln(64), gh(64);
        des_gpu_partition_info(b->h, &nparts, &ps, &us, le.data(), ln.data(), gh.data());
        if (nparts > 1) {
            std::printf("  [B200] mesh partitioned over %d devices: partition %.3f s, upload %.3f s; elements per part",
                        nparts, ps, us);
            for (int r = 0; r < nparts; ++r) std::printf(" %d", le[r]);
            std::printf(" (%d in the mesh), ghost nodes", var.nelem);
            for (int r = 0; r < nparts; ++r) std::printf(" %d", gh[r]);
            std::printf("\n");
        }
    }
    flatten_markers(b, param, var);
    if ((rc = des_gpu_upload_elemmarkers(b->h, b->flat_markers.data()))) return fail(b, rc, "upload_elemmarkers");
    des_gpu_set_compensation_pressure(b->h, var.compensation_pressure);
    const int64_t t2 = get_nanoseconds();
    DesState s;
    state_pointers(s, var, true);
    if ((rc = des_gpu_upload_state(b->h, &s))) return fail(b, rc, "upload_state");
    std::printf("  [B200] upload of %d elements: mesh %.4f s, markers %.4f s, state %.4f s\n", var.nelem,
                (t1 - t0) * 1e-9, (t2 - t1) * 1e-9, (get_nanoseconds() - t2) * 1e-9);
    return 0;
}

DesBridge *des_bridge_create(const Param &param, const Variables &var, int device)
{
    DesBridge *b = new DesBridge;
    fill_param(b->p, param, var);
    // DES_BRIDGE_NDEV=n: one handle over n devices of the box (DES_BRIDGE_DEVICES=0,0,1: which ones;
    // an index may repeat) -- partition, ghost exchange and gather/scatter happen inside the library
    int ndev = 1;
    std::vector<int> devs;
    if (const char *s = std::getenv("DES_BRIDGE_NDEV")) ndev = std::max(1, std::atoi(s));
    if (const char *s = std::getenv("DES_BRIDGE_DEVICES")) {
        for (const char *q = s; *q; ) { devs.push_back(std::atoi(q)); while (*q && *q != ',') ++q; if (*q) ++q; }
        ndev = int(devs.size());
    }
    b->ndev = (ndev > 1 || !devs.empty()) ? std::max(ndev, 2) : 1;   // > 1: a group handle (even over one device)
    int rc = (ndev > 1 || !devs.empty())
        ? des_gpu_create_multi(&b->p, ndev, devs.empty() ? nullptr : devs.data(), &b->h)
        : des_gpu_create(&b->p, device, &b->h);
    // storage order inside the library: unstructured (tetgen) meshes step ~9 % faster with nodes
    // and elements in Morton order (profiles/r02_order_ab_tetgen_1M.json); the regular box
    // generator's own order is better left alone (measured 3 % slower in Morton order at 51M tets).
    // DES_REORDER in the environment overrides.
    if (!rc && !std::getenv("DES_REORDER"))
        des_gpu_set_reorder(b->h, param.mesh.meshing_option == 1 ? 0 : 2);
    if (rc) {
        std::fprintf(stderr, "des_bridge: des_gpu_create failed (code %d): %s\n", rc,
                     b->h ? des_gpu_last_error(b->h) : "no CUDA device");
        if (b->h) des_gpu_destroy(b->h);
        delete b;
        return nullptr;     // there is no CPU fallback: the caller must stop
    }
    if (des_bridge_upload(b, param, var)) { des_bridge_destroy(b); return nullptr; }
    return b;
}

int des_bridge_step(DesBridge *b, const Param &param, Variables &var)
{
    // On quality-check steps surface_processes() also moves markers on the host, in the middle
    // of update_mesh (bc.cxx:1829-1843): break the step there.
    const bool marker_step = param.control.has_moving_mesh && var.steps != 0 &&
        var.steps % param.mesh.quality_check_step_interval == 0;
    if (!marker_step) {
        int rc = des_gpu_step(b->h, var.dt, var.time, var.steps);
        if (rc) return fail(b, rc, "step");
        return 0;
    }
    int rc;
    for (int seg = 0; seg <= 4; ++seg)
        if ((rc = des_gpu_step_segment(b->h, seg, var.dt, var.time))) return fail(b, rc, "step_segment");
    DesState s;
    std::memset(&s, 0, sizeof(s));
    s.coord = var.coord->data();
    s.volume = var.volume->data();          // pre-swap: top elements already hold the new volume
    s.dhacc = var.surfinfo.dhacc->data();
    s.edvacc_surf = var.surfinfo.edvacc_surf->data();
    if ((rc = des_gpu_download_state(b->h, &s))) return fail(b, rc, "download_state");
    var.markersets[0]->correct_surface_marker(param, var, *var.surfinfo.dhacc, *var.elemmarkers,
                                              *var.markers_in_elem);
    for (int i = 0; i < var.surfinfo.ntop; i++)
        (*var.surfinfo.dhacc)[(*var.surfinfo.top_nodes)[i]] = 0.;
    var.markersets[0]->set_surface_marker(param, var, param.mesh.smallest_size, param.mat.mattype_sed,
                                          *var.surfinfo.edvacc_surf, *var.elemmarkers, *var.markers_in_elem);
    std::memset(&s, 0, sizeof(s));
    s.dhacc = var.surfinfo.dhacc->data();
    s.edvacc_surf = var.surfinfo.edvacc_surf->data();
    if ((rc = des_gpu_upload_state(b->h, &s))) return fail(b, rc, "upload_state");
    if ((rc = des_bridge_markers_changed(b, param, var))) return rc;
    if ((rc = des_gpu_step_segment(b->h, 5, var.dt, var.time))) return fail(b, rc, "step_segment");
    return 0;
}

// ---- control.has_PT: the step as the reference's un-fused call sequence (the twins) -----------
namespace {

#define TW(call) do { int rc_ = (call); if (rc_) return fail(b, rc_, #call); } while (0)

// the surface-marker work surface_processes() does on the host on quality-check steps
// (bc.cxx:1829-1843), between the two halves of update_mesh
int marker_interlude(DesBridge *b, const Param &param, Variables &var)
{
    DesState s;
    std::memset(&s, 0, sizeof(s));
    s.coord = var.coord->data();
    s.volume = var.volume->data();
    s.dhacc = var.surfinfo.dhacc->data();
    s.edvacc_surf = var.surfinfo.edvacc_surf->data();
    TW(des_gpu_download_state(b->h, &s));
    var.markersets[0]->correct_surface_marker(param, var, *var.surfinfo.dhacc, *var.elemmarkers,
                                              *var.markers_in_elem);
    for (int i = 0; i < var.surfinfo.ntop; i++)
        (*var.surfinfo.dhacc)[(*var.surfinfo.top_nodes)[i]] = 0.;
    var.markersets[0]->set_surface_marker(param, var, param.mesh.smallest_size, param.mat.mattype_sed,
                                          *var.surfinfo.edvacc_surf, *var.elemmarkers, *var.markers_in_elem);
    std::memset(&s, 0, sizeof(s));
    s.dhacc = var.surfinfo.dhacc->data();
    s.edvacc_surf = var.surfinfo.edvacc_surf->data();
    TW(des_gpu_upload_state(b->h, &s));
    return des_bridge_markers_changed(b, param, var);
}

// the body both pseudo-transient loops share (dynearthsol.cxx:562-584 and 812-860)
int pt_iterations(DesBridge *b, Param &param, Variables &var, double &residual_old, bool with_remesh)
{
    TW(des_gpu_set_pt_jump(b->h, 1));
    param.control.PT_jump = true;
    for (int pt_step = 0; pt_step < param.control.PT_max_iter; ++pt_step) {
        TW(des_gpu_apply_vbcs(b->h, var.time));
        if (param.control.has_moving_mesh) TW(des_gpu_update_mesh(b->h, var.dt));
        TW(des_gpu_update_strain_rate(b->h));
        TW(des_gpu_compute_dvoldt(b->h));
        TW(des_gpu_compute_edvoldt(b->h));
        TW(des_gpu_update_stress(b->h, var.dt));
        TW(des_gpu_update_force(b->h));
        TW(des_gpu_update_velocity(b->h, var.dt));
        TW(des_gpu_residual_force(b->h, &var.l2_residual));
        const double relative_change = std::fabs((var.l2_residual - residual_old) / residual_old);
        if (relative_change < param.control.PT_relative_tolerance) break;
        residual_old = var.l2_residual;
        if (with_remesh && pt_step % param.mesh.quality_check_step_interval == 0 && param.control.has_moving_mesh) {
            int index; double min_quality;
            const int bad = des_bridge_bad_mesh_quality(b, param, var, index, min_quality);
            if (bad < 0) return 1;
            if (bad) {
                des_bridge_output_barrier();
                if (des_bridge_download(b, var, DES_DL_ALL)) return 1;
                if (param.sim.has_output_during_remeshing) var.output->write_exact(var);
                remesh(param, var, bad);
                if (des_bridge_upload(b, param, var)) return 1;
                TW(des_gpu_set_pt_jump(b->h, 1));
                if (param.sim.has_output_during_remeshing) var.output->write_exact(var);
            }
        }
    }
    param.control.PT_jump = false;
    TW(des_gpu_set_pt_jump(b->h, 0));
    return 0;
}

} // namespace

int des_bridge_step_pt(DesBridge *b, Param &param, Variables &var)
{
    // dynearthsol.cxx:776-876 (no hydraulics): every call goes to its fine-grained twin
    if (param.control.has_thermal_diffusion) TW(des_gpu_update_temperature(b->h, var.dt));
    TW(des_gpu_update_strain_rate(b->h));
    TW(des_gpu_compute_dvoldt(b->h));
    TW(des_gpu_compute_edvoldt(b->h));
    TW(des_gpu_update_stress(b->h, var.dt));
    if (param.control.is_using_mixed_stress) TW(des_gpu_nmd_stress(b->h));
    TW(des_gpu_update_force(b->h));
    TW(des_gpu_update_velocity(b->h, var.dt));
    TW(des_gpu_residual_force(b->h, &var.l2_residual));
    double residual_old = var.l2_residual;
    if (param.control.has_PT) { int rc = pt_iterations(b, param, var, residual_old, true); if (rc) return rc; }
    TW(des_gpu_apply_vbcs(b->h, var.time));
    if (param.control.has_moving_mesh) {
        const bool marker_step = var.steps != 0 && var.steps % param.mesh.quality_check_step_interval == 0;
        if (!marker_step) TW(des_gpu_update_mesh(b->h, var.dt));
        else {
            TW(des_gpu_update_mesh_part(b->h, var.dt, 0));
            int rc = marker_interlude(b, param, var); if (rc) return rc;
            TW(des_gpu_update_mesh_part(b->h, var.dt, 1));
        }
    }
    if (param.mat.rheol_type & MatProps::rh_elastic) TW(des_gpu_rotate_stress(b->h, var.dt));
    return 0;
}

int des_bridge_initial_body_force_adjustment(DesBridge *b, Param &param, Variables &var)
{
    // dynearthsol.cxx:546-591, 753-761: "only works with the PT loop"; the prescribed tractions
    // stay off until it is done (fields.cxx:690)
    TW(des_gpu_residual_force(b->h, &var.l2_residual));
    double residual_old = var.l2_residual;
    if (param.control.has_PT) { int rc = pt_iterations(b, param, var, residual_old, false); if (rc) return rc; }
    param.ic.has_body_force_adjustment = false;
    TW(des_gpu_set_body_force_adjustment(b->h, 0));
    return 0;
}

double des_bridge_compute_dt(DesBridge *b, const Param &param, Variables &var)
{
    (void)param;
    DesDtReductions r;
    int rc = des_gpu_compute_dt(b->h, &r);
    if (rc) { fail(b, rc, "compute_dt"); return -1; }
    var.surfinfo.max_surf_vel = r.max_surf_vel;
    var.global_dt_min = r.global_dt_min;
    return des_dt_from_reductions(&b->p, &r, &var.max_global_vel_mag);
}

int des_bridge_download(DesBridge *b, Variables &var, int what)
{
    DesState s;
    if (what & DES_DL_ALL) state_pointers(s, var, true);
    else if (what & DES_DL_OUTPUT) state_pointers(s, var, false);
    else {
        std::memset(&s, 0, sizeof(s));
        s.coord = var.coord->data(); s.volume = var.volume->data();
        s.temperature = var.temperature->data();
    }
    int rc = des_gpu_download_state(b->h, &s);
    if (rc) return fail(b, rc, "download_state");
    return 0;
}

// phase_changes (dynearthsol.cxx:885-887) with the marker lookups on the device: the markers'
// element / shape-function weights / rock type stay resident between host changes
// (des_gpu_upload_markers), des_gpu_phase_changes reads the resident coordinates and temperature,
// keeps the device's rock types and elemmarkers current and returns the markers that changed; here
// they are written into the reference's MarkerSet and elemmarkers exactly as phasechanges.cxx:143-147
// does, and the hydration block (:160-176, host marker bookkeeping) follows.  No geometry
// download, no elemmarkers upload.  A handle over several devices, or DES_BRIDGE_PHASE=ref, takes
// the reference's own phase_changes on downloaded fields.
int des_bridge_phase_changes(DesBridge *b, const Param &param, Variables &var)
{
    if (param.mat.nmat == 1 || param.mat.phase_change_option == 0) return 0;      // phasechanges.cxx:111
    static const bool want_ref = std::getenv("DES_BRIDGE_PHASE") && std::strcmp(std::getenv("DES_BRIDGE_PHASE"), "ref") == 0;
    const int opt = param.mat.phase_change_option;
    if (want_ref || b->ndev > 1 || (opt != 1 && opt != 101)) {
        int rc = des_bridge_download(b, var, DES_DL_GEOMETRY);
        if (rc) return rc;
        phase_changes(param, var);
        return des_bridge_markers_changed(b, param, var);
    }
    MarkerSet &ms = *var.markersets[0];
    const int n = ms.get_nmarkers();
    int rc;
    if (b->markers_stale) {
        b->mk_elem.resize(n); b->mk_mattype.resize(n); b->mk_eta.resize(size_t(4) * std::max(n, 1));
        for (int m = 0; m < n; ++m) {
            b->mk_elem[m] = ms.get_elem(m);
            b->mk_mattype[m] = ms.get_mattype(m);
            for (int i = 0; i < NODES_PER_ELEM; ++i) b->mk_eta[size_t(i) * n + m] = ms.get_eta(m)[i];
        }
        if ((rc = des_gpu_upload_markers(b->h, n, n, b->mk_elem.data(), b->mk_eta.data(), b->mk_mattype.data())))
            return fail(b, rc, "upload_markers");
        b->markers_stale = false;
    }
    const int *hyd = nullptr;
    if (var.hydrous_elemmarkers) {
        b->hyd_count.resize(var.nelem);
        for (int e = 0; e < var.nelem; ++e) b->hyd_count[e] = (*var.hydrous_elemmarkers)[e][0];
        hyd = b->hyd_count.data();
    }
    b->mk_changed.resize(std::max(n, 1)); b->mk_new.resize(std::max(n, 1)); b->mk_hyd.resize(std::max(n, 1));
    int nchanged = 0;
    if ((rc = des_gpu_phase_changes(b->h, opt, hyd, n, &nchanged, b->mk_changed.data(), b->mk_new.data(),
                                    b->mk_hyd.data())))
        return fail(b, rc, "phase_changes");
    for (int k = 0; k < nchanged; ++k) {            // phasechanges.cxx:143-147
        const int m = b->mk_changed[k], e = ms.get_elem(m), current_mt = ms.get_mattype(m), new_mt = b->mk_new[k];
        ms.set_mattype(m, new_mt);
        --(*var.elemmarkers)[e][current_mt];
        ++(*var.elemmarkers)[e][new_mt];
    }
    if (opt == 1)
        for (int m = 0; m < n; ++m) ms.set_tmp(m, b->mk_hyd[m]);                  // :131
    if (param.control.has_hydration_processes) {    // :160-176: a hydrous marker where one dehydrated
        MarkerSet &hydms = *var.markersets[var.hydrous_marker_index];
        auto &hydem = *var.hydrous_elemmarkers;
        for (int e = 0; e < var.nelem; ++e) {
            const int nmarkers = (*var.markers_in_elem)[e].size();
            for (int i = 0; i < nmarkers; ++i) {
                const int m = (*var.markers_in_elem)[e][i];
                if (ms.get_tmp(m) == 0) continue;
                const int el = ms.get_elem(m);
                ConstShapefnAccessor eta = ms.get_eta(m);
                hydms.append_marker(eta, el, 0, 0., 0., 0., 0., 0.);
                ++hydem[el][0];
            }
        }
    }
    return 0;
}

int des_bridge_markers_changed(DesBridge *b, const Param &param, const Variables &var)
{
    b->markers_stale = true;
    flatten_markers(b, param, var);
    int rc = des_gpu_upload_elemmarkers(b->h, b->flat_markers.data());
    if (rc) return fail(b, rc, "upload_elemmarkers");
    return 0;
}

int des_bridge_isostasy_adjustment(DesBridge *b, const Param &param, Variables &var)
{
    std::cout << "Adjusting isostasy for " << param.ic.isostasy_adjustment_time_in_yr << " yrs...\n";
    var.dt = des_bridge_compute_dt(b, param, var);
    if (!(var.dt > 0)) return fail(b, DES_E_NAN, "isostasy: dt <= 0");
    int iso_steps = param.ic.isostasy_adjustment_time_in_yr*YEAR2SEC / var.dt;
    for (int n = 0; n < iso_steps; n++) {
        int rc;
        // update_strain_rate ... update_mesh, dynearthsol.cxx:507-536 (no vbcs: free boundary)
        if ((rc = des_gpu_update_strain_rate(b->h))) return fail(b, rc, "update_strain_rate");
        if ((rc = des_gpu_compute_dvoldt(b->h))) return fail(b, rc, "compute_dvoldt");
        if ((rc = des_gpu_compute_edvoldt(b->h))) return fail(b, rc, "compute_edvoldt");
        if ((rc = des_gpu_update_stress(b->h, var.dt))) return fail(b, rc, "update_stress");
        if ((rc = des_gpu_update_force(b->h))) return fail(b, rc, "update_force");
        if ((rc = des_gpu_update_velocity(b->h, var.dt))) return fail(b, rc, "update_velocity");
        if ((rc = des_gpu_isostasy_constrain_velocity(b->h))) return fail(b, rc, "isostasy_constrain_velocity");
        if ((rc = des_gpu_update_mesh(b->h, var.dt))) return fail(b, rc, "update_mesh");
    }
    std::cout << "Adjusted isostasy for " << iso_steps << " steps.\n";
    return 0;
}

int des_bridge_bad_mesh_quality(DesBridge *b, const Param &param, const Variables &var,
                                int &index, double &min_quality)
{
    DesQualityParam qp;
    qp.smallest_size = param.mesh.smallest_size;
    qp.resolution = param.mesh.resolution;
    qp.max_boundary_distortion = param.mesh.max_boundary_distortion;
    qp.min_quality = param.mesh.min_quality;
    qp.xlength = param.mesh.xlength; qp.ylength = param.mesh.ylength; qp.zlength = param.mesh.zlength;
    qp.remeshing_option = param.mesh.remeshing_option;
    DesQuality q;
    int rc = des_gpu_mesh_quality(b->h, &qp, &q);
    if (rc) { fail(b, rc, "mesh_quality"); return -1; }
    const int code = des_quality_verdict(&qp, &q, &index, &min_quality);
    // the reference's diagnostics (remeshing.cxx:2769, 2789, 2804-2826, 2850-2855)
    if (code == 3)
        std::cout << "    The size of element #" << index << " is too small.\n";
    else if (code == 2 && q.first_bottom_node >= 0)
        std::cout << "    Node #" << index << " is too far from the bottm: z = " << q.bottom_value << "\n";
    else if (code == 2) {
        const uint flag = (*var.bcflag)[index];
        const char *side = (flag & BOUNDX0) ? "x0 side: x = " : (flag & BOUNDX1) ? "x1 side: x = " :
                           (flag & BOUNDY0) ? "y0 side: y = " : "y1 side: y = ";
        std::cout << "    Node #" << index << " is too far from the " << side << q.side_value << "\n";
    }
    else if (code == 1) {
        std::cout << "    Element #" << index << " has mesh quality = " << min_quality << " (bcflag: ";
        for (int i = 0; i < NODES_PER_ELEM; i++) {
            std::cout << (*var.bcflag)[(*var.connectivity)[index][i]];
            std::cout << (i < NODES_PER_ELEM - 1 ? ", " : ").\n");
        }
    }
    return code;
}

int des_bridge_average_fields(DesBridge *b, const Param &param, const Variables &var)
{
    const int interval = param.mesh.quality_check_step_interval;    // Output::average_interval
    const int first = (var.steps % interval == 1);
    if (first) b->avg_time0 = var.time;
    int rc = des_gpu_average_fields(b->h, first);
    if (rc) return fail(b, rc, "average_fields");
    return 0;
}

int des_bridge_flush_averages(DesBridge *b, const Param &param, Variables &var)
{
    // Output keeps time0 / coord0 / strain0 / stress_avg / delta_plstrain_avg private.  Its own
    // first-step branch (output.cxx:317-339) copies exactly those five from `var`, so it is
    // shown a `var` whose time, coord, strain, stress and delta_plstrain ARE the accumulators.
    array_t coord0(var.nnode);
    tensor_t strain0(var.nelem), stress_avg(var.nelem);
    double_vec dpl(var.nelem);
    int rc = des_gpu_download_averages(b->h, coord0.data(), strain0.data(), stress_avg.data(), dpl.data());
    if (rc) return fail(b, rc, "download_averages");
    array_t *coord = var.coord; tensor_t *strain = var.strain, *stress = var.stress;
    double_vec *delta_plstrain = var.delta_plstrain;
    const double time = var.time; const int steps = var.steps;
    var.coord = &coord0; var.strain = &strain0; var.stress = &stress_avg; var.delta_plstrain = &dpl;
    var.time = b->avg_time0;
    var.steps = param.mesh.quality_check_step_interval + 1;     // steps % average_interval == 1
    var.output->average_fields(var);
    var.coord = coord; var.strain = strain; var.stress = stress; var.delta_plstrain = delta_plstrain;
    var.time = time; var.steps = steps;
    return 0;
}

DesGpu *des_bridge_handle(DesBridge *b) { return b ? b->h : nullptr; }
double des_bridge_avg_time0(const DesBridge *b) { return b->avg_time0; }

const char *des_bridge_last_error(const DesBridge *b) { return b ? b->err.c_str() : "null bridge"; }

void des_bridge_destroy(DesBridge *b)
{
    if (!b) return;
    if (b->h) des_gpu_destroy(b->h);
    delete b;
}
