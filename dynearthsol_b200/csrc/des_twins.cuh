// Fine-grained twins of the reference calls (include/des_gpu.h, last block): the
// un-fused loop nests, on the same device state, built from the same device functions and
// kernels as the fused step (run with a subset of their flags).  They exist for the
// reference's PT loop (dynearthsol.cxx:803-864), isostasy_adjustment (:506-539), the
// post-remesh calls (remeshing.cxx:3133-3168) and the per-function parity tests.
#pragma once

namespace {

// compute_dvoldt element loop on the STORED strain rate, geometry.cxx:217-225
__global__ void k_tw_dvol_elem(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const size_t E = D.nelem;
    const double dj = D.strain_rate[e] + D.strain_rate[E + e] + D.strain_rate[2 * E + e];
    D.ca_hi[size_t(e) * 4] = dj * D.volume[e];
}

// compute_edvoldt, geometry.cxx:263-272
__global__ void k_tw_edvoldt(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const int4 c = D.conn[e];
    double dj = 0;
    dj += D.ntmp[c.x]; dj += D.ntmp[c.y]; dj += D.ntmp[c.z]; dj += D.ntmp[c.w];
    D.edvoldt[e] = dj / 4;
}

// NMD_stress first loop, geometry.cxx:293-295
__global__ void k_tw_nmd_pre(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    D.etmp[e] = D.dpressure[e] * D.volume[e];
}

// NMD_stress third loop, geometry.cxx:315-331
__global__ void k_tw_nmd_elem(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const size_t E = D.nelem;
    const int4 c = D.conn[e];
    double dp = 0;
    dp += D.ntmp[c.x]; dp += D.ntmp[c.y]; dp += D.ntmp[c.z]; dp += D.ntmp[c.w];
    const double dp_el = dp / 4;
    const double ddp = (-D.dpressure[e] + dp_el) / 3;
    for (int i = 0; i < 3; ++i) D.stress[size_t(i) * E + e] += ddp;
}

// compute_volume, geometry.cxx:180-197
__global__ void k_tw_volume(const __grid_constant__ Dev D)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const int4 c = D.conn[e];
    D.volume[e] = tet_volume(D.xT[c.x], D.xT[c.y], D.xT[c.z], D.xT[c.w]);
}

// calculate_residual_force, fields.cxx:700-722.  force_residual[n] is the contribution of
// the LAST support element (fields.cxx:673 assigns, it does not accumulate).  The sum over
// nodes is an unordered reduction in the reference too (OpenMP reduction(+)).
__global__ void k_tw_residual(const __grid_constant__ Dev D, double *acc)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    double l2 = 0;
    if (n < D.nnode) {
        const double num = double(D.nnode) * 3;
        int end = D.sup_idx[n + 1];
        while (end > D.sup_idx[n] && D.sup_ent[end - 1] < 0) --end;     // padding
        if (end > D.sup_idx[n]) {
            const int ent = D.sup_ent[end - 1];
            const double *tr = D.fc + DES_FC_OFF(ent);
            for (int j = 0; j < 3; ++j) l2 += tr[j] * tr[j] / num;
        }
    }
    for (int o = 16; o > 0; o >>= 1) l2 += __shfl_xor_sync(0xffffffffu, l2, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(acc, l2);
}

int twin_ready(DesGpu *h)
{
    if (!h) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "the fine-grained twins run on a single-device handle");
    CK(cudaSetDevice(h->device));
    return flush_tail(h);
}

} // namespace

extern "C" {

int des_gpu_update_temperature(DesGpu *h, double dt)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, (k_elem_a<false, true>), h->D.nelem, 256, h->P, h->D, 0.0, 0, 1);
    LAUNCH(KC_TWIN, (k_node_a<false, true>), h->D.nnode, 256, h->P, h->D, dt, 1);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_update_strain_rate(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, (k_elem_a<false, true>), h->D.nelem, 256, h->P, h->D, 0.0, 0, 2);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_compute_dvoldt(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_tw_dvol_elem, h->D.nelem, 256, h->D);
    LAUNCH(KC_TWIN, (k_node_a<false, true>), h->D.nnode, 256, h->P, h->D, 0.0, 4);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_compute_edvoldt(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_tw_edvoldt, h->D.nelem, 256, h->D);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_update_stress(DesGpu *h, double dt)
{
    int rc = twin_ready(h); if (rc) return rc;
    switch (h->hp.rheol_type) {     // E-B itself on the stored strain rate and edvoldt
    case DES_RH_ELASTIC: LAUNCH(KC_TWIN, (k2_elem_b<false, false, DES_RH_ELASTIC, false, true>), h->D.nelem, 128, h->P, h->D, dt, 0.0, 0, 0); break;
    case DES_RH_MAXWELL: LAUNCH(KC_TWIN, (k2_elem_b<false, false, DES_RH_MAXWELL, false, true>), h->D.nelem, 128, h->P, h->D, dt, 0.0, 0, 0); break;
    case DES_RH_EP:      LAUNCH(KC_TWIN, (k2_elem_b<false, false, DES_RH_EP, false, true>), h->D.nelem, 128, h->P, h->D, dt, 0.0, 0, 0); break;
    default:             LAUNCH(KC_TWIN, (k2_elem_b<false, false, DES_RH_EVP, false, true>), h->D.nelem, 128, h->P, h->D, dt, 0.0, 0, 0); break;
    }
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_nmd_stress(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_tw_nmd_pre, h->D.nelem, 256, h->D);
    LAUNCH(KC_TWIN, k_node_b, h->D.nnode, 256, h->D);
    LAUNCH(KC_TWIN, k_tw_nmd_elem, h->D.nelem, 256, h->D);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_update_force(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_sbc_facets, h->D.nsbc_facets, 128, h->P, h->D, h->compensation_pressure);
    LAUNCH(KC_TWIN, k_nbc_facets, h->D.nnbc_facets, 128, h->P, h->D);
    LAUNCH_M(KC_TWIN, k_elem_c, false, h->D.nelem, 256, h->P, h->D, 0);
    LAUNCH(KC_TWIN, k_node_c, h->D.nnode, 256, h->P, h->D, 0.0, 0, 1);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_update_velocity(DesGpu *h, double dt)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_node_c, h->D.nnode, 256, h->P, h->D, dt, 0, 2);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_apply_vbcs(DesGpu *h, double time)
{
    int rc = twin_ready(h); if (rc) return rc;
    int bc_z1 = h->hp.vbc_types[5];
    if (time > h->hp.vbc_val_z1_loading_period) bc_z1 = 0;
    DevParams P = h->P;
    if (h->pt_jump) for (int i = 0; i < 6; ++i) P.vbc_values[i] = 0.0;      // bc.cxx:330-344
    LAUNCH(KC_TWIN, k_node_c, h->D.nnode, 256, P, h->D, 0.0, bc_z1, 4);
    CK(cudaGetLastError());
    return DES_OK;
}

// update_mesh, dynearthsol.cxx:448-492: update_coordinate, surface_processes (not while
// PT_jump is set), volume swap, compute_volume, compute_mass.  part 0 = up to and including
// surface_processes (the host's surface-marker routines run after it on quality-check steps,
// bc.cxx:1829-1843), part 1 = the rest, -1 = all of it.
static int update_mesh_impl(DesGpu *h, double dt, int part)
{
    int rc = twin_ready(h); if (rc) return rc;
    if (part != 1) {
        LAUNCH(KC_TWIN, k_node_c, h->D.nnode, 256, h->P, h->D, dt, 0, 8);
        if (!h->pt_jump && (rc = run_surface(h, dt, true))) return rc;
        h->last_dt = dt;
    }
    if (part != 0) {
        std::swap(h->D.volume, h->D.volume_old);
        LAUNCH(KC_TWIN, (k_elem_a<true, false>), h->D.nelem, 256, h->P, h->D, 0.0, 3, 0);
        LAUNCH(KC_TWIN, (k_node_a<true, false>), h->D.nnode, 256, h->P, h->D, 0.0, 0);
        if ((rc = ymass_update(h))) return rc;
    }
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_update_mesh(DesGpu *h, double dt) { return update_mesh_impl(h, dt, -1); }

int des_gpu_update_mesh_part(DesGpu *h, double dt, int part)
{
    if (part != 0 && part != 1) return DES_E_ARG;
    return update_mesh_impl(h, dt, part);
}

int des_gpu_isostasy_constrain_velocity(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_isostasy_velocity, h->D.nnode, 256, h->D, h->hp.has_winkler_foundation ? 0 : 1);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_compute_volume(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, k_tw_volume, h->D.nelem, 256, h->D);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_compute_mass(DesGpu *h)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, (k_elem_a<true, false>), h->D.nelem, 256, h->P, h->D, 0.0, 2, 0);
    LAUNCH(KC_TWIN, (k_node_a<true, false>), h->D.nnode, 256, h->P, h->D, 0.0, 0);
    if ((rc = ymass_update(h))) return rc;
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_rotate_stress(DesGpu *h, double dt)
{
    int rc = twin_ready(h); if (rc) return rc;
    LAUNCH(KC_TWIN, (k_elem_a<true, false>), h->D.nelem, 256, h->P, h->D, dt, 4, 0);
    CK(cudaGetLastError());
    return DES_OK;
}

int des_gpu_residual_force(DesGpu *h, double *l2)
{
    int rc = twin_ready(h); if (rc) return rc;
    if (!l2) return DES_E_ARG;
    if ((rc = ensure_stage(h, sizeof(double)))) return rc;
    double *acc = static_cast<double *>(h->stage);
    CK(cudaMemsetAsync(acc, 0, sizeof(double), h->stream));
    LAUNCH(KC_TWIN, k_tw_residual, h->D.nnode, 256, h->D, acc);
    double v = 0;
    CK(cudaMemcpyAsync(&v, acc, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *l2 = std::sqrt(v);
    return DES_OK;
}

} // extern "C"
