// phase_changes (phasechanges.cxx:109-158) on device-resident markers (SURVEY.md section 8f row 2;
// include/des_gpu.h, "phase changes" block).  The reference walks every marker of every element
// each 10 steps (dynearthsol.cxx:885): depth and temperature at the marker from the element's
// nodes and the marker's shape-function weights (MarkerSet::get_ZPT, markerset.cxx:973-986),
// pressure from ref_pressure (matprops.cxx:153-174), then the rock-type table of
// simple_subduction (phasechanges.cxx:10-84).  Here the marker arrays of the material marker set
// (element, weights, rock type) stay on the device between the points where host code changes
// them, the lookup reads the resident coordinates and temperature, applies the changes to the
// resident rock types and elemmarkers counters (integer atomics: the counts do not depend on the
// order) and hands back only the markers that changed, in ascending marker order.
#pragma once

namespace {

// One thread per marker.  Same statements as get_ZPT / simple_subduction, same order.
__global__ void k_phase_changes(int nmarkers, int nmat, DevParams P, Dev D, const int *__restrict__ einv,
                                const int *__restrict__ m_elem, const double *__restrict__ m_eta, int eta_stride,
                                int *__restrict__ m_mattype, const int *__restrict__ hydrous_count,
                                int *__restrict__ elemmarkers, int *__restrict__ changed, int *__restrict__ hyd_inc_out)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= nmarkers) return;
    const int el = m_elem[m];                       // the caller's element id
    const int ei = einv ? einv[el] : el;
    const int4 c4 = D.conn[ei];
    const int conn[4] = {c4.x, c4.y, c4.z, c4.w};
    double Z = 0, T = 0;
    for (int i = 0; i < 4; ++i) {
        const double4 x = D.xT[conn[i]];
        const double eta = m_eta[size_t(i) * eta_stride + m];
        Z += x.z * eta;
        T += x.w * eta;
    }
    const double Pm = ref_pressure(P, Z);
    const int mt_mantle = 0, mt_serpentinized_mantle = 1, mt_oceanic_crust = 2, mt_eclogite = 3,
              mt_sediment = 4, mt_schist = 5;
    const int current_mt = m_mattype[m];
    int new_mt = current_mt, hyd_inc = 0;
    switch (current_mt) {
    case mt_oceanic_crust: {                        // basalt -> eclogite (phasechanges.cxx:28-41)
        const double min_eclogite_T = 500 + 273;
        const double transition_pressure = -0.3e9 + 2.2e6 * T;
        const double dehydration_T = 150 + 273;
        if (T > min_eclogite_T && Pm > transition_pressure) new_mt = mt_eclogite;
        else if (T > dehydration_T) hyd_inc = 1;
    } break;
    case mt_sediment: {                             // sediment -> schist (:43-56)
        const double min_schist_T = 650 + 273;
        const double min_schist_Z = -20e3;
        const double dehydration_T = 150 + 273;
        if (T > min_schist_T && Z < min_schist_Z) new_mt = mt_schist;
        else if (T > dehydration_T) hyd_inc = 1;
    } break;
    case mt_serpentinized_mantle: {                 // serpentinite -> mantle (:58-69)
        const double transition_pressure = 2.1e9 + (7.5e9 - 2.1e9) * (T - (730 + 273)) / (500 - 730);
        const double min_serpentine_T = 550 + 273;
        if (T > min_serpentine_T && Pm > transition_pressure) { new_mt = mt_mantle; hyd_inc = 1; }
    } break;
    case mt_mantle: {                               // mantle -> serpentinite where hydrous markers are (:71-80)
        const double min_serpentine_T = 550 + 273;
        if (T <= min_serpentine_T && hydrous_count && hydrous_count[el]) new_mt = mt_serpentinized_mantle;
    } break;
    default: break;
    }
    if (hyd_inc_out) hyd_inc_out[m] = hyd_inc;
    const int ch = (new_mt != current_mt);
    changed[m] = ch;
    if (ch) {
        m_mattype[m] = new_mt;                      // phasechanges.cxx:143-147
        atomicSub(elemmarkers + size_t(ei) * nmat + current_mt, 1);
        atomicAdd(elemmarkers + size_t(ei) * nmat + new_mt, 1);
    }
}

__global__ void k_phase_collect(int n, const int *__restrict__ changed, const int *__restrict__ pos,
                                const int *__restrict__ mattype, int cap, int *__restrict__ out_marker,
                                int *__restrict__ out_mattype)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < n && changed[m] && pos[m] < cap) { out_marker[pos[m]] = m; out_mattype[pos[m]] = mattype[m]; }
}
__global__ void k_marker_range(int n, int nelem, int nmat, const int *__restrict__ elem,
                               const int *__restrict__ mattype, int *__restrict__ bad)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < n && (elem[m] < 0 || elem[m] >= nelem || mattype[m] < 0 || mattype[m] >= nmat)) *bad = 1;
}

} // namespace

extern "C" {

int des_gpu_upload_markers(DesGpu *h, int nmarkers, int stride, const int *elem, const double *eta,
                           const int *mattype)
{
    if (!h || nmarkers < 0 || stride < nmarkers || (nmarkers && (!elem || !eta || !mattype))) return DES_E_ARG;
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "device-resident markers: single-device handles only");
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    CK(cudaSetDevice(h->device));
    const size_t n = nmarkers;
    if (n > h->mk_cap) {                            // mesh-lifetime allocations, grown when needed
        int rc;
        const size_t cap = n + n / 8 + 1024;
        void *old[] = { h->mk_elem, h->mk_mattype, h->mk_eta, h->mk_changed, h->mk_pos, h->mk_hyd, h->mk_out,
                        h->mk_scan_tmp };
        for (void *q : old) {                       // the smaller set is released now, not at the next upload_mesh
            if (!q) continue;
            auto it = std::find(h->mesh_allocs.begin(), h->mesh_allocs.end(), q);
            if (it != h->mesh_allocs.end()) { cudaFree(q); h->mesh_allocs.erase(it); }
        }
        h->mk_cap = 0; h->mk_n = 0;
        if ((rc = dalloc(h, &h->mk_elem, cap))) return rc;
        if ((rc = dalloc(h, &h->mk_mattype, cap))) return rc;
        if ((rc = dalloc(h, &h->mk_eta, cap * 4))) return rc;
        if ((rc = dalloc(h, &h->mk_changed, cap))) return rc;
        if ((rc = dalloc(h, &h->mk_pos, cap))) return rc;
        if ((rc = dalloc(h, &h->mk_hyd, cap))) return rc;
        if ((rc = dalloc(h, &h->mk_out, cap * 2 + 2))) return rc;
        size_t tmp = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, h->mk_changed, h->mk_pos, int(cap)));
        if ((rc = dalloc(h, &h->mk_scan_tmp, tmp + 16))) return rc;
        h->mk_scan_bytes = tmp + 16;
        h->mk_cap = cap;
    }
    h->mk_n = nmarkers;
    if (nmarkers == 0) return DES_OK;
    CK(cudaMemcpyAsync(h->mk_elem, elem, n * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->mk_mattype, mattype, n * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    for (int i = 0; i < 4; ++i)                     // shapefn is Array2D<double, 4> SoA with its own stride
        CK(cudaMemcpyAsync(h->mk_eta + size_t(i) * h->mk_cap, eta + size_t(i) * stride, n * sizeof(double),
                           cudaMemcpyHostToDevice, h->stream));
    int *bad = h->mk_out;
    CK(cudaMemsetAsync(bad, 0, sizeof(int), h->stream));
    LAUNCH(KC_PACK, k_marker_range, nmarkers, 256, nmarkers, h->D.nelem, h->hp.nmat, h->mk_elem, h->mk_mattype, bad);
    int hb = 0;
    CK(cudaMemcpyAsync(&hb, bad, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (hb) { h->mk_n = 0; return fail(h, DES_E_ARG, "upload_markers: element or rock type out of range"); }
    return DES_OK;
}

int des_gpu_phase_changes(DesGpu *h, int option, const int *hydrous_count, int cap, int *nchanged,
                          int *changed_marker, int *changed_mattype, int *hyd_inc)
{
    if (!h || !nchanged || cap < 0 || (cap && (!changed_marker || !changed_mattype))) return DES_E_ARG;
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "device-resident markers: single-device handles only");
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    // phasechanges.cxx:128-141: 1 simple_subduction, 101 the custom template (which changes
    // nothing); anything else is the reference's EXIT_CONFIG_VALUE
    if (option != 1 && option != 101) return fail(h, DES_E_UNSUPPORTED, "unknown phase_change_option");
    if (option == 1 && h->hp.nmat < 6) return fail(h, DES_E_ARG, "simple_subduction needs rock types 0..5");
    *nchanged = 0;
    const int n = h->mk_n;
    if (option == 101 || n == 0) {
        if (hyd_inc) memset(hyd_inc, 0, size_t(n) * sizeof(int));
        return DES_OK;
    }
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);     // the pending mass computation must see the OLD markers
    if (rc) return rc;
    int *d_hyd = nullptr;
    if (hydrous_count) {
        if ((rc = ensure_stage(h, size_t(h->D.nelem) * sizeof(int)))) return rc;
        d_hyd = static_cast<int *>(h->stage);
        if ((rc = h2d(h, d_hyd, hydrous_count, size_t(h->D.nelem)))) return rc;
    }
    LAUNCH(KC_PACK, k_phase_changes, n, 128, n, h->hp.nmat, h->P, h->D, h->einv, h->mk_elem, h->mk_eta,
           int(h->mk_cap), h->mk_mattype, d_hyd, const_cast<int *>(h->D.elemmarkers), h->mk_changed,
           hyd_inc ? h->mk_hyd : nullptr);
    size_t tmp = h->mk_scan_bytes;
    CK(cub::DeviceScan::ExclusiveSum(h->mk_scan_tmp, tmp, h->mk_changed, h->mk_pos, n, h->stream));
    LAUNCH(KC_PACK, k_phase_collect, n, 256, n, h->mk_changed, h->mk_pos, h->mk_mattype, cap, h->mk_out,
           h->mk_out + h->mk_cap);
    int last[2] = {0, 0};
    CK(cudaMemcpyAsync(&last[0], h->mk_pos + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&last[1], h->mk_changed + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    const int total = last[0] + last[1];
    *nchanged = total;
    const int take = std::min(total, cap);
    if (take) {
        CK(cudaMemcpyAsync(changed_marker, h->mk_out, size_t(take) * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(changed_mattype, h->mk_out + h->mk_cap, size_t(take) * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    }
    if (hyd_inc) CK(cudaMemcpyAsync(hyd_inc, h->mk_hyd, size_t(n) * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DES_OK;
}

int des_gpu_download_elemmarkers(DesGpu *h, int *elemmarkers)
{
    if (!h || !elemmarkers) return DES_E_ARG;
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "download_elemmarkers: single-device handles only");
    if (!h->have_mesh) return fail(h, DES_E_STATE, "upload_mesh first");
    CK(cudaSetDevice(h->device));
    const int E = h->D.nelem, nmat = h->hp.nmat;
    std::vector<int> tmp(size_t(E) * nmat);
    CK(cudaMemcpyAsync(tmp.data(), h->D.elemmarkers, tmp.size() * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->h_einv.empty()) { memcpy(elemmarkers, tmp.data(), tmp.size() * sizeof(int)); return DES_OK; }
    for (int e = 0; e < E; ++e)                     // h_einv: caller -> internal
        memcpy(elemmarkers + size_t(e) * nmat, tmp.data() + size_t(h->h_einv[e]) * nmat, nmat * sizeof(int));
    return DES_OK;
}

} // extern "C"
