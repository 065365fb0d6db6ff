// des_gpu_download_frame (include/des_gpu.h): the arrays of one output
// frame, Output::_write (output.cxx:95-245), leave the device already in the layout BinaryOutput
// puts on disk -- Array2D::pack_to's [n][components] (binaryio.cxx:151-166) -- and with
// _write's own host loops applied: the averaged variants (output.cxx:131-190), density
// (MatProps::rho, :192-196), mesh quality (elem_quality, :198-202) and the most abundant
// marker type (:214-220).  The host then only has to fwrite the buffers.
// All of it is HBM-bound transposition: 8 B read + 8 B written per value, then PCIe.
#pragma once

namespace {

// node records {x, y, z, T} -> [n][3] in the caller's node order
__global__ void k_frame_vec3(int n, const double4 *__restrict__ in, const int *__restrict__ nperm,
                             double *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = nperm ? nperm[i] : i;
    const double4 v = in[i];
    out[o * 3] = v.x; out[o * 3 + 1] = v.y; out[o * 3 + 2] = v.z;
}

// "velocity averaged", output.cxx:131-135: (coord - coord0) * inv_dt; coord0 is the snapshot
// of des_gpu_average_fields, [3][n] in the caller's node order
__global__ void k_frame_vel_avg(int n, const double4 *__restrict__ xT, const int *__restrict__ nperm,
                                const double *__restrict__ c0, double inv_dt, double *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = nperm ? nperm[i] : i;
    const double4 v = xT[i];
    out[o * 3] = (v.x - c0[o]) * inv_dt;
    out[o * 3 + 1] = (v.y - c0[size_t(n) + o]) * inv_dt;
    out[o * 3 + 2] = (v.z - c0[size_t(2) * n + o]) * inv_dt;
}

// rows [rows][n] in the internal order -> [n][rows] in the caller's order.
// MODE 0: a;  1: a * scale (output.cxx:152, 186);  2: (a - b) * scale (output.cxx:164)
template <int MODE>
__global__ void k_frame_rows(int n, int rows, const double *__restrict__ a, const double *__restrict__ b,
                             double scale, const int *__restrict__ perm, double *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = perm ? perm[i] : i;
    for (int r = 0; r < rows; ++r) {
        double v = a[size_t(r) * n + i];
        if (MODE == 1) v = v * scale;
        if (MODE == 2) v = (v - b[size_t(r) * n + i]) * scale;
        out[o * rows + r] = v;
    }
}

// density, mesh quality and material of every element (output.cxx:192-220), caller's order.
// out = three consecutive [nelem] arrays.
template <bool SINGLE>
__global__ void k_frame_elem_derived(const __grid_constant__ DevParams P, const __grid_constant__ Dev D,
                                     const int *__restrict__ eperm, double normf, double *__restrict__ out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D.nelem) return;
    const int nmat = SINGLE ? 1 : P.nmat;
    const size_t E = D.nelem, o = eperm ? eperm[e] : e;
    const int4 c = D.conn[e];
    const double4 a = D.xT[c.x], b = D.xT[c.y], cc = D.xT[c.z], d = D.xT[c.w];
    Markers mk;
    load_markers(D.elemmarkers, e, nmat, mk);
    out[o] = mat_rho(P, nmat, mk, a.w, b.w, cc.w, d.w);
    // elem_quality, geometry.cxx:1893-1899
    const double vol = D.volume[e];
    const double area_sum = tri_area(a, b, cc) + tri_area(a, b, d) + tri_area(cc, d, a) + tri_area(cc, d, b);
    out[E + o] = normf * vol * vol / (area_sum * area_sum * area_sum);
    // std::max_element: the first of equal maxima
    int best = 0;
    for (int m = 1; m < nmat; ++m) if (mk.cnt[m] > mk.cnt[best]) best = m;
    out[2 * E + o] = double(best);
}

} // namespace

extern "C" {

int des_host_alloc(size_t bytes, void **out)
{
    if (!out) return DES_E_ARG;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        *out = nullptr;
        return DES_E_CUDA;
    }
    return DES_OK;
}

int des_host_free(void *p)
{
    if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return DES_E_CUDA; }
    return DES_OK;
}

int des_gpu_download_frame(DesGpu *h, const DesFrame *f)
{
    if (!h || !f) return DES_E_ARG;
    if (!h->have_mesh || !h->have_state) return fail(h, DES_E_STATE, "upload mesh and state first");
    if (h->multi) return fail(h, DES_E_UNSUPPORTED, "download_frame on a group handle: use des_gpu_download_state");
    const bool avg = f->averaged != 0;
    if (avg && !h->avg_valid) return fail(h, DES_E_STATE, "averaged frame without des_gpu_average_fields");
    if (!avg && (f->velocity_averaged || f->stress_averaged))
        return fail(h, DES_E_ARG, "velocity_averaged / stress_averaged need averaged = 1");
    CK(cudaSetDevice(h->device));
    int rc = flush_tail(h);         // resolves the deferred end-of-step work and a pending average
    if (rc) return rc;
    Dev &D = h->D;
    const size_t N = D.nnode, E = D.nelem;
    // one staging area, one stream: transpose, copy, transpose, copy ... and a single
    // synchronisation at the end (the kernels are ~1 % of the PCIe time of their arrays)
    const size_t area = std::max(E * 6, N * 3);
    if ((rc = ensure_stage(h, area * sizeof(double)))) return rc;
    double *st = static_cast<double *>(h->stage);
#define FRAME_OUT(dst, count) CK(cudaMemcpyAsync(dst, st, (count) * sizeof(double), cudaMemcpyDeviceToHost, h->stream))
    const int n = int(N), e = int(E);
    if (f->coordinate) {
        LAUNCH(KC_PACK, k_frame_vec3, n, 256, n, D.xT, h->nperm, st);
        FRAME_OUT(f->coordinate, N * 3);
    }
    if (f->velocity) {
        LAUNCH(KC_PACK, k_frame_vec3, n, 256, n, D.vv, h->nperm, st);
        FRAME_OUT(f->velocity, N * 3);
    }
    if (f->velocity_averaged) {
        LAUNCH(KC_PACK, k_frame_vel_avg, n, 256, n, D.xT, h->nperm, h->avg_coord0, f->inv_dt, st);
        FRAME_OUT(f->velocity_averaged, N * 3);
    }
    if (f->temperature) {
        LAUNCH(KC_PACK, k_unpack_w, n, 256, n, D.xT, h->nperm, st);
        FRAME_OUT(f->temperature, N);
    }
    if (f->force) {
        LAUNCH(KC_PACK, k_frame_rows<0>, n, 256, n, 3, D.force, nullptr, 0.0, h->nperm, st);
        FRAME_OUT(f->force, N * 3);
    }
#define FRAME_ROWS(dst, MODE, rows, a, b, scale) if (dst) { \
        LAUNCH(KC_PACK, k_frame_rows<MODE>, e, 256, e, rows, a, b, scale, h->eperm, st); \
        FRAME_OUT(dst, E * (rows)); }
    FRAME_ROWS(f->plstrain, 0, 1, D.plstrain, nullptr, 0.0)
    if (avg) { FRAME_ROWS(f->plstrain_rate, 1, 1, D.avg_dpl, nullptr, f->inv_dt) }
    else     { FRAME_ROWS(f->plstrain_rate, 0, 1, D.delta_plstrain, nullptr, 0.0) }
    if (avg) { FRAME_ROWS(f->strain_rate, 2, 6, D.strain, D.avg_strain0, f->inv_dt) }
    else     { FRAME_ROWS(f->strain_rate, 0, 6, D.strain_rate, nullptr, 0.0) }
    FRAME_ROWS(f->strain, 0, 6, D.strain, nullptr, 0.0)
    FRAME_ROWS(f->stress, 0, 6, D.stress, nullptr, 0.0)
    FRAME_ROWS(f->viscosity, 0, 1, D.viscosity, nullptr, 0.0)
    FRAME_ROWS(f->stress_averaged, 1, 6, D.avg_stress, nullptr, f->stress_avg_scale)
#undef FRAME_ROWS
    if (f->density || f->mesh_quality || f->material) {
        const double normf = 216 * std::sqrt(3);        // geometry.cxx:1886
        if (h->hp.nmat == 1) LAUNCH(KC_PACK, k_frame_elem_derived<true>, e, 256, h->P, D, h->eperm, normf, st);
        else LAUNCH(KC_PACK, k_frame_elem_derived<false>, e, 256, h->P, D, h->eperm, normf, st);
        const size_t bytes = E * sizeof(double);
        if (f->density) CK(cudaMemcpyAsync(f->density, st, bytes, cudaMemcpyDeviceToHost, h->stream));
        if (f->mesh_quality) CK(cudaMemcpyAsync(f->mesh_quality, st + E, bytes, cudaMemcpyDeviceToHost, h->stream));
        if (f->material) CK(cudaMemcpyAsync(f->material, st + 2 * E, bytes, cudaMemcpyDeviceToHost, h->stream));
    }
#undef FRAME_OUT
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    return DES_OK;
}

} // extern "C"
