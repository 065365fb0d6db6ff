// Link-time swap of the two field-transfer calls of remesh().  The
// bridge binary is linked with --wrap on
//   nearest_neighbor_interpolation   (nn-interpolation.cxx:661-702)
//   barycentric_node_interpolation   (brc-interpolation.cxx:345-423)
// so the calls in remesh() (remeshing.cxx:3058-3069) -- the reference's own, unmodified object
// code -- arrive here and go to the library's device implementation:
//   des_remesh_transfer_cells   nearest old centroid, volume-ratio mapping through sample points
//                               located among the 32 nearest old cells, inject_field per field
//   des_remesh_transfer_nodes   enclosing old element of every new node (nearest node, one ring,
//                               capped breadth-first search, fallbacks), interpolate_field per field
// What this file repeats from the reference is only which members of `Variables` each call
// replaces (nn-interpolation.cxx:526-650, brc-interpolation.cxx:353-420): every one of them is
// re-newed at the new size and the old array deleted, exactly as the reference does, so the rest
// of remesh() finds what it expects.  Same bits as the reference's own transfer
// (tests/test_transfer_gpu.py; the remeshing bridge tests compare the frames after the remesh
// with the reference binary's).  DES_BRIDGE_TRANSFER=ref sends both calls on to the reference's
// own functions (__real_).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "constants.hpp"
#include "parameters.hpp"
#include "barycentric-fn.hpp"
#include "geometry.hpp"
#include "utils.hpp"
#include "../include/des_gpu.h"

#ifndef THREED
#error "3D only"
#endif

#define DES_NN  _Z30nearest_neighbor_interpolationRK5ParamR9VariablesRK26Barycentric_transformationRK7Array2DIdLi3EERKS7_IiLi4EEb
#define DES_BRC _Z30barycentric_node_interpolationRK5ParamR9VariablesRK26Barycentric_transformationRK7Array2DIdLi3EERKS7_IiLi4EE
#define DES_BRCT _Z35barycentric_node_interpolation_forTRK5ParamRK9VariablesRK26Barycentric_transformationRK7Array2DIdLi3EERKS8_IiLi4EERK7SupportRKSt6vectorIdSaIdEERSK_
#define DES_CAT2(a, b) a##b
#define DES_CAT(a, b) DES_CAT2(a, b)
extern "C" void DES_CAT(__real_, DES_NN)(const Param &, Variables &, const Barycentric_transformation &,
                                         const array_t &, const conn_t &, const bool);
extern "C" void DES_CAT(__real_, DES_BRC)(const Param &, Variables &, const Barycentric_transformation &,
                                          const array_t &, const conn_t &);

extern "C" void DES_CAT(__real_, DES_BRCT)(const Param &, const Variables &, const Barycentric_transformation &,
                                           const array_t &, const conn_t &, const Support &, const double_vec &,
                                           double_vec &);

namespace {
bool use_reference()
{
    const char *s = std::getenv("DES_BRIDGE_TRANSFER");
    return s && std::strcmp(s, "ref") == 0;
}
int transfer_device()
{
    const char *s = std::getenv("DES_BRIDGE_DEVICE");
    return s ? std::atoi(s) : 0;
}
[[noreturn]] void transfer_die(const char *what, int rc)
{
    std::fprintf(stderr, "des_bridge: %s failed (code %d); there is no CPU fallback\n", what, rc);
    std::exit(1);
}

// the members one call replaces: old array in, freshly allocated array of the new size out
struct Swap {
    std::vector<DesTransferField> fields;
    std::vector<double_vec **> scalars; std::vector<double_vec *> new_scalars;
    std::vector<array_t **> vectors;    std::vector<array_t *> new_vectors;
    std::vector<tensor_t **> tensors;   std::vector<tensor_t *> new_tensors;
    int n_new;
    explicit Swap(int n) : n_new(n) {}
    void add(double_vec *&v)
    {
        if (!v) return;                         // switched off by remesh() (the SPR chain)
        double_vec *t = new double_vec(n_new);
        scalars.push_back(&v); new_scalars.push_back(t);
        fields.push_back(DesTransferField{ v->data(), t->data(), 1 });
    }
    void add(array_t *&v)
    {
        if (!v) return;
        array_t *t = new array_t(n_new);
        vectors.push_back(&v); new_vectors.push_back(t);
        fields.push_back(DesTransferField{ v->data(), t->data(), NDIMS });
    }
    void add(tensor_t *&v)
    {
        if (!v) return;
        tensor_t *t = new tensor_t(n_new);
        tensors.push_back(&v); new_tensors.push_back(t);
        fields.push_back(DesTransferField{ v->data(), t->data(), NSTR });
    }
    void commit()
    {
        for (size_t i = 0; i < scalars.size(); ++i) { delete *scalars[i]; *scalars[i] = new_scalars[i]; }
        for (size_t i = 0; i < vectors.size(); ++i) { delete *vectors[i]; *vectors[i] = new_vectors[i]; }
        for (size_t i = 0; i < tensors.size(); ++i) { delete *tensors[i]; *tensors[i] = new_tensors[i]; }
    }
};

const bool verbose = std::getenv("DES_BRIDGE_TRANSFER_TIMING") != nullptr;
double g_transfer_seconds = 0;      // device field transfers since the last des_bridge_take_transfer_seconds()
void account(const DesTransferStats &st) { g_transfer_seconds += st.search_seconds + st.acm_seconds + st.field_seconds; }
} // namespace

// seconds the device field transfers took inside the remesh() that just returned (0 with
// DES_BRIDGE_TRANSFER=ref); the driver prints them in its remesh-cycle line
double des_bridge_take_transfer_seconds()
{
    const double s = g_transfer_seconds;
    g_transfer_seconds = 0;
    return s;
}

extern "C" void DES_CAT(__wrap_, DES_NN)(const Param &param, Variables &var,
                                         const Barycentric_transformation &bary,
                                         const array_t &old_coord, const conn_t &old_conn,
                                         const bool is_surface)
{
    if (use_reference()) { DES_CAT(__real_, DES_NN)(param, var, bary, old_coord, old_conn, is_surface); return; }
    DesTransferMeshes m;
    std::memset(&m, 0, sizeof(m));
    m.old_nnode = old_coord.size(); m.old_coord = old_coord.data();
    m.old_ncell = old_conn.size();  m.old_conn = old_conn.data();
    m.new_nnode = var.nnode;        m.new_coord = var.coord->data();
    std::vector<double> area;
    DesTransferStats st;
    if (is_surface) {
        // the areas remesh() gave bary_surface (remeshing.cxx:2889-2892): the reference's own
        // function on the same coordinates
        const int n_old = int(old_conn.size());
        area.resize(n_old);
        for (int i = 0; i < n_old; ++i)
            area[i] = compute_area_facet(old_coord.view_const(old_conn[i]));
        m.old_size = area.data();
        m.new_ncell = var.connectivity_surface->size(); m.new_conn = var.connectivity_surface->data();
        Swap sw(m.new_ncell);
        sw.add(var.surfinfo.edvacc_surf);                               // nn-interpolation.cxx:537-545
        const int rc = des_remesh_transfer_cells(transfer_device(), &m, 1, int(sw.fields.size()), sw.fields.data(),
                                                 nullptr, &st);
        if (rc) transfer_die("des_remesh_transfer_cells (top surface)", rc);
        sw.commit();
        account(st);
    } else {
        m.old_size = var.volume->data();                                // still the old mesh's volumes
        m.new_ncell = var.nelem; m.new_conn = var.connectivity->data();
        Swap sw(m.new_ncell);
        sw.add(var.plstrain); sw.add(var.delta_plstrain); sw.add(var.strain);      // :549-556
        sw.add(var.radiogenic_source); sw.add(var.dyn_fric_coeff); sw.add(var.state_variable);
        sw.add(var.volume_old);                                                    // :560-569
        sw.add(var.stress); sw.add(var.stressyy);                                  // :584-588
        sw.add(var.spr_blend_weight); sw.add(var.spr_p_ref_old);                   // :593-603
        const int rc = des_remesh_transfer_cells(transfer_device(), &m, 0, int(sw.fields.size()), sw.fields.data(),
                                                 var.remesh_is_changed->data(), &st);
        if (rc) transfer_die("des_remesh_transfer_cells (elements)", rc);
        sw.commit();
        account(st);
    }
    if (verbose)
        std::printf("    [B200] %s transfer on the device: %d of %d changed, search %.3f s, ratios %.3f s, fields %.3f s\n",
                    is_surface ? "top-surface" : "element", st.nchanged, m.new_ncell, st.search_seconds,
                    st.acm_seconds, st.field_seconds);
}

extern "C" void DES_CAT(__wrap_, DES_BRC)(const Param &param, Variables &var,
                                          const Barycentric_transformation &bary,
                                          const array_t &old_coord, const conn_t &old_conn)
{
    if (use_reference()) { DES_CAT(__real_, DES_BRC)(param, var, bary, old_coord, old_conn); return; }
    DesTransferMeshes m;
    std::memset(&m, 0, sizeof(m));
    m.old_nnode = old_coord.size(); m.old_coord = old_coord.data();
    m.old_ncell = old_conn.size();  m.old_conn = old_conn.data();
    m.old_size = var.volume->data();
    m.new_nnode = var.nnode;        m.new_coord = var.coord->data();
    Swap sw(var.nnode);
    sw.add(var.temperature); sw.add(var.ppressure); sw.add(var.dppressure);        // brc-interpolation.cxx:357-364
    sw.add(var.vel); sw.add(var.coord0);                                           // :373-377
    sw.add(var.stress_n); sw.add(var.stressyy_n);                                  // :380-390
    DesTransferStats st;
    // var.support still holds the OLD mesh's lists here (remesh() rebuilds them after this call)
    const int rc = des_remesh_transfer_nodes(transfer_device(), &m, var.support.idx_data.data(),
                                             var.support.arr_data.data(), int(sw.fields.size()), sw.fields.data(),
                                             nullptr, nullptr, &st);
    if (rc) transfer_die("des_remesh_transfer_nodes", rc);
    sw.commit();
    account(st);
    if (verbose)
        std::printf("    [B200] node transfer on the device: %d of %d by the wider search (%d on a face, %d outside), "
                    "search %.3f s + %.3f s, fields %.3f s\n", st.nchanged, var.nnode, st.n_on_face, st.n_outside,
                    st.search_seconds, st.acm_seconds, st.field_seconds);
}

// barycentric_node_interpolation_forT (brc-interpolation.cxx:425-443; called by ic-read-temp.cxx:154
// to bring an external temperature field onto the mesh): the same search, one field, the input
// mesh's own support lists.  The volumes are the ones the caller gave its Barycentric_transformation
// (compute_volume on the input mesh, ic-read-temp.cxx:149-153): the reference's own function again.
extern "C" void DES_CAT(__wrap_, DES_BRCT)(const Param &param, const Variables &var,
                                           const Barycentric_transformation &bary,
                                           const array_t &input_coord, const conn_t &input_conn,
                                           const Support &input_support, const double_vec &input_temperature,
                                           double_vec &output_temperature)
{
    if (use_reference()) {
        DES_CAT(__real_, DES_BRCT)(param, var, bary, input_coord, input_conn, input_support, input_temperature,
                                   output_temperature);
        return;
    }
    double_vec volume(input_conn.size());
    compute_volume(input_coord, input_conn, volume);
    DesTransferMeshes m;
    std::memset(&m, 0, sizeof(m));
    m.old_nnode = input_coord.size(); m.old_coord = input_coord.data();
    m.old_ncell = input_conn.size();  m.old_conn = input_conn.data();
    m.old_size = volume.data();
    m.new_nnode = var.nnode;          m.new_coord = var.coord->data();
    DesTransferField f = { input_temperature.data(), output_temperature.data(), 1 };
    if (int(output_temperature.size()) < var.nnode) transfer_die("barycentric_node_interpolation_forT: output too short", 3);
    const int rc = des_remesh_transfer_nodes(transfer_device(), &m, input_support.idx_data.data(),
                                             input_support.arr_data.data(), 1, &f, nullptr, nullptr, nullptr);
    if (rc) transfer_die("des_remesh_transfer_nodes (external temperature)", rc);
}
