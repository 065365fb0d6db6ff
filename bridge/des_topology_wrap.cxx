// Link-time swap of the two topology builders that keep the reference's driver from starting a
// large mesh.  The bridge binary is linked with
//   -Wl,--wrap=_Z22create_boundary_facetsR9Variables -Wl,--wrap=_Z14create_supportR9Variables
// so the calls in init() (dynearthsol.cxx:169-170) and remesh() (remeshing.cxx:3022-3100) --
// the reference's own, unmodified object code -- arrive here instead of in mesh.cxx:
//   create_boundary_facets (mesh.cxx:3161-3283)  O(nseg * nelem) scan of every element for every
//        boundary segment  ->  des_gpu_build_support + des_gpu_find_boundary_facets (one thread
//        per segment over the support of its smallest node) + des_host_bucket_facets
//   create_support (mesh.cxx:3287-3329)  ->  des_gpu_build_support
// Same arrays bit for bit (tests/test_topology_gpu.py; the bridge frame tests run through it).
// DES_BRIDGE_TOPOLOGY=ref sends both calls on to the reference's own functions (__real_).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "constants.hpp"
#include "parameters.hpp"
#include "utils.hpp"
#include "../include/des_gpu.h"

#ifndef THREED
#error "3D only"
#endif

extern "C" void __real__Z22create_boundary_facetsR9Variables(Variables &var);
extern "C" void __real__Z14create_supportR9Variables(Variables &var);

namespace {
bool use_reference()
{
    const char *s = std::getenv("DES_BRIDGE_TOPOLOGY");
    return s && std::strcmp(s, "ref") == 0;
}
int topo_device()
{
    const char *s = std::getenv("DES_BRIDGE_DEVICE");
    return s ? std::atoi(s) : 0;
}
[[noreturn]] void topo_die(const char *what, int rc)
{
    std::fprintf(stderr, "des_bridge: %s failed (code %d); there is no CPU fallback\n", what, rc);
    std::exit(1);
}
void build_support(Variables &var)
{
    Support &s = var.support;
    s.idx_data.assign(size_t(var.nnode) + 1, 0);
    s.arr_data.resize(size_t(var.nelem) * NODES_PER_ELEM);
    s.lidx_data.resize(size_t(var.nelem) * NODES_PER_ELEM);
    const int rc = des_gpu_build_support(topo_device(), var.nnode, var.nelem, var.connectivity->data(),
                                         s.idx_data.data(), s.arr_data.data(), s.lidx_data.data());
    if (rc) topo_die("des_gpu_build_support", rc);
    s.rebind();
}
} // namespace

extern "C" void __wrap__Z14create_supportR9Variables(Variables &var)
{
    if (use_reference()) { __real__Z14create_supportR9Variables(var); return; }
    build_support(var);
}

extern "C" void __wrap__Z22create_boundary_facetsR9Variables(Variables &var)
{
    if (use_reference()) { __real__Z22create_boundary_facetsR9Variables(var); return; }
    // the search walks the support lists of the NEW mesh; var.support cannot be used or touched
    // here: init() builds it only after this call (dynearthsol.cxx:169-170), and inside remesh()
    // it still holds the OLD mesh's lists, which barycentric_node_interpolation reads after this
    // call (remeshing.cxx:3020-3037, brc-interpolation.cxx:357)
    std::vector<int> sidx(size_t(var.nnode) + 1), sarr(size_t(var.nelem) * NODES_PER_ELEM), slidx(sarr.size());
    {
        const int rc0 = des_gpu_build_support(topo_device(), var.nnode, var.nelem, var.connectivity->data(),
                                              sidx.data(), sarr.data(), slidx.data());
        if (rc0) topo_die("des_gpu_build_support", rc0);
    }
    const int nseg = var.nseg;
    std::vector<int> se(nseg), sf(nseg);
    int rc = des_gpu_find_boundary_facets(topo_device(), var.nnode, var.nelem, nseg, var.connectivity->data(),
                                          var.segment->data(), var.segflag->data(), var.bcflag->data(),
                                          sidx.data(), sarr.data(), se.data(), sf.data());
    if (rc) topo_die("des_gpu_find_boundary_facets", rc);
    for (int i = 0; i < nseg; ++i)
        if (se[i] == -2) {
            std::fprintf(stderr, "Error: %d-th segment is not on any element\n", i);     // mesh.cxx:3218
            std::exit(1);
        }
    int nb[DES_NBDRY], off[DES_NBDRY + 1];
    std::vector<int> pairs(size_t(nseg) * 2 + 2);
    // first pass for the counts (conn_surface needs etop), second with the surface connectivity
    rc = des_host_bucket_facets(var.nelem, nseg, var.segflag->data(), se.data(), sf.data(),
                                var.connectivity->data(), nb, off, pairs.data(), nullptr);
    if (rc) topo_die("des_host_bucket_facets", rc);
    const int etop = nb[iboundz1];
    var.connectivity_surface = new conn_t(etop);
    rc = des_host_bucket_facets(var.nelem, nseg, var.segflag->data(), se.data(), sf.data(),
                                var.connectivity->data(), nb, off, pairs.data(), var.connectivity_surface->data());
    if (rc) topo_die("des_host_bucket_facets", rc);
    for (int k = 0; k < nbdrytypes; ++k) {
        int_pair_vec &v = *var.bfacets[k];
        for (int i = 0; i < nb[k]; ++i)
            v.emplace_back(pairs[2 * (off[k] + i)], pairs[2 * (off[k] + i) + 1]);
    }
}
