/* des_gpu.h -- C-ABI of the B200-native explicit time-step core for DynEarthSol (3D).
 *
 * This is the drop-in boundary for the reference's step loop (dynearthsol.cxx:768-894):
 * plain pointers and sizes, no C++/torch types.  The reference has no plugin ABI of its
 * own (its boundary is free C++ functions taking `const Param&` / `Variables&`,
 * fields.hpp / rheology.hpp / geometry.hpp / bc.hpp), so each entry point below names the
 * reference call it replaces.  All host arrays use the reference's in-memory layout:
 * Array2D<T,N> is SoA, element [i][d] at a[d*n + i] (array2d.hpp:410-425, built -DSOA);
 * double_vec / int_vec are plain contiguous arrays.  Indices are int32, fields are FP64.
 *
 * Ownership: host pointers are BORROWED for the duration of one call and never cached
 * (remesh() re-news nearly every array, fields.cxx:125-194).  The handle owns all device
 * memory.  Errors: every call returns 0 or a DES_E_* code; the library never exit()s.
 * Threading: call from one host thread (the driver is single-threaded at call level).
 *
 * There is NO CPU fallback: without a CUDA device des_gpu_create() fails with DES_E_NODEVICE.
 */
#ifndef DES_GPU_H
#define DES_GPU_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DES_MAX_MAT 16      /* materials per model (reference: mat.num_materials) */
#define DES_NBDRY 10        /* constants.hpp:27-37: x0 x1 y0 y1 z0 z1 n0 n1 n2 n3 */

/* status codes; the numeric values of the runtime ones follow utils.hpp:20-44 in spirit */
enum {
    DES_OK = 0,
    DES_E_NODEVICE = 1,     /* no CUDA device / CUDA runtime failure at create */
    DES_E_CUDA = 2,         /* a CUDA call failed; see des_gpu_last_error() */
    DES_E_ARG = 3,          /* bad argument (null pointer, size mismatch) */
    DES_E_UNSUPPORTED = 4,  /* option outside the hot-path scope */
    DES_E_STATE = 5,        /* call out of order (e.g. step before upload) */
    DES_E_NAN = 11          /* runtime NaN / non-positive dt (EXIT_RUNTIME_NAN analogue) */
};

/* rheology bit flags, matprops.hpp:86-95 */
enum {
    DES_RH_ELASTIC = 1, DES_RH_VISCOUS = 2, DES_RH_MAXWELL = 3,
    DES_RH_EP = 5, DES_RH_EVP = 7
};

typedef struct DesGpu DesGpu;   /* opaque */

/* The scalars of Param.{control,bc,mat,mesh} and Variables the kernels read
 * (parameters.hpp:160-560; input.cxx defaults).  Filled by the bridge from `Param`. */
typedef struct DesParam {
    /* control */
    double gravity;                 /* control.gravity */
    double inertial_scaling;        /* control.inertial_scaling */
    double damping_factor;          /* control.damping_factor */
    double surface_diffusivity;     /* var.surfinfo.surf_diff (= control.surface_diffusivity) */
    double surf_base_level;         /* control.surf_base_level */
    double dt_fraction;             /* control.dt_fraction */
    double fixed_dt;                /* control.fixed_dt (0 = dynamic) */
    double characteristic_speed;    /* control.characteristic_speed (0 = use max_vbc_val) */
    int damping_option;             /* control.damping_option 0..4 (4 = Rayleigh), fields.cxx:490-575 */
    int is_quasi_static;            /* control.is_quasi_static */
    int has_thermal_diffusion;      /* control.has_thermal_diffusion */
    int is_using_mixed_stress;      /* control.is_using_mixed_stress -> NMD_stress */
    int surface_process_option;     /* 0 or 1 (simple_diffusion) */
    int ref_pressure_option;        /* 0, 1 (PREM), 2 (modified PREM); matprops.cxx:153-174 */
    int has_moving_mesh;            /* control.has_moving_mesh */
    int has_body_force_adjustment;  /* ic.has_body_force_adjustment (gates neumann bcs) */
    /* bc */
    int has_winkler_foundation, has_water_loading, has_elastic_foundation;
    double winkler_delta_rho, elastic_foundation_constant, sea_water_density;
    double surface_temperature;
    int vbc_types[DES_NBDRY];       /* var.vbc_types, dynearthsol.cxx:62-71 */
    double vbc_values[DES_NBDRY];   /* var.vbc_values */
    double vbc_val_lateral[4];      /* bc.vbc_val_{x0,x1,y0,y1}_l (type 6) */
    double vbc_val_z1_loading_period;
    int stress_bc_types[6];         /* var.stress_bc_types (dynearthsol.cxx:110-115): 0 none, 1 / 2 / 3 = traction along
                                     * x / y / z on that boundary -- apply_stress_bcs_neumann, bc.cxx:829-912 */
    double max_vbc_val;             /* var.max_vbc_val (find_max_vbc, bc.cxx:66-91) */
    /* mesh */
    double zlength;                 /* mesh.zlength (Winkler datum) */
    /* mat */
    int rheol_type;                 /* DES_RH_* */
    int nmat;
    int mattype_ref;
    double visc_min, visc_max, tension_max, therm_diff_max;
    double rho0[DES_MAX_MAT], alpha[DES_MAX_MAT];
    double bulk_modulus[DES_MAX_MAT], shear_modulus[DES_MAX_MAT];
    double visc_exponent[DES_MAX_MAT], visc_coefficient[DES_MAX_MAT];
    double visc_activation_energy[DES_MAX_MAT], visc_activation_volume[DES_MAX_MAT];
    double heat_capacity[DES_MAX_MAT], therm_cond[DES_MAX_MAT];
    double pls0[DES_MAX_MAT], pls1[DES_MAX_MAT];
    double cohesion0[DES_MAX_MAT], cohesion1[DES_MAX_MAT];
    double friction_angle0[DES_MAX_MAT], friction_angle1[DES_MAX_MAT];
    double dilation_angle0[DES_MAX_MAT], dilation_angle1[DES_MAX_MAT];
    double porosity[DES_MAX_MAT];
    double stress_bc_values[6];     /* var.stress_bc_values (dynearthsol.cxx:117-122) */
} DesParam;

/* Topology produced by the reference's host mesh code (mesh.cxx), immutable between
 * remeshes.  Replaces nothing: it is what the kernels gather through. */
typedef struct DesMesh {
    int nnode, nelem;
    const int *connectivity;        /* conn_t, [4*nelem] SoA */
    const unsigned *bcflag;         /* [nnode] */
    const int *support_idx;         /* Support.idx_data [nnode+1]  (mesh.cxx:3287-3329) */
    const int *support_arr;         /* Support.arr_data [4*nelem], ascending element ids */
    const int *support_lidx;        /* Support.lidx_data [4*nelem] */
    int nbfacets[DES_NBDRY];
    const int *bfacets[DES_NBDRY];  /* (elem, facet) pairs, [2*nbfacets[i]] */
    int nbnodes[DES_NBDRY];
    const int *bnodes[DES_NBDRY];
    const double *bnormals;         /* array_t(10): [3*10] SoA stride 10 (bc.cxx:94-224) */
    int n_edge_vec;                 /* number of doubles in edge_vec */
    const double *edge_vec;
    const int *edge_slot;           /* [100], -1 = no shared edge */
    int ntop_elems;
    const int *top_elems;           /* var.top_elems */
    int ntop;                       /* surfinfo.ntop */
    const int *top_nodes;           /* surfinfo.top_nodes [ntop] */
    int etop;                       /* surfinfo.etop == nbfacets[z1] */
    const int *support_surf_idx;    /* surfinfo.support_surf.idx_data [ntop+1] */
    const int *support_surf_arr;    /* surfinfo.support_surf.arr_data */
    const int *elem_and_nodes;      /* segment_t [3*etop] SoA: top-node LOCAL index per facet node */
    const int *connectivity_surface;/* conn_t(etop): [4*etop] SoA, first 3 columns used */
    const double *coord;            /* var.coord [3*nnode] SoA: read only by a handle that spans several
                                     * devices (recursive coordinate bisection of the element centroids) */
} DesMesh;

/* Field arrays.  NULL members are skipped, so the same struct serves as an upload list,
 * a download list and a download mask.  Classification: SURVEY.md section 8b. */
typedef struct DesState {
    double *coord, *vel, *coord0;                   /* [3*nnode] SoA */
    double *temperature;                            /* [nnode] */
    double *stress, *strain, *strain_rate;          /* [6*nelem] SoA {XX,YY,ZZ,XY,XZ,YZ} */
    double *plstrain, *delta_plstrain;              /* [nelem] */
    double *volume, *volume_old;                    /* [nelem] */
    double *viscosity, *dpressure, *edvoldt;        /* [nelem] */
    double *radiogenic_source;                      /* [nelem] */
    double *force;                                  /* [3*nnode] (derived) */
    double *volume_n, *mass, *tmass;                /* [nnode] (derived) */
    double *ntmp;                                   /* [nnode] scratch (debug/parity) */
    double *dhacc;                                  /* surfinfo.dhacc [nnode] */
    double *edvacc_surf;                            /* surfinfo.edvacc_surf [etop] */
} DesState;

/* What compute_dt()'s element loop reduces (geometry.cxx:1506-1593) plus the surface
 * velocity it reads; the scalar algebra after the loop stays on the host
 * (des_dt_from_reductions). */
typedef struct DesDtReductions {
    double minl, dt_maxwell, dt_diffusion, global_dt_min, global_max_vem, max_surf_vel;
} DesDtReductions;

/* ---- lifetime ------------------------------------------------------------------------ */
/* after init()/restart() (dynearthsol.cxx:634-650); replaces init_offload_device() */
int des_gpu_create(const DesParam *param, int device, DesGpu **out);
/* The same handle over `ndev` devices of one box (devices == NULL: 0..ndev-1; a device may be
 * listed more than once).  des_gpu_upload_mesh then partitions the mesh (RCB of the element
 * centroids, node owner = lowest part, one redundant element layer), des_gpu_step exchanges the
 * ghost-node values between the parts after each nodal barrier (peer copies, no reduction), and
 * des_gpu_upload_state / _download_state / _upload_elemmarkers scatter and gather the host's
 * global arrays, so the driver's call sites do not change and results are bit-identical to one
 * device.  After every remesh the next des_gpu_upload_mesh repartitions; des_gpu_partition_info
 * reports what that cost.  Not available on a group handle: the fine-grained twins, segment parts,
 * des_gpu_check_nan, des_gpu_device_ptr, kernel timing (DES_E_UNSUPPORTED). */
int des_gpu_create_multi(const DesParam *param, int ndev, const int *devices, DesGpu **out);
/* the last des_gpu_upload_mesh: number of parts, host seconds spent partitioning (RCB, owners,
 * local meshes, exchange plans) and uploading; per part (arrays of nparts ints, may be NULL) the
 * local element / node counts including the redundant layer, and the ghost nodes it receives */
int des_gpu_partition_info(DesGpu *h, int *nparts, double *partition_seconds, double *upload_seconds,
                           int *local_elems, int *local_nodes, int *ghost_nodes);
/* Internal storage order, effective at the next des_gpu_upload_mesh: 0 = the caller's (default),
 * 1 = elements by smallest node index, 2 = nodes in the Morton order of their coordinates and
 * elements in the Morton order of their centroids (needs DesMesh.coord).  The reference
 * renumbers along the longest axis (mesh.cxx:2696-2821); on unstructured meshes the Morton
 * order is ~11 % faster.  Every array still crosses this ABI in the caller's order and each
 * node's support keeps the caller's element order, so results do not change by a bit.  The
 * environment variable DES_REORDER=morton selects 2 at create.  des_gpu_device_ptr then
 * returns arrays in the internal order. */
int des_gpu_set_reorder(DesGpu *h, int mode);
/* the same partitioner without a device (pure host integer work): RCB part of every element
 * [nelem], owner of every node [nnode], and per part the local element / node / ghost-node counts
 * des_gpu_upload_mesh would build ([nparts] each; any output may be NULL) */
int des_host_partition(int nnode, int nelem, const int *connectivity, const double *coord, int nparts,
                       int *elem_part, int *node_owner, int *local_elems, int *local_nodes, int *ghost_nodes);
/* end() (dynearthsol.cxx:440-445) */
int des_gpu_destroy(DesGpu *h);
const char *des_gpu_last_error(const DesGpu *h);

/* ---- data movement ------------------------------------------------------------------- */
/* after init, restart and every remesh() (dynearthsol.cxx:950): new topology; drops all
 * device fields, so an upload_state must follow. */
int des_gpu_upload_mesh(DesGpu *h, const DesMesh *mesh);
/* H->D of the members that are non-NULL */
int des_gpu_upload_state(DesGpu *h, const DesState *s);
/* D->H into the reference's own Variables storage, before output (dynearthsol.cxx:926),
 * phase_changes (:885), bad_mesh_quality (:942), remesh (:950), average_fields (:903).
 * Completes any deferred end-of-step work first. */
int des_gpu_download_state(DesGpu *h, const DesState *s);
/* the four refresh_elem_cache() sites (dynearthsol.cxx:206,479,776,887): flat
 * [nelem*nmat] copy of elemmarkers; replaces MatProps::refresh_elem_cache (matprops.cxx:259) */
int des_gpu_upload_elemmarkers(DesGpu *h, const int *elemmarkers);
/* scalars of Variables that change while running */
int des_gpu_set_compensation_pressure(DesGpu *h, double p);

/* ---- the step ------------------------------------------------------------------------ */
/* One explicit step == dynearthsol.cxx:778-876 for a run without PT loop / hydraulics:
 * update_temperature, update_strain_rate, compute_dvoldt, compute_edvoldt, update_stress,
 * NMD_stress, update_force (+apply_stress_bcs, apply_damping), update_velocity, apply_vbcs,
 * update_mesh (update_coordinate, surface_processes, volume swap, compute_volume,
 * compute_mass), rotate_stress.  `time` and `steps` are var.time / var.steps AFTER the
 * increment at dynearthsol.cxx:773-774.  Asynchronous; the end-of-step element work
 * (volume, mass, rotation) may be deferred and fused with the next step's first kernel. */
int des_gpu_step(DesGpu *h, double dt, double time, int steps);
/* The same step in six segments, for hosts that exchange ghost-node values between the
 * nodal barriers of a step (multi-GPU, see des_gpu_set_halo): 0 = temperature/strain-rate
 * trace and nodal gathers up to compute_dvoldt (fields.cxx:197-262, geometry.cxx:203-246);
 * 1 = compute_edvoldt, update_stress and NMD_stress's nodal loop (rheology.cxx:703,
 * geometry.cxx:282-309); 2 = NMD correction, update_force, update_velocity, apply_vbcs,
 * update_coordinate (fields.cxx:609-784); 3 = simple_diffusion (bc.cxx:916-1112), which
 * reads the neighbours' new coordinates; 4 = edvacc_surf and correct_surface_element
 * (bc.cxx:1655-1707,1787-1797) -- on quality-check steps the host's surface-marker routines
 * (bc.cxx:1829-1843) run after it on downloaded coord/volume/dhacc/edvacc_surf; 5 = the
 * volume swap (dynearthsol.cxx:466-470), no kernel.
 * des_gpu_step(h, dt, time, steps) == segments 0..5 in order. */
int des_gpu_step_segment(DesGpu *h, int segment, double dt, double time);
/* compute_dt (dynearthsol.cxx:643,893; geometry.cxx:1480): device reductions only.  This is the
 * step loop's sync point (des_gpu_step is asynchronous): it returns DES_E_NAN when a step since
 * the last upload computed a non-positive or NaN element volume (an inverted or degenerate
 * tetrahedron: the host then writes its error frame -- downloads still work -- and dies).
 * The element loop rides on the next step's first kernel, so the call costs no extra pass. */
int des_gpu_compute_dt(DesGpu *h, DesDtReductions *out);
/* No counterpart in the reference: added for callers that
 * may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The same reductions as des_gpu_compute_dt -- same inputs (the handle's resident
 * mesh and state), same outputs, bit-identical values -- but the per-block partial results live
 * in a file-scope static __device__ array of the library (655 400 bytes on every GPU that loads
 * it, held until the process exits) and the call allocates nothing: no cudaMalloc, cudaFree or
 * allocating library routine.  It runs as its own pass over the elements (it does not ride on the
 * next step's first kernel) and leaves the handle's step schedule untouched.
 * Limit: meshes of at most DES_STATIC_DT_MAX_ELEMS elements; larger ones are refused with
 * DES_E_ARG and a message naming the limit.  Single-device handles only (DES_E_UNSUPPORTED for a
 * group).  Every call on a device shares the one array: calls that overlap on the same GPU
 * (two handles on different streams, two host threads) race -- serialise them. */
#define DES_STATIC_DT_MAX_ELEMS 16777216
int des_gpu_compute_dt_static(DesGpu *h, DesDtReductions *out);
/* host scalar algebra of geometry.cxx:1597-1630 (no GPU work); returns dt, <=0 on error */
double des_dt_from_reductions(const DesParam *param, const DesDtReductions *r,
                              double *max_global_vel_mag);
int des_gpu_sync(DesGpu *h);

/* ---- host touch points of the step loop kept on the device -- */
/* bad_mesh_quality (dynearthsol.cxx:942; remeshing.cxx:2752-2866): the members of Param.mesh
 * it reads ... */
typedef struct DesQualityParam {
    double smallest_size, resolution, max_boundary_distortion, min_quality;
    double xlength, ylength, zlength;
    int remeshing_option;
} DesQualityParam;
/* ... and what its loops find.  Ids are the caller's; -1 = none.  `value` members are the
 * coordinate the reference prints for the offending node. */
typedef struct DesQuality {
    int first_small_elem;       /* lowest e with volume[e] < smallest_vol (return 3) */
    int first_bottom_node;      /* lowest bottom node too far from -zlength (return 2) */
    int first_side_node;        /* remeshing_option 13: lowest side node off its plane (return 2) */
    int worst_elem;             /* worst_elem_quality's argmin (first minimum), geometry.cxx:1914 */
    double worst_quality;       /* its elem_quality, BEFORE the cube root of remeshing.cxx:2845 */
    double bottom_value, side_value;
} DesQuality;
/* device reductions only; no field leaves the GPU */
int des_gpu_mesh_quality(DesGpu *h, const DesQualityParam *qp, DesQuality *out);
/* the decision ladder of bad_mesh_quality on those reductions (host libm for the cube root):
 * returns 0..3 like the reference, sets index and min_quality the same way (min_quality is
 * only written when the ladder reaches the distortion check) */
int des_quality_verdict(const DesQualityParam *qp, const DesQuality *q, int *index, double *min_quality);

/* check_nan (utils.hpp:323-385, called by Output::write_exact_error): the number of NaN entries
 * in volume, dpressure, viscosity, stress, temperature, tmass, force, vel and coord */
int des_gpu_check_nan(DesGpu *h, int *count);

/* Output::average_fields (dynearthsol.cxx:903; output.cxx:312-352) with device-resident
 * accumulators.  first != 0 (steps % average_interval == 1): snapshot coord, strain, stress,
 * delta_plstrain; otherwise stress_avg += stress, delta_plstrain_avg += delta_plstrain. */
int des_gpu_average_fields(DesGpu *h, int first);
/* D->H of the accumulators in the reference's layouts ([3*nnode], [6*nelem], [6*nelem],
 * [nelem]); NULL members are skipped.  DES_E_STATE if no snapshot was taken since the mesh
 * was uploaded. */
int des_gpu_download_averages(DesGpu *h, double *coord0, double *strain0, double *stress_avg,
                              double *delta_plstrain_avg);

/* ---- phase_changes on device-resident markers -------------------
 * The arrays of the material marker set, markersets[0] (markerset.hpp:106-110): element, the
 * four shape-function weights (shapefn SoA, component i of marker m at eta[i*stride + m]) and
 * rock type of every marker.  Upload after init / restart / remesh and whenever host code has
 * moved, added or removed markers (surface processes, replenishment); between those points the
 * device keeps them, and des_gpu_phase_changes keeps the rock types current.  Single-device
 * handles (a group handle answers DES_E_UNSUPPORTED: the host path stays). */
int des_gpu_upload_markers(DesGpu *h, int nmarkers, int stride, const int *elem, const double *eta,
                           const int *mattype);
/* phase_changes (dynearthsol.cxx:885; phasechanges.cxx:109-158) with MarkerSet::get_ZPT
 * (markerset.cxx:973-986) and ref_pressure (matprops.cxx:153-174) read from the resident
 * coordinates and temperature: option 1 = simple_subduction (phasechanges.cxx:10-84), 101 = the
 * custom template (which changes nothing); anything else DES_E_UNSUPPORTED (the reference's
 * EXIT_CONFIG_VALUE).  hydrous_count [nelem] = (*var.hydrous_elemmarkers)[e][0], NULL = all zero.
 * The device applies the changes to its rock types and elemmarkers counters (no
 * des_gpu_upload_elemmarkers needed afterwards) and returns the markers that changed in ascending
 * marker order: *nchanged of them, the first min(*nchanged, cap) in changed_marker /
 * changed_mattype (the host then does set_mattype and moves its own elemmarkers counts,
 * phasechanges.cxx:143-147).  hyd_inc [nmarkers] (may be NULL) receives what the reference stores
 * with set_tmp (:131) for its hydration block (:160-176), which stays host code. */
int des_gpu_phase_changes(DesGpu *h, int option, const int *hydrous_count, int cap, int *nchanged,
                          int *changed_marker, int *changed_mattype, int *hyd_inc);
/* the device's elemmarkers counters, [nelem][nmat] in the caller's element order */
int des_gpu_download_elemmarkers(DesGpu *h, int *elemmarkers);

/* ---- one output frame straight from the device ----------------
 * The arrays Output::_write (output.cxx:95-245) hands to BinaryOutput, already in the layout
 * BinaryOutput::write_array puts on disk (Array2D::pack_to, binaryio.cxx:151-166: [n][components])
 * and with _write's own host loops applied on the device: the averaged variants
 * (output.cxx:131-190), density = MatProps::rho(e) (:192-196), mesh quality = elem_quality(e)
 * (:198-202), material = the most abundant marker type (:214-220).  The caller fwrites them.
 * Every pointer is a HOST buffer (des_host_alloc gives pinned ones: the copies then run at PCIe
 * speed); NULL = not wanted.  Values are bit-identical to what the reference writes. */
typedef struct DesFrame {
    int averaged;               /* !disable_averaging && is_averaged, output.cxx:103 */
    double inv_dt;              /* 1 / (time - time0), output.cxx:105 (averaged only) */
    double stress_avg_scale;    /* 1 / (average_interval + 1), output.cxx:185 (averaged only) */
    double *coordinate;         /* [nnode][3] */
    double *velocity;           /* [nnode][3] */
    double *velocity_averaged;  /* [nnode][3]  (coord - coord0) * inv_dt; averaged only */
    double *temperature;        /* [nnode] */
    double *plstrain;           /* [nelem] */
    double *plstrain_rate;      /* [nelem]     delta_plstrain, or its running sum * inv_dt */
    double *strain_rate;        /* [nelem][6]  strain_rate, or (strain - strain0) * inv_dt */
    double *strain;             /* [nelem][6] */
    double *stress;             /* [nelem][6] */
    double *viscosity;          /* [nelem] */
    double *stress_averaged;    /* [nelem][6]  running sum * stress_avg_scale; averaged only */
    double *density;            /* [nelem] */
    double *mesh_quality;       /* [nelem] */
    double *material;           /* [nelem]     (a double, as the reference writes it) */
    double *force;              /* [nnode][3] */
} DesFrame;
/* Single-device handles (a group handle answers DES_E_UNSUPPORTED: use des_gpu_download_state).
 * averaged = 1 needs des_gpu_average_fields since the interval began (DES_E_STATE otherwise). */
int des_gpu_download_frame(DesGpu *h, const DesFrame *frame);
/* page-locked host memory for callers without the CUDA runtime (cudaHostAlloc / cudaFreeHost) */
int des_host_alloc(size_t bytes, void **out);
int des_host_free(void *p);

/* ---- fine-grained twins of the reference calls (PT loop dynearthsol.cxx:803-864,
 * isostasy_adjustment :506-539, remeshing.cxx:3133-3168; also the per-function parity
 * tests).  Same device state; each is the un-fused reference loop nest. --------------- */
int des_gpu_update_temperature(DesGpu *h, double dt);               /* fields.cxx:197 */
int des_gpu_update_strain_rate(DesGpu *h);                          /* fields.cxx:405 */
int des_gpu_compute_dvoldt(DesGpu *h);                              /* geometry.cxx:203 */
int des_gpu_compute_edvoldt(DesGpu *h);                             /* geometry.cxx:249 */
int des_gpu_update_stress(DesGpu *h, double dt);                    /* rheology.cxx:703 */
int des_gpu_nmd_stress(DesGpu *h);                                  /* geometry.cxx:282 */
int des_gpu_update_force(DesGpu *h);                                /* fields.cxx:609 */
int des_gpu_update_velocity(DesGpu *h, double dt);                  /* fields.cxx:725 */
int des_gpu_apply_vbcs(DesGpu *h, double time);                     /* bc.cxx:227 */
int des_gpu_update_mesh(DesGpu *h, double dt);                      /* dynearthsol.cxx:448 */
/* the same in two parts: 0 = update_coordinate + surface_processes, 1 = volume swap, compute_volume,
 * compute_mass; on quality-check steps the host's surface-marker routines run between them */
int des_gpu_update_mesh_part(DesGpu *h, double dt, int part);
/* param.control.PT_jump of the pseudo-transient loops (dynearthsol.cxx:561,811): while set,
 * des_gpu_apply_vbcs holds the six face values at zero (bc.cxx:330-344) and des_gpu_update_mesh
 * runs without surface_processes (dynearthsol.cxx:456) */
int des_gpu_set_pt_jump(DesGpu *h, int on);
/* param.ic.has_body_force_adjustment (dynearthsol.cxx:753-761): while set, update_force skips the
 * prescribed tractions (fields.cxx:690) */
int des_gpu_set_body_force_adjustment(DesGpu *h, int on);
int des_gpu_compute_volume(DesGpu *h);                              /* geometry.cxx:170 */
int des_gpu_compute_mass(DesGpu *h);                                /* geometry.cxx:1743 */
int des_gpu_rotate_stress(DesGpu *h, double dt);                    /* fields.cxx:827 */
int des_gpu_residual_force(DesGpu *h, double *l2);                  /* fields.cxx:700 */
/* the velocity constraint inside isostasy_adjustment's loop (dynearthsol.cxx:522-533): vx = vy = 0,
 * and vz = 0 on the bottom when there is no Winkler foundation */
int des_gpu_isostasy_constrain_velocity(DesGpu *h);

/* ---- multi-GPU: one handle per device, elements partitioned, ghost nodes exchanged --------
 * The reference has no distributed mode; this is the boundary a host (one process per GPU
 * over torch.distributed/NCCL in dynearthsol_b200/multigpu.py, or one thread per device in
 * a C++ driver) uses.  Each handle holds a sub-mesh: the elements touching the nodes the rank
 * owns, plus the other ("ghost") nodes of those elements.  Owned nodes see their complete
 * support in the reference's ascending element order, so results are bit-identical to one
 * GPU; ghost nodes receive their values from the owner after each nodal barrier.
 *   des_gpu_set_stream   own == 0: run this handle's kernels on the caller's CUDA stream
 *                        (cudaStream_t; 0 is the legacy default stream), so the caller's
 *                        collectives order with them; own != 0: back to a private stream
 *   des_gpu_set_halo     local indices of owned nodes other ranks need (send) and of ghost
 *                        nodes (recv), each concatenated over neighbours in message order
 *   des_gpu_halo_buffers device pointers of the packed send / receive buffers
 *   des_gpu_halo_width   doubles per entry in phase 0 {T, dvoldt}, 1 {dp_nd},
 *                        2 {x, y, z, vx, vy, vz}, 3 {z, dh}
 *   des_gpu_halo_pack / _unpack   gather into / scatter from the buffers, after segment
 *                        `phase` of des_gpu_step_segment */
int des_gpu_set_stream(DesGpu *h, void *cuda_stream, int own);
/* Overlap of a ghost exchange with the element kernel that follows it.  The host orders its
 * local elements so that the first n_interior touch no ghost node (des_gpu_set_interior) and
 * runs segments 0..2 in parts: 0 = element kernel on those elements, on `cuda_stream` (NULL =
 * the handle's stream) -- it needs nothing the exchange delivers, so it runs while the messages
 * are in flight; 1 = element kernel on the remaining elements, after the unpack; 2 = the rest
 * of the segment (its node gather), after both.  Ordering between streams is the host's job
 * (events).  Parts 0,1,2 in order == des_gpu_step_segment(seg). */
int des_gpu_set_interior(DesGpu *h, int n_interior);
int des_gpu_step_segment_part(DesGpu *h, int seg, int part, double dt, double time, void *cuda_stream);
int des_gpu_set_halo(DesGpu *h, int nsend, const int *send_nodes, int nrecv, const int *recv_nodes);
int des_gpu_halo_buffers(DesGpu *h, void **send_buf, void **recv_buf);
int des_gpu_halo_width(int phase);
int des_gpu_halo_pack(DesGpu *h, int phase);
int des_gpu_halo_unpack(DesGpu *h, int phase);

/* ---- host-side topology helpers (pure integer work, no GPU) ---------------------------- */
/* create_support (mesh.cxx:3287-3329): node -> (element, local node) CSR with ascending
 * element ids per node.  connectivity is conn_t SoA [4*nelem]; idx [nnode+1], arr and lidx
 * [4*nelem] are caller-allocated.  Bit-identical to the reference's arrays. */
int des_host_build_support(int nnode, int nelem, const int *connectivity,
                           int *idx, int *arr, int *lidx);
/* ---- topology builders: the integer tables of DesMesh from
 * connectivity + boundary segments, bit-identical to mesh.cxx:2824-3329, without the
 * reference's O(nseg*nelem) facet search.  Arrays are the reference's layouts (conn_t /
 * segment_t SoA); outputs are caller-allocated.  The two O(nelem) builders run on the GPU,
 * the O(nnode) / O(surface) ones on the host with the reference's own std::sort calls. ---- */
/* create_support (mesh.cxx:3287-3329) on device `device`; outputs as des_host_build_support */
int des_gpu_build_support(int device, int nnode, int nelem, const int *connectivity,
                          int *idx, int *arr, int *lidx);
/* the search loop of create_boundary_facets (mesh.cxx:3176-3225), through the support of each
 * segment's smallest node: seg_elem[i] = element (-1: not a boundary segment, -2: on no
 * element), seg_facet[i] = its facet 0..3 */
int des_gpu_find_boundary_facets(int device, int nnode, int nelem, int nseg, const int *connectivity,
                                 const int *segment, const int *segflag, const unsigned *bcflag,
                                 const int *support_idx, const int *support_arr,
                                 int *seg_elem, int *seg_facet);
/* create_boundary_flags2 (mesh.cxx:2824-2835); bcflag zeroed by the caller */
int des_host_boundary_flags(int nnode, int nseg, const int *segment, const int *segflag, unsigned *bcflag);
/* create_boundary_nodes (mesh.cxx:2854-2880); bnodes == NULL: counts only */
int des_host_boundary_nodes(int nnode, const unsigned *bcflag, int nbnodes[DES_NBDRY],
                            int *const bnodes[DES_NBDRY]);
/* bucketing + per-boundary sort + connectivity_surface of create_boundary_facets
 * (mesh.cxx:3205-3262): pairs [2*nseg], bucket k at pairs + 2*offset[k] */
int des_host_bucket_facets(int nelem, int nseg, const int *segflag, const int *seg_elem, const int *seg_facet,
                           const int *connectivity, int nbfacets[DES_NBDRY], int offset[DES_NBDRY + 1],
                           int *pairs, int *conn_surface);
/* create_top_elems (mesh.cxx:2882-2931), create_surface_info / create_support_surf integer
 * tables (mesh.cxx:3031-3098, 2937-2964) */
int des_host_surface_topology(int nnode, int ntop, const int *bnodes_z1, const double *coord_x,
                              const int *support_idx, const int *support_arr,
                              int etop, const int *conn_surface,
                              int *top_nodes, int max_top_elems, int *ntop_elems, int *top_elems,
                              int *elem_and_nodes, int *ssup_idx, int *ssup_arr);
/* ---- remesh field transfer: what remesh() asks of
 * nearest_neighbor_interpolation and barycentric_node_interpolation (remeshing.cxx:3058-3069),
 * on device `device`, between an OLD and a NEW mesh.  No handle: remesh() runs between a full
 * download and a fresh upload.  Every pointer is a HOST array in the reference's own layout
 * (Array2D SoA: component d of item i at [d*n + i]; conn_t has 4 components, a surface facet
 * uses the first 3).  The k-nearest-neighbour searches are exact, ascending by (distance^2,
 * index): nanoflann's result order with equal distances resolved by the lower index (nanoflann
 * resolves them by its tree traversal order); every arithmetic statement keeps the reference's
 * operation order, so integer outputs and transferred fields are bit-identical. ------------- */
typedef struct DesTransferMeshes {
    int old_nnode;              /* old_coord.size() */
    const double *old_coord;    /* [3][old_nnode] */
    int old_ncell;              /* old elements, or old top-surface facets */
    const int *old_conn;        /* [4][old_ncell]  old_connectivity / old_connectivity_surface */
    const double *old_size;     /* [old_ncell]     the volumes (areas) given to Barycentric_transformation */
    int new_nnode;
    const double *new_coord;    /* [3][new_nnode]  *var.coord */
    int new_ncell;              /* new elements / facets (unused by the node transfer) */
    const int *new_conn;        /* [4][new_ncell]  (unused by the node transfer) */
} DesTransferMeshes;
typedef struct DesTransferField {
    const double *src;          /* [ncomp][old items] */
    double *dst;                /* [ncomp][new items], caller-allocated */
    int ncomp;                  /* 1 (double_vec), 3 (array_t) or 6 (tensor_t) */
} DesTransferField;
typedef struct DesTransferStats {
    int nchanged;               /* cells: entries with is_changed > 0 after the nearest-neighbour
                                 * pass; nodes: new nodes the one-ring search did not place */
    int nsample;                /* cells: sample points per changed cell */
    int n_on_face, n_outside;   /* nodes: taken as lying on a face / outside the old domain */
    double search_seconds, acm_seconds, field_seconds;   /* wall clock, copies included */
} DesTransferStats;
/* nearest_neighbor_interpolation (nn-interpolation.cxx:661-702): nearest old centroid of every
 * new cell (find_nearest_neighbor, :16-75), volume-ratio mapping of the changed ones through
 * sample points located in the 32 nearest old cells (find_acm_elem_ratios, :77-318), then
 * inject_field (:433-523) for every field.  is_surface = 0: elements (old_size = volume);
 * 1: the facets of the top surface (old_size = area; centroids compared in x, y).
 * is_changed (may be NULL) receives var.remesh_is_changed: 0 same centroid, 1 remapped by
 * ratios, -1 changed but kept on its nearest neighbour. */
int des_remesh_transfer_cells(int device, const DesTransferMeshes *m, int is_surface, int nfields,
                              const DesTransferField *fields, int *is_changed, DesTransferStats *stats);
/* barycentric_node_interpolation (brc-interpolation.cxx:345-423): the old element around every
 * new node -- nearest old node, its one ring, then the capped breadth-first search and the two
 * fallbacks of prepare_interpolation (:98-341) -- and interpolate_field (:18-96) for every field.
 * old_support_idx / _arr: create_support's lists of the OLD mesh.  el [new_nnode] and
 * brc [new_nnode][4] (may be NULL) receive the enclosing element and the four weights.
 * DES_E_UNSUPPORTED: a new node's search met more than 16 384 old elements within three rings
 * of its nearest old node (per-node scratch of the search). */
int des_remesh_transfer_nodes(int device, const DesTransferMeshes *m, const int *old_support_idx,
                              const int *old_support_arr, int nfields, const DesTransferField *fields,
                              int *el, double *brc, DesTransferStats *stats);

/* sizeof() of the ABI structs as compiled, so a foreign-language binding can verify its
 * own layout: which = 0 DesParam, 1 DesMesh, 2 DesState, 3 DesDtReductions, 4 DesQualityParam, 5 DesQuality,
 * 6 DesFrame, 7 DesTransferMeshes, 8 DesTransferField, 9 DesTransferStats */
int des_abi_sizeof(int which);

/* ---- instrumentation ----------------------------------------------------------------- */
/* kernels launched by this handle since create (bench.py's gpu_launches) */
long long des_gpu_launch_count(const DesGpu *h);
/* cumulative device time (ms, CUDA events on the launching stream) of the kernel class
 * `which` (0..DES_NKCLASS-1) and its launch count, when timing is enabled */
#define DES_NKCLASS 16
int des_gpu_enable_kernel_timing(DesGpu *h, int on);
int des_gpu_kernel_time(DesGpu *h, int which, double *ms, long long *launches, const char **name);
/* CUDA events on the handle's stream (the stream the kernels are launched on): record
 * slot 0..7, then elapsed milliseconds between two recorded slots (synchronises). */
int des_gpu_event_record(DesGpu *h, int slot);
int des_gpu_event_elapsed_ms(DesGpu *h, int slot_a, int slot_b, double *ms);
/* raw device pointer of a named internal array (for zero-copy harnesses that keep
 * inputs resident in HBM); NULL if unknown */
void *des_gpu_device_ptr(DesGpu *h, const char *name);

#ifdef __cplusplus
}
#endif
#endif /* DES_GPU_H */
