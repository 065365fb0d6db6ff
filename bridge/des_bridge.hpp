// Bridge between the reference's C++ driver and the C-ABI of include/des_gpu.h.
//
// This is the file a DynEarthSol maintainer adds next to dynearthsol.cxx: it marshals
// `Param` / `Variables` (parameters.hpp:160-832) into the POD structs of des_gpu.h and
// back.  It is compiled against the reference's own headers and re-reads every host pointer
// from `Variables` on each call (remesh() re-news nearly every array, fields.cxx:125-194,
// and update_mesh swaps volume/volume_old each step, dynearthsol.cxx:466-470).
// See INTEGRATION.md for the call-site swaps in dynearthsol.cxx.
#ifndef DES_BRIDGE_HPP
#define DES_BRIDGE_HPP

#include "parameters.hpp"
#include "../include/des_gpu.h"

// what des_bridge_download brings back into `Variables`
enum {
    DES_DL_OUTPUT = 1,      // everything Output::_write reads (output.cxx:95-274)
    DES_DL_GEOMETRY = 2,    // coord, volume, temperature: bad_mesh_quality, phase_changes
    DES_DL_ALL = 4          // every evolving + derived array (before remesh / checkpoint)
};

struct DesBridge;

// after init() / restart(): creates the device model (device `device`) and uploads mesh + state
DesBridge *des_bridge_create(const Param &param, const Variables &var, int device);
// after every remesh() (dynearthsol.cxx:950): new topology and all fields
int des_bridge_upload(DesBridge *b, const Param &param, const Variables &var);
// the block dynearthsol.cxx:776-876 (refresh_elem_cache ... rotate_stress); var.steps,
// var.time already incremented by the caller as in the reference (:773-774)
int des_bridge_step(DesBridge *b, const Param &param, Variables &var);
// control.has_PT: the same block as the reference's un-fused call sequence through the fine-grained
// twins, with the pseudo-transient loop (dynearthsol.cxx:803-864) between update_velocity and
// apply_vbcs; sets var.l2_residual.  (param is not const: the loop toggles control.PT_jump.)
int des_bridge_step_pt(DesBridge *b, Param &param, Variables &var);
// initial_body_force_adjustment (dynearthsol.cxx:546-591, 753-761), before the step loop
int des_bridge_initial_body_force_adjustment(DesBridge *b, Param &param, Variables &var);
// compute_dt (dynearthsol.cxx:893): device reductions + the reference's scalar algebra;
// also sets var.max_global_vel_mag, var.global_dt_min, var.surfinfo.max_surf_vel
double des_bridge_compute_dt(DesBridge *b, const Param &param, Variables &var);
// device -> the reference's own arrays, before host code reads them
int des_bridge_download(DesBridge *b, Variables &var, int what);
// seconds of the device field transfers (bridge/des_transfer_wrap.cxx) inside the last remesh()
double des_bridge_take_transfer_seconds();
// phase_changes + refresh_elem_cache (dynearthsol.cxx:885-887) with the marker lookups on the
// device (des_gpu_phase_changes); the changed rock types are written back into the MarkerSet
int des_bridge_phase_changes(DesBridge *b, const Param &param, Variables &var);
// elemmarkers changed on the host (phase_changes, surface markers): dynearthsol.cxx:887
int des_bridge_markers_changed(DesBridge *b, const Param &param, const Variables &var);
// isostasy_adjustment (dynearthsol.cxx:496-539), between init() and the first compute_dt: the
// same call sequence through the fine-grained twins; sets var.dt like the reference
int des_bridge_isostasy_adjustment(DesBridge *b, const Param &param, Variables &var);
// bad_mesh_quality (dynearthsol.cxx:942; remeshing.cxx:2752-2866) from device reductions: same
// return code, index, min_quality and diagnostic line as the reference, no field download
int des_bridge_bad_mesh_quality(DesBridge *b, const Param &param, const Variables &var,
                                int &index, double &min_quality);
// Output::average_fields (dynearthsol.cxx:903; output.cxx:312-352) with the accumulators kept
// on the device: replaces the per-step download of stress and delta_plstrain
int des_bridge_average_fields(DesBridge *b, const Param &param, const Variables &var);
// before Output::write on a frame step: hands the device accumulators to the Output object
// (through its own average_fields, so output.cxx stays untouched)
int des_bridge_flush_averages(DesBridge *b, const Param &param, Variables &var);
// Output::write for a regular frame (dynearthsol.cxx:928) without the trip through `Variables`:
// the device delivers each array in the on-disk layout (des_gpu_download_frame) and the file is
// written through the reference's BinaryOutput (bridge/des_frame_writer.cxx), in the background
// when sim.has_marker_output = no.  Replaces des_bridge_download + des_bridge_flush_averages +
// var.output->write(var).
int des_bridge_write_frame(DesBridge *b, const Param &param, Variables &var);
bool des_bridge_direct_output_ok(DesBridge *b, const Variables &var);
// waits for a frame that is still being written; before any other use of var.output, before
// remesh() and before the run ends
void des_bridge_output_barrier();
void des_bridge_frame_timing(double *download_s, double *write_s);
DesGpu *des_bridge_handle(DesBridge *b);
double des_bridge_avg_time0(const DesBridge *b);
const char *des_bridge_last_error(const DesBridge *b);
void des_bridge_destroy(DesBridge *b);

#endif
