// dynearthsol3d.b200 -- the reference's driver with the step-loop call sites swapped for the
// C-ABI (through des_bridge).  Everything except main() is the reference's own object code
// (oracle/_ref/obj/*.o, dynearthsol.cxx compiled with -Dmain=desref_unused_main): input
// parsing, init(), mesh generation, markers, phase changes, remeshing, output.
//
// main() below follows dynearthsol.cxx:593-1030 for the configurations in scope (isostasy
// spin-up, restart and the pseudo-transient loop included; no GoSPL, hydraulics, rate-and-state/ATS, monitor): the statements that remain are
// the reference's own calls; the lines marked [B200] are the swaps INTEGRATION.md lists.
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <iomanip>
#include <limits>

#include "constants.hpp"
#include "parameters.hpp"
#include "geometry.hpp"
#include "input.hpp"
#include "matprops.hpp"
#include "markerset.hpp"
#include "output.hpp"
#include "phasechanges.hpp"
#include "remeshing.hpp"
#include "utils.hpp"

#include "des_bridge.hpp"

void init_var(const Param& param, Variables& var);      // dynearthsol.cxx:36
void init(const Param& param, Variables& var);          // dynearthsol.cxx:159
void restart(const Param& param, Variables& var);       // dynearthsol.cxx:231
void end(Variables& var);                               // dynearthsol.cxx:440

static int unsupported(const char *what)
{
    std::cerr << "dynearthsol3d.b200: '" << what << "' is outside the B200 explicit core "
                 "(SURVEY.md section 2); use the reference binary for it.\n";
    return 2;
}

int main(int argc, const char* argv[])
{
    if (argc != 2) {
        std::cout << "Usage: " << argv[0] << " config_file\n";
        return -1;
    }
    Param param;
    get_input_parameters(argv[1], param);
    if (param.control.has_hydraulic_diffusion) return unsupported("control.has_hydraulic_diffusion");
    if (param.control.use_global_velocity_scaling) return unsupported("control.use_global_velocity_scaling");

    static Variables var;
    init_var(param, var);
    var.output = new Output(param, var.func_time.start_time,
                            (param.sim.is_restarting) ? param.sim.restarting_from_frame : 0);
    DesBridge *gpu = nullptr;
    if (! param.sim.is_restarting) {
        init(param, var);
        var.info_display_next_step = param.sim.info_display_step_interval;

        // [B200] device model: after init() (dynearthsol.cxx:634), replaces init_offload_device()
        gpu = des_bridge_create(param, var, 0);
        if (!gpu) return 1;
        if (param.ic.isostasy_adjustment_time_in_yr > 0) {
            // [B200] isostasy_adjustment(param, var);   (dynearthsol.cxx:638-641) through the twins,
            // then the adjusted state back into Variables for the first frame
            if (des_bridge_isostasy_adjustment(gpu, param, var)) return 1;
            if (des_bridge_download(gpu, var, DES_DL_ALL)) return 1;
        }
        // [B200] var.dt = compute_dt(param, var);   (dynearthsol.cxx:643)
        var.dt = des_bridge_compute_dt(gpu, param, var);
        if (!(var.dt > 0)) { std::cerr << "dt <= 0\n"; return 11; }

        if (param.sim.has_initial_checkpoint)
            var.output->write_checkpoint(param, var);
    }
    else {
        // restart() (dynearthsol.cxx:231-437) rebuilds mesh, topology, fields, mass and dt on the
        // host from the checkpoint; the device needs nothing extra: the normal upload
        restart(param, var);
        gpu = des_bridge_create(param, var, 0);
        if (!gpu) return 1;
    }

    int64_t init_time = get_nanoseconds() - var.func_time.start_time;
    // (benchmark runs of multi-million-element boxes skip the initial frame: tens of GB with markers)
    if (!std::getenv("DES_BRIDGE_NO_INITIAL_FRAME"))
        var.output->write_exact(var);

    const double starting_time = var.reference_frame_time;
    var.reference_frame_time = starting_time + param.sim.output_time_interval_in_yr * YEAR2SEC;
    const double starting_step = var.steps;
    int next_regular_frame = 1;

    if (param.ic.has_body_force_adjustment) {
        // [B200] initial_body_force_adjustment(param, var);   (dynearthsol.cxx:753-761)
        if (des_bridge_initial_body_force_adjustment(gpu, param, var)) return 1;
    }

    std::cout << "Starting simulation (explicit core on the GPU)...\n";
    do {
        var.steps ++;
        var.time += var.dt;
        // [B200] dynearthsol.cxx:776-876: refresh_elem_cache, update_temperature,
        // update_strain_rate, compute_dvoldt, compute_edvoldt, update_stress, NMD_stress,
        // update_force, update_velocity, apply_vbcs, update_mesh, rotate_stress
        // (with control.has_PT the block runs as the reference's own call sequence through the
        // fine-grained twins, the pseudo-transient loop of dynearthsol.cxx:803-864 in the middle)
        if (param.control.has_PT ? des_bridge_step_pt(gpu, param, var) : des_bridge_step(gpu, param, var)) return 1;

        const int slow_updates_interval = 10;
        if (var.steps % slow_updates_interval == 0) {
            // [B200] phase_changes(param, var); var.mat->refresh_elem_cache();   (dynearthsol.cxx:885-887)
            // -- the marker lookups run on the device-resident markers, coordinates and temperature
            if (des_bridge_phase_changes(gpu, param, var)) return 1;
            if (param.control.has_hydration_processes) {
                // advect_hydrous_markers (dynearthsol.cxx:889-892) stays host code: it walks the
                // hydrous marker set through coord, volume and the support lists
                if (des_bridge_download(gpu, var, DES_DL_GEOMETRY)) return 1;
                advect_hydrous_markers(param, var, 10*var.dt, *var.markersets[var.hydrous_marker_index],
                                       *var.hydrous_elemmarkers);
            }
            // [B200] var.dt = compute_dt(param, var);   (dynearthsol.cxx:893)
            var.dt = des_bridge_compute_dt(gpu, param, var);
            if (!(var.dt > 0)) { std::cerr << "dt <= 0 at step " << var.steps << "\n"; return 11; }
        }

        if (param.sim.is_outputting_averaged_fields) {
            // [B200] var.output->average_fields(var);   (dynearthsol.cxx:903) -- the running sums of
            // stress and delta_plstrain stay on the device, nothing crosses PCIe here
            if (des_bridge_average_fields(gpu, param, var)) return 1;
        }

        if (( (param.sim.output_step_interval != std::numeric_limits<int>::max() &&
               (var.steps - starting_step) == next_regular_frame * param.sim.output_step_interval)
              ||
              (param.sim.output_time_interval_in_yr != std::numeric_limits<double>::max() &&
               (var.time - starting_time) > next_regular_frame * param.sim.output_time_interval_in_yr * YEAR2SEC)
             )
            &&
            ((! param.sim.is_outputting_averaged_fields) ||
                (param.sim.is_outputting_averaged_fields &&
                    (var.steps % param.mesh.quality_check_step_interval == 0)))
            ) {
                const bool checkpoint = next_regular_frame % param.sim.checkpoint_frame_interval == 0;
                if (des_bridge_direct_output_ok(gpu, var)) {
                    // [B200] var.output->write(var);   (dynearthsol.cxx:928) -- the frame's arrays come
                    // from the device in the on-disk layout, nothing passes through Variables
                    if (checkpoint) {
                        des_bridge_output_barrier();
                        des_bridge_download(gpu, var, DES_DL_ALL);
                        var.output->write_checkpoint(param, var);
                    }
                    if (des_bridge_write_frame(gpu, param, var)) return 1;
                }
                else {
                    // [B200] bring the output fields back into Variables before Output::write
                    des_bridge_download(gpu, var, DES_DL_ALL);
                    if (param.sim.is_outputting_averaged_fields)
                        if (des_bridge_flush_averages(gpu, param, var)) return 1;
                    if (checkpoint)
                        var.output->write_checkpoint(param, var);
                    var.output->write(var);
                }
                next_regular_frame ++;
                var.reference_frame_time = starting_time +
                    next_regular_frame * param.sim.output_time_interval_in_yr * YEAR2SEC;
        }

        if (var.steps % param.mesh.quality_check_step_interval == 0) {
            double min_quality = 1.0;
            if (param.control.has_moving_mesh) {
                // [B200] quality_is_bad = bad_mesh_quality(param, var, bad_quality_index, min_quality);
                // (dynearthsol.cxx:942) -- its loops over volume, coord and bcflag run on the device
                int quality_is_bad, bad_quality_index;
                quality_is_bad = des_bridge_bad_mesh_quality(gpu, param, var, bad_quality_index, min_quality);
                if (quality_is_bad < 0) return 1;
                if (quality_is_bad) {
                    // [B200] full state to the host, host remesh, full re-upload (cost reported)
                    des_bridge_output_barrier();    // a frame still being written reads the old mesh
                    int64_t t0 = get_nanoseconds();
                    des_bridge_download(gpu, var, DES_DL_ALL);
                    int64_t t1 = get_nanoseconds();
                    if (param.sim.has_output_during_remeshing) var.output->write_exact(var);
                    int64_t t2 = get_nanoseconds();
                    remesh(param, var, quality_is_bad);
                    int64_t t3 = get_nanoseconds();
                    if (des_bridge_upload(gpu, param, var)) return 1;
                    int64_t t4 = get_nanoseconds();
                    // the reference's "Compute" line excludes remesh() (remeshing.cxx:3185 adds its own
                    // time); the device side of the cycle goes into the same bucket and is printed
                    var.func_time.remesh_time += (t1 - t0) + (t4 - t3);
                    std::cout << "  [B200] download + host remesh + re-upload: ";
                    print_time_ns(t4 - t0 - (t2 - t1));
                    const double transfer_s = des_bridge_take_transfer_seconds();
                    std::cout << "  (download " << (t1 - t0) * 1e-9 << " s, host remesh " << (t3 - t2) * 1e-9
                              << " s";
                    if (transfer_s > 0) std::cout << " of which field transfer on the device " << transfer_s << " s";
                    std::cout << ", re-upload " << (t4 - t3) * 1e-9 << " s)\n";
                    if (param.sim.has_output_during_remeshing) var.output->write_exact(var);
                }
            }
            if (var.steps >= var.info_display_next_step) {
                std::cout << "              Step = " << var.steps
                    << ", time = " << std::scientific << std::setprecision(5)
                    << var.time / YEAR2SEC << " yr"
                    << ", vmax = " << var.max_global_vel_mag << " m/s\n";
                var.info_display_next_step = var.steps + param.sim.info_display_step_interval;
            }
        }
    } while (var.steps < param.sim.max_steps &&
             (var.time <= param.sim.max_time_in_yr * YEAR2SEC ||
              (param.sim.is_outputting_averaged_fields &&
               var.steps % param.mesh.quality_check_step_interval != 0)));

    des_bridge_output_barrier();
    des_bridge_download(gpu, var, DES_DL_ALL);
    des_bridge_destroy(gpu);
    end(var);

    std::cout << "Ending simulation.\n";
    int64_t duration_ns = get_nanoseconds() - var.func_time.start_time;
    int64_t computation_time = duration_ns - var.func_time.remesh_time - var.func_time.output_time - init_time;
    std::cout << "Time summary...\n  Execute : ";
    print_time_ns(duration_ns);
    std::cout << "\n  Compute : ";
    print_time_ns(computation_time);
    std::cout << " / " << var.steps << " = " << std::setprecision(6)
              << computation_time * 1.e-9 / var.steps << " s/step\n";
    return 0;
}
