// Regular output frames written straight from the device.
//
// Output::write (output.cxx:277) needs every field back in `Variables` (14 GB through pageable
// memory at 51 M tets), transposes each Array2D on one host thread (binaryio.cxx:151-166) and
// runs host loops for the averaged fields, the density and the mesh quality
// (output.cxx:131-202).  Here the device hands over each array already in the on-disk layout
// with those loops applied (des_gpu_download_frame), into page-locked buffers, and this file only
// fwrites them -- through the reference's own BinaryOutput object, so the header, the scalars, the
// host-side arrays (connectivity, pore pressure, radiogenic source, coord0, bcflag), the marker
// sets and the .info line are the reference's code and the file is byte-identical to what
// Output::write would have produced.  With sim.has_marker_output = no the fwrites run on a
// background thread while the device goes on stepping.
//
// Output and BinaryOutput keep everything private and the reference's objects stay unmodified,
// so this one translation unit reads their headers with `private` opened up (the same kind of
// seam as the --wrap'ed topology builders: no reference source is changed).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <iostream>
#include <iomanip>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "constants.hpp"
#include "parameters.hpp"
#include "array2d.hpp"
#include "matprops.hpp"
#include "markerset.hpp"
#include "utils.hpp"
#define private public
#include "binaryio.hpp"
#include "output.hpp"
#undef private

#include "des_bridge.hpp"

namespace {

const char *frame_fields[] = {
    "coordinate", "velocity", "velocity_averaged", "temperature", "plstrain", "plstrain_rate",
    "strain_rate", "strain", "stress", "viscosity", "stress_averaged", "density", "mesh_quality",
    "material", "force" };
enum { F_COORD, F_VEL, F_VEL_AVG, F_TEMP, F_PLS, F_PLS_RATE, F_SRATE, F_STRAIN, F_STRESS, F_VISC,
       F_STRESS_AVG, F_DENSITY, F_QUALITY, F_MATERIAL, F_FORCE, F_COUNT };
// components per node / element; > 0 per element, < 0 per node
const int frame_width[F_COUNT] = { -3, -3, -3, -1, 1, 1, 6, 6, 6, 1, 6, 1, 1, 1, -3 };

struct FrameWriter {
    double *buf[F_COUNT] = {};
    size_t cap[F_COUNT] = {};
    bool pinned[F_COUNT] = {};
    std::thread th;
    double last_download_s = 0, last_write_s = 0;

    void join() { if (th.joinable()) th.join(); }
    int reserve(int i, size_t count)
    {
        if (count <= cap[i]) return 0;
        release(i);
        void *p = nullptr;
        if (des_host_alloc(count * sizeof(double), &p) == DES_OK) pinned[i] = true;
        else { p = std::malloc(count * sizeof(double)); pinned[i] = false; }    // pageable: slower copies
        if (!p) return 1;
        buf[i] = static_cast<double *>(p); cap[i] = count;
        return 0;
    }
    void release(int i)
    {
        if (buf[i]) { if (pinned[i]) des_host_free(buf[i]); else std::free(buf[i]); }
        buf[i] = nullptr; cap[i] = 0;
    }
    // the buffers are left to the process exit: this object outlives the CUDA runtime
    ~FrameWriter() { join(); }
};

FrameWriter g_writer;

// what BinaryOutput::write_array does after pack_to (binaryio.cxx:151-176) for a buffer that is
// already packed
void write_packed(BinaryOutput &bin, const char *name, const double *p, size_t count)
{
    bin.write_header(name);
    const size_t n = std::fwrite(p, sizeof(double), count, bin.f);
    if (n != count) std::cerr << "Error: cannot write array: " << name << '\n';
    bin.eof_pos += n * sizeof(double);
}

struct FrameJob {
    std::string filename;
    bool rename_if_exists, averaged, markers;
    int steps, nnode, nelem, nseg;
    double walltime, time, dt;
    const Variables *var;
};

// the body of Output::_write between the BinaryOutput constructor and write_info
// (output.cxx:119-243), array for array in the same order
void write_frame_file(const FrameJob &j, const FrameWriter &w)
{
    const Variables &var = *j.var;
    const size_t N = j.nnode, E = j.nelem;
    BinaryOutput bin(j.filename.c_str(), j.rename_if_exists);
    write_packed(bin, "coordinate", w.buf[F_COORD], N * 3);
    bin.write_array(*var.connectivity, "connectivity", var.connectivity->size());
    bin.write_scalar(j.steps, "steps");
    bin.write_scalar(j.walltime, "walltime_sec");
    bin.write_scalar(j.nnode, "nnode");
    bin.write_scalar(j.nelem, "nelem");
    bin.write_scalar(j.time, "time_sec");
    bin.write_scalar(j.dt, "dt_sec");
    bin.write_scalar(j.nseg, "nseg");
    write_packed(bin, "velocity", w.buf[F_VEL], N * 3);
    if (j.averaged) write_packed(bin, "velocity averaged", w.buf[F_VEL_AVG], N * 3);
    write_packed(bin, "temperature", w.buf[F_TEMP], N);
    bin.write_array(*var.ppressure, "pore pressure", var.ppressure->size());
    bin.write_array(*var.radiogenic_source, "radiogenic source", var.radiogenic_source->size());
    write_packed(bin, "plastic strain", w.buf[F_PLS], E);
    write_packed(bin, "plastic strain-rate", w.buf[F_PLS_RATE], E);
    write_packed(bin, "strain-rate", w.buf[F_SRATE], E * 6);
    write_packed(bin, "strain", w.buf[F_STRAIN], E * 6);
    write_packed(bin, "stress", w.buf[F_STRESS], E * 6);
    write_packed(bin, "viscosity", w.buf[F_VISC], E);
    if (j.averaged) write_packed(bin, "stress averaged", w.buf[F_STRESS_AVG], E * 6);
    write_packed(bin, "density", w.buf[F_DENSITY], E);
    write_packed(bin, "mesh quality", w.buf[F_QUALITY], E);
    write_packed(bin, "material", w.buf[F_MATERIAL], E);
    write_packed(bin, "force", w.buf[F_FORCE], N * 3);
    bin.write_array(*var.coord0, "coord0", var.coord0->size());
    bin.write_array(*var.bcflag, "bcflag", var.bcflag->size());
    if (j.markers)
        for (auto ms = var.markersets.begin(); ms != var.markersets.end(); ++ms)
            (*ms)->write_save_file(var, bin);
}

} // namespace

void des_bridge_output_barrier() { g_writer.join(); }

// DES_BRIDGE_DIRECT_OUTPUT=0 keeps Output::write; a group handle gathers through
// des_gpu_download_state; rate-and-state friction writes two more arrays (output.cxx:204-207)
bool des_bridge_direct_output_ok(DesBridge *b, const Variables &var)
{
    static const bool on = !(std::getenv("DES_BRIDGE_DIRECT_OUTPUT") && std::atoi(std::getenv("DES_BRIDGE_DIRECT_OUTPUT")) == 0);
    if (!on || (var.mat->rheol_type & MatProps::rh_rsf)) return false;
    int nparts = 1;
    des_gpu_partition_info(des_bridge_handle(b), &nparts, nullptr, nullptr, nullptr, nullptr, nullptr);
    return nparts == 1;
}

void des_bridge_frame_timing(double *download_s, double *write_s)
{
    if (download_s) *download_s = g_writer.last_download_s;
    if (write_s) *write_s = g_writer.last_write_s;
}

int des_bridge_write_frame(DesBridge *b, const Param &param, Variables &var)
{
    const int64_t t_begin = get_nanoseconds();
    FrameWriter &w = g_writer;
    w.join();                       // the buffers and `var.output` are single-user
    Output &out = *var.output;
    DesGpu *h = des_bridge_handle(b);

    // output.cxx:99-106
    out.run_time_ns = get_nanoseconds() - out.start_time;
    const bool averaged = out.is_averaged;
    double dt = var.dt, inv_dt = 0;
    const double time0 = des_bridge_avg_time0(b);
    if (averaged) {
        dt = (var.time - time0) / out.average_interval;
        inv_dt = 1.0 / (var.time - time0);
    }

    const size_t N = var.nnode, E = var.nelem;
    DesFrame f;
    std::memset(&f, 0, sizeof(f));
    f.averaged = averaged ? 1 : 0;
    f.inv_dt = inv_dt;
    f.stress_avg_scale = 1.0 / (out.average_interval + 1);
    double **slot[F_COUNT] = { &f.coordinate, &f.velocity, &f.velocity_averaged, &f.temperature, &f.plstrain,
        &f.plstrain_rate, &f.strain_rate, &f.strain, &f.stress, &f.viscosity, &f.stress_averaged,
        &f.density, &f.mesh_quality, &f.material, &f.force };
    for (int i = 0; i < F_COUNT; ++i) {
        if (!averaged && (i == F_VEL_AVG || i == F_STRESS_AVG)) continue;
        const size_t count = frame_width[i] > 0 ? E * frame_width[i] : N * size_t(-frame_width[i]);
        if (w.reserve(i, count)) { std::cerr << "frame buffer '" << frame_fields[i] << "': out of host memory\n"; return 1; }
        *slot[i] = w.buf[i];
    }
    if (des_gpu_download_frame(h, &f)) {
        std::cerr << "des_gpu_download_frame: " << des_gpu_last_error(h) << "\n";
        return 1;
    }
    const bool markers = out.has_marker_output;
    if (markers) {
        // MarkerSet::calculate_marker_coord (markerset.cxx:989) reads var.coord
        DesState s;
        std::memset(&s, 0, sizeof(s));
        s.coord = var.coord->data();
        if (des_gpu_download_state(h, &s)) { std::cerr << des_gpu_last_error(h) << "\n"; return 1; }
    }
    const int64_t t_down = get_nanoseconds();
    w.last_download_s = (t_down - t_begin) * 1e-9;
    static const bool verbose = std::getenv("DES_BRIDGE_FRAME_TIMING") != nullptr;

    FrameJob j;
    char filename[256];
    std::snprintf(filename, 255, "%s.save.%06d", out.modelname.c_str(), out.frame);
    j.filename = filename;
    j.rename_if_exists = out.may_overwrite_ && (out.frame == out.start_frame_);
    j.averaged = averaged; j.markers = markers;
    j.steps = var.steps; j.nnode = var.nnode; j.nelem = var.nelem; j.nseg = var.nseg;
    j.walltime = double(out.run_time_ns) * 1e-9; j.time = var.time; j.dt = dt;
    j.var = &var;

    static const bool async_ok = !(std::getenv("DES_BRIDGE_ASYNC_OUTPUT") && std::atoi(std::getenv("DES_BRIDGE_ASYNC_OUTPUT")) == 0);
    if (markers || !async_ok) {
        // the marker arrays change on the host between frames (phase changes, surface markers)
        write_frame_file(j, w);
        w.last_write_s = (get_nanoseconds() - t_down) * 1e-9;
    }
    else {
        w.th = std::thread([j, &w]() {
            const int64_t t0 = get_nanoseconds();
            write_frame_file(j, w);
            w.last_write_s = (get_nanoseconds() - t0) * 1e-9;
        });
    }

    // output.cxx:245-274
    out.write_info(var, dt);
    std::cout << "  Output # " << out.frame << ", step = " << var.steps;
    if (dt / YEAR2SEC > 0.001)
        std::cout << ", time = " << std::scientific << std::setprecision(5) << var.time / YEAR2SEC << " yr"
                  << ", vmax = " << var.max_global_vel_mag << " m/s"
                  << ", dt = " << std::scientific << std::setprecision(5) << dt / YEAR2SEC << " yr";
    else
        std::cout << ", time = " << std::scientific << std::setprecision(5) << var.time << " sec"
                  << ", vmax = " << var.max_global_vel_mag << " m/s"
                  << ", dt = " << std::scientific << std::setprecision(5) << dt << " sec";
    std::cout << ", wt = ";
    print_time_ns(out.run_time_ns);
    std::cout << "\n";
    out.frame++;
    if (verbose)
        std::cout << "  [B200] frame from the device: download " << w.last_download_s << " s, "
                  << ((markers || !async_ok) ? "file written in " : "file being written in the background; previous one took ")
                  << w.last_write_s << " s\n";

    // Output::write, output.cxx:283-285
    var.noutput += 1;
    var.func_time.output_time += get_nanoseconds() - t_begin;
    return 0;
}
