// Dump helpers of a libgran driver program -- the reference keeps them next to main.cpp (writer.h:16-23, writer.cpp:13-170).
//
// write_particles(dir, x, theta, r_part) has the reference's signature and file numbering (particles_1.vtk, particles_2.vtk,
// ...; writer.cpp:9,51-56) and writes the same bytes; it formats host vectors and never touches the device.
// lgb200::async_writer is the device-side pipeline for time loops: write_particles(system) / write_necks(system) /
// write_spring_connectors(system) snapshot what the file needs on the system's stream and return; a copy stream and a
// background thread produce the file (include/lgb200/lgb200.h: lg_dumper_*).  The reference's write_necks /
// write_spring_connectors read the functor's dense N x N tables; here the system's bond table takes their place.
#ifndef LIBGRAN_IO_WRITER_H
#define LIBGRAN_IO_WRITER_H

#include <string>
#include <vector>

#include <libgran/backend/lgb200_backend.h>

namespace lgb200 {
inline std::size_t& particle_file_counter() {
    static std::size_t counter = 1;  // writer.cpp:9
    return counter;
}

class async_writer {
public:
    template <typename system_t>
    async_writer(system_t const& system, std::string const& dir) : handle(system.native_handle()) {
        LGB200_CHECK(lg_dumper_create(handle, dir.c_str(), &dumper));
    }
    async_writer(async_writer const&) = delete;
    async_writer& operator=(async_writer const&) = delete;
    ~async_writer() { lg_dumper_destroy(dumper); }

    void write_particles() { LGB200_CHECK(lg_dump_particles(handle, dumper)); }
    void write_necks() { LGB200_CHECK(lg_dump_necks(handle, dumper)); }
    void write_spring_connectors() { LGB200_CHECK(lg_dump_spring_connectors(handle, dumper)); }
    // the tutorial's dump_particle_positions(dir, count, x, r_part) (Writerside/topics/Custom-unary-force-model.md:339-355)
    void dump_particle_positions(std::size_t count) { LGB200_CHECK(lg_dump_particles_csv(handle, dumper, static_cast<int64_t>(count))); }
    void wait() { LGB200_CHECK(lg_dumper_wait(dumper)); }

private:
    lg_system* handle;
    lg_dumper* dumper = nullptr;
};
}  // namespace lgb200

template <typename field_value_t>
void write_particles(std::string const& dir, std::vector<field_value_t> const& x, std::vector<field_value_t> const& theta,
                     double r_part) {
    std::vector<double> bx, bt;
    lgb200::pack(x, bx);
    lgb200::pack(theta, bt);
    const std::string path = dir + "/particles_" + std::to_string(lgb200::particle_file_counter()++) + ".vtk";
    LGB200_CHECK(lg_write_particles(path.c_str(), static_cast<int64_t>(x.size()), bx.data(), bt.data(), r_part));
}

#endif  // LIBGRAN_IO_WRITER_H
