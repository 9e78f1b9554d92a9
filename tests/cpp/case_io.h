// Case file IO shared by the C++ host drivers (template_driver.cpp, variants_driver.cpp); format in template_driver.cpp.
#pragma once
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <vector>

#include "vec3.h"

enum { M_CONTACT = 1, M_HAMAKER = 2, M_SINTER = 4, M_FACET = 8 };

struct Case {
    int64_t n, mask, use_nl, n_steps, rebuild_every, move_facets, n_uv;
    double r_part, mass, inertia, r_verlet, dt, contact[15], A, h0, k_d, gamma_d, crit, fv[9], fvel[3], sc[15], sA, sh0;
    std::vector<vec3> x, v, th, w, uv;
};

static void read_exact(FILE* f, void* p, size_t bytes) {
    if (fread(p, 1, bytes, f) != bytes) throw std::runtime_error("short read");
}

static std::vector<vec3> read_vecs(FILE* f, int64_t count) {
    std::vector<double> raw(3 * (size_t)count);
    if (count) read_exact(f, raw.data(), raw.size() * sizeof(double));
    std::vector<vec3> out((size_t)count);
    for (size_t i = 0; i < out.size(); i++) out[i] = vec3(raw[3 * i], raw[3 * i + 1], raw[3 * i + 2]);
    return out;
}

static Case load(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) throw std::runtime_error("cannot open case file");
    Case c;
    read_exact(f, &c.n, 7 * sizeof(int64_t));
    read_exact(f, &c.r_part, (5 + 15 + 5 + 9 + 3 + 15 + 2) * sizeof(double));
    c.x = read_vecs(f, c.n);
    c.v = read_vecs(f, c.n);
    c.th = read_vecs(f, c.n);
    c.w = read_vecs(f, c.n);
    c.uv = read_vecs(f, c.n_uv);
    fclose(f);
    return c;
}

template <typename System>
static void dump(const char* path, System& sys) {
    FILE* f = fopen(path, "wb");
    if (!f) throw std::runtime_error("cannot open output file");
    const std::vector<vec3>* fields[6] = {&sys.get_x(), &sys.get_v(), &sys.get_theta(), &sys.get_omega(), &sys.get_a(), &sys.get_alpha()};
    for (auto* fld : fields)
        for (auto const& q : *fld) fwrite(q.c, sizeof(double), 3, f);
    fclose(f);
}

