// generators.cu -- host-side initial-condition tooling behind the C ABI.  Pure host code
// (this translation unit launches nothing); it lives in the library so that a C++ driver and the Python benches build
// the SAME synthetic inputs for BASELINE.json's configs 2, 3 and 5 in seconds instead of minutes of interpreter loops.
// What a libgran driver does before the time loop (main.cpp:56-93, test/hamaker_3.cpp:60-87), at scale.
//
// Randomness: std::mt19937_64 (the generator the reference's drivers use, test/sintered.cpp:30-39) consumed through
// explicit integer -> double conversions, so a (seed, n) pair produces the same coordinates on every platform.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <atomic>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/lgb200/lgb200.h"

namespace lg {
int fail_message(int code, const char* msg);  // api.cu
}

namespace {

struct Rng {
    std::mt19937_64 g;
    explicit Rng(uint64_t seed) : g(seed) {}
    double uniform() { return (double)(g() >> 11) * (1.0 / 9007199254740992.0); }  // [0, 1)
    uint64_t below(uint64_t n) { return g() % n; }                                 // n << 2^64: bias is negligible
    void direction(double d[3]) {  // uniform on the sphere (Marsaglia 1972)
        for (;;) {
            const double a = 2.0 * uniform() - 1.0, b = 2.0 * uniform() - 1.0;
            const double s = a * a + b * b;
            if (s >= 1.0 || s == 0.0) continue;
            const double t = 2.0 * std::sqrt(1.0 - s);
            d[0] = a * t;
            d[1] = b * t;
            d[2] = 1.0 - 2.0 * s;
            return;
        }
    }
};

inline uint64_t cell_key(int64_t cx, int64_t cy, int64_t cz) {
    const uint64_t m = (1ull << 21) - 1ull;
    return (((uint64_t)(cx + (1 << 20)) & m) << 42) | (((uint64_t)(cy + (1 << 20)) & m) << 21) | ((uint64_t)(cz + (1 << 20)) & m);
}

}  // namespace

extern "C" {

// Particle-cluster sticking (SURVEY.md 8d cfg 2 / 3): particle k is placed at exactly 2 r from a uniformly chosen
// earlier particle in a uniform random direction and rejected if it comes within 2 r (1 + 1e-9) of any OTHER particle.
int lg_gen_fractal_aggregate(int64_t n, uint64_t seed, double r_part, double* x_out) {
    if (n <= 0 || !x_out || !(r_part > 0.0)) return lg::fail_message(LG_EINVAL, "bad arguments to lg_gen_fractal_aggregate");
    Rng rng(seed);
    const double cell = 2.0 * r_part * 1.001;
    const double lim2 = std::pow(2.0 * r_part * (1.0 + 1e-9), 2.0);
    std::unordered_map<uint64_t, std::vector<int32_t>> grid;
    grid.reserve((size_t)n * 2);
    auto key_of = [&](const double* p, int64_t c[3]) {
        for (int k = 0; k < 3; k++) c[k] = (int64_t)std::floor(p[k] / cell);
    };
    x_out[0] = x_out[1] = x_out[2] = 0.0;
    {
        int64_t c[3];
        key_of(x_out, c);
        grid[cell_key(c[0], c[1], c[2])].push_back(0);
    }
    int64_t count = 1;
    while (count < n) {
        const int64_t parent = (int64_t)rng.below((uint64_t)count);
        double d[3];
        rng.direction(d);
        double cand[3];
        for (int k = 0; k < 3; k++) cand[k] = x_out[3 * parent + k] + 2.0 * r_part * d[k];
        int64_t c[3];
        key_of(cand, c);
        bool ok = true;
        for (int64_t dx = -1; dx <= 1 && ok; dx++)
            for (int64_t dy = -1; dy <= 1 && ok; dy++)
                for (int64_t dz = -1; dz <= 1 && ok; dz++) {
                    auto it = grid.find(cell_key(c[0] + dx, c[1] + dy, c[2] + dz));
                    if (it == grid.end()) continue;
                    for (int32_t j : it->second) {
                        if (j == parent) continue;
                        const double ex = x_out[3 * (size_t)j] - cand[0], ey = x_out[3 * (size_t)j + 1] - cand[1],
                                     ez = x_out[3 * (size_t)j + 2] - cand[2];
                        if (ex * ex + ey * ey + ez * ez < lim2) {
                            ok = false;
                            break;
                        }
                    }
                }
        if (!ok) continue;
        std::memcpy(x_out + 3 * count, cand, sizeof(cand));
        grid[cell_key(c[0], c[1], c[2])].push_back((int32_t)count);
        count++;
    }
    return LG_OK;
}

// Random sequential addition of non-overlapping spheres (SURVEY.md 8d cfg 5) to `volume_fraction` in a box of
// aspect : 1 : 1 (elongated in x, the slab axis).  min_gap_factor scales the hard-core diameter: 1.0 = no overlaps,
// slightly below 1 admits shallow initial contacts (overlap <= (1 - factor) * 2 r).  box_out = edge lengths.
int lg_gen_random_packing(int64_t n, double volume_fraction, double aspect, double min_gap_factor, uint64_t seed, double r_part,
                          double* x_out, double box_out[3]) {
    if (n <= 0 || !x_out || !(r_part > 0.0) || !(volume_fraction > 0.0) || volume_fraction > 0.36 || !(aspect > 0.0) ||
        !(min_gap_factor > 0.5) || min_gap_factor > 1.0)
        return lg::fail_message(LG_EINVAL, "bad arguments to lg_gen_random_packing (volume_fraction <= 0.36: RSA jams at 0.38)");
    const double pi = 3.14159265358979323846;
    const double vol = (double)n * (4.0 / 3.0 * pi * r_part * r_part * r_part) / volume_fraction;
    const double ly = std::cbrt(vol / aspect);
    const double box[3] = {aspect * ly, ly, ly};
    const double dmin = 2.0 * r_part * min_gap_factor;
    int64_t dim[3];
    double csz[3];
    for (int k = 0; k < 3; k++) {
        dim[k] = std::max<int64_t>(1, (int64_t)std::floor(box[k] / dmin));
        csz[k] = box[k] / (double)dim[k];
    }
    const int64_t n_cells = dim[0] * dim[1] * dim[2];
    std::vector<int32_t> head((size_t)n_cells, -1), next((size_t)n, -1);
    const double lim2 = dmin * dmin;
    // The box is cut into K chunks along x (whole cell columns, at least 3 cells wide); chunk c receives the ids
    // [c n / K, (c + 1) n / K) and its own generator (seed + c).  Even chunks are filled first, concurrently -- their
    // neighbours are empty -- then the odd chunks between them, whose neighbours are complete and only read.  The result
    // depends on (seed, n, K) but not on the number of threads; ids come out coarsely ordered in x.
    int64_t K = std::min<int64_t>(std::max<int64_t>(1, n / 40000), dim[0] / 3);
    if (K < 1) K = 1;
    std::atomic<int> failed{0};
    auto fill_chunk = [&](int64_t c) {
        Rng rng(seed + (uint64_t)c);
        const int64_t cx0 = c * dim[0] / K, cx1 = (c + 1) * dim[0] / K;
        const double x_lo = (double)cx0 * csz[0], x_w = (double)(cx1 - cx0) * csz[0];
        const int64_t id0 = c * n / K, id1 = (c + 1) * n / K;
        uint64_t tries = 0;
        const uint64_t max_tries = (uint64_t)(id1 - id0 + 1) * 4000ull;
        int64_t count = id0;
        while (count < id1) {
            if (++tries > max_tries) {
                failed.store(1);
                return;
            }
            double p[3];
            int64_t cc[3];
            p[0] = x_lo + rng.uniform() * x_w;
            p[1] = rng.uniform() * box[1];
            p[2] = rng.uniform() * box[2];
            for (int k = 0; k < 3; k++) cc[k] = std::min<int64_t>((int64_t)(p[k] / csz[k]), dim[k] - 1);
            cc[0] = std::min<int64_t>(std::max<int64_t>(cc[0], cx0), cx1 - 1);
            bool ok = true;
            for (int64_t dx = -1; dx <= 1 && ok; dx++) {
                const int64_t cx = cc[0] + dx;
                if (cx < 0 || cx >= dim[0]) continue;
                for (int64_t dy = -1; dy <= 1 && ok; dy++) {
                    const int64_t cy = cc[1] + dy;
                    if (cy < 0 || cy >= dim[1]) continue;
                    for (int64_t dz = -1; dz <= 1 && ok; dz++) {
                        const int64_t cz = cc[2] + dz;
                        if (cz < 0 || cz >= dim[2]) continue;
                        for (int32_t j = head[(size_t)((cx * dim[1] + cy) * dim[2] + cz)]; j >= 0; j = next[(size_t)j]) {
                            const double ex = x_out[3 * (size_t)j] - p[0], ey = x_out[3 * (size_t)j + 1] - p[1],
                                         ez = x_out[3 * (size_t)j + 2] - p[2];
                            if (ex * ex + ey * ey + ez * ez < lim2) {
                                ok = false;
                                break;
                            }
                        }
                    }
                }
            }
            if (!ok) continue;
            std::memcpy(x_out + 3 * count, p, sizeof(p));
            const size_t cl = (size_t)((cc[0] * dim[1] + cc[1]) * dim[2] + cc[2]);
            next[(size_t)count] = head[cl];
            head[cl] = (int32_t)count;
            count++;
        }
    };
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    for (int phase = 0; phase < 2; phase++) {
        std::vector<int64_t> todo;
        for (int64_t c = phase; c < K; c += 2) todo.push_back(c);
        std::atomic<size_t> cursor{0};
        std::vector<std::thread> pool;
        const unsigned nt = (unsigned)std::min<size_t>(hw, todo.size());
        for (unsigned t = 0; t < nt; t++)
            pool.emplace_back([&] {
                for (size_t q = cursor.fetch_add(1); q < todo.size(); q = cursor.fetch_add(1)) fill_chunk(todo[q]);
            });
        for (auto& th : pool) th.join();
    }
    if (failed.load()) return lg::fail_message(LG_ECAPACITY, "random packing: a chunk of the box jammed before its particles were placed");
    if (box_out)
        for (int k = 0; k < 3; k++) box_out[k] = box[k];
    return LG_OK;
}

}  // extern "C"
