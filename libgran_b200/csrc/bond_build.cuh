// bond_build.cuh -- the construction of ONE sinter bridge, shared by the host scan (api.cu: scan_bonds) and the device scan
// (kernels_bonds.cu) so that both produce the reference's bits (LG/sinter_bridge/sinter_bridge.h:160-206):
//   n_i = normalized(x_j - x_i), n_j = -n_i; db = v x n_j (the caller redraws v while |db| == 0, :171-175);
//   ds = normalize(v - (v . n_i) n_i), db = normalize(db);
//   connectors of i: x_i +- ds r, x_i +- db r, x_i + n_i r; of j: x_j +- ds r, x_j +- db r, x_j + n_j r; rest lengths 2r x 4, 0.
#pragma once
#include "vec3.cuh"

namespace lg {

// returns false when v is collinear with the bond axis (the reference would draw again)
LG_HD bool build_bridge(d3 part_i, d3 part_j, d3 tv, double r, double* ci15, double* cj15) {
    const d3 n_i = normalized(part_j - part_i);
    const d3 n_j = -n_i;
    d3 db = cross(tv, n_j);
    if (norm(db) == 0.0) return false;
    d3 ds = tv - dot(tv, n_i) * n_i;
    db = db / sqrt(dot(db, db));
    ds = ds / sqrt(dot(ds, ds));
    const d3 ci[5] = {part_i + ds * r, part_i - ds * r, part_i + db * r, part_i - db * r, part_i + n_i * r};
    const d3 cj[5] = {part_j + ds * r, part_j - ds * r, part_j + db * r, part_j - db * r, part_j + n_j * r};
    for (int k = 0; k < 5; k++) {
        store3(ci15 + 3 * k, ci[k]);
        store3(cj15 + 3 * k, cj[k]);
    }
    return true;
}

// the bond predicate of sinter_bridge.h:155-157: NOT (|dist - 2 r| > crit), dist measured from the lower index
LG_HD bool bridge_candidate(d3 part_lo, d3 part_hi, double r, double crit) {
    const double separation = norm(part_hi - part_lo) - 2.0 * r;
    return !(fabs(separation) > crit);
}

}  // namespace lg
