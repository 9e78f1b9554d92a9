"""A/B of the neighbor-row rebuild kernels (round-2 row_sweep2_kernel + tag_clear_kernel against the round-1 kernels,
LGB200_REBUILD_V1=1): the rows must be identical arrays, and the wall-clock per lg_update_neighbor_list is reported.

    python scripts/rebuild_ab.py            # runs itself twice (v2, v1) and compares
The round-1 kernels are bit-equal to the oracle / reference (tests/test_gpu_parity.py), so equality with them is parity."""
import hashlib
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cases():
    """(name, x, r_verlet, owned mask or None); AB_CASES=name,name restricts the list (profiling runs)."""
    from libgran_b200 import scenarios as S

    r = S.R_PART
    rng = np.random.default_rng(5)
    want = [c for c in os.environ.get("AB_CASES", "").split(",") if c]

    def packing():
        return S.random_packing(1_250_000, volume_fraction=0.30).state[0]

    def lattice():
        return S.lattice_block(40, 40, 40, spacing=2.0, jitter=0.02).state[0]

    table = [
        ("packing_1.25M_rv5r", lambda: (packing(), 5.0 * r, None)),
        ("packing_300k_shuffled_ids", lambda: (np.ascontiguousarray(packing()[:300_000][rng.permutation(300_000)]), 5.0 * r, None)),
        ("cfg4_1M_aggregates", lambda: (S.multi_aggregate_collision(6667).state[0], 5.0 * r, None)),
        # rows of ~80 in the bulk, shorter on the faces: both the staged and the direct path of the fill pass
        ("lattice_40^3_long_rows", lambda: (lattice(), 5.2 * r, None)),
        ("lattice_40^3_owned_mask", lambda: (lattice(), 3.0 * r, (rng.random(64000) < 0.7).astype(np.uint8))),
        ("three_particles", lambda: (np.array([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0], [50.0, 0.0, 0.0]]) * r, 5.0 * r, None)),
    ]
    for name, make in table:
        if not want or name in want:
            yield (name,) + make()


def child():
    import torch

    from libgran_b200 import scenarios as S
    from libgran_b200.system import GranularSystem

    res = {}
    for name, x, rv, owned in cases():
        n = len(x)
        z = np.zeros((n, 3))
        s = GranularSystem(x, z, z, z, S.R_PART, S.MASS, S.INERTIA, device=0, r_verlet=rv)
        if owned is not None:
            d_owned = torch.from_numpy(owned).cuda()
            torch.cuda.synchronize()
            s._check(s.lib.lg_set_owned_device(s.h, d_owned.data_ptr()))
        s.update_neighbor_list()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = int(os.environ.get("AB_REPS", "5"))
        for _ in range(reps):
            s.update_neighbor_list()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / reps * 1e3
        ptr, cols = s.rows()
        h = hashlib.sha256(ptr.tobytes() + cols.tobytes()).hexdigest()[:16]
        res[name] = {"n": n, "pairs": int(len(cols)), "max_row": int(np.diff(ptr).max()), "ms_per_rebuild": round(ms, 4), "sha": h}
        s.close()
    print("RESULT " + json.dumps(res))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        return child()
    got = {}
    for mode in ("v2", "v1"):
        env = dict(os.environ)
        env["LGB200_REBUILD_V1"] = "1" if mode == "v1" else "0"
        out = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True, timeout=280)
        line = [l for l in out.stdout.splitlines() if l.startswith("RESULT ")]
        if not line:
            print(out.stdout[-2000:], out.stderr[-4000:])
            raise SystemExit(f"{mode}: child failed rc={out.returncode}")
        got[mode] = json.loads(line[0][7:])
    ok = True
    for name in got["v2"]:
        a, b = got["v2"][name], got["v1"][name]
        same = a["sha"] == b["sha"] and a["pairs"] == b["pairs"]
        ok &= same
        print(json.dumps({"case": name, "n": a["n"], "pairs": a["pairs"], "max_row": a["max_row"], "rows_identical": same,
                          "ms_v2": a["ms_per_rebuild"], "ms_v1": b["ms_per_rebuild"]}))
    raise SystemExit(0 if ok else 1)


if __name__ == "__main__":
    main()
