"""lg_diag_sums_static: the allocation-free twin of lg_diagnostics_partial (no counterpart in the reference; added for
callers that cannot call cudaMalloc).  Its per-block partial sums live in a file-scope __device__ array of the library, so
the test runs it twice in a row -- a large problem that fills every block's slot, then a small one that must not see the
first call's leftovers -- and compares each against the existing entry point on an lg_system holding the same state.
Both paths launch the same two kernels with the same grid, so the sums agree to the last bit; the assertion uses 1e-10,
the tolerance of the other force / diagnostics tests.

This file is named to be collected last."""
import ctypes as C

import numpy as np
import pytest

LG_OK, LG_EINVAL, LG_ESIZE = 0, -1, -5
R_PART, MASS, INERTIA = 1.4e-8, 1.7e-20, 1.3e-36


def test_rejects_sizes_beyond_the_maximum(lib):
    """Needs no GPU: the size check comes before any CUDA call."""
    n_max = lib.lg_diag_sums_static_max_particles()
    assert n_max == 1 << 26
    rc = lib.lg_diag_sums_static(n_max + 1, None, None, None, MASS, INERTIA, 0, None, None, None)
    assert rc == LG_ESIZE
    msg = lib.lg_last_error().decode()
    assert "exceeds" in msg and str(n_max) in msg
    assert lib.lg_diag_sums_static(0, None, None, None, MASS, INERTIA, 0, None, None, None) == LG_EINVAL
    assert lib.lg_diag_sums_static(10, None, None, None, MASS, INERTIA, 0, None, None, None) == LG_EINVAL  # null pointers
    assert lib.lg_diag_sums_static(10, None, None, None, MASS, INERTIA, 2, None, None, None) == LG_EINVAL  # bad pass


def _state(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1.0, 1.0, size=(n, 3)) * 200.0 * R_PART
    v = rng.normal(size=(n, 3)) * 0.3
    th = rng.normal(size=(n, 3))
    w = rng.normal(size=(n, 3)) * 1.0e6
    return x, v, th, w


def _existing(lib, state):
    """lg_diagnostics_partial on a system holding `state`: the 12 sums (pass 0 then pass 1 about the centre of mass)."""
    import torch

    from libgran_b200.slab import _view
    from libgran_b200.system import GranularSystem

    x, v, th, w = state
    n = len(x)
    sys_ = GranularSystem(x, v, th, w, R_PART, MASS, INERTIA, device=0)
    p = C.c_void_p()
    assert lib.lg_diagnostics_partial(sys_.h, 0, None, C.byref(p)) == LG_OK, lib.lg_last_error()
    torch.cuda.synchronize()  # the sums are queued on the system's own stream
    sums = _view(p.value, (12,), torch.float64, torch.device("cuda", 0)).clone()
    com = (sums[8:11] / n).contiguous()
    torch.cuda.synchronize()
    assert lib.lg_diagnostics_partial(sys_.h, 1, com.data_ptr(), C.byref(p)) == LG_OK, lib.lg_last_error()
    torch.cuda.synchronize()
    sums[11] = _view(p.value, (12,), torch.float64, torch.device("cuda", 0))[11]
    out = sums.cpu().numpy()
    sys_.close()
    return out, com


def _static(lib, dev_state, com, out):
    """The new entry point on caller-owned device tensors, on torch's current stream."""
    import torch

    x, v, w = dev_state
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.lg_diag_sums_static(x.shape[0], x.data_ptr(), v.data_ptr(), w.data_ptr(), MASS, INERTIA, 0, None, out.data_ptr(), st)
    assert rc == LG_OK, lib.lg_last_error()
    rc = lib.lg_diag_sums_static(x.shape[0], x.data_ptr(), None, None, MASS, INERTIA, 1, com.data_ptr(), out.data_ptr(), st)
    assert rc == LG_OK, lib.lg_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _close(got, want):
    # 1e-10 relative to each sum, with an absolute floor for the sums that cancel (sum v, sum x are ~ sqrt(n) * scale)
    floor = 1e-10 * np.array([want[0]] * 2 + [np.abs(want[2:5]).max()] * 3 + [np.abs(want[5:8]).max()] * 3 +
                             [np.abs(want[8:11]).max()] * 3 + [want[11]])
    return np.all(np.abs(got - want) <= 1e-10 * np.abs(want) + floor)


@pytest.mark.gpu
def test_static_scratch_sums_match_the_system_entry_point_twice_in_a_row(lib):
    import torch

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    big, small = _state(150_000, 11), _state(1_000, 12)  # 512 blocks (grid-stride loop), then 4 blocks
    want_big, com_big = _existing(lib, big)
    want_small, com_small = _existing(lib, small)

    def to_dev(state):
        x, v, _, w = state
        return tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x, v, w))

    d_big, d_small = to_dev(big), to_dev(small)
    out = torch.full((12,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    got_big = _static(lib, d_big, com_big, out)  # first call: resolves the array's device address
    assert _close(got_big, want_big), f"first call\n got {got_big}\nwant {want_big}"

    out.fill_(float("nan"))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    got_small = _static(lib, d_small, com_small, out)  # second call: the array still holds the first call's 512 partials
    free_after = torch.cuda.mem_get_info(0)[0]
    assert _close(got_small, want_small), f"second call\n got {got_small}\nwant {want_small}"
    assert free_after >= free_before, "lg_diag_sums_static must not allocate device memory"

    # and once more on the large input: the result does not depend on what the previous call left behind
    out.fill_(float("nan"))
    again = _static(lib, d_big, com_big, out)
    assert np.array_equal(again, got_big)

    # a host pointer is refused with a message instead of faulting inside the kernel
    host = np.zeros((16, 3))
    rc = lib.lg_diag_sums_static(16, host.ctypes.data, host.ctypes.data, host.ctypes.data, MASS, INERTIA, 0, None,
                                 out.data_ptr(), None)
    assert rc == LG_EINVAL and "device memory" in lib.lg_last_error().decode()
