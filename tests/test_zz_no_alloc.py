"""vf_dem_minmax_noalloc_async (include/vf_cuda.h): the allocation-free variant of vf_dem_minmax, through the C ABI.

Its scratch is a static device array inside libvfcuda.so that every call shares and that keeps the previous call's
contents, so the checks run on calls IN A ROW whose sizes and contents differ: a stale accumulator or a stale sample would
show.  The bar is the one of tests/test_gpu_parity.py for min / max / percentiles: bit-exact against the existing entry
point (vf_dem_minmax on the same values in host memory).

Kept in a file of its own that pytest collects last; passes alone in a fresh process
and after the other tests.  Device buffers come from torch (test plumbing only: the entry point itself allocates nothing).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(vf_build):
    L = C.CDLL(vf_build.lib_path())
    L.vf_last_error.restype = C.c_char_p
    L.vf_dem_minmax_noalloc_async.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
    L.vf_dem_minmax.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    assert L.vf_init(-1) == 0, L.vf_last_error()
    return L


def _existing(lib, a, clamp):
    lo, hi = C.c_float(), C.c_float()
    assert lib.vf_dem_minmax(a.ctypes.data, a.size, clamp, C.byref(lo), C.byref(hi)) == 0, lib.vf_last_error()
    return np.array([lo.value, hi.value], np.float32)


def _noalloc(lib, torch, a, clamp, stream):
    d = torch.from_numpy(a).cuda()
    out = torch.full((2,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()                       # the uploads above ran on torch's stream
    rc = lib.vf_dem_minmax_noalloc_async(d.data_ptr(), a.size, clamp, out.data_ptr(), stream)
    assert rc == 0, lib.vf_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_noalloc_minmax_equals_existing_entry_point_on_calls_in_a_row(lib):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(11)
    big = (rng.normal(size=300_000) * rng.choice([0.1, 1.0, 40.0], size=300_000)).astype(np.float32)      # step 4, 75 000 samples
    mid = rng.normal(size=131_071).astype(np.float32)                                                    # step 1: the largest sample there is
    small = rng.uniform(5.0, 6.0, size=1000).astype(np.float32)                                           # far fewer samples than the call before
    zeros = np.abs(rng.normal(size=70_001)).astype(np.float32) + np.float32(0.25)
    zeros[4321] = -0.0                                                                                    # min is a zero: its sign is that of the first zero
    zeros[60_000] = 0.0
    one = np.array([-7.5], np.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rep in range(2):                                                                                  # the whole sequence twice, back to back
        for name, a in (("big", big), ("small", small), ("mid", mid), ("zeros", zeros), ("one", one), ("big", big)):
            for clamp in (1, 0, 1):
                want = _existing(lib, a, clamp)
                got = _noalloc(lib, torch, a, clamp, stream)
                assert got.tobytes() == want.tobytes(), (rep, name, clamp, got, want)
    lo0 = _noalloc(lib, torch, zeros, 0, stream)[0]
    assert lo0 == 0.0 and np.signbit(lo0)


def test_noalloc_rejects_bad_sizes_without_touching_memory(lib):
    torch = pytest.importorskip("torch")
    buf = torch.zeros(4, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.float32, device="cuda")
    assert lib.vf_dem_minmax_noalloc_async(buf.data_ptr(), 0, 1, out.data_ptr(), None) != 0
    assert b"empty" in lib.vf_last_error()
    assert lib.vf_dem_minmax_noalloc_async(buf.data_ptr(), 1 << 32, 1, out.data_ptr(), None) != 0
    assert b"exceeds the maximum" in lib.vf_last_error()
    assert lib.vf_dem_minmax_noalloc_async(None, 4, 1, out.data_ptr(), None) != 0
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0           # nothing was launched
