"""`ssb_dedup_static`: de-duplication for callers that are not allowed to allocate device memory
(a host process that owns every byte of HBM through its own pool and hands the library raw
buffers, or a latency-critical path where cudaMalloc would stall the device).  Its scratch hash
set is a fixed array inside the library instead of a caller-provided buffer.

The entry point is driven through the C ABI exactly as such a caller would (raw pointers, raw
stream) and compared with the ordinary `ssb_dedup` on the same keys — twice in a row, the second
time with a smaller, different batch, because every call shares the one built-in array and must
cope with what the previous call left behind.  Collected last (file name) so a run that stops at
the first failure has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_KEYS = 65536  # SSB_DEDUP_STATIC_MAX_KEYS in include/ssb200.h


def _lib_handle():
    from swiftsnails_b200 import _lib

    lib = C.CDLL(_lib.kernels_path())
    vp, i64, u32 = C.c_void_p, C.c_int64, C.c_uint32
    lib.ssb_dedup_static.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.ssb_dedup_static.restype = C.c_int
    lib.ssb_dedup.argtypes = [vp, i64, vp, vp, u32, u32, vp, vp, vp, C.c_int, vp]
    lib.ssb_dedup.restype = C.c_int
    return lib


def _run_both(lib, keys_np, device, epoch):
    n = len(keys_np)
    stream = torch.cuda.current_stream().cuda_stream
    keys = torch.from_numpy(keys_np.view(np.int32)).to(device)
    out = {}
    # the new entry point: inputs and outputs only
    uniq = torch.full((n,), -7, dtype=torch.int32, device=device)
    inv = torch.full((n,), -7, dtype=torch.int32, device=device)
    cnt = torch.full((1,), 12345, dtype=torch.int32, device=device)   # stale value: the call must reset it
    rc = lib.ssb_dedup_static(keys.data_ptr(), n, uniq.data_ptr(), inv.data_ptr(), cnt.data_ptr(), stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    out["static"] = (uniq.cpu().numpy().view(np.uint32), inv.cpu().numpy().view(np.uint32), int(cnt.item()))
    # the existing entry point with caller-provided scratch
    cap = 1 << 18
    sset = torch.zeros(cap, dtype=torch.int64, device=device)
    sids = torch.empty(cap, dtype=torch.int32, device=device)
    uniq2 = torch.empty(n, dtype=torch.int32, device=device)
    inv2 = torch.empty(n, dtype=torch.int32, device=device)
    cnt2 = torch.zeros(1, dtype=torch.int32, device=device)
    rc = lib.ssb_dedup(keys.data_ptr(), n, sset.data_ptr(), sids.data_ptr(), cap, epoch, uniq2.data_ptr(),
                       inv2.data_ptr(), cnt2.data_ptr(), 1, stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    out["ref"] = (uniq2.cpu().numpy().view(np.uint32), inv2.cpu().numpy().view(np.uint32), int(cnt2.item()))
    return out


def _check(keys_np, res):
    valid = keys_np != 0xFFFFFFFF
    (u1, i1, n1), (u2, i2, n2) = res["static"], res["ref"]
    assert n1 == n2 == len(np.unique(keys_np[valid]))
    # ids are handed out in a race-dependent order in both entry points: compare as sets and
    # through the reconstruction unique[inverse] == keys (exact, integers)
    assert np.array_equal(np.sort(u1[:n1]), np.sort(u2[:n2]))
    assert np.array_equal(u1[i1[valid]], keys_np[valid])
    assert np.array_equal(u2[i2[valid]], keys_np[valid])
    assert np.all(i1[~valid] == 0xFFFFFFFF) and np.all(i2[~valid] == 0xFFFFFFFF)


def test_static_scratch_dedup_matches_ssb_dedup_twice(device):
    lib = _lib_handle()
    rng = np.random.default_rng(7)
    # call 1: a full-size batch with heavy duplication and a few padding keys
    k1 = rng.integers(0, 20_000, size=MAX_KEYS, dtype=np.int64).astype(np.uint32)
    k1[::997] = 0xFFFFFFFF
    _check(k1, _run_both(lib, k1, device, epoch=1))
    # call 2, straight after: fewer keys from a disjoint range — anything the first call left in
    # the shared scratch must not leak into this result
    k2 = (rng.integers(0, 3_000, size=10_001, dtype=np.int64) + 1_000_000).astype(np.uint32)
    _check(k2, _run_both(lib, k2, device, epoch=2))


def test_static_scratch_dedup_rejects_oversize(device):
    lib = _lib_handle()
    n = MAX_KEYS + 1
    keys = torch.zeros(n, dtype=torch.int32, device=device)
    uniq = torch.empty_like(keys)
    inv = torch.empty_like(keys)
    cnt = torch.zeros(1, dtype=torch.int32, device=device)
    rc = lib.ssb_dedup_static(keys.data_ptr(), n, uniq.data_ptr(), inv.data_ptr(), cnt.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 1  # cudaErrorInvalidValue; the reason is printed on stderr
    torch.cuda.synchronize()
