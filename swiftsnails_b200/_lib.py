"""ctypes bindings for the native libraries (C ABI in ``include/ssb200.h``).

The device library is loaded from ``swiftsnails_b200/lib`` (built in-tree by
``swiftsnails_b200.build``).  If a CUDA device is present and the library is missing, loading
fails loudly — there is no silent PyTorch fallback for the hot ops.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_HERE, "lib")

MAX_SHARDS = 16
EMPTY_KEY = 0xFFFFFFFF
LOCKED_KEY = 0xFFFFFFFE
INVALID_SLOT = 0xFFFFFFFF
BUCKET = 32

_lock = threading.Lock()
_kernels = None
_core = None

# number of kernel launches issued through this module (bench.py reports it as gpu_launches)
launch_count = 0


class SsbTableView(C.Structure):
    _fields_ = [
        ("keys", C.c_void_p * MAX_SHARDS),
        ("rows", C.c_void_p * MAX_SHARDS),
        ("state", C.c_void_p * MAX_SHARDS),
        ("shadow", C.c_void_p * MAX_SHARDS),
        ("frag_map", C.c_void_p),
        ("frag_num", C.c_uint32),
        ("capacity", C.c_uint32),
        ("dim", C.c_int32),
        ("num_shards", C.c_int32),
        ("seed", C.c_uint64),
        ("init_scale", C.c_float),
        ("state_words", C.c_int32),
        ("init_kind", C.c_int32),
        ("pad_", C.c_int32),
    ]


class SsbInboxView(C.Structure):
    _fields_ = [
        ("rows", C.c_void_p * MAX_SHARDS),
        ("slots", C.c_void_p * MAX_SHARDS),
        ("count", C.c_void_p * MAX_SHARDS),
        ("flags", C.c_void_p * MAX_SHARDS),
        ("capacity", C.c_uint32),
        ("pad_", C.c_uint32),
    ]


class SsbChunkInboxView(C.Structure):
    _fields_ = [
        ("rows", C.c_void_p * MAX_SHARDS),
        ("slots", C.c_void_p * MAX_SHARDS),
        ("ready", C.c_void_p * MAX_SHARDS),
        ("total", C.c_void_p * MAX_SHARDS),
        ("seg_rows", C.c_uint32),
        ("max_chunks", C.c_uint32),
    ]


_vp, _i64, _i32, _u32, _u64, _f32 = C.c_void_p, C.c_int64, C.c_int32, C.c_uint32, C.c_uint64, C.c_float

# name -> (argtypes, kernel launches per call)
_KERNEL_SIGS = {
    "ssb_fmix64": ([_vp, _vp, _i64, _vp], 1),
    "ssb_hash_route": ([_vp, _i64, _vp, _u32, _i32, _vp, _vp, _vp], 1),
    "ssb_bucket_scatter": ([_vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp], 1),
    "ssb_dedup": ([_vp, _i64, _vp, _vp, _u32, _u32, _vp, _vp, _vp, _i32, _vp], 2),
    "ssb_pull": ([C.POINTER(SsbTableView), _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp], 1),
    "ssb_pull_cached": ([C.POINTER(SsbTableView), _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp], 1),
    "ssb_gather_slots": ([C.POINTER(SsbTableView), _vp, _i64, _vp, _vp, _vp], 1),
    "ssb_push_sgd": ([C.POINTER(SsbTableView), _vp, _vp, _i64, _vp, _f32, _i32, _vp], 1),
    "ssb_push_adagrad": ([C.POINTER(SsbTableView), _vp, _vp, _i64, _vp, _f32, _f32, _i32, _vp], 1),
    "ssb_push_send": ([C.POINTER(SsbTableView), C.POINTER(SsbInboxView), _vp, _vp, _i64, _vp, _i32, _f32, _f32, _i32, _i32, _vp, _vp], 1),
    "ssb_push_send2": ([C.POINTER(SsbTableView), C.POINTER(SsbInboxView), _vp, _vp, _i64, _vp, _i32, _f32, _f32, _i32, _i32, _vp, _vp, _vp], 1),
    "ssb_push_send_chunks": ([C.POINTER(SsbTableView), C.POINTER(SsbChunkInboxView), _vp, _vp, _i64, _vp, _i32, _f32, _f32, _i32, _u32, _vp, _vp, _vp, _vp, _vp], 1),
    "ssb_push_apply_chunks": ([C.POINTER(SsbTableView), _vp, _vp, _vp, _vp, _u32, _i32, _i32, _u32, _f32, _f32, _i32, _u32, _vp, _vp], 1),
    "ssb_inbox_publish": ([C.POINTER(SsbInboxView), _i32, _i32, _vp, _u32, _vp], 1),
    "ssb_push_bf16_seg": ([C.POINTER(SsbTableView), _vp, _vp, _i64, _i32, _vp, _f32, _f32, _i32, _vp], 1),
    "ssb_push_bf16": ([C.POINTER(SsbTableView), _vp, _vp, _i64, _vp, _f32, _f32, _i32, _vp], 1),
    "ssb_flags_signal": ([C.POINTER(SsbInboxView), _i32, _i32, _u32, _vp], 1),
    "ssb_flags_wait": ([_vp, _i32, _u32, _vp], 1),
    "ssb_flags_wait2": ([_vp, _i32, _u32, _u32, _vp, _vp], 1),
    "ssb_dump_compact_range": ([_vp, _vp, _vp, _u32, _i32, _u32, _u32, _vp, _vp, _vp, _vp, _u32, _i32, _vp], 1),
    "ssb_red_probe": ([_vp, _u32, _i32, _i32, _i32, _vp], 1),
    "ssb_build_keys": ([_vp, _i64, _u32, _vp, _vp], 1),
    "ssb_neg_sample": ([_vp, _vp, _u32, _u64, _u64, _u32, _vp, _i64, _vp], 1),
    "ssb_sgns_make_pairs": ([_vp, _vp, _i64, _i32, _i32, _u64, _u64, _vp, _vp, _vp, _vp], 1),
    "ssb_sgns_tiles": ([_vp, _vp, _i32, _vp, _vp, _vp, _i64, _vp, _i32, _f32, _vp, _vp], 1),
    "ssb_sgns_tc_set_ctas": ([_i32], 0),
    "ssb_sgns_block_tc": ([_vp, _vp, _i64, _i32, _vp, _vp, _i64, _i32, _i32, _u64, _u64, _vp, _f32, _vp, _i32, _i32, _vp, _vp], 1),
    "ssb_sgns_block_tc_direct": ([_vp, _u64, _u32, _u32, _u32, _vp, _vp, _i32, _vp, _vp, _i64, _i32, _i32, _u64, _u64, _vp, _f32,
                                  _vp, _i32, _vp], 1),
    "ssb_dedup_static": ([_vp, _i64, _vp, _vp, _vp, _vp], 2),
    "ssb_lr_fwd_bwd": ([_vp, _vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp], 1),
    "ssb_lr_fused_step": ([C.POINTER(SsbTableView), _vp, _vp, _vp, _i64, _i32, _vp, _u64, _vp, _vp, _vp, _vp, _u32, _vp, _u32, _vp,
                           _i32, _vp, _i32, _vp], 1),
    "ssb_dump_compact": ([_vp, _vp, _u32, _i32, _vp, _vp, _vp, _u32, _vp], 1),
    "ssb_table_load": ([C.POINTER(SsbTableView), _vp, _vp, _vp, _i64, _vp, _vp], 1),
    "ssb_p2p_copy": ([_vp, _vp, _i64, _i32, _vp], 1),
    "ssb_fill_u32": ([_vp, _u32, _i64, _vp], 1),
}


class NativeError(RuntimeError):
    pass


def kernels_path():
    return os.path.join(LIB_DIR, "libssb200_kernels.so")


def core_path():
    return os.path.join(LIB_DIR, "libssb200_core.so")


def kernels():
    """Load (once) and return the CUDA kernel library."""
    global _kernels
    if _kernels is not None:
        return _kernels
    with _lock:
        if _kernels is not None:
            return _kernels
        path = kernels_path()
        if not os.path.exists(path):
            # in-tree build on first use (nvcc is part of the image); never a PyTorch fallback
            try:
                from . import build as _build

                _build.build_kernels()
            except Exception as e:  # noqa: BLE001
                raise NativeError("native kernel library missing and the in-tree build failed: %s" % (e,))
        if not os.path.exists(path):
            raise NativeError(
                "native kernel library missing: %s — run `python -m swiftsnails_b200.build` "
                "(there is no PyTorch fallback for the hot ops)" % path)
        lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
        for name, (argtypes, _) in _KERNEL_SIGS.items():
            if not hasattr(lib, name):
                continue
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = C.c_int
        lib.ssb_error_string.argtypes = [C.c_int]
        lib.ssb_error_string.restype = C.c_char_p
        lib.ssb_version.restype = C.c_int
        _kernels = lib
    return _kernels


def call(name, *args):
    """Call a kernel entry point, raise on a CUDA error, count launches."""
    global launch_count
    lib = kernels()
    err = getattr(lib, name)(*args)
    if err != 0:
        raise NativeError("%s failed: cudaError %d (%s)" % (name, err, lib.ssb_error_string(err).decode()))
    launch_count += _KERNEL_SIGS.get(name, (None, 1))[1]


def register_sig(name, argtypes, launches=1, restype=C.c_int):
    """Register a late-added entry point (e.g. from an optional translation unit)."""
    _KERNEL_SIGS[name] = (argtypes, launches)
    lib = kernels()
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


def has_symbol(name):
    try:
        return hasattr(kernels(), name)
    except NativeError:
        return False


def core():
    """Load (once) and return the C++ host core library (CPU, no CUDA needed)."""
    global _core
    if _core is not None:
        return _core
    with _lock:
        if _core is not None:
            return _core
        path = core_path()
        if not os.path.exists(path):
            try:
                from . import build as _build

                _build.build_core()
            except Exception as e:  # noqa: BLE001
                raise NativeError("host core library missing and the in-tree build failed: %s" % (e,))
        if not os.path.exists(path):
            raise NativeError("host core library missing: %s — run `python -m swiftsnails_b200.build`" % path)
        _core = C.CDLL(path, mode=C.RTLD_GLOBAL)
    return _core


def ptr(t):
    """Device/host pointer of a torch tensor (or None)."""
    return None if t is None else t.data_ptr()


def stream_ptr(stream=None):
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream
