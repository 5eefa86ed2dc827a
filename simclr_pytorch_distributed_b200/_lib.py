"""Loader for libsimclr_b200.so, the C-ABI library declared in include/simclr_b200.h.

The library is built in-tree (csrc/libsimclr_b200.so) by `build()` with nvcc for sm_100a and
opened with ctypes.  There is no fallback of any kind: if the library cannot be built or
loaded, `load()` raises, and every compute entry point raises when it fails.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import threading
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, 'csrc')
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(CSRC, 'libsimclr_b200.so')
_DEFAULT_LIB_PATH = LIB_PATH
SOURCES = ['api.cu', 'bn.cu', 'bn_cluster.cu', 'bn_ring.cu', 'head_tc.cu', 'supcon_generic.cu', 'supcon_tc.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr']

# dtype / enum codes of include/simclr_b200.h
F32, BF16, F16 = 0, 1, 2
MASK_SIMCLR, MASK_LABELS, MASK_EXPLICIT = 0, 1, 2
MODE_ALL, MODE_ONE = 0, 1
ENGINE_FP32, ENGINE_BF16, ENGINE_BF16X3 = 0, 1, 2
ROWS_VIEW_MAJOR, ROWS_SAMPLE_MAJOR = 0, 1
BN_RELU = 1
BN_ADD = 2
BN_PDL = 4
BN_NOALLOC_MAX_CHANNELS = 4096      # SCB_BN_NOALLOC_MAX_CHANNELS

_lock = threading.Lock()
_lib = None


def _nvcc():
    for cand in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    return None


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cu', '.cuh'))]
    deps.append(os.path.join(ROOT, 'include', 'simclr_b200.h'))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into csrc/libsimclr_b200.so (no GPU needed)."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError('libsimclr_b200.so is missing or stale and nvcc was not found; '
                           'build it with `python -c "import __graft_entry__ as g; g.build()"`')
    # One builder at a time: the ranks of a torchrun job import this module together, and two compilers writing the
    # same object files would corrupt them.  Whoever gets the lock second finds the library fresh and returns.
    import fcntl
    with open(os.path.join(CSRC, '.build.lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():
                return LIB_PATH
            return _build_locked(nvcc, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(nvcc, verbose):
    objs = []

    def compile_one(src):
        obj = os.path.join(CSRC, src[:-3] + '.o')
        cmd = [nvcc] + NVCC_FLAGS + ['-c', os.path.join(CSRC, src), '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s\n%s' % (src, r.stdout, r.stderr))
        if verbose and r.stderr:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    tmp = LIB_PATH + '.tmp.%d' % os.getpid()
    r = subprocess.run([nvcc, '-shared', '-o', tmp] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a'],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


class PeerChannel(ctypes.Structure):
    """`scb_peer_channel` of include/simclr_b200.h."""
    _fields_ = [('dst', ctypes.c_void_p * 16), ('flag', ctypes.c_void_p * 16), ('state', ctypes.c_void_p),
                ('n_dst', ctypes.c_int32)]


class SupconProblem(ctypes.Structure):
    """`scb_supcon_problem` of include/simclr_b200.h."""
    _fields_ = [
        ('k', ctypes.c_void_p), ('k_lo', ctypes.c_void_p),
        ('labels', ctypes.c_void_p), ('mask', ctypes.c_void_p),
        ('k_dtype', ctypes.c_int32), ('n_k', ctypes.c_int32), ('dim', ctypes.c_int32),
        ('ldk', ctypes.c_int32), ('n_views', ctypes.c_int32), ('seg', ctypes.c_int32),
        ('n_seg', ctypes.c_int32), ('mask_kind', ctypes.c_int32), ('mode', ctypes.c_int32),
        ('engine', ctypes.c_int32), ('scalar_dtype', ctypes.c_int32), ('inv_temperature', ctypes.c_float),
        ('loss_scale', ctypes.c_float), ('row_layout', ctypes.c_int32), ('label_cache', ctypes.c_void_p),
    ]


_vp, _i, _i64, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
_pp = ctypes.POINTER(SupconProblem)

# name -> (restype, argtypes): exactly the declarations of include/simclr_b200.h
PROTOTYPES = {
    'scb_last_error': (ctypes.c_char_p, []),
    'scb_version': (_i, []),
    'scb_device_cc': (_i, []),
    'scb_launch_count': (_i64, []),
    'scb_stream_capture_id': (_i64, [_vp]),
    'scb_debug_set_trace': (_i, [_vp]),
    'scb_debug_set_events': (_i, [_vp, _vp]),
    'scb_supcon_prep': (_i, [_vp, _i, _i, _i, _i, _i64, _i64, _i, _f, _vp, _vp, _i, _i, _vp, _vp]),
    'scb_supcon_workspace_bytes': (_i64, [_pp, _i]),
    'scb_supcon_label_cache_bytes': (_i64, [_pp]),
    'scb_supcon_forward': (_i, [_pp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    'scb_supcon_backward': (_i, [_pp, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _vp, _i64, _vp]),
    'scb_supcon_backward_features': (_i, [_pp, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i64, _i64, _i, _vp, _vp, _i64,
                                          _i64, _vp, _i64, _vp]),
    'scb_supcon_finish': (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _i64, _i64, _vp]),
    'scb_feature_norm_terms': (_i, [_vp, _i, _vp, _f, _f, _f, _vp, _vp]),
    'scb_feature_norm_backward': (_i, [_vp, _i, _i, _i, _i, _i64, _i64, _vp, _vp, _f, _f, _vp, _i, _vp, _i64, _i64, _vp]),
    'scb_head_forward': (_i, [_vp, _i64, _vp, _i64, _vp, _i64, _i, _i, _i, _vp, _i, _i64, _vp, _i, _f, _vp]),
    'scb_head_normalize_backward': (_i, [_vp, _i, _i64, _vp, _i, _i64, _vp, _i64, _i, _vp, _i, _i64, _i, _vp]),
    'scb_peer_bcast': (_i, [_vp, _vp, _i, _vp, _i, _vp]),
    'scb_supcon_prep_peers': (_i, [_vp, _i, _i, _i, _i, _i64, _i64, _i, _f, _vp, _i64, _i, _i, _vp, _vp]),
    'scb_supcon_forward_peers': (_i, [_pp, _i, _i, _i, _vp, _vp, _vp, _i64, _vp]),
    'scb_peer_wait': (_i, [_vp, _vp, _i, _vp, _vp, _i, _vp]),
    'scb_bn_workspace_bytes': (_i64, [_i]),
    'scb_bn_stats': (_i, [_vp, _i, _i, _i64, _i, _i64, _vp, _vp, _i64, _vp]),
    'scb_bn_stats_noalloc': (_i, [_vp, _i, _i, _i64, _i, _i64, _vp, _vp]),
    'scb_bn_apply': (_i, [_vp, _vp, _vp, _i, _i, _i64, _i, _i64, _vp, _i, _vp, _vp, _vp, _vp, _f, _f,
                          _vp, _vp, _vp, _vp, _i, _i, _vp]),
    'scb_bn_backward_reduce': (_i, [_vp, _vp, _vp, _i, _i, _i64, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                                    _vp, _i, _i, _vp, _i64, _vp]),
    'scb_bn_backward_apply': (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i64, _i, _i64, _vp, _vp, _vp, _vp, _vp,
                                   _vp, _i, _i, _vp]),
    'scb_bn_fused_supported': (_i, [_i, _i, _i64, _i, _i64, _i, _i]),
    'scb_bn_cluster_occupancy': (_i, [_i]),
    'scb_bn_forward_fused': (_i, [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _i, _i,
                                  _vp, _i64, _vp]),
    'scb_bn_backward_fused': (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                   _i, _i, _vp, _i64, _vp]),
    'scb_bn_stats_fused': (_i, [_vp, _i, _i64, _i, _vp, _vp]),
    'scb_bn_backward_reduce_fused': (_i, [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i,
                                          _i, _vp]),
}


def load():
    """Return the ctypes handle of the library (building it first if it is stale)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        # a library put in place by hand (tools/build_trace_lib.sh points LIB_PATH at its instrumented build) is
        # loaded as it is: rebuilding it here would silently replace it by the plain one
        if LIB_PATH == _DEFAULT_LIB_PATH and _stale():
            build()
        try:
            lib = ctypes.CDLL(LIB_PATH)
        except OSError as e:   # no silent fallback
            raise RuntimeError('cannot load %s: %s' % (LIB_PATH, e)) from e
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)      # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str = ''):
    """Turn a negative return code into an exception carrying scb_last_error()."""
    if rc == 0:
        return
    msg = load().scb_last_error().decode('utf-8', 'replace')
    if rc == -1:
        raise ValueError('%s: %s' % (what, msg) if what else msg)
    raise RuntimeError('%s failed (code %d): %s' % (what or 'libsimclr_b200', rc, msg))


def dtype_code(dtype):
    import torch
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    if dtype == torch.float16:
        return F16
    raise TypeError('unsupported dtype %s (float32, bfloat16 and float16 are supported)' % dtype)
