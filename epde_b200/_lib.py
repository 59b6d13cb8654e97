"""ctypes binding of libepde_b200.so (C ABI declared in include/epde_b200.h).

There is no CPU fallback: if the library is missing, cannot be built or cannot be loaded the
import fails with a clear error.
"""
import ctypes as C
import os

from . import build as _build

c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_f64p = C.POINTER(C.c_double)
vp = C.c_void_p

# name -> (restype, argtypes); must list every symbol of include/epde_b200.h
_ENC = [C.c_int, c_i32p, c_i32p, c_i32p, c_f64p, c_i32p, c_f64p, C.c_int, c_i32p, c_i32p]
_OUT = [c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_i32p, c_i32p, c_i32p]
SIGNATURES = {
    'epde_last_error': (C.c_char_p, []),
    'epde_version': (C.c_char_p, []),
    'epde_ctx_create': (C.c_int, [C.c_int, C.c_int, c_i64p, C.POINTER(vp)]),
    'epde_ctx_destroy': (C.c_int, [vp]),
    'epde_ctx_set_stream': (C.c_int, [vp, vp]),
    'epde_ctx_set_shard': (C.c_int, [vp, C.c_int64, C.c_int64]),
    'epde_cache_upload': (C.c_int, [vp, C.c_int, c_f64p, C.c_int64]),
    'epde_cache_upload_device': (C.c_int, [vp, C.c_int, vp, C.c_int64]),
    'epde_cache_download': (C.c_int, [vp, C.c_int, c_f64p, C.c_int64]),
    'epde_cache_contains': (C.c_int, [vp, C.c_int, C.POINTER(C.c_int)]),
    'epde_cache_delete': (C.c_int, [vp, C.c_int]),
    'epde_cache_bytes': (C.c_int, [vp, c_i64p]),
    'epde_ctx_set_sample_weights': (C.c_int, [vp, c_f64p, C.c_int64]),
    'epde_ctx_num_windows': (C.c_int, [vp, C.POINTER(C.c_int)]),
    'epde_ctx_num_segments': (C.c_int, [vp, C.POINTER(C.c_int)]),
    'epde_evaluate_batch': (C.c_int, [vp] + _ENC + _OUT),
    'epde_batch_create': (C.c_int, [vp] + _ENC + [C.POINTER(vp)]),
    'epde_batch_destroy': (C.c_int, [vp]),
    'epde_batch_set_given': (C.c_int, [vp, c_i32p, c_f64p, c_f64p]),
    'epde_batch_set_norms': (C.c_int, [vp, C.c_int]),
    'epde_batch_info': (C.c_int, [vp, c_i64p]),
    'epde_batch_gram_size': (C.c_int, [vp, c_i64p]),
    'epde_batch_residual_size': (C.c_int, [vp, c_i64p]),
    'epde_batch_gram': (C.c_int, [vp, vp]),
    'epde_batch_solve': (C.c_int, [vp, vp]),
    'epde_batch_residual': (C.c_int, [vp, vp]),
    'epde_batch_solution_size': (C.c_int, [vp, C.c_int, c_i64p]),
    'epde_batch_solve_part': (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    'epde_batch_set_solution': (C.c_int, [vp, vp, C.c_int]),
    'epde_batch_fetch': (C.c_int, [vp, vp] + _OUT),
    'epde_ctx_synchronize': (C.c_int, [vp]),
    'epde_fd_gradient': (C.c_int, [vp, vp, vp, C.c_int, c_i64p, C.c_int, c_f64p]),
    'epde_cache_upload_interior': (C.c_int, [vp, C.c_int, vp, C.c_int, c_i64p, c_i64p]),
    'epde_poly_derivative': (C.c_int, [vp, vp, vp, C.c_int, c_i64p, C.c_int, C.c_int, C.c_int, c_f64p, c_f64p]),
    'epde_poly_derivative_noalloc': (C.c_int, [vp, vp, vp, C.c_int, c_i64p, C.c_int, C.c_int, C.c_int, c_f64p, c_f64p]),
    'epde_measure_fp64_peak': (C.c_int, [C.c_int, C.c_double, c_f64p, c_f64p]),
    'epde_ctx_launch_count': (C.c_int, [vp, c_i64p]),
    'epde_ctx_set_timing': (C.c_int, [vp, C.c_int]),
    'epde_ctx_kernel_time': (C.c_int, [vp, C.c_int, c_f64p, c_i64p]),
}

_lib = None
ABI_VERSION = 'epde_b200 0.3.0'      # must be a prefix of epde_version() (csrc/epde_b200.cu)


class EpdeB200Error(RuntimeError):
    pass


def library_path():
    return _build.LIB


def load():
    """Load (building first when stale and nvcc is present) and return the ctypes library."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if _build.is_stale():
        try:
            _build.build()
        except Exception as exc:          # noqa: BLE001
            # never bind a stale binary against newer ctypes signatures: argument lists may have changed
            state = 'missing' if not os.path.exists(path) else 'older than its sources'
            raise EpdeB200Error(
                f'epde_b200: CUDA library {path} is {state} and could not be built ({exc}); '
                'there is no CPU fallback') from exc
    try:
        lib = C.CDLL(path)
    except OSError as exc:
        raise EpdeB200Error(f'epde_b200: cannot load {path}: {exc}; there is no CPU fallback') from exc
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError when a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    got = lib.epde_version().decode('utf-8', 'replace')
    if ABI_VERSION not in got:
        raise EpdeB200Error(f'epde_b200: {path} reports "{got}" but this package binds ABI "{ABI_VERSION}"; '
                            'rebuild with python -m epde_b200.build --force')
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise EpdeB200Error(load().epde_last_error().decode('utf-8', 'replace'))
