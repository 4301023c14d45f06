"""ctypes binding of libdxgpu.so (include/dxgpu.h).

The CUDA library is the product; there is no Python or CPU implementation behind it.  If the
shared object is missing or cannot be loaded, everything here raises — by design.
"""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
# DXGPU_LIB: development override used by the tuning sweeps (tools/build_variant.sh) to load a differently tuned build
# of the same library; unset, the in-tree build is the only thing ever loaded
LIB_PATH = os.environ.get("DXGPU_LIB") or os.path.join(_HERE, "libdxgpu.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "dxgpu.h")

_lib = None


class DxError(RuntimeError):
    """Mirror of the reference's EngineException (datax-core exception/EngineException.scala:7-8)."""


class GenParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_keys", C.c_int64), ("n_homes", C.c_int64), ("base_ms", C.c_int64),
                ("interval_ms", C.c_int64), ("watermark_ms", C.c_int64), ("shape", C.c_int32),
                ("late_permille", C.c_int32), ("future_permille", C.c_int32), ("pad", C.c_int32)]


def declared_symbols():
    """Every function name declared in include/dxgpu.h."""
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dx_[a-z0-9_]+)\s*\(", text)))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DxError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i64, i32, cp = C.c_void_p, C.c_int64, C.c_int, C.c_char_p
    P = C.POINTER
    sig = {
        "dx_ctx_create": (i32, [i32, P(vp)]),
        "dx_ctx_destroy": (None, [vp]),
        "dx_ctx_last_error": (cp, [vp]),
        "dx_ctx_set_timing": (i32, [vp, i32]),
        "dx_ctx_take_stage_ms": (i32, [vp, cp, P(C.c_double), P(i64)]),
        "dx_ctx_take_launches": (i64, [vp]),
        "dx_ctx_stream": (vp, [vp]),
        "dx_ctx_synchronize": (i32, [vp]),
        "dx_dev_alloc": (i32, [vp, i64, P(vp)]),
        "dx_dev_free": (None, [vp, vp]),
        "dx_copy_h2d": (i32, [vp, vp, vp, i64]),
        "dx_copy_d2h": (i32, [vp, vp, vp, i64]),
        "dx_decoder_create": (i32, [vp, cp, P(vp)]),
        "dx_decoder_destroy": (None, [vp]),
        "dx_decoder_decode": (i32, [vp, vp, i64, vp, i64, i32, P(vp)]),
        "dx_decoder_stats": (i32, [vp, P(i64), P(i64), P(i64)]),
        "dx_table_num_rows": (i64, [vp]),
        "dx_table_num_columns": (i32, [vp]),
        "dx_table_column_name": (cp, [vp, i32]),
        "dx_table_column_type": (i32, [vp, i32]),
        "dx_table_column_nesting": (i32, [vp, i32, P(cp)]),
        "dx_table_column_path": (cp, [vp, i32, i32]),
        "dx_table_find_column": (i32, [vp, cp]),
        "dx_table_flag_words": (i32, [vp, i32]),
        "dx_table_column_device": (i32, [vp, i32, P(vp), P(vp), P(i32), P(i32)]),
        "dx_table_column_to_host": (i32, [vp, i32, vp, vp]),
        "dx_table_dict_size": (i32, [vp, i32, P(i64)]),
        "dx_table_dict_string": (i32, [vp, i32, i64, P(vp), P(i64)]),
        "dx_table_release": (None, [vp]),
        "dx_gen_events": (i32, [vp, P(GenParams), i64, i64, P(vp), P(i64), P(vp)]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(L, name):
            continue  # optional until declared; test_abi checks header <-> exports
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def bind(name, restype, argtypes):
    """Late binding for entry points added after the table above."""
    fn = getattr(lib(), name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn
