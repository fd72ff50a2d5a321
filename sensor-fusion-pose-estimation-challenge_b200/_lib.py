"""ctypes binding of libsfpe.so (C ABI in include/sfpe.h).

There is no fallback: if the shared library is missing or does not export a
declared symbol, importing the product raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SFPE_LIB: another build of the same library (tuning experiments, tools/variants.sh)
LIB_PATH = os.environ.get("SFPE_LIB") or os.path.join(_HERE, "libsfpe.so")

SFPE_F32, SFPE_F64 = 0, 1
RANGE_CHW, RANGE_HWC = 0, 1
FRAME_ROW_OOB, FRAME_COL_OOB, FRAME_NAN = 1, 2, 4
MAX_SLICES = 8
STATIC_MAX_CELLS = 1 << 18   # sfpe_range_scan_static: largest H*W
ERR_TOO_LARGE = -5
ABI_VERSION = 1


class Params(C.Structure):
    _fields_ = [
        ("res_v_deg", C.c_double), ("res_h_deg", C.c_double),
        ("vfov_lo_deg", C.c_double), ("vfov_hi_deg", C.c_double),
        ("height_bg", C.c_double),
        ("bev_range", C.c_double), ("bev_res_x", C.c_double), ("bev_res_y", C.c_double),
        ("enable_range", C.c_int32), ("enable_bev", C.c_int32),
        ("use_extrinsic", C.c_int32), ("n_slices", C.c_int32),
        ("extrinsic", C.c_double * 12),
        ("slice_edges", C.c_double * (MAX_SLICES + 1)),
        ("key_prefix_bits", C.c_int32), ("path", C.c_int32),
        ("ablation", C.c_int32), ("firing_stride", C.c_int32),
        ("reserved", C.c_int32 * 4),
    ]


class Points(C.Structure):
    _fields_ = [
        ("data", C.c_void_p), ("stride_bytes", C.c_int64), ("dtype", C.c_int32),
        ("off_x", C.c_int32), ("off_y", C.c_int32), ("off_z", C.c_int32), ("off_i", C.c_int32),
        ("reserved", C.c_int32),
    ]


class Outputs(C.Structure):
    _fields_ = [
        ("range", C.c_void_p), ("range_dtype", C.c_int32), ("range_layout", C.c_int32),
        ("range_winner", C.c_void_p),
        ("bev", C.c_void_p), ("bev_dtype", C.c_int32), ("reserved0", C.c_int32),
        ("bev_counts", C.c_void_p), ("bev_height", C.c_void_p), ("bev_intensity", C.c_void_p),
        ("status", C.c_void_p), ("dropped", C.c_void_p),
        ("reserved", C.c_void_p * 4),
    ]


# name -> (restype, argtypes); must list every function include/sfpe.h declares
SIGNATURES = {
    "sfpe_abi_version": (C.c_int, []),
    "sfpe_strerror": (C.c_char_p, [C.c_int]),
    "sfpe_default_params": (None, [C.POINTER(Params)]),
    "sfpe_grid_dims": (C.c_int, [C.POINTER(Params)] + [C.POINTER(C.c_int32)] * 4),
    "sfpe_create": (C.c_int, [C.c_int, C.POINTER(Params), C.POINTER(C.c_void_p)]),
    "sfpe_destroy": (None, [C.c_void_p]),
    "sfpe_last_error": (C.c_char_p, [C.c_void_p]),
    "sfpe_project_batch": (C.c_int, [C.c_void_p, C.POINTER(Points), C.c_void_p, C.c_int64, C.c_int64,
                                     C.c_int32, C.POINTER(Outputs), C.c_void_p]),
    "sfpe_range_scan_static": (C.c_int, [C.c_int, C.POINTER(Params), C.POINTER(Points), C.c_void_p, C.c_int64,
                                         C.c_int64, C.c_int32, C.POINTER(Outputs), C.c_void_p]),
    "sfpe_project_sweep_host": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Points), C.c_int64, C.c_void_p,
                                          C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                          C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "sfpe_decode_pointcloud2": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_int32] * 6 +
                                [C.c_void_p, C.c_void_p]),
    "sfpe_colormap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_double,
                                C.c_double, C.c_void_p, C.c_void_p]),
    "sfpe_colormap_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                     C.c_double, C.c_double, C.c_void_p]),
    "sfpe_invariant_mask_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                                             C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "sfpe_roll_width": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                  C.c_void_p, C.c_void_p]),
    "sfpe_density_lut": (C.c_int, [C.c_void_p, C.c_void_p]),
    "sfpe_bev_pack_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sfpe_bev_unpack_host": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "sfpe_bev_unpack_rgb_host": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32]),
    "sfpe_back_project": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32,
                                    C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double,
                                    C.c_void_p, C.c_void_p, C.c_void_p]),
    "sfpe_find_bbox_3d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_void_p,
                                    C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_double,
                                    C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sfpe_find_obstacle": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sfpe_launch_count": (C.c_int64, [C.c_void_p]),
    "sfpe_last_path": (C.c_char_p, [C.c_void_p]),
    "sfpe_phase_profile": (C.c_int, [C.c_void_p, C.c_int32]),
    "sfpe_phase_cycles": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "sfpe_selftest_fastpath": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_double)]),
}

_lib = None


def load():
    """dlopen libsfpe.so and bind every declared symbol.  Raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            "libsfpe.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C %s/csrc`).  There is no CPU fallback." % (LIB_PATH, _HERE))
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    v = lib.sfpe_abi_version()
    if v != ABI_VERSION:
        raise ImportError("libsfpe.so ABI version %d, binding expects %d" % (v, ABI_VERSION))
    _lib = lib
    return lib


class SfpeError(RuntimeError):
    pass


def check(rc, ctx=None):
    if rc == 0:
        return
    lib = load()
    msg = lib.sfpe_strerror(rc).decode()
    if ctx is not None and rc == -2:
        msg += ": " + lib.sfpe_last_error(ctx).decode()
    if rc == -1:
        raise ValueError("sfpe: " + msg)
    raise SfpeError("sfpe: " + msg)


def default_params():
    p = Params()
    load().sfpe_default_params(C.byref(p))
    return p


def grid_dims(p):
    h, w, nx, ny = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    check(load().sfpe_grid_dims(C.byref(p), C.byref(h), C.byref(w), C.byref(nx), C.byref(ny)))
    return h.value, w.value, nx.value, ny.value
