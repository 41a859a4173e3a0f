"""ctypes loader for ekat_b200/lib/libekat_b200.so. There is no fallback: if the CUDA
library is missing or a symbol is absent, importing fails loudly."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libekat_b200.so")
HEADER = os.path.join(HERE, "..", "include", "ekat_b200", "ekat_b200.h")

_lib = None

i64, cint, cuint, u64, vp = ctypes.c_int64, ctypes.c_int, ctypes.c_uint, ctypes.c_uint64, ctypes.c_void_p

_TRIDIAG = [vp, vp, vp, vp, i64, cint, cint, cint, i64, i64, cint, cuint, vp]
_SIGS = {}
for _s in ("thomas", "bfb", "cr"):
    for _t in ("f64", "f32"):
        _SIGS[f"ekat_b200_tridiag_{_s}_{_t}"] = _TRIDIAG
for _t in ("f64", "f32"):
    _SIGS[f"ekat_b200_lin_interp_setup_{_t}"] = [vp, vp, vp, i64, cint, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_lin_interp_apply_{_t}"] = [vp, vp, vp, vp, vp, i64, cint, cint, i64, i64, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_lin_interp_{_t}"] = [vp, vp, vp, vp, vp, i64, cint, cint, i64, i64, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_lin_interp_static_map_{_t}"] = [vp, vp, vp, vp, i64, cint, cint, i64, i64, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_lin_interp_lm_{_t}"] = [vp, vp, vp, vp, i64, cint, cint, i64, cint, vp]
    _SIGS[f"ekat_b200_tridiag_lin_interp_{_t}"] = [vp, vp, vp, vp, vp, vp, vp, vp, i64, cint, cint, cint, i64, vp]
    _SIGS[f"ekat_b200_col_to_lev_{_t}"] = [vp, vp, i64, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_lev_to_col_{_t}"] = [vp, vp, i64, cint, i64, i64, vp]
    _SIGS[f"ekat_b200_adj_diff_{_t}"] = [vp, vp, i64, cint, i64, i64, cint, cint, vp]
    _SIGS[f"ekat_b200_diffusion_solve_{_t}"] = [vp, vp, vp, i64, cint, i64, vp]
    _SIGS[f"ekat_b200_diffusion_matrix_{_t}"] = [vp, vp, vp, vp, vp, i64, cint, i64, vp]
    _SIGS[f"ekat_b200_shift_{_t}"] = [vp, vp, i64, cint, i64, i64, cint, cint, ctypes.c_double if _t == "f64" else ctypes.c_float, vp]
    _SIGS[f"ekat_b200_index_and_shift_{_t}"] = [vp, vp, vp, vp, i64, cint, cint, i64, i64, i64, cint, cint, vp]
    _SIGS[f"ekat_b200_transpose3_{_t}"] = [vp, vp, i64, cint, cint, i64, i64, cint, vp]
    _SIGS[f"ekat_b200_repack_{_t}"] = [vp, vp, i64, cint, i64, i64, cint, vp]
    _SIGS[f"ekat_b200_view_reduction_{_t}"] = [vp, vp, i64, cint, cint, i64, cint, cint, vp]
    _SIGS[f"ekat_b200_checksum_{_t}"] = [vp, i64, i64, i64, vp, vp]
    _SIGS[f"ekat_b200_tridiag_residual_{_t}"] = [vp, vp, vp, vp, vp, i64, cint, cint, cint, i64, vp, vp]
    _SIGS[f"ekat_b200_fill_tridiag_{_t}"] = [vp, vp, vp, vp, i64, i64, cint, cint, cint, i64, i64, cint, cint, vp]
    _SIGS[f"ekat_b200_fill_lin_interp_{_t}"] = [vp, vp, vp, i64, i64, cint, cint, i64, i64, cint, u64, vp]
    _SIGS[f"ekat_b200_tridiag_host_{_t}"] = [vp, cint, vp, vp, vp, vp, i64, cint, cint, cint, i64, i64, cint]
    _SIGS[f"ekat_b200_lin_interp_host_{_t}"] = [vp, vp, vp, vp, vp, i64, cint, cint, i64, i64, cint, i64, i64]
_SIGS["ekat_b200_pipeline_create"] = [ctypes.POINTER(vp), ctypes.c_size_t, cint]
_SIGS["ekat_b200_pipeline_destroy"] = [vp]
_SIGS["ekat_b200_pipeline_set_deferred"] = [vp, cint]
_SIGS["ekat_b200_pipeline_sync"] = [vp]
_SIGS["ekat_b200_pipeline_stats"] = [vp, ctypes.POINTER(u64), ctypes.POINTER(u64), ctypes.POINTER(u64)]
_SIGS["ekat_b200_selftest_div_f64"] = [u64, u64, cint, vp, vp]
_SIGS["ekat_b200_selftest_div_f32"] = [u64, u64, cint, vp, vp]
_SIGS["ekat_b200_device_alloc"] = [ctypes.POINTER(vp), ctypes.c_size_t]
_SIGS["ekat_b200_device_free"] = [vp]
_SIGS["ekat_b200_version"] = []
_SIGS["ekat_b200_last_cuda_error"] = []
_SIGS["ekat_b200_launch_count"] = []
_SIGS["ekat_b200_last_kernel"] = []
_SIGS["ekat_b200_set_tuning"] = [cint, cint, cint]
_SIGS["ekat_b200_debug_trace"] = [vp, cint]

STATUS = {0: "OK", -1: "ERR_INVALID", -2: "ERR_UNSUPPORTED", -3: "ERR_CUDA", -4: "ERR_TOO_LARGE"}


class EkatB200Error(RuntimeError):
    def __init__(self, fn, rc, detail=""):
        self.rc = rc
        super().__init__(f"{fn} -> {STATUS.get(rc, rc)}{(': ' + detail) if detail else ''}")


def declared_symbols():
    """Every function name include/ekat_b200/ekat_b200.h declares (macro-generated ones expanded)."""
    text = open(HEADER).read()
    names = set(re.findall(r"\b(ekat_b200_[a-z0-9_]+)\s*\(", text))
    names |= set(re.findall(r"EKAT_B200_DECL_TRIDIAG\((ekat_b200_[a-z0-9_]+),", text))
    out = set()
    for n in names:
        if n.endswith("_##sfx") or "##" in n:
            continue
        out.add(n)
    for base in re.findall(r"\b(ekat_b200_[a-z0-9_]+_)##sfx", text):
        out |= {base + "f64", base + "f32"}
    return sorted(out)


def lib():
    """The loaded library with argtypes set. Raises if it was not built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m ekat_b200.build` (needs nvcc). "
                "ekat_b200 has no CPU or PyTorch fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, args in _SIGS.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.argtypes = args
            fn.restype = cint
        L.ekat_b200_last_cuda_error.restype = ctypes.c_char_p
        L.ekat_b200_last_kernel.restype = ctypes.c_char_p
        L.ekat_b200_launch_count.restype = u64
        L.ekat_b200_set_tuning.restype = None
        _lib = L
    return _lib


def check(fn, rc):
    if rc != 0:
        detail = lib().ekat_b200_last_cuda_error().decode() if rc == -3 else ""
        raise EkatB200Error(fn, rc, detail)
