"""ctypes loader for libdmnerf_b200.so (the C-ABI in include/dmnerf_b200.h).

There is no fallback: if the shared library has not been built, or a call returns a non-zero
code, a RuntimeError is raised.  Build with ``python -m dm_nerf_b200._lib`` (or
``__graft_entry__.build()``): plain ``nvcc -gencode arch=compute_100a,code=sm_100a``.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
SO_PATH = os.path.join(PKG_DIR, "libdmnerf_b200.so")

MODE_FP32, MODE_BF16, MODE_F16, MODE_TC32 = 0, 1, 2, 3
COMP_DROP_EMPTY, COMP_RAW_INS = 1, 2
PEN_ACCUMULATE = 2
NUM_PARAMS = 30

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]
OBJ_DIR = os.path.join(PKG_DIR, "build")


def experimental() -> bool:
    """The parked forward variants (csrc/experimental/) are compiled only when DMNERF_EXPERIMENTAL=1."""
    return os.environ.get("DMNERF_EXPERIMENTAL", "0") not in ("", "0")


def sources():
    src = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    if experimental():
        src += sorted(glob.glob(os.path.join(CSRC, "experimental", "*.cu")))
    return src


def _headers():
    return (glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "experimental", "*.cuh"))
            + [os.path.join(PKG_DIR, "..", "include", "dmnerf_b200.h")])


def _flavour() -> str:
    return "exp" if experimental() else "prod"


def needs_build() -> bool:
    if not os.path.exists(SO_PATH):
        return True
    tag = os.path.join(OBJ_DIR, "flavour")
    if os.path.exists(tag) and open(tag).read().strip() != _flavour():
        return True
    t = os.path.getmtime(SO_PATH)
    return any(os.path.getmtime(d) > t for d in sources() + _headers() if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a (one nvcc per file, in parallel; cross-compiles without a GPU) and
    link them into one shared library.  Objects are cached under build/ and rebuilt when their source or any header
    is newer."""
    if not force and not needs_build():
        return SO_PATH
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.environ.get("NVCC", "nvcc")
    odir = os.path.join(OBJ_DIR, _flavour())
    os.makedirs(odir, exist_ok=True)
    flags = NVCC_FLAGS + (["-DDMNERF_EXPERIMENTAL"] if experimental() else []) + (["-Xptxas", "-v"] if verbose else [])
    hdr_t = max(os.path.getmtime(h) for h in _headers() if os.path.exists(h))

    def compile_one(src):
        obj = os.path.join(odir, os.path.basename(src)[:-3] + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), hdr_t):
            return obj, ""
        res = subprocess.run([nvcc] + flags + ["-c", "-o", obj, src], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + res.stdout + res.stderr)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, sources()))
    if verbose:
        print("".join(r[1] for r in results))
    res = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", SO_PATH]
                         + [r[0] for r in results], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    with open(os.path.join(OBJ_DIR, "flavour"), "w") as f:
        f.write(_flavour())
    return SO_PATH


_p = C.c_void_p
_i64 = C.c_int64
_int = C.c_int
_f = C.c_float

_SIGNATURES = {
    "dmnerf_version": (_int, []),
    "dmnerf_built_arch": (_int, []),
    "dmnerf_has_experimental": (_int, []),
    "dmnerf_last_error": (_int, [C.c_char_p, _int]),
    "dmnerf_packed_bytes": (_i64, [_int, _int]),
    "dmnerf_weights_pack": (_int, [C.POINTER(_p), _int, _int, _p, _p]),
    "dmnerf_mlp_fwd": (_int, [_p, _p, _p, _i64, _int, _p, _int, _int, _p, _p, _p]),
    "dmnerf_mlp_fwd_embedded": (_int, [_p, _i64, _p, _int, _int, _p, _p, _p]),
    "dmnerf_mlp_density": (_int, [_p, _p, _p, _i64, _int, _p, _int, _int, _p, _p]),
    "dmnerf_weights_from_sigma": (_int, [_p, _p, _p, _i64, _int, _int, _p, _p]),
    "dmnerf_acts_bytes": (_i64, [_i64, _int, _int]),
    "dmnerf_mlp_bwd": (_int, [_p, _p, _p, _i64, _int, _int, C.POINTER(_p), _p, _p]),
    "dmnerf_bwd_workspace_bytes": (_i64, [_i64, _int, _int]),
    "dmnerf_mlp_bwd_bf16_noalloc": (_int, [_p, _p, _p, _i64, _int, C.POINTER(_p), _p]),
    "dmnerf_noalloc_max_points": (_i64, []),
    "dmnerf_composite_fwd": (_int, [_p, _p, _p, _i64, _int, _int, _int, _p, _p, _p, _p, _p, _p]),
    "dmnerf_composite_bwd": (_int, [_p, _p, _p, _i64, _int, _int, _int, _int, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "dmnerf_sample_pdf": (_int, [_p, _p, _p, _i64, _i64, _int, _int, _p, _p, _p]),
    "dmnerf_sample_cdf": (_int, [_p, _p, _p, _i64, _i64, _int, _int, _p, _p, _p]),
    "dmnerf_fine_z": (_int, [_p, _p, _p, _i64, _i64, _int, _int, _p, _p, _p]),
    "dmnerf_coarse_tail": (_int, [_p, _p, _i64, _p, _p, _i64, _i64, _int, _int, _int, _p, _p, _p]),
    "dmnerf_render_workspace_bytes": (_i64, [_i64, _i64, _int, _int, _int, _int]),
    "dmnerf_render_full": (_int, [_int, _int, _p, _p, _i64, _i64, _p, _p, _int, _int, _p, _p, _int, _int, _i64, _p, _i64,
                                  _p, _p, _p, _p, _p]),
    "dmnerf_sort_cat": (_int, [C.POINTER(_p), C.POINTER(_int), _int, _i64, _p, _p]),
    "dmnerf_jitter_z": (_int, [_p, _p, _i64, _int, _p, _p]),
    "dmnerf_penalizer_fwd": (_int, [_p, _p, _p, _p, _i64, _int, _int, _f, _f, _int, _p, _p, _p]),
    "dmnerf_penalizer_bwd": (_int, [_p, _p, _p, _p, _i64, _int, _int, _f, _f, _int, _p, _p, _p, _p]),
    "dmnerf_exchange": (_int, [_p, _p, _p, _p, _int, _int, _i64, _int, _int, _p, _p, _p]),
    "dmnerf_rays_k": (_int, [_int, _int, C.POINTER(_f), C.POINTER(_f), _i64, _p, _i64, _p, _p, _p]),
    "dmnerf_rays_k_dev": (_int, [_int, _int, _p, _p, _i64, _p, _i64, _p, _p, _p]),
    "dmnerf_embed": (_int, [_p, _i64, _int, _p, _p]),
    "dmnerf_label_argmax": (_int, [_p, _i64, _int, _p, _p]),
    "dmnerf_selftest_umma": (_int, [_p, _p, _int, _int, _p, _p]),
    "dmnerf_selftest_umma_mn": (_int, [_p, _p, _int, _p, _p]),
    "dmnerf_selftest_umma2": (_int, [_p, _p, _int, _int, _p, _p]),
    "dmnerf_selftest_tmem_read_bw": (_int, [_int, _int, _int, _p, _p]),
    "dmnerf_selftest_umma_rate": (_int, [_int, _int, _p, _p]),
    "dmnerf_debug_counters": (_int, [_p]),
    "dmnerf_debug_counters_tc32": (_int, [_p]),
    "dmnerf_hungarian": (_int, [_p, _int, _i64, _int, _p, _int, _p, _p]),
    "dmnerf_hungarian_infeasible": (_int, [C.POINTER(_int), _p]),
    "dmnerf_segment_sum": (_int, [_p, _p, _i64, _int, _int, _p, _p]),
    "dmnerf_segment_sum3": (_int, [_p, _p, _p, _i64, _int, _int, _p, _p]),
    "dmnerf_assign_scores": (_int, [_p, _p, _i64, _int, _p, _p, _p, _p, _p]),
    "dmnerf_assign_loss_grad": (_int, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _int, _p, _p, _p, _p]),
    "dmnerf_selftest_issue_cost": (_int, [_int, _int, _p, _p]),
    "dmnerf_debug_flags": (_int, [_int]),
    "dmnerf_reserve_sms": (_int, [_int]),
    "dmnerf_draw_pixels": (_int, [_p, _i64, _i64, _p, C.c_uint64, _p]),
    "dmnerf_adam_step": (_int, [_p, _p, _p, _p, _i64, _p, _f, _p, _f, _f, _f, _f, _p]),
}

EXPORTS = tuple(_SIGNATURES)

_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library.  Raises if it is missing -- no CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                f"dm_nerf_b200: {SO_PATH} is missing. Build it with `python -m dm_nerf_b200._lib` "
                "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path.")
        handle = C.CDLL(SO_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error() -> str:
    buf = C.create_string_buffer(512)
    lib().dmnerf_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed with code {rc}: {last_error()}")


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
