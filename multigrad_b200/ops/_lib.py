"""Build and load ``libmgb200.so`` (the sm_100a kernel library) and mirror its
C structs with ``ctypes`` (binding layer N2 of SURVEY.md).

The library is built **in tree** (``multigrad_b200/lib/libmgb200.so``) with
``nvcc -gencode arch=compute_100a,code=sm_100a`` so that it travels with the
source snapshot to the GPU box.  It exposes a plain C ABI
(``include/multigrad_b200.h``); Python passes raw device pointers
(``tensor.data_ptr()``) and the current CUDA stream.

On a GPU box a missing/unloadable library is a hard error -- there is no
silent PyTorch fallback for the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Optional

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = Path(__file__).resolve().parents[1] / "lib"
LIB_PATH = LIB_DIR / "libmgb200.so"
BUILD_DIR = ROOT / "build" / "obj"
HASH_PATH = LIB_DIR / "libmgb200.hash"

SOURCES = ["workspace.cu", "erfbin.cu", "allreduce.cu", "adam.cu",
           "linlink.cu", "group.cu", "nvls.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
    "-std=c++17", "-Xcompiler", "-fPIC",
]

MGB_MAX_WORLD = 16
MGB_MAX_BINS = 32
MGB_MAX_PARAMS = 8
MGB_MAX_BATCH = 4096


class MgbPeerTable(C.Structure):
    _fields_ = [
        ("world", C.c_int32), ("rank", C.c_int32),
        ("ll_units", C.c_uint32), ("spin_timeout_ms", C.c_uint32),
        ("bulk_bytes", C.c_uint64),
        ("base", C.c_void_p * MGB_MAX_WORLD),
    ]


class MgbErfBinModel(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("n", C.c_int64),
        ("nbins", C.c_int32), ("nparams", C.c_int32),
        ("mu_idx", C.c_int32), ("sig_idx", C.c_int32),
        ("sigma_link", C.c_int32), ("loss_kind", C.c_int32),
        ("loss_eps", C.c_float), ("x0", C.c_float),
        ("step_flags", C.c_int32), ("reserved0", C.c_int32),
        ("edges", C.c_double * (MGB_MAX_BINS + 1)),
        ("norm", C.c_float * MGB_MAX_BINS),
        ("target", C.c_float * MGB_MAX_BINS),
        ("inv_err", C.c_float * MGB_MAX_BINS),
    ]


class MgbStepArgs(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("batch", C.c_int32), ("reduce", C.c_int32),
        ("jac_at_params", C.c_int32),
        ("params", C.c_void_p), ("gate", C.c_void_p),
        ("out_sumstats", C.c_void_p), ("out_jac", C.c_void_p),
        ("out_loss", C.c_void_p), ("out_grad", C.c_void_p),
        ("opt_p", C.c_void_p), ("opt_u", C.c_void_p),
        ("opt_m", C.c_void_p), ("opt_v", C.c_void_p),
        ("opt_step", C.c_void_p),
        ("traj", C.c_void_p), ("loss_hist", C.c_void_p),
        ("traj_rows", C.c_int32), ("step_base", C.c_int32),
        ("lr", C.c_float), ("b1", C.c_float), ("b2", C.c_float),
        ("eps", C.c_float),
        ("bound_kind", C.c_int32 * MGB_MAX_PARAMS),
        ("bound_lo", C.c_float * MGB_MAX_PARAMS),
        ("bound_hi", C.c_float * MGB_MAX_PARAMS),
        ("wait_flag", C.c_void_p), ("wait_value", C.c_uint32),
        ("wait_timeout_ms", C.c_uint32),
        ("nsteps", C.c_int32), ("flags", C.c_int32),
        ("trace", C.c_void_p),
        ("rng_step", C.c_uint32), ("reserved1", C.c_uint32),
    ]


MGB_MAX_GROUP_MODELS = 8


class MgbGroupArgs(C.Structure):
    _fields_ = [
        ("n_models", C.c_int32), ("nparams", C.c_int32),
        ("active", C.c_int32 * MGB_MAX_GROUP_MODELS),
        ("loss", C.c_void_p * MGB_MAX_GROUP_MODELS),
        ("grad", C.c_void_p * MGB_MAX_GROUP_MODELS),
    ]


MODE_SUMSTATS, MODE_SUMSTATS_JAC, MODE_LOSS_GRAD, MODE_ADAM, MODE_GD = range(5)
STEP_STATIC_SPLIT = 1
STEP_CONST_KEY = 2


def nvcc_path() -> Optional[str]:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"),
                 "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def _source_hash() -> str:
    """Content hash of everything the library is built from."""
    import hashlib
    h = hashlib.sha256()
    deps = [CSRC / s for s in SOURCES if (CSRC / s).exists()]
    deps += sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.inc"))
    deps += sorted(INCLUDE.glob("*.h"))
    for d in deps:
        h.update(d.name.encode())
        h.update(d.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _stale() -> bool:
    """True when the built library does not match the sources (by content,
    not mtime: the snapshot that travels to the GPU box rewrites mtimes)."""
    if not LIB_PATH.exists() or not HASH_PATH.exists():
        return True
    return HASH_PATH.read_text().strip() != _source_hash()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libmgb200.so``.

    nvcc cross-compiles, so this also works on a box without a GPU.
    """
    if not force and not _stale():
        return LIB_PATH
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libmgb200.so")
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    srcs = [s for s in SOURCES if (CSRC / s).exists()]
    hdr_t = max([p.stat().st_mtime for p in
                 list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inc")) +
                 list(INCLUDE.glob("*.h"))] + [0.0])

    def compile_one(src: str) -> Path:
        obj = BUILD_DIR / (Path(src).stem + ".o")
        s = CSRC / src
        if (not force and obj.exists()
                and obj.stat().st_mtime > max(s.stat().st_mtime, hdr_t)):
            return obj
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(s), "-o",
               str(obj)]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    tmp = LIB_PATH.with_suffix(".so.tmp%d" % os.getpid())
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
           "-o", str(tmp), *map(str, objs)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}{r.stderr}")
    os.replace(tmp, LIB_PATH)
    HASH_PATH.write_text(_source_hash())
    return LIB_PATH


_LIB: Optional[C.CDLL] = None


def _declare(lib: C.CDLL) -> None:
    vp, i32, u32, u64, i64, f32 = (C.c_void_p, C.c_int32, C.c_uint32,
                                   C.c_uint64, C.c_int64, C.c_float)
    lib.mgb_last_error.restype = C.c_char_p
    lib.mgb_version.restype = C.c_int
    lib.mgb_ws_bytes.restype = u64
    lib.mgb_ws_bytes.argtypes = [C.c_int, u32, u64]
    lib.mgb_ws_alloc.argtypes = [u64, C.POINTER(vp)]
    lib.mgb_ws_free.argtypes = [vp]
    lib.mgb_ws_export.argtypes = [vp, C.c_char_p]
    lib.mgb_ws_import.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.mgb_ws_unimport.argtypes = [vp]
    lib.mgb_ws_bulk_offset.restype = u64
    lib.mgb_ws_bulk_offset.argtypes = [C.c_int, u32]
    lib.mgb_peer_copy.restype = C.c_int
    lib.mgb_peer_copy.argtypes = [vp, vp, u64, vp]
    lib.mgb_ws_read_ctrl.argtypes = [vp, C.c_int, C.POINTER(u32), vp]
    lib.mgb_allreduce_sum.argtypes = [C.POINTER(MgbPeerTable), vp, u64,
                                      C.c_int, C.c_int, vp]
    lib.mgb_bcast_f32.argtypes = [C.POINTER(MgbPeerTable), vp, u32, C.c_int,
                                  vp]
    lib.mgb_erfbin_scratch_bytes.restype = u64
    lib.mgb_erfbin_scratch_bytes.argtypes = [C.c_int, C.c_int]
    lib.mgb_erfbin_grid.restype = C.c_int
    lib.mgb_erfbin_grid.argtypes = [C.c_int]
    lib.mgb_erfbin_cluster.restype = C.c_int
    lib.mgb_erfbin_cluster.argtypes = [C.c_int, C.c_int]
    lib.mgb_erfbin_step.argtypes = [C.POINTER(MgbErfBinModel),
                                    C.POINTER(MgbStepArgs),
                                    C.POINTER(MgbPeerTable), vp, u64, vp]
    lib.mgb_erfbin_eval_host.argtypes = [C.POINTER(MgbErfBinModel),
                                         C.POINTER(MgbPeerTable), vp, vp, vp,
                                         vp, vp, u64, vp]
    lib.mgb_erfbin_eval_host.restype = C.c_int
    lib.mgb_erfbin_eval_host_zc.argtypes = [C.POINTER(MgbErfBinModel),
                                            C.POINTER(MgbPeerTable), vp, vp,
                                            vp, u64, vp, u32]
    lib.mgb_erfbin_eval_host_zc.restype = C.c_int
    lib.mgb_erfbin_adam_host.argtypes = [
        C.POINTER(MgbErfBinModel), C.POINTER(MgbPeerTable), vp, vp, vp, vp,
        i32, i32, f32, f32, f32, f32, vp, vp, vp, u64, vp, u32, vp, u64,
        C.POINTER(C.c_double)]
    lib.mgb_erfbin_adam_host.restype = C.c_int
    lib.mgb_erfbin_host_session.argtypes = [
        C.POINTER(MgbErfBinModel), C.POINTER(MgbPeerTable), vp, vp, vp, u64,
        vp, u32, i32, C.POINTER(u32)]
    lib.mgb_erfbin_host_session.restype = C.c_int
    lib.mgb_erfbin_step_static.argtypes = [C.POINTER(MgbErfBinModel),
                                           C.POINTER(MgbStepArgs),
                                           C.POINTER(MgbPeerTable), vp]
    lib.mgb_erfbin_step_static.restype = C.c_int
    if hasattr(lib, "mgb_linlink_selftest"):
        lib.mgb_linlink_selftest.argtypes = [vp, vp, vp, vp, vp, vp, C.c_int,
                                             vp]
        lib.mgb_linlink_selftest.restype = C.c_int
    lib.mgb_mc_supported.argtypes = [C.POINTER(C.c_int)]
    lib.mgb_mc_round_size.argtypes = [C.c_int, u64, C.POINTER(u64)]
    lib.mgb_mc_create.argtypes = [C.c_int, u64, C.POINTER(u64),
                                  C.POINTER(C.c_int)]
    lib.mgb_mc_import.argtypes = [C.c_int, C.POINTER(u64)]
    lib.mgb_mc_add_device.argtypes = [u64]
    lib.mgb_mc_bind.argtypes = [u64, u64, C.POINTER(vp), C.POINTER(vp),
                                C.POINTER(u64)]
    lib.mgb_mc_release.argtypes = [u64, u64, vp, vp, u64]
    lib.mgb_allreduce_sum_nvls.argtypes = [C.POINTER(MgbPeerTable), vp, vp,
                                           u64, vp, u64, vp]
    for name in ("mgb_mc_supported", "mgb_mc_round_size", "mgb_mc_create",
                 "mgb_mc_import", "mgb_mc_add_device", "mgb_mc_bind",
                 "mgb_mc_release", "mgb_allreduce_sum_nvls"):
        getattr(lib, name).restype = C.c_int
    lib.mgb_group_step.argtypes = [C.POINTER(MgbGroupArgs),
                                   C.POINTER(MgbStepArgs),
                                   C.POINTER(MgbPeerTable), vp]
    lib.mgb_group_step.restype = C.c_int
    lib.mgb_adam_update.argtypes = [vp, vp, vp, vp, i32, i32, f32, f32, f32,
                                    f32, vp]
    lib.mgb_pointer_jump.argtypes = [vp, vp, i64, i32, C.POINTER(i32), vp, vp]
    for name in ("mgb_ws_alloc", "mgb_ws_free", "mgb_ws_export",
                 "mgb_ws_import", "mgb_ws_unimport", "mgb_ws_read_ctrl",
                 "mgb_allreduce_sum", "mgb_bcast_f32", "mgb_erfbin_step",
                 "mgb_adam_update", "mgb_pointer_jump"):
        getattr(lib, name).restype = C.c_int


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first when the sources are newer) the kernel library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if build_if_missing and nvcc_path() is not None and _stale():
        build()
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ "
            "as g; g.build()'` (needs nvcc) before using the CUDA path")
    override = os.environ.get("MULTIGRAD_B200_LIB")   # tuning experiments
    lib = C.CDLL(override if override else str(LIB_PATH))
    _declare(lib)
    _check_abi(lib)
    _LIB = lib
    return lib


def _check_abi(lib: C.CDLL) -> None:
    """The ctypes mirrors must have the size the library was compiled with:
    a stale mirror would silently corrupt kernel arguments."""
    if not hasattr(lib, "mgb_abi_sizes"):
        return                      # experimental build without the query
    from . import linlink as _ll    # mirrors of the linear-link structs
    mirrors = [("MgbPeerTable", MgbPeerTable),
               ("MgbErfBinModel", MgbErfBinModel),
               ("MgbStepArgs", MgbStepArgs), ("MgbGroupArgs", MgbGroupArgs),
               ("MgbLinModel", _ll.MgbLinModel),
               ("MgbLinArgs", _ll.MgbLinArgs)]
    out = (C.c_uint32 * len(mirrors))()
    lib.mgb_abi_sizes.argtypes = [C.POINTER(C.c_uint32), C.c_int]
    lib.mgb_abi_sizes.restype = C.c_int
    n = lib.mgb_abi_sizes(out, len(mirrors))
    for i, (name, cls) in enumerate(mirrors[:n]):
        if int(out[i]) != C.sizeof(cls):
            raise MgbError(
                f"ABI mismatch: sizeof({name}) is {int(out[i])} in "
                f"libmgb200.so but {C.sizeof(cls)} in the Python mirror "
                "(rebuild the library or update the mirror)")


def available() -> bool:
    return LIB_PATH.exists() or nvcc_path() is not None


class MgbError(RuntimeError):
    pass


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = load().mgb_last_error().decode(errors="replace")
        raise MgbError(f"{what or 'libmgb200'} failed (code {code}): {msg}")


def stream_ptr(device=None) -> C.c_void_p:
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
