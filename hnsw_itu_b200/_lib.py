"""ctypes binding of libhnswb200.so (the C ABI in include/hnsw_b200.h).

The CUDA extension is the product: if the shared object is missing, or no CUDA
device is usable, calls fail loudly -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HB_LIB_PATH") or os.path.join(_HERE, "libhnswb200.so")  # HB_LIB_PATH: A/B builds (tools/ab.py)
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "hnsw_b200.h")

HB_OK = 0
HB_ERR_INVALID, HB_ERR_IO, HB_ERR_FORMAT, HB_ERR_CUDA, HB_ERR_NO_DEVICE = 1, 2, 3, 4, 5
HB_ERR_OOM, HB_ERR_EMPTY_RESULT, HB_ERR_UNSUPPORTED, HB_ERR_INTERNAL = 6, 7, 8, 9
HB_PAD_LIKE_CLI = 1
HB_EXACT_VISITED = 2
HB_NOALLOC_MAX_QUERIES = 65536  # hb_knns_device_noalloc: queries per call its static scratch holds
ALGO_NAMES = ["bruteforce", "nsw", "hnsw"]

_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)


class HbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"hnsw_b200 error {code}: {msg}")
        self.code = code


class IndexInfo(C.Structure):
    _fields_ = [("algo", C.c_int32), ("device", C.c_int32), ("n_graphs", C.c_uint32), ("row_stride", C.c_uint32),
                ("size", C.c_uint64), ("n_nodes_total", C.c_uint64), ("n_edges_total", C.c_uint64),
                ("entry_point", C.c_int64), ("device_bytes", C.c_uint64), ("build_prunes", C.c_uint64),
                ("build_unlinked", C.c_uint64), ("build_dropped_appends", C.c_uint64)]


class SynthParamsC(C.Structure):
    _fields_ = [("seed_planes", C.c_uint64), ("seed_centres", C.c_uint64), ("seed_rows", C.c_uint64),
                ("latent_dim", C.c_uint32), ("n_clusters", C.c_uint32), ("centre_scale", C.c_uint32),
                ("noise_scale", C.c_uint32), ("distinct_planes", C.c_uint32)]


class BuildParamsC(C.Structure):
    _fields_ = [("ef_construction", C.c_uint32), ("connections", C.c_uint32), ("max_connections", C.c_uint32),
                ("rayon_threads", C.c_uint32), ("serial_prefix", C.c_uint32), ("nsw", C.c_uint32)]


def build_library(force: bool = False) -> str:
    """Compile csrc/ for sm_100a into hnsw_itu_b200/libhnswb200.so (nvcc cross-compiles without a GPU)."""
    src_dir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir)] + [HEADER_PATH]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    subprocess.run(["sh", os.path.join(src_dir, "build.sh"), LIB_PATH], check=True)
    return LIB_PATH


_lib = None


def _prefer_bundled_nccl():
    """libhnswb200 resolves NCCL with dlopen("libnccl.so.2") at first multi-GPU use.  In a process that also imports
    PyTorch, both must end up with the SAME libnccl: the dynamic linker keys loaded objects by SONAME, so if the system
    libnccl (older) were loaded first, `import torch` would later bind to it and fail on symbols it lacks.  Point the
    library at the NCCL wheel PyTorch itself loads (site-packages/nvidia/nccl/lib) when there is one."""
    if os.environ.get("HB_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for base in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
            cand = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["HB_NCCL_LIB"] = cand
                return
    except Exception:
        pass


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(hnsw_itu_b200 has no CPU fallback)")
    _prefer_bundled_nccl()
    L = C.CDLL(LIB_PATH)
    if os.environ.get("HB_LIB_PATH"):
        # an A/B build (tools/ab.py) may predate the newest entry points: binding a missing symbol is skipped, calling it fails
        class _Tolerant:
            def __init__(self, lib):
                object.__setattr__(self, "_lib", lib)

            def __getattr__(self, name):
                try:
                    return getattr(self._lib, name)
                except AttributeError:
                    return C.CFUNCTYPE(C.c_int)(lambda: HB_ERR_UNSUPPORTED)
        L = _Tolerant(L)
    vp = C.c_void_p
    L.hb_last_error.restype = C.c_char_p
    L.hb_version.restype = C.c_int
    L.hb_device_count.argtypes = [C.POINTER(C.c_int)]
    L.hb_index_load_idx.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(vp)]
    L.hb_index_from_parts.argtypes = [C.c_int, C.c_uint32, _u64p, C.POINTER(_u64p), C.POINTER(_u64p), C.POINTER(_u64p),
                                      C.POINTER(vp), C.c_int, C.c_int64, C.c_int, C.c_int, C.POINTER(vp)]
    L.hb_index_save_idx.argtypes = [vp, C.c_char_p]
    L.hb_index_set_attrs.argtypes = [vp, C.c_double, C.c_char_p]
    L.hb_index_get_attrs.argtypes = [vp, C.POINTER(C.c_int), _u64p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_char_p,
                                     C.c_uint64, C.c_char_p, C.c_uint64]
    L.hb_index_set_format_size.argtypes = [vp, C.c_int]
    L.hb_index_compact.argtypes = [vp]
    L.hb_index_free.argtypes = [vp]
    L.hb_index_free.restype = None
    L.hb_index_size.argtypes = [vp]
    L.hb_index_size.restype = C.c_uint64
    L.hb_index_get_info.argtypes = [vp, C.POINTER(IndexInfo)]
    L.hb_index_graph_dims.argtypes = [vp, C.c_uint32, _u64p, _u64p]
    L.hb_index_graph_export.argtypes = [vp, C.c_uint32, _u64p, _u64p, _u64p, _u64p]
    L.hb_knns.argtypes = [vp, _u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _u64p, _u64p, _u32p, vp]
    L.hb_search.argtypes = [vp, _u64p, C.c_uint32, C.c_uint32, C.c_uint32, _u64p, _u64p, _u32p]
    L.hb_knns_device.argtypes = [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp]
    L.hb_knns_device_noalloc.argtypes = [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, vp, vp]
    L.hb_index_last_search_ms.argtypes = [vp]
    L.hb_index_last_search_ms.restype = C.c_double
    L.hb_kernel_launches.restype = C.c_uint64
    L.hb_ground_truth.argtypes = [_u64p, C.c_uint64, _u64p, C.c_uint64, C.c_uint32, C.c_int, _u64p, _u64p]
    L.hb_ground_truth_device.argtypes = [vp, C.c_uint64, C.c_uint64, vp, C.c_uint64, C.c_uint32, vp, vp]
    L.hb_ground_truth_algo.argtypes = [_u64p, C.c_uint64, _u64p, C.c_uint64, C.c_uint32, C.c_int, C.c_int, _u64p, _u64p]
    L.hb_ground_truth_device_algo.argtypes = [vp, C.c_uint64, C.c_uint64, vp, C.c_uint64, C.c_uint32, C.c_int, vp, vp]
    L.hb_topk_merge_device.argtypes = [vp, C.c_uint32, C.c_uint64, C.c_uint32, vp, vp]
    L.hb_unpack_device.argtypes = [vp, C.c_uint64, vp, vp, vp]
    L.hb_pack_device.argtypes = [vp, vp, vp, C.c_uint64, C.c_uint32, C.c_uint64, vp, vp]
    L.hb_synth_device.argtypes = [C.POINTER(SynthParamsC), C.c_uint64, C.c_uint64, vp, vp]
    L.hb_build_hnsw.argtypes = [_u64p, C.c_uint64, C.POINTER(BuildParamsC), C.c_int, C.c_int, C.POINTER(vp)]
    L.hb_tensor_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double)]
    # multi-GPU (multi.cu)
    u8p, dbl3 = C.POINTER(C.c_uint8), C.POINTER(C.c_double)
    L.hb_comm_unique_id.argtypes = [u8p]
    L.hb_comm_init_rank.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.hb_comm_free.argtypes = [vp]
    L.hb_comm_free.restype = None
    L.hb_comm_phase_ms.argtypes = [vp, dbl3]
    L.hb_ground_truth_sharded_device.argtypes = [vp, vp, C.c_uint64, C.c_uint64, vp, C.c_uint64, C.c_uint32, vp, vp]
    L.hb_knns_sharded_device.argtypes = [vp, vp, C.c_uint64, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, vp]
    L.hb_index_broadcast.argtypes = [vp, vp, C.c_int, C.POINTER(vp)]
    L.hb_multi_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    L.hb_multi_free.argtypes = [vp]
    L.hb_multi_free.restype = None
    L.hb_multi_device_count.argtypes = [vp]
    L.hb_multi_size.argtypes = [vp]
    L.hb_multi_size.restype = C.c_uint64
    L.hb_multi_load_idx.argtypes = [vp, C.c_char_p]
    L.hb_multi_replicate.argtypes = [vp, vp]
    L.hb_multi_build_sharded.argtypes = [vp, _u64p, C.c_uint64, C.POINTER(BuildParamsC)]
    L.hb_multi_load_shards.argtypes = [vp, C.POINTER(C.c_char_p), _u64p]
    L.hb_multi_knns.argtypes = [vp, _u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _u64p, _u64p, _u32p]
    L.hb_multi_ground_truth.argtypes = [vp, _u64p, C.c_uint64, _u64p, C.c_uint64, C.c_uint32, _u64p, _u64p]
    L.hb_multi_phase_ms.argtypes = [vp, dbl3]
    L.hb_multi_index.argtypes = [vp, C.c_int]
    L.hb_multi_index.restype = vp
    _lib = L
    return L


def check(rc: int):
    if rc != HB_OK:
        raise HbError(rc, lib().hb_last_error().decode(errors="replace"))


def declared_symbols() -> list[str]:
    """Names of every function include/hnsw_b200.h declares (used by the symbol-export test)."""
    import re
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", text)))
