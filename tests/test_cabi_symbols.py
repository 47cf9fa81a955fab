"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/hnsw_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os

import numpy as np
import pytest

from hnsw_itu_b200 import _lib


def test_library_loads_and_exports_every_declared_symbol():
    L = _lib.lib()
    names = _lib.declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/hnsw_b200.h but not exported"
    assert L.hb_version() == 100


def test_status_codes_and_error_string():
    L = _lib.lib()
    h = C.c_void_p()
    rc = L.hb_index_load_idx(None, 0, 0, C.byref(h))
    assert rc == _lib.HB_ERR_INVALID and L.hb_last_error()
    info = _lib.IndexInfo()
    assert L.hb_index_get_info(None, C.byref(info)) == _lib.HB_ERR_INVALID
    assert L.hb_index_size(None) == 0
    L.hb_index_free(None)  # no-op


def _no_gpu():
    n = C.c_int(-1)
    assert _lib.lib().hb_device_count(C.byref(n)) == 0
    return n.value == 0


def test_no_cpu_fallback_without_a_device():
    if not _no_gpu():
        pytest.skip("a CUDA device is present")
    import hnsw_itu_b200 as hb
    data = np.zeros((4, 16), dtype=np.uint64)
    with pytest.raises(hb.HbError) as e:
        hb.ground_truth(data, data[:1], 2)
    assert e.value.code == _lib.HB_ERR_NO_DEVICE
    with pytest.raises(hb.HbError) as e:
        hb.index_from_parts("bruteforce", [{"sketches": data}], None)
    assert e.value.code == _lib.HB_ERR_NO_DEVICE
    with pytest.raises(hb.HbError) as e:
        hb.read_index(os.path.join(os.path.dirname(__file__), "golden", "tiny_hnsw.idx"))
    assert e.value.code == _lib.HB_ERR_NO_DEVICE


def test_product_never_touches_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "hnsw_itu_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh", ".sh")):
                text = open(os.path.join(d, f), errors="replace").read()
                for needle in ("import oracle", "from oracle", "liboracle", "hnsw_oracle.hpp", "oracle_capi"):
                    assert needle not in text, f"{f} references the oracle ({needle})"
