"""Host-side constants mirror the C header (no GPU): the Python enum of ground-truth kernels and the error codes the
wrapper raises are the values include/hnsw_b200.h declares."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header():
    return open(os.path.join(ROOT, "include", "hnsw_b200.h")).read()


def test_gt_algo_enum_matches_header():
    from hnsw_itu_b200 import index
    m = re.search(r"typedef enum hb_gt_algo \{([^}]*)\}", _header())
    assert m, "hb_gt_algo not found in the header"
    vals = {k.strip(): int(v) for k, v in (e.split("=") for e in m.group(1).split(","))}
    assert vals == {"HB_GT_AUTO": 0, "HB_GT_SCAN": 1, "HB_GT_MMA": 2, "HB_GT_MXF4": 3}
    assert index.GT_ALGOS == {"auto": vals["HB_GT_AUTO"], "scan": vals["HB_GT_SCAN"], "mma": vals["HB_GT_MMA"],
                              "mxf4": vals["HB_GT_MXF4"]}


def test_unknown_gt_algo_is_rejected_before_any_device_work():
    import numpy as np
    import pytest
    import hnsw_itu_b200 as hb
    d = np.zeros((4, 16), dtype=np.uint64)
    with pytest.raises(KeyError):
        hb.ground_truth(d, d, 1, algo="tensor")
