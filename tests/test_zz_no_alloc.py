"""hb_knns_device_noalloc (scratch in a static device array, no allocation of any kind) against hb_knns_device, through
the C ABI.  No counterpart in the reference: the entry point was added at the request of the run's operators.
In a file of its own that pytest collects last (the round's end runs pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N, K, EF = 30_000, 10, 64


def _search(lib, _lib, entry, idx, dq, nq, k, ef, outs, stream):
    dk, dd, df = outs
    if entry == "noalloc":
        rc = lib.hb_knns_device_noalloc(idx._h, dq.data_ptr(), nq, k, ef, 0, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), stream)
    else:
        rc = lib.hb_knns_device(idx._h, dq.data_ptr(), nq, k, ef, 0, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), None, stream)
    _lib.check(rc)
    return dk[:nq].cpu().numpy().copy(), dd[:nq].cpu().numpy().copy(), df[:nq].cpu().numpy().copy()


def test_no_alloc_entry_point_matches_knns_device_twice_in_a_row():
    import torch
    import hnsw_itu_b200 as hb
    from hnsw_itu_b200 import _lib
    lib = _lib.lib()
    assert isinstance(lib.hb_knns_device_noalloc.argtypes, (list, tuple))
    torch.cuda.set_device(0)
    data = torch.empty((N, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(data.data_ptr(), N, 1003)
    nq_a, nq_b = 3000, 1237  # the second call is smaller: stale status words of the first stay behind it in the array
    qa = torch.empty((nq_a, 16), dtype=torch.int64, device="cuda")
    qb = torch.empty((nq_b, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(qa.data_ptr(), nq_a, 1004)
    hb.synth_sketches(qb.data_ptr(), nq_b, 1005)
    torch.cuda.synchronize()
    idx = hb.build_hnsw(data.cpu().numpy().view(np.uint64), efc=48, m=12, M=48, keep_host=False)
    # every buffer the calls touch exists before the first of them
    outs = [tuple(torch.empty(s, dtype=t, device="cuda") for s, t in (((nq_a, K), torch.int64), ((nq_a, K), torch.int64), ((nq_a,), torch.int32)))
            for _ in range(2)]
    stream = torch.cuda.Stream()
    s = C.c_void_p(stream.cuda_stream)
    torch.cuda.synchronize()

    # the very first search of the process goes through the new entry point (scratch address resolved on a cold library)
    first = _search(lib, _lib, "noalloc", idx, qa, nq_a, K, EF, outs[0], s)
    ref_a = _search(lib, _lib, "alloc", idx, qa, nq_a, K, EF, outs[1], s)
    ref_b = _search(lib, _lib, "alloc", idx, qb, nq_b, K, EF, outs[1], s)
    for got, ref in zip(first, ref_a):
        assert np.array_equal(got, ref)  # integer work: bit-exact, the tolerance every parity test here uses
    assert (ref_a[2] == K).all()

    # two calls in a row, nothing allocated on the device in between (the driver's free-memory figure does not move)
    for o in outs[0]:
        o.fill_(-1)  # (before the measurement: torch loads its fill kernel's module on first use, which takes device memory)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    a = _search(lib, _lib, "noalloc", idx, qa, nq_a, K, EF, outs[0], s)
    b = _search(lib, _lib, "noalloc", idx, qb, nq_b, K, EF, outs[0], s)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    for got, ref in zip(a, ref_a):
        assert np.array_equal(got, ref)
    for got, ref in zip(b, ref_b):
        assert np.array_equal(got, ref)
    assert free1 == free0, (free0, free1)

    # other k / ef on the same array, and the NSW-style single call of one query
    for k, ef, nq in ((1, 1, 1), (10, 16, 500), (25, 128, 700), (100, 256, 300)):
        o2 = tuple(torch.empty(sh, dtype=t, device="cuda") for sh, t in (((nq, k), torch.int64), ((nq, k), torch.int64), ((nq,), torch.int32)))
        o3 = tuple(torch.empty_like(x) for x in o2)
        got = _search(lib, _lib, "noalloc", idx, qa, nq, k, ef, o2, s)
        ref = _search(lib, _lib, "alloc", idx, qa, nq, k, ef, o3, s)
        for g, r in zip(got, ref):
            assert np.array_equal(g, r), (k, ef, nq)

    # limits: refused with a clear error, nothing launched
    big = _lib.HB_NOALLOC_MAX_QUERIES + 1
    with pytest.raises(_lib.HbError) as e:
        _lib.check(lib.hb_knns_device_noalloc(idx._h, qa.data_ptr(), big, K, EF, 0, outs[0][0].data_ptr(), outs[0][1].data_ptr(),
                                              outs[0][2].data_ptr(), s))
    assert e.value.code == _lib.HB_ERR_UNSUPPORTED and "HB_NOALLOC_MAX_QUERIES" in str(e.value)
    with pytest.raises(_lib.HbError) as e:
        _lib.check(lib.hb_knns_device_noalloc(idx._h, qa.data_ptr(), 10, K, 300, 0, outs[0][0].data_ptr(), outs[0][1].data_ptr(),
                                              outs[0][2].data_ptr(), s))
    assert e.value.code == _lib.HB_ERR_UNSUPPORTED
    # and the array still works after the refusals
    again = _search(lib, _lib, "noalloc", idx, qb, nq_b, K, EF, outs[0], s)
    for got, ref in zip(again, ref_b):
        assert np.array_equal(got, ref)
