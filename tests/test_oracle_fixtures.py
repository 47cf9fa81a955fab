"""CPU: the oracle reproduces the committed fixtures from the committed .idx bytes, and its
.idx reader/writer round-trips byte for byte (bincode layout)."""
import filecmp
import os
import struct

import numpy as np

G = os.path.join(os.path.dirname(__file__), "golden")
CASES = [(10, 64), (10, 16), (1, 1), (10, 4), (25, 40), (100, 100)]


def _queries():
    return np.fromfile(os.path.join(G, "tiny_queries.u64"), dtype=np.uint64).reshape(-1, 16)


def test_oracle_matches_fixtures(oracle):
    exp = np.load(os.path.join(G, "tiny_expected.npz"))
    q = _queries()
    for name in ("hnsw", "nsw"):
        idx = oracle.load_idx(os.path.join(G, f"tiny_{name}.idx"))
        assert idx.algo == name
        for k, ef in CASES:
            keys, dists, found = idx.knns(q, k, ef)
            assert np.array_equal(keys, exp[f"{name}_k{k}_ef{ef}_keys"])
            assert np.array_equal(dists, exp[f"{name}_k{k}_ef{ef}_dists"])
            assert np.array_equal(found, exp[f"{name}_k{k}_ef{ef}_found"])
    data = oracle.load_idx(os.path.join(G, "tiny_hnsw.idx")).graph(0)["sketches"]
    gk, gd = oracle.ground_truth(data, q, 10)
    assert np.array_equal(gk, exp["gt_k10_keys"]) and np.array_equal(gd, exp["gt_k10_dists"])


def test_idx_roundtrip_is_byte_exact(oracle, tmp_path):
    for name in ("hnsw", "nsw", "bruteforce"):
        src = os.path.join(G, f"tiny_{name}.idx")
        dst = str(tmp_path / f"{name}.idx")
        oracle.load_idx(src).save_idx(dst)
        assert filecmp.cmp(src, dst, shallow=False)


def test_idx_header_layout():
    # ResultAttrs (cli/src/main.rs:311-320) then the Indexes tag (main.rs:441-446): bincode fixed-width LE
    b = open(os.path.join(G, "tiny_hnsw.idx"), "rb").read(256)
    assert b[0] == 1                                          # format_size: bool
    assert struct.unpack_from("<Q", b, 1)[0] == 7 and b[9:16] == b"hamming"   # data: String
    assert struct.unpack_from("<Q", b, 16)[0] == 600          # size
    assert struct.unpack_from("<I", b, 24)[0] == 2            # algo: Hnsw
    plen = struct.unpack_from("<Q", b, 44)[0]                 # after buildtime f64, querytime f64
    assert b[52:52 + plen] == b"index=(efc=24,m=6,M=16)"
    assert struct.unpack_from("<I", b, 52 + plen)[0] == 2     # Indexes::HNSW


def test_padding_and_k_above_ef(oracle):
    idx = oracle.load_idx(os.path.join(G, "tiny_hnsw.idx"))
    q = _queries()[:5]
    keys, dists, found = idx.knns(q, 10, 4)           # k > ef: at most ef results (cli/src/main.rs:190-195)
    assert (found == 4).all() and (keys[:, 4:] == 0).all()
    pk, pd, pf = idx.knns(q, 10, 4, pad=True)         # Indexes::search pads with the first result (main.rs:467-474)
    assert (pf == 4).all()
    assert np.array_equal(pk[:, :4], keys[:, :4])
    assert (pk[:, 4:] == pk[:, :1]).all() and (pd[:, 4:] == pd[:, :1]).all()


def test_recall_definition():
    from oracle.oracle import recall_at_k
    ids = np.array([[1, 2, 3], [4, 5, 6]]); gt = np.array([[3, 2, 9], [7, 8, 9]])
    assert recall_at_k(ids, gt, 3) == 2 / 6   # cli/eval/eval.py:48-56
