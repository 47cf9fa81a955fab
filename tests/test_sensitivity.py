"""SURVEY 8(c) sensitivity test: how much of the reference's OWN output depends on neighbour iteration order.

The reference iterates `HashSet<usize>` adjacency lists (src/collections/simplegraph.rs:10-14,111-117) whose order is
re-randomised in every process, and the accept test reads the current worst element per
neighbour (src/index/nsw.rs:156-177), so two runs of the real binary on the same .idx can return different id lists at
distance ties.  Parity here is defined against ONE canonical order (the .idx file order); this test quantifies the band
the reference itself moves in -- the oracle re-run with randomly permuted neighbour order, several seeds -- and checks
that the canonical-order result (which the CUDA path reproduces bit for bit) sits inside it.
"""
import numpy as np

from oracle import oracle as orc


def permuted(graphs, seed):
    """The same graphs with every adjacency list in a random order (what a fresh process of the reference sees)."""
    rng = np.random.default_rng(seed)
    out = []
    for g in graphs:
        ids = np.array(g["ids"], dtype=np.uint64, copy=True)
        off = np.asarray(g["off"], dtype=np.uint64)
        # one random key per edge, sorted within its row: a uniform permutation of every row
        row = np.repeat(np.arange(len(off) - 1), np.diff(off).astype(np.int64))
        order = np.lexsort((rng.random(len(ids)), row))
        out.append({"sketches": g["sketches"], "down": g["down"], "off": g["off"], "ids": ids[order]})
    return out


def sensitivity(o, data, queries, k, ef, seeds, efc=48, m=12, M=48):
    oidx = o.build(data, "hnsw", efc=efc, m=m, M=M, parallel=False)
    graphs = oidx.graphs()
    gk, _ = o.ground_truth(data, queries, k)
    ck, cd, _ = oidx.knns(queries, k, ef)
    canon_recall = orc.recall_at_k(ck, gk, k)
    recalls, changed_ids, changed_sets = [], [], []
    for s in seeds:
        p = o.from_parts("hnsw", permuted(graphs, s), oidx.ep)
        pk, pd, _ = p.knns(queries, k, ef)
        recalls.append(orc.recall_at_k(pk, gk, k))
        changed_ids.append(float((pk != ck).any(axis=1).mean()))
        changed_sets.append(float(np.mean([set(a.tolist()) != set(b.tolist()) for a, b in zip(pk, ck)])))
    return {"canonical_recall": canon_recall, "permuted_recalls": recalls, "frac_queries_id_list_changed": changed_ids,
            "frac_queries_id_set_changed": changed_sets}


def test_permutation_keeps_the_graph():
    o = orc.get()
    data = o.synth(500, 5)
    g = o.build(data, "hnsw", efc=16, m=4, M=8, parallel=False).graphs()
    p = permuted(g, 1)
    for a, b in zip(g, p):
        off = np.asarray(a["off"], dtype=np.int64)
        for v in range(len(off) - 1):
            assert sorted(np.asarray(a["ids"])[off[v]:off[v + 1]].tolist()) == sorted(np.asarray(b["ids"])[off[v]:off[v + 1]].tolist())
    assert any(not np.array_equal(np.asarray(a["ids"]), np.asarray(b["ids"])) for a, b in zip(g, p))


def test_canonical_order_sits_inside_the_references_own_band():
    o = orc.get()
    # tie-heavy data (64 distinct hyperplanes: distances are multiples of 16) makes the order matter visibly
    for planes, tol in ((1024, 0.01), (64, 0.03)):
        p = orc.SynthParams(latent_dim=16, n_clusters=64, centre_scale=1, noise_scale=1, distinct_planes=planes)
        data, q = o.synth(8000, 41, p), o.synth(400, 42, p)
        r = sensitivity(o, data, q, 10, 32, seeds=range(6))
        lo, hi = min(r["permuted_recalls"]), max(r["permuted_recalls"])
        # the canonical run is one more draw from the same family: inside the band up to its own spread
        assert lo - tol <= r["canonical_recall"] <= hi + tol, r
        if planes == 64:
            # the reference is NOT reproducible run to run on tie-heavy data: some id lists change with the order
            assert max(r["frac_queries_id_list_changed"]) > 0.0, r
