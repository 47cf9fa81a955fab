"""The C++ stand-in for the reference CLI's `query-index` / `ground-truth` / `index` subcommands (cli/hnsw_itu.cpp):
flags, exit codes and result conventions (1-based keys, padding, attrs) -- cli/src/main.rs:583-754."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")
BIN = os.path.join(ROOT, "cli", "hnsw-itu")


@pytest.fixture(scope="module")
def cli():
    if not os.path.exists(os.path.join(ROOT, "hnsw_itu_b200", "libhnswb200.so")):
        from hnsw_itu_b200 import _lib
        _lib.build_library()
    subprocess.run(["sh", os.path.join(ROOT, "cli", "build.sh")], check=True, capture_output=True)
    return BIN


def run(cli, *args):
    return subprocess.run([cli, *args], capture_output=True, text=True)


def test_help_and_flag_errors(cli):
    assert run(cli, "--help").returncode == 0
    assert run(cli).returncode == 2
    r = run(cli, "query-index", "-Q", "x.u64")
    assert r.returncode == 1 and "--indexfile" in r.stderr
    assert run(cli, "query-index", "--bogus", "1").returncode == 1
    assert run(cli, "frobnicate").returncode == 2


def test_no_gpu_is_an_error_not_a_fallback(cli, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = run(cli, "query-index", "-i", os.path.join(G, "tiny_hnsw.idx"), "-Q", os.path.join(G, "tiny_queries.u64"), "-o", str(tmp_path / "r"))
    assert r.returncode == 1 and ("read_index" in r.stderr or "knns" in r.stderr)


@pytest.mark.gpu
def test_query_index_matches_golden(cli, tmp_path):
    exp = np.load(os.path.join(G, "tiny_expected.npz"))
    out = str(tmp_path / "res")
    r = run(cli, "query-index", "-i", os.path.join(G, "tiny_hnsw.idx"), "-Q", os.path.join(G, "tiny_queries.u64"), "-o", out, "-k", "10", "-e", "64")
    assert r.returncode == 0, r.stderr
    knns = np.fromfile(out + ".knns.u64", dtype=np.uint64).reshape(-1, 10)
    dists = np.fromfile(out + ".dists.u64", dtype=np.uint64).reshape(-1, 10)
    assert np.array_equal(knns, exp["hnsw_k10_ef64_keys"] + np.uint64(1))   # main.rs:277
    assert np.array_equal(dists, exp["hnsw_k10_ef64_dists"])
    attrs = json.load(open(out + ".attrs.json"))
    assert attrs["algo"] == "Hnsw" and attrs["data"] == "hamming" and attrs["params"].endswith(",query=(ef=64)")
    # -S (one search per query) returns the same rows; k > ef pads with the first result (main.rs:467-474)
    r = run(cli, "query-index", "-i", os.path.join(G, "tiny_hnsw.idx"), "-Q", os.path.join(G, "tiny_queries.u64"), "-o", out + "s", "-k", "10", "-e", "4", "-S")
    assert r.returncode == 0 and "greater than `ef`" in r.stderr
    ks = np.fromfile(out + "s.knns.u64", dtype=np.uint64).reshape(-1, 10)
    assert np.array_equal(ks[:, :4], exp["hnsw_k10_ef4_keys"][:, :4] + np.uint64(1))
    assert (ks[:, 4:] == ks[:, :1]).all()


@pytest.mark.gpu
def test_ground_truth_and_index_round_trip(cli, tmp_path, oracle):
    from oracle import oracle as orc
    data = oracle.synth(5000, 71, orc.SynthParams(latent_dim=8, n_clusters=16))
    q = oracle.synth(40, 72, orc.SynthParams(latent_dim=8, n_clusters=16))
    dpath, qpath = str(tmp_path / "d.npy"), str(tmp_path / "q.u64")
    np.save(dpath, data)
    q.tofile(qpath)
    out = str(tmp_path / "gt")
    r = run(cli, "ground-truth", "-d", dpath, "-Q", qpath, "-o", out, "-k", "20", "-b", "1000", "-l", "3000")
    assert r.returncode == 0, r.stderr
    gk, gd = oracle.ground_truth(data[1000:4000], q, 20)
    assert np.array_equal(np.fromfile(out + ".knns.u64", dtype=np.uint64).reshape(-1, 20), gk + np.uint64(1))   # main.rs:745
    assert np.array_equal(np.fromfile(out + ".dists.u64", dtype=np.uint64).reshape(-1, 20), gd)
    # index -> .idx in the reference's bincode format -> query-index; the oracle reads the same file and must agree
    ipath = str(tmp_path / "i.idx")
    r = run(cli, "index", "-d", dpath, "-o", ipath, "-c", "32", "-m", "8", "-M", "32")
    assert r.returncode == 0, r.stderr
    r = run(cli, "inspect", ipath)
    assert r.returncode == 0 and "size: 5000" in r.stdout and "algo: Hnsw" in r.stdout
    assert "base has 5000 nodes" in r.stdout and "connection distribution:" in r.stdout and "p100 " in r.stdout
    assert "query on whole index returned 5000/5000 elements" in r.stdout     # main.rs:806-815
    r = run(cli, "query-index", "-i", ipath, "-Q", qpath, "-o", out + "q", "-k", "10", "-e", "64")
    assert r.returncode == 0, r.stderr
    oidx = oracle.load_idx(ipath)
    ok, od, _ = oidx.knns(q, 10, 64)
    assert np.array_equal(np.fromfile(out + "q.knns.u64", dtype=np.uint64).reshape(-1, 10), ok + np.uint64(1))
    assert json.load(open(out + "q.attrs.json"))["params"] == "index=(efc=32,m=8,M=32),query=(ef=64)"


@pytest.mark.gpu
def test_hdf5_in_and_out_query_subcommand_and_algorithms(cli, tmp_path, oracle):
    """`query` (main.rs:550-571) end to end on HDF5 containers: data and queries in the "hamming" dataset, result in
    "knns" + the six root attributes (main.rs:287-306); NSW and Bruteforce variants (main.rs:441-446); `--len` past the
    end clamps with a warning (main.rs:125-133); ground-truth with fewer than k rows pads with the first (main.rs:467-474)."""
    from oracle import oracle as orc
    from tests import h5ref
    tool = os.path.join(ROOT, "cli", "h5tool")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", tool, os.path.join(ROOT, "cli", "h5tool.cpp")], check=True)
    sp = orc.SynthParams(latent_dim=8, n_clusters=16)
    data, q = oracle.synth(4000, 81, sp), oracle.synth(30, 82, sp)
    data.tofile(tmp_path / "d.u64")
    dh5, qh5 = str(tmp_path / "data.h5"), str(tmp_path / "q.h5")
    assert run(tool, "pack", dh5, "hamming", "16", str(tmp_path / "d.u64")).returncode == 0
    h5ref.write_chunked(qh5, "hamming", q, 8)           # a chunked query file (what h5py writes with chunks=...)
    tk, td = oracle.ground_truth(data, q, 10)
    for algo, name in (("hnsw", "Hnsw"), ("nsw", "Nsw"), ("bruteforce", "Bruteforce")):
        out, ipath = str(tmp_path / f"res_{algo}.h5"), str(tmp_path / f"{algo}.idx")
        r = run(cli, "query", "-d", dh5, "-Q", qh5, "-o", out, "-i", ipath, "-a", algo, "-c", "48", "-m", "12", "-M", "48", "-e", "64", "-S")
        assert r.returncode == 0, r.stderr
        ds, attrs = h5ref.read_file(out)
        assert set(ds) == {"knns"} and ds["knns"].shape == (30, 10)
        assert attrs["algo"] == name and attrs["data"] == "hamming" and attrs["size"] == "4000"
        assert attrs["params"] == "index=(efc=48,m=12,M=48),query=(ef=64)" and attrs["buildtime"] > 0 and attrs["querytime"] > 0
        # the written .idx is the reference's bincode: the oracle reads it and returns the same rows
        ok, _, _ = oracle.load_idx(ipath).knns(q, 10, 64)
        if algo == "bruteforce":
            # Bruteforce::search returns MinK's heap order (bruteforce.rs:37-47) unless -s is given; the device path
            # always returns the rows ascending: same set per row, and equal to the sorted ground truth
            assert np.array_equal(np.sort(ds["knns"], axis=1), np.sort(ok + np.uint64(1), axis=1))
            assert np.array_equal(ds["knns"], tk + np.uint64(1))
        else:
            assert np.array_equal(ds["knns"], ok + np.uint64(1)), algo
            assert orc.recall_at_k(ds["knns"] - np.uint64(1), tk, 10) > 0.8
    # ground truth to HDF5; --len past the end clamps; fewer rows than k pads with the first result
    gout = str(tmp_path / "gt.h5")
    r = run(cli, "ground-truth", "-d", dh5, "-Q", qh5, "-o", gout, "-k", "10", "-b", "3995", "-l", "100")
    assert r.returncode == 0 and "smaller than specified" in r.stderr, r.stderr
    ds, attrs = h5ref.read_file(gout)
    assert attrs == {} and set(ds) == {"knns", "dists"}
    ek, ed = oracle.ground_truth(data[3995:], q, 5)
    assert np.array_equal(ds["knns"][:, :5], ek + np.uint64(1)) and np.array_equal(ds["dists"][:, :5], ed)
    assert (ds["knns"][:, 5:] == ds["knns"][:, :1]).all() and (ds["dists"][:, 5:] == ds["dists"][:, :1]).all()
    # index sub-range: format_size is false, so `size` is the plain number (main.rs:122,282-286)
    r = run(cli, "index", "-d", dh5, "-o", str(tmp_path / "sub.idx"), "-b", "1000", "-l", "2000", "-c", "32", "-m", "8", "-M", "32")
    assert r.returncode == 0, r.stderr
    r = run(cli, "inspect", str(tmp_path / "sub.idx"))
    assert "format_size: false" in r.stdout and "size: 2000" in r.stdout
