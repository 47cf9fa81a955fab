"""cli/h5min.hpp (the CLI stand-in's HDF5 subset) against tests/h5ref.py, an independent Python restatement of the same
sections of the HDF5 File Format Specification, and by round trip.  CPU only.  Neither side can be checked against
libhdf5 / h5py here (absent from this environment) -- said plainly in DESIGN.md."""
import os
import subprocess

import numpy as np
import pytest

from tests import h5ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "cli", "h5tool")


@pytest.fixture(scope="module")
def tool():
    src = [os.path.join(ROOT, "cli", f) for f in ("h5tool.cpp", "h5min.hpp")]
    if not os.path.exists(TOOL) or any(os.path.getmtime(s) > os.path.getmtime(TOOL) for s in src):
        subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-o", TOOL, src[0]], check=True)
    return TOOL


def _run(*a):
    return subprocess.run(list(a), capture_output=True, text=True)


def test_writer_output_parses_with_the_independent_reader(tool, tmp_path):
    rng = np.random.default_rng(7)
    knns = rng.integers(1, 2**63, size=(37, 10), dtype=np.uint64)
    dists = rng.integers(0, 1025, size=(37, 10), dtype=np.uint64)
    knns.tofile(tmp_path / "k.u64")
    dists.tofile(tmp_path / "d.u64")
    out = str(tmp_path / "res.h5")
    params = "index=(efc=96,m=24,M=256),query=(ef=64)"
    r = _run(tool, "pack", out, "knns", "10", str(tmp_path / "k.u64"), "dists", "10", str(tmp_path / "d.u64"),
             "--attr-str", "data", "hamming", "--attr-str", "size", "10M", "--attr-str", "algo", "Hnsw",
             "--attr-f64", "buildtime", "575.21", "--attr-f64", "querytime", "0.378697", "--attr-str", "params", params)
    assert r.returncode == 0, r.stderr
    ds, attrs = h5ref.read_file(out)
    assert set(ds) == {"knns", "dists"}
    assert np.array_equal(ds["knns"], knns) and np.array_equal(ds["dists"], dists)
    # the six attributes write_result attaches (cli/src/main.rs:301-306)
    assert attrs == {"data": "hamming", "size": "10M", "algo": "Hnsw", "buildtime": 575.21, "querytime": 0.378697, "params": params}


def test_round_trip_through_the_cpp_reader(tool, tmp_path):
    a = (np.arange(1000 * 16, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)).reshape(1000, 16)
    a.tofile(tmp_path / "a.u64")
    out = str(tmp_path / "data.h5")
    assert _run(tool, "pack", out, "hamming", "16", str(tmp_path / "a.u64"), "--attr-str", "note", "üñí").returncode == 0
    r = _run(tool, "dump", out, "hamming", str(tmp_path / "b.u64"))
    assert r.returncode == 0 and r.stdout.split() == ["1000", "16"], r.stderr
    assert np.array_equal(np.fromfile(tmp_path / "b.u64", dtype=np.uint64).reshape(1000, 16), a)
    assert "note\tstr\tüñí" in _run(tool, "attrs", out).stdout
    r = _run(tool, "dump", out, "nope", str(tmp_path / "c.u64"))
    assert r.returncode == 1 and "no dataset" in r.stderr


@pytest.mark.parametrize("rows,chunk_rows", [(100, 100), (1000, 64), (129, 128), (5, 8)])
def test_cpp_reader_on_a_chunked_file_from_the_python_writer(tool, tmp_path, rows, chunk_rows):
    a = np.random.default_rng(rows).integers(0, 2**64 - 1, size=(rows, 16), dtype=np.uint64)
    path = str(tmp_path / "chunked.h5")
    h5ref.write_chunked(path, "hamming", a, chunk_rows)
    r = _run(tool, "dump", path, "hamming", str(tmp_path / "o.u64"))
    assert r.returncode == 0 and r.stdout.split() == [str(rows), "16"], r.stderr
    assert np.array_equal(np.fromfile(tmp_path / "o.u64", dtype=np.uint64).reshape(rows, 16), a)


def test_rejects_what_it_cannot_read(tool, tmp_path):
    p = tmp_path / "x.h5"
    p.write_bytes(b"not an hdf5 file at all" * 10)
    r = _run(tool, "dump", str(p), "hamming", str(tmp_path / "o.u64"))
    assert r.returncode == 1 and "not an HDF5 file" in r.stderr
    p.write_bytes(h5ref.SIG + bytes([9]) + b"\0" * 200)
    r = _run(tool, "dump", str(p), "hamming", str(tmp_path / "o.u64"))
    assert r.returncode == 1 and "superblock version" in r.stderr
