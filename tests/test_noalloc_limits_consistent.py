"""hb_knns_device_noalloc's stated limit is one number: the header, the ctypes mirror and the Rust -sys crate agree, and
the header says what the entry point is (no counterpart in the reference).  CPU only; the behaviour itself is checked on
the GPU by tests/test_zz_no_alloc.py."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_limit_is_the_same_everywhere():
    from hnsw_itu_b200 import _lib
    header = open(os.path.join(ROOT, "include", "hnsw_b200.h")).read()
    m = re.search(r"#define\s+HB_NOALLOC_MAX_QUERIES\s+(\d+)u", header)
    assert m, "HB_NOALLOC_MAX_QUERIES missing from the header"
    limit = int(m.group(1))
    assert _lib.HB_NOALLOC_MAX_QUERIES == limit
    rust = open(os.path.join(ROOT, "rust", "hnsw-b200-sys", "src", "lib.rs")).read()
    r = re.search(r"pub const HB_NOALLOC_MAX_QUERIES: u32 = (\d+);", rust)
    assert r and int(r.group(1)) == limit
    assert "pub fn hb_knns_device_noalloc(" in rust
    # the static scratch is "modest": control block + one word per query
    assert 256 + 4 * limit <= 1 << 20


def test_header_states_what_it_is_and_its_limits():
    header = open(os.path.join(ROOT, "include", "hnsw_b200.h")).read()
    i = header.index("int hb_knns_device_noalloc(")
    comment = header[header.rfind("/*", 0, i):i]
    for phrase in ("No counterpart in the reference", "request of the run's operators", "HB_NOALLOC_MAX_QUERIES", "race", "ef <= 256"):
        assert phrase in comment, phrase
    for doc in ("DESIGN.md", "INTEGRATION.md"):
        assert "hb_knns_device_noalloc" in open(os.path.join(ROOT, doc)).read(), doc
