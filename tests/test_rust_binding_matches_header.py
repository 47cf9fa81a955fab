"""rust/hnsw-b200-sys/src/lib.rs cannot be compiled here (no Rust toolchain); this keeps it from drifting: every function
the header declares is bound, nothing else is, and the argument counts agree."""
import os
import re

from hnsw_itu_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _c_decls():
    text = re.sub(r"/\*.*?\*/", "", open(_lib.HEADER_PATH).read(), flags=re.S)
    out = {}
    for m in re.finditer(r"\b(hb_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def _rust_decls():
    text = open(os.path.join(ROOT, "rust", "hnsw-b200-sys", "src", "lib.rs")).read()
    out = {}
    for m in re.finditer(r"pub fn (hb_[a-z0-9_]+)\s*\(([^)]*)\)", text, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if not args else args.count(":")
    return out


def test_every_header_function_is_bound_with_the_same_arity():
    c, r = _c_decls(), _rust_decls()
    assert set(c) == set(_lib.declared_symbols())
    assert set(r) == set(c), (sorted(set(c) - set(r)), sorted(set(r) - set(c)))
    assert {k: v for k, v in r.items() if c[k] != v} == {}


def test_build_params_struct_matches():
    text = re.sub(r"/\*.*?\*/", "", open(_lib.HEADER_PATH).read(), flags=re.S)
    c_fields = re.findall(r"uint32_t (\w+);", re.search(r"typedef struct hb_build_params \{(.*?)\}", text, flags=re.S).group(1))
    rust = open(os.path.join(ROOT, "rust", "hnsw-b200-sys", "src", "lib.rs")).read()
    r_fields = re.findall(r"pub (\w+): u32", re.search(r"pub struct hb_build_params \{(.*?)\}", rust, flags=re.S).group(1))
    assert c_fields == r_fields == [f for f, _ in _lib.BuildParamsC._fields_]
