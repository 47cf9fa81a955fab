"""hnsw_itu_b200 -- B200-native batched HNSW query engine for hnsw-itu's binary-sketch path.

Host-side mirror of the reference's interface for this path (Rust cannot be
compiled in this environment, so the mirror is Python over the C ABI):

    Sketch / Point::distance        cli/src/sketch.rs:7-29      -> Sketch
    Distance{distance,key,point}    src/index/mod.rs:49-97      -> Distance
    Index::{size,search,knns}       src/index/mod.rs:21-43      -> Index.size/search/knns
    HNSW / NSW / Bruteforce         src/index/*.rs              -> Index (algo attribute)
    read_index / write_index        cli/src/main.rs:224-248     -> read_index / Index.write_index
    ground-truth                    cli/src/main.rs:715-754     -> ground_truth
"""
from .index import (Distance, Index, Sketch, build_hnsw, ground_truth, index_from_parts, read_index, synth_sketches)  # noqa: F401
from ._lib import HbError, build_library  # noqa: F401
from .multi import Comm, MultiIndex  # noqa: F401,E402
