//! Raw bindings to `libhnswb200` -- mirrors `include/hnsw_b200.h` one to one.
//! UNCOMPILED (no Rust toolchain in the build image); see rust/README.md.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_double, c_int, c_void};

#[repr(C)]
pub struct hb_index {
    _private: [u8; 0],
}
#[repr(C)]
pub struct hb_comm {
    _private: [u8; 0],
}
#[repr(C)]
pub struct hb_multi {
    _private: [u8; 0],
}

pub const HB_OK: c_int = 0;
pub const HB_ERR_INVALID: c_int = 1;
pub const HB_ERR_IO: c_int = 2;
pub const HB_ERR_FORMAT: c_int = 3;
pub const HB_ERR_CUDA: c_int = 4;
pub const HB_ERR_NO_DEVICE: c_int = 5;
pub const HB_ERR_OOM: c_int = 6;
pub const HB_ERR_EMPTY_RESULT: c_int = 7;
pub const HB_ERR_UNSUPPORTED: c_int = 8;
pub const HB_ERR_INTERNAL: c_int = 9;

/// `Indexes<P>` variants, same numbering as the bincode enum tag (cli/src/main.rs:441-446)
pub const HB_ALGO_BRUTEFORCE: c_int = 0;
pub const HB_ALGO_NSW: c_int = 1;
pub const HB_ALGO_HNSW: c_int = 2;
/// `Indexes::search`: repeat the first result up to k (cli/src/main.rs:467-474)
pub const HB_PAD_LIKE_CLI: u32 = 1;
/// run the kernel with the reference's exact visited set whatever ef is (its dist_evals counter = |visited|)
pub const HB_EXACT_VISITED: u32 = 2;
pub const HB_MAX_EF: u32 = 4096;
pub const HB_NOALLOC_MAX_QUERIES: u32 = 65536;
pub const HB_MAX_K: u32 = 4096;
pub const HB_COMM_ID_BYTES: usize = 128;

#[repr(C)]
#[derive(Default, Clone, Copy, Debug)]
pub struct hb_index_info {
    pub algo: i32,
    pub device: i32,
    pub n_graphs: u32,
    pub row_stride: u32,
    pub size: u64,
    pub n_nodes_total: u64,
    pub n_edges_total: u64,
    pub entry_point: i64,
    pub device_bytes: u64,
    pub build_prunes: u64,
    pub build_unlinked: u64,
    pub build_dropped_appends: u64,
}

#[repr(C)]
#[derive(Default, Clone, Copy, Debug)]
pub struct hb_counters {
    pub expansions: u32,
    pub ids_scanned: u32,
    pub dist_evals: u32,
    pub accepted: u32,
}

#[repr(C)]
#[derive(Default, Clone, Copy, Debug)]
pub struct hb_build_params {
    pub ef_construction: u32,
    pub connections: u32,
    pub max_connections: u32,
    pub rayon_threads: u32,
    pub serial_prefix: u32,
    pub nsw: u32,
}

#[repr(C)]
#[derive(Default, Clone, Copy, Debug)]
pub struct hb_synth_params {
    pub seed_planes: u64,
    pub seed_centres: u64,
    pub seed_rows: u64,
    pub latent_dim: u32,
    pub n_clusters: u32,
    pub centre_scale: u32,
    pub noise_scale: u32,
    pub distinct_planes: u32,
}

extern "C" {
    pub fn hb_last_error() -> *const c_char;
    pub fn hb_version() -> c_int;
    pub fn hb_device_count(count: *mut c_int) -> c_int;

    // ---- index life cycle
    pub fn hb_index_load_idx(path: *const c_char, device: c_int, keep_host: c_int, out: *mut *mut hb_index) -> c_int;
    pub fn hb_index_from_parts(algo: c_int, n_graphs: u32, n_nodes: *const u64, sketches: *const *const u64, down: *const *const u64, off: *const *const u64, ids: *const *const c_void, id_bytes: c_int, entry_point: i64, device: c_int, keep_host: c_int, out: *mut *mut hb_index) -> c_int;
    pub fn hb_index_save_idx(idx: *const hb_index, path: *const c_char) -> c_int;
    pub fn hb_index_set_attrs(idx: *mut hb_index, buildtime: c_double, params: *const c_char) -> c_int;
    pub fn hb_index_set_format_size(idx: *mut hb_index, format_size: c_int) -> c_int;
    pub fn hb_index_get_attrs(idx: *const hb_index, format_size: *mut c_int, size: *mut u64, buildtime: *mut c_double, querytime: *mut c_double, data: *mut c_char, data_cap: u64, params: *mut c_char, params_cap: u64) -> c_int;
    pub fn hb_index_compact(idx: *mut hb_index) -> c_int;
    pub fn hb_index_free(idx: *mut hb_index);
    pub fn hb_index_size(idx: *const hb_index) -> u64;
    pub fn hb_index_get_info(idx: *const hb_index, info: *mut hb_index_info) -> c_int;
    pub fn hb_index_graph_dims(idx: *const hb_index, graph: u32, n_nodes: *mut u64, n_edges: *mut u64) -> c_int;
    pub fn hb_index_graph_export(idx: *const hb_index, graph: u32, sketches: *mut u64, down: *mut u64, off: *mut u64, ids: *mut u64) -> c_int;

    // ---- queries
    pub fn hb_knns(idx: *const hb_index, queries: *const u64, nq: u64, k: u32, ef: u32, flags: u32, keys: *mut u64, dists: *mut u64, found: *mut u32, counters: *mut hb_counters) -> c_int;
    pub fn hb_search(idx: *const hb_index, query: *const u64, k: u32, ef: u32, flags: u32, keys: *mut u64, dists: *mut u64, found: *mut u32) -> c_int;
    pub fn hb_knns_device(idx: *const hb_index, d_queries: *const c_void, nq: u64, k: u32, ef: u32, flags: u32, d_keys: *mut c_void, d_dists: *mut c_void, d_found: *mut c_void, d_counters: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_knns_device_noalloc(idx: *const hb_index, d_queries: *const c_void, nq: u64, k: u32, ef: u32, flags: u32, d_keys: *mut c_void, d_dists: *mut c_void, d_found: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_index_last_search_ms(idx: *const hb_index) -> c_double;
    pub fn hb_kernel_launches() -> u64;

    // ---- brute force / ground truth
    pub fn hb_ground_truth(data: *const u64, n: u64, queries: *const u64, nq: u64, k: u32, device: c_int, keys: *mut u64, dists: *mut u64) -> c_int;
    pub fn hb_ground_truth_device(d_data: *const c_void, n: u64, key_offset: u64, d_queries: *const c_void, nq: u64, k: u32, d_packed: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_ground_truth_algo(data: *const u64, n: u64, queries: *const u64, nq: u64, k: u32, device: c_int, algo: c_int, keys: *mut u64, dists: *mut u64) -> c_int;
    pub fn hb_ground_truth_device_algo(d_data: *const c_void, n: u64, key_offset: u64, d_queries: *const c_void, nq: u64, k: u32, algo: c_int, d_packed: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_tensor_peak(device: c_int, kind: c_int, tops: *mut c_double) -> c_int;
    pub fn hb_topk_merge_device(d_lists: *const c_void, n_lists: u32, nq: u64, k: u32, d_packed_out: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_pack_device(d_keys: *const c_void, d_dists: *const c_void, d_found: *const c_void, nq: u64, k: u32, key_offset: u64, d_packed: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_unpack_device(d_packed: *const c_void, n: u64, d_keys: *mut c_void, d_dists: *mut c_void, stream: *mut c_void) -> c_int;

    // ---- synthetic sketches, index construction
    pub fn hb_synth_device(p: *const hb_synth_params, start_row: u64, n_rows: u64, d_out: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_build_hnsw(data: *const u64, n: u64, p: *const hb_build_params, device: c_int, keep_host: c_int, out: *mut *mut hb_index) -> c_int;

    // ---- multi-GPU: one rank per process
    pub fn hb_comm_unique_id(id: *mut u8) -> c_int;
    pub fn hb_comm_init_rank(id: *const u8, nranks: c_int, rank: c_int, device: c_int, out: *mut *mut hb_comm) -> c_int;
    pub fn hb_comm_free(c: *mut hb_comm);
    pub fn hb_comm_phase_ms(c: *mut hb_comm, ms: *mut c_double) -> c_int;
    pub fn hb_ground_truth_sharded_device(c: *mut hb_comm, d_shard: *const c_void, n_local: u64, key_offset: u64, d_queries: *const c_void, nq: u64, k: u32, d_packed: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_knns_sharded_device(c: *mut hb_comm, idx: *const hb_index, key_offset: u64, d_queries: *const c_void, nq: u64, k: u32, ef: u32, d_packed: *mut c_void, stream: *mut c_void) -> c_int;
    pub fn hb_index_broadcast(c: *mut hb_comm, src: *mut hb_index, root: c_int, out: *mut *mut hb_index) -> c_int;

    // ---- multi-GPU: one process, N devices
    pub fn hb_multi_create(devices: *const c_int, ndev: c_int, out: *mut *mut hb_multi) -> c_int;
    pub fn hb_multi_free(m: *mut hb_multi);
    pub fn hb_multi_device_count(m: *const hb_multi) -> c_int;
    pub fn hb_multi_size(m: *const hb_multi) -> u64;
    pub fn hb_multi_load_idx(m: *mut hb_multi, path: *const c_char) -> c_int;
    pub fn hb_multi_replicate(m: *mut hb_multi, src: *const hb_index) -> c_int;
    pub fn hb_multi_build_sharded(m: *mut hb_multi, data: *const u64, n: u64, p: *const hb_build_params) -> c_int;
    pub fn hb_multi_load_shards(m: *mut hb_multi, paths: *const *const c_char, key_offsets: *const u64) -> c_int;
    pub fn hb_multi_knns(m: *mut hb_multi, queries: *const u64, nq: u64, k: u32, ef: u32, flags: u32, keys: *mut u64, dists: *mut u64, found: *mut u32) -> c_int;
    pub fn hb_multi_ground_truth(m: *mut hb_multi, data: *const u64, n: u64, queries: *const u64, nq: u64, k: u32, keys: *mut u64, dists: *mut u64) -> c_int;
    pub fn hb_multi_phase_ms(m: *mut hb_multi, ms: *mut c_double) -> c_int;
    pub fn hb_multi_index(m: *const hb_multi, r: c_int) -> *const hb_index;
}
