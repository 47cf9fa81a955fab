/* hnsw_b200.h -- C ABI of the B200-native batched HNSW query engine for the
 * binary-sketch path of JonasUJ/hnsw-itu (1024-bit sketches, Hamming distance).
 *
 * This header is the drop-in boundary: plain pointers and sizes, no C++ or
 * torch types.  Each entry point names the reference interface it replaces
 * (paths relative to the reference checkout).  The reference is Rust with no
 * FFI today; INTEGRATION.md shows the `-sys` binding a maintainer would add.
 *
 * Conventions
 *   - a sketch is 16 little-endian uint64 words = 128 bytes (cli/src/sketch.rs:7-10);
 *     `queries` / `data` arrays are row-major (n, 16) uint64.
 *   - keys are 0-based node indices as in the .idx file (the CLI adds +1 only
 *     when writing HDF5, cli/src/main.rs:277,745).
 *   - results are ordered by the reference's total order (distance, key)
 *     (src/index/mod.rs:90-97).
 *   - neighbour iteration order is the order ids appear in the .idx file
 *     (the reference's own order is re-randomised per process).
 *   - every function returns an hb_status; hb_last_error() gives a
 *     thread-local message.  No exception or abort crosses this boundary.
 *   - there is no CPU fallback: without a CUDA device every compute entry
 *     point fails with HB_ERR_NO_DEVICE.
 */
#ifndef HNSW_B200_H_
#define HNSW_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HB_VERSION 100 /* 0.1.0 */
#define HB_SKETCH_WORDS 16

typedef enum hb_status {
  HB_OK = 0,
  HB_ERR_INVALID = 1,      /* bad argument */
  HB_ERR_IO = 2,           /* file could not be opened / short read or write */
  HB_ERR_FORMAT = 3,       /* .idx bytes or graph arrays are inconsistent */
  HB_ERR_CUDA = 4,         /* CUDA runtime error (message has the detail) */
  HB_ERR_NO_DEVICE = 5,    /* no usable CUDA device: there is no CPU fallback */
  HB_ERR_OOM = 6,          /* host or device allocation failed */
  HB_ERR_EMPTY_RESULT = 7, /* padding requested but a search found nothing (the reference panics, main.rs:472) */
  HB_ERR_UNSUPPORTED = 8,  /* e.g. ef above HB_MAX_EF */
  HB_ERR_INTERNAL = 9
} hb_status;

/* Indexes<P> variants, same numbering as the bincode enum tag (cli/src/main.rs:441-446) */
typedef enum hb_algo { HB_ALGO_BRUTEFORCE = 0, HB_ALGO_NSW = 1, HB_ALGO_HNSW = 2 } hb_algo;

/* flags for hb_knns / hb_search */
#define HB_PAD_LIKE_CLI 1u /* Indexes::search: repeat the first result up to k (main.rs:467-474) */
#define HB_EXACT_VISITED 2u /* run the kernel that keeps the reference's exact visited set (nsw.rs:143,157-161) whatever ef is:
                              same results, slower; its dist_evals counter then equals the reference's |visited| */

#define HB_MAX_EF 4096u
#define HB_MAX_K 4096u

typedef struct hb_index hb_index; /* opaque: device-resident index + optional host copy */

typedef struct hb_index_info {
  int32_t algo;        /* hb_algo */
  int32_t device;      /* CUDA ordinal the index lives on */
  uint32_t n_graphs;   /* 0 bruteforce, 1 NSW, 1 + layers for HNSW (graph 0 = base) */
  uint32_t row_stride; /* 32-bit words per adjacency row in HBM (word 0 = degree) */
  uint64_t size;       /* Index::size(): number of base points */
  uint64_t n_nodes_total;
  uint64_t n_edges_total; /* directed */
  int64_t entry_point;    /* as stored in the file (index into the top graph), -1 = None */
  uint64_t device_bytes;
  /* hb_build_hnsw statistics (0 for loaded indexes): M-prunings done (nsw.rs:109-130), edges removed on the far
   * side by them, and back edges dropped because a row was physically full (should be 0) */
  uint64_t build_prunes, build_unlinked, build_dropped_appends;
} hb_index_info;

/* per-query work counters (feature `instrument` of the reference traces V as
 * "visited", src/index/nsw.rs:180-181; X/E/A complete the roofline byte model) */
typedef struct hb_counters {
  uint32_t expansions;   /* X: candidates expanded */
  uint32_t ids_scanned;  /* E: neighbour ids iterated */
  uint32_t dist_evals;   /* V: distance evaluations performed; >= |visited| of the reference (the ef <= 256 path keeps a
                            recently-seen cache instead of an exact visited set; results, X, E and A are unaffected) */
  uint32_t accepted;     /* A: pushes into the beam */
} hb_counters;

const char* hb_last_error(void);
int hb_version(void);
int hb_device_count(int* count);

/* ---- index life cycle ------------------------------------------------- */

/* read_index (cli/src/main.rs:224-234): parse a bincode 1.3.3 IndexFile<Sketch>
 * and place it in HBM of `device`.  keep_host != 0 retains the host copy needed
 * by hb_index_save_idx / hb_index_graph_export. */
int hb_index_load_idx(const char* path, int device, int keep_host, hb_index** out);

/* Build an index handle from arrays (what a Rust wrapper holding an
 * `HNSW<Sketch>` would pass: hnsw.rs:255-271 `layers()`/`base()`).
 * Graph 0 is the base, graph g>0 is layers[g-1].  For each graph:
 * sketches (n,16) u64, down n u64 (index into the graph below; ignored for
 * graph 0), off n+1 u64 CSR offsets, ids u32 or u64 (id_bytes = 4 | 8) in
 * canonical neighbour order.  entry_point indexes the top graph, -1 = None. */
int hb_index_from_parts(int algo, uint32_t n_graphs, const uint64_t* n_nodes, const uint64_t* const* sketches,
                        const uint64_t* const* down, const uint64_t* const* off, const void* const* ids, int id_bytes,
                        int64_t entry_point, int device, int keep_host, hb_index** out);

/* write_index (main.rs:236-248): needs keep_host. */
int hb_index_save_idx(const hb_index* idx, const char* path);
/* ResultAttrs carried by the .idx (main.rs:311-320); params is "index=(efc=..,m=..,M=..)" */
int hb_index_set_attrs(hb_index* idx, double buildtime, const char* params);
/* ResultAttrs.format_size: false when the index covers a `--start/--len` sub-range (main.rs:122) */
int hb_index_set_format_size(hb_index* idx, int format_size);
/* ResultAttrs as loaded / set (main.rs:311-320): what write_result copies into the result file (main.rs:287-306).
 * data/params are copied into caller buffers (NUL-terminated, truncated to cap); any pointer may be NULL. */
int hb_index_get_attrs(const hb_index* idx, int* format_size, uint64_t* size, double* buildtime, double* querytime,
                       char* data, uint64_t data_cap, char* params, uint64_t params_cap);
/* Re-lay the adjacency of a loaded / built index from fixed-stride rows to line-granular rows + one 32-bit offset per node
 * (about 5x less HBM at M = 256: 16 GB -> 3 GB for 10 M points; one more dependent 4-byte read per expansion).  Results are
 * unchanged.  Also applied automatically when HB_INDEX_LAYOUT=compact is set.  hb_index_info.row_stride reads 0 afterwards. */
int hb_index_compact(hb_index* idx);
void hb_index_free(hb_index* idx);
uint64_t hb_index_size(const hb_index* idx); /* Index::size (src/index/mod.rs:22) */
int hb_index_get_info(const hb_index* idx, hb_index_info* info);
/* host copy of one graph (needs keep_host); any pointer may be NULL.  ids are returned as u64. */
int hb_index_graph_dims(const hb_index* idx, uint32_t graph, uint64_t* n_nodes, uint64_t* n_edges);
int hb_index_graph_export(const hb_index* idx, uint32_t graph, uint64_t* sketches, uint64_t* down, uint64_t* off, uint64_t* ids);

/* ---- queries ------------------------------------------------------------ */

/* Index::knns (src/index/mod.rs:28-42) through Indexes::search
 * (cli/src/main.rs:457-477): one batched launch, results in input order.
 * HOST buffers; the transfers happen inside (page-locked caller memory is used in
 * place; pageable memory goes through page-locked staging owned by the index,
 * copied by a few helper threads).  keys/dists are (nq, k) u64,
 * found (nq) u32 = results before padding (<= min(k, ef)); rows are
 * zero-filled past found[i] unless HB_PAD_LIKE_CLI.  counters may be NULL. */
int hb_knns(const hb_index* idx, const uint64_t* queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags,
            uint64_t* keys, uint64_t* dists, uint32_t* found, hb_counters* counters);

/* Index::search (src/index/mod.rs:23): a batch of one. */
int hb_search(const hb_index* idx, const uint64_t* query, uint32_t k, uint32_t ef, uint32_t flags, uint64_t* keys,
              uint64_t* dists, uint32_t* found);

/* Same with DEVICE buffers on the index's device; `stream` is a cudaStream_t
 * (NULL = default stream).  Synchronises `stream` before returning (the
 * exact-visited-set overflow check needs one word back from the device). */
int hb_knns_device(const hb_index* idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags,
                   void* d_keys, void* d_dists, void* d_found, void* d_counters, void* stream);

/* No counterpart in the reference: added at the request of the run's operators, for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  hb_knns_device for
 * ef <= 256 that allocates NOTHING: no cudaMalloc / cudaFree / stream-ordered or pooled allocation, no event creation.
 * The kernel's scratch (work counter, overflow flag, one status word per query) lives in a static device array of
 * 256 B + 4 * HB_NOALLOC_MAX_QUERIES bytes that every GPU that loads the library holds until exit.  Same results as
 * hb_knns_device (same kernel).  Limits: nq <= HB_NOALLOC_MAX_QUERIES and ef <= 256, no HB_EXACT_VISITED, no counters
 * (HB_ERR_UNSUPPORTED otherwise); a batch in which a query overflows its boundary-tie stack (tie-heavy data; never
 * seen on the synthetic sets) also returns HB_ERR_UNSUPPORTED, with found[] = 0 for those queries, because the rerun
 * needs allocated tables -- use hb_knns_device for it.  All calls on a device share the one array: calls that overlap
 * on the GPU (two streams, two threads) race; serialise them.  Synchronises `stream` before returning. */
#define HB_NOALLOC_MAX_QUERIES 65536u
int hb_knns_device_noalloc(const hb_index* idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags,
                           void* d_keys, void* d_dists, void* d_found, void* stream);

/* Device time (ms, CUDA events on the launching stream) of the search kernel(s) of the last hb_knns* call on this
 * index; used for the roofline figure. */
double hb_index_last_search_ms(const hb_index* idx);
/* Number of CUDA kernels this library has launched in this process so far. */
uint64_t hb_kernel_launches(void);

/* ---- brute force / ground truth ----------------------------------------------- */

/* `ground-truth` (cli/src/main.rs:715-754): Bruteforce::search
 * (src/index/bruteforce.rs:37-47, MinK src/collections/mod.rs:52-77) for every
 * query followed by the CLI's sort: the exact k smallest (distance, key),
 * ascending.  HOST buffers; keys/dists (nq, k) u64.  n < k pads with
 * (distance, key) = (UINT64_MAX, UINT64_MAX). */
int hb_ground_truth(const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k, int device,
                    uint64_t* keys, uint64_t* dists);

/* DEVICE buffers: d_data (n,16) u64 is one shard whose first row has global
 * key `key_offset`; writes (nq, k) packed u64 = (distance << 32) | key,
 * ascending; missing entries are UINT64_MAX. */
int hb_ground_truth_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq,
                           uint32_t k, void* d_packed, void* stream);

/* Kernel choice for the two calls above.  HB_GT_AUTO picks the tcgen05 GEMM on E2M1 operands (gt_mxf4.cu) for batches
 * of at least 32 queries over at least 4096 rows and the XOR+POPC scan (bruteforce.cu) otherwise; HB_GT_MMA is the
 * same GEMM on int8 operands (gt_tcgen05.cu).  All three are exact and return identical lists.  The *_algo variants force one (HB_ERR_UNSUPPORTED if it cannot run the request). */
typedef enum hb_gt_algo { HB_GT_AUTO = 0, HB_GT_SCAN = 1, HB_GT_MMA = 2, HB_GT_MXF4 = 3 } hb_gt_algo;
int hb_ground_truth_algo(const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k, int device,
                         int algo, uint64_t* keys, uint64_t* dists);
int hb_ground_truth_device_algo(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq,
                                uint32_t k, int algo, void* d_packed, void* stream);

/* Measured peak of the tensor pipe for the instruction the ground-truth kernels issue (same tile shape and shared-memory
 * layout, nothing else in the kernel): kind 0 = tcgen05.mma kind::mxf4.block_scale (E2M1, M128 N240 K64), 1 = kind::i8
 * (M128 N256 K32).  *tops = 10^12 operations/s (2*M*N*K per instruction) over all SMs.  The roofline denominator of
 * the `ground-truth` path (new: MEASURED_PEAKS.json only carries a bf16 cuBLAS figure). */
int hb_tensor_peak(int device, int kind, double* tops);

/* Exact (distance, key) merge of n_lists per-shard lists laid out
 * (n_lists, nq, k) packed u64 -> (nq, k) packed (new for sharding; the order
 * is total so the merge equals a global Bruteforce + sort). */
int hb_topk_merge_device(const void* d_lists, uint32_t n_lists, uint64_t nq, uint32_t k, void* d_packed_out, void* stream);
/* Per-shard graph-search results (hb_knns_device outputs on a shard whose first row has global key
 * `key_offset`, the reference's `index -b start -l len` sub-range, cli/src/main.rs:596-602,136-140) -> packed
 * ascending lists with global keys, ready for an all-gather + hb_topk_merge_device.  Entries past found[i] become
 * UINT64_MAX. */
int hb_pack_device(const void* d_keys, const void* d_dists, const void* d_found, uint64_t nq, uint32_t k, uint64_t key_offset,
                   void* d_packed, void* stream);
/* packed (nq,k) -> keys, dists u64 (UINT64_MAX entries stay UINT64_MAX in both) */
int hb_unpack_device(const void* d_packed, uint64_t n, void* d_keys, void* d_dists, void* stream);

/* ---- synthetic sketches (bench/test input; integer-exact SimHash mixture) ---- */
typedef struct hb_synth_params {
  uint64_t seed_planes, seed_centres, seed_rows;
  uint32_t latent_dim, n_clusters, centre_scale, noise_scale, distinct_planes;
} hb_synth_params;
int hb_synth_device(const hb_synth_params* p, uint64_t start_row, uint64_t n_rows, void* d_out, void* stream);

/* ---- index construction (HNSWBuilder::extend_parallel, src/index/hnsw.rs:44-169) ---- */
typedef struct hb_build_params {
  uint32_t ef_construction; /* -c, default 96 (main.rs:519-521) */
  uint32_t connections;     /* -m, default 24 */
  uint32_t max_connections; /* -M, default 256 */
  uint32_t rayon_threads;   /* stands in for rayon::current_num_threads(): chunk = 32 * this, and it is mixed into the level seed */
  uint32_t serial_prefix;   /* points inserted one at a time before chunking starts; 0 = the reference's 50 000 (hnsw.rs:54) */
  uint32_t nsw;             /* != 0: NSWBuilder::extend_parallel (src/index/nsw.rs:224-322): one graph, no upper layers; the handle is an NSW index */
} hb_build_params;
int hb_build_hnsw(const uint64_t* data, uint64_t n, const hb_build_params* p, int device, int keep_host, hb_index** out);

/* ---- multi-GPU (new: the reference is one process on one host and its only scaling hook is the
 * `--start/--len` row sub-range of `index` / `ground-truth`, cli/src/main.rs:596-602,698-704,136-140) ---------------
 * REPLICATE: every device holds the whole index, Index::knns (src/index/mod.rs:28-42) splits the batch contiguously,
 *            outputs go to disjoint slices in input order; no data-path collective.
 * SHARD:     device r holds rows [start_r, start_r+len_r) (`-b start_r -l len_r`), searches them for all queries;
 *            per-shard (distance, key) lists are exchanged with one ncclAllGather and merged exactly. */
typedef enum hb_multi_mode { HB_MULTI_REPLICATE = 0, HB_MULTI_SHARD = 1 } hb_multi_mode;

/* One rank of an NCCL communicator bound to one device; for hosts that run one process per GPU. */
typedef struct hb_comm hb_comm;
#define HB_COMM_ID_BYTES 128
int hb_comm_unique_id(uint8_t id[HB_COMM_ID_BYTES]); /* rank 0 makes it; the host distributes it to every rank */
int hb_comm_init_rank(const uint8_t id[HB_COMM_ID_BYTES], int nranks, int rank, int device, hb_comm** out);
void hb_comm_free(hb_comm* c);
/* device time (ms) of the phases of the rank's last sharded call: [0] search, [1] all-gather, [2] merge */
int hb_comm_phase_ms(hb_comm* c, double ms[3]);
/* collective: `ground-truth` over row shards.  d_shard = this rank's rows (first global key = key_offset), d_queries
 * = the full batch on every rank; d_packed (nq, k) receives the GLOBAL result on every rank.  Asynchronous on stream. */
int hb_ground_truth_sharded_device(hb_comm* c, const void* d_shard, uint64_t n_local, uint64_t key_offset, const void* d_queries,
                                   uint64_t nq, uint32_t k, void* d_packed, void* stream);
/* collective: sharded HNSW.  idx = the index over this rank's rows (local keys). */
int hb_knns_sharded_device(hb_comm* c, const hb_index* idx, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                           uint32_t ef, void* d_packed, void* stream);
/* collective: replicate rank `root`'s index on every rank GPU-to-GPU (ncclBroadcast of the device arrays).  The root
 * passes src and gets it back in *out; the others pass NULL and own the replica they receive (no host copy). */
int hb_index_broadcast(hb_comm* c, hb_index* src, int root, hb_index** out);

/* One process driving N devices (what a Rust/C caller of this ABI uses: ncclCommInitAll + one host thread per device
 * inside the library). */
typedef struct hb_multi hb_multi;
int hb_multi_create(const int* devices, int ndev, hb_multi** out);
void hb_multi_free(hb_multi* m);
int hb_multi_device_count(const hb_multi* m);
uint64_t hb_multi_size(const hb_multi* m); /* Index::size over all devices' rows */
/* REPLICATE: read_index (main.rs:224-234) once onto devices[0], then broadcast over NVLink */
int hb_multi_load_idx(hb_multi* m, const char* path);
/* REPLICATE an index living on devices[0]; `src` stays owned by the caller and must outlive `m` */
int hb_multi_replicate(hb_multi* m, const hb_index* src);
/* SHARD: device r builds an HNSW over rows [r*n/N, (r+1)*n/N) of the host array (`index -b -l`), concurrently */
int hb_multi_build_sharded(hb_multi* m, const uint64_t* data, uint64_t n, const hb_build_params* p);
/* SHARD from files: paths[r] = shard r's .idx, key_offsets[r] = global key of its first row */
int hb_multi_load_shards(hb_multi* m, const char* const* paths, const uint64_t* key_offsets);
/* Index::knns through Indexes::search over all devices; HOST buffers, same contract as hb_knns */
int hb_multi_knns(hb_multi* m, const uint64_t* queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, uint64_t* keys,
                  uint64_t* dists, uint32_t* found);
/* `ground-truth` with the rows sharded over the devices; HOST buffers, same contract as hb_ground_truth */
int hb_multi_ground_truth(hb_multi* m, const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k,
                          uint64_t* keys, uint64_t* dists);
int hb_multi_phase_ms(hb_multi* m, double ms[3]);
const hb_index* hb_multi_index(const hb_multi* m, int r); /* device r's replica / shard, borrowed */

#ifdef __cplusplus
}
#endif
#endif /* HNSW_B200_H_ */
