/*
 * dgnn_cabi.h -- C ABI of the B200-native crystal-graph builder (libdgnn_b200.so).
 *
 * This is the drop-in boundary for the reference's graph layer.  The reference has no FFI for this path; its
 * boundary is the C++ class API in include/graph/ (SURVEY.md section 8b).  Each entry point below names the
 * reference interface it replaces (paths relative to the reference repository):
 *
 *   dgnn_build            NeighborList::NeighborList / build_with_pbc      src/graph/neighbor_list.cpp:14-66
 *                         + compute_num_images / create_image_cloud        src/graph/neighbor_list.cpp:68-94
 *                         + CrystalGraph ctor, edge part                   src/graph/crystal_graph.cpp:23-40
 *                         (batched over many structures: new surface, the reference is one structure per call)
 *   dgnn_rbf_expand       graph::gaussian_rbf                              src/graph/edge_features.cpp:7-24
 *   dgnn_rbf_expand_f64   CrystalGraph::edge_attr() storage (col-major)    src/graph/crystal_graph.cpp:30,37
 *   dgnn_copy_to_host     NeighborList::neighbors(i) / edge_index()        src/graph/neighbor_list.cpp:23-25,
 *                                                                          src/graph/crystal_graph.cpp:49-55
 *
 * The C++ facade that keeps the reference's class API (defect_gnn::graph::NeighborList, gaussian_rbf,
 * CrystalGraph) on top of this ABI lives in defect-gnn-cpp_b200/host/.  No torch / C++ types cross this
 * boundary: plain pointers, sizes and int status codes; errors never propagate as exceptions.
 *
 * Threading: a dgnn_ctx belongs to one CUDA device and one host thread at a time.  Results returned by
 * dgnn_build are device buffers owned by the ctx and stay valid until the next dgnn_build / dgnn_destroy on it.
 * There is no CPU fallback: every call fails with DGNN_ERR_CUDA if no sm_100 device is usable.
 */
#ifndef DGNN_CABI_H
#define DGNN_CABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGNN_ABI_VERSION 1

/* status codes (0 = ok, < 0 = error) */
enum {
    DGNN_OK = 0,
    DGNN_ERR_INVALID_ARG = -1,   /* null pointer, negative size, r_cutoff <= 0, non-monotone offsets ... */
    DGNN_ERR_CUDA = -2,          /* CUDA runtime error (message in dgnn_last_error) */
    DGNN_ERR_OOM = -3,           /* device allocation failed */
    DGNN_ERR_SINGULAR = -4,      /* a lattice matrix is singular / non-finite */
    DGNN_ERR_IMAGE_RANGE = -5,   /* reference image range n = ceil(r_c / l_min) + 1 exceeds 127 */
    DGNN_ERR_GEOMETRY = -6,      /* cell so thin that the stencil reach exceeds the supported 15 cells */
    DGNN_ERR_UNWRAPPED = -7,     /* atoms lie absurdly far outside the unit cell (halo grid too large) */
    DGNN_ERR_DEGREE = -8,        /* a row has more neighbours than the sorter supports (see DGNN_MAX_DEGREE) */
    DGNN_ERR_TOO_LARGE = -9,     /* batch exceeds 2^31-1 atoms / image-cloud points / cells: split it */
    DGNN_ERR_UNSUPPORTED = -10,
    DGNN_ERR_BUSY = -11          /* DLPack exports of this ctx's buffers are still alive (see dgnn_collated_dlpack) */
};

#define DGNN_MAX_DEGREE 4096
#define DGNN_UNCAPPED UINT64_MAX /* max_neighbors = SIZE_MAX: topology caller, betti_features.cpp:107 */

/* dgnn_batch.flags */
#define DGNN_INPUT_DEVICE 0x1u /* atom_offsets/x/y/z/lattice are device pointers on the ctx's device */
#define DGNN_SKIP_RBF 0x2u     /* build the neighbour list only (the viewer path never asks for features) */

typedef struct dgnn_ctx dgnn_ctx; /* opaque, one per CUDA device */

/* arena_bytes: initial size hint for the edge arena (0 = grow on demand). */
int dgnn_create(int device, size_t arena_bytes, dgnn_ctx** out);
void dgnn_destroy(dgnn_ctx* ctx);
const char* dgnn_last_error(const dgnn_ctx* ctx);
const char* dgnn_strerror(int code);
int dgnn_abi_version(void);

/* A batch of structures in SoA form (the repack of crystal::Structure, include/crystal/structure.hpp:12-38). */
typedef struct {
    int64_t n_structs;
    const int64_t* atom_offsets; /* [n_structs+1], atom_offsets[0] = 0 */
    const double* x;             /* [n_atoms] Cartesian Angstrom = Structure::atoms()[i].position, SoA */
    const double* y;
    const double* z;
    const double* lattice;       /* [n_structs*9] row-major, rows = a, b, c (vasp_parser.cpp:32-38) */
    double r_cutoff;             /* NeighborList r_cutoff   (neighbor_list.hpp:20, default 10.0) */
    uint64_t max_neighbors;      /* NeighborList max_neighbors (default 20; DGNN_UNCAPPED = no cap) */
    double epsilon;              /* NeighborList epsilon    (default 1e-10) */
    double rbf_r_cutoff;         /* CrystalGraph r_cutoff   (crystal_graph.hpp:18, default 10) */
    double dr;                   /* CrystalGraph dr (default 0.1); K = floor(rbf_r_cutoff / dr); dr <= 0: no RBF */
    uint32_t flags;
} dgnn_batch;

/* Result: device pointers owned by ctx.  CSR by source atom (global atom index across the batch); rows are
 * sorted by the canonical key (distance, dst, n_a, n_b, n_c) and truncated to max_neighbors
 * (neighbor_list.cpp:56-62; the reference's tie order is unspecified).
 * Edge order == CrystalGraph edge order (crystal_graph.cpp:32-40): source ascending, then row order. */
typedef struct {
    int64_t n_structs;
    int64_t n_atoms;
    int64_t n_edges;
    int32_t K;                     /* RBF bins (0 if skipped) */
    int32_t max_degree;            /* largest uncapped row in the batch */
    const int64_t* row_ptr;        /* [n_atoms+1] */
    const int32_t* dst;            /* [E] Neighbor::idx, local to its structure (edge_index(1,e)) */
    const int32_t* image;          /* [E] periodic image of the neighbour: int8 n_a | n_b<<8 | n_c<<16 */
    const double* dist;            /* [E] Neighbor::distance (fp64, bit-exact with the reference arithmetic) */
    const float* rbf;              /* [E*K] row-major fp32 edge_attr (consumer layout, utilities.py:69-71) */
    const int64_t* struct_edge_offsets; /* [n_structs+1] first edge of every structure */
    int64_t n_near_cutoff;         /* candidate pairs with |d - r_c| <= 1e-6 A (either side), parity report */
    int64_t n_tie_rows;            /* rows truncated through a distance tie (capped mode; R8) */
} dgnn_graphs;

int dgnn_build(dgnn_ctx* ctx, const dgnn_batch* batch, dgnn_graphs* out);

/* Host sinks for the offsets of every following dgnn_build on this ctx (NULL, NULL switches them off): row_ptr
 * [n_atoms+1] and struct_edge_offsets [n_structs+1] -- the counts / offsets a host-side consumer needs while the graphs
 * stay on the device -- are copied back on the ctx's copy stream as soon as the scan has produced them, i.e. WHILE the
 * fill pass runs, and have arrived when dgnn_build returns.  The buffers must hold the next build's sizes (pinned
 * memory for a true overlap). */
int dgnn_set_offset_sinks(dgnn_ctx* ctx, int64_t* host_row_ptr, int64_t* host_struct_edge_offsets);

/* Copy results of the last build to host buffers (any pointer may be NULL to skip it).  Sizes as in dgnn_graphs. */
int dgnn_copy_to_host(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t* row_ptr, int32_t* dst, int32_t* image,
                      double* dist, float* rbf, int64_t* struct_edge_offsets);

/* Copy the edges [e0, e1) of the last build (for chunked / per-structure host materialisation). */
int dgnn_copy_edges_to_host(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t e0, int64_t e1, int32_t* dst,
                            int32_t* image, double* dist, float* rbf);

/* Stand-alone fused Gaussian RBF expansion (edge_features.cpp:7-24) of n distances.
 * dist / out may be host or device pointers independently (device pointers must live on the ctx's device).
 * out: n * K floats row-major; returns K through *K_out. */
int dgnn_rbf_expand(dgnn_ctx* ctx, const double* dist, int64_t n, double r_cutoff, double dr, float* out,
                    int32_t* K_out);
/* fp64 variant writing the reference's own storage: column-major E x K doubles (crystal_graph.cpp:30,37). */
int dgnn_rbf_expand_f64(dgnn_ctx* ctx, const double* dist, int64_t n, double r_cutoff, double dr, double* out,
                        int32_t* K_out);

/* ---- Consumers of the last build's device-resident CSR (the callers either side of the path) --------
 * All of them take the dgnn_graphs of the LAST dgnn_build on this ctx; when that build read device inputs
 * (DGNN_INPUT_DEVICE) the caller's coordinate buffers must still be alive. */

/* f1: PyG-style batch collation, left in HBM for the training consumer.  Replaces CrystalGraph's node / edge-index
 * assembly (crystal_graph.cpp:15-21, 32-37) plus the planned train::BatchedGraph collation (PROJECT_ROADMAP.md:
 * 892-904) with the dtypes of the Python consumer (utilities.py:69-71: x float, edge_index long, edge_attr float). */
typedef struct {
    int64_t n_nodes;
    int64_t n_edges;
    int32_t F;                  /* node feature width (0: no embedding table given) */
    int32_t K;                  /* edge feature width (= dgnn_graphs::K) */
    const float* x;             /* [n_nodes, F] row-major fp32: x[i] = atom_embeddings.at(node_keys[i]) */
    const int64_t* edge_index;  /* [2, n_edges] row-major int64: row 0 = source node, row 1 = target node, both GLOBAL
                                 * (structure s's nodes start at atom_offsets[s]) */
    const int64_t* batch;       /* [n_nodes] structure index of every node */
    const float* edge_attr;     /* [n_edges, K] = dgnn_graphs::rbf */
} dgnn_collated;
/* node_keys: [n_atoms] int32, the key CrystalGraph passes to atom_embeddings.at() -- the reference passes
 * Atom::element, the POSCAR species index (structure.cpp:15), against a table keyed by atomic number
 * (json_parser.cpp:22-30); pass whichever the caller means, the lookup itself is the same.  Host pointer, or device
 * when flags has DGNN_INPUT_DEVICE.  emb_keys [n_keys] (host, distinct, 0 <= key < 65536) and emb_table
 * [n_keys, F] (host, row-major fp64, converted to fp32 once) are the std::map.  A node whose key is absent fails
 * with DGNN_ERR_INVALID_ARG (std::map::at would throw std::out_of_range, crystal_graph.cpp:20).
 * node_keys / emb_keys / emb_table may all be NULL (F = 0): only edge_index and batch are produced. */
int dgnn_collate(dgnn_ctx* ctx, const dgnn_graphs* g, const int32_t* node_keys, const int32_t* emb_keys,
                 const double* emb_table, int32_t n_keys, int32_t F, uint32_t flags, dgnn_collated* out);
int dgnn_copy_collated_to_host(dgnn_ctx* ctx, const dgnn_collated* c, float* x, int64_t* edge_index, int64_t* batch);

/* f1, zero-copy hand-over: one array of the collated batch as a DLPack tensor (DLManagedTensor*, dlpack.h ABI of the
 * "dltensor" capsule that torch.from_dlpack / cupy.from_dlpack / jax consume; kDLCUDA, row-major, no strides).
 * which: 0 = x [n_nodes, F] f32, 1 = edge_index [2, n_edges] i64, 2 = batch [n_nodes] i64, 3 = edge_attr [n_edges, K] f32.
 * The tensor BORROWS the ctx arena: while any export is alive (its deleter has not run) dgnn_build / dgnn_collate on
 * this ctx fail with DGNN_ERR_BUSY instead of overwriting memory a consumer still reads, and dgnn_destroy defers the
 * release of the context to the last deleter.  dgnn_live_exports tells how many are outstanding. */
int dgnn_collated_dlpack(dgnn_ctx* ctx, const dgnn_collated* c, int32_t which, void** managed_tensor);
int dgnn_live_exports(const dgnn_ctx* ctx);
/* runs the tensor's deleter (for producers whose capsule was never consumed) */
void dgnn_dlpack_release(void* managed_tensor);

/* f4: the viewer's flat getters (wasm_bindings.cpp:206-294) in one kernel + one D2H per array: sources / targets
 * [E] int32 (atom indices local to the structure), distances [E] float32, displacements [3E] float32
 * (static_cast<float> of Neighbor::displacement = cloud_pos - query_pos, neighbor_list.cpp:51).  Host outputs,
 * any may be NULL. */
int dgnn_edge_views(dgnn_ctx* ctx, const dgnn_graphs* g, int32_t* sources, int32_t* targets, float* distances,
                    float* displacements);

/* f2: per-atom neighbourhood point clouds of the Betti caller (betti_features.cpp:65-75) for the atoms [a0, a1):
 * atom a owns the points [cloud_ptr[a - a0], cloud_ptr[a - a0 + 1]) of `cloud` ([n_points, 3] fp64): first its own
 * position, then position + displacement of every neighbour in row order.  n_points = (row_ptr[a1] - row_ptr[a0])
 * + (a1 - a0).  Host outputs. */
int dgnn_neighborhood_clouds(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t a0, int64_t a1, double* cloud,
                             int64_t* cloud_ptr);

/* f2: the pairwise-distance tables the Betti caller builds from every cloud (src/topology/ripser_wrapper.cpp:60-66:
 * sq_norms + sq_norms^T - 2 P P^T, cwiseMax(0), cwiseSqrt, in Eigen's FMA-free operation order), for the atoms [a0, a1):
 *   mode 0: the full (deg+1) x (deg+1) fp64 table, column-major -- the Eigen::MatrixXd `dist`;
 *   mode 1: its compressed lower triangle in fp32, entries (i, 0..i-1) for i = 1..deg -- the `distances` vector
 *           compute_persistence_from_distances hands to Ripser (:18-24).
 * Atom a0 + i owns values [table_ptr[i], table_ptr[i+1]).  Outputs are DEVICE pointers owned by the ctx (a C2 atom's
 * full table is 207 kB: chunk the atom range), valid until the next call on this ctx; copy helper below. */
typedef struct {
    int64_t n_atoms;
    int64_t n_values;
    const void* values;        /* double (mode 0) or float (mode 1), device */
    const int64_t* table_ptr;  /* [n_atoms + 1], device */
    int32_t mode;
} dgnn_distance_tables_out;
int dgnn_distance_tables(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t a0, int64_t a1, int32_t mode,
                         dgnn_distance_tables_out* out);
int dgnn_copy_distance_tables_to_host(dgnn_ctx* ctx, const dgnn_distance_tables_out* t, void* values,
                                      int64_t* table_ptr);

/* Device time (CUDA events on the ctx stream, kernels only -- no H2D / D2H) and algorithmic bytes of the last
 * dgnn_collate / dgnn_edge_views / dgnn_neighborhood_clouds call. */
typedef struct {
    float device_ms;
    int32_t kernel_launches;
    int64_t bytes_read;    /* algorithmic bytes the kernels read ... */
    int64_t bytes_written; /* ... and wrote */
} dgnn_consumer_info;
int dgnn_last_consumer_info(const dgnn_ctx* ctx, dgnn_consumer_info* out);

/* Device timings of the phases of the last dgnn_build (CUDA events on the ctx stream), milliseconds:
 * bin = geometry + cell-list binning (K1), dist = count + scan + fill/sort (K2), rbf = K3, h2d = input upload,
 * total = first to last event. */
typedef struct {
    float h2d_ms, bin_ms, count_ms, fill_ms, rbf_ms, total_ms;
    int32_t kernel_launches; /* kernels launched by the last build */
    int32_t n_items;         /* work items (structure x source-cell block) */
    int64_t cloud_points;    /* pruned periodic image cloud size */
    int32_t staged_tile_cap; /* shared-memory tile capacity used (points) */
    int32_t row_cap;         /* shared-memory row buffer capacity used */
    int32_t n_cell_structs;  /* structures routed to the cell-list path (the rest took the direct path) */
    int32_t direct_warps;    /* warps per CTA of the direct fill pass */
    int32_t rbf_fused;       /* 1: features were written by the fill pass's fused epilogue (rbf_ms is then 0) */
    int32_t hit_cap;         /* per-row hit-list slots the count pass wrote (0: the fill pass repeated the traversal) */
    int32_t onepass;         /* 1: single-pass build (no count pass; row offsets by look-back), 2: its estimated edge
                              * arena was too small and the fill pass rewrote every row, 3: its look-back timed out (SMs
                              * held by other work) and the batch was rebuilt with count + fill, 0: count + fill */
    int32_t onepass_fallback_rows; /* rows that outgrew the single-pass row buffer and were completed by a fill pass */
    float tie_run_mean;      /* sampled rows: mean number of entries sharing a truncated sort key (1.0: none; 0: the
                              * build did not probe).  Ideal lattices put whole coordination shells on one key */
    int32_t full_tie_sort;   /* 1: the uncapped fill pass ran with the warp-cooperative tie ranking (tie_run_mean >= 3) */
    int32_t split_big_rows;  /* > 0: skewed degrees -- this many rows beyond the full-occupancy row buffer (row_cap) were
                              * filled by a second launch with a buffer for the largest row; 0: one fill launch */
} dgnn_build_info;
int dgnn_last_build_info(const dgnn_ctx* ctx, dgnn_build_info* out);

/* ---- streamed build: jobs larger than one edge arena (BASELINE config 5: 10^6 structures, ~4 TB of graphs) ---------
 * Replaces the reference's per-file driver loop (src/preprocess/preprocess_betti.cpp:37-79: parse -> Structure ->
 * NeighborList per file, nothing retained) for batched jobs.  The job is cut into chunks by the caller's SOURCE; chunk c
 * is built into arena c & 1 (the ctx and an internal twin with the same options) by a producer thread while the SINK
 * consumes chunk c - 1 on the calling thread, so upload + count + fill of the next chunk overlap the consumer.
 *   source(user, c, batch): fill *batch (host or device inputs, as for dgnn_build) for chunk c = 0, 1, ...; return 1, or
 *                           0 when the job is finished, < 0 on error.  Called on the producer thread; the buffers of
 *                           chunk c must stay valid until source is asked for chunk c + 1.
 *   sink(user, c, graphs, arena): the chunk's graphs (device pointers, as from dgnn_build) and the context that owns
 *                           them -- pass `arena` (not the ctx given to dgnn_build_stream) to dgnn_copy_to_host,
 *                           dgnn_collate, dgnn_distance_tables ... for this chunk; the other arena is busy building the
 *                           next one.  Valid until sink returns: the arena is then rebuilt for chunk c + 2.  Called on
 *                           the calling thread, in chunk order.  Return 0 to continue.  May be NULL (build and drop).
 * Chunk size is the caller's: what one arena holds (176 B per edge at K = 40: 10 k C2 structures = 45 GB). */
typedef int (*dgnn_stream_source)(void* user, int64_t chunk, dgnn_batch* batch);
typedef int (*dgnn_stream_sink)(void* user, int64_t chunk, const dgnn_graphs* graphs, dgnn_ctx* arena);
typedef struct {
    int64_t n_chunks, n_structs, n_atoms, n_edges;
    double wall_ms;          /* first source call to last sink return (host clock) */
    double build_device_ms;  /* sum of the chunks' device build times (CUDA events) */
    double h2d_ms;           /* of which input upload */
    double sink_ms;          /* host time spent inside the sink */
    int64_t kernel_launches;
} dgnn_stream_info;
int dgnn_build_stream(dgnn_ctx* ctx, dgnn_stream_source source, void* source_user, dgnn_stream_sink sink,
                      void* sink_user, dgnn_stream_info* info);

/* Number of successful dgnn_build calls on this ctx so far: a caller that kept device pointers of a build can tell
 * whether they are still the current result (the facade's CrystalGraph reuses its NeighborList's device distances). */
uint64_t dgnn_build_serial(const dgnn_ctx* ctx);

/* Page-locked host memory for input staging (the chunked upload of dgnn_build only overlaps with the count pass when
 * the host buffers are pinned).  Used by the facade's bulk POSCAR ingest, which parses straight into such buffers. */
int dgnn_host_alloc(size_t bytes, void** out);
void dgnn_host_free(void* p);

/* ---- allocation-free row offsets ------------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  A variant of the degree scan that
 * dgnn_build runs between its count and fill passes: row_ptr[i] = degrees[0] + ... + degrees[i - 1] for i = 0 ... n_rows
 * (row_ptr has n_rows + 1 entries; row_ptr[n_rows] = number of edges).  `degrees` and `row_ptr` are DEVICE pointers on
 * `device`; the kernels are issued on `cuda_stream` (a cudaStream_t, NULL = default stream) and the call returns
 * without synchronising.  It needs no dgnn_ctx and calls no cudaMalloc / cudaFree / allocating routine: the scan's
 * per-tile sums live in a fixed __device__ array of 65 544 bytes inside the library (every GPU that loads the library
 * holds it until the process exits).
 *   Limit: n_rows <= DGNN_NOALLOC_MAX_ROWS (2^24); more is rejected with DGNN_ERR_TOO_LARGE.
 *   All calls on one device share that array: calls that overlap on the GPU (different streams, concurrent host
 *   threads) race -- issue them on one stream or order them with events.  Nothing in it needs clearing between calls.
 *   err / err_len: optional buffer for the message of a failing call (may be NULL / 0).
 * Makes `device` the calling thread's current device, like every other entry point here. */
#define DGNN_NOALLOC_MAX_ROWS (1LL << 24)
int dgnn_row_offsets_noalloc(int device, const int32_t* degrees, int64_t n_rows, int64_t* row_ptr, void* cuda_stream,
                             char* err, size_t err_len);

/* The CUDA stream (cudaStream_t) all work of this ctx is issued on, for event timing by callers. */
void* dgnn_stream(dgnn_ctx* ctx);

/* Options (all optional; defaults are the measured best):
 *   cell_frac         target grid-cell width of the cell-list path as a fraction of r_cutoff (default 0.5)
 *   direct_max_atoms  structures with more atoms use the cell-list path instead of the direct path (default 2048)
 *   hit_cap           per-row hit-list slots recorded by the count pass (-1 auto from the density, 0 off)
 *   onepass           single-pass build (no count pass) of batches without cell-list structures: -1 auto (small
 *                     batches), 0 off (count + fill), 1 wherever legal
 *   fuse_rbf          1 (default): features written by the fill pass's epilogue; 0: separate RBF launch
 *   upload_chunks     host inputs cross PCIe in this many structure-aligned chunks behind the count pass (default 4)
 *   cell32            1 (default): cell-list structures use fp32 tiles staged in shared memory with the fp64 decision in
 *                     a guard band; 0: the fp64 traversal (A/B reference)
 *   sym_count         1 (default): the count pass sweeps j >= i and decides both directions of a pair
 * Tuning and testing knobs (tile sizes, sort buckets, look-back spin limit ...) are spelled "tune.<name>"; they are
 * listed next to dgnn_set_option in csrc/dgnn_options.cu and are not part of the interface. */
int dgnn_set_option(dgnn_ctx* ctx, const char* name, double value);

#ifdef __cplusplus
}
#endif
#endif /* DGNN_CABI_H */
