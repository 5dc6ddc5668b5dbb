/* kdtree_b200.h — C ABI of libkdtree_b200.so (host core + sm_100a kernels).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; kd_last_error()
 *     gives the message of the last failure on the calling thread;
 *   - points are row-major float32 [n][dim]; point index i has the reference's
 *     ID i + 1 (+ id_offset in forest mode, kdtree_mpi.cpp:223);
 *   - a tree is `perm` (int32[n], IN-ORDER layout, values = point indices):
 *     the node of range [lo, lo+n) is perm[lo + n/2], left = [lo, lo+n/2),
 *     right = [lo+n/2+1, lo+n)  (shape rule of kdtree_sequential.cpp:51-56);
 *   - device entry points take raw device pointers and a cudaStream_t passed as
 *     void*; they never allocate: scratch comes from the caller (sizes from the
 *     *_workspace_bytes functions) or, for kd_build_gpu_noalloc only, from an
 *     array built into the library; they do not synchronise the stream unless
 *     stated.
 */
#ifndef KDTREE_B200_H_
#define KDTREE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* kd_last_error(void);
int kd_version(void);
/* -DKD_DEBUG builds of the library (libkdtree_b200_dbg.so) check every computed store offset of the build kernels on
 * the device; kd_debug_bounds_status() synchronises the device and fails with the code of the first assertion that
 * failed since the last call.  Release builds: kd_debug_enabled() == 0 and the status is always 0. */
int kd_debug_enabled(void);
int kd_debug_bounds_status(void);
void kd_free(void* p); /* releases buffers returned by the kd_format_* functions */

/* ---- problem generation (reference Utility.cpp:6-18, kdtree_mpi.cpp:19-41) ---- */
/* rows [n_rows][dim] of mt19937(seed) -> uniform_real_distribution<float>(-100,100), bit-exact. */
int kd_generate_uniform(int seed, int dim, int64_t n_rows, float* out);
/* rows [first_row, first_row + n_rows) of the same stream (O(first_row) skip-ahead). */
int kd_generate_uniform_rows(int seed, int dim, int64_t first_row, int64_t n_rows, float* out);
/* number of mismatches between the library's per-draw arithmetic and libstdc++'s distribution object (0 expected) */
int64_t kd_generator_selftest(int seed, int64_t count);
/* seeded Gaussian blobs (not in the reference; BASELINE.json config 4). */
int kd_generate_clustered(int seed, int dim, int64_t n_rows, int n_clusters, float sigma, float* out);

/* ---- CPU build / query / node materialisation (kdtree_sequential.cpp:30-136) ---- */
/* select-based build under the total order (coord[axis], index); threads <= 0 -> hardware count. */
int kd_build_cpu(const float* points, int64_t n, int dim, int32_t* perm, int threads);
/* the reference's formulation (sort first n-1, take n/2) with the same total order. */
int kd_build_cpu_sortref(const float* points, int64_t n, int dim, int32_t* perm);
/* reference DFS (near side first, strict <, far side iff da*da < best).  Outputs are optional (NULL). */
int kd_nearest_cpu(const float* points, int64_t n, int dim, const int32_t* perm, const float* queries,
                   int64_t nq, int32_t* out_index, float* out_dist2, float* out_dist, int64_t* out_visited);
/* Node{point,left,right} (Node.hpp:35-45) as three int32[n] arrays indexed by in-order slot; -1 = null. */
int kd_materialize_nodes_cpu(int64_t n, const int32_t* perm, int32_t* node_point, int32_t* left, int32_t* right);
int64_t kd_root_slot(int64_t n);
/* checks shape-independent invariants of a perm built with the reference rule; returns #violations in *bad. */
int kd_check_tree_cpu(const float* points, int64_t n, int dim, const int32_t* perm, int64_t* bad);

/* the same check with ties allowed on either side: 0 violations = a possible output of the reference's unstable sort */
int kd_check_tree_ties_cpu(const float* points, int64_t n, int dim, const int32_t* perm, int64_t* bad);

/* ---- text protocol (Utility.cpp:21-63,122-124; Node.cpp:16-28) ---- */
int kd_format_float(float v, char* buf, int cap);
int kd_format_result_line(int id, float distance, char* buf, int cap);
int kd_format_point(const float* coords, int dim, int id, char** out, size_t* len);
int kd_format_tree(const float* points, int64_t n, int dim, const int32_t* perm, int id_offset, char** out,
                   size_t* len);
int kd_format_head_and_leaves(const float* points, int64_t n, int dim, const int32_t* perm, int id_offset,
                              char** out, size_t* len);

/* ---- single-GPU build (new; semantics of kdtree_sequential.cpp:30-66) ---- */
size_t kd_build_workspace_bytes(int64_t n, int dim);
/* points: device float32 [n][dim]; perm: device int32[n]; tree_points (optional, may be NULL): device
 * float32 [n][dim] receiving the coordinates in in-order slot order; workspace: 256-byte aligned device memory
 * of kd_build_workspace_bytes(n, dim) bytes (the TMA bulk copies of the finisher need aligned columns). */
int kd_build_gpu(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points, void* workspace,
                 size_t workspace_bytes, void* stream);
/* The same build for callers that cannot allocate device memory (code running inside someone else's fixed
 * memory plan, or where cudaMalloc and its implicit synchronisation are off limits): the scratch is a 16 MiB
 * array built into the library, so nothing is allocated and no workspace is passed.  Accepts only problems with
 * kd_build_workspace_bytes(n, dim) <= 16 MiB (about 400 000 3-D points) and fails with a message otherwise.
 * The array is shared by all calls on a device and keeps the previous call's contents (the build re-zeroes what
 * it needs): calls that may overlap on the GPU race, serialise them.  The first call on each device resolves
 * the array's address; make it before any stream capture.  Every GPU that loads the library holds the 16 MiB
 * until exit.  No counterpart in the reference. */
int kd_build_gpu_noalloc(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points, void* stream);
/* kd_build_status for the built-in scratch: synchronises the stream and reports a device-side failure of the last
 * kd_build_gpu_noalloc(n, dim) on the current device (that caller has no workspace pointer to pass).  Allocates
 * nothing.  No counterpart in the reference. */
int kd_build_status_noalloc(int64_t n, int dim, void* stream);
/* synchronises the stream and reports a device-side failure of the last build that used `workspace` */
int kd_build_status(const void* workspace, int64_t n, int dim, void* stream);
/* number of kernel launches issued by the last kd_build_gpu call on this thread */
int64_t kd_last_launch_count(void);

/* ---- single-GPU query (new; semantics of kdtree_sequential.cpp:75-136) ---- */
/* tree_points: [n][dim] in slot order (from kd_build_gpu or kd_gather_rows_gpu). Outputs optional. */
int kd_nearest_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                   int64_t nq, int32_t* out_index, float* out_dist2, float* out_dist, int32_t* out_visited,
                   void* stream);
/* The same answers for high-dimensional trees, where the reference's pruning never fires and every query visits all
 * n nodes (its default 500000 x 128 problem, Utility.cpp:98-99): node-parallel exact distances (every sum in the
 * reference's order, kdtree_sequential.cpp:14-25), a per-query proof that nothing is pruned, the thread-per-query
 * traversal for the queries where the proof fails.  workspace: kd_nearest_dense_workspace_bytes(n, nq) bytes. */
size_t kd_nearest_dense_workspace_bytes(int64_t n, int64_t nq);
int kd_nearest_dense_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                         int64_t nq, int32_t* out_index, float* out_dist2, float* out_dist, int32_t* out_visited,
                         void* workspace, size_t workspace_bytes, void* stream);
/* Batched k-NN (extension; the reference has k = 1): the reference traversal with a bounded sorted list of the k
 * best nodes; outputs [nq][k] ascending, -1 / +inf padded; 1 <= k <= 32; k = 1 equals kd_nearest_*. */
int kd_knn_cpu(const float* points, int64_t n, int dim, const int32_t* perm, const float* queries, int64_t nq, int k,
               int32_t* out_index, float* out_dist2);
int kd_knn_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries, int64_t nq,
               int k, int32_t* out_index, float* out_dist2, void* stream);
/* Coherent batches: kd_query_order_gpu groups the queries by the sub-tree (top `levels` <= 16 levels) their search
 * enters first (own counting sort: key + histogram + scan + scatter kernels); the *_ordered entry points hand query
 * order[t] to thread t, so a warp's threads walk the same nodes.  Outputs are indexed by the original query number and
 * are identical to the unordered calls.  key_scratch: nq uint32; bins_scratch: (1 << levels) uint32. */
int kd_query_order_gpu(const float* tree_points, int64_t n, int dim, const float* queries, int64_t nq, int levels,
                       int32_t* order, void* key_scratch, void* bins_scratch, void* stream);
int kd_nearest_ordered_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                           int64_t nq, const int32_t* order, int32_t* out_index, float* out_dist2, float* out_dist,
                           int32_t* out_visited, void* stream);
int kd_knn_ordered_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                       int64_t nq, int k, const int32_t* order, int32_t* out_index, float* out_dist2, void* stream);
int kd_gather_rows_gpu(const float* points, const int32_t* perm, int64_t n, int dim, float* out, void* stream);
int kd_materialize_nodes_gpu(int64_t n, const int32_t* perm, int32_t* node_point, int32_t* left, int32_t* right,
                             void* stream);
/* per-axis min/max of a point set: out_min[dim], out_max[dim] (device). */
int kd_bbox_gpu(const float* points, int64_t n, int dim, float* out_min, float* out_max, void* stream);

/* ---- distributed build of ONE global tree over peer memory (new; oracle = the reference's sequential tree,
 * kdtree_sequential.cpp:30-66; the reference's own multi-rank program only builds W local trees) ----
 * Every rank owns a "symmetric heap" of kd_dist_heap_bytes() bytes with the same layout; heap[r] is rank r's
 * heap as addressable from the calling rank (CUDA IPC mapping, or plain pointers for W virtual ranks on one GPU).
 * The tree arrays live inside the heap at the offsets reported by kd_dist_offsets().  world must be 1, 2, 4 or 8.
 * dim <= 8: the coordinates travel with the points; larger dim (the reference's default is 128): every rank receives all
 * points once and only indices move (the heap then holds n_total x dim floats per rank). */
typedef struct {
    int world, rank, dim, with_tree_points;
    int64_t n_total;
    void* heap[8];
    void* mc_heap; /* optional NVLS multicast mapping of the same heaps (one store reaches every GPU); NULL = none */
} kd_dist_ctx;
size_t kd_dist_heap_bytes(int64_t n_total, int dim, int world, int with_tree_points);
int kd_dist_offsets(int64_t n_total, int dim, int world, int with_tree_points, size_t* perm_off, size_t* tp_off);
/* loads this rank's slice (rows [first_row, first_row + local_n) of the global point array) into its heap */
int kd_dist_init(const kd_dist_ctx* ctx, const float* local_points, int64_t local_n, int64_t first_row, void* stream);
/* one rank's whole build in program order (one process per GPU); `serial` must grow with every build on a heap */
int kd_dist_run(const kd_dist_ctx* ctx, uint64_t serial, void* stream);
/* the same build phase by phase, for W virtual ranks driven by one process (posts of all ranks before waits) */
int kd_dist_phase(const kd_dist_ctx* ctx, int level, int phase, int pass, uint64_t serial, void* stream);
/* the (phase, pass) pairs of one distributed level in launch order (sample bracket, digit passes, counts, partition,
 * left-maximum fix-up); drivers of virtual ranks launch each pair for every rank before the next pair */
int kd_dist_schedule(int* phases, int* passes, int cap);
/* synchronises the stream; out[8] = this rank's counters of the level that ran last: local count, elements going left,
 * holds-the-node flag, first left / right output offset, elements below the bracket, candidates, bracket-used flag */
int kd_dist_level_counts(const kd_dist_ctx* ctx, uint32_t* out, void* stream);
int kd_dist_local_build(const kd_dist_ctx* ctx, void* stream);
int kd_dist_done(const kd_dist_ctx* ctx, int post, uint64_t serial, void* stream);
/* last step after the done-barrier (gather mode, dim > 8: tree_points from the full point array; otherwise nothing) */
int kd_dist_finish(const kd_dist_ctx* ctx, void* stream);
/* synchronises the stream; non-zero if a device-side check or a peer wait failed */
int kd_dist_status(const kd_dist_ctx* ctx, void* stream);

/* ---- forest mode (kdtree_mpi.cpp:204-253): per-query minimum of the ranks' distances, folded into rank 0's
 * buffer with system-scope atomicMin over peer memory (replaces MPI_Reduce(MPI_MIN), kdtree_mpi.cpp:252-253).
 * Buffers are kd_forest_buf_bytes(nq) bytes from kd_symm_alloc; `epoch` must grow with every call. ---- */
size_t kd_forest_buf_bytes(int nq);
int kd_forest_min_reduce(void* root_buf, void* my_buf, int rank, int world, const float* dist, int nq, float* out,
                         uint64_t epoch, int* err, void* stream);

/* ---- bounding box over ranks (K12): per-axis min / max of a point set spread over W GPUs, all-reduced with system-scope
 * atomicMin / atomicMax into every rank's small symmetric buffer over peer memory (no collective library call).
 * bufs[r] = rank r's buffer (kd_peer_minmax_buf_bytes(dim) bytes from kd_symm_alloc) as addressable from this rank;
 * `epoch` must grow with every call.  No counterpart in the reference. ---- */
size_t kd_peer_minmax_buf_bytes(int dim);
int kd_peer_minmax_reduce(void* const* bufs, int rank, int world, const float* local_min, const float* local_max, int dim,
                          float* out_min, float* out_max, uint64_t epoch, int* err, void* stream);

/* ---- symmetric-heap runtime: allocation and CUDA IPC mapping (handle exchange is the caller's job) ---- */
int kd_symm_alloc(size_t bytes, void** ptr, void* handle64);
int kd_symm_open(const void* handle64, void** ptr);
int kd_symm_close(void* ptr);
int kd_symm_free(void* ptr);
int kd_symm_memset(void* ptr, size_t bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KDTREE_B200_H_ */
