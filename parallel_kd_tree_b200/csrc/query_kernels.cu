// Query-side kernels for sm_100a: 1-NN traversal (K10), exact distance (K11),
// node materialisation (K9) and bounding box (K12).
//
// The traversal reproduces the reference's `nearest` (kdtree_sequential.cpp:75-136)
// node for node: near side first (da < 0 -> left, else right), strict `<` updates,
// far side visited iff da*da < best.  The reference's trees are not valid kd-trees
// (SURVEY.md §0 quirk 1), so the visiting order is part of the result and must not
// be rearranged.  Distances follow kdtree_sequential.cpp:14-25: one sequential float
// sum, separate multiply and add (__fmul_rn/__fadd_rn, never contracted), sqrt by
// __fsqrt_rn (Node.cpp:36-38).
//
// The tree is the in-order permutation, so children are computed, never loaded:
// node(lo, n) = slot lo + n/2, left = (lo, n/2), right = (lo + n/2 + 1, n - n/2 - 1).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <string>

#include "../../include/kdtree_b200.h"
#include "kd_common.h"
#include "kd_error.h"

namespace kd {

__device__ __forceinline__ float dist2_exact(const float* __restrict__ a, const float* __restrict__ b, int dim) {
    float dist = 0.f;
    for (int i = 0; i < dim; ++i) {
        float t = __fsub_rn(a[i], b[i]);
        dist = __fadd_rn(dist, __fmul_rn(t, t));
    }
    return dist;
}

constexpr int kMaxDepth = 33;

// Row access for the traversal kernels.  DIM > 0: the query lives in registers, a node row is fetched with the widest
// aligned vector load and the split-axis coordinate is picked with a select chain (no dynamic register index);
// DIM == 0: any dimension, coordinates are re-read from memory.
template <int DIM>
struct RowOps {
    float q[DIM];
    __device__ __forceinline__ void load_query(const float* __restrict__ qp, int) {
#pragma unroll
        for (int i = 0; i < DIM; ++i) q[i] = qp[i];
    }
    __device__ __forceinline__ void visit(const float* __restrict__ pp, int, int axis, float& de, float& da) const {
        float p[DIM];
        if (DIM % 4 == 0) {
#pragma unroll
            for (int i = 0; i < DIM; i += 4) {
                const float4 v = __ldg(reinterpret_cast<const float4*>(pp + i));
                p[i] = v.x; p[i + 1] = v.y; p[i + 2] = v.z; p[i + 3] = v.w;
            }
        } else if (DIM % 2 == 0) {
#pragma unroll
            for (int i = 0; i < DIM; i += 2) {
                const float2 v = __ldg(reinterpret_cast<const float2*>(pp + i));
                p[i] = v.x; p[i + 1] = v.y;
            }
        } else {
#pragma unroll
            for (int i = 0; i < DIM; ++i) p[i] = __ldg(pp + i);
        }
        float dist = 0.f, pa = p[0], qa = q[0];
#pragma unroll
        for (int i = 0; i < DIM; ++i) {
            const float t = __fsub_rn(p[i], q[i]);
            dist = __fadd_rn(dist, __fmul_rn(t, t));
            if (i > 0 && axis == i) { pa = p[i]; qa = q[i]; }
        }
        de = dist;
        da = __fsub_rn(qa, pa);
    }
};
template <>
struct RowOps<0> {
    const float* qp_;
    __device__ __forceinline__ void load_query(const float* __restrict__ qp, int) { qp_ = qp; }
    __device__ __forceinline__ void visit(const float* __restrict__ pp, int dim, int axis, float& de, float& da) const {
        de = dist2_exact(pp, qp_, dim);
        da = __fsub_rn(qp_[axis], pp[axis]);
    }
};

// one thread per query; explicit stack of postponed far sides.  `order` (optional) lists the queries so that the
// threads of a warp hold queries that walk the same part of the tree (k_qkey / k_qscatter below): their loads then
// hit the same lines and their loops stay in step.  The split axis is carried along (axis + 1 mod dim per level)
// instead of being recomputed with an integer division per node.
template <int DIM>
__global__ void __launch_bounds__(128) k_nearest(const float* __restrict__ tp, const int32_t* __restrict__ perm,
                                                 unsigned N, int dim, const float* __restrict__ queries, long long nq,
                                                 int32_t* __restrict__ out_index, float* __restrict__ out_d2,
                                                 float* __restrict__ out_d, int32_t* __restrict__ out_visited,
                                                 const int* __restrict__ need, const int32_t* __restrict__ order) {
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    const long long q = order ? (long long)order[t] : t;
    if (need && !need[q]) return;   // already answered by the dense path
    if (DIM > 0) dim = DIM;
    RowOps<DIM> R;
    R.load_query(queries + (size_t)q * dim, dim);
    unsigned st_lo[kMaxDepth], st_n[kMaxDepth];
    float st_da2[kMaxDepth];
    unsigned short st_axis[kMaxDepth];
    int sp = 0;
    unsigned best = N >> 1;  // nearest_neighbor(): best starts at the root (kdtree_sequential.cpp:133-136)
    float best_d, da_unused;
    R.visit(tp + (size_t)best * dim, dim, 0, best_d, da_unused);
    unsigned lo = 0, cnt = N;
    int axis = 0, visited = 0;
    for (;;) {
        while (cnt > 0) {
            unsigned m = cnt >> 1, slot = lo + m;
            float de, da;
            R.visit(tp + (size_t)slot * dim, dim, axis, de, da);
            ++visited;
            if (de < best_d) { best_d = de; best = slot; }
            unsigned rlo = slot + 1, rn = cnt - m - 1;
            float da2 = __fmul_rn(da, da);
            axis = axis + 1 == dim ? 0 : axis + 1;
            st_da2[sp] = da2;
            st_axis[sp] = (unsigned short)axis;
            if (da < 0.f) {
                st_lo[sp] = rlo; st_n[sp] = rn;
                cnt = m;
            } else {
                st_lo[sp] = lo; st_n[sp] = m;
                lo = rlo; cnt = rn;
            }
            ++sp;
        }
        bool resumed = false;
        while (sp > 0) {
            --sp;
            if (st_n[sp] > 0 && st_da2[sp] < best_d) {
                lo = st_lo[sp]; cnt = st_n[sp]; axis = st_axis[sp];
                resumed = true;
                break;
            }
        }
        if (!resumed) break;
    }
    if (out_index) out_index[q] = perm[best];
    if (out_d2) out_d2[q] = best_d;
    if (out_d) out_d[q] = __fsqrt_rn(best_d);
    if (out_visited) out_visited[q] = visited;
}

// ---- query ordering: key = the near-side path through the top `levels` levels of the tree ----
// (queries with the same key descend into the same sub-tree first; a counting sort on the key groups them)
__global__ void __launch_bounds__(256) k_qkey(const float* __restrict__ tp, unsigned N, int dim,
                                              const float* __restrict__ queries, long long nq, int levels,
                                              unsigned* __restrict__ key, unsigned* __restrict__ hist) {
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float* qp = queries + (size_t)q * dim;
    unsigned lo = 0, cnt = N, k = 0;
    int axis = 0;
    for (int l = 0; l < levels; ++l) {
        k <<= 1;
        if (cnt > 0) {
            const unsigned m = cnt >> 1, slot = lo + m;
            const float da = __fsub_rn(qp[axis], tp[(size_t)slot * dim + axis]);
            if (da < 0.f) cnt = m; else { k |= 1u; lo = slot + 1; cnt = cnt - m - 1; }
            axis = axis + 1 == dim ? 0 : axis + 1;
        }
    }
    key[q] = k;
    atomicAdd(&hist[k], 1u);
}

// exclusive scan of hist[0 .. bins) in place, one CTA (bins <= 65536)
__global__ void __launch_bounds__(1024) k_qscan(unsigned* __restrict__ hist, unsigned bins) {
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry;
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid == 0) carry = 0u;
    __syncthreads();
    for (unsigned b0 = 0; b0 < bins; b0 += 1024u) {
        const unsigned i = b0 + tid, v = i < bins ? hist[i] : 0u;
        unsigned inc = v;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= (unsigned)o) inc += t; }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        unsigned base = carry;
        for (unsigned w = 0; w < warp; ++w) base += wsum[w];
        if (i < bins) hist[i] = base + inc - v;
        __syncthreads();
        if (tid == 1023) carry = base + inc;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_qscatter(const unsigned* __restrict__ key, unsigned* __restrict__ cursor,
                                                  long long nq, int32_t* __restrict__ order) {
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    order[atomicAdd(&cursor[key[q]], 1u)] = (int32_t)q;
}

// K9: Node{point,left,right} arrays indexed by in-order slot (Node.hpp:35-45); -1 = nullptr
__global__ void k_materialize(unsigned N, const int32_t* __restrict__ perm, int32_t* __restrict__ node_point,
                              int32_t* __restrict__ left, int32_t* __restrict__ right) {
    unsigned slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= N) return;
    unsigned lo = 0, n = N;
    for (;;) {
        unsigned m = n >> 1, node = lo + m, rn = n - m - 1;
        if (slot == node) {
            node_point[slot] = perm[slot];
            left[slot] = m ? (int32_t)(lo + (m >> 1)) : -1;
            right[slot] = rn ? (int32_t)(node + 1 + (rn >> 1)) : -1;
            return;
        }
        if (slot < node) { n = m; } else { lo = node + 1; n = rn; }
    }
}

// K12: per-axis min/max.  Thread t always sees axis t % dim because the grid stride is a multiple of dim.
__global__ void k_bbox_init(unsigned* mn, unsigned* mx, int dim) {
    int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d < dim) { mn[d] = 0xffffffffu; mx[d] = 0u; }
}

__global__ void __launch_bounds__(256) k_bbox(const float* __restrict__ pts, size_t total, int dim, size_t stride,
                                              unsigned* mn, unsigned* mx) {
    // per-thread running min / max, then ONE shared-memory atomic pair per thread and one global atomic pair per
    // CTA and axis (round 1 issued a global atomic pair per thread)
    __shared__ unsigned s_mn[256], s_mx[256];
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool small = dim <= 256;
    if (small) { s_mn[threadIdx.x] = 0xffffffffu; s_mx[threadIdx.x] = 0u; }
    __syncthreads();
    unsigned lo = 0xffffffffu, hi = 0u;
    if (t < stride)
        for (size_t e = t; e < total; e += stride) {
            const unsigned k = ordered_key(pts[e]);
            lo = min(lo, k);
            hi = max(hi, k);
        }
    const int axis = (int)(t % (size_t)dim);
    if (lo <= hi) {
        if (small) { atomicMin(&s_mn[axis], lo); atomicMax(&s_mx[axis], hi); }
        else { atomicMin(&mn[axis], lo); atomicMax(&mx[axis], hi); }
    }
    __syncthreads();
    if (small && (int)threadIdx.x < dim && s_mn[threadIdx.x] <= s_mx[threadIdx.x]) {
        atomicMin(&mn[threadIdx.x], s_mn[threadIdx.x]);
        atomicMax(&mx[threadIdx.x], s_mx[threadIdx.x]);
    }
}

__global__ void k_bbox_finish(unsigned* mn, unsigned* mx, int dim) {
    int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= dim) return;
    auto back = [](unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); };
    reinterpret_cast<float*>(mn)[d] = back(mn[d]);
    reinterpret_cast<float*>(mx)[d] = back(mx[d]);
}


// ------------------------------------------------------------------------------------------------
// K10/K11, high-dimensional path.  In high dimension the reference's pruning test da*da < best
// (kdtree_sequential.cpp:118) never fails: one squared coordinate difference is far below any sum of `dim`
// of them, so every query walks all N nodes (500 000 of 500 000 at dim 128) and a thread per
// query would compute N dim-long sums one after the other.  The traversal ORDER, however, only matters for
// (a) which subtrees are pruned and (b) which of several equally distant nodes is met first.  So:
//   k_nn_dense    node-parallel: one thread per node runs the full sequential dim-term sum (the reference's
//                 rounding, kdtree_sequential.cpp:14-25) for up to 16 queries at once, tiles of 256 nodes x 32
//                 coordinates staged in shared memory; per query the smallest distance and the largest da*da
//                 over ALL nodes meet in two global atomics per CTA;
//   k_nn_collect  the nodes that attain the smallest distance (normally one);
//   k_nn_resolve  if max da*da < min distance, then best >= min distance > da*da at every test the reference
//                 would make: nothing is pruned, all N nodes are visited, and the answer is the node attaining
//                 the minimum that the near-first DFS meets first (strict <, kdtree_sequential.cpp:91); ties are
//                 ordered by walking down from the root to where two candidates part.  Otherwise the query is
//                 handed to the DFS kernel (k_nearest), which is exact in every case.
// ------------------------------------------------------------------------------------------------
constexpr int kDenseQB = 16, kDenseRows = 256, kDenseChunk = 32, kDenseMaxCand = 64;

struct DenseWs {
    float* dall;          // [kDenseQB][N] distances of the current batch
    unsigned* gmin;       // [kDenseQB] smallest distance (float bits; distances are >= 0)
    unsigned* gmaxda;     // [kDenseQB] largest da*da
    unsigned* ccnt;       // [kDenseQB] nodes attaining the minimum
    unsigned* cand;       // [kDenseQB][kDenseMaxCand]
    int* need;            // [nq] 1 = the DFS kernel has to answer this query
};

__device__ __forceinline__ int slot_depth(unsigned N, unsigned slot) {
    unsigned lo = 0, n = N;
    for (int d = 0;; ++d) {
        const unsigned m = n >> 1, s = lo + m;
        if (slot == s) return d;
        if (slot < s) n = m; else { lo = s + 1; n = n - m - 1; }
    }
}

__global__ void k_nn_init(DenseWs W) {
    const int t = threadIdx.x;
    if (t < kDenseQB) { W.gmin[t] = 0xffffffffu; W.gmaxda[t] = 0u; W.ccnt[t] = 0u; }
}

__global__ void __launch_bounds__(kDenseRows) k_nn_dense(const float* __restrict__ tp, unsigned N, int dim,
                                                         const float* __restrict__ queries, int nqb, DenseWs W) {
    __shared__ float tile[kDenseRows][kDenseChunk + 1];
    __shared__ float qs[kDenseQB][kDenseChunk];
    __shared__ unsigned wmin[kDenseRows / 32][kDenseQB], wmax[kDenseRows / 32][kDenseQB];
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const unsigned s0 = blockIdx.x * kDenseRows, slot = s0 + tid;
    const unsigned rows = min((unsigned)kDenseRows, N - s0);
    float acc[kDenseQB];
#pragma unroll
    for (int q = 0; q < kDenseQB; ++q) acc[q] = 0.f;
    for (int c0 = 0; c0 < dim; c0 += kDenseChunk) {
        const int cw = min(kDenseChunk, dim - c0);
        __syncthreads();
        for (unsigned r = warp; r < rows; r += kDenseRows / 32)   // one 128-byte run of a row per warp
            if ((int)lane < cw) tile[r][lane] = tp[(size_t)(s0 + r) * dim + c0 + lane];
        for (unsigned e = tid; e < (unsigned)(kDenseQB * kDenseChunk); e += kDenseRows) {
            const int q = e / kDenseChunk, d = e % kDenseChunk;
            qs[q][d] = (q < nqb && d < cw) ? queries[(size_t)q * dim + c0 + d] : 0.f;
        }
        __syncthreads();
        if (tid < rows) {
            for (int d = 0; d < cw; ++d) {     // coordinates in order: the sum is the reference's, term by term
                const float x = tile[tid][d];
#pragma unroll
                for (int q = 0; q < kDenseQB; ++q) {
                    const float t = __fsub_rn(x, qs[q][d]);
                    acc[q] = __fadd_rn(acc[q], __fmul_rn(t, t));
                }
            }
        }
    }
    unsigned da2[kDenseQB];
    if (tid < rows) {
        const int axis = slot_depth(N, slot) % dim;
        const float xa = tp[(size_t)slot * dim + axis];
#pragma unroll
        for (int q = 0; q < kDenseQB; ++q) {
            float v = 0.f;
            if (q < nqb) {
                const float da = __fsub_rn(queries[(size_t)q * dim + axis], xa);
                v = __fmul_rn(da, da);
                W.dall[(size_t)q * N + slot] = acc[q];
            }
            da2[q] = __float_as_uint(v);
        }
    }
#pragma unroll
    for (int q = 0; q < kDenseQB; ++q) {
        const bool on = tid < rows && q < nqb;
        const unsigned mn = __reduce_min_sync(0xffffffffu, on ? __float_as_uint(acc[q]) : 0xffffffffu);
        const unsigned mx = __reduce_max_sync(0xffffffffu, on ? da2[q] : 0u);
        if (lane == 0) { wmin[warp][q] = mn; wmax[warp][q] = mx; }
    }
    __syncthreads();
    if (tid < (unsigned)nqb) {
        unsigned mn = 0xffffffffu, mx = 0u;
        for (int w = 0; w < kDenseRows / 32; ++w) { mn = min(mn, wmin[w][tid]); mx = max(mx, wmax[w][tid]); }
        atomicMin(&W.gmin[tid], mn);
        atomicMax(&W.gmaxda[tid], mx);
    }
}

__global__ void __launch_bounds__(256) k_nn_collect(unsigned N, int nqb, DenseWs W) {
    const unsigned slot = blockIdx.x * 256u + threadIdx.x;
    if (slot >= N) return;
    for (int q = 0; q < nqb; ++q)
        if (__float_as_uint(W.dall[(size_t)q * N + slot]) == W.gmin[q]) {
            const unsigned p = atomicAdd(&W.ccnt[q], 1u);
            if (p < (unsigned)kDenseMaxCand) W.cand[q * kDenseMaxCand + p] = slot;
        }
}

// true if the near-first DFS of query qp meets slot a before slot b (a != b)
__device__ bool dfs_before(const float* __restrict__ tp, unsigned N, int dim, const float* __restrict__ qp, unsigned a,
                           unsigned b) {
    unsigned lo = 0, n = N;
    for (int depth = 0;; ++depth) {
        const unsigned m = n >> 1, s = lo + m;
        if (a == s) return true;
        if (b == s) return false;
        const bool al = a < s, bl = b < s;
        if (al != bl) {
            const int axis = depth % dim;
            const float da = __fsub_rn(qp[axis], tp[(size_t)s * dim + axis]);
            return al == (da < 0.f);   // the near side (left iff da < 0) is walked first
        }
        if (al) n = m; else { lo = s + 1; n = n - m - 1; }
    }
}

__global__ void k_nn_resolve(const float* __restrict__ tp, const int32_t* __restrict__ perm, unsigned N, int dim,
                             const float* __restrict__ queries, int nqb, DenseWs W, int* __restrict__ need,
                             int32_t* __restrict__ out_index, float* __restrict__ out_d2, float* __restrict__ out_d,
                             int32_t* __restrict__ out_visited) {
    const int q = threadIdx.x;
    if (q >= nqb) return;
    const float dmin = __uint_as_float(W.gmin[q]), damax = __uint_as_float(W.gmaxda[q]);
    const unsigned c = W.ccnt[q];
    if (!(damax < dmin) || c == 0u || c > (unsigned)kDenseMaxCand) { need[q] = 1; return; }
    const float* qp = queries + (size_t)q * dim;
    unsigned best = W.cand[q * kDenseMaxCand];
    for (unsigned k = 1; k < c; ++k) {
        const unsigned s = W.cand[q * kDenseMaxCand + k];
        if (dfs_before(tp, N, dim, qp, s, best)) best = s;
    }
    need[q] = 0;
    if (out_index) out_index[q] = perm[best];
    if (out_d2) out_d2[q] = dmin;
    if (out_d) out_d[q] = __fsqrt_rn(dmin);
    if (out_visited) out_visited[q] = (int32_t)N;
}

}  // namespace kd

extern "C" {

size_t kd_nearest_dense_workspace_bytes(int64_t n, int64_t nq) {
    if (n <= 0 || nq < 0) return 0;
    return (size_t)kd::kDenseQB * (size_t)n * 4 + (size_t)kd::kDenseQB * (3 + kd::kDenseMaxCand) * 4 + (size_t)nq * 4 + 1024;
}

/* 1-NN with the reference's answers (kdtree_sequential.cpp:75-136) for workloads where pruning does not fire
 * (high dimension): node-parallel exact distances, a proof per query that the reference visits every node, and the
 * thread-per-query DFS for whatever cannot be proven.  Same outputs as kd_nearest_gpu. */
int kd_nearest_dense_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                         int64_t nq, int32_t* out_index, float* out_dist2, float* out_dist, int32_t* out_visited,
                         void* workspace, size_t workspace_bytes, void* stream) {
    if (!tree_points || !perm || !queries || !workspace || n <= 0 || dim <= 0 || nq < 0 || n >= (1ll << 31))
        return kd::fail("kd_nearest_dense_gpu: bad arguments");
    if (workspace_bytes < kd_nearest_dense_workspace_bytes(n, nq)) return kd::fail("kd_nearest_dense_gpu: workspace too small");
    if (nq == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;
    kd::DenseWs W;
    W.dall = (float*)ws;
    ws += (size_t)kd::kDenseQB * (size_t)n * 4;
    W.gmin = (unsigned*)ws;
    W.gmaxda = W.gmin + kd::kDenseQB;
    W.ccnt = W.gmaxda + kd::kDenseQB;
    W.cand = W.ccnt + kd::kDenseQB;
    W.need = (int*)(W.cand + kd::kDenseQB * kd::kDenseMaxCand);
    const unsigned N = (unsigned)n;
    for (int64_t q0 = 0; q0 < nq; q0 += kd::kDenseQB) {
        const int nqb = (int)std::min<int64_t>(kd::kDenseQB, nq - q0);
        const float* qb = queries + (size_t)q0 * dim;
        kd::k_nn_init<<<1, 32, 0, st>>>(W);
        kd::k_nn_dense<<<(N + kd::kDenseRows - 1) / kd::kDenseRows, kd::kDenseRows, 0, st>>>(tree_points, N, dim, qb, nqb, W);
        kd::k_nn_collect<<<(N + 255) / 256, 256, 0, st>>>(N, nqb, W);
        kd::k_nn_resolve<<<1, 32, 0, st>>>(tree_points, perm, N, dim, qb, nqb, W, W.need + q0,
                                           out_index ? out_index + q0 : nullptr, out_dist2 ? out_dist2 + q0 : nullptr,
                                           out_dist ? out_dist + q0 : nullptr, out_visited ? out_visited + q0 : nullptr);
    }
    kd::k_nearest<0><<<(unsigned)((nq + 127) / 128), 128, 0, st>>>(tree_points, perm, N, dim, queries, nq, out_index,
                                                                 out_dist2, out_dist, out_visited, W.need, nullptr);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_nearest_dense_gpu: ") + cudaGetErrorString(e));
}

}  // extern "C"

namespace kd {
}  // namespace kd

extern "C" {

// dim 1..8 run the register-resident specialisations (vector loads need rows aligned to their width: the tree
// arrays of this library are; foreign pointers fall back to the generic kernel)
#define KD_DIM_DISPATCH(dim, aligned, CALL)                 \
    switch ((aligned) ? (dim) : 0) {                        \
        case 1: { constexpr int D = 1; CALL; } break;       \
        case 2: { constexpr int D = 2; CALL; } break;       \
        case 3: { constexpr int D = 3; CALL; } break;       \
        case 4: { constexpr int D = 4; CALL; } break;       \
        case 5: { constexpr int D = 5; CALL; } break;       \
        case 6: { constexpr int D = 6; CALL; } break;       \
        case 7: { constexpr int D = 7; CALL; } break;       \
        case 8: { constexpr int D = 8; CALL; } break;       \
        default: { constexpr int D = 0; CALL; } break;      \
    }

#define KD_RET_LAUNCH(what)                                                                       \
    do {                                                                                          \
        cudaError_t e_ = cudaGetLastError();                                                      \
        return e_ == cudaSuccess ? 0 : kd::fail(std::string(what) + ": " + cudaGetErrorString(e_)); \
    } while (0)

int kd_nearest_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries, int64_t nq,
                   int32_t* out_index, float* out_dist2, float* out_dist, int32_t* out_visited, void* stream) {
    if (!tree_points || !perm || !queries || n <= 0 || dim <= 0 || nq < 0 || n >= (1ll << 31))
        return kd::fail("kd_nearest_gpu: bad arguments");
    if (nq == 0) return 0;
    return kd_nearest_ordered_gpu(tree_points, perm, n, dim, queries, nq, nullptr, out_index, out_dist2, out_dist,
                                  out_visited, stream);
}

/* Ordering of a query batch for coherent traversal: order[t] = index of the t-th query when the queries are grouped
 * by the sub-tree (of depth `levels` <= 16) their search descends into first.  key: nq uint32 of scratch; bins:
 * (1 << levels) uint32 of scratch.  Results of the *_ordered entry points do not depend on the order. */
int kd_query_order_gpu(const float* tree_points, int64_t n, int dim, const float* queries, int64_t nq, int levels,
                       int32_t* order, void* key_scratch, void* bins_scratch, void* stream) {
    if (!tree_points || !queries || !order || !key_scratch || !bins_scratch || n <= 0 || dim <= 0 || nq < 0 || levels < 1 ||
        levels > 16 || n >= (1ll << 31) || nq >= (1ll << 31))
        return kd::fail("kd_query_order_gpu: bad arguments (1 <= levels <= 16)");
    if (nq == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned bins = 1u << levels;
    if (cudaMemsetAsync(bins_scratch, 0, (size_t)bins * 4, st) != cudaSuccess) return kd::fail("kd_query_order_gpu: memset failed");
    kd::k_qkey<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(tree_points, (unsigned)n, dim, queries, nq, levels,
                                                           (unsigned*)key_scratch, (unsigned*)bins_scratch);
    kd::k_qscan<<<1, 1024, 0, st>>>((unsigned*)bins_scratch, bins);
    kd::k_qscatter<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>((const unsigned*)key_scratch, (unsigned*)bins_scratch, nq, order);
    KD_RET_LAUNCH("kd_query_order_gpu");
}

int kd_nearest_ordered_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                           int64_t nq, const int32_t* order, int32_t* out_index, float* out_dist2, float* out_dist,
                           int32_t* out_visited, void* stream) {
    if (!tree_points || !perm || !queries || n <= 0 || dim <= 0 || nq < 0 || n >= (1ll << 31))
        return kd::fail("kd_nearest_ordered_gpu: bad arguments");
    if (nq == 0) return 0;
    const bool aligned = ((uintptr_t)tree_points & 15u) == 0;
    KD_DIM_DISPATCH(dim, aligned, (kd::k_nearest<D><<<(unsigned)((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        tree_points, perm, (unsigned)n, dim, queries, nq, out_index, out_dist2, out_dist, out_visited, nullptr, order)));
    KD_RET_LAUNCH("kd_nearest_ordered_gpu");
}

int kd_materialize_nodes_gpu(int64_t n, const int32_t* perm, int32_t* node_point, int32_t* left, int32_t* right,
                             void* stream) {
    if (!perm || !node_point || !left || !right || n <= 0 || n >= (1ll << 31))
        return kd::fail("kd_materialize_nodes_gpu: bad arguments");
    kd::k_materialize<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>((unsigned)n, perm, node_point, left,
                                                                                  right);
    KD_RET_LAUNCH("kd_materialize_nodes_gpu");
}

int kd_bbox_gpu(const float* points, int64_t n, int dim, float* out_min, float* out_max, void* stream) {
    if (!points || !out_min || !out_max || n <= 0 || dim <= 0) return kd::fail("kd_bbox_gpu: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    unsigned* mn = reinterpret_cast<unsigned*>(out_min);
    unsigned* mx = reinterpret_cast<unsigned*>(out_max);
    size_t total = (size_t)n * dim;
    size_t threads = (size_t)148 * 8 * 256;
    size_t stride = (threads / dim) * dim;
    if (stride == 0) stride = dim;  // dim larger than the grid: one thread per axis, still correct
    size_t blocks = (stride + 255) / 256;
    kd::k_bbox_init<<<(dim + 255) / 256, 256, 0, s>>>(mn, mx, dim);
    kd::k_bbox<<<(unsigned)blocks, 256, 0, s>>>(points, total, dim, stride, mn, mx);
    kd::k_bbox_finish<<<(dim + 255) / 256, 256, 0, s>>>(mn, mx, dim);
    KD_RET_LAUNCH("kd_bbox_gpu");
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Batched k-NN (extension: the reference answers k = 1 only).  Definition: the reference's traversal
// (same visiting order, near side first) with a bounded list of the k best nodes seen so far instead of
// one; a node enters the list iff the list is not full or its distance is strictly smaller than the
// current k-th best; the far side is visited iff the list is not full or da*da < current k-th best.
// For k = 1 this is exactly kdtree_sequential.cpp:75-136.  The list is kept sorted (ascending, earlier
// visit first among equals); k <= kMaxK.  kd_knn_cpu (host_core.cpp) is the same definition on the host.
// ------------------------------------------------------------------------------------------------
namespace kd {
constexpr int kMaxK = 32;

template <int DIM>
__global__ void __launch_bounds__(128) k_knn(const float* __restrict__ tp, const int32_t* __restrict__ perm, unsigned N,
                                             int dim, const float* __restrict__ queries, long long nq, int k,
                                             int32_t* __restrict__ out_index, float* __restrict__ out_d2,
                                             const int32_t* __restrict__ order) {
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    const long long q = order ? (long long)order[t] : t;
    if (DIM > 0) dim = DIM;
    RowOps<DIM> R;
    R.load_query(queries + (size_t)q * dim, dim);
    unsigned st_lo[kMaxDepth], st_n[kMaxDepth];
    float st_da2[kMaxDepth];
    unsigned short st_axis[kMaxDepth];
    float bd[kMaxK];
    unsigned bs[kMaxK];
    int have = 0, sp = 0;
    unsigned lo = 0, cnt = N;
    int axis = 0;
    for (;;) {
        while (cnt > 0) {
            unsigned m = cnt >> 1, slot = lo + m;
            float de, da;
            R.visit(tp + (size_t)slot * dim, dim, axis, de, da);
            if (have < k || de < bd[have - 1]) {
                int p = have < k ? have : k - 1;        // insertion, stable for equal distances
                while (p > 0 && bd[p - 1] > de) { bd[p] = bd[p - 1]; bs[p] = bs[p - 1]; --p; }
                bd[p] = de;
                bs[p] = slot;
                if (have < k) ++have;
            }
            unsigned rlo = slot + 1, rn = cnt - m - 1;
            axis = axis + 1 == dim ? 0 : axis + 1;
            st_da2[sp] = __fmul_rn(da, da);
            st_axis[sp] = (unsigned short)axis;
            if (da < 0.f) { st_lo[sp] = rlo; st_n[sp] = rn; cnt = m; }
            else { st_lo[sp] = lo; st_n[sp] = m; lo = rlo; cnt = rn; }
            ++sp;
        }
        bool resumed = false;
        while (sp > 0) {
            --sp;
            if (st_n[sp] > 0 && (have < k || st_da2[sp] < bd[have - 1])) {
                lo = st_lo[sp]; cnt = st_n[sp]; axis = st_axis[sp];
                resumed = true;
                break;
            }
        }
        if (!resumed) break;
    }
    for (int i = 0; i < k; ++i) {
        out_index[(size_t)q * k + i] = i < have ? perm[bs[i]] : -1;
        out_d2[(size_t)q * k + i] = i < have ? bd[i] : __int_as_float(0x7f800000);
    }
}
}  // namespace kd

extern "C" int kd_knn_ordered_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                                  int64_t nq, int k, const int32_t* order, int32_t* out_index, float* out_dist2,
                                  void* stream) {
    if (!tree_points || !perm || !queries || !out_index || !out_dist2 || n <= 0 || dim <= 0 || nq < 0 || k < 1 ||
        k > kd::kMaxK || n >= (1ll << 31))
        return kd::fail("kd_knn_gpu: bad arguments (1 <= k <= 32)");
    if (nq == 0) return 0;
    const bool aligned = ((uintptr_t)tree_points & 15u) == 0;
    KD_DIM_DISPATCH(dim, aligned, (kd::k_knn<D><<<(unsigned)((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        tree_points, perm, (unsigned)n, dim, queries, nq, k, out_index, out_dist2, order)));
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_knn_gpu: ") + cudaGetErrorString(e));
}

extern "C" int kd_knn_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                          int64_t nq, int k, int32_t* out_index, float* out_dist2, void* stream) {
    return kd_knn_ordered_gpu(tree_points, perm, n, dim, queries, nq, k, nullptr, out_index, out_dist2, stream);
}
