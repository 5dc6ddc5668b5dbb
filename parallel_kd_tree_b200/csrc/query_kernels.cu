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

// one thread per query; explicit stack of postponed far sides
__global__ void __launch_bounds__(128) k_nearest(const float* __restrict__ tp, const int32_t* __restrict__ perm,
                                                 unsigned N, int dim, const float* __restrict__ queries, long long nq,
                                                 int32_t* __restrict__ out_index, float* __restrict__ out_d2,
                                                 float* __restrict__ out_d, int32_t* __restrict__ out_visited) {
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float* qp = queries + (size_t)q * dim;
    unsigned st_lo[kMaxDepth], st_n[kMaxDepth];
    float st_da2[kMaxDepth];
    unsigned char st_depth[kMaxDepth];
    int sp = 0;
    unsigned best = N >> 1;  // nearest_neighbor(): best starts at the root (kdtree_sequential.cpp:133-136)
    float best_d = dist2_exact(tp + (size_t)best * dim, qp, dim);
    unsigned lo = 0, cnt = N;
    int depth = 0, visited = 0;
    for (;;) {
        while (cnt > 0) {
            unsigned m = cnt >> 1, slot = lo + m;
            const float* pp = tp + (size_t)slot * dim;
            int axis = depth % dim;
            float de = dist2_exact(pp, qp, dim);
            float da = __fsub_rn(qp[axis], pp[axis]);
            ++visited;
            if (de < best_d) { best_d = de; best = slot; }
            unsigned rlo = slot + 1, rn = cnt - m - 1;
            float da2 = __fmul_rn(da, da);
            st_da2[sp] = da2;
            st_depth[sp] = (unsigned char)(depth + 1);
            if (da < 0.f) {
                st_lo[sp] = rlo; st_n[sp] = rn;
                cnt = m;
            } else {
                st_lo[sp] = lo; st_n[sp] = m;
                lo = rlo; cnt = rn;
            }
            ++sp;
            ++depth;
        }
        bool resumed = false;
        while (sp > 0) {
            --sp;
            if (st_n[sp] > 0 && st_da2[sp] < best_d) {
                lo = st_lo[sp]; cnt = st_n[sp]; depth = st_depth[sp];
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
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= stride) return;
    unsigned lo = 0xffffffffu, hi = 0u;
    for (size_t e = t; e < total; e += stride) {
        unsigned k = ordered_key(pts[e]);
        lo = min(lo, k);
        hi = max(hi, k);
    }
    int axis = (int)(t % (size_t)dim);
    if (lo <= hi) { atomicMin(&mn[axis], lo); atomicMax(&mx[axis], hi); }
}

__global__ void k_bbox_finish(unsigned* mn, unsigned* mx, int dim) {
    int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= dim) return;
    auto back = [](unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); };
    reinterpret_cast<float*>(mn)[d] = back(mn[d]);
    reinterpret_cast<float*>(mx)[d] = back(mx[d]);
}

}  // namespace kd

extern "C" {

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
    kd::k_nearest<<<(unsigned)((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        tree_points, perm, (unsigned)n, dim, queries, nq, out_index, out_dist2, out_dist, out_visited);
    KD_RET_LAUNCH("kd_nearest_gpu");
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

__global__ void __launch_bounds__(128) k_knn(const float* __restrict__ tp, const int32_t* __restrict__ perm, unsigned N,
                                             int dim, const float* __restrict__ queries, long long nq, int k,
                                             int32_t* __restrict__ out_index, float* __restrict__ out_d2) {
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float* qp = queries + (size_t)q * dim;
    unsigned st_lo[kMaxDepth], st_n[kMaxDepth];
    float st_da2[kMaxDepth];
    unsigned char st_depth[kMaxDepth];
    float bd[kMaxK];
    unsigned bs[kMaxK];
    int have = 0, sp = 0;
    unsigned lo = 0, cnt = N;
    int depth = 0;
    for (;;) {
        while (cnt > 0) {
            unsigned m = cnt >> 1, slot = lo + m;
            const float* pp = tp + (size_t)slot * dim;
            int axis = depth % dim;
            float de = dist2_exact(pp, qp, dim);
            float da = __fsub_rn(qp[axis], pp[axis]);
            if (have < k || de < bd[have - 1]) {
                int p = have < k ? have : k - 1;        // insertion, stable for equal distances
                while (p > 0 && bd[p - 1] > de) { bd[p] = bd[p - 1]; bs[p] = bs[p - 1]; --p; }
                bd[p] = de;
                bs[p] = slot;
                if (have < k) ++have;
            }
            unsigned rlo = slot + 1, rn = cnt - m - 1;
            st_da2[sp] = __fmul_rn(da, da);
            st_depth[sp] = (unsigned char)(depth + 1);
            if (da < 0.f) { st_lo[sp] = rlo; st_n[sp] = rn; cnt = m; }
            else { st_lo[sp] = lo; st_n[sp] = m; lo = rlo; cnt = rn; }
            ++sp;
            ++depth;
        }
        bool resumed = false;
        while (sp > 0) {
            --sp;
            if (st_n[sp] > 0 && (have < k || st_da2[sp] < bd[have - 1])) {
                lo = st_lo[sp]; cnt = st_n[sp]; depth = st_depth[sp];
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

extern "C" int kd_knn_gpu(const float* tree_points, const int32_t* perm, int64_t n, int dim, const float* queries,
                          int64_t nq, int k, int32_t* out_index, float* out_dist2, void* stream) {
    if (!tree_points || !perm || !queries || !out_index || !out_dist2 || n <= 0 || dim <= 0 || nq < 0 || k < 1 ||
        k > kd::kMaxK || n >= (1ll << 31))
        return kd::fail("kd_knn_gpu: bad arguments (1 <= k <= 32)");
    if (nq == 0) return 0;
    kd::k_knn<<<(unsigned)((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(tree_points, perm, (unsigned)n, dim, queries,
                                                                             nq, k, out_index, out_dist2);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_knn_gpu: ") + cudaGetErrorString(e));
}
