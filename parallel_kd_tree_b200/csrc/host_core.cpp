// Native host core: problem generation, CPU build / query, node materialisation,
// tree checker and the reference's text protocol.  No CUDA in this file.
//
// Parity map (reference file:line -> here)
//   Utility.cpp:6-18  generate_problem            -> kd_generate_uniform
//   kdtree_mpi.cpp:19-41 slice generator           -> kd_generate_uniform_rows (row window of the same stream)
//   kdtree_sequential.cpp:30-70 build_tree         -> kd_build_cpu (select + pin) / kd_build_cpu_sortref
//   kdtree_sequential.cpp:75-136 nearest           -> kd_nearest_cpu (explicit stack, same visiting order)
//   Node.hpp:35-45 Node                            -> kd_materialize_nodes_cpu
//   Node.cpp:16-28, Utility.cpp:21-63,122-124      -> kd_format_*
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kdtree_b200.h"
#include "kd_common.h"
#include "kd_error.h"

namespace {

struct AxisLess {
    const float* pts;
    int dim, axis;
    // total order (coord[axis] with -0 == +0, then index): Node.cpp:46-48 plus the canonical tie-break
    bool operator()(int32_t a, int32_t b) const {
        float fa = pts[(size_t)a * dim + axis], fb = pts[(size_t)b * dim + axis];
        if (fa < fb) return true;
        if (fb < fa) return false;
        return a < b;
    }
};

// One range of the recursion of kdtree_sequential.cpp:30-66, restated without the sort:
// the last slot is pinned, the node is the rank-n/2 element of the other n-1, and the
// largest element of the left part is moved to the left part's last slot (it is what a
// sort would have left there, and it becomes the next level's pinned element).
template <bool kSort>
void split_range(int32_t* idx, int64_t lo, int64_t n, const AxisLess& less) {
    int64_t m = n / 2;
    int32_t* first = idx + lo;
    if (kSort) {
        std::sort(first, first + (n - 1), less);
    } else {
        std::nth_element(first, first + m, first + (n - 1), less);
        if (m > 1) {
            int32_t* mx = std::max_element(first, first + m, less);
            std::swap(*mx, first[m - 1]);
        }
    }
}

template <bool kSort>
void build_subtree(const float* pts, int dim, int32_t* idx, int64_t lo, int64_t n, int depth) {
    struct Frame { int64_t lo, n; int depth; };
    std::vector<Frame> st;
    st.push_back({lo, n, depth});
    while (!st.empty()) {
        Frame f = st.back();
        st.pop_back();
        if (f.n < 2) continue;
        AxisLess less{pts, dim, f.depth % dim};
        split_range<kSort>(idx, f.lo, f.n, less);
        int64_t m = f.n / 2;
        st.push_back({f.lo + m + 1, f.n - m - 1, f.depth + 1});
        st.push_back({f.lo, m, f.depth + 1});
    }
}

std::string fmt_float(float v) {
    std::ostringstream os;  // default ostream formatting == what the reference's operator<< produces
    os << v;
    return os.str();
}

void append_point(std::ostringstream& os, const float* c, int dim, int id) {
    // Node.cpp:16-28 including the ", , ..., " artefact for dim > 5
    os << "Point(ID=" << id << ", dimension=" << dim << ", coordinates=[";
    for (int d = 0; d < dim - 1; ++d) {
        os << c[d] << ", ";
        if (d >= 4) { os << ", ..., "; break; }
    }
    os << c[dim - 1] << "])";
}

int dup_out(const std::string& s, char** out, size_t* len) {
    char* p = (char*)std::malloc(s.size() + 1);
    if (!p) return kd::fail("out of memory");
    std::memcpy(p, s.data(), s.size());
    p[s.size()] = 0;
    *out = p;
    if (len) *len = s.size();
    return 0;
}

}  // namespace

extern "C" {

const char* kd_last_error(void) { return kd::error_slot().c_str(); }
int kd_version(void) { return 100; }
void kd_free(void* p) { std::free(p); }

// The arithmetic of std::uniform_real_distribution<float>(-100, 100) on one 32-bit engine draw in libstdc++
// (generate_canonical<float, 24>: float(draw) / 2^32, clamped below 1; then u * (b - a) + a as two separately
// rounded float operations — SURVEY.md §3.5).  kd_generator_selftest() checks it against the library objects.
static inline float uniform_from_draw(uint32_t draw) {
    float u = (float)draw / 4294967296.0f;
    if (u >= 1.0f) u = std::nextafter(1.0f, 0.0f);
    return u * 200.0f + -100.0f;  // no contraction: this file is built with -ffp-contract=off
}

int kd_generate_uniform_rows(int seed, int dim, int64_t first_row, int64_t n_rows, float* out) {
    if (dim <= 0 || n_rows < 0 || first_row < 0 || (!out && n_rows > 0)) return kd::fail("kd_generate_uniform_rows: bad arguments");
    std::mt19937 rng(seed);
    // exactly one engine draw per float, so a row window is a plain skip-ahead
    unsigned long long skip = (unsigned long long)first_row * (unsigned long long)dim;
    rng.discard(skip);
    const size_t total = (size_t)n_rows * (size_t)dim;
    // The engine is inherently sequential and dominates (3.7 of ~5 ns per value): a producer/consumer split of
    // draw and conversion was measured and gained nothing, so this stays one loop.
    for (size_t i = 0; i < total; ++i) out[i] = uniform_from_draw((uint32_t)rng());
    return 0;
}

int64_t kd_generator_selftest(int seed, int64_t count) {
    // mismatches between uniform_from_draw and libstdc++'s own distribution object, on a random stream and on
    // the draws where rounding matters (ends of the range, the 2^24 boundary of exact int -> float conversion)
    struct Fixed {
        using result_type = uint32_t;
        uint32_t v;
        static constexpr result_type min() { return 0; }
        static constexpr result_type max() { return 0xffffffffu; }
        result_type operator()() { return v; }
    };
    int64_t bad = 0;
    std::uniform_real_distribution<float> dist(-100, 100);
    std::mt19937 a(seed), b(seed);
    for (int64_t i = 0; i < count; ++i) {
        float x = dist(a), y = uniform_from_draw((uint32_t)b());
        if (std::memcmp(&x, &y, 4) != 0) ++bad;
    }
    const uint32_t edges[] = {0u, 1u, 2u, 0xffffffffu, 0xfffffffeu, 0xffffff80u, 0xffffff7fu, 0xffffff00u, 0x80000000u,
                              0x7fffffffu, 0x01000000u, 0x01000001u, 0x00ffffffu, 0x80000080u, 0x80000081u};
    for (uint32_t d : edges) {
        Fixed f{d};
        float x = dist(f), y = uniform_from_draw(d);
        if (std::memcmp(&x, &y, 4) != 0) ++bad;
    }
    return bad;
}

int kd_generate_uniform(int seed, int dim, int64_t n_rows, float* out) {
    return kd_generate_uniform_rows(seed, dim, 0, n_rows, out);
}

int kd_generate_clustered(int seed, int dim, int64_t n_rows, int n_clusters, float sigma, float* out) {
    if (dim <= 0 || n_rows < 0 || n_clusters <= 0 || !out) return kd::fail("kd_generate_clustered: bad arguments");
    // Definition (ours): centres ~ U(-80,80)^dim, relative weights ~ U(0.2,1), then for every row a
    // weighted cluster pick and centre + sigma * N(0,1) per coordinate, clamped to [-100,100].
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> cdist(-80.f, 80.f), wdist(0.2f, 1.f);
    std::vector<float> centres((size_t)n_clusters * dim);
    for (auto& c : centres) c = cdist(rng);
    std::vector<double> cum(n_clusters);
    double acc = 0;
    for (int k = 0; k < n_clusters; ++k) { acc += wdist(rng); cum[k] = acc; }
    std::uniform_real_distribution<double> u01(0.0, 1.0);
    std::normal_distribution<float> nd(0.f, 1.f);
    for (int64_t i = 0; i < n_rows; ++i) {
        double r = u01(rng) * acc;
        int k = (int)(std::lower_bound(cum.begin(), cum.end(), r) - cum.begin());
        if (k >= n_clusters) k = n_clusters - 1;
        for (int d = 0; d < dim; ++d) {
            float v = centres[(size_t)k * dim + d] + sigma * nd(rng);
            out[(size_t)i * dim + d] = std::min(100.f, std::max(-100.f, v));
        }
    }
    return 0;
}

static int build_cpu_impl(const float* points, int64_t n, int dim, int32_t* perm, int threads, bool sortref) {
    if (!points || !perm || n <= 0 || dim <= 0) return kd::fail("kd_build_cpu: bad arguments");
    if (n > 0x7fffffffLL) return kd::fail("kd_build_cpu: n must fit in int32");
    for (int64_t i = 0; i < n; ++i) perm[i] = (int32_t)i;
    if (sortref) {
        build_subtree<true>(points, dim, perm, 0, n, 0);
        return 0;
    }
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    if (threads == 1 || n < (1 << 16)) {
        build_subtree<false>(points, dim, perm, 0, n, 0);
        return 0;
    }
    // top levels level by level (ranges of one level in parallel), then one task per sub-tree
    struct Range { int64_t lo, n; int depth; };
    std::vector<Range> level{{0, n, 0}};
    while ((int)level.size() < threads * 4 && level[0].n > (1 << 14)) {
        std::vector<Range> next;
        std::atomic<size_t> cur{0};
        auto work = [&]() {
            for (size_t i; (i = cur.fetch_add(1)) < level.size();) {
                const Range& r = level[i];
                if (r.n < 2) continue;
                split_range<false>(perm, r.lo, r.n, AxisLess{points, dim, r.depth % dim});
            }
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < std::min<int>(threads, (int)level.size()); ++t) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
        for (const Range& r : level) {
            if (r.n < 2) continue;
            int64_t m = r.n / 2;
            next.push_back({r.lo, m, r.depth + 1});
            next.push_back({r.lo + m + 1, r.n - m - 1, r.depth + 1});
        }
        level.swap(next);
        if (level.empty()) return 0;
    }
    std::atomic<size_t> cur{0};
    auto work = [&]() {
        for (size_t i; (i = cur.fetch_add(1)) < level.size();)
            build_subtree<false>(points, dim, perm, level[i].lo, level[i].n, level[i].depth);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    return 0;
}

int kd_build_cpu(const float* points, int64_t n, int dim, int32_t* perm, int threads) {
    return build_cpu_impl(points, n, dim, perm, threads, false);
}
int kd_build_cpu_sortref(const float* points, int64_t n, int dim, int32_t* perm) {
    return build_cpu_impl(points, n, dim, perm, 1, true);
}

int kd_nearest_cpu(const float* points, int64_t n, int dim, const int32_t* perm, const float* queries, int64_t nq,
                   int32_t* out_index, float* out_dist2, float* out_dist, int64_t* out_visited) {
    if (!points || !perm || !queries || n <= 0 || dim <= 0 || nq < 0) return kd::fail("kd_nearest_cpu: bad arguments");
    struct Far { int64_t lo, n; int depth; float da2; };
    auto d2 = [&](const float* a, const float* b) {
        // kdtree_sequential.cpp:14-25: strictly sequential float sum, separate multiply and add
        float dist = 0;
        for (int i = 0; i < dim; ++i) {
            float tmp = a[i] - b[i];
            dist += tmp * tmp;  // built with -ffp-contract=off: no FMA, like the reference's -mavx build
        }
        return dist;
    };
    // queries are independent: one contiguous block of them per hardware thread
    const int workers = (int)std::max<int64_t>(1, std::min<int64_t>(std::thread::hardware_concurrency(), nq / 4));
    auto run = [&](int64_t q0, int64_t q1) {
    std::vector<Far> st;
    st.reserve(64);
    for (int64_t q = q0; q < q1; ++q) {
        const float* qp = queries + (size_t)q * dim;
        int64_t root = n / 2;
        int32_t best = perm[root];
        float best_d = d2(points + (size_t)best * dim, qp);  // nearest_neighbor(): best starts at the root
        int64_t visited = 0;
        int64_t lo = 0, cnt = n;
        int depth = 0;
        st.clear();
        for (;;) {
            while (cnt > 0) {
                int64_t m = cnt / 2;
                int32_t p = perm[lo + m];
                const float* pp = points + (size_t)p * dim;
                int axis = depth % dim;
                float de = d2(pp, qp);
                float da = qp[axis] - pp[axis];
                ++visited;
                if (de < best_d) { best_d = de; best = p; }
                int64_t llo = lo, ln = m, rlo = lo + m + 1, rn = cnt - m - 1;
                float da2 = da * da;
                if (da < 0) {
                    st.push_back({rlo, rn, depth + 1, da2});
                    lo = llo; cnt = ln;
                } else {
                    st.push_back({llo, ln, depth + 1, da2});
                    lo = rlo; cnt = rn;
                }
                ++depth;
            }
            bool resumed = false;
            while (!st.empty()) {
                Far f = st.back();
                st.pop_back();
                if (f.n > 0 && f.da2 < best_d) {
                    lo = f.lo; cnt = f.n; depth = f.depth;
                    resumed = true;
                    break;
                }
            }
            if (!resumed) break;
        }
        if (out_index) out_index[q] = best;
        if (out_dist2) out_dist2[q] = best_d;
        if (out_dist) out_dist[q] = std::sqrt(best_d);
        if (out_visited) out_visited[q] = visited;
    }
    };
    std::vector<std::thread> pool;
    for (int w = 1; w < workers; ++w) pool.emplace_back(run, nq * w / workers, nq * (w + 1) / workers);
    run(0, nq / workers);
    for (auto& t : pool) t.join();
    return 0;
}

int kd_knn_cpu(const float* points, int64_t n, int dim, const int32_t* perm, const float* queries, int64_t nq, int k,
               int32_t* out_index, float* out_dist2) {
    // k-NN extension: the reference traversal with a bounded sorted list (definition in query_kernels.cu)
    if (!points || !perm || !queries || !out_index || !out_dist2 || n <= 0 || dim <= 0 || nq < 0 || k < 1 || k > 32)
        return kd::fail("kd_knn_cpu: bad arguments (1 <= k <= 32)");
    struct Far { int64_t lo, n; int depth; float da2; };
    std::vector<Far> st;
    std::vector<float> bd(k);
    std::vector<int32_t> bp(k);
    for (int64_t q = 0; q < nq; ++q) {
        const float* qp = queries + (size_t)q * dim;
        int have = 0;
        int64_t lo = 0, cnt = n;
        int depth = 0;
        st.clear();
        for (;;) {
            while (cnt > 0) {
                int64_t m = cnt / 2;
                int32_t p = perm[lo + m];
                const float* pp = points + (size_t)p * dim;
                int axis = depth % dim;
                float de = 0;
                for (int i = 0; i < dim; ++i) { float t = pp[i] - qp[i]; de += t * t; }
                float da = qp[axis] - pp[axis];
                if (have < k || de < bd[have - 1]) {
                    int at = have < k ? have : k - 1;
                    while (at > 0 && bd[at - 1] > de) { bd[at] = bd[at - 1]; bp[at] = bp[at - 1]; --at; }
                    bd[at] = de;
                    bp[at] = p;
                    if (have < k) ++have;
                }
                float da2 = da * da;
                if (da < 0) { st.push_back({lo + m + 1, cnt - m - 1, depth + 1, da2}); cnt = m; }
                else { st.push_back({lo, m, depth + 1, da2}); lo = lo + m + 1; cnt = cnt - m - 1; }
                ++depth;
            }
            bool resumed = false;
            while (!st.empty()) {
                Far f = st.back();
                st.pop_back();
                if (f.n > 0 && (have < k || f.da2 < bd[have - 1])) { lo = f.lo; cnt = f.n; depth = f.depth; resumed = true; break; }
            }
            if (!resumed) break;
        }
        for (int i = 0; i < k; ++i) {
            out_index[(size_t)q * k + i] = i < have ? bp[i] : -1;
            out_dist2[(size_t)q * k + i] = i < have ? bd[i] : INFINITY;
        }
    }
    return 0;
}

int64_t kd_root_slot(int64_t n) { return n > 0 ? n / 2 : -1; }

int kd_materialize_nodes_cpu(int64_t n, const int32_t* perm, int32_t* node_point, int32_t* left, int32_t* right) {
    if (n <= 0 || !perm || !node_point || !left || !right) return kd::fail("kd_materialize_nodes_cpu: bad arguments");
    struct R { int64_t lo, n; };
    std::vector<R> st{{0, n}};
    while (!st.empty()) {
        R r = st.back();
        st.pop_back();
        int64_t m = r.n / 2, slot = r.lo + m, rn = r.n - m - 1;
        node_point[slot] = perm[slot];
        left[slot] = m > 0 ? (int32_t)(r.lo + m / 2) : -1;
        right[slot] = rn > 0 ? (int32_t)(slot + 1 + rn / 2) : -1;
        if (m > 0) st.push_back({r.lo, m});
        if (rn > 0) st.push_back({slot + 1, rn});
    }
    return 0;
}

int kd_check_tree_cpu(const float* points, int64_t n, int dim, const int32_t* perm, int64_t* bad) {
    if (!points || !perm || !bad || n <= 0 || dim <= 0) return kd::fail("kd_check_tree_cpu: bad arguments");
    int64_t v = 0;
    std::vector<uint8_t> seen((size_t)n, 0);
    for (int64_t i = 0; i < n; ++i) {
        if (perm[i] < 0 || perm[i] >= n || seen[perm[i]]) ++v; else seen[perm[i]] = 1;
    }
    if (v) { *bad = v; return 0; }
    struct R { int64_t lo, n; int depth; };
    std::vector<R> st{{0, n, 0}};
    while (!st.empty()) {
        R r = st.back();
        st.pop_back();
        if (r.n < 3) continue;  // n == 2: the node is the pinned element itself, nothing is ordered
        AxisLess less{points, dim, r.depth % dim};
        int64_t m = r.n / 2;
        int32_t node = perm[r.lo + m];
        // the pinned element of a range ends in the range's last slot (right-most leaf)
        for (int64_t i = r.lo; i < r.lo + m; ++i)
            if (!less(perm[i], node)) ++v;
        for (int64_t i = r.lo + m + 1; i < r.lo + r.n - 1; ++i)
            if (!less(node, perm[i])) ++v;
        // the left part's pinned element (its last slot) is its maximum on this axis
        for (int64_t i = r.lo; i + 1 < r.lo + m; ++i)
            if (!less(perm[i], perm[r.lo + m - 1])) ++v;
        st.push_back({r.lo, m, r.depth + 1});
        st.push_back({r.lo + m + 1, r.n - m - 1, r.depth + 1});
    }
    *bad = v;
    return 0;
}

int kd_check_tree_ties_cpu(const float* points, int64_t n, int dim, const int32_t* perm, int64_t* bad) {
    // Tie-tolerant form of kd_check_tree_cpu: equal coordinates may fall on either side (what an unstable
    // std::sort is free to do).  A permutation with 0 violations is a possible output of
    // the reference's build_tree on this input.
    if (!points || !perm || !bad || n <= 0 || dim <= 0) return kd::fail("kd_check_tree_ties_cpu: bad arguments");
    int64_t v = 0;
    std::vector<uint8_t> seen((size_t)n, 0);
    for (int64_t i = 0; i < n; ++i) {
        if (perm[i] < 0 || perm[i] >= n || seen[perm[i]]) ++v; else seen[perm[i]] = 1;
    }
    if (v) { *bad = v; return 0; }
    struct R { int64_t lo, n; int depth; };
    std::vector<R> st{{0, n, 0}};
    while (!st.empty()) {
        R r = st.back();
        st.pop_back();
        if (r.n < 3) continue;
        const int axis = r.depth % dim;
        auto key = [&](int64_t slot) { return points[(size_t)perm[slot] * dim + axis]; };
        int64_t m = r.n / 2;
        float node = key(r.lo + m);
        for (int64_t i = r.lo; i < r.lo + m; ++i)
            if (key(i) > node) ++v;
        for (int64_t i = r.lo + m + 1; i < r.lo + r.n - 1; ++i)
            if (key(i) < node) ++v;
        float lmax = key(r.lo + m - 1);
        for (int64_t i = r.lo; i + 1 < r.lo + m; ++i)
            if (key(i) > lmax) ++v;
        st.push_back({r.lo, m, r.depth + 1});
        st.push_back({r.lo + m + 1, r.n - m - 1, r.depth + 1});
    }
    *bad = v;
    return 0;
}

int kd_format_float(float v, char* buf, int cap) {
    std::string s = fmt_float(v);
    if ((int)s.size() + 1 > cap) return kd::fail("kd_format_float: buffer too small");
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 0;
}

int kd_format_result_line(int id, float distance, char* buf, int cap) {
    std::ostringstream os;  // Utility.cpp:122-124
    os << "ID: " << id << " \t DISTANCE: " << distance;
    std::string s = os.str();
    if ((int)s.size() + 1 > cap) return kd::fail("kd_format_result_line: buffer too small");
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 0;
}

int kd_format_point(const float* coords, int dim, int id, char** out, size_t* len) {
    if (!coords || dim <= 0 || !out) return kd::fail("kd_format_point: bad arguments");
    std::ostringstream os;
    append_point(os, coords, dim, id);
    return dup_out(os.str(), out, len);
}

int kd_format_tree(const float* points, int64_t n, int dim, const int32_t* perm, int id_offset, char** out,
                   size_t* len) {
    if (!points || !perm || !out || n <= 0 || dim <= 0) return kd::fail("kd_format_tree: bad arguments");
    std::ostringstream os;  // Utility.cpp:49-63: preorder, `depth` tabs, NODE(@depth=d): <point>
    struct R { int64_t lo, n; int depth; };
    std::vector<R> st{{0, n, 0}};
    while (!st.empty()) {
        R r = st.back();
        st.pop_back();
        if (r.n <= 0) continue;
        int64_t m = r.n / 2;
        int32_t p = perm[r.lo + m];
        for (int d = 0; d < r.depth; ++d) os << "\t";
        os << "NODE(@depth=" << r.depth << "): ";
        append_point(os, points + (size_t)p * dim, dim, p + 1 + id_offset);
        os << "\n";
        st.push_back({r.lo + m + 1, r.n - m - 1, r.depth + 1});
        st.push_back({r.lo, m, r.depth + 1});
    }
    return dup_out(os.str(), out, len);
}

int kd_format_head_and_leaves(const float* points, int64_t n, int dim, const int32_t* perm, int id_offset,
                              char** out, size_t* len) {
    if (!points || !perm || !out || n <= 0 || dim <= 0) return kd::fail("kd_format_head_and_leaves: bad arguments");
    // Utility.cpp:21-36.  In in-order layout the left-most leaf is slot 0 and the right-most slot n-1.
    int64_t lo = 0, cnt = n;
    while (cnt / 2 > 0) cnt = cnt / 2;             // follow ->left
    int32_t leftleaf = perm[lo + cnt / 2];
    lo = 0; cnt = n;
    while (cnt - cnt / 2 - 1 > 0) { lo = lo + cnt / 2 + 1; cnt = cnt - cnt / 2 - 1; }  // follow ->right
    int32_t rightleaf = perm[lo + cnt / 2];
    int32_t head = perm[n / 2];
    std::ostringstream os;
    os << "\n" << "Head: ";
    append_point(os, points + (size_t)head * dim, dim, head + 1 + id_offset);
    os << "\n" << "Left Leaf: ";
    append_point(os, points + (size_t)leftleaf * dim, dim, leftleaf + 1 + id_offset);
    os << "\n" << "Right Leaf: ";
    append_point(os, points + (size_t)rightleaf * dim, dim, rightleaf + 1 + id_offset);
    os << "\n\n";
    return dup_out(os.str(), out, len);
}

}  // extern "C"
