// Single-GPU kd-tree build for sm_100a.
//
// What is built: exactly the tree of the reference's build_tree_rec
// (kdtree_sequential.cpp:30-66) under the total order (coord[axis], index):
// for a range of n points whose LAST slot is "pinned" (never sorted,
// kdtree_sequential.cpp:46-48), the node is the rank-n/2 element of the other
// n-1, the n/2 smaller ones form the left range with their maximum in the left
// range's last slot (so it is pinned one level down), the larger ones plus the
// old pinned element form the right range.  How it is built is new: a
// level-synchronous select + partition, no sort, no recursion.
//
//   K1  k_init / k_gather_keys          SoA working columns (+ per-level key gather for dim > kMaxCarry)
//   K2  k_sample + k_bracket            large ranges: a sorted sample brackets the rank, one pass compacts the
//       k_hist + k_scan                 ~10 % candidates, MSB radix select on the composite (key, index) over them
//   K3  k_partition + k_fixup           large ranges: unstable three-way partition, node emission, left-max -> last slot
//   K2+K3  k_medium                     ranges of 513..16384 points: one CTA selects and partitions in one kernel
//   K4  k_subtree                       windows of <= kFinishT slots: whole sub-trees inside shared memory
//                                       (TMA-staged, per-level radix select + partition in registers / shuffles)
//
// Ranges are never stored: range s of level l is found by walking the shape
// rule (left n/2, right n-n/2-1) along the bits of s.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>

#include "../../include/kdtree_b200.h"
#include "build_plan.h"
#include "kd_common.h"
#include "kd_error.h"

namespace kd {

// ------------------------------------------------------------------------------------------
// K1: working columns
// ------------------------------------------------------------------------------------------
// AoS rows -> SoA columns through a shared-memory tile: the rows are read as one flat coalesced run and
// every column is written as a coalesced run (a thread-per-row version touches each 128-byte line `dim` times)
constexpr int kInitRows = 1024, kInitBlock = 256;
__global__ void __launch_bounds__(kInitBlock) k_init(const float* __restrict__ pts, unsigned N, int dim, int ncol,
                                                     Cols dst) {
    __shared__ float tile[kInitRows * kMaxCarry];
    const unsigned row0 = blockIdx.x * kInitRows, tid = threadIdx.x;
    const unsigned rows = min((unsigned)kInitRows, N - row0);
    for (unsigned r = tid; r < rows; r += kInitBlock) dst.idx[row0 + r] = (int32_t)(row0 + r);
    if (ncol == 0) return;
    const unsigned total = rows * (unsigned)dim;
    const float* __restrict__ src = pts + (size_t)row0 * dim;
#pragma unroll 4
    for (unsigned e = tid; e < total; e += kInitBlock) tile[e] = src[e];
    __syncthreads();
    for (int c = 0; c < ncol; ++c)
#pragma unroll 4
        for (unsigned r = tid; r < rows; r += kInitBlock) dst.c[c][row0 + r] = tile[r * dim + c];
}

__global__ void k_state_init(SegState* st, unsigned N, int level, unsigned nseg) {
    unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    SegState S;
    S.prefix = 0ull;
    S.rank = n >> 1;  // node = rank n/2 among the n-1 unpinned elements
    S.cnt = n - 1;
    S.curL = 0;
    S.curR = 0;
    S.shift = 64;
    S.done = 0;
    S.lowP = 0ull;
    S.highP = ~0ull;
    S.cL = 0;
    S.cand_cnt = 0;
    S.use_cand = 0;
    S.pad_ = 0;
    st[s] = S;
}

__global__ void k_gather_keys(const float* __restrict__ pts, int dim, int axis, const int32_t* __restrict__ idx,
                              float* __restrict__ keybuf, unsigned N, int level, int cps) {
    unsigned s = blockIdx.x / cps, j = blockIdx.x % cps;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    unsigned chunk = (n + cps - 1) / cps;
    unsigned b = lo + j * chunk, e = min(b + chunk, lo + n);
    for (unsigned i = b + threadIdx.x; i < e; i += blockDim.x)
        keybuf[i] = pts[(size_t)idx[i] * dim + axis];
}

// ------------------------------------------------------------------------------------------
// K2: radix select.  Pass d histograms digit d of the composite over the candidates that
// still match the prefix chosen by passes 0..d-1.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void warp_hist_add(unsigned* sh, unsigned digit, bool valid) {
    // one shared atomic per distinct digit in the warp: constant cost under any key distribution
    unsigned peers = __match_any_sync(0xffffffffu, valid ? digit : 0xffffffffu);
    if (valid && (__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&sh[digit], __popc(peers));
}

// ---- bracket select for large ranges -------------------------------------------------------------
// A sorted sample of the range brackets the wanted rank between two pivots (k_sample); ONE pass over the
// keys counts what lies below the bracket and copies what lies inside it (about a tenth of the range) to the
// candidate arrays (k_bracket); the digit passes then run on the candidates only.  The result is exact: the
// counts prove whether the wanted rank is inside the bracket, and if it is not (or for adversarial orderings
// of the input) the range simply falls back to digit passes over all of its keys.
// Sample size per range: 1024, or 4096 for ranges of at least 2^20 points, where halving the bracket (14 % -> 7 % of
// the range) saves more in the candidate passes than the larger sort costs.  The half-width of the bracket is 4.5
// sigma of a random sample's rank (sigma = sqrt(S) / 2): 72 of 1024, 144 of 4096.
template <int SN>
__global__ void __launch_bounds__(512) k_sample(const float* __restrict__ keys, const int32_t* __restrict__ idx,
                                                SegState* st, unsigned N, int level) {
    constexpr int kDelta = SN == 1024 ? 72 : 144;
    __shared__ unsigned long long sm[SN];
    unsigned lo, n;
    seg_range(N, level, blockIdx.x, lo, n);
    const unsigned free_n = n - 1, m = n >> 1, tid = threadIdx.x;
    for (unsigned k = tid; k < (unsigned)SN; k += 512) {
        const unsigned pos = lo + (unsigned)(((unsigned long long)k * free_n) / SN);  // systematic sample
        sm[k] = composite(ordered_key(keys[pos]), (unsigned)idx[pos]);
    }
    __syncthreads();
    for (unsigned k = 2; k <= (unsigned)SN; k <<= 1) {
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            for (unsigned t = tid; t < (unsigned)SN / 2; t += 512) {
                const unsigned i = ((t & ~(j - 1u)) << 1) | (t & (j - 1u)), l = i | j;  // t-th pair at distance j
                const bool asc = (i & k) == 0u;
                const unsigned long long a = sm[i], b = sm[l];
                if ((a > b) == asc) { sm[i] = b; sm[l] = a; }
            }
            __syncthreads();
        }
    }
    if (tid == 0) {
        const int r = (int)(((unsigned long long)m * SN) / free_n);
        const int li = r - kDelta, hi = r + kDelta;
        st[blockIdx.x].lowP = li < 0 ? 0ull : sm[li];
        st[blockIdx.x].highP = hi >= SN ? ~0ull : sm[hi];
    }
}

// One streaming pass, no block-wide synchronisation: every CTA compacts the candidates of its chunk to the
// chunk's own first slots of the candidate arrays (they can never overflow it) and publishes its count; the
// digit passes then visit (chunk start, count) per CTA.  The pass is issue-bound, so the per-element work is
// two float compares (float order == key order, -0 == +0 included); candidate slots are handed out once per
// tile with a warp scan and one shared-memory atomic per warp.
template <int BLOCK, int IPT>
__global__ void __launch_bounds__(BLOCK) k_bracket(const float* __restrict__ keys, const int32_t* __restrict__ idx,
                                                   SegState* st, float* __restrict__ cand_key,
                                                   int32_t* __restrict__ cand_idx, unsigned* __restrict__ cand_cta,
                                                   unsigned N, int level, int cps) {
    __shared__ unsigned s_cur, s_less;
    const unsigned s = blockIdx.x / cps, j = blockIdx.x % cps;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned long long lowP = st[s].lowP, highP = st[s].highP;
    const unsigned lowK = (unsigned)(lowP >> 32), highK = (unsigned)(highP >> 32);
    auto key_float = [](unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); };
    // an open end of the bracket (only for ranks near the ends of a range) becomes an infinite pivot
    const float lowF = lowP == 0ull ? __int_as_float(0xff800000) : key_float(lowK);
    const float highF = highP == ~0ull ? __int_as_float(0x7f800000) : key_float(highK);
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    const unsigned free_n = n - 1;
    const unsigned chunk = (free_n + cps - 1) / cps;
    const unsigned b = min(lo + j * chunk, lo + free_n), e = min(b + chunk, lo + free_n);
    unsigned less = 0;
    if (tid == 0) { s_cur = 0; s_less = 0; }
    __syncthreads();
    for (unsigned tb = b; tb < e; tb += BLOCK * IPT) {
        float f[IPT];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const unsigned i = tb + k * BLOCK + tid;
            f[k] = i < e ? keys[i] : __int_as_float(0x7fc00000);  // NaN: neither below nor inside the bracket
        }
        unsigned cmask = 0;
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const float v = f[k];
            if (v < lowF) ++less;
            else if (v <= highF) {
                bool cand = true;
                if (v == lowF || v == highF) {  // only on the bracket's edge keys does the index decide
                    const unsigned i = tb + k * BLOCK + tid;
                    const unsigned long long c = composite(ordered_key(v), (unsigned)idx[i]);
                    if (c < lowP) { ++less; cand = false; }
                    else cand = c <= highP;
                }
                cmask |= cand ? (1u << k) : 0u;
            }
        }
        // candidate slots: exclusive warp scan of the per-thread counts, one shared atomic per warp
        const unsigned mine = __popc(cmask);
        unsigned inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        const unsigned total = __shfl_sync(0xffffffffu, inc, 31);
        unsigned base = 0;
        if (lane == 31 && total) base = atomicAdd(&s_cur, total);
        base = __shfl_sync(0xffffffffu, base, 31);
        unsigned p = b + base + inc - mine;
#pragma unroll
        for (int k = 0; k < IPT; ++k)
            if (cmask & (1u << k)) {
                if (KD_INB(p >= b && p < e, 903)) {   // a chunk's candidates can never outgrow the chunk
                    cand_key[p] = f[k];
                    cand_idx[p] = (int32_t)(tb + k * BLOCK + tid);  // the candidate's slot; its index is fetched on demand
                }
                ++p;
            }
    }
    for (int o = 16; o > 0; o >>= 1) less += __shfl_xor_sync(0xffffffffu, less, o);
    if (lane == 0 && less) atomicAdd(&s_less, less);
    __syncthreads();
    if (tid == 0) {
        cand_cta[blockIdx.x] = s_cur;
        if (s_cur) atomicAdd(&st[s].cand_cnt, s_cur);
        if (s_less) atomicAdd(&st[s].cL, s_less);
    }
}

__global__ void k_bracket_finish(SegState* st, unsigned N, int level, unsigned nseg) {
    const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    const unsigned m = n >> 1, cL = st[s].cL, cc = st[s].cand_cnt;
    if (m >= cL && m - cL < cc) {  // the wanted rank is inside the bracket: select rank m - cL among the candidates
        st[s].use_cand = 1;
        st[s].rank = m - cL;
        st[s].cnt = cc;
        // every candidate lies in [lowP, highP]: the bits the two pivots share are resolved already, so the first
        // digit pass starts right below them instead of re-counting eleven bits that cannot differ
        const unsigned long long lowP = st[s].lowP, highP = st[s].highP;
        if (lowP != 0ull && highP != ~0ull && lowP != highP) {
            const int hi = 64 - __clzll((long long)(lowP ^ highP));
            st[s].shift = hi;
            st[s].prefix = hi >= 64 ? 0ull : ((lowP >> hi) << hi);
        }
    }
}

template <int BLOCK, int UNROLL>
__global__ void __launch_bounds__(BLOCK) k_hist(const float* __restrict__ keys_all, const int32_t* __restrict__ idx_all,
                                                float* cand_key, int32_t* cand_idx, unsigned* cand_cta, SegState* st,
                                                unsigned* __restrict__ ghist, unsigned N, int level, int cps, int d) {
    __shared__ unsigned sh[kDigitBins];
    __shared__ unsigned s_w;   // survivors re-compacted by this CTA (see below)
    unsigned s = blockIdx.x / cps, j = blockIdx.x % cps;
    SegState S = st[s];
    if (S.done == 1) return;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    // the pinned last slot never takes part (kdtree_sequential.cpp:46-48); after a successful bracket the
    // passes run on the range's candidates instead of on all of its keys
    const float* keys = S.use_cand ? cand_key : keys_all;
    const bool via = S.use_cand != 0;  // candidates carry their slot; the point index is read through it
    auto point_index = [&](unsigned i) -> unsigned { return (unsigned)(via ? idx_all[cand_idx[i]] : idx_all[i]); };
    const unsigned free_n = n - 1;
    const unsigned chunk = (free_n + cps - 1) / cps;
    const unsigned b = min(lo + j * chunk, lo + free_n);
    // (k_bracket left this CTA's candidates at the first slots of the same chunk)
    const unsigned e = S.use_cand ? b + cand_cta[blockIdx.x] : min(b + chunk, lo + free_n);
    if (b >= e) return;
    if (S.done == 2) {
        // exactly one candidate matches the prefix, but elements outside the bracket may share those bits:
        // read the candidate back so that the partition compares complete composites
        const int psh = S.shift;
        const int ksh = psh >= 32 ? psh - 32 : 0;
        const unsigned wantk = (unsigned)(S.prefix >> 32) >> ksh;
        for (unsigned i = b + threadIdx.x; i < e; i += BLOCK) {
            const unsigned key = ordered_key(keys[i]);
            if ((key >> ksh) != wantk) continue;  // the index is fetched only for keys that can match
            const unsigned long long c = composite(key, point_index(i));
            if ((c >> psh) == (S.prefix >> psh)) {
                st[s].prefix = c;
                st[s].shift = 0;
                __threadfence();
                st[s].done = 1;
            }
        }
        return;
    }
    for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) sh[t] = 0;
    if (threadIdx.x == 0) s_w = 0;
    __syncthreads();
    // bits [hi, 64) of the composite are resolved; this pass counts the digit right below them (at most 11 bits,
    // never straddling the key | index boundary)
    const int hi = S.shift, dsh = select_digit_shift(hi);
    const unsigned mask = (1u << (hi - dsh)) - 1u;
    const unsigned iters = (e - b + BLOCK * UNROLL - 1) / (BLOCK * UNROLL);
    (void)d;
    // Once a digit has been chosen only a sliver of the candidates still matches the prefix (1 / 2048 after the
    // first digit), yet every later pass would stream all of them again.  So a pass that sees few matches re-compacts
    // them, in place, to the front of this CTA's own run (the writes of a tile land below the next tile's reads) and
    // shortens the run: the following passes and the read-back touch the survivors only.
    const bool compact = via && S.cnt * 8u < S.cand_cnt;
    if (hi > 32) {
        const int ks = dsh - 32, ps = hi - 32;
        const unsigned want = hi >= 64 ? 0u : (unsigned)(S.prefix >> hi);
        for (unsigned it = 0; it < iters; ++it) {
            unsigned i0 = b + it * (BLOCK * UNROLL) + threadIdx.x;
            float f[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) f[u] = (i0 + u * BLOCK) < e ? keys[i0 + u * BLOCK] : 0.f;
            unsigned vm = 0u;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                bool valid = (i0 + u * BLOCK) < e;
                unsigned key = ordered_key(f[u]);
                if (hi < 64) valid = valid && ((key >> ps) == want);
                warp_hist_add(sh, (key >> ks) & mask, valid);
                vm |= valid ? (1u << u) : 0u;
            }
            if (compact) {
                int32_t slot[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) slot[u] = ((vm >> u) & 1u) ? cand_idx[i0 + u * BLOCK] : 0;
                __syncthreads();   // every read of this tile is done before its range may be overwritten
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if ((vm >> u) & 1u) {
                        const unsigned p = b + atomicAdd(&s_w, 1u);
                        cand_key[p] = f[u];
                        cand_idx[p] = slot[u];
                    }
            }
        }
    } else {
        const unsigned wantkey = (unsigned)(S.prefix >> 32);
        const unsigned wantidx = hi >= 32 ? 0u : ((unsigned)(S.prefix & 0xffffffffull) >> hi);
        for (unsigned it = 0; it < iters; ++it) {
            unsigned i0 = b + it * (BLOCK * UNROLL) + threadIdx.x;
            float f[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) f[u] = (i0 + u * BLOCK) < e ? keys[i0 + u * BLOCK] : 0.f;
            unsigned vm = 0u;
            int32_t slot[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                unsigned i = i0 + u * BLOCK;
                bool valid = i < e && ordered_key(f[u]) == wantkey;
                slot[u] = (valid && via) ? cand_idx[i] : (int32_t)i;
                unsigned id = valid ? (unsigned)idx_all[slot[u]] : 0u;
                if (hi < 32) valid = valid && ((id >> hi) == wantidx);
                warp_hist_add(sh, (id >> dsh) & mask, valid);
                vm |= valid ? (1u << u) : 0u;
            }
            if (compact) {
                __syncthreads();
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if ((vm >> u) & 1u) {
                        const unsigned p = b + atomicAdd(&s_w, 1u);
                        cand_key[p] = f[u];
                        cand_idx[p] = slot[u];
                    }
            }
        }
    }
    __syncthreads();
    if (compact && threadIdx.x == 0) cand_cta[blockIdx.x] = s_w;
    unsigned* gh = ghist + (size_t)s * kDigitBins;
    for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) {
        unsigned v = sh[t];
        if (v) atomicAdd(&gh[t], v);
    }
}

// one CTA of 256 threads per range: locate the bin holding the wanted rank, extend the prefix, clear bins
__global__ void __launch_bounds__(256) k_scan(SegState* st, unsigned* ghist, int d, int* err) {
    __shared__ unsigned wsum[8];
    __shared__ int fbin;
    __shared__ unsigned fcum, fcnt;
    unsigned s = blockIdx.x;
    SegState S = st[s];
    if (S.done) return;  // finished, or waiting for the candidate read-back of the next k_hist launch
    unsigned* h = ghist + (size_t)s * kDigitBins;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) fbin = -1;
    unsigned v[8], sum = 0;
    uint4 a = reinterpret_cast<uint4*>(h)[tid * 2], bq = reinterpret_cast<uint4*>(h)[tid * 2 + 1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = bq.x; v[5] = bq.y; v[6] = bq.z; v[7] = bq.w;
#pragma unroll
    for (int k = 0; k < 8; ++k) sum += v[k];
    unsigned inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    unsigned base = 0;
    for (int k = 0; k < w; ++k) base += wsum[k];
    unsigned c = base + inc - sum;
    const unsigned r = S.rank;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (r >= c && r < c + v[k]) { fbin = tid * 8 + k; fcum = c; fcnt = v[k]; }
        c += v[k];
    }
    reinterpret_cast<uint4*>(h)[tid * 2] = make_uint4(0, 0, 0, 0);
    reinterpret_cast<uint4*>(h)[tid * 2 + 1] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    if (tid == 0) {
        if (fbin < 0) {
            atomicExch(err, 1 + d);  // the histogram does not contain the wanted rank: internal error
            S.done = 1;
        } else {
            const int dsh = select_digit_shift(S.shift);
            S.prefix |= (unsigned long long)(unsigned)fbin << dsh;
            S.rank = r - fcum;
            S.cnt = fcnt;
            S.shift = dsh;
            S.done = dsh == 0 ? 1 : (fcnt == 1u ? (S.use_cand ? 2 : 1) : 0);
        }
        st[s] = S;
    }
}

// ------------------------------------------------------------------------------------------
// K3: partition around the selected element.  Order inside the left / right parts is free
// (the next level selects under a total order), so output slots are claimed with atomics.
// ------------------------------------------------------------------------------------------

template <int NCOL, int BLOCK, int IPT>
__global__ void __launch_bounds__(BLOCK) k_partition(Cols src, Cols dst, const float* __restrict__ keys, int axis,
                                                     SegState* st, MaxRec* rec, OutPtrs out, unsigned out_lo, int dim,
                                                     unsigned N, int level, int cps) {
    __shared__ unsigned s_cnt[2], s_base[2];
    __shared__ unsigned long long s_wcomp[BLOCK / 32];
    __shared__ unsigned s_wpos[BLOCK / 32];
    const unsigned s = blockIdx.x / cps, j = blockIdx.x % cps;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned long long prefix = st[s].prefix;
    const int shift = st[s].shift;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    const unsigned m = n >> 1, free_n = n - 1;
    const unsigned chunk = (free_n + cps - 1) / cps;
    const unsigned b = lo + j * chunk, e = min(b + chunk, lo + free_n);
    const unsigned long long want = prefix >> shift;
    unsigned long long best = 0ull;
    unsigned bestpos = 0, bestvalid = 0;
    bool has_node = false;
    int node_id = 0;
    float node_val[NCOL > 0 ? NCOL : 1];

    if (j == (unsigned)cps - 1 && tid == 0) {  // the pinned element keeps its slot: last of the right range
        unsigned p = lo + n - 1;
#pragma unroll
        for (int c = 0; c < NCOL; ++c) dst.c[c][p] = src.c[c][p];
        dst.idx[p] = src.idx[p];
    }
    for (unsigned tb = b; tb < e; tb += BLOCK * IPT) {
        if (tid < 2) s_cnt[tid] = 0;
        __syncthreads();
        // every column of the tile goes to registers first: the loads overlap the ranking below
        float val[IPT][NCOL > 0 ? NCOL : 1];
        float kf[IPT];
        int id[IPT];
        unsigned rk[IPT];
        signed char side[IPT];  // -1 left, 0 node, 1 right, 2 nothing
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            unsigned i = tb + k * BLOCK + tid;
            bool valid = i < e;
            id[k] = valid ? src.idx[i] : 0;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) val[k][c] = valid ? src.c[c][i] : 0.f;
            if (NCOL == 0) kf[k] = valid ? keys[i] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            unsigned i = tb + k * BLOCK + tid;
            if (NCOL > 0) {
                kf[k] = val[k][0];
#pragma unroll
                for (int c = 1; c < NCOL; ++c) kf[k] = (c == axis) ? val[k][c] : kf[k];
            }
            side[k] = 2;
            if (i < e) {
                unsigned long long c = composite(ordered_key(kf[k]), (unsigned)id[k]) >> shift;
                side[k] = c < want ? -1 : (c > want ? 1 : 0);
            }
        }
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            unsigned bl = __ballot_sync(0xffffffffu, side[k] == -1);
            unsigned br = __ballot_sync(0xffffffffu, side[k] == 1);
            unsigned wl = 0, wr = 0;
            if (lane == 0) {
                if (bl) wl = atomicAdd(&s_cnt[0], __popc(bl));
                if (br) wr = atomicAdd(&s_cnt[1], __popc(br));
            }
            wl = __shfl_sync(0xffffffffu, wl, 0);
            wr = __shfl_sync(0xffffffffu, wr, 0);
            unsigned lt = (1u << lane) - 1u;
            rk[k] = side[k] == -1 ? wl + __popc(bl & lt) : wr + __popc(br & lt);
        }
        __syncthreads();
        if (tid < 2) {
            unsigned c = s_cnt[tid];
            s_base[tid] = c ? atomicAdd(tid == 0 ? &st[s].curL : &st[s].curR, c) : 0u;
        }
        __syncthreads();
        const unsigned baseL = lo + s_base[0], baseR = lo + m + 1 + s_base[1];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            if (side[k] == 2) continue;
            if (side[k] == 0) {  // the node of this range: remembered, stored once after the sweep
                node_id = id[k];
                has_node = true;
#pragma unroll
                for (int c = 0; c < NCOL; ++c) node_val[c] = val[k][c];
                continue;
            }
            unsigned p = (side[k] == -1 ? baseL : baseR) + rk[k];
            // left elements fill [lo, lo+m), right ones [lo+m+1, lo+n-1) (the last slot belongs to the pinned element)
            if (!KD_INB(side[k] == -1 ? (p >= lo && p < lo + m) : (p > lo + m && p < lo + n - 1u), 901)) continue;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) dst.c[c][p] = val[k][c];
            dst.idx[p] = id[k];
            if (side[k] == -1) {
                unsigned long long c = composite(ordered_key(kf[k]), (unsigned)id[k]);
                if (!bestvalid || c > best) { best = c; bestpos = p; bestvalid = 1; }
            }
        }
    }
    if (has_node) {
        // the node's final slot is lo + n/2 (kdtree_sequential.cpp:51,65) in every replica of the tree arrays
        // (the all-gather of a distributed build is part of this store)
        for (int r = 0; r < out.n; ++r) {
            out.perm[r][out_lo + lo + m] = node_id;
            if (out.tp[r]) {
#pragma unroll
                for (int c = 0; c < NCOL; ++c) out.tp[r][(size_t)(out_lo + lo + m) * dim + c] = node_val[c];
            }
        }
    }
    // block-wide arg-max of the left-going composites -> one record per CTA
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long oc = __shfl_xor_sync(0xffffffffu, best, o);
        unsigned op = __shfl_xor_sync(0xffffffffu, bestpos, o);
        unsigned ov = __shfl_xor_sync(0xffffffffu, bestvalid, o);
        if (ov && (!bestvalid || oc > best)) { best = oc; bestpos = op; bestvalid = 1; }
    }
    if (lane == 0) { s_wcomp[tid >> 5] = bestvalid ? best : 0ull; s_wpos[tid >> 5] = bestvalid ? bestpos : 0xffffffffu; }
    __syncthreads();
    if (tid == 0) {
        MaxRec r{0ull, 0u, 0u};
        for (int w = 0; w < BLOCK / 32; ++w)
            if (s_wpos[w] != 0xffffffffu && (!r.valid || s_wcomp[w] > r.comp)) { r.comp = s_wcomp[w]; r.pos = s_wpos[w]; r.valid = 1; }
        rec[blockIdx.x] = r;
    }
}

// one warp per range: the largest left element goes to the left range's last slot (pinned next level)
__global__ void __launch_bounds__(256) k_fixup(Cols dst, int ncol, const MaxRec* __restrict__ rec, unsigned N,
                                               int level, unsigned nseg, int cps) {
    unsigned s = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (s >= nseg) return;
    const int lane = threadIdx.x & 31;
    unsigned long long best = 0ull;
    unsigned bestpos = 0, bestvalid = 0;
    for (int t = lane; t < cps; t += 32) {
        MaxRec r = rec[(size_t)s * cps + t];
        if (r.valid && (!bestvalid || r.comp > best)) { best = r.comp; bestpos = r.pos; bestvalid = 1; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long oc = __shfl_xor_sync(0xffffffffu, best, o);
        unsigned op = __shfl_xor_sync(0xffffffffu, bestpos, o);
        unsigned ov = __shfl_xor_sync(0xffffffffu, bestvalid, o);
        if (ov && (!bestvalid || oc > best)) { best = oc; bestpos = op; bestvalid = 1; }
    }
    if (!bestvalid) return;
    unsigned lo, n;
    seg_range(N, level, s, lo, n);
    unsigned last = lo + (n >> 1) - 1;
    if (bestpos == last) return;
    if (!KD_INB(bestpos >= lo && bestpos < last, 904)) return;
    if (lane < ncol) {
        float a = dst.c[lane][bestpos], b = dst.c[lane][last];
        dst.c[lane][bestpos] = b;
        dst.c[lane][last] = a;
    } else if (lane == ncol) {
        int32_t a = dst.idx[bestpos], b = dst.idx[last];
        dst.idx[bestpos] = b;
        dst.idx[last] = a;
    }
}

// ------------------------------------------------------------------------------------------
// K2+K3 for medium ranges (kFinishT < n <= kMediumMax): ONE CTA owns the range for the whole
// level.  The split-axis keys are staged in shared memory once, every digit pass of the select
// runs on them (shared-memory histogram, no global atomics, no global histogram), then a single
// sweep moves the points: the node to its final slot, the left range's maximum straight to the left
// range's last slot (found before the sweep, so no fix-up kernel), everything else to slots claimed
// with warp-aggregated shared-memory cursors.  HBM traffic per level: keys once + points once in,
// points once out.
// ------------------------------------------------------------------------------------------
template <int NCOL, int DBITS>
__global__ void __launch_bounds__(512, (NCOL <= 4 ? 3 : 1)) k_medium(Cols src, Cols dst, const float* __restrict__ keys, int axis,
                                                OutPtrs out, unsigned out_lo, int dim, unsigned N, int level,
                                                unsigned cap, int* err) {
    // Digits of DBITS bits (11 for the larger ranges, 8 below 4096 points: a 256-bin histogram is cheaper to
    // clear and scan than the few extra passes over keys that already sit in shared memory).  The digits do
    // not start at bit 63: the range's smallest and largest key are found while the keys are staged, the bits
    // they share are the initial prefix, and the first digit sits right below them, so the first pass already
    // splits the range into up to 2^DBITS bins.
    constexpr int kBins = 1 << DBITS;
    extern __shared__ __align__(16) unsigned char msm[];
    unsigned* skey = reinterpret_cast<unsigned*>(msm);   // [cap] ordered keys of the unpinned elements
    unsigned* hist = skey + cap;                         // [kBins]
    __shared__ unsigned s_w[16];
    __shared__ unsigned long long s_prefix, s_bcomp[16];
    __shared__ unsigned s_rank, s_cnt, s_done, s_cur[2], s_cum, s_min, s_max, s_bpos[16];
    __shared__ int s_shift, s_bin;
    const unsigned tid = threadIdx.x, B = blockDim.x, lane = tid & 31, wid = tid >> 5;
    unsigned lo, n;
    seg_range(N, level, blockIdx.x, lo, n);
    const unsigned m = n >> 1, free_n = n - 1;
    const float* __restrict__ kcol = NCOL > 0 ? src.c[axis] : keys;
    if (tid == 0) { s_rank = m; s_done = 0; s_shift = 64; s_cur[0] = s_cur[1] = 0; s_min = 0xffffffffu; s_max = 0u; }
    __syncthreads();
    unsigned kmin = 0xffffffffu, kmax = 0u;
    for (unsigned i0 = tid; i0 < free_n; i0 += 4 * B) {  // four loads in flight per thread
        float f[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) f[u] = (i0 + u * B) < free_n ? kcol[lo + i0 + u * B] : 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if ((i0 + u * B) < free_n) {
                const unsigned key = ordered_key(f[u]);
                skey[i0 + u * B] = key;
                kmin = min(kmin, key);
                kmax = max(kmax, key);
            }
    }
    for (int o = 16; o > 0; o >>= 1) {
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    }
    if (lane == 0) { atomicMin(&s_min, kmin); atomicMax(&s_max, kmax); }
    __syncthreads();
    const unsigned vary = s_min ^ s_max;
    const int top = vary ? 32 - __clz(vary) : 0;  // the low `top` key bits differ inside the range
    if (tid == 0) s_prefix = top >= 32 ? 0ull : ((unsigned long long)((s_min >> top) << top) << 32);
    __syncthreads();
    const unsigned iters = (free_n + B - 1) / B;
    // ---- select: rank m of the free_n composites (key, index); bits [hi, 64) are resolved ----
    int hi = 32 + top;
    for (int pass = 0; pass < 8; ++pass) {
        for (unsigned t = tid; t < (unsigned)kBins; t += B) hist[t] = 0;
        __syncthreads();
        if (s_done) break;
        const unsigned long long prefix = s_prefix;
        const int dsh = hi > 32 ? max(hi - DBITS, 32) : max(hi - DBITS, 0);  // digits never straddle key | index
        const unsigned mask = (1u << (hi - dsh)) - 1u;
        for (unsigned it = 0; it < iters; ++it) {
            const unsigned i = it * B + tid;
            bool valid = i < free_n;
            const unsigned key = valid ? skey[i] : 0u;
            unsigned digit;
            if (dsh >= 32) {   // a digit of the key
                if (hi < 64) valid = valid && ((key >> (hi - 32)) == (unsigned)(prefix >> hi));
                digit = (key >> (dsh - 32)) & mask;
            } else {           // a digit of the point index, among the elements whose key equals the node's
                valid = valid && key == (unsigned)(prefix >> 32);
                const unsigned id = valid ? (unsigned)src.idx[lo + i] : 0u;
                if (hi < 32) valid = valid && ((id >> hi) == ((unsigned)(prefix & 0xffffffffull) >> hi));
                digit = (id >> dsh) & mask;
            }
            warp_hist_add(hist, digit, valid);
        }
        __syncthreads();
        // locate the bin of the wanted rank: each of the first kBins / per threads owns `per` consecutive bins
        const unsigned per = (unsigned)kBins >= B ? (unsigned)kBins / B : 1u;
        const bool owner = tid * per < (unsigned)kBins;
        unsigned sum = 0;
        if (owner)
            for (unsigned k = 0; k < per; ++k) sum += hist[tid * per + k];
        unsigned inc = sum;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (unsigned)o) inc += t;
        }
        if (lane == 31) s_w[wid] = inc;
        if (tid == 0) s_bin = -1;
        __syncthreads();
        unsigned base = 0;
        for (unsigned k = 0; k < wid; ++k) base += s_w[k];
        unsigned c = base + inc - sum;
        const unsigned r = s_rank;
        if (owner)
            for (unsigned k = 0; k < per; ++k) {
                const unsigned v = hist[tid * per + k];
                if (r >= c && r < c + v) { s_bin = (int)(tid * per + k); s_cnt = v; s_cum = c; }
                c += v;
            }
        __syncthreads();
        if (tid == 0) {
            if (s_bin < 0) { atomicExch(err, 11 + pass); s_done = 1; }
            else {
                s_prefix = prefix | ((unsigned long long)(unsigned)s_bin << dsh);
                s_rank = r - s_cum;
                s_shift = dsh;
                if (s_cnt == 1u || dsh == 0) s_done = 1;
            }
        }
        hi = dsh;
        __syncthreads();
    }
    __syncthreads();
    const int shift = s_shift;
    const unsigned long long want = s_prefix >> shift;
    // ---- move the points: loads of four elements per thread are issued before any of them is ranked ----
    if (tid == 0) {  // the pinned element keeps its slot: last of the right range
        const unsigned p = lo + n - 1;
#pragma unroll
        for (int c = 0; c < NCOL; ++c) dst.c[c][p] = src.c[c][p];
        dst.idx[p] = src.idx[p];
    }
    unsigned long long best = 0ull;   // largest left-going composite seen by this thread, and where it went
    unsigned bestpos = 0xffffffffu;
    constexpr int U = 4;
    for (unsigned it = 0; it < iters; it += U) {
        float val[U][NCOL > 0 ? NCOL : 1];
        int id[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned i = (it + u) * B + tid;
            const bool valid = i < free_n;
            id[u] = valid ? src.idx[lo + i] : 0;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) val[u][c] = valid ? src.c[c][lo + i] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned i = (it + u) * B + tid;
            int side = 2;
            unsigned long long comp = 0ull;
            if (i < free_n) {
                comp = composite(skey[i], (unsigned)id[u]);
                const unsigned long long c = comp >> shift;
                side = c < want ? -1 : (c > want ? 1 : 0);
            }
            const unsigned bl = __ballot_sync(0xffffffffu, side == -1);
            const unsigned br = __ballot_sync(0xffffffffu, side == 1);
            unsigned wl = 0, wr = 0;
            if (lane == 0) {
                if (bl) wl = atomicAdd(&s_cur[0], __popc(bl));
                if (br) wr = atomicAdd(&s_cur[1], __popc(br));
            }
            wl = __shfl_sync(0xffffffffu, wl, 0);
            wr = __shfl_sync(0xffffffffu, wr, 0);
            const unsigned lt = (1u << lane) - 1u;
            if (side == 2) continue;
            if (side == 0) {  // the node: final slot lo + n/2 in every replica of the tree arrays
                for (int r = 0; r < out.n; ++r) {
                    out.perm[r][out_lo + lo + m] = id[u];
                    if (out.tp[r]) {
#pragma unroll
                        for (int c = 0; c < NCOL; ++c) out.tp[r][(size_t)(out_lo + lo + m) * dim + c] = val[u][c];
                    }
                }
                continue;
            }
            const unsigned p = side == -1 ? lo + wl + __popc(bl & lt) : lo + m + 1 + wr + __popc(br & lt);
            if (!KD_INB(side == -1 ? p < lo + m : (p > lo + m && p < lo + n - 1u), 902)) continue;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) dst.c[c][p] = val[u][c];
            dst.idx[p] = id[u];
            if (side == -1 && (bestpos == 0xffffffffu || comp > best)) { best = comp; bestpos = p; }
        }
    }
    // ---- the left range's maximum goes to the left range's last slot (it is pinned one level down) ----
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long oc = __shfl_xor_sync(0xffffffffu, best, o);
        const unsigned op = __shfl_xor_sync(0xffffffffu, bestpos, o);
        if (op != 0xffffffffu && (bestpos == 0xffffffffu || oc > best)) { best = oc; bestpos = op; }
    }
    if (lane == 0) { s_bcomp[wid] = best; s_bpos[wid] = bestpos; }
    __threadfence();
    __syncthreads();   // every store of the sweep is visible to the thread that swaps
    if (tid == 0) {
        best = 0ull;
        bestpos = 0xffffffffu;
        for (unsigned w = 0; w < (B >> 5); ++w)
            if (s_bpos[w] != 0xffffffffu && (bestpos == 0xffffffffu || s_bcomp[w] > best)) { best = s_bcomp[w]; bestpos = s_bpos[w]; }
        const unsigned last = lo + m - 1;
        if (bestpos != 0xffffffffu && bestpos != last) {
#pragma unroll
            for (int c = 0; c < NCOL; ++c) {
                const float a = __ldcg(&dst.c[c][bestpos]), b2 = __ldcg(&dst.c[c][last]);
                dst.c[c][bestpos] = b2;
                dst.c[c][last] = a;
            }
            const int32_t a = __ldcg(&dst.idx[bestpos]), b2 = __ldcg(&dst.idx[last]);
            dst.idx[bestpos] = b2;
            dst.idx[last] = a;
        }
    }
}

// TMA (bulk async copy engine) helpers: 1-D global -> shared copies completing on an mbarrier.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// bounded wait: a copy that never completes must not hang the device (the caller falls back to plain loads)
__device__ __forceinline__ bool mbar_wait(unsigned long long* bar, unsigned parity) {
    for (unsigned i = 0; i < (1u << 20); ++i)
        if (mbar_try_wait(bar, parity)) return true;
    return false;
}

// ------------------------------------------------------------------------------------------
// K4: shared-memory sub-tree builder.  One CTA owns a window of n <= kFinishT points and builds
// its whole sub-tree (up to 12 levels) with ONE trip through HBM: the window is staged with the TMA
// engine, every level is a select + partition on (key, index) composites that never leaves the SM,
// and the finished window is written out once, coalesced, to every replica of the tree arrays.
//
// What a level has to produce (kdtree_sequential.cpp:46-56, restated without the sort): the element
// of rank n/2 among the sub-range's first n-1 positions (the node), the smaller ones on its left with
// their maximum in the left part's last position (it is pinned one level down), the larger ones on its
// right, the old pinned element untouched.  Points never move: `sord` maps position -> staged slot and
// is permuted in place (every team reads its positions into registers before it writes them).
//
// Virtual layout: sub-range q of sub-level j owns the P-aligned block q of P0 = pow2 >= n virtual
// positions (P = P0 >> j >= its size), 8 consecutive virtual positions per thread, so the threads of
// sub-range q are an aligned group of P/8 threads, a "team":
//   P >= 512       team = 2..16 warps  -> counts meet in shared memory, __syncthreads between phases
//   16 <= P <= 256 team = 2..32 lanes  -> every warp runs on its own (no CTA barrier until the write-out);
//                                         counts are warp reductions over the team's lanes (REDUX)
//   P <= 8         a thread owns whole sub-ranges -> all-pairs ranks in registers (that IS the sort)
// Select of a team: MSB radix select with 4-bit digits aligned to the range of the keys that are still
// alive (key - min, so the first digit already splits >= 8 ways; index digits only among equal keys).
// No shared-memory atomics per element: each thread counts its 8 elements into sixteen 4-bit fields of
// a 64-bit register, the fields are widened to 16 bits and summed with 8 warp reductions.  As soon as
// at most `cap` elements are alive they are compacted into a team list (ballot ranks) and resolved by
// all-pairs counting, which yields the node's composite X and, if the node is not the smallest
// candidate, the left maximum M; otherwise M is a max-reduction over the elements below X.
// ------------------------------------------------------------------------------------------
constexpr int kSubBlock = 512;
constexpr int kSubWarps = kSubBlock / 32;
static_assert(kSubBlock * 8 == kFinishT, "8 virtual positions per thread");
constexpr unsigned kSubListPerWarp = 128;                 // candidate slots (u64) per warp
constexpr unsigned kSubWarpScratch = kSubListPerWarp + 32; // + (X, M) for up to 16 teams of the warp
constexpr unsigned kSubCap = 64;                           // candidates resolved by all-pairs counting
constexpr int kSubTS = kFinishT + 8;                       // column stride (the staged window starts 16-byte aligned)

struct SubTeam { unsigned kmin, kmax, lcnt, pad; };       // shared state of a multi-warp team

struct SubCtx {
    float* scol;                 // [ncol][kSubTS] coordinates by slot
    int32_t* sidx;               // [kSubTS]       point index by slot
    unsigned long long* slist;   // [kSubWarps][kSubWarpScratch]
    SubTeam* steam;              // [8]
    unsigned* sbins;             // [3][8 teams][8 words]
    unsigned* swtot;             // [kSubWarps]
    unsigned* stab;              // [2][kFinishT / 2]  sub-range table (lo << 16 | n), by level parity
    unsigned short* sord;        // [kFinishT]     slot by position
    const float* pts;            // gather mode (ncol == 0): keys come from the point array
    int ncol, dim;
    int* err;
};

__device__ __forceinline__ void sub_load(const SubCtx& C, int axis, unsigned slo, unsigned sn, unsigned i0,
                                         unsigned long long (&comp)[8], unsigned (&slot)[8], unsigned& valid) {
    const float* __restrict__ kcol = C.ncol ? C.scol + (size_t)axis * kSubTS : nullptr;
    valid = 0u;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const unsigned i = i0 + (unsigned)e;
        comp[e] = ~0ull;
        slot[e] = 0u;
        if (sn >= 3u && i + 1u < sn) {   // position i of the sub-range's first n-1 (the last one is pinned)
            const unsigned s = C.sord[slo + i];
            const unsigned id = (unsigned)C.sidx[s];
            const float f = kcol ? kcol[s] : C.pts[(size_t)id * C.dim + axis];
            comp[e] = composite(ordered_key(f), id);
            slot[e] = s;
            valid |= 1u << e;
        }
    }
}

// Team-wide reductions.  Multi-warp teams (kCta) reduce per warp here and meet in shared memory afterwards;
// sub-warp teams are aligned groups of tpt lanes: xor butterflies (a REDUX with a per-lane member mask would
// be emulated with a loop), whole-warp teams use one REDUX.
template <bool kCta>
__device__ __forceinline__ unsigned team_add(unsigned v, unsigned tpt) {
    if (kCta || tpt == 32u) return __reduce_add_sync(0xffffffffu, v);
    for (unsigned o = tpt >> 1; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <bool kCta>
__device__ __forceinline__ unsigned team_min(unsigned v, unsigned tpt) {
    if (kCta || tpt == 32u) return __reduce_min_sync(0xffffffffu, v);
    for (unsigned o = tpt >> 1; o; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <bool kCta>
__device__ __forceinline__ unsigned team_max(unsigned v, unsigned tpt) {
    if (kCta || tpt == 32u) return __reduce_max_sync(0xffffffffu, v);
    for (unsigned o = tpt >> 1; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Team levels.  kCta: the team spans whole warps (P >= 512); otherwise it is an aligned group of lanes.
template <bool kCta>
__device__ __forceinline__ void sub_team_level(const SubCtx& C, int j, int logP, int axis) {
    constexpr unsigned FULL = 0xffffffffu;
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const int logT = logP - 3;                                     // log2(threads per team)
    const unsigned tpt = 1u << logT, q = tid >> logT, tl = tid & (tpt - 1u);
    const unsigned* tab = C.stab + (j & 1) * (kFinishT / 2);
    unsigned* tabn = C.stab + ((j + 1) & 1) * (kFinishT / 2);
    const bool in_range = q < (1u << j);
    // table slot of sub-range q = its first virtual position / 4: the slots of a warp's sub-ranges are then private
    // to the warp at every level (warps run unsynchronised below P = 512)
    const unsigned ent = in_range ? tab[q << (logP - 2)] : 0u;
    const unsigned slo = ent >> 16, sn = ent & 0xffffu, m = sn >> 1;
    bool active = sn >= 3u;
    const unsigned tmask = (kCta || tpt == 32u) ? FULL : (((1u << tpt) - 1u) << (lane & ~(tpt - 1u)));
    const unsigned team_w0 = kCta ? ((q << logT) >> 5) : warp;
    const unsigned tiw = kCta ? 0u : (lane >> logT);               // team index inside the warp
    const unsigned tpw = kCta ? 1u : (32u >> logT);                // teams per warp
    const unsigned lcap = kCta ? kSubCap : min(kSubCap, 4u * tpt);
    unsigned long long* L = C.slist + (size_t)team_w0 * kSubWarpScratch + tiw * (kSubListPerWarp / tpw);
    unsigned long long* XM = C.slist + (size_t)team_w0 * kSubWarpScratch + kSubListPerWarp + 2u * tiw;
    auto tsync = [] { if (kCta) __syncthreads(); else __syncwarp(); };

    unsigned long long comp[8];
    unsigned slot[8], valid;
    sub_load(C, axis, slo, sn, tl * 8u, comp, slot, valid);

    // ---- range of the keys: the digits are aligned to it ----
    unsigned kmn = 0xffffffffu, kmx = 0u;
#pragma unroll
    for (int e = 0; e < 8; ++e)
        if ((valid >> e) & 1u) { const unsigned k = (unsigned)(comp[e] >> 32); kmn = min(kmn, k); kmx = max(kmx, k); }
    kmn = team_min<kCta>(kmn, tpt);
    kmx = team_max<kCta>(kmx, tpt);
    if (kCta) {
        SubTeam& T = C.steam[q];
        if (lane == 0 && kmn <= kmx) { atomicMin(&T.kmin, kmn); atomicMax(&T.kmax, kmx); }
        __syncthreads();
        kmn = T.kmin;
        kmx = T.kmax;
    }
    int mode = 0, sh = 0;      // mode 0: digits of (key - base); mode 1: digits of the point index (equal keys)
    unsigned base = kmn, r = m, acnt = active ? sn - 1u : 0u, alive = valid;
    if (active) {
        if (kmn == kmx) { mode = 1; sh = 28; }
        else { const int t = 31 - __clz(kmx - kmn); sh = t > 3 ? t - 3 : 0; }
    }
    // ---- radix select until at most lcap elements are alive ----
    for (unsigned it = 0;; ++it) {
        if (it >= 20u && active && acnt > lcap) {   // 16 digits always suffice: never spin on a logic error
            if (tl == 0) atomicExch(C.err, 60 + j);
            active = false; alive = 0u; valid = 0u; acnt = 0u;
        }
        const bool need = active && acnt > lcap;
        if (kCta) { if (!__syncthreads_or(need ? 1 : 0)) break; }
        else if (!__any_sync(FULL, need)) break;
        // per-thread counts: sixteen 4-bit fields; digit d lives in the low (d < 8) or high nibble of byte d & 7,
        // so that the two nibble planes are bins 0..7 and 8..15 in order
        unsigned long long acc = 0ull;
        unsigned dig = 0u;
        if (need) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const unsigned u = mode ? (unsigned)comp[e] : (unsigned)(comp[e] >> 32) - base;
                const unsigned d = (u >> sh) & 15u;
                if ((alive >> e) & 1u) { acc += 1ull << (((d & 7u) << 3) | ((d >> 3) << 2)); dig |= d << (4 * e); }
            }
        }
        const unsigned long long pa = acc & 0x0f0f0f0f0f0f0f0full, pb = (acc >> 4) & 0x0f0f0f0f0f0f0f0full;
        unsigned tb = 16u, cbase = 0u, ccnt = 0u;
        if (kCta) {
            // counts of up to 4095: 16-bit fields, word k = bins 2k (low half) and 2k+1 (high half)
            unsigned w[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const unsigned c16 = (unsigned)((k < 4 ? pa : pb) >> (16 * (k & 3))) & 0xffffu;
                w[k] = __reduce_add_sync(FULL, (c16 & 0xffu) | ((c16 >> 8) << 16));
            }
            unsigned* B = C.sbins + (it % 3u) * 64u;
            if (need && lane < 8u) {
                unsigned v = w[0];
#pragma unroll
                for (int k = 1; k < 8; ++k) v = lane == (unsigned)k ? w[k] : v;
                if (v) atomicAdd(&B[q * 8u + lane], v);
            }
            if (tid < 64u) C.sbins[((it + 2u) % 3u) * 64u + tid] = 0u;   // read two barriers ago, used two from now
            __syncthreads();
            if (need) {
                unsigned cum = 0u;
#pragma unroll
                for (int b = 0; b < 16; ++b) {
                    const unsigned c = (B[q * 8u + (b >> 1)] >> (16 * (b & 1))) & 0xffffu;
                    if (tb == 16u && r < cum + c) { tb = (unsigned)b; cbase = cum; ccnt = c; }
                    cum += c;
                }
            }
        } else {
            // at most 255 elements per team: byte fields, word k = bins 4k .. 4k+3
            unsigned w[4] = {(unsigned)pa, (unsigned)(pa >> 32), (unsigned)pb, (unsigned)(pb >> 32)};
#pragma unroll
            for (int k = 0; k < 4; ++k) w[k] = team_add<false>(w[k], tpt);
            // inclusive byte prefixes by one multiply per word (no byte overflows: the total is below 256)
            const unsigned p0 = w[0] * 0x01010101u, p1 = w[1] * 0x01010101u, p2 = w[2] * 0x01010101u, p3 = w[3] * 0x01010101u;
            const unsigned c1 = p0 >> 24, c2 = c1 + (p1 >> 24), c3 = c2 + (p2 >> 24), c4 = c3 + (p3 >> 24);
            if (r < c4) {
                const unsigned kk = (r >= c1 ? 1u : 0u) + (r >= c2 ? 1u : 0u) + (r >= c3 ? 1u : 0u);
                const unsigned pk = kk == 0u ? p0 : kk == 1u ? p1 : kk == 2u ? p2 : p3;
                const unsigned wb = kk == 0u ? 0u : kk == 1u ? c1 : kk == 2u ? c2 : c3;
                const unsigned rr = r - wb;
                const unsigned b0 = pk & 0xffu, b1 = (pk >> 8) & 0xffu, b2 = (pk >> 16) & 0xffu, b3 = pk >> 24;
                const unsigned bi = (rr >= b0 ? 1u : 0u) + (rr >= b1 ? 1u : 0u) + (rr >= b2 ? 1u : 0u);
                const unsigned below = bi == 0u ? 0u : bi == 1u ? b0 : bi == 2u ? b1 : b2;
                const unsigned upto = bi == 0u ? b0 : bi == 1u ? b1 : bi == 2u ? b2 : b3;
                tb = 4u * kk + bi;
                cbase = wb + below;
                ccnt = upto - below;
            }
        }
        if (need) {
            if (tb == 16u) {   // the counts lost the wanted rank: internal error, leave the sub-range as it is
                if (tl == 0) atomicExch(C.err, 40 + j);
                active = false; alive = 0u; valid = 0u; acnt = 0u;
            } else {
                r -= cbase;
                acnt = ccnt;
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (((dig >> (4 * e)) & 15u) != tb) alive &= ~(1u << e);
                if (mode == 0) {
                    base += tb << sh;
                    if (sh == 0) { mode = 1; sh = 28; } else sh = sh > 4 ? sh - 4 : 0;
                } else {
                    sh = sh >= 4 ? sh - 4 : 0;   // after the last index digit a single element is alive
                }
            }
        }
    }
    // ---- the alive elements -> the team's list (ballot ranks; one atomic per warp for multi-warp teams) ----
    {
        unsigned run = 0u;
        if (kCta) {
            const unsigned total = __reduce_add_sync(FULL, (unsigned)__popc(alive));
            if (lane == 0 && total) run = atomicAdd(&C.steam[q].lcnt, total);
            run = __shfl_sync(FULL, run, 0);
        }
        const unsigned lt = (1u << lane) - 1u;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const bool a = (alive >> e) & 1u;
            const unsigned b = __ballot_sync(FULL, a) & tmask;
            if (a && KD_INB(run + __popc(b & lt) < (kCta ? kSubListPerWarp : kSubListPerWarp / tpw), 907)) L[run + __popc(b & lt)] = comp[e];
            run += __popc(b);
        }
        if (active && tl == 0 && r == 0u) XM[1] = 0ull;
    }
    tsync();
    // ---- all-pairs ranks among the candidates: rank r is the node, rank r-1 the left maximum ----
    for (unsigned ci = tl; ci < acnt; ci += tpt) {
        const unsigned long long x = L[ci];
        unsigned rk = 0u;
        for (unsigned k = 0; k < acnt; ++k) rk += L[k] < x ? 1u : 0u;
        if (rk == r) XM[0] = x;
        if (rk + 1u == r) XM[1] = x;
    }
    tsync();
    const unsigned long long X = active ? XM[0] : 0ull;
    unsigned long long M = (active && r > 0u) ? XM[1] : 0ull;
    // side of every element, once: bit e of lessX / isX
    unsigned lessX = 0u, isX = 0u;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        lessX |= (comp[e] < X ? 1u : 0u) << e;
        isX |= (comp[e] == X ? 1u : 0u) << e;
    }
    lessX &= valid;
    {   // the node was the smallest candidate: the left maximum is the largest element below it
        const bool needM = active && r == 0u;
        const bool anyM = kCta ? (__syncthreads_or(needM ? 1 : 0) != 0) : (__any_sync(FULL, needM) != 0);
        if (anyM) {
            unsigned long long best = 0ull;
            if (needM) {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (((lessX >> e) & 1u) && comp[e] > best) best = comp[e];
            }
            const unsigned hk = team_max<kCta>((unsigned)(best >> 32), tpt);
            unsigned li = (best != 0ull && (unsigned)(best >> 32) == hk) ? (unsigned)best + 1u : 0u;
            li = team_max<kCta>(li, tpt);
            if (kCta) {
                if (needM && lane == 0 && li) atomicMax(&XM[1], ((unsigned long long)hk << 32) | (li - 1u));
                __syncthreads();
                if (needM) M = XM[1];
            } else if (needM) {
                M = ((unsigned long long)hk << 32) | (li - 1u);
            }
        }
    }
    // ---- new positions: left part (its maximum last), node, right part; the pinned element stays ----
    unsigned isM = 0u;
#pragma unroll
    for (int e = 0; e < 8; ++e) isM |= (comp[e] == M ? 1u : 0u) << e;
    isM &= lessX;
    const unsigned leftm = lessX & ~isM, rightm = valid & ~lessX & ~isX;
    const unsigned v = (unsigned)__popc(leftm) | ((unsigned)__popc(rightm) << 16);
    unsigned incl = v;
    const unsigned sw = kCta ? 32u : tpt;
    for (unsigned o = 1; o < sw; o <<= 1) {
        const unsigned t = __shfl_up_sync(FULL, incl, o, sw);
        if ((lane & (sw - 1u)) >= o) incl += t;
    }
    unsigned ex = incl - v;
    if (kCta) {
        if (lane == 31) C.swtot[warp] = incl;
        __syncthreads();
        for (unsigned w2 = team_w0; w2 < warp; ++w2) ex += C.swtot[w2];
    }
    unsigned exL = ex & 0xffffu, exR = (ex >> 16) + m + 1u;
#pragma unroll
    for (int e = 0; e < 8; ++e)
        if ((valid >> e) & 1u) {
            unsigned pos;
            if ((lessX >> e) & 1u) pos = ((isM >> e) & 1u) ? m - 1u : exL++;
            else pos = ((isX >> e) & 1u) ? m : exR++;
            if (KD_INB(pos + 1u < sn && slo + pos < (unsigned)kFinishT, 905)) C.sord[slo + pos] = (unsigned short)slot[e];
        }
    if (in_range && tl == 0) {
        tabn[(2u * q) << (logP - 3)] = (slo << 16) | m;
        tabn[(2u * q + 1u) << (logP - 3)] = sn ? (((slo + m + 1u) << 16) | (sn - m - 1u)) : (slo << 16);
    }
    if (kCta) {
        if (tid < 8u) { C.steam[tid].kmin = 0xffffffffu; C.steam[tid].kmax = 0u; C.steam[tid].lcnt = 0u; }
        if (tid < 192u) C.sbins[tid] = 0u;
        __syncthreads();
    } else {
        __syncwarp();
    }
}

// Thread levels (P = 8 or 4): a thread owns 8 / P whole sub-ranges; the ranks of a sub-range's free elements
// among themselves are their new positions (the reference's sort of the first n-1, done by counting).
template <int LOGP>
__device__ __forceinline__ void sub_thread_level(const SubCtx& C, int j, int axis) {
    constexpr int P = 1 << LOGP, G = 8 / P;
    const unsigned tid = threadIdx.x, nsub = 1u << j;
    const unsigned* tab = C.stab + (j & 1) * (kFinishT / 2);
    unsigned* tabn = C.stab + ((j + 1) & 1) * (kFinishT / 2);
    const float* __restrict__ kcol = C.ncol ? C.scol + (size_t)axis * kSubTS : nullptr;
    unsigned long long comp[8];
    unsigned slot[8], rank[8], glo[G], gn[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const unsigned q = tid * G + g;
        const unsigned ent = q < nsub ? tab[q << (LOGP - 2)] : 0u;
        glo[g] = ent >> 16;
        gn[g] = ent & 0xffffu;
        if (LOGP == 3 && q < nsub) {   // (the children of a P = 4 level are never looked up)
            const unsigned sn = gn[g], m = sn >> 1;
            tabn[2u * q] = (glo[g] << 16) | m;
            tabn[2u * q + 1u] = sn ? (((glo[g] + m + 1u) << 16) | (sn - m - 1u)) : (glo[g] << 16);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const unsigned sn = gn[e >> LOGP], slo = glo[e >> LOGP], i = (unsigned)(e & (P - 1));
        comp[e] = ~0ull;
        slot[e] = 0u;
        rank[e] = 0u;
        if (sn >= 3u && i + 1u < sn) {
            const unsigned s = C.sord[slo + i];
            const unsigned id = (unsigned)C.sidx[s];
            const float f = kcol ? kcol[s] : C.pts[(size_t)id * C.dim + axis];
            comp[e] = composite(ordered_key(f), id);
            slot[e] = s;
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e)
#pragma unroll
        for (int f = e + 1; f < 8; ++f)
            if ((e >> LOGP) == (f >> LOGP)) {   // same sub-range (padding is larger than every element)
                const unsigned lt = comp[f] < comp[e] ? 1u : 0u;
                rank[e] += lt;
                rank[f] += 1u - lt;
            }
#pragma unroll
    for (int e = 0; e < 8; ++e)
        if (comp[e] != ~0ull && KD_INB(rank[e] + 1u < gn[e >> LOGP], 906)) C.sord[glo[e >> LOGP] + rank[e]] = (unsigned short)slot[e];
    __syncwarp();
}

__global__ void __launch_bounds__(kSubBlock, 2) k_subtree(Cols src, int ncol, const float* __restrict__ pts, int dim,
                                                          unsigned N, int level, int depth0, OutPtrs out,
                                                          unsigned out_lo, int* err, int max_levels) {
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned lo, n;
    seg_range(N, level, blockIdx.x, lo, n);
    if (n == 0) return;
    SubCtx C;
    C.scol = reinterpret_cast<float*>(smem);
    C.sidx = reinterpret_cast<int32_t*>(C.scol + (size_t)ncol * kSubTS);
    C.slist = reinterpret_cast<unsigned long long*>(C.sidx + kSubTS);
    C.steam = reinterpret_cast<SubTeam*>(C.slist + (size_t)kSubWarps * kSubWarpScratch);
    C.sbins = reinterpret_cast<unsigned*>(C.steam + 8);
    C.swtot = C.sbins + 192;
    C.stab = C.swtot + kSubWarps;
    C.sord = reinterpret_cast<unsigned short*>(C.stab + kFinishT);
    C.pts = pts;
    C.ncol = ncol;
    C.dim = dim;
    C.err = err;
    __shared__ __align__(8) unsigned long long fin_bar;
    const unsigned tid = threadIdx.x;
    // Stage the window with the TMA engine: one bulk copy per column (cp.async.bulk, completion counted in
    // bytes on an mbarrier) instead of 512 threads issuing scalar loads.  Bulk copies need 16-byte aligned
    // addresses and sizes, so the copy starts `delta` elements early and slots are offsets into that window.
    // (all columns share one alignment phase: their bases are 256-byte aligned plus a common element offset)
    const unsigned delta = (unsigned)((reinterpret_cast<uintptr_t>(src.idx + lo) >> 2) & 3u);
    const ptrdiff_t alo = (ptrdiff_t)lo - (ptrdiff_t)delta;
    const unsigned bytes = ((delta + n + 3u) & ~3u) * 4u;
    if (tid == 0) mbar_init(&fin_bar, 1);
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(&fin_bar, bytes * (unsigned)(ncol + 1));
        for (int c = 0; c < ncol; ++c) tma_load_1d(C.scol + (size_t)c * kSubTS, src.c[c] + alo, bytes, &fin_bar);
        tma_load_1d(C.sidx, src.idx + alo, bytes, &fin_bar);
    }
    for (unsigned p = tid; p < n; p += kSubBlock) C.sord[p] = (unsigned short)(p + delta);
    if (tid < 8u) { C.steam[tid].kmin = 0xffffffffu; C.steam[tid].kmax = 0u; C.steam[tid].lcnt = 0u; }
    if (tid < 192u) C.sbins[tid] = 0u;
    if (tid == 0) C.stab[0] = n;   // sub-range 0 of sub-level 0: lo = 0, size n
    if (!__syncthreads_and(mbar_wait(&fin_bar, 0) ? 1 : 0)) {  // never expected; same data through the LSU instead
        for (unsigned p = tid; p < n; p += kSubBlock) {
            C.sidx[delta + p] = src.idx[lo + p];
            for (int c = 0; c < ncol; ++c) C.scol[(size_t)c * kSubTS + delta + p] = src.c[c][lo + p];
        }
        __syncthreads();
    }
    int logP0 = 3;
    while ((1u << logP0) < n) ++logP0;
    int j = 0;
    const int d0 = depth0 + level;
    for (; (n >> j) >= 3u && logP0 - j >= 9 && j < max_levels; ++j) sub_team_level<true>(C, j, logP0 - j, (d0 + j) % dim);
    // from here on every warp owns its sub-ranges: no CTA-wide barrier until the write-out
    for (; (n >> j) >= 3u && j < max_levels; ++j) {   // (max_levels: debugging aid, KD_B200_SUBTREE_MAXLEVELS)
        const int logP = logP0 - j, axis = (d0 + j) % dim;
        if (logP >= 4) sub_team_level<false>(C, j, logP, axis);
        else if (logP == 3) sub_thread_level<3>(C, j, axis);
        else sub_thread_level<2>(C, j, axis);
    }
    __syncthreads();
    // Every replica of the tree arrays receives this window: the sub-tree all-gather is fused into the store.
    // Both arrays are written as flat runs (consecutive threads -> consecutive 4-byte words), so a warp emits
    // whole 128-byte lines, which is what NVLink wants for the remote replicas.
    const unsigned total = ncol ? n * (unsigned)dim : 0u;
    for (int r = 0; r < out.n; ++r) {
        int32_t* __restrict__ perm = out.perm[r] + out_lo + lo;
        float* __restrict__ tree_pts = out.tp[r] ? out.tp[r] + (size_t)(out_lo + lo) * dim : nullptr;
        for (unsigned p = tid; p < n; p += kSubBlock) perm[p] = C.sidx[C.sord[p]];
        if (tree_pts) {
            for (unsigned f = tid; f < total; f += kSubBlock) {
                const unsigned p = f / (unsigned)dim, c = f - p * (unsigned)dim;
                tree_pts[f] = C.scol[(size_t)c * kSubTS + C.sord[p]];
            }
        }
    }
}

__global__ void k_gather_rows(const float* __restrict__ pts, const int32_t* __restrict__ perm, unsigned N, int dim,
                              float* __restrict__ out) {
    size_t total = (size_t)N * dim;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        size_t row = e / dim;
        int c = (int)(e - row * dim);
        out[e] = pts[(size_t)perm[row] * dim + c];
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static thread_local int64_t g_launches = 0;

BuildPlan make_plan(int64_t n, int dim) {
    BuildPlan p{};
    p.n = n;
    p.dim = dim;
    p.ncol = dim <= kMaxCarry ? dim : 0;
    // levels handled with several CTAs per range and a global histogram: while the largest range
    // (the left-most one, n >> level) exceeds kMediumMax
    int lv = 0;
    while ((n >> lv) > kMediumMax) ++lv;
    p.finish_level = lv;
    p.max_seg = lv == 0 ? 1 : (1ll << (lv - 1));  // ranges of the deepest such level
    p.max_rec = p.max_seg + kTargetCtas;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    for (int b = 0; b < 2; ++b) {
        for (int c = 0; c < p.ncol; ++c) p.off_col[b][c] = take((size_t)n * 4);
        p.off_idx[b] = take((size_t)n * 4);
    }
    p.off_key = p.ncol ? 0 : take((size_t)n * 4);
    p.off_state = take((size_t)p.max_seg * sizeof(SegState));
    p.off_hist = take((size_t)p.max_seg * kDigitBins * 4);
    p.off_rec = take((size_t)p.max_rec * 16);
    p.off_err = take(256);
    // candidate arrays of the bracket select, only when there are large-range levels
    p.off_ckey = lv > 0 ? take((size_t)n * 4) : 0;
    p.off_cidx = lv > 0 ? take((size_t)n * 4) : 0;
    p.off_ccta = take((size_t)p.max_rec * 4);
    p.bytes = off;
    return p;
}

static inline int ctas_per_seg(int64_t nseg, int64_t seg_n, int64_t min_chunk) {
    int64_t want = (kTargetCtas + nseg - 1) / nseg;
    int64_t cap = std::max<int64_t>(1, (seg_n + min_chunk - 1) / min_chunk);
    return (int)std::max<int64_t>(1, std::min(want, cap));
}

#define KD_LAUNCH_CHECK(what)                                                              \
    do {                                                                                   \
        ++g_launches;                                                                      \
        cudaError_t e_ = cudaGetLastError();                                               \
        if (e_ != cudaSuccess) return fail(std::string(what) + ": " + cudaGetErrorString(e_)); \
    } while (0)

static int subtree_max_levels() {
    static const int v = [] { const char* e = std::getenv("KD_B200_SUBTREE_MAXLEVELS"); return e ? std::atoi(e) : 64; }();
    return v;
}

// Opt-in shared-memory sizes are a property of the (function, device) pair: set them once per device, not per build.
static int set_func_attributes_once() {
    static std::mutex mu;
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(std::string("kd build: cudaGetDevice: ") + cudaGetErrorString(e));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && done[dev]) return 0;
    e = cudaFuncSetAttribute(k_subtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)subtree_smem_bytes(kMaxCarry));
#define KD_MED_ATTR(NC) if (e == cudaSuccess) e = cudaFuncSetAttribute(k_medium<NC, 11>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)medium_smem_bytes())
    KD_MED_ATTR(0); KD_MED_ATTR(1); KD_MED_ATTR(2); KD_MED_ATTR(3); KD_MED_ATTR(4);
    KD_MED_ATTR(5); KD_MED_ATTR(6); KD_MED_ATTR(7); KD_MED_ATTR(8);
#undef KD_MED_ATTR
    if (e != cudaSuccess) return fail(std::string("kd build: cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    if (dev >= 0 && dev < 64) done[dev] = true;
    return 0;
}

// Builds the sub-tree of one range whose points already sit in buf[cur] (slots 0..n-1): select/partition
// levels, then the shared-memory sub-tree builder.  `depth0` is the depth of the range's node in the whole tree (it fixes the
// split axes), `out_lo` the slot of the range's first point in the tree arrays, `out` every replica of
// the tree arrays that receives the result.
int build_core(Cols buf[2], int cur, int64_t n, int dim, int ncol, int depth0, unsigned out_lo, OutPtrs out,
               const float* points, BuildScratch sc, cudaStream_t stream, int stop_level, int* cur_out) {
    if (int rc = set_func_attributes_once()) return rc;
    const size_t sub_smem = subtree_smem_bytes(ncol);
    const unsigned N = (unsigned)n;
    // Level schedule: select/partition with several CTAs per range while ranges are large, one CTA per
    // range (k_medium) while they exceed kFinishT, then the shared-memory sub-tree builder takes every
    // range of that level down to its leaves.
    int finish_level = 0;
    while ((n >> finish_level) > kFinishT) ++finish_level;
    for (int level = 0;; ++level) {
        const int64_t nseg = 1ll << level;
        if (level == stop_level) {   // the caller continues range by range (build_core on every range of this level)
            if (cur_out) *cur_out = cur;
            return 0;
        }
        if (level == finish_level) {
            k_subtree<<<(unsigned)nseg, kSubBlock, sub_smem, stream>>>(buf[cur], ncol, points, dim, N, level, depth0,
                                                                       out, out_lo, sc.err, subtree_max_levels());
            KD_LAUNCH_CHECK("k_subtree");
            break;
        }
        if ((n >> level) <= kMediumMax) {
            const int axis = (depth0 + level) % dim;
            const float* keys = nullptr;
            if (!ncol) {
                k_gather_keys<<<(unsigned)nseg, 256, 0, stream>>>(points, dim, axis, buf[cur].idx, sc.keybuf, N, level, 1);
                KD_LAUNCH_CHECK("k_gather_keys");
                keys = sc.keybuf;
            }
            const int64_t seg_n = n >> level;
            const int block = seg_n > 4096 ? 512 : (seg_n > 1024 ? 256 : 128);
            const unsigned cap = (unsigned)((seg_n + 31) & ~31ll);   // shared memory sized to the level's ranges
            const bool narrow = seg_n <= 4096;   // 8-bit digits for the small ranges
            const size_t msmem = (size_t)cap * 4 + (size_t)(narrow ? 256 : kDigitBins) * 4;
#define KD_MED(NC)                                                                                                  \
    do {                                                                                                            \
        if (narrow)                                                                                                 \
            k_medium<NC, 8><<<(unsigned)nseg, block, msmem, stream>>>(buf[cur], buf[cur ^ 1], keys, axis, out, out_lo, \
                                                                      dim, N, level, cap, sc.err);                  \
        else                                                                                                        \
            k_medium<NC, 11><<<(unsigned)nseg, block, msmem, stream>>>(buf[cur], buf[cur ^ 1], keys, axis, out,      \
                                                                       out_lo, dim, N, level, cap, sc.err);         \
    } while (0)
            switch (ncol) {
                case 0: KD_MED(0); break;
                case 1: KD_MED(1); break;
                case 2: KD_MED(2); break;
                case 3: KD_MED(3); break;
                case 4: KD_MED(4); break;
                case 5: KD_MED(5); break;
                case 6: KD_MED(6); break;
                case 7: KD_MED(7); break;
                default: KD_MED(8); break;
            }
#undef KD_MED
            KD_LAUNCH_CHECK("k_medium");
            cur ^= 1;
            continue;
        }
        const int64_t seg_n = n >> level;
        const int axis = (depth0 + level) % dim;
        const int cps_h = ctas_per_seg(nseg, seg_n, 8192);
        const int cps_p = ctas_per_seg(nseg, seg_n, 2048);
        k_state_init<<<(unsigned)((nseg + 255) / 256), 256, 0, stream>>>(sc.st, N, level, (unsigned)nseg);
        KD_LAUNCH_CHECK("k_state_init");
        const float* keys;
        if (ncol) {
            keys = buf[cur].c[axis];
        } else {
            k_gather_keys<<<(unsigned)(nseg * cps_p), 256, 0, stream>>>(points, dim, axis, buf[cur].idx, sc.keybuf, N,
                                                                        level, cps_p);
            KD_LAUNCH_CHECK("k_gather_keys");
            keys = sc.keybuf;
        }
        const bool bracket = sc.cand_key != nullptr;
        if (bracket) {
            if (seg_n >= (1ll << 20)) k_sample<4096><<<(unsigned)nseg, 512, 0, stream>>>(keys, buf[cur].idx, sc.st, N, level);
            else k_sample<1024><<<(unsigned)nseg, 512, 0, stream>>>(keys, buf[cur].idx, sc.st, N, level);
            KD_LAUNCH_CHECK("k_sample");
            k_bracket<512, 8><<<(unsigned)(nseg * cps_p), 512, 0, stream>>>(keys, buf[cur].idx, sc.st, sc.cand_key,
                                                                           sc.cand_idx, sc.cand_cta, N, level, cps_p);
            KD_LAUNCH_CHECK("k_bracket");
            k_bracket_finish<<<(unsigned)((nseg + 255) / 256), 256, 0, stream>>>(sc.st, N, level, (unsigned)nseg);
            KD_LAUNCH_CHECK("k_bracket_finish");
        }
        // (a candidate read-back requested by pass d's scan is served by pass d+1's k_hist; pass 5 never asks)
        for (int d = 0; d < kNumDigits; ++d) {
            // with a bracket the passes use the bracket's CTA geometry (per-CTA candidate runs)
            const int cps_d = bracket ? cps_p : cps_h;
            k_hist<512, 4><<<(unsigned)(nseg * cps_d), 512, 0, stream>>>(keys, buf[cur].idx, sc.cand_key, sc.cand_idx,
                                                                        sc.cand_cta, sc.st, sc.ghist, N, level, cps_d, d);
            KD_LAUNCH_CHECK("k_hist");
            k_scan<<<(unsigned)nseg, 256, 0, stream>>>(sc.st, sc.ghist, d, sc.err);
            KD_LAUNCH_CHECK("k_scan");
        }
        {
            constexpr int PB = 128, PI = 4;
            const unsigned grid = (unsigned)(nseg * cps_p);
#define KD_PART(NC)                                                                                                \
    k_partition<NC, PB, PI><<<grid, PB, 0, stream>>>(buf[cur], buf[cur ^ 1], keys, axis, sc.st, sc.rec, out, out_lo, \
                                                     dim, N, level, cps_p)
            switch (ncol) {
                case 0: KD_PART(0); break;
                case 1: KD_PART(1); break;
                case 2: KD_PART(2); break;
                case 3: KD_PART(3); break;
                case 4: KD_PART(4); break;
                case 5: KD_PART(5); break;
                case 6: KD_PART(6); break;
                case 7: KD_PART(7); break;
                default: KD_PART(8); break;
            }
#undef KD_PART
        }
        KD_LAUNCH_CHECK("k_partition");
        k_fixup<<<(unsigned)((nseg + 7) / 8), 256, 0, stream>>>(buf[cur ^ 1], ncol, sc.rec, N, level, (unsigned)nseg,
                                                               cps_p);
        KD_LAUNCH_CHECK("k_fixup");
        cur ^= 1;
    }
    return 0;
}

int build_gpu(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points, void* workspace,
              size_t workspace_bytes, cudaStream_t stream) {
    if (!points || !perm || !workspace || n <= 0 || dim <= 0) return fail("kd_build_gpu: bad arguments");
    if (n >= (1ll << 31)) return fail("kd_build_gpu: n must be below 2^31");
    BuildPlan p = make_plan(n, dim);
    if (workspace_bytes < p.bytes) return fail("kd_build_gpu: workspace too small");
    if (((uintptr_t)workspace & 255u) != 0) return fail("kd_build_gpu: the workspace must be 256-byte aligned");
    g_launches = 0;
    unsigned char* ws = (unsigned char*)workspace;
    Cols buf[2];
    for (int b = 0; b < 2; ++b) {
        for (int c = 0; c < kMaxCarry; ++c) buf[b].c[c] = c < p.ncol ? (float*)(ws + p.off_col[b][c]) : nullptr;
        buf[b].idx = (int32_t*)(ws + p.off_idx[b]);
    }
    BuildScratch sc;
    sc.keybuf = p.ncol ? nullptr : (float*)(ws + p.off_key);
    sc.st = (SegState*)(ws + p.off_state);
    sc.ghist = (unsigned*)(ws + p.off_hist);
    sc.rec = (MaxRec*)(ws + p.off_rec);
    sc.err = (int*)(ws + p.off_err);
    sc.cand_key = p.finish_level > 0 ? (float*)(ws + p.off_ckey) : nullptr;
    sc.cand_idx = p.finish_level > 0 ? (int32_t*)(ws + p.off_cidx) : nullptr;
    sc.cand_cta = (unsigned*)(ws + p.off_ccta);
    const unsigned N = (unsigned)n;

    if (cudaMemsetAsync(sc.ghist, 0, (size_t)p.max_seg * kDigitBins * 4, stream) != cudaSuccess ||
        cudaMemsetAsync(sc.err, 0, 256, stream) != cudaSuccess)
        return fail("kd_build_gpu: cudaMemsetAsync failed");
    k_init<<<(unsigned)((n + kInitRows - 1) / kInitRows), kInitBlock, 0, stream>>>(points, N, dim, p.ncol, buf[0]);
    KD_LAUNCH_CHECK("k_init");
    OutPtrs out{};
    out.n = 1;
    out.perm[0] = perm;
    out.tp[0] = tree_points;
    if (int rc = build_core(buf, 0, n, dim, p.ncol, 0, 0u, out, points, sc, stream, -1, nullptr)) return rc;
    if (tree_points && !p.ncol) {
        k_gather_rows<<<148 * 8, 256, 0, stream>>>(points, perm, N, dim, tree_points);
        KD_LAUNCH_CHECK("k_gather_rows");
    }
    return 0;
}

void count_launch() { ++g_launches; }
void reset_launches() { g_launches = 0; }
int64_t last_launches() { return g_launches; }

}  // namespace kd

extern "C" {

size_t kd_build_workspace_bytes(int64_t n, int dim) {
    if (n <= 0 || dim <= 0) return 0;
    return kd::make_plan(n, dim).bytes;
}

int kd_build_gpu(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points, void* workspace,
                 size_t workspace_bytes, void* stream) {
    return kd::build_gpu(points, n, dim, perm, tree_points, workspace, workspace_bytes, (cudaStream_t)stream);
}

int64_t kd_last_launch_count(void) { return kd::last_launches(); }

int kd_build_status(const void* workspace, int64_t n, int dim, void* stream) {
    if (!workspace || n <= 0 || dim <= 0) return kd::fail("kd_build_status: bad arguments");
    kd::BuildPlan p = kd::make_plan(n, dim);
    int err = 0;
    cudaError_t e = cudaMemcpyAsync(&err, (const unsigned char*)workspace + p.off_err, sizeof(int),
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess) return kd::fail(std::string("kd_build_status: ") + cudaGetErrorString(e));
    if (err) return kd::fail("kd_build_gpu: select lost its rank in digit pass " + std::to_string(err - 1));
    return 0;
}

/* -DKD_DEBUG builds: device-side bounds assertions (see kd_common.h).  Returns 0 if none failed since the last call
 * (and in release builds, where kd_debug_enabled() is 0), else fails with the code of the first one. */
int kd_debug_enabled(void) {
#if defined(KD_DEBUG)
    return 1;
#else
    return 0;
#endif
}
int kd_dist_debug_err(void);
int kd_debug_bounds_status(void) {
#if defined(KD_DEBUG)
    int a = 0, zero = 0;
    if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpyFromSymbol(&a, kd_dbg_err, sizeof(int)) != cudaSuccess ||
        cudaMemcpyToSymbol(kd_dbg_err, &zero, sizeof(int)) != cudaSuccess)
        return kd::fail("kd_debug_bounds_status: could not read the assertion word");
    const int b = kd_dist_debug_err();
    if (a || b) return kd::fail("bounds assertion failed on the device, code " + std::to_string(a ? a : b));
#endif
    return 0;
}

int kd_gather_rows_gpu(const float* points, const int32_t* perm, int64_t n, int dim, float* out, void* stream) {
    if (!points || !perm || !out || n <= 0 || dim <= 0) return kd::fail("kd_gather_rows_gpu: bad arguments");
    kd::k_gather_rows<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(points, perm, (unsigned)n, dim, out);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_gather_rows_gpu: ") + cudaGetErrorString(e));
}

}  // extern "C"
