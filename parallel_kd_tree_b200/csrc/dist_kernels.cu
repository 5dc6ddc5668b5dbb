// Distributed (multi-GPU) build of ONE global tree over peer memory (NVLink 5 / NVSwitch), no NCCL on
// the data path.  New scope with respect to the reference: kdtree_mpi.cpp builds W independent local
// trees (kdtree_mpi.cpp:204-230); here the W GPUs build the tree the reference's *sequential* program
// builds over all N points (kdtree_sequential.cpp:30-66), bit for bit.
//
// The top log2(W) levels span GPUs.  At level l the W ranks form W >> l groups; group g owns range g of
// that level, spread over its members.  Per level (level_schedule() lists the steps):
//   K5  k_dsample / k_dpivots /     sample bracket across the group: sorted samples to every member (one-shot P2P),
//       k_dbracket(_wait)           identical pivots everywhere, ONE counting / compaction pass over the local keys;
//       k_hist_dist + k_scan_dist   radix select over the ~5 % candidates; the last CTA of a rank's histogram kernel
//                                   publishes its 2048 bins with NVLS multimem.red into the group's accumulator on
//                                   every GPU (or stores them into a slot in every member's memory where no
//                                   multicast mapping exists) and bumps an arrival counter / flag; every member
//                                   picks the same bin.  No collective call, no host round trip.
//   K6  k_partition_dist            stream-compaction partition whose stores go straight into the owner GPUs'
//                                   next-level buffers (staged runs, 16-byte P2P stores); members' runs are ordered so
//                                   that every GPU keeps one of its halves at home; the node goes to every GPU's
//                                   tree arrays; per-rank counts are exchanged one-shot.
//       k_max_*                     the left range's largest element is swapped into its last slot
//                                   (it is the next pinned element) by the GPU that owns that slot.
// After the last distributed level every GPU owns one complete sub-range and finishes it with the
// single-GPU kernels, whose stores are replicated into every GPU's tree arrays (K7: the sub-tree
// all-gather is fused into k_subtree's write-out; k_replicate serves the optional sliced local phase).
// dim > 8: gather mode (k_rows_push once, k_dist_gather_keys per level, only indices move).
//
// Every wait on a peer is a separate tiny kernel from the matching post, so the same code also runs
// as W *virtual* ranks of one process on one GPU (posts of all ranks are launched before the waits);
// that is how the logic is tested without a multi-GPU box.  Flags carry a monotonically increasing
// epoch and are never reset.  Waits give up after 10 s of wall time, record an error, and every later kernel of the
// build returns at once (kd_dist_status reports it).  TMA is used for loads of this GPU's own memory only.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/kdtree_b200.h"
#include "build_plan.h"
#include "kd_common.h"
#include "kd_error.h"

namespace kd {

constexpr int kMaxWorld = 8;
enum FlagKind { F_HIST = 0, F_CNT, F_PART, F_MAX, F_FIX, F_DONE, F_SAMP, F_BRK, F_ROWS, kNumFlagKinds };
constexpr int kDSample = 1024;          // sample composites per rank per distributed level
constexpr int kRedStride = kDigitBins + 32;   // one NVLS accumulator: the bins + its arrival counter (+ padding)
constexpr int kRedPasses = 8;

struct DistCtl {
    unsigned long long prefix;
    unsigned rank_wanted, cnt_global;
    int shift, done;
    unsigned my_less;      // local elements already known to sort before the node
    unsigned mine;         // 1 if this rank holds the node
    unsigned cnt;          // local element count of the current level
    unsigned Loff, Roff;   // first left / right output offset of this rank inside the child ranges
    unsigned curL, curR;   // local partition cursors
    unsigned ticket;       // last-CTA election of k_hist_dist
    int err;
    // sample bracket (all members derive the same pivots from the union of the ranks' samples)
    unsigned cL;           // local elements below lowP
    unsigned cand_cnt;     // local candidates (lowP <= composite <= highP), compacted to the candidate arrays
    int use_cand;          // 1: the digit passes run on the candidates
    unsigned long long lowP, highP;
    unsigned bticket;      // last-CTA election of k_dbracket
    unsigned pad[3];
};

struct MaxSlot { unsigned long long comp; unsigned off; unsigned valid; };

struct DistLayout {
    size_t ctl, flags, hist_slots, lhist, own_hist, samp_slots, brk_slots, racc, cnt_slots, max_slots, rec, col[2][kMaxCarry],
        idx[2], perm, tp;
    size_t st, ghist, lrec, lerr, ckey, cidx, ccta;  // scratch of the local (single-GPU) phase
    size_t pts_all, keybuf;                          // dim > kMaxCarry: all points on every rank, split-axis keys per level
    size_t bytes;
    int64_t cap, max_seg, max_rec;
};

static DistLayout make_layout(int64_t n_total, int dim, int world, bool with_tp) {
    DistLayout L{};
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    L.cap = n_total / world + n_total % world + world + 8;
    L.ctl = take(sizeof(DistCtl));
    L.flags = take(sizeof(unsigned long long) * kNumFlagKinds * kMaxWorld);
    L.hist_slots = take((size_t)2 * kMaxWorld * kDigitBins * 4);
    L.lhist = take((size_t)kDigitBins * 4);
    L.own_hist = take((size_t)2 * kDigitBins * 4);
    L.samp_slots = take((size_t)kMaxWorld * kDSample * 8);
    L.brk_slots = take((size_t)kMaxWorld * 4 * 4);
    // NVLS accumulators of the histogram reduction: [build parity][level 0: 1 | level 1: 2 | level 2: 4 groups][pass]
    L.racc = take((size_t)2 * 7 * kRedPasses * kRedStride * 4);
    L.cnt_slots = take((size_t)kMaxWorld * 4 * 4);
    L.max_slots = take(sizeof(MaxSlot) * kMaxWorld);
    L.rec = take(sizeof(MaxRec) * (size_t)(kTargetCtas + 8));
    // dim <= kMaxCarry: the coordinates travel with the points in SoA columns.  Larger dim (the reference's default
    // is 128, Utility.cpp:98): every rank receives ALL points once (the ranks' slices are pushed to every peer before
    // level 0), only the index column moves afterwards and the split-axis key is gathered per level.
    const int ncol = dim <= kMaxCarry ? dim : 0;
    for (int b = 0; b < 2; ++b) {
        for (int c = 0; c < ncol; ++c) L.col[b][c] = take((size_t)L.cap * 4);
        L.idx[b] = take((size_t)L.cap * 4);
    }
    L.pts_all = ncol ? 0 : take((size_t)n_total * dim * 4);
    L.keybuf = ncol ? 0 : take((size_t)L.cap * 4);
    L.perm = take((size_t)n_total * 4);
    L.tp = with_tp ? take((size_t)n_total * dim * 4) : 0;
    int lv = 0;
    while ((L.cap >> lv) > kMediumMax) ++lv;
    L.max_seg = lv == 0 ? 1 : (1ll << (lv - 1));
    L.max_rec = L.max_seg + kTargetCtas;
    L.st = take((size_t)L.max_seg * sizeof(SegState));
    L.ghist = take((size_t)L.max_seg * kDigitBins * 4);
    L.lrec = take((size_t)(L.max_rec + 8 * kTargetCtas) * sizeof(MaxRec));   // (+ room for 8 concurrent slices)
    L.lerr = take(256);
    L.ckey = take((size_t)L.cap * 4);
    L.cidx = take((size_t)L.cap * 4);
    L.ccta = take((size_t)(L.max_rec + 8 * kTargetCtas) * 4);
    L.bytes = off;
    return L;
}

struct DistParams {
    int world, rank, dim, ncol;
    unsigned N;
    unsigned char* heap[kMaxWorld];  // this rank's view of every rank's heap
    unsigned char* mc;               // NVLS multicast view of the heaps (a store lands on every GPU) or null
    DistLayout L;
};

// geometry of one distributed level, identical on all members of a group
struct LevelGeo {
    int level, G, gb, j;   // group size, first rank of my group, my member index
    unsigned lo, n;        // my group's range
};

__host__ __device__ inline void host_seg_range(unsigned N, int level, unsigned s, unsigned& lo, unsigned& n) {
    unsigned l = 0, cnt = N;
    for (int b = level - 1; b >= 0; --b) {
        unsigned m = cnt >> 1;
        if ((s >> b) & 1u) { l += m + 1; cnt = cnt - m - 1; } else { cnt = m; }
    }
    lo = l;
    n = cnt;
}

__host__ __device__ inline unsigned share_of(unsigned n, unsigned G, unsigned j) { return n / G + (j < n % G ? 1u : 0u); }

// offset o of a range of n elements spread evenly over G members -> (member, local position)
__device__ __forceinline__ void owner_of(unsigned o, unsigned n, unsigned G, unsigned& member, unsigned& pos) {
    const unsigned base = n / G, rem = n % G, big = rem * (base + 1u);
    if (o < big) { member = o / (base + 1u); pos = o - member * (base + 1u); }
    else { const unsigned o2 = o - big; const unsigned q = o2 / base; member = rem + q; pos = o2 - q * base; }
}

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Peer waits give up after kWaitBudgetNs of WALL time (the global nanosecond timer, independent of the SM clock) and
// report an error instead of hanging the device; every kernel that follows a failed wait returns at once (ctl->err).
constexpr unsigned long long kWaitBudgetNs = 10ull * 1000 * 1000 * 1000;   // 10 s: far beyond any lazy module load or peer mapping
__device__ __forceinline__ bool wait_flag(const unsigned long long* p, unsigned long long epoch) {
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(p) < epoch) {
        if (global_ns() - t0 > kWaitBudgetNs) return false;
        __nanosleep(64);
    }
    return true;
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool wait_count(const unsigned* p, unsigned want) {
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys_u32(p) < want) {
        if (global_ns() - t0 > kWaitBudgetNs) return false;
        __nanosleep(64);
    }
    return true;
}
// NVLS: one reduction store on the multicast address adds `v` into the same word of every GPU's heap (the switch
// fans it out); the release form orders this thread's earlier multicast reductions before it at every destination.
__device__ __forceinline__ void multimem_red_add(unsigned* mc, unsigned v) {
    asm volatile("multimem.red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_release(unsigned* mc, unsigned v) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
// accumulator of (build parity, level, group, pass) inside a heap (byte offset)
__host__ __device__ inline size_t racc_off(const DistLayout& L, unsigned parity, int level, int group, int pass) {
    const int area = (level == 0 ? 0 : level == 1 ? 1 : 3) + group;
    return L.racc + ((size_t)(parity * 7 + area) * kRedPasses + (size_t)pass) * kRedStride * 4;
}

#define DCTL(P, r) (reinterpret_cast<DistCtl*>((P).heap[r] + (P).L.ctl))
#define DFLAG(P, r, kind, src) (reinterpret_cast<unsigned long long*>((P).heap[r] + (P).L.flags) + (kind) * kMaxWorld + (src))

__device__ __forceinline__ Cols cols_of(const DistParams& P, int r, int b) {
    Cols c;
    for (int k = 0; k < kMaxCarry; ++k) c.c[k] = k < P.ncol ? reinterpret_cast<float*>(P.heap[r] + P.L.col[b][k]) : nullptr;
    c.idx = reinterpret_cast<int32_t*>(P.heap[r] + P.L.idx[b]);
    return c;
}

// ------------------------------------------------------------------------------------------------
// AoS rows -> SoA columns of the rank's slice through a shared-memory tile (rows read as one flat coalesced run,
// every column written as a coalesced run; same scheme as k_init of the single-GPU build)
constexpr int kDInitRows = 1024, kDInitBlock = 256;
__global__ void __launch_bounds__(kDInitBlock) k_dist_init(DistParams P, const float* __restrict__ pts, unsigned local_n,
                                                          unsigned first_row) {
    __shared__ float tile[kDInitRows * kMaxCarry];
    const unsigned row0 = blockIdx.x * kDInitRows, tid = threadIdx.x;
    Cols dst = cols_of(P, P.rank, 0);
    if (blockIdx.x == 0 && tid == 0) {
        DistCtl* c = DCTL(P, P.rank);
        c->cnt = local_n;
        c->err = 0;
        c->ticket = 0;
    }
    if (row0 >= local_n) return;
    const unsigned rows = min((unsigned)kDInitRows, local_n - row0);
    for (unsigned r = tid; r < rows; r += kDInitBlock) dst.idx[row0 + r] = (int32_t)(first_row + row0 + r);
    const unsigned total = rows * (unsigned)P.dim;
    const float* __restrict__ src = pts + (size_t)row0 * P.dim;
#pragma unroll 4
    for (unsigned e = tid; e < total; e += kDInitBlock) tile[e] = src[e];
    __syncthreads();
    for (int c = 0; c < P.ncol; ++c)
#pragma unroll 4
        for (unsigned r = tid; r < rows; r += kDInitBlock) dst.c[c][row0 + r] = tile[r * P.dim + c];
}

// gather mode (dim > kMaxCarry): the rank's rows go to its OWN copy of the full point array, indices to column 0
__global__ void __launch_bounds__(256) k_dist_init_rows(DistParams P, const float* __restrict__ pts, unsigned local_n,
                                                        unsigned first_row) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    if (tid == 0) {
        DistCtl* c = DCTL(P, P.rank);
        c->cnt = local_n;
        c->err = 0;
        c->ticket = 0;
    }
    int32_t* idx = reinterpret_cast<int32_t*>(P.heap[P.rank] + P.L.idx[0]);
    float* all = reinterpret_cast<float*>(P.heap[P.rank] + P.L.pts_all) + (size_t)first_row * P.dim;
    for (size_t i = tid; i < local_n; i += nth) idx[i] = (int32_t)(first_row + i);
    const size_t total = (size_t)local_n * P.dim;
    for (size_t e = tid; e < total; e += nth) all[e] = pts[e];
}

// split-axis keys of the rank's current elements (gather mode): keybuf[i] = points[idx[i]][axis]
__global__ void __launch_bounds__(256) k_dist_gather_keys(DistParams P, LevelGeo g, int cur, int axis) {
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    const unsigned local_n = g.level > 0 ? share_of(g.n, (unsigned)g.G, (unsigned)g.j) : ctl->cnt;
    const int32_t* __restrict__ idx = reinterpret_cast<const int32_t*>(P.heap[P.rank] + P.L.idx[cur]);
    const float* __restrict__ all = reinterpret_cast<const float*>(P.heap[P.rank] + P.L.pts_all);
    float* __restrict__ kb = reinterpret_cast<float*>(P.heap[P.rank] + P.L.keybuf);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < local_n; i += gridDim.x * blockDim.x)
        kb[i] = all[(size_t)idx[i] * P.dim + axis];
}

// the key column of the current level: a carried coordinate column, or the gathered keys
__device__ __forceinline__ const float* key_column(const DistParams& P, int cur, int axis) {
    return P.ncol ? reinterpret_cast<const float*>(P.heap[P.rank] + P.L.col[cur][axis])
                  : reinterpret_cast<const float*>(P.heap[P.rank] + P.L.keybuf);
}

__global__ void k_dist_state_init(DistParams P, LevelGeo g) {
    if (threadIdx.x || blockIdx.x) return;
    DistCtl* c = DCTL(P, P.rank);
    c->prefix = 0ull;
    c->rank_wanted = g.n >> 1;
    c->cnt_global = g.n - 1;
    c->shift = 64;
    c->done = 0;
    c->my_less = 0;
    c->mine = 0;
    c->curL = c->curR = 0;
    c->ticket = 0;
    c->cL = 0;
    c->cand_cnt = 0;
    c->use_cand = 0;
    c->lowP = 0ull;
    c->highP = ~0ull;
    c->bticket = 0;
    if (g.level > 0) c->cnt = share_of(g.n, (unsigned)g.G, (unsigned)g.j);
}

// ---- K5, sample bracket across the group ----
// Every member posts a sorted systematic sample of its keys to all members (one-shot P2P stores); every member then
// ranks the union and picks the same two pivots around the wanted rank; ONE pass over the local keys counts what
// lies below the bracket and compacts what lies inside it (a few per cent); the digit passes run on those
// candidates, starting below the bits the pivots share.  Exact: the counts prove whether the rank is inside the
// bracket, otherwise the level falls back to digit passes over all keys.
__global__ void __launch_bounds__(512) k_dsample(DistParams P, LevelGeo g, int cur, int axis, unsigned long long epoch) {
    __shared__ unsigned long long sm[kDSample];
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    const unsigned local_n = ctl->cnt, tid = threadIdx.x;
    const unsigned free_n = (g.j == g.G - 1) ? local_n - 1u : local_n;
    const Cols src = cols_of(P, P.rank, cur);
    for (unsigned k = tid; k < (unsigned)kDSample; k += 512) {
        const unsigned pos = (unsigned)(((unsigned long long)k * free_n) / kDSample);
        sm[k] = composite(ordered_key(key_column(P, cur, axis)[pos]), (unsigned)src.idx[pos]);
    }
    __syncthreads();
    for (unsigned k = 2; k <= (unsigned)kDSample; k <<= 1)
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            const unsigned i = ((tid & ~(j - 1u)) << 1) | (tid & (j - 1u)), l = i | j;
            const bool asc = (i & k) == 0u;
            const unsigned long long a = sm[i], b = sm[l];
            if ((a > b) == asc) { sm[i] = b; sm[l] = a; }
            __syncthreads();
        }
    for (int m = 0; m < g.G; ++m) {
        unsigned long long* slot = reinterpret_cast<unsigned long long*>(P.heap[g.gb + m] + P.L.samp_slots) + (size_t)g.j * kDSample;
        for (unsigned k = tid; k < (unsigned)kDSample; k += 512) slot[k] = sm[k];
    }
    __threadfence_system();
    __syncthreads();
    if ((int)tid < g.G) st_release_sys(DFLAG(P, g.gb + tid, F_SAMP, g.j), epoch);
}

__device__ __forceinline__ unsigned lower_bound_u64(const unsigned long long* a, unsigned n, unsigned long long x) {
    unsigned lo = 0, hi = n;
    while (lo < hi) {
        const unsigned mid = (lo + hi) >> 1;
        if (a[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(1024) k_dpivots(DistParams P, LevelGeo g, unsigned long long epoch) {
    extern __shared__ __align__(16) unsigned char pv_smem[];
    unsigned long long* all = reinterpret_cast<unsigned long long*>(pv_smem);   // [G][kDSample] sorted lists
    __shared__ unsigned long long s_low, s_high;
    DistCtl* ctl = DCTL(P, P.rank);
    const unsigned tid = threadIdx.x;
    if (ctl->err) return;
    if ((int)tid < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, F_SAMP, tid), epoch)) atomicExch(&ctl->err, 600);
    }
    if (tid == 0) { s_low = 0ull; s_high = ~0ull; }
    __syncthreads();
    if (ctl->err) return;
    const unsigned S = (unsigned)g.G * kDSample;
    const unsigned long long* slots = reinterpret_cast<const unsigned long long*>(P.heap[P.rank] + P.L.samp_slots);
    for (unsigned e = tid; e < S; e += 1024) all[e] = __ldcg(&slots[e]);
    __syncthreads();
    const unsigned m = g.n >> 1, free_total = g.n - 1u;
    const long long R = (long long)(((unsigned long long)m * S) / free_total);
    const long long delta = g.G >= 8 ? 204 : g.G >= 4 ? 144 : g.G >= 2 ? 102 : 72;   // 72 * sqrt(G): 4.5 sigma of a sample rank
    const long long rlo = R - delta, rhi = R + delta;
    for (unsigned e = tid; e < S; e += 1024) {
        const unsigned l = e / kDSample, i = e % kDSample;
        const unsigned long long x = all[e];
        long long rank = i;                        // rank of x in the union = elements below it in every list
        for (int l2 = 0; l2 < g.G; ++l2)
            if ((unsigned)l2 != l) rank += lower_bound_u64(all + (size_t)l2 * kDSample, kDSample, x);
        if (rank <= rlo) atomicMax(&s_low, x);
        if (rank >= rhi) atomicMin(&s_high, x);
    }
    __syncthreads();
    if (tid == 0) { ctl->lowP = s_low; ctl->highP = s_high; }
}

template <int BLOCK, int IPT>
__global__ void __launch_bounds__(BLOCK) k_dbracket(DistParams P, LevelGeo g, int cur, int axis, unsigned long long epoch) {
    __shared__ unsigned s_cur, s_less;
    __shared__ int s_last;
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned long long lowP = ctl->lowP, highP = ctl->highP;
    const unsigned lowK = (unsigned)(lowP >> 32), highK = (unsigned)(highP >> 32);
    auto key_float = [](unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); };
    const float lowF = lowP == 0ull ? __int_as_float(0xff800000) : key_float(lowK);
    const float highF = highP == ~0ull ? __int_as_float(0x7f800000) : key_float(highK);
    const unsigned local_n = ctl->cnt;
    const unsigned free_n = (g.j == g.G - 1) ? local_n - 1u : local_n;
    const Cols src = cols_of(P, P.rank, cur);
    const float* __restrict__ keys = key_column(P, cur, axis);
    const int32_t* __restrict__ idx = src.idx;
    float* __restrict__ cand_key = reinterpret_cast<float*>(P.heap[P.rank] + P.L.ckey);
    int32_t* __restrict__ cand_idx = reinterpret_cast<int32_t*>(P.heap[P.rank] + P.L.cidx);
    unsigned* __restrict__ cand_cta = reinterpret_cast<unsigned*>(P.heap[P.rank] + P.L.ccta);
    const unsigned chunk = (free_n + gridDim.x - 1) / gridDim.x;
    const unsigned b = min(blockIdx.x * chunk, free_n), e = min(b + chunk, free_n);
    unsigned less = 0;
    if (tid == 0) { s_cur = 0; s_less = 0; }
    __syncthreads();
    for (unsigned tb = b; tb < e; tb += BLOCK * IPT) {
        float f[IPT];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const unsigned i = tb + k * BLOCK + tid;
            f[k] = i < e ? keys[i] : __int_as_float(0x7fc00000);
        }
        unsigned cmask = 0;
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const float v = f[k];
            if (v < lowF) ++less;
            else if (v <= highF) {
                bool cand = true;
                if (v == lowF || v == highF) {
                    const unsigned i = tb + k * BLOCK + tid;
                    const unsigned long long c = composite(ordered_key(v), (unsigned)idx[i]);
                    if (c < lowP) { ++less; cand = false; }
                    else cand = c <= highP;
                }
                cmask |= cand ? (1u << k) : 0u;
            }
        }
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
                cand_key[p] = f[k];
                cand_idx[p] = (int32_t)(tb + k * BLOCK + tid);
                ++p;
            }
    }
    for (int o = 16; o > 0; o >>= 1) less += __shfl_xor_sync(0xffffffffu, less, o);
    if (lane == 0 && less) atomicAdd(&s_less, less);
    __syncthreads();
    if (tid == 0) {
        cand_cta[blockIdx.x] = s_cur;
        if (s_cur) atomicAdd(&ctl->cand_cnt, s_cur);
        if (s_less) atomicAdd(&ctl->cL, s_less);
        __threadfence();
        s_last = atomicAdd(&ctl->bticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < g.G) {   // my (below, inside) counts -> every member, one-shot
        unsigned* slot = reinterpret_cast<unsigned*>(P.heap[g.gb + tid] + P.L.brk_slots) + g.j * 4;
        slot[0] = __ldcg(&ctl->cL);
        slot[1] = __ldcg(&ctl->cand_cnt);
        __threadfence_system();
        st_release_sys(DFLAG(P, g.gb + tid, F_BRK, g.j), epoch);
    }
}

__global__ void k_dbracket_wait(DistParams P, LevelGeo g, unsigned long long epoch) {
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    if ((int)threadIdx.x < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, F_BRK, threadIdx.x), epoch)) atomicExch(&ctl->err, 610);
    }
    __syncthreads();
    if (threadIdx.x || ctl->err) return;
    const unsigned* slots = reinterpret_cast<const unsigned*>(P.heap[P.rank] + P.L.brk_slots);
    unsigned cL = 0, cc = 0;
    for (int m = 0; m < g.G; ++m) { cL += __ldcg(&slots[m * 4]); cc += __ldcg(&slots[m * 4 + 1]); }
    const unsigned m = g.n >> 1;
    if (m >= cL && m - cL < cc) {   // the wanted rank is inside the bracket on the union of the members' counts
        ctl->use_cand = 1;
        ctl->rank_wanted = m - cL;
        ctl->cnt_global = cc;
        ctl->my_less = ctl->cL;
        const unsigned long long lowP = ctl->lowP, highP = ctl->highP;
        if (lowP != 0ull && highP != ~0ull && lowP != highP) {
            const int hi = 64 - __clzll((long long)(lowP ^ highP));
            ctl->shift = hi;
            ctl->prefix = hi >= 64 ? 0ull : ((lowP >> hi) << hi);
        }
    }
}

// K5 (post side): local histogram of digit d; the last CTA to finish publishes the rank's bins to every
// member of the group and raises their flag.
__device__ __forceinline__ void warp_hist_add_d(unsigned* sh, unsigned digit, bool valid) {
    unsigned peers = __match_any_sync(0xffffffffu, valid ? digit : 0xffffffffu);
    if (valid && (__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&sh[digit], __popc(peers));
}

template <int BLOCK, int UNROLL>
__global__ void __launch_bounds__(BLOCK) k_hist_dist(DistParams P, LevelGeo g, int cur, int axis, int d,
                                                     unsigned long long epoch, unsigned parity, unsigned nchunk) {
    __shared__ unsigned sh[kDigitBins];
    __shared__ int s_last;
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->done || ctl->err) return;
    const unsigned long long prefix = ctl->prefix;
    const unsigned local_n = ctl->cnt;
    const unsigned free_n = (g.j == g.G - 1) ? local_n - 1u : local_n;  // the range's pinned last element sits on the last member
    const Cols src = cols_of(P, P.rank, cur);
    const bool via = ctl->use_cand != 0;   // after a successful bracket: the compacted candidates of this CTA's chunk
    const float* __restrict__ keys = via ? reinterpret_cast<const float*>(P.heap[P.rank] + P.L.ckey) : key_column(P, cur, axis);
    const int32_t* __restrict__ cand_idx = reinterpret_cast<const int32_t*>(P.heap[P.rank] + P.L.cidx);
    const int32_t* __restrict__ idx = src.idx;
    // the keys are cut into `nchunk` chunks (the geometry k_dbracket compacted the candidates with); a CTA walks
    // several of them, so that the fixed cost of a pass (clearing and flushing 2048 bins) is paid by few CTAs
    const unsigned chunk = (free_n + nchunk - 1) / nchunk;
    for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) sh[t] = 0;
    __syncthreads();
    // bits [hi, 64) of the composite are resolved; this pass counts the (at most 11-bit) digit right below them
    const int hi = ctl->shift, dsh = select_digit_shift(hi);
    const unsigned mask = (1u << (hi - dsh)) - 1u;
    const unsigned wantkey = (unsigned)(prefix >> 32);
    for (unsigned ck = blockIdx.x; ck < nchunk; ck += gridDim.x) {
    const unsigned b = min(ck * chunk, free_n);
    const unsigned e = via ? b + reinterpret_cast<const unsigned*>(P.heap[P.rank] + P.L.ccta)[ck] : min(b + chunk, free_n);
    const unsigned iters = e > b ? (e - b + BLOCK * UNROLL - 1) / (BLOCK * UNROLL) : 0u;
    for (unsigned it = 0; it < iters; ++it) {
        unsigned i0 = b + it * (BLOCK * UNROLL) + threadIdx.x;
        float f[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) f[u] = (i0 + u * BLOCK) < e ? keys[i0 + u * BLOCK] : 0.f;
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const unsigned i = i0 + u * BLOCK;
            bool valid = i < e;
            const unsigned key = ordered_key(f[u]);
            unsigned digit;
            if (hi > 32) {
                if (hi < 64) valid = valid && ((key >> (hi - 32)) == (unsigned)(prefix >> hi));
                digit = (key >> (dsh - 32)) & mask;
            } else {
                valid = valid && key == wantkey;
                const unsigned id = valid ? (unsigned)(via ? idx[cand_idx[i]] : idx[i]) : 0u;
                if (hi < 32) valid = valid && ((id >> hi) == ((unsigned)(prefix & 0xffffffffull) >> hi));
                digit = (id >> dsh) & mask;
            }
            warp_hist_add_d(sh, digit, valid);
        }
    }
    }
    __syncthreads();
    unsigned* lh = reinterpret_cast<unsigned*>(P.heap[P.rank] + P.L.lhist);
    for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) {
        unsigned v = sh[t];
        if (v) atomicAdd(&lh[t], v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&ctl->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // The last CTA publishes the rank's bins.  NVLS: ONE multimem.red per non-empty bin adds it into the group's
    // accumulator on every GPU (the switch performs the reduction's fan-out), then one release-ordered reduction
    // bumps the accumulator's arrival counter.  Without a multicast mapping: one-shot slot stores to every member.
    unsigned* own = reinterpret_cast<unsigned*>(P.heap[P.rank] + P.L.own_hist) + (size_t)(d & 1) * kDigitBins;
    if (P.mc) {
        unsigned* acc = reinterpret_cast<unsigned*>(P.mc + racc_off(P.L, parity, g.level, P.rank / g.G, d));
        for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) {
            const unsigned v = __ldcg(&lh[t]);
            lh[t] = 0;
            own[t] = v;
            if (v) multimem_red_add(acc + t, v);
        }
        if (threadIdx.x == 0) ctl->ticket = 0;
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) multimem_red_add_release(acc + kDigitBins, 1u);
        return;
    }
    for (int t = threadIdx.x; t < kDigitBins; t += BLOCK) {
        const unsigned v = __ldcg(&lh[t]);
        lh[t] = 0;
        own[t] = v;
        for (int m = 0; m < g.G; ++m) {
            unsigned* slot = reinterpret_cast<unsigned*>(P.heap[g.gb + m] + P.L.hist_slots) +
                             ((size_t)(d & 1) * kMaxWorld + g.j) * kDigitBins;
            slot[t] = v;
        }
    }
    if (threadIdx.x == 0) ctl->ticket = 0;
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < g.G) st_release_sys(DFLAG(P, g.gb + threadIdx.x, F_HIST, g.j), epoch);
}

// K5 (wait side): the summed bins (NVLS accumulator, or the members' slots), the bin of the wanted rank, the prefix.
__global__ void __launch_bounds__(256) k_scan_dist(DistParams P, LevelGeo g, int d, unsigned long long epoch, unsigned parity) {
    __shared__ unsigned wsum[8];
    __shared__ int fbin;
    __shared__ unsigned fcum, fcnt, s_own_less, s_own_at;
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->done || ctl->err) return;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) { fbin = -1; s_own_less = 0; s_own_at = 0; }
    const unsigned* acc = reinterpret_cast<const unsigned*>(P.heap[P.rank] + racc_off(P.L, parity, g.level, P.rank / g.G, d));
    if (P.mc) {
        if (tid == 0 && !wait_count(acc + kDigitBins, (unsigned)g.G)) atomicExch(&ctl->err, 100 + d);
    } else if (tid < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, F_HIST, tid), epoch)) atomicExch(&ctl->err, 100 + d);
    }
    __syncthreads();
    if (ctl->err) return;
    const unsigned* slots = reinterpret_cast<const unsigned*>(P.heap[P.rank] + P.L.hist_slots) +
                            (size_t)(d & 1) * kMaxWorld * kDigitBins;
    const unsigned* ownh = reinterpret_cast<const unsigned*>(P.heap[P.rank] + P.L.own_hist) + (size_t)(d & 1) * kDigitBins;
    unsigned v[8], own[8], sum = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        unsigned a = 0;
        if (P.mc) a = __ldcg(&acc[tid * 8 + k]);
        else for (int m = 0; m < g.G; ++m) a += __ldcg(&slots[(size_t)m * kDigitBins + tid * 8 + k]);
        v[k] = a;
        own[k] = __ldcg(&ownh[tid * 8 + k]);
        sum += a;
    }
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
    const unsigned r = ctl->rank_wanted;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (r >= c && r < c + v[k]) { fbin = tid * 8 + k; fcum = c; fcnt = v[k]; }
        c += v[k];
    }
    __syncthreads();
    const int fb = fbin;
    unsigned less = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (tid * 8 + k < fb) less += own[k];
        if (tid * 8 + k == fb) s_own_at = own[k];
    }
    for (int o = 16; o > 0; o >>= 1) less += __shfl_xor_sync(0xffffffffu, less, o);
    if (lane == 0 && less) atomicAdd(&s_own_less, less);
    __syncthreads();
    if (tid == 0) {
        if (fb < 0) {
            atomicExch(&ctl->err, 1 + d);
            ctl->done = 1;
        } else {
            const int dsh = select_digit_shift(ctl->shift);
            ctl->prefix |= (unsigned long long)(unsigned)fb << dsh;
            ctl->rank_wanted = r - fcum;
            ctl->cnt_global = fcnt;
            ctl->shift = dsh;
            ctl->my_less += s_own_less;
            // one element left (inside the bracket, if there is one): the partition tells it from the elements
            // outside the bracket that share its resolved bits by comparing with the pivots first
            if (fcnt == 1u || dsh == 0) { ctl->done = 1; ctl->mine = s_own_at; }
        }
    }
}

// per-rank left/right counts -> every member (post), then exclusive scan over members (wait)
__global__ void k_counts_post(DistParams P, LevelGeo g, unsigned long long epoch) {
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    const unsigned local_n = ctl->cnt;
    const unsigned free_n = (g.j == g.G - 1) ? local_n - 1u : local_n;
    const unsigned lc = ctl->my_less, rc = free_n - lc - ctl->mine;
    if ((int)threadIdx.x < g.G) {
        unsigned* slot = reinterpret_cast<unsigned*>(P.heap[g.gb + threadIdx.x] + P.L.cnt_slots) + g.j * 4;
        slot[0] = lc;
        slot[1] = rc;
        __threadfence_system();
        st_release_sys(DFLAG(P, g.gb + threadIdx.x, F_CNT, g.j), epoch);
    }
}

__global__ void k_counts_wait(DistParams P, LevelGeo g, unsigned long long epoch) {
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    if ((int)threadIdx.x < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, F_CNT, threadIdx.x), epoch)) atomicExch(&ctl->err, 200);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned* slots = reinterpret_cast<const unsigned*>(P.heap[P.rank] + P.L.cnt_slots);
        // Order of the members' runs inside the child ranges: (0, G/2, 1, G/2+1, ...) instead of rank order.  Owner
        // k of the left child then receives the left parts of members k and k+G/2, owner G/2+k of the right child
        // the right parts of the same two: every member keeps one of its two halves at home, and only half of the
        // points cross NVLink at every level (rank order moved 7/8 of them at G = 8, 3/4 at G = 4).  The order
        // inside a child range is free (the next level selects under a total order).
        const int Gh = g.G >> 1;
        auto order_key = [Gh](int m) { return Gh > 0 ? (m % Gh) * 2 + m / Gh : m; };
        const int mykey = order_key(g.j);
        unsigned lo = 0, ro = 0, lt = 0, rt = 0;
        for (int m = 0; m < g.G; ++m) {
            const unsigned lc = __ldcg(&slots[m * 4]), rc = __ldcg(&slots[m * 4 + 1]);
            if (order_key(m) < mykey) { lo += lc; ro += rc; }
            lt += lc;
            rt += rc;
        }
        ctl->Loff = lo;
        ctl->Roff = ro;
        if (lt != (g.n >> 1) || rt != g.n - (g.n >> 1) - 2u) atomicExch(&ctl->err, 201);
    }
}

// K6: partition with P2P stores into the owners' next-level buffers
template <int NCOL, int BLOCK, int IPT>
__global__ void __launch_bounds__(BLOCK) k_partition_dist(DistParams P, LevelGeo g, int cur, int axis, bool with_tp) {
    extern __shared__ __align__(16) unsigned char pd_smem[];
    float* stage = reinterpret_cast<float*>(pd_smem);                                  // [NCOL][BLOCK * IPT]
    int* stage_idx = reinterpret_cast<int*>(stage + (size_t)NCOL * BLOCK * IPT);       // [BLOCK * IPT]
    __shared__ unsigned s_cnt[2], s_base[2];
    __shared__ unsigned long long s_wcomp[BLOCK / 32];
    __shared__ unsigned s_wpos[BLOCK / 32];
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;   // a peer wait failed earlier: the counts below cannot be trusted, store nothing anywhere
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned long long prefix = ctl->prefix;
    const int shift = ctl->shift;
    const unsigned long long want = shift >= 64 ? 0ull : (prefix >> shift);
    const bool bracketed = ctl->use_cand != 0;
    const unsigned long long lowP = bracketed ? ctl->lowP : 0ull, highP = bracketed ? ctl->highP : ~0ull;
    const unsigned local_n = ctl->cnt;
    const bool last_member = g.j == g.G - 1;
    const unsigned free_n = last_member ? local_n - 1u : local_n;
    const unsigned m = g.n >> 1, nr = g.n - m - 1u;      // sizes of the child ranges
    const unsigned Gh = (unsigned)g.G >> 1;              // members per child range
    const unsigned Loff = ctl->Loff, Roff = ctl->Roff;
    const Cols src = cols_of(P, P.rank, cur);
    const unsigned chunk = (free_n + gridDim.x - 1) / gridDim.x;
    const unsigned b = min(blockIdx.x * chunk, free_n), e = min(b + chunk, free_n);
    unsigned long long best = 0ull;
    unsigned bestpos = 0, bestvalid = 0;
    bool has_node = false;
    int node_id = 0;
    constexpr int NC1 = NCOL > 0 ? NCOL : 1;   // NCOL == 0 (gather mode): only the index column moves
    const float* __restrict__ keycol = key_column(P, cur, axis);
    float node_val[NC1];

    if (last_member && blockIdx.x == gridDim.x - 1 && tid == 0) {
        // the range's pinned element: last element of the right child range, i.e. of the group's last member
        const unsigned p = local_n - 1u;
        unsigned mem, pos;
        owner_of(nr - 1u, nr, Gh, mem, pos);
        const Cols d = cols_of(P, g.gb + Gh + mem, cur ^ 1);
#pragma unroll
        for (int c = 0; c < NCOL; ++c) d.c[c][pos] = src.c[c][p];
        d.idx[pos] = src.idx[p];
    }
    for (unsigned tb = b; tb < e; tb += BLOCK * IPT) {
        if (tid < 2) s_cnt[tid] = 0;
        __syncthreads();
        float val[IPT][NC1];
        float kf0[IPT];
        int id[IPT];
        unsigned rk[IPT], key[IPT];
        signed char side[IPT];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            unsigned i = tb + k * BLOCK + tid;
            bool valid = i < e;
            id[k] = valid ? src.idx[i] : 0;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) val[k][c] = valid ? src.c[c][i] : 0.f;
            kf0[k] = (NCOL == 0 && valid) ? keycol[i] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            unsigned i = tb + k * BLOCK + tid;
            float kf = NCOL > 0 ? val[k][0] : kf0[k];
#pragma unroll
            for (int c = 1; c < NCOL; ++c) kf = (c == axis) ? val[k][c] : kf;
            key[k] = ordered_key(kf);
            side[k] = 2;
            if (i < e) {
                const unsigned long long cf = composite(key[k], (unsigned)id[k]);
                const unsigned long long c = cf >> shift;
                // outside the bracket the pivots decide (such elements may share the node's resolved high bits)
                side[k] = cf < lowP ? -1 : (cf > highP ? 1 : (c < want ? -1 : (c > want ? 1 : 0)));
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
            s_base[tid] = c ? atomicAdd(tid == 0 ? &ctl->curL : &ctl->curR, c) : 0u;
        }
        __syncthreads();
        const unsigned baseL = Loff + s_base[0], baseR = Roff + s_base[1];
        const unsigned nl = s_cnt[0], nrt = s_cnt[1];
        // stage the tile in shared memory in output order (left run, then right run) ...
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
            const bool left = side[k] == -1;
            const unsigned si = left ? rk[k] : nl + rk[k];
#pragma unroll
            for (int c = 0; c < NCOL; ++c) stage[c * (BLOCK * IPT) + si] = val[k][c];
            stage_idx[si] = id[k];
            if (left) {
                unsigned long long c = composite(key[k], (unsigned)id[k]);
                if (!bestvalid || c > best) { best = c; bestpos = baseL + rk[k]; bestvalid = 1; }
            }
        }
        __syncthreads();
        // ... and push it out as contiguous runs with 16-byte stores (a run = the part of the left or right
        // output that lands on one owner GPU; there are at most a few per tile).  Every warp takes whole columns
        // (warp w: columns w, w + warps, ...; the index column counts as one), so all columns stream at once and
        // a warp emits 512 contiguous bytes per store instruction, which is what the NVLink egress path wants.
        for (int side_i = 0; side_i < 2; ++side_i) {
            const unsigned cnt = side_i ? nrt : nl, s_first = side_i ? nl : 0u;
            const unsigned n_c = side_i ? nr : m;
            unsigned done = 0;
            while (done < cnt) {
                const unsigned o = (side_i ? baseR : baseL) + done;
                unsigned mem, pos;
                owner_of(o, n_c, Gh, mem, pos);
                const unsigned room = share_of(n_c, Gh, mem) - pos;   // slots left on this owner
                const unsigned k = min(cnt - done, room);
                unsigned char* dheap = P.heap[g.gb + (side_i ? Gh + mem : mem)];
                const unsigned s0 = s_first + done;
                const unsigned head = min(k, (4u - (pos & 3u)) & 3u);  // elements before the first 16-byte boundary
                const unsigned body = (k - head) >> 2;                 // whole 16-byte chunks
                const unsigned tail0 = head + (body << 2);
                // the run [pos, pos + k) must lie inside the owner's share of the child range and inside its columns
                if (!KD_INB(mem < Gh && pos + k <= share_of(n_c, Gh, mem) && (size_t)pos + k <= (size_t)P.L.cap, 951)) { done += k; continue; }
                for (int c = tid >> 5; c <= NCOL; c += BLOCK / 32) {
                    const float* __restrict__ sc = c < NCOL ? stage + (size_t)c * (BLOCK * IPT) + s0
                                                           : reinterpret_cast<const float*>(stage_idx) + s0;
                    float* __restrict__ dc = reinterpret_cast<float*>(dheap + (c < NCOL ? P.L.col[cur ^ 1][c] : P.L.idx[cur ^ 1])) + pos;
                    if ((unsigned)lane < head) dc[lane] = sc[lane];
                    for (unsigned q = lane; q < body; q += 32) {
                        const unsigned e0 = head + (q << 2);
                        *reinterpret_cast<float4*>(dc + e0) = make_float4(sc[e0], sc[e0 + 1], sc[e0 + 2], sc[e0 + 3]);
                    }
                    if (tail0 + lane < k) dc[tail0 + lane] = sc[tail0 + lane];
                }
                done += k;
            }
        }
    }
    if (has_node) {
        // the node of this range: slot lo + n/2 of every GPU's tree arrays
        // (with an NVLS multicast mapping one store is replicated by the switch: multimem.st lowers to a plain
        // STG on the multicast address, so the same code path serves both)
        const int nrep = P.mc ? 1 : P.world;
        for (int r = 0; r < nrep; ++r) {
            unsigned char* base = P.mc ? P.mc : P.heap[r];
            reinterpret_cast<int32_t*>(base + P.L.perm)[g.lo + m] = node_id;
            if (with_tp) {
                float* tp = reinterpret_cast<float*>(base + P.L.tp);
#pragma unroll
                for (int c = 0; c < NCOL; ++c) tp[(size_t)(g.lo + m) * P.dim + c] = node_val[c];
            }
        }
    }
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
        reinterpret_cast<MaxRec*>(P.heap[P.rank] + P.L.rec)[blockIdx.x] = r;
    }
    __threadfence_system();  // my peer stores are ordered before anything this GPU signals later
}

// plain group barrier, split in two kernels (see the header comment)
__global__ void k_flag_post(DistParams P, LevelGeo g, int kind, unsigned long long epoch) {
    __threadfence_system();
    if ((int)threadIdx.x < g.G) st_release_sys(DFLAG(P, g.gb + threadIdx.x, kind, g.j), epoch);
}
__global__ void k_flag_wait(DistParams P, LevelGeo g, int kind, unsigned long long epoch) {
    if ((int)threadIdx.x < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, kind, threadIdx.x), epoch)) atomicExch(&DCTL(P, P.rank)->err, 300 + kind);
    }
}

// the left child's largest element -> its last slot.  Post: my best candidate to every member.
__global__ void __launch_bounds__(256) k_max_post(DistParams P, LevelGeo g, int nrec, unsigned long long epoch) {
    __shared__ unsigned long long s_c[8];
    __shared__ unsigned s_p[8], s_v[8];
    const MaxRec* rec = reinterpret_cast<const MaxRec*>(P.heap[P.rank] + P.L.rec);
    if (DCTL(P, P.rank)->err) return;
    unsigned long long best = 0ull;
    unsigned bestpos = 0, bestvalid = 0;
    for (int t = threadIdx.x; t < nrec; t += blockDim.x) {
        MaxRec r = rec[t];
        if (r.valid && (!bestvalid || r.comp > best)) { best = r.comp; bestpos = r.pos; bestvalid = 1; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long oc = __shfl_xor_sync(0xffffffffu, best, o);
        unsigned op = __shfl_xor_sync(0xffffffffu, bestpos, o);
        unsigned ov = __shfl_xor_sync(0xffffffffu, bestvalid, o);
        if (ov && (!bestvalid || oc > best)) { best = oc; bestpos = op; bestvalid = 1; }
    }
    if ((threadIdx.x & 31) == 0) { s_c[threadIdx.x >> 5] = best; s_p[threadIdx.x >> 5] = bestpos; s_v[threadIdx.x >> 5] = bestvalid; }
    __syncthreads();
    if ((int)threadIdx.x < g.G) {
        MaxSlot ms{0ull, 0u, 0u};
        for (int w = 0; w < 8; ++w)
            if (s_v[w] && (!ms.valid || s_c[w] > ms.comp)) { ms.comp = s_c[w]; ms.off = s_p[w]; ms.valid = 1; }
        MaxSlot* slot = reinterpret_cast<MaxSlot*>(P.heap[g.gb + threadIdx.x] + P.L.max_slots) + g.j;
        *slot = ms;
        __threadfence_system();
        st_release_sys(DFLAG(P, g.gb + threadIdx.x, F_MAX, g.j), epoch);
    }
}

// Wait: the owner of the left child's last slot exchanges it with the winner.
__global__ void k_max_wait_swap(DistParams P, LevelGeo g, int cur, unsigned long long epoch) {
    DistCtl* ctl = DCTL(P, P.rank);
    if (ctl->err) return;
    if ((int)threadIdx.x < g.G) {
        if (!wait_flag(DFLAG(P, P.rank, F_MAX, threadIdx.x), epoch)) atomicExch(&ctl->err, 400);
    }
    __syncthreads();
    if (ctl->err) return;
    const unsigned m = g.n >> 1, Gh = (unsigned)g.G >> 1;
    unsigned lastmem, lastpos;
    owner_of(m - 1u, m, Gh, lastmem, lastpos);
    if ((unsigned)g.j != lastmem) return;  // only the owner of slot m-1 of the left child acts
    const MaxSlot* slots = reinterpret_cast<const MaxSlot*>(P.heap[P.rank] + P.L.max_slots);
    MaxSlot bestm{0ull, 0u, 0u};
    for (int k = 0; k < g.G; ++k) {
        MaxSlot s = slots[k];
        if (s.valid && (!bestm.valid || s.comp > bestm.comp)) bestm = s;
    }
    if (!bestm.valid) { if (threadIdx.x == 0) atomicExch(&ctl->err, 401); return; }
    if (bestm.off == m - 1u) return;
    unsigned wm, wp;
    owner_of(bestm.off, m, Gh, wm, wp);
    Cols a = cols_of(P, g.gb + wm, cur ^ 1), b = cols_of(P, P.rank, cur ^ 1);
    const int lane = threadIdx.x;
    if (lane < P.ncol) {
        float x = a.c[lane][wp], y = b.c[lane][lastpos];
        a.c[lane][wp] = y;
        b.c[lane][lastpos] = x;
    } else if (lane == P.ncol) {
        int32_t x = a.idx[wp], y = b.idx[lastpos];
        a.idx[wp] = y;
        b.idx[lastpos] = x;
    }
    __threadfence_system();
}

__global__ void k_dist_read_err(DistParams P, int* out) { *out = DCTL(P, P.rank)->err; }

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct HostCtx {
    DistParams P;
    bool with_tp;
};

static LevelGeo make_geo(const DistParams& P, int level) {
    LevelGeo g;
    g.level = level;
    g.G = P.world >> level;
    g.gb = (P.rank / g.G) * g.G;
    g.j = P.rank % g.G;
    host_seg_range(P.N, level, (unsigned)(P.rank / g.G), g.lo, g.n);
    return g;
}

#define KD_DLAUNCH(what)                                                                      \
    do {                                                                                      \
        count_launch();                                                                       \
        cudaError_t e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) return fail(std::string(what) + ": " + cudaGetErrorString(e_)); \
    } while (0)

enum Phase { PH_STATE = 0, PH_HIST, PH_SCAN, PH_CNT_POST, PH_CNT_WAIT, PH_PART, PH_PART_POST, PH_PART_WAIT, PH_MAX_POST,
             PH_MAX_WAIT, PH_FIX_POST, PH_FIX_WAIT, PH_SAMPLE_POST, PH_PIVOT_WAIT, PH_BRACKET_POST, PH_BRACKET_WAIT,
             PH_ROWS_PUSH, PH_ROWS_POST, PH_ROWS_WAIT, PH_KEYS, kNumPhases };

// One distributed level in launch order: (phase, pass) pairs.  Drivers of virtual ranks launch every entry for all
// ranks before the next entry (posts before waits); kd_dist_run walks the same list for its own rank.
struct Step { int phase, pass; };
static int level_schedule(Step* out) {
    int n = 0;
    out[n++] = {PH_STATE, 0};
    out[n++] = {PH_ROWS_PUSH, 4};   // gather mode, level 0 only: every rank's rows to every rank (no-ops otherwise)
    out[n++] = {PH_ROWS_POST, 4};
    out[n++] = {PH_ROWS_WAIT, 4};
    out[n++] = {PH_KEYS, 4};        // gather mode: the level's split-axis keys
    out[n++] = {PH_SAMPLE_POST, 6};
    out[n++] = {PH_PIVOT_WAIT, 6};
    out[n++] = {PH_BRACKET_POST, 5};
    out[n++] = {PH_BRACKET_WAIT, 5};
    for (int d = 0; d < kNumDigits; ++d) { out[n++] = {PH_HIST, d}; out[n++] = {PH_SCAN, d}; }
    for (int ph = PH_CNT_POST; ph <= PH_FIX_WAIT; ++ph) out[n++] = {ph, 7};
    return n;
}

static int dist_grid(unsigned local_n, int min_chunk) {
    int64_t cap = std::max<int64_t>(1, ((int64_t)local_n + min_chunk - 1) / min_chunk);
    return (int)std::min<int64_t>(kTargetCtas, cap);
}

// K7, bulk part: a finished slice of this rank's tree arrays -> every other GPU's tree arrays.  With an NVLS
// multicast mapping ONE 16-byte store per chunk is fanned out by the switch (multimem.st lowers to a plain STG on
// the multicast address); otherwise one store per peer.  Runs on a side stream while the next slice is built.
struct RepDst { int32_t* perm[kMaxWorld]; float* tp[kMaxWorld]; int n; };

__device__ __forceinline__ void rep_copy(const unsigned* __restrict__ src, unsigned* const* dst, int nd, size_t words,
                                         size_t tid, size_t nth) {
    // words [0, words) of src -> the same offsets of every dst; src and dst share their alignment phase
    const size_t head = min(words, (size_t)((4u - ((reinterpret_cast<uintptr_t>(src) >> 2) & 3u)) & 3u));
    const size_t body = (words - head) >> 2;
    for (size_t i = tid; i < head; i += nth)
        for (int d = 0; d < nd; ++d) dst[d][i] = src[i];
    const uint4* __restrict__ s4 = reinterpret_cast<const uint4*>(src + head);
    for (size_t i = tid; i < body; i += nth) {
        const uint4 v = __ldcs(&s4[i]);
        for (int d = 0; d < nd; ++d) reinterpret_cast<uint4*>(dst[d] + head)[i] = v;
    }
    for (size_t i = head + (body << 2) + tid; i < words; i += nth)
        for (int d = 0; d < nd; ++d) dst[d][i] = src[i];
}

__global__ void __launch_bounds__(512) k_replicate(const int32_t* __restrict__ perm, const float* __restrict__ tp,
                                                   RepDst dst, size_t first, size_t count, int dim) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    unsigned* d[kMaxWorld];
    for (int r = 0; r < dst.n; ++r) d[r] = reinterpret_cast<unsigned*>(dst.perm[r] + first);
    rep_copy(reinterpret_cast<const unsigned*>(perm + first), d, dst.n, count, tid, nth);
    if (tp) {
        for (int r = 0; r < dst.n; ++r) d[r] = reinterpret_cast<unsigned*>(dst.tp[r] + first * dim);
        rep_copy(reinterpret_cast<const unsigned*>(tp + first * dim), d, dst.n, count * dim, tid, nth);
    }
}

// gather mode: a run of floats of this rank's point array -> the same offsets of the other ranks' copies
__global__ void __launch_bounds__(512) k_rows_push(const float* __restrict__ all, RepDst dst, size_t first, size_t count) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    unsigned* d[kMaxWorld];
    for (int r = 0; r < dst.n; ++r) d[r] = reinterpret_cast<unsigned*>(dst.tp[r] + first);
    rep_copy(reinterpret_cast<const unsigned*>(all + first), d, dst.n, count, tid, nth);
}

// CTA width of k_partition_dist (KD_DIST_PART_BLOCK = 128 | 512).  128-thread CTAs with 512-point tiles: many
// independent tiles per SM hide each tile's load -> rank -> claim (a global atomic round trip) -> stage -> push chain
// (measured on 2 B200, 2^25 points per GPU: 0.68 ms with 128 threads, 0.98 ms with 512; unstaged stores straight
// from the registers, two short runs per warp instruction, were slower than either over the link: 0.95 ms)
static int part_block() {
    static const int v = [] { const char* e = std::getenv("KD_DIST_PART_BLOCK"); int x = e ? std::atoi(e) : 128; return x == 512 ? 512 : 128; }();
    return v;
}

// opt-in shared-memory sizes of the distributed kernels, once per device, every result checked
static int dist_func_attributes_once() {
    static std::mutex mu;
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(std::string("kd_dist: cudaGetDevice: ") + cudaGetErrorString(e));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && done[dev]) return 0;
    e = cudaFuncSetAttribute(k_dpivots, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxWorld * kDSample * 8);
#define KD_PD_ATTR(NC) if (e == cudaSuccess) e = cudaFuncSetAttribute(k_partition_dist<NC, 512, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, ((NC) + 1) * 2048 * 4)
    KD_PD_ATTR(0); KD_PD_ATTR(1); KD_PD_ATTR(2); KD_PD_ATTR(3); KD_PD_ATTR(4); KD_PD_ATTR(5); KD_PD_ATTR(6); KD_PD_ATTR(7); KD_PD_ATTR(8);
#undef KD_PD_ATTR
    if (e != cudaSuccess) return fail(std::string("kd_dist: cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    if (dev >= 0 && dev < 64) done[dev] = true;
    return 0;
}

static int run_phase(const HostCtx& H, int level, int phase, int pass, unsigned long long serial, cudaStream_t st) {
    if (int rc = dist_func_attributes_once()) return rc;
    const DistParams& P = H.P;
    const LevelGeo g = make_geo(P, level);
    const int cur = level & 1;
    const int axis = level % P.dim;
    const unsigned long long epoch = serial * 64ull + (unsigned long long)level * 8ull + (unsigned long long)pass + 1ull;
    // upper bound of my local count at this level (the device reads the exact value from its control block)
    const unsigned local_cap = (unsigned)H.P.L.cap;
    const unsigned parity = (unsigned)(serial & 1ull);
    switch (phase) {
        case PH_STATE:
            if (level == 0 && P.mc) {
                // NVLS accumulators of the NEXT build (other parity): the build that used them last is complete on
                // every rank (its done-barrier precedes this build), the next build starts after this one's barrier
                if (cudaMemsetAsync(P.heap[P.rank] + racc_off(P.L, parity ^ 1u, 0, 0, 0), 0,
                                    (size_t)7 * kRedPasses * kRedStride * 4, st) != cudaSuccess)
                    return fail("kd_dist_phase: cudaMemsetAsync failed");
            }
            k_dist_state_init<<<1, 32, 0, st>>>(P, g);
            KD_DLAUNCH("k_dist_state_init");
            break;
        case PH_ROWS_PUSH: {
            if (P.ncol != 0 || level != 0 || P.world == 1) break;
            // my slice of the point array (already in my own copy) -> the same rows of every other rank's copy
            unsigned first = (unsigned)((P.N / P.world) * P.rank);
            unsigned count = P.N / P.world + (P.rank == P.world - 1 ? P.N % P.world : 0u);   // kdtree_mpi.cpp:208-216
            RepDst rep{};
            if (P.mc) { rep.n = 1; rep.tp[0] = (float*)(P.mc + P.L.pts_all); }
            else for (int r = 0; r < P.world; ++r)
                if (r != P.rank) rep.tp[rep.n++] = (float*)(P.heap[r] + P.L.pts_all);
            k_rows_push<<<148, 512, 0, st>>>((const float*)(P.heap[P.rank] + P.L.pts_all), rep, (size_t)first * P.dim,
                                             (size_t)count * P.dim);
            KD_DLAUNCH("k_rows_push");
            break;
        }
        case PH_ROWS_POST:
            if (P.ncol != 0 || level != 0 || P.world == 1) break;
            k_flag_post<<<1, 32, 0, st>>>(P, g, F_ROWS, epoch);
            KD_DLAUNCH("k_flag_post");
            break;
        case PH_ROWS_WAIT:
            if (P.ncol != 0 || level != 0 || P.world == 1) break;
            k_flag_wait<<<1, 32, 0, st>>>(P, g, F_ROWS, epoch);
            KD_DLAUNCH("k_flag_wait");
            break;
        case PH_KEYS:
            if (P.ncol != 0) break;
            k_dist_gather_keys<<<dist_grid(local_cap, 2048), 256, 0, st>>>(P, g, cur, axis);
            KD_DLAUNCH("k_dist_gather_keys");
            break;
        case PH_SAMPLE_POST:
            k_dsample<<<1, 512, 0, st>>>(P, g, cur, axis, epoch);
            KD_DLAUNCH("k_dsample");
            break;
        case PH_PIVOT_WAIT:
            k_dpivots<<<1, 1024, (size_t)g.G * kDSample * 8, st>>>(P, g, epoch);
            KD_DLAUNCH("k_dpivots");
            break;
        case PH_BRACKET_POST:   // same grid as the digit passes: they read the candidates chunk by chunk
            k_dbracket<512, 8><<<dist_grid(local_cap, 8192), 512, 0, st>>>(P, g, cur, axis, epoch);
            KD_DLAUNCH("k_dbracket");
            break;
        case PH_BRACKET_WAIT:
            k_dbracket_wait<<<1, 32, 0, st>>>(P, g, epoch);
            KD_DLAUNCH("k_dbracket_wait");
            break;
        case PH_HIST:
            {   // full-key passes need every SM; passes over the candidates are tiny
                const int nchunk = dist_grid(local_cap, 8192);
                k_hist_dist<512, 4><<<std::min(nchunk, 296), 512, 0, st>>>(P, g, cur, axis, pass, epoch, parity, (unsigned)nchunk);
            }
            KD_DLAUNCH("k_hist_dist");
            break;
        case PH_SCAN:
            k_scan_dist<<<1, 256, 0, st>>>(P, g, pass, epoch, parity);
            KD_DLAUNCH("k_scan_dist");
            break;
        case PH_CNT_POST:
            k_counts_post<<<1, 32, 0, st>>>(P, g, epoch);
            KD_DLAUNCH("k_counts_post");
            break;
        case PH_CNT_WAIT:
            k_counts_wait<<<1, 32, 0, st>>>(P, g, epoch);
            KD_DLAUNCH("k_counts_wait");
            break;
        case PH_PART: {
            const int grid = dist_grid(local_cap, 2048);
            const int pblock = part_block();
#define KD_PD(NC)                                                                                              \
    do {                                                                                                       \
        if (pblock == 128)                                                                                     \
            k_partition_dist<NC, 128, 4><<<grid, 128, (size_t)((NC) + 1) * 512 * 4, st>>>(P, g, cur, axis, H.with_tp);  \
        else                                                                                                   \
            k_partition_dist<NC, 512, 4><<<grid, 512, (size_t)((NC) + 1) * 2048 * 4, st>>>(P, g, cur, axis, H.with_tp); \
    } while (0)
            switch (P.ncol) {
                case 0: KD_PD(0); break;
                case 1: KD_PD(1); break;
                case 2: KD_PD(2); break;
                case 3: KD_PD(3); break;
                case 4: KD_PD(4); break;
                case 5: KD_PD(5); break;
                case 6: KD_PD(6); break;
                case 7: KD_PD(7); break;
                default: KD_PD(8); break;
            }
            KD_DLAUNCH("k_partition_dist");
            break;
        }
        case PH_PART_POST:
            k_flag_post<<<1, 32, 0, st>>>(P, g, F_PART, epoch);
            KD_DLAUNCH("k_flag_post");
            break;
        case PH_PART_WAIT:
            k_flag_wait<<<1, 32, 0, st>>>(P, g, F_PART, epoch);
            KD_DLAUNCH("k_flag_wait");
            break;
        case PH_MAX_POST:
            k_max_post<<<1, 256, 0, st>>>(P, g, dist_grid(local_cap, 2048), epoch);
            KD_DLAUNCH("k_max_post");
            break;
        case PH_MAX_WAIT:
            k_max_wait_swap<<<1, 32, 0, st>>>(P, g, cur, epoch);
            KD_DLAUNCH("k_max_wait_swap");
            break;
        case PH_FIX_POST:
            k_flag_post<<<1, 32, 0, st>>>(P, g, F_FIX, epoch);
            KD_DLAUNCH("k_flag_post");
            break;
        case PH_FIX_WAIT:
            k_flag_wait<<<1, 32, 0, st>>>(P, g, F_FIX, epoch);
            KD_DLAUNCH("k_flag_wait");
            break;
        default:
            return fail("kd_dist_phase: unknown phase");
    }
    return 0;
}

static int dist_levels(int world) {
    int l = 0;
    while ((1 << l) < world) ++l;
    return l;
}

// streams + events of the sliced local phase, one set per device: slice c is built and replicated on stream c,
// earlier slices have the higher priority, so they finish first and their replication runs next to the others' work
constexpr int kMaxSlices = 8;
struct SideStream { cudaStream_t s[kMaxSlices]; cudaEvent_t fork, join[kMaxSlices]; bool ok; };
static int side_stream(SideStream** out) {
    static std::mutex mu;
    static SideStream tab[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return fail("kd_dist: cudaGetDevice failed");
    std::lock_guard<std::mutex> lock(mu);
    SideStream& S = tab[dev];
    if (!S.ok) {
        int least = 0, greatest = 0;
        cudaDeviceGetStreamPriorityRange(&least, &greatest);   // numerically lower = higher priority
        for (int c = 0; c < kMaxSlices; ++c) {
            const int prio = std::max(greatest, std::min(least, greatest + c));
            if (cudaStreamCreateWithPriority(&S.s[c], cudaStreamNonBlocking, prio) != cudaSuccess)
                return fail("kd_dist: cudaStreamCreateWithPriority failed");
            if (cudaEventCreateWithFlags(&S.join[c], cudaEventDisableTiming) != cudaSuccess) return fail("kd_dist: cudaEventCreate failed");
        }
        if (cudaEventCreateWithFlags(&S.fork, cudaEventDisableTiming) != cudaSuccess) return fail("kd_dist: cudaEventCreate failed");
        S.ok = true;
    }
    *out = &S;
    return 0;
}

static int replicate_grid() {   // CTAs of k_replicate (KD_DIST_REP_GRID)
    static const int v = [] { const char* e = std::getenv("KD_DIST_REP_GRID"); int x = e ? std::atoi(e) : 64; return x < 1 ? 1 : x; }();
    return v;
}
constexpr unsigned kChunkMinLocal = 1u << 21;  // the local phase is finished in slices when a GPU holds at least this many points
// 1 << chunk_level() slices (KD_DIST_CHUNK_LEVEL; default 0 = the whole slice at once, stores replicated by the kernels
// themselves).  Measured on 8 B200 at 2^28 x 3: 19.3 ms unsliced, 20.2 ms with 4 slices on prioritized streams, 23.1 ms
// with 4 slices one after the other.  The replication does overlap (each slice's k_replicate runs at 716 GB/s next to
// the later slices' kernels), but every slice repeats the latency-bound top levels (~17 dependent launches per level)
// and the slices do not fill each other's gaps: four 2^23 builds take 11.3 ms where one 2^25 build takes 8.9 ms.
static int chunk_level() {
    static const int v = [] { const char* e = std::getenv("KD_DIST_CHUNK_LEVEL"); int x = e ? std::atoi(e) : 0; return x < 0 ? 0 : (x > 3 ? 3 : x); }();
    return v;
}

static int local_build(const HostCtx& H, cudaStream_t st) {
    const DistParams& P = H.P;
    const int L = dist_levels(P.world);
    unsigned lo, n;
    host_seg_range(P.N, L, (unsigned)P.rank, lo, n);
    unsigned char* base = P.heap[P.rank];
    Cols buf[2];
    for (int b = 0; b < 2; ++b) {
        for (int c = 0; c < kMaxCarry; ++c) buf[b].c[c] = c < P.ncol ? (float*)(base + P.L.col[b][c]) : nullptr;
        buf[b].idx = (int32_t*)(base + P.L.idx[b]);
    }
    BuildScratch sc;
    const float* all_pts = P.ncol ? nullptr : (const float*)(base + P.L.pts_all);   // gather mode: keys come from here
    sc.keybuf = P.ncol ? nullptr : (float*)(base + P.L.keybuf);
    sc.st = (SegState*)(base + P.L.st);
    sc.ghist = (unsigned*)(base + P.L.ghist);
    sc.rec = (MaxRec*)(base + P.L.lrec);
    sc.err = (int*)(base + P.L.lerr);
    sc.cand_key = (float*)(base + P.L.ckey);
    sc.cand_idx = (int32_t*)(base + P.L.cidx);
    sc.cand_cta = (unsigned*)(base + P.L.ccta);
    if (cudaMemsetAsync(sc.ghist, 0, (size_t)P.L.max_seg * kDigitBins * 4, st) != cudaSuccess ||
        cudaMemsetAsync(sc.err, 0, 256, st) != cudaSuccess)
        return fail("kd_dist_local_build: cudaMemsetAsync failed");
    OutPtrs out{};
    RepDst rep{};
    if (P.mc) {  // one replica: the multicast mapping, the switch fans the stores out to every GPU
        out.n = 1;
        out.perm[0] = (int32_t*)(P.mc + P.L.perm);
        out.tp[0] = H.with_tp ? (float*)(P.mc + P.L.tp) : nullptr;
        rep.n = 1;
        rep.perm[0] = out.perm[0];
        rep.tp[0] = out.tp[0];
    } else {
        out.n = P.world;
        for (int r = 0; r < P.world; ++r) {
            out.perm[r] = (int32_t*)(P.heap[r] + P.L.perm);
            out.tp[r] = H.with_tp ? (float*)(P.heap[r] + P.L.tp) : nullptr;
            if (r != P.rank) { rep.perm[rep.n] = out.perm[r]; rep.tp[rep.n] = out.tp[r]; ++rep.n; }
        }
    }
    const int kChunkLevel = chunk_level();
    if (P.world == 1 || n < kChunkMinLocal || kChunkLevel == 0 || P.ncol == 0) {
        if (P.ncol == 0)   // tree_points are gathered from the full point array once the permutation is complete
            for (int r = 0; r < out.n; ++r) out.tp[r] = nullptr;
        return build_core(buf, L & 1, (int64_t)n, P.dim, P.ncol, L, lo, out, all_pts, sc, st, -1, nullptr);
    }

    // Overlap of the sub-tree all-gather with the local phase: the top kChunkLevel local levels run on the whole
    // slice (their few nodes go to every replica directly); then the slice is finished range by range into this
    // GPU's OWN tree arrays, every range on a stream of its own (own scratch), earlier ranges at higher priority.
    // The ranges fill each other's launch-latency gaps, finish one after the other, and each is pushed to the peers
    // (k_replicate) on its stream while the later ones are still being built.  3.3 GB arrive at every GPU in an
    // 8-GPU build of 2^28 x 3: >= 4.3 ms of link time that used to follow the local phase.
    SideStream* S = nullptr;
    if (int rc = side_stream(&S)) return rc;
    int cur = L & 1;
    if (int rc = build_core(buf, cur, (int64_t)n, P.dim, P.ncol, L, lo, out, nullptr, sc, st, kChunkLevel, &cur)) return rc;
    OutPtrs own{};
    own.n = 1;
    own.perm[0] = (int32_t*)(base + P.L.perm);
    own.tp[0] = H.with_tp ? (float*)(base + P.L.tp) : nullptr;
    if (cudaEventRecord(S->fork, st) != cudaSuccess) return fail("kd_dist_local_build: cudaEventRecord failed");
    const unsigned nslice = 1u << kChunkLevel;
    // KD_DIST_TRACE=1: timed events per slice, printed (with a device synchronisation) at the end of the call
    static const bool trace = std::getenv("KD_DIST_TRACE") != nullptr;
    cudaEvent_t tev[2 * kMaxSlices + 1] = {};
    if (trace) {
        for (auto& e : tev) cudaEventCreate(&e);
        cudaEventRecord(tev[2 * kMaxSlices], st);
    }
    const size_t seg_per = (size_t)std::max<int64_t>(1, P.L.max_seg >> kChunkLevel), rec_per = seg_per + kTargetCtas;
    for (unsigned c = 0; c < nslice; ++c) {
        cudaStream_t sc_st = S->s[c];
        if (cudaStreamWaitEvent(sc_st, S->fork, 0) != cudaSuccess) return fail("kd_dist_local_build: cudaStreamWaitEvent failed");
        unsigned clo, cn;
        host_seg_range(n, kChunkLevel, c, clo, cn);
        Cols sb[2];
        for (int b = 0; b < 2; ++b) {
            for (int k = 0; k < kMaxCarry; ++k) sb[b].c[k] = buf[b].c[k] ? buf[b].c[k] + clo : nullptr;
            sb[b].idx = buf[b].idx + clo;
        }
        BuildScratch s2 = sc;   // private select state per slice (the slices run concurrently)
        s2.st = sc.st + c * seg_per;
        s2.ghist = sc.ghist + c * seg_per * kDigitBins;
        s2.rec = sc.rec + c * rec_per;
        s2.cand_cta = sc.cand_cta + c * rec_per;
        s2.cand_key = sc.cand_key + clo;
        s2.cand_idx = sc.cand_idx + clo;
        if (int rc = build_core(sb, cur, (int64_t)cn, P.dim, P.ncol, L + kChunkLevel, lo + clo, own, nullptr, s2, sc_st, -1, nullptr))
            return rc;
        if (trace) cudaEventRecord(tev[2 * c], sc_st);
        k_replicate<<<replicate_grid(), 512, 0, sc_st>>>(own.perm[0], own.tp[0], rep, (size_t)lo + clo, (size_t)cn, P.dim);
        KD_DLAUNCH("k_replicate");
        if (trace) cudaEventRecord(tev[2 * c + 1], sc_st);
        if (cudaEventRecord(S->join[c], sc_st) != cudaSuccess || cudaStreamWaitEvent(st, S->join[c], 0) != cudaSuccess)
            return fail("kd_dist_local_build: joining a slice stream failed");
    }
    if (trace) {
        cudaDeviceSynchronize();
        std::string line = "[kd trace] rank " + std::to_string(P.rank) + " slices (ms after the fork: built / replicated):";
        for (unsigned c = 0; c < nslice; ++c) {
            float a = 0.f, b = 0.f;
            cudaEventElapsedTime(&a, tev[2 * kMaxSlices], tev[2 * c]);
            cudaEventElapsedTime(&b, tev[2 * kMaxSlices], tev[2 * c + 1]);
            char buf2[64];
            snprintf(buf2, sizeof buf2, " %.2f/%.2f", a, b);
            line += buf2;
        }
        fprintf(stderr, "%s\n", line.c_str());
        for (auto& e : tev) cudaEventDestroy(e);
    }
    return 0;
}

}  // namespace kd

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

static int fill_ctx(const kd_dist_ctx* c, kd::HostCtx& H) {
    if (!c) return kd::fail("kd_dist: null context");
    if (c->world < 1 || c->world > kd::kMaxWorld || (c->world & (c->world - 1))) return kd::fail("kd_dist: world must be 1, 2, 4 or 8");
    if (c->rank < 0 || c->rank >= c->world) return kd::fail("kd_dist: bad rank");
    if (c->dim < 1 || c->dim > 4096) return kd::fail("kd_dist: dim must be in [1, 4096]");
    if (c->n_total < 64ll * c->world || c->n_total >= (1ll << 31))
        return kd::fail("kd_dist: need 64 * world <= n_total < 2^31");
    H.P.world = c->world;
    H.P.rank = c->rank;
    H.P.dim = c->dim;
    H.P.ncol = c->dim <= kd::kMaxCarry ? c->dim : 0;   // larger dim: gather mode
    H.P.N = (unsigned)c->n_total;
    H.with_tp = c->with_tree_points != 0;
    H.P.L = kd::make_layout(c->n_total, c->dim, c->world, H.with_tp);
    for (int r = 0; r < kd::kMaxWorld; ++r) H.P.heap[r] = r < c->world ? (unsigned char*)c->heap[r] : nullptr;
    H.P.mc = (unsigned char*)c->mc_heap;
    for (int r = 0; r < c->world; ++r)
        if (!H.P.heap[r]) return kd::fail("kd_dist: missing heap pointer");
    return 0;
}

size_t kd_dist_heap_bytes(int64_t n_total, int dim, int world, int with_tree_points) {
    if (n_total <= 0 || dim <= 0 || dim > 4096 || world < 1 || world > kd::kMaxWorld) return 0;
    return kd::make_layout(n_total, dim, world, with_tree_points != 0).bytes;
}

int kd_dist_offsets(int64_t n_total, int dim, int world, int with_tree_points, size_t* perm_off, size_t* tp_off) {
    if (n_total <= 0 || dim <= 0 || dim > 4096 || world < 1 || world > kd::kMaxWorld)
        return kd::fail("kd_dist_offsets: bad arguments");
    kd::DistLayout L = kd::make_layout(n_total, dim, world, with_tree_points != 0);
    if (perm_off) *perm_off = L.perm;
    if (tp_off) *tp_off = L.tp;
    return 0;
}

int kd_dist_init(const kd_dist_ctx* ctx, const float* local_points, int64_t local_n, int64_t first_row, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    if (!local_points || local_n <= 0 || local_n > H.P.L.cap) return kd::fail("kd_dist_init: bad local slice");
    cudaStream_t st = (cudaStream_t)stream;
    // control words and flags start at zero once per heap; epochs only grow afterwards
    kd::reset_launches();
    if (H.P.ncol)
        kd::k_dist_init<<<(unsigned)((local_n + kd::kDInitRows - 1) / kd::kDInitRows), kd::kDInitBlock, 0, st>>>(
            H.P, local_points, (unsigned)local_n, (unsigned)first_row);
    else
        kd::k_dist_init_rows<<<148 * 4, 256, 0, st>>>(H.P, local_points, (unsigned)local_n, (unsigned)first_row);
    kd::count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_dist_init: ") + cudaGetErrorString(e));
}

int kd_dist_phase(const kd_dist_ctx* ctx, int level, int phase, int pass, uint64_t serial, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    if (level < 0 || level >= kd::dist_levels(H.P.world)) return kd::fail("kd_dist_phase: bad level");
    return kd::run_phase(H, level, phase, pass, serial, (cudaStream_t)stream);
}

int kd_dist_local_build(const kd_dist_ctx* ctx, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    return kd::local_build(H, (cudaStream_t)stream);
}

int kd_dist_done(const kd_dist_ctx* ctx, int post, uint64_t serial, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    kd::LevelGeo g = kd::make_geo(H.P, 0);  // the whole world is one group
    const unsigned long long epoch = serial * 64ull + 63ull;
    if (post) kd::k_flag_post<<<1, 32, 0, (cudaStream_t)stream>>>(H.P, g, kd::F_DONE, epoch);
    else kd::k_flag_wait<<<1, 32, 0, (cudaStream_t)stream>>>(H.P, g, kd::F_DONE, epoch);
    kd::count_launch();
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_dist_done: ") + cudaGetErrorString(e));
}

/* after the done-barrier: in gather mode (dim > 8) every rank fills its tree_points from its own copy of the point
 * array and the complete permutation (nothing to do otherwise) */
int kd_dist_finish(const kd_dist_ctx* ctx, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    if (H.P.ncol != 0 || !H.with_tp) return 0;
    unsigned char* base = H.P.heap[H.P.rank];
    return kd_gather_rows_gpu((const float*)(base + H.P.L.pts_all), (const int32_t*)(base + H.P.L.perm), (int64_t)H.P.N,
                              H.P.dim, (float*)(base + H.P.L.tp), stream);
}

// one rank's whole build, in program order (one process per GPU)
int kd_dist_run(const kd_dist_ctx* ctx, uint64_t serial, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int L = kd::dist_levels(H.P.world);
    kd::Step steps[64];
    const int ns = kd::level_schedule(steps);
    for (int level = 0; level < L; ++level)
        for (int k = 0; k < ns; ++k)
            if (int rc = kd::run_phase(H, level, steps[k].phase, steps[k].pass, serial, st)) return rc;
    if (int rc = kd::local_build(H, st)) return rc;
    if (int rc = kd_dist_done(ctx, 1, serial, stream)) return rc;
    if (int rc = kd_dist_done(ctx, 0, serial, stream)) return rc;
    return kd_dist_finish(ctx, stream);
}

/* Synchronises the stream and copies this rank's counters of the level that ran last to the host:
 * out[0..7] = local count, elements going left, 1 if the rank holds the node, first left / right output offset,
 * elements below the bracket, candidates inside it, 1 if the bracket was used (load-balance reports, tools). */
int kd_dist_level_counts(const kd_dist_ctx* ctx, uint32_t* out, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    if (!out) return kd::fail("kd_dist_level_counts: null output");
    kd::DistCtl c;
    cudaError_t e = cudaMemcpyAsync(&c, H.P.heap[H.P.rank] + H.P.L.ctl, sizeof c, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess) return kd::fail(std::string("kd_dist_level_counts: ") + cudaGetErrorString(e));
    out[0] = c.cnt; out[1] = c.my_less; out[2] = c.mine; out[3] = c.Loff; out[4] = c.Roff;
    out[5] = c.cL; out[6] = c.cand_cnt; out[7] = (uint32_t)c.use_cand;
    return 0;
}

/* first failed bounds assertion of the distributed kernels (-DKD_DEBUG builds; 0 otherwise); clears it */
int kd_dist_debug_err(void) {
#if defined(KD_DEBUG)
    int a = 0, zero = 0;
    if (cudaMemcpyFromSymbol(&a, kd_dbg_err, sizeof(int)) != cudaSuccess) return -1;
    cudaMemcpyToSymbol(kd_dbg_err, &zero, sizeof(int));
    return a;
#else
    return 0;
#endif
}

/* the (phase, pass) pairs of one distributed level in launch order; returns their number (<= cap) */
int kd_dist_schedule(int* phases, int* passes, int cap) {
    kd::Step steps[64];
    const int ns = kd::level_schedule(steps);
    if (!phases || !passes || cap < ns) return kd::fail("kd_dist_schedule: need room for " + std::to_string(ns) + " steps"), -1;
    for (int k = 0; k < ns; ++k) { phases[k] = steps[k].phase; passes[k] = steps[k].pass; }
    return ns;
}

int kd_dist_status(const kd_dist_ctx* ctx, void* stream) {
    kd::HostCtx H;
    if (int rc = fill_ctx(ctx, H)) return rc;
    int host_err[2] = {0, 0};
    int* dev_err = (int*)(H.P.heap[H.P.rank] + H.P.L.lerr) + 8;
    kd::k_dist_read_err<<<1, 1, 0, (cudaStream_t)stream>>>(H.P, dev_err);
    cudaError_t e = cudaMemcpyAsync(&host_err[0], dev_err, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(&host_err[1], H.P.heap[H.P.rank] + H.P.L.lerr, sizeof(int), cudaMemcpyDeviceToHost,
                            (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess) return kd::fail(std::string("kd_dist_status: ") + cudaGetErrorString(e));
    if (host_err[0]) return kd::fail("distributed build failed on the device, code " + std::to_string(host_err[0]));
    if (host_err[1]) return kd::fail("local phase of the distributed build failed, digit pass " + std::to_string(host_err[1] - 1));
    return 0;
}

// ---- symmetric heap runtime (H6): allocation + CUDA IPC handles; exchange of the handles is the caller's
// job (torch.distributed object collectives), nothing else goes through a collective library ----
int kd_symm_alloc(size_t bytes, void** ptr, void* handle64) {
    if (!ptr || !handle64 || bytes == 0) return kd::fail("kd_symm_alloc: bad arguments");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return kd::fail(std::string("kd_symm_alloc: cudaMalloc: ") + cudaGetErrorString(e));
    e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return kd::fail(std::string("kd_symm_alloc: ") + cudaGetErrorString(e)); }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    std::memcpy(handle64, &h, 64);
    *ptr = p;
    return 0;
}

int kd_symm_open(const void* handle64, void** ptr) {
    if (!handle64 || !ptr) return kd::fail("kd_symm_open: bad arguments");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return kd::fail(std::string("kd_symm_open: cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
    *ptr = p;
    return 0;
}

int kd_symm_close(void* ptr) {
    cudaError_t e = cudaIpcCloseMemHandle(ptr);
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_symm_close: ") + cudaGetErrorString(e));
}

int kd_symm_free(void* ptr) {
    cudaError_t e = cudaFree(ptr);
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_symm_free: ") + cudaGetErrorString(e));
}

int kd_symm_memset(void* ptr, size_t bytes, void* stream) {
    cudaError_t e = cudaMemsetAsync(ptr, 0, bytes, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_symm_memset: ") + cudaGetErrorString(e));
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// K8: forest min-reduce (the reference's MPI_Reduce(MPI_MIN) of the per-rank query distances,
// kdtree_mpi.cpp:252-253) as one kernel per rank over peer memory: every rank folds its distances into
// rank 0's buffer with system-scope atomicMin (distances are >= 0, so their float bits order like
// unsigned integers) and raises a flag there; rank 0 waits for all flags and decodes.
// Buffer layout per rank (a small symmetric allocation): u64 ready[8] | u64 done[8] | u32 val[nq].
// ------------------------------------------------------------------------------------------------
namespace kd {
struct ForestBuf {
    unsigned long long ready[kMaxWorld];
    unsigned long long done[kMaxWorld];
    unsigned val[1];
};

__global__ void k_forest_min(ForestBuf* root, ForestBuf* mine, int rank, int world, const float* __restrict__ dist,
                             int nq, float* __restrict__ out, unsigned long long epoch, int* err) {
    const int tid = threadIdx.x;
    if (rank == 0) {
        for (int q = tid; q < nq; q += blockDim.x) root->val[q] = 0xffffffffu;
        __threadfence_system();
        __syncthreads();
        // tell every rank (through its own buffer) that the accumulator is initialised
        // (peers poll their local `ready[0]`, which rank 0 cannot reach here, so they poll root->ready[rank])
        if (tid < world) st_release_sys(&root->ready[tid], epoch);
    } else {
        __shared__ int s_ok;
        if (tid == 0) {
            s_ok = wait_flag(&root->ready[rank], epoch) ? 1 : 0;
            if (!s_ok) atomicExch(err, 500);
        }
        __syncthreads();
        if (!s_ok) return;   // the accumulator may not be initialised: fold nothing into it, signal nothing
    }
    for (int q = tid; q < nq; q += blockDim.x) atomicMin_system(&root->val[q], __float_as_uint(dist[q]));
    __threadfence_system();
    __syncthreads();
    if (tid == 0) st_release_sys(&root->done[rank], epoch);
    if (rank == 0) {
        if (tid < world && !wait_flag(&root->done[tid], epoch)) atomicExch(err, 501);
        __syncthreads();
        for (int q = tid; q < nq; q += blockDim.x) out[q] = __uint_as_float(__ldcg(&root->val[q]));
    }
    (void)mine;
}
}  // namespace kd

extern "C" {
size_t kd_forest_buf_bytes(int nq) { return sizeof(kd::ForestBuf) + (size_t)(nq > 0 ? nq : 1) * 4 + 64; }

/* dist: this rank's nq distances (device); out: rank 0 receives the per-query minimum (device, may be NULL
 * elsewhere); root_buf: rank 0's buffer as addressable from this rank; err: device int. */
int kd_forest_min_reduce(void* root_buf, void* my_buf, int rank, int world, const float* dist, int nq, float* out,
                         uint64_t epoch, int* err, void* stream) {
    if (!root_buf || !dist || nq <= 0 || world < 1 || world > kd::kMaxWorld || rank < 0 || rank >= world || !err ||
        (rank == 0 && !out))
        return kd::fail("kd_forest_min_reduce: bad arguments");
    kd::k_forest_min<<<1, 256, 0, (cudaStream_t)stream>>>((kd::ForestBuf*)root_buf, (kd::ForestBuf*)my_buf, rank, world,
                                                         dist, nq, out, epoch, err);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_forest_min_reduce: ") + cudaGetErrorString(e));
}
}

// ------------------------------------------------------------------------------------------------
// K12 over ranks: per-axis min / max of a point set spread over W GPUs, as ONE kernel per rank over peer memory
// (an all-reduce: every rank ends up with the result).  Every rank initialises its own accumulator and tells every
// peer so; folds its local bounds into EVERY rank's accumulator with system-scope atomicMin / atomicMax on the
// order-preserving keys; signals every peer; waits for all signals and decodes.  No collective library call.
// Buffer per rank (a small symmetric allocation): u64 ready[8] | u64 done[8] | u32 mn[dim] | u32 mx[dim].
// ------------------------------------------------------------------------------------------------
namespace kd {
struct MinMaxBuf {
    unsigned long long ready[kMaxWorld];
    unsigned long long done[kMaxWorld];
    unsigned v[1];   // mn[dim], mx[dim]
};
struct MinMaxPeers { MinMaxBuf* b[kMaxWorld]; };

__global__ void __launch_bounds__(256) k_peer_minmax(MinMaxPeers peers, int rank, int world, const float* __restrict__ lmin,
                                                     const float* __restrict__ lmax, int dim, float* __restrict__ out_min,
                                                     float* __restrict__ out_max, unsigned long long epoch, int* err) {
    __shared__ int s_ok;
    const int tid = threadIdx.x;
    MinMaxBuf* mine = peers.b[rank];
    if (tid == 0) s_ok = 1;
    for (int d = tid; d < dim; d += blockDim.x) { mine->v[d] = 0xffffffffu; mine->v[dim + d] = 0u; }
    __threadfence_system();
    __syncthreads();
    if (tid < world) st_release_sys(&peers.b[tid]->ready[rank], epoch);          // "my accumulator is ready", to everyone
    if (tid < world && !wait_flag(&mine->ready[tid], epoch)) { atomicExch(err, 700); s_ok = 0; }
    __syncthreads();
    if (!s_ok) return;   // some accumulator may not be initialised: fold nothing, signal nothing
    for (int r = 0; r < world; ++r)
        for (int d = tid; d < dim; d += blockDim.x) {
            atomicMin_system(&peers.b[r]->v[d], ordered_key(lmin[d]));
            atomicMax_system(&peers.b[r]->v[dim + d], ordered_key(lmax[d]));
        }
    __threadfence_system();
    __syncthreads();
    if (tid < world) st_release_sys(&peers.b[tid]->done[rank], epoch);
    if (tid < world && !wait_flag(&mine->done[tid], epoch)) { atomicExch(err, 701); s_ok = 0; }
    __syncthreads();
    if (!s_ok) return;
    auto back = [](unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); };
    for (int d = tid; d < dim; d += blockDim.x) {
        out_min[d] = back(__ldcg(&mine->v[d]));
        out_max[d] = back(__ldcg(&mine->v[dim + d]));
    }
}
}  // namespace kd

extern "C" {
size_t kd_peer_minmax_buf_bytes(int dim) { return sizeof(kd::MinMaxBuf) + (size_t)2 * (dim > 0 ? dim : 1) * 4 + 64; }

/* bufs[r]: rank r's buffer (kd_peer_minmax_buf_bytes, from kd_symm_alloc, zeroed) as addressable from this rank;
 * local_min / local_max: this rank's per-axis bounds (device, e.g. from kd_bbox_gpu); out_*: the bounds over all ranks
 * (device); epoch must grow with every call; err: device int. */
int kd_peer_minmax_reduce(void* const* bufs, int rank, int world, const float* local_min, const float* local_max, int dim,
                          float* out_min, float* out_max, uint64_t epoch, int* err, void* stream) {
    if (!bufs || !local_min || !local_max || !out_min || !out_max || !err || dim <= 0 || world < 1 ||
        world > kd::kMaxWorld || rank < 0 || rank >= world)
        return kd::fail("kd_peer_minmax_reduce: bad arguments");
    kd::MinMaxPeers P{};
    for (int r = 0; r < world; ++r) {
        if (!bufs[r]) return kd::fail("kd_peer_minmax_reduce: missing buffer pointer");
        P.b[r] = (kd::MinMaxBuf*)bufs[r];
    }
    kd::k_peer_minmax<<<1, 256, 0, (cudaStream_t)stream>>>(P, rank, world, local_min, local_max, dim, out_min, out_max, epoch, err);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : kd::fail(std::string("kd_peer_minmax_reduce: ") + cudaGetErrorString(e));
}
}
