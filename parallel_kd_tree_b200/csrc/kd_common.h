// Shared host/device definitions for the B200 kd-tree library.
//
// Tree representation: the tree *shape* depends on N only
// (reference kdtree_sequential.cpp:51-56: node = list[n/2], left = n/2 points,
// right = n - n/2 - 1 points), so a tree is one permutation `perm[N]` in
// IN-ORDER layout: the node of the range [lo, lo+n) sits at slot lo + n/2, its
// left sub-tree owns [lo, lo+n/2) and its right sub-tree [lo+n/2+1, lo+n).
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define KD_HD __host__ __device__ __forceinline__
#else
#define KD_HD inline
#endif

namespace kd {

// Order-preserving map float -> uint32 with -0 folded onto +0, so that the
// reference comparator `a < b` (Node.cpp:46-48) becomes an unsigned compare
// and equal floats give equal keys (ties are then broken by point index).
KD_HD uint32_t ordered_key_bits(uint32_t u) {
    if ((u << 1) == 0u) u = 0u;  // +-0 -> +0
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
#if !defined(__CUDA_ARCH__)
inline uint32_t ordered_key(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return ordered_key_bits(u);
}
#else
__device__ __forceinline__ uint32_t ordered_key(float f) { return ordered_key_bits(__float_as_uint(f)); }
#endif

// 64-bit composite used by every select: (key on the split axis, point index).
KD_HD uint64_t composite(uint32_t key, uint32_t idx) { return ((uint64_t)key << 32) | idx; }

// Children of a range of n points in in-order layout.
KD_HD int64_t left_size(int64_t n) { return n / 2; }
KD_HD int64_t right_size(int64_t n) { return n - n / 2 - 1; }

// Number of levels of the tree over n points (max node depth + 1).
KD_HD int tree_levels(int64_t n) {
    int l = 0;
    while (n > 0) { n >>= 1; ++l; }
    return l;
}

// Radix digit schedule over the 64-bit composite, most significant first:
// key bits 31..21, 20..10, 9..0, then index bits 31..21, 20..10, 9..0.
constexpr int kNumDigits = 6;
constexpr int kDigitBins = 2048;
KD_HD int digit_shift(int d) { return d == 0 ? 53 : d == 1 ? 42 : d == 2 ? 32 : d == 3 ? 21 : d == 4 ? 10 : 0; }
KD_HD int digit_bits(int d) { return (d == 2 || d == 5) ? 10 : 11; }
// Single-GPU select: bits [hi, 64) of the composite are resolved; the next digit is the (at most) 11 bits below
// them and never straddles the key | index boundary at bit 32.  From hi = 64 this is the schedule above.
KD_HD int select_digit_shift(int hi) { return hi > 32 ? (hi - 11 > 32 ? hi - 11 : 32) : (hi - 11 > 0 ? hi - 11 : 0); }

}  // namespace kd

// Sanitizer substitute (compute-sanitizer is closed on the development pool): a -DKD_DEBUG build checks every computed
// store offset of the build kernels against the range it must fall into.  A failed check records its code (first one
// wins), SKIPS the store and lets the kernel run on, so barriers stay balanced; kd_debug_bounds_status() reports it.
// In release builds KD_INB(cond, code) is the constant `true` and costs nothing.
#if defined(KD_DEBUG) && defined(__CUDACC__)
static __device__ int kd_dbg_err;   // one per translation unit
#define KD_INB(cond, code) ((cond) ? true : (atomicCAS(&kd_dbg_err, 0, (code)), false))
#else
#define KD_INB(cond, code) true
#endif
