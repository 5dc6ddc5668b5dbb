// Workspace layout and small PODs shared by the build kernels and their host drivers.
#pragma once
#include <cstddef>
#include <cstdint>

namespace kd {

constexpr int kMaxCarry = 8;      // dim <= kMaxCarry: coordinates travel with the points (SoA columns);
                                  // larger dim: only indices move and keys are gathered per level
constexpr int kFinishT = 4096;    // ranges of at most this many points are finished in shared memory (k_subtree)
constexpr int kMediumMax = 16384;   // ranges up to this size are handled by one CTA per level (k_medium)
constexpr int kTargetCtas = 148 * 16;

struct SegState {
    unsigned long long prefix;  // composite bits resolved so far (bits below `shift` are zero)
    unsigned int rank;          // 0-based rank still wanted among the candidates
    unsigned int cnt;           // candidates matching the prefix
    unsigned int curL, curR;    // output cursors of the partition
    int shift;                  // compare (composite >> shift) with (prefix >> shift)
    int done;                   // 0 selecting, 1 finished, 2 the unique candidate still has to be read back
    // bracket select (large ranges): candidates are the elements with lowP <= composite <= highP
    unsigned long long lowP, highP;
    unsigned int cL;            // elements below lowP
    unsigned int cand_cnt;      // candidates written to the candidate arrays (at the range's own slots)
    int use_cand;               // 1: the digit passes run on the candidate arrays
    int pad_;
};

struct Cols {
    float* c[kMaxCarry];
    int32_t* idx;
};

struct MaxRec {
    unsigned long long comp;
    unsigned pos;
    unsigned valid;
};

constexpr int kMaxReplicas = 8;
// Every copy of the tree arrays that receives the build's output (one per GPU in a distributed build).
struct OutPtrs {
    int32_t* perm[kMaxReplicas];
    float* tp[kMaxReplicas];  // may be null
    int n;
};

struct BuildScratch {
    SegState* st;
    unsigned* ghist;
    MaxRec* rec;
    int* err;
    float* keybuf;     // gather mode only
    float* cand_key;   // bracket select: candidate keys (raw float bits) ...
    int32_t* cand_idx; // ... and point indices, n entries each; null disables the bracket
    unsigned* cand_cta; // candidates found by every CTA of k_bracket
};

struct BuildPlan {
    int64_t n;
    int dim, ncol, finish_level;
    int64_t max_seg, max_rec;
    size_t off_col[2][kMaxCarry], off_idx[2], off_key, off_state, off_hist, off_rec, off_err, off_ckey, off_cidx, off_ccta, bytes;
};

BuildPlan make_plan(int64_t n, int dim);

#if defined(__CUDACC__)
// Range s of level `level` of a tree over N points: walk the shape rule along the bits of s.
__device__ __forceinline__ void seg_range(unsigned N, int level, unsigned s, unsigned& lo, unsigned& n) {
    unsigned l = 0, cnt = N;
    for (int b = level - 1; b >= 0; --b) {
        unsigned m = cnt >> 1;
        if ((s >> b) & 1u) { l += m + 1; cnt = cnt - m - 1; } else { cnt = m; }
    }
    lo = l;
    n = cnt;
}

// stop_level >= 0: only the select/partition levels [0, stop_level) run (stop_level must lie above the level of
// the shared-memory builder, i.e. (n >> stop_level) > kFinishT) and *cur_out receives the buffer holding the result.
int build_core(Cols buf[2], int cur, int64_t n, int dim, int ncol, int depth0, unsigned out_lo, OutPtrs out,
               const float* points, BuildScratch sc, cudaStream_t stream, int stop_level, int* cur_out);
void count_launch();
void reset_launches();
#endif

inline size_t medium_smem_bytes() { return (size_t)kMediumMax * 4 + (size_t)2048 * 4; }

inline size_t subtree_smem_bytes(int ncol) {
    // scol[ncol][T+8] f32, sidx[T+8] i32, slist[16 warps][160] u64, steam[8] x 16 B, sbins[192] u32, swtot[16] u32,
    // stab[2][T/2] u32, sord[T] u16  (k_subtree carves them in this order)
    return (size_t)(ncol + 1) * (kFinishT + 8) * 4 + (size_t)16 * 160 * 8 + 8 * 16 + 192 * 4 + 16 * 4 +
           (size_t)kFinishT * 4 + (size_t)kFinishT * 2;
}

}  // namespace kd
