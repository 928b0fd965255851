// sample.cuh — device-resident form of one sample's top-gene transcripts.
//
// Layout in HBM (all arrays live for the lifetime of the rcq_sample):
//   fwd[arena/16]   uint32  2-bit bases (A0 C1 G2 T3), base p in bits 2*(p%16) of word p/16
//   rc[arena/16]    uint32  per-transcript reverse complement, transcript t at the same offset
//   mask[arena/32], mask_rc[arena/32]  uint32, bit set = not ACGT (k-mers over it are not seeds,
//                   it mismatches everything in extension)
//   off[n_seq]      int32   arena offset of transcript t (multiple of 32 bases), len[n_seq]
//   gene/iso[n_seq] int32
//   index[n_index]  uint64  sorted minimizer entries  key<<32 | pos<<1 | orient, key = a second
//                   bijective mix of the canonical 16-mer hash: minimizer hashes crowd towards 0, the
//                   remix spreads them evenly over the directory buckets
//   fctx[n_index]   uint64  flank context of each entry: the 13 bases before / after the 16-mer on its
//                   canonical strand (low / high word; bit 31 = unavailable, bit 30 low = palindrome)
//   dir[2^DIRBITS+1] uint32 first index entry of each hash-prefix bucket
//   sel[arena/32]   uint32  bitmap of indexed positions;  gidx[n_seq] dense gene index
//   smask[arena/32], smask_rc[arena/32]  uint32  seeding mask = mask | symmetric-DUST low-complexity mask
//                   (dust_kernels.cuh): 16-mers touching it are not indexed, seeds need a clean 28-base window;
//                   extensions keep reading `mask` (soft masking)
//   blk2tx[arena/32] int32  transcript owning each 32-base block (transcripts start on block
//                   boundaries);  ol[n_seq] int2 {off, len} for one-load transcript bounds
#pragma once
#include "common.cuh"

#define RCQ_DIRBITS 22

struct rcq_sample {
    rcq_ctx* ctx = nullptr;
    int32_t n_seq = 0;
    int64_t total = 0;      // sum of transcript lengths
    int64_t arena = 0;      // padded arena length in bases (multiple of 64, + guard)
    int64_t n_index = 0;
    int64_t n_masked = 0;
    int64_t n_dust = 0;     // bases masked by DUST and not by `mask`
    int32_t max_len = 0;
    uint32_t *smask = nullptr, *smask_rc = nullptr;
    uint32_t *fwd = nullptr, *rc = nullptr, *mask = nullptr, *mask_rc = nullptr;
    int32_t *off = nullptr, *len = nullptr, *gene = nullptr, *iso = nullptr;
    uint64_t* index = nullptr;
    uint64_t* fctx = nullptr;  // flank context of index[i] (sample.cu: flank_ctx)
    uint32_t* dir = nullptr;
    uint32_t* sel = nullptr;   // arena/32 words: bit p set = position p holds an index entry
    int32_t* gidx = nullptr;   // dense index of gene[t] among the sample's sorted distinct genes
    int32_t* blk2tx = nullptr; // arena/32 entries
    int2* ol = nullptr;        // {off, len} per transcript
    int32_t n_genes = 0;
};

// A strand view of a sample: which arena to read for subject-side bases.
struct SeqArena {
    const uint32_t* bits;
    const uint32_t* mask;  // nullptr when the sample has no masked base
};

// 32 consecutive bases starting at arena position p (may be unaligned), as 64 bits.
__device__ __forceinline__ uint64_t fetch32(const uint32_t* __restrict__ bits, int64_t p) {
    int64_t w = p >> 4;
    int sh = (int)(p & 15) * 2;
    uint32_t a = __ldg(bits + w), b = __ldg(bits + w + 1), c = __ldg(bits + w + 2);
    uint64_t lo = ((uint64_t)b << 32) | a;
    if (sh == 0) return lo;
    return (lo >> sh) | ((uint64_t)c << (64 - sh));
}
// 32 mask bits starting at arena position p
__device__ __forceinline__ uint32_t fetchmask32(const uint32_t* __restrict__ mask, int64_t p) {
    int64_t w = p >> 5;
    int sh = (int)(p & 31);
    uint32_t a = __ldg(mask + w), b = __ldg(mask + w + 1);
    return sh ? ((a >> sh) | (b << (32 - sh))) : a;
}
// one bit per base: 1 where the 2-bit codes of x and y are equal (bit i <-> base i), 32 bases
__device__ __forceinline__ uint32_t eq32(uint64_t x, uint64_t y) {
    uint64_t d = x ^ y;
    d = (d | (d >> 1)) & 0x5555555555555555ull;  // bit 2i set = mismatch at base i
    // compress even bits to 32 bits
    d = (d | (d >> 1)) & 0x3333333333333333ull;
    d = (d | (d >> 2)) & 0x0f0f0f0f0f0f0f0full;
    d = (d | (d >> 4)) & 0x00ff00ff00ff00ffull;
    d = (d | (d >> 8)) & 0x0000ffff0000ffffull;
    d = (d | (d >> 16)) & 0x00000000ffffffffull;
    return ~(uint32_t)d;
}
// match bits for 32 aligned base pairs starting at (pa, pb); masked bases never match
__device__ __forceinline__ uint32_t match32(const SeqArena& A, int64_t pa, const SeqArena& B,
                                            int64_t pb) {
    uint32_t m = eq32(fetch32(A.bits, pa), fetch32(B.bits, pb));
    if (A.mask) m &= ~fetchmask32(A.mask, pa);
    if (B.mask) m &= ~fetchmask32(B.mask, pb);
    return m;
}
// same, for 32 base pairs ENDING at (pa, pb) inclusive, bit i <-> pair (pa-i, pb-i).
// Callers guarantee pa, pb >= 0; positions below 0 of the arena are never inspected because
// callers clamp the usable bit count by the distance to the transcript start.
__device__ __forceinline__ uint32_t match32_back(const SeqArena& A, int64_t pa, const SeqArena& B,
                                                 int64_t pb) {
    // arenas carry a 32-base guard in front (see sample.cu), so pa-31 >= -32 is addressable
    uint32_t m = match32(A, pa - 31, B, pb - 31);
    return __brev(m);
}
