// search.cuh — records passed between the kernels of one pair search.
#pragma once
#include "sample.cuh"

struct Seed { int32_t ta, tb, minus, rs, re, diag; };               // exact run [rs, re) on ta
struct Ungapped { int32_t ta, tb, minus, diag, a0, a1, score, i0; };  // i0 = gapped start on ta
// Gapped HSP, 12 int32 (layout of rcq_debug_last_hsps in include/rcq.h); ta < 0 = empty slot
struct Gapped { int32_t ta, tb, minus, a0, a1, b0, b1, score2, mism, gaps, gapopen, nident; };

struct PairView {
    SeqArena Af, Ar, Bf, Br;     // A forward / reverse-complement, B forward / reverse-complement
    const int32_t *offA, *lenA, *offB, *lenB;
    int32_t nA, nB;
    const uint32_t *selA, *selB;
    const int32_t *blkA, *blkB;   // 32-base block -> transcript
    const int2 *olA, *olB;        // {off, len} per transcript
    const uint32_t *smA, *smB, *smBr;   // seeding masks (N | DUST): A forward, B forward, B reverse complement
};

struct rcq_table {
    rcq_ctx* ctx = nullptr;
    int64_t n = 0;
    int32_t* cols = nullptr;  // 17 column-major int32 arrays of n rows (order of RCQ_COL_*)
};
enum {
    RCQ_COL_LENGTH = 0, RCQ_COL_MISMATCH, RCQ_COL_GAPOPEN, RCQ_COL_QSTART, RCQ_COL_QEND,
    RCQ_COL_SSTART, RCQ_COL_SEND, RCQ_COL_SCORE, RCQ_COL_GAPS, RCQ_COL_NIDENT, RCQ_COL_QGENE,
    RCQ_COL_QISO, RCQ_COL_SGENE, RCQ_COL_SISO, RCQ_COL_QLEN, RCQ_COL_MINUS, RCQ_COL_REVERSE,
    RCQ_NCOLS
};
