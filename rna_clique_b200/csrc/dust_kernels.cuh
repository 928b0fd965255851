// dust_kernels.cuh — symmetric DUST (blastn's default query filter "-dust 20 64 1": level 20, window 64,
// linker 1; Morgulis, Gertz, Schaffer, Agarwala 2006) as part of the sample build.  What blastn does to every
// query before find_homologs.py:124 sees a hit: low-complexity stretches are SOFT-masked, i.e. kept out of
// seeding only.  The mask joins the N mask in the seeding mask `smask` read by k_minimizers (masked 16-mers
// are not indexed) and by k_seed_verify (a seed needs a 28-base window clean in both samples).
//
// Definition (the oracle states it the same way, oracle/rcq_oracle.c: dust_mask): over the triplets of a
// maximal ACGT run an interval of l >= 2 triplets scores r / (l - 1), r = sum over triplet values of
// c (c - 1) / 2; it is perfect when l <= 62, the score exceeds 2 and no interval inside it scores higher;
// the mask is the union of all perfect intervals.  A pure, order-independent function of the sequence.
//
//   k_dust_triplets  one byte per position: the triplet starting there (0..63) or 0xFF (touches a masked base)
//   k_dust_flag      sliding pair counts over the last 16 / 32 / 62 triplets, one thread per 256 positions;
//                    an interval of l triplets ending at e lies inside the window of the next scale up, so
//                    r16 <= 8 and r32 <= 32 and r62 <= 64 proves that no interval ending at e scores above 2
//                    (l <= 16 needs r > 2 (5 - 1), 16 < l <= 32 needs r > 32, l > 32 needs r > 64).  Random
//                    sequence passes this at well under 1 % of the positions.
//   k_dust_eval      for every flagged end position: the 62-step suffix scan, and where a suffix scores
//                    above 2 the rows of the triangle below it (best score of any contained interval);
//                    perfect suffixes set their bases in the mask.
//   k_dust_finish    smask = N mask | DUST, and its per-transcript reversal for the reverse-complement arena.
// Bound: integer issue on L1/shared-memory resident data; HBM traffic is one byte per base each way.
#pragma once
#include "sample.cuh"

#define DUST_MAXT 62
#define DUST_SEG 256          // end positions per thread of k_dust_flag
#define DUST_FLAG_THREADS 64
#define DUST_EVAL_THREADS 128

// 16 positions per thread: triplets starting at p .. p+15 need bases p .. p+17
__global__ void k_dust_triplets(const uint32_t* __restrict__ fwd, const uint32_t* __restrict__ mask, int64_t arena,
                                uint8_t* __restrict__ trip) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w * 16 >= arena) return;
    const int64_t last = arena / 16 - 1;
    const uint32_t a = __ldg(fwd + w), b = w < last ? __ldg(fwd + w + 1) : 0u;
    const uint32_t m0 = __ldg(mask + (w >> 1)), m1 = (w >> 1) + 1 < arena / 32 ? __ldg(mask + (w >> 1) + 1) : 0xFFFFFFFFu;
    const uint64_t m = ((uint64_t)m1 << 32 | m0) >> ((w & 1) * 16);     // mask bits of bases 16w ..
    uint32_t out[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 16; j++) {
        uint32_t t = __funnelshift_r(a, b, 2 * j) & 63u;   // any one-to-one code of the three bases will do
        if ((uint32_t)(m >> j) & 7u) t = 0xFFu;
        out[j >> 2] |= t << (8 * (j & 3));
    }
    *(uint4*)(trip + w * 16) = make_uint4(out[0], out[1], out[2], out[3]);
}

// A block covers DUST_FLAG_THREADS x DUST_SEG consecutive positions.  Its triplet bytes (plus the 64 before them)
// are staged in shared memory with coalesced loads, rows of 256 bytes padded to 260 so that the lanes of a warp,
// which walk 256 bytes apart, read different banks; a thread's four counters of one word (triplet & 3) share a
// 32-bit word whose bank is the thread's own.  (Reading the bytes straight from global memory cost 3.7 ms per
// sample: 32 sectors per warp load, four loads per step.)
#define DUST_ROW 260
__global__ void __launch_bounds__(DUST_FLAG_THREADS) k_dust_flag(const uint8_t* __restrict__ trip, int64_t arena,
                                                                  uint32_t* __restrict__ flag) {
    // one counter array per window length, [triplet >> 2][thread], byte = triplet & 3: separate arrays let the
    // three read-modify-write chains of a step run side by side
    __shared__ uint32_t cnt16[16 * DUST_FLAG_THREADS], cnt32[16 * DUST_FLAG_THREADS], cnt62[16 * DUST_FLAG_THREADS];
    __shared__ __align__(16) uint8_t tb[(DUST_FLAG_THREADS + 1) * DUST_ROW];   // positions block_start - 64 ..
    const int tid = threadIdx.x;
    const int64_t block_start = (int64_t)blockIdx.x * DUST_FLAG_THREADS * DUST_SEG;
    for (int i = tid; i < 16 * DUST_FLAG_THREADS; i += DUST_FLAG_THREADS) { cnt16[i] = 0u; cnt32[i] = 0u; cnt62[i] = 0u; }
    // stage: word w of the region holds positions block_start - 64 + 4w .. +3
    const int n_words = (DUST_FLAG_THREADS * DUST_SEG + 64) / 4;
    for (int w = tid; w < n_words; w += DUST_FLAG_THREADS) {
        const int64_t p = block_start - 64 + 4ll * w;
        uint32_t v = 0xFFFFFFFFu;
        if (p >= 0 && p + 4 <= arena) v = __ldg((const uint32_t*)(trip + p));
        const int rel = 4 * w;
        *(uint32_t*)(tb + (rel >> 8) * DUST_ROW + (rel & 255)) = v;
    }
    __syncthreads();
    const int64_t start = block_start + (int64_t)tid * DUST_SEG;
    if (start >= arena) return;
    const int64_t end = min(start + DUST_SEG, arena);
    int32_t r16 = 0, r32 = 0, r62 = 0;
    uint32_t bits = 0;
    uint8_t* const cb0 = (uint8_t*)cnt16; uint8_t* const cb1 = (uint8_t*)cnt32; uint8_t* const cb2 = (uint8_t*)cnt62;
#define TB(rel) tb[((rel) >> 8) * DUST_ROW + ((rel) & 255)]
#define CNT(s, t) ((s) == 0 ? cb0 : (s) == 1 ? cb1 : cb2)[((((t) >> 2)) * DUST_FLAG_THREADS + tid) * 4 + ((t) & 3)]
    // rel = position - (block_start - 64); the walk starts DUST_MAXT positions early to fill the windows
    const int rel0 = 64 + tid * DUST_SEG;
    const int n = (int)(end - start);
    for (int x = -DUST_MAXT; x < n; x++) {
        const int rel = rel0 + x;
        const uint32_t t = TB(rel);
        if (t != 0xFFu) { r16 += CNT(0, t)++; r32 += CNT(1, t)++; r62 += CNT(2, t)++; }
        if (x - 16 >= -DUST_MAXT) { const uint32_t o = TB(rel - 16); if (o != 0xFFu) r16 -= --CNT(0, o); }
        if (x - 32 >= -DUST_MAXT) { const uint32_t o = TB(rel - 32); if (o != 0xFFu) r32 -= --CNT(1, o); }
        if (x >= 0) {
            const uint32_t o = TB(rel - DUST_MAXT);          // x - 62 >= -62
            if (o != 0xFFu) r62 -= --CNT(2, o);
            if (t != 0xFFu && (r16 > 8 || r32 > 32 || r62 > 64)) bits |= 1u << (x & 31);
            if ((x & 31) == 31) { flag[(start + x) >> 5] = bits; bits = 0; }
        }
    }
#undef CNT
#undef TB
}

// Warp per 1024 positions (grid-stride): the flagged end positions of the chunk are listed in shared memory,
// every lane scans the suffixes of its positions, the ones with a suffix above 2 go to a second list, and the
// triangles of that list are evaluated.  Per-thread scratch in shared memory: the window's triplets (newest
// first) and 64 byte counters, 33 words apart so that lanes never share a bank.
#define DUST_SCR_WORDS 33
__global__ void __launch_bounds__(DUST_EVAL_THREADS) k_dust_eval(const uint8_t* __restrict__ trip, const uint32_t* __restrict__ flag,
                                                                  int64_t arena, uint32_t* __restrict__ dust,
                                                                  unsigned long long* __restrict__ counters /* [2]=flagged [3]=evaluated */) {
    __shared__ uint32_t scr_all[DUST_EVAL_THREADS * DUST_SCR_WORDS];
    __shared__ uint16_t list_all[DUST_EVAL_THREADS / 32][1024];
    __shared__ uint16_t list2_all[DUST_EVAL_THREADS / 32][1024];
    __shared__ uint8_t lo_all[DUST_EVAL_THREADS / 32][1024];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    uint8_t* tw = (uint8_t*)(scr_all + threadIdx.x * DUST_SCR_WORDS);        // 64 bytes of triplets
    uint8_t* c = tw + 64;                                                     // 64 counters
    uint32_t* cw = (uint32_t*)c;
    uint16_t* list = list_all[wib];
    uint16_t* list2 = list2_all[wib];
    uint8_t* los = lo_all[wib];
    const int64_t n_chunks = (arena + 1023) >> 10;
    unsigned long long n_flag = 0, n_eval = 0;
    for (int64_t chunk = (int64_t)blockIdx.x * (DUST_EVAL_THREADS / 32) + wib; chunk < n_chunks;
         chunk += (int64_t)gridDim.x * (DUST_EVAL_THREADS / 32)) {
        const int64_t wi = chunk * 32 + lane;
        uint32_t f = wi < arena / 32 ? __ldg(flag + wi) : 0u;
        const int cntf = __popc(f);
        int incl = cntf;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        if (!total) continue;
        __syncwarp();
        for (int slot = incl - cntf; f; f &= f - 1) list[slot++] = (uint16_t)(lane * 32 + __ffs(f) - 1);
        __syncwarp();
        n_flag += (lane == 0) ? total : 0;
        // ---- pass 1: suffix scans
        int n2 = 0;
        for (int base = 0; base < total; base += 32) {
            const int k = base + lane;
            int lo = 0;
            if (k < total) {
                const int64_t e = chunk * 1024 + list[k];
                int nwin = 0;
                for (; nwin < DUST_MAXT && e - nwin >= 0; nwin++) {
                    const uint8_t t = trip[e - nwin];
                    if (t == 0xFFu) break;
                    tw[nwin] = t;
                }
#pragma unroll
                for (int i = 0; i < 16; i++) cw[i] = 0u;
                int32_t r = 0;
                for (int x = 0; x < nwin; x++) {
                    r += c[tw[x]]++;
                    if (r > 2 * x) lo = x;       // l - 1 = x; x = 0 never qualifies (r = 0)
                }
            }
            const unsigned q = __ballot_sync(0xffffffffu, lo > 0);
            if (lo > 0) {
                const int slot = n2 + __popc(q & ((1u << lane) - 1u));
                list2[slot] = list[k];
                los[slot] = (uint8_t)lo;
            }
            n2 += __popc(q);
        }
        __syncwarp();
        n_eval += (lane == 0) ? n2 : 0;
        // ---- pass 2: triangles.  Rows i = e - xi, xi = 1 .. lo: best (r, l-1) of any interval inside [i, e]
        for (int base = 0; base < n2; base += 32) {
            const int k = base + lane;
            if (k < n2) {
                const int64_t e = chunk * 1024 + list2[k];
                const int lo = los[k];
                for (int x = 0; x <= lo; x++) tw[x] = trip[e - x];
                int32_t best_r = 0, best_l = 1, far = -1;
                for (int xi = 1; xi <= lo; xi++) {
#pragma unroll
                    for (int i = 0; i < 16; i++) cw[i] = 0u;
                    int32_t r = 0;
                    for (int xj = xi; xj >= 0; xj--) {
                        r += c[tw[xj]]++;
                        const int32_t l1 = xi - xj;                       // (#triplets of [i, j]) - 1
                        if (l1 > 0 && r * best_l > best_r * l1) { best_r = r; best_l = l1; }
                    }
                    if (r > 2 * xi && r * best_l >= best_r * xi) far = xi;
                }
                if (far > 0) {
                    const int64_t b0 = e - far, b1 = e + 3;               // bases [b0, b1)
                    for (int64_t w = b0 >> 5; w <= (b1 - 1) >> 5; w++) {
                        const int64_t lo_b = max(b0, w << 5), hi_b = min(b1, (w + 1) << 5);
                        const uint32_t m = (hi_b - lo_b == 32) ? 0xFFFFFFFFu : (((1u << (hi_b - lo_b)) - 1u) << (lo_b & 31));
                        atomicOr(dust + w, m);
                    }
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0) {
        if (n_flag) atomicAdd(counters + 2, n_flag);
        if (n_eval) atomicAdd(counters + 3, n_eval);
    }
}

// smask = mask | dust (in place over `dust`), counting the bases DUST adds
__global__ void k_dust_or(const uint32_t* __restrict__ mask, uint32_t* __restrict__ smask, int64_t words,
                          unsigned long long* __restrict__ counters /* [4]=DUST-only bases */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t extra = 0;
    if (i < words) {
        const uint32_t m = mask[i], d = smask[i];
        extra = __popc(d & ~m);
        smask[i] = m | d;
    }
    for (int o = 16; o; o >>= 1) extra += __shfl_down_sync(0xffffffffu, extra, o);
    if ((threadIdx.x & 31) == 0 && extra) atomicAdd(counters + 4, (unsigned long long)extra);
}

// per-transcript reversal of a bitmap (the mask half of k_revcomp)
__global__ void k_revmask(const uint32_t* __restrict__ mask, const int32_t* __restrict__ aoff, const int32_t* __restrict__ len,
                          int32_t n_seq, uint32_t* __restrict__ mrc) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < n_seq; t += nwarps) {
        const int64_t off = aoff[t];
        const int32_t n = len[t];
        const int32_t nm = (n + 31) >> 5;
        for (int32_t w = lane; w < nm; w += 32) {
            const int32_t q = n - 32 * w - 32;
            uint32_t x = fetchmask32(mask, off + max(q, 0));
            int v = 32;
            if (q < 0) { v = 32 + q; x <<= (32 - v); }
            uint32_t r = __brev(x);
            if (v < 32) r |= ~((1u << v) - 1u);
            mrc[(off >> 5) + w] = r;
        }
    }
}
