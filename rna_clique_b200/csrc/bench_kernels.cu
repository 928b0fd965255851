// bench_kernels.cu — measurement helpers for bench.py (not on the product path).
//
// rcq_int32_issue_peak: the INT32 instruction-issue ceiling of this GPU, measured in the same run as the kernels
// that are held against it (k_gapped is bound by integer issue, not by memory: DESIGN.md section 4).  Every
// thread runs eight independent chains of add / xor / funnel-shift on registers; with 148 SMs x 2048 resident
// threads the schedulers always have an eligible warp, so the rate is the issue limit of the integer pipe.
// Work unit: one lane-operation = one thread executing one 32-bit integer instruction, the unit in which
// bench.py states the algorithmic work of k_gapped (12 lane-operations per extension step).
#include "common.cuh"

#define PEAK_CHAINS 8
#define PEAK_OPS_PER_ITER (3 * PEAK_CHAINS)

__global__ void __launch_bounds__(256) k_int32_issue(uint32_t* __restrict__ out, int iters, uint32_t seed) {
    uint32_t x[PEAK_CHAINS];
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; c++) x[c] = seed + threadIdx.x * 2654435761u + c * 40503u + blockIdx.x;
    const uint32_t k = seed | 1u;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < PEAK_CHAINS; c++) {
            // three dependent integer instructions per chain and iteration: IADD3, LOP3, SHF
            asm volatile("add.u32 %0, %0, %1;\n\txor.b32 %0, %0, %2;\n\tshf.l.wrap.b32 %0, %0, %0, 7;"
                         : "+r"(x[c]) : "r"(k), "r"(x[(c + 1) % PEAK_CHAINS]));
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; c++) acc ^= x[c];
    if (acc == 0x12345678u) out[0] = acc;      // keeps the chains alive; practically never taken
}

extern "C" int rcq_int32_issue_peak(rcq_ctx* ctx, int32_t reps, double* lane_ops_per_second) {
    if (!ctx || !lane_ops_per_second || reps < 1) return rcq_fail(RCQ_ERR_ARG, "rcq_int32_issue_peak: bad argument");
    CU(cudaSetDevice(ctx->device));
    uint32_t* out = nullptr;
    RC(arena_reset(ctx));
    RC(talloc(ctx, &out, 4));
    const int blocks = RCQ_SMS * 8, threads = 256, iters = 4096;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    k_int32_issue<<<blocks, threads, 0, ctx->stream>>>(out, iters, 12345u);     // warm-up
    double best = 0;
    for (int r = 0; r < reps; r++) {
        CU(cudaEventRecord(e0, ctx->stream));
        k_int32_issue<<<blocks, threads, 0, ctx->stream>>>(out, iters, 777u + r);
        CU(cudaEventRecord(e1, ctx->stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double rate = (double)blocks * threads * iters * PEAK_OPS_PER_ITER / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    CU(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *lane_ops_per_second = best;
    return 0;
}
