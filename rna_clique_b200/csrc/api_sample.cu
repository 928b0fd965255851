// api_sample.cu — context lifetime and rcq_sample_* entry points of the C ABI (include/rcq.h).
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include "sample.cuh"

thread_local std::string g_rcq_err;

int rcq_sample_build(rcq_ctx* ctx, rcq_sample* s, const uint8_t* d_ascii, int64_t ascii_len,
                     const int64_t* d_starts);  // sample.cu

// small per-sample arrays out of mapped pinned host memory (no copy engine, see common.cuh)
static __global__ void k_unstage(const char* __restrict__ h, int32_t n, int64_t* __restrict__ starts,
                                 int32_t* __restrict__ gidx, int32_t* __restrict__ off, int32_t* __restrict__ len,
                                 int32_t* __restrict__ gene, int32_t* __restrict__ iso) {
    const int64_t* hs = (const int64_t*)h;
    const int32_t* hi = (const int32_t*)(h + 8 * (size_t)n);
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        starts[i] = hs[i];
        gidx[i] = hi[i]; off[i] = hi[(size_t)n + i]; len[i] = hi[2 * (size_t)n + i];
        gene[i] = hi[3 * (size_t)n + i]; iso[i] = hi[4 * (size_t)n + i];
    }
}

extern "C" {

const char* rcq_last_error(void) { return g_rcq_err.c_str(); }
int rcq_version(void) { return 1; }

int rcq_ctx_create(int device, rcq_ctx** out) {
    if (!out) return rcq_fail(RCQ_ERR_ARG, "rcq_ctx_create: out is NULL");
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return rcq_fail(RCQ_ERR_ARG, "rcq_ctx_create: no such device");
    CU(cudaSetDevice(device));
    rcq_ctx* c = new rcq_ctx();
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaHostAlloc((void**)&c->h_pin, 64 * sizeof(int64_t), cudaHostAllocMapped));
    CU(cudaHostGetDevicePointer((void**)&c->h_pin_dev, c->h_pin, 0));
    CU(cudaMalloc((void**)&c->d_ctr, 64 * sizeof(int64_t)));
    for (int i = 0; i < 16; i++) CU(cudaEventCreate(&c->ev[i]));
    // one explicit pool per context: blocks freed on this stream are reused by this stream at once;
    // with the device's default pool a block freed on another stream forces a fresh (slow, device-wide)
    // physical allocation whenever that stream has not caught up yet
    cudaMemPoolProps props;
    memset(&props, 0, sizeof props);
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    CU(cudaMemPoolCreate(&c->pool, &props));
    uint64_t thr = UINT64_MAX;  // keep freed blocks in the pool: no OS round trips per pair
    CU(cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &thr));
    *out = c;
    return 0;
}

void rcq_ctx_destroy(rcq_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->dbg_hsp) cudaFree(c->dbg_hsp);
    if (c->gs_rows) cudaFree(c->gs_rows);
    for (void* p : c->arena_overflow) cudaFreeAsync(p, c->stream);
    if (c->arena) cudaFree(c->arena);
    if (c->gs_locks) cudaFree(c->gs_locks);
    for (int i = 0; i < 16; i++) cudaEventDestroy(c->ev[i]);
    cudaFreeHost(c->h_pin);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    cudaFree(c->d_ctr);
    cudaStreamDestroy(c->stream);
    if (c->pool) cudaMemPoolDestroy(c->pool);
    delete c;
}

int rcq_ctx_sync(rcq_ctx* c) {
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}
void* rcq_ctx_stream(rcq_ctx* c) { return (void*)c->stream; }
int64_t rcq_ctx_launches(rcq_ctx* c) { return c->launches; }
int rcq_ctx_stage_times(rcq_ctx* c, double ms[8], int64_t calls[8], int reset) {
    CU(cudaStreamSynchronize(c->stream));
    stage_collect(c);
    for (int i = 0; i < 8; i++) {
        if (ms) ms[i] = c->stage_ms[i];
        if (calls) calls[i] = c->stage_calls[i];
        if (reset) { c->stage_ms[i] = 0; c->stage_calls[i] = 0; }
    }
    return 0;
}

int rcq_sample_create(rcq_ctx* ctx, const uint8_t* ascii, int64_t ascii_len, int ascii_on_device,
                      const int64_t* starts, const int32_t* lengths, const int32_t* gene,
                      const int32_t* iso, int32_t n_seq, rcq_sample** out) {
    if (!ctx || !out || n_seq < 0 || ascii_len < 0 || (n_seq > 0 && (!ascii || !starts || !lengths || !gene || !iso)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_sample_create: bad argument");
    CU(cudaSetDevice(ctx->device));
    static const bool trace = getenv("RCQ_TRACE") != nullptr;
    auto T0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        auto T1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[rcq_sample_create] %-12s %.3f ms\n", what, std::chrono::duration<double, std::milli>(T1 - T0).count());
        T0 = T1;
    };
    rcq_sample* s = new rcq_sample();
    s->ctx = ctx;
    s->n_seq = n_seq;
    // arena offsets on the host (n_seq is small): every transcript starts on a 32-base boundary
    std::vector<int32_t> aoff(n_seq > 0 ? n_seq : 1), len(n_seq > 0 ? n_seq : 1);
    int64_t pos = 64, total = 0;
    int32_t max_len = 0;
    for (int32_t t = 0; t < n_seq; t++) {
        int64_t l = lengths[t];
        if (l < 0 || l > 1000000 || starts[t] < 0 || starts[t] + l > ascii_len) {
            delete s;
            return rcq_fail(RCQ_ERR_ARG, "rcq_sample_create: transcript outside ascii[] or longer than 1,000,000 bases");
        }
        aoff[t] = (int32_t)pos;
        len[t] = (int32_t)l;
        if (l > max_len) max_len = (int32_t)l;
        total += l;
        pos = (pos + l + 31) & ~31ll;
        if (pos > 0x7ffffff0ll - 256) { delete s; return rcq_fail(RCQ_ERR_ARG, "rcq_sample_create: sample exceeds 2^31 bases"); }
    }
    lap("offsets");
    s->total = total;
    s->max_len = max_len;
    s->arena = ((pos + 63) & ~63ll) + 128;
    const uint8_t* d_ascii = ascii;
    uint8_t* own_ascii = nullptr;
    int64_t* d_starts = nullptr;
    RC(arena_reset(ctx));
    RC(talloc(ctx, &d_starts, n_seq));
    if (!ascii_on_device && n_seq) {
        RC(talloc(ctx, &own_ascii, ascii_len + 64));
        CU(cudaMemcpyAsync(own_ascii, ascii, ascii_len, cudaMemcpyHostToDevice, ctx->stream));
        d_ascii = own_ascii;
    }
    RC(dalloc(ctx, &s->off, n_seq));
    RC(dalloc(ctx, &s->len, n_seq));
    RC(dalloc(ctx, &s->gene, n_seq));
    RC(dalloc(ctx, &s->iso, n_seq));
    RC(dalloc(ctx, &s->gidx, n_seq));
    std::vector<int32_t> gidx(n_seq > 0 ? n_seq : 1);
    {
        // dense rank of every gene id among the sample's distinct ids: counting pass when the ids are
        // small non-negative integers (rnaSPAdes gene numbers), sort otherwise
        int32_t gmin = 0, gmax = -1;
        for (int32_t t = 0; t < n_seq; t++) {
            if (t == 0 || gene[t] < gmin) gmin = gene[t];
            if (t == 0 || gene[t] > gmax) gmax = gene[t];
        }
        if (n_seq > 0 && gmin >= 0 && (int64_t)gmax < 8ll * n_seq + 1024) {
            std::vector<int32_t> rank((size_t)gmax + 2, 0);
            for (int32_t t = 0; t < n_seq; t++) rank[gene[t]] = 1;
            int32_t acc = 0;
            for (int64_t g = 0; g <= gmax; g++) { int32_t present = rank[g]; rank[g] = acc; acc += present; }
            s->n_genes = acc;
            for (int32_t t = 0; t < n_seq; t++) gidx[t] = rank[gene[t]];
        } else {
            std::vector<int32_t> u(gene, gene + n_seq);
            std::sort(u.begin(), u.end());
            u.erase(std::unique(u.begin(), u.end()), u.end());
            s->n_genes = (int32_t)u.size();
            for (int32_t t = 0; t < n_seq; t++)
                gidx[t] = (int32_t)(std::lower_bound(u.begin(), u.end(), gene[t]) - u.begin());
        }
    }
    lap("gene index");
    if (n_seq) {
        // starts | gidx | off | len | gene | iso through mapped pinned staging, read by one kernel
        const size_t n = (size_t)n_seq;
        RC(stage_reserve(ctx, 28 * n));
        char* h = ctx->h_stage;
        memcpy(h, starts, 8 * n);
        memcpy(h + 8 * n, gidx.data(), 4 * n);
        memcpy(h + 12 * n, aoff.data(), 4 * n);
        memcpy(h + 16 * n, len.data(), 4 * n);
        memcpy(h + 20 * n, gene, 4 * n);
        memcpy(h + 24 * n, iso, 4 * n);
        k_unstage<<<RCQ_SMS, 256, 0, ctx->stream>>>(ctx->h_stage_dev, n_seq, d_starts, s->gidx, s->off, s->len, s->gene, s->iso);
        CU(cudaGetLastError());
        ctx->launches++;
    }
    lap("h2d small");
    stage_tic(ctx, 0);
    int rc = rcq_sample_build(ctx, s, d_ascii, ascii_len, d_starts);
    stage_toc(ctx, 0);
    lap("build launch");
    // the host staging vectors (aoff, len) must outlive the async copies
    cudaStreamSynchronize(ctx->stream);
    stage_collect(ctx);
    lap("final sync");
    if (rc) { rcq_sample_free(s); return rc; }
    *out = s;
    return 0;
}

void rcq_sample_free(rcq_sample* s) {
    if (!s) return;
    rcq_ctx* c = s->ctx;
    cudaSetDevice(c->device);
    dfree(c, s->fwd); dfree(c, s->rc); dfree(c, s->mask); dfree(c, s->mask_rc);
    dfree(c, s->off); dfree(c, s->len); dfree(c, s->gene); dfree(c, s->iso);
    dfree(c, s->index); dfree(c, s->fctx); dfree(c, s->dir); dfree(c, s->sel); dfree(c, s->gidx);
    dfree(c, s->blk2tx); dfree(c, s->ol); dfree(c, s->smask); dfree(c, s->smask_rc);
    delete s;
}

int rcq_sample_info(const rcq_sample* s, int64_t info[5]) {
    if (!s || !info) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_info: bad argument");
    info[0] = s->n_seq; info[1] = s->total; info[2] = s->arena; info[3] = s->n_index; info[4] = s->n_masked;
    return 0;
}

// ---- multi-GPU: a built sample as one flat device buffer (the rank that built it broadcasts it) ----
// meta: [0]=n_seq [1]=total [2]=arena [3]=n_index [4]=n_masked [5]=max_len [6]=n_genes [7]=flat bytes
struct FlatPart { void** ptr; int64_t bytes; };
static int flat_parts(rcq_sample* s, FlatPart* out) {
    const int64_t w = s->arena / 16 * 4, m = s->arena / 32 * 4, n = (int64_t)s->n_seq * 4;
    int k = 0;
    out[k++] = {(void**)&s->fwd, w}; out[k++] = {(void**)&s->rc, w};
    out[k++] = {(void**)&s->mask, m}; out[k++] = {(void**)&s->mask_rc, m};
    out[k++] = {(void**)&s->sel, m}; out[k++] = {(void**)&s->blk2tx, m};
    out[k++] = {(void**)&s->smask, m}; out[k++] = {(void**)&s->smask_rc, m};
    out[k++] = {(void**)&s->off, n}; out[k++] = {(void**)&s->len, n}; out[k++] = {(void**)&s->gene, n};
    out[k++] = {(void**)&s->iso, n}; out[k++] = {(void**)&s->gidx, n}; out[k++] = {(void**)&s->ol, 2 * n};
    out[k++] = {(void**)&s->index, s->n_index * 8}; out[k++] = {(void**)&s->fctx, s->n_index * 8};
    out[k++] = {(void**)&s->dir, ((1ll << RCQ_DIRBITS) + 1) * 4};
    return k;
}
static inline int64_t pad256(int64_t b) { return (b + 255) & ~255ll; }

int rcq_sample_meta(const rcq_sample* s, int64_t meta[8]) {
    if (!s || !meta) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_meta: bad argument");
    meta[0] = s->n_seq; meta[1] = s->total; meta[2] = s->arena; meta[3] = s->n_index; meta[4] = s->n_masked;
    meta[5] = s->max_len; meta[6] = s->n_genes;
    FlatPart parts[24];
    const int k = flat_parts(const_cast<rcq_sample*>(s), parts);
    int64_t tot = 0;
    for (int i = 0; i < k; i++) tot += pad256(parts[i].bytes);
    meta[7] = tot;
    return 0;
}

int rcq_sample_export(rcq_ctx* ctx, const rcq_sample* s, uint8_t* dev_flat) {
    if (!ctx || !s || !dev_flat) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_export: bad argument");
    CU(cudaSetDevice(ctx->device));
    FlatPart parts[24];
    const int k = flat_parts(const_cast<rcq_sample*>(s), parts);
    int64_t at = 0;
    for (int i = 0; i < k; i++) {
        if (parts[i].bytes) CU(cudaMemcpyAsync(dev_flat + at, *parts[i].ptr, parts[i].bytes, cudaMemcpyDeviceToDevice, ctx->stream));
        at += pad256(parts[i].bytes);
    }
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rcq_sample_import(rcq_ctx* ctx, const int64_t meta[8], const uint8_t* dev_flat, rcq_sample** out) {
    if (!ctx || !meta || !dev_flat || !out) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_import: bad argument");
    CU(cudaSetDevice(ctx->device));
    rcq_sample* s = new rcq_sample();
    s->ctx = ctx;
    s->n_seq = (int32_t)meta[0]; s->total = meta[1]; s->arena = meta[2]; s->n_index = meta[3]; s->n_masked = meta[4];
    s->max_len = (int32_t)meta[5]; s->n_genes = (int32_t)meta[6];
    FlatPart parts[24];
    const int k = flat_parts(s, parts);
    int64_t at = 0;
    for (int i = 0; i < k; i++) {
        char* p = nullptr;
        int rc = dalloc(ctx, &p, parts[i].bytes);
        if (rc) { rcq_sample_free(s); return rc; }
        *parts[i].ptr = p;
        if (parts[i].bytes) CU(cudaMemcpyAsync(p, dev_flat + at, parts[i].bytes, cudaMemcpyDeviceToDevice, ctx->stream));
        at += pad256(parts[i].bytes);
    }
    if (at != meta[7]) { rcq_sample_free(s); return rcq_fail(RCQ_ERR_ARG, "rcq_sample_import: meta does not match the flat size"); }
    CU(cudaStreamSynchronize(ctx->stream));
    *out = s;
    return 0;
}

int rcq_ctx_set_dust(rcq_ctx* ctx, int on) {
    if (!ctx) return 0;
    const int before = ctx->dust_on;
    ctx->dust_on = on ? 1 : 0;
    return before;
}

int rcq_sample_seed_mask(const rcq_sample* s, uint32_t* smask, int64_t* n_dust) {
    if (!s) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_seed_mask: bad argument");
    rcq_ctx* c = s->ctx;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    if (smask) CU(cudaMemcpy(smask, s->smask, s->arena / 32 * 4, cudaMemcpyDeviceToHost));
    if (n_dust) *n_dust = s->n_dust;
    return 0;
}

int rcq_sample_download(const rcq_sample* s, uint32_t* fwd, uint32_t* rc, uint32_t* mask,
                        int64_t* arena_off, uint64_t* index) {
    if (!s) return rcq_fail(RCQ_ERR_ARG, "rcq_sample_download: bad argument");
    rcq_ctx* c = s->ctx;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    if (fwd) CU(cudaMemcpy(fwd, s->fwd, s->arena / 16 * 4, cudaMemcpyDeviceToHost));
    if (rc) CU(cudaMemcpy(rc, s->rc, s->arena / 16 * 4, cudaMemcpyDeviceToHost));
    if (mask) CU(cudaMemcpy(mask, s->mask, s->arena / 32 * 4, cudaMemcpyDeviceToHost));
    if (arena_off && s->n_seq) {
        std::vector<int32_t> t(s->n_seq);
        CU(cudaMemcpy(t.data(), s->off, s->n_seq * 4, cudaMemcpyDeviceToHost));
        for (int i = 0; i < s->n_seq; i++) arena_off[i] = t[i];
    }
    if (index && s->n_index) CU(cudaMemcpy(index, s->index, s->n_index * 8, cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"
