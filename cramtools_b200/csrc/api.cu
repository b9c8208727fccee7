// api.cu -- context management and the host-memory entry points of the C ABI (include/cramgpu.h).
#include "internal.cuh"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

static thread_local std::string g_init_error;

int ctx_fail(cramgpu_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_init_error = buf;
    return code;
}

int ctx_cuda(cramgpu_ctx *ctx, cudaError_t e, const char *what)
{
    return ctx_fail(ctx, CRAMGPU_ERR_CUDA, "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
}

int pool_reserve(cramgpu_ctx *ctx, DevPool &pool, size_t bytes)
{
    if (bytes <= pool.cap) return CRAMGPU_OK;
    // the old buffer may still be in use by enqueued work
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    if (pool.p) CUDA_TRY(ctx, cudaFree(pool.p));
    pool.p = nullptr; pool.cap = 0;
    size_t want = rup(bytes + bytes / 8, 1 << 20);
    cudaError_t e = cudaMalloc(&pool.p, want);
    if (e != cudaSuccess) { want = rup(bytes, 1 << 20); e = cudaMalloc(&pool.p, want); }
    if (e != cudaSuccess) return ctx_cuda(ctx, e, "cudaMalloc(scratch)");
    pool.cap = want;
    return CRAMGPU_OK;
}

int pinned_reserve(cramgpu_ctx *ctx, void **p, size_t *cap, size_t bytes)
{
    if (bytes <= *cap) return CRAMGPU_OK;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (*p) CUDA_TRY(ctx, cudaFreeHost(*p));
    *p = nullptr; *cap = 0;
    size_t want = rup(bytes + bytes / 8, 1 << 16);
    CUDA_TRY(ctx, cudaHostAlloc(p, want, cudaHostAllocDefault));
    *cap = want;
    return CRAMGPU_OK;
}

static cudaEvent_t ev_get(cramgpu_ctx *ctx)
{
    if (!ctx->ev_pool.empty()) { cudaEvent_t e = ctx->ev_pool.back(); ctx->ev_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
void prof_begin(cramgpu_ctx *ctx, cudaStream_t st, const char *name)
{
    cramgpu_ctx::Span s{name, ev_get(ctx), ev_get(ctx)};
    cudaEventRecord(s.a, st);
    ctx->spans.push_back(s);
}
void prof_end(cramgpu_ctx *ctx, cudaStream_t st) { cudaEventRecord(ctx->spans.back().b, st); }

// ---- host-side copies --------------------------------------------------------------------------------------
// Staging copies between caller memory and the context's pinned buffers run on a few threads (one core moves
// ~8 GB/s, PCIe 5 x16 moves ~50 GB/s); memory the caller already pinned (cudaHostAlloc / cudaHostRegister) is
// handed to the DMA engine directly, without staging.
#include <thread>

cudaError_t memcpy_pieces(void *dst, const void *src, size_t bytes, cudaMemcpyKind kind, cudaStream_t st)
{
    const size_t P = 16u << 20;
    const int WINDOW = 3;
    if (bytes <= (size_t)(WINDOW + 1) * P) {
        for (size_t o = 0; o < bytes; o += P) {
            cudaError_t e = cudaMemcpyAsync((uint8_t *)dst + o, (const uint8_t *)src + o, std::min(P, bytes - o), kind, st);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
    // A long copy is PACED: at most WINDOW pieces are handed to the copy engine ahead of the one that is moving.  The engine
    // serves what it has been given in order, so a multi-gigabyte copy enqueued at once makes every other context's copy --
    // however small -- wait for all of it (measured: the 0.4 GB of compressed blocks of one context waited 60 ms behind the
    // 3.3 GB of reads of another).  The calling thread waits here instead; the host entry points are synchronous anyway.
    cudaEvent_t ev[WINDOW];
    for (int k = 0; k < WINDOW; k++) {
        cudaError_t e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming | cudaEventBlockingSync);
        if (e != cudaSuccess) { for (int j = 0; j < k; j++) cudaEventDestroy(ev[j]); return e; }
    }
    cudaError_t rc = cudaSuccess;
    size_t i = 0;
    for (size_t o = 0; o < bytes && rc == cudaSuccess; o += P, i++) {
        if (i >= (size_t)WINDOW) rc = cudaEventSynchronize(ev[i % WINDOW]);
        if (rc == cudaSuccess) rc = cudaMemcpyAsync((uint8_t *)dst + o, (const uint8_t *)src + o, std::min(P, bytes - o), kind, st);
        if (rc == cudaSuccess) rc = cudaEventRecord(ev[i % WINDOW], st);
    }
    for (int k = 0; k < WINDOW; k++) cudaEventDestroy(ev[k]);
    return rc;
}

bool host_is_pinned(const void *p)
{
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// cuPointerGetAttribute through the runtime's driver entry point lookup (no link-time dependency on libcuda)
#include <cuda.h>
typedef CUresult (*PointerAttrFn)(void *, CUpointer_attribute, CUdeviceptr);
static PointerAttrFn pointer_attr_fn()
{
    static PointerAttrFn fn = []() -> PointerAttrFn {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return (PointerAttrFn)f;
    }();
    return fn;
}

MemKind MemMap::kind(const void *p)
{
    const uintptr_t a = (uintptr_t)p;
    if (!p) return MEM_PAGEABLE;
    if (last < ranges.size() && a >= ranges[last].lo && a < ranges[last].hi) return ranges[last].kind;
    for (size_t i = 0; i < ranges.size(); i++)
        if (a >= ranges[i].lo && a < ranges[i].hi) { last = i; return ranges[i].kind; }
    cudaPointerAttributes at;
    MemKind k = MEM_PAGEABLE;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) cudaGetLastError();
    else if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) k = MEM_DEVICE;
    else if (at.type == cudaMemoryTypeHost) k = MEM_PINNED;
    Range r{a, a + 1, k};
    if (k == MEM_PAGEABLE) {
        // memory CUDA does not know: a 2 MB window of host address space never holds device memory
        r.lo = a & ~(uintptr_t)((2u << 20) - 1); r.hi = r.lo + (2u << 20);
    } else if (PointerAttrFn fn = pointer_attr_fn()) {
        CUdeviceptr start = 0; size_t size = 0;
        if (fn(&start, CU_POINTER_ATTRIBUTE_RANGE_START_ADDR, (CUdeviceptr)a) == CUDA_SUCCESS &&
            fn(&size, CU_POINTER_ATTRIBUTE_RANGE_SIZE, (CUdeviceptr)a) == CUDA_SUCCESS && size) {
            r.lo = (uintptr_t)start; r.hi = r.lo + size;
        }
    }
    ranges.push_back(r);
    last = ranges.size() - 1;
    return k;
}

void par_memcpy(std::vector<CopyJob> &jobs)
{
    size_t total = 0;
    for (auto &j : jobs) total += j.n;
    unsigned hw = std::thread::hardware_concurrency();
    unsigned nt = total < (8u << 20) ? 1 : std::min(8u, hw ? hw : 1u);
    if (nt <= 1) { for (auto &j : jobs) if (j.n) memcpy(j.dst, j.src, j.n); jobs.clear(); return; }
    // cut into pieces of <= 4 MB and deal them round-robin
    std::vector<CopyJob> pieces;
    const size_t P = 4u << 20;
    for (auto &j : jobs)
        for (size_t o = 0; o < j.n; o += P) pieces.push_back(CopyJob{(uint8_t *)j.dst + o, (const uint8_t *)j.src + o, std::min(P, j.n - o)});
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([&pieces, t, nt]() { for (size_t i = t; i < pieces.size(); i += nt) memcpy(pieces[i].dst, pieces[i].src, pieces[i].n); });
    for (auto &x : th) x.join();
    jobs.clear();
}

extern "C" {

int cramgpu_profile_enable(cramgpu_ctx *ctx, int on)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    ctx->profiling = on != 0;
    return CRAMGPU_OK;
}

// Drain the recorded spans: the caller must have synchronised the stream(s) the work ran on.
// Writes up to cap entries; returns the number of spans recorded since the last call.
int cramgpu_profile_read(cramgpu_ctx *ctx, uint32_t cap, const char **names, float *ms)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    uint32_t n = (uint32_t)ctx->spans.size();
    for (uint32_t i = 0; i < n; i++) {
        cramgpu_ctx::Span &s = ctx->spans[i];
        float t = 0.f;
        cudaEventElapsedTime(&t, s.a, s.b);
        if (i < cap) { if (names) names[i] = s.name; if (ms) ms[i] = t; }
        ctx->ev_pool.push_back(s.a); ctx->ev_pool.push_back(s.b);
    }
    ctx->spans.clear();
    return (int)n;
}

int cramgpu_init(int device, cramgpu_ctx **out)
{
    if (!out) return ctx_fail(nullptr, CRAMGPU_ERR_ARG, "null ctx pointer");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return ctx_fail(nullptr, CRAMGPU_ERR_CUDA, "no usable CUDA device (%s); libcramgpu has no CPU fallback",
                        e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= n) return ctx_fail(nullptr, CRAMGPU_ERR_ARG, "device %d out of range [0,%d)", device, n);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return ctx_cuda(nullptr, e, "cudaSetDevice");
    cramgpu_ctx *ctx = new cramgpu_ctx();
    ctx->device = device;
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_slice, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_frame, cudaEventDisableTiming);
    if (e == cudaSuccess && !getenv("CRAMGPU_NO_HP_STREAM")) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        e = cudaStreamCreateWithPriority(&ctx->hp_stream, cudaStreamNonBlocking, hi);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->hp_in, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->hp_out, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) { delete ctx; return ctx_cuda(nullptr, e, "stream/event creation"); }
    *out = ctx;
    return CRAMGPU_OK;
}

void cramgpu_destroy(cramgpu_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaDeviceSynchronize();
    if (ctx->scratch.p) cudaFree(ctx->scratch.p);
    if (ctx->dev_in.p) cudaFree(ctx->dev_in.p);
    if (ctx->dev_out.p) cudaFree(ctx->dev_out.p);
    if (ctx->rec_scratch.p) cudaFree(ctx->rec_scratch.p);
    if (ctx->rec_out.p) cudaFree(ctx->rec_out.p);
    if (ctx->rec_unp.p) cudaFree(ctx->rec_unp.p);
    if (ctx->hs_scratch.p) cudaFree(ctx->hs_scratch.p);
    if (ctx->slice_scratch.p) cudaFree(ctx->slice_scratch.p);
    if (ctx->mate_scratch.p) cudaFree(ctx->mate_scratch.p);
    if (ctx->trk_scratch.p) cudaFree(ctx->trk_scratch.p);
    if (ctx->trk_own.p) cudaFree(ctx->trk_own.p);
    if (ctx->trk_ext.p) cudaFree(ctx->trk_ext.p);
    if (ctx->pin_trk) cudaFreeHost(ctx->pin_trk);
    if (ctx->frame_scratch.p) cudaFree(ctx->frame_scratch.p);
    if (ctx->frame_out.p) cudaFree(ctx->frame_out.p);
    if (ctx->pin_frame) cudaFreeHost(ctx->pin_frame);
    if (ctx->ev_frame) cudaEventDestroy(ctx->ev_frame);
    if (ctx->enc_big.p) cudaFree(ctx->enc_big.p);
    if (ctx->unp_in.p) cudaFree(ctx->unp_in.p);
    if (ctx->bam_scratch.p) cudaFree(ctx->bam_scratch.p);
    if (ctx->bam_cigar.p) cudaFree(ctx->bam_cigar.p);
    if (ctx->bam_out.p) cudaFree(ctx->bam_out.p);
    if (ctx->bam_unp.p) cudaFree(ctx->bam_unp.p);
    if (ctx->bam_pk_scratch.p) cudaFree(ctx->bam_pk_scratch.p);
    if (ctx->bam_unp_scratch.p) cudaFree(ctx->bam_unp_scratch.p);
    if (ctx->pin_bam) cudaFreeHost(ctx->pin_bam);
    if (ctx->slice_in.p) cudaFree(ctx->slice_in.p);
    if (ctx->pin_slice) cudaFreeHost(ctx->pin_slice);
    if (ctx->ev_slice) cudaEventDestroy(ctx->ev_slice);
    if (ctx->hs_out.p) cudaFree(ctx->hs_out.p);
    if (ctx->hs_in.p) cudaFree(ctx->hs_in.p);
    if (ctx->pin_hs) cudaFreeHost(ctx->pin_hs);
    if (ctx->pin_hs2) cudaFreeHost(ctx->pin_hs2);
    if (ctx->pin_rec) cudaFreeHost(ctx->pin_rec);
    free(ctx->rec_fault);
    if (ctx->pin_feat) cudaFreeHost(ctx->pin_feat);
    if (ctx->pin_in) cudaFreeHost(ctx->pin_in);
    if (ctx->pin_out) cudaFreeHost(ctx->pin_out);
    if (ctx->pin_meta) cudaFreeHost(ctx->pin_meta);
    if (ctx->ev) cudaEventDestroy(ctx->ev);
    if (ctx->hp_in) cudaEventDestroy(ctx->hp_in);
    if (ctx->hp_out) cudaEventDestroy(ctx->hp_out);
    if (ctx->hp_stream) cudaStreamDestroy(ctx->hp_stream);
    for (auto &s : ctx->spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
    for (auto e : ctx->ev_pool) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *cramgpu_last_error(const cramgpu_ctx *ctx)
{
    return ctx ? ctx->err.c_str() : g_init_error.c_str();
}

int cramgpu_sync(cramgpu_ctx *ctx)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CRAMGPU_OK;
}

uint64_t cramgpu_launch_count(const cramgpu_ctx *ctx) { return ctx ? ctx->launches : 0; }

int cramgpu_set_option(cramgpu_ctx *ctx, const char *name, uint64_t value)
{
    if (!ctx || !name) return CRAMGPU_ERR_ARG;
    if (!strcmp(name, "dec_wide_pool_blocks")) { ctx->dec_wide_pool_blocks = (uint32_t)std::min<uint64_t>(value, 1u << 20); return CRAMGPU_OK; }
    return ctx_fail(ctx, CRAMGPU_ERR_ARG, "unknown option %s", name);
}

int cramgpu_copy_to_host(cramgpu_ctx *ctx, void *dst_host, const void *src_dev, uint64_t bytes, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!bytes) return CRAMGPU_OK;
    if (!dst_host || !src_dev) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    CUDA_TRY(ctx, cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CRAMGPU_OK;
}

int cramgpu_rans_encode_dev(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                            const uint32_t *in_len, const uint8_t *order, uint8_t *d_out, uint64_t out_cap,
                            uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!d_out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return rans_encode_device(ctx, nblocks, d_in, in_off, in_len, order, d_out, out_cap, d_out_off, d_out_len, d_status,
                              stream ? (cudaStream_t)stream : ctx->stream);
}

int cramgpu_rans_decode_dev(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                            const uint32_t *in_len, const uint8_t *in_prefix, uint8_t *d_out, uint64_t out_cap,
                            uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return rans_decode_device(ctx, nblocks, d_in, in_off, in_len, in_prefix, d_out, out_cap, d_out_off, d_out_len, d_status,
                              stream ? (cudaStream_t)stream : ctx->stream);
}

int cramgpu_rans_raw_sizes(uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len, uint32_t *raw_len,
                           int32_t *status)
{
    int first = CRAMGPU_OK;
    for (uint32_t i = 0; i < nblocks; i++) {
        uint32_t ord;
        int rc = rans_parse_prefix(in[i], in_len[i], &ord, &raw_len[i]);
        if (rc != CRAMGPU_OK) raw_len[i] = 0;
        if (status) status[i] = rc;
        if (rc != CRAMGPU_OK && first == CRAMGPU_OK) first = rc;
    }
    return first;
}

static const char *status_text(int rc)
{
    switch (rc) {
    case CRAMGPU_ERR_INPUT_LENGTH: return "Incorrect input length.";
    case CRAMGPU_ERR_OUT_TOO_SMALL: return "Output buffer too small";
    case CRAMGPU_ERR_UNKNOWN_ORDER: return "Unknown rANS order";
    case CRAMGPU_ERR_CORRUPT: return "Corrupt rANS block";
    default: return "error";
    }
}

// Shared by the two host entry points.  Inputs: blocks in device memory are coded where they are (16-byte aligned ones;
// others are copied next to the staged blocks); large blocks in memory the caller pinned go to the device directly; the
// rest is staged in pinned memory, every run of consecutive staged blocks travelling as one copy.  Results the other way
// round; the decoder writes blocks whose destination is device memory straight to that address.
static int host_batch(cramgpu_ctx *ctx, bool encode, uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len,
                      const uint8_t *order, uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len,
                      int32_t *status)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (nblocks == 0) return CRAMGPU_OK;
    if (!in || !in_len || !out || !out_cap || !out_len || (encode && !order)) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;

    const uint32_t BIG = 64u << 10;
    enum { STAGED = 0, PINNED = 1, IN_PLACE = 2, DEV_COPY = 3 };
    MemMap mm;
    std::vector<uint64_t> off(nblocks), poff(nblocks), addr(nblocks);      // place in the device input buffer / in pinned staging / final device address
    std::vector<uint8_t> how(nblocks, STAGED);
    std::vector<uint8_t> prefix(encode ? 0 : 9 * (size_t)nblocks, 0);
    std::vector<uint64_t> dst_addr(encode ? 0 : nblocks, 0);               // decoder: device destinations
    uint64_t total_in = 0, total_pin = 0, total_out_cap = 0;
    for (uint32_t i = 0; i < nblocks; i++) {
        if (in_len[i] && !in[i]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "in[%u] is null", i);
        off[i] = total_in; poff[i] = total_pin;
        if (!in_len[i]) continue;
        const MemKind k = mm.kind(in[i]);
        if (k == MEM_DEVICE) {
            if (!encode) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "in[%u]: compressed blocks in device memory go through cramgpu_rans_decode_dev", i);
            how[i] = ((uintptr_t)in[i] & 15) ? DEV_COPY : IN_PLACE;
        } else if (k == MEM_PINNED && in_len[i] >= BIG) how[i] = PINNED;
        if (how[i] != IN_PLACE) total_in += rup(in_len[i], 16);
        if (how[i] == STAGED) total_pin += rup(in_len[i], 16);
        if (!encode) {
            memcpy(&prefix[9 * (size_t)i], in[i], in_len[i] < 9 ? in_len[i] : 9);
            uint32_t ord, raw;
            if (rans_parse_prefix(in[i], in_len[i], &ord, &raw) == CRAMGPU_OK) {
                if (out[i] && mm.kind(out[i]) == MEM_DEVICE) dst_addr[i] = (uint64_t)(uintptr_t)out[i];
                else total_out_cap += rup(raw, 16);
            }
        }
    }
    // results header: out_off (nblocks+1 u64) | out_len (u32) | status (i32)
    size_t hdr_bytes = 8 * (size_t)(nblocks + 1) + 8 * (size_t)nblocks;
    if (int rc = pinned_reserve(ctx, &ctx->pin_in, &ctx->pin_in_cap, total_pin + 16)) return rc;
    if (int rc = pool_reserve(ctx, ctx->dev_in, total_in + hdr_bytes + 64)) return rc;
    uint8_t *pin = (uint8_t *)ctx->pin_in;
    uint8_t *d_in = (uint8_t *)ctx->dev_in.p;
    uint8_t *d_hdr = d_in + rup(total_in, 16) + 16;
    uint64_t *d_out_off = (uint64_t *)d_hdr;
    uint32_t *d_out_len = (uint32_t *)(d_hdr + 8 * (size_t)(nblocks + 1));
    int32_t *d_status = (int32_t *)(d_hdr + 8 * (size_t)(nblocks + 1) + 4 * (size_t)nblocks);
    {
        std::vector<CopyJob> jobs;
        for (uint32_t i = 0; i < nblocks; i++) {
            addr[i] = how[i] == IN_PLACE ? (uint64_t)(uintptr_t)in[i] : (uint64_t)(uintptr_t)(d_in + off[i]);
            if (in_len[i] && how[i] == STAGED) jobs.push_back(CopyJob{pin + poff[i], in[i], in_len[i]});
        }
        par_memcpy(jobs);
        uint32_t i = 0;
        while (i < nblocks) {
            if (how[i] == IN_PLACE || !in_len[i]) { i++; continue; }
            if (how[i] == PINNED) { CUDA_TRY(ctx, memcpy_pieces(d_in + off[i], in[i], in_len[i], cudaMemcpyHostToDevice, st)); i++; continue; }
            if (how[i] == DEV_COPY) { CUDA_TRY(ctx, cudaMemcpyAsync(d_in + off[i], in[i], in_len[i], cudaMemcpyDeviceToDevice, st)); i++; continue; }
            // a run of staged blocks (blocks coded in place and empty ones take no room): contiguous in both buffers
            uint32_t j = i;
            uint64_t end = 0;
            while (j < nblocks && (how[j] == STAGED || how[j] == IN_PLACE || !in_len[j])) { if (how[j] == STAGED && in_len[j]) end = off[j] + in_len[j]; j++; }
            if (end > off[i]) CUDA_TRY(ctx, memcpy_pieces(d_in + off[i], pin + poff[i], end - off[i], cudaMemcpyHostToDevice, st));
            i = j;
        }
    }

    if (int rc = pinned_reserve(ctx, &ctx->pin_out, &ctx->pin_out_cap, hdr_bytes + 64)) return rc;
    if (encode) {
        if (int rc = rans_encode_device(ctx, nblocks, nullptr, addr.data(), in_len, order, nullptr, ~0ull, d_out_off, d_out_len,
                                        d_status, st))
            return rc;
        // the packed size is only known on the device: fetch it, then size the output buffers
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pin_out, d_out_off + nblocks, 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
        total_out_cap = *(uint64_t *)ctx->pin_out;
        if (int rc = pool_reserve(ctx, ctx->dev_out, total_out_cap + 64)) return rc;
        if (int rc = rans_encode_pack(ctx, (uint8_t *)ctx->dev_out.p, st)) return rc;
    } else {
        if (int rc = pool_reserve(ctx, ctx->dev_out, total_out_cap + 64)) return rc;
        if (int rc = rans_decode_device(ctx, nblocks, nullptr, addr.data(), in_len, prefix.data(), (uint8_t *)ctx->dev_out.p,
                                        total_out_cap, d_out_off, d_out_len, d_status, st, dst_addr.data(), out_cap))
            return rc;
    }
    if (int rc = pinned_reserve(ctx, &ctx->pin_out, &ctx->pin_out_cap, hdr_bytes + 64)) return rc;
    uint8_t *pout = (uint8_t *)ctx->pin_out;
    CUDA_TRY(ctx, cudaMemcpyAsync(pout, d_hdr, hdr_bytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    // the header moves aside: pin_out may be re-allocated for the staged results
    std::vector<uint8_t> hdr(pout, pout + hdr_bytes);
    const uint64_t *h_off = (const uint64_t *)hdr.data();
    const uint32_t *h_len = (const uint32_t *)(hdr.data() + 8 * (size_t)(nblocks + 1));
    const int32_t *h_status = (const int32_t *)(hdr.data() + 8 * (size_t)(nblocks + 1) + 4 * (size_t)nblocks);

    // results: 0 = through the staging buffer, 1 = straight into pinned caller memory, 2 = device copy, 3 = already in place
    const uint8_t *d_out = (const uint8_t *)ctx->dev_out.p;
    std::vector<uint8_t> dest(nblocks, 0);
    std::vector<uint64_t> pso(nblocks, 0);
    uint64_t total_ps = 0;
    for (uint32_t i = 0; i < nblocks; i++) {
        const bool okb = h_status[i] == CRAMGPU_OK && h_len[i] && h_len[i] <= out_cap[i] && out[i];
        if (!okb) { dest[i] = 0; continue; }
        if (!encode && dst_addr[i]) { dest[i] = 3; continue; }
        const MemKind k = mm.kind(out[i]);
        dest[i] = k == MEM_DEVICE ? 2 : (k == MEM_PINNED && h_len[i] >= BIG) ? 1 : 0;
        if (dest[i] == 0) { pso[i] = total_ps; total_ps += rup(h_len[i], 16); }
    }
    if (int rc = pinned_reserve(ctx, &ctx->pin_out, &ctx->pin_out_cap, total_ps + 64)) return rc;
    uint8_t *pdata = (uint8_t *)ctx->pin_out;
    {
        uint32_t i = 0;
        while (i < nblocks) {
            const bool okb = h_status[i] == CRAMGPU_OK && h_len[i] && h_len[i] <= out_cap[i] && out[i];
            if (!okb || dest[i] == 3) { i++; continue; }
            if (dest[i]) {
                CUDA_TRY(ctx, cudaMemcpyAsync(out[i], d_out + h_off[i], h_len[i], dest[i] == 2 ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
                i++;
                continue;
            }
            // a run of staged results that are neighbours on the device (16-byte aligned back to back there and in staging)
            uint32_t j = i + 1;
            uint64_t end = h_off[i] + h_len[i];
            while (j < nblocks && dest[j] == 0 && h_status[j] == CRAMGPU_OK && h_len[j] && h_len[j] <= out_cap[j] && out[j] &&
                   h_off[j] == rup(end, 16) && pso[j] == pso[i] + (h_off[j] - h_off[i])) { end = h_off[j] + h_len[j]; j++; }
            CUDA_TRY(ctx, memcpy_pieces(pdata + pso[i], d_out + h_off[i], end - h_off[i], cudaMemcpyDeviceToHost, st));
            i = j;
        }
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    int first = CRAMGPU_OK;
    std::vector<CopyJob> jobs;
    for (uint32_t i = 0; i < nblocks; i++) {
        int rc = h_status[i];
        out_len[i] = 0;
        if (rc == CRAMGPU_OK) {
            if (h_len[i] > out_cap[i]) rc = CRAMGPU_ERR_OUT_TOO_SMALL;
            else {
                if (h_len[i] && dest[i] == 0 && out[i]) jobs.push_back(CopyJob{out[i], pdata + pso[i], h_len[i]});
                out_len[i] = h_len[i];
            }
        }
        if (status) status[i] = rc;
        if (rc != CRAMGPU_OK && first == CRAMGPU_OK) {
            first = rc;
            // the compressor's wording is RANS.allocateIfNeeded's [RANS.java:79-85], the decompressor's RANS.uncompress's [:44]
            if (rc == CRAMGPU_ERR_OUT_TOO_SMALL && encode) ctx_fail(ctx, rc, "Insufficient buffer size. (block %u needs %u bytes)", i, h_len[i]);
            else if (rc == CRAMGPU_ERR_OUT_TOO_SMALL) ctx_fail(ctx, rc, "Output buffer too small to fit %u bytes. (block %u)", h_len[i], i);
            else if (rc == CRAMGPU_ERR_UNKNOWN_ORDER) ctx_fail(ctx, rc, "Unknown rANS order: %d (block %u)", encode ? (int)order[i] : (int)in[i][0], i);
            else ctx_fail(ctx, rc, "%s (block %u)", status_text(rc), i);
        }
    }
    par_memcpy(jobs);
    return first;
}

int cramgpu_rans_encode_batch(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len,
                              const uint8_t *order, uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len,
                              int32_t *status)
{
    return host_batch(ctx, true, nblocks, in, in_len, order, out, out_cap, out_len, status);
}

int cramgpu_rans_decode_batch(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len,
                              uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len, int32_t *status)
{
    return host_batch(ctx, false, nblocks, in, in_len, nullptr, out, out_cap, out_len, status);
}

// ---- multi-GPU: output offsets and record counters of this rank's containers -----------------------------------------
// Containers are independent and dealt round-robin (container with global ordinal g goes to rank g % world).  What the
// writer keeps as running sums -- container.offset (CRAMContainerStreamWriter.java:229-230) and globalRecordCounter
// (ContainerFactory [ContainerFactory.java:77,105]) -- becomes an exclusive scan over the all-gathered per-container
// {compressed bytes, records, bases}.  The exchange itself is the caller's: NCCL / gloo through torch.distributed in
// cramtools_b200/sharding.py, MPI or sockets from a JVM.  Host-only.
int cramgpu_allgather_sizes(cramgpu_ctx *ctx, uint32_t n_local, const cramgpu_container_size *local, uint32_t rank, uint32_t world,
                            cramgpu_allgather_fn gather, void *user, uint64_t base_offset, uint64_t base_record,
                            uint64_t *offset, uint64_t *record_counter, cramgpu_container_size *total)
{
    if (world == 0 || rank >= world || (n_local && (!local || !offset || !record_counter)) || (world > 1 && !gather))
        return ctx ? ctx_fail(ctx, CRAMGPU_ERR_ARG, "bad argument") : CRAMGPU_ERR_ARG;
    // every rank contributes the same number of slots: the largest local count (ranks may differ by one container)
    uint64_t nmax = n_local;
    std::vector<uint64_t> counts(world, n_local);
    if (world > 1) {
        uint64_t mine = n_local;
        if (int rc = gather(user, &mine, counts.data(), sizeof mine)) return ctx ? ctx_fail(ctx, CRAMGPU_ERR_ARG, "all-gather callback failed (%d)", rc) : CRAMGPU_ERR_ARG;
        for (uint64_t c : counts) nmax = std::max(nmax, c);
    }
    std::vector<cramgpu_container_size> send(nmax), all((size_t)nmax * world);
    for (uint64_t i = 0; i < nmax; i++) send[i] = i < n_local ? local[i] : cramgpu_container_size{0, 0, 0, 0};
    if (world > 1) {
        if (nmax)
            if (int rc = gather(user, send.data(), all.data(), nmax * sizeof(cramgpu_container_size)))
                return ctx ? ctx_fail(ctx, CRAMGPU_ERR_ARG, "all-gather callback failed (%d)", rc) : CRAMGPU_ERR_ARG;
    } else all = send;
    uint64_t off = base_offset, rec = base_record, bases = 0, bytes = 0, recs = 0;
    for (uint64_t j = 0; j < nmax; j++)              // global ordinal g = j * world + r
        for (uint32_t r = 0; r < world; r++) {
            if (j >= counts[r]) continue;
            const cramgpu_container_size &c = all[(size_t)r * nmax + j];
            if (r == rank) { offset[j] = off; record_counter[j] = rec; }
            off += c.compressed_bytes; rec += c.n_records;
            bytes += c.compressed_bytes; recs += c.n_records; bases += c.n_bases;
        }
    if (total) { total->compressed_bytes = bytes; total->n_bases = bases; total->n_records = (uint32_t)std::min<uint64_t>(recs, 0xFFFFFFFFu); total->reserved = 0; }
    return CRAMGPU_OK;
}

}  // extern "C"
