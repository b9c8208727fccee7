// internal.cuh -- shared declarations of libcramgpu (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <vector>
#include <algorithm>

#include "../../include/cramgpu.h"

#define TF_SHIFT 12
#define TOTFREQ 4096u
#define RANS_L (1u << 23)

#define HIST_CHUNK (128u * 1024u)   // bytes of one block handled by one histogram CTA
#define HIST1_CHUNK (512u * 1024u)  // the same for the order-1 pair counts (their shared-memory tables cost more to clear and add up)
#define PACK_CHUNK (64u * 1024u)    // bytes of one block copied by one pack CTA
#define O0_TAB_CAP 1040u            // >= 256*(1+1+2)+1
#define O1_TAB_CAP 263936u          // >= 256*(2+1025)+1, multiple of 256
#define O1_SMEM_ENTRIES 12032u      // compact (ctx x sym) histogram cells kept in shared memory per CTA

static inline uint64_t rup(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

// One rANS block of the batch, as the kernels see it.
struct BlockDesc {
    uint64_t in_off;      // offset of the raw (encode) or compressed (decode) bytes in the input buffer
    uint64_t work_off;    // encode: offset of this block's table+payload work area in scratch
    uint64_t out_off;     // decode: offset of the decoded bytes in the output buffer
    uint32_t in_len;      // encode: raw size; decode: compressed size
    uint32_t out_len;     // decode: raw size
    uint32_t work_cap;    // encode: bytes of work area (table region first, payload grows down from the end)
    uint32_t tab_cap;     // encode: bytes reserved for the table at the start of the work area
    uint32_t order;       // effective order (0 when n < 4), or 0xFF = skip (empty / rejected on host)
    uint32_t o1_index;    // rank among the order-1 blocks of this call (selects F1 / syms1 / lut1 slabs)
    uint32_t pad0, pad1;
};

// Alphabet of an order-1 block: contexts = bytes that occur, plus 0; symbols = bytes that occur.
struct O1Meta {
    uint8_t cmap[256];    // byte -> context rank
    uint8_t smap[256];    // byte -> symbol rank
    uint8_t calpha[256];  // context rank -> byte
    uint8_t salpha[256];  // symbol rank -> byte
    uint32_t nctx, nsym;
    uint64_t cell_off;    // where the block's nctx x nsym cells start in the count / coder-cell arrays (k_enc_layout)
};

struct DevPool {
    void *p = nullptr;
    size_t cap = 0;
};

struct EncodePending {     // what k_pack needs once the output buffer exists
    const BlockDesc *desc = nullptr;
    const uint8_t *work = nullptr;
    const uint32_t *tlen = nullptr, *plen = nullptr;
    const uint64_t *out_off = nullptr;
    const uint32_t *out_len = nullptr;
    uint32_t nblocks = 0;
    uint64_t max_out = 0;
};

struct cramgpu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    DevPool scratch;          // carved anew by every call
    DevPool enc_big;          // encoder: pair counts, coder cells and work areas, sized on the device (k_enc_layout)
    DevPool dev_in, dev_out;  // device staging of the host entry points
    DevPool rec_scratch, rec_out;   // record-feature work: counters / results (results stay valid until the next record call)
    DevPool unp_in;                 // staged host series of cramgpu_features_unpack_resident
    DevPool rec_unp;                // results of cramgpu_features_unpack (its input series may be an encode result in rec_out)
    DevPool hs_scratch, hs_out, hs_in;   // cramgpu_host_series_pack: tables, result, staged inputs of the host variant
    DevPool bam_scratch, bam_cigar, bam_out, bam_unp, bam_pk_scratch, bam_unp_scratch;   // bam.cu: sizes/offsets, rebuilt CIGARs, packed records, unpacked arrays
    void *pin_bam = nullptr; size_t pin_bam_cap = 0;
    DevPool frame_scratch, frame_out;    // blocks.cu
    void *pin_frame = nullptr; size_t pin_frame_cap = 0; cudaEvent_t ev_frame = nullptr;
    DevPool mate_scratch;                // cramgpu_mate_link: name tables
    DevPool trk_scratch, trk_own, trk_ext;   // tracks.cu: tables, per-container count arrays, the per-base selector bytes
    void *pin_trk = nullptr; size_t pin_trk_cap = 0;
    bool trk_armed = false;              // trk_ext holds the bytes of the reads the next feature call with X / P selectors takes
    uint64_t trk_reads = 0; int trk_nsel = 0; int32_t trk_kind[7] = {0}, trk_param[7] = {0};
    DevPool slice_scratch, slice_in;     // cramgpu_slice_headers: accumulators, staged inputs of the host variant
    void *pin_slice = nullptr; size_t pin_slice_cap = 0; cudaEvent_t ev_slice = nullptr;
    void *pin_hs = nullptr, *pin_hs2 = nullptr; size_t pin_hs_cap = 0, pin_hs2_cap = 0;
    void *pin_rec = nullptr; size_t pin_rec_cap = 0;
    void *rec_fault = nullptr;      // records.cu: arguments of the last feature call with a BIN treatment (malloc)
    void *pin_feat = nullptr; size_t pin_feat_cap = 0;   // host copy of the last cramgpu_features_encode result
    void *pin_in = nullptr, *pin_out = nullptr, *pin_meta = nullptr;
    size_t pin_in_cap = 0, pin_out_cap = 0, pin_meta_cap = 0;
    cudaEvent_t ev = nullptr;
    // The coder kernels are latency-bound (a few warps per SM for tens of milliseconds).  They run on a high-priority
    // side stream so that, when several contexts share a GPU, their CTAs are placed ahead of the queued CTAs of the
    // throughput kernels of other streams instead of waiting behind them.
    cudaStream_t hp_stream = nullptr;
    cudaEvent_t hp_in = nullptr, hp_out = nullptr;
    EncodePending pend;
    uint32_t dec_wide_pool_blocks = 2048;   // rans_decode.cu: pool of wide-alphabet table images (cramgpu_set_option)
    bool enc_attr_done = false, dec_attr_done = false;   // per context (= per device): dynamic shared-memory limits of the coder kernels set
    // optional per-kernel timing (cramgpu_profile_*): CUDA events around every launch
    bool profiling = false;
    struct Span { const char *name; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> ev_pool;
};

// LAUNCH(ctx, stream, "name", kernel<<<...>>>(...)) counts the launch and, when profiling, brackets it with events
void prof_begin(cramgpu_ctx *ctx, cudaStream_t st, const char *name);
void prof_end(cramgpu_ctx *ctx, cudaStream_t st);
// run `...` (kernel launches that use the stream variable `hs`) on the context's high-priority stream, ordered after
// everything enqueued on `st` so far and before everything enqueued on `st` afterwards
#define ON_HP_STREAM(ctx, st, ...)                                              \
    do {                                                                        \
        cudaStream_t hs = (ctx)->hp_stream ? (ctx)->hp_stream : (st);           \
        if (hs != (st)) { cudaEventRecord((ctx)->hp_in, (st)); cudaStreamWaitEvent(hs, (ctx)->hp_in, 0); } \
        __VA_ARGS__;                                                            \
        if (hs != (st)) { cudaEventRecord((ctx)->hp_out, hs); cudaStreamWaitEvent((st), (ctx)->hp_out, 0); } \
    } while (0)

#define LAUNCH(ctx, st, name, ...)              \
    do {                                        \
        if ((ctx)->profiling) prof_begin((ctx), (st), (name)); \
        __VA_ARGS__;                            \
        (ctx)->launches++;                      \
        if ((ctx)->profiling) prof_end((ctx), (st)); \
        LAUNCH_CHECK((st), (name));             \
    } while (0)

// Debug builds (-DCRAMGPU_SYNC_LAUNCHES): wait for every kernel and name the one that faulted.  compute-sanitizer is not
// available on every pool; this at least says where to look.
#ifdef CRAMGPU_SYNC_LAUNCHES
#include <stdio.h>
#define LAUNCH_CHECK(st, name)                                                                    \
    do {                                                                                          \
        cudaError_t le__ = cudaStreamSynchronize(st);                                             \
        if (le__ != cudaSuccess) fprintf(stderr, "[cramgpu] kernel %s: %s\n", (name), cudaGetErrorString(le__)); \
    } while (0)
#else
#define LAUNCH_CHECK(st, name) do { } while (0)
#endif

int ctx_fail(cramgpu_ctx *ctx, int code, const char *fmt, ...);
int ctx_cuda(cramgpu_ctx *ctx, cudaError_t e, const char *what);
int pool_reserve(cramgpu_ctx *ctx, DevPool &pool, size_t bytes);
int pinned_reserve(cramgpu_ctx *ctx, void **p, size_t *cap, size_t bytes);
int tracks_selectors(const cramgpu_qs_policy *pol, int n_pol, int32_t *kind, int32_t *param);   // tracks.cu

#define CUDA_TRY(ctx, call)                                            \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) return ctx_cuda((ctx), e__, #call);    \
    } while (0)

// bump allocator over the context scratch pool (sizes first, then pointers)
struct Carver {
    size_t off = 0;
    size_t take(size_t bytes, size_t align = 256) {
        off = rup(off, align);
        size_t o = off;
        off += bytes;
        return o;
    }
};

// Host <-> device copies are issued in pieces: a copy engine serves the streams' copies one whole operation after the
// other, and a multi-gigabyte copy of one context would hold up the small copies of every other context for its duration.
cudaError_t memcpy_pieces(void *dst, const void *src, size_t bytes, cudaMemcpyKind kind, cudaStream_t st);

struct CopyJob { void *dst; const void *src; size_t n; };
void par_memcpy(std::vector<CopyJob> &jobs);      // multi-threaded staging copies (clears the list)
bool host_is_pinned(const void *p);               // page-locked host memory known to CUDA?

// Where a caller's pointer lives.  The host entry points take host OR device pointers for their bulk arrays (unified
// virtual addressing tells them apart): device memory is used in place, page-locked host memory goes to the DMA engine
// as it is, pageable host memory is staged.  One MemMap per call remembers the ranges it has already asked about.
enum MemKind { MEM_PAGEABLE = 0, MEM_PINNED = 1, MEM_DEVICE = 2 };
struct MemMap {
    struct Range { uintptr_t lo, hi; MemKind kind; };
    std::vector<Range> ranges;
    size_t last = 0;
    MemKind kind(const void *p);
};

// rans.cu
int rans_encode_device(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                       const uint32_t *in_len, const uint8_t *order, uint8_t *d_out, uint64_t out_cap,
                       uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, cudaStream_t st);
int rans_encode_pack(cramgpu_ctx *ctx, uint8_t *d_out, cudaStream_t st);
int rans_decode_device(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                       const uint32_t *in_len, const uint8_t *in_prefix, uint8_t *d_out, uint64_t out_cap,
                       uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, cudaStream_t st,
                       const uint64_t *dst_addr = nullptr, const uint32_t *dst_cap = nullptr);
int rans_parse_prefix(const uint8_t *p, uint32_t len, uint32_t *order, uint32_t *raw);
