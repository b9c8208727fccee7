// series.cu -- the per-record fields Writer.write emits BEFORE the read features, packed on the device.
//
// Restatement of Writer.write [Writer.java:139-172] (htsjdk 2.4.1) for the series the Java host used to build itself:
//   BF  ITF8  CramCompressionRecord.flags            Sam2CramRecordFactory.createCramRecord [:122-131,143-152]
//   CF  byte  getCompressionFlags()                  bit 0 force-preserve scores, 1 detached, 2 mate downstream, 3 unknown bases
//   RI  ITF8  sequenceId, multi-reference slices only [:141-142]
//   RL  ITF8  readLength
//   AP  ITF8  alignmentStart, or its delta to the previous record's start when CompressionHeader.APDelta is set
//             (ContainerFactory.buildSlice [:179-182]: the first record against Slice.alignmentStart)
//   RG  ITF8  readGroupID (-1 = none: 5 bytes)
//   RN  name + TAB (ByteArrayStopEncoding) for every record when CompressionHeader.readNamesIncluded, else for detached ones
//   MF  byte  getMateFlags()         )
//   NS  ITF8  mateSequenceID         )  detached records only [:158-167]
//   NP  ITF8  mateAlignmentStart     )
//   TS  ITF8  templateSize           )
//   NF  ITF8  recordsToNextFragment, records whose mate follows in the slice [:168-169]
//   TL  ITF8  tagIdsIndex [:172]
// Pinned to the jar: tests/golden/streams_jvm.json holds what Writer.write produced in the mini-JVM (oracle/make_golden_streams.py).
//
//   k_hs_count   CTA per container: bytes of each series
//   k_hs_layout  one CTA: region offsets, series-major, every (container, series) region at a multiple of 16
//   k_hs_write   CTA per container, 256 records per round: three block scans over packed byte counts, then every thread
//                writes its record's fields
#include "internal.cuh"
#include <string.h>

#define HS_N CRAMGPU_HS_COUNT

struct HsArgs {
    uint64_t n_reads; uint32_t n_containers;
    const uint64_t *cont_off; const int32_t *slice_start; const uint8_t *multi_ref;      // device copies
    uint32_t ap_delta, names_included;
    const uint16_t *flag; const uint8_t *cf, *mate_cf, *force_qs, *rflags;
    const int32_t *ref_id, *pos; const uint64_t *read_off; const int32_t *read_group;
    const uint8_t *names; const uint64_t *name_off;
    const int32_t *mate_ref_id, *mate_pos, *tlen, *nf, *tl;
    uint64_t *totals;          // [nc][HS_N]
    uint64_t *stream_off;      // [nc][HS_N]
    uint8_t *streams;
};

__device__ __forceinline__ uint32_t hs_itf8_len(uint32_t v)     // ITF8.writeUnsignedITF8 [ITF8.java:53-84]
{
    return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : v < (1u << 28) ? 4 : 5;
}
__device__ __forceinline__ void hs_itf8_put(uint8_t *p, uint32_t v)
{
    if (v < (1u << 7)) { p[0] = (uint8_t)v; }
    else if (v < (1u << 14)) { p[0] = (uint8_t)((v >> 8) | 0x80); p[1] = (uint8_t)v; }
    else if (v < (1u << 21)) { p[0] = (uint8_t)((v >> 16) | 0xC0); p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)v; }
    else if (v < (1u << 28)) { p[0] = (uint8_t)((v >> 24) | 0xE0); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }
    else { p[0] = (uint8_t)((v >> 28) | 0xF0); p[1] = (uint8_t)(v >> 20); p[2] = (uint8_t)(v >> 12); p[3] = (uint8_t)(v >> 4); p[4] = (uint8_t)v; }
}

// the values of one record, as Writer.write reads them
struct HsRec {
    uint32_t bf, cf, ri, rl, ap, rg, mf, ns, np, ts, nf, tl, name_len;
    uint64_t name_at;
    bool detached, downstream, named;
};

__device__ __forceinline__ HsRec hs_load(const HsArgs &a, uint64_t r, uint32_t c, uint64_t first)
{
    HsRec v;
    const uint32_t f = a.flag[r];
    // BF: the SAM flag without the two mate bits; an unpaired read also loses proper-pair and first / last segment [:122-131,143-152]
    v.bf = (f & 1u) ? (f & (0x1u | 0x2u | 0x4u | 0x10u | 0x40u | 0x80u | 0x100u | 0x200u | 0x400u | 0x800u))
                    : (f & (0x4u | 0x10u | 0x100u | 0x200u | 0x400u | 0x800u));
    // MF: bit 0 mate on the reverse strand (SAM 0x20), bit 1 mate unmapped (SAM 0x8) [:114-115]
    v.mf = (f & 1u) ? (((f & 0x20u) ? 1u : 0u) | ((f & 0x8u) ? 2u : 0u)) : 0u;
    if (a.cf) v.cf = a.cf[r];
    else v.cf = (a.force_qs ? (a.force_qs[r] & 1u) : 0u) | (a.mate_cf ? (a.mate_cf[r] & 6u) : 2u) |
                ((a.rflags && (a.rflags[r] & CRAMGPU_RF_NO_BASES)) ? 8u : 0u);
    v.detached = (v.cf & 2u) != 0;
    v.downstream = !v.detached && (v.cf & 4u) != 0;
    v.ri = a.ref_id ? (uint32_t)a.ref_id[r] : 0u;
    v.rl = (uint32_t)(a.read_off[r + 1] - a.read_off[r]);
    const int32_t p = a.pos[r];
    v.ap = a.ap_delta ? (uint32_t)(p - (r == first ? a.slice_start[c] : a.pos[r - 1])) : (uint32_t)p;
    v.rg = a.read_group ? (uint32_t)a.read_group[r] : 0xFFFFFFFFu;
    v.named = a.names && (a.names_included || v.detached);
    v.name_at = v.named ? a.name_off[r] : 0;
    v.name_len = v.named ? (uint32_t)(a.name_off[r + 1] - a.name_off[r]) : 0u;
    v.ns = a.mate_ref_id ? (uint32_t)a.mate_ref_id[r] : 0xFFFFFFFFu;
    v.np = a.mate_pos ? (uint32_t)a.mate_pos[r] : 0u;
    v.ts = a.tlen ? (uint32_t)a.tlen[r] : 0u;
    v.nf = a.nf ? (uint32_t)a.nf[r] : 0u;
    v.tl = a.tl ? (uint32_t)a.tl[r] : 0u;
    return v;
}

// byte counts of one record, packed for the block scans: 12 bits per series (a round of 256 records stays below 4096
// bytes per series), the name bytes in a word of their own
struct HsSizes { unsigned long long a, b; uint32_t c, rn; };
__device__ __forceinline__ HsSizes hs_sizes(const HsRec &v, bool multi)
{
    HsSizes s;
    s.a = (unsigned long long)hs_itf8_len(v.bf) | (1ull << 12) | ((unsigned long long)(multi ? hs_itf8_len(v.ri) : 0u) << 24) |
          ((unsigned long long)hs_itf8_len(v.rl) << 36) | ((unsigned long long)hs_itf8_len(v.ap) << 48);
    s.b = (unsigned long long)hs_itf8_len(v.rg) | ((unsigned long long)(v.detached ? 1u : 0u) << 12) |
          ((unsigned long long)(v.detached ? hs_itf8_len(v.ns) : 0u) << 24) | ((unsigned long long)(v.detached ? hs_itf8_len(v.np) : 0u) << 36) |
          ((unsigned long long)(v.detached ? hs_itf8_len(v.ts) : 0u) << 48);
    s.c = (v.downstream ? hs_itf8_len(v.nf) : 0u) | (hs_itf8_len(v.tl) << 12);
    s.rn = v.named ? v.name_len + 1u : 0u;
    return s;
}

template <typename T>
__device__ __forceinline__ T hs_block_scan(T v, T *wsum, T &total)      // exclusive scan over the 256 threads of the CTA
{
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T t = __shfl_up_sync(0xFFFFFFFFu, s, d);
        if (lane >= (uint32_t)d) s += t;
    }
    __syncthreads();                     // wsum may still be read from the scan before
    if (lane == 31) wsum[w] = s;
    __syncthreads();
    T base = 0, tot = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) { if (k < w) base += wsum[k]; tot += wsum[k]; }
    total = tot;
    return base + s - v;
}

__global__ void __launch_bounds__(256) k_hs_count(HsArgs a)
{
    const uint32_t c = blockIdx.x;
    const uint64_t r0 = a.cont_off[c], r1 = a.cont_off[c + 1];
    const bool multi = a.multi_ref && a.multi_ref[c];
    unsigned long long t[HS_N];
#pragma unroll
    for (int i = 0; i < HS_N; i++) t[i] = 0;
    for (uint64_t r = r0 + threadIdx.x; r < r1; r += 256) {
        const HsRec v = hs_load(a, r, c, r0);
        const HsSizes s = hs_sizes(v, multi);
        t[CRAMGPU_HS_BF] += s.a & 0xFFF; t[CRAMGPU_HS_CF] += (s.a >> 12) & 0xFFF; t[CRAMGPU_HS_RI] += (s.a >> 24) & 0xFFF;
        t[CRAMGPU_HS_RL] += (s.a >> 36) & 0xFFF; t[CRAMGPU_HS_AP] += (s.a >> 48) & 0xFFF;
        t[CRAMGPU_HS_RG] += s.b & 0xFFF; t[CRAMGPU_HS_MF] += (s.b >> 12) & 0xFFF; t[CRAMGPU_HS_NS] += (s.b >> 24) & 0xFFF;
        t[CRAMGPU_HS_NP] += (s.b >> 36) & 0xFFF; t[CRAMGPU_HS_TS] += (s.b >> 48) & 0xFFF;
        t[CRAMGPU_HS_NF] += s.c & 0xFFF; t[CRAMGPU_HS_TL] += (s.c >> 12) & 0xFFF; t[CRAMGPU_HS_RN] += s.rn;
    }
    __shared__ unsigned long long sm[8][HS_N];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < HS_N; i++) {
        unsigned long long v = t[i];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
        if (lane == 0) sm[w][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < HS_N) {
        unsigned long long v = 0;
        for (int k = 0; k < 8; k++) v += sm[k][threadIdx.x];
        a.totals[(size_t)c * HS_N + threadIdx.x] = v;
    }
}

// series-major: the BF regions of all containers, then CF, ...; out_total[0] = bytes used
__global__ void __launch_bounds__(1024) k_hs_layout(const uint64_t *__restrict__ totals, uint32_t nc, uint64_t *__restrict__ off, uint64_t *__restrict__ out_total)
{
    __shared__ uint64_t wsum[32];
    __shared__ uint64_t carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (uint32_t s = 0; s < HS_N; s++)
        for (uint32_t base = 0; base < nc; base += 1024) {
            const uint32_t c = base + tid;
            const uint64_t len = c < nc ? (totals[(size_t)c * HS_N + s] + 15) & ~15ull : 0;
            uint64_t x = len;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint64_t t = __shfl_up_sync(0xFFFFFFFFu, x, d); if (lane >= (uint32_t)d) x += t; }
            if (lane == 31) wsum[w] = x;
            __syncthreads();
            uint64_t o = carry;
            for (uint32_t k = 0; k < w; k++) o += wsum[k];
            if (c < nc) off[(size_t)c * HS_N + s] = o + x - len;
            __syncthreads();
            if (tid == 1023) carry = o + x;
            __syncthreads();
        }
    if (tid == 0) out_total[0] = carry;
}

__global__ void __launch_bounds__(256) k_hs_write(HsArgs a)
{
    const uint32_t c = blockIdx.x;
    const uint64_t r0 = a.cont_off[c], r1 = a.cont_off[c + 1];
    const bool multi = a.multi_ref && a.multi_ref[c];
    __shared__ unsigned long long wsA[8];
    __shared__ uint32_t wsC[8];
    __shared__ uint64_t base[HS_N];
    if (threadIdx.x < HS_N) base[threadIdx.x] = a.stream_off[(size_t)c * HS_N + threadIdx.x];
    __syncthreads();
    uint8_t *S = a.streams;
    for (uint64_t t0 = r0; t0 < r1; t0 += 256) {
        const uint64_t r = t0 + threadIdx.x;
        const bool live = r < r1;
        HsRec v = {};
        HsSizes s = {0, 0, 0, 0};
        if (live) { v = hs_load(a, r, c, r0); s = hs_sizes(v, multi); }
        unsigned long long totA, totB; uint32_t totC, totN;
        const unsigned long long pa = hs_block_scan(s.a, wsA, totA);
        const unsigned long long pb = hs_block_scan(s.b, wsA, totB);
        const uint32_t pc = hs_block_scan(s.c, wsC, totC);
        const uint32_t pn = hs_block_scan(s.rn, wsC, totN);
        if (live) {
            hs_itf8_put(S + base[CRAMGPU_HS_BF] + (pa & 0xFFF), v.bf);
            S[base[CRAMGPU_HS_CF] + ((pa >> 12) & 0xFFF)] = (uint8_t)v.cf;
            if (multi) hs_itf8_put(S + base[CRAMGPU_HS_RI] + ((pa >> 24) & 0xFFF), v.ri);
            hs_itf8_put(S + base[CRAMGPU_HS_RL] + ((pa >> 36) & 0xFFF), v.rl);
            hs_itf8_put(S + base[CRAMGPU_HS_AP] + ((pa >> 48) & 0xFFF), v.ap);
            hs_itf8_put(S + base[CRAMGPU_HS_RG] + (pb & 0xFFF), v.rg);
            if (v.detached) {
                S[base[CRAMGPU_HS_MF] + ((pb >> 12) & 0xFFF)] = (uint8_t)v.mf;
                hs_itf8_put(S + base[CRAMGPU_HS_NS] + ((pb >> 24) & 0xFFF), v.ns);
                hs_itf8_put(S + base[CRAMGPU_HS_NP] + ((pb >> 36) & 0xFFF), v.np);
                hs_itf8_put(S + base[CRAMGPU_HS_TS] + ((pb >> 48) & 0xFFF), v.ts);
            } else if (v.downstream) {
                hs_itf8_put(S + base[CRAMGPU_HS_NF] + (pc & 0xFFF), v.nf);
            }
            hs_itf8_put(S + base[CRAMGPU_HS_TL] + ((pc >> 12) & 0xFFF), v.tl);
        }
        // names: the warp copies one record's name after the other, lanes over consecutive bytes (a thread copying its own
        // name byte by byte touches one cache line per lane and instruction)
        {
            const uint32_t lane = threadIdx.x & 31;
            const bool named = live && v.named;
            const uint64_t my_dst = base[CRAMGPU_HS_RN] + pn, my_src = named ? v.name_at : 0;
            const uint32_t my_len = named ? v.name_len : 0xFFFFFFFFu;
            uint32_t todo = __ballot_sync(0xFFFFFFFFu, named);
            while (todo) {
                const int i = __ffs(todo) - 1;
                todo &= todo - 1;
                const uint64_t dst = __shfl_sync(0xFFFFFFFFu, my_dst, i), src = __shfl_sync(0xFFFFFFFFu, my_src, i);
                const uint32_t len = __shfl_sync(0xFFFFFFFFu, my_len, i);
                for (uint32_t b = lane; b <= len; b += 32) S[dst + b] = b < len ? a.names[src + b] : (uint8_t)'\t';
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            base[CRAMGPU_HS_BF] += totA & 0xFFF; base[CRAMGPU_HS_CF] += (totA >> 12) & 0xFFF; base[CRAMGPU_HS_RI] += (totA >> 24) & 0xFFF;
            base[CRAMGPU_HS_RL] += (totA >> 36) & 0xFFF; base[CRAMGPU_HS_AP] += (totA >> 48) & 0xFFF;
            base[CRAMGPU_HS_RG] += totB & 0xFFF; base[CRAMGPU_HS_MF] += (totB >> 12) & 0xFFF; base[CRAMGPU_HS_NS] += (totB >> 24) & 0xFFF;
            base[CRAMGPU_HS_NP] += (totB >> 36) & 0xFFF; base[CRAMGPU_HS_TS] += (totB >> 48) & 0xFFF;
            base[CRAMGPU_HS_NF] += totC & 0xFFF; base[CRAMGPU_HS_TL] += (totC >> 12) & 0xFFF; base[CRAMGPU_HS_RN] += totN;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------ host side

static int hs_pack_impl(cramgpu_ctx *ctx, const cramgpu_host_fields *in, cramgpu_host_series *out, uint64_t *h_off, uint64_t *h_len,
                        cudaStream_t st)
{
    if (!in || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    memset(out, 0, sizeof *out);
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->cont_off || !in->flag || !in->pos || !in->read_off || (in->ap_delta && !in->slice_start) || (in->names && !in->name_off))
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->cont_off[0] != 0 || in->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    bool any_multi = false;
    if (in->multi_ref)
        for (uint32_t c = 0; c < nc; c++) any_multi = any_multi || in->multi_ref[c];
    if (any_multi && !in->ref_id) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "multi-reference slice without ref_id");
    Carver cv;
    size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_ss = cv.take(4 * (size_t)nc), o_mr = cv.take(nc);
    size_t o_tot = cv.take(8 * (size_t)nc * HS_N), o_off = cv.take(8 * (size_t)nc * HS_N), o_sum = cv.take(16);
    if (int rc = pool_reserve(ctx, ctx->hs_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->hs_scratch.p;
    const size_t span = o_sum + 16 - o_tot;                    // totals, offsets and the grand total lie one after the other
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, 16 * (size_t)(nc + 1) + span + 1024)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));
    uint8_t *pm = (uint8_t *)ctx->pin_meta;
    size_t po = 0;
    auto up = [&](const void *src, size_t bytes, void *dst) -> cudaError_t {
        if (!bytes || !src) return cudaSuccess;
        memcpy(pm + po, src, bytes);
        cudaError_t e = cudaMemcpyAsync(dst, pm + po, bytes, cudaMemcpyHostToDevice, st);
        po += rup(bytes, 16);
        return e;
    };
    CUDA_TRY(ctx, up(in->cont_off, 8 * (size_t)(nc + 1), S + o_cont));
    CUDA_TRY(ctx, up(in->slice_start, 4 * (size_t)nc, S + o_ss));
    CUDA_TRY(ctx, up(in->multi_ref, nc, S + o_mr));
    HsArgs a;
    memset(&a, 0, sizeof a);
    a.n_reads = n; a.n_containers = nc;
    a.cont_off = (const uint64_t *)(S + o_cont); a.slice_start = in->slice_start ? (const int32_t *)(S + o_ss) : nullptr;
    a.multi_ref = in->multi_ref ? S + o_mr : nullptr;
    a.ap_delta = in->ap_delta; a.names_included = in->names_included;
    a.flag = in->flag; a.cf = in->cf; a.mate_cf = in->mate_cf; a.force_qs = in->force_qs; a.rflags = in->rflags;
    a.ref_id = in->ref_id; a.pos = in->pos; a.read_off = in->read_off; a.read_group = in->read_group;
    a.names = in->names; a.name_off = in->name_off;
    a.mate_ref_id = in->mate_ref_id; a.mate_pos = in->mate_pos; a.tlen = in->tlen; a.nf = in->nf; a.tl = in->tl;
    a.totals = (uint64_t *)(S + o_tot); a.stream_off = (uint64_t *)(S + o_off);
    LAUNCH(ctx, st, "k_hs_count", k_hs_count<<<nc, 256, 0, st>>>(a));
    LAUNCH(ctx, st, "k_hs_layout", k_hs_layout<<<1, 1024, 0, st>>>(a.totals, nc, a.stream_off, (uint64_t *)(S + o_sum)));
    // the host learns the layout (it hands the series on to the block codec) and the buffer is sized from the total
    uint8_t *pr = pm + rup(po, 64);
    CUDA_TRY(ctx, cudaMemcpyAsync(pr, S + o_tot, span, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t *t_len = (const uint64_t *)pr, *t_off = (const uint64_t *)(pr + (o_off - o_tot));
    const uint64_t total = *(const uint64_t *)(pr + (o_sum - o_tot));
    if (h_len) memcpy(h_len, t_len, 8 * (size_t)nc * HS_N);
    if (h_off) memcpy(h_off, t_off, 8 * (size_t)nc * HS_N);
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));
    if (int rc = pool_reserve(ctx, ctx->hs_out, total + 64)) return rc;
    a.streams = (uint8_t *)ctx->hs_out.p;
    LAUNCH(ctx, st, "k_hs_write", k_hs_write<<<nc, 256, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    out->streams = a.streams; out->streams_len = total;
    out->stream_off = a.stream_off; out->stream_len = a.totals;
    return CRAMGPU_OK;
}

extern "C" int cramgpu_host_series_pack_dev(cramgpu_ctx *ctx, const cramgpu_host_fields *in, cramgpu_host_series *out,
                                            uint64_t *h_stream_off, uint64_t *h_stream_len, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return hs_pack_impl(ctx, in, out, h_stream_off, h_stream_len, stream ? (cudaStream_t)stream : ctx->stream);
}

// host variants: every per-record array in host memory.  host_series = bit mask (1u << CRAMGPU_HS_x) of the series that come
// back to the host; with dev_out the result also stays on the device for cramgpu_rans_encode_batch to code in place.
static int hs_pack_host(cramgpu_ctx *ctx, const cramgpu_host_fields *in, uint32_t host_series, cramgpu_host_series *out,
                        cramgpu_host_series *dev_out)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    memset(out, 0, sizeof *out);
    if (dev_out) memset(dev_out, 0, sizeof *dev_out);
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->flag || !in->pos || !in->read_off || (in->names && !in->name_off)) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    // stage the per-record arrays (arrays that already live on the device are used where they are)
    MemMap mm;
    struct Item { const void *src; size_t bytes; size_t off; };
    std::vector<Item> items;
    Carver cv;
    auto add = [&](const void *p, size_t bytes) -> size_t {
        if (!p || mm.kind(p) == MEM_DEVICE) return (size_t)-1;
        size_t o = cv.take(bytes + 16); items.push_back({p, bytes, o}); return o;
    };
    MemKind nok = in->name_off ? mm.kind(in->name_off) : MEM_PAGEABLE;
    if (in->names && nok == MEM_DEVICE) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "name_off must be host memory in the host variant");
    const uint64_t name_bytes = in->names ? in->name_off[n] : 0;
    size_t i_flag = add(in->flag, 2 * n), i_cf = add(in->cf, n), i_mcf = add(in->mate_cf, n), i_fq = add(in->force_qs, n), i_rfl = add(in->rflags, n);
    size_t i_ref = add(in->ref_id, 4 * n), i_pos = add(in->pos, 4 * n), i_ro = add(in->read_off, 8 * (n + 1)), i_rg = add(in->read_group, 4 * n);
    size_t i_nm = add(in->names, name_bytes), i_no = add(in->name_off, in->names ? 8 * (n + 1) : 0);
    size_t i_mr = add(in->mate_ref_id, 4 * n), i_mp = add(in->mate_pos, 4 * n), i_ts = add(in->tlen, 4 * n), i_nf = add(in->nf, 4 * n), i_tl = add(in->tl, 4 * n);
    if (int rc = pool_reserve(ctx, ctx->hs_in, cv.off + 16)) return rc;
    uint8_t *D = (uint8_t *)ctx->hs_in.p;
    for (const Item &it : items)
        if (it.bytes) CUDA_TRY(ctx, memcpy_pieces(D + it.off, it.src, it.bytes, cudaMemcpyHostToDevice, st));
    cramgpu_host_fields d = *in;
    auto at = [&](size_t o, const void *orig) -> const void * { return o == (size_t)-1 ? orig : (const void *)(D + o); };
    d.flag = (const uint16_t *)at(i_flag, in->flag); d.cf = (const uint8_t *)at(i_cf, in->cf); d.mate_cf = (const uint8_t *)at(i_mcf, in->mate_cf);
    d.force_qs = (const uint8_t *)at(i_fq, in->force_qs); d.rflags = (const uint8_t *)at(i_rfl, in->rflags);
    d.ref_id = (const int32_t *)at(i_ref, in->ref_id); d.pos = (const int32_t *)at(i_pos, in->pos);
    d.read_off = (const uint64_t *)at(i_ro, in->read_off); d.read_group = (const int32_t *)at(i_rg, in->read_group);
    d.names = (const uint8_t *)at(i_nm, in->names); d.name_off = (const uint64_t *)at(i_no, in->name_off);
    d.mate_ref_id = (const int32_t *)at(i_mr, in->mate_ref_id); d.mate_pos = (const int32_t *)at(i_mp, in->mate_pos);
    d.tlen = (const int32_t *)at(i_ts, in->tlen); d.nf = (const int32_t *)at(i_nf, in->nf); d.tl = (const int32_t *)at(i_tl, in->tl);
    cramgpu_host_series dev;
    const size_t tab = 8 * (size_t)nc * HS_N;
    if (int rc = pinned_reserve(ctx, &ctx->pin_hs, &ctx->pin_hs_cap, 2 * tab + 64)) return rc;
    uint64_t *h_off = (uint64_t *)ctx->pin_hs, *h_len = h_off + (size_t)nc * HS_N;
    if (int rc = hs_pack_impl(ctx, &d, &dev, h_off, h_len, st)) return rc;
    if (int rc = pinned_reserve(ctx, &ctx->pin_hs2, &ctx->pin_hs2_cap, dev.streams_len + 64)) return rc;
    // the layout is series-major: a series of all containers is one contiguous piece
    uint8_t *hstreams = (uint8_t *)ctx->pin_hs2;
    for (uint32_t k = 0; k < HS_N; k++) {
        if (!(host_series & (1u << k))) continue;
        const uint64_t lo = h_off[k], hi = k + 1 < HS_N ? h_off[k + 1] : dev.streams_len;      // container 0's region of series k .. of series k + 1
        if (hi > lo) CUDA_TRY(ctx, memcpy_pieces(hstreams + lo, dev.streams + lo, hi - lo, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    out->streams = hstreams; out->streams_len = dev.streams_len;
    out->stream_off = h_off; out->stream_len = h_len;
    if (dev_out) *dev_out = dev;
    return CRAMGPU_OK;
}

extern "C" int cramgpu_host_series_pack(cramgpu_ctx *ctx, const cramgpu_host_fields *in, cramgpu_host_series *out)
{
    return hs_pack_host(ctx, in, 0xFFFFFFFFu, out, nullptr);
}

extern "C" int cramgpu_host_series_pack_resident(cramgpu_ctx *ctx, const cramgpu_host_fields *in, uint32_t host_series,
                                                 cramgpu_host_series *out, cramgpu_host_series *dev_out)
{
    if (!dev_out) return ctx ? ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument") : CRAMGPU_ERR_ARG;
    return hs_pack_host(ctx, in, host_series, out, dev_out);
}
