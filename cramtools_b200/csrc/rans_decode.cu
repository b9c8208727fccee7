// rans_decode.cu -- htsjdk-compatible rANS order-0 / order-1 DECODE for sm_100a.
//
// Restatement of RANS.uncompress [RANS.java:29-55] as a batch pipeline:
//
//   (host)        9-byte prefix: order, compressed size check, raw size        [RANS.java:31-44]
//   k_dec_tables  frequency table -> slot->symbol LUT + {freq, start} per symbol Frequencies.readStats_o0 [:13-46], _o1 [:50-106]
//   k_dec0        4-way interleaved decoder, one lane per state                 D04.uncompress [D04.java:9-85]
//   k_dec1        4 quarter-lanes, context = previously decoded byte            D14.uncompress [D14.java:9-70]
//
// The four states of a block renormalise from ONE byte cursor in the order 0,1,2,3
// [D04.java:36-39, Decoding.java:84-90]; per step each lane needs 0..2 bytes, and its read position is
// the cursor plus the bytes taken by the lower lanes of its group (two ballots + popc).
#include "internal.cuh"

struct DecMeta {
    uint32_t pay_off;   // offset of the 16 state bytes inside the compressed block
};

// grid nblocks, 128 threads.  Thread 0 walks the table bytes (inherently sequential: every field's
// position depends on the previous ones); all threads then expand each row.
//   lut [ctx][4096] u8   slot -> symbol                (Arrays.fill(R, x, x+F, j), [Frequencies.java:31])
//   fc  [ctx][256]  u32  freq << 16 | cumulative start
__global__ void __launch_bounds__(128) k_dec_tables(const BlockDesc *__restrict__ desc, const uint8_t *__restrict__ d_in,
                                                    uint8_t *__restrict__ lut0, uint32_t *__restrict__ fc0,
                                                    uint8_t *__restrict__ lut1, uint32_t *__restrict__ fc1,
                                                    DecMeta *__restrict__ meta, int32_t *__restrict__ status)
{
    uint32_t blk = blockIdx.x;
    const BlockDesc d = desc[blk];
    if (d.order > 1) return;
    const uint8_t *in = d_in + d.in_off;
    const uint32_t len = d.in_len;
    const bool o1 = d.order == 1;
    uint8_t *lut = o1 ? lut1 + (size_t)d.o1_index * (256 * 4096) : lut0 + (size_t)blk * 4096;
    uint32_t *fc = o1 ? fc1 + (size_t)d.o1_index * 65536 : fc0 + (size_t)blk * 256;

    __shared__ uint16_t F[256];
    __shared__ uint16_t C[257];
    __shared__ uint8_t symlist[256];
    __shared__ uint32_t s_nsym, s_pos, s_ctx, s_more, s_bad;
    uint32_t tid = threadIdx.x;
    // parser state (thread 0 only)
    uint32_t pos = 9, ctx = 0;
    int rle_i = 0;
    bool bad = false;
    auto getb = [&]() -> uint32_t { if (pos >= len) { bad = true; return 0; } return in[pos++]; };
    auto peekb = [&]() -> int { if (pos >= len) { bad = true; return -1; } return in[pos]; };
    if (tid == 0 && o1) ctx = getb();

    for (;;) {
        F[tid] = 0; F[tid + 128] = 0;
        __syncthreads();
        if (tid == 0) {
            // one row: readStats_o0 [:13-46] / inner loop of readStats_o1 [:62-94]
            uint32_t x = 0, ns = 0;
            int rle = 0;
            uint32_t j = getb();
            do {
                uint32_t f = getb();
                if (f >= 128) f = ((f & 127) << 8) | getb();
                if (o1 && f == 0) f = TOTFREQ;                                   // [:68-69]
                if (x + f > TOTFREQ || ns >= 256) { bad = true; break; }          // Java: AIOOBE in Arrays.fill
                F[j] = (uint16_t)f; C[j] = (uint16_t)x; symlist[ns++] = (uint8_t)j;
                x += f;
                if (rle == 0 && peekb() == (int)j + 1) { j = getb(); rle = (int)getb(); }
                else if (rle != 0) { rle--; j++; if (j > 255) { bad = true; break; } }
                else j = getb();
                if (bad) break;
            } while (j != 0);
            s_nsym = ns; s_ctx = ctx;
            if (o1 && !bad) {
                // next context [:96-105]
                if (rle_i == 0 && peekb() == (int)ctx + 1) { ctx = getb(); rle_i = (int)getb(); }
                else if (rle_i != 0) { rle_i--; ctx++; if (ctx > 255) bad = true; }
                else ctx = getb();
            }
            s_more = (o1 && !bad && ctx != 0) ? 1u : 0u;
            s_bad = bad ? 1u : 0u;
            s_pos = pos;
        }
        __syncthreads();
        if (s_bad) break;
        // expand the row: every symbol fills its slots
        uint8_t *row_lut = lut + (size_t)s_ctx * 4096 * (o1 ? 1 : 0);
        uint32_t *row_fc = fc + (size_t)s_ctx * 256 * (o1 ? 1 : 0);
        uint32_t ns = s_nsym;
        for (uint32_t a = tid >> 5; a < ns; a += 4) {       // one warp per symbol
            uint32_t j = symlist[a], f = F[j], c = C[j];
            if ((tid & 31) == 0) row_fc[j] = (f << 16) | c;
            for (uint32_t s = tid & 31; s < f; s += 32) row_lut[c + s] = (uint8_t)j;
        }
        __syncthreads();
        if (!s_more) break;
    }
    if (tid == 0) {
        meta[blk].pay_off = s_pos;
        if (s_bad) status[blk] = CRAMGPU_ERR_CORRUPT;
    }
}

__device__ __forceinline__ uint32_t ld_byte(const uint8_t *p, uint32_t pos, uint32_t len, bool &under)
{
    if (pos >= len) { under = true; return 0; }
    return p[pos];
}

// Renormalise the 4 states of every group from the group's shared cursor, lanes in the order 0,1,2,3.
__device__ __forceinline__ uint32_t renorm_step(uint32_t x, bool act, const uint8_t *in, uint32_t len, uint32_t &cur,
                                                uint32_t lane, bool &under)
{
    bool n1 = act && x < RANS_L;
    bool n2 = act && x < (1u << 15);       // after one byte still below 2^23 -> a second byte
    uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1), b2 = __ballot_sync(0xFFFFFFFFu, n2);
    uint32_t g4 = lane & ~3u, k = lane & 3u;
    uint32_t gm = 0xFu << g4;
    uint32_t lm = ((1u << k) - 1u) << g4;  // lower lanes of the group
    uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);
    if (n1) {
        x = (x << 8) | ld_byte(in, at, len, under);
        if (n2) x = (x << 8) | ld_byte(in, at + 1, len, under);
    }
    cur += __popc(b1 & gm) + __popc(b2 & gm);
    return x;
}

// k_dec0: 1 warp per CTA, 8 blocks per warp.
__global__ void __launch_bounds__(32) k_dec0(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                             uint32_t nslots, const uint8_t *__restrict__ d_in,
                                             const uint8_t *__restrict__ lut0, const uint32_t *__restrict__ fc0,
                                             const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                             int32_t *__restrict__ status)
{
    uint32_t lane = threadIdx.x, k = lane & 3u;
    uint32_t slot = blockIdx.x * 8 + (lane >> 2);
    bool have = slot < nslots;
    uint32_t blk = have ? slots[slot] : 0;
    BlockDesc d = desc[blk];
    have = have && status[blk] == 0;
    uint32_t n = have ? d.out_len : 0;
    const uint8_t *in = d_in + d.in_off;
    uint32_t len = d.in_len;
    const uint8_t *lut = lut0 + (size_t)blk * 4096;
    const uint32_t *fc = fc0 + (size_t)blk * 256;
    uint8_t *out = d_out + d.out_off;
    bool under = false;

    uint32_t cur = have ? meta[blk].pay_off : 0;
    uint32_t x = 0;
    if (have) {
        x = ld_byte(in, cur + 4 * k, len, under) | (ld_byte(in, cur + 4 * k + 1, len, under) << 8) |
            (ld_byte(in, cur + 4 * k + 2, len, under) << 16) | (ld_byte(in, cur + 4 * k + 3, len, under) << 24);
        cur += 16;
    }
    uint32_t steps = n >> 2;
    uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, steps);
    for (uint32_t t = 0; t < max_steps; t++) {
        bool act = t < steps;
        if (act) {
            uint32_t m = x & (TOTFREQ - 1);
            uint32_t s = lut[m];
            uint32_t e = fc[s];
            out[(uint64_t)t * 4 + k] = (uint8_t)s;
            x = (e >> 16) * (x >> TF_SHIFT) + m - (e & 0xFFFF);                 // [Decoding.java:35-48]
        }
        x = renorm_step(x, act, in, len, cur, lane, under);
    }
    if (have && k < (n & 3u)) out[(uint64_t)steps * 4 + k] = lut[x & (TOTFREQ - 1)];   // tail [D04.java:60-83]
    if (have && under) status[blk] = CRAMGPU_ERR_CORRUPT;
}

// k_dec1: lane k decodes the quarter starting at k*(n>>2); lane 3 also the remainder [D14.java:62-68].
__global__ void __launch_bounds__(32) k_dec1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                             uint32_t nslots, const uint8_t *__restrict__ d_in,
                                             const uint8_t *__restrict__ lut1, const uint32_t *__restrict__ fc1,
                                             const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                             int32_t *__restrict__ status)
{
    uint32_t lane = threadIdx.x, k = lane & 3u;
    uint32_t slot = blockIdx.x * 8 + (lane >> 2);
    bool have = slot < nslots;
    uint32_t blk = have ? slots[slot] : 0;
    BlockDesc d = desc[blk];
    have = have && status[blk] == 0;
    uint32_t n = have ? d.out_len : 0;
    const uint8_t *in = d_in + d.in_off;
    uint32_t len = d.in_len;
    const uint8_t *lut = lut1 + (size_t)d.o1_index * (256 * 4096);
    const uint32_t *fc = fc1 + (size_t)d.o1_index * 65536;
    uint32_t q = n >> 2;
    uint8_t *out = d_out + d.out_off + (uint64_t)k * q;
    uint32_t mylen = (k == 3) ? n - 3 * q : q;
    bool under = false;

    uint32_t cur = have ? meta[blk].pay_off : 0;
    uint32_t x = 0;
    if (have) {
        x = ld_byte(in, cur + 4 * k, len, under) | (ld_byte(in, cur + 4 * k + 1, len, under) << 8) |
            (ld_byte(in, cur + 4 * k + 2, len, under) << 16) | (ld_byte(in, cur + 4 * k + 3, len, under) << 24);
        cur += 16;
    }
    uint32_t steps = n - 3 * q;
    uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, steps);
    uint32_t l = 0;
    for (uint32_t t = 0; t < max_steps; t++) {
        bool act = t < mylen;
        if (act) {
            uint32_t m = x & (TOTFREQ - 1);
            uint32_t s = lut[l * 4096 + m];
            uint32_t e = fc[l * 256 + s];
            out[t] = (uint8_t)s;
            x = (e >> 16) * (x >> TF_SHIFT) + m - (e & 0xFFFF);
            l = s;
        }
        x = renorm_step(x, act, in, len, cur, lane, under);
    }
    if (have && under) status[blk] = CRAMGPU_ERR_CORRUPT;
}

// ------------------------------------------------------------------------------------------------
// host

// RANS.uncompress prefix handling [RANS.java:29-44]
int rans_parse_prefix(const uint8_t *p, uint32_t len, uint32_t *order, uint32_t *raw)
{
    *order = 0; *raw = 0;
    if (len == 0) return CRAMGPU_OK;
    if (len < 9) return CRAMGPU_ERR_CORRUPT;                 // Java: BufferUnderflowException
    uint32_t in_sz = p[1] | (p[2] << 8) | (p[3] << 16) | ((uint32_t)p[4] << 24);
    if (in_sz != len - 9) return CRAMGPU_ERR_INPUT_LENGTH;   // in_sz != remaining - RAW_BYTE_LENGTH
    *order = p[0];
    *raw = p[5] | (p[6] << 8) | (p[7] << 16) | ((uint32_t)p[8] << 24);
    if (*order > 1) return CRAMGPU_ERR_UNKNOWN_ORDER;
    return CRAMGPU_OK;
}

#include <algorithm>

int rans_decode_device(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                       const uint32_t *in_len, const uint8_t *in_prefix, uint8_t *d_out, uint64_t out_cap,
                       uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, cudaStream_t st)
{
    if (nblocks == 0) return CRAMGPU_OK;
    if (!d_in || !in_off || !in_len || !in_prefix || !d_out || !d_out_off || !d_out_len || !d_status)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");

    std::vector<BlockDesc> desc(nblocks);
    std::vector<int32_t> hstatus(nblocks, 0);
    std::vector<uint64_t> hoff(nblocks + 1);
    std::vector<uint32_t> hlen(nblocks);
    std::vector<uint32_t> s0, s1;
    uint64_t out_total = 0;
    uint32_t n1 = 0;
    for (uint32_t i = 0; i < nblocks; i++) {
        BlockDesc &d = desc[i];
        d = BlockDesc();
        d.in_off = in_off[i];
        d.in_len = in_len[i];
        uint32_t ord, raw;
        int rc = rans_parse_prefix(in_prefix + 9 * (size_t)i, in_len[i], &ord, &raw);
        hstatus[i] = rc;
        hoff[i] = out_total;
        d.order = 0xFF;
        if (rc != CRAMGPU_OK || in_len[i] == 0) { hlen[i] = 0; continue; }
        hlen[i] = raw;
        if (out_total + rup(raw, 16) > out_cap) {
            hstatus[i] = CRAMGPU_ERR_OUT_TOO_SMALL; hlen[i] = 0;
            continue;
        }
        d.order = ord;
        d.out_len = raw;
        d.out_off = out_total;
        out_total += rup(raw, 16);
        if (ord) { d.o1_index = n1++; s1.push_back(i); } else s0.push_back(i);
    }
    hoff[nblocks] = out_total;
    auto by_len = [&](uint32_t a, uint32_t b) { return desc[a].out_len > desc[b].out_len; };
    std::stable_sort(s0.begin(), s0.end(), by_len);
    std::stable_sort(s1.begin(), s1.end(), by_len);
    const uint32_t n0 = (uint32_t)s0.size();

    Carver c;
    size_t o_desc = c.take(sizeof(BlockDesc) * nblocks);
    size_t o_s0 = c.take(4 * (size_t)std::max(n0, 1u)), o_s1 = c.take(4 * (size_t)std::max(n1, 1u));
    size_t o_meta = c.take(sizeof(DecMeta) * (size_t)nblocks);
    size_t o_lut0 = c.take(4096 * (size_t)nblocks), o_fc0 = c.take(1024 * (size_t)nblocks);
    size_t o_lut1 = c.take((size_t)n1 * 256 * 4096 + 16), o_fc1 = c.take((size_t)n1 * 65536 * 4 + 16);
    if (int rc = pool_reserve(ctx, ctx->scratch, c.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->scratch.p;

    size_t meta_bytes = sizeof(BlockDesc) * nblocks + 4 * (size_t)(n0 + n1) + 16 * (size_t)(nblocks + 1) + 128;
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, meta_bytes)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));
    uint8_t *pm = (uint8_t *)ctx->pin_meta;
    size_t po = 0;
    auto up = [&](const void *src, size_t bytes, void *dst) -> cudaError_t {
        if (!bytes) return cudaSuccess;
        memcpy(pm + po, src, bytes);
        cudaError_t e = cudaMemcpyAsync(dst, pm + po, bytes, cudaMemcpyHostToDevice, st);
        po += rup(bytes, 16);
        return e;
    };
    CUDA_TRY(ctx, up(desc.data(), sizeof(BlockDesc) * nblocks, S + o_desc));
    CUDA_TRY(ctx, up(s0.data(), 4 * (size_t)n0, S + o_s0));
    CUDA_TRY(ctx, up(s1.data(), 4 * (size_t)n1, S + o_s1));
    CUDA_TRY(ctx, up(hstatus.data(), 4 * (size_t)nblocks, d_status));
    CUDA_TRY(ctx, up(hoff.data(), 8 * (size_t)(nblocks + 1), d_out_off));
    CUDA_TRY(ctx, up(hlen.data(), 4 * (size_t)nblocks, d_out_len));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));

    const BlockDesc *dd = (const BlockDesc *)(S + o_desc);
    DecMeta *meta = (DecMeta *)(S + o_meta);
    uint8_t *lut0 = S + o_lut0, *lut1 = S + o_lut1;
    uint32_t *fc0 = (uint32_t *)(S + o_fc0), *fc1 = (uint32_t *)(S + o_fc1);
    if (n0 + n1) {
        LAUNCH(ctx, st, "k_dec_tables", k_dec_tables<<<nblocks, 128, 0, st>>>(dd, d_in, lut0, fc0, lut1, fc1, meta, d_status));
        if (n1) {
            LAUNCH(ctx, st, "k_dec1", k_dec1<<<(n1 + 7) / 8, 32, 0, st>>>(dd, (const uint32_t *)(S + o_s1), n1, d_in, lut1, fc1, meta, d_out, d_status));
        }
        if (n0) {
            LAUNCH(ctx, st, "k_dec0", k_dec0<<<(n0 + 7) / 8, 32, 0, st>>>(dd, (const uint32_t *)(S + o_s0), n0, d_in, lut0, fc0, meta, d_out, d_status));
        }
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}
