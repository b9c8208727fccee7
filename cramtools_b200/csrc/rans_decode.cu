// rans_decode.cu -- htsjdk-compatible rANS order-0 / order-1 DECODE for sm_100a.
//
// Restatement of RANS.uncompress [RANS.java:29-55] as a batch pipeline:
//
//   (host)        9-byte prefix: order, compressed size check, raw size        [RANS.java:31-44]
//   k_dec_tables  frequency table -> decoder tables                             Frequencies.readStats_o0 [:13-46], _o1 [:50-106]
//   k_dec0        4-way interleaved decoder, one lane per state                 D04.uncompress [D04.java:9-85]
//   k_dec1        4 quarter-lanes, context = previously decoded byte            D14.uncompress [D14.java:9-70]
//   k_dec1_big    same, for blocks whose alphabet does not fit the shared-memory tables
//
// The four states of a block renormalise from ONE byte cursor in the order 0,1,2,3
// [D04.java:36-39, Decoding.java:84-90]; per step each lane needs 0..2 bytes, and its read position is
// the cursor plus the bytes taken by the lower lanes of its group (two ballots + popc).
//
// Every table lookup of the decoder depends on the state, so nothing can be prefetched: the tables must be
// on-chip.  Order 0: the 4 KB slot->symbol table and 1 KB {freq,start} table of each of the warp's 8 blocks
// are copied to shared memory.  Order 1 would need 4 KB per context (168 KB for a 42-symbol quality block),
// so it uses a compact form instead (232 B per context, blocks with <= 48 distinct bytes): a 128-entry table
// slot >> 5 -> first candidate symbol rank, plus the cumulative frequencies of the ranks; the symbol is the
// candidate advanced by the number of range ends <= slot among the next three (a loop finishes the rare longer
// runs, out of line).  Profiles shaped the kernel (profiles/r1_k_dec1_v*.txt): a warp alone on its SM sub-partition
// issues one dependent instruction every ~4-5 cycles, so the decoder warp keeps only the recurrence (48 instructions per
// step, a chain of 13 ALU operations and 2 shared-memory loads) and a helper warp per CTA stages the compressed stream
// (cp.async into a byte ring per group) and writes the output (aligned 16-byte stores); see k_dec1.
#include "internal.cuh"

#define DEC_MAXU 48          // alphabet limit of the shared-memory order-1 path
#define DEC_ROWB 232         // bytes per context row: coarse[128] u8 | cum[52] u16 (cum[0] = 0, cum[r+1] = end of rank r, rest 0xFFFF)
#define DEC_CHUNK 16         // decoder steps per stream-staging round (<= 8 bytes per step and group)
#define DEC_RING 128         // stream ring words per group

struct DecMeta {
    uint32_t pay_off;   // offset of the 16 state bytes inside the compressed block
    uint32_t small;     // order-1: 1 when the compact tables were built
};

// compact order-1 decoder tables of one block (global scratch; copied to shared memory by k_dec1)
struct DecCompact {
    uint8_t rows[DEC_MAXU][DEC_ROWB];   // per context rank: coarse | cum
    uint8_t ualpha[DEC_MAXU];           // rank -> byte
    uint32_t nU, pad[3];
};

__device__ __forceinline__ void cp_async4(void *smem, const void *g)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(g));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// grid nblocks, 128 threads.  Thread 0 walks the table bytes (inherently sequential: every field's
// position depends on the previous ones); all threads then expand each row.
//   order 0 and large-alphabet order 1:
//     lut [ctx][4096] u8   slot -> symbol                (Arrays.fill(R, x, x+F, j), [Frequencies.java:31])
//     fc  [ctx][256]  u32  freq << 16 | cumulative start
//   order 1 with <= 64 distinct bytes (contexts and symbols together): DecCompact
__global__ void __launch_bounds__(128) k_dec_tables(const BlockDesc *__restrict__ desc, const uint8_t *__restrict__ d_in,
                                                    uint8_t *__restrict__ lut0, uint32_t *__restrict__ fc0,
                                                    uint8_t *__restrict__ lut1, uint32_t *__restrict__ fc1,
                                                    DecCompact *__restrict__ compact,
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
    DecCompact *cp = compact + d.o1_index;

    __shared__ uint16_t F[256];
    __shared__ uint16_t C[257];
    __shared__ uint8_t symlist[256];
    __shared__ uint8_t present[256];
    __shared__ uint8_t rank[256];
    __shared__ uint32_t wcnt[4];
    __shared__ uint32_t s_nsym, s_pos, s_ctx, s_more, s_bad, s_nU;
    uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // parser state (thread 0 only)
    uint32_t pos = 9, ctx = 0;
    int rle_i = 0;
    bool bad = false;
    auto getb = [&]() -> uint32_t { if (pos >= len) { bad = true; return 0; } return in[pos++]; };
    auto peekb = [&]() -> int { if (pos >= len) { bad = true; return -1; } return in[pos]; };
    // one row: readStats_o0 [:13-46] / inner loop of readStats_o1 [:62-94].  collect: only mark the bytes seen.
    auto parse_row = [&](bool collect) -> uint32_t {
        uint32_t x = 0, ns = 0;
        int rle = 0;
        uint32_t j = getb();
        do {
            uint32_t f = getb();
            if (f >= 128) f = ((f & 127) << 8) | getb();
            if (o1 && f == 0) f = TOTFREQ;                                   // [:68-69]
            if (x + f > TOTFREQ || ns >= 256) { bad = true; break; }          // Java: AIOOBE in Arrays.fill
            if (collect) present[j] = 1;
            else { F[j] = (uint16_t)f; C[j] = (uint16_t)x; symlist[ns] = (uint8_t)j; }
            ns++;
            x += f;
            if (rle == 0 && peekb() == (int)j + 1) { j = getb(); rle = (int)getb(); }
            else if (rle != 0) { rle--; j++; if (j > 255) { bad = true; break; } }
            else j = getb();
            if (bad) break;
        } while (j != 0);
        return ns;
    };
    auto next_ctx = [&]() {                                                   // [:96-105]
        if (rle_i == 0 && peekb() == (int)ctx + 1) { ctx = getb(); rle_i = (int)getb(); }
        else if (rle_i != 0) { rle_i--; ctx++; if (ctx > 255) bad = true; }
        else ctx = getb();
    };

    // ---- order 1, pass A: which bytes occur (as context or symbol)?  -> ranks
    bool small = false;
    if (o1) {
        present[tid] = 0; present[tid + 128] = 0;
        __syncthreads();
        if (tid == 0) {
            ctx = getb();
            do {
                present[ctx] = 1;
                parse_row(true);
                if (bad) break;
                next_ctx();
            } while (!bad && ctx != 0);
            s_bad = bad ? 1u : 0u;
            pos = 9; ctx = 0; rle_i = 0;      // rewind for pass B
        }
        __syncthreads();
        if (!s_bad) {
            // rank of byte b = number of present bytes below b; thread handles bytes 2*tid, 2*tid+1
            uint32_t p0 = present[2 * tid], p1 = present[2 * tid + 1];
            uint32_t c = p0 + p1;
            uint32_t ex = c;
            for (int dlt = 1; dlt < 32; dlt <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, ex, dlt); if (lane >= (uint32_t)dlt) ex += t; }
            if (lane == 31) wcnt[w] = ex;
            __syncthreads();
            uint32_t b0 = ex - c;
            for (uint32_t k = 0; k < w; k++) b0 += wcnt[k];
            rank[2 * tid] = (uint8_t)b0; rank[2 * tid + 1] = (uint8_t)(b0 + p0);
            if (tid == 127) s_nU = b0 + c;
            __syncthreads();
            small = s_nU <= DEC_MAXU;
            if (small) {
                if (p0) cp->ualpha[b0] = (uint8_t)(2 * tid);
                if (p1) cp->ualpha[b0 + p0] = (uint8_t)(2 * tid + 1);
                if (tid == 0) cp->nU = s_nU;
                // contexts that never get a row: coarse 0, cumulative ends 0xFFFF except cum[0] = 0 (only corrupt streams get here)
                for (uint32_t i = tid; i < DEC_MAXU * DEC_ROWB; i += 128) {
                    uint32_t o = i % DEC_ROWB;
                    (&cp->rows[0][0])[i] = (o < 130) ? 0 : 0xFF;
                }
            }
        }
        __syncthreads();
        if (tid == 0 && !s_bad) ctx = getb();
    }

    // ---- pass B: parse each row, expand it
    if (!(o1 && s_bad)) for (;;) {
        F[tid] = 0; F[tid + 128] = 0;
        __syncthreads();
        if (tid == 0) {
            uint32_t ns = parse_row(false);
            s_nsym = ns; s_ctx = ctx;
            if (o1 && !bad) next_ctx();
            s_more = (o1 && !bad && ctx != 0) ? 1u : 0u;
            s_bad = bad ? 1u : 0u;
            s_pos = pos;
        }
        __syncthreads();
        if (s_bad) break;
        if (small) {
            // cumulative ends over byte ranks (warp 0, lane owns ranks 2*lane, 2*lane+1), then the coarse table
            uint8_t *row = cp->rows[rank[s_ctx]];
            uint16_t *cum = reinterpret_cast<uint16_t *>(row + 128);
            if (w == 0) {
                uint32_t r0 = 2 * lane, r1 = r0 + 1, nU = s_nU;
                uint32_t f0 = r0 < nU ? F[cp->ualpha[r0]] : 0, f1 = r1 < nU ? F[cp->ualpha[r1]] : 0;
                uint32_t inc = f0 + f1;
                for (int dlt = 1; dlt < 32; dlt <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, dlt); if (lane >= (uint32_t)dlt) inc += t; }
                if (r0 < nU) { cum[r0 + 1] = (uint16_t)(inc - f1); C[r0] = (uint16_t)(inc - f1); }
                if (r1 < nU) { cum[r1 + 1] = (uint16_t)inc; C[r1] = (uint16_t)inc; }     // C[] reused: range ends by rank
                if (lane == 0) cum[0] = 0;
                for (uint32_t r = nU + 1 + lane; r < 52; r += 32) cum[r] = 0xFFFF;
            }
            __syncthreads();
            {
                // coarse[b] = first rank whose range reaches beyond slot 32*b (clamped to the last rank)
                uint32_t nU = s_nU, slot = 32 * tid, r = 0;
                while (r + 1 < nU && C[r] <= slot) r++;
                row[tid] = (uint8_t)r;
            }
        } else {
            uint8_t *row_lut = lut + (size_t)s_ctx * 4096 * (o1 ? 1 : 0);
            uint32_t *row_fc = fc + (size_t)s_ctx * 256 * (o1 ? 1 : 0);
            uint32_t ns = s_nsym;
            for (uint32_t a = w; a < ns; a += 4) {       // one warp per symbol
                uint32_t j = symlist[a], f = F[j], c = C[j];
                if (lane == 0) row_fc[j] = (f << 16) | c;
                for (uint32_t s = lane; s < f; s += 32) row_lut[c + s] = (uint8_t)j;
            }
        }
        __syncthreads();
        if (!s_more) break;
    }
    if (tid == 0) {
        meta[blk].pay_off = s_pos;
        meta[blk].small = small ? 1u : 0u;
        if (s_bad) status[blk] = CRAMGPU_ERR_CORRUPT;
    }
}

// ---- the compressed stream of a group, staged in shared memory -----------------------------------------
// Byte ring per group (DEC1_RINGB bytes), addressed by the byte offset inside the block, filled by 4-byte cp.async.
struct Stream {
    const uint8_t *base;      // block start in global memory
    uint32_t len4;            // block length rounded up to 4
    uint32_t wfetch;          // first word not yet requested
};

// k_dec1: lane k decodes the quarter starting at k*(n>>2); lane 3 also the remainder [D14.java:62-68].
// Shared-memory path for blocks with <= DEC_MAXU distinct bytes.  Two warps per CTA:
//   warp 0  decoder: lane = state, 8 blocks; per step only the recurrence itself -- table search, state update,
//           renormalisation from the group's byte ring -- and one byte store (the symbol RANK) into the out ring;
//   warp 1  helper: keeps the 8 byte rings filled (cp.async), and turns the ranks of finished chunks into bytes that
//           go to global memory as aligned 16-byte stores.
// The warps meet once per chunk of DEC1_CH steps through two counters in shared memory (release / acquire).
// Dynamic shared memory, see Dec1Smem.
#define DEC1_CH 32                   // steps per chunk (<= 8 compressed bytes per step and group: 256 bytes per chunk)
#define DEC1_SLOTS 4                 // out-ring chunks
#define DEC1_RINGB 1024              // compressed-byte ring per group
struct __align__(16) Dec1Smem {
    uint8_t ring[8][DEC1_RINGB];                         // compressed bytes, per group (ring-size aligned)
    uint8_t outr[DEC1_SLOTS][DEC1_CH][32];               // symbol ranks: [chunk slot][step][decoder lane]
    uint8_t rows[8][(DEC_MAXU + 1) * DEC_ROWB + 8];      // +1 row: a corrupt table can point one row past the alphabet; 16-byte multiple
    uint8_t alpha[8][DEC_MAXU + 16];
    uint32_t cur[8];                                     // group cursors at the end of the last finished chunk
    int32_t ready, done;                                 // last chunk decoded / last helper round finished
};
static_assert(((DEC_MAXU + 1) * DEC_ROWB + 8) % 16 == 0 && (DEC_MAXU * DEC_ROWB) % 16 == 0, "table rows are copied as 16-byte vectors");
#define DEC1_SMEM (sizeof(Dec1Smem) + DEC1_RINGB)

__device__ __forceinline__ void st_release_s32(int32_t *p, int32_t v)
{
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_s32(const int32_t *p)
{
    int32_t v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint32_t prmt_raw(uint32_t a, uint32_t b, uint32_t sel)     // bytes of {b:a} picked by the nibbles of sel
{
    uint32_t v;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(v) : "r"(a), "r"(b), "r"(sel));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}

// table reads (immutable while the kernel runs: the compiler may schedule them freely)
__device__ __forceinline__ uint32_t lds_tab_u8(uint32_t saddr)
{
    uint32_t v;
    asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint32_t lds_tab_u32(uint32_t saddr)
{
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint32_t lds_tab_u16(uint32_t saddr)
{
    uint32_t v;
    asm("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}

// ---- the two rare cases of a decoder step, out of line so that the common path is straight-line code ----
// More than three short ranges in a row: lanes with a3 know their rank is >= first + 3 (C = start of first + 3) and
// search on, four cumulative ends per round; the rows end in 0xFFFF padding.  -> C | E << 16 | rank << 32
__device__ __noinline__ uint64_t dec1_long_search(uint32_t ca, uint32_t m, bool a3, uint32_t C, uint32_t E, uint32_t r)
{
    uint32_t pa = ca + 128 + 8;               // cum[first + 4]
    bool go = a3;
    while (__any_sync(0xFFFFFFFFu, go)) {
        const uint32_t e0 = lds_tab_u16(pa), e1 = lds_tab_u16(pa + 2), e2 = lds_tab_u16(pa + 4), e3 = lds_tab_u16(pa + 6);
        if (go) {
            const bool g0 = m >= e0, g1 = m >= e1, g2 = m >= e2, g3 = m >= e3;     // sorted: a prefix of them holds
            r += (uint32_t)g0 + (uint32_t)g1 + (uint32_t)g2 + (uint32_t)g3;
            C = g3 ? e3 : g2 ? e2 : g1 ? e1 : g0 ? e0 : C;
            E = g3 ? 0u : g2 ? e3 : g1 ? e2 : g0 ? e1 : e0;
            go = g3;
        }
        pa += 8;
    }
    return (uint64_t)(C & 0xFFFFu) | ((uint64_t)(E & 0xFFFFu) << 16) | ((uint64_t)r << 32);
}
// a lane needs two bytes: the general renormalisation.  -> x | cursor << 32
__device__ __noinline__ uint64_t dec1_renorm2(uint32_t x, bool n1, bool n2, uint32_t b1, uint32_t cur, uint32_t lm, uint32_t gm, uint32_t rb)
{
    const uint32_t b2 = __ballot_sync(0xFFFFFFFFu, n2);
    const uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);
    const uint32_t v0 = lds_u8(rb | (at & (DEC1_RINGB - 1))), v1 = lds_u8(rb | ((at + 1) & (DEC1_RINGB - 1)));
    const uint32_t x1 = (x << 8) | v0;
    x = n2 ? ((x1 << 8) | v1) : (n1 ? x1 : x);
    cur += __popc(b1 & gm) + __popc(b2 & gm);
    return (uint64_t)x | ((uint64_t)cur << 32);
}

__global__ void __launch_bounds__(64) k_dec1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                             uint32_t nslots, const uint8_t *__restrict__ d_in,
                                             const DecCompact *__restrict__ compact,
                                             const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                             int32_t *__restrict__ status)
{
    extern __shared__ uint8_t smem_raw[];
    Dec1Smem &S = *reinterpret_cast<Dec1Smem *>(smem_raw + ((DEC1_RINGB - ((uint32_t)__cvta_generic_to_shared(smem_raw) & (DEC1_RINGB - 1))) & (DEC1_RINGB - 1)));
    const uint32_t lane = threadIdx.x & 31, k = lane & 3u, g = lane >> 2;
    // warp role as a value the compiler knows to be warp-uniform (votes inside the role branches then need no divergence checks)
    const bool helper = __reduce_max_sync(0xFFFFFFFFu, threadIdx.x >> 5) != 0;
    const uint32_t slot = blockIdx.x * 8 + g;
    bool have = slot < nslots;
    const uint32_t blk = have ? slots[slot] : 0;
    const BlockDesc d = desc[blk];
    have = have && status[blk] == 0 && meta[blk].small == 1;
    if (__ballot_sync(0xFFFFFFFFu, have) == 0) return;            // both warps see the same flags
    const uint32_t n = have ? d.out_len : 0;
    const uint32_t q = n >> 2;
    uint8_t *out = d_out + d.out_off + (uint64_t)k * q;
    const uint32_t mylen = (k == 3) ? n - 3 * q : q;
    const uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, n - 3 * q);
    const int32_t nchunks = (int32_t)((max_steps + DEC1_CH - 1) / DEC1_CH);

    // tables of the 8 blocks (both warps copy), first ring fill and the 4 states (decoder)
    for (uint32_t gg = 0; gg < 8; gg++) {
        const uint32_t hv = __shfl_sync(0xFFFFFFFFu, (uint32_t)have, gg * 4);
        const uint32_t oi = __shfl_sync(0xFFFFFFFFu, d.o1_index, gg * 4);
        if (!hv) {
            // no block in this group: its lanes still run the loop, on empty contexts (coarse 0, cum[0] = 0, the ends 0xFFFF)
            for (uint32_t i = threadIdx.x; i < DEC_MAXU * DEC_ROWB; i += 64) S.rows[gg][i] = (i % DEC_ROWB) < 130 ? 0 : 0xFF;
            continue;
        }
        const DecCompact *cp = compact + oi;
        const uint4 *src = reinterpret_cast<const uint4 *>(&cp->rows[0][0]);
        uint4 *dst = reinterpret_cast<uint4 *>(&S.rows[gg][0]);
        for (uint32_t i = threadIdx.x; i < DEC_MAXU * DEC_ROWB / 16; i += 64) dst[i] = __ldg(src + i);
        for (uint32_t i = threadIdx.x; i < DEC_MAXU; i += 64) S.alpha[gg][i] = cp->ualpha[i];
    }
    // the row past the alphabet: an empty context (coarse 0, cum[0] = 0, the ends 0xFFFF), like the unused rows k_dec_tables writes
    for (uint32_t i = threadIdx.x; i < 8 * DEC_ROWB; i += 64) S.rows[i / DEC_ROWB][DEC_MAXU * DEC_ROWB + i % DEC_ROWB] = (i % DEC_ROWB) < 130 ? 0 : 0xFF;
    if (threadIdx.x == 0) { S.ready = -1; S.done = -1; }
    Stream st;
    st.base = d_in + d.in_off;
    st.len4 = have ? (d.in_len + 3) & ~3u : 0;
    const uint32_t cur0 = have ? meta[blk].pay_off : 0;
    st.wfetch = cur0 >> 2;
    uint8_t *ring = S.ring[g];
    if (helper) {
        // the ring holds the words [cur / 4, (cur + ring) / 4) of its group
        const uint32_t tgt = (cur0 + DEC1_RINGB) >> 2, lim = st.len4 >> 2;
        for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
            if (w < lim) cp_async4(ring + ((4 * w) & (DEC1_RINGB - 1)), st.base + 4ull * w);
        st.wfetch = tgt;
        cp_async_commit();
        cp_async_wait_all();
    }
    __syncthreads();

    if (!helper) {
        // ------------------------------------------------------------------------------ decoder warp
        uint32_t x = ring[(cur0 + 4 * k) & (DEC1_RINGB - 1)] | (ring[(cur0 + 4 * k + 1) & (DEC1_RINGB - 1)] << 8) |
                     (ring[(cur0 + 4 * k + 2) & (DEC1_RINGB - 1)] << 16) | ((uint32_t)ring[(cur0 + 4 * k + 3) & (DEC1_RINGB - 1)] << 24);
        uint32_t cur = cur0 + 16;
        // Wc = the next 4 bytes of the group's stream (bytes cur .. cur + 3), in a register
        uint32_t Wc = __funnelshift_r(lds_u32((uint32_t)__cvta_generic_to_shared(ring) | (cur & (DEC1_RINGB - 4))),
                                      lds_u32((uint32_t)__cvta_generic_to_shared(ring) | ((cur + 4) & (DEC1_RINGB - 4))), cur << 3);
        const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
        const uint32_t rb = (uint32_t)__cvta_generic_to_shared(ring);            // aligned to the ring size
        const uint32_t rows = (uint32_t)__cvta_generic_to_shared(&S.rows[g][0]);
        uint32_t row = rows;                  // context 0 is byte 0, rank 0 whenever the block has a context-0 row
        for (int32_t c = 0; c < nchunks; c++) {
            // the helper must be done with round c-2: stream bytes of this chunk staged (<= 256 per chunk), out slot free
            if (c >= 2) while (ld_acquire_s32(&S.done) < c - 2) {}
            uint8_t *orow = &S.outr[c & (DEC1_SLOTS - 1)][0][lane];
            const uint32_t t0 = (uint32_t)c * DEC1_CH;
            uint32_t rem = mylen > t0 ? mylen - t0 : 0;                          // steps this lane still owes
            for (uint32_t uo = 0; uo < DEC1_CH; uo += 8, rem = rem > 8 ? rem - 8 : 0, orow += 8 * 32) {
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) {
                    const bool act = u < rem;
                    const uint32_t m = x & (TOTFREQ - 1);
                    uint32_t r = lds_tab_u8(row + (m >> 5));                        // first candidate rank
                    const uint32_t ca = row + 2 * r;
                    const uint32_t c0 = lds_tab_u16(ca + 128), c1 = lds_tab_u16(ca + 130);   // start of r, end of r
                    const uint32_t c2 = lds_tab_u16(ca + 132), c3 = lds_tab_u16(ca + 134);   // ends of r+1, r+2
                    const bool a1 = m >= c1, a2 = m >= c2, a3 = m >= c3;
                    uint32_t C = a2 ? (a3 ? c3 : c2) : (a1 ? c1 : c0);
                    uint32_t E = a2 ? c3 : (a1 ? c2 : c1);
                    r += a2 ? (a3 ? 3u : 2u) : (a1 ? 1u : 0u);
                    uint32_t xn = (E - C) * (x >> TF_SHIFT) + m - C;                // [Decoding.java:35-48]; lanes past their end idle on garbage
                    // rare (out of line, tested after the common result is on its way): more than three short ranges in a row
                    if (__builtin_expect(__any_sync(0xFFFFFFFFu, a3), 0)) {
                        const uint64_t fix = dec1_long_search(ca, m, a3, C, E, r);
                        C = (uint32_t)fix & 0xFFFFu; E = ((uint32_t)fix >> 16) & 0xFFFFu; r = (uint32_t)(fix >> 32);
                        xn = (E - C) * (x >> TF_SHIFT) + m - C;
                    }
                    orow[u * 32] = (uint8_t)r;
                    row = rows + r * DEC_ROWB;                                      // context of the next symbol
                    // renormalise, lanes of a group in the order 0,1,2,3 [Decoding.java:84-90]; one byte is the common case
                    const bool n1 = act && xn < RANS_L;
                    const bool n2 = act && xn < (1u << 15);      // after one byte still below 2^23 -> a second byte
                    const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1);
                    // this lane's byte is byte (bytes taken by the lower lanes of the group) of the window Wc; one byte-permute
                    // shifts it into the state, or leaves the state alone (selector 0x3210) when the lane does not renormalise
                    x = prmt_raw(xn, Wc, (n1 ? 0x2104u : 0x3210u) + __popc(b1 & (n1 ? lm : 0u)));
                    if (__builtin_expect(__any_sync(0xFFFFFFFFu, n2), 0)) {
                        const uint64_t two = dec1_renorm2(xn, n1, n2, b1, cur, lm, gm, rb);
                        x = (uint32_t)two; cur = (uint32_t)(two >> 32);
                    } else {
                        cur += __popc(b1 & gm);
                    }
                    // the window for the next step (not needed before its renormalisation, a whole table search away)
                    Wc = __funnelshift_r(lds_u32(rb | (cur & (DEC1_RINGB - 4))), lds_u32(rb | ((cur + 4) & (DEC1_RINGB - 4))), cur << 3);
                }
            }
            if (k == 0) S.cur[g] = cur;
            __syncwarp();
            if (lane == 0) st_release_s32(&S.ready, c);
        }
        if (have && cur > d.in_len) status[blk] = CRAMGPU_ERR_CORRUPT;       // ran past the compressed bytes
    } else {
        // ------------------------------------------------------------------------------ helper warp
        const uint32_t ph = (uint32_t)((uintptr_t)out & 15);
        const uint8_t *alpha = S.alpha[g];
        const uint8_t *oring = &S.outr[0][0][lane];
        const uint32_t lim = st.len4 >> 2;
        for (int32_t c = 0; c <= nchunks; c++) {
            // round c: chunk c is decoded (round nchunks only flushes the bytes held back for alignment)
            if (c < nchunks) {
                while (ld_acquire_s32(&S.ready) < c) __nanosleep(100);
                // stream: the ring may now hold [cur, cur + ring) of its group
                const uint32_t tgt = (S.cur[g] + DEC1_RINGB) >> 2;
                for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
                    if (w < lim) cp_async4(ring + ((4 * w) & (DEC1_RINGB - 1)), st.base + 4ull * w);
                if (tgt > st.wfetch) st.wfetch = tgt;
                cp_async_commit();
            }
            // output: steps [32c - ph, 32c + 32 - ph) of this lane's quarter, as two aligned 16-byte granules
            const int64_t tb = (int64_t)c * DEC1_CH - ph;
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int64_t t0 = tb + 16 * j;
                if (t0 + 16 <= 0 || t0 >= (int64_t)mylen) continue;
                uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const int64_t t = t0 + i;
                    const uint32_t tt = (uint32_t)t & (DEC1_SLOTS * DEC1_CH - 1);
                    const uint32_t b = (t >= 0 && t < (int64_t)mylen) ? alpha[oring[tt * 32]] : 0;
                    w[i >> 2] |= b << (8 * (i & 3));
                }
                if (t0 >= 0 && t0 + 16 <= (int64_t)mylen) *reinterpret_cast<uint4 *>(out + t0) = make_uint4(w[0], w[1], w[2], w[3]);
                else {
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const int64_t t = t0 + i;
                        if (t >= 0 && t < (int64_t)mylen) out[t] = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
                    }
                }
            }
            if (c < nchunks) {
                cp_async_wait_all();
                __syncwarp();
                if (lane == 0) st_release_s32(&S.done, c);
            }
        }
    }
}

// k_dec0: the order-0 decoder in the two-warp form of k_dec1 (decoder warp + helper warp, 8 blocks per CTA).
// The four states of a block take the bytes 4t + k of the output in turn [D04.java:36-58], so a group's ranks of one
// step are four consecutive output bytes and the helper writes one contiguous stream per group.
struct __align__(16) Dec0Smem {
    uint8_t ring[8][DEC1_RINGB];                         // compressed bytes, per group (ring-size aligned)
    uint8_t outr[DEC1_SLOTS][DEC1_CH][32];               // decoded bytes: [chunk slot][step][decoder lane]
    uint8_t lut[8][4096];                                // slot -> symbol
    uint32_t fc[8][256];                                 // symbol -> freq << 16 | start
    uint32_t cur[8];
    int32_t ready, done;
};
#define DEC0_SMEM (sizeof(Dec0Smem) + DEC1_RINGB)

__global__ void __launch_bounds__(64) k_dec0(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                              uint32_t nslots, const uint8_t *__restrict__ d_in,
                                              const uint8_t *__restrict__ lut0, const uint32_t *__restrict__ fc0,
                                              const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                              int32_t *__restrict__ status)
{
    extern __shared__ uint8_t smem_raw[];
    Dec0Smem &S = *reinterpret_cast<Dec0Smem *>(smem_raw + ((DEC1_RINGB - ((uint32_t)__cvta_generic_to_shared(smem_raw) & (DEC1_RINGB - 1))) & (DEC1_RINGB - 1)));
    const uint32_t lane = threadIdx.x & 31, k = lane & 3u, g = lane >> 2;
    const bool helper = __reduce_max_sync(0xFFFFFFFFu, threadIdx.x >> 5) != 0;       // warp-uniform role (see k_dec1)
    const uint32_t slot = blockIdx.x * 8 + g;
    bool have = slot < nslots;
    const uint32_t blk = have ? slots[slot] : 0;
    const BlockDesc d = desc[blk];
    have = have && status[blk] == 0;
    if (__ballot_sync(0xFFFFFFFFu, have) == 0) return;
    const uint32_t n = have ? d.out_len : 0;
    const uint32_t steps = n >> 2;                                     // whole steps; the last n & 3 bytes come from the final states
    uint8_t *out = d_out + d.out_off;
    const uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, steps);
    const int32_t nchunks = (int32_t)((max_steps + DEC1_CH - 1) / DEC1_CH);

    for (uint32_t gg = 0; gg < 8; gg++) {
        const uint32_t hv = __shfl_sync(0xFFFFFFFFu, (uint32_t)have, gg * 4);
        const uint32_t b = __shfl_sync(0xFFFFFFFFu, blk, gg * 4);
        if (!hv) {
            // no block in this group: its lanes idle on a table that maps every slot to symbol 0 with frequency 4096
            for (uint32_t i = threadIdx.x; i < 1024; i += 64) reinterpret_cast<uint32_t *>(S.lut[gg])[i] = 0;
            for (uint32_t i = threadIdx.x; i < 256; i += 64) S.fc[gg][i] = 4096u << 16;
            continue;
        }
        const uint4 *src = reinterpret_cast<const uint4 *>(lut0 + (size_t)b * 4096);
        uint4 *dst = reinterpret_cast<uint4 *>(S.lut[gg]);
        for (uint32_t i = threadIdx.x; i < 256; i += 64) dst[i] = __ldg(src + i);
        for (uint32_t i = threadIdx.x; i < 256; i += 64) S.fc[gg][i] = __ldg(fc0 + (size_t)b * 256 + i);
    }
    if (threadIdx.x == 0) { S.ready = -1; S.done = -1; }
    Stream st;
    st.base = d_in + d.in_off;
    st.len4 = have ? (d.in_len + 3) & ~3u : 0;
    const uint32_t cur0 = have ? meta[blk].pay_off : 0;
    st.wfetch = cur0 >> 2;
    uint8_t *ring = S.ring[g];
    if (helper) {
        const uint32_t tgt = (cur0 + DEC1_RINGB) >> 2, lim = st.len4 >> 2;
        for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
            if (w < lim) cp_async4(ring + ((4 * w) & (DEC1_RINGB - 1)), st.base + 4ull * w);
        st.wfetch = tgt;
        cp_async_commit();
        cp_async_wait_all();
    }
    __syncthreads();

    if (!helper) {
        // ------------------------------------------------------------------------------ decoder warp
        uint32_t x = ring[(cur0 + 4 * k) & (DEC1_RINGB - 1)] | (ring[(cur0 + 4 * k + 1) & (DEC1_RINGB - 1)] << 8) |
                     (ring[(cur0 + 4 * k + 2) & (DEC1_RINGB - 1)] << 16) | ((uint32_t)ring[(cur0 + 4 * k + 3) & (DEC1_RINGB - 1)] << 24);
        uint32_t cur = cur0 + 16;
        const uint32_t rb = (uint32_t)__cvta_generic_to_shared(ring);
        uint32_t Wc = __funnelshift_r(lds_u32(rb | (cur & (DEC1_RINGB - 4))), lds_u32(rb | ((cur + 4) & (DEC1_RINGB - 4))), cur << 3);
        const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
        const uint32_t lutb = (uint32_t)__cvta_generic_to_shared(&S.lut[g][0]), fcb = (uint32_t)__cvta_generic_to_shared(&S.fc[g][0]);
        for (int32_t c = 0; c < nchunks; c++) {
            if (c >= 2) while (ld_acquire_s32(&S.done) < c - 2) {}
            uint8_t *orow = &S.outr[c & (DEC1_SLOTS - 1)][0][lane];
            const uint32_t t0 = (uint32_t)c * DEC1_CH;
            uint32_t rem = steps > t0 ? steps - t0 : 0;
            for (uint32_t uo = 0; uo < DEC1_CH; uo += 8, rem = rem > 8 ? rem - 8 : 0, orow += 8 * 32) {
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) {
                    const bool act = u < rem;
                    const uint32_t xold = x;
                    const uint32_t m = x & (TOTFREQ - 1);
                    const uint32_t s = lds_tab_u8(lutb + m);
                    const uint32_t e = lds_tab_u32(fcb + 4 * s);
                    orow[u * 32] = (uint8_t)s;
                    const uint32_t xn = (e >> 16) * (x >> TF_SHIFT) + m - (e & 0xFFFFu);      // [Decoding.java:35-48]
                    const bool n1 = act && xn < RANS_L;
                    const bool n2 = act && xn < (1u << 15);
                    const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1);
                    x = prmt_raw(xn, Wc, (n1 ? 0x2104u : 0x3210u) + __popc(b1 & (n1 ? lm : 0u)));
                    if (__builtin_expect(__any_sync(0xFFFFFFFFu, n2), 0)) {
                        const uint64_t two = dec1_renorm2(xn, n1, n2, b1, cur, lm, gm, rb);
                        x = (uint32_t)two; cur = (uint32_t)(two >> 32);
                    } else {
                        cur += __popc(b1 & gm);
                    }
                    x = act ? x : xold;                                     // a finished block keeps its final states: they give the tail bytes
                    Wc = __funnelshift_r(lds_u32(rb | (cur & (DEC1_RINGB - 4))), lds_u32(rb | ((cur + 4) & (DEC1_RINGB - 4))), cur << 3);
                }
            }
            if (k == 0) S.cur[g] = cur;
            __syncwarp();
            if (lane == 0) st_release_s32(&S.ready, c);
        }
        if (have && k < (n & 3u)) out[(uint64_t)steps * 4 + k] = S.lut[g][x & (TOTFREQ - 1)];   // tail [D04.java:60-83]
        if (have && cur > d.in_len) status[blk] = CRAMGPU_ERR_CORRUPT;
    } else {
        // ------------------------------------------------------------------------------ helper warp
        // output: the group's bytes [0, 4 * steps) are one stream; lane k of the group writes the k-th quarter of every
        // chunk's 128 bytes, shifted so that its stores are aligned 16-byte granules
        const uint32_t nbytes = steps * 4;
        const uint32_t ph = (uint32_t)((uintptr_t)out & 15);
        const uint8_t *oring = &S.outr[0][0][4 * g];
        const uint32_t lim = st.len4 >> 2;
        for (int32_t c = 0; c <= nchunks; c++) {
            if (c < nchunks) {
                while (ld_acquire_s32(&S.ready) < c) __nanosleep(100);
                const uint32_t tgt = (S.cur[g] + DEC1_RINGB) >> 2;
                for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
                    if (w < lim) cp_async4(ring + ((4 * w) & (DEC1_RINGB - 1)), st.base + 4ull * w);
                if (tgt > st.wfetch) st.wfetch = tgt;
                cp_async_commit();
            }
            // stream bytes [128 c - ph, 128 c + 128 - ph): 8 granules, two per lane of the group
            const int64_t tb = (int64_t)c * (4 * DEC1_CH) - ph + 32 * (int64_t)k;
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int64_t p0 = tb + 16 * j;
                if (p0 + 16 <= 0 || p0 >= (int64_t)nbytes) continue;
                uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const int64_t p = p0 + i;
                    // byte p of the stream = state p & 3 of step p >> 2
                    const uint32_t stp = (uint32_t)(p >> 2) & (DEC1_SLOTS * DEC1_CH - 1);
                    const uint32_t b = (p >= 0 && p < (int64_t)nbytes) ? oring[stp * 32 + ((uint32_t)p & 3u)] : 0;
                    w[i >> 2] |= b << (8 * (i & 3));
                }
                if (p0 >= 0 && p0 + 16 <= (int64_t)nbytes) *reinterpret_cast<uint4 *>(out + p0) = make_uint4(w[0], w[1], w[2], w[3]);
                else {
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const int64_t p = p0 + i;
                        if (p >= 0 && p < (int64_t)nbytes) out[p] = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
                    }
                }
            }
            if (c < nchunks) {
                cp_async_wait_all();
                __syncwarp();
                if (lane == 0) st_release_s32(&S.done, c);
            }
        }
    }
}

__device__ __forceinline__ uint32_t ld_byte(const uint8_t *p, uint32_t pos, uint32_t len, bool &under)
{
    if (pos >= len) { under = true; return 0; }
    return p[pos];
}

// k_dec1_big: general order-1 path (tables in global memory / L2), for alphabets above DEC_MAXU.
__global__ void __launch_bounds__(32) k_dec1_big(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
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
    have = have && status[blk] == 0 && meta[blk].small == 0;
    if (__ballot_sync(0xFFFFFFFFu, have) == 0) return;
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
        bool n1 = act && x < RANS_L;
        bool n2 = act && x < (1u << 15);
        uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1), b2 = __ballot_sync(0xFFFFFFFFu, n2);
        uint32_t g4 = lane & ~3u;
        uint32_t gm = 0xFu << g4;
        uint32_t lm = ((1u << k) - 1u) << g4;
        uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);
        if (n1) {
            x = (x << 8) | ld_byte(in, at, len, under);
            if (n2) x = (x << 8) | ld_byte(in, at + 1, len, under);
        }
        cur += __popc(b1 & gm) + __popc(b2 & gm);
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
                       uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, cudaStream_t st,
                       const uint64_t *dst_addr, const uint32_t *dst_cap)
{
    if (nblocks == 0) return CRAMGPU_OK;
    if (!in_off || !in_len || !in_prefix || !d_out || !d_out_off || !d_out_len || !d_status)   // d_in may be NULL: absolute addresses in in_off
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
        const bool placed = dst_addr && dst_addr[i];            // decoded straight to a device address of the caller's
        if (placed ? (dst_cap && raw > dst_cap[i]) : (out_total + rup(raw, 16) > out_cap)) {
            hstatus[i] = CRAMGPU_ERR_OUT_TOO_SMALL; hlen[i] = 0;
            continue;
        }
        d.order = ord;
        d.out_len = raw;
        if (placed) { d.out_off = dst_addr[i] - (uint64_t)(uintptr_t)d_out; hoff[i] = d.out_off; }       // d_out + out_off = the address
        else { d.out_off = out_total; out_total += rup(raw, 16); }
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
    size_t o_compact = c.take(sizeof(DecCompact) * (size_t)std::max(n1, 1u));
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
    DecCompact *compact = (DecCompact *)(S + o_compact);
    if (!ctx->dec_attr_done) {
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_dec1, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC1_SMEM));
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_dec0, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC0_SMEM));
        ctx->dec_attr_done = true;
    }
    if (n0 + n1) {
        LAUNCH(ctx, st, "k_dec_tables", k_dec_tables<<<nblocks, 128, 0, st>>>(dd, d_in, lut0, fc0, lut1, fc1, compact, meta, d_status));
        if (n1) {
            ON_HP_STREAM(ctx, st, LAUNCH(ctx, hs, "k_dec1", k_dec1<<<(n1 + 7) / 8, 64, DEC1_SMEM, hs>>>(dd, (const uint32_t *)(S + o_s1), n1, d_in, compact, meta, d_out, d_status)));
            LAUNCH(ctx, st, "k_dec1_big", k_dec1_big<<<(n1 + 7) / 8, 32, 0, st>>>(dd, (const uint32_t *)(S + o_s1), n1, d_in, lut1, fc1, meta, d_out, d_status));
        }
        if (n0) {
            LAUNCH(ctx, st, "k_dec0", k_dec0<<<(n0 + 7) / 8, 64, DEC0_SMEM, st>>>(dd, (const uint32_t *)(S + o_s0), n0, d_in, lut0, fc0, meta, d_out, d_status));
        }
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}
