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
#include <stdlib.h>

#define DEC_MAXU 64          // alphabet limit of the shared-memory order-1 path (distinct bytes, contexts and symbols together)
#define DEC_NB 32            // buckets per context row, 128 slots each
#define DEC_ROWE (DEC_NB * 8)                                  // bytes of bucket entries per context row
#define DEC_IMG_CUM (DEC_MAXU * DEC_ROWE)                      // table image in global scratch: entries | cumulative rows | alphabet
#define DEC_CUMB_MAX (2 * (DEC_MAXU + 8))
#define DEC_IMG_ALPHA (DEC_IMG_CUM + DEC_MAXU * DEC_CUMB_MAX)
#define DEC_IMG (DEC_IMG_ALPHA + 256)                          // bytes per order-1 block, a multiple of 256
#define DEC_CHUNK 16         // order 0: decoder steps per stream-staging round (<= 8 bytes per step and group)
#define DEC_RING 128         // order 0: stream ring words per group

// Order-1 decoder tables (built by k_dec_tables, copied to shared memory by k_dec1), per context row (row = rank of the
// context byte among the bytes that occur in the block):
//   32 bucket entries of 8 bytes, bucket = slot >> 7:
//     c0  start of the range that covers the bucket's first slot        f0  its frequency
//     f1  frequency of the next symbol that occurs in this context (0 = none)
//     r0, r1  ranks of the two symbols
//   so one group of loads at an address that depends on the state only finds the symbol whenever it is one of the first two
//   candidates of its bucket (99.2 % of the steps on quality scores); otherwise
//   cum[rank] u16: cumulative frequencies over ALL ranks (cum[0] = 0, cum[r + 1] = end of rank r), followed by >= 4
//   entries 0xFFFF, is searched from r1 + 1 on.
struct DecMeta {
    uint32_t pay_off;   // offset of the 16 state bytes inside the compressed block
    uint32_t nU;        // order 1: number of distinct bytes (contexts and symbols together), 0 = no tables (corrupt)
    uint32_t img;       // order 1: where the block's table image starts, in 256-byte units from `images`; DEC_NO_IMG = none
    uint32_t pad;
};
#define DEC_NO_IMG 0xFFFFFFFFu
// Table images of the blocks with more than DEC_MAXU distinct bytes ("wide": ITF8 series such as TS and NP of paired reads,
// names, tags) have nU rows of 256 FINE buckets (16 slots each, one byte = the rank whose range holds the bucket's first
// slot) | cumulative rows | alphabet, up to 201 KB, and live in a pool behind the fixed slots, handed out by an atomic
// cursor while k_dec_tables runs.  They are decoded from global memory by k_dec1s<true>.  A call with more wide blocks than
// the pool holds leaves the rest to k_dec1_big.
__host__ __device__ __forceinline__ uint32_t dec_wide_img_bytes(uint32_t nU) { return (nU * (DEC_ROWE + ((2u * (nU + 5u) + 7u) & ~7u)) + 256u + 255u) & ~255u; }
#define DEC_WIDE_IMG_MAX 200960u     // dec_wide_img_bytes(256)
// the pool holds ctx->dec_wide_pool_blocks (2048) images of the largest size (410 MB), or as many as the call has order-1 blocks

__host__ __device__ __forceinline__ uint32_t dec_cumb(uint32_t nU) { return (2u * (nU + 5u) + 7u) & ~7u; }          // bytes of one cumulative row
__host__ __device__ __forceinline__ uint32_t dec_tab_bytes(uint32_t nU) { return (nU * (DEC_ROWE + dec_cumb(nU)) + 255u) & ~255u; }

__device__ __forceinline__ void cp_async4(void *smem, const void *g)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(g));
}
__device__ __forceinline__ void cp_async4s(uint32_t saddr, const void *g)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
// Sequential reader of a frequency table (one thread walks the bytes: every field's position depends on the fields
// before it).  Frequencies.readStats_o0 [:13-46] / readStats_o1 [:50-106].
struct TabParser {
    const uint8_t *in;
    uint32_t len, pos, ctx;
    int rle_i;
    bool bad, o1;
    __device__ uint32_t getb() { if (pos >= len) { bad = true; return 0; } return in[pos++]; }
    __device__ int peekb() { if (pos >= len) { bad = true; return -1; } return in[pos]; }
    // one row: readStats_o0 [:13-46] / inner loop of readStats_o1 [:62-94].  present != nullptr: only mark the bytes seen.
    __device__ uint32_t parse_row(uint8_t *present, uint16_t *F, uint16_t *C, uint8_t *symlist)
    {
        uint32_t x = 0, ns = 0;
        int rle = 0;
        uint32_t j = getb();
        do {
            uint32_t f = getb();
            if (f >= 128) f = ((f & 127) << 8) | getb();
            if (o1 && f == 0) f = TOTFREQ;                                   // [:68-69]
            if (x + f > TOTFREQ || ns >= 256) { bad = true; break; }          // Java: AIOOBE in Arrays.fill
            if (present) present[j] = 1;
            else { F[j] = (uint16_t)f; C[j] = (uint16_t)x; symlist[ns] = (uint8_t)j; }
            ns++;
            x += f;
            if (rle == 0 && peekb() == (int)j + 1) { j = getb(); rle = (int)getb(); }
            else if (rle != 0) { rle--; j++; if (j > 255) { bad = true; break; } }
            else j = getb();
            if (bad) break;
        } while (j != 0);
        return ns;
    }
    __device__ void next_ctx()                                                // [:96-105]
    {
        if (rle_i == 0 && peekb() == (int)ctx + 1) { ctx = getb(); rle_i = (int)getb(); }
        else if (rle_i != 0) { rle_i--; ctx++; if (ctx > 255) bad = true; }
        else ctx = getb();
    }
};

// grid nblocks, 128 threads.  Thread 0 walks the table bytes (inherently sequential: every field's
// position depends on the previous ones); all threads then expand each row.
//   order 0:
//     lut [4096] u8   slot -> symbol                (Arrays.fill(R, x, x+F, j), [Frequencies.java:31])
//     fc  [256]  u32  freq << 16 | cumulative start
//   order 1 with <= DEC_MAXU distinct bytes (contexts and symbols together): the table image described above;
//   order 1 with more: only the alphabet size and the payload offset (k_dec1_big builds its tables itself)
__global__ void __launch_bounds__(128) k_dec_tables(const BlockDesc *__restrict__ desc, const uint8_t *__restrict__ d_in,
                                                    uint8_t *__restrict__ lut0, uint32_t *__restrict__ fc0,
                                                    uint8_t *__restrict__ images, uint32_t pool_first, uint32_t pool_units,
                                                    uint32_t *__restrict__ pool_cursor,
                                                    DecMeta *__restrict__ meta, int32_t *__restrict__ status)
{
    uint32_t blk = blockIdx.x;
    const BlockDesc d = desc[blk];
    if (d.order > 1) return;
    const uint8_t *in = d_in + d.in_off;
    const uint32_t len = d.in_len;
    const bool o1 = d.order == 1;
    uint8_t *lut = lut0 + (size_t)d.o1_index * 4096;          // order 0: o1_index = rank among the order-0 blocks
    uint32_t *fc = fc0 + (size_t)d.o1_index * 256;
    uint8_t *img = images + (size_t)d.o1_index * DEC_IMG;
    uint32_t img_units = d.o1_index * (DEC_IMG / 256u), cum_off = DEC_IMG_CUM, alpha_off = DEC_IMG_ALPHA;

    __shared__ uint16_t F[256];
    __shared__ uint16_t C[257];
    __shared__ uint8_t symlist[256];
    __shared__ uint8_t present[256];
    __shared__ uint8_t rank[256];
    __shared__ uint8_t ualpha[256];
    __shared__ uint32_t wcnt[4];
    __shared__ uint32_t s_nsym, s_pos, s_ctx, s_more, s_bad, s_nU, s_img;
    uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    TabParser P = {in, len, 9u, 0u, 0, false, o1};      // thread 0 only
    // ---- order 1, pass A: which bytes occur (as context or symbol)?  -> ranks
    bool small = false;
    uint32_t cumb = 0;
    if (o1) {
        present[tid] = 0; present[tid + 128] = 0;
        __syncthreads();
        if (tid == 0) {
            P.ctx = P.getb();
            do {
                present[P.ctx] = 1;
                P.parse_row(present, nullptr, nullptr, nullptr);
                if (P.bad) break;
                P.next_ctx();
            } while (!P.bad && P.ctx != 0);
            s_bad = P.bad ? 1u : 0u;
            s_pos = P.pos;                    // the payload follows the table's final 0
            P.pos = 9; P.ctx = 0; P.rle_i = 0;      // rewind for pass B
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
            small = true;
            const uint32_t nU = s_nU;
            cumb = dec_cumb(nU);
            if (nU > DEC_MAXU) {
                // a wide alphabet: its image comes out of the pool (or the block is left to k_dec1_big)
                if (tid == 0) {
                    const uint32_t units = dec_wide_img_bytes(nU) / 256u;
                    const uint32_t at = atomicAdd(pool_cursor, units);
                    s_img = at + units <= pool_units ? pool_first + at : DEC_NO_IMG;
                }
                __syncthreads();
                img_units = s_img;
                small = img_units != DEC_NO_IMG;
                img = images + (size_t)img_units * 256u;
                cum_off = nU * DEC_ROWE; alpha_off = cum_off + nU * cumb;
            }
            if (small) {
                if (p0) { ualpha[b0] = (uint8_t)(2 * tid); img[alpha_off + b0] = (uint8_t)(2 * tid); }
                if (p1) { ualpha[b0 + p0] = (uint8_t)(2 * tid + 1); img[alpha_off + b0 + p0] = (uint8_t)(2 * tid + 1); }
                // contexts that never get a row (only corrupt streams reach them): empty buckets, cum[0] = 0, the ends 0xFFFF
                uint32_t *e32 = reinterpret_cast<uint32_t *>(img);
                for (uint32_t i = tid; i < nU * (DEC_ROWE / 4); i += 128) e32[i] = 0;
                uint16_t *c16 = reinterpret_cast<uint16_t *>(img + cum_off);
                const uint32_t cw = cumb / 2;
                for (uint32_t i = tid; i < nU * cw; i += 128) c16[i] = (i % cw) ? 0xFFFF : 0;
            }
        }
        __syncthreads();
        if (tid == 0 && !s_bad) P.ctx = P.getb();
    }

    if (!o1) {
        // slots and symbols that a table does not cover decode as Java's zero-filled arrays do (symbol 0, frequency 0), not
        // as whatever an earlier call left in the scratch
        uint32_t *l32 = reinterpret_cast<uint32_t *>(lut);
        for (uint32_t i = tid; i < 1024; i += 128) l32[i] = 0;
        for (uint32_t i = tid; i < 256; i += 128) fc[i] = 0;
        __syncthreads();
    }
    // ---- pass B: parse each row, expand it (a wide block that found no room in the pool stops here: pass A found its payload)
    if (o1 && !small) {
        if (tid == 0) {
            meta[blk].pay_off = s_pos;
            meta[blk].nU = s_bad ? 0u : s_nU;
            meta[blk].img = DEC_NO_IMG;
            if (s_bad) status[blk] = CRAMGPU_ERR_CORRUPT;
        }
        return;
    }
    for (;;) {
        F[tid] = 0; F[tid + 128] = 0;
        __syncthreads();
        if (tid == 0) {
            uint32_t ns = P.parse_row(nullptr, F, C, symlist);
            s_nsym = ns; s_ctx = P.ctx;
            if (o1 && !P.bad) P.next_ctx();
            s_more = (o1 && !P.bad && P.ctx != 0) ? 1u : 0u;
            s_bad = P.bad ? 1u : 0u;
            s_pos = P.pos;
        }
        __syncthreads();
        if (s_bad) break;
        if (small) {
            // cumulative ends over byte ranks (warp 0, lane owns ranks 2*lane, 2*lane+1), then the bucket entries
            const uint32_t rc = rank[s_ctx], nU = s_nU;
            uint16_t *cum = reinterpret_cast<uint16_t *>(img + cum_off + (size_t)rc * cumb);
            if (w == 0) {
                uint32_t carry = 0;
                for (uint32_t rb = 0; rb < nU; rb += 64) {                   // 64 ranks per round, lane owns two
                    uint32_t r0 = rb + 2 * lane, r1 = r0 + 1;
                    uint32_t f0 = r0 < nU ? F[ualpha[r0]] : 0, f1 = r1 < nU ? F[ualpha[r1]] : 0;
                    uint32_t inc = f0 + f1;
                    for (int dlt = 1; dlt < 32; dlt <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, dlt); if (lane >= (uint32_t)dlt) inc += t; }
                    inc += carry;
                    if (r0 < nU) { cum[r0 + 1] = (uint16_t)(inc - f1); C[r0] = (uint16_t)(inc - f1); }
                    if (r1 < nU) { cum[r1 + 1] = (uint16_t)inc; C[r1] = (uint16_t)inc; }     // C[] reused: range ends by rank
                    carry = __shfl_sync(0xFFFFFFFFu, inc, 31);
                }
                if (lane == 0) cum[0] = 0;
            }
            __syncthreads();
            if (nU > DEC_MAXU) {
                // wide image: 256 fine buckets of 16 slots, one byte each = the rank whose range holds the bucket's first slot
                // (first r with C[r] > slot, at most nU - 1); the decoder walks the cumulative row from there
                uint8_t *fine = img + (size_t)rc * DEC_ROWE;
                for (uint32_t b = tid; b < 256; b += 128) {
                    const uint32_t slot = 16 * b;
                    uint32_t lo = 0, hi = nU - 1;
                    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (C[mid] <= slot) lo = mid + 1; else hi = mid; }
                    fine[b] = (uint8_t)lo;
                }
            } else if (tid < DEC_NB) {
                // bucket tid: r = first rank whose range reaches beyond slot 128 * tid (clamped to the last rank), q = the next rank that occurs
                uint32_t slot = 128 * tid, r = 0;
                while (r + 1 < nU && C[r] <= slot) r++;
                const uint32_t s0 = r ? C[r - 1] : 0, e0 = C[r];
                uint32_t q = r, f1 = 0;
                for (uint32_t t = r + 1; t < nU; t++) {
                    const uint32_t ft = (uint32_t)C[t] - C[t - 1];
                    if (ft) { q = t; f1 = ft; break; }
                }
                uint32_t *e = reinterpret_cast<uint32_t *>(img + (size_t)rc * DEC_ROWE + 8 * tid);
                e[0] = s0 | ((e0 - s0) << 16);
                e[1] = f1 | (r << 16) | (q << 24);
            }
        } else {
            uint8_t *row_lut = lut;
            uint32_t *row_fc = fc;
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
        meta[blk].nU = (small && !s_bad) ? s_nU : 0u;
        meta[blk].img = img_units;
        if (s_bad) status[blk] = CRAMGPU_ERR_CORRUPT;
    }
}

// ---- shared-memory access by 32-bit shared-window address ------------------------------------------------
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t saddr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void sts_u8(uint32_t saddr, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(saddr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_v4(uint32_t saddr, uint4 v)
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts_v2(uint32_t saddr, uint2 v) { asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(saddr), "r"(v.x), "r"(v.y) : "memory"); }
__device__ __forceinline__ void sts_u32(uint32_t saddr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory"); }
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
__device__ __forceinline__ uint32_t prmt_raw(uint32_t a, uint32_t b, uint32_t sel)     // bytes of {b:a} picked by the nibbles of sel
{
    uint32_t v;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(v) : "r"(a), "r"(b), "r"(sel));
    return v;
}
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

// ---- order 1 -----------------------------------------------------------------------------------------------
// Each of a block's 4 states is a strict recurrence, and every table lookup depends on the state: the time of a block is
// (steps) x (length of the dependent chain of one step), whatever else the GPU does.  k_dec1 therefore keeps that chain
// as short as it gets -- address from the state (2 ALU), ONE round of shared-memory loads (the bucket entry), candidate
// choice and state update (4 ALU), vote, population count and one byte-permute for the renormalisation -- and spends
// lanes freely: a warp decodes only `ng` blocks (4 lanes each), ng chosen by the host so that about one warp runs per SM
// sub-partition (a lone warp issues a dependent instruction every ~4 cycles; several warps per scheduler would contend for
// issue slots, fewer blocks per warp make the out-of-line cases rarer).  No helper warps: after every chunk of 32 steps
// the lanes write their own output (aligned 16-byte stores from a small ring) and refill their block's stream ring
// (cp.async, two chunks ahead).
//
// A CTA is only a container of shared memory: its warps never meet.  Which blocks a CTA holds is decided on the device
// (k_dec_plan), because the table size of a block (its alphabet) is known only after k_dec_tables: blocks are taken in
// the host's order (longest first) until the CTA's arena is full.
#define DEC1_CH 64                   // steps per chunk
#define DEC1_RING 512                // compressed-byte ring per block (a state takes at most 12 bits per step for long: a chunk stays below 400 bytes)
#define DEC1_NEED 416                // bytes that must be staged beyond the cursor when a chunk starts
#define DEC1_OUTH (DEC1_CH + 16)     // half of an output ring: one chunk + up to 15 bytes held back for alignment
#define DEC1_OUTQ (2 * DEC1_OUTH)    // output ring per quarter
#define DEC1_FIX (DEC1_RING + 4 * DEC1_OUTQ + DEC_MAXU)   // per block: stream ring, 4 output rings, alphabet
#define DEC1_ARENA 113664            // shared memory of one CTA: two CTAs per SM
#define DEC1_SMEM (DEC1_ARENA + 512)
#define DEC1_WARPS 4
#define PLAN_TILE 8192
static_assert(DEC1_FIX % 64 == 0, "per-block fixed part in 64-byte units");
static_assert(4 * (DEC1_FIX + DEC_MAXU * (DEC_ROWE + DEC_CUMB_MAX) + 256) + 272 <= DEC1_ARENA, "four blocks of any alphabet fit one CTA");

// k_dec_plan: one CTA.  Packs the order-1 blocks (in the order of `slots`) into CTAs of k_dec1:
//   cta_first[c]  first slot of CTA c          cta_cfg[c] = number of slots | blocks per warp << 8
// A CTA takes consecutive slots while their tables fit the arena and there are at most DEC1_WARPS * ng * ns of them
// (ns = blocks per lane of k_dec1); ng = ng_long for CTAs that start with a block of at least len_long bytes, 8 / ns for the others.
__global__ void __launch_bounds__(256) k_dec_plan(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots, uint32_t n,
                                                  const DecMeta *__restrict__ meta, const int32_t *__restrict__ status,
                                                  uint32_t ng_long, uint32_t len_long, uint32_t ns,
                                                  uint32_t *__restrict__ cta_first, uint32_t *__restrict__ cta_cfg, uint32_t *__restrict__ ncta)
{
    __shared__ uint16_t s_need[PLAN_TILE];       // shared memory of the block's rings and tables, in 64-byte units
    __shared__ uint8_t s_long[PLAN_TILE];
    __shared__ uint32_t s_i, s_c;
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) { s_i = 0; s_c = 0; }
    __syncthreads();
    for (;;) {
        const uint32_t base = s_i;
        if (base >= n) break;
        const uint32_t tile_n = min((uint32_t)PLAN_TILE, n - base);
        for (uint32_t j = tid; j < tile_n; j += 256) {
            const uint32_t blk = slots[base + j];
            const uint32_t nU = status[blk] == 0 ? meta[blk].nU : 0u;
            // every slot of a CTA has its rings, whether or not the block is decoded here (large alphabet, rejected)
            s_need[j] = (uint16_t)((DEC1_FIX + ((nU && nU <= DEC_MAXU) ? dec_tab_bytes(nU) : 0u)) / 64);
            s_long[j] = desc[blk].out_len >= len_long;
        }
        __syncthreads();
        if (tid < 32) {
            const bool last = base + tile_n >= n;
            uint32_t i = 0, c = s_c;
            while (i < tile_n && (last || i + 32 <= tile_n)) {
                const uint32_t idx = i + lane;
                const bool valid = idx < tile_n;
                const uint32_t need = valid ? s_need[idx] : 0u;
                const uint32_t ng = s_long[i] ? ng_long : 8u / ns;
                uint32_t incl = need;
#pragma unroll
                for (int dlt = 1; dlt < 32; dlt <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, dlt); if (lane >= (uint32_t)dlt) incl += t; }
                const bool ok = valid && incl * 64 + 272 <= DEC1_ARENA && lane < DEC1_WARPS * ng * ns;
                const uint32_t b = __ballot_sync(0xFFFFFFFFu, ok);
                uint32_t take = b == 0xFFFFFFFFu ? 32u : (uint32_t)__ffs(~b) - 1u;
                if (take == 0) take = 1;                   // cannot happen: any single block fits
                if (lane == 0) { cta_first[c] = base + i; cta_cfg[c] = take | (ng << 8); }
                c++;
                i += take;
            }
            if (lane == 0) { s_i = base + i; s_c = c; }
        }
        __syncthreads();
    }
    if (tid == 0) *ncta = s_c;
}

// the state of one decoder lane that a step changes
struct D1State {
    uint32_t x;        // rANS state
    uint32_t cur;      // the group's byte cursor
    uint32_t r;        // rank of the symbol just decoded = context of the next step
};

// The general step, out of line: any alphabet position, 0-2 renormalisation bytes, lanes past their end (act = false)
// idle.  Used for the steps where some lane of the warp misses both bucket candidates or needs two bytes, and for the
// tail chunks where the lanes' lengths differ.  -> x | cur << 32, and the rank through *rank_out in a register pair.
struct D1Tab { uint32_t ebase, cbase, cumb, nU, ring; };
__device__ __noinline__ uint4 dec1_general_step(uint32_t x, uint32_t cur, uint32_t r, bool act, uint32_t ebase, uint32_t cbase, uint32_t cumb,
                                                uint32_t nU, uint32_t ring, uint32_t lm, uint32_t gm)
{
    const uint32_t m = x & (TOTFREQ - 1);
    const uint32_t ea = ebase + (r << 8) + ((m >> 7) << 3);
    const uint32_t c0 = lds_tab_u16(ea), f0 = lds_tab_u16(ea + 2), f1 = lds_tab_u16(ea + 4);
    const uint32_t r0 = lds_tab_u8(ea + 6), r1 = lds_tab_u8(ea + 7);
    uint32_t C = c0, Fq = f0, rn = r0, bad = 0;
    if (m - c0 >= f0) {
        C = c0 + f0; Fq = f1; rn = r1;
        if (m - C >= f1) {
            // beyond both candidates: search the cumulative row from the rank after r1, four ends per round (the row ends in
            // 0xFFFF padding, and the ends are sorted: the ones <= m form a prefix)
            const uint32_t ca = cbase + r * cumb;
            uint32_t j = r1 + 1;
            for (;;) {
                const uint32_t pa = ca + 2 * j;
                const uint32_t e1 = lds_tab_u16(pa + 2), e2 = lds_tab_u16(pa + 4), e3 = lds_tab_u16(pa + 6), e4 = lds_tab_u16(pa + 8);
                j += (uint32_t)(m >= e1) + (uint32_t)(m >= e2) + (uint32_t)(m >= e3) + (uint32_t)(m >= e4);
                if (m < e4) break;
            }
            C = lds_tab_u16(ca + 2 * j);
            Fq = lds_tab_u16(ca + 2 * j + 2) - C;
            if (j >= nU || Fq == 0 || Fq > TOTFREQ) { bad = 1; j = min(j, nU - 1); Fq &= 0x1FFF; }     // only a corrupt stream gets here
            rn = j;
        }
    }
    __syncwarp();
    uint32_t xn = Fq * (x >> TF_SHIFT) + m - C;                          // [Decoding.java:35-48]
    const bool n1 = act && xn < RANS_L;
    const bool n2 = act && xn < (1u << 15);                              // after one byte still below 2^23 -> a second byte
    const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1), b2 = __ballot_sync(0xFFFFFFFFu, n2);
    const uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);         // lanes of a group take their bytes in the order 0,1,2,3 [Decoding.java:84-90]
    const uint32_t v0 = lds_u8(ring | (at & (DEC1_RING - 1))), v1 = lds_u8(ring | ((at + 1) & (DEC1_RING - 1)));
    const uint32_t x1 = (xn << 8) | v0;
    xn = n2 ? ((x1 << 8) | v1) : (n1 ? x1 : xn);
    cur += __popc(b1 & gm) + __popc(b2 & gm);
    return make_uint4(act ? xn : x, cur, act ? rn : r, bad & (uint32_t)act);
}

// NS = blocks per lane: a lane can carry the same state index k of NS different blocks as NS independent instruction
// streams in one basic block.  Measured (profiles/r2_k_dec1_v6.txt): a lone warp takes ~3.3 cycles per instruction on this
// code whether or not neighbouring instructions depend on each other (two interleaved streams: 302 cycles per step pair
// against 168 per step), so a step costs what its instruction count costs and NS = 1 is what ships.
#define DEC1_NS 1
template <int NS>
__global__ void __launch_bounds__(32 * DEC1_WARPS) k_dec1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                                          const uint32_t *__restrict__ cta_first, const uint32_t *__restrict__ cta_cfg,
                                                          const uint32_t *__restrict__ ncta, const uint8_t *__restrict__ d_in,
                                                          const uint8_t *__restrict__ images, const DecMeta *__restrict__ meta,
                                                          uint8_t *__restrict__ d_out, int32_t *__restrict__ status)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t c = blockIdx.x;
    if (c >= *ncta) return;
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 511u) & ~511u;
    const uint32_t first = cta_first[c], cfg = cta_cfg[c], cnt = cfg & 0xFFu, ng = cfg >> 8;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, k = lane & 3u, g = lane >> 2;
    const uint32_t bpw = ng * NS;                                  // blocks of one warp
    if (w * bpw >= cnt) return;                                    // warp-uniform

    // the CTA's blocks: lane i looks at slot first + i; tables are laid out back to back in slot order
    uint32_t blk_i = 0, nU_i = 0;
    if (lane < cnt) {
        blk_i = slots[first + lane];
        nU_i = status[blk_i] == 0 ? meta[blk_i].nU : 0u;
        if (nU_i > DEC_MAXU) nU_i = 0;                             // large alphabet: k_dec1_big
    }
    const uint32_t tab_i = nU_i ? dec_tab_bytes(nU_i) : 0u;
    uint32_t excl = tab_i;
#pragma unroll
    for (int dlt = 1; dlt < 32; dlt <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, excl, dlt); if (lane >= (uint32_t)dlt) excl += t; }
    excl -= tab_i;
    const uint32_t tab0 = sbase + ((cnt * DEC1_FIX + 16u + 255u) & ~255u);

    // this lane's blocks: stream s of group g is block (w * ng + g) * NS + s of the CTA.  A stream without a block becomes
    // an exact clone of the warp's first live stream with the same k (same block, hence the same state, table and ring
    // addresses, and the same bytes stored to the same places): the hot loop then needs no "is this lane real" predicate.
    uint32_t bi[NS];
    {
        bool own[NS];
        uint32_t livem[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const uint32_t b = (w * ng + g) * NS + s;
            const uint32_t nUb = __shfl_sync(0xFFFFFFFFu, nU_i, b & 31);          // every lane takes part in the shuffle
            own[s] = g < ng && b < cnt && nUb != 0;
            bi[s] = b;
            livem[s] = __ballot_sync(0xFFFFFFFFu, own[s]);
        }
        uint32_t sh_bi = 0;
        bool found = false;
#pragma unroll
        for (int s = 0; s < NS; s++)
            if (!found && livem[s]) { sh_bi = (w * ng + ((uint32_t)(__ffs(livem[s]) - 1) >> 2)) * NS + s; found = true; }
        if (!found) return;
#pragma unroll
        for (int s = 0; s < NS; s++)
            if (!own[s]) bi[s] = sh_bi;
    }
    uint32_t blk[NS], nU[NS], ebase[NS], cbase[NS], cumb[NS], ring[NS], outq[NS], alpha[NS];
    uint8_t *out[NS];
    const uint8_t *sbytes[NS];
    uint32_t mylen[NS], lim[NS], ph[NS], in_len[NS], qlen[NS], rest[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        blk[s] = __shfl_sync(0xFFFFFFFFu, blk_i, bi[s]);
        nU[s] = __shfl_sync(0xFFFFFFFFu, nU_i, bi[s]);
        ebase[s] = tab0 + __shfl_sync(0xFFFFFFFFu, excl, bi[s]);
        cumb[s] = dec_cumb(nU[s]);
        cbase[s] = ebase[s] + nU[s] * DEC_ROWE;
        ring[s] = sbase + bi[s] * DEC1_RING;
        outq[s] = sbase + cnt * DEC1_RING + 16 + bi[s] * (4 * DEC1_OUTQ) + k * DEC1_OUTQ;
        alpha[s] = sbase + cnt * (DEC1_RING + 4 * DEC1_OUTQ) + 16 + bi[s] * DEC_MAXU;
        const BlockDesc d = desc[blk[s]];
        const uint32_t n = d.out_len;
        const uint32_t q = n >> 2;
        out[s] = d_out + d.out_off + (uint64_t)k * q;
        mylen[s] = (k == 3) ? n - 3 * q : q;                       // lane 3 also decodes the remainder [D14.java:62-68]
        qlen[s] = q;
        rest[s] = n - 3 * q;
        ph[s] = (uint32_t)((uintptr_t)out[s] & 15);                // the output ring is shifted so that global 16-byte granules are ring granules
        sbytes[s] = d_in + d.in_off;
        in_len[s] = d.in_len;
        lim[s] = (d.in_len + 3) >> 2;                              // words of the block
    }

    // tables of the warp's blocks: global image -> shared memory
    for (uint32_t bb = 0; bb < bpw; bb++) {
        const uint32_t b = w * bpw + bb;
        if (b >= cnt) break;
        const uint32_t nUg = __shfl_sync(0xFFFFFFFFu, nU_i, b);
        if (!nUg) continue;
        const uint32_t eb = tab0 + __shfl_sync(0xFFFFFFFFu, excl, b);
        const uint8_t *img = images + (size_t)desc[__shfl_sync(0xFFFFFFFFu, blk_i, b)].o1_index * DEC_IMG;
        const uint4 *se = reinterpret_cast<const uint4 *>(img);
        for (uint32_t i = lane; i < nUg * (DEC_ROWE / 16); i += 32) sts_v4(eb + 16 * i, __ldg(se + i));
        const uint2 *sc = reinterpret_cast<const uint2 *>(img + DEC_IMG_CUM);
        const uint32_t cb = eb + nUg * DEC_ROWE;
        for (uint32_t i = lane; i < nUg * dec_cumb(nUg) / 8; i += 32) sts_v2(cb + 8 * i, __ldg(sc + i));
        const uint32_t al = sbase + cnt * (DEC1_RING + 4 * DEC1_OUTQ) + 16 + b * DEC_MAXU;
        if (lane < DEC_MAXU / 4) sts_u32(al + 4 * lane, __ldg(reinterpret_cast<const uint32_t *>(img + DEC_IMG_ALPHA) + lane));
    }

    uint32_t nfull = 0xFFFFFFFFu, nmax = 0;
#pragma unroll
    for (int s = 0; s < NS; s++) {
        nfull = min(nfull, __reduce_min_sync(0xFFFFFFFFu, qlen[s]));          // steps every live lane takes
        nmax = max(nmax, __reduce_max_sync(0xFFFFFFFFu, rest[s]));
    }
    const uint32_t nchunks = (nmax + DEC1_CH - 1) / DEC1_CH;

    // the compressed streams: ring of DEC1_RING bytes per block addressed by the byte offset inside the block, filled by
    // 4-byte cp.async, lane k of the group takes every fourth word
    uint32_t wfetch[NS], cur[NS], x[NS], Wc[NS], rowa[NS], r[NS];
    uint32_t hzd[NS], hzi[NS];      // bytes below hzd are in the ring once every copy group but the newest has landed; below hzi: once all have
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const uint32_t cur0 = meta[blk[s]].pay_off;
        wfetch[s] = cur0 >> 2;
        const uint32_t tgt = (cur0 + DEC1_RING) >> 2;
        for (uint32_t wd = wfetch[s] + k; wd < tgt; wd += 4)
            if (wd < lim[s]) cp_async4s(ring[s] + ((4 * wd) & (DEC1_RING - 1)), sbytes[s] + 4ull * wd);
        wfetch[s] = tgt;
        cur[s] = cur0;
    }
    cp_async_commit();
    cp_async_wait_all();
    __syncwarp();
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const uint32_t c0 = cur[s] + 4 * k;
        x[s] = lds_u8(ring[s] + (c0 & (DEC1_RING - 1))) | (lds_u8(ring[s] + ((c0 + 1) & (DEC1_RING - 1))) << 8) |
               (lds_u8(ring[s] + ((c0 + 2) & (DEC1_RING - 1))) << 16) | (lds_u8(ring[s] + ((c0 + 3) & (DEC1_RING - 1))) << 24);
        cur[s] += 16;
        // Wc = the next 4 bytes of the group's stream (bytes cur .. cur + 3), in a register
        Wc[s] = __funnelshift_r(lds_u32(ring[s] | (cur[s] & (DEC1_RING - 4))), lds_u32(ring[s] | ((cur[s] + 4) & (DEC1_RING - 4))), cur[s] << 3);
        rowa[s] = ebase[s];                   // context 0 is byte 0, rank 0 whenever the block has a context-0 row
        r[s] = 0;
        hzd[s] = hzi[s] = wfetch[s] >= lim[s] ? 0xFFFFFFFFu : 4 * wfetch[s];
    }
    const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
    uint32_t bad = 0;
    uint32_t bprev[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) bprev[s] = 0;

    for (uint32_t ch = 0; ch < nchunks; ch++) {
        // stream bytes of this chunk: normally everything but the newest copy group suffices
        bool starved = false;
#pragma unroll
        for (int s = 0; s < NS; s++) starved = starved || cur[s] + DEC1_NEED > hzd[s];
        if (__any_sync(0xFFFFFFFFu, starved)) {
            cp_async_wait_all();
#pragma unroll
            for (int s = 0; s < NS; s++) hzd[s] = hzi[s];
        } else cp_async_wait_1();
        __syncwarp();
        const uint32_t t0 = ch * DEC1_CH;
        // output ring half of this chunk; the bytes the previous chunk held back move in front of it
        uint32_t oh[NS], ow[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) {
            oh[s] = outq[s] + (ch & 1u) * DEC1_OUTH;
            const uint32_t ohp = outq[s] + ((ch & 1u) ^ 1u) * DEC1_OUTH;
            sts_v4(oh[s], lds_v4(ohp + DEC1_CH));
            ow[s] = oh[s] + ph[s];            // step t0 + u of this lane goes to ow + u
        }

        if (t0 + DEC1_CH <= nfull) {
            // -------------------------------------------------------------------- all lanes active: the short chain
            // Written in phases over the NS streams so that the loads of all streams are issued before anything waits for
            // them.  The byte of step u is stored during step u + 1 (its table lookup has long returned by then: a store
            // right behind its own load would hold up every instruction after it, the warp issues in order); bprev lives
            // across chunks, so the store of step 0 rewrites the last byte of the previous chunk where the copy above
            // put it (or lands in the unused byte in front of the ring half when nothing is held back).
#pragma unroll 8
            for (uint32_t u = 0; u < DEC1_CH; u++) {
                uint32_t m[NS], xs[NS], c0[NS], f0[NS], w1[NS], xn[NS], b1[NS], xf[NS], cf[NS], rf[NS], dl[NS];
                bool a[NS], n1[NS];
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    m[s] = x[s] & (TOTFREQ - 1);
                    xs[s] = x[s] >> TF_SHIFT;
                    const uint32_t ea = rowa[s] | ((x[s] >> 4) & 0xF8u);           // bucket = slot >> 7, 8 bytes each; rows are 256-byte aligned
                    c0[s] = lds_tab_u16(ea); f0[s] = lds_tab_u16(ea + 2);
                    w1[s] = lds_tab_u32(ea + 4);                                   // f1 | r0 << 16 | r1 << 24
                }
#pragma unroll
                for (int s = 0; s < NS; s++) sts_u8(ow[s] + u - 1, bprev[s]);
                bool rl = false;
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const uint32_t f1 = w1[s] & 0xFFFFu;
                    const uint32_t d0 = m[s] - c0[s], d1 = d0 - f0[s];
                    a[s] = d0 >= f0[s];                                            // not the first candidate
                    dl[s] = d1;                                                    // negative exactly when it is the first
                    const uint32_t p0 = f0[s] * xs[s] + d0, p1 = f1 * xs[s] + d1;  // [Decoding.java:35-48] for either candidate
                    xn[s] = a[s] ? p1 : p0;
                    n1[s] = xn[s] < RANS_L;
                    rl = rl || (a[s] && d1 >= f1) || xn[s] < (1u << 15);           // beyond both candidates, or two bytes to take
                }
#pragma unroll
                for (int s = 0; s < NS; s++) b1[s] = __ballot_sync(0xFFFFFFFFu, n1[s]);
                const bool rare = __any_sync(0xFFFFFFFFu, rl);
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    // this lane's byte is byte (bytes taken by the lower lanes of the group) of the window Wc; one byte-permute,
                    // predicated on the lane renormalising at all, shifts it into the state
                    const uint32_t cnt = __popc(b1[s] & lm);
                    const uint32_t px = prmt_raw(xn[s], Wc[s], 0x2104u + cnt);
                    xf[s] = n1[s] ? px : xn[s];
                    cf[s] = cur[s] + __popc(b1[s] & gm);
                    // r1 or r0: byte 3 or 2 of w1; the selector comes from the sign of d1 in one instruction (LEA.HI.SX32)
                    rf[s] = prmt_raw(w1[s], 0u, 0x4443u + (uint32_t)((int32_t)dl[s] >> 31));
                }
                if (__builtin_expect(rare, 0)) {
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        const uint4 gs = dec1_general_step(x[s], cur[s], r[s], true, ebase[s], cbase[s], cumb[s], nU[s], ring[s], lm, gm);
                        xf[s] = gs.x; cf[s] = gs.y; rf[s] = gs.z; bad |= gs.w;
                    }
                }
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    x[s] = xf[s]; cur[s] = cf[s]; r[s] = rf[s];
                    Wc[s] = __funnelshift_r(lds_u32(ring[s] | (cur[s] & (DEC1_RING - 4))), lds_u32(ring[s] | ((cur[s] + 4) & (DEC1_RING - 4))), cur[s] << 3);
                    rowa[s] = ebase[s] + (r[s] << 8);
                    bprev[s] = lds_tab_u8(alpha[s] + r[s]);
                }
            }
#pragma unroll
            for (int s = 0; s < NS; s++) sts_u8(ow[s] + DEC1_CH - 1, bprev[s]);
        } else {
            // -------------------------------------------------------------------- tail: lanes end at different steps
            for (uint32_t u = 0; u < DEC1_CH; u++) {
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const bool act = t0 + u < mylen[s];
                    const uint4 gs = dec1_general_step(x[s], cur[s], r[s], act, ebase[s], cbase[s], cumb[s], nU[s], ring[s], lm, gm);
                    x[s] = gs.x; cur[s] = gs.y; r[s] = gs.z; bad |= gs.w;
                    if (act) sts_u8(ow[s] + u, lds_tab_u8(alpha[s] + r[s]));
                }
            }
#pragma unroll
            for (int s = 0; s < NS; s++) {
                Wc[s] = __funnelshift_r(lds_u32(ring[s] | (cur[s] & (DEC1_RING - 4))), lds_u32(ring[s] | ((cur[s] + 4) & (DEC1_RING - 4))), cur[s] << 3);
                rowa[s] = ebase[s] + (r[s] << 8);
            }
        }

#pragma unroll
        for (int s = 0; s < NS; s++) {
            // output: steps [t0 - ph, t0 + DEC1_CH - ph) of this lane's quarter are the ring half's bytes [0, DEC1_CH): aligned granules
            {
                const int32_t tb = (int32_t)t0 - (int32_t)ph[s], ml = (int32_t)mylen[s];       // a quarter is < 2^30 + 4 bytes
#pragma unroll
                for (int j = 0; j < DEC1_CH / 16; j++) {
                    const int32_t ts = tb + 16 * j;
                    if (ts >= 0 && ts + 16 <= ml) *reinterpret_cast<uint4 *>(out[s] + ts) = lds_v4(oh[s] + 16 * j);
                    else if (ts + 16 > 0 && ts < ml) {
                        const uint4 v = lds_v4(oh[s] + 16 * j);
                        const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t t = ts + i;
                            if (t >= 0 && t < ml) out[s][t] = (uint8_t)(wv[i >> 2] >> (8 * (i & 3)));
                        }
                    }
                }
            }
            // stream: request everything the ring can hold beyond the cursor
            const uint32_t tgt = (cur[s] + DEC1_RING) >> 2;
            for (uint32_t wd = wfetch[s] + k; wd < tgt; wd += 4)
                if (wd < lim[s]) cp_async4s(ring[s] + ((4 * wd) & (DEC1_RING - 1)), sbytes[s] + 4ull * wd);
            if (tgt > wfetch[s]) wfetch[s] = tgt;
            hzd[s] = hzi[s];
            hzi[s] = wfetch[s] >= lim[s] ? 0xFFFFFFFFu : 4 * wfetch[s];
        }
        cp_async_commit();
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
        // the bytes still held back for alignment: steps [DEC1_CH * nchunks - ph, DEC1_CH * nchunks) sit behind the last chunk's half
        if (ph[s]) {
            const uint32_t ohp = outq[s] + (((nchunks & 1u) ^ 1u)) * DEC1_OUTH;
            const int64_t ts = (int64_t)nchunks * DEC1_CH - ph[s];
            for (uint32_t i = 0; i < ph[s]; i++) {
                const int64_t t = ts + i;
                if (t >= 0 && t < (int64_t)mylen[s]) out[s][t] = (uint8_t)lds_u8(ohp + DEC1_CH + i);
            }
        }
        if (cur[s] > in_len[s] || bad) status[blk[s]] = CRAMGPU_ERR_CORRUPT;       // ran past the compressed bytes, or left the tables
    }
}

__device__ __forceinline__ uint32_t ld_byte(const uint8_t *p, uint32_t pos, uint32_t len, bool &under)
{
    if (pos >= len) { under = true; return 0; }
    return p[pos];
}

// k_dec1s: the SHORT order-1 blocks of a call whose long blocks keep the machine busy (a slice has one 1.5 MB quality
// block and a dozen series of 10-50 KB).  k_dec1 holds every SM's shared memory for as long as the longest block
// takes, so short blocks queued behind it would only start when it ends (8 of 45 ms in the benchmark).  This kernel
// needs no shared memory at all -- bucket entries, cumulative rows and the alphabet are read from the table images in
// global memory (L2), the stream and the output byte by byte -- so its one-warp CTAs fit beside k_dec1's and the short
// blocks are decoded in its shadow: a step costs ~3 L2 round trips instead of ~170 cycles, which nobody waits for.
// One warp = 8 blocks, lane = state, every step in the general form of dec1_general_step.
//
// WIDE = true: the decoder of the blocks with more than DEC_MAXU distinct bytes, whose tables (up to 201 KB) fit no CTA's
// shared memory: `slots` is the list k_dec_wide_plan made of them, wide[0] their number, wide[1] the blocks per warp (1, 2,
// 4 or 8: as few as keep the machine's warp slots filled).  A step is a chain of three global-memory round trips (fine
// bucket -> five cumulative entries -> renormalisation byte); every wide block of the call is in flight at once.
template <bool WIDE>
__global__ void __launch_bounds__(32) k_dec1s(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots, uint32_t nslots,
                                              const uint32_t *__restrict__ wide,
                                              const uint8_t *__restrict__ d_in, const uint8_t *__restrict__ images,
                                              const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out, int32_t *__restrict__ status)
{
    const uint32_t lane = threadIdx.x, k = lane & 3u;
    uint32_t gpw = 8;
    if (WIDE) { nslots = wide[0]; gpw = wide[1]; }
    const uint32_t slot = blockIdx.x * gpw + (lane >> 2);
    bool have = slot < nslots && (lane >> 2) < gpw;
    const uint32_t blk = have ? slots[slot] : 0;
    const BlockDesc d = desc[blk];
    const DecMeta mt = meta[blk];
    const uint32_t nU = have && status[blk] == 0 ? mt.nU : 0u;
    have = have && nU != 0 && mt.img != DEC_NO_IMG && (WIDE ? nU > DEC_MAXU : nU <= DEC_MAXU);
    if (__ballot_sync(0xFFFFFFFFu, have) == 0) return;
    const uint8_t *img = images + (size_t)(have ? mt.img : 0u) * 256u;
    const uint32_t cumb = dec_cumb(nU);
    const uint32_t cum_off = WIDE ? nU * DEC_ROWE : DEC_IMG_CUM, alpha_off = WIDE ? cum_off + nU * cumb : DEC_IMG_ALPHA;
    const uint32_t n = have ? d.out_len : 0, q = n >> 2;
    uint8_t *out = d_out + d.out_off + (uint64_t)k * q;
    const uint32_t mylen = (k == 3) ? n - 3 * q : q;
    const uint8_t *in = d_in + d.in_off;
    const uint32_t len = d.in_len;
    bool under = false;
    uint32_t bad = 0;
    uint32_t cur = have ? mt.pay_off : 0, x = 0;
    if (have) {
        x = ld_byte(in, cur + 4 * k, len, under) | (ld_byte(in, cur + 4 * k + 1, len, under) << 8) |
            (ld_byte(in, cur + 4 * k + 2, len, under) << 16) | (ld_byte(in, cur + 4 * k + 3, len, under) << 24);
        cur += 16;
    }
    const uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, n - 3 * q);
    const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
    uint32_t r = 0;
    for (uint32_t t = 0; t < max_steps; t++) {
        const bool act = have && t < mylen;
        uint32_t xn = x, rn = r;
        if (act && WIDE) {
            // fine bucket -> first candidate rank, then one round of five cumulative entries: the start of the candidate and the
            // next four ends (sorted; the ones <= m form a prefix; the row ends in 0xFFFF padding)
            const uint32_t m = x & (TOTFREQ - 1);
            const uint16_t *cum = reinterpret_cast<const uint16_t *>(img + cum_off + (size_t)r * cumb);
            uint32_t j = __ldg(img + (size_t)r * DEC_ROWE + (m >> 4));
            uint32_t C, Fq;
            for (;;) {
                const uint32_t e0 = __ldg(cum + j), e1 = __ldg(cum + j + 1), e2 = __ldg(cum + j + 2), e3 = __ldg(cum + j + 3), e4 = __ldg(cum + j + 4);
                const uint32_t c = (uint32_t)(m >= e1) + (uint32_t)(m >= e2) + (uint32_t)(m >= e3) + (uint32_t)(m >= e4);
                j += c;
                if (c < 4) {
                    C = c == 0 ? e0 : c == 1 ? e1 : c == 2 ? e2 : e3;
                    Fq = (c == 0 ? e1 : c == 1 ? e2 : c == 2 ? e3 : e4) - C;
                    break;
                }
                if (j >= nU) { C = e4; Fq = 0; break; }                      // ran off the row: corrupt
            }
            if (j >= nU || Fq == 0 || Fq > TOTFREQ) { bad = 1; j = min(j, nU - 1); Fq &= 0x1FFF; }         // only a corrupt stream gets here
            rn = j;
            xn = Fq * (x >> TF_SHIFT) + m - C;                                   // [Decoding.java:35-48]
            out[t] = __ldg(img + alpha_off + rn);
        } else if (act) {
            const uint32_t m = x & (TOTFREQ - 1);
            const uint2 e = __ldg(reinterpret_cast<const uint2 *>(img + (size_t)r * DEC_ROWE + ((m >> 7) << 3)));
            const uint32_t c0 = e.x & 0xFFFFu, f0 = e.x >> 16, f1 = e.y & 0xFFFFu, r0 = (e.y >> 16) & 0xFFu, r1 = e.y >> 24;
            uint32_t C = c0, Fq = f0;
            rn = r0;
            if (m - c0 >= f0) {
                C = c0 + f0; Fq = f1; rn = r1;
                if (m - C >= f1) {
                    const uint16_t *cum = reinterpret_cast<const uint16_t *>(img + cum_off + (size_t)r * cumb);
                    uint32_t j = r1 + 1;
                    while (m >= __ldg(cum + j + 1)) j++;
                    C = __ldg(cum + j);
                    Fq = __ldg(cum + j + 1) - C;
                    if (j >= nU || Fq == 0 || Fq > TOTFREQ) { bad = 1; j = min(j, nU - 1); Fq &= 0x1FFF; }     // only a corrupt stream gets here
                    rn = j;
                }
            }
            xn = Fq * (x >> TF_SHIFT) + m - C;                                   // [Decoding.java:35-48]
            out[t] = __ldg(img + alpha_off + rn);
        }
        __syncwarp();
        const bool n1 = act && xn < RANS_L;
        const bool n2 = act && xn < (1u << 15);
        const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1), b2 = __ballot_sync(0xFFFFFFFFu, n2);
        const uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);             // lanes of a group take their bytes in the order 0,1,2,3
        if (n1) {
            xn = (xn << 8) | ld_byte(in, at, len, under);
            if (n2) xn = (xn << 8) | ld_byte(in, at + 1, len, under);
        }
        cur += __popc(b1 & gm) + __popc(b2 & gm);
        x = xn; r = rn;
    }
    if (have && (under || bad)) status[blk] = CRAMGPU_ERR_CORRUPT;
}

// k_dec_wide_plan: one CTA.  The order-1 blocks with a pooled table image, in the order of `slots` (longest first) -> wide_slots;
// wide[0] = their number, wide[1] = blocks per warp of k_dec1s<true>: the smallest of 1, 2, 4, 8 that needs at most
// `warp_budget` warps.
__global__ void __launch_bounds__(256) k_dec_wide_plan(const uint32_t *__restrict__ slots, uint32_t n, const DecMeta *__restrict__ meta,
                                                       const int32_t *__restrict__ status, uint32_t warp_budget,
                                                       uint32_t *__restrict__ wide_slots, uint32_t *__restrict__ wide)
{
    __shared__ uint32_t s_w[8], s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (uint32_t i0 = 0; i0 < n; i0 += 256) {
        const uint32_t i = i0 + tid;
        bool is = false;
        uint32_t blk = 0;
        if (i < n) { blk = slots[i]; const DecMeta m = meta[blk]; is = status[blk] == 0 && m.nU > DEC_MAXU && m.img != DEC_NO_IMG; }
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, is);
        if (lane == 0) s_w[w] = __popc(b);
        __syncthreads();
        uint32_t at = s_base + __popc(b & ((1u << lane) - 1u));
        for (uint32_t k = 0; k < w; k++) at += s_w[k];
        if (is) wide_slots[at] = blk;
        __syncthreads();
        if (tid == 0) { uint32_t t = 0; for (int k = 0; k < 8; k++) t += s_w[k]; s_base += t; }
        __syncthreads();
    }
    if (tid == 0) {
        const uint32_t nw = s_base;
        uint32_t g = 1;
        while (g < 8 && (nw + g - 1) / g > warp_budget) g *= 2;
        wide[0] = nw; wide[1] = g;
    }
}

// k_dec0: the order-0 decoder in the two-warp form of k_dec1 (decoder warp + helper warp, 8 blocks per CTA).
// The four states of a block take the bytes 4t + k of the output in turn [D04.java:36-58], so a group's ranks of one
// step are four consecutive output bytes and the helper writes one contiguous stream per group.
#define DEC0_CH 32                   // steps per chunk (<= 8 compressed bytes per step and group: 256 bytes per chunk)
#define DEC0_SLOTS 4                 // out-ring chunks
#define DEC0_RINGB 1024              // compressed-byte ring per group
struct Stream {
    const uint8_t *base;      // block start in global memory
    uint32_t len4;            // block length rounded up to 4
    uint32_t wfetch;          // first word not yet requested
};
// a lane needs two bytes: the general renormalisation.  -> x | cursor << 32
__device__ __noinline__ uint64_t dec0_renorm2(uint32_t x, bool n1, bool n2, uint32_t b1, uint32_t cur, uint32_t lm, uint32_t gm, uint32_t rb)
{
    const uint32_t b2 = __ballot_sync(0xFFFFFFFFu, n2);
    const uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);
    const uint32_t v0 = lds_u8(rb | (at & (DEC0_RINGB - 1))), v1 = lds_u8(rb | ((at + 1) & (DEC0_RINGB - 1)));
    const uint32_t x1 = (x << 8) | v0;
    x = n2 ? ((x1 << 8) | v1) : (n1 ? x1 : x);
    cur += __popc(b1 & gm) + __popc(b2 & gm);
    return (uint64_t)x | ((uint64_t)cur << 32);
}
struct __align__(16) Dec0Smem {
    uint8_t ring[8][DEC0_RINGB];                         // compressed bytes, per group (ring-size aligned)
    uint8_t outr[DEC0_SLOTS][DEC0_CH][32];               // decoded bytes: [chunk slot][step][decoder lane]
    uint8_t lut[8][4096];                                // slot -> symbol
    uint32_t fc[8][256];                                 // symbol -> freq << 16 | start
    uint32_t cur[8];
    int32_t ready, done;
};
#define DEC0_SMEM (sizeof(Dec0Smem) + DEC0_RINGB)

__global__ void __launch_bounds__(64) k_dec0(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                              uint32_t nslots, const uint8_t *__restrict__ d_in,
                                              const uint8_t *__restrict__ lut0, const uint32_t *__restrict__ fc0,
                                              const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                              int32_t *__restrict__ status)
{
    extern __shared__ uint8_t smem_raw[];
    Dec0Smem &S = *reinterpret_cast<Dec0Smem *>(smem_raw + ((DEC0_RINGB - ((uint32_t)__cvta_generic_to_shared(smem_raw) & (DEC0_RINGB - 1))) & (DEC0_RINGB - 1)));
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
    const int32_t nchunks = (int32_t)((max_steps + DEC0_CH - 1) / DEC0_CH);

    for (uint32_t gg = 0; gg < 8; gg++) {
        const uint32_t hv = __shfl_sync(0xFFFFFFFFu, (uint32_t)have, gg * 4);
        const uint32_t oi = __shfl_sync(0xFFFFFFFFu, d.o1_index, gg * 4);      // rank among the order-0 blocks
        if (!hv) {
            // no block in this group: its lanes idle on a table that maps every slot to symbol 0 with frequency 4096
            for (uint32_t i = threadIdx.x; i < 1024; i += 64) reinterpret_cast<uint32_t *>(S.lut[gg])[i] = 0;
            for (uint32_t i = threadIdx.x; i < 256; i += 64) S.fc[gg][i] = 4096u << 16;
            continue;
        }
        const uint4 *src = reinterpret_cast<const uint4 *>(lut0 + (size_t)oi * 4096);
        uint4 *dst = reinterpret_cast<uint4 *>(S.lut[gg]);
        for (uint32_t i = threadIdx.x; i < 256; i += 64) dst[i] = __ldg(src + i);
        for (uint32_t i = threadIdx.x; i < 256; i += 64) S.fc[gg][i] = __ldg(fc0 + (size_t)oi * 256 + i);
    }
    if (threadIdx.x == 0) { S.ready = -1; S.done = -1; }
    Stream st;
    st.base = d_in + d.in_off;
    st.len4 = have ? (d.in_len + 3) & ~3u : 0;
    const uint32_t cur0 = have ? meta[blk].pay_off : 0;
    st.wfetch = cur0 >> 2;
    uint8_t *ring = S.ring[g];
    if (helper) {
        const uint32_t tgt = (cur0 + DEC0_RINGB) >> 2, lim = st.len4 >> 2;
        for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
            if (w < lim) cp_async4(ring + ((4 * w) & (DEC0_RINGB - 1)), st.base + 4ull * w);
        st.wfetch = tgt;
        cp_async_commit();
        cp_async_wait_all();
    }
    __syncthreads();

    if (!helper) {
        // ------------------------------------------------------------------------------ decoder warp
        uint32_t x = ring[(cur0 + 4 * k) & (DEC0_RINGB - 1)] | (ring[(cur0 + 4 * k + 1) & (DEC0_RINGB - 1)] << 8) |
                     (ring[(cur0 + 4 * k + 2) & (DEC0_RINGB - 1)] << 16) | ((uint32_t)ring[(cur0 + 4 * k + 3) & (DEC0_RINGB - 1)] << 24);
        uint32_t cur = cur0 + 16;
        const uint32_t rb = (uint32_t)__cvta_generic_to_shared(ring);
        uint32_t Wc = __funnelshift_r(lds_u32(rb | (cur & (DEC0_RINGB - 4))), lds_u32(rb | ((cur + 4) & (DEC0_RINGB - 4))), cur << 3);
        const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
        const uint32_t lutb = (uint32_t)__cvta_generic_to_shared(&S.lut[g][0]), fcb = (uint32_t)__cvta_generic_to_shared(&S.fc[g][0]);
        for (int32_t c = 0; c < nchunks; c++) {
            if (c >= 2) while (ld_acquire_s32(&S.done) < c - 2) {}
            uint8_t *orow = &S.outr[c & (DEC0_SLOTS - 1)][0][lane];
            const uint32_t t0 = (uint32_t)c * DEC0_CH;
            uint32_t rem = steps > t0 ? steps - t0 : 0;
            for (uint32_t uo = 0; uo < DEC0_CH; uo += 8, rem = rem > 8 ? rem - 8 : 0, orow += 8 * 32) {
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
                        const uint64_t two = dec0_renorm2(xn, n1, n2, b1, cur, lm, gm, rb);
                        x = (uint32_t)two; cur = (uint32_t)(two >> 32);
                    } else {
                        cur += __popc(b1 & gm);
                    }
                    x = act ? x : xold;                                     // a finished block keeps its final states: they give the tail bytes
                    Wc = __funnelshift_r(lds_u32(rb | (cur & (DEC0_RINGB - 4))), lds_u32(rb | ((cur + 4) & (DEC0_RINGB - 4))), cur << 3);
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
                const uint32_t tgt = (S.cur[g] + DEC0_RINGB) >> 2;
                for (uint32_t w = st.wfetch + k; w < tgt; w += 4)
                    if (w < lim) cp_async4(ring + ((4 * w) & (DEC0_RINGB - 1)), st.base + 4ull * w);
                if (tgt > st.wfetch) st.wfetch = tgt;
                cp_async_commit();
            }
            // stream bytes [128 c - ph, 128 c + 128 - ph): 8 granules, two per lane of the group
            const int64_t tb = (int64_t)c * (4 * DEC0_CH) - ph + 32 * (int64_t)k;
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int64_t p0 = tb + 16 * j;
                if (p0 + 16 <= 0 || p0 >= (int64_t)nbytes) continue;
                uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const int64_t p = p0 + i;
                    // byte p of the stream = state p & 3 of step p >> 2
                    const uint32_t stp = (uint32_t)(p >> 2) & (DEC0_SLOTS * DEC0_CH - 1);
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


// k_dec1_big: order-1 blocks with more than DEC_MAXU distinct bytes (template sizes of detached mates, tag bytes,
// binary data).  Tables in global memory: one slab per CTA -- lut [256 contexts][4096] u8 slot -> symbol, fc [256][256]
// u32 freq << 16 | start (Frequencies.readStats_o1 [:50-106]) -- that the CTA fills for one block after the other, so the
// scratch is bounded by the grid, not by the number of blocks.  128 threads build, warp 0 decodes (lanes 0-3).
#define BIG_LUT (256u * 4096u)
#define BIG_FC (65536u * 4u)
#define BIG_SLAB (BIG_LUT + BIG_FC)
__global__ void __launch_bounds__(128) k_dec1_big(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                                  uint32_t nslots, const uint8_t *__restrict__ d_in, uint8_t *slabs,
                                                  const DecMeta *__restrict__ meta, uint8_t *__restrict__ d_out,
                                                  int32_t *__restrict__ status)
{
    uint8_t *lut = slabs + (size_t)blockIdx.x * BIG_SLAB;
    uint32_t *fc = reinterpret_cast<uint32_t *>(lut + BIG_LUT);
    __shared__ uint16_t F[256];
    __shared__ uint16_t C[257];
    __shared__ uint8_t symlist[256];
    __shared__ uint32_t s_nsym, s_ctx, s_more, s_bad;
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5, k = lane & 3u;
    for (uint32_t slot = blockIdx.x; slot < nslots; slot += gridDim.x) {
        const uint32_t blk = slots[slot];
        if (status[blk] != 0 || meta[blk].nU <= DEC_MAXU || meta[blk].img != DEC_NO_IMG) continue;       // same answer for the whole CTA
        const BlockDesc d = desc[blk];
        const uint8_t *in = d_in + d.in_off;
        const uint32_t len = d.in_len;
        // contexts and slots the table does not cover decode as symbol 0 with frequency 0, as the zero-filled Java arrays do
        {
            uint4 *z = reinterpret_cast<uint4 *>(lut);
            for (uint32_t i = tid; i < BIG_SLAB / 16; i += 128) z[i] = make_uint4(0, 0, 0, 0);
        }
        TabParser P = {in, len, 9u, 0u, 0, false, true};      // thread 0 only
        if (tid == 0) P.ctx = P.getb();
        __syncthreads();
        for (;;) {
            F[tid] = 0; F[tid + 128] = 0;
            __syncthreads();
            if (tid == 0) {
                s_nsym = P.parse_row(nullptr, F, C, symlist);
                s_ctx = P.ctx;
                if (!P.bad) P.next_ctx();
                s_more = (!P.bad && P.ctx != 0) ? 1u : 0u;
                s_bad = P.bad ? 1u : 0u;
            }
            __syncthreads();
            if (s_bad) break;
            uint8_t *row_lut = lut + (size_t)s_ctx * 4096;
            uint32_t *row_fc = fc + (size_t)s_ctx * 256;
            const uint32_t ns = s_nsym;
            for (uint32_t a = w; a < ns; a += 4) {       // one warp per symbol
                const uint32_t j = symlist[a], f = F[j], c = C[j];
                if (lane == 0) row_fc[j] = (f << 16) | c;
                for (uint32_t sl = lane; sl < f; sl += 32) row_lut[c + sl] = (uint8_t)j;
            }
            __syncthreads();
            if (!s_more) break;
        }
        __threadfence_block();
        __syncthreads();
        if (s_bad) { if (tid == 0) status[blk] = CRAMGPU_ERR_CORRUPT; __syncthreads(); continue; }
        if (w == 0) {
            // D14.uncompress [D14.java:9-70]: lane k decodes the quarter starting at k * (n >> 2), lane 3 also the remainder
            const bool have = lane < 4;
            const uint32_t n = have ? d.out_len : 0;
            const uint32_t q = n >> 2;
            uint8_t *out = d_out + d.out_off + (uint64_t)k * q;
            const uint32_t mylen = (k == 3) ? n - 3 * q : q;
            bool under = false;
            uint32_t cur = meta[blk].pay_off;
            uint32_t x = 0;
            if (have) {
                x = ld_byte(in, cur + 4 * k, len, under) | (ld_byte(in, cur + 4 * k + 1, len, under) << 8) |
                    (ld_byte(in, cur + 4 * k + 2, len, under) << 16) | (ld_byte(in, cur + 4 * k + 3, len, under) << 24);
                cur += 16;
            }
            const uint32_t max_steps = __reduce_max_sync(0xFFFFFFFFu, n - 3 * q);
            const uint32_t gm = 0xFu << (lane & ~3u), lm = ((1u << k) - 1u) << (lane & ~3u);
            uint32_t l = 0;
            for (uint32_t t = 0; t < max_steps; t++) {
                const bool act = have && t < mylen;
                if (act) {
                    const uint32_t m = x & (TOTFREQ - 1);
                    const uint32_t sy = lut[l * 4096 + m];
                    const uint32_t e = fc[l * 256 + sy];
                    out[t] = (uint8_t)sy;
                    x = (e >> 16) * (x >> TF_SHIFT) + m - (e & 0xFFFF);               // [Decoding.java:35-48]
                    l = sy;
                }
                const bool n1 = act && x < RANS_L;
                const bool n2 = act && x < (1u << 15);
                const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, n1), b2 = __ballot_sync(0xFFFFFFFFu, n2);
                const uint32_t at = cur + __popc(b1 & lm) + __popc(b2 & lm);
                if (n1) {
                    x = (x << 8) | ld_byte(in, at, len, under);
                    if (n2) x = (x << 8) | ld_byte(in, at + 1, len, under);
                }
                cur += __popc(b1 & gm) + __popc(b2 & gm);
            }
            if (have && under) status[blk] = CRAMGPU_ERR_CORRUPT;
        }
        __syncthreads();
    }
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
    // the decoders fetch the stream as aligned words (cp.async): a misaligned block would fault the device, not return an error
    for (uint32_t i = 0; i < nblocks; i++)
        if (in_len[i] && ((in_off[i] + (d_in ? (uint64_t)(uintptr_t)d_in : 0)) & 15))
            return ctx_fail(ctx, CRAMGPU_ERR_ARG, "block %u: compressed bytes must start at a multiple of 16", i);

    std::vector<BlockDesc> desc(nblocks);
    std::vector<int32_t> hstatus(nblocks, 0);
    std::vector<uint64_t> hoff(nblocks + 1);
    std::vector<uint32_t> hlen(nblocks);
    std::vector<uint32_t> s0, s1;
    uint64_t out_total = 0;
    uint32_t n1 = 0, n0 = 0, max1 = 0;
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
        // o1_index: rank among the blocks of the same order (selects the block's table scratch)
        if (ord) { d.o1_index = n1++; s1.push_back(i); max1 = std::max(max1, raw); } else { d.o1_index = n0++; s0.push_back(i); }
    }
    hoff[nblocks] = out_total;
    auto by_len = [&](uint32_t a, uint32_t b) { return desc[a].out_len > desc[b].out_len; };
    std::stable_sort(s0.begin(), s0.end(), by_len);
    std::stable_sort(s1.begin(), s1.end(), by_len);

    // Blocks per warp of the order-1 decoder.  The long blocks set the time of the call (one block = one dependent chain),
    // so they get as few neighbours in their warp as the machine affords: about one decoder warp per SM sub-partition.
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    const uint32_t len_long = std::max(max1 / 4, 1u);
    uint32_t n_long = 0;
    for (uint32_t i : s1) n_long += desc[i].out_len >= len_long;
    const uint32_t NS = DEC1_NS;
    uint32_t ng_long = 1;
    while (ng_long * NS < 8 && (n_long + ng_long * NS - 1) / (ng_long * NS) > (uint32_t)sms * 5) ng_long *= 2;
    // When the long blocks fill the machine (every SM's shared memory for the whole pass), the short ones ride along in
    // k_dec1s instead of queueing behind them.  s1 is sorted by length: the long blocks are its first n_long entries.
    const bool ride = ctx->hp_stream && n_long >= (uint32_t)sms && max1 >= (256u << 10) && n1 > n_long;
    const uint32_t n1_main = ride ? n_long : n1, n1_ride = n1 - n1_main;
    const uint32_t nslab = std::min(n1, 16u);                  // CTAs (= 1.25 MB table slabs) of the overflow path for wide alphabets
    const uint64_t pool_bytes = (uint64_t)std::min(n1, ctx->dec_wide_pool_blocks) * DEC_WIDE_IMG_MAX;
    const uint32_t ncta_max = (n1_main + 3) / 4 + 1;           // a CTA of k_dec1 takes at least 4 blocks (or all that are left)

    Carver c;
    size_t o_desc = c.take(sizeof(BlockDesc) * nblocks);
    size_t o_s0 = c.take(4 * (size_t)std::max(n0, 1u)), o_s1 = c.take(4 * (size_t)std::max(n1, 1u));
    size_t o_meta = c.take(sizeof(DecMeta) * (size_t)nblocks);
    size_t o_lut0 = c.take(4096 * (size_t)n0 + 16), o_fc0 = c.take(1024 * (size_t)n0 + 16);
    size_t o_img = c.take((size_t)n1 * DEC_IMG + pool_bytes + 16);          // fixed slots, then the pool of wide images
    size_t o_slab = c.take((size_t)nslab * BIG_SLAB + 16);
    size_t o_wslots = c.take(4 * (size_t)std::max(n1, 1u)), o_wide = c.take(16);
    size_t o_first = c.take(4 * (size_t)ncta_max), o_cfg = c.take(4 * (size_t)ncta_max), o_ncta = c.take(16);
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
    uint8_t *lut0 = S + o_lut0;
    uint32_t *fc0 = (uint32_t *)(S + o_fc0);
    uint8_t *images = S + o_img, *slabs = S + o_slab;
    uint32_t *cta_first = (uint32_t *)(S + o_first), *cta_cfg = (uint32_t *)(S + o_cfg), *ncta = (uint32_t *)(S + o_ncta);
    const uint32_t *slots0 = (const uint32_t *)(S + o_s0), *slots1 = (const uint32_t *)(S + o_s1);
    if (!ctx->dec_attr_done) {
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_dec1<DEC1_NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC1_SMEM));
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_dec0, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC0_SMEM));
        ctx->dec_attr_done = true;
    }
    if (n0 + n1) {
        uint32_t *wide_slots = (uint32_t *)(S + o_wslots), *wide = (uint32_t *)(S + o_wide);
        CUDA_TRY(ctx, cudaMemsetAsync(wide, 0, 16, st));                    // [2] = pool cursor
        LAUNCH(ctx, st, "k_dec_tables", k_dec_tables<<<nblocks, 128, 0, st>>>(dd, d_in, lut0, fc0, images, n1 * (DEC_IMG / 256u),
                                                                              (uint32_t)(pool_bytes / 256u), wide + 2, meta, d_status));
        // k_dec1 goes to the context's high-priority stream; the kernels that may run beside it (short blocks, large
        // alphabets, order 0) are enqueued on `st` before `st` waits for it
        cudaStream_t hs = ctx->hp_stream ? ctx->hp_stream : st;
        if (n1) {
            LAUNCH(ctx, st, "k_dec_plan", k_dec_plan<<<1, 256, 0, st>>>(dd, slots1, n1_main, meta, d_status, ng_long, len_long, NS, cta_first, cta_cfg, ncta));
            if (hs != st) { cudaEventRecord(ctx->hp_in, st); cudaStreamWaitEvent(hs, ctx->hp_in, 0); }
            LAUNCH(ctx, hs, "k_dec1", k_dec1<DEC1_NS><<<ncta_max, 32 * DEC1_WARPS, DEC1_SMEM, hs>>>(dd, slots1, cta_first, cta_cfg, ncta, d_in, images, meta, d_out, d_status));
            if (hs != st) cudaEventRecord(ctx->hp_out, hs);
            if (n1_ride) LAUNCH(ctx, st, "k_dec1s", k_dec1s<false><<<(n1_ride + 7) / 8, 32, 0, st>>>(dd, slots1 + n1_main, n1_ride, nullptr, d_in, images, meta, d_out, d_status));
            // wide alphabets: from their pooled images in global memory, every block of the call at once; what found no room in
            // the pool goes through k_dec1_big's slabs
            LAUNCH(ctx, st, "k_dec_wide_plan", k_dec_wide_plan<<<1, 256, 0, st>>>(slots1, n1, meta, d_status, (uint32_t)sms * 16u, wide_slots, wide));
            LAUNCH(ctx, st, "k_dec1w", k_dec1s<true><<<n1, 32, 0, st>>>(dd, wide_slots, 0, wide, d_in, images, meta, d_out, d_status));
            LAUNCH(ctx, st, "k_dec1_big", k_dec1_big<<<nslab, 128, 0, st>>>(dd, slots1, n1, d_in, slabs, meta, d_out, d_status));
        }
        if (n0) {
            LAUNCH(ctx, st, "k_dec0", k_dec0<<<(n0 + 7) / 8, 64, DEC0_SMEM, st>>>(dd, slots0, n0, d_in, lut0, fc0, meta, d_out, d_status));
        }
        if (n1 && hs != st) cudaStreamWaitEvent(st, ctx->hp_out, 0);
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}
