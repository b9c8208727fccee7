// rans_encode.cu -- htsjdk-compatible rANS order-0 / order-1 ENCODE for sm_100a.
//
// Byte-identical restatement of RANS.compress [RANS.java:60-73] as a batch pipeline:
//
//   k_hist0    byte histogram of every block                  Frequencies.calcFrequencies_o0 [:109-117]
//   k_prep_o1  alphabet of each order-1 block, zero its (ctx x sym) cells
//   k_hist1    order-1 pair counts                            Frequencies.calcFrequencies_o1 [:149-166]
//   k_tables0  normalise to 4095, symbol table, freq table    [:118-145], buildSyms_o0 [:197-211], writeFrequencies_o0 [:215-244]
//   k_tables1  same per context (IEEE double scaling)         [:167-193], buildSyms_o1 [:248-265], writeFrequencies_o1 [:269-328]
//   k_enc0     4-way interleaved coder, one lane per state    E04.compress [E04.java:12-57], RansEncPutSymbol [Encoding.java:46-71]
//   k_enc1     4 quarter-lanes, context = previous byte       E14.compress [E14.java:13-86]
//   k_scan_out output offsets (16-byte aligned, back to back)
//   k_pack     prefix + table + payload into the output       RANS.finalizeCompressed [RANS.java:132-142]
//
// The coder is a strict recurrence per state, so a block offers exactly 4 lanes of parallelism;
// the kernels pack 8 blocks into a warp (one lane per state) and rely on many blocks per call.
// Renormalisation bytes of the 4 lanes share one output cursor: per step the lanes emit in the order
// 3,2,1,0 [E04.java:41-44] and the whole buffer is reversed at the end [rans/Utils.java:7-33]; here the
// bytes are written downwards from the end of the block's work area, positions coming from two
// ballots + popc per step, so no reversal pass is needed.
#include "internal.cuh"
#include <stdlib.h>
#include <stdio.h>

// ------------------------------------------------------------------------------------------------
// device helpers

// Encoding.RansEncSymbolInit [Encoding.java:21-43]: everything of a coder symbol except rcp_freq, in one word
//   bias (13 bits) | freq << 13 (13 bits) | (rcp_shift - 32) << 26
// x_max = freq << 19 and cmpl_freq = 4096 - freq are recomputed in the coder (off the critical chain); rcp_freq
// depends on freq alone and comes from a 4096-entry table the coder kernel builds in shared memory.
__device__ __forceinline__ uint32_t enc_cell(uint32_t start, uint32_t freq)
{
    uint32_t sh = freq >= 2 ? 31 - __clz(freq - 1) : 0;             // shift - 1, shift = ceil(log2 freq)
    uint32_t bias = freq >= 2 ? start : start + TOTFREQ - 1;
    return bias | (freq << 13) | (sh << 26);
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v)
{
    return __reduce_add_sync(0xFFFFFFFFu, v);
}
__device__ __forceinline__ uint32_t warp_excl_scan(uint32_t v, uint32_t lane)
{
    uint32_t s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xFFFFFFFFu, s, d);
        if (lane >= (uint32_t)d) s += t;
    }
    return s - v;
}
__device__ __forceinline__ unsigned long long warp_max64(unsigned long long v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long t = __shfl_xor_sync(0xFFFFFFFFu, v, d);
        v = t > v ? t : v;
    }
    return v;
}

// One warp serialises one frequency row (Frequencies.writeFrequencies_o0 [:215-244]; rows of _o1 [:297-324]).
//   rowF   256 normalised frequencies in shared memory (0 = symbol absent)
//   mask   8 words of shared memory scratch private to this warp
//   dst    destination, or nullptr to only measure
// Run-length rule restated position-locally: inside a maximal run [a..b] of consecutive present symbols,
// a writes its symbol byte, a+1 writes its symbol byte and the count b-a-1 of further symbols, the rest
// write no symbol byte.  Every present symbol then writes its frequency in 1 byte (<128) or 2 bytes.
__device__ uint32_t warp_write_row(const uint16_t *rowF, uint32_t *mask, uint8_t *dst, uint32_t lane)
{
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t b = __ballot_sync(0xFFFFFFFFu, rowF[w * 32 + lane] != 0);
        if (lane == 0) mask[w] = b;
    }
    __syncwarp();
    uint32_t nb = 0;
    uint8_t buf[8 * 4];
    for (int i = 0; i < 8; i++) {
        int j = lane * 8 + i;
        uint32_t f = rowF[j];
        if (!f) continue;
        bool p1 = j > 0 && ((mask[(j - 1) >> 5] >> ((j - 1) & 31)) & 1);
        bool p2 = j > 1 && ((mask[(j - 2) >> 5] >> ((j - 2) & 31)) & 1);
        if (!p1) {
            buf[nb++] = (uint8_t)j;
        } else if (!p2) {
            // end of the run that contains j
            int w = j >> 5;
            uint32_t inv = ~mask[w] & (0xFFFFFFFFu << (j & 31));
            while (inv == 0 && w < 7) { w++; inv = ~mask[w]; }
            int end = inv ? (w * 32 + __ffs(inv) - 2) : 255;
            buf[nb++] = (uint8_t)j;
            buf[nb++] = (uint8_t)(end - j);
        }
        if (f < 128) buf[nb++] = (uint8_t)f;
        else { buf[nb++] = (uint8_t)(128 | (f >> 8)); buf[nb++] = (uint8_t)(f & 0xFF); }
    }
    uint32_t off = warp_excl_scan(nb, lane);
    uint32_t total = __shfl_sync(0xFFFFFFFFu, off + nb, 31);
    if (dst) {
        for (uint32_t i = 0; i < nb; i++) dst[off + i] = buf[i];
        if (lane == 0) dst[total] = 0;
    }
    __syncwarp();
    return total + 1;
}

// ------------------------------------------------------------------------------------------------
// k_hist0: grid (nblocks, chunks); 256 threads; per-warp private 256-bin histograms in shared memory.
__global__ void __launch_bounds__(256) k_hist0(const BlockDesc *__restrict__ desc, const uint8_t *__restrict__ d_in,
                                               uint32_t *__restrict__ F0)
{
    const BlockDesc d = desc[blockIdx.x];
    uint64_t start = (uint64_t)blockIdx.y * HIST_CHUNK;
    if (d.order > 1 || start >= d.in_len) return;
    uint32_t len = min((uint64_t)HIST_CHUNK, d.in_len - start);
    const uint8_t *in = d_in + d.in_off + start;

    __shared__ uint32_t h[8][256];
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&h[0][0])[i] = 0;
    __syncthreads();
    uint32_t *hw = h[threadIdx.x >> 5];
    uint32_t nvec = len >> 4;
    const uint4 *v = reinterpret_cast<const uint4 *>(in);
    for (uint32_t i = threadIdx.x; i < nvec; i += 256) {
        uint4 q = __ldg(v + i);
        uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            atomicAdd(&hw[w[k] & 0xFF], 1u);
            atomicAdd(&hw[(w[k] >> 8) & 0xFF], 1u);
            atomicAdd(&hw[(w[k] >> 16) & 0xFF], 1u);
            atomicAdd(&hw[w[k] >> 24], 1u);
        }
    }
    for (uint32_t i = (nvec << 4) + threadIdx.x; i < len; i += 256) atomicAdd(&hw[in[i]], 1u);
    __syncthreads();
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) s += h[w][threadIdx.x];
    if (s) atomicAdd(&F0[(size_t)blockIdx.x * 256 + threadIdx.x], s);
}

// ------------------------------------------------------------------------------------------------
// k_prep_o1: one CTA (256 threads) per order-1 block.
__global__ void __launch_bounds__(256) k_prep_o1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ o1_blocks,
                                                 const uint32_t *__restrict__ F0, O1Meta *__restrict__ meta)
{
    uint32_t blk = o1_blocks[blockIdx.x];
    uint32_t j = threadIdx.x, lane = j & 31, w = j >> 5;
    bool present = F0[(size_t)blk * 256 + j] != 0;
    bool isctx = present || j == 0;
    __shared__ uint32_t cntS[8], cntC[8];
    uint32_t bs = __ballot_sync(0xFFFFFFFFu, present), bc = __ballot_sync(0xFFFFFFFFu, isctx);
    if (lane == 0) { cntS[w] = __popc(bs); cntC[w] = __popc(bc); }
    __syncthreads();
    uint32_t baseS = 0, baseC = 0, totS = 0, totC = 0;
    for (int k = 0; k < 8; k++) {
        if (k < (int)w) { baseS += cntS[k]; baseC += cntC[k]; }
        totS += cntS[k]; totC += cntC[k];
    }
    uint32_t rs = baseS + __popc(bs & ((1u << lane) - 1)), rc = baseC + __popc(bc & ((1u << lane) - 1));
    O1Meta *m = meta + blockIdx.x;
    m->smap[j] = present ? (uint8_t)rs : 0;
    m->cmap[j] = isctx ? (uint8_t)rc : 0;
    if (present) m->salpha[rs] = (uint8_t)j;
    if (isctx) m->calpha[rc] = (uint8_t)j;
    if (j == 0) { m->nsym = totS; m->nctx = totC; }
    (void)desc;
}

// ------------------------------------------------------------------------------------------------
// k_enc_layout: one CTA.  Now that every order-1 block's alphabet is known, size its tables by what it needs instead of
// the worst case (256 x 256 cells, a 258 KB frequency table): cell offsets of the order-1 blocks, and every block's work
// area (table bytes first, payload growing down from the end).  totals[0] = cells, totals[1] = work bytes: the host sizes
// the buffers from them.  A context row of the table costs at most 2 (context, run) + 4 per symbol (symbol, run, 1-2
// bytes of frequency) + 1 (end) bytes [Frequencies.java:258-324].
__device__ __forceinline__ unsigned long long block_excl_scan_u64(unsigned long long v, unsigned long long *s_warp, unsigned long long &total)
{
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long inc = v;
    for (int d = 1; d < 32; d <<= 1) { unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= (uint32_t)d) inc += t; }
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    unsigned long long base = 0, tot = 0;
    for (uint32_t k = 0; k < blockDim.x / 32; k++) { if (k < w) base += s_warp[k]; tot += s_warp[k]; }
    __syncthreads();
    total = tot;
    return base + inc - v;
}
__global__ void __launch_bounds__(1024) k_enc_layout(BlockDesc *__restrict__ desc, uint32_t nblocks, O1Meta *__restrict__ meta, uint32_t n1,
                                                     unsigned long long *__restrict__ totals)
{
    __shared__ unsigned long long s_warp[32];
    unsigned long long carry = 0;
    for (uint32_t i0 = 0; i0 < n1; i0 += 1024) {
        const uint32_t i = i0 + threadIdx.x;
        const unsigned long long need = i < n1 ? (((unsigned long long)meta[i].nctx * meta[i].nsym + 3ull) & ~3ull) : 0ull;
        unsigned long long tot;
        const unsigned long long at = block_excl_scan_u64(need, s_warp, tot);
        if (i < n1) meta[i].cell_off = carry + at;
        carry += tot;
    }
    if (threadIdx.x == 0) totals[0] = carry;
    carry = 0;
    for (uint32_t i0 = 0; i0 < nblocks; i0 += 1024) {
        const uint32_t i = i0 + threadIdx.x;
        unsigned long long need = 0;
        uint32_t tab_cap = 0, work_cap = 0;
        if (i < nblocks && desc[i].order <= 1) {
            const BlockDesc d = desc[i];
            if (d.order == 1) { const O1Meta &m = meta[d.o1_index]; tab_cap = (m.nctx * (3u + 4u * m.nsym) + 1u + 31u) & ~15u; }
            else tab_cap = O0_TAB_CAP;
            const unsigned long long pay_cap = ((unsigned long long)d.in_len + d.in_len / 4 + 64 + 15) & ~15ull;
            work_cap = (uint32_t)(tab_cap + pay_cap);
            need = ((unsigned long long)work_cap + 255ull) & ~255ull;
        }
        unsigned long long tot;
        const unsigned long long at = block_excl_scan_u64(need, s_warp, tot);
        if (i < nblocks && desc[i].order <= 1) { desc[i].tab_cap = tab_cap; desc[i].work_cap = work_cap; desc[i].work_off = carry + at; }
        carry += tot;
    }
    if (threadIdx.x == 0) totals[1] = carry;
}

// ------------------------------------------------------------------------------------------------
// k_hist1: grid (n_o1, chunks); 256 threads.  Counts (previous byte, byte) pairs of one chunk into a compact
// (ctx rank x sym rank) table in shared memory, then adds the non-zero cells to the block's global table.
__global__ void __launch_bounds__(256) k_hist1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ o1_blocks,
                                               const uint8_t *__restrict__ d_in, const O1Meta *__restrict__ meta,
                                               uint32_t *__restrict__ F1)
{
    const BlockDesc d = desc[o1_blocks[blockIdx.x]];
    uint64_t start = (uint64_t)blockIdx.y * HIST1_CHUNK;
    if (start >= d.in_len) return;
    uint32_t len = min((uint64_t)HIST1_CHUNK, d.in_len - start);
    const uint8_t *base = d_in + d.in_off;
    const uint8_t *in = base + start;
    const O1Meta *m = meta + blockIdx.x;
    uint32_t *g = F1 + m->cell_off;

    // cs[b] = (context row of byte b: cmap[b] * nsym) | (symbol rank of b: smap[b]) << 16 -- one lookup per byte gives the
    // column of this byte and the row of the next one
    __shared__ uint32_t cs[256];
    __shared__ uint32_t cells[O1_SMEM_ENTRIES];
    const uint32_t ns = m->nsym, ncell = m->nsym * m->nctx;
    cs[threadIdx.x] = (uint32_t)m->cmap[threadIdx.x] * ns | ((uint32_t)m->smap[threadIdx.x] << 16);
    const bool in_smem = ncell <= O1_SMEM_ENTRIES;
    // Quality streams put a third of all pairs into one cell: lanes of a warp hitting the same word are served one after
    // the other.  As many copies of the table as fit (up to 8): lane l counts into copy l % copies, added up at the end.
    const uint32_t ncopy = in_smem ? min(8u, O1_SMEM_ENTRIES / max(ncell, 1u)) : 1u;
    if (in_smem) for (uint32_t i = threadIdx.x; i < ncell * ncopy; i += 256) cells[i] = 0;
    __syncthreads();
    uint32_t *t = in_smem ? cells + (threadIdx.x % ncopy) * ncell : g;

    uint32_t nvec = len >> 4;
    const uint4 *v = reinterpret_cast<const uint4 *>(in);
    for (uint32_t i = threadIdx.x; i < nvec; i += 256) {
        uint4 q = __ldg(v + i);
        uint32_t w[4] = {q.x, q.y, q.z, q.w};
        uint64_t gpos = start + ((uint64_t)i << 4);
        uint32_t prev = gpos ? in[((int64_t)i << 4) - 1] : 0;   // last = 0 before the first byte [:152]
        uint32_t row = cs[prev] & 0xFFFFu;
#pragma unroll
        for (int k = 0; k < 4; k++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const uint32_t e = cs[(w[k] >> (8 * b)) & 0xFF];
                atomicAdd(&t[row + (e >> 16)], 1u);
                row = e & 0xFFFFu;
            }
        }
    }
    for (uint32_t i = (nvec << 4) + threadIdx.x; i < len; i += 256) {
        uint64_t gpos = start + i;
        uint32_t prev = gpos ? base[gpos - 1] : 0;
        atomicAdd(&t[(cs[prev] & 0xFFFFu) + (cs[in[i]] >> 16)], 1u);
    }
    if (blockIdx.y == 0 && threadIdx.x < 3) {
        // start symbols of lanes 1..3 are coded in context 0 [:163-166]
        uint32_t q = d.in_len >> 2;
        atomicAdd(&t[(cs[0] & 0xFFFFu) + (cs[base[(uint64_t)(threadIdx.x + 1) * q]] >> 16)], 1u);
    }
    if (in_smem) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < ncell; i += 256) {
            uint32_t c = cells[i];
            for (uint32_t k = 1; k < ncopy; k++) c += cells[k * ncell + i];
            if (c) atomicAdd(&g[i], c);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_tables0: one CTA (256 threads, thread j = symbol j) per order-0 block.
__global__ void __launch_bounds__(256) k_tables0(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ o0_blocks,
                                                 const uint32_t *__restrict__ F0, uint32_t *__restrict__ tab0,
                                                 uint8_t *__restrict__ work, uint32_t *__restrict__ tlen)
{
    uint32_t blk = o0_blocks[blockIdx.x];
    const BlockDesc d = desc[blk];
    uint32_t j = threadIdx.x, lane = j & 31, w = j >> 5;
    uint32_t f = F0[(size_t)blk * 256 + j];
    uint64_t T = d.in_len;
    uint64_t tr = ((uint64_t)TOTFREQ << 31) / T + (1ull << 30) / T;              // [:118]

    __shared__ unsigned long long redk[8];
    __shared__ uint32_t reds[8];
    __shared__ uint16_t rowF[256];
    __shared__ uint32_t mask[8];
    // M = first index of the maximum raw count [:123-126]
    unsigned long long key = f ? (((unsigned long long)f << 32) | (255u - j)) : 0ull;
    key = warp_max64(key);
    uint32_t nf = f ? (uint32_t)(((uint64_t)f * tr) >> 31) : 0;                  // [:127]
    if (f && nf == 0) nf = 1;
    uint32_t s = warp_sum(nf);
    if (lane == 0) { redk[w] = key; reds[w] = s; }
    __syncthreads();
    unsigned long long bk = 0; uint32_t fsum = 1;                                // fsum++ [:131]
    for (int k = 0; k < 8; k++) { bk = redk[k] > bk ? redk[k] : bk; fsum += reds[k]; }
    uint32_t M = 255u - (uint32_t)(bk & 0xFFFFFFFFu);
    if (j == M) nf = (uint32_t)((int)nf + ((int)TOTFREQ - (int)fsum));           // [:132-135]
    // cumulative start [buildSyms_o0 :197-211]
    uint32_t ex = warp_excl_scan(nf, lane);
    uint32_t wtot = __shfl_sync(0xFFFFFFFFu, ex + nf, 31);
    __syncthreads();
    if (lane == 0) reds[w] = wtot;
    __syncthreads();
    uint32_t base = 0;
    for (int k = 0; k < (int)w; k++) base += reds[k];
    tab0[(size_t)blk * 256 + j] = nf ? enc_cell(base + ex, nf) : 0u;         // [buildSyms_o0 :197-211]
    rowF[j] = (uint16_t)nf;
    __syncthreads();
    if (w == 0) {
        uint32_t n = warp_write_row(rowF, mask, work + d.work_off, lane);
        if (lane == 0) tlen[blk] = n;
    }
}

// ------------------------------------------------------------------------------------------------
// k_tables1: one CTA (8 warps) per order-1 block; a warp owns one context row at a time.
// Pass 1 normalises each row, writes its coder symbols and measures its serialised length; thread 0 then
// lays out the table (context bytes use the same run-length rule over contexts with T != 0 [:275-296]);
// pass 2 serialises the rows in place.
__global__ void __launch_bounds__(256) k_tables1(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ o1_blocks,
                                                 const O1Meta *__restrict__ meta, uint32_t *__restrict__ F1,
                                                 uint32_t *__restrict__ tab1, uint8_t *__restrict__ work,
                                                 uint32_t *__restrict__ tlen)
{
    uint32_t blk = o1_blocks[blockIdx.x];
    const BlockDesc d = desc[blk];
    const O1Meta *m = meta + blockIdx.x;
    uint32_t *g = F1 + m->cell_off;
    uint32_t *sy = tab1 + m->cell_off;                         // compact [ctx rank][sym rank] {freq | start << 12}
    uint8_t *tab = work + d.work_off;
    uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;

    __shared__ uint8_t salpha[256], calpha[256];
    __shared__ uint16_t rowF[8][256];
    __shared__ uint32_t mask[8][8];
    __shared__ uint32_t rowlen[256];   // 0 = context absent (T == 0)
    __shared__ uint32_t rowoff[256];
    __shared__ uint32_t total;
    salpha[tid] = m->salpha[tid];
    calpha[tid] = m->calpha[tid];
    const uint32_t ns = m->nsym, nc = m->nctx;
    __syncthreads();

    for (int pass = 0; pass < 2; pass++) {
        for (uint32_t r = w; r < nc; r += 8) {
            if (pass == 1 && rowlen[r] == 0) continue;
            for (int i = lane; i < 256; i += 32) rowF[w][i] = 0;
            __syncwarp();
            uint32_t *row = g + (size_t)r * ns;
            uint32_t f[8];
            uint32_t a0 = lane * 8;
#pragma unroll
            for (int i = 0; i < 8; i++) f[i] = (a0 + i < ns) ? row[a0 + i] : 0;
            if (pass == 0) {
                uint32_t T = 0;
                unsigned long long key = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    T += f[i];
                    unsigned long long k = f[i] ? (((unsigned long long)f[i] << 32) | (0xFFFFu - (a0 + i))) : 0ull;
                    key = k > key ? k : key;
                }
                T = warp_sum(T);
                if (T == 0) { if (lane == 0) rowlen[r] = 0; continue; }
                key = warp_max64(key);
                uint32_t M = 0xFFFFu - (uint32_t)(key & 0xFFFFFFFFu);
                // double p = 4096.0 / T; F = (int)(F * p), IEEE double, no contraction [:170,181]
                double p = __ddiv_rn(4096.0, (double)T);
                uint32_t t2 = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (f[i]) {
                        uint32_t nf = (uint32_t)__double2int_rz(__dmul_rn((double)f[i], p));
                        f[i] = nf ? nf : 1;
                        t2 += f[i];
                    }
                }
                t2 = warp_sum(t2) + 1;                                            // t2++ [:186]
                if (M >= a0 && M < a0 + 8) {
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        if (a0 + i == M) f[i] = (uint32_t)((int)f[i] + ((int)TOTFREQ - (int)t2));
                }
                uint32_t ls = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) ls += f[i];
                uint32_t x = warp_excl_scan(ls, lane);
                uint32_t *srow = sy + (size_t)r * ns;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (a0 + i < ns) {
                        row[a0 + i] = f[i];
                        srow[a0 + i] = f[i] ? enc_cell(x, f[i]) : 0u;                   // [buildSyms_o1 :248-265]
                        x += f[i];
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (a0 + i < ns && f[i]) rowF[w][salpha[a0 + i]] = (uint16_t)f[i];
            __syncwarp();
            uint32_t n = warp_write_row(rowF[w], mask[w], pass ? tab + rowoff[r] : nullptr, lane);
            if (pass == 0 && lane == 0) rowlen[r] = n;
        }
        __syncthreads();
        if (pass == 0 && tid == 0) {
            uint32_t off = 0;
            int rle = 0;
            for (uint32_t r = 0; r < nc; r++) {
                if (!rowlen[r]) continue;
                uint32_t i = calpha[r];
                if (rle > 0) rle--;
                else {
                    tab[off++] = (uint8_t)i;
                    bool prev = r > 0 && rowlen[r - 1] && calpha[r - 1] + 1u == i;
                    if (i > 0 && prev) {
                        uint32_t e = r;
                        while (e + 1 < nc && rowlen[e + 1] && calpha[e + 1] == calpha[e] + 1u) e++;
                        rle = (int)(e - r);
                        tab[off++] = (uint8_t)rle;
                    }
                }
                rowoff[r] = off;
                off += rowlen[r];
            }
            tab[off++] = 0;
            total = off;
        }
        __syncthreads();
    }
    if (tid == 0) tlen[blk] = total;
}

// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// The coder kernel.  A CTA of 12 warps codes 8 blocks (32 states) as a software pipeline over chunks of 32 coder
// steps; the warps hand work over through shared-memory rings and meet at one __syncthreads per chunk:
//
//   8 PRODUCER warps (one block = 4 states each)  for every (state, step) of chunk I: input byte(s) -> {freq, start}
//                     cell of the block's table (shared memory when it fits, else global) -> packed coder descriptor
//                     {rcp_freq, bias | freq << 13 | (rcp_shift - 32) << 26}  [Encoding.RansEncSymbolInit, Encoding.java:21-43].
//                     Threads run over the 32 steps of one state at a time, so input reads are contiguous; the input
//                     words themselves arrive two chunks ahead by cp.async.  rcp_freq comes from a 4096-entry table
//                     built in shared memory when the CTA starts.
//   1 CODER warp      lane = state: 32 dependent steps of chunk I-1; nothing but the recurrence of
//                     RansEncPutSymbol [Encoding.java:46-71], ~25 instructions per step, no global memory.  It shares
//                     its SM sub-partition with one producer and one emitter.
//   3 EMITTER warps (3, 3 and 2 blocks)  chunk I-2: positions of the renormalisation bytes (lanes emit in the order 3,2,1,0
//                     within a step, [E04.java:41-44]) by a prefix sum over the steps; bytes stored downwards from the
//                     end of the block's work area == the reversed buffer of [rans/Utils.java:7-33]
//
// Why: a warp alone on an SM issues one dependent instruction every ~4 cycles and a sub-partition sustains ~0.6 IPC on
// this code (ncu: profiles/r1_k_enc1_v2_hotloop.txt, profiles/r1_k_enc_v6_roles.txt), so the states' recurrence is kept
// as short as possible and everything that does not depend on the state runs beside it on other warps.
// A step in which a state has no symbol (tails, shorter blocks of the CTA) gets freq = 4096: x_max = 2^31 (never
// renormalises), cmpl_freq = 0, bias = 0 -- the recurrence maps x to x without any special case.
#define EC 32
#define ENC_TAB_ENTRIES 1792u     // order-1 {freq,start} cells per block kept in shared memory (ctx x sym <= 1792, e.g. 42 x 42)
#define ENC_THREADS 480               // 15 warps: 1 coder, 8 producers, 3 emitters and 3 that only help with the set-up
#define DR_STRIDE 33              // descriptor ring: uint2 per (state row, step), padded: conflict-free for both sides
#define OR_STRIDE 36              // byte ring: u32 per (step row, state), padded so the emitter's 16-byte reads spread over the banks

__device__ __forceinline__ void cp_async4(void *smem, const void *g)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(g));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

struct EncStream {                // one per state lane of the CTA, in shared memory (32 bytes)
    const uint8_t *base;          // first byte of the block (16-byte aligned)
    const uint32_t *gtab;         // block's compact table in global memory
    uint32_t off;                 // o1: offset of the lane's quarter inside the block; o0: 0
    uint32_t len;                 // o1: symbols owned by this lane; o0: block length
    uint32_t steps;               // steps of the lane's block (longest lane)
    uint32_t ns_smem;             // o1: symbols per context row of the compact table | in-shared-memory flag << 31
};

#define NULL_DESC make_uint2(0u, 4096u << 13)
#ifdef CRAMGPU_TIMING_EXPERIMENTS
__device__ unsigned long long g_enc_dbg[16];   // CRAMGPU_DBG bit 4: busy cycles per warp of CTA 0
#define ENC_DBG(x) x
#else
#define ENC_DBG(x)
#endif

template <int ORDER>
__global__ void __launch_bounds__(ENC_THREADS) k_enc(const BlockDesc *__restrict__ desc, const uint32_t *__restrict__ slots,
                                                     uint32_t nslots, const uint8_t *__restrict__ d_in,
                                                     const uint32_t *__restrict__ tabs, const O1Meta *__restrict__ meta,
                                                     uint8_t *__restrict__ work, uint32_t *__restrict__ plen,
                                                     int32_t *__restrict__ status, int dbg)
{
    extern __shared__ __align__(16) uint8_t smem[];
    constexpr uint32_t TABE = ORDER ? ENC_TAB_ENTRIES : 256u;
    constexpr uint32_t RW = ORDER ? 32u : 128u;                                   // input ring words per state (o1) / per block (o0)
    // Every ring is addressed as an offset from `smem` so that the compiler keeps the accesses in the shared state
    // space (pointers that travel through lambdas become generic and cost an S2R + address rebuild per access).
    constexpr uint32_t OFF_DR = 0;                                                // uint2 [2][32 states][DR_STRIDE]
    constexpr uint32_t OFF_OR = OFF_DR + 2 * 32 * DR_STRIDE * 8;                  // u32   [2][32 steps][OR_STRIDE]
    constexpr uint32_t OFF_RCP = OFF_OR + 2 * 32 * OR_STRIDE * 4;                 // u32   [4096] ceil(2^(31+shift) / F)
    constexpr uint32_t OFF_IR = OFF_RCP + 4096 * 4;                               // u32   [1024] input staging
    constexpr uint32_t OFF_TAB = OFF_IR + 1024 * 4;                               // u32   [8][TABE]
    constexpr uint32_t OFF_MAPS = OFF_TAB + 8 * TABE * 4;                         // u8    o1: [8][2][256] cmap | smap
#define DRING(i) (reinterpret_cast<uint2 *>(smem + OFF_DR)[i])
#define ORING(i) (reinterpret_cast<uint32_t *>(smem + OFF_OR)[i])
#define RCPLUT(i) (reinterpret_cast<uint32_t *>(smem + OFF_RCP)[i])
#define IRING(i) (reinterpret_cast<uint32_t *>(smem + OFF_IR)[i])
#define TAB(i) (reinterpret_cast<uint32_t *>(smem + OFF_TAB)[i])
#define MAPS(i) (smem[OFF_MAPS + (i)])
#define RING_BYTE(r, o) ((uint32_t)smem[OFF_IR + (r) * (RW * 4) + ((o) & (RW * 4 - 1))])     /* input byte at offset o of the block */
    __shared__ EncStream strm[32];
    __shared__ uint8_t *g_end[8];
    __shared__ uint32_t g_cap[8], g_p[8], g_blk[8], g_ovf[8], xfin[32];
    __shared__ uint32_t s_max_steps;

    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // Roles by SM sub-partition (warp w runs on sub-partition w % 4).  The coder's recurrence is the critical path of the
    // whole kernel and a sub-partition issues in order, one instruction per cycle at best: every instruction another warp
    // issues there is a cycle the coder may have to wait (ncu, profiles/r1_k_enc_v6_roles.txt: 52 cycles per step alone,
    // 97 when it shared its sub-partition with a producer and an emitter).  So sub-partition 0 runs the coder and nothing
    // else -- warps 4, 8 and 12 help with the set-up and leave -- and the 11 helpers share the other three:
    //   1: emitter(1) + producers(5, 9, 13)    2: emitter(2) + producers(6, 10, 14)    3: emitter(3) + producers(7, 11)
    const bool is_coder = wid == 0, is_emit = wid >= 1 && wid <= 3, is_prod = wid > 4 && (wid & 3) != 0;
    const bool is_idle = wid != 0 && (wid & 3) == 0;
    const uint32_t prod_id = wid - 5 - (wid >> 2) + 1;                            // 5,6,7,9,10,11,13,14 -> 0..7
    const uint32_t emit_first = wid == 1 ? 0 : wid == 2 ? 3 : 5, emit_last = wid == 1 ? 3 : wid == 2 ? 5 : 8;

    // ---- set-up: every warp helps
    if (tid < 32) {
        uint32_t g = tid >> 2, k = tid & 3;
        uint32_t slot = blockIdx.x * 8 + g;
        bool have = slot < nslots;
        uint32_t blk = have ? slots[slot] : 0;
        BlockDesc d = desc[blk];
        uint32_t n = have ? d.in_len : 0;
        EncStream s;
        s.base = d_in + d.in_off;
        if (ORDER) {
            uint32_t q = n >> 2;
            s.off = k * q;
            s.len = (k == 3) ? n - 3 * q : q;
            s.steps = n - 3 * q;
            const O1Meta *m = meta + d.o1_index;
            s.gtab = tabs + m->cell_off;
            s.ns_smem = (have ? m->nsym : 1) | ((have && m->nsym * m->nctx <= ENC_TAB_ENTRIES) ? 0x80000000u : 0u);
        } else {
            s.off = 0;
            s.len = n;
            s.steps = (n + 3) >> 2;
            s.gtab = tabs + (size_t)blk * 256;
            s.ns_smem = 0x80000000u;
        }
        strm[tid] = s;
        if (k == 0) {
            g_end[g] = work + d.work_off + d.work_cap;
            g_cap[g] = have && n ? d.work_cap - d.tab_cap : 0;
            g_p[g] = 0; g_blk[g] = have && n ? blk : 0xFFFFFFFFu; g_ovf[g] = 0;
        }
        uint32_t ms = __reduce_max_sync(0xFFFFFFFFu, s.steps);
        if (tid == 0) s_max_steps = ms;
    }
    // rcp_freq of every frequency: ((1 << (shift+31)) + F - 1) / F == ceil(2^(31+shift) / F).  The double division
    // rounded up, then ceil, is exact: the quotient is < 2^32 and its fractional part is a multiple of 1/F (>= 2^-12),
    // far above the 2^-21 granularity of a double near 2^32.
    for (uint32_t F = tid; F < 4096; F += ENC_THREADS) {
        uint32_t r = 0xFFFFFFFFu;
        if (F >= 2) {
            uint32_t shift = 32 - __clz(F - 1);
            double num = __hiloint2double((int)((1023u + 31u + shift) << 20), 0);
            r = __double2uint_ru(__ddiv_ru(num, (double)F));
        }
        RCPLUT(F) = r;
    }
    __syncthreads();
    for (uint32_t g = 0; g < 8; g++) {
        const EncStream &s = strm[g * 4];
        if (g_blk[g] == 0xFFFFFFFFu) continue;
        if (ORDER) {
            const O1Meta *m = meta + desc[g_blk[g]].o1_index;
            for (uint32_t i = tid; i < 256; i += ENC_THREADS) { MAPS(g * 512 + i) = m->cmap[i]; MAPS(g * 512 + 256 + i) = m->smap[i]; }
            if (s.ns_smem >> 31) {
                uint32_t cells = m->nsym * m->nctx;
                for (uint32_t i = tid; i < cells; i += ENC_THREADS) TAB(g * TABE + i) = __ldg(s.gtab + i);
            }
        } else {
            for (uint32_t i = tid; i < 256; i += ENC_THREADS) TAB(g * TABE + i) = __ldg(s.gtab + i);
        }
    }
    __syncthreads();
    if (is_idle) return;                       // (a warp that has exited no longer takes part in the CTA's barriers)

    const uint32_t max_steps = s_max_steps;
    const uint32_t nchunks = (max_steps + EC - 1) / EC;
    uint32_t x = RANS_L;                       // coder state

    // Input staging (producer warps).  Order 1: one ring of 32 words per state lane (a chunk reads 33 consecutive
    // bytes of the lane's quarter); order 0: one ring of 128 words per block (a chunk reads 128 consecutive bytes).
    // Slots are addressed by the word offset inside the block, so the three chunks in flight never collide.
    EncStream mine[4];                         // a producer warp owns the 4 states of one block
    const uint32_t g_mine = prod_id & 7;
    if (is_prod) {
#pragma unroll
        for (int si = 0; si < 4; si++) mine[si] = strm[g_mine * 4 + si];
    }
    // Input prefetch of one chunk for this producer warp's block, two cp.async instructions per warp:
    // order 1: lane -> (state = lane >> 3, word = lane & 7) and a second round for words 8, 9 (33 bytes span <= 10 words);
    // order 0: lane -> word of step t.  The lane's own state parameters stay in registers for the whole kernel.
    uint32_t pf_off = 0, pf_len = 0, pf_steps = 0, pf_slot = 0;
    const uint8_t *pf_base = nullptr;
    if (is_prod) {
        const EncStream &ps = strm[g_mine * 4 + (ORDER ? (lane >> 3) : 0)];
        pf_off = ps.off; pf_len = ps.len; pf_steps = ps.steps; pf_base = ps.base;
        pf_slot = ORDER ? (g_mine * 4 + (lane >> 3)) * RW : g_mine * RW;
    }
#define PREFETCH_CHUNK(c)                                                                                          \
    do {                                                                                                           \
        const uint32_t t0_ = (c) * EC;                                                                             \
        if (ORDER) {                                                                                               \
            if (t0_ < pf_steps && pf_len) {                                                                        \
                uint32_t jhi_ = pf_steps - 1 - t0_;                   /* first symbol of the chunk */               \
                uint32_t jlo_ = jhi_ >= EC ? jhi_ - EC : 0;           /* down to the last symbol's context byte */  \
                if (jhi_ > pf_len - 1) jhi_ = pf_len - 1;                                                          \
                if (jhi_ >= jlo_) {                                                                                \
                    uint32_t w0_ = (pf_off + jlo_) >> 2, w1_ = (pf_off + jhi_) >> 2;                               \
                    uint32_t w_ = w0_ + (lane & 7);                                                                \
                    if (w_ <= w1_) cp_async4(&IRING(pf_slot + (w_ & (RW - 1))), pf_base + 4ull * w_);              \
                    w_ += 8;                                                                                       \
                    if ((lane & 7) < 2 && w_ <= w1_) cp_async4(&IRING(pf_slot + (w_ & (RW - 1))), pf_base + 4ull * w_); \
                }                                                                                                  \
            }                                                                                                      \
        } else {                                                                                                   \
            uint32_t t_ = t0_ + lane;                                                                              \
            if (t_ < pf_steps) {                                                                                   \
                uint32_t w_ = pf_steps - 1 - t_;                      /* word of step t */                          \
                cp_async4(&IRING(pf_slot + (w_ & (RW - 1))), pf_base + 4ull * w_);                                 \
            }                                                                                                      \
        }                                                                                                          \
    } while (0)
    if (is_prod) {
        PREFETCH_CHUNK(0u);
        cp_async_commit();
        if (nchunks > 1) PREFETCH_CHUNK(1u);
        cp_async_commit();
    }

    ENC_DBG(unsigned long long busy = 0; unsigned long long t_start = clock64();)
    for (uint32_t I = 0; I < nchunks + 2; I++) {
        ENC_DBG(unsigned long long t_in = clock64();)
        if (is_prod) {
            // ================= PRODUCERS: descriptors of chunk I; input words of chunk I+2 on their way
            // one cp.async group per chunk; chunk I was requested two iterations ago, so only it has to have landed
            if (I + 2 < nchunks && !(dbg & 8)) PREFETCH_CHUNK(I + 2);
            cp_async_commit();
            asm volatile("cp.async.wait_group 2;" ::: "memory");
            __syncwarp();
            if (I < nchunks && !(dbg & 2)) {
                const uint32_t dbase = (I & 1) * 32 * DR_STRIDE + lane;
                const uint32_t t = I * EC + lane;               // this thread's step
#pragma unroll
                for (uint32_t si = 0; si < 4; si++) {
                    const uint32_t s = g_mine * 4 + si;
                    const EncStream &st = mine[si];
                    uint32_t e = 0;
                    if (t < st.steps) {
                        if (ORDER) {
                            uint32_t j = st.steps - 1 - t;
                            if (j < st.len) {
                                uint32_t o = st.off + j;
                                uint32_t l = RING_BYTE(s, o);
                                uint32_t c = j ? RING_BYTE(s, o - 1) : 0;            // context 0 at the lane's first byte
                                uint32_t cell = MAPS(g_mine * 512 + c) * (st.ns_smem & 0xFFFFu) + MAPS(g_mine * 512 + 256 + l);
                                e = (st.ns_smem >> 31) ? TAB(g_mine * TABE + cell) : __ldg(st.gtab + cell);
                            }
                        } else {
                            uint32_t w = st.steps - 1 - t;
                            if ((uint64_t)w * 4 + si < st.len) e = TAB(g_mine * TABE + RING_BYTE(g_mine, w * 4 + si));
                        }
                    }
                    // e = packed symbol word (0 = no symbol for this state in this step)
                    DRING(dbase + s * DR_STRIDE) = e ? make_uint2(RCPLUT((e >> 13) & 0xFFFu), e) : NULL_DESC;
                }
            }
        } else if (is_coder) {
            // ================= CODER: chunk I-1
            if (I >= 1 && I <= nchunks && !(dbg & 4)) {
                const uint32_t dbase = ((I - 1) & 1) * 32 * DR_STRIDE + lane * DR_STRIDE;
                const uint32_t obase = ((I - 1) & 1) * 32 * OR_STRIDE + lane;
                uint2 d = DRING(dbase);
#pragma unroll
                for (uint32_t u = 0; u < EC; u++) {
                    uint2 dn = DRING(dbase + u + 1);           // next step's descriptor (slot 32 is padding)
                    uint32_t F = (d.y >> 13) & 0x1FFFu;
                    uint32_t x_max = F << 19;
                    uint32_t x8 = x >> 8, x16 = x >> 16;
                    bool c1 = x >= x_max, c2 = x8 >= x_max;
                    uint32_t xn = c2 ? x16 : (c1 ? x8 : x);
                    uint32_t q = __umulhi(xn, d.x) >> (d.y >> 26);
                    ORING(obase + u * OR_STRIDE) = (x & 0xFFFFu) | (((uint32_t)c1 + (uint32_t)c2) << 16);
                    x = xn + (d.y & 0x1FFFu) + q * (TOTFREQ - F);
                    d = dn;
                }
            }
        } else if (is_emit) {
            // ================= EMITTERS: chunk I-2
            if (I >= 2 && !(dbg & 1)) {
                const uint32_t obase = (I & 1) * 32 * OR_STRIDE + lane * OR_STRIDE;
                const uint32_t lt = (1u << lane) - 1u;
                for (uint32_t g = emit_first; g < emit_last; g++) {
                    uint4 v = *reinterpret_cast<const uint4 *>(&ORING(obase + 4 * g));
                    uint32_t nb0 = v.x >> 16, nb1 = v.y >> 16, nb2 = v.z >> 16, nb3 = v.w >> 16;
                    uint32_t tot = nb0 + nb1 + nb2 + nb3;                           // 0..8 bytes in this step
                    // prefix over the 32 steps from four ballots of the bit planes (short dependency chain)
                    uint32_t b0 = __ballot_sync(0xFFFFFFFFu, tot & 1), b1 = __ballot_sync(0xFFFFFFFFu, tot & 2);
                    uint32_t b2 = __ballot_sync(0xFFFFFFFFu, tot & 4), b3 = __ballot_sync(0xFFFFFFFFu, tot & 8);
                    uint32_t ex = __popc(b0 & lt) + 2 * __popc(b1 & lt) + 4 * __popc(b2 & lt) + 8 * __popc(b3 & lt);
                    uint32_t total = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2) + 8 * __popc(b3);
                    uint32_t p0 = g_p[g], cap = g_cap[g];
                    bool ovf = p0 + total > cap;
                    // this step's bytes in emission order (lanes 3,2,1,0; low byte of each state first), branch-free
                    unsigned long long pk = v.w & ((1u << (8 * nb3)) - 1u);
                    uint32_t acc = nb3;
                    pk |= (unsigned long long)(v.z & ((1u << (8 * nb2)) - 1u)) << (8 * acc); acc += nb2;
                    pk |= (unsigned long long)(v.y & ((1u << (8 * nb1)) - 1u)) << (8 * acc); acc += nb1;
                    pk |= (unsigned long long)(v.x & ((1u << (8 * nb0)) - 1u)) << (8 * acc);
                    uint8_t *dst = g_end[g] - 1 - (p0 + ex);                         // emission index e lives at end - 1 - e
                    uint32_t n = ovf ? 0 : tot;
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        if ((uint32_t)i < n) dst[-i] = (uint8_t)(pk >> (8 * i));
                    __syncwarp();
                    if (lane == 0) { if (ovf) g_ovf[g] = 1; else g_p[g] = p0 + total; }
                    __syncwarp();
                }
            }
        }
        ENC_DBG(busy += clock64() - t_in;)
        __syncthreads();
    }
    ENC_DBG(if ((dbg & 16) && blockIdx.x == 0 && lane == 0) { g_enc_dbg[wid] = busy; if (wid == 0) g_enc_dbg[12] = clock64() - t_start; })
    if (is_coder) xfin[lane] = x;
    __syncthreads();
    if (wid == 2) {
        // states 3,2,1,0 big-endian then reversed == states 0..3 little-endian in front of the payload [E04.java:46-50]
        uint32_t g = lane >> 2, k = lane & 3;
        uint32_t blk = g_blk[g];
        if (blk != 0xFFFFFFFFu) {
            uint32_t p = g_p[g], cap = g_cap[g];
            bool ovf = g_ovf[g] || p + 16 > cap;
            if (!ovf) {
                uint32_t xs = xfin[lane];
                uint8_t *s = g_end[g] - p - 16 + 4 * k;
                s[0] = (uint8_t)xs; s[1] = (uint8_t)(xs >> 8); s[2] = (uint8_t)(xs >> 16); s[3] = (uint8_t)(xs >> 24);
            }
            if (k == 0) {
                plen[blk] = ovf ? 0 : p + 16;
                if (ovf) status[blk] = CRAMGPU_ERR_OUT_TOO_SMALL;
            }
        }
    }
}
#define ENC_SMEM(ORDER) (2 * 32 * DR_STRIDE * 8 + 2 * 32 * OR_STRIDE * 4 + 4096 * 4 + 4096 + 8 * ((ORDER) ? ENC_TAB_ENTRIES : 256u) * 4 + ((ORDER) ? 8 * 512 : 0))

// ------------------------------------------------------------------------------------------------
// k_scan_out: one CTA; out_off[i] = sum of roundup16(len) of the blocks before i.
__global__ void __launch_bounds__(1024) k_scan_out(const BlockDesc *__restrict__ desc, uint32_t nblocks,
                                                   const uint32_t *__restrict__ tlen, const uint32_t *__restrict__ plen,
                                                   uint64_t out_cap, uint64_t *__restrict__ out_off,
                                                   uint32_t *__restrict__ out_len, int32_t *__restrict__ status)
{
    __shared__ uint64_t wsum[32];
    __shared__ uint64_t carry;
    uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += 1024) {
        uint32_t i = base + tid;
        uint32_t len = 0;
        if (i < nblocks && desc[i].order <= 1 && status[i] == 0) len = 9 + tlen[i] + plen[i];
        uint64_t a = (len + 15u) & ~15ull;
        uint64_t s = a;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            uint64_t t = __shfl_up_sync(0xFFFFFFFFu, s, dlt);
            if (lane >= (uint32_t)dlt) s += t;
        }
        if (lane == 31) wsum[w] = s;
        __syncthreads();
        uint64_t off = carry;
        for (uint32_t k = 0; k < w; k++) off += wsum[k];
        off += s - a;
        if (i < nblocks) {
            if (off + len > out_cap && len) { status[i] = CRAMGPU_ERR_OUT_TOO_SMALL; len = 0; }      // the bytes themselves must fit; the rounding only places the next block
            out_off[i] = off;
            out_len[i] = len;
        }
        __syncthreads();
        if (tid == 1023) carry = off + a;
        __syncthreads();
    }
    if (tid == 0) out_off[nblocks] = carry;
}

// k_pack: grid (nblocks, chunks): prefix [RANS.java:132-142] + table + payload -> output.
__global__ void __launch_bounds__(256) k_pack(const BlockDesc *__restrict__ desc, const uint8_t *__restrict__ work,
                                              const uint32_t *__restrict__ tlen, const uint32_t *__restrict__ plen,
                                              const uint64_t *__restrict__ out_off, const uint32_t *__restrict__ out_len,
                                              uint8_t *__restrict__ d_out)
{
    uint32_t blk = blockIdx.x;
    uint32_t total = out_len[blk];
    uint64_t start = (uint64_t)blockIdx.y * PACK_CHUNK;
    if (start >= total) return;
    const BlockDesc d = desc[blk];
    uint32_t tl = tlen[blk], pl = plen[blk];
    const uint8_t *tab = work + d.work_off;
    const uint8_t *pay = work + d.work_off + d.work_cap - pl;
    uint8_t *out = d_out + out_off[blk];
    uint32_t stop = min((uint64_t)total, start + PACK_CHUNK);
    uint32_t csize = tl + pl;
    for (uint32_t i = start + threadIdx.x; i < stop; i += 256) {
        uint8_t b;
        if (i >= 9 + tl) b = pay[i - 9 - tl];
        else if (i >= 9) b = tab[i - 9];
        else if (i == 0) b = (uint8_t)d.order;
        else if (i < 5) b = (uint8_t)(csize >> (8 * (i - 1)));
        else b = (uint8_t)(d.in_len >> (8 * (i - 5)));
        out[i] = b;
    }
}

// ------------------------------------------------------------------------------------------------
// host launcher

uint32_t cramgpu_rans_bound(uint32_t n)
{
    // (int)(1.05 * n + 257*257*3 + 4) with Java's saturating double -> int conversion
    const double b = 1.05 * (double)n + 257 * 257 * 3 + 4;
    return b >= 2147483647.0 ? 0x7FFFFFFFu : (uint32_t)(int64_t)b;
}

#include <algorithm>

int rans_encode_device(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_in, const uint64_t *in_off,
                       const uint32_t *in_len, const uint8_t *order, uint8_t *d_out, uint64_t out_cap,
                       uint64_t *d_out_off, uint32_t *d_out_len, int32_t *d_status, cudaStream_t st)
{
    if (nblocks == 0) return CRAMGPU_OK;
    if (!in_off || !in_len || !order || !d_out_off || !d_out_len || !d_status)   // d_in may be NULL: in_off then holds absolute device addresses
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");

    // ---- plan (host): effective order, work areas, length-sorted slot lists
    std::vector<BlockDesc> desc(nblocks);
    std::vector<int32_t> hstatus(nblocks, 0);
    std::vector<uint32_t> o0, o1;
    uint32_t max_len = 0;
    for (uint32_t i = 0; i < nblocks; i++) {
        BlockDesc &d = desc[i];
        d = BlockDesc();
        d.in_off = in_off[i];
        d.in_len = in_len[i];
        if (in_off[i] & 15) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "in_off[%u] is not a multiple of 16", i);
        uint32_t ord = order[i];
        if (d.in_len == 0) { d.order = 0xFF; continue; }                       // empty -> empty [RANS.java:60-61]
        if (ord > 1) { d.order = 0xFF; hstatus[i] = CRAMGPU_ERR_UNKNOWN_ORDER; continue; }
        if (d.in_len < 4) ord = 0;                                             // [RANS.java:63-64]
        d.order = ord;                       // tab_cap, work_cap, work_off: k_enc_layout, once the alphabets are known
        if (ord) { d.o1_index = (uint32_t)o1.size(); o1.push_back(i); } else o0.push_back(i);
        max_len = std::max(max_len, d.in_len);
    }
    auto by_len = [&](uint32_t a, uint32_t b) { return desc[a].in_len > desc[b].in_len; };
    std::vector<uint32_t> s0 = o0, s1 = o1;
    std::stable_sort(s0.begin(), s0.end(), by_len);
    std::stable_sort(s1.begin(), s1.end(), by_len);
    const uint32_t n0 = (uint32_t)o0.size(), n1 = (uint32_t)o1.size();

    // ---- carve scratch
    Carver c;
    size_t o_desc = c.take(sizeof(BlockDesc) * nblocks);
    size_t o_o0 = c.take(4 * (size_t)std::max(n0, 1u)), o_o1 = c.take(4 * (size_t)std::max(n1, 1u));
    size_t o_s0 = c.take(4 * (size_t)std::max(n0, 1u)), o_s1 = c.take(4 * (size_t)std::max(n1, 1u));
    size_t o_F0 = c.take(1024 * (size_t)nblocks);
    size_t o_tlen = c.take(4 * (size_t)nblocks), o_plen = c.take(4 * (size_t)nblocks);
    size_t o_zero_end = c.off;                                                  // [o_F0, here) is zeroed
    size_t o_syms0 = c.take(1024 * (size_t)nblocks);
    size_t o_meta = c.take(sizeof(O1Meta) * (size_t)std::max(n1, 1u));
    size_t o_totals = c.take(16);
    if (int rc = pool_reserve(ctx, ctx->scratch, c.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->scratch.p;

    // ---- upload the plan through pinned memory
    size_t meta_bytes = sizeof(BlockDesc) * nblocks + 4 * (size_t)(2 * n0 + 2 * n1) + 4 * (size_t)nblocks + 512;
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, meta_bytes)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));       // previous call may still be reading the pinned plan
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
    CUDA_TRY(ctx, up(o0.data(), 4 * (size_t)n0, S + o_o0));
    CUDA_TRY(ctx, up(o1.data(), 4 * (size_t)n1, S + o_o1));
    CUDA_TRY(ctx, up(s0.data(), 4 * (size_t)n0, S + o_s0));
    CUDA_TRY(ctx, up(s1.data(), 4 * (size_t)n1, S + o_s1));
    CUDA_TRY(ctx, up(hstatus.data(), 4 * (size_t)nblocks, d_status));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_F0, 0, o_zero_end - o_F0, st));

    const BlockDesc *dd = (const BlockDesc *)(S + o_desc);
    uint32_t *F0 = (uint32_t *)(S + o_F0), *tlen = (uint32_t *)(S + o_tlen), *plen = (uint32_t *)(S + o_plen);
    uint32_t *syms0 = (uint32_t *)(S + o_syms0);
#ifdef CRAMGPU_TIMING_EXPERIMENTS
    const int dbg = getenv("CRAMGPU_DBG") ? atoi(getenv("CRAMGPU_DBG")) : 0;   // role-by-role timing (scripts/time_enc.py): wrong output, never in a shipped build
#else
    const int dbg = 0;
#endif
    if (!ctx->enc_attr_done) {
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_enc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ENC_SMEM(1)));
        CUDA_TRY(ctx, cudaFuncSetAttribute(k_enc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, ENC_SMEM(0)));
        ctx->enc_attr_done = true;
    }
    O1Meta *meta = (O1Meta *)(S + o_meta);
    uint32_t *F1 = nullptr, *syms1 = nullptr;
    uint8_t *work = nullptr;
    const uint32_t *d_o0 = (const uint32_t *)(S + o_o0), *d_o1 = (const uint32_t *)(S + o_o1);
    const uint32_t *d_s0 = (const uint32_t *)(S + o_s0), *d_s1 = (const uint32_t *)(S + o_s1);

    if (n0 + n1) {
        uint32_t chunks = (max_len + HIST_CHUNK - 1) / HIST_CHUNK;
        LAUNCH(ctx, st, "k_hist0", k_hist0<<<dim3(nblocks, chunks), 256, 0, st>>>(dd, d_in, F0));
        if (n1) LAUNCH(ctx, st, "k_prep_o1", k_prep_o1<<<n1, 256, 0, st>>>(dd, d_o1, F0, meta));
        // The alphabets are known now: lay the tables and work areas out by what the blocks need (a quality block: 41 x 41
        // cells and a 7 KB table, not 65 536 cells and 258 KB) and size the buffers from the totals.  One synchronisation.
        {
            unsigned long long *d_tot = (unsigned long long *)(S + o_totals);
            LAUNCH(ctx, st, "k_enc_layout", k_enc_layout<<<1, 1024, 0, st>>>((BlockDesc *)(S + o_desc), nblocks, meta, n1, d_tot));
            CUDA_TRY(ctx, cudaMemcpyAsync(pm + rup(po, 64), d_tot, 16, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(ctx, cudaStreamSynchronize(st));
            const uint64_t cells = ((const uint64_t *)(pm + rup(po, 64)))[0], work_total = ((const uint64_t *)(pm + rup(po, 64)))[1];
            Carver cb;
            const size_t b_F1 = cb.take(4 * cells + 16), b_sy = cb.take(4 * cells + 16), b_work = cb.take(work_total + 256);
            if (int rc = pool_reserve(ctx, ctx->enc_big, cb.off)) return rc;
            uint8_t *B = (uint8_t *)ctx->enc_big.p;
            F1 = (uint32_t *)(B + b_F1); syms1 = (uint32_t *)(B + b_sy); work = B + b_work;
            if (cells) CUDA_TRY(ctx, cudaMemsetAsync(F1, 0, 4 * cells, st));
        }
        if (n1) {
            uint32_t max1 = desc[s1[0]].in_len;
            LAUNCH(ctx, st, "k_hist1", k_hist1<<<dim3(n1, (max1 + HIST1_CHUNK - 1) / HIST1_CHUNK), 256, 0, st>>>(dd, d_o1, d_in, meta, F1));
            LAUNCH(ctx, st, "k_tables1", k_tables1<<<n1, 256, 0, st>>>(dd, d_o1, meta, F1, syms1, work, tlen));
            ON_HP_STREAM(ctx, st, LAUNCH(ctx, hs, "k_enc1", k_enc<1><<<(n1 + 7) / 8, ENC_THREADS, ENC_SMEM(1), hs>>>(dd, d_s1, n1, d_in, syms1, meta, work, plen, d_status, dbg)));
        }
        if (n0) {
            LAUNCH(ctx, st, "k_tables0", k_tables0<<<n0, 256, 0, st>>>(dd, d_o0, F0, syms0, work, tlen));
            ON_HP_STREAM(ctx, st, LAUNCH(ctx, hs, "k_enc0", k_enc<0><<<(n0 + 7) / 8, ENC_THREADS, ENC_SMEM(0), hs>>>(dd, d_s0, n0, d_in, syms0, meta, work, plen, d_status, dbg)));
        }
    }
    LAUNCH(ctx, st, "k_scan_out", k_scan_out<<<1, 1024, 0, st>>>(dd, nblocks, tlen, plen, out_cap, d_out_off, d_out_len, d_status));
#ifdef CRAMGPU_TIMING_EXPERIMENTS
    if (dbg & 16) {
        unsigned long long h[16];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, g_enc_dbg, sizeof h);
        fprintf(stderr, "k_enc busy cycles per warp (CTA 0), total %llu:", h[12]);
        for (int w = 0; w < 12; w++) fprintf(stderr, " %llu", h[w]);
        fprintf(stderr, "\n");
    }
#endif
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->pend = EncodePending{dd, work, tlen, plen, d_out_off, d_out_len, nblocks,
                              (n0 + n1) ? 9 + (uint64_t)(n1 ? O1_TAB_CAP : O0_TAB_CAP) + max_len + max_len / 4 + 80 : 0};
    if (d_out) return rans_encode_pack(ctx, d_out, st);
    return CRAMGPU_OK;
}

// second half of the encode pipeline: needs the output buffer (the host entry point sizes it from the
// device-computed total before calling this)
int rans_encode_pack(cramgpu_ctx *ctx, uint8_t *d_out, cudaStream_t st)
{
    const EncodePending &p = ctx->pend;
    if (p.max_out) {
        LAUNCH(ctx, st, "k_pack", k_pack<<<dim3(p.nblocks, (uint32_t)((p.max_out + PACK_CHUNK - 1) / PACK_CHUNK)), 256, 0, st>>>(
            p.desc, p.work, p.tlen, p.plen, p.out_off, p.out_len, d_out));
        CUDA_TRY(ctx, cudaGetLastError());
    }
    return CRAMGPU_OK;
}

// ------------------------------------------------------------------------------------------------
// cramgpu_rans0_encode_static: order-0 compression of ONE block for callers that must not allocate.
//
// Some hosts cannot ask the CUDA runtime for memory while they work: a JVM-side or serving-side memory manager
// may already own every byte of HBM, so a cudaMalloc would fail or break its accounting, and latency-sensitive
// callers cannot accept the device-wide synchronisation a cudaMalloc / cudaFree implies.  For them this entry
// point runs the same kernels as cramgpu_rans_encode_dev, but every byte of scratch (block descriptor, histogram,
// coder table, table bytes, payload work area) lives in the file-scope device array below, which each GPU that
// loads the library holds from load to exit.  Nothing is allocated or freed here, on the device or through torch.
//
//  - limit: at most CRAMGPU_STATIC_MAX_RAW (1 MiB) raw bytes per call; larger inputs are rejected.
//  - the first call on a device looks up the array's device address and raises the coder kernel's shared-memory
//    limit; make that first call before any stream capture begins.
//  - all calls share the one array and it keeps the previous call's contents; the counters that must start at zero
//    are cleared on `stream` at the start of every call.  Two calls whose work overlaps on the GPU (different
//    streams, no ordering between them) race on the array: order them.
#define CRAMGPU_STATIC_MAX_RAW (1u << 20)
#define NA_OFF_DESC 0u
#define NA_OFF_SLOT 256u
#define NA_OFF_ZERO 512u                 // [F0 1024 B][tlen 4][plen 4][out_off 2 x 8]  -- cleared every call
#define NA_ZERO_BYTES 1088u
#define NA_OFF_TAB0 2048u
#define NA_OFF_WORK 4096u
#define NA_WORK_MAX (O0_TAB_CAP + ((CRAMGPU_STATIC_MAX_RAW + CRAMGPU_STATIC_MAX_RAW / 4 + 64 + 15) & ~15u))
#define NA_BYTES ((NA_OFF_WORK + NA_WORK_MAX + 255u) & ~255u)

static __device__ __align__(256) unsigned char g_static_scratch[NA_BYTES];

__global__ void k_static_setup(BlockDesc d, BlockDesc *desc, uint32_t *slot)
{
    *desc = d;
    *slot = 0;
}

#include <mutex>

extern "C" int cramgpu_rans0_encode_static(const uint8_t *d_in, uint32_t n, uint8_t *d_out, uint32_t out_cap,
                                           uint32_t *d_out_len, int32_t *d_status, void *stream)
{
    if (!d_out || !d_out_len || !d_status || (n && !d_in)) return ctx_fail(nullptr, CRAMGPU_ERR_ARG, "cramgpu_rans0_encode_static: null argument");
    if (n > CRAMGPU_STATIC_MAX_RAW)
        return ctx_fail(nullptr, CRAMGPU_ERR_ARG, "cramgpu_rans0_encode_static: %u raw bytes exceed the %u-byte limit of the static scratch; "
                        "use cramgpu_rans_encode_dev for larger blocks", n, (unsigned)CRAMGPU_STATIC_MAX_RAW);
    if (reinterpret_cast<uintptr_t>(d_in) & 15) return ctx_fail(nullptr, CRAMGPU_ERR_ARG, "cramgpu_rans0_encode_static: d_in must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;

    // once per device, with that device current: the array's name is only a host-side handle, kernels need its device address
    static unsigned char *dev_ptr[64];
    static bool ready[64];
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) return ctx_fail(nullptr, CRAMGPU_ERR_CUDA, "cramgpu_rans0_encode_static: no current CUDA device");
    unsigned char *S = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!ready[dev]) {
            unsigned char *p = nullptr;
            e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
            if (e != cudaSuccess || !p) return ctx_cuda(nullptr, e, "cudaGetSymbolAddress(g_static_scratch)");
            e = cudaFuncSetAttribute(k_enc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, ENC_SMEM(0));
            if (e != cudaSuccess) return ctx_cuda(nullptr, e, "cudaFuncSetAttribute(k_enc<0>)");
            dev_ptr[dev] = p;
            ready[dev] = true;
        }
        S = dev_ptr[dev];
    }
    e = cudaMemsetAsync(d_status, 0, sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_out_len, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return ctx_cuda(nullptr, e, "cudaMemsetAsync(outputs)");
    if (n == 0) return CRAMGPU_OK;                                   // empty in, empty out [RANS.java:60-61]

    BlockDesc d = BlockDesc();
    d.in_off = 0; d.in_len = n; d.order = 0; d.tab_cap = O0_TAB_CAP;
    d.work_cap = (uint32_t)(O0_TAB_CAP + (((uint64_t)n + n / 4 + 64 + 15) & ~15ull));
    d.work_off = 0;
    BlockDesc *desc = (BlockDesc *)(S + NA_OFF_DESC);
    uint32_t *slot = (uint32_t *)(S + NA_OFF_SLOT);
    uint32_t *F0 = (uint32_t *)(S + NA_OFF_ZERO), *tlen = F0 + 256, *plen = F0 + 257;
    uint64_t *out_off = (uint64_t *)(S + NA_OFF_ZERO + 1032);
    uint32_t *tab0 = (uint32_t *)(S + NA_OFF_TAB0);
    uint8_t *work = S + NA_OFF_WORK;

    e = cudaMemsetAsync(S + NA_OFF_ZERO, 0, NA_ZERO_BYTES, st);      // what must start at zero
    if (e != cudaSuccess) return ctx_cuda(nullptr, e, "cudaMemsetAsync(static scratch)");
    k_static_setup<<<1, 1, 0, st>>>(d, desc, slot);
    k_hist0<<<dim3(1, (n + HIST_CHUNK - 1) / HIST_CHUNK), 256, 0, st>>>(desc, d_in, F0);
    k_tables0<<<1, 256, 0, st>>>(desc, slot, F0, tab0, work, tlen);
    k_enc<0><<<1, ENC_THREADS, ENC_SMEM(0), st>>>(desc, slot, 1, d_in, tab0, nullptr, work, plen, d_status, 0);
    k_scan_out<<<1, 1024, 0, st>>>(desc, 1, tlen, plen, out_cap, out_off, d_out_len, d_status);
    uint64_t max_out = 9 + (uint64_t)O0_TAB_CAP + n + n / 4 + 80;
    k_pack<<<dim3(1, (uint32_t)((max_out + PACK_CHUNK - 1) / PACK_CHUNK)), 256, 0, st>>>(desc, work, tlen, plen, out_off, d_out_len, d_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return ctx_cuda(nullptr, e, "cramgpu_rans0_encode_static launch");
    return CRAMGPU_OK;
}
