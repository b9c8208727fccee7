// blocks.cu -- the CRAM v3 block layer around the coded series (the block part).
//
// Block.write(3, os) [htsjdk Block.java:272-294]: method | content type | ITF8 content id | ITF8 compressed size | ITF8 raw
// size | data | CRC32 of all of it, little endian [CRC32OutputStream.java:50-51].  cramgpu_blocks_frame_dev wraps every
// compressed block the codec left in device memory into that form, back to back, so that one copy brings a slice's blocks to
// the Java host ready to be written -- the host neither touches the bytes again nor runs java.util.zip.CRC32 over them.
//
// "sizes, scan, write" like bam.cu.  The CRC of a block (up to megabytes) is not one thread's job: the framed block is cut into
// 16 KB chunks counted from its END, a chunk into 64-byte pieces counted from its end, so that every piece but a block's first
// and every chunk but its first is full.  A thread computes the raw CRC register of its piece (initial value 0, or ~0 for the
// block's first piece); registers combine by crc(A || B) = shift_|B|(crc(A)) ^ crc_0(B), where shift_n -- running the register
// through n zero bytes -- is linear: 32 column vectors per n, and only n = 64 and n = 16 384 occur.
#include "internal.cuh"
#include <string.h>

namespace {

#define FR_CHUNK 16384u
#define FR_PIECE 64u
#define FR_THREADS (FR_CHUNK / FR_PIECE)          // 256

struct FrameArgs {
    uint32_t nblocks;
    const uint8_t *comp;              // base of the compressed bytes (or NULL: comp_off holds absolute addresses)
    const uint64_t *comp_off;
    const uint32_t *comp_len, *raw_len;
    const int32_t *content_id;
    const uint8_t *method, *content_type;
    uint32_t *size;                   // framed size per block
    const uint64_t *out_off;          // write pass
    uint8_t *out;
    uint32_t *chunk_crc;              // [sum of chunks] raw registers
    const uint64_t *chunk_first;      // first chunk slot of every block
    uint32_t shiftp[8][32], shift16k[32];      // shift by 64 << k bytes (k = 0..7), by 16 384 bytes
};

__device__ __forceinline__ uint32_t itf8_bytes(uint32_t v) { return v < (1u << 7) ? 1u : v < (1u << 14) ? 2u : v < (1u << 21) ? 3u : v < (1u << 28) ? 4u : 5u; }
// ITF8.writeUnsignedITF8 [ITF8.java:53-84]
__device__ __forceinline__ uint32_t itf8_write(uint8_t *p, uint32_t v)
{
    if (v < (1u << 7)) { p[0] = (uint8_t)v; return 1; }
    if (v < (1u << 14)) { p[0] = (uint8_t)((v >> 8) | 0x80); p[1] = (uint8_t)v; return 2; }
    if (v < (1u << 21)) { p[0] = (uint8_t)((v >> 16) | 0xC0); p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)v; return 3; }
    if (v < (1u << 28)) { p[0] = (uint8_t)((v >> 24) | 0xE0); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; return 4; }
    p[0] = (uint8_t)(0xF0 | ((v >> 28) & 0xFF)); p[1] = (uint8_t)(v >> 20); p[2] = (uint8_t)(v >> 12); p[3] = (uint8_t)(v >> 4); p[4] = (uint8_t)v;
    return 5;
}
__device__ __forceinline__ uint32_t header_bytes(const FrameArgs &a, uint32_t b) { return 2u + itf8_bytes((uint32_t)a.content_id[b]) + itf8_bytes(a.comp_len[b]) + itf8_bytes(a.raw_len[b]); }

__global__ void __launch_bounds__(256) k_frame_size(FrameArgs a, uint32_t *nchunks)
{
    const uint32_t b = blockIdx.x * 256 + threadIdx.x;
    if (b >= a.nblocks) return;
    const uint32_t body = header_bytes(a, b) + a.comp_len[b];
    a.size[b] = body + 4;
    nchunks[b] = (body + FR_CHUNK - 1) / FR_CHUNK;
}

__device__ __forceinline__ uint32_t shift_reg(uint32_t r, const uint32_t *cols)
{
    uint32_t v = 0;
#pragma unroll
    for (int j = 0; j < 32; j++) v ^= ((r >> j) & 1u) ? cols[j] : 0u;
    return v;
}

// one CTA per chunk of any block (chunk_first[] says whose: a grid of blocks x chunks-of-the-longest would be mostly CTAs
// with nothing to do); chunk index counts from the block's start, boundaries from its end
__global__ void __launch_bounds__(FR_THREADS) k_frame_write(FrameArgs a)
{
    __shared__ uint32_t T[256];
    __shared__ uint8_t hdr[20];
    __shared__ uint32_t part[FR_THREADS];
    __shared__ uint32_t s_hlen;
    const uint32_t tid = threadIdx.x;
    uint32_t b;
    {
        uint32_t lo_b = 0, hi_b = a.nblocks;                          // largest b with chunk_first[b] <= blockIdx.x
        while (hi_b - lo_b > 1) { const uint32_t mid = (lo_b + hi_b) >> 1; if (a.chunk_first[mid] <= blockIdx.x) lo_b = mid; else hi_b = mid; }
        b = lo_b;
    }
    const uint32_t ck = blockIdx.x - (uint32_t)a.chunk_first[b];
    const uint32_t clen = a.comp_len[b];
    if (tid == 0) {
        uint32_t h = 0;
        hdr[h++] = a.method[b]; hdr[h++] = a.content_type[b];
        h += itf8_write(hdr + h, (uint32_t)a.content_id[b]); h += itf8_write(hdr + h, clen); h += itf8_write(hdr + h, a.raw_len[b]);
        s_hlen = h;
    }
    {
        uint32_t c = tid;
#pragma unroll
        for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        T[tid] = c;
    }
    __syncthreads();
    const uint32_t hlen = s_hlen, body = hlen + clen;
    const uint32_t nch = (body + FR_CHUNK - 1) / FR_CHUNK;
    if (ck >= nch) return;
    // chunk ck covers [lo, hi) of the body; the first chunk is the short one
    const uint32_t first_len = body - (nch - 1) * FR_CHUNK;
    const uint32_t lo = ck == 0 ? 0u : first_len + (ck - 1) * FR_CHUNK, hi = ck == 0 ? first_len : lo + FR_CHUNK;
    const uint8_t *src = (a.comp ? a.comp : (const uint8_t *)0) + a.comp_off[b];
    uint8_t *dst = a.out + a.out_off[b];
    // pieces of 64 bytes counted from the chunk's end: piece t of np covers [hi - (np - t) * 64, ...), the first one short.
    // The chunk passes through shared memory on its way out (contiguous reads and writes), a piece per 68-byte row so that the
    // threads' byte-by-byte walks down their rows hit different banks.
    __shared__ uint8_t stage[FR_THREADS * 68];
    const uint32_t np = (hi - lo + FR_PIECE - 1) / FR_PIECE;
    for (uint32_t p = lo + tid; p < hi; p += FR_THREADS) {
        const uint8_t byte = p < hlen ? hdr[p] : src[p - hlen];
        dst[p] = byte;
        const uint32_t d = hi - 1 - p;                                     // distance from the chunk's end
        stage[(np - 1 - d / FR_PIECE) * 68 + (FR_PIECE - 1 - d % FR_PIECE)] = byte;
    }
    __syncthreads();
    uint32_t reg = 0;
    if (tid < np) {
        const uint32_t first = tid == 0 ? FR_PIECE - ((hi - lo) - (np - 1) * FR_PIECE) : 0u;     // the first piece starts inside its row
        reg = (ck == 0 && tid == 0) ? ~0u : 0u;
        for (uint32_t i = first; i < FR_PIECE; i++) reg = T[(reg ^ stage[tid * 68 + i]) & 0xFF] ^ (reg >> 8);
    }
    // combine as a tree over e = distance from the chunk's end: at level k the piece run ending at e (a multiple of 2^(k+1))
    // takes in the run of 64 << k bytes before it, which is (e + 2^k); a missing left run is the empty one (register 0)
    part[tid < np ? (np - 1 - tid) : tid] = reg;                          // part[] indexed by e; slots >= np hold 0
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t stride = 1u << k;
        uint32_t v = 0;
        const bool mine = (tid & (2 * stride - 1)) == 0 && tid + stride < FR_THREADS;
        if (mine) v = shift_reg(part[tid + stride], a.shiftp[k]) ^ part[tid];
        __syncthreads();
        if (mine) part[tid] = v;
        __syncthreads();
    }
    if (tid == 0) a.chunk_crc[blockIdx.x] = part[0];
}

__global__ void __launch_bounds__(256) k_frame_crc(FrameArgs a, const uint32_t *nchunks)
{
    const uint32_t b = blockIdx.x * 256 + threadIdx.x;
    if (b >= a.nblocks) return;
    const uint32_t nch = nchunks[b];
    const uint32_t *c = a.chunk_crc + a.chunk_first[b];
    uint32_t r = nch ? c[0] : ~0u;
    for (uint32_t k = 1; k < nch; k++) r = shift_reg(r, a.shift16k) ^ c[k];
    const uint32_t crc = ~r;
    uint8_t *p = a.out + a.out_off[b] + a.size[b] - 4;
    p[0] = (uint8_t)crc; p[1] = (uint8_t)(crc >> 8); p[2] = (uint8_t)(crc >> 16); p[3] = (uint8_t)(crc >> 24);
}

// exclusive scans (single CTA; block counts are thousands): sizes rounded up to 16 -> out_off, chunk counts -> chunk_first
__global__ void __launch_bounds__(1024) k_frame_scan(const uint32_t *size, const uint32_t *nchunks, uint32_t n, uint64_t *out_off,
                                                     uint64_t *chunk_first, uint32_t *max_chunks)
{
    uint32_t mx = 0;
    for (uint32_t i = threadIdx.x; i < n; i += 1024) mx = max(mx, nchunks[i]);
    mx = __reduce_max_sync(0xFFFFFFFFu, mx);
    if ((threadIdx.x & 31) == 0) atomicMax(max_chunks, mx);
    __shared__ unsigned long long sw[2][32];
    __shared__ unsigned long long carry[2];
    if (threadIdx.x < 2) carry[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t i0 = 0; i0 < n; i0 += 1024) {
        const uint32_t i = i0 + threadIdx.x;
        unsigned long long v0 = i < n ? (((unsigned long long)size[i] + 15ull) & ~15ull) : 0ull, v1 = i < n ? nchunks[i] : 0ull;
        unsigned long long s0 = v0, s1 = v1;
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long t0 = __shfl_up_sync(0xFFFFFFFFu, s0, d), t1 = __shfl_up_sync(0xFFFFFFFFu, s1, d);
            if (lane >= (uint32_t)d) { s0 += t0; s1 += t1; }
        }
        if (lane == 31) { sw[0][w] = s0; sw[1][w] = s1; }
        __syncthreads();
        unsigned long long b0 = carry[0], b1 = carry[1], tot0 = 0, tot1 = 0;
        for (uint32_t k = 0; k < 32; k++) { if (k < w) { b0 += sw[0][k]; b1 += sw[1][k]; } tot0 += sw[0][k]; tot1 += sw[1][k]; }
        if (i < n) { out_off[i] = b0 + s0 - v0; chunk_first[i] = b1 + s1 - v1; }
        __syncthreads();
        if (threadIdx.x == 0) { carry[0] += tot0; carry[1] += tot1; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out_off[n] = carry[0]; chunk_first[n] = carry[1]; }
}

// columns of shift_n: the register 1 << j run through n zero bytes
void shift_columns(uint32_t n, uint32_t *cols)
{
    uint32_t T[256];
    for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; T[i] = c; }
    for (int j = 0; j < 32; j++) {
        uint32_t r = 1u << j;
        for (uint32_t k = 0; k < n; k++) r = T[r & 0xFF] ^ (r >> 8);
        cols[j] = r;
    }
}


// ---- the container around the blocks (container part) -------------------------------------------------------
// ContainerIO.writeContainer [ContainerIO.java:152-209] for containers of one slice: container header | compression header
// block | slice header block | core block | external blocks, tightly packed.  The framing kernels above run over an EXTENDED
// block list (two more blocks per container: the compression header the Java host built, and the slice header content
// k_cont_prepare serialises from the device-side slice header); k_cont_layout places them, k_cont_header writes the headers.
#define CT_TAGS 1u

struct ContArgs {
    uint32_t n;                        // containers
    uint32_t flags;
    const cramgpu_slice_header *slices;
    const uint64_t *grc;               // global record counter per container
    const uint64_t *ch_off;            // compression header bytes per container (offsets into ch)
    const uint8_t *ch;
    const uint32_t *block_first;       // [n + 1] into the caller's block table
    const uint8_t *comp; const uint64_t *comp_off; const uint32_t *comp_len, *raw_len;      // the caller's block table
    const int32_t *content_id; const uint8_t *method, *content_type;
    uint8_t *sh; uint32_t sh_stride;   // slice header content slots
    // the extended table (absolute addresses in x_off)
    uint64_t *x_off; uint32_t *x_len, *x_raw; int32_t *x_id; uint8_t *x_method, *x_type;
    uint8_t sha1[20], sha512[64];
};

__device__ __forceinline__ uint32_t ltf8_bytes(uint64_t v)
{
    for (uint32_t n = 1; n < 9; n++) if ((v >> (7 * n)) == 0) return n;
    return 9;
}
// LTF8.writeUnsignedLTF8 [LTF8.java:104-190]
__device__ __forceinline__ uint32_t ltf8_write(uint8_t *p, uint64_t v)
{
    const uint32_t n = ltf8_bytes(v);
    if (n == 9) { p[0] = 0xFF; for (int k = 0; k < 8; k++) p[1 + k] = (uint8_t)(v >> (8 * (7 - k))); return 9; }
    for (uint32_t k = 0; k < n; k++) p[k] = (uint8_t)(v >> (8 * (n - 1 - k)));
    p[0] |= (uint8_t)(0xFFu << (9 - n));
    return n;
}
__device__ __forceinline__ uint32_t put_tag(uint8_t *p, char t0, char t1, const uint8_t *val, uint32_t len)
{
    // BinaryTagCodec.writeTag of a byte[] [BinaryTagCodec.java:185-228]: tag, 'B', 'c', int32 count, the bytes
    p[0] = (uint8_t)t0; p[1] = (uint8_t)t1; p[2] = 'B'; p[3] = 'c';
    p[4] = (uint8_t)len; p[5] = (uint8_t)(len >> 8); p[6] = (uint8_t)(len >> 16); p[7] = (uint8_t)(len >> 24);
    for (uint32_t i = 0; i < len; i++) p[8 + i] = val[i];
    return 8 + len;
}

// one thread per container: slice header content [SliceIO.java:76-107] and the container's rows of the extended table
__global__ void __launch_bounds__(128) k_cont_prepare(ContArgs a)
{
    const uint32_t c = blockIdx.x * 128 + threadIdx.x;
    if (c >= a.n) return;
    const uint32_t b0 = a.block_first[c], b1 = a.block_first[c + 1], nb = b1 - b0;     // core + external
    const cramgpu_slice_header h = a.slices[c];
    uint8_t *p = a.sh + (size_t)c * a.sh_stride;
    uint32_t o = 0;
    o += itf8_write(p + o, (uint32_t)h.sequence_id); o += itf8_write(p + o, (uint32_t)h.alignment_start);
    o += itf8_write(p + o, (uint32_t)h.alignment_span); o += itf8_write(p + o, h.n_records);
    o += ltf8_write(p + o, a.grc[c]);
    o += itf8_write(p + o, nb);                                            // nofBlocks [SliceIO.java:132]
    const uint32_t next = nb ? nb - 1 : 0;
    o += itf8_write(p + o, next);                                          // CramArray.write: count, elements
    for (uint32_t j = 1; j < nb; j++) o += itf8_write(p + o, (uint32_t)a.content_id[b0 + j]);
    o += itf8_write(p + o, 0xFFFFFFFFu);                                   // embeddedRefBlockContentID = -1
    for (int i = 0; i < 16; i++) p[o++] = h.ref_md5[i];
    if (a.flags & CT_TAGS) {                                               // the slice's tag list: B1 S1 B5 S5 BD SD
        uint8_t w[4];
        o += put_tag(p + o, 'B', '1', a.sha1, 20); o += put_tag(p + o, 'S', '1', a.sha1, 20);
        o += put_tag(p + o, 'B', '5', a.sha512, 64); o += put_tag(p + o, 'S', '5', a.sha512, 64);
        w[0] = (uint8_t)h.bd; w[1] = (uint8_t)(h.bd >> 8); w[2] = (uint8_t)(h.bd >> 16); w[3] = (uint8_t)(h.bd >> 24);
        o += put_tag(p + o, 'B', 'D', w, 4);
        w[0] = (uint8_t)h.sd; w[1] = (uint8_t)(h.sd >> 8); w[2] = (uint8_t)(h.sd >> 16); w[3] = (uint8_t)(h.sd >> 24);
        o += put_tag(p + o, 'S', 'D', w, 4);
    }
    const uint32_t e0 = b0 + 2 * c;
    const uint32_t chl = (uint32_t)(a.ch_off[c + 1] - a.ch_off[c]);
    a.x_off[e0] = (uint64_t)(uintptr_t)(a.ch + a.ch_off[c]); a.x_len[e0] = chl; a.x_raw[e0] = chl;
    a.x_id[e0] = 0; a.x_method[e0] = 0; a.x_type[e0] = 1;                   // RAW, COMPRESSION_HEADER [ContainerIO.java:171-181]
    a.x_off[e0 + 1] = (uint64_t)(uintptr_t)p; a.x_len[e0 + 1] = o; a.x_raw[e0 + 1] = o;
    a.x_id[e0 + 1] = 0; a.x_method[e0 + 1] = 0; a.x_type[e0 + 1] = 2;       // RAW, MAPPED_SLICE [Block.buildNewSliceHeaderBlock]
    for (uint32_t j = 0; j < nb; j++) {
        const uint32_t e = e0 + 2 + j, b = b0 + j;
        a.x_off[e] = (uint64_t)(uintptr_t)(a.comp ? a.comp : (const uint8_t *)0) + a.comp_off[b];
        a.x_len[e] = a.comp_len[b]; a.x_raw[e] = a.raw_len[b];
        a.x_id[e] = a.content_id[b]; a.x_method[e] = a.method[b]; a.x_type[e] = a.content_type[b];
    }
}

struct ContLayout {
    uint32_t n;
    const cramgpu_slice_header *slices;
    const uint64_t *grc;
    const uint32_t *block_first;
    const uint32_t *x_size;            // framed size per extended block
    uint64_t *x_out_off;               // [next + 1]
    uint64_t *cont_off;                // [n + 1]
    uint32_t *hdr_len;                 // [n]
};

__device__ __forceinline__ void cont_fields(const cramgpu_slice_header &h, int32_t &seq, int32_t &start, int32_t &span)
{
    // ContainerFactory.buildContainer [:93-94] and calculateAlignmentBoundaries [:108-123] for one slice: a container of an
    // unplaced slice keeps its initial -1 / -1 / 0 [Container.java:28-30]
    seq = h.sequence_id;
    if (seq != -1) { start = h.alignment_start; span = h.alignment_span; } else { start = -1; span = 0; }
}

// single CTA: container sizes (header + blocks, tight), their exclusive scan, then every block's place
__global__ void __launch_bounds__(1024) k_cont_layout(ContLayout a)
{
    __shared__ unsigned long long sw[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t c0 = 0; c0 < a.n; c0 += 1024) {
        const uint32_t c = c0 + threadIdx.x;
        unsigned long long tot = 0;
        uint32_t hl = 0;
        if (c < a.n) {
            const uint32_t e0 = a.block_first[c] + 2 * c, e1 = a.block_first[c + 1] + 2 * (c + 1);
            unsigned long long body = 0;
            for (uint32_t e = e0; e < e1; e++) body += a.x_size[e];
            const cramgpu_slice_header h = a.slices[c];
            int32_t seq, start, span;
            cont_fields(h, seq, start, span);
            hl = 4 + itf8_bytes((uint32_t)seq) + itf8_bytes((uint32_t)start) + itf8_bytes((uint32_t)span) + itf8_bytes(h.n_records) +
                 ltf8_bytes(a.grc[c]) + ltf8_bytes(h.bases) + itf8_bytes(e1 - e0) + 1 + itf8_bytes(a.x_size[e0]) + 4;
            a.hdr_len[c] = hl;
            tot = hl + body;
        }
        unsigned long long s = tot;
        for (int d = 1; d < 32; d <<= 1) { unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, s, d); if (lane >= (uint32_t)d) s += t; }
        if (lane == 31) sw[w] = s;
        __syncthreads();
        unsigned long long base = carry, all = 0;
        for (uint32_t k = 0; k < 32; k++) { if (k < w) base += sw[k]; all += sw[k]; }
        if (c < a.n) {
            unsigned long long at = base + s - tot;
            a.cont_off[c] = at;
            at += hl;
            const uint32_t e0 = a.block_first[c] + 2 * c, e1 = a.block_first[c + 1] + 2 * (c + 1);
            for (uint32_t e = e0; e < e1; e++) { a.x_out_off[e] = at; at += a.x_size[e]; }
        }
        __syncthreads();
        if (threadIdx.x == 0) carry += all;
        __syncthreads();
    }
    if (threadIdx.x == 0) { a.cont_off[a.n] = carry; a.x_out_off[a.block_first[a.n] + 2 * a.n] = carry; }
}

// one thread per container: ContainerHeaderIO.writeContainerHeader(3, ...) [ContainerHeaderIO.java:69-84]
__global__ void __launch_bounds__(128) k_cont_header(ContLayout a, uint8_t *out)
{
    const uint32_t c = blockIdx.x * 128 + threadIdx.x;
    if (c >= a.n) return;
    const uint32_t e0 = a.block_first[c] + 2 * c, e1 = a.block_first[c + 1] + 2 * (c + 1);
    const cramgpu_slice_header h = a.slices[c];
    int32_t seq, start, span;
    cont_fields(h, seq, start, span);
    uint8_t *p = out + a.cont_off[c];
    const uint32_t hl = a.hdr_len[c];
    const uint32_t bytes = (uint32_t)(a.cont_off[c + 1] - a.cont_off[c] - hl);               // containerByteSize: the blocks
    uint32_t o = 0;
    p[o++] = (uint8_t)bytes; p[o++] = (uint8_t)(bytes >> 8); p[o++] = (uint8_t)(bytes >> 16); p[o++] = (uint8_t)(bytes >> 24);
    o += itf8_write(p + o, (uint32_t)seq); o += itf8_write(p + o, (uint32_t)start); o += itf8_write(p + o, (uint32_t)span);
    o += itf8_write(p + o, h.n_records); o += ltf8_write(p + o, a.grc[c]); o += ltf8_write(p + o, h.bases);
    o += itf8_write(p + o, e1 - e0);                                       // blockCount [ContainerIO.java:183-193]
    o += itf8_write(p + o, 1u); o += itf8_write(p + o, a.x_size[e0]);      // landmarks: the slice starts behind the compression header block
    uint32_t r = ~0u;
    for (uint32_t i = 0; i < o; i++) { r ^= p[i]; for (int k = 0; k < 8; k++) r = (r & 1) ? 0xEDB88320u ^ (r >> 1) : r >> 1; }
    r = ~r;
    p[o++] = (uint8_t)r; p[o++] = (uint8_t)(r >> 8); p[o++] = (uint8_t)(r >> 16); p[o++] = (uint8_t)(r >> 24);
}

}  // namespace

extern "C" int cramgpu_blocks_frame_dev(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_comp, const uint64_t *d_comp_off,
                                        const uint32_t *d_comp_len, const uint32_t *d_raw_len, const int32_t *content_id,
                                        const uint8_t *method, const uint8_t *content_type, uint8_t **d_out, uint64_t *out_len,
                                        uint64_t **d_out_off, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!d_out || !out_len || !d_out_off) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    *d_out = nullptr; *out_len = 0; *d_out_off = nullptr;
    if (nblocks == 0) return CRAMGPU_OK;
    if (!d_comp_off || !d_comp_len || !d_raw_len || !content_id || !method || !content_type) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    const size_t n = nblocks;
    Carver cv;
    const size_t o_id = cv.take(4 * n), o_m = cv.take(n), o_t = cv.take(n), o_size = cv.take(4 * n), o_nch = cv.take(4 * n);
    const size_t o_off = cv.take(8 * (n + 1)), o_cf = cv.take(8 * (n + 1)), o_mx = cv.take(16);
    if (int rc = pool_reserve(ctx, ctx->frame_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->frame_scratch.p;
    const size_t up_bytes = rup(4 * n, 16) + 2 * rup(n, 16);
    if (int rc = pinned_reserve(ctx, &ctx->pin_frame, &ctx->pin_frame_cap, up_bytes + 128)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_frame));
    uint8_t *pm = (uint8_t *)ctx->pin_frame;
    memcpy(pm, content_id, 4 * n); memcpy(pm + rup(4 * n, 16), method, n); memcpy(pm + rup(4 * n, 16) + rup(n, 16), content_type, n);
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_id, pm, 4 * n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_m, pm + rup(4 * n, 16), n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_t, pm + rup(4 * n, 16) + rup(n, 16), n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_frame, st));
    FrameArgs a;
    memset(&a, 0, sizeof a);
    a.nblocks = nblocks; a.comp = d_comp; a.comp_off = d_comp_off; a.comp_len = d_comp_len; a.raw_len = d_raw_len;
    a.content_id = (const int32_t *)(S + o_id); a.method = S + o_m; a.content_type = S + o_t;
    a.size = (uint32_t *)(S + o_size); a.out_off = (const uint64_t *)(S + o_off); a.chunk_first = (const uint64_t *)(S + o_cf);
    struct Cols { uint32_t p[8][32], c16k[32]; Cols() { for (int k = 0; k < 8; k++) shift_columns(FR_PIECE << k, p[k]); shift_columns(FR_CHUNK, c16k); } };
    static const Cols cols;                                          // computed once (thread-safe static)
    memcpy(a.shiftp, cols.p, sizeof cols.p); memcpy(a.shift16k, cols.c16k, sizeof cols.c16k);
    uint32_t *d_nch = (uint32_t *)(S + o_nch);
    LAUNCH(ctx, st, "k_frame_size", k_frame_size<<<(nblocks + 255) / 256, 256, 0, st>>>(a, d_nch));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_mx, 0, 16, st));
    LAUNCH(ctx, st, "k_frame_scan", k_frame_scan<<<1, 1024, 0, st>>>(a.size, d_nch, nblocks, (uint64_t *)(S + o_off), (uint64_t *)(S + o_cf), (uint32_t *)(S + o_mx)));
    // totals size the result buffers: bytes, chunk slots; the longest block sets the grid
    uint64_t *tot = (uint64_t *)(pm + up_bytes);
    CUDA_TRY(ctx, cudaMemcpyAsync(tot, S + o_off + 8 * n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(tot + 1, S + o_cf + 8 * n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(tot + 2, S + o_mx, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t total = tot[0], nchunks_total = tot[1], max_chunks = tot[2] & 0xFFFFFFFFull;
    Carver co;
    const size_t q_out = co.take(total + 16), q_crc = co.take(4 * nchunks_total + 16);
    if (int rc = pool_reserve(ctx, ctx->frame_out, co.off)) return rc;
    uint8_t *O = (uint8_t *)ctx->frame_out.p;
    a.out = O + q_out; a.chunk_crc = (uint32_t *)(O + q_crc);
    (void)max_chunks;
    if (nchunks_total) LAUNCH(ctx, st, "k_frame_write", k_frame_write<<<(uint32_t)nchunks_total, FR_THREADS, 0, st>>>(a));
    LAUNCH(ctx, st, "k_frame_crc", k_frame_crc<<<(nblocks + 255) / 256, 256, 0, st>>>(a, d_nch));
    CUDA_TRY(ctx, cudaGetLastError());
    *d_out = a.out; *out_len = total; *d_out_off = (uint64_t *)(S + o_off);
    return CRAMGPU_OK;
}

extern "C" int cramgpu_containers_assemble_dev(cramgpu_ctx *ctx, const cramgpu_container_in *in, uint8_t **d_out, uint64_t *out_len,
                                               uint64_t **d_out_off, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !d_out || !out_len || !d_out_off) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    *d_out = nullptr; *out_len = 0; *d_out_off = nullptr;
    const uint32_t n = in->n_containers, nblocks = in->nblocks;
    if (n == 0) return CRAMGPU_OK;
    if (!in->d_slices || !in->global_record_counter || !in->comp_header || !in->comp_header_off || !in->block_first ||
        !in->d_comp_off || !in->d_comp_len || !in->d_raw_len || !in->content_id || !in->method || !in->content_type)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->block_first[0] != 0 || in->block_first[n] != nblocks) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "block_first does not cover the block table");
    uint32_t max_ids = 0;
    for (uint32_t c = 0; c < n; c++) {
        if (in->block_first[c + 1] <= in->block_first[c]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "a container needs its core block");
        if (in->comp_header_off[c + 1] < in->comp_header_off[c]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "comp_header_off decreases");
        max_ids = std::max(max_ids, in->block_first[c + 1] - in->block_first[c] - 1);
    }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    const size_t next = (size_t)nblocks + 2 * (size_t)n, nb = nblocks;
    const size_t ch_bytes = in->comp_header_off[n] - in->comp_header_off[0];
    const uint32_t sh_stride = (uint32_t)rup(288 + 5 * (size_t)max_ids, 16);
    Carver cv;
    const size_t o_id = cv.take(4 * nb), o_m = cv.take(nb), o_t = cv.take(nb), o_bf = cv.take(4 * ((size_t)n + 1)), o_grc = cv.take(8 * (size_t)n);
    const size_t o_cho = cv.take(8 * ((size_t)n + 1)), o_ch = cv.take(ch_bytes + 16), o_sh = cv.take((size_t)n * sh_stride);
    const size_t x_off = cv.take(8 * next), x_len = cv.take(4 * next), x_raw = cv.take(4 * next), x_id = cv.take(4 * next);
    const size_t x_m = cv.take(next), x_t = cv.take(next), o_size = cv.take(4 * next), o_nch = cv.take(4 * next);
    const size_t o_off = cv.take(8 * (next + 1)), o_dummy = cv.take(8 * (next + 1)), o_cf = cv.take(8 * (next + 1)), o_mx = cv.take(16);
    const size_t o_coff = cv.take(8 * ((size_t)n + 1)), o_hl = cv.take(4 * (size_t)n);
    if (int rc = pool_reserve(ctx, ctx->frame_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->frame_scratch.p;
    // small host tables -> device through page-locked staging
    Carver pc;
    const size_t p_id = pc.take(4 * nb, 16), p_m = pc.take(nb, 16), p_t = pc.take(nb, 16), p_bf = pc.take(4 * ((size_t)n + 1), 16);
    const size_t p_grc = pc.take(8 * (size_t)n, 16), p_cho = pc.take(8 * ((size_t)n + 1), 16), p_ch = pc.take(ch_bytes, 16), p_tot = pc.take(32, 16);
    if (int rc = pinned_reserve(ctx, &ctx->pin_frame, &ctx->pin_frame_cap, pc.off + 128)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_frame));
    uint8_t *pm = (uint8_t *)ctx->pin_frame;
    memcpy(pm + p_id, in->content_id, 4 * nb); memcpy(pm + p_m, in->method, nb); memcpy(pm + p_t, in->content_type, nb);
    memcpy(pm + p_bf, in->block_first, 4 * ((size_t)n + 1)); memcpy(pm + p_grc, in->global_record_counter, 8 * (size_t)n);
    {
        uint64_t *co = (uint64_t *)(pm + p_cho);
        for (uint32_t c = 0; c <= n; c++) co[c] = in->comp_header_off[c] - in->comp_header_off[0];
        memcpy(pm + p_ch, in->comp_header + in->comp_header_off[0], ch_bytes);
    }
    auto up = [&](size_t dev_o, size_t pin_o, size_t bytes) { return bytes ? cudaMemcpyAsync(S + dev_o, pm + pin_o, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess; };
    CUDA_TRY(ctx, up(o_id, p_id, 4 * nb)); CUDA_TRY(ctx, up(o_m, p_m, nb)); CUDA_TRY(ctx, up(o_t, p_t, nb));
    CUDA_TRY(ctx, up(o_bf, p_bf, 4 * ((size_t)n + 1))); CUDA_TRY(ctx, up(o_grc, p_grc, 8 * (size_t)n));
    CUDA_TRY(ctx, up(o_cho, p_cho, 8 * ((size_t)n + 1))); CUDA_TRY(ctx, up(o_ch, p_ch, ch_bytes));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_frame, st));

    ContArgs ca;
    memset(&ca, 0, sizeof ca);
    ca.n = n; ca.flags = in->flags; ca.slices = in->d_slices; ca.grc = (const uint64_t *)(S + o_grc);
    ca.ch_off = (const uint64_t *)(S + o_cho); ca.ch = S + o_ch; ca.block_first = (const uint32_t *)(S + o_bf);
    ca.comp = in->d_comp; ca.comp_off = in->d_comp_off; ca.comp_len = in->d_comp_len; ca.raw_len = in->d_raw_len;
    ca.content_id = (const int32_t *)(S + o_id); ca.method = S + o_m; ca.content_type = S + o_t;
    ca.sh = S + o_sh; ca.sh_stride = sh_stride;
    ca.x_off = (uint64_t *)(S + x_off); ca.x_len = (uint32_t *)(S + x_len); ca.x_raw = (uint32_t *)(S + x_raw);
    ca.x_id = (int32_t *)(S + x_id); ca.x_method = S + x_m; ca.x_type = S + x_t;
    cramgpu_slice_sha_tags(ca.sha1, ca.sha512);
    LAUNCH(ctx, st, "k_cont_prepare", k_cont_prepare<<<(n + 127) / 128, 128, 0, st>>>(ca));

    FrameArgs a;
    memset(&a, 0, sizeof a);
    a.nblocks = (uint32_t)next; a.comp = nullptr; a.comp_off = ca.x_off; a.comp_len = ca.x_len; a.raw_len = ca.x_raw;
    a.content_id = ca.x_id; a.method = ca.x_method; a.content_type = ca.x_type;
    a.size = (uint32_t *)(S + o_size); a.out_off = (const uint64_t *)(S + o_off); a.chunk_first = (const uint64_t *)(S + o_cf);
    struct Cols { uint32_t p[8][32], c16k[32]; Cols() { for (int k = 0; k < 8; k++) shift_columns(FR_PIECE << k, p[k]); shift_columns(FR_CHUNK, c16k); } };
    static const Cols cols;
    memcpy(a.shiftp, cols.p, sizeof cols.p); memcpy(a.shift16k, cols.c16k, sizeof cols.c16k);
    uint32_t *d_nch = (uint32_t *)(S + o_nch);
    LAUNCH(ctx, st, "k_frame_size", k_frame_size<<<((uint32_t)next + 255) / 256, 256, 0, st>>>(a, d_nch));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_mx, 0, 16, st));
    // chunk slots of the CRC work as in cramgpu_blocks_frame_dev (its padded offsets go to a dummy); the real places are tight
    LAUNCH(ctx, st, "k_frame_scan", k_frame_scan<<<1, 1024, 0, st>>>(a.size, d_nch, (uint32_t)next, (uint64_t *)(S + o_dummy), (uint64_t *)(S + o_cf), (uint32_t *)(S + o_mx)));
    ContLayout cl;
    cl.n = n; cl.slices = in->d_slices; cl.grc = ca.grc; cl.block_first = ca.block_first; cl.x_size = a.size;
    cl.x_out_off = (uint64_t *)(S + o_off); cl.cont_off = (uint64_t *)(S + o_coff); cl.hdr_len = (uint32_t *)(S + o_hl);
    LAUNCH(ctx, st, "k_cont_layout", k_cont_layout<<<1, 1024, 0, st>>>(cl));
    uint64_t *tot = (uint64_t *)(pm + p_tot);
    CUDA_TRY(ctx, cudaMemcpyAsync(tot, S + o_coff + 8 * (size_t)n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(tot + 1, S + o_cf + 8 * next, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t total = tot[0], nchunks_total = tot[1];
    Carver co;
    const size_t q_out = co.take(total + 16), q_crc = co.take(4 * nchunks_total + 16);
    if (int rc = pool_reserve(ctx, ctx->frame_out, co.off)) return rc;
    uint8_t *O = (uint8_t *)ctx->frame_out.p;
    a.out = O + q_out; a.chunk_crc = (uint32_t *)(O + q_crc);
    if (nchunks_total) LAUNCH(ctx, st, "k_frame_write", k_frame_write<<<(uint32_t)nchunks_total, FR_THREADS, 0, st>>>(a));
    LAUNCH(ctx, st, "k_frame_crc", k_frame_crc<<<((uint32_t)next + 255) / 256, 256, 0, st>>>(a, d_nch));
    LAUNCH(ctx, st, "k_cont_header", k_cont_header<<<(n + 127) / 128, 128, 0, st>>>(cl, a.out));
    CUDA_TRY(ctx, cudaGetLastError());
    *d_out = a.out; *out_len = total; *d_out_off = cl.cont_off;
    return CRAMGPU_OK;
}
