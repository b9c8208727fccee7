// bam.cu -- CIGAR rebuild and BAM record (de)serialisation on the device.
//
//   cramgpu_cigar_rebuild_dev   Cram2SamRecordFactory.getCigar2 [htsjdk Cram2SamRecordFactory.java:119-225]: read features -> CIGAR
//   cramgpu_bam_pack_dev        the record layout of cramtools' flat-buffer BAM writer
//                               (src/main/java/htsjdk/samtools/cram/encoding/reader/BAMRecordView.java:70-241), index bin as
//                               SAMRecord.computeIndexingBin [SAMRecord.java:1524-1531] / GenomicIndexUtil.regionToBin [:63-70]
//   cramgpu_bam_unpack_dev      the inverse: a stripe of raw BAM records as Bam_OBA_Supplier cuts them
//                               (src/main/java/htsjdk/samtools/cram/paralell/Bam_OBA_Supplier.java:45-82) -> the structure of
//                               arrays the other entry points take (4-bit bases -> ASCII, pos + 1, ...)
//
// All three are "sizes, scan, write": a record's output position depends on the sizes of the records before it.  Sizes are one
// thread per record; the scan is three small kernels over 4096-item tiles; the writers are one warp per record with the lanes
// over the bytes, so that loads and stores of a record are contiguous.  HBM-bound byte shuffling, no shared memory to speak of.
#include "internal.cuh"
#include <string.h>

namespace {

// ---------------------------------------------------------------------------------------------- exclusive scan u32 -> u64
#define SCAN_TILE 4096u
__global__ void __launch_bounds__(256) k_scan_tile_sums(const uint32_t *__restrict__ in, uint64_t n, uint64_t *__restrict__ tile_sum)
{
    __shared__ unsigned long long s[8];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    unsigned long long v = 0;
    for (uint32_t k = threadIdx.x; k < SCAN_TILE; k += 256) if (base + k < n) v += in[base + k];
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long t = 0; for (int k = 0; k < 8; k++) t += s[k]; tile_sum[blockIdx.x] = t; }
}
// one CTA: tile sums -> exclusive tile offsets (in place); total -> out[n]
__global__ void __launch_bounds__(256) k_scan_tiles(uint64_t *__restrict__ tile_sum, uint64_t ntiles, uint64_t *__restrict__ total_at)
{
    __shared__ unsigned long long s[256], carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint64_t t0 = 0; t0 < ntiles; t0 += 256) {
        const uint64_t i = t0 + threadIdx.x;
        const unsigned long long v = i < ntiles ? tile_sum[i] : 0;
        s[threadIdx.x] = v;
        __syncthreads();
        for (uint32_t d = 1; d < 256; d <<= 1) {
            const unsigned long long a = threadIdx.x >= d ? s[threadIdx.x - d] : 0;
            __syncthreads();
            s[threadIdx.x] += a;
            __syncthreads();
        }
        if (i < ntiles) tile_sum[i] = carry + s[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 255) carry += s[255];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_at = carry;
}
__global__ void __launch_bounds__(256) k_scan_apply(const uint32_t *__restrict__ in, uint64_t n, const uint64_t *__restrict__ tile_off,
                                                    uint64_t *__restrict__ out)
{
    // thread t owns the 16 consecutive items 16 t .. 16 t + 15 of the tile
    __shared__ unsigned long long s[256];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + 16ull * threadIdx.x;
    uint32_t v[16];
    unsigned long long sum = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) { v[k] = base + k < n ? in[base + k] : 0u; sum += v[k]; }
    s[threadIdx.x] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < 256; d <<= 1) {
        const unsigned long long a = threadIdx.x >= d ? s[threadIdx.x - d] : 0;
        __syncthreads();
        s[threadIdx.x] += a;
        __syncthreads();
    }
    unsigned long long at = tile_off[blockIdx.x] + s[threadIdx.x] - sum;
#pragma unroll
    for (int k = 0; k < 16; k++) { if (base + k < n) out[base + k] = at; at += v[k]; }
}

}  // namespace

// sizes[n] (u32) -> off[n + 1] (u64), off[n] = total.  `tiles` = scratch of at least (n / 4096 + 1) u64.
static int scan_sizes(cramgpu_ctx *ctx, const uint32_t *d_sizes, uint64_t n, uint64_t *d_off, uint64_t *d_tiles, cudaStream_t st)
{
    const uint64_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (ntiles == 0) { CUDA_TRY(ctx, cudaMemsetAsync(d_off, 0, 8, st)); return CRAMGPU_OK; }
    LAUNCH(ctx, st, "k_scan_tile_sums", k_scan_tile_sums<<<(uint32_t)ntiles, 256, 0, st>>>(d_sizes, n, d_tiles));
    LAUNCH(ctx, st, "k_scan_tiles", k_scan_tiles<<<1, 256, 0, st>>>(d_tiles, ntiles, d_off + n));
    LAUNCH(ctx, st, "k_scan_apply", k_scan_apply<<<(uint32_t)ntiles, 256, 0, st>>>(d_sizes, n, d_tiles, d_off));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}

// ---------------------------------------------------------------------------------------------- CIGAR from read features
namespace {

enum { CG_M = 0, CG_I, CG_D, CG_N, CG_S, CG_H, CG_P };
__device__ __forceinline__ bool cg_consumes_read(int op) { return op == CG_M || op == CG_I || op == CG_S; }

struct CigarArgs {
    uint64_t n;
    const uint16_t *flag;
    const uint64_t *read_off, *feat_off;
    const uint8_t *f_op;
    const int32_t *f_pos, *f_val, *f_aux;
    uint32_t *count;              // elements per record
    const uint64_t *cigar_off;    // write pass
    uint32_t *cigar;
};

// getCigar2 [Cram2SamRecordFactory.java:119-225], one thread per record; emit(length, op) in element order.
// A record with the unmapped flag has no CIGAR (create() sets "*" [:68-69]).
template <typename E>
__device__ __forceinline__ void cigar_walk(const CigarArgs &a, uint64_t r, E emit)
{
    if (a.flag[r] & 4) return;
    const int32_t read_length = (int32_t)(a.read_off[r + 1] - a.read_off[r]);
    const uint64_t f0 = a.feat_off[r], f1 = a.feat_off[r + 1];
    int last_op = CG_M;
    int32_t last_len = 0, last_pos = 1;
    bool any = false, emitted = false;
    for (uint64_t f = f0; f < f1; f++) {
        const uint8_t op = a.f_op[f];
        int co; int32_t rl;
        switch (op) {                                        // [:148-183]; everything else (quality scores) is skipped before
        case 'I': co = CG_I; rl = a.f_aux[f]; break;         // the position is even looked at? no: the gap comes first [:134-146]
        case 'S': co = CG_S; rl = a.f_aux[f]; break;
        case 'H': co = CG_H; rl = a.f_val[f]; break;
        case 'i': co = CG_I; rl = 1; break;
        case 'D': co = CG_D; rl = a.f_val[f]; break;
        case 'N': co = CG_N; rl = a.f_val[f]; break;
        case 'P': co = CG_P; rl = a.f_val[f]; break;
        case 'X': case 'B': co = CG_M; rl = 1; break;
        default: co = -1; rl = 0; break;
        }
        any = true;
        const int32_t pos = a.f_pos[f];
        const int32_t gap = pos - (last_pos + last_len);
        if (gap > 0) {
            if (last_op != CG_M) { emit(last_len, last_op); emitted = true; last_pos += last_len; last_len = gap; }
            else last_len += gap;
            last_op = CG_M;
        }
        if (co < 0) continue;
        if (last_op != co) {
            if (last_len > 0) { emit(last_len, last_op); emitted = true; }
            last_op = co; last_len = rl; last_pos = pos;
        } else last_len += rl;
        if (!cg_consumes_read(co)) last_pos -= rl;
    }
    if (!any) { emit(read_length, CG_M); return; }           // no features at all [:119-121]
    if (last_op != CG_M) {
        emit(last_len, last_op); emitted = true;
        if (read_length >= last_pos + last_len) emit(read_length - (last_len + last_pos) + 1, CG_M);
    } else if (read_length == 0 || read_length > last_pos - 1) {
        emit(read_length == 0 ? last_len : read_length - last_pos + 1, CG_M); emitted = true;
    }
    if (!emitted) emit(read_length, CG_M);                   // empty list [:220-222]
}

template <bool WRITE>
__global__ void __launch_bounds__(128) k_cigar(CigarArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
    if (r >= a.n) return;
    uint32_t k = 0;
    uint32_t *dst = WRITE ? a.cigar + a.cigar_off[r] : nullptr;
    cigar_walk(a, r, [&](int32_t len, int op) { if (WRITE) dst[k] = ((uint32_t)len << 4) | (uint32_t)op; k++; });
    if (!WRITE) a.count[r] = k;
}

}  // namespace

extern "C" int cramgpu_cigar_rebuild_dev(cramgpu_ctx *ctx, uint64_t n_reads, const uint16_t *d_flag, const uint64_t *d_read_off,
                                         const uint64_t *d_feat_off, const uint8_t *d_f_op, const int32_t *d_f_pos, const int32_t *d_f_val,
                                         const int32_t *d_f_aux, uint32_t **d_cigar, uint64_t **d_cigar_off, uint64_t *cigar_len, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!d_cigar || !d_cigar_off || !cigar_len) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    *d_cigar = nullptr; *d_cigar_off = nullptr; *cigar_len = 0;
    if (n_reads == 0) return CRAMGPU_OK;
    if (!d_flag || !d_read_off || !d_feat_off) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    const uint64_t n = n_reads;
    Carver cv;
    const size_t o_cnt = cv.take(4 * n), o_off = cv.take(8 * (n + 1)), o_tiles = cv.take(8 * (n / SCAN_TILE + 2));
    if (int rc = pool_reserve(ctx, ctx->bam_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->bam_scratch.p;
    CigarArgs a;
    memset(&a, 0, sizeof a);
    a.n = n; a.flag = d_flag; a.read_off = d_read_off; a.feat_off = d_feat_off; a.f_op = d_f_op; a.f_pos = d_f_pos; a.f_val = d_f_val; a.f_aux = d_f_aux;
    a.count = (uint32_t *)(S + o_cnt); a.cigar_off = (const uint64_t *)(S + o_off);
    const uint32_t grid = (uint32_t)((n + 127) / 128);
    LAUNCH(ctx, st, "k_cigar_count", k_cigar<false><<<grid, 128, 0, st>>>(a));
    if (int rc = scan_sizes(ctx, a.count, n, (uint64_t *)(S + o_off), (uint64_t *)(S + o_tiles), st)) return rc;
    if (int rc = pinned_reserve(ctx, &ctx->pin_bam, &ctx->pin_bam_cap, 64)) return rc;
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pin_bam, S + o_off + 8 * n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));                 // the total sizes the result buffer
    const uint64_t total = *(const uint64_t *)ctx->pin_bam;
    if (int rc = pool_reserve(ctx, ctx->bam_cigar, 4 * total + 16)) return rc;
    a.cigar = (uint32_t *)ctx->bam_cigar.p;
    LAUNCH(ctx, st, "k_cigar_write", k_cigar<true><<<grid, 128, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    *d_cigar = a.cigar; *d_cigar_off = (uint64_t *)(S + o_off); *cigar_len = total;
    return CRAMGPU_OK;
}

// ---------------------------------------------------------------------------------------------- BAM records: pack
namespace {

// GenomicIndexUtil.regionToBin [GenomicIndexUtil.java:63-70]
__device__ __forceinline__ uint32_t reg2bin(int32_t beg, int32_t end)
{
    --end;
    if (beg >> 14 == end >> 14) return 4681 + (beg >> 14);
    if (beg >> 17 == end >> 17) return 585 + (beg >> 17);
    if (beg >> 20 == end >> 20) return 73 + (beg >> 20);
    if (beg >> 23 == end >> 23) return 9 + (beg >> 23);
    if (beg >> 26 == end >> 26) return 1 + (beg >> 26);
    return 0;
}

// BAMRecordView.charToCompressedBaseLow / High [BAMRecordView.java:360-530]: 16 = IllegalArgumentException
__device__ __forceinline__ uint32_t base_nybble(uint8_t b)
{
    switch (b) {
    case '=': return 0;
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'M': case 'm': return 3;
    case 'G': case 'g': return 4;
    case 'R': case 'r': return 5;
    case 'S': case 's': return 6;
    case 'V': case 'v': return 7;
    case 'T': case 't': return 8;
    case 'W': case 'w': return 9;
    case 'Y': case 'y': return 10;
    case 'H': case 'h': return 11;
    case 'K': case 'k': return 12;
    case 'D': case 'd': return 13;
    case 'B': case 'b': return 14;
    case 'N': case 'n': case '.': return 15;
    default: return 16;
    }
}

struct PackArgs {
    cramgpu_bam_records in;
    uint32_t *size;               // bytes of every record, with the block_size field
    const uint64_t *rec_off;      // write pass
    uint8_t *out;
    int32_t *status;
};

__global__ void __launch_bounds__(128) k_bam_size(PackArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
    if (r >= a.in.n_reads) return;
    const uint64_t lname = a.in.name_off[r + 1] - a.in.name_off[r] + 1, ncig = a.in.cigar_off[r + 1] - a.in.cigar_off[r];
    const uint64_t lseq = a.in.read_off[r + 1] - a.in.read_off[r], ltag = a.in.tags ? a.in.tag_off[r + 1] - a.in.tag_off[r] : 0;
    a.size[r] = (uint32_t)(36 + lname + 4 * ncig + (lseq + 1) / 2 + lseq + ltag);
}


// copy len bytes src -> dst with the lanes of a warp, B x 32 bytes per round: all loads of a round are issued before its
// stores (a store may alias the next load as far as the compiler knows, and a load that waits behind it costs a round trip
// to memory -- the kernel's time is those round trips, not its bytes)
// a byte load that is predicated instead of branched around, at a constant offset from a pointer that is computed once per
// round: @p LDG.U8 [R + imm].  These kernels are bound by instruction issue -- 486 warp instructions per record when every
// byte paid a 64-bit address add and every bounds check a branch with its reconvergence point (profiles/).
template <int OFF>
__device__ __forceinline__ uint32_t ldg_u8_if(const uint8_t *base, bool p)
{
    uint32_t v = 0;
    asm("{\n .reg .pred q;\n setp.ne.u32 q, %2, 0;\n @q ld.global.nc.u8 %0, [%1+%3];\n}" : "+r"(v) : "l"(base), "r"((uint32_t)p), "n"(OFF));
    return v;
}
template <int B, int J = 0>
struct WarpCopyRound {
    static __device__ __forceinline__ void load(uint32_t *v, const uint8_t *s, uint32_t kb, uint32_t len)
    {
        v[J] = ldg_u8_if<32 * J>(s, kb + 32 * J < len);
        WarpCopyRound<B, J + 1>::load(v, s, kb, len);
    }
};
template <int B>
struct WarpCopyRound<B, B> { static __device__ __forceinline__ void load(uint32_t *, const uint8_t *, uint32_t, uint32_t) {} };

// copy len bytes src -> dst with the lanes of a warp, B x 32 bytes per round: all loads of a round are issued before its
// stores (a store may alias the next load as far as the compiler knows)
template <int B>
__device__ __forceinline__ void warp_copy(uint8_t *dst, const uint8_t *src, uint32_t len, uint32_t lane)
{
    for (uint32_t k0 = 0; k0 < len; k0 += 32 * B) {
        const uint32_t kb = k0 + lane;
        uint32_t v[B];
        WarpCopyRound<B>::load(v, src + kb, kb, len);
        uint8_t *d = dst + kb;
#pragma unroll
        for (int j = 0; j < B; j++) if (kb + 32 * j < len) d[32 * j] = (uint8_t)v[j];
    }
}

// one warp per record, lanes over the bytes.  Everything the record needs is loaded before the first byte is written.
__global__ void __launch_bounds__(128) k_bam_write(PackArgs a)
{
    __shared__ uint8_t nyb[256];                                            // base -> nybble, 16 = not a BAM base
    for (uint32_t i = threadIdx.x; i < 256; i += 128) nyb[i] = (uint8_t)base_nybble((uint8_t)i);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const cramgpu_bam_records &R = a.in;
    const uint8_t *const R_bases = R.bases, *const R_quals = R.quals, *const R_names = R.names;      // kernel parameters: once, not per record
    const uint32_t *const R_cigar = R.cigar;
    uint8_t *const out = a.out;
    // A CTA takes tile after tile of 32 consecutive records (grid-stride: the grid fills the machine once).  The tile's offsets and
    // scalar fields come in with one round of coalesced loads by the whole CTA, so a record costs its warp ONE trip to memory --
    // its data -- instead of two.
    __shared__ uint64_t s_off[5][33];                                       // name, CIGAR, read, tag, output offsets
    __shared__ int32_t s_i32[5][32];                                        // pos, ref id, mate ref id, mate pos, template size
    __shared__ uint16_t s_flag[32];
    __shared__ uint8_t s_mapq[32], s_rfl[32];
    const uint64_t nrec = a.in.n_reads, ntiles = (nrec + 31) / 32;
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const uint64_t r0 = tile * 32;
    const uint32_t cnt = (uint32_t)(nrec - r0 < 32 ? nrec - r0 : 32);
    __syncthreads();                                                        // the previous tile is done with the tables
    {
        const uint32_t t = threadIdx.x;
        if (t <= cnt && t < 33) {
            s_off[0][t] = __ldg(R.name_off + r0 + t); s_off[1][t] = __ldg(R.cigar_off + r0 + t); s_off[2][t] = __ldg(R.read_off + r0 + t);
            s_off[3][t] = R.tags ? __ldg(R.tag_off + r0 + t) : 0ull; s_off[4][t] = __ldg(a.rec_off + r0 + t);
        } else if (t >= 64 && t - 64 < cnt) {
            const uint32_t u = t - 64;
            s_i32[0][u] = __ldg(R.pos + r0 + u); s_i32[1][u] = R.ref_id ? __ldg(R.ref_id + r0 + u) : 0;
            s_i32[2][u] = R.mate_ref_id ? __ldg(R.mate_ref_id + r0 + u) : -1; s_i32[3][u] = R.mate_pos ? __ldg(R.mate_pos + r0 + u) : 0;
            s_i32[4][u] = R.tlen ? __ldg(R.tlen + r0 + u) : 0;
            s_flag[u] = __ldg(R.flag + r0 + u); s_mapq[u] = __ldg(R.mapq + r0 + u); s_rfl[u] = R.rflags ? __ldg(R.rflags + r0 + u) : (uint8_t)0;
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x >> 5; i < cnt; i += 4) {
    const uint64_t r = r0 + i;
    const uint64_t n0 = s_off[0][i], c0 = s_off[1][i], b0 = s_off[2][i], t0 = s_off[3][i], o0 = s_off[4][i];
    const uint32_t lname = (uint32_t)(s_off[0][i + 1] - n0), ncig = (uint32_t)(s_off[1][i + 1] - c0), lseq = (uint32_t)(s_off[2][i + 1] - b0);
    const uint32_t ltag = (uint32_t)(s_off[3][i + 1] - t0), total = (uint32_t)(s_off[4][i + 1] - o0);
    const int32_t pos = s_i32[0][i], ref_id = s_i32[1][i], mref = s_i32[2][i], mpos = s_i32[3][i], tlen = s_i32[4][i];
    const uint32_t flag = s_flag[i], mapq = s_mapq[i];
    const bool noq = (s_rfl[i] & CRAMGPU_RF_NO_QUALS) != 0;
    uint8_t *o = out + o0;
    const uint32_t nsq = (lseq + 1) / 2;
    // ---- round 0 of every field: all of a record's loads are in flight before its first byte is written (a record then costs
    //      two trips to memory -- offsets, data -- instead of one per field); what is longer than a round goes on below
    const uint32_t cw0 = lane < ncig ? __ldg(R_cigar + c0 + lane) : 0u;
    const uint32_t nm0 = ldg_u8_if<0>(R_names + n0 + lane, lane < lname);
    const uint8_t *bp0 = R_bases + b0 + 2 * lane;
    uint32_t h[3], l[3], q[5];
    h[0] = ldg_u8_if<0>(bp0, 2 * lane < lseq);        l[0] = ldg_u8_if<1>(bp0, 2 * lane + 1 < lseq);
    h[1] = ldg_u8_if<64>(bp0, 2 * lane + 64 < lseq);  l[1] = ldg_u8_if<65>(bp0, 2 * lane + 65 < lseq);
    h[2] = ldg_u8_if<128>(bp0, 2 * lane + 128 < lseq); l[2] = ldg_u8_if<129>(bp0, 2 * lane + 129 < lseq);
    WarpCopyRound<5>::load(q, R_quals + b0 + lane, lane, noq ? 0u : lseq);
    // CIGAR: reference length -> alignment end -> bin [SAMRecord.java:591-598,1524-1531]
    int32_t reflen = 0;
    for (uint32_t k0 = 0; k0 < ncig; k0 += 32) {
        const uint32_t k = k0 + lane;
        const uint32_t w = k0 == 0 ? cw0 : (k < ncig ? __ldg(R_cigar + c0 + k) : 0u), op = w & 15u;
        if (k < ncig) {
            if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) reflen += (int32_t)(w >> 4);
            uint8_t *qq = o + 36 + lname + 1 + 4 * k;
            qq[0] = (uint8_t)w; qq[1] = (uint8_t)(w >> 8); qq[2] = (uint8_t)(w >> 16); qq[3] = (uint8_t)(w >> 24);
        }
    }
    for (int d = 16; d; d >>= 1) reflen += __shfl_xor_sync(0xFFFFFFFFu, reflen, d);
    {
        // the 36 fixed bytes: nine little-endian words, lane k writes byte k (lanes 0-3 also bytes 32-35)
        const int32_t beg = pos - 1;
        int32_t end = (flag & 4) ? 0 : pos + reflen - 1;
        if (end <= 0) end = beg + 1;
        const uint32_t W3 = (lname + 1) | (mapq << 8) | ((uint32_t)reg2bin(beg, end) << 16), W4 = (ncig & 0xFFFFu) | (flag << 16);
        const uint32_t wi = lane >> 2;
        const uint32_t w = wi == 0 ? total - 4 : wi == 1 ? (uint32_t)ref_id : wi == 2 ? (uint32_t)beg : wi == 3 ? W3 : wi == 4 ? W4 : wi == 5 ? lseq
                           : wi == 6 ? (uint32_t)mref : (uint32_t)(mpos - 1);
        o[lane] = (uint8_t)(w >> (8 * (lane & 3)));
        if (lane < 4) o[32 + lane] = (uint8_t)((uint32_t)tlen >> (8 * lane));
        if (lane == 4) o[36 + lname] = 0;
    }
    if (lane < lname) o[36 + lane] = (uint8_t)nm0;
    if (lname > 32) warp_copy<1>(o + 36 + 32, R_names + n0 + 32, lname - 32, lane);
    uint8_t *sq = o + 36 + lname + 1 + 4 * ncig;
    bool bad = false;
    for (uint32_t k0 = 0; k0 < nsq; k0 += 96) {
        const uint32_t kb = k0 + lane;
        if (k0) {
            const uint8_t *bp = R_bases + b0 + 2 * kb;
            h[0] = ldg_u8_if<0>(bp, 2 * kb < lseq);        l[0] = ldg_u8_if<1>(bp, 2 * kb + 1 < lseq);
            h[1] = ldg_u8_if<64>(bp, 2 * kb + 64 < lseq);  l[1] = ldg_u8_if<65>(bp, 2 * kb + 65 < lseq);
            h[2] = ldg_u8_if<128>(bp, 2 * kb + 128 < lseq); l[2] = ldg_u8_if<129>(bp, 2 * kb + 129 < lseq);
        }
        uint8_t *d = sq + kb;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const uint32_t k = kb + 32 * j;
            uint32_t hi = nyb[h[j]], lo = 2 * k + 1 < lseq ? nyb[l[j]] : 0u;
            bad |= k < nsq && (hi > 15 || lo > 15);
            hi = min(hi, 15u); lo = min(lo, 15u);
            if (k < nsq) d[32 * j] = (uint8_t)((hi << 4) | lo);
        }
    }
    uint8_t *ql = sq + nsq;
    if (noq) { for (uint32_t k = lane; k < lseq; k += 32) ql[k] = 0xFF; }
    else {
        uint8_t *d = ql + lane;
#pragma unroll
        for (int j = 0; j < 5; j++) if (lane + 32 * j < lseq) d[32 * j] = (uint8_t)q[j];
        if (lseq > 160) warp_copy<5>(ql + 160, R_quals + b0 + 160, lseq - 160, lane);
    }
    if (ltag) warp_copy<4>(ql + lseq, R.tags + t0, ltag, lane);
    bad = __any_sync(0xFFFFFFFFu, bad);
    if (lane == 0) a.status[r] = bad ? CRAMGPU_ERR_ARG : CRAMGPU_OK;
    }
    }
}

}  // namespace

extern "C" int cramgpu_bam_pack_dev(cramgpu_ctx *ctx, const cramgpu_bam_records *in, uint8_t **d_bam, uint64_t *bam_len, uint64_t **d_rec_off,
                                    int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !d_bam || !bam_len || !d_rec_off) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    *d_bam = nullptr; *bam_len = 0; *d_rec_off = nullptr;
    const uint64_t n = in->n_reads;
    if (n == 0) return CRAMGPU_OK;
    if (!in->pos || !in->flag || !in->mapq || !in->name_off || !in->cigar_off || !in->read_off || !in->bases || !in->quals || !d_status ||
        (in->tags && !in->tag_off))
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    Carver cv;
    const size_t o_size = cv.take(4 * n), o_off = cv.take(8 * (n + 1)), o_tiles = cv.take(8 * (n / SCAN_TILE + 2));
    if (int rc = pool_reserve(ctx, ctx->bam_pk_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->bam_pk_scratch.p;
    PackArgs a;
    memset(&a, 0, sizeof a);
    a.in = *in; a.size = (uint32_t *)(S + o_size); a.rec_off = (const uint64_t *)(S + o_off); a.status = d_status;
    LAUNCH(ctx, st, "k_bam_size", k_bam_size<<<(uint32_t)((n + 127) / 128), 128, 0, st>>>(a));
    if (int rc = scan_sizes(ctx, a.size, n, (uint64_t *)(S + o_off), (uint64_t *)(S + o_tiles), st)) return rc;
    if (int rc = pinned_reserve(ctx, &ctx->pin_bam, &ctx->pin_bam_cap, 64)) return rc;
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pin_bam, S + o_off + 8 * n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t total = *(const uint64_t *)ctx->pin_bam;
    if (int rc = pool_reserve(ctx, ctx->bam_out, total + 16)) return rc;
    a.out = (uint8_t *)ctx->bam_out.p;
    LAUNCH(ctx, st, "k_bam_write", k_bam_write<<<(uint32_t)std::min<uint64_t>((n + 31) / 32, 148ull * 16), 128, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    *d_bam = a.out; *bam_len = total; *d_rec_off = (uint64_t *)(S + o_off);
    return CRAMGPU_OK;
}

// ---------------------------------------------------------------------------------------------- BAM records: unpack
namespace {

struct UnpackArgs {
    const uint8_t *bam;
    const uint64_t *rec_off;
    uint64_t n;
    uint32_t *sz;                  // [4][n]: name bytes, cigar elements, bases, tag bytes
    const uint64_t *name_off, *cigar_off, *read_off, *tag_off;      // write pass
    int32_t *ref_id, *pos, *mate_ref_id, *mate_pos, *tlen;
    uint16_t *flag, *bin;
    uint8_t *mapq, *rflags, *names, *bases, *quals, *tags;
    uint32_t *cigar;
    int32_t *status;
};

__device__ __forceinline__ uint32_t get_le32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
__device__ __forceinline__ uint32_t get_le16(const uint8_t *p) { return p[0] | (p[1] << 8); }

__global__ void __launch_bounds__(128) k_bam_scan(UnpackArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
    if (r >= a.n) return;
    const uint64_t o = a.rec_off[r], len = a.rec_off[r + 1] - o;
    uint32_t lname = 0, ncig = 0, lseq = 0, ltag = 0;
    int32_t st = CRAMGPU_ERR_CORRUPT;
    if (len >= 36) {
        const uint8_t *p = a.bam + o;
        const uint64_t bs = get_le32(p);
        lname = p[12]; ncig = get_le16(p + 16); lseq = get_le32(p + 20);
        const uint64_t fixed = 32ull + lname + 4ull * ncig + ((uint64_t)lseq + 1) / 2 + lseq;
        if (bs + 4 == len && lname >= 1 && fixed <= bs) { ltag = (uint32_t)(bs - fixed); lname -= 1; st = CRAMGPU_OK; }
        else { lname = ncig = lseq = 0; }
    }
    a.sz[r] = lname; a.sz[a.n + r] = ncig; a.sz[2 * a.n + r] = lseq; a.sz[3 * a.n + r] = ltag;
    a.status[r] = st;
}

__global__ void __launch_bounds__(128) k_bam_unpack(UnpackArgs a)
{
    const uint32_t lane = threadIdx.x & 31;
    // tiles of 32 records, their offsets through shared memory with one round of coalesced loads (see k_bam_write)
    __shared__ uint64_t s_off[5][33];                                       // record, name, CIGAR, read, tag offsets
    __shared__ int32_t s_st[32];
    const uint64_t ntiles = (a.n + 31) / 32;
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const uint64_t r0 = tile * 32;
    const uint32_t cnt = (uint32_t)(a.n - r0 < 32 ? a.n - r0 : 32);
    __syncthreads();
    {
        const uint32_t t = threadIdx.x;
        if (t <= cnt && t < 33) {
            s_off[0][t] = __ldg(a.rec_off + r0 + t); s_off[1][t] = __ldg(a.name_off + r0 + t); s_off[2][t] = __ldg(a.cigar_off + r0 + t);
            s_off[3][t] = __ldg(a.read_off + r0 + t); s_off[4][t] = __ldg(a.tag_off + r0 + t);
        } else if (t >= 64 && t - 64 < cnt) s_st[t - 64] = __ldg(a.status + r0 + t - 64);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x >> 5; i < cnt; i += 4) {
    const uint64_t r = r0 + i;
    const uint8_t *p = a.bam + s_off[0][i];
    const bool ok = s_st[i] == CRAMGPU_OK;
    const uint64_t n0 = s_off[1][i], c0 = s_off[2][i], b0 = s_off[3][i], t0 = s_off[4][i];
    const uint32_t lname = (uint32_t)(s_off[1][i + 1] - n0), ncig = (uint32_t)(s_off[2][i + 1] - c0);
    const uint32_t lseq = (uint32_t)(s_off[3][i + 1] - b0), ltag = (uint32_t)(s_off[4][i + 1] - t0);
    const uint8_t *nm = p + 36, *cg = nm + lname + 1, *sq = cg + 4 * ncig, *ql = sq + (lseq + 1) / 2, *tg = ql + lseq;
    if (!ok) {
        if (lane == 0) {
            a.ref_id[r] = -1; a.pos[r] = 0; a.mapq[r] = 0; a.bin[r] = 0; a.flag[r] = 4; a.mate_ref_id[r] = -1; a.mate_pos[r] = 0; a.tlen[r] = 0;
            a.rflags[r] = CRAMGPU_RF_NO_BASES | CRAMGPU_RF_NO_QUALS;
        }
        continue;
    }
    // ---- every first-round load of the record before its first store (see k_bam_write)
    const uint32_t hb = __ldg(p + lane), hb2 = lane < 4 ? __ldg(p + 32 + lane) : 0u;
    const uint32_t q0 = lseq ? __ldg(ql) : 0u;
    const uint32_t nm0 = ldg_u8_if<0>(nm + lane, lane < lname);
    const uint8_t *cq0 = cg + 4 * lane;
    const bool hasc = lane < ncig;
    const uint32_t cb0 = ldg_u8_if<0>(cq0, hasc), cb1 = ldg_u8_if<1>(cq0, hasc), cb2 = ldg_u8_if<2>(cq0, hasc), cb3 = ldg_u8_if<3>(cq0, hasc);
    uint32_t sb[5], qv[5];
    {
        const uint8_t *sp = sq + (lane >> 1), *qp = ql + lane;             // (lane + 32 j) >> 1 = (lane >> 1) + 16 j
        sb[0] = ldg_u8_if<0>(sp, lane < lseq);        qv[0] = ldg_u8_if<0>(qp, lane < lseq);
        sb[1] = ldg_u8_if<16>(sp, lane + 32 < lseq);  qv[1] = ldg_u8_if<32>(qp, lane + 32 < lseq);
        sb[2] = ldg_u8_if<32>(sp, lane + 64 < lseq);  qv[2] = ldg_u8_if<64>(qp, lane + 64 < lseq);
        sb[3] = ldg_u8_if<48>(sp, lane + 96 < lseq);  qv[3] = ldg_u8_if<96>(qp, lane + 96 < lseq);
        sb[4] = ldg_u8_if<64>(sp, lane + 128 < lseq); qv[4] = ldg_u8_if<128>(qp, lane + 128 < lseq);
    }
    {
        // the 36 fixed bytes: lane k holds byte k, lanes 0-3 also bytes 32-35; words by shuffles, and the lane that holds a
        // field's word writes the field
        uint32_t w = hb | (__shfl_down_sync(0xFFFFFFFFu, hb, 1) << 8) | (__shfl_down_sync(0xFFFFFFFFu, hb, 2) << 16) | (__shfl_down_sync(0xFFFFFFFFu, hb, 3) << 24);
        const uint32_t w8 = hb2 | (__shfl_down_sync(0xFFFFFFFFu, hb2, 1) << 8) | (__shfl_down_sync(0xFFFFFFFFu, hb2, 2) << 16) | (__shfl_down_sync(0xFFFFFFFFu, hb2, 3) << 24);
        if (lane == 4) a.ref_id[r] = (int32_t)w;
        if (lane == 8) a.pos[r] = (int32_t)w + 1;
        if (lane == 12) { a.mapq[r] = (uint8_t)(w >> 8); a.bin[r] = (uint16_t)(w >> 16); }
        if (lane == 16) a.flag[r] = (uint16_t)(w >> 16);
        if (lane == 24) a.mate_ref_id[r] = (int32_t)w;
        if (lane == 28) a.mate_pos[r] = (int32_t)w + 1;
        if (lane == 0) {
            a.tlen[r] = (int32_t)w8;
            a.rflags[r] = (uint8_t)((lseq == 0 ? CRAMGPU_RF_NO_BASES : 0u) | ((lseq == 0 || q0 == 0xFF) ? CRAMGPU_RF_NO_QUALS : 0u));
        }
    }
    if (lane < lname) a.names[n0 + lane] = (uint8_t)nm0;
    if (lname > 32) warp_copy<1>(a.names + n0 + 32, nm + 32, lname - 32, lane);
    uint32_t *oc = a.cigar + c0;
    if (hasc) oc[lane] = cb0 | (cb1 << 8) | (cb2 << 16) | (cb3 << 24);
    for (uint32_t k = 32 + lane; k < ncig; k += 32) {
        const uint8_t *q = cg + 4 * k;
        const uint32_t b0_ = __ldg(q), b1 = __ldg(q + 1), b2 = __ldg(q + 2), b3 = __ldg(q + 3);
        oc[k] = b0_ | (b1 << 8) | (b2 << 16) | (b3 << 24);
    }
    uint8_t *ob = a.bases + b0, *oq = a.quals + b0;
    for (uint32_t k0 = 0; k0 < lseq; k0 += 160) {
        const uint32_t kb = k0 + lane;
        if (k0) {
            const uint8_t *sp = sq + (kb >> 1), *qp = ql + kb;
            sb[0] = ldg_u8_if<0>(sp, kb < lseq);        qv[0] = ldg_u8_if<0>(qp, kb < lseq);
            sb[1] = ldg_u8_if<16>(sp, kb + 32 < lseq);  qv[1] = ldg_u8_if<32>(qp, kb + 32 < lseq);
            sb[2] = ldg_u8_if<32>(sp, kb + 64 < lseq);  qv[2] = ldg_u8_if<64>(qp, kb + 64 < lseq);
            sb[3] = ldg_u8_if<48>(sp, kb + 96 < lseq);  qv[3] = ldg_u8_if<96>(qp, kb + 96 < lseq);
            sb[4] = ldg_u8_if<64>(sp, kb + 128 < lseq); qv[4] = ldg_u8_if<128>(qp, kb + 128 < lseq);
        }
        uint8_t *obp = ob + kb, *oqp = oq + kb;
        const uint32_t sh = (kb & 1) ? 0u : 4u;                             // the same for every j: 32 j is even
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const uint32_t ny = (sb[j] >> sh) & 15u;
            const unsigned long long tab = ny < 8 ? 0x565352474D43413Dull : 0x4E42444B48595754ull;       // "=ACMGRSV" / "TWYHKDBN" in registers
            if (kb + 32 * j < lseq) { obp[32 * j] = (uint8_t)(tab >> (8 * (ny & 7))); oqp[32 * j] = (uint8_t)qv[j]; }
        }
    }
    if (ltag) warp_copy<4>(a.tags + t0, tg, ltag, lane);
    }
    }
}

}  // namespace

extern "C" int cramgpu_bam_unpack_dev(cramgpu_ctx *ctx, const uint8_t *d_bam, const uint64_t *d_rec_off, uint64_t n_reads,
                                      cramgpu_bam_records *out, int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    memset(out, 0, sizeof *out);
    const uint64_t n = n_reads;
    if (n == 0) return CRAMGPU_OK;
    if (!d_bam || !d_rec_off || !d_status) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    Carver cv;
    const size_t o_sz = cv.take(16 * n), o_tiles = cv.take(8 * (n / SCAN_TILE + 2));
    size_t o_off[4];
    for (int k = 0; k < 4; k++) o_off[k] = cv.take(8 * (n + 1));
    const size_t o_i32 = cv.take(20 * n), o_u16 = cv.take(4 * n), o_u8 = cv.take(2 * n);
    if (int rc = pool_reserve(ctx, ctx->bam_unp_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->bam_unp_scratch.p;
    UnpackArgs a;
    memset(&a, 0, sizeof a);
    a.bam = d_bam; a.rec_off = d_rec_off; a.n = n; a.sz = (uint32_t *)(S + o_sz); a.status = d_status;
    LAUNCH(ctx, st, "k_bam_scan", k_bam_scan<<<(uint32_t)((n + 127) / 128), 128, 0, st>>>(a));
    for (int k = 0; k < 4; k++)
        if (int rc = scan_sizes(ctx, a.sz + (size_t)k * n, n, (uint64_t *)(S + o_off[k]), (uint64_t *)(S + o_tiles), st)) return rc;
    if (int rc = pinned_reserve(ctx, &ctx->pin_bam, &ctx->pin_bam_cap, 64)) return rc;
    uint64_t *tot = (uint64_t *)ctx->pin_bam;
    for (int k = 0; k < 4; k++) CUDA_TRY(ctx, cudaMemcpyAsync(tot + k, S + o_off[k] + 8 * n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t t_name = tot[0], t_cig = tot[1], t_seq = tot[2], t_tag = tot[3];
    Carver co;
    const size_t q_names = co.take(t_name + 16), q_cig = co.take(4 * t_cig + 16), q_bases = co.take(t_seq + 16), q_quals = co.take(t_seq + 16),
                 q_tags = co.take(t_tag + 16);
    if (int rc = pool_reserve(ctx, ctx->bam_unp, co.off)) return rc;
    uint8_t *O = (uint8_t *)ctx->bam_unp.p;
    a.name_off = (const uint64_t *)(S + o_off[0]); a.cigar_off = (const uint64_t *)(S + o_off[1]);
    a.read_off = (const uint64_t *)(S + o_off[2]); a.tag_off = (const uint64_t *)(S + o_off[3]);
    int32_t *i32 = (int32_t *)(S + o_i32);
    a.ref_id = i32; a.pos = i32 + n; a.mate_ref_id = i32 + 2 * n; a.mate_pos = i32 + 3 * n; a.tlen = i32 + 4 * n;
    a.flag = (uint16_t *)(S + o_u16); a.bin = a.flag + n; a.mapq = S + o_u8; a.rflags = S + o_u8 + n;
    a.names = O + q_names; a.cigar = (uint32_t *)(O + q_cig); a.bases = O + q_bases; a.quals = O + q_quals; a.tags = O + q_tags;
    LAUNCH(ctx, st, "k_bam_unpack", k_bam_unpack<<<(uint32_t)std::min<uint64_t>((n + 31) / 32, 148ull * 16), 128, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    out->n_reads = n; out->ref_id = a.ref_id; out->pos = a.pos; out->mate_ref_id = a.mate_ref_id; out->mate_pos = a.mate_pos; out->tlen = a.tlen;
    out->flag = a.flag; out->bin = a.bin; out->mapq = a.mapq; out->rflags = a.rflags; out->names = a.names; out->name_off = a.name_off;
    out->cigar = a.cigar; out->cigar_off = a.cigar_off; out->bases = a.bases; out->quals = a.quals; out->read_off = a.read_off;
    out->tags = a.tags; out->tag_off = a.tag_off; out->names_len = t_name; out->cigar_len = t_cig; out->bases_len = t_seq; out->tags_len = t_tag;
    return CRAMGPU_OK;
}
