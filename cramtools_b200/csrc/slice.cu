// slice.cu -- slice header fields and content digests of a batch of slices.
//
// What ContainerFactory.buildSlice [htsjdk ContainerFactory.java:138-176] derives from the records of a slice: record count,
// base count, reference id (or MULTI_REFERENCE), alignment start / span, and the BD / SD digests: CRC32 of every record's
// base array / score array, summed as ints (cram/digest/Crc32Hasher, IntegerSumCombine).  The four SHA tags of
// ContentDigests.ALL are constants in htsjdk 2.4.1 (include/cramgpu.h says why).
//
// One thread per record.  CRC32 one byte per table lookup, the table replicated per lane in shared memory (32 KB: no bank
// conflicts); a record's bytes are read as aligned 16-byte loads, so records may start at any byte.  The per-slice sums go
// through a warp reduction and one atomic per warp and field.  2 bytes in per base, nothing out; bound by the shared-memory
// lookups (one per byte and lane), not by HBM.
#include "internal.cuh"
#include <string.h>

namespace {

struct SliceAcc {                      // zero-initialised
    unsigned long long bases;
    uint32_t bd, sd;
    int32_t inv_min_start;             // max of (INT32_MAX - start) over the placed records: 0 = none
    int32_t max_end;
    uint32_t multi;                    // a record's reference differs from the first record's
    uint32_t pad;
};

struct SliceArgs {
    uint64_t n_reads; uint32_t n_containers;
    const uint64_t *cont_off;          // device copy
    const uint8_t *bases, *scores, *rflags, *f_op, *force_qs, *streams;
    const uint64_t *read_off, *feat_off, *read_qs_off;
    const int32_t *ref_id, *pos, *f_pos, *f_val, *f_aux;
    const uint16_t *flag;
    SliceAcc *acc;
    cramgpu_slice_header *out;
    const uint8_t *ref;                // optional: slice reference MD5
    const uint64_t *ref_off, *ref_len; // device copies
};

// CRC32 one byte at a time, T = the 256-entry table replicated 32 times: entry i of lane l at T[32 * i + l], so that the 32
// lanes of a warp, each looking up its own index, always hit 32 different banks (a table shared by the lanes would serialise
// 3-4 ways on random indices, and the lookups are the whole cost of this kernel).
__device__ __forceinline__ uint32_t crc_step(uint32_t crc, const uint32_t *Tl) { return Tl[(crc & 0xFFu) << 5] ^ (crc >> 8); }
__device__ __forceinline__ uint32_t crc_byte(uint32_t crc, uint32_t b, const uint32_t *Tl) { return crc_step(crc ^ b, Tl); }

// java.util.zip.CRC32 of len bytes at p (any alignment).  The bytes come in as aligned 16-byte loads; only the aligned chunks
// that hold bytes of the span are read, and of the first and the last chunk only the words / bytes inside it count.
// (32-byte pieces were slower: more registers, fewer records in flight; the kernel lives on occupancy.)
__device__ __forceinline__ uint32_t crc32_span(const uint8_t *p, uint32_t len, const uint32_t *Tl)
{
    if (!len) return 0;
    const uint32_t mis = (uint32_t)((uintptr_t)p & 15);
    const uint4 *q = (const uint4 *)((uintptr_t)p - mis);
    const uint32_t end = mis + len;                       // span = [mis, end) in bytes from q
    uint32_t crc = ~0u;
    for (uint32_t base = 0; base < end; base += 16, q++) {
        const uint4 v = __ldg(q);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (uint32_t k = 0; k < 4; k++) {
            const uint32_t wb = base + 4 * k;             // first byte of this word
            if (wb >= mis && wb + 4 <= end) {             // the common case: the whole word
                crc ^= w[k];
                crc = crc_step(crc, Tl); crc = crc_step(crc, Tl); crc = crc_step(crc, Tl); crc = crc_step(crc, Tl);
            } else if (wb + 4 > mis && wb < end) {        // first / last word of the span: bytes [lo, hi) of it
                const uint32_t lo = wb < mis ? mis - wb : 0u, hi = min(4u, end - wb);
                uint32_t x = w[k] >> (8 * lo);
                for (uint32_t b = lo; b < hi; b++) { crc = crc_byte(crc, x & 0xFFu, Tl); x >>= 8; }
            }
        }
    }
    return ~crc;
}

// Two spans of one length whose addresses agree modulo 16 (a record's bases and scores): the two dependent chains of table
// lookups interleave, which is worth as much as twice the occupancy.  -> (crc of p1, crc of p2)
__device__ __forceinline__ uint2 crc32_span2(const uint8_t *p1, const uint8_t *p2, uint32_t len, const uint32_t *Tl)
{
    if (!len) return make_uint2(0u, 0u);
    const uint32_t mis = (uint32_t)((uintptr_t)p1 & 15);
    const uint4 *q1 = (const uint4 *)((uintptr_t)p1 - mis), *q2 = (const uint4 *)((uintptr_t)p2 - mis);
    const uint32_t end = mis + len;
    uint32_t c1 = ~0u, c2 = ~0u;
    for (uint32_t base = 0; base < end; base += 16, q1++, q2++) {
        const uint4 v1 = __ldg(q1), v2 = __ldg(q2);
        const uint32_t w1[4] = {v1.x, v1.y, v1.z, v1.w}, w2[4] = {v2.x, v2.y, v2.z, v2.w};
#pragma unroll
        for (uint32_t k = 0; k < 4; k++) {
            const uint32_t wb = base + 4 * k;
            if (wb >= mis && wb + 4 <= end) {
                c1 ^= w1[k]; c2 ^= w2[k];
#pragma unroll
                for (int t = 0; t < 4; t++) { c1 = crc_step(c1, Tl); c2 = crc_step(c2, Tl); }
            } else if (wb + 4 > mis && wb < end) {
                const uint32_t lo = wb < mis ? mis - wb : 0u, hi = min(4u, end - wb);
                uint32_t x1 = w1[k] >> (8 * lo), x2 = w2[k] >> (8 * lo);
                for (uint32_t b = lo; b < hi; b++) { c1 = crc_byte(c1, x1 & 0xFFu, Tl); c2 = crc_byte(c2, x2 & 0xFFu, Tl); x1 >>= 8; x2 >>= 8; }
            }
        }
    }
    return make_uint2(~c1, ~c2);
}

__device__ __forceinline__ uint32_t container_of(const uint64_t *cont_off, uint32_t nc, uint64_t r)
{
    uint32_t lo = 0, hi = nc;                       // largest c with cont_off[c] <= r
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cont_off[mid] <= r) lo = mid; else hi = mid; }
    return lo;
}

#define SLICE_THREADS 512
__global__ void __launch_bounds__(SLICE_THREADS) k_slice_digest(SliceArgs a)
{
    __shared__ uint32_t Trep[256 * 32];
    for (uint32_t i = threadIdx.x; i < 256 * 32; i += SLICE_THREADS) {
        uint32_t c = i >> 5;
#pragma unroll
        for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        Trep[i] = c;
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31;
    const uint32_t *T = Trep + lane;                      // this lane's copy of the table
    uint64_t r; uint32_t c; bool live;
    if (gridDim.y > 1 || a.n_containers == 1) {
        c = blockIdx.y;
        r = a.cont_off[c] + (uint64_t)blockIdx.x * SLICE_THREADS + threadIdx.x;
        live = r < a.cont_off[c + 1];
    } else {
        r = (uint64_t)blockIdx.x * SLICE_THREADS + threadIdx.x;
        live = r < a.n_reads;
        c = live ? container_of(a.cont_off, a.n_containers, r) : 0xFFFFFFFFu;
    }
    uint32_t bd = 0, sd = 0, rlen = 0, multi = 0;
    int32_t inv_start = 0, end = 0;
    if (live) {
        const uint64_t lo = a.read_off[r];
        rlen = (uint32_t)(a.read_off[r + 1] - lo);
        const bool no_quals = a.rflags && (a.rflags[r] & CRAMGPU_RF_NO_QUALS);
        const uint64_t f0 = a.feat_off[r], f1 = a.feat_off[r + 1];
        const bool both = !no_quals && a.scores && ((((uintptr_t)a.bases ^ (uintptr_t)a.scores) & 15) == 0);
        if (both) { const uint2 c2 = crc32_span2(a.bases + lo, a.scores + lo, rlen, T); bd = c2.x; sd = c2.y; }
        else bd = crc32_span(a.bases + lo, rlen, T);
        if (both) { }
        else if (no_quals) sd = 0;
        else if (a.scores) sd = crc32_span(a.scores + lo, rlen, T);
        else {
            const uint8_t fq = a.force_qs[r];
            if (!(fq & 2)) sd = 0;                                            // no score array at all
            else if ((fq & 1) && a.read_qs_off[r] != ~0ull) sd = crc32_span(a.streams + a.read_qs_off[r], rlen, T);
            else {
                // -1 everywhere except at the record's Q features (they are in position order)
                uint32_t crc = ~0u;
                uint64_t f = f0;
                int32_t next_q = -1, q_val = 0;
                auto advance = [&]() {
                    next_q = -1;
                    while (f < f1) { const uint64_t g = f++; if (a.f_op[g] == 'Q') { next_q = a.f_pos[g] - 1; q_val = a.f_val[g]; break; } }
                };
                advance();
                for (uint32_t i = 0; i < rlen; i++) {
                    uint32_t b = 0xFF;
                    if ((int32_t)i == next_q) { b = (uint32_t)q_val & 0xFF; advance(); }
                    crc = crc_byte(crc, b, T);
                }
                sd = rlen ? ~crc : 0;
            }
        }
        // placement [ContainerFactory.java:156-162]; getAlignmentEnd [CramCompressionRecord.java:160-194]
        const int32_t first = a.ref_id ? a.ref_id[a.cont_off[c]] : 0, mine = a.ref_id ? a.ref_id[r] : 0;
        multi = mine != first;
        const int32_t pos = a.pos[r];
        if (pos != 0) {
            if (!(a.flag[r] & 4)) {
                int32_t span = (int32_t)rlen;
                for (uint64_t f = f0; f < f1; f++) {
                    const uint8_t op = a.f_op[f];
                    if (op == 'i') span -= 1;
                    else if (op == 'S') span -= a.f_aux[f];
                    else if (op == 'D') span += a.f_val[f];
                }
                end = pos + span - 1;
            }
            inv_start = pos > 0 ? INT32_MAX - pos : INT32_MAX;          // a negative start cannot come out of a SAMRecord
        }
    }
    // per-slice sums: one atomic per warp when the warp is inside one slice
    const uint32_t c0 = __shfl_sync(0xFFFFFFFFu, c, 0);
    const bool same = __all_sync(0xFFFFFFFFu, !live || c == c0);
    if (same) {
        uint32_t len_sum = rlen;
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            bd += __shfl_xor_sync(0xFFFFFFFFu, bd, d); sd += __shfl_xor_sync(0xFFFFFFFFu, sd, d);
            len_sum += __shfl_xor_sync(0xFFFFFFFFu, len_sum, d);
            inv_start = max(inv_start, __shfl_xor_sync(0xFFFFFFFFu, inv_start, d));
            end = max(end, __shfl_xor_sync(0xFFFFFFFFu, end, d));
            multi |= __shfl_xor_sync(0xFFFFFFFFu, multi, d);
        }
        if (lane == 0 && c0 != 0xFFFFFFFFu) {
            SliceAcc *s = a.acc + c0;
            atomicAdd(&s->bd, bd); atomicAdd(&s->sd, sd); atomicAdd(&s->bases, (unsigned long long)len_sum);
            if (inv_start) atomicMax(&s->inv_min_start, inv_start);
            if (end > 0) atomicMax(&s->max_end, end);
            if (multi) atomicOr(&s->multi, 1u);
        }
    } else if (live) {
        SliceAcc *s = a.acc + c;
        atomicAdd(&s->bd, bd); atomicAdd(&s->sd, sd); atomicAdd(&s->bases, (unsigned long long)rlen);
        if (inv_start) atomicMax(&s->inv_min_start, inv_start);
        if (end > 0) atomicMax(&s->max_end, end);
        if (multi) atomicOr(&s->multi, 1u);
    }
}

__global__ void k_slice_finish(SliceArgs a)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= a.n_containers) return;
    const SliceAcc s = a.acc[c];
    const uint64_t n = a.cont_off[c + 1] - a.cont_off[c];
    cramgpu_slice_header h;
    h.n_records = (uint32_t)n; h.bases = s.bases; h.bd = s.bd; h.sd = s.sd;
    h.sequence_id = n == 0 ? -1 : s.multi ? -2 : (a.ref_id ? a.ref_id[a.cont_off[c]] : 0);
    if (s.multi || s.inv_min_start == 0) { h.alignment_start = -1; h.alignment_span = 0; }        // [:169-172]
    else {
        h.alignment_start = INT32_MAX - s.inv_min_start;
        h.alignment_span = s.max_end - h.alignment_start + 1;                                       // [:174-175]
        if (h.alignment_span < 0) h.alignment_span = 0;                                             // CramSerilization.java:184-186
    }
    for (int k = 0; k < 16; k++) h.ref_md5[k] = 0;
    a.out[c] = h;
}

// MD5 (RFC 1321) of the slice's reference span, one thread per slice: Slice.setRefMD5 [Slice.java:145-159] ->
// SequenceUtil.calculateMD5 [SequenceUtil.java:962-966].  A slice's span is tens of kilobases: a few hundred 64-byte blocks, a
// strictly sequential hash -- all slices of the batch run side by side.
__constant__ uint32_t MD5_K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af, 0xffff5bb1,
    0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453,
    0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a, 0xfffa3942,
    0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70, 0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05,
    0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d,
    0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
__constant__ uint8_t MD5_S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
                                  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};

__device__ __forceinline__ void md5_block(uint32_t st[4], const uint32_t w[16])
{
    uint32_t A = st[0], B = st[1], Cc = st[2], D = st[3];
#pragma unroll
    for (int i = 0; i < 64; i++) {
        uint32_t F; int g;
        if (i < 16) { F = (B & Cc) | (~B & D); g = i; }
        else if (i < 32) { F = (D & B) | (~D & Cc); g = (5 * i + 1) & 15; }
        else if (i < 48) { F = B ^ Cc ^ D; g = (3 * i + 5) & 15; }
        else { F = Cc ^ (B | ~D); g = (7 * i) & 15; }
        F = F + A + MD5_K[i] + w[g];
        A = D; D = Cc; Cc = B;
        B = B + ((F << MD5_S[i]) | (F >> (32 - MD5_S[i])));
    }
    st[0] += A; st[1] += B; st[2] += Cc; st[3] += D;
}

__global__ void __launch_bounds__(64) k_slice_md5(SliceArgs a)
{
    const uint32_t c = blockIdx.x * 64 + threadIdx.x;
    if (c >= a.n_containers) return;
    const cramgpu_slice_header h = a.out[c];
    if (h.sequence_id < 0 || h.alignment_start < 1) return;               // never set / "Empty slice ref md5" [:147-151]: zeros
    const int64_t len_ref = (int64_t)a.ref_len[c];
    int64_t span = min((int64_t)h.alignment_span, len_ref - h.alignment_start + 1);           // [:154]
    if (span < 0) span = 0;
    const uint8_t *p = a.ref + a.ref_off[c] + (h.alignment_start - 1);
    uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    uint32_t w[16];
    int64_t done = 0;
    for (; done + 64 <= span; done += 64) {
#pragma unroll
        for (int k = 0; k < 16; k++)
            w[k] = (uint32_t)p[done + 4 * k] | ((uint32_t)p[done + 4 * k + 1] << 8) | ((uint32_t)p[done + 4 * k + 2] << 16) | ((uint32_t)p[done + 4 * k + 3] << 24);
        md5_block(st, w);
    }
    // the tail: remaining bytes, 0x80, zeros, the bit length (one or two blocks)
    uint8_t tail[128];
    const int rem = (int)(span - done);
    for (int k = 0; k < 128; k++) tail[k] = 0;
    for (int k = 0; k < rem; k++) tail[k] = p[done + k];
    tail[rem] = 0x80;
    const int tl = rem < 56 ? 64 : 128;
    const uint64_t bits = (uint64_t)span * 8;
    for (int k = 0; k < 8; k++) tail[tl - 8 + k] = (uint8_t)(bits >> (8 * k));
    for (int b = 0; b < tl; b += 64) {
        for (int k = 0; k < 16; k++)
            w[k] = (uint32_t)tail[b + 4 * k] | ((uint32_t)tail[b + 4 * k + 1] << 8) | ((uint32_t)tail[b + 4 * k + 2] << 16) | ((uint32_t)tail[b + 4 * k + 3] << 24);
        md5_block(st, w);
    }
    for (int k = 0; k < 16; k++) a.out[c].ref_md5[k] = (uint8_t)(st[k >> 2] >> (8 * (k & 3)));
}

}  // namespace

static int slice_impl(cramgpu_ctx *ctx, const cramgpu_slice_in *in, cramgpu_slice_header *d_out, cudaStream_t st)
{
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (nc == 0) return CRAMGPU_OK;
    if (!in->cont_off || !d_out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->cont_off[0] != 0 || in->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    uint64_t max_reads = 0;
    for (uint32_t c = 0; c < nc; c++) {
        if (in->cont_off[c] > in->cont_off[c + 1]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off not monotonic");
        max_reads = std::max(max_reads, in->cont_off[c + 1] - in->cont_off[c]);
    }
    if (n && (!in->bases || !in->read_off || !in->pos || !in->flag || !in->feat_off))
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (n && !in->scores && (!in->force_qs || !in->read_qs_off || !in->streams))
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "scores == NULL needs force_qs, read_qs_off and streams");
    Carver cv;
    const bool md5 = in->ref != nullptr;
    if (md5 && (!in->ref_off || !in->ref_len)) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "ref needs ref_off and ref_len");
    if (md5)
        for (uint32_t c = 0; c < nc; c++)
            if (in->ref_off[c] + in->ref_len[c] > in->ref_total) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "reference sequence of container %u outside ref", c);
    const size_t up_bytes = 8 * (size_t)(nc + 1) + 16 * (size_t)nc;
    const size_t o_cont = cv.take(up_bytes), o_acc = cv.take(sizeof(SliceAcc) * (size_t)nc);
    if (int rc = pool_reserve(ctx, ctx->slice_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->slice_scratch.p;
    if (int rc = pinned_reserve(ctx, &ctx->pin_slice, &ctx->pin_slice_cap, rup(up_bytes, 64) + sizeof(cramgpu_slice_header) * (size_t)nc + 64)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_slice));              // the previous call's upload has left the pinned buffer
    memcpy(ctx->pin_slice, in->cont_off, 8 * (size_t)(nc + 1));
    if (md5) {
        memcpy((uint8_t *)ctx->pin_slice + 8 * (size_t)(nc + 1), in->ref_off, 8 * (size_t)nc);
        memcpy((uint8_t *)ctx->pin_slice + 8 * (size_t)(nc + 1) + 8 * (size_t)nc, in->ref_len, 8 * (size_t)nc);
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_cont, ctx->pin_slice, md5 ? up_bytes : 8 * (size_t)(nc + 1), cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_slice, st));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_acc, 0, sizeof(SliceAcc) * (size_t)nc, st));
    SliceArgs a;
    memset(&a, 0, sizeof a);
    a.n_reads = n; a.n_containers = nc; a.cont_off = (const uint64_t *)(S + o_cont);
    a.bases = in->bases; a.scores = in->scores; a.rflags = in->rflags; a.f_op = in->f_op; a.force_qs = in->force_qs; a.streams = in->streams;
    a.read_off = in->read_off; a.feat_off = in->feat_off; a.read_qs_off = in->read_qs_off;
    a.ref_id = in->ref_id; a.pos = in->pos; a.f_pos = in->f_pos; a.f_val = in->f_val; a.f_aux = in->f_aux; a.flag = in->flag;
    a.acc = (SliceAcc *)(S + o_acc); a.out = d_out;
    a.ref = in->ref; a.ref_off = (const uint64_t *)(S + o_cont + 8 * (size_t)(nc + 1)); a.ref_len = a.ref_off + nc;
    if (n) {
        const dim3 grid = nc <= 65535 ? dim3((uint32_t)((max_reads + SLICE_THREADS - 1) / SLICE_THREADS), nc)
                                      : dim3((uint32_t)((n + SLICE_THREADS - 1) / SLICE_THREADS), 1);
        LAUNCH(ctx, st, "k_slice_digest", k_slice_digest<<<grid, SLICE_THREADS, 0, st>>>(a));
    }
    LAUNCH(ctx, st, "k_slice_finish", k_slice_finish<<<(nc + 127) / 128, 128, 0, st>>>(a));
    if (md5) LAUNCH(ctx, st, "k_slice_md5", k_slice_md5<<<(nc + 63) / 64, 64, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}

extern "C" int cramgpu_slice_headers_dev(cramgpu_ctx *ctx, const cramgpu_slice_in *in, cramgpu_slice_header *d_out, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return slice_impl(ctx, in, d_out, stream ? (cudaStream_t)stream : ctx->stream);
}

extern "C" int cramgpu_slice_headers(cramgpu_ctx *ctx, const cramgpu_slice_in *in, cramgpu_slice_header *out)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (nc == 0) return CRAMGPU_OK;
    if (n && (!in->read_off || !in->feat_off)) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    MemMap mm;
    // sizes of the variable-length arrays: the last entries of the offset tables (host memory, or fetched from the device)
    auto last_u64 = [&](const uint64_t *p, uint64_t idx, uint64_t *v) -> int {
        if (mm.kind(p) == MEM_DEVICE) {
            CUDA_TRY(ctx, cudaMemcpyAsync(v, p + idx, 8, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(ctx, cudaStreamSynchronize(st));
        } else *v = p[idx];
        return CRAMGPU_OK;
    };
    uint64_t nb = 0, nf = 0;
    if (n) {
        if (int rc = last_u64(in->read_off, n, &nb)) return rc;
        if (int rc = last_u64(in->feat_off, n, &nf)) return rc;
    }
    struct Item { const void *src; size_t bytes; size_t off; };
    std::vector<Item> items;
    Carver cv;
    auto add = [&](const void *p, size_t bytes) -> size_t {
        if (!p || mm.kind(p) == MEM_DEVICE) return (size_t)-1;
        size_t o = cv.take(bytes + 16); items.push_back({p, bytes, o}); return o;
    };
    if (!in->scores && in->streams && mm.kind(in->streams) != MEM_DEVICE)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "scores == NULL: streams must be the device result of a feature call");
    const size_t i_b = add(in->bases, nb), i_s = add(in->scores, nb), i_ro = add(in->read_off, 8 * (n + 1)), i_rf = add(in->rflags, n);
    const size_t i_id = add(in->ref_id, 4 * n), i_pos = add(in->pos, 4 * n), i_fl = add(in->flag, 2 * n);
    const size_t i_op = add(in->f_op, nf), i_fp = add(in->f_pos, 4 * nf), i_fv = add(in->f_val, 4 * nf), i_fa = add(in->f_aux, 4 * nf);
    const size_t i_fo = add(in->feat_off, 8 * (n + 1)), i_fq = add(in->force_qs, n), i_qo = add(in->read_qs_off, 8 * n);
    const size_t i_ref = add(in->ref, in->ref ? in->ref_total : 0);
    const size_t o_out = cv.take(sizeof(cramgpu_slice_header) * (size_t)nc);
    if (int rc = pool_reserve(ctx, ctx->slice_in, cv.off + 16)) return rc;
    uint8_t *D = (uint8_t *)ctx->slice_in.p;
    for (const Item &it : items)
        if (it.bytes) CUDA_TRY(ctx, memcpy_pieces(D + it.off, it.src, it.bytes, cudaMemcpyHostToDevice, st));
    auto at = [&](size_t o, const void *orig) -> const void * { return o == (size_t)-1 ? orig : (const void *)(D + o); };
    cramgpu_slice_in d = *in;
    d.bases = (const uint8_t *)at(i_b, in->bases); d.scores = (const uint8_t *)at(i_s, in->scores);
    d.read_off = (const uint64_t *)at(i_ro, in->read_off); d.rflags = (const uint8_t *)at(i_rf, in->rflags);
    d.ref_id = (const int32_t *)at(i_id, in->ref_id); d.pos = (const int32_t *)at(i_pos, in->pos); d.flag = (const uint16_t *)at(i_fl, in->flag);
    d.f_op = (const uint8_t *)at(i_op, in->f_op); d.f_pos = (const int32_t *)at(i_fp, in->f_pos); d.f_val = (const int32_t *)at(i_fv, in->f_val);
    d.f_aux = (const int32_t *)at(i_fa, in->f_aux); d.feat_off = (const uint64_t *)at(i_fo, in->feat_off);
    d.force_qs = (const uint8_t *)at(i_fq, in->force_qs); d.read_qs_off = (const uint64_t *)at(i_qo, in->read_qs_off);
    d.ref = (const uint8_t *)at(i_ref, in->ref);
    cramgpu_slice_header *d_out = (cramgpu_slice_header *)(D + o_out);
    if (int rc = slice_impl(ctx, &d, d_out, st)) return rc;
    uint8_t *ph = (uint8_t *)ctx->pin_slice + rup(8 * (size_t)(nc + 1) + 16 * (size_t)nc, 64);
    CUDA_TRY(ctx, cudaMemcpyAsync(ph, d_out, sizeof(cramgpu_slice_header) * (size_t)nc, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    memcpy(out, ph, sizeof(cramgpu_slice_header) * (size_t)nc);
    return CRAMGPU_OK;
}

extern "C" void cramgpu_slice_sha_tags(uint8_t sha1[20], uint8_t sha512[64])
{
    // SHA-1 and SHA-512 of the empty message
    static const uint8_t S1[20] = {0xda, 0x39, 0xa3, 0xee, 0x5e, 0x6b, 0x4b, 0x0d, 0x32, 0x55, 0xbf, 0xef, 0x95, 0x60, 0x18, 0x90, 0xaf, 0xd8, 0x07, 0x09};
    static const uint8_t S5[64] = {0xcf, 0x83, 0xe1, 0x35, 0x7e, 0xef, 0xb8, 0xbd, 0xf1, 0x54, 0x28, 0x50, 0xd6, 0x6d, 0x80, 0x07,
                                   0xd6, 0x20, 0xe4, 0x05, 0x0b, 0x57, 0x15, 0xdc, 0x83, 0xf4, 0xa9, 0x21, 0xd3, 0x6c, 0xe9, 0xce,
                                   0x47, 0xd0, 0xd1, 0x3c, 0x5d, 0x85, 0xf2, 0xb0, 0xff, 0x83, 0x18, 0xd2, 0x87, 0x7e, 0xec, 0x2f,
                                   0x63, 0xb9, 0x31, 0xbd, 0x47, 0x41, 0x7a, 0x81, 0xa5, 0x38, 0x32, 0x7a, 0xf9, 0x27, 0xda, 0x3e};
    if (sha1) memcpy(sha1, S1, 20);
    if (sha512) memcpy(sha512, S5, 64);
}
