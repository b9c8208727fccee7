// tracks.cu -- the coverage and mismatch tracks behind the lossy selectors X and P.
//
// Reference: CramSerilization.convert keeps one ReferenceTracks object per sequence (ref/ReferenceTracks.java) and, for
// every container in file order, (1) takes the container's alignment span (getSpans, CramSerilization.java:60-86),
// (2) tracks.ensureRange(start, span) [:125], (3) updateTracks [:225-257]: +1 coverage for every reference position an
// alignment covers, +1 mismatch where a base of an M / = / X operator differs from the reference (with the loop indices
// as they are written there), and then (4) looks at the records one by one: addQS asks coverageAt(pos) < n (selector X,
// lossy/QualityScorePreservation.java:447-465) and mismatchesAt(alignment start + i) > n (selector P, :466-470).
// So container c sees the counts of containers 0..c of this call and of every earlier call, and none of c+1...
//
// Here a call takes a batch of containers at once.  Every container gets its OWN count arrays over its span (coverage as
// a difference array that one CTA per container scans; mismatches by atomics), and what container c sees at a position is
//     carry(pos) + sum over j = c, c-1, ... of own_j(pos)        as long as max query end of containers 0..j >= pos
// (a prefix maximum, so the walk stops after one or two containers on sorted input), evaluated once per position a
// container's records can ask about (k_trk_bits) and spread to the bases by a CIGAR walk per record (k_trk_masks); at the end
// of the call the own arrays are folded into the carry arrays, which live as long as the tracks object.  The window of ReferenceTracks (1 000 000
// positions that slide forward) only shows as cut-offs: a position at or beyond window start + length is neither counted
// nor read (0), and baseAt there is 'N'; the window start per container is a scan over the containers' spans that the
// host does between two kernels (the only synchronisation).  What is NOT modelled: a window that moves while the records
// of a container are looked at (a container whose reads reach beyond the window although the window was moved to its
// start, i.e. a span above 1 Mbp) -- such a container, and with it the rest of the call, is refused with the message of
// the case; the tracks object then stays unusable, as the reference's conversion ends with an exception there too.
#include "internal.cuh"
#include <string.h>

struct cramgpu_tracks {
    int device = 0;
    uint64_t ref_len = 0;
    int64_t window = 0;                 // ReferenceTracks.bases.length = min(window size, reference length)
    int64_t position = 1;               // 1-based start of the window
    bool poisoned = false;
    uint8_t *d_ref = nullptr;           // raw reference bases
    int32_t *d_cov = nullptr, *d_mm = nullptr;       // carry: counts of all containers of earlier calls, index = position (1-based)
};

namespace {

#define OP_M 0
#define OP_I 1
#define OP_D 2
#define OP_N 3
#define OP_S 4
#define OP_EQ 7
#define OP_X 8
#define SAM_UNMAPPED 4
__device__ __forceinline__ bool t_consumes_read(uint32_t op) { return op == OP_M || op == OP_I || op == OP_S || op == OP_EQ || op == OP_X; }
__device__ __forceinline__ bool t_consumes_ref(uint32_t op) { return op == OP_M || op == OP_D || op == OP_N || op == OP_EQ || op == OP_X; }

// per container, filled by k_trk_ends (device) and read by the host
struct TrkSum {
    int32_t min_start;      // smallest alignment start of the placed records (INT_MAX: none)
    int32_t max_end;        // largest SAMRecord.getAlignmentEnd() (0 for a record with the unmapped flag); INT_MIN: none
    int32_t max_cov;        // largest reference position a CIGAR reaches (updateTracks does not look at the unmapped flag)
    int32_t max_query;      // largest position + 1 any record can ask the tracks about
    int32_t max_start;      // largest alignment start
    uint32_t flags;         // 1: a record without bases whose CIGAR has M / = / X (updateTracks indexes an empty array)
};
// per container, written by the host
struct TrkCont {
    int64_t position;       // window start while the container is converted
    int32_t start;          // first position of the own arrays
    int32_t own_len;        // positions in them (0: nothing counted)
    int64_t own_off;        // their place in the own buffers (own_len + 1 slots: the difference array needs one more)
    int32_t reach;          // prefix maximum of the last position the own arrays of containers 0..c hold
    int32_t ok;             // 0: refused, the records get no ext bytes
    int64_t q_off;          // the container's selector bytes: one per position start .. start + q_len - 1 (what its records can ask
    int32_t q_len;          //   about inside the window); bit 1 + e = selector e holds there, as this container sees the tracks
    int32_t pad;
};

struct TrkArgs {
    uint64_t n_reads; uint32_t n_containers;
    const uint64_t *cont_off;
    const uint8_t *bases; const uint64_t *read_off;
    const uint32_t *cigar; const uint64_t *cigar_off;
    const int32_t *pos; const uint16_t *flag; const uint8_t *rflags;
    TrkSum *sum; const TrkCont *cont;
    const uint8_t *ref; int64_t ref_len, window;
    int32_t *own_cov, *own_mm;
    const int32_t *carry_cov, *carry_mm;
    uint8_t *ext;
    uint8_t *qbits;                    // selector bytes per (container, position), see TrkCont::q_off
    int n_sel; int32_t sel_kind[7], sel_param[7];        // the X (5) / P (6) selectors in policy order: bit 1 + e of the ext byte
};

__device__ __forceinline__ uint32_t trk_container_of(const uint64_t *cont_off, uint32_t nc, uint64_t r)
{
    uint32_t lo = 0, hi = nc;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cont_off[mid] <= r) lo = mid; else hi = mid; }
    return lo;
}

// one thread per record: alignment end, furthest query, per-container extremes
__global__ void __launch_bounds__(256) k_trk_ends(TrkArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= a.n_reads) return;
    const int32_t pos = a.pos[r];
    if (pos == 0) return;                                                   // unplaced: not in getSpans, not in updateTracks
    const uint32_t c = trk_container_of(a.cont_off, a.n_containers, r);
    int64_t ref_l = 0, all_l = 0;
    bool has_m = false;
    for (uint64_t ci = a.cigar_off[r]; ci < a.cigar_off[r + 1]; ci++) {
        const uint32_t cw = a.cigar[ci], op = cw & 15, len = cw >> 4;
        if (t_consumes_ref(op)) ref_l += len;
        if (t_consumes_ref(op) || t_consumes_read(op)) all_l += len;
        has_m |= (op == OP_M || op == OP_EQ || op == OP_X) && len;
    }
    const uint64_t rlen = a.read_off[r + 1] - a.read_off[r];
    const int32_t end = (a.flag[r] & SAM_UNMAPPED) ? 0 : (int32_t)(pos + ref_l - 1);           // SAMRecord.getAlignmentEnd
    const int64_t q = (int64_t)pos + (all_l > (int64_t)rlen ? all_l : (int64_t)rlen);             // exclusive bound of every position asked
    TrkSum *s = a.sum + c;
    const int32_t cov_end = (int32_t)(pos + ref_l - 1), qq = (int32_t)(q > 0x7FFFFFFF ? 0x7FFFFFFF : q);
    const uint32_t fl = (has_m && rlen == 0) ? 1u : 0u;
    // the records of a warp are nearly always of one container: one set of atomics per warp then (10 000 records share five words)
    const uint32_t act = __activemask();
    const uint32_t c0 = __shfl_sync(act, c, __ffs(act) - 1);
    if (__all_sync(act, c == c0)) {
        const int32_t mn = __reduce_min_sync(act, pos), mx = __reduce_max_sync(act, pos), me = __reduce_max_sync(act, end);
        const int32_t mc = __reduce_max_sync(act, cov_end), mq = __reduce_max_sync(act, qq);
        const uint32_t f = __reduce_or_sync(act, fl);
        if ((threadIdx.x & 31) == (uint32_t)(__ffs(act) - 1)) {
            atomicMin(&s->min_start, mn); atomicMax(&s->max_start, mx); atomicMax(&s->max_end, me);
            atomicMax(&s->max_cov, mc); atomicMax(&s->max_query, mq);
            if (f) atomicOr(&s->flags, f);
        }
    } else {
        atomicMin(&s->min_start, pos); atomicMax(&s->max_start, pos); atomicMax(&s->max_end, end);
        atomicMax(&s->max_cov, cov_end); atomicMax(&s->max_query, qq);
        if (fl) atomicOr(&s->flags, fl);
    }
}

// one thread per record: its counts into the container's own arrays (updateTracks, CramSerilization.java:225-257)
__global__ void __launch_bounds__(256) k_trk_own(TrkArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= a.n_reads) return;
    const int32_t pos = a.pos[r];
    if (pos == 0) return;
    const uint32_t c = trk_container_of(a.cont_off, a.n_containers, r);
    const TrkCont k = a.cont[c];
    if (!k.ok || k.own_len == 0) return;
    const int64_t limit = k.position + a.window;                            // positions at or beyond it are outside the window
    const uint8_t *rb = a.bases + a.read_off[r];
    const int64_t rlen = (int64_t)(a.read_off[r + 1] - a.read_off[r]);
    int32_t *cov = a.own_cov + k.own_off, *mm = a.own_mm + k.own_off;
    int64_t ref_pos = pos, read_pos = 0;
    for (uint64_t ci = a.cigar_off[r]; ci < a.cigar_off[r + 1]; ci++) {
        const uint32_t cw = a.cigar[ci], op = cw & 15;
        const int64_t len = cw >> 4;
        if (t_consumes_ref(op) && len) {
            // addCoverage(refPos + e) for e < len [:231-234]: a difference array, clipped to the window
            const int64_t lo = ref_pos, hi = (ref_pos + len < limit ? ref_pos + len : limit);      // [lo, hi)
            if (hi > lo) { atomicAdd(cov + (lo - k.start), 1); atomicAdd(cov + (hi - k.start), -1); }
        }
        if (op == OP_M || op == OP_X || op == OP_EQ) {
            // `for (pos = readPos; pos < length; pos++)` with read index readPos + pos and reference position refPos + pos [:239-244]
            for (int64_t p = read_pos; p < len; p++) {
                if (read_pos + p >= rlen) break;                            // (cannot happen when the CIGAR fits the read)
                const int64_t rp = ref_pos + p;
                const uint8_t ref_base = rp < limit ? a.ref[rp - 1] : (uint8_t)'N';            // baseAt [ReferenceTracks.java:141-146]
                if (rb[read_pos + p] != ref_base && rp < limit) atomicAdd(mm + (rp - k.start), 1);
            }
        }
        if (t_consumes_read(op)) read_pos += len;
        if (t_consumes_ref(op)) ref_pos += len;
    }
}

// one CTA per container: the coverage difference array becomes the coverage (inclusive prefix sum over own_len slots)
__global__ void __launch_bounds__(1024) k_trk_scan(TrkArgs a)
{
    const TrkCont k = a.cont[blockIdx.x];
    if (!k.ok || k.own_len == 0) return;
    int32_t *cov = a.own_cov + k.own_off;
    __shared__ int32_t sw[32];
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int32_t i0 = 0; i0 < k.own_len; i0 += 1024) {
        const int32_t i = i0 + (int32_t)threadIdx.x;
        const int32_t v = i < k.own_len ? cov[i] : 0;
        int32_t s = v;
        for (int d = 1; d < 32; d <<= 1) { const int32_t t = __shfl_up_sync(0xFFFFFFFFu, s, d); if (lane >= (uint32_t)d) s += t; }
        if (lane == 31) sw[w] = s;
        __syncthreads();
        int32_t base = carry, all = 0;
        for (uint32_t j = 0; j < 32; j++) { if (j < w) base += sw[j]; all += sw[j]; }
        if (i < k.own_len) cov[i] = base + s;
        __syncthreads();
        if (threadIdx.x == 0) carry += all;
        __syncthreads();
    }
}

// what container c sees at `pos` (inside its window): carry + the own arrays of containers c, c-1, ... whose reach covers pos
__device__ __forceinline__ int32_t trk_seen(const TrkArgs &a, const int32_t *carry, const int32_t *own, uint32_t c, int64_t pos)
{
    int32_t v = carry[pos];
    for (int64_t j = c; j >= 0; j--) {
        const TrkCont &kj = a.cont[j];
        if (kj.reach < pos) break;                                          // prefix maximum: nothing further back holds pos either
        if (kj.own_len && pos >= kj.start && pos < (int64_t)kj.start + kj.own_len) v += own[kj.own_off + (pos - kj.start)];
    }
    return (int32_t)(int16_t)v;                                             // the tracks are Java shorts
}

// one CTA per container: the selectors at every position its records can ask about -- coverageAt(pos) < n [:447-465],
// mismatchesAt(pos) > n [:466-470] -- evaluated once per position instead of once per base (45 reads cover a position)
__global__ void __launch_bounds__(256) k_trk_bits(TrkArgs a)
{
    const uint32_t c = blockIdx.x;
    const TrkCont k = a.cont[c];
    if (!k.ok) return;
    bool any_p = false, any_x = false;
    for (int s = 0; s < a.n_sel; s++) { any_p |= a.sel_kind[s] == 6; any_x |= a.sel_kind[s] == 5; }
    uint8_t *q = a.qbits + k.q_off;
    for (int32_t i = threadIdx.x; i < k.q_len; i += 256) {
        const int64_t pos = (int64_t)k.start + i;
        const int32_t cov = any_x ? trk_seen(a, a.carry_cov, a.own_cov, c, pos) : 0;
        const int32_t mm = any_p ? trk_seen(a, a.carry_mm, a.own_mm, c, pos) : 0;
        uint8_t e = 0;
        for (int s = 0; s < a.n_sel; s++)
            if (a.sel_kind[s] == 5 ? cov < a.sel_param[s] : mm > a.sel_param[s]) e |= (uint8_t)(2u << s);
        q[i] = e;
    }
}

// one warp per record: the ext byte of every base -- bit 0 = base of an M / = / X operator, bit 1 + e = selector e holds.
// Selector P asks at alignment start + index in the read, whatever the CIGAR says; selector X at the reference position
// of the CIGAR walk, for the bases of M / = / X only.  A position beyond the window reads as 0 counts [ReferenceTracks.java:148-161].
__global__ void __launch_bounds__(128) k_trk_masks(TrkArgs a)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t r = (uint64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (r >= a.n_reads) return;
    const uint64_t lo = a.read_off[r];
    const int64_t rlen = (int64_t)(a.read_off[r + 1] - lo);
    uint8_t *ext = a.ext + lo;
    const int32_t pos = a.pos[r];
    const uint32_t c = pos ? trk_container_of(a.cont_off, a.n_containers, r) : 0u;
    if (pos == 0 || !a.cont[c].ok) {                                        // addQS returns before the masks [:394-396] / refused container
        for (int64_t i = lane; i < rlen; i += 32) ext[i] = 0;
        return;
    }
    const TrkCont k = a.cont[c];
    const uint8_t *q = a.qbits + k.q_off;
    uint8_t pmask = 0, xmask = 0, xbeyond = 0;                              // beyond the window: coverage 0 < n, mismatches 0 > n never
    for (int s = 0; s < a.n_sel; s++) {
        if (a.sel_kind[s] == 6) pmask |= (uint8_t)(2u << s);
        else { xmask |= (uint8_t)(2u << s); if (0 < a.sel_param[s]) xbeyond |= (uint8_t)(2u << s); }
    }
    int64_t z = 0, ref_pos = pos;
    for (uint64_t ci = a.cigar_off[r]; ci < a.cigar_off[r + 1]; ci++) {
        const uint32_t cw = a.cigar[ci], op = cw & 15;
        const int64_t len = cw >> 4;
        if (t_consumes_read(op)) {
            const bool is_m = op == OP_M || op == OP_EQ || op == OP_X;
            for (int64_t i = lane; i < len && z + i < rlen; i += 32) {
                const int64_t pi = (int64_t)pos + z + i - k.start;          // selector P's position
                uint8_t e = (pi >= 0 && pi < k.q_len) ? (q[pi] & pmask) : (uint8_t)0;
                if (is_m) {
                    const int64_t xi = ref_pos + i - k.start;
                    e |= 1 | ((xi >= 0 && xi < k.q_len) ? (q[xi] & xmask) : xbeyond);
                }
                ext[z + i] = e;
            }
            z += len;
        }
        if (t_consumes_ref(op)) ref_pos += len;
    }
    for (int64_t i = z + lane; i < rlen; i += 32) {                         // bases the CIGAR does not reach: selector P only
        const int64_t pi = (int64_t)pos + i - k.start;
        ext[i] = (pi >= 0 && pi < k.q_len) ? (q[pi] & pmask) : (uint8_t)0;
    }
}

// one CTA per container: its own arrays into the carry arrays (containers overlap: atomics)
__global__ void __launch_bounds__(256) k_trk_fold(TrkArgs a, int32_t *carry_cov, int32_t *carry_mm)
{
    const TrkCont k = a.cont[blockIdx.x];
    if (!k.ok) return;
    for (int32_t i = threadIdx.x; i < k.own_len; i += 256) {
        const int32_t v = a.own_cov[k.own_off + i], m = a.own_mm[k.own_off + i];
        if (v) atomicAdd(carry_cov + k.start + i, v);
        if (m) atomicAdd(carry_mm + k.start + i, m);
    }
}

}  // namespace

// the X / P selectors of a policy set, in policy order: bit 1 + e of the ext byte (records.cu assigns the same bits)
int tracks_selectors(const cramgpu_qs_policy *pol, int n_pol, int32_t *kind, int32_t *param)
{
    int n = 0;
    for (int k = 0; k < n_pol; k++)
        for (int b = 0; b < pol[k].n_base_cats && b < 8; b++) {
            const int cat = pol[k].base_cat[b] & 0xFF;
            if (cat != 5 && cat != 6) continue;
            if (n < 7) { kind[n] = cat; param[n] = pol[k].base_cat[b] >> 8; }
            n++;
        }
    return n;
}

extern "C" int cramgpu_tracks_create(cramgpu_ctx *ctx, const uint8_t *ref, uint64_t ref_len, uint32_t window, cramgpu_tracks **out)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!ref || !out || ref_len == 0 || ref_len > 0x7FFFFFF0ull) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "tracks: a reference of 1 .. 2^31 - 17 bases");
    *out = nullptr;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cramgpu_tracks *t = new cramgpu_tracks();
    t->device = ctx->device; t->ref_len = ref_len;
    const uint64_t w = window ? window : 1000000u;                          // ReferenceTracks.DEFAULT_WINDOW_SIZE
    t->window = (int64_t)(w < ref_len ? w : ref_len);
    cudaError_t e = cudaMalloc((void **)&t->d_ref, ref_len + 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&t->d_cov, 4 * (ref_len + 2));
    if (e == cudaSuccess) e = cudaMalloc((void **)&t->d_mm, 4 * (ref_len + 2));
    if (e == cudaSuccess) e = cudaMemcpy(t->d_ref, ref, ref_len, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(t->d_cov, 0, 4 * (ref_len + 2));
    if (e == cudaSuccess) e = cudaMemset(t->d_mm, 0, 4 * (ref_len + 2));
    if (e != cudaSuccess) {
        cudaFree(t->d_ref); cudaFree(t->d_cov); cudaFree(t->d_mm); delete t;
        return ctx_cuda(ctx, e, "cramgpu_tracks_create");
    }
    *out = t;
    return CRAMGPU_OK;
}

extern "C" void cramgpu_tracks_destroy(cramgpu_ctx *ctx, cramgpu_tracks *t)
{
    if (!t) return;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFree(t->d_ref); cudaFree(t->d_cov); cudaFree(t->d_mm);
    delete t;
}

extern "C" int64_t cramgpu_tracks_position(const cramgpu_tracks *t) { return t ? t->position : 0; }

extern "C" int cramgpu_tracks_update_dev(cramgpu_ctx *ctx, cramgpu_tracks *t, const cramgpu_reads *reads, const cramgpu_qs_policy *pol,
                                         int n_pol, int32_t *container_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    ctx->trk_armed = false;
    if (!t || !reads) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    const uint32_t nc = reads->n_containers;
    const uint64_t n = reads->n_reads;
    if (container_status) for (uint32_t c = 0; c < nc; c++) container_status[c] = CRAMGPU_OK;
    TrkArgs a;
    memset(&a, 0, sizeof a);
    a.n_sel = (n_pol > 0 && pol) ? tracks_selectors(pol, n_pol, a.sel_kind, a.sel_param) : 0;
    if (a.n_sel == 0) return CRAMGPU_OK;                                    // areReferenceTracksRequired() is false: the tracks are not touched
    if (a.n_sel > 7) return ctx_fail(ctx, CRAMGPU_ERR_UNSUPPORTED, "more than 7 X / P selectors in a lossy model");
    if (t->poisoned) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "tracks: an earlier call failed, the tracks are undefined (as after the reference's exception)");
    if (t->device != ctx->device) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "tracks belong to another device");
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!reads->cont_off || !reads->bases || !reads->read_off || !reads->cigar || !reads->cigar_off || !reads->pos || !reads->flag)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;

    Carver cv;
    const size_t o_co = cv.take(8 * ((size_t)nc + 1)), o_sum = cv.take(sizeof(TrkSum) * (size_t)nc), o_cont = cv.take(sizeof(TrkCont) * (size_t)nc);
    if (int rc = pool_reserve(ctx, ctx->trk_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->trk_scratch.p;
    Carver pc;
    const size_t p_co = pc.take(8 * ((size_t)nc + 1), 16), p_sum = pc.take(sizeof(TrkSum) * (size_t)nc, 16), p_cont = pc.take(sizeof(TrkCont) * (size_t)nc, 16);
    const size_t p_tot = pc.take(16, 16);
    if (int rc = pinned_reserve(ctx, &ctx->pin_trk, &ctx->pin_trk_cap, pc.off + 64)) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));                               // the page-locked staging of an earlier call is free again
    uint8_t *pm = (uint8_t *)ctx->pin_trk;
    memcpy(pm + p_co, reads->cont_off, 8 * ((size_t)nc + 1));
    TrkSum *hs = (TrkSum *)(pm + p_sum);
    for (uint32_t c = 0; c < nc; c++) { hs[c].min_start = 0x7FFFFFFF; hs[c].max_end = hs[c].max_cov = hs[c].max_query = hs[c].max_start = (int32_t)0x80000000; hs[c].flags = 0; }
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_co, pm + p_co, 8 * ((size_t)nc + 1), cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_sum, pm + p_sum, sizeof(TrkSum) * (size_t)nc, cudaMemcpyHostToDevice, st));
    a.n_reads = n; a.n_containers = nc; a.cont_off = (const uint64_t *)(S + o_co);
    a.bases = reads->bases; a.read_off = reads->read_off; a.cigar = reads->cigar; a.cigar_off = reads->cigar_off;
    a.pos = reads->pos; a.flag = reads->flag; a.rflags = reads->rflags;
    a.sum = (TrkSum *)(S + o_sum); a.cont = (const TrkCont *)(S + o_cont);
    a.ref = t->d_ref; a.ref_len = (int64_t)t->ref_len; a.window = t->window;
    a.carry_cov = t->d_cov; a.carry_mm = t->d_mm;
    LAUNCH(ctx, st, "k_trk_ends", k_trk_ends<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a));
    CUDA_TRY(ctx, cudaMemcpyAsync(pm + p_sum, S + o_sum, sizeof(TrkSum) * (size_t)nc, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(pm + p_tot, reads->read_off + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t bases_len = *(const uint64_t *)(pm + p_tot);

    // ---- the window per container: CramSerilization.convert [:117-127] + ReferenceTracks.ensureRange / moveForwardTo
    TrkCont *hc = (TrkCont *)(pm + p_cont);
    int64_t position = t->position, own_total = 0, q_total = 0;
    const int64_t W = t->window, L = (int64_t)t->ref_len;
    int32_t reach = (int32_t)0x80000000;
    int fail_code = CRAMGPU_OK;
    uint32_t fail_at = nc;
    const char *fail_msg = "";
    for (uint32_t c = 0; c < nc; c++) {
        TrkCont &k = hc[c];
        memset(&k, 0, sizeof k);
        k.position = position; k.reach = reach;
        if (fail_code) { if (container_status) container_status[c] = fail_code; continue; }
        const TrkSum &s = hs[c];
        if (s.min_start == 0x7FFFFFFF) { k.ok = 1; continue; }             // no placed record: the sequence is not in the spans
        const int64_t start = s.min_start, span = (int64_t)s.max_end - start;                  // AlignmentSpan.add: union [AlignmentSpan.java:50-61]
        if (start < position) { fail_code = CRAMGPU_ERR_ARG; fail_msg = "Cannot move the window backwards"; }
        else if (start + span > position + W) {                            // moveForwardTo(start)
            if (start - 1 >= L) { fail_code = CRAMGPU_ERR_ARG; fail_msg = "New position is beyond the reference"; }
            else position = start > L - W + 1 ? L - W + 1 : start;
        }
        const bool at_end = position + W - 1 >= L;                          // beyond the window = beyond the reference
        if (!fail_code && s.flags & 1) { fail_code = CRAMGPU_ERR_UNSUPPORTED; fail_msg = "a record without bases under the X / P selectors (the reference indexes an empty array)"; }
        if (!fail_code && !at_end && (int64_t)s.max_query > position + W) {
            fail_code = CRAMGPU_ERR_UNSUPPORTED; fail_msg = "a container reaches beyond the tracks' window: the window would move between its records";
        }
        if (!fail_code && at_end && (int64_t)s.max_start > L && (int64_t)s.max_end > position + W) {
            fail_code = CRAMGPU_ERR_ARG; fail_msg = "New position is beyond the reference";
        }
        if (fail_code) { fail_at = c; if (container_status) container_status[c] = fail_code; k.position = position; continue; }
        k.position = position; k.start = (int32_t)start; k.ok = 1;
        int64_t last = s.max_cov < position + W - 1 ? s.max_cov : position + W - 1;                // counts stop at the window's end
        k.own_len = last >= start ? (int32_t)(last - start + 1) : 0;
        k.own_off = own_total;
        own_total += (int64_t)k.own_len + 1;
        int64_t qlast = (int64_t)s.max_query - 1 < position + W - 1 ? (int64_t)s.max_query - 1 : position + W - 1;     // asked about, and inside the window
        k.q_len = qlast >= start ? (int32_t)(qlast - start + 1) : 0;
        k.q_off = q_total;
        q_total += ((int64_t)k.q_len + 15) & ~15ll;
        if (k.own_len && (int32_t)last > reach) reach = (int32_t)last;
        k.reach = reach;
    }
    if (own_total > (int64_t)1 << 31) return ctx_fail(ctx, CRAMGPU_ERR_UNSUPPORTED, "tracks: the containers of one call span more than 2^31 positions in all");
    Carver co;
    const size_t q_cov = co.take(4 * (size_t)(own_total + 1)), q_mm = co.take(4 * (size_t)(own_total + 1));
    const size_t zero_end = co.off;
    const size_t q_bits = co.take((size_t)q_total + 16);
    if (int rc = pool_reserve(ctx, ctx->trk_own, co.off)) return rc;
    if (int rc = pool_reserve(ctx, ctx->trk_ext, bases_len + 16)) return rc;
    uint8_t *O = (uint8_t *)ctx->trk_own.p;
    a.own_cov = (int32_t *)(O + q_cov); a.own_mm = (int32_t *)(O + q_mm); a.ext = (uint8_t *)ctx->trk_ext.p; a.qbits = O + q_bits;
    CUDA_TRY(ctx, cudaMemsetAsync(O, 0, zero_end, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_cont, pm + p_cont, sizeof(TrkCont) * (size_t)nc, cudaMemcpyHostToDevice, st));
    LAUNCH(ctx, st, "k_trk_own", k_trk_own<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a));
    LAUNCH(ctx, st, "k_trk_scan", k_trk_scan<<<nc, 1024, 0, st>>>(a));
    LAUNCH(ctx, st, "k_trk_bits", k_trk_bits<<<nc, 256, 0, st>>>(a));
    LAUNCH(ctx, st, "k_trk_masks", k_trk_masks<<<(unsigned)((n + 3) / 4), 128, 0, st>>>(a));
    LAUNCH(ctx, st, "k_trk_fold", k_trk_fold<<<nc, 256, 0, st>>>(a, t->d_cov, t->d_mm));
    CUDA_TRY(ctx, cudaGetLastError());
    t->position = position;
    if (fail_code) {
        t->poisoned = true;
        return ctx_fail(ctx, fail_code, "container %u: %s", fail_at, fail_msg);
    }
    ctx->trk_armed = true; ctx->trk_reads = n; ctx->trk_nsel = a.n_sel;
    memcpy(ctx->trk_kind, a.sel_kind, sizeof a.sel_kind); memcpy(ctx->trk_param, a.sel_param, sizeof a.sel_param);
    return CRAMGPU_OK;
}
