// mates.cu -- mate linking of the records of a container.
//
// CramSerilization.setMateInfo / detach (src/main/java/htsjdk/samtools/cram/CramSerilization.java:258-373) decides, for a
// coordinate-sorted container, which records are written as linked mates (CF bit "mate downstream" + the NF series =
// recordsToNextFragment) and which are "detached" (mate fields written in full: MF, NS, NP, TS).  The reference walks the
// records with two TreeMaps by read name (primary / secondary alignments).  What it computes is a function of each name
// group alone:
//   a multi-fragment record whose (class, name) occurs once in the container           -> detached
//   occurs twice: the first is linked to the second (NF = distance - 1) if the pair is "predictable" [:318-349]:
//        first-of-pair / last-of-pair flags fit, both template sizes equal +-computeInsertSize, each record's mate fields
//        describe the other; else both detached
//   occurs three times or more: the third arrival detaches the chain, later ones find it linked and detach again  -> all detached
//   a record without the paired flag                                                     -> detached
// so one CTA per container builds a hash table of the names (open addressing, exact name comparison on every probe, a table
// slot holds the two smallest record indices of its group and the group size) and every record reads its verdict from its
// group's slot.  The tables live in global scratch, one per resident CTA.
#include "internal.cuh"
#include <string.h>

namespace {

#define MATE_EMPTY 0xFFFFFFFFu

struct MateArgs {
    uint64_t n_reads; uint32_t n_containers; uint32_t sorted;
    const uint64_t *cont_off;                 // device copy
    const uint16_t *flag;
    const uint8_t *names; const uint64_t *name_off;
    const int32_t *ref_id, *pos, *mate_ref_id, *mate_pos, *tlen;
    const uint64_t *read_off, *feat_off;
    const uint8_t *f_op; const int32_t *f_val, *f_aux;
    uint8_t *mate_cf; int32_t *nf;
    uint32_t *tables;                         // per CTA: holder | first | second | count, `tsize` entries each
    uint32_t tsize;
};

__device__ __forceinline__ bool same_name(const MateArgs &a, uint64_t i, uint64_t j)
{
    const uint64_t ai = a.name_off[i], aj = a.name_off[j];
    const uint64_t li = a.name_off[i + 1] - ai;
    if (li != a.name_off[j + 1] - aj) return false;
    for (uint64_t k = 0; k < li; k++) if (a.names[ai + k] != a.names[aj + k]) return false;
    return true;
}

// CramCompressionRecord.getAlignmentEnd [CramCompressionRecord.java:160-194]
__device__ __forceinline__ int32_t alignment_end(const MateArgs &a, uint64_t r)
{
    if (a.flag[r] & 4) return 0;
    int32_t span = (int32_t)(a.read_off[r + 1] - a.read_off[r]);
    for (uint64_t f = a.feat_off[r]; f < a.feat_off[r + 1]; f++) {
        const uint8_t op = a.f_op[f];
        if (op == 'i') span -= 1;
        else if (op == 'S') span -= a.f_aux[f];
        else if (op == 'D') span += a.f_val[f];
    }
    return a.pos[r] + span - 1;
}

// CramNormalizer.computeInsertSize (src/main/java/htsjdk/samtools/cram/build/CramNormalizer.java:455-470)
__device__ __forceinline__ int32_t insert_size(const MateArgs &a, uint64_t x, uint64_t y)
{
    const uint16_t fx = a.flag[x], fy = a.flag[y];
    if ((fx & 4) || (fy & 4)) return 0;
    if ((a.ref_id ? a.ref_id[x] : 0) != (a.ref_id ? a.ref_id[y] : 0)) return 0;
    const int32_t px = (fx & 0x10) ? alignment_end(a, x) : a.pos[x];
    const int32_t py = (fy & 0x10) ? alignment_end(a, y) : a.pos[y];
    return py - px + (py >= px ? 1 : -1);
}

// the record fields as Sam2CramRecordFactory.createCramRecord sets them [Sam2CramRecordFactory.java:112-146]
__device__ __forceinline__ bool is_first(uint16_t f) { return (f & 1) && (f & 0x40); }
__device__ __forceinline__ bool is_last(uint16_t f) { return (f & 1) && !(f & 0x40) && (f & 0x80); }

__global__ void __launch_bounds__(256) k_mate_link(MateArgs a)
{
    uint32_t *holder = a.tables + (size_t)blockIdx.x * 4 * a.tsize, *first = holder + a.tsize, *second = first + a.tsize, *count = second + a.tsize;
    const uint32_t tid = threadIdx.x;
    for (uint32_t c = blockIdx.x; c < a.n_containers; c += gridDim.x) {
        const uint64_t i0 = a.cont_off[c], i1 = a.cont_off[c + 1];
        const uint32_t n = (uint32_t)(i1 - i0);
        if (!a.sorted) {                                   // every record detached [CramSerilization.java:175-179]
            for (uint64_t i = i0 + tid; i < i1; i += 256) { a.mate_cf[i] = 2; a.nf[i] = -1; }
            continue;
        }
        uint32_t size = 64;
        while (size < 2 * n) size <<= 1;                   // <= tsize by construction
        const uint32_t mask = size - 1;
        for (uint32_t s = tid; s < size; s += 256) { holder[s] = MATE_EMPTY; first[s] = MATE_EMPTY; second[s] = MATE_EMPTY; count[s] = 0; }
        __syncthreads();
        // ---- every paired record finds the slot of its (class, name) group
        for (uint64_t i = i0 + tid; i < i1; i += 256) {
            const uint16_t f = a.flag[i];
            if (!(f & 1)) { a.mate_cf[i] = 2; a.nf[i] = -1; continue; }
            uint32_t h = 2166136261u ^ ((f & 0x100) ? 0x9E3779B9u : 0u);
            for (uint64_t k = a.name_off[i]; k < a.name_off[i + 1]; k++) h = (h ^ a.names[k]) * 16777619u;
            h ^= h >> 15;
            const uint32_t me = (uint32_t)(i - i0);
            uint32_t s = h & mask;
            for (;;) {
                uint32_t cur = holder[s];
                if (cur == MATE_EMPTY) {
                    cur = atomicCAS(&holder[s], MATE_EMPTY, me);
                    if (cur == MATE_EMPTY) break;          // claimed
                }
                if (((a.flag[i0 + cur] ^ f) & 0x100) == 0 && same_name(a, i0 + cur, i)) break;
                s = (s + 1) & mask;
            }
            a.nf[i] = (int32_t)s;                          // kept until the verdict
            atomicMin(&first[s], me);
            atomicAdd(&count[s], 1u);
        }
        __syncthreads();
        for (uint64_t i = i0 + tid; i < i1; i += 256) {
            if (!(a.flag[i] & 1)) continue;
            const uint32_t s = (uint32_t)a.nf[i], me = (uint32_t)(i - i0);
            if (me != first[s]) atomicMin(&second[s], me);
        }
        __syncthreads();
        // ---- verdicts
        for (uint64_t i = i0 + tid; i < i1; i += 256) {
            const uint16_t f = a.flag[i];
            if (!(f & 1)) continue;
            const uint32_t s = (uint32_t)a.nf[i];
            uint8_t cf = 2; int32_t nf = -1;
            if (count[s] == 2) {
                const uint64_t r = i0 + first[s], l = i0 + second[s];
                const uint16_t fr = a.flag[r], fl = a.flag[l];
                bool ok = (is_first(fr) && is_last(fl)) || (is_first(fl) && is_last(fr));           // [:326]
                if (ok) {
                    const int32_t tl = insert_size(a, r, l);
                    ok = a.tlen[r] == tl && a.tlen[l] == -tl;                                           // [:328-341]
                }
                const int32_t ref_r = a.ref_id ? a.ref_id[r] : 0, ref_l = a.ref_id ? a.ref_id[l] : 0;
                ok = ok && a.mate_ref_id[r] == ref_l && ref_r == a.mate_ref_id[l] &&                   // [:345-349]
                     (((fr & 0x20) != 0) == ((fl & 0x10) != 0)) && (((fl & 0x20) != 0) == ((fr & 0x10) != 0)) &&
                     a.mate_pos[r] == a.pos[l] && a.mate_pos[l] == a.pos[r];
                if (ok) {
                    if (i == r) { cf = 4; nf = (int32_t)(l - r) - 1; }                                  // [:300-306]
                    else cf = 0;
                }
            }
            a.mate_cf[i] = cf; a.nf[i] = nf;
        }
        __syncthreads();
    }
}


// ---- decode: the pairing part of CramNormalizer.normalize (CramNormalizer.java:67-102, 236-270), the inverse of the above ----
// next / previous links from the compression flags and NF; restoreMateInfo along every chain that starts at a record nobody
// points to.  The reference does it record by record; what comes out is order-free except where several records point to
// one: `previous` is then the LAST of them, and where several chains end in one record its mate is the head that came
// last -- both are maxima over record indices.
struct RestArgs {
    uint64_t n_reads; uint32_t n_containers;
    const uint64_t *cont_off;
    const uint16_t *flag; const uint8_t *cf; uint8_t *mate_flags; int32_t *nf;
    const int32_t *ref_id; int32_t *pos, *mate_ref_id, *mate_pos, *tlen;
    int32_t *next, *prev, *winner;            // scratch, record indices (-1 = none)
    int32_t *status;                          // per container
    MateArgs m;                               // read-only view for getAlignmentEnd / computeInsertSize
};
__device__ __forceinline__ bool rest_linked(const RestArgs &a, uint64_t r) { return (a.flag[r] & 1) && !(a.cf[r] & 2); }   // isMultiFragment() && !isDetached()
__device__ __forceinline__ uint32_t rest_container_of(const uint64_t *cont_off, uint32_t nc, uint64_t r)
{
    uint32_t lo = 0, hi = nc;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cont_off[mid] <= r) lo = mid; else hi = mid; }
    return lo;
}

__global__ void __launch_bounds__(256) k_mate_rest_links(RestArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= a.n_reads) return;
    if (a.ref_id && a.ref_id[r] == -1) a.pos[r] = 0;                        // NO_ALIGNMENT_START for an unplaced record [:69-72]
    if (!rest_linked(a, r)) { a.nf[r] = -1; a.next[r] = -1; return; }      // [:81-87] (its `previous` is dropped too: see rest_prev)
    a.next[r] = -1;
    if (a.cf[r] & 4) {                                                      // isHasMateDownStream() [:88-93]
        const uint32_t c = rest_container_of(a.cont_off, a.n_containers, r);
        const int64_t j = (int64_t)r + 1 + a.nf[r];
        if (a.nf[r] < 0 || j >= (int64_t)a.cont_off[c + 1]) { a.status[c] = CRAMGPU_ERR_CORRUPT; return; }      // records.get() throws
        a.next[r] = (int32_t)j;
        atomicMax(a.prev + j, (int32_t)r);
    }
}
__device__ __forceinline__ int32_t rest_prev(const RestArgs &a, uint64_t r) { return rest_linked(a, r) ? a.prev[r] : -1; }

__global__ void __launch_bounds__(256) k_mate_rest_last(RestArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= a.n_reads || rest_prev(a, r) != -1 || a.next[r] == -1) return;
    int32_t cur = (int32_t)r;
    while (a.next[cur] != -1) cur = a.next[cur];                            // next > cur: the walk ends
    atomicMax(a.winner + cur, (int32_t)r);
}

// setNextMate [:259-266]
__device__ __forceinline__ void rest_set_next_mate(const RestArgs &a, int32_t r, int32_t m)
{
    const uint16_t fm = a.flag[m];
    uint8_t f = a.mate_flags[r];                                            // CramCompressionRecord.mateFlags [:270-283]
    f = (fm & 0x4) ? (f | 2) : (f & ~2);                                    // setMateUnmapped(next.isSegmentUnmapped())
    f = (fm & 0x10) ? (f | 1) : (f & ~1);                                   // setMateNegativeStrand(next.isNegativeStrand())
    a.mate_flags[r] = f;
    const int32_t ms = a.ref_id ? a.ref_id[m] : 0;
    a.mate_ref_id[r] = ms;
    a.mate_pos[r] = ms == -1 ? 0 : a.pos[m];
}

__global__ void __launch_bounds__(256) k_mate_rest_apply(RestArgs a)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= a.n_reads || rest_prev(a, r) != -1 || a.next[r] == -1) return;
    int32_t cur = (int32_t)r;
    while (a.next[cur] != -1) { rest_set_next_mate(a, cur, a.next[cur]); cur = a.next[cur]; }      // restoreMateInfo [:236-257]
    const int32_t tl = insert_size(a.m, r, (uint64_t)cur);
    a.tlen[r] = tl;
    if (a.winner[cur] == (int32_t)r) { rest_set_next_mate(a, cur, (int32_t)r); a.tlen[cur] = -tl; }
}

}  // namespace

static int mate_impl(cramgpu_ctx *ctx, const cramgpu_mate_in *in, uint8_t *d_cf, int32_t *d_nf, cudaStream_t st)
{
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->cont_off || !d_cf || !d_nf || !in->flag) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->coordinate_sorted && (!in->names || !in->name_off || !in->pos || !in->mate_ref_id || !in->mate_pos || !in->tlen || !in->read_off ||
                                  !in->feat_off))
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->cont_off[0] != 0 || in->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    uint64_t max_reads = 0;
    for (uint32_t c = 0; c < nc; c++) {
        if (in->cont_off[c] > in->cont_off[c + 1]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off not monotonic");
        max_reads = std::max(max_reads, in->cont_off[c + 1] - in->cont_off[c]);
    }
    if (max_reads > (1ull << 30)) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "container too large");
    uint32_t tsize = 64;
    while (tsize < 2 * max_reads) tsize <<= 1;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    // as many CTAs as 256 MB of tables allow, at most four per SM
    uint32_t grid = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(nc, 4ull * sms), std::max<uint64_t>(1, (256ull << 20) / (16ull * tsize)));
    Carver cv;
    const size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_tab = cv.take(16 * (size_t)tsize * grid);
    if (int rc = pool_reserve(ctx, ctx->mate_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->mate_scratch.p;
    if (int rc = pinned_reserve(ctx, &ctx->pin_slice, &ctx->pin_slice_cap, 8 * (size_t)(nc + 1) + 64)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_slice));
    memcpy(ctx->pin_slice, in->cont_off, 8 * (size_t)(nc + 1));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_cont, ctx->pin_slice, 8 * (size_t)(nc + 1), cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_slice, st));
    MateArgs a;
    memset(&a, 0, sizeof a);
    a.n_reads = n; a.n_containers = nc; a.sorted = in->coordinate_sorted; a.cont_off = (const uint64_t *)(S + o_cont);
    a.flag = in->flag; a.names = in->names; a.name_off = in->name_off; a.ref_id = in->ref_id; a.pos = in->pos;
    a.mate_ref_id = in->mate_ref_id; a.mate_pos = in->mate_pos; a.tlen = in->tlen; a.read_off = in->read_off; a.feat_off = in->feat_off;
    a.f_op = in->f_op; a.f_val = in->f_val; a.f_aux = in->f_aux; a.mate_cf = d_cf; a.nf = d_nf;
    a.tables = (uint32_t *)(S + o_tab); a.tsize = tsize;
    LAUNCH(ctx, st, "k_mate_link", k_mate_link<<<grid, 256, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}

extern "C" int cramgpu_mate_link_dev(cramgpu_ctx *ctx, const cramgpu_mate_in *in, uint8_t *d_mate_cf, int32_t *d_nf, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return mate_impl(ctx, in, d_mate_cf, d_nf, stream ? (cudaStream_t)stream : ctx->stream);
}

extern "C" int cramgpu_mate_link(cramgpu_ctx *ctx, const cramgpu_mate_in *in, uint8_t *mate_cf, int32_t *nf)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !mate_cf || !nf) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const uint64_t n = in->n_reads;
    if (n == 0 || in->n_containers == 0) return CRAMGPU_OK;
    MemMap mm;
    auto last_u64 = [&](const uint64_t *p, uint64_t *v) -> int {
        *v = 0;
        if (!p) return CRAMGPU_OK;
        if (mm.kind(p) == MEM_DEVICE) {
            CUDA_TRY(ctx, cudaMemcpyAsync(v, p + n, 8, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(ctx, cudaStreamSynchronize(st));
        } else *v = p[n];
        return CRAMGPU_OK;
    };
    uint64_t name_bytes = 0, nfeat = 0;
    if (int rc = last_u64(in->name_off, &name_bytes)) return rc;
    if (int rc = last_u64(in->feat_off, &nfeat)) return rc;
    struct Item { const void *src; size_t bytes; size_t off; };
    std::vector<Item> items;
    Carver cv;
    auto add = [&](const void *p, size_t bytes) -> size_t {
        if (!p || mm.kind(p) == MEM_DEVICE) return (size_t)-1;
        size_t o = cv.take(bytes + 16); items.push_back({p, bytes, o}); return o;
    };
    const size_t i_fl = add(in->flag, 2 * n), i_nm = add(in->names, name_bytes), i_no = add(in->name_off, 8 * (n + 1));
    const size_t i_id = add(in->ref_id, 4 * n), i_pos = add(in->pos, 4 * n), i_mr = add(in->mate_ref_id, 4 * n), i_mp = add(in->mate_pos, 4 * n);
    const size_t i_tl = add(in->tlen, 4 * n), i_ro = add(in->read_off, 8 * (n + 1)), i_fo = add(in->feat_off, 8 * (n + 1));
    const size_t i_op = add(in->f_op, nfeat), i_fv = add(in->f_val, 4 * nfeat), i_fa = add(in->f_aux, 4 * nfeat);
    const size_t o_cf = cv.take(n + 16), o_nf = cv.take(4 * n + 16);
    if (int rc = pool_reserve(ctx, ctx->slice_in, cv.off + 16)) return rc;
    uint8_t *D = (uint8_t *)ctx->slice_in.p;
    for (const Item &it : items)
        if (it.bytes) CUDA_TRY(ctx, memcpy_pieces(D + it.off, it.src, it.bytes, cudaMemcpyHostToDevice, st));
    auto at = [&](size_t o, const void *orig) -> const void * { return o == (size_t)-1 ? orig : (const void *)(D + o); };
    cramgpu_mate_in d = *in;
    d.flag = (const uint16_t *)at(i_fl, in->flag); d.names = (const uint8_t *)at(i_nm, in->names); d.name_off = (const uint64_t *)at(i_no, in->name_off);
    d.ref_id = (const int32_t *)at(i_id, in->ref_id); d.pos = (const int32_t *)at(i_pos, in->pos);
    d.mate_ref_id = (const int32_t *)at(i_mr, in->mate_ref_id); d.mate_pos = (const int32_t *)at(i_mp, in->mate_pos);
    d.tlen = (const int32_t *)at(i_tl, in->tlen); d.read_off = (const uint64_t *)at(i_ro, in->read_off);
    d.feat_off = (const uint64_t *)at(i_fo, in->feat_off); d.f_op = (const uint8_t *)at(i_op, in->f_op);
    d.f_val = (const int32_t *)at(i_fv, in->f_val); d.f_aux = (const int32_t *)at(i_fa, in->f_aux);
    if (int rc = mate_impl(ctx, &d, D + o_cf, (int32_t *)(D + o_nf), st)) return rc;
    CUDA_TRY(ctx, cudaMemcpyAsync(mate_cf, D + o_cf, n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(nf, D + o_nf, 4 * n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CRAMGPU_OK;
}


extern "C" int cramgpu_mate_restore_dev(cramgpu_ctx *ctx, const cramgpu_mate_restore *io, int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!io) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    const uint64_t n = io->n_reads;
    const uint32_t nc = io->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!io->cont_off || !io->flag || !io->cf || !io->mate_flags || !io->nf || !io->pos || !io->mate_ref_id || !io->mate_pos || !io->tlen || !io->read_off ||
        !io->feat_off || !d_status)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (n > 0x7FFFFFF0ull) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "too many records");
    if (io->cont_off[0] != 0 || io->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    Carver cv;
    const size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_next = cv.take(4 * n), o_prev = cv.take(4 * n), o_win = cv.take(4 * n);
    if (int rc = pool_reserve(ctx, ctx->mate_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->mate_scratch.p;
    if (int rc = pinned_reserve(ctx, &ctx->pin_slice, &ctx->pin_slice_cap, 8 * (size_t)(nc + 1) + 64)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_slice));
    memcpy(ctx->pin_slice, io->cont_off, 8 * (size_t)(nc + 1));
    CUDA_TRY(ctx, cudaMemcpyAsync(S + o_cont, ctx->pin_slice, 8 * (size_t)(nc + 1), cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev_slice, st));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_prev, 0xFF, 4 * n, st));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_win, 0xFF, 4 * n, st));
    CUDA_TRY(ctx, cudaMemsetAsync(d_status, 0, 4 * (size_t)nc, st));
    RestArgs a;
    memset(&a, 0, sizeof a);
    a.n_reads = n; a.n_containers = nc; a.cont_off = (const uint64_t *)(S + o_cont);
    a.flag = io->flag; a.cf = io->cf; a.mate_flags = io->mate_flags; a.nf = io->nf; a.ref_id = io->ref_id; a.pos = io->pos;
    a.mate_ref_id = io->mate_ref_id; a.mate_pos = io->mate_pos; a.tlen = io->tlen;
    a.next = (int32_t *)(S + o_next); a.prev = (int32_t *)(S + o_prev); a.winner = (int32_t *)(S + o_win); a.status = d_status;
    a.m.n_reads = n; a.m.flag = io->flag; a.m.ref_id = io->ref_id; a.m.pos = io->pos; a.m.read_off = io->read_off; a.m.feat_off = io->feat_off;
    a.m.f_op = io->f_op; a.m.f_val = io->f_val; a.m.f_aux = io->f_aux;
    const unsigned grid = (unsigned)((n + 255) / 256);
    LAUNCH(ctx, st, "k_mate_rest_links", k_mate_rest_links<<<grid, 256, 0, st>>>(a));
    LAUNCH(ctx, st, "k_mate_rest_last", k_mate_rest_last<<<grid, 256, 0, st>>>(a));
    LAUNCH(ctx, st, "k_mate_rest_apply", k_mate_rest_apply<<<grid, 256, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}
