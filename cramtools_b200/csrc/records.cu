// records.cu -- per-record work of the CRAM slice hot path for sm_100a.
//
// Encode: two passes over the records (count, then write) around per-container scans; in each pass a record is handled by the
// cheapest kernel that can handle it:
//   k_ref_norm           normalizeBase over the reference windows, once per call
//   k_rec_simple<W, PB>  one THREAD per record, in order: classifies the record; `<len>M` records are compared with the
//                        normalised reference as aligned words; records with several CIGAR operators and the odd ones go to lists
//   k_rec_multi<W, PB>   the listed records, one thread per entry, all CIGAR operators (PB: quality models with base categories,
//                        evaluated bit-parallel on groups of four bases: pb_mask / pb_planes)
//   k_rec_walk<W>        one WARP per record (ballots, warp compaction): the fallback for region overhang and odd CIGARs
//   k_rec_matrix         SubstitutionMatrix ranking per container            [SubstitutionMatrix.java:42-57,141-169]
//   k_rec_scan / k_rec_layout   per-container exclusive scans of the 14 per-record counters; series-major region offsets
//   k_rec_forced_flat / k_rec_forced_scores   the preserved score arrays under a lossy model
//   k_rec_bin_fault      (only after a BIN treatment met a score >= 110) marks the records the reference throws for
// All of them restate
//   Sam2CramRecordFactory.createVariations / addSubstitutionsAndMaskedBases [Sam2CramRecordFactory.java:205-335]
//   QualityScorePreservation.addQualityScores / addQS     (lossy/QualityScorePreservation.java:276-301,332-489)
//   Writer.write, feature part                            [Writer.java:180-258]
//
// Decode:
//   k_unpack             series -> feature arrays: the feature part of CramRecordReader.read (row P)
//   k_norm_simple        thread per record: stored score arrays and bases of records without features (warp-wide copies when
//                        32 neighbouring records are contiguous); the others go to a list
//   k_norm               the list: CramNormalizer.restoreReadBases / restoreQualityScores (build/CramNormalizer.java:287-434)
#include "internal.cuh"

#define NCNT 14
enum { C_FN = 0, C_FC, C_FP, C_BS, C_DL, C_BA, C_SC, C_QS, C_RS, C_HC, C_PD, C_MQ, C_NFEAT, C_PAYLOAD };
enum { OP_M = 0, OP_I, OP_D, OP_N, OP_S, OP_H, OP_P, OP_EQ, OP_X };
#define SAM_UNMAPPED 4
#define MAX_POL 16

struct PolSet {
    int n;                       // < 0: lossless
    int per_base;                // any policy with base categories
    int any_bin;                 // any policy with the BIN treatment (the only one that can fault, see bin_fault)
    int n_ext;                   // X / P selectors in all policies together (each has a bit in the per-base ext bytes)
    cramgpu_qs_policy p[MAX_POL];
    uint8_t ext_bit[MAX_POL][8]; // policy k, base category b (an X or a P) -> its bit in the ext byte (1..7; bit 0 = "base of an M/=/X operator")
};

struct RecArgs {
    // inputs
    const uint8_t *ref, *bases, *quals;
    const uint8_t *qs_ext;       // X / P selectors: one byte per base (same offsets as bases) from tracks.cu, else NULL
    const uint64_t *read_off, *cigar_off;
    const uint32_t *cigar;
    const int32_t *pos;
    const uint16_t *flag;
    const uint8_t *mapq, *rflags;
    const uint64_t *cont_off, *ref_off, *ref_len;    // device copies of the host tables
    uint32_t n_containers;
    uint64_t n_reads;
    // pass-1 outputs / pass-2 inputs
    uint32_t *cnt;               // [NCNT][n_reads]: counts, then container-local exclusive offsets
    uint8_t *force;              // [n_reads]
    unsigned long long *subst;   // [n_containers][25]
    const uint8_t *codes;        // [n_containers][32]: 25 codes + pad
    const uint64_t *base;        // [n_containers][NCNT] absolute bases (series: byte offset in streams)
    int32_t *status;             // [1] first error
    // pass-2 outputs
    uint8_t *f_op; int32_t *f_pos, *f_val, *f_aux; uint64_t *feat_off;
    uint8_t *payload, *streams;
    uint64_t *read_qs_off, *read_ba_off;
    // fast path (k_rec_simple): normalised copy of the reference, per-container totals, and the list of the reads
    // it leaves to the general walk
    const uint8_t *nref;         // normalizeBase(ref[i]) at the same offsets as ref
    const uint64_t *totals;      // [n_containers][NCNT] (write pass)
    uint64_t *clist_r; uint32_t *clist_c; unsigned long long *clist_n;      // reads of the general warp-per-read walk
    uint64_t *tlist_r; uint32_t *tlist_c; unsigned long long *tlist_n;      // reads with several CIGAR operators (k_rec_multi)
    uint64_t *wlist_r; uint32_t *wlist_c; unsigned long long *wlist_n;      // single-operator reads that have features (write pass: k_rec_multi)
};

__device__ __forceinline__ uint8_t norm_base(uint8_t b)      // cram/build/Utils.normalizeBase [Utils.java:31-49]
{
    uint8_t u = b & 0xDF;                                     // upper-case letters
    bool letter = (b | 0x20) == 'a' || (b | 0x20) == 'c' || (b | 0x20) == 'g' || (b | 0x20) == 't';
    return letter ? u : (uint8_t)'N';
}
__device__ __forceinline__ int base_index(uint8_t b)
{
    return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : b == 'N' ? 4 : -1;
}
__device__ __forceinline__ uint32_t itf8_len(uint32_t v)     // ITF8.writeUnsignedITF8 [ITF8.java:53-84]
{
    return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : v < (1u << 28) ? 4 : 5;
}
__device__ __forceinline__ void itf8_put(uint8_t *p, uint32_t v)
{
    if (v < (1u << 7)) { p[0] = (uint8_t)v; }
    else if (v < (1u << 14)) { p[0] = (uint8_t)((v >> 8) | 0x80); p[1] = (uint8_t)v; }
    else if (v < (1u << 21)) { p[0] = (uint8_t)((v >> 16) | 0xC0); p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)v; }
    else if (v < (1u << 28)) { p[0] = (uint8_t)((v >> 24) | 0xE0); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }
    else { p[0] = (uint8_t)((v >> 28) | 0xF0); p[1] = (uint8_t)(v >> 20); p[2] = (uint8_t)(v >> 12); p[3] = (uint8_t)(v >> 4); p[4] = (uint8_t)v; }
}
__device__ __forceinline__ bool consumes_read(int op) { return op == OP_M || op == OP_I || op == OP_S || op == OP_EQ || op == OP_X; }
__device__ __forceinline__ bool consumes_ref(int op) { return op == OP_M || op == OP_D || op == OP_N || op == OP_EQ || op == OP_X; }

// Binning.Illumina_binning_matrix (lossy/Binning.java:63-71), as the closed form of its 110 entries
__device__ __forceinline__ uint8_t illumina_bin(uint8_t q)
{
    return q < 2 ? q : q < 10 ? 6 : q < 20 ? 15 : q < 25 ? 22 : q < 30 ? 27 : q < 35 ? 33 : q < 40 ? 37 : 40;
}
// The table has 110 entries and is indexed with a signed byte: a score of 110..255 makes the reference throw
// ArrayIndexOutOfBoundsException out of addQualityScores [QualityScorePreservation.java:258-266].  The kernels note that in
// the call's status word; k_rec_bin_fault then marks the records (bit 7 of force_qs).
#define BIN_FAULT_Q 110
__device__ __forceinline__ void bin_fault(int32_t *status) { atomicCAS(status, 0, CRAMGPU_ERR_UNSUPPORTED); }
__device__ __forceinline__ uint8_t treat(uint8_t q, int treatment, int32_t *status)
{
    if (treatment == 1 && q >= BIN_FAULT_Q) bin_fault(status);
    return treatment == 1 ? illumina_bin(q) : treatment == 2 ? (uint8_t)0xFF : q;
}
__device__ __forceinline__ bool read_in_category(const cramgpu_qs_policy &p, uint16_t flag, uint8_t mapq)
{
    switch (p.read_cat) {
    case 0: case 1: return true;
    case 2: return (flag & SAM_UNMAPPED) != 0;
    case 3: return (int)mapq < p.read_param;
    default: return (int)mapq > p.read_param;
    }
}
// mask value of one per-base policy at one base (QualityScorePreservation.java:398-476)
// ext = the base's byte from tracks.cu (0 without X / P selectors), eb = the policy's row of PolSet::ext_bit
__device__ __forceinline__ bool base_mask(const cramgpu_qs_policy &p, bool match, bool is_ins, bool dmark, uint32_t ext, const uint8_t *eb)
{
    bool m = false;
    for (int b = 0; b < p.n_base_cats; b++) {
        switch (p.base_cat[b] & 0xFF) {
        case 1: m = match; break;          // MATCH: assignment for every read-consuming position
        case 2: m = !match; break;         // MISMATCH
        case 3: m = m || dmark; break;     // FLANKING_DELETION: only sets
        case 4: m = m || is_ins; break;    // INSERTION: only sets
        case 5: if (ext & 1) m = (ext >> eb[b]) & 1; break;       // LOWER_COVERAGE: assignment for the bases of M / = / X [:447-465]
        case 6: m = m || ((ext >> eb[b]) & 1); break;             // PILEUP: only sets [:466-470]
        }
    }
    return m;
}
// the score addQualityScores leaves at one base (0xFF = -1 = none), policies in their sorted order
__device__ __forceinline__ uint8_t score_at(const PolSet &ps, uint8_t q, uint16_t flag, uint8_t mapq, int32_t pos,
                                            bool match, bool is_ins, bool dmark, uint32_t ext, int32_t *status)
{
    uint8_t s = 0xFF;
    for (int k = 0; k < ps.n; k++) {
        const cramgpu_qs_policy &p = ps.p[k];
        if (!read_in_category(p, flag, mapq)) continue;
        if (p.n_base_cats == 0) { if (p.treatment != 2) s = treat(q, p.treatment, status); }
        else if (pos != 0 && base_mask(p, match, is_ins, dmark, ext, ps.ext_bit[k])) s = treat(q, p.treatment, status);
    }
    return s;
}

__device__ __forceinline__ uint32_t warp_excl_scan_u32(uint32_t v, uint32_t lane, uint32_t &total)
{
    uint32_t s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xFFFFFFFFu, s, d);
        if (lane >= (uint32_t)d) s += t;
    }
    total = __shfl_sync(0xFFFFFFFFu, s, 31);
    return s - v;
}

// container of read r: binary search over cont_off
__device__ __forceinline__ uint32_t container_of(const uint64_t *cont_off, uint32_t nc, uint64_t r)
{
    uint32_t lo = 0, hi = nc;              // cont_off[lo] <= r < cont_off[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (cont_off[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- byte copies by one thread -------------------------------------------------------------------
// 4 source bytes at any address: two aligned words and a funnel shift.  The second word is read only when the address
// is not aligned, so no word is touched that does not hold at least one byte of [p, p + 4).
__device__ __forceinline__ uint32_t ld_u32_any(const uint8_t *p)
{
    const uint32_t sh = (uint32_t)((uintptr_t)p & 3);
    const uint32_t *w = reinterpret_cast<const uint32_t *>(p - sh);
    const uint32_t lo = w[0], hi = sh ? w[1] : 0u;
    return __funnelshift_r(lo, hi, sh * 8);
}
// copy n bytes src -> dst, any alignment on both sides: 16-byte stores where dst allows, 4-byte loads.  xw maps a
// word of 4 source bytes, xb a single byte (the same function on one byte).
template <class FW, class FB>
__device__ __forceinline__ void copy_bytes(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, uint32_t n, FW xw, FB xb)
{
    uint32_t i = 0;
    while (i < n && ((uintptr_t)(dst + i) & 3)) { dst[i] = xb(src[i]); i++; }
    while (i + 4 <= n && ((uintptr_t)(dst + i) & 15)) { *reinterpret_cast<uint32_t *>(dst + i) = xw(ld_u32_any(src + i)); i += 4; }
    if (i + 16 <= n) {
        const uint8_t *s = src + i;
        const uint32_t sh = (uint32_t)((uintptr_t)s & 3), sh8 = sh * 8;
        const uint32_t *w = reinterpret_cast<const uint32_t *>(s - sh);
        uint32_t w0 = *w;
        for (; i + 16 <= n; i += 16) {
            const uint32_t w1 = w[1], w2 = w[2], w3 = w[3];
            const uint32_t w4 = sh ? w[4] : 0u;
            uint4 o;
            o.x = xw(__funnelshift_r(w0, w1, sh8)); o.y = xw(__funnelshift_r(w1, w2, sh8));
            o.z = xw(__funnelshift_r(w2, w3, sh8)); o.w = xw(__funnelshift_r(w3, w4, sh8));
            *reinterpret_cast<uint4 *>(dst + i) = o;
            w0 = sh ? w4 : ((i + 32 <= n) ? w[4] : 0u);
            w += 4;
        }
    }
    while (i + 4 <= n) { *reinterpret_cast<uint32_t *>(dst + i) = xw(ld_u32_any(src + i)); i += 4; }
    while (i < n) { dst[i] = xb(src[i]); i++; }
}
struct IdW { __device__ __forceinline__ uint32_t operator()(uint32_t w) const { return w; } };
struct IdB { __device__ __forceinline__ uint8_t operator()(uint8_t b) const { return b; } };

// ---- byte copies by one warp ---------------------------------------------------------------------
// 16 source bytes at any address as four words (aligned loads, funnel shift; the fifth word only when it holds data)
__device__ __forceinline__ uint4 ld_u128_any(const uint8_t *p)
{
    const uint32_t sh = (uint32_t)((uintptr_t)p & 3), sh8 = sh * 8;
    const uint32_t *w = reinterpret_cast<const uint32_t *>(p - sh);
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = sh ? w[4] : 0u;
    return make_uint4(__funnelshift_r(w0, w1, sh8), __funnelshift_r(w1, w2, sh8), __funnelshift_r(w2, w3, sh8), __funnelshift_r(w3, w4, sh8));
}
// [src, src + n) -> [dst, dst + n), both contiguous, lanes over consecutive 16-byte granules of dst: every load and store
// of the warp falls into a handful of cache lines (a thread-per-read copy touches one line per lane).
// Returns, per lane, whether it saw a byte 0xFF.
template <class FW, class FB>
__device__ __forceinline__ bool warp_copy_x(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, uint32_t n, uint32_t lane, FW xw, FB xb)
{
    bool ff = false;
    uint32_t h = (16u - (uint32_t)((uintptr_t)dst & 15)) & 15u;
    if (h > n) h = n;
    if (lane < h) { const uint8_t b = src[lane]; dst[lane] = xb(b); ff = b == 0xFF; }
    const uint32_t ng = (n - h) >> 4;
    for (uint32_t g = lane; g < ng; g += 32) {
        uint4 v = ld_u128_any(src + h + 16 * g);
        ff = ff || (__vcmpeq4(v.x, 0xFFFFFFFFu) | __vcmpeq4(v.y, 0xFFFFFFFFu) | __vcmpeq4(v.z, 0xFFFFFFFFu) | __vcmpeq4(v.w, 0xFFFFFFFFu)) != 0;
        v.x = xw(v.x); v.y = xw(v.y); v.z = xw(v.z); v.w = xw(v.w);
        *reinterpret_cast<uint4 *>(dst + h + 16 * g) = v;
    }
    const uint32_t t0 = h + 16 * ng;
    if (t0 + lane < n) { const uint8_t b = src[t0 + lane]; dst[t0 + lane] = xb(b); ff = ff || b == 0xFF; }
    return ff;
}
__device__ __forceinline__ bool warp_copy(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, uint32_t n, uint32_t lane)
{
    return warp_copy_x(dst, src, n, lane, IdW(), IdB());
}
// the bases of 32 consecutive reads of one length L (16 <= L <= 8192), read j = bytes [o_j, o_j + L) of `ref`, into the
// contiguous [dst, dst + 32 L): lanes over the aligned 16-byte granules of dst; a granule takes its bytes from one read
// or from the end of one and the start of the next.  o = this lane's o_j; `slack` / `avail` = readable bytes in front of / from ref.
// Called by all 32 lanes.
__device__ __forceinline__ void warp_gather_reads(uint8_t *__restrict__ dst, const uint8_t *__restrict__ ref, uint64_t slack, uint64_t avail,
                                                  uint32_t o, uint32_t L, uint32_t lane)
{
    const uint32_t n = 32 * L;
    const uint32_t h = (16u - (uint32_t)((uintptr_t)dst & 15)) & 15u;            // < 16 <= L: head bytes belong to read 0
    const uint32_t ng = (n - h) >> 4, t0 = h + 16 * ng;                          // tail bytes (< 16) belong to read 31
    if (lane == 0) for (uint32_t i = 0; i < h; i++) dst[i] = ref[o + i];
    if (lane == 31) for (uint32_t i = t0; i < n; i++) dst[i] = ref[o + (i - 31 * L)];
    const uint32_t inv = 0xFFFFFFFFu / L + 1;                                    // floor(x / L) = umulhi(x, inv) for x < 2^32 / L
    for (uint32_t g0 = 0; g0 < ng; g0 += 32) {
        const uint32_t g = g0 + lane;
        const bool valid = g < ng;
        const uint32_t off = h + 16 * (valid ? g : 0);                           // first byte of the granule inside the span
        const uint32_t j = __umulhi(off, inv), within = off - j * L;
        const uint32_t nA = min(16u, L - within);                                // bytes of the granule that come from read j
        const uint32_t oA = __shfl_sync(0xFFFFFFFFu, o, j), oB = __shfl_sync(0xFFFFFFFFu, o, min(j + 1, 31u));
        if (!valid) continue;
        // vector loads may touch up to 19 bytes from their start: stay inside the reference array at both ends
        const bool vec = (uint64_t)oA + within + 20 <= avail && (nA == 16 || ((uint64_t)oB + slack >= nA && (uint64_t)oB + 20 <= avail));
        if (!vec) {
            for (uint32_t i = 0; i < 16; i++) dst[off + i] = i < nA ? ref[oA + within + i] : ref[oB + i - nA];
            continue;
        }
        uint4 v = ld_u128_any(ref + oA + within);
        if (nA < 16) {
            // granule byte i >= nA is byte i - nA of read j + 1: 16 bytes from nA bytes in front of it line up with the granule
            const uint4 b = ld_u128_any(ref + oB - nA);
            const uint32_t m0 = nA >= 4 ? 0xFFFFFFFFu : ((1u << (8 * nA)) - 1u);
            const uint32_t m1 = nA >= 8 ? 0xFFFFFFFFu : (nA <= 4 ? 0u : ((1u << (8 * (nA - 4))) - 1u));
            const uint32_t m2 = nA >= 12 ? 0xFFFFFFFFu : (nA <= 8 ? 0u : ((1u << (8 * (nA - 8))) - 1u));
            const uint32_t m3 = nA <= 12 ? 0u : ((1u << (8 * (nA - 12))) - 1u);
            v.x = (v.x & m0) | (b.x & ~m0); v.y = (v.y & m1) | (b.y & ~m1);
            v.z = (v.z & m2) | (b.z & ~m2); v.w = (v.w & m3) | (b.w & ~m3);
        }
        *reinterpret_cast<uint4 *>(dst + off) = v;
    }
}


// normalizeBase over the reference (windows or whole), 16 bytes per thread: grid (chunks, windows)
__global__ void __launch_bounds__(256) k_ref_norm(const uint8_t *__restrict__ ref, uint8_t *__restrict__ nref,
                                                  const uint64_t *__restrict__ win_off, const uint64_t *__restrict__ win_len,
                                                  uint64_t whole_len)
{
    uint64_t lo, hi;
    if (win_off) { lo = win_off[blockIdx.y]; hi = lo + win_len[blockIdx.y]; } else { lo = 0; hi = whole_len; }
    // 16-byte granules of the absolute offsets, clipped to the window
    const uint64_t g0 = (lo & ~15ull) + ((uint64_t)blockIdx.x * 256 + threadIdx.x) * 16;
    if (g0 >= hi) return;
    if (g0 >= lo && g0 + 16 <= hi && (((uintptr_t)ref | (uintptr_t)nref) & 15) == 0) {
        uint4 v = *reinterpret_cast<const uint4 *>(ref + g0);
        uint32_t *w = &v.x;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t x = w[k], o = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) o |= (uint32_t)norm_base((uint8_t)(x >> (8 * j))) << (8 * j);
            w[k] = o;
        }
        *reinterpret_cast<uint4 *>(nref + g0) = v;
    } else {
        for (uint64_t i = g0 < lo ? lo : g0; i < g0 + 16 && i < hi; i++) nref[i] = norm_base(ref[i]);
    }
}

// ---- k_rec_simple: one THREAD per read ------------------------------------------------------------
// Every read under a quality model without base categories (lossless included): force-preserve is then a per-read
// decision and no sparse Q features occur, so the walk over the CIGAR needs no cross-lane work at all.  Left to the list
// of the general warp-per-read walk: models with base categories, reads that overhang their reference window, CIGARs
// that do not add up to the read length, more than SIMPLE_MAX_OPS operators, unknown operators.
// Neighbouring threads own neighbouring reads: the per-read tables are read and the per-read results written coalesced,
// and M segments are compared as aligned words (one funnel shift re-aligns the reference against the read).
#define SIMPLE_MAX_OPS 12
struct SimpleRead {
    bool live, simple, mapped, forced, walk;      // walk: there is a CIGAR to walk
    bool pb;                                      // model with base categories and a record they apply to: perbase_read
    uint8_t force, rfl, mapq;
    uint16_t flag;
    uint32_t rlen, nops;
    int32_t pos;
    uint64_t lo, c0;
};
__device__ __forceinline__ SimpleRead classify_read(const RecArgs &a, const PolSet &ps, uint64_t r, uint32_t c, bool live)
{
    SimpleRead s;
    s.live = live; s.simple = false; s.mapped = false; s.forced = false; s.walk = false; s.pb = false; s.flag = 0;
    s.force = 0; s.rfl = 0; s.mapq = 0; s.rlen = 0; s.nops = 0; s.pos = 0; s.lo = 0; s.c0 = 0;
    if (!live) return s;
    s.lo = a.read_off[r];
    const uint64_t rlen = a.read_off[r + 1] - s.lo;
    s.c0 = a.cigar_off[r];
    const uint64_t nops = a.cigar_off[r + 1] - s.c0;
    s.pos = a.pos[r];
    const uint16_t flag = a.flag[r];
    s.flag = flag;
    s.mapq = a.mapq[r];
    s.rfl = a.rflags ? a.rflags[r] : 0;
    const bool has_quals = !(s.rfl & CRAMGPU_RF_NO_QUALS);
    s.mapped = !(flag & SAM_UNMAPPED);
    if (rlen >= (1ull << 31)) return s;
    // base categories only matter for records with scores, an alignment start and a CIGAR (QualityScorePreservation :332-340)
    const bool pb = ps.per_base && has_quals && s.pos != 0 && nops > 0;
    bool qposs = false;
    if (ps.n < 0) s.force = has_quals ? 3 : 0;
    else if (has_quals) {
        bool f = false;
        for (int k = 0; k < ps.n; k++) {
            const cramgpu_qs_policy &p = ps.p[k];
            if (!read_in_category(p, flag, s.mapq) || p.n_base_cats) continue;        // (with pb the final word is perbase_read's)
            f = p.treatment != 2;
            qposs = qposs || p.treatment != 2;
        }
        s.force = (f ? 1 : 0) | 2;
    }
    s.forced = s.force & 1;
    s.rlen = (uint32_t)rlen;
    if (!s.mapped || s.pos == 0) { s.simple = !pb; return s; }
    if (!pb && !s.forced && qposs) return s;                            // sparse Q features from read-level policies alone: general walk
    if (nops == 0) { s.simple = true; return s; }
    if (nops > SIMPLE_MAX_OPS || rlen == 0 || s.pos < 0) return s;
    uint64_t nread = 0, nref = 0;
    for (uint32_t i = 0; i < (uint32_t)nops; i++) {
        const uint32_t cw = a.cigar[s.c0 + i], op = cw & 15, len = cw >> 4;
        if (op > OP_X) return s;
        if (consumes_read(op)) nread += len;
        if (consumes_ref(op)) nref += len;
    }
    if (nread != rlen || (uint64_t)(s.pos - 1) + nref > a.ref_len[c]) return s;   // odd CIGAR / overhang ('N' fill): general walk
    s.simple = true; s.walk = true; s.nops = (uint32_t)nops; s.pb = pb;
    return s;
}

__device__ __forceinline__ void rec_list_append(uint64_t *lr, uint32_t *lc, unsigned long long *ln, bool put, uint64_t r, uint32_t c, uint32_t lane)
{
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, put);
    if (!m) return;
    unsigned long long base = 0;
    if (lane == (uint32_t)(__ffs(m) - 1)) base = atomicAdd(ln, (unsigned long long)__popc(m));      // one atomic per warp
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
    if (put) { const uint64_t k = base + __popc(m & ((1u << lane) - 1u)); lr[k] = r; lc[k] = c; }
}

template <bool WRITE>
__device__ __forceinline__ uint32_t simple_read(const RecArgs &a, const PolSet &ps, const SimpleRead &s, const uint64_t r, const uint32_t c);
template <bool WRITE>
__device__ __forceinline__ uint32_t perbase_read(const RecArgs &a, const PolSet &ps, const SimpleRead &s, const uint64_t r, const uint32_t c);
// PB: the model has base categories (a separate instantiation of the kernels, so that the others do not carry its registers)
template <bool WRITE, bool PB>
__device__ __forceinline__ uint32_t thread_read(const RecArgs &a, const PolSet &ps, const SimpleRead &s, const uint64_t r, const uint32_t c)
{
    if (PB && s.pb) return perbase_read<WRITE>(a, ps, s, r, c);
    return simple_read<WRITE>(a, ps, s, r, c);
}

// all reads in order; reads with several CIGAR operators are only classified here and done by k_rec_multi, so that the
// lanes of a warp do the same kind of work
template <bool WRITE, bool PB>
__global__ void __launch_bounds__(128) k_rec_simple(RecArgs a, PolSet ps)
{
    const uint32_t lane = threadIdx.x & 31;
    uint64_t r; uint32_t c; bool live;
    if (gridDim.y > 1 || a.n_containers == 1) {
        c = blockIdx.y;
        r = a.cont_off[c] + (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.cont_off[c + 1];
    } else {
        r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.n_reads;
        c = live ? container_of(a.cont_off, a.n_containers, r) : 0;
    }
    const SimpleRead s = classify_read(a, ps, r, c, live);
    const bool multi = s.simple && s.nops > 1;
    if (!WRITE) {
        rec_list_append(a.clist_r, a.clist_c, a.clist_n, live && !s.simple, r, c, lane);
        rec_list_append(a.tlist_r, a.tlist_c, a.tlist_n, multi, r, c, lane);
        uint32_t nf = 0;
        if (s.simple && !multi) nf = thread_read<false, PB>(a, ps, s, r, c);
        // the write pass of a read with features is done from a list as well: most reads have none, and a warp of the
        // in-order kernel would otherwise wait for its few that do
        rec_list_append(a.wlist_r, a.wlist_c, a.wlist_n, nf != 0, r, c, lane);
    } else if (s.simple && !multi) {
        const uint32_t here = a.cnt[(size_t)C_NFEAT * a.n_reads + r];
        const uint64_t next = (r + 1 < a.cont_off[c + 1]) ? (uint64_t)a.cnt[(size_t)C_NFEAT * a.n_reads + r + 1] : a.totals[(size_t)c * NCNT + C_NFEAT];
        if (next == here) thread_read<true, PB>(a, ps, s, r, c);
    }
    if (WRITE && ps.n < 0) {
        // lossless: the preserved array is the record's own qualities [Writer.java:250-251,257-258] (multi-operator reads
        // included).  The reads of a warp lie back to back in `quals`; when their arrays do so in QS too, one warp-wide copy
        const bool cp = s.simple && s.forced;
        const uint64_t qs = cp ? a.base[(size_t)c * NCNT + C_QS] + a.cnt[(size_t)C_QS * a.n_reads + r] : 0;
        const uint64_t qs0 = __shfl_sync(0xFFFFFFFFu, qs, 0), lo0 = __shfl_sync(0xFFFFFFFFu, s.lo, 0);
        const uint32_t L0 = __shfl_sync(0xFFFFFFFFu, s.rlen, 0);
        const bool flat = cp && s.rlen == L0 && qs == qs0 + (uint64_t)lane * L0 && s.lo == lo0 + (uint64_t)lane * L0;
        if (__all_sync(0xFFFFFFFFu, flat) && L0 <= (1u << 20)) warp_copy(a.streams + qs0, a.quals + lo0, 32 * L0, lane);
        else if (cp) copy_bytes(a.streams + qs, a.quals + s.lo, s.rlen, IdW(), IdB());
    }
}

template <bool WRITE, bool PB>
__global__ void __launch_bounds__(128) k_rec_multi(RecArgs a, PolSet ps, const uint64_t *__restrict__ lr, const uint32_t *__restrict__ lc,
                                                   const unsigned long long *__restrict__ ln)
{
    const uint64_t n = *ln, step = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
        const uint64_t r = lr[i];
        const uint32_t c = lc[i];
        thread_read<WRITE, PB>(a, ps, classify_read(a, ps, r, c, true), r, c);
    }
}

// ---- quality models with base categories, one thread per read --------------------------------------
// The per-base part of QualityScorePreservation.addQualityScores / addQS [:332-489] works on masks over the bases of a
// record; here a mask is a machine word, one bit per base of a group of up to four bases (the four bytes of an aligned
// word of the record's bases), and the policies are evaluated with bitwise operations on all bases of the group at once.
//   match / ins / dm: base equals the RAW reference base at its position / lies in an insertion / is at or right after
//   the start of one of the last two deletions.
// mask of one per-base policy over a group (base_mask above, bit-parallel)
// one bit per byte of a word: bit j = byte j of v is non-zero in its low bit (v holds 0x00 / 0xFF per byte)
__device__ __forceinline__ uint32_t nibble_of(uint32_t v) { return (((v & 0x01010101u) * 0x01020408u) >> 24) & 0xFu; }
// ext: the group's four ext bytes (byte j = base j), eb: the policy's row of PolSet::ext_bit
__device__ __forceinline__ uint32_t pb_mask(const cramgpu_qs_policy &p, uint32_t match, uint32_t ins, uint32_t dm, uint32_t ext, const uint8_t *eb)
{
    uint32_t m = 0;
    for (int b = 0; b < p.n_base_cats; b++) {
        switch (p.base_cat[b] & 0xFF) {
        case 1: m = match; break;            // MATCH: assignment
        case 2: m = ~match; break;           // MISMATCH
        case 3: m |= dm; break;              // FLANKING_DELETION: only sets
        case 4: m |= ins; break;             // INSERTION: only sets
        case 5: { const uint32_t xv = nibble_of(ext); m = (m & ~xv) | (nibble_of(ext >> eb[b]) & xv); } break;   // LOWER_COVERAGE: assignment where valid
        case 6: m |= nibble_of(ext >> eb[b]); break;                                                              // PILEUP: only sets
        }
    }
    return m;
}
// score_at above, bit-parallel: `has` = bases that end up with a score, `bin` = those whose score is binned.
// appl: bit k = policy k applies to the record (its read category)
// binany = bases that any applicable policy bins, overridden later or not (those are the ones the reference indexes the
// binning table with)
__device__ __forceinline__ void pb_planes(const PolSet &ps, uint32_t appl, uint32_t match, uint32_t ins, uint32_t dm, uint32_t ext, uint32_t &has,
                                          uint32_t &bin, uint32_t &binany)
{
    has = 0; bin = 0; binany = 0;
    for (int k = 0; k < ps.n; k++) {
        if (!((appl >> k) & 1)) continue;
        const cramgpu_qs_policy &p = ps.p[k];
        uint32_t hit;
        if (p.n_base_cats == 0) { if (p.treatment == 2) continue; hit = ~0u; }
        else hit = pb_mask(p, match, ins, dm, ext, ps.ext_bit[k]);
        has = (has & ~hit) | (p.treatment != 2 ? hit : 0u);
        bin = (bin & ~hit) | (p.treatment == 1 ? hit : 0u);
        binany |= p.treatment == 1 ? hit : 0u;
    }
}
// bits of the deletion-flank positions d_a, d_a + 1, d_b, d_b + 1 inside the group that starts at read offset z0
__device__ __forceinline__ uint32_t pb_dm(int32_t z0, int32_t d_a, int32_t d_b)
{
    uint32_t m = 0;
    if (d_a < 0) return 0;
    int32_t o = d_a - z0; if (o >= 0 && o < 4) m |= 1u << o;
    o++; if (o >= 0 && o < 4) m |= 1u << o;
    o = d_b - z0; if (o >= 0 && o < 4 && d_b >= 0) m |= 1u << o;
    o++; if (o >= 0 && o < 4 && d_b >= 0) m |= 1u << o;
    return m;
}

// a group of up to four bases of one read-consuming operator
struct PbGroup {
    int32_t z0;          // read offset of bit 0 (negative by up to 3 for the first word of an M segment)
    int32_t r0;          // reference offset (from the alignment start) that goes with bit 0
    uint32_t valid;      // bits that are bases of this operator
    uint32_t match, ins, dm;
    uint32_t op, len;    // the operator and its length
    uint32_t rel;        // offset of bit 0 inside the operator (I and S: the base's index)
    uint32_t av;         // the bases: aligned word (M / = / X) or one byte (I, S)
    uint32_t ext;        // the bases' ext bytes (X / P selectors), byte j = bit j; 0 without
};
// Walks the operators of a record: visit(PbGroup) for the read-consuming ones in groups of up to four bases, other(op, len, z)
// for the rest (and for an empty soft clip).  M / = / X segments go by aligned words of the bases, I and S base by base
// (their "reference position" is where the walk stands, [:398-476], and may lie beyond the window).
template <class V, class O>
__device__ __forceinline__ void pb_walk(const RecArgs &a, uint64_t lo, uint64_t c0, uint32_t nops, int32_t pos, uint32_t c, V visit, O other)
{
    const uint8_t *rb = a.bases + lo;
    const uint8_t *ref = a.ref + a.ref_off[c];
    const int64_t ref_len = (int64_t)a.ref_len[c];
    uint32_t z = 0, aoff = 0;
    int32_t d_a = -9, d_b = -9;
    for (uint32_t ci = 0; ci < nops; ci++) {
        const uint32_t cw = a.cigar[c0 + ci], op = cw & 15, len = cw >> 4;
        if (op == OP_D) { if (d_a != (int32_t)z) { d_b = d_a; d_a = (int32_t)z; } }
        PbGroup g;
        g.op = op; g.len = len;
        if (!consumes_read(op) || (op == OP_S && len == 0)) other(op, len, z);
        else if (op == OP_I || op == OP_S) {
            for (uint32_t i = 0; i < len; i++) {
                const int64_t rp = (int64_t)pos - 1 + aoff + i;
                const uint8_t b = rb[z + i], tb = (rp >= 0 && rp < ref_len) ? ref[rp] : (uint8_t)'N';       // baseAt beyond the window's end [ReferenceTracks.java:141-146]
                g.z0 = (int32_t)(z + i); g.r0 = (int32_t)(aoff + i); g.valid = 1u; g.match = b == tb ? 1u : 0u; g.ins = op == OP_I ? 1u : 0u;
                g.ext = a.qs_ext ? a.qs_ext[lo + z + i] : 0u;
                g.dm = pb_dm(g.z0, d_a, d_b) & 1u; g.rel = i; g.av = b;
                visit(g);
            }
        } else if (len) {
            const uint8_t *A = rb + z;
            const uint32_t sa = (uint32_t)((uintptr_t)A & 3);
            const uint32_t *aw = reinterpret_cast<const uint32_t *>(A - sa);
            const uint8_t *Rs = ref + (int64_t)pos - 1 + aoff;                 // raw reference at the segment's first base (inside the window)
            const uint32_t nwords = (sa + len + 3) >> 2, tail = (sa + len) & 3;
            for (uint32_t k = 0; k < nwords; k++) {
                g.av = aw[k];
                g.valid = 0xFu;
                if (k == 0) g.valid &= 0xFu << sa;
                if (k == nwords - 1 && tail) g.valid &= (1u << tail) - 1u;
                // the raw reference bytes of the word's valid bases (bytes in front of / behind the segment are not touched)
                uint32_t rv = 0;
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if ((g.valid >> j) & 1) rv |= (uint32_t)Rs[(int64_t)4 * k + j - sa] << (8 * j);
                g.z0 = (int32_t)(z + 4 * k) - (int32_t)sa; g.r0 = (int32_t)(aoff + 4 * k) - (int32_t)sa;
                g.match = nibble_of(__vcmpeq4(g.av, rv)) & g.valid; g.ins = 0; g.dm = pb_dm(g.z0, d_a, d_b); g.rel = 4 * k - sa;
                g.ext = 0;
                if (a.qs_ext) {
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if ((g.valid >> j) & 1) g.ext |= (uint32_t)a.qs_ext[lo + z + 4 * k + j - sa] << (8 * j);
                }
                visit(g);
            }
        }
        if (consumes_read(op)) z += len;
        if (consumes_ref(op)) aoff += len;
    }
}

// force-preserve decision of a record under a model with base categories: the safety latch [:487-488] needs the number of
// bases every per-base policy masks
__device__ __forceinline__ uint8_t pb_force(const RecArgs &a, const PolSet &ps, uint64_t lo, uint64_t c0, uint32_t nops, int32_t pos, uint32_t rlen,
                                            uint32_t c, uint32_t appl)
{
    uint32_t masked[MAX_POL];
#pragma unroll
    for (int k = 0; k < MAX_POL; k++) masked[k] = 0;
    pb_walk(a, lo, c0, nops, pos, c,
            [&](const PbGroup &g) {
#pragma unroll
                for (int k = 0; k < MAX_POL; k++)
                    if (k < ps.n && ps.p[k].n_base_cats) masked[k] += __popc(pb_mask(ps.p[k], g.match, g.ins, g.dm, g.ext, ps.ext_bit[k]) & g.valid);
            },
            [&](uint32_t, uint32_t, uint32_t) {});
    bool f = false;
#pragma unroll
    for (int k = 0; k < MAX_POL; k++) {
        if (k >= ps.n || !((appl >> k) & 1)) continue;
        if (ps.p[k].n_base_cats == 0) f = ps.p[k].treatment != 2;                  // BIN / PRESERVE force, DROP clears
        else if (masked[k] > rlen / 2) f = true;
    }
    return (uint8_t)((f ? 1 : 0) | 2);
}

// A record with quality scores, an alignment start and a regular CIGAR under a model with base categories, by one thread:
// force decision, features merged with the sparse Q features in position order (a base's own feature first), series
// bytes.  Same results as walk_read.  -> number of read features
template <bool WRITE>
__device__ __forceinline__ uint32_t perbase_read(const RecArgs &a, const PolSet &ps, const SimpleRead &s, const uint64_t r, const uint32_t c)
{
    const uint64_t N = a.n_reads;
    const uint32_t rlen = s.rlen;
    uint32_t appl = 0;
    for (int k = 0; k < ps.n; k++) if (read_in_category(ps.p[k], s.flag, s.mapq)) appl |= 1u << k;
    const uint8_t force = WRITE ? a.force[r] : pb_force(a, ps, s.lo, s.c0, s.nops, s.pos, rlen, c, appl);
    const bool forced = force & 1, qf = !forced;
    uint32_t n_feat = 0, n_fp = 0, n_bs = 0, n_dl = 0, n_ba = 0, n_sc = 0, n_qs = 0, n_rs = 0, n_hc = 0, n_pd = 0, n_pay = 0;
    const uint64_t *B = WRITE ? a.base + (size_t)c * NCNT : nullptr;
    uint64_t fbase = 0, paybase = 0;
    uint8_t *sFC = nullptr, *sFP = nullptr, *sBS = nullptr, *sDL = nullptr, *sBA = nullptr, *sSC = nullptr, *sQS = nullptr, *sRS = nullptr,
            *sHC = nullptr, *sPD = nullptr;
    const uint8_t *codes = nullptr;
    uint32_t nf = 0;
    if (WRITE) {
        auto at = [&](int k) -> uint64_t { return B[k] + a.cnt[(size_t)k * N + r]; };
        const uint32_t here = a.cnt[(size_t)C_NFEAT * N + r];
        const uint64_t next = (r + 1 < a.cont_off[c + 1]) ? (uint64_t)a.cnt[(size_t)C_NFEAT * N + r + 1] : a.totals[(size_t)c * NCNT + C_NFEAT];
        nf = (uint32_t)(next - here);
        fbase = B[C_NFEAT] + here;
        a.feat_off[r] = fbase;
        a.read_ba_off[r] = ~0ull;
        itf8_put(a.streams + at(C_FN), nf);
        itf8_put(a.streams + at(C_MQ), s.mapq);
        sFC = a.streams + at(C_FC); sFP = a.streams + at(C_FP); sBS = a.streams + at(C_BS); sDL = a.streams + at(C_DL);
        sBA = a.streams + at(C_BA); sSC = a.streams + at(C_SC); sQS = a.streams + at(C_QS); sRS = a.streams + at(C_RS);
        sHC = a.streams + at(C_HC); sPD = a.streams + at(C_PD); paybase = at(C_PAYLOAD);
        codes = a.codes + (size_t)c * 32;
    }
    if (!WRITE || nf || s.nops > 1) {
        const uint8_t *rb = a.bases + s.lo, *rq = a.quals + s.lo;
        const uint8_t *nref = a.nref + a.ref_off[c] + (uint64_t)(s.pos - 1);
        int32_t prev = 0;
        auto feature = [&](uint8_t opc, int32_t fpos, int32_t val, int32_t aux) {
            const uint32_t d = (uint32_t)(fpos - prev);
            prev = fpos;
            if (WRITE) {
                const uint64_t fi = fbase + n_feat;
                a.f_op[fi] = opc; a.f_pos[fi] = fpos; a.f_val[fi] = val; a.f_aux[fi] = aux;
                sFC[n_feat] = opc;
                itf8_put(sFP + n_fp, d);
            }
            n_feat++; n_fp += itf8_len(d);
        };
        pb_walk(a, s.lo, s.c0, s.nops, s.pos, c,
                [&](const PbGroup &g) {
                    const uint32_t op = g.op, av = g.av, valid = g.valid;
                    const int32_t z0 = g.z0;
                    const bool isM = op != OP_I && op != OP_S;
                    uint32_t fbits = 0, rv = 0;
                    if (isM) {
                        // normalised reference bytes of the valid bases: substitutions [:306-330]
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if ((valid >> j) & 1) rv |= (uint32_t)nref[g.r0 + j] << (8 * j);
                        fbits = ~nibble_of(__vcmpeq4(av, rv)) & valid;
                    } else if (op == OP_I) fbits = 1;
                    else fbits = g.rel == 0 ? 1u : 0u;                            // S: the clip's feature sits at its first base
                    uint32_t has = 0, bin = 0;
                    if (qf) {
                        uint32_t binany;
                        pb_planes(ps, appl, g.match, g.ins, g.dm, g.ext, has, bin, binany);
                        has &= valid;
                        if (ps.any_bin)
                            for (uint32_t m = binany & valid; m; m &= m - 1)
                                if (rq[z0 + __ffs(m) - 1] >= BIN_FAULT_Q) bin_fault(a.status);
                    }
                    for (uint32_t todo = fbits | has; todo; todo &= todo - 1) {
                        const uint32_t j = (uint32_t)__ffs(todo) - 1;
                        const int32_t zi = z0 + (int32_t)j;
                        if ((fbits >> j) & 1) {
                            if (isM) {
                                const uint8_t b = (uint8_t)(av >> (8 * j)), refb = (uint8_t)(rv >> (8 * j));
                                const int ri = base_index(refb), bi = base_index(b);
                                uint8_t code = 0;
                                if (!WRITE) { if (ri >= 0 && bi >= 0) atomicAdd(&a.subst[(size_t)c * 25 + ri * 5 + bi], 1ull); }
                                else { code = (ri >= 0 && bi >= 0) ? codes[ri * 5 + bi] : 0; sBS[n_bs] = code; }
                                feature('X', zi + 1, code, (int32_t)b | ((int32_t)refb << 8));
                                n_bs++;
                            } else if (op == OP_I) {
                                const uint8_t b = (uint8_t)av;
                                if (WRITE) sBA[n_ba] = b;
                                feature('i', zi + 1, b, 0);
                                n_ba++;
                            } else {
                                // SoftClip(pos, bases[pos .. pos+len)) [:268-272]
                                const uint32_t len = g.len;
                                if (WRITE) {
                                    for (uint32_t i = 0; i < len; i++) { const uint8_t b = rb[zi + i]; sSC[n_sc + i] = b; a.payload[paybase + n_pay + i] = b; }
                                    sSC[n_sc + len] = '\t';
                                }
                                feature('S', zi + 1, (int32_t)(paybase + n_pay), (int32_t)len);
                                n_sc += len + 1; n_pay += len;
                            }
                        }
                        if ((has >> j) & 1) {
                            const uint8_t q = rq[zi], sc = ((bin >> j) & 1) ? illumina_bin(q) : q;
                            if (sc < 128) {                                          // scores[i] > -1 [:290]
                                if (WRITE) sQS[n_qs] = sc;
                                feature('Q', zi + 1, sc, 0);
                                n_qs++;
                            }
                        }
                    }
                },
                [&](uint32_t op, uint32_t len, uint32_t z) {
                    if (op == OP_S) { if (WRITE) sSC[n_sc] = '\t'; n_sc++; return; }       // empty clip: only its terminator
                    // D, N, H, P: one feature at the next read position [:228-240]
                    const uint32_t vl = itf8_len(len);
                    if (WRITE) itf8_put(op == OP_D ? sDL + n_dl : op == OP_N ? sRS + n_rs : op == OP_H ? sHC + n_hc : sPD + n_pd, len);
                    feature(op == OP_D ? 'D' : op == OP_N ? 'N' : op == OP_H ? 'H' : 'P', (int32_t)z + 1, (int32_t)len, 0);
                    if (op == OP_D) n_dl += vl; else if (op == OP_N) n_rs += vl; else if (op == OP_H) n_hc += vl; else n_pd += vl;
                });
    }
    if (WRITE) {
        a.read_qs_off[r] = forced ? (uint64_t)(sQS + n_qs - a.streams) : ~0ull;     // the array itself: k_rec_forced_scores
    } else {
        a.force[r] = force;
        uint32_t *cn = a.cnt + r;
        cn[C_FN * N] = itf8_len(n_feat); cn[C_MQ * N] = itf8_len(s.mapq);
        if (forced) n_qs += rlen;
        if (n_qs) cn[C_QS * N] = n_qs;
        if (n_feat) { cn[C_FC * N] = n_feat; cn[C_FP * N] = n_fp; cn[C_NFEAT * N] = n_feat; }
        if (n_bs) cn[C_BS * N] = n_bs;
        if (n_ba) cn[C_BA * N] = n_ba;
        if (n_dl) cn[C_DL * N] = n_dl;
        if (n_sc) { cn[C_SC * N] = n_sc; cn[C_PAYLOAD * N] = n_pay; }
        if (n_rs) cn[C_RS * N] = n_rs;
        if (n_hc) cn[C_HC * N] = n_hc;
        if (n_pd) cn[C_PD * N] = n_pd;
    }
    return n_feat;
}

// -> number of read features
template <bool WRITE>
__device__ __forceinline__ uint32_t simple_read(const RecArgs &a, const PolSet &ps, const SimpleRead &s, const uint64_t r, const uint32_t c)
{
    const uint64_t N = a.n_reads;
    const uint32_t rlen = s.rlen;
    uint32_t n_feat = 0, n_fp = 0, n_bs = 0, n_dl = 0, n_ba = 0, n_sc = 0, n_rs = 0, n_hc = 0, n_pd = 0, n_pay = 0;
    // write-pass bases
    const uint64_t *B = WRITE ? a.base + (size_t)c * NCNT : nullptr;
    uint64_t fbase = 0, paybase = 0;
    uint32_t nf = 0;
    uint8_t *sFC = nullptr, *sFP = nullptr, *sBS = nullptr, *sDL = nullptr, *sBA = nullptr, *sSC = nullptr, *sRS = nullptr,
            *sHC = nullptr, *sPD = nullptr;
    const uint8_t *codes = nullptr;
    if (WRITE) {
        auto at = [&](int k) -> uint64_t { return B[k] + a.cnt[(size_t)k * N + r]; };
        const uint32_t here = a.cnt[(size_t)C_NFEAT * N + r];
        const uint64_t next = (r + 1 < a.cont_off[c + 1]) ? (uint64_t)a.cnt[(size_t)C_NFEAT * N + r + 1] : a.totals[(size_t)c * NCNT + C_NFEAT];
        nf = (uint32_t)(next - here);
        fbase = B[C_NFEAT] + here;
        a.feat_off[r] = fbase;
        a.read_qs_off[r] = s.forced ? at(C_QS) : ~0ull;
        const bool ba = !s.mapped && !(s.rfl & CRAMGPU_RF_NO_BASES);
        const uint64_t bao = ba ? at(C_BA) : ~0ull;
        a.read_ba_off[r] = bao;
        if (ba) copy_bytes(a.streams + bao, a.bases + s.lo, rlen, IdW(), IdB());          // [Writer.java:254-256]
        if (s.mapped) {
            itf8_put(a.streams + at(C_FN), nf);
            itf8_put(a.streams + at(C_MQ), s.mapq);
        }
        if (nf || s.nops > 1) {
            sFC = a.streams + at(C_FC); sFP = a.streams + at(C_FP); sBS = a.streams + at(C_BS); sDL = a.streams + at(C_DL);
            sBA = a.streams + at(C_BA); sSC = a.streams + at(C_SC); sRS = a.streams + at(C_RS); sHC = a.streams + at(C_HC);
            sPD = a.streams + at(C_PD); paybase = at(C_PAYLOAD);
            codes = a.codes + (size_t)c * 32;
        }
    }

    if (s.walk && (!WRITE || nf || s.nops > 1)) {
        // createVariations [Sam2CramRecordFactory.java:205-296], one operator after the other
        const uint8_t *rb = a.bases + s.lo;
        const uint8_t *nref = a.nref + a.ref_off[c] + (uint64_t)(s.pos - 1);          // nref has slack on both sides
        uint32_t z = 0, aoff = 0;
        int32_t prev = 0;
        // a feature without series bytes of its own
        auto feature = [&](uint8_t opc, int32_t fpos, int32_t val, int32_t aux) {
            const uint32_t d = (uint32_t)(fpos - prev);
            prev = fpos;
            if (WRITE) {
                const uint64_t fi = fbase + n_feat;
                a.f_op[fi] = opc; a.f_pos[fi] = fpos; a.f_val[fi] = val; a.f_aux[fi] = aux;
                sFC[n_feat] = opc;
                itf8_put(sFP + n_fp, d);
            }
            n_feat++; n_fp += itf8_len(d);
        };
        for (uint32_t ci = 0; ci < s.nops; ci++) {
            const uint32_t cw = a.cigar[s.c0 + ci], op = cw & 15, len = cw >> 4;
            if (!consumes_read(op)) {
                // D, N, H, P: one feature at the next read position [:228-240]
                const uint32_t vl = itf8_len(len);
                if (WRITE) itf8_put(op == OP_D ? sDL + n_dl : op == OP_N ? sRS + n_rs : op == OP_H ? sHC + n_hc : sPD + n_pd, len);
                feature(op == OP_D ? 'D' : op == OP_N ? 'N' : op == OP_H ? 'H' : 'P', (int32_t)z + 1, (int32_t)len, 0);
                if (op == OP_D) n_dl += vl; else if (op == OP_N) n_rs += vl; else if (op == OP_H) n_hc += vl; else n_pd += vl;
            } else if (op == OP_I) {
                // one 'i' feature per inserted base [:262-266]
                for (uint32_t i = 0; i < len; i++) {
                    const uint8_t b = rb[z + i];
                    if (WRITE) sBA[n_ba] = b;
                    feature('i', (int32_t)(z + i) + 1, b, 0);
                    n_ba++;
                }
            } else if (op == OP_S) {
                // SoftClip(pos, bases[pos .. pos+len)) [:268-272]; bytes + TAB into SC, bytes into the payload
                if (len) {
                    if (WRITE) {
                        for (uint32_t i = 0; i < len; i++) { const uint8_t b = rb[z + i]; sSC[n_sc + i] = b; a.payload[paybase + n_pay + i] = b; }
                    }
                    feature('S', (int32_t)z + 1, (int32_t)(paybase + n_pay), (int32_t)len);
                }
                if (WRITE) sSC[n_sc + len] = '\t';
                n_sc += len + 1; n_pay += len;
            } else if (len) {
                // M, =, X: read words (aligned) against the reference bytes at the same read offsets
                const uint8_t *A = rb + z;
                const uint32_t sa = (uint32_t)((uintptr_t)A & 3);
                const uint32_t *aw = reinterpret_cast<const uint32_t *>(A - sa);
                const uint8_t *Bs = nref + aoff - sa;
                const uint32_t sb = (uint32_t)((uintptr_t)Bs & 3), sb8 = sb * 8;
                const uint32_t *bw = reinterpret_cast<const uint32_t *>(Bs - sb);
                const uint32_t nwords = (sa + len + 3) >> 2, tail = (sa + len) & 3;
                uint32_t bprev = bw[0];
                for (uint32_t k = 0; k < nwords; k++) {
                    const uint32_t av = aw[k], bnext = bw[k + 1];
                    const uint32_t bv = __funnelshift_r(bprev, bnext, sb8);
                    bprev = bnext;
                    uint32_t x = av ^ bv;
                    if (k == 0) x &= 0xFFFFFFFFu << (8 * sa);
                    if (k == nwords - 1 && tail) x &= (1u << (8 * tail)) - 1u;
                    while (x) {                                              // a substitution per differing byte [:306-330]
                        const uint32_t j = (uint32_t)(__ffs(x) - 1) >> 3;
                        x &= ~(0xFFu << (8 * j));
                        const uint8_t b = (uint8_t)(av >> (8 * j)), refb = (uint8_t)(bv >> (8 * j));
                        const int ri = base_index(refb), bi = base_index(b);
                        uint8_t code = 0;
                        if (!WRITE) {
                            if (ri >= 0 && bi >= 0) atomicAdd(&a.subst[(size_t)c * 25 + ri * 5 + bi], 1ull);
                        } else {
                            code = (ri >= 0 && bi >= 0) ? codes[ri * 5 + bi] : 0;           // codes[][] default 0 [:39]
                            sBS[n_bs] = code;
                        }
                        feature('X', (int32_t)(z + 4 * k + j - sa) + 1, code, (int32_t)b | ((int32_t)refb << 8));
                        n_bs++;
                    }
                }
            }
            if (consumes_read(op)) z += len;
            if (consumes_ref(op)) aoff += len;
        }
    }

    if (!WRITE) {
        a.force[r] = s.force;
        uint32_t *cn = a.cnt + r;
        if (s.mapped) { cn[C_FN * N] = itf8_len(n_feat); cn[C_MQ * N] = itf8_len(s.mapq); }
        else if (!(s.rfl & CRAMGPU_RF_NO_BASES) && rlen) n_ba = rlen;
        if (s.forced && rlen) cn[C_QS * N] = rlen;
        if (n_feat) { cn[C_FC * N] = n_feat; cn[C_FP * N] = n_fp; cn[C_NFEAT * N] = n_feat; }
        if (n_bs) cn[C_BS * N] = n_bs;
        if (n_ba) cn[C_BA * N] = n_ba;
        if (n_dl) cn[C_DL * N] = n_dl;
        if (n_sc) { cn[C_SC * N] = n_sc; cn[C_PAYLOAD * N] = n_pay; }
        if (n_rs) cn[C_RS * N] = n_rs;
        if (n_hc) cn[C_HC * N] = n_hc;
        if (n_pd) cn[C_PD * N] = n_pd;
    }
    return n_feat;
}

// the general walk of one read by one warp (any CIGAR, any policy)
template <bool WRITE>
__device__ __forceinline__ void walk_read(const RecArgs &a, const PolSet &ps, const uint64_t r, const uint32_t c, const uint32_t lane)
{
    const uint8_t *ref = a.ref + a.ref_off[c];
    const int64_t ref_len = (int64_t)a.ref_len[c];
    const uint64_t lo = a.read_off[r];
    const int64_t rlen = (int64_t)(a.read_off[r + 1] - lo);
    const uint8_t *rb = a.bases + lo, *rq = a.quals + lo;
    const uint64_t c0 = a.cigar_off[r], c1 = a.cigar_off[r + 1];
    const int32_t pos = a.pos[r];
    const uint16_t flag = a.flag[r];
    const uint8_t mapq = a.mapq[r];
    const uint8_t rfl = a.rflags ? a.rflags[r] : 0;
    const bool has_quals = !(rfl & CRAMGPU_RF_NO_QUALS);
    const bool mapped = !(flag & SAM_UNMAPPED);
    const bool hasfeat = mapped && pos != 0;

    // ---- force-preserve flag (count pass); bit1 = record carries a score array
    uint8_t force;
    if (!WRITE) {
        if (ps.n < 0) force = has_quals ? 3 : 0;                                   // CramSerilization.java:160-161
        else if (!has_quals) force = 0;                                            // QualityScorePreservation.java:277-281
        else {
            bool f = false;
            uint32_t masked[MAX_POL];
            for (int k = 0; k < MAX_POL; k++) masked[k] = 0;
            if (ps.per_base && pos != 0) {
                int64_t z = 0, aoff = 0, d_a = -9, d_b = -9;
                for (uint64_t ci = c0; ci < c1; ci++) {
                    uint32_t cw = a.cigar[ci]; int op = cw & 15; int64_t len = cw >> 4;
                    if (op == OP_D) { if (d_a != z) { d_b = d_a; d_a = z; } }
                    if (consumes_read(op)) {
                        for (int64_t i0 = 0; i0 < len; i0 += 32) {
                            int64_t i = i0 + lane;
                            bool in = i < len && z + i < rlen;
                            int64_t zi = z + i;
                            uint8_t b = in ? rb[zi] : 0;
                            int64_t rp = (int64_t)pos - 1 + aoff + i;
                            uint8_t tb = (in && rp >= 0 && rp < ref_len) ? ref[rp] : (uint8_t)'N';   // tracks.baseAt(refPos + i)
                            bool dm = zi == d_a || zi == d_a + 1 || zi == d_b || zi == d_b + 1;
                            const uint32_t ext = (in && a.qs_ext) ? a.qs_ext[lo + zi] : 0u;
                            for (int k = 0; k < ps.n; k++)
                                if (ps.p[k].n_base_cats) {
                                    bool m = in && base_mask(ps.p[k], b == tb, op == OP_I, dm, ext, ps.ext_bit[k]);
                                    masked[k] += __popc(__ballot_sync(0xFFFFFFFFu, m));
                                }
                        }
                        z += len;
                    }
                    if (consumes_ref(op)) aoff += len;
                }
            }
            for (int k = 0; k < ps.n; k++) {
                const cramgpu_qs_policy &p = ps.p[k];
                if (!read_in_category(p, flag, mapq)) continue;
                if (p.n_base_cats == 0) f = p.treatment != 2;                      // BIN / PRESERVE force, DROP clears
                else if (pos != 0 && (int64_t)masked[k] > rlen / 2) f = true;      // safety latch [:487-488]
            }
            force = (f ? 1 : 0) | 2;
        }
        if (lane == 0) a.force[r] = force;
    } else {
        force = a.force[r];
    }
    const bool forced = force & 1;
    const bool q_feats = !forced && (force & 2) && ps.n >= 0;        // sparse Q features possible

    // ---- running series cursors (uniform across the warp)
    uint32_t n_feat = 0, n_fp = 0, n_bs = 0, n_dl = 0, n_ba = 0, n_sc = 0, n_qs = 0, n_rs = 0, n_hc = 0, n_pd = 0, n_pay = 0;
    int32_t prev = 0;
    // write-pass bases
    uint64_t fbase = 0, paybase = 0;
    uint8_t *sFC = nullptr, *sFP = nullptr, *sBS = nullptr, *sDL = nullptr, *sBA = nullptr, *sSC = nullptr, *sQS = nullptr,
            *sRS = nullptr, *sHC = nullptr, *sPD = nullptr;
    if (WRITE) {
        const uint64_t *B = a.base + (size_t)c * NCNT;
        auto at = [&](int k) -> uint64_t { return B[k] + a.cnt[(size_t)k * a.n_reads + r]; };
        fbase = at(C_NFEAT);
        sQS = a.streams + at(C_QS);
        // the other series are rare: fetch their offsets only for reads that have features (or unmapped bases) at all
        const uint32_t nf_here = a.cnt[(size_t)C_NFEAT * a.n_reads + r];
        const uint64_t nf_next = (r + 1 < a.cont_off[c + 1]) ? (uint64_t)a.cnt[(size_t)C_NFEAT * a.n_reads + r + 1] : ~0ull;
        if (nf_next != nf_here || !mapped) {
            paybase = at(C_PAYLOAD);
            sFC = a.streams + at(C_FC); sFP = a.streams + at(C_FP); sBS = a.streams + at(C_BS); sDL = a.streams + at(C_DL);
            sBA = a.streams + at(C_BA); sSC = a.streams + at(C_SC); sRS = a.streams + at(C_RS);
            sHC = a.streams + at(C_HC); sPD = a.streams + at(C_PD);
        }
        if (lane == 0) {
            a.feat_off[r] = fbase;                      // feat_off[n_reads] is filled in by the launcher
            a.read_qs_off[r] = ~0ull;
            a.read_ba_off[r] = ~0ull;
        }
    }

    if (!mapped) {
        // unmapped: bases one by one into BA (unless unknown), then the preserved scores [Writer.java:254-258]
        if (!(rfl & CRAMGPU_RF_NO_BASES)) {
            if (WRITE) {
                for (int64_t i = lane; i < rlen; i += 32) sBA[i] = rb[i];
                if (lane == 0) a.read_ba_off[r] = (uint64_t)(sBA - a.streams);
            }
            n_ba += (uint32_t)rlen;
        }
    } else if (hasfeat) {
        const uint8_t *codes = WRITE ? a.codes + (size_t)c * 32 : nullptr;
        int64_t z = 0, aoff = 0, d_a = -9, d_b = -9;
        for (uint64_t ci = c0; ci < c1; ci++) {
            uint32_t cw = a.cigar[ci]; int op = cw & 15; int64_t len = cw >> 4;
            if (op > OP_X) { if (lane == 0) atomicExch(a.status, CRAMGPU_ERR_ARG); break; }   // "Unsupported cigar operator"
            if (op == OP_D) { if (d_a != z) { d_b = d_a; d_a = z; } }
            if (!consumes_read(op)) {
                // D, N, H, P: one feature at the next read position [:228-240]
                int32_t fpos = (int32_t)z + 1;
                uint32_t dl = itf8_len((uint32_t)(fpos - prev)), vl = itf8_len((uint32_t)len);
                if (WRITE && lane == 0) {
                    uint64_t fi = fbase + n_feat;
                    uint8_t opc = op == OP_D ? 'D' : op == OP_N ? 'N' : op == OP_H ? 'H' : 'P';
                    a.f_op[fi] = opc; a.f_pos[fi] = fpos; a.f_val[fi] = (int32_t)len; a.f_aux[fi] = 0;
                    sFC[n_feat] = opc;
                    itf8_put(sFP + n_fp, (uint32_t)(fpos - prev));
                    uint8_t *dst = op == OP_D ? sDL + n_dl : op == OP_N ? sRS + n_rs : op == OP_H ? sHC + n_hc : sPD + n_pd;
                    itf8_put(dst, (uint32_t)len);
                }
                n_feat++; n_fp += dl; prev = fpos;
                if (op == OP_D) n_dl += vl; else if (op == OP_N) n_rs += vl; else if (op == OP_H) n_hc += vl; else n_pd += vl;
            } else {
                const bool isM = op == OP_M || op == OP_EQ || op == OP_X;
                if (op == OP_S) {
                    // SoftClip(pos, bases[pos .. pos+len)) [:268-272]; bytes + TAB into SC, bytes into the payload
                    if (WRITE) {
                        for (int64_t i = lane; i < len; i += 32) {
                            uint8_t b = rlen ? rb[z + i] : (uint8_t)'N';
                            sSC[n_sc + i] = b; a.payload[paybase + n_pay + i] = b;
                        }
                        if (lane == 0) sSC[n_sc + len] = '\t';
                    }
                }
                for (int64_t i0 = 0; i0 < len; i0 += 32) {
                    int64_t i = i0 + lane;
                    bool in = i < len;
                    int64_t zi = z + i;
                    uint8_t b = in ? (rlen ? rb[zi] : (uint8_t)'N') : 0;                      // empty bases -> 'N' x cigarLen [:218-220]
                    int64_t rp = (int64_t)pos - 1 + aoff + i;
                    uint8_t rraw = (in && rp >= 0 && rp < ref_len) ? ref[rp] : (uint8_t)'N';     // baseAt: 'N' beyond the window's end
                    uint8_t refb = norm_base((in && rp < ref_len) ? rraw : (uint8_t)'N');     // [:306-308]
                    bool isX = in && isM && b != refb;
                    bool hasF = in && (isX || op == OP_I || (op == OP_S && i == 0));
                    uint8_t sc = 0xFF;
                    if (q_feats && in && zi < rlen) {
                        bool dm = zi == d_a || zi == d_a + 1 || zi == d_b || zi == d_b + 1;
                        sc = score_at(ps, rq[zi], flag, mapq, pos, b == rraw, op == OP_I, dm, a.qs_ext ? a.qs_ext[lo + zi] : 0u, a.status);
                    }
                    bool hasQ = sc < 128;                                                     // scores[i] > -1 [:290]
                    uint32_t bF = __ballot_sync(0xFFFFFFFFu, hasF), bQ = __ballot_sync(0xFFFFFFFFu, hasQ);
                    uint32_t bAny = bF | bQ;
                    if (bAny) {
                        uint32_t lt = (1u << lane) - 1u;
                        int32_t mypos = (int32_t)zi + 1;
                        // position of the previous feature: nearest lower lane with a feature, else the carried one
                        uint32_t below = bAny & lt;
                        int32_t pprev = below ? (int32_t)(z + i0 + (31 - __clz(below))) + 1 : prev;
                        uint32_t dF = hasF ? (uint32_t)(mypos - pprev) : 0;
                        uint32_t dQ = hasQ ? (hasF ? 0u : (uint32_t)(mypos - pprev)) : 0;
                        uint32_t lF = hasF ? itf8_len(dF) : 0, lQ = hasQ ? itf8_len(dQ) : 0;
                        uint32_t fp_tot, fp_off = warp_excl_scan_u32(lF + lQ, lane, fp_tot);
                        uint32_t fidx = n_feat + __popc(bF & lt) + __popc(bQ & lt);
                        uint32_t bX = __ballot_sync(0xFFFFFFFFu, isX), bI = __ballot_sync(0xFFFFFFFFu, in && op == OP_I);
                        if (!WRITE) {
                            if (isX) {
                                int ri = base_index(refb), bi = base_index(b);
                                if (ri >= 0 && bi >= 0) atomicAdd(&a.subst[(size_t)c * 25 + ri * 5 + bi], 1ull);
                            }
                        } else {
                            if (hasF) {
                                uint64_t fi = fbase + fidx;
                                uint8_t opc = isX ? 'X' : op == OP_I ? 'i' : 'S';
                                a.f_op[fi] = opc; a.f_pos[fi] = mypos;
                                sFC[fidx] = opc;
                                itf8_put(sFP + n_fp + fp_off, dF);
                                if (isX) {
                                    int ri = base_index(refb), bi = base_index(b);
                                    uint8_t code = (ri >= 0 && bi >= 0) ? codes[ri * 5 + bi] : 0;  // codes[][] default 0 [:39]
                                    a.f_val[fi] = code; a.f_aux[fi] = (int32_t)b | ((int32_t)refb << 8);
                                    sBS[n_bs + __popc(bX & lt)] = code;
                                } else if (op == OP_I) {
                                    a.f_val[fi] = b; a.f_aux[fi] = 0;
                                    sBA[n_ba + __popc(bI & lt)] = b;
                                } else {
                                    a.f_val[fi] = (int32_t)(paybase + n_pay); a.f_aux[fi] = (int32_t)len;
                                }
                            }
                            if (hasQ) {
                                uint64_t fi = fbase + fidx + (hasF ? 1 : 0);
                                a.f_op[fi] = 'Q'; a.f_pos[fi] = mypos; a.f_val[fi] = sc; a.f_aux[fi] = 0;
                                sFC[fidx + (hasF ? 1 : 0)] = 'Q';
                                itf8_put(sFP + n_fp + fp_off + lF, dQ);
                                sQS[n_qs + __popc(bQ & lt)] = sc;
                            }
                        }
                        n_feat += __popc(bF) + __popc(bQ);
                        n_fp += fp_tot;
                        n_bs += __popc(bX); n_ba += __popc(bI); n_qs += __popc(bQ);
                        prev = (int32_t)(z + i0 + (31 - __clz(bAny))) + 1;
                    }
                }
                if (op == OP_S) { n_sc += (uint32_t)len + 1; n_pay += (uint32_t)len; }
                z += len;
            }
            if (consumes_ref(op)) aoff += len;
        }
    }
    // preserved score array after the features (mapped) or after the bases (unmapped) [Writer.java:250-251,257-258]
    if (forced) {
        if (WRITE) {
            // lossless: the array is the record's own qualities; under a lossy model k_rec_forced_scores fills it in
            if (ps.n < 0) for (int64_t i = lane; i < rlen; i += 32) sQS[n_qs + i] = rq[i];
            if (lane == 0) a.read_qs_off[r] = (uint64_t)(sQS + n_qs - a.streams);
        }
        n_qs += (uint32_t)rlen;
    }
    if (!WRITE && lane == 0) {
        // the counter arrays were cleared before the pass: store only what is not zero (most reads: FN, QS, MQ)
        uint32_t *cn = a.cnt + r;
        const uint64_t N = a.n_reads;
        if (mapped) { cn[C_FN * N] = itf8_len(n_feat); cn[C_MQ * N] = itf8_len(mapq); }
        if (n_qs) cn[C_QS * N] = n_qs;
        if (n_feat) { cn[C_FC * N] = n_feat; cn[C_FP * N] = n_fp; cn[C_NFEAT * N] = n_feat; }
        if (n_bs) cn[C_BS * N] = n_bs;
        if (n_ba) cn[C_BA * N] = n_ba;
        if (n_dl) cn[C_DL * N] = n_dl;
        if (n_sc) { cn[C_SC * N] = n_sc; cn[C_PAYLOAD * N] = n_pay; }
        if (n_rs) cn[C_RS * N] = n_rs;
        if (n_hc) cn[C_HC * N] = n_hc;
        if (n_pd) cn[C_PD * N] = n_pd;
    }
    if (WRITE && lane == 0 && mapped) {
        const uint64_t *B = a.base + (size_t)c * NCNT;
        itf8_put(a.streams + B[C_FN] + a.cnt[(size_t)C_FN * a.n_reads + r], n_feat);
        itf8_put(a.streams + B[C_MQ] + a.cnt[(size_t)C_MQ * a.n_reads + r], mapq);
    }
}

// general walk over the reads k_rec_simple left aside: warps stride over the list
template <bool WRITE>
__global__ void __launch_bounds__(128) k_rec_walk(RecArgs a, PolSet ps)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t n = *a.clist_n, step = (uint64_t)gridDim.x * 4;
    for (uint64_t i = (uint64_t)blockIdx.x * 4 + (threadIdx.x >> 5); i < n; i += step)
        walk_read<WRITE>(a, ps, a.clist_r[i], a.clist_c[i], lane);
}

// Score arrays under a quality model WITHOUT base categories: every score of a force-preserved record gets the treatment
// of the last policy that applies to the record (QualityScorePreservation.addQualityScores [:276-301] with only read
// categories in play) -- preserved as it is, or binned [Binning.java:63-71].  One thread per read, in order; 32 records of
// one length and one treatment whose arrays lie back to back are done as one warp-wide copy.
struct BinW {
    int32_t *status;
    __device__ __forceinline__ uint32_t operator()(uint32_t w) const
    {
        // any byte of w >= 110: bit 7 set, or the low seven bits + 18 carry into bit 7
        if (((w | ((w & 0x7F7F7F7Fu) + 0x12121212u)) & 0x80808080u)) bin_fault(status);
        return (uint32_t)illumina_bin((uint8_t)w) | ((uint32_t)illumina_bin((uint8_t)(w >> 8)) << 8) |
               ((uint32_t)illumina_bin((uint8_t)(w >> 16)) << 16) | ((uint32_t)illumina_bin((uint8_t)(w >> 24)) << 24);
    }
};
struct BinB {
    int32_t *status;
    __device__ __forceinline__ uint8_t operator()(uint8_t b) const { if (b >= BIN_FAULT_Q) bin_fault(status); return illumina_bin(b); }
};
struct FillW { __device__ __forceinline__ uint32_t operator()(uint32_t) const { return 0xFFFFFFFFu; } };
struct FillB { __device__ __forceinline__ uint8_t operator()(uint8_t) const { return 0xFF; } };

__global__ void __launch_bounds__(128) k_rec_forced_flat(RecArgs a, PolSet ps)
{
    const uint32_t lane = threadIdx.x & 31;
    uint64_t r; uint32_t c; bool live;
    if (gridDim.y > 1 || a.n_containers == 1) {
        c = blockIdx.y;
        r = a.cont_off[c] + (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.cont_off[c + 1];
    } else {
        r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.n_reads;
    }
    uint64_t qoff = ~0ull, lo = 0, rlen = 0;
    int treatment = 2;                                   // 2: no policy sets a score (bytes stay "none")
    if (live && (a.force[r] & 1)) {
        qoff = a.read_qs_off[r];
        lo = a.read_off[r];
        rlen = a.read_off[r + 1] - lo;
        const uint16_t flag = a.flag[r];
        const uint8_t mapq = a.mapq[r];
        bool binned = false;
        for (int k = 0; k < ps.n; k++)
            if (read_in_category(ps.p[k], flag, mapq) && ps.p[k].treatment != 2) { treatment = ps.p[k].treatment; binned |= treatment == 1; }
        if (binned && treatment != 1)                    // binned by an earlier policy, then overwritten: the lookup still happened
            for (uint64_t i = 0; i < rlen; i++) if (a.quals[lo + i] >= BIN_FAULT_Q) bin_fault(a.status);
    }
    const bool has = qoff != ~0ull && rlen < (1ull << 31);
    const uint64_t q0 = __shfl_sync(0xFFFFFFFFu, qoff, 0), lo0 = __shfl_sync(0xFFFFFFFFu, lo, 0), rl0 = __shfl_sync(0xFFFFFFFFu, rlen, 0);
    const int t0 = __shfl_sync(0xFFFFFFFFu, treatment, 0);
    const uint32_t L0 = (uint32_t)rl0;
    const bool flat = has && rlen == rl0 && treatment == t0 && qoff == q0 + (uint64_t)lane * L0 && lo == lo0 + (uint64_t)lane * L0;
    if (__all_sync(0xFFFFFFFFu, flat) && rl0 <= (1u << 20)) {
        if (t0 == 0) warp_copy(a.streams + q0, a.quals + lo0, 32 * L0, lane);
        else if (t0 == 1) warp_copy_x(a.streams + q0, a.quals + lo0, 32 * L0, lane, BinW{a.status}, BinB{a.status});
        else warp_copy_x(a.streams + q0, a.quals + lo0, 32 * L0, lane, FillW(), FillB());
    } else if (has) {
        if (treatment == 0) copy_bytes(a.streams + qoff, a.quals + lo, (uint32_t)rlen, IdW(), IdB());
        else if (treatment == 1) copy_bytes(a.streams + qoff, a.quals + lo, (uint32_t)rlen, BinW{a.status}, BinB{a.status});
        else copy_bytes(a.streams + qoff, a.quals + lo, (uint32_t)rlen, FillW(), FillB());
    }
}

// Forced arrays under a lossy model need the per-base scores again; done by a second small kernel so that the walk above
// stays simple: one warp per read, same CIGAR walk, writes into the already reserved QS slot.
__global__ void __launch_bounds__(128) k_rec_forced_scores(RecArgs a, PolSet ps)
{
    const uint32_t lane = threadIdx.x & 31;
    uint64_t r; uint32_t c;
    if (gridDim.y > 1 || a.n_containers == 1) {
        c = blockIdx.y;
        r = a.cont_off[c] + (uint64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
        if (r >= a.cont_off[c + 1]) return;
    } else {
        r = (uint64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
        if (r >= a.n_reads) return;
        c = container_of(a.cont_off, a.n_containers, r);
    }
    if (!(a.force[r] & 1)) return;
    uint64_t qoff = a.read_qs_off[r];
    if (qoff == ~0ull) return;
    const uint8_t *ref = a.ref + a.ref_off[c];
    const int64_t ref_len = (int64_t)a.ref_len[c];
    const uint64_t lo = a.read_off[r];
    const int64_t rlen = (int64_t)(a.read_off[r + 1] - lo);
    const uint8_t *rb = a.bases + lo, *rq = a.quals + lo;
    const int32_t pos = a.pos[r];
    const uint16_t flag = a.flag[r];
    const uint8_t mapq = a.mapq[r];
    uint8_t *dst = a.streams + qoff;
    // read-level part of every base (bases not covered by a read-consuming CIGAR op see only these policies)
    for (int64_t i = lane; i < rlen; i += 32) dst[i] = score_at(ps, rq[i], flag, mapq, 0, false, false, false, 0u, a.status);
    if (!ps.per_base || pos == 0) return;
    __syncwarp();
    int64_t z = 0, aoff = 0, d_a = -9, d_b = -9;
    for (uint64_t ci = a.cigar_off[r]; ci < a.cigar_off[r + 1]; ci++) {
        uint32_t cw = a.cigar[ci]; int op = cw & 15; int64_t len = cw >> 4;
        if (op == OP_D) { if (d_a != z) { d_b = d_a; d_a = z; } }
        if (consumes_read(op)) {
            for (int64_t i = lane; i < len; i += 32) {
                int64_t zi = z + i;
                if (zi >= rlen) break;
                int64_t rp = (int64_t)pos - 1 + aoff + i;
                uint8_t tb = (rp >= 0 && rp < ref_len) ? ref[rp] : (uint8_t)'N';
                bool dm = zi == d_a || zi == d_a + 1 || zi == d_b || zi == d_b + 1;
                dst[zi] = score_at(ps, rq[zi], flag, mapq, pos, rb[zi] == tb, op == OP_I, dm, a.qs_ext ? a.qs_ext[lo + zi] : 0u, a.status);
            }
            z += len;
        }
        if (consumes_ref(op)) aoff += len;
    }
}

// After a call whose status word says that a binned score was outside the binning table: which records?  One warp per
// record, every base against every BIN policy that applies to it (the places where score_at calls treat with BIN), bit 7
// of force_qs set for the records the reference throws on.  Runs only then.
__device__ __forceinline__ bool bin_hits(const PolSet &ps, uint16_t flag, uint8_t mapq, int32_t pos, bool match, bool is_ins, bool dmark, uint32_t ext)
{
    for (int k = 0; k < ps.n; k++) {
        const cramgpu_qs_policy &p = ps.p[k];
        if (p.treatment != 1 || !read_in_category(p, flag, mapq)) continue;
        if (p.n_base_cats == 0 || (pos != 0 && base_mask(p, match, is_ins, dmark, ext, ps.ext_bit[k]))) return true;
    }
    return false;
}
__global__ void __launch_bounds__(128) k_rec_bin_fault(RecArgs a, PolSet ps)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t r = (uint64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (r >= a.n_reads) return;
    if (a.rflags && (a.rflags[r] & CRAMGPU_RF_NO_QUALS)) return;                  // addQualityScores returns first [:276-280]
    const uint32_t c = container_of(a.cont_off, a.n_containers, r);
    const uint8_t *ref = a.ref + a.ref_off[c];
    const int64_t ref_len = (int64_t)a.ref_len[c];
    const uint64_t lo = a.read_off[r];
    const int64_t rlen = (int64_t)(a.read_off[r + 1] - lo);
    const uint8_t *rb = a.bases + lo, *rq = a.quals + lo;
    const int32_t pos = a.pos[r];
    const uint16_t flag = a.flag[r];
    const uint8_t mapq = a.mapq[r];
    bool fault = false;
    for (int64_t i = lane; i < rlen; i += 32) fault |= rq[i] >= BIN_FAULT_Q && bin_hits(ps, flag, mapq, 0, false, false, false, 0u);
    if (ps.per_base && pos != 0) {
        int64_t z = 0, aoff = 0, d_a = -9, d_b = -9;
        for (uint64_t ci = a.cigar_off[r]; ci < a.cigar_off[r + 1]; ci++) {
            uint32_t cw = a.cigar[ci]; int op = cw & 15; int64_t len = cw >> 4;
            if (op == OP_D) { if (d_a != z) { d_b = d_a; d_a = z; } }
            if (consumes_read(op)) {
                for (int64_t i = lane; i < len; i += 32) {
                    int64_t zi = z + i;
                    if (zi >= rlen) break;
                    int64_t rp = (int64_t)pos - 1 + aoff + i;
                    uint8_t tb = (rp >= 0 && rp < ref_len) ? ref[rp] : (uint8_t)'N';
                    bool dm = zi == d_a || zi == d_a + 1 || zi == d_b || zi == d_b + 1;
                    fault |= rq[zi] >= BIN_FAULT_Q && bin_hits(ps, flag, mapq, pos, rb[zi] == tb, op == OP_I, dm, a.qs_ext ? a.qs_ext[lo + zi] : 0u);
                }
                z += len;
            }
            if (consumes_ref(op)) aoff += len;
        }
    }
    fault = __any_sync(0xFFFFFFFFu, fault);
    if (lane == 0 && fault) a.force[r] |= 0x80;
}

// SubstitutionMatrix.<init>(long[][]) / rank [SubstitutionMatrix.java:42-57,129-169]: one thread per container
__global__ void k_rec_matrix(const unsigned long long *__restrict__ subst, uint32_t nc, uint8_t *__restrict__ codes,
                             uint8_t *__restrict__ matrix)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    const unsigned long long *cnt = subst + (size_t)c * 25;
    for (int r = 0; r < 5; r++) {
        int base[4]; long long freq[4]; int k = 0;
        for (int b = 0; b < 5; b++) if (b != r) { base[k] = b; freq[k] = (long long)cnt[r * 5 + b]; k++; }
        for (int i = 1; i < 4; i++) {                       // stable insertion sort with the reference's comparator
            int bb = base[i]; long long ff = freq[i]; int j = i - 1;
            while (j >= 0) {
                int cmp = (freq[j] != ff) ? (int)(unsigned int)(unsigned long long)(ff - freq[j]) : (base[j] - bb);
                if (cmp > 0) { base[j + 1] = base[j]; freq[j + 1] = freq[j]; j--; } else break;
            }
            base[j + 1] = bb; freq[j + 1] = ff;
        }
        int rank_of[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 4; i++) rank_of[base[i]] = i;
        uint8_t byte = 0;
        for (int b = 0; b < 5; b++) {
            codes[(size_t)c * 32 + r * 5 + b] = (b == r) ? 0 : (uint8_t)rank_of[b];
            if (b != r) byte = (uint8_t)((byte << 2) | rank_of[b]);
        }
        matrix[(size_t)c * 8 + r] = byte;
    }
    matrix[(size_t)c * 8 + 5] = matrix[(size_t)c * 8 + 6] = matrix[(size_t)c * 8 + 7] = 0;
}

// per-container exclusive scans of the per-read counters: grid (n_containers, NCNT), 256 threads
__global__ void __launch_bounds__(256) k_rec_scan(uint32_t *__restrict__ cnt, const uint64_t *__restrict__ cont_off,
                                                  uint64_t n_reads, uint64_t *__restrict__ totals)
{
    uint32_t c = blockIdx.x, k = blockIdx.y;
    uint64_t r0 = cont_off[c], r1 = cont_off[c + 1];
    uint32_t *v = cnt + (size_t)k * n_reads;
    __shared__ uint32_t wsum[8];
    __shared__ unsigned long long carry;
    uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (uint64_t base = r0; base < r1; base += 256) {
        uint64_t r = base + tid;
        uint32_t x = r < r1 ? v[r] : 0;
        uint32_t tot, ex = warp_excl_scan_u32(x, lane, tot);
        if (lane == 0) wsum[w] = tot;
        __syncthreads();
        uint32_t off = 0, all = 0;
        for (uint32_t j = 0; j < 8; j++) { if (j < w) off += wsum[j]; all += wsum[j]; }
        unsigned long long cbase = carry;
        if (r < r1) v[r] = (uint32_t)(cbase + off + ex);
        __syncthreads();
        if (tid == 0) carry = cbase + all;
        __syncthreads();
    }
    if (tid == 0) totals[(size_t)c * NCNT + k] = carry;
}

// absolute bases: one CTA; thread = container.  The streams buffer is SERIES-major: all containers' FN regions, then all
// FC regions, ... so that a whole series of the call is one contiguous piece of memory (one copy takes it to the host, or
// leaves it out).  Region starts are multiples of 16, series starts multiples of 256.
// grand[0] = bytes of streams, grand[1] = features, grand[2] = payload bytes, grand[4 + 2k] / [5 + 2k] = start / extent of series k
#define NSCAN (CRAMGPU_S_COUNT + 2)
__global__ void __launch_bounds__(1024) k_rec_layout(const uint64_t *__restrict__ totals, uint32_t nc, uint64_t *__restrict__ base,
                                                     uint64_t *__restrict__ stream_off, uint64_t *__restrict__ stream_len,
                                                     uint64_t *__restrict__ grand)
{
    __shared__ unsigned long long wsum[NSCAN][32];
    __shared__ unsigned long long carry[NSCAN], start[NSCAN];
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    auto val = [&](uint32_t c, int q) -> unsigned long long {
        if (c >= nc) return 0;
        if (q < CRAMGPU_S_COUNT) return (totals[(size_t)c * NCNT + q] + 15) & ~15ull;
        return totals[(size_t)c * NCNT + (q == CRAMGPU_S_COUNT ? C_NFEAT : C_PAYLOAD)];
    };
    // pass 1: extent of every series over all containers -> where each series starts
    {
        unsigned long long acc[NSCAN];
#pragma unroll
        for (int q = 0; q < NSCAN; q++) acc[q] = 0;
        for (uint32_t c = tid; c < nc; c += 1024)
#pragma unroll
            for (int q = 0; q < NSCAN; q++) acc[q] += val(c, q);
#pragma unroll
        for (int q = 0; q < NSCAN; q++) {
            unsigned long long v = acc[q];
            for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
            if (lane == 0) wsum[q][w] = v;
        }
        __syncthreads();
        if (tid == 0) {
            unsigned long long at = 0;
            for (int q = 0; q < NSCAN; q++) {
                unsigned long long t = 0;
                for (int j = 0; j < 32; j++) t += wsum[q][j];
                if (q < CRAMGPU_S_COUNT) { start[q] = at; grand[4 + 2 * q] = at; grand[5 + 2 * q] = t; at += (t + 255) & ~255ull; }
                else { start[q] = 0; grand[q - CRAMGPU_S_COUNT + 1] = t; }
                carry[q] = 0;
            }
            grand[0] = at;
        }
        __syncthreads();
    }
    // pass 2: exclusive scans over the containers, one quantity after the other
    for (uint32_t c0 = 0; c0 < nc; c0 += 1024) {
        const uint32_t c = c0 + tid;
        for (int q = 0; q < NSCAN; q++) {
            const unsigned long long v = val(c, q);
            unsigned long long sc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, sc, d);
                if (lane >= (uint32_t)d) sc += t;
            }
            if (lane == 31) wsum[0][w] = sc;
            __syncthreads();
            unsigned long long ex = carry[q] + sc - v;
            for (uint32_t j = 0; j < w; j++) ex += wsum[0][j];
            if (c < nc) {
                if (q < CRAMGPU_S_COUNT) {
                    const unsigned long long sb = start[q] + ex;
                    base[(size_t)c * NCNT + q] = sb;
                    stream_off[(size_t)c * CRAMGPU_S_COUNT + q] = sb;
                    stream_len[(size_t)c * CRAMGPU_S_COUNT + q] = totals[(size_t)c * NCNT + q];
                } else {
                    base[(size_t)c * NCNT + (q == CRAMGPU_S_COUNT ? C_NFEAT : C_PAYLOAD)] = ex;
                }
            }
            __syncthreads();
            if (tid == 1023) carry[q] = ex + v;
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------ decode

struct NormArgs {
    uint64_t n_reads; uint32_t n_containers;
    const uint64_t *cont_off, *win_off, *win_len; const int64_t *win_start0;      // device copies
    const uint8_t *ref, *matrix;
    const int32_t *pos; const uint64_t *read_off; const uint8_t *rflags; const uint64_t *feat_off;
    const uint8_t *f_op; const int32_t *f_pos, *f_val, *f_aux; const uint8_t *payload;
    const uint8_t *qs; const uint64_t *read_qs_off; const uint8_t *ba; const uint64_t *read_ba_off;
    uint8_t *bases, *quals, *qual_star; int32_t *status;
    uint64_t *clist_r; uint32_t *clist_c; unsigned long long *clist_n;            // records left to k_norm
    const uint8_t *nref;                                                         // normalizeBase(ref[i]) over the regions, same offsets as ref
    uint64_t ref_total;
    uint64_t payload_len;                                                        // bytes in payload (soft-clip / insertion offsets are checked against it)
};

// base with rank `code` for normalised reference base r: SubstitutionMatrix(byte[]) / base() [:59-..,177]
__device__ __forceinline__ uint8_t matrix_base(const uint8_t *m, uint8_t refb, int code)
{
    int r = base_index(refb);
    const char B[5] = {'A', 'C', 'G', 'T', 'N'};
    int k = 0;
    uint8_t out = 'N';
    for (int b = 0; b < 5; b++) {
        if (b == r) continue;
        int rank = (m[r] >> (2 * (3 - k))) & 3;
        if (rank == code) out = (uint8_t)B[b];
        k++;
    }
    return out;
}

// Decode runs one THREAD per record, in two kernels:
//   k_norm_simple  every record, in order (coalesced tables and results): the qualities of records with a stored score
//                  array, and the bases of records whose restoration is a plain copy -- no read features and inside the
//                  reference region (the region bytes as they are, :346-362), unmapped (bases from BA), unknown bases.
//                  The others (read features, region overhang, no stored array) go to a list;
//   k_norm         the list, one thread per entry, so that all lanes of a warp do the same kind of work: the
//                  feature-by-feature restoration [CramNormalizer.java:335-434] over a normalised copy of the region.
__device__ __forceinline__ void append_list(const NormArgs &a, bool cx, uint64_t r, uint32_t c, uint32_t lane)
{
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, cx);
    if (!m) return;
    unsigned long long base = 0;
    if (lane == (uint32_t)(__ffs(m) - 1)) base = atomicAdd(a.clist_n, (unsigned long long)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
    if (cx) { const uint64_t k = base + __popc(m & ((1u << lane) - 1u)); a.clist_r[k] = r; a.clist_c[k] = c; }
}

__global__ void __launch_bounds__(128) k_norm_simple(NormArgs a)
{
    const uint32_t lane = threadIdx.x & 31;
    uint64_t r; uint32_t c; bool live;
    if (gridDim.y > 1 || a.n_containers == 1) {
        c = blockIdx.y;
        r = a.cont_off[c] + (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.cont_off[c + 1];
    } else {
        r = (uint64_t)blockIdx.x * 128 + threadIdx.x;
        live = r < a.n_reads;
        c = live ? container_of(a.cont_off, a.n_containers, r) : 0;
    }
    bool listed = false, stored = false, inwin = false;
    uint64_t lo = 0, rlen = 0, qo = ~0ull;
    int64_t o = 0;                             // offset of the record inside its reference region
    const uint8_t *src = nullptr;              // where the bases come from (nullptr: nothing to write here)
    if (live) {
        lo = a.read_off[r];
        rlen = a.read_off[r + 1] - lo;
        const uint8_t rfl = a.rflags[r];
        qo = a.read_qs_off[r];
        stored = (rfl & CRAMGPU_NF_FORCE_QS) != 0;
        listed = !stored || rlen >= (1ull << 31);
        if (rfl & CRAMGPU_NF_UNKNOWN_BASES) {
        } else if (rfl & CRAMGPU_NF_UNMAPPED) {
            const uint64_t bo = a.read_ba_off[r];
            if (bo != ~0ull) src = a.ba + bo;
        } else if (rlen > 0) {
            o = (int64_t)a.pos[r] - 1 - a.win_start0[c];
            inwin = o >= 0 && (uint64_t)o + rlen <= a.win_len[c];
            if (a.feat_off[r] != a.feat_off[r + 1] || !inwin) listed = true;
            else src = a.ref + a.win_off[c] + o;
        }
    }
    append_list(a, listed, r, c, lane);
    const uint64_t lo0 = __shfl_sync(0xFFFFFFFFu, lo, 0), rl0 = __shfl_sync(0xFFFFFFFFu, rlen, 0);
    const uint32_t L0 = (uint32_t)rl0, c0 = __shfl_sync(0xFFFFFFFFu, c, 0);
    // ---- bases.  32 mapped records of one length inside one region: the warp fills their (contiguous) bases together;
    //      records with read features get the plain region bytes too and are redone by k_norm, which runs afterwards.
    const bool fb = live && inwin && rlen == rl0 && c == c0 && o < (int64_t)0xFFFF0000ll;
    if (__all_sync(0xFFFFFFFFu, fb) && rl0 >= 16 && rl0 <= 8192)
        warp_gather_reads(a.bases + lo0, a.ref + a.win_off[c], a.win_off[c], a.ref_total - a.win_off[c], (uint32_t)o, L0, lane);
    else if (src && rlen < (1ull << 31)) copy_bytes(a.bases + lo, src, (uint32_t)rlen, IdW(), IdB());
    // ---- qualities: restoreQualityScores with a stored array (:306-333): 0xFF = missing -> default score
    const uint32_t dflt = '?' - '!';
    const uint64_t qo0 = __shfl_sync(0xFFFFFFFFu, qo, 0);
    const bool fq = live && stored && qo != ~0ull && rlen == rl0 && qo == qo0 + (uint64_t)lane * L0;
    if (__all_sync(0xFFFFFFFFu, fq) && rl0 >= 1 && rl0 <= (1u << 20)) {
        // the 32 arrays lie back to back in QS as the records do in `quals`: one warp-wide copy, unless a score is missing
        const bool ff = warp_copy(a.quals + lo0, a.qs + qo0, 32 * L0, lane);
        if (!__any_sync(0xFFFFFFFFu, ff)) {
            a.qual_star[r] = 0;
            if (!listed) a.status[r] = CRAMGPU_OK;
            return;
        }
    }
    if (!live || !stored || rlen >= (1ull << 31)) return;
    uint32_t present = 0;
    if (qo != ~0ull) {
        auto xw = [&](uint32_t w) -> uint32_t {
            const uint32_t miss = __vcmpeq4(w, 0xFFFFFFFFu);
            present |= ~miss;
            return (w & ~miss) | (miss & (dflt * 0x01010101u));
        };
        auto xb = [&](uint8_t b) -> uint8_t { if (b == 0xFF) return (uint8_t)dflt; present = 1; return b; };
        copy_bytes(a.quals + lo, a.qs + qo, (uint32_t)rlen, xw, xb);
    } else {
        copy_bytes(a.quals + lo, a.quals + lo, (uint32_t)rlen, [&](uint32_t) { return dflt * 0x01010101u; }, [&](uint8_t) { return (uint8_t)dflt; });
    }
    a.qual_star[r] = present == 0;             // missingScores == scores.length
    if (!listed) a.status[r] = CRAMGPU_OK;
}

__global__ void __launch_bounds__(128) k_norm(NormArgs a)
{
    const uint64_t n = *a.clist_n, step = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t li = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; li < n; li += step) {
        const uint64_t r = a.clist_r[li];
        const uint32_t c = a.clist_c[li];
        const uint8_t *ref = a.ref + a.win_off[c];             // 'N' fill / region checks use the raw region ...
        const uint8_t *nref = a.nref + a.win_off[c];           // ... feature records see normalizeBase(region)
        const int64_t ref_len = (int64_t)a.win_len[c], off0 = a.win_start0[c];
        const uint8_t *mtx = a.matrix + (size_t)c * 8;
        const uint64_t lo = a.read_off[r];
        const int64_t rlen = (int64_t)(a.read_off[r + 1] - lo);
        const uint8_t rfl = a.rflags[r];
        const uint64_t f0 = a.feat_off[r], f1 = a.feat_off[r + 1];
        uint8_t *bases = a.bases + lo, *quals = a.quals + lo;
        int bad = 0;
        // a run of n region bases (normalised) into the read at 0-based offset `at`
        auto run = [&](int64_t at, int64_t rp, int64_t cnt) {
            if (rp + cnt <= ref_len && cnt < (1ll << 31)) copy_bytes(bases + at, nref + rp, (uint32_t)cnt, IdW(), IdB());
            else for (int64_t i = 0; i < cnt; i++) bases[at + i] = rp + i >= ref_len ? (uint8_t)'N' : nref[rp + i];
        };

        // ---- bases: restoreBases / restoreReadBases (build/CramNormalizer.java:208-244,335-434); the plain cases were done by k_norm_simple
        if (!(rfl & (CRAMGPU_NF_UNKNOWN_BASES | CRAMGPU_NF_UNMAPPED)) && rlen > 0) {
            const int64_t al0 = (int64_t)a.pos[r] - 1;
            if (f0 == f1) {
                // no features: straight copy, 'N' fill when the record overhangs the region (:346-362); no normalizeBase here
                if (ref_len + off0 < al0 + rlen) {
                    const int64_t o = al0 - off0, avail = ref_len - o;
                    const int64_t ncopy = (o > 0 && avail > 0) ? (rlen < avail ? rlen : avail) : 0;
                    for (int64_t i = 0; i < rlen; i++) bases[i] = i < ncopy ? ref[o + i] : (uint8_t)'N';
                } else if (al0 - off0 < 0) {
                    bad = 1;
                } else if (rlen >= (1ll << 31)) {
                    for (int64_t i = 0; i < rlen; i++) bases[i] = ref[al0 - off0 + i];
                }
            } else {
                int64_t pir = 1, pis = 0;      // posInRead (1-based), posInSeq
                for (uint64_t f = f0; f < f1 && !bad; f++) {
                    const int64_t fpos = a.f_pos[f];
                    const uint8_t op = a.f_op[f];
                    const int64_t cnt = fpos - pir;            // reference bases up to the feature (:365-373)
                    if (cnt > 0) {
                        if (al0 + pis - off0 < 0 || fpos - 1 > rlen || pir < 1) { bad = 1; break; }
                        run(pir - 1, al0 + pis - off0, cnt);
                        pir += cnt; pis += cnt;
                    }
                    switch (op) {
                    case 'X': {
                        const int64_t rp = al0 + pis - off0;
                        if (rp < 0 || pir > rlen || pir < 1) { bad = 1; break; }
                        const uint8_t refb = rp >= ref_len ? (uint8_t)'N' : nref[rp];
                        bases[pir - 1] = norm_base(matrix_base(mtx, refb, a.f_val[f]));
                        pir++; pis++;
                        break;
                    }
                    case 'S': case 'I': {
                        const int64_t ln = a.f_aux[f], po = a.f_val[f];
                        // Java: ArrayIndexOutOfBoundsException / NegativeArraySizeException on a read or payload overrun
                        if (ln < 0 || pir < 1 || pir - 1 + ln > rlen || po < 0 || (uint64_t)(po + ln) > a.payload_len) { bad = 1; break; }
                        const uint8_t *pay = a.payload + po;
                        for (int64_t i = 0; i < ln; i++) bases[pir - 1 + i] = norm_base(pay[i]);
                        pir += ln;
                        break;
                    }
                    case 'D': case 'N': pis += a.f_val[f]; break;
                    case 'i':
                        if (pir > rlen || pir < 1) { bad = 1; break; }
                        bases[pir - 1] = norm_base((uint8_t)a.f_val[f]);
                        pir++;
                        break;
                    default: break;            // H, P, Q, B consume nothing here
                    }
                }
                if (!bad) {
                    // tail from the reference while it lasts (:412-415); what is left stays 0 -> normalizeBase -> 'N'
                    const int64_t cnt = rlen - pir + 1;
                    if (cnt > 0 && al0 + pis - off0 < 0) bad = 1;
                    else if (cnt > 0) run(pir - 1, al0 + pis - off0, cnt);
                    if (!bad)
                        for (uint64_t f = f0; f < f1; f++)          // ReadBase overwrites (:418-427)
                            if (a.f_op[f] == 'B') {
                                const int64_t fp = a.f_pos[f];
                                if (fp < 1 || fp > rlen) { bad = 1; break; }       // Java: ArrayIndexOutOfBoundsException
                                bases[fp - 1] = norm_base((uint8_t)a.f_val[f]);
                            }
                }
            }
        }

        // ---- qualities without a stored array: restoreQualityScores (:287-305)
        if (!(rfl & CRAMGPU_NF_FORCE_QS)) {
            const uint8_t dflt = '?' - '!';
            for (int64_t i = 0; i < rlen; i++) quals[i] = dflt;
            uint32_t any = 0;
            for (uint64_t f = f0; f < f1; f++) {
                const uint8_t op = a.f_op[f];
                if (op != 'Q' && op != 'B') continue;
                const int64_t fp = a.f_pos[f];
                if (fp < 1 || fp > rlen) { bad = 1; break; }                       // Java: ArrayIndexOutOfBoundsException (:296-302)
                quals[fp - 1] = (uint8_t)(op == 'Q' ? a.f_val[f] : a.f_aux[f]);
                any = 1;
            }
            a.qual_star[r] = any == 0;
        } else if (rlen >= (1ll << 31)) {
            // (k_norm_simple leaves reads of 2 GiB and more alone)
            const uint64_t qo = a.read_qs_off[r];
            uint32_t present = 0;
            for (int64_t i = 0; i < rlen; i++) {
                uint8_t sc = qo != ~0ull ? a.qs[qo + i] : (uint8_t)0xFF;
                if (sc == 0xFF) sc = '?' - '!'; else present = 1;
                quals[i] = sc;
            }
            a.qual_star[r] = present == 0;
        }
        a.status[r] = bad ? CRAMGPU_ERR_CORRUPT : CRAMGPU_OK;
    }
}


// flag[0] = 1 when any record failed (the host then looks for the first one)
__global__ void __launch_bounds__(256) k_any_status(const int32_t *__restrict__ status, uint64_t n, int32_t *__restrict__ flag)
{
    bool bad = false;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) bad = bad || status[i] != 0;
    if (__any_sync(0xFFFFFFFFu, bad) && (threadIdx.x & 31) == 0) flag[0] = 1;
}

// ------------------------------------------------------------------------------------------------ host side

#include <string.h>

// QualityScorePreservation.parsePolicies / parseSinglePolicy / readParam / readTreatment
// (src/main/java/htsjdk/samtools/cram/lossy/QualityScorePreservation.java:141-256)
extern "C" int cramgpu_parse_lossy_spec(const char *spec, cramgpu_qs_policy *out, int cap)
{
    if (!spec || !out || cap <= 0) return CRAMGPU_ERR_ARG;
    int n = 0;
    size_t pos = 0, total = strlen(spec);
    while (pos <= total) {
        const char *dash = strchr(spec + pos, '-');
        size_t end = dash ? (size_t)(dash - spec) : total;
        if (end > pos) {
            if (n == cap) return CRAMGPU_ERR_ARG;
            cramgpu_qs_policy p;
            memset(&p, 0, sizeof p);
            bool have_treatment = false;
            size_t i = pos;
            auto read_param = [&]() { int v = 0; while (i < end && spec[i] >= '0' && spec[i] <= '9') v = v * 10 + (spec[i++] - '0'); return v; };
            auto read_treatment = [&]() {
                int prm = read_param();
                have_treatment = true;
                p.treat_param = prm;
                p.treatment = prm == 0 ? 2 : prm == 40 ? 0 : 1;        // 0 -> drop, 40 -> preserve, else bin(param)
            };
            auto add_cat = [&](int cat) { if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = cat; };
            while (i < end) {
                char code = spec[i++];
                switch (code) {
                case 'R': add_cat(1); read_treatment(); break;
                case 'N': add_cat(2); read_treatment(); break;
                case 'D': add_cat(3); read_treatment(); break;
                case 'I': add_cat(4); read_treatment(); break;
                case 'M': p.read_cat = 4; p.read_param = read_param(); break;
                case 'm': p.read_cat = 3; p.read_param = read_param(); break;
                case 'U': p.read_cat = 2; read_treatment(); break;
                case '_': read_treatment(); break;
                case '*': p.read_cat = 1; read_treatment(); break;
                case 'X': add_cat(5 | (read_param() << 8)); break;                     // lowerThanCoverage(n): reads no treatment [:210-213]
                case 'P': add_cat(6 | (read_param() << 8)); read_treatment(); break;   // pileup(n) [:230-234]
                default: return CRAMGPU_ERR_ARG;                        // "Unknown read or base category: " + code
                }
                if (!have_treatment) { p.treatment = 0; p.treat_param = 40; have_treatment = true; }
            }
            out[n++] = p;
        }
        if (!dash) break;
        pos = end + 1;
    }
    // Collections.sort by treatment ordinal, descending (stable): DROP, BIN, PRESERVE
    for (int i = 1; i < n; i++) {
        cramgpu_qs_policy p = out[i];
        int j = i - 1;
        while (j >= 0 && out[j].treatment < p.treatment) { out[j + 1] = out[j]; j--; }
        out[j + 1] = p;
    }
    return n;
}

struct RecFault { RecArgs a; PolSet ps; };       // what k_rec_bin_fault needs, kept from the call (ctx->rec_fault)

static int make_polset(cramgpu_ctx *ctx, const cramgpu_qs_policy *pol, int n, PolSet &ps)
{
    memset(&ps, 0, sizeof ps);
    ps.n = n < 0 ? -1 : n;
    if (n > MAX_POL) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "more than %d lossy policies", MAX_POL);
    for (int k = 0; k < n; k++) {
        if (!pol) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null policies");
        ps.p[k] = pol[k];
        if (pol[k].n_base_cats < 0 || pol[k].n_base_cats > 8) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "bad policy");
        if (pol[k].n_base_cats) ps.per_base = 1;
        if (pol[k].treatment == 1) ps.any_bin = 1;
        for (int b = 0; b < pol[k].n_base_cats; b++) {
            const int cat = pol[k].base_cat[b] & 0xFF;
            if (cat < 1 || cat > 6) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "bad policy");
            if (cat >= 5) ps.ext_bit[k][b] = (uint8_t)(1 + ps.n_ext++);       // the order of tracks_selectors (tracks.cu)
        }
    }
    if (ps.n_ext > 7) return ctx_fail(ctx, CRAMGPU_ERR_UNSUPPORTED, "more than 7 X / P selectors in a lossy model");
    return CRAMGPU_OK;
}

// upload a small host table through the pinned plan buffer
struct Uploader {
    cramgpu_ctx *ctx; cudaStream_t st; size_t po = 0;
    cudaError_t up(const void *src, size_t bytes, void *dst)
    {
        if (!bytes) return cudaSuccess;
        memcpy((uint8_t *)ctx->pin_meta + po, src, bytes);
        cudaError_t e = cudaMemcpyAsync(dst, (uint8_t *)ctx->pin_meta + po, bytes, cudaMemcpyHostToDevice, st);
        po += rup(bytes, 16);
        return e;
    }
};

// device-pointer implementation shared by both encode entry points.  R holds device pointers for the big arrays.
static int features_encode_impl(cramgpu_ctx *ctx, const cramgpu_reads *R, const cramgpu_qs_policy *pol, int npol,
                                cramgpu_features *out, uint64_t *h_stream_off, uint64_t *h_stream_len, uint64_t *series_span,
                                cudaStream_t st)
{
    if (!R || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    memset(out, 0, sizeof *out);
    const uint64_t n = R->n_reads;
    const uint32_t nc = R->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!R->cont_off || !R->ref_off || !R->ref_len || !R->ref || !R->bases || !R->quals || !R->read_off || !R->cigar_off ||
        !R->pos || !R->flag || !R->mapq)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (R->cont_off[0] != 0 || R->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    for (uint32_t c = 0; c < nc; c++) {
        if (R->cont_off[c] > R->cont_off[c + 1]) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off not monotonic");
        if (R->ref_off[c] + R->ref_len[c] > R->ref_total) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "reference window of container %u outside ref", c);
    }
    PolSet ps;
    if (int rc = make_polset(ctx, pol, npol, ps)) return rc;
    // X / P selectors: the per-base bytes cramgpu_tracks_update_dev left for exactly these reads and selectors
    const uint8_t *qs_ext = nullptr;
    if (ps.n_ext) {
        int32_t kind[7], param[7];
        const int ns = tracks_selectors(pol, npol, kind, param);
        bool same = ctx->trk_armed && ctx->trk_reads == n && ctx->trk_nsel == ns;
        for (int e = 0; same && e < ns; e++) same = kind[e] == ctx->trk_kind[e] && param[e] == ctx->trk_param[e];
        if (!same) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "The quality score lossy model requires reference tracks.");   // CramSerilization.java:121
        qs_ext = (const uint8_t *)ctx->trk_ext.p;
        ctx->trk_armed = false;
    }

    Carver cv;
    size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_roff = cv.take(8 * (size_t)nc), o_rlen = cv.take(8 * (size_t)nc);
    size_t o_cnt = cv.take(4 * (size_t)NCNT * n), o_force = cv.take(n);
    size_t o_subst = cv.take(8 * 25 * (size_t)nc), o_status = cv.take(16), o_cln = cv.take(16), o_tln = cv.take(16), o_wln = cv.take(16);
    size_t o_zero_end = cv.off;
    size_t o_clr = cv.take(8 * (size_t)n), o_clc = cv.take(4 * (size_t)n), o_tlr = cv.take(8 * (size_t)n), o_tlc = cv.take(4 * (size_t)n),
           o_wlr = cv.take(8 * (size_t)n), o_wlc = cv.take(4 * (size_t)n);
    size_t o_nref = cv.take(R->ref_total + 512) + 256;                          // slack on both sides of the normalised copy
    size_t o_codes = cv.take(32 * (size_t)nc), o_matrix = cv.take(8 * (size_t)nc);
    size_t o_totals = cv.take(8 * (size_t)NCNT * nc), o_base = cv.take(8 * (size_t)NCNT * nc);
    size_t o_soff = cv.take(8 * (size_t)CRAMGPU_S_COUNT * nc), o_slen = cv.take(8 * (size_t)CRAMGPU_S_COUNT * nc);
    size_t o_grand = cv.take(512);
    if (int rc = pool_reserve(ctx, ctx->rec_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->rec_scratch.p;

    size_t plan_bytes = 8 * (size_t)(3 * nc + 1) + 256;
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, plan_bytes)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));
    Uploader U{ctx, st};
    CUDA_TRY(ctx, U.up(R->cont_off, 8 * (size_t)(nc + 1), S + o_cont));
    CUDA_TRY(ctx, U.up(R->ref_off, 8 * (size_t)nc, S + o_roff));
    CUDA_TRY(ctx, U.up(R->ref_len, 8 * (size_t)nc, S + o_rlen));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_cnt, 0, o_zero_end - o_cnt, st));      // counters, flags, histogram, status

    RecArgs a;
    memset(&a, 0, sizeof a);
    a.ref = R->ref; a.bases = R->bases; a.quals = R->quals; a.read_off = R->read_off; a.cigar_off = R->cigar_off;
    a.cigar = R->cigar; a.pos = R->pos; a.flag = R->flag; a.mapq = R->mapq; a.rflags = R->rflags;
    a.qs_ext = qs_ext;
    a.cont_off = (const uint64_t *)(S + o_cont); a.ref_off = (const uint64_t *)(S + o_roff); a.ref_len = (const uint64_t *)(S + o_rlen);
    a.n_containers = nc; a.n_reads = n;
    a.cnt = (uint32_t *)(S + o_cnt); a.force = S + o_force; a.subst = (unsigned long long *)(S + o_subst);
    a.codes = S + o_codes; a.base = (const uint64_t *)(S + o_base); a.status = (int32_t *)(S + o_status);
    uint64_t *d_totals = (uint64_t *)(S + o_totals), *d_soff = (uint64_t *)(S + o_soff), *d_slen = (uint64_t *)(S + o_slen);
    uint64_t *d_grand = (uint64_t *)(S + o_grand);
    a.nref = S + o_nref; a.totals = d_totals;
    a.clist_r = (uint64_t *)(S + o_clr); a.clist_c = (uint32_t *)(S + o_clc); a.clist_n = (unsigned long long *)(S + o_cln);
    a.tlist_r = (uint64_t *)(S + o_tlr); a.tlist_c = (uint32_t *)(S + o_tlc); a.tlist_n = (unsigned long long *)(S + o_tln);
    a.wlist_r = (uint64_t *)(S + o_wlr); a.wlist_c = (uint32_t *)(S + o_wlc); a.wlist_n = (unsigned long long *)(S + o_wln);

    uint64_t max_reads = 0, max_win = 0, sum_win = 0;
    for (uint32_t c = 0; c < nc; c++) {
        max_reads = std::max(max_reads, R->cont_off[c + 1] - R->cont_off[c]);
        max_win = std::max(max_win, R->ref_len[c]); sum_win += R->ref_len[c];
    }
    // normalised reference: the windows, or the whole array when the windows add up to more than that
    {
        if (nc <= 65535 && sum_win < R->ref_total)
            LAUNCH(ctx, st, "k_ref_norm", k_ref_norm<<<dim3((uint32_t)((max_win + 31) / 4096 + 1), nc), 256, 0, st>>>(R->ref, S + o_nref, a.ref_off, a.ref_len, 0));
        else
            LAUNCH(ctx, st, "k_ref_norm", k_ref_norm<<<dim3((uint32_t)(R->ref_total / 4096 + 1), 1), 256, 0, st>>>(R->ref, S + o_nref, nullptr, nullptr, R->ref_total));
    }
    const dim3 grid = nc <= 65535 ? dim3((uint32_t)((max_reads + 3) / 4), nc) : dim3((uint32_t)((n + 3) / 4), 1);      // warp per read
    const dim3 sgrid = nc <= 65535 ? dim3((uint32_t)((max_reads + 127) / 128), nc) : dim3((uint32_t)((n + 127) / 128), 1);   // thread per read
    const uint32_t wgrid = (uint32_t)std::min<uint64_t>((n + 3) / 4, 148 * 16);
    const uint32_t mgrid = (uint32_t)std::min<uint64_t>((n + 127) / 128, 148 * 8);
    if (ps.per_base) {
        LAUNCH(ctx, st, "k_rec_simple_count", k_rec_simple<false, true><<<sgrid, 128, 0, st>>>(a, ps));
        LAUNCH(ctx, st, "k_rec_multi_count", k_rec_multi<false, true><<<mgrid, 128, 0, st>>>(a, ps, a.tlist_r, a.tlist_c, a.tlist_n));
    } else {
        LAUNCH(ctx, st, "k_rec_simple_count", k_rec_simple<false, false><<<sgrid, 128, 0, st>>>(a, ps));
        LAUNCH(ctx, st, "k_rec_multi_count", k_rec_multi<false, false><<<mgrid, 128, 0, st>>>(a, ps, a.tlist_r, a.tlist_c, a.tlist_n));
    }
    LAUNCH(ctx, st, "k_rec_walk_count", k_rec_walk<false><<<wgrid, 128, 0, st>>>(a, ps));
    LAUNCH(ctx, st, "k_rec_matrix", k_rec_matrix<<<(nc + 127) / 128, 128, 0, st>>>(a.subst, nc, S + o_codes, S + o_matrix));
    LAUNCH(ctx, st, "k_rec_scan", k_rec_scan<<<dim3(nc, NCNT), 256, 0, st>>>(a.cnt, a.cont_off, n, d_totals));
    LAUNCH(ctx, st, "k_rec_layout", k_rec_layout<<<1, 1024, 0, st>>>(d_totals, nc, (uint64_t *)(S + o_base), d_soff, d_slen, d_grand));
    CUDA_TRY(ctx, cudaGetLastError());

    // totals (and the per-container tables the caller needs) come back to the host
    size_t tab = 8 * (size_t)CRAMGPU_S_COUNT * nc;
    if (int rc = pinned_reserve(ctx, &ctx->pin_rec, &ctx->pin_rec_cap, 2 * tab + 512)) return rc;
    uint8_t *pr = (uint8_t *)ctx->pin_rec;
    CUDA_TRY(ctx, cudaMemcpyAsync(pr, d_grand, 256, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(pr + 256, S + o_status, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(pr + 320, d_soff, tab, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(pr + 320 + tab, d_slen, tab, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint64_t *grand = (const uint64_t *)pr;
    int32_t dstatus = *(const int32_t *)(pr + 256);
    if (dstatus == CRAMGPU_ERR_ARG) return ctx_fail(ctx, dstatus, "Unsupported cigar operator");    // else: a binning fault, reported at the end
    const uint64_t sbytes = grand[0], nfeat = grand[1], paylen = grand[2];
    if (h_stream_off) memcpy(h_stream_off, pr + 320, tab);
    if (h_stream_len) memcpy(h_stream_len, pr + 320 + tab, tab);
    if (series_span) memcpy(series_span, grand + 4, 16 * CRAMGPU_S_COUNT);      // (start, extent) of every series in streams

    Carver co;
    size_t q_op = co.take(nfeat + 16), q_pos = co.take(4 * nfeat + 16), q_val = co.take(4 * nfeat + 16), q_aux = co.take(4 * nfeat + 16);
    size_t q_foff = co.take(8 * (size_t)(n + 1)), q_pay = co.take(paylen + 16), q_str = co.take(sbytes + 16);
    size_t q_qso = co.take(8 * (size_t)n), q_bao = co.take(8 * (size_t)n);
    if (int rc = pool_reserve(ctx, ctx->rec_out, co.off)) return rc;
    uint8_t *O = (uint8_t *)ctx->rec_out.p;
    a.f_op = O + q_op; a.f_pos = (int32_t *)(O + q_pos); a.f_val = (int32_t *)(O + q_val); a.f_aux = (int32_t *)(O + q_aux);
    a.feat_off = (uint64_t *)(O + q_foff); a.payload = O + q_pay; a.streams = O + q_str;
    a.read_qs_off = (uint64_t *)(O + q_qso); a.read_ba_off = (uint64_t *)(O + q_bao);
    if (ps.per_base) {
        LAUNCH(ctx, st, "k_rec_simple_write", k_rec_simple<true, true><<<sgrid, 128, 0, st>>>(a, ps));
        LAUNCH(ctx, st, "k_rec_multi_write", k_rec_multi<true, true><<<mgrid, 128, 0, st>>>(a, ps, a.tlist_r, a.tlist_c, a.tlist_n));
        LAUNCH(ctx, st, "k_rec_multi_write", k_rec_multi<true, true><<<mgrid, 128, 0, st>>>(a, ps, a.wlist_r, a.wlist_c, a.wlist_n));
    } else {
        LAUNCH(ctx, st, "k_rec_simple_write", k_rec_simple<true, false><<<sgrid, 128, 0, st>>>(a, ps));
        LAUNCH(ctx, st, "k_rec_multi_write", k_rec_multi<true, false><<<mgrid, 128, 0, st>>>(a, ps, a.tlist_r, a.tlist_c, a.tlist_n));
        LAUNCH(ctx, st, "k_rec_multi_write", k_rec_multi<true, false><<<mgrid, 128, 0, st>>>(a, ps, a.wlist_r, a.wlist_c, a.wlist_n));
    }
    LAUNCH(ctx, st, "k_rec_walk_write", k_rec_walk<true><<<wgrid, 128, 0, st>>>(a, ps));
    if (ps.n >= 0 && !ps.per_base) LAUNCH(ctx, st, "k_rec_forced_flat", k_rec_forced_flat<<<sgrid, 128, 0, st>>>(a, ps));
    else if (ps.n >= 0) LAUNCH(ctx, st, "k_rec_forced_scores", k_rec_forced_scores<<<grid, 128, 0, st>>>(a, ps));
    CUDA_TRY(ctx, cudaMemcpyAsync(a.feat_off + n, d_grand + 1, 8, cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(ctx, cudaGetLastError());
    *(volatile int32_t *)(pr + 264) = 0;
    if (ps.any_bin) {
        // a score outside the binning table: the status word says so once the stream has drained (the callers synchronise and
        // then run features_encode_fault)
        CUDA_TRY(ctx, cudaMemcpyAsync(pr + 264, S + o_status, 4, cudaMemcpyDeviceToHost, st));
        if (!ctx->rec_fault) ctx->rec_fault = malloc(sizeof(RecFault));
        if (!ctx->rec_fault) return ctx_fail(ctx, CRAMGPU_ERR_CUDA, "out of host memory");
        ((RecFault *)ctx->rec_fault)->a = a; ((RecFault *)ctx->rec_fault)->ps = ps;
    }

    out->n_feat = nfeat; out->f_op = a.f_op; out->f_pos = a.f_pos; out->f_val = a.f_val; out->f_aux = a.f_aux;
    out->feat_off = a.feat_off; out->payload = a.payload; out->payload_len = paylen; out->force_qs = a.force;
    out->matrix = S + o_matrix; out->streams = a.streams; out->streams_len = sbytes; out->stream_off = d_soff;
    out->stream_len = d_slen; out->read_qs_off = a.read_qs_off; out->read_ba_off = a.read_ba_off;
    return CRAMGPU_OK;
}

// after the stream has drained: did a BIN treatment meet a score of 110 or more?  Then mark the records and fail the call the
// way the reference does (ArrayIndexOutOfBoundsException out of Binning.Illumina_binning_matrix).  The results stay valid
// for the other records.
static int features_encode_fault(cramgpu_ctx *ctx, cudaStream_t st)
{
    if (*(volatile int32_t *)((uint8_t *)ctx->pin_rec + 264) == 0) return CRAMGPU_OK;
    const RecFault *f = (const RecFault *)ctx->rec_fault;
    LAUNCH(ctx, st, "k_rec_bin_fault", k_rec_bin_fault<<<(uint32_t)((f->a.n_reads + 3) / 4), 128, 0, st>>>(f->a, f->ps));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_ERR_UNSUPPORTED;
}
#define BIN_FAULT_MSG "ArrayIndexOutOfBoundsException: quality score >= 110 under a binning treatment (bit 7 of force_qs marks the records)"

extern "C" int cramgpu_features_encode_dev(cramgpu_ctx *ctx, const cramgpu_reads *reads, const cramgpu_qs_policy *policies,
                                           int n_policies, cramgpu_features *out, uint64_t *h_stream_off,
                                           uint64_t *h_stream_len, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    if (int rc = features_encode_impl(ctx, reads, policies, n_policies, out, h_stream_off, h_stream_len, nullptr, st)) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (features_encode_fault(ctx, st)) {
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
        return ctx_fail(ctx, CRAMGPU_ERR_UNSUPPORTED, BIN_FAULT_MSG);
    }
    return CRAMGPU_OK;
}

// copy helper for the host entry points: place the caller's arrays back to back in one device buffer.  Arrays that are in
// device memory already are used where they are; arrays in memory the caller pinned go to the DMA engine directly; the
// others are staged through `pin`.
struct Stager {
    MemMap mm;
    std::vector<const void *> src; std::vector<size_t> bytes, off, poff; std::vector<uint8_t> kind;
    size_t total = 0, ptotal = 0;          // bytes of the device buffer / of the pinned staging buffer
    size_t add(const void *p, size_t b)
    {
        MemKind k = (p && b) ? mm.kind(p) : MEM_PAGEABLE;
        if (k == MEM_PINNED && b < (64u << 10)) k = MEM_PAGEABLE;          // small: cheaper to stage than to copy on its own
        off.push_back(total); poff.push_back(ptotal); src.push_back(p); bytes.push_back(b); kind.push_back((uint8_t)k);
        if (k != MEM_DEVICE) total += rup(b ? b : 1, 256);
        if (k == MEM_PAGEABLE && p) ptotal += rup(b ? b : 1, 256);
        return off.size() - 1;
    }
    size_t reserve(size_t b) { off.push_back(total); poff.push_back(ptotal); src.push_back(nullptr); bytes.push_back(0); kind.push_back(MEM_PAGEABLE); total += rup(b ? b : 1, 256); return off.size() - 1; }
    bool on_device(size_t k) const { return kind[k] == MEM_DEVICE; }
    // where entry k is (or will be, once to_device has run) in device memory
    uint8_t *at(uint8_t *D, size_t k) const { return kind[k] == MEM_DEVICE ? (uint8_t *)src[k] : D + off[k]; }
    int to_device(cramgpu_ctx *ctx, uint8_t *pin, uint8_t *D, cudaStream_t st)
    {
        std::vector<CopyJob> jobs;
        for (size_t k = 0; k < src.size(); k++)
            if (bytes[k] && src[k] && kind[k] == MEM_PAGEABLE) jobs.push_back(CopyJob{pin + poff[k], src[k], bytes[k]});
        par_memcpy(jobs);
        for (size_t k = 0; k < src.size(); k++) {
            if (!bytes[k] || !src[k] || kind[k] == MEM_DEVICE) continue;
            CUDA_TRY(ctx, memcpy_pieces(D + off[k], kind[k] == MEM_PINNED ? src[k] : (const void *)(pin + poff[k]), bytes[k], cudaMemcpyHostToDevice, st));
        }
        return CRAMGPU_OK;
    }
};

// host-memory encode: copy in, run, copy out.  host_series: bit k = series k comes back to the host; dev_out (optional)
// receives the same results as device pointers.
static int features_encode_host(cramgpu_ctx *ctx, const cramgpu_reads *R, const cramgpu_qs_policy *policies, int n_policies,
                                uint32_t host_series, cramgpu_features *out, cramgpu_features *dev_out)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!R || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    memset(out, 0, sizeof *out);
    if (dev_out) memset(dev_out, 0, sizeof *dev_out);
    const uint64_t n = R->n_reads;
    const uint32_t nc = R->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!R->read_off || !R->cigar_off || !R->bases || !R->quals || !R->pos || !R->flag || !R->mapq || !R->ref)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    const uint64_t nb = R->read_off[n], ncig = R->cigar_off[n];           // the two offset tables are read here: host memory
    Stager sg;
    size_t i_ref = sg.add(R->ref, R->ref_total), i_bases = sg.add(R->bases, nb), i_quals = sg.add(R->quals, nb);
    size_t i_roff = sg.add(R->read_off, 8 * (n + 1)), i_cig = sg.add(R->cigar, 4 * ncig), i_coff = sg.add(R->cigar_off, 8 * (n + 1));
    size_t i_pos = sg.add(R->pos, 4 * n), i_flag = sg.add(R->flag, 2 * n), i_mapq = sg.add(R->mapq, n);
    size_t i_rfl = R->rflags ? sg.add(R->rflags, n) : 0;
    if (int rc = pinned_reserve(ctx, &ctx->pin_in, &ctx->pin_in_cap, sg.ptotal + 16)) return rc;
    if (int rc = pool_reserve(ctx, ctx->dev_in, sg.total)) return rc;
    uint8_t *pin = (uint8_t *)ctx->pin_in, *D = (uint8_t *)ctx->dev_in.p;
    if (int rc = sg.to_device(ctx, pin, D, st)) return rc;
    cramgpu_reads dev = *R;
    dev.ref = sg.at(D, i_ref); dev.bases = sg.at(D, i_bases); dev.quals = sg.at(D, i_quals); dev.read_off = (const uint64_t *)sg.at(D, i_roff);
    dev.cigar = (const uint32_t *)sg.at(D, i_cig); dev.cigar_off = (const uint64_t *)sg.at(D, i_coff); dev.pos = (const int32_t *)sg.at(D, i_pos);
    dev.flag = (const uint16_t *)sg.at(D, i_flag); dev.mapq = sg.at(D, i_mapq); dev.rflags = R->rflags ? sg.at(D, i_rfl) : nullptr;
    cramgpu_features df;
    uint64_t span[2 * CRAMGPU_S_COUNT];
    if (int rc = features_encode_impl(ctx, &dev, policies, n_policies, &df, nullptr, nullptr, span, st)) return rc;
    if (dev_out) *dev_out = df;

    // results to pinned host memory owned by the context.  The streams buffer keeps its device layout; a series is one
    // contiguous piece of it, copied only when asked for.
    const size_t tab = 8 * (size_t)CRAMGPU_S_COUNT * nc;
    Stager so;
    size_t r_op = so.reserve(df.n_feat), r_pos = so.reserve(4 * df.n_feat), r_val = so.reserve(4 * df.n_feat);
    size_t r_aux = so.reserve(4 * df.n_feat), r_foff = so.reserve(8 * (n + 1)), r_pay = so.reserve(df.payload_len);
    size_t r_force = so.reserve(n), r_mtx = so.reserve(8 * (size_t)nc), r_str = so.reserve(df.streams_len);
    size_t r_soff = so.reserve(tab), r_slen = so.reserve(tab);
    size_t r_qso = so.reserve(8 * n), r_bao = so.reserve(8 * n);
    // results get a pinned buffer of their own: they must survive the following block-codec / normaliser calls,
    // which recycle pin_in / pin_out
    if (int rc = pinned_reserve(ctx, &ctx->pin_feat, &ctx->pin_feat_cap, so.total)) return rc;
    uint8_t *po = (uint8_t *)ctx->pin_feat;
    auto down = [&](size_t slot, const void *d, size_t bytes) -> cudaError_t {
        return bytes ? memcpy_pieces(po + so.off[slot], d, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    // CRAMGPU_FEATURES_ON_DEVICE_ONLY: a host that hands the series straight to the block layer has no use for the feature arrays
    const bool arrays = !(dev_out && (host_series & CRAMGPU_FEATURES_ON_DEVICE_ONLY));
    if (arrays) {
        CUDA_TRY(ctx, down(r_op, df.f_op, df.n_feat)); CUDA_TRY(ctx, down(r_pos, df.f_pos, 4 * df.n_feat));
        CUDA_TRY(ctx, down(r_val, df.f_val, 4 * df.n_feat)); CUDA_TRY(ctx, down(r_aux, df.f_aux, 4 * df.n_feat));
        CUDA_TRY(ctx, down(r_foff, df.feat_off, 8 * (n + 1))); CUDA_TRY(ctx, down(r_pay, df.payload, df.payload_len));
        CUDA_TRY(ctx, down(r_qso, df.read_qs_off, 8 * n)); CUDA_TRY(ctx, down(r_bao, df.read_ba_off, 8 * n));
    }
    CUDA_TRY(ctx, down(r_force, df.force_qs, n)); CUDA_TRY(ctx, down(r_mtx, df.matrix, 8 * (size_t)nc));
    CUDA_TRY(ctx, down(r_soff, df.stream_off, tab)); CUDA_TRY(ctx, down(r_slen, df.stream_len, tab));
    for (int k = 0; k < CRAMGPU_S_COUNT; k++)
        if (((host_series >> k) & 1) && span[2 * k + 1])
            CUDA_TRY(ctx, memcpy_pieces(po + so.off[r_str] + span[2 * k], df.streams + span[2 * k], span[2 * k + 1], cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const bool fault = features_encode_fault(ctx, st) != CRAMGPU_OK;
    if (fault) {
        CUDA_TRY(ctx, down(r_force, df.force_qs, n));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    *out = df;
    out->f_op = po + so.off[r_op]; out->f_pos = (const int32_t *)(po + so.off[r_pos]); out->f_val = (const int32_t *)(po + so.off[r_val]);
    out->f_aux = (const int32_t *)(po + so.off[r_aux]); out->feat_off = (const uint64_t *)(po + so.off[r_foff]); out->payload = po + so.off[r_pay];
    out->force_qs = po + so.off[r_force]; out->matrix = po + so.off[r_mtx]; out->streams = po + so.off[r_str];
    out->stream_off = (const uint64_t *)(po + so.off[r_soff]); out->stream_len = (const uint64_t *)(po + so.off[r_slen]);
    out->read_qs_off = (const uint64_t *)(po + so.off[r_qso]); out->read_ba_off = (const uint64_t *)(po + so.off[r_bao]);
    if (!arrays) { out->f_op = nullptr; out->f_pos = out->f_val = out->f_aux = nullptr; out->feat_off = nullptr; out->payload = nullptr;
                   out->read_qs_off = out->read_ba_off = nullptr; }
    if (fault) return ctx_fail(ctx, CRAMGPU_ERR_UNSUPPORTED, BIN_FAULT_MSG);
    return CRAMGPU_OK;
}

extern "C" int cramgpu_features_encode(cramgpu_ctx *ctx, const cramgpu_reads *R, const cramgpu_qs_policy *policies,
                                       int n_policies, cramgpu_features *out)
{
    return features_encode_host(ctx, R, policies, n_policies, 0xFFFFFFFFu, out, nullptr);
}

extern "C" int cramgpu_features_encode_resident(cramgpu_ctx *ctx, const cramgpu_reads *R, const cramgpu_qs_policy *policies,
                                                int n_policies, uint32_t host_series, cramgpu_features *out,
                                                cramgpu_features *dev_out)
{
    if (!dev_out) return ctx ? ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument") : CRAMGPU_ERR_ARG;
    return features_encode_host(ctx, R, policies, n_policies, host_series, out, dev_out);
}

static int normalize_impl(cramgpu_ctx *ctx, const cramgpu_norm_in *in, uint8_t *d_bases, uint8_t *d_quals,
                          uint8_t *d_star, int32_t *d_status, cudaStream_t st)
{
    if (!in) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->cont_off || !in->win_off || !in->win_len || !in->win_start0 || !in->pos || !in->read_off || !in->rflags ||
        !in->feat_off || !d_bases || !d_quals || !d_star || !d_status)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->cont_off[0] != 0 || in->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    for (uint32_t c = 0; c < nc; c++)
        if (in->win_off[c] + in->win_len[c] > in->ref_total) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "reference region of container %u outside ref", c);
    Carver cv;
    size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_woff = cv.take(8 * (size_t)nc), o_wlen = cv.take(8 * (size_t)nc), o_ws = cv.take(8 * (size_t)nc);
    size_t o_cln = cv.take(16), o_clr = cv.take(8 * (size_t)n), o_clc = cv.take(4 * (size_t)n);
    size_t o_nref = cv.take(in->ref_total + 512) + 256;
    if (int rc = pool_reserve(ctx, ctx->rec_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->rec_scratch.p;
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, 8 * (size_t)(4 * nc + 1) + 256)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));
    Uploader U{ctx, st};
    CUDA_TRY(ctx, U.up(in->cont_off, 8 * (size_t)(nc + 1), S + o_cont));
    CUDA_TRY(ctx, U.up(in->win_off, 8 * (size_t)nc, S + o_woff));
    CUDA_TRY(ctx, U.up(in->win_len, 8 * (size_t)nc, S + o_wlen));
    CUDA_TRY(ctx, U.up(in->win_start0, 8 * (size_t)nc, S + o_ws));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));
    NormArgs a;
    memset(&a, 0, sizeof a);
    a.n_reads = n; a.n_containers = nc;
    a.cont_off = (const uint64_t *)(S + o_cont); a.win_off = (const uint64_t *)(S + o_woff); a.win_len = (const uint64_t *)(S + o_wlen);
    a.win_start0 = (const int64_t *)(S + o_ws);
    a.ref = in->ref; a.matrix = in->matrix; a.pos = in->pos; a.read_off = in->read_off; a.rflags = in->rflags; a.feat_off = in->feat_off;
    a.f_op = in->f_op; a.f_pos = in->f_pos; a.f_val = in->f_val; a.f_aux = in->f_aux; a.payload = in->payload;
    a.qs = in->qs; a.read_qs_off = in->read_qs_off; a.ba = in->ba; a.read_ba_off = in->read_ba_off;
    a.bases = d_bases; a.quals = d_quals; a.qual_star = d_star; a.status = d_status;
    uint64_t max_reads = 0, max_win = 0, sum_win = 0;
    for (uint32_t c = 0; c < nc; c++) {
        max_reads = std::max(max_reads, in->cont_off[c + 1] - in->cont_off[c]);
        max_win = std::max(max_win, in->win_len[c]); sum_win += in->win_len[c];
    }
    a.clist_r = (uint64_t *)(S + o_clr); a.clist_c = (uint32_t *)(S + o_clc); a.clist_n = (unsigned long long *)(S + o_cln);
    a.nref = S + o_nref; a.ref_total = in->ref_total; a.payload_len = in->payload ? in->payload_len : 0;
    CUDA_TRY(ctx, cudaMemsetAsync(S + o_cln, 0, 16, st));
    if (nc <= 65535 && sum_win < in->ref_total)
        LAUNCH(ctx, st, "k_ref_norm", k_ref_norm<<<dim3((uint32_t)((max_win + 31) / 4096 + 1), nc), 256, 0, st>>>(in->ref, S + o_nref, a.win_off, a.win_len, 0));
    else
        LAUNCH(ctx, st, "k_ref_norm", k_ref_norm<<<dim3((uint32_t)(in->ref_total / 4096 + 1), 1), 256, 0, st>>>(in->ref, S + o_nref, nullptr, nullptr, in->ref_total));
    const dim3 sgrid = nc <= 65535 ? dim3((uint32_t)((max_reads + 127) / 128), nc) : dim3((uint32_t)((n + 127) / 128), 1);
    LAUNCH(ctx, st, "k_norm_simple", k_norm_simple<<<sgrid, 128, 0, st>>>(a));
    LAUNCH(ctx, st, "k_norm", k_norm<<<(uint32_t)std::min<uint64_t>((n + 127) / 128, 148 * 8), 128, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    return CRAMGPU_OK;
}

extern "C" int cramgpu_normalize_decode_dev(cramgpu_ctx *ctx, const cramgpu_norm_in *in, uint8_t *d_bases, uint8_t *d_quals,
                                            uint8_t *d_qual_star, int32_t *d_status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return normalize_impl(ctx, in, d_bases, d_quals, d_qual_star, d_status, stream ? (cudaStream_t)stream : ctx->stream);
}

extern "C" int cramgpu_normalize_decode(cramgpu_ctx *ctx, const cramgpu_norm_in *in, uint8_t *bases, uint8_t *quals,
                                        uint8_t *qual_star, int32_t *status)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !bases || !quals || !qual_star || !status) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->read_off || !in->feat_off) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    Stager sg;
    // read_off is read here (host memory); feat_off / read_qs_off / read_ba_off may also be the device-resident result of
    // cramgpu_features_unpack_resident, in which case the feature arrays are device memory too and need no extents
    const bool foff_dev = sg.mm.kind(in->feat_off) == MEM_DEVICE;
    const bool qso_dev = in->read_qs_off && sg.mm.kind(in->read_qs_off) == MEM_DEVICE, bao_dev = in->read_ba_off && sg.mm.kind(in->read_ba_off) == MEM_DEVICE;
    const uint64_t nb = in->read_off[n], nf = foff_dev ? 1 : in->feat_off[n];      // device arrays are used where they are: any non-zero extent
    // extents of the side buffers in host memory: the furthest byte any record refers to (buffers in device memory are
    // used where they are, whatever their extent)
    uint64_t qs_len = 0, ba_len = 0, pay_len = 0;
    const bool qs_dev = in->qs && sg.mm.kind(in->qs) == MEM_DEVICE, ba_dev = in->ba && sg.mm.kind(in->ba) == MEM_DEVICE;
    const bool pay_dev = in->payload && sg.mm.kind(in->payload) == MEM_DEVICE;
    if ((in->read_qs_off && !qs_dev && !qso_dev) || (in->read_ba_off && !ba_dev && !bao_dev))
        for (uint64_t r = 0; r < n; r++) {
            uint64_t len = in->read_off[r + 1] - in->read_off[r];
            if (in->read_qs_off && !qs_dev && !qso_dev && in->read_qs_off[r] != ~0ull) qs_len = std::max(qs_len, in->read_qs_off[r] + len);
            if (in->read_ba_off && !ba_dev && !bao_dev && in->read_ba_off[r] != ~0ull) ba_len = std::max(ba_len, in->read_ba_off[r] + len);
        }
    if (!pay_dev) pay_len = in->payload ? in->payload_len : 0;
    size_t i_ref = sg.add(in->ref, in->ref_total), i_mtx = sg.add(in->matrix, 8 * (size_t)nc), i_pos = sg.add(in->pos, 4 * n);
    size_t i_roff = sg.add(in->read_off, 8 * (n + 1)), i_rfl = sg.add(in->rflags, n), i_foff = sg.add(in->feat_off, 8 * (n + 1));
    size_t i_op = sg.add(in->f_op, nf), i_fp = sg.add(in->f_pos, 4 * nf), i_fv = sg.add(in->f_val, 4 * nf), i_fa = sg.add(in->f_aux, 4 * nf);
    size_t i_pay = sg.add(in->payload, pay_dev ? 1 : pay_len), i_qs = sg.add(in->qs, qs_dev ? 1 : qs_len);
    size_t i_qso = in->read_qs_off ? sg.add(in->read_qs_off, 8 * n) : sg.reserve(8 * n);
    size_t i_ba = sg.add(in->ba, ba_dev ? 1 : ba_len);
    size_t i_bao = in->read_ba_off ? sg.add(in->read_ba_off, 8 * n) : sg.reserve(8 * n);
    // results: into the caller's buffers when those are device memory, else into slots of the staging buffer
    const bool b_dev = sg.mm.kind(bases) == MEM_DEVICE, q_dev = sg.mm.kind(quals) == MEM_DEVICE;
    size_t o_bases = sg.reserve(b_dev ? 1 : nb + 16), o_quals = sg.reserve(q_dev ? 1 : nb + 16);
    size_t o_star = sg.reserve(n), o_stat = sg.reserve(4 * n), o_flag = sg.reserve(16);
    if (int rc = pinned_reserve(ctx, &ctx->pin_in, &ctx->pin_in_cap, sg.ptotal + 16)) return rc;
    if (int rc = pool_reserve(ctx, ctx->dev_in, sg.total)) return rc;
    uint8_t *pin = (uint8_t *)ctx->pin_in, *D = (uint8_t *)ctx->dev_in.p;
    if (int rc = sg.to_device(ctx, pin, D, st)) return rc;
    // records without a stored array / BA bases still need valid offset tables
    if (!in->read_qs_off) CUDA_TRY(ctx, cudaMemsetAsync(sg.at(D, i_qso), 0xFF, 8 * n, st));
    if (!in->read_ba_off) CUDA_TRY(ctx, cudaMemsetAsync(sg.at(D, i_bao), 0xFF, 8 * n, st));
    cramgpu_norm_in dev = *in;
    dev.ref = sg.at(D, i_ref); dev.matrix = sg.at(D, i_mtx); dev.pos = (const int32_t *)sg.at(D, i_pos); dev.read_off = (const uint64_t *)sg.at(D, i_roff);
    dev.rflags = sg.at(D, i_rfl); dev.feat_off = (const uint64_t *)sg.at(D, i_foff); dev.f_op = sg.at(D, i_op); dev.f_pos = (const int32_t *)sg.at(D, i_fp);
    dev.f_val = (const int32_t *)sg.at(D, i_fv); dev.f_aux = (const int32_t *)sg.at(D, i_fa); dev.payload = sg.at(D, i_pay); dev.qs = sg.at(D, i_qs);
    dev.read_qs_off = (const uint64_t *)sg.at(D, i_qso); dev.ba = sg.at(D, i_ba); dev.read_ba_off = (const uint64_t *)sg.at(D, i_bao);
    if (!in->qs && qso_dev) dev.qs = nullptr;                       // absolute device addresses in read_qs_off / read_ba_off
    if (!in->ba && bao_dev) dev.ba = nullptr;
    uint8_t *d_b = b_dev ? bases : sg.at(D, o_bases), *d_q = q_dev ? quals : sg.at(D, o_quals);
    if (int rc = normalize_impl(ctx, &dev, d_b, d_q, sg.at(D, o_star), (int32_t *)sg.at(D, o_stat), st)) return rc;
    int32_t *d_flag = (int32_t *)sg.at(D, o_flag);
    CUDA_TRY(ctx, cudaMemsetAsync(d_flag, 0, 4, st));
    LAUNCH(ctx, st, "k_any_status", k_any_status<<<(uint32_t)std::min<uint64_t>((n + 255) / 256, 148 * 8), 256, 0, st>>>((const int32_t *)sg.at(D, o_stat), n, d_flag));
    // staging for what does not go straight into the caller's memory
    const bool pb = nb && !b_dev && host_is_pinned(bases), pq = nb && !q_dev && host_is_pinned(quals);
    const size_t s_b = (nb && !b_dev && !pb) ? rup(nb + 16, 256) : 0, s_q = (nb && !q_dev && !pq) ? rup(nb + 16, 256) : 0;
    if (int rc = pinned_reserve(ctx, &ctx->pin_out, &ctx->pin_out_cap, s_b + s_q + rup(n, 256) + rup(4 * n, 256) + 256)) return rc;
    uint8_t *po = (uint8_t *)ctx->pin_out;
    size_t p_q = s_b, p_s = s_b + s_q, p_t = p_s + rup(n, 256), p_f = p_t + rup(4 * n, 256);
    if (nb && !b_dev) CUDA_TRY(ctx, memcpy_pieces(pb ? bases : po, d_b, nb, cudaMemcpyDeviceToHost, st));
    if (nb && !q_dev) CUDA_TRY(ctx, memcpy_pieces(pq ? quals : po + p_q, d_q, nb, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(po + p_s, sg.at(D, o_star), n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(po + p_t, sg.at(D, o_stat), 4 * n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(po + p_f, d_flag, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    {
        std::vector<CopyJob> jobs;
        if (nb && !b_dev && !pb) jobs.push_back(CopyJob{bases, po, (size_t)nb});
        if (nb && !q_dev && !pq) jobs.push_back(CopyJob{quals, po + p_q, (size_t)nb});
        jobs.push_back(CopyJob{qual_star, po + p_s, (size_t)n});
        jobs.push_back(CopyJob{status, po + p_t, (size_t)(4 * n)});
        par_memcpy(jobs);
    }
    if (*(const int32_t *)(po + p_f))
        for (uint64_t r = 0; r < n; r++)
            if (status[r]) return ctx_fail(ctx, status[r], "record %llu lies outside its reference region (ArrayIndexOutOfBoundsException in the reference)", (unsigned long long)r);
    return CRAMGPU_OK;
}

// ------------------------------------------------------------------------------------------------ series -> features
// Row P, the mirror of the stream packing above: the feature part of CramRecordReader.read
// [encoding/reader/CramRecordReader.java:59-...] for the series cramgpu_features_encode writes.  Per mapped record:
// FN, then per feature FC (operator), FP (position delta) and the operator's value -- X: BS code, D: DL, N: RS, H: HC,
// P: PD, i: BA byte, Q: QS byte, S: SC bytes up to the TAB (ByteArrayStopEncoding) --, then MQ, then the score array
// (readLength bytes of QS) when the record carries one.  Unmapped: readLength bytes of BA unless the bases are
// unknown, then the score array.  Every integer is ITF8 [ITF8.java:21-51].
// The walk is sequential by nature (every field's position depends on the ones before it): one warp per container with
// one working lane (containers in the lanes of one warp would wait for each other's cache misses at every load), all
// containers of the call side by side.
struct UnpackArgs {
    uint32_t n_containers;
    const uint64_t *cont_off, *soff, *slen, *fbase, *pbase;     // device copies of the per-container tables
    const uint8_t *streams;
    const uint64_t *read_off; const uint8_t *rflags;
    uint8_t *f_op; int32_t *f_pos, *f_val, *f_aux; uint64_t *feat_off; uint8_t *payload;
    uint64_t *read_qs_off, *read_ba_off; uint8_t *mapq; int32_t *status;
};

// cursors of one container's walk (lane 0 owns them; the other lanes see the values lane 0 publishes per batch)
struct UnpackCur {
    uint64_t p[CRAMGPU_S_COUNT], e[CRAMGPU_S_COUNT];
    const uint8_t *S;
    int bad;
    __device__ __forceinline__ uint32_t byte(int k) { if (p[k] >= e[k]) { bad |= 1; return 0; } return S[p[k]++]; }
    __device__ __forceinline__ int32_t itf8(int k)                                   // ITF8.readUnsignedITF8 [ITF8.java:21-51]
    {
        const uint32_t b1 = byte(k);
        if (b1 < 0x80) return (int32_t)b1;
        if (b1 < 0xC0) return (int32_t)(((b1 & 0x7F) << 8) | byte(k));
        if (b1 < 0xE0) { const uint32_t b2 = byte(k); return (int32_t)(((b1 & 0x3F) << 16) | (b2 << 8) | byte(k)); }
        if (b1 < 0xF0) { const uint32_t b2 = byte(k), b3 = byte(k); return (int32_t)(((b1 & 0x1F) << 24) | (b2 << 16) | (b3 << 8) | byte(k)); }
        const uint32_t b2 = byte(k), b3 = byte(k), b4 = byte(k);
        return (int32_t)(((b1 & 0x0F) << 28) | (b2 << 20) | (b3 << 12) | (b4 << 4) | (byte(k) & 0x0F));
    }
};
// the nf features of one record, in the order the writer put them; fi = first free feature slot; -> features written
__device__ __forceinline__ uint32_t unpack_features(const UnpackArgs &a, UnpackCur &cu, uint32_t nf, uint64_t fi, uint64_t f_end, uint64_t &pay)
{
    int32_t prev = 0;
    uint32_t done = 0;
    for (; done < nf && !cu.bad; done++, fi++) {
        if (fi >= f_end) { cu.bad |= 1; break; }                       // more features than operator bytes
        const uint8_t op = (uint8_t)cu.byte(CRAMGPU_S_FC);
        prev += cu.itf8(CRAMGPU_S_FP);
        int32_t val = 0, aux = 0;
        switch (op) {
        case 'X': val = (int32_t)cu.byte(CRAMGPU_S_BS); break;
        case 'D': val = cu.itf8(CRAMGPU_S_DL); break;
        case 'N': val = cu.itf8(CRAMGPU_S_RS); break;
        case 'H': val = cu.itf8(CRAMGPU_S_HC); break;
        case 'P': val = cu.itf8(CRAMGPU_S_PD); break;
        case 'i': val = (int32_t)cu.byte(CRAMGPU_S_BA); break;
        case 'Q': val = (int32_t)cu.byte(CRAMGPU_S_QS); break;
        case 'S': {
            val = (int32_t)pay;
            for (;;) {
                if (cu.p[CRAMGPU_S_SC] >= cu.e[CRAMGPU_S_SC]) { cu.bad |= 1; break; }
                const uint8_t b = cu.S[cu.p[CRAMGPU_S_SC]++];
                if (b == '\t') break;
                a.payload[pay++] = b; aux++;
            }
            break;
        }
        default: cu.bad |= 2; break;                                  // an operator this library does not write (B, I, b, q)
        }
        a.f_op[fi] = op; a.f_pos[fi] = prev; a.f_val[fi] = val; a.f_aux[fi] = aux;
    }
    return done;
}

// One warp per container.  32 records per round: their per-record fields (FN, MQ -- one byte each in all but rare
// records --, flags, lengths, the offsets of the score arrays) are handled by the 32 lanes at once; lane 0 walks the few
// features of the round one after the other (every field's position depends on the fields before it).  A round with a
// multi-byte FN or MQ is walked by lane 0 alone.
__global__ void __launch_bounds__(32) k_unpack(UnpackArgs a)
{
    const uint32_t c = blockIdx.x, lane = threadIdx.x, lt = (1u << lane) - 1u;
    if (c >= a.n_containers) return;
    __shared__ uint32_t s_nf[32], s_cnt[32], s_arrq[32], s_arrb[32], s_q[32], s_i[32], s_bad;
    __shared__ uint64_t s_p[CRAMGPU_S_COUNT], s_pay;
    UnpackCur cu;
#pragma unroll
    for (int k = 0; k < CRAMGPU_S_COUNT; k++) { cu.p[k] = a.soff[(size_t)c * CRAMGPU_S_COUNT + k]; cu.e[k] = cu.p[k] + a.slen[(size_t)c * CRAMGPU_S_COUNT + k]; }
    cu.S = a.streams; cu.bad = 0;
    const uint64_t fb = a.fbase[c], f_cap = a.slen[(size_t)c * CRAMGPU_S_COUNT + CRAMGPU_S_FC];
    uint64_t nf_total = 0, pay = a.pbase[c];                           // kept identical in all lanes
    const uint64_t r_end = a.cont_off[c + 1];
    for (uint64_t r0 = a.cont_off[c]; r0 < r_end; r0 += 32) {
        const uint64_t r = r0 + lane;
        const bool live = r < r_end;
        const uint64_t rlen = live ? a.read_off[r + 1] - a.read_off[r] : 0;
        const uint8_t fl = live ? a.rflags[r] : (uint8_t)(CRAMGPU_NF_UNMAPPED | CRAMGPU_NF_UNKNOWN_BASES);
        const bool mapped = live && !(fl & CRAMGPU_NF_UNMAPPED);
        const uint32_t mb = __ballot_sync(0xFFFFFFFFu, mapped), rank = __popc(mb & lt), nm = __popc(mb);
        // lane 0's cursors, as of the start of the round
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < CRAMGPU_S_COUNT; k++) s_p[k] = cu.p[k];
        }
        __syncwarp();
        const uint64_t pFN = s_p[CRAMGPU_S_FN], pMQ = s_p[CRAMGPU_S_MQ], pQS = s_p[CRAMGPU_S_QS], pBA = s_p[CRAMGPU_S_BA];
        // FN and MQ of record `rank` among the mapped ones, if every such field of the round is a single byte
        uint32_t bFN = 0, bMQ = 0;
        bool ok = true;
        if (mapped) {
            ok = pFN + rank < cu.e[CRAMGPU_S_FN] && pMQ + rank < cu.e[CRAMGPU_S_MQ];
            if (ok) { bFN = cu.S[pFN + rank]; bMQ = cu.S[pMQ + rank]; ok = bFN < 0x80 && bMQ < 0x80; }
        }
        const bool fast = __all_sync(0xFFFFFFFFu, ok);
        const uint32_t arrq = (live && (fl & CRAMGPU_NF_FORCE_QS)) ? (uint32_t)rlen : 0u;                              // score array
        const uint32_t arrb = (live && !mapped && !(fl & CRAMGPU_NF_UNKNOWN_BASES)) ? (uint32_t)rlen : 0u;             // unmapped bases
        if (fast) {
            const uint32_t nf = bFN;
            uint32_t tot, ex = warp_excl_scan_u32(nf, lane, tot);
            if (live) { a.feat_off[r] = fb + nf_total + ex; a.mapq[r] = (uint8_t)bMQ; }
            // where the arrays of the records before this one end (arrays only; the Q / i bytes of features come on top)
            uint32_t totq, totb;
            const uint32_t exq = warp_excl_scan_u32(arrq, lane, totq), exb = warp_excl_scan_u32(arrb, lane, totb);
            s_nf[lane] = ex; s_cnt[lane] = nf; s_arrq[lane] = exq; s_arrb[lane] = exb; s_q[lane] = 0; s_i[lane] = 0;
            const uint32_t fm = __ballot_sync(0xFFFFFFFFu, nf != 0);
            __syncwarp();
            if (lane == 0) {
                cu.p[CRAMGPU_S_FN] += nm; cu.p[CRAMGPU_S_MQ] += nm;
                // only the records that have features are walked, in turn (their features may take bytes of QS / BA)
                uint32_t qrun = 0, irun = 0;
                for (uint32_t mk = fm; mk && !cu.bad; mk &= mk - 1) {
                    const uint32_t l = (uint32_t)__ffs(mk) - 1;
                    const uint32_t n = s_cnt[l];
                    const uint64_t q0 = pQS + s_arrq[l] + qrun, b0 = pBA + s_arrb[l] + irun;
                    cu.p[CRAMGPU_S_QS] = q0; cu.p[CRAMGPU_S_BA] = b0;
                    unpack_features(a, cu, n, fb + nf_total + s_nf[l], fb + f_cap, pay);
                    const uint32_t q = (uint32_t)(cu.p[CRAMGPU_S_QS] - q0), i = (uint32_t)(cu.p[CRAMGPU_S_BA] - b0);
                    s_q[l] = q; s_i[l] = i; qrun += q; irun += i;
                }
                cu.p[CRAMGPU_S_QS] = pQS + totq + qrun; cu.p[CRAMGPU_S_BA] = pBA + totb + irun;
                if (cu.p[CRAMGPU_S_QS] > cu.e[CRAMGPU_S_QS] || cu.p[CRAMGPU_S_BA] > cu.e[CRAMGPU_S_BA]) cu.bad |= 1;
                s_pay = pay; s_bad = (uint32_t)cu.bad;
            }
            __syncwarp();
            {
                // a record's arrays start after the arrays of the records before it and after every feature byte up to its own
                uint32_t tq, ti;
                const uint32_t q = s_q[lane], i = s_i[lane];
                const uint32_t qi = warp_excl_scan_u32(q, lane, tq) + q, ii = warp_excl_scan_u32(i, lane, ti) + i;
                if (live) {
                    a.read_qs_off[r] = (fl & CRAMGPU_NF_FORCE_QS) ? pQS + exq + qi : ~0ull;
                    a.read_ba_off[r] = (!mapped && !(fl & CRAMGPU_NF_UNKNOWN_BASES)) ? pBA + exb + ii : ~0ull;
                }
            }
            nf_total += tot; pay = s_pay;
            if (s_bad) break;
            __syncwarp();
        } else {
            // a multi-byte FN or MQ somewhere in the round: lane 0 walks its records one by one
            if (lane == 0) {
                uint64_t nft = nf_total;
                for (uint64_t rr = r0; rr < r0 + 32 && rr < r_end && !cu.bad; rr++) {
                    const uint64_t len = a.read_off[rr + 1] - a.read_off[rr];
                    const uint8_t f2 = a.rflags[rr];
                    a.feat_off[rr] = fb + nft;
                    uint64_t qo = ~0ull, bo = ~0ull;
                    uint32_t mq = 0;
                    if (!(f2 & CRAMGPU_NF_UNMAPPED)) {
                        const uint32_t n = (uint32_t)cu.itf8(CRAMGPU_S_FN);
                        nft += unpack_features(a, cu, n, fb + nft, fb + f_cap, pay);
                        mq = (uint32_t)cu.itf8(CRAMGPU_S_MQ);
                    } else if (!(f2 & CRAMGPU_NF_UNKNOWN_BASES)) {
                        bo = cu.p[CRAMGPU_S_BA]; cu.p[CRAMGPU_S_BA] += len;
                        if (cu.p[CRAMGPU_S_BA] > cu.e[CRAMGPU_S_BA]) cu.bad |= 1;
                    }
                    if (f2 & CRAMGPU_NF_FORCE_QS) {
                        qo = cu.p[CRAMGPU_S_QS]; cu.p[CRAMGPU_S_QS] += len;
                        if (cu.p[CRAMGPU_S_QS] > cu.e[CRAMGPU_S_QS]) cu.bad |= 1;
                    }
                    a.read_qs_off[rr] = qo; a.read_ba_off[rr] = bo; a.mapq[rr] = (uint8_t)mq;
                }
                s_p[0] = nft; s_pay = pay; s_bad = (uint32_t)cu.bad;
            }
            __syncwarp();
            nf_total = s_p[0]; pay = s_pay;
            if (s_bad) break;
            __syncwarp();
        }
    }
    __syncwarp();
    const int bad = (int)__shfl_sync(0xFFFFFFFFu, (uint32_t)cu.bad, 0);
    // the operator bytes a corrupt container does not account for still get defined feature slots, and its records defined offsets
    for (uint64_t i = (bad ? 0 : nf_total) + lane; i < f_cap; i += 32) { a.f_op[fb + i] = 0; a.f_pos[fb + i] = 0; a.f_val[fb + i] = 0; a.f_aux[fb + i] = 0; }
    if (bad) {
        // nothing of a corrupt container is handed on: its records get no features and no arrays
        for (uint64_t r = a.cont_off[c] + lane; r < r_end; r += 32) { a.feat_off[r] = fb; a.read_qs_off[r] = ~0ull; a.read_ba_off[r] = ~0ull; a.mapq[r] = 0; }
    }
    if (lane == 0) a.status[c] = (bad & 2) ? CRAMGPU_ERR_UNSUPPORTED : bad ? CRAMGPU_ERR_CORRUPT : CRAMGPU_OK;
}

static int features_unpack_impl(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *out, uint8_t **d_mapq,
                                int32_t *h_status, cudaStream_t st)
{
    if (!in || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    memset(out, 0, sizeof *out);
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (!in->cont_off || !in->stream_off || !in->stream_len || !in->read_off || !in->rflags)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    if (in->cont_off[0] != 0 || in->cont_off[nc] != n) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "cont_off does not cover the reads");
    // one operator byte per feature: the FC lengths give every container's place in the feature arrays up front
    std::vector<uint64_t> fbase(nc), pbase(nc);
    uint64_t nfeat = 0, paylen = 0;
    for (uint32_t c = 0; c < nc; c++) {
        fbase[c] = nfeat; nfeat += in->stream_len[(size_t)c * CRAMGPU_S_COUNT + CRAMGPU_S_FC];
        pbase[c] = paylen; paylen += in->stream_len[(size_t)c * CRAMGPU_S_COUNT + CRAMGPU_S_SC];   // soft-clip bytes incl. their TABs: an upper bound
        for (int k = 0; k < CRAMGPU_S_COUNT; k++)
            if (in->streams && in->stream_off[(size_t)c * CRAMGPU_S_COUNT + k] + in->stream_len[(size_t)c * CRAMGPU_S_COUNT + k] > in->streams_len)
                return ctx_fail(ctx, CRAMGPU_ERR_ARG, "series %d of container %u lies outside streams", k, c);
    }
    const size_t tab = 8 * (size_t)CRAMGPU_S_COUNT * nc;
    Carver cv;
    size_t o_cont = cv.take(8 * (size_t)(nc + 1)), o_soff = cv.take(tab), o_slen = cv.take(tab), o_fb = cv.take(8 * (size_t)nc), o_pb = cv.take(8 * (size_t)nc);
    size_t o_stat = cv.take(4 * (size_t)nc);
    if (int rc = pool_reserve(ctx, ctx->rec_scratch, cv.off)) return rc;
    uint8_t *S = (uint8_t *)ctx->rec_scratch.p;
    if (int rc = pinned_reserve(ctx, &ctx->pin_meta, &ctx->pin_meta_cap, 2 * tab + 8 * (size_t)(3 * nc + 1) + 512)) return rc;
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev));
    Uploader U{ctx, st};
    CUDA_TRY(ctx, U.up(in->cont_off, 8 * (size_t)(nc + 1), S + o_cont));
    CUDA_TRY(ctx, U.up(in->stream_off, tab, S + o_soff));
    CUDA_TRY(ctx, U.up(in->stream_len, tab, S + o_slen));
    CUDA_TRY(ctx, U.up(fbase.data(), 8 * (size_t)nc, S + o_fb));
    CUDA_TRY(ctx, U.up(pbase.data(), 8 * (size_t)nc, S + o_pb));
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev, st));
    Carver co;
    size_t q_op = co.take(nfeat + 16), q_pos = co.take(4 * nfeat + 16), q_val = co.take(4 * nfeat + 16), q_aux = co.take(4 * nfeat + 16);
    size_t q_foff = co.take(8 * (size_t)(n + 1)), q_pay = co.take(paylen + 16);
    size_t q_qso = co.take(8 * (size_t)n), q_bao = co.take(8 * (size_t)n), q_mq = co.take(n);
    if (int rc = pool_reserve(ctx, ctx->rec_unp, co.off)) return rc;
    uint8_t *O = (uint8_t *)ctx->rec_unp.p;
    UnpackArgs a;
    memset(&a, 0, sizeof a);
    a.n_containers = nc;
    a.cont_off = (const uint64_t *)(S + o_cont); a.soff = (const uint64_t *)(S + o_soff); a.slen = (const uint64_t *)(S + o_slen);
    a.fbase = (const uint64_t *)(S + o_fb); a.pbase = (const uint64_t *)(S + o_pb);
    a.streams = in->streams; a.read_off = in->read_off; a.rflags = in->rflags;
    a.f_op = O + q_op; a.f_pos = (int32_t *)(O + q_pos); a.f_val = (int32_t *)(O + q_val); a.f_aux = (int32_t *)(O + q_aux);
    a.feat_off = (uint64_t *)(O + q_foff); a.payload = O + q_pay; a.read_qs_off = (uint64_t *)(O + q_qso); a.read_ba_off = (uint64_t *)(O + q_bao);
    a.mapq = O + q_mq; a.status = (int32_t *)(S + o_stat);
    LAUNCH(ctx, st, "k_unpack", k_unpack<<<nc, 32, 0, st>>>(a));
    CUDA_TRY(ctx, cudaGetLastError());
    // feat_off[n] = all features; per-container status back to the host
    if (int rc = pinned_reserve(ctx, &ctx->pin_rec, &ctx->pin_rec_cap, 4 * (size_t)nc + 64)) return rc;
    uint64_t *pr = (uint64_t *)ctx->pin_rec;
    pr[0] = nfeat;
    CUDA_TRY(ctx, cudaMemcpyAsync(a.feat_off + n, pr, 8, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync((uint8_t *)pr + 16, a.status, 4 * (size_t)nc, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const int32_t *hs = (const int32_t *)((uint8_t *)pr + 16);
    int first = CRAMGPU_OK;
    for (uint32_t c = 0; c < nc; c++) {
        if (h_status) h_status[c] = hs[c];
        if (hs[c] && !first) {
            first = hs[c];
            ctx_fail(ctx, first, first == CRAMGPU_ERR_UNSUPPORTED ? "container %u: read feature operator outside {X,D,N,H,P,i,Q,S}" : "container %u: series end before the records do", c);
        }
    }
    out->n_feat = nfeat; out->f_op = a.f_op; out->f_pos = a.f_pos; out->f_val = a.f_val; out->f_aux = a.f_aux;
    out->feat_off = a.feat_off; out->payload = a.payload; out->payload_len = paylen;
    out->streams = in->streams; out->streams_len = in->streams_len;
    out->read_qs_off = a.read_qs_off; out->read_ba_off = a.read_ba_off;
    if (d_mapq) *d_mapq = a.mapq;
    return first;
}

extern "C" int cramgpu_features_unpack_dev(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *out, uint8_t **d_mapq,
                                           int32_t *status, void *stream)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return features_unpack_impl(ctx, in, out, d_mapq, status, stream ? (cudaStream_t)stream : ctx->stream);
}

// Host variant that leaves the features on the device: in->streams must be NULL and stream_off holds ABSOLUTE addresses, each
// series in host memory (staged: pieces that lie next to each other travel as one copy) or in device memory (used in place).
extern "C" int cramgpu_features_unpack_resident(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *dev_out, uint8_t *mapq,
                                                int32_t *status)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !dev_out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    memset(dev_out, 0, sizeof *dev_out);
    const uint64_t n = in->n_reads;
    const uint32_t nc = in->n_containers;
    if (n == 0 || nc == 0) return CRAMGPU_OK;
    if (in->streams || !in->stream_off || !in->stream_len || !in->read_off || !in->rflags)
        return ctx_fail(ctx, CRAMGPU_ERR_ARG, "streams must be NULL (absolute addresses in stream_off); null argument");
    MemMap mm;
    struct Piece { uint64_t addr, len; size_t idx; };
    std::vector<Piece> pieces;
    const size_t ncell = (size_t)nc * CRAMGPU_S_COUNT;
    for (size_t i = 0; i < ncell; i++)
        if (in->stream_len[i] && mm.kind((const void *)(uintptr_t)in->stream_off[i]) != MEM_DEVICE) pieces.push_back({in->stream_off[i], in->stream_len[i], i});
    std::sort(pieces.begin(), pieces.end(), [](const Piece &x, const Piece &y) { return x.addr < y.addr; });
    // runs of neighbouring pieces (gaps of alignment padding ride along)
    struct Run { uint64_t lo, hi; size_t off; };
    std::vector<Run> runs;
    Carver cv;
    for (const Piece &p : pieces) {
        if (!runs.empty() && p.addr <= runs.back().hi + 256) runs.back().hi = std::max(runs.back().hi, p.addr + p.len);
        else runs.push_back({p.addr, p.addr + p.len, 0});
    }
    for (Run &r : runs) r.off = cv.take(r.hi - r.lo + 16);
    const bool roff_host = mm.kind(in->read_off) != MEM_DEVICE, rfl_host = mm.kind(in->rflags) != MEM_DEVICE;
    const size_t o_roff = roff_host ? cv.take(8 * (n + 1)) : 0, o_rfl = rfl_host ? cv.take(n) : 0;
    if (int rc = pool_reserve(ctx, ctx->unp_in, cv.off + 16)) return rc;
    uint8_t *D = (uint8_t *)ctx->unp_in.p;
    for (const Run &r : runs) CUDA_TRY(ctx, memcpy_pieces(D + r.off, (const void *)(uintptr_t)r.lo, r.hi - r.lo, cudaMemcpyHostToDevice, st));
    if (roff_host) CUDA_TRY(ctx, memcpy_pieces(D + o_roff, in->read_off, 8 * (n + 1), cudaMemcpyHostToDevice, st));
    if (rfl_host) CUDA_TRY(ctx, memcpy_pieces(D + o_rfl, in->rflags, n, cudaMemcpyHostToDevice, st));
    std::vector<uint64_t> addr(in->stream_off, in->stream_off + ncell);
    {
        size_t ri = 0;
        for (const Piece &p : pieces) {                       // pieces and runs are both in address order
            while (p.addr >= runs[ri].hi) ri++;
            addr[p.idx] = (uint64_t)(uintptr_t)(D + runs[ri].off + (p.addr - runs[ri].lo));
        }
    }
    cramgpu_series dev = *in;
    dev.stream_off = addr.data();
    dev.read_off = roff_host ? (const uint64_t *)(D + o_roff) : in->read_off;
    dev.rflags = rfl_host ? D + o_rfl : in->rflags;
    uint8_t *d_mq = nullptr;
    const int rc0 = features_unpack_impl(ctx, &dev, dev_out, &d_mq, status, st);
    if (rc0 != CRAMGPU_OK && rc0 != CRAMGPU_ERR_CORRUPT && rc0 != CRAMGPU_ERR_UNSUPPORTED) return rc0;
    if (mapq) {
        CUDA_TRY(ctx, cudaMemcpyAsync(mapq, d_mq, n, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    return rc0;
}

extern "C" int cramgpu_features_unpack(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *out, uint8_t *mapq,
                                       int32_t *status)
{
    if (!ctx) return CRAMGPU_ERR_ARG;
    if (!in || !out) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    memset(out, 0, sizeof *out);
    const uint64_t n = in->n_reads;
    if (n == 0 || in->n_containers == 0) return CRAMGPU_OK;
    if (!in->read_off || !in->rflags || !in->streams) return ctx_fail(ctx, CRAMGPU_ERR_ARG, "null argument");
    Stager sg;
    size_t i_str = sg.add(in->streams, in->streams_len), i_roff = sg.add(in->read_off, 8 * (n + 1)), i_rfl = sg.add(in->rflags, n);
    if (int rc = pinned_reserve(ctx, &ctx->pin_in, &ctx->pin_in_cap, sg.ptotal + 16)) return rc;
    if (int rc = pool_reserve(ctx, ctx->dev_in, sg.total)) return rc;
    uint8_t *pin = (uint8_t *)ctx->pin_in, *D = (uint8_t *)ctx->dev_in.p;
    if (int rc = sg.to_device(ctx, pin, D, st)) return rc;
    cramgpu_series dev = *in;
    dev.streams = sg.at(D, i_str); dev.read_off = (const uint64_t *)sg.at(D, i_roff); dev.rflags = sg.at(D, i_rfl);
    cramgpu_features df;
    uint8_t *d_mq = nullptr;
    const int rc0 = features_unpack_impl(ctx, &dev, &df, &d_mq, status, st);
    if (rc0 != CRAMGPU_OK && rc0 != CRAMGPU_ERR_CORRUPT && rc0 != CRAMGPU_ERR_UNSUPPORTED) return rc0;
    Stager so;
    size_t r_op = so.reserve(df.n_feat), r_pos = so.reserve(4 * df.n_feat), r_val = so.reserve(4 * df.n_feat), r_aux = so.reserve(4 * df.n_feat);
    size_t r_foff = so.reserve(8 * (n + 1)), r_pay = so.reserve(df.payload_len), r_qso = so.reserve(8 * n), r_bao = so.reserve(8 * n), r_mq = so.reserve(n);
    if (int rc = pinned_reserve(ctx, &ctx->pin_feat, &ctx->pin_feat_cap, so.total)) return rc;
    uint8_t *po = (uint8_t *)ctx->pin_feat;
    auto down = [&](size_t slot, const void *d, size_t bytes) -> cudaError_t {
        return bytes ? memcpy_pieces(po + so.off[slot], d, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    CUDA_TRY(ctx, down(r_op, df.f_op, df.n_feat)); CUDA_TRY(ctx, down(r_pos, df.f_pos, 4 * df.n_feat));
    CUDA_TRY(ctx, down(r_val, df.f_val, 4 * df.n_feat)); CUDA_TRY(ctx, down(r_aux, df.f_aux, 4 * df.n_feat));
    CUDA_TRY(ctx, down(r_foff, df.feat_off, 8 * (n + 1))); CUDA_TRY(ctx, down(r_pay, df.payload, df.payload_len));
    CUDA_TRY(ctx, down(r_qso, df.read_qs_off, 8 * n)); CUDA_TRY(ctx, down(r_bao, df.read_ba_off, 8 * n)); CUDA_TRY(ctx, down(r_mq, d_mq, n));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    *out = df;
    out->f_op = po + so.off[r_op]; out->f_pos = (const int32_t *)(po + so.off[r_pos]); out->f_val = (const int32_t *)(po + so.off[r_val]);
    out->f_aux = (const int32_t *)(po + so.off[r_aux]); out->feat_off = (const uint64_t *)(po + so.off[r_foff]); out->payload = po + so.off[r_pay];
    out->read_qs_off = (const uint64_t *)(po + so.off[r_qso]); out->read_ba_off = (const uint64_t *)(po + so.off[r_bao]);
    out->streams = in->streams;
    if (mapq) memcpy(mapq, po + so.off[r_mq], n);
    return rc0;
}
