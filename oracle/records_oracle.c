/*
 * oracle/records_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, single-threaded restatement of the per-record part of the CRAM slice hot path:
 *
 *   F2-F4  Sam2CramRecordFactory.createCramRecord / createVariations / addSubstitutionsAndMaskedBases
 *          (htsjdk 2.4.1 bytecode, [Sam2CramRecordFactory.java:108-335])
 *   F5     CompressionHeaderFactory.buildFrequencies / updateSubstitutionCodes [:171-209],
 *          SubstitutionMatrix.<init>(long[][]) / rank [SubstitutionMatrix.java:42-57,141-169]
 *   Q      QualityScorePreservation.parsePolicies / addQualityScores / addQS, Binning
 *          (src/main/java/htsjdk/samtools/cram/lossy/QualityScorePreservation.java:167-301,332-489)
 *   W      Writer.write [Writer.java:139-261] for the series that come out of the feature work, ITF8 [ITF8.java:53-84]
 *   N1,N2  CramNormalizer.restoreReadBases / restoreQualityScores
 *          (src/main/java/htsjdk/samtools/cram/build/CramNormalizer.java:287-434)
 *
 * Parity pin: tests/golden/records_jvm.json was produced by executing those very methods out of
 * /root/reference/cramtools-3.0.jar with oracle/minijvm.py (oracle/make_golden_records.py);
 * tests/test_oracle_records.py checks this file against it.  The Writer stream assembly (row W) is pinned by
 * tests/golden/streams_jvm.json: the raw series ContainerFactory.buildSlice / Writer.write produced in the mini-JVM
 * (oracle/make_golden_streams.py), compared byte for byte by tests/test_oracle_streams.py.
 *
 * The per-base lossy selectors X and P read ReferenceTracks coverage / pile-up counters: ReferenceTracks, getSpans and
 * updateTracks are restated literally further down (tracks_t) and pinned by tests/golden/tracks_jvm.json
 * (oracle/make_golden_tracks.py, tests/test_oracle_tracks.py).  Without a tracks object R/N/D/I use "tracks present" semantics
 * over the whole reference: baseAt(p) = ref[p-1], 'N' beyond its end.
 *
 * Data layout = structure of arrays, identical to the C ABI (include/cramgpu.h):
 *   reads:  bases/quals u8 concatenated, read_off[n+1]; cigar u32 BAM-packed, cigar_off[n+1];
 *           pos i32 (1-based), flag u16 (SAM), mapq u8, rflags u8 (bit0: no quality scores '*')
 *   features: f_op u8, f_pos i32 (1-based in read), f_val i32, f_aux i32; feat_off[n+1]; payload bytes
 *           X: val = code (-1 until coded), aux = base | refBase << 8
 *           i: val = base      D/N/H/P: val = length      Q: val = score
 *           S: val = offset into payload, aux = length
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { OP_M = 0, OP_I, OP_D, OP_N, OP_S, OP_H, OP_P, OP_EQ, OP_X };
#define RF_NO_QUALS 1
#define SAM_UNMAPPED 4

/* cram/build/Utils.normalizeBase [Utils.java:31-49] */
uint8_t rec_oracle_normalize_base(uint8_t b)
{
    switch (b) {
    case 'a': case 'A': return 'A';
    case 'c': case 'C': return 'C';
    case 'g': case 'G': return 'G';
    case 't': case 'T': return 'T';
    default: return 'N';
    }
}
#define NB rec_oracle_normalize_base

/* ------------------------------------------------------------------ ITF8 [ITF8.java:53-84] */
int rec_oracle_itf8(int32_t value, uint8_t *out)
{
    uint32_t v = (uint32_t)value;
    if ((v >> 7) == 0) { out[0] = (uint8_t)v; return 1; }
    if ((v >> 14) == 0) { out[0] = (uint8_t)((v >> 8) | 0x80); out[1] = (uint8_t)v; return 2; }
    if ((v >> 21) == 0) { out[0] = (uint8_t)((v >> 16) | 0xC0); out[1] = (uint8_t)(v >> 8); out[2] = (uint8_t)v; return 3; }
    if ((v >> 28) == 0) {
        out[0] = (uint8_t)((v >> 24) | 0xE0); out[1] = (uint8_t)(v >> 16); out[2] = (uint8_t)(v >> 8); out[3] = (uint8_t)v;
        return 4;
    }
    out[0] = (uint8_t)((v >> 28) | 0xF0); out[1] = (uint8_t)(v >> 20); out[2] = (uint8_t)(v >> 12);
    out[3] = (uint8_t)(v >> 4); out[4] = (uint8_t)v;
    return 5;
}

/* ------------------------------------------------------------------ feature extraction (F2-F4) */

typedef struct {
    uint8_t *op; int32_t *pos, *val, *aux; int64_t n, cap;
    uint8_t *payload; int64_t pay_n, pay_cap;
} featbuf_t;

static void fb_push(featbuf_t *f, uint8_t op, int32_t pos, int32_t val, int32_t aux)
{
    if (f->n == f->cap) {
        f->cap = f->cap ? f->cap * 2 : 1024;
        f->op = realloc(f->op, (size_t)f->cap);
        f->pos = realloc(f->pos, (size_t)f->cap * 4);
        f->val = realloc(f->val, (size_t)f->cap * 4);
        f->aux = realloc(f->aux, (size_t)f->cap * 4);
    }
    f->op[f->n] = op; f->pos[f->n] = pos; f->val[f->n] = val; f->aux[f->n] = aux; f->n++;
}
static int32_t fb_payload(featbuf_t *f, const uint8_t *p, int32_t len)
{
    if (f->pay_n + len > f->pay_cap) {
        f->pay_cap = (f->pay_n + len) * 2 + 1024;
        f->payload = realloc(f->payload, (size_t)f->pay_cap);
    }
    memcpy(f->payload + f->pay_n, p, (size_t)len);
    f->pay_n += len;
    return (int32_t)(f->pay_n - len);
}

static int consumes_read(int op) { return op == OP_M || op == OP_I || op == OP_S || op == OP_EQ || op == OP_X; }
static int consumes_ref(int op) { return op == OP_M || op == OP_D || op == OP_N || op == OP_EQ || op == OP_X; }

/*
 * One container.  Returns number of features, fills caller arrays via the out pointers (malloc'ed; free with
 * rec_oracle_free).  subst_counts[5][5] (ref x read, ACGTN order) accumulates Substitution features.
 */
typedef struct {
    int64_t n_feat; uint8_t *f_op; int32_t *f_pos, *f_val, *f_aux; int64_t *feat_off;
    uint8_t *payload; int64_t payload_len;
    int64_t subst_counts[5][5];
    uint8_t matrix[5];
} rec_features_t;

static int base_index(uint8_t b)
{
    switch (b) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; case 'N': return 4; default: return -1; }
}

void rec_oracle_free(rec_features_t *r)
{
    free(r->f_op); free(r->f_pos); free(r->f_val); free(r->f_aux); free(r->feat_off); free(r->payload);
    memset(r, 0, sizeof *r);
}

/* SubstitutionMatrix.rank + constructor [SubstitutionMatrix.java:42-57,129-169]: counts[ref][read], ACGTN order.
 * codes[ref][read] = rank; matrix byte per ref base = ranks of the 4 other bases, in base order, 2 bits each. */
void rec_oracle_subst_matrix(const int64_t counts[5][5], uint8_t matrix[5], uint8_t codes[5][5])
{
    for (int r = 0; r < 5; r++) {
        int base[4]; int64_t freq[4]; int k = 0;
        for (int b = 0; b < 5; b++) if (b != r) { base[k] = b; freq[k] = counts[r][b]; k++; }
        /* stable sort by the comparator [:129-135]: freq differs -> (int)(f2 - f1); else base order */
        for (int i = 1; i < 4; i++) {
            int bb = base[i]; int64_t ff = freq[i]; int j = i - 1;
            while (j >= 0) {
                int32_t c = (freq[j] != ff) ? (int32_t)(uint32_t)(uint64_t)(ff - freq[j]) : (base[j] - bb);
                /* compare(a=elem j, b=new): result>0 means a after b */
                if (c > 0) { base[j + 1] = base[j]; freq[j + 1] = freq[j]; j--; } else break;
            }
            base[j + 1] = bb; freq[j + 1] = ff;
        }
        int rank_of[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 4; i++) rank_of[base[i]] = i;
        uint8_t byte = 0;
        for (int b = 0; b < 5; b++) {
            codes[r][b] = 0;
            if (b == r) continue;
            byte = (uint8_t)((byte << 2) | rank_of[b]);
            codes[r][b] = (uint8_t)rank_of[b];
        }
        matrix[r] = byte;
    }
}

/* SubstitutionMatrix(byte[]) decode side: base with rank `code` for reference base r (ACGTN index) */
static uint8_t matrix_base(const uint8_t matrix[5], int r, int code)
{
    static const char B[] = "ACGTN";
    int k = 0;
    for (int b = 0; b < 5; b++) {
        if (b == r) continue;
        int rank = (matrix[r] >> (2 * (3 - k))) & 3;
        if (rank == code) return (uint8_t)B[b];
        k++;
    }
    return 'N';
}

int rec_oracle_extract(int64_t n_reads, const uint8_t *bases, const int64_t *read_off, const uint32_t *cigar,
                       const int64_t *cigar_off, const int32_t *pos, const uint16_t *flag, const uint8_t *ref,
                       int64_t ref_len, rec_features_t *out)
{
    featbuf_t fb;
    memset(&fb, 0, sizeof fb);
    memset(out, 0, sizeof *out);
    out->feat_off = malloc((size_t)(n_reads + 1) * 8);
    for (int64_t r = 0; r < n_reads; r++) {
        out->feat_off[r] = fb.n;
        /* features only for mapped reads with a position [Sam2CramRecordFactory.java:149-151] */
        if ((flag[r] & SAM_UNMAPPED) || pos[r] == 0) continue;
        const uint8_t *rb = bases + read_off[r];
        int64_t rlen = read_off[r + 1] - read_off[r];
        uint8_t *fill = NULL;
        int64_t cig_read_len = 0;
        for (int64_t c = cigar_off[r]; c < cigar_off[r + 1]; c++)
            if (consumes_read(cigar[c] & 15)) cig_read_len += cigar[c] >> 4;
        if (rlen == 0) {                          /* bases = 'N' x cigarLen [:218-220] */
            fill = malloc((size_t)cig_read_len + 1);
            memset(fill, 'N', (size_t)cig_read_len);
            rb = fill;
        }
        int32_t zpos = 0, aoff = 0;               /* zeroBasedPositionInRead, alignmentStartOffset */
        for (int64_t c = cigar_off[r]; c < cigar_off[r + 1]; c++) {
            int op = cigar[c] & 15; int32_t len = (int32_t)(cigar[c] >> 4);
            switch (op) {
            case OP_D: fb_push(&fb, 'D', zpos + 1, len, 0); break;
            case OP_N: fb_push(&fb, 'N', zpos + 1, len, 0); break;
            case OP_P: fb_push(&fb, 'P', zpos + 1, len, 0); break;
            case OP_H: fb_push(&fb, 'H', zpos + 1, len, 0); break;
            case OP_S: fb_push(&fb, 'S', zpos + 1, fb_payload(&fb, rb + zpos, len), len); break;
            case OP_I:                             /* one InsertBase per base [:282-292] */
                for (int32_t i = 0; i < len; i++) fb_push(&fb, 'i', zpos + 1 + i, rb[zpos + i], 0);
                break;
            case OP_M: case OP_EQ: case OP_X:     /* addSubstitutionsAndMaskedBases [:297-335] */
                for (int32_t i = 0; i < len; i++) {
                    int64_t rc = (int64_t)pos[r] + i + aoff - 1;
                    uint8_t refb = rc >= ref_len ? 'N' : ref[rc];
                    refb = NB(refb);
                    uint8_t b = rb[zpos + i];
                    if (b != refb) {
                        fb_push(&fb, 'X', zpos + i + 1, -1, b | (refb << 8));
                        int ri = base_index(refb), bi = base_index(b);
                        if (ri >= 0 && bi >= 0) out->subst_counts[ri][bi]++;   /* other bytes land outside ACGTN rows */
                    }
                }
                break;
            default: free(fill); return -1;       /* "Unsupported cigar operator" */
            }
            if (consumes_read(op)) zpos += len;
            if (consumes_ref(op)) aoff += len;
        }
        free(fill);
    }
    out->feat_off[n_reads] = fb.n;
    /* F5: rank, then code every substitution [CompressionHeaderFactory.java:171-185] */
    uint8_t codes[5][5];
    rec_oracle_subst_matrix((const int64_t(*)[5])out->subst_counts, out->matrix, codes);
    for (int64_t i = 0; i < fb.n; i++) {
        if (fb.op[i] != 'X') continue;
        int ri = base_index((uint8_t)(fb.aux[i] >> 8)), bi = base_index((uint8_t)(fb.aux[i] & 0xFF));
        fb.val[i] = (ri >= 0 && bi >= 0) ? codes[ri][bi] : 0;   /* codes[][] default 0 outside ACGTN */
    }
    out->n_feat = fb.n; out->f_op = fb.op; out->f_pos = fb.pos; out->f_val = fb.val; out->f_aux = fb.aux;
    out->payload = fb.payload; out->payload_len = fb.pay_n;
    return 0;
}

/* ------------------------------------------------------------------ quality preservation (Q) */

/* Binning.Illumina_binning_matrix (src/main/java/htsjdk/samtools/cram/lossy/Binning.java:63-71) */
static const int8_t ILLUMINA_BIN[110] = {
    0, 1, 6, 6, 6, 6, 6, 6, 6, 6, 15, 15, 15, 15, 15, 15, 15, 15, 15, 15, 22, 22, 22, 22, 22, 27, 27, 27, 27, 27,
    33, 33, 33, 33, 33, 37, 37, 37, 37, 37, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40,
    40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40,
    40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40, 40};

enum { RC_NONE = 0, RC_ALL, RC_UNPLACED, RC_LOWER_MAPQ, RC_HIGHER_MAPQ };
enum { BC_MATCH = 1, BC_MISMATCH, BC_FLANKING_DELETION, BC_INSERTION, BC_LOWER_COVERAGE, BC_PILEUP };   /* the last two: parameter in bits 8..31 */
enum { T_PRESERVE = 0, T_BIN = 1, T_DROP = 2 };

typedef struct {
    int32_t read_cat, read_param;
    int32_t n_base_cats, base_cat[8];
    int32_t treatment, treat_param;
} qs_policy_t;

/* parsePolicies / parseSinglePolicy / readTreatment [QualityScorePreservation.java:141-256].
 * Returns the number of policies (sorted: DROP, BIN, PRESERVE; stable), -1 = unknown category.
 * X<n> (coverage below n; reads no treatment) and P<n> (more than n mismatches piled up; then a treatment) [:210-233]
 * become base categories BC_LOWER_COVERAGE / BC_PILEUP with n in bits 8..31. */
int rec_oracle_parse_policies(const char *spec, qs_policy_t *out, int cap)
{
    int n = 0;
    const char *s = spec;
    while (*s) {
        const char *e = strchr(s, '-');
        size_t len = e ? (size_t)(e - s) : strlen(s);
        if (len) {
            if (n == cap) return -1;
            qs_policy_t p;
            memset(&p, 0, sizeof p);
            int have_t = 0;
            size_t i = 0;
#define READ_PARAM(v) do { v = 0; while (i < len && s[i] >= '0' && s[i] <= '9') v = v * 10 + (s[i++] - '0'); } while (0)
#define READ_TREAT() do { int prm; READ_PARAM(prm); have_t = 1; \
        if (prm == 0) { p.treatment = T_DROP; p.treat_param = 0; } else if (prm == 40) { p.treatment = T_PRESERVE; p.treat_param = 40; } \
        else { p.treatment = T_BIN; p.treat_param = prm; } } while (0)
            while (i < len) {
                char code = s[i++];
                switch (code) {
                case 'R': if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_MATCH; READ_TREAT(); break;
                case 'N': if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_MISMATCH; READ_TREAT(); break;
                case 'D': if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_FLANKING_DELETION; READ_TREAT(); break;
                case 'I': if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_INSERTION; READ_TREAT(); break;
                case 'M': p.read_cat = RC_HIGHER_MAPQ; READ_PARAM(p.read_param); break;
                case 'm': p.read_cat = RC_LOWER_MAPQ; READ_PARAM(p.read_param); break;
                case 'U': p.read_cat = RC_UNPLACED; READ_TREAT(); break;
                case '_': READ_TREAT(); break;
                case '*': p.read_cat = RC_ALL; READ_TREAT(); break;
                case 'X': { int v; READ_PARAM(v); if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_LOWER_COVERAGE | (v << 8); } break;
                case 'P': { int v; READ_PARAM(v); if (p.n_base_cats < 8) p.base_cat[p.n_base_cats++] = BC_PILEUP | (v << 8); READ_TREAT(); } break;
                default: return -1;
                }
                if (!have_t) { p.treatment = T_PRESERVE; have_t = 1; }   /* null treatment -> preserve [:251-252] */
            }
            out[n++] = p;
        }
        if (!e) break;
        s = e + 1;
    }
    /* stable sort by treatment ordinal, descending [:176-188] */
    for (int i = 1; i < n; i++) {
        qs_policy_t p = out[i]; int j = i - 1;
        while (j >= 0 && out[j].treatment < p.treatment) { out[j + 1] = out[j]; j--; }
        out[j + 1] = p;
    }
    return n;
}

/* ------------------------------------------------------------------ ReferenceTracks (the X and P selectors)
 * src/main/java/htsjdk/samtools/cram/ref/ReferenceTracks.java, literally: a window of `len` positions starting at the 1-based
 * `position`; coverage and mismatches are Java shorts (they wrap); an index below 0 is Java's ArrayIndexOutOfBoundsException
 * and a RuntimeException of ensureRange / moveForwardTo ends the conversion: both set `thrown` (1 = AIOOBE, 2 = "Cannot move
 * the window backwards", 3 = "New position is beyond the reference"), after which nothing else is defined. */
typedef struct {
    const uint8_t *reference; int64_t ref_len;
    int32_t position, len, thrown;
    uint8_t *bases; int16_t *coverage, *mismatches;
} tracks_t;

tracks_t *rec_oracle_tracks_new(const uint8_t *reference, int64_t ref_len, int32_t window)        /* [:36-50] */
{
    tracks_t *t = calloc(1, sizeof *t);
    t->reference = reference; t->ref_len = ref_len;
    t->len = (int32_t)(window < ref_len ? window : ref_len);
    t->bases = malloc((size_t)t->len + 1); t->coverage = calloc((size_t)t->len + 1, 2); t->mismatches = calloc((size_t)t->len + 1, 2);
    t->position = 1;
    memcpy(t->bases, reference, (size_t)t->len);                                                    /* reset() [:127-131] */
    return t;
}
void rec_oracle_tracks_free(tracks_t *t) { if (t) { free(t->bases); free(t->coverage); free(t->mismatches); free(t); } }
int32_t rec_oracle_tracks_position(const tracks_t *t) { return t->position; }
int32_t rec_oracle_tracks_len(const tracks_t *t) { return t->len; }
int32_t rec_oracle_tracks_thrown(const tracks_t *t) { return t->thrown; }
const int16_t *rec_oracle_tracks_coverage(const tracks_t *t) { return t->coverage; }
const int16_t *rec_oracle_tracks_mismatches(const tracks_t *t) { return t->mismatches; }

static void tracks_move_forward_to(tracks_t *t, int32_t new_pos)                                    /* [:87-125] */
{
    if ((int64_t)new_pos - 1 >= t->ref_len) { t->thrown = 3; return; }
    if (new_pos < t->position) { t->thrown = 2; return; }
    if ((int64_t)new_pos > t->ref_len - t->len + 1) new_pos = (int32_t)(t->ref_len - t->len + 1);
    if (new_pos == t->position) return;
    int64_t n = t->ref_len - new_pos + 1; if (n > t->len) n = t->len;
    memcpy(t->bases, t->reference + new_pos - 1, (size_t)n);
    if (new_pos > t->position && t->position + t->len - new_pos > 0) {
        for (int32_t i = 0; i < t->len; i++) {
            if (i + new_pos - t->position < t->len) { t->coverage[i] = t->coverage[i + new_pos - t->position]; t->mismatches[i] = t->mismatches[i + new_pos - t->position]; }
            else { t->coverage[i] = 0; t->mismatches[i] = 0; }
        }
    } else { memset(t->coverage, 0, 2 * (size_t)t->len); memset(t->mismatches, 0, 2 * (size_t)t->len); }
    t->position = new_pos;
}
void rec_oracle_tracks_ensure_range(tracks_t *t, int32_t start, int32_t length)                     /* [:133-139] */
{
    if (t->thrown) return;
    if (start < t->position) { t->thrown = 2; return; }
    if ((int64_t)start + length > (int64_t)t->position + t->len) tracks_move_forward_to(t, start);
}
static uint8_t tracks_base_at(tracks_t *t, int32_t pos)                                             /* [:141-146] */
{
    if (pos - t->position < t->len) { if (pos - t->position < 0) { t->thrown = 1; return 0; } return t->bases[pos - t->position]; }
    return 'N';
}
static int16_t tracks_coverage_at(tracks_t *t, int32_t pos)                                         /* [:148-154] */
{
    if (pos - t->position >= t->len) return 0;
    if (pos - t->position < 0) { t->thrown = 1; return 0; }
    return t->coverage[pos - t->position];
}
static int16_t tracks_mismatches_at(tracks_t *t, int32_t pos)                                       /* [:156-161] */
{
    if (pos - t->position >= t->len) return 0;
    if (pos - t->position < 0) { t->thrown = 1; return 0; }
    return t->mismatches[pos - t->position];
}
static void tracks_add(tracks_t *t, int16_t *track, int32_t pos)                                    /* addCoverage / addMismatches [:163-171] */
{
    if (pos - t->position < t->len) { if (pos - t->position < 0) { t->thrown = 1; return; } track[pos - t->position] = (int16_t)(track[pos - t->position] + 1); }
}
/* alignment end as SAMRecord.getAlignmentEnd gives it: 0 for a record with the unmapped flag */
static int32_t sam_alignment_end(int32_t pos, uint16_t flag, const uint32_t *cigar, int64_t ncig)
{
    if (flag & SAM_UNMAPPED) return 0;
    int64_t rl = 0;
    for (int64_t c = 0; c < ncig; c++) if (consumes_ref(cigar[c] & 15)) rl += cigar[c] >> 4;
    return (int32_t)(pos + rl - 1);
}
/* CramSerilization.getSpans for the records of one sequence (ref_id != -1 is the caller's business) [CramSerilization.java:60-86,
 * AlignmentSpan.add :50-61]: union of (start, end - start).  n = 0 leaves *start = *span = 0 and returns 0. */
int rec_oracle_span(int64_t n, const int32_t *pos, const uint16_t *flag, const uint32_t *cigar, const uint64_t *cigar_off, int32_t *start, int32_t *span)
{
    int have = 0; int32_t s0 = 0, sp0 = 0;
    for (int64_t r = 0; r < n; r++) {
        int32_t s = pos[r], sp = sam_alignment_end(pos[r], flag[r], cigar + cigar_off[r], (int64_t)(cigar_off[r + 1] - cigar_off[r])) - s;
        if (!have) { s0 = s; sp0 = sp; have = 1; }
        else if (s0 > s) { int32_t a = s0 + sp0, b = s + sp; sp0 = (a > b ? a : b) - s; s0 = s; }
        else if (s0 < s) { int32_t a = s0 + sp0, b = s + sp; sp0 = (a > b ? a : b) - s0; }
        else sp0 = sp0 > sp ? sp0 : sp;
    }
    *start = s0; *span = sp0;
    return have;
}
/* CramSerilization.updateTracks for one record [CramSerilization.java:225-257], the inner loop's indices as they are written
 * there: `pos` runs from readPos, and indexes the read at readPos + pos and the reference at refPos + pos */
void rec_oracle_tracks_update(tracks_t *t, const uint8_t *rb, int64_t rlen, const uint32_t *cigar, int64_t ncig, int32_t pos)
{
    if (t->thrown || pos == 0) return;
    int32_t ref_pos = pos; int64_t read_pos = 0;
    for (int64_t c = 0; c < ncig && !t->thrown; c++) {
        int op = cigar[c] & 15; int64_t len = cigar[c] >> 4;
        if (consumes_ref(op)) for (int64_t e = 0; e < len; e++) tracks_add(t, t->coverage, (int32_t)(ref_pos + e));
        if (op == OP_M || op == OP_X || op == OP_EQ) {
            for (int64_t p = read_pos; p < len; p++) {
                if (read_pos + p >= rlen) { t->thrown = 1; return; }
                uint8_t read_base = rb[read_pos + p], ref_base = tracks_base_at(t, (int32_t)(ref_pos + p));
                if (read_base != ref_base) tracks_add(t, t->mismatches, (int32_t)(ref_pos + p));
            }
        }
        if (consumes_read(op)) read_pos += len;
        if (consumes_ref(op)) ref_pos += (int32_t)len;
    }
}

/*
 * addQualityScores for one read [QualityScorePreservation.java:276-301, 332-489].
 * scores[rlen] receives the array the reference leaves in record.qualityScores (0xFF = -1 = absent).
 * Returns: bit0 = forcePreserveQualityScores, bit1 = "record has a score array" (quals were present),
 * bit7 = the reference throws ArrayIndexOutOfBoundsException for this record: a BIN treatment indexed
 * Binning.Illumina_binning_matrix (110 entries, signed byte index) with a score of 110..255 [:258-266; Binning.java:63-71];
 * the restatement goes on with entry 109 so that the other records of the batch can still be compared.
 * lossless (n_pol < 0): CramSerilization.java:160-161.
 */
int rec_oracle_quality_t(const uint8_t *rb, const uint8_t *qs, int64_t rlen, int has_quals, const uint32_t *cigar, int64_t ncig,
                         int32_t pos, uint16_t flag, uint8_t mapq, const uint8_t *ref, int64_t ref_len,
                         const qs_policy_t *pol, int n_pol, uint8_t *scores, tracks_t *t);
int rec_oracle_quality(const uint8_t *rb, const uint8_t *qs, int64_t rlen, int has_quals, const uint32_t *cigar, int64_t ncig,
                       int32_t pos, uint16_t flag, uint8_t mapq, const uint8_t *ref, int64_t ref_len,
                       const qs_policy_t *pol, int n_pol, uint8_t *scores)
{
    return rec_oracle_quality_t(rb, qs, rlen, has_quals, cigar, ncig, pos, flag, mapq, ref, ref_len, pol, n_pol, scores, NULL);
}
/* the same with the ReferenceTracks the conversion carries (t may be NULL: then X and P mask nothing); the per-base part asks
 * the tracks for the record's range first [:397], which can move the window.  Check rec_oracle_tracks_thrown afterwards. */
int rec_oracle_quality_t(const uint8_t *rb, const uint8_t *qs, int64_t rlen, int has_quals, const uint32_t *cigar, int64_t ncig,
                         int32_t pos, uint16_t flag, uint8_t mapq, const uint8_t *ref, int64_t ref_len,
                         const qs_policy_t *pol, int n_pol, uint8_t *scores, tracks_t *t)
{
    if (n_pol < 0) {
        if (has_quals) memcpy(scores, qs, (size_t)rlen);
        return has_quals ? 3 : 0;
    }
    if (!has_quals) return 0;
    int force = 0, fault = 0;
    memset(scores, 0xFF, (size_t)rlen);
    uint8_t *mask = malloc((size_t)rlen + 1);
    for (int k = 0; k < n_pol; k++) {
        const qs_policy_t *p = &pol[k];
        if (p->read_cat != RC_NONE) {
            int proper = 0;
            switch (p->read_cat) {
            case RC_ALL: proper = 1; break;
            case RC_UNPLACED: proper = (flag & SAM_UNMAPPED) != 0; break;
            case RC_LOWER_MAPQ: proper = mapq < p->read_param; break;
            case RC_HIGHER_MAPQ: proper = mapq > p->read_param; break;
            }
            if (!proper) continue;
        }
        if (p->n_base_cats == 0) {
            switch (p->treatment) {
            case T_BIN:
                for (int64_t i = 0; i < rlen; i++) { if (qs[i] >= 110) fault = 0x80; scores[i] = (uint8_t)ILLUMINA_BIN[qs[i] < 110 ? qs[i] : 109]; }
                force = 1; break;
            case T_PRESERVE: memcpy(scores, qs, (size_t)rlen); force = 1; break;
            case T_DROP: force = 0; break;
            }
            continue;
        }
        if (pos == 0) continue;                                      /* NO_ALIGNMENT_START [:394-396] */
        if (t) rec_oracle_tracks_ensure_range(t, pos, sam_alignment_end(pos, flag, cigar, ncig) - pos);     /* alSpan [:334, :397] */
        memset(mask, 0, (size_t)rlen + 1);
        for (int b = 0; b < p->n_base_cats; b++) {
            int64_t rp = 0, refp;
            switch (p->base_cat[b] & 0xFF) {
            case BC_FLANKING_DELETION:
                for (int64_t c = 0; c < ncig; c++) {
                    int op = cigar[c] & 15; int64_t len = cigar[c] >> 4;
                    if (op == OP_D) { if (rp < rlen) mask[rp] = 1; if (rp + 1 < rlen) mask[rp + 1] = 1; }
                    if (consumes_read(op)) rp += len;
                }
                break;
            case BC_MATCH: case BC_MISMATCH:
                refp = pos;
                for (int64_t c = 0; c < ncig; c++) {
                    int op = cigar[c] & 15; int64_t len = cigar[c] >> 4;
                    if (consumes_read(op)) {
                        for (int64_t i = 0; i < len && rp + i < rlen; i++) {
                            int64_t z = refp + i - 1;                 /* tracks.baseAt(refPos + i) */
                            uint8_t tb = t ? tracks_base_at(t, (int32_t)(refp + i)) : (z >= 0 && z < ref_len) ? ref[z] : (uint8_t)'N';   /* beyond the window: 'N' [ReferenceTracks.java:141-146] */
                            int match = rb[rp + i] == tb;
                            mask[rp + i] = ((p->base_cat[b] & 0xFF) == BC_MATCH) ? match : !match;
                        }
                        rp += len;
                    }
                    if (consumes_ref(op)) refp += len;
                }
                break;
            case BC_INSERTION:
                for (int64_t c = 0; c < ncig; c++) {
                    int op = cigar[c] & 15; int64_t len = cigar[c] >> 4;
                    if (op == OP_I) for (int64_t i = 0; i < len && rp + i < rlen; i++) mask[rp + i] = 1;
                    if (consumes_read(op)) rp += len;
                }
                break;
            case BC_LOWER_COVERAGE:                                  /* [:447-465]: pos starts at 1, the mask index is pos + i - 1 */
                refp = pos;
                for (int64_t c = 0; c < ncig && t; c++) {
                    int op = cigar[c] & 15; int64_t len = cigar[c] >> 4;
                    if (op == OP_M || op == OP_EQ || op == OP_X)
                        for (int64_t i = 0; i < len && rp + i < rlen; i++) mask[rp + i] = tracks_coverage_at(t, (int32_t)(refp + i)) < (p->base_cat[b] >> 8);
                    if (consumes_read(op)) rp += len;
                    if (consumes_ref(op)) refp += len;
                }
                break;
            case BC_PILEUP:                                          /* [:466-470]: position = alignment start + index in the read */
                for (int64_t i = 0; i < rlen && t; i++) if (tracks_mismatches_at(t, (int32_t)(pos + i)) > (p->base_cat[b] >> 8)) mask[i] = 1;
                break;
            }
        }
        int64_t masked = 0;
        for (int64_t i = 0; i < rlen; i++) {
            if (!mask[i]) continue;
            switch (p->treatment) {
            case T_BIN: if (qs[i] >= 110) fault = 0x80; scores[i] = (uint8_t)ILLUMINA_BIN[qs[i] < 110 ? qs[i] : 109]; break;
            case T_DROP: scores[i] = 0xFF; break;
            default: scores[i] = qs[i]; break;
            }
            masked++;
        }
        if (masked > rlen / 2) force = 1;                             /* safety latch [:487-488] */
    }
    free(mask);
    return force | 2 | fault;
}

/* ------------------------------------------------------------------ normaliser (N1, N2) */

/* CramNormalizer.restoreReadBases [CramNormalizer.java:335-434] for one mapped record.
 * ref = region array, ref_off0 = refOffsetZeroBased.  Returns 0, or -1 where Java would throw AIOOBE. */
int rec_oracle_restore_bases(int32_t pos, int64_t rlen, int64_t nfeat, const uint8_t *f_op, const int32_t *f_pos,
                             const int32_t *f_val, const int32_t *f_aux, const uint8_t *payload, const uint8_t *ref,
                             int64_t ref_len, int64_t ref_off0, const uint8_t matrix[5], uint8_t *bases)
{
    int64_t al0 = (int64_t)pos - 1;
    if (rlen == 0) return 0;                            /* unknownBases || readLength == 0 -> NULL_SEQUENCE [:337-338] */
    memset(bases, 0, (size_t)rlen);
    if (nfeat == 0) {
        if (ref_len + ref_off0 < al0 + rlen) {
            memset(bases, 'N', (size_t)rlen);
            int64_t off = al0 - ref_off0, avail = ref_len - off;
            if (off > 0 && avail > 0) memcpy(bases, ref + off, (size_t)(rlen < avail ? rlen : avail));
        } else {
            if (al0 - ref_off0 < 0) return -1;
            memcpy(bases, ref + (al0 - ref_off0), (size_t)rlen);
        }
        return 0;                                       /* note: no normalizeBase on this path */
    }
    int64_t pir = 1, pis = 0;                           /* posInRead (1-based), posInSeq */
    for (int64_t i = 0; i < nfeat; i++) {
        for (; pir < f_pos[i]; pir++) {
            int64_t rp = al0 + pis++ - ref_off0;
            if (rp < 0 || pir > rlen) return -1;
            bases[pir - 1] = rp >= ref_len ? 'N' : ref[rp];
        }
        switch (f_op[i]) {
        case 'X': {
            int64_t rp = al0 + pis - ref_off0;
            if (rp < 0 || pir > rlen) return -1;
            uint8_t refb = NB(rp >= ref_len ? 'N' : ref[rp]);
            bases[pir++ - 1] = matrix_base(matrix, base_index(refb), f_val[i]);
            pis++;
            break;
        }
        case 'S': case 'I':
            for (int32_t k = 0; k < f_aux[i]; k++) { if (pir > rlen) return -1; bases[pir++ - 1] = payload[f_val[i] + k]; }
            break;
        case 'D': case 'N': pis += f_val[i]; break;
        case 'i': if (pir > rlen) return -1; bases[pir++ - 1] = (uint8_t)f_val[i]; break;
        default: break;                                  /* H, P, Q, B: no base consumed here */
        }
    }
    for (; pir <= rlen && al0 + pis - ref_off0 < ref_len; pir++, pis++) {
        if (al0 + pis - ref_off0 < 0) return -1;
        bases[pir - 1] = ref[al0 + pis - ref_off0];
    }
    for (int64_t i = 0; i < nfeat; i++)
        if (f_op[i] == 'B') bases[f_pos[i] - 1] = (uint8_t)f_val[i];
    for (int64_t i = 0; i < rlen; i++) bases[i] = NB(bases[i]);
    return 0;
}

/* CramNormalizer.restoreQualityScores [CramNormalizer.java:287-333].
 * force: scores[] holds the stored array (0xFF = -1).  Returns 1 when the result is NULL_QUALS ('*'). */
int rec_oracle_restore_quals(int force, int64_t rlen, int64_t nfeat, const uint8_t *f_op, const int32_t *f_pos,
                             const int32_t *f_val, const int32_t *f_aux, uint8_t *scores)
{
    const uint8_t dflt = '?' - '!';
    if (!force) {
        int star = 1;
        memset(scores, dflt, (size_t)rlen);
        for (int64_t i = 0; i < nfeat; i++) {
            if (f_op[i] == 'Q') { scores[f_pos[i] - 1] = (uint8_t)f_val[i]; star = 0; }
            else if (f_op[i] == 'B') { scores[f_pos[i] - 1] = (uint8_t)f_aux[i]; star = 0; }
        }
        return star;
    }
    int64_t missing = 0;
    for (int64_t i = 0; i < rlen; i++) if (scores[i] == 0xFF) { scores[i] = dflt; missing++; }
    return missing == rlen;
}

/* ------------------------------------------------------------------ series streams (W) */
/*
 * Writer.write [Writer.java:180-258] restricted to the series that the feature work feeds.  Stream ids are
 * EncodingKey ordinals: FN 15, FC 3, FP 2, BS 16, DL 5, BA 6, SC 27, QS 4, RS 26, HC 28, PD 29, MQ 11.
 * `streams[k]` must have room (caller sizes them generously); lens[k] receives the byte counts.
 * q_feats: per read, the Q features addQualityScores added (sorted in with the others by position, stable).
 */
enum { S_FN = 0, S_FC, S_FP, S_BS, S_DL, S_BA, S_SC, S_QS, S_RS, S_HC, S_PD, S_MQ, S_COUNT };

void rec_oracle_write_streams(int64_t n_reads, const uint8_t *bases, const int64_t *read_off, const uint16_t *flag,
                              const uint8_t *mapq, const int32_t *pos, const rec_features_t *ft, const uint8_t *force_qs,
                              const uint8_t *scores /* same offsets as bases */, uint8_t **streams, int64_t *lens)
{
    for (int k = 0; k < S_COUNT; k++) lens[k] = 0;
#define PUT(k, b) (streams[k][lens[k]++] = (uint8_t)(b))
#define PUT_ITF8(k, v) (lens[k] += rec_oracle_itf8((int32_t)(v), streams[k] + lens[k]))
    for (int64_t r = 0; r < n_reads; r++) {
        int64_t rlen = read_off[r + 1] - read_off[r];
        int forced = force_qs[r] & 1, has_scores = (force_qs[r] & 2) != 0;
        const uint8_t *sc = scores + read_off[r];
        if (!(flag[r] & SAM_UNMAPPED)) {
            /* merge the record's features with its Q features (position order, originals first on ties) */
            int64_t a = ft->feat_off[r], ae = ft->feat_off[r + 1];
            int64_t nq = 0;
            if (!forced && has_scores) for (int64_t i = 0; i < rlen; i++) if (sc[i] != 0xFF && !(sc[i] & 0x80)) nq++;
            PUT_ITF8(S_FN, (ae - a) + nq);
            int32_t prev = 0;
            int64_t qi = 0;                                   /* next read position (0-based) to scan for a Q score */
            for (;;) {
                int64_t qpos = -1;
                if (!forced && has_scores) {
                    while (qi < rlen && (sc[qi] == 0xFF || (sc[qi] & 0x80))) qi++;
                    if (qi < rlen) qpos = qi + 1;
                }
                int take_q;
                if (a < ae && qpos > 0) take_q = qpos < ft->f_pos[a];
                else if (a < ae) take_q = 0;
                else if (qpos > 0) take_q = 1;
                else break;
                if (take_q) {
                    PUT(S_FC, 'Q'); PUT_ITF8(S_FP, (int32_t)qpos - prev); prev = (int32_t)qpos;
                    PUT(S_QS, sc[qi]); qi++;
                    continue;
                }
                uint8_t op = ft->f_op[a];
                PUT(S_FC, op); PUT_ITF8(S_FP, ft->f_pos[a] - prev); prev = ft->f_pos[a];
                switch (op) {
                case 'X': PUT(S_BS, ft->f_val[a]); break;
                case 'D': PUT_ITF8(S_DL, ft->f_val[a]); break;
                case 'N': PUT_ITF8(S_RS, ft->f_val[a]); break;
                case 'H': PUT_ITF8(S_HC, ft->f_val[a]); break;
                case 'P': PUT_ITF8(S_PD, ft->f_val[a]); break;
                case 'i': PUT(S_BA, ft->f_val[a]); break;
                case 'S':
                    for (int32_t k = 0; k < ft->f_aux[a]; k++) PUT(S_SC, ft->payload[ft->f_val[a] + k]);
                    PUT(S_SC, '\t');                           /* ByteArrayStopEncoding */
                    break;
                }
                a++;
            }
            PUT_ITF8(S_MQ, mapq[r]);
            if (forced) for (int64_t i = 0; i < rlen; i++) PUT(S_QS, sc[i]);
        } else {
            for (int64_t i = 0; i < rlen; i++) PUT(S_BA, bases[read_off[r] + i]);
            if (forced) for (int64_t i = 0; i < rlen; i++) PUT(S_QS, sc[i]);
        }
    }
    (void)pos;
}


/* ------------------------------------------------------------------ the record fields in front of the read features (row W, first half)
 * Writer.write [Writer.java:139-172] driven from ContainerFactory.buildSlice [:179-184]; BF / MF from the SAM flag as
 * Sam2CramRecordFactory.createCramRecord sets them [:114-131,143-152].  Series order of out / lens:
 * BF CF RI RL AP RG RN MF NS NP TS NF TL.  cf = CramCompressionRecord.getCompressionFlags() (bit 1 detached, bit 2 mate
 * downstream).  Optional arrays may be NULL (read_group: -1, mate_ref: -1, mate_pos / tlen / nf / tl: 0, names: RN empty). */
enum { H_BF = 0, H_CF, H_RI, H_RL, H_AP, H_RG, H_RN, H_MF, H_NS, H_NP, H_TS, H_NF, H_TL, H_COUNT };
void rec_oracle_write_host_streams(int64_t n, const uint16_t *flag, const uint8_t *cf, const int32_t *ref_id, const int64_t *read_off,
                                   const int32_t *pos, const int32_t *read_group, const uint8_t *names, const int64_t *name_off,
                                   const int32_t *mate_ref, const int32_t *mate_pos, const int32_t *tlen, const int32_t *nf,
                                   const int32_t *tl, int ap_delta, int multi_ref, int32_t slice_start, int names_included,
                                   uint8_t **out, int64_t *lens)
{
    for (int k = 0; k < H_COUNT; k++) lens[k] = 0;
#define HPUT(k, b) (out[k][lens[k]++] = (uint8_t)(b))
#define HPUT_ITF8(k, v) (lens[k] += rec_oracle_itf8((int32_t)(v), out[k] + lens[k]))
    int32_t prev = slice_start;
    for (int64_t r = 0; r < n; r++) {
        const uint32_t f = flag[r];
        const uint32_t bf = (f & 1u) ? (f & (0x1u | 0x2u | 0x4u | 0x10u | 0x40u | 0x80u | 0x100u | 0x200u | 0x400u | 0x800u))
                                     : (f & (0x4u | 0x10u | 0x100u | 0x200u | 0x400u | 0x800u));
        const uint32_t mf = (f & 1u) ? (((f & 0x20u) ? 1u : 0u) | ((f & 0x8u) ? 2u : 0u)) : 0u;
        HPUT_ITF8(H_BF, bf);                                             /* :139 */
        HPUT(H_CF, cf[r]);                                               /* :140 */
        if (multi_ref) HPUT_ITF8(H_RI, ref_id[r]);                       /* :141-142 */
        HPUT_ITF8(H_RL, read_off[r + 1] - read_off[r]);                  /* :144 */
        if (ap_delta) { HPUT_ITF8(H_AP, pos[r] - prev); prev = pos[r]; } /* :146-147, ContainerFactory [:181-182] */
        else HPUT_ITF8(H_AP, pos[r]);                                    /* :149 */
        HPUT_ITF8(H_RG, read_group ? read_group[r] : -1);                /* :151 */
        const int detached = (cf[r] & 2) != 0;
        const int named = names && (names_included || detached);         /* :153-154 / :160-161 */
        if (detached) HPUT(H_MF, mf);                                    /* :159 */
        if (named) {
            for (int64_t i = name_off[r]; i < name_off[r + 1]; i++) HPUT(H_RN, names[i]);
            HPUT(H_RN, '\t');
        }
        if (detached) {
            HPUT_ITF8(H_NS, mate_ref ? mate_ref[r] : -1);                /* :163 */
            HPUT_ITF8(H_NP, mate_pos ? mate_pos[r] : 0);                 /* :164 */
            HPUT_ITF8(H_TS, tlen ? tlen[r] : 0);                         /* :165 */
        } else if (cf[r] & 4) {
            HPUT_ITF8(H_NF, nf ? nf[r] : 0);                             /* :168-169 */
        }
        HPUT_ITF8(H_TL, tl ? tl[r] : 0);                                 /* :172 */
    }
#undef HPUT
#undef HPUT_ITF8
}

/* ------------------------------------------------------------------ slice header fields and content digests
 *
 * ContainerFactory.buildSlice [ContainerFactory.java:138-176], what it derives from the records before any byte is written:
 *   nofRecords, bases = sum of readLength, sequenceId (first record's, -2 = MULTI_REFERENCE as soon as one differs),
 *   alignmentStart = min over the records with alignmentStart != 0, alignmentSpan = max(getAlignmentEnd()) - start + 1
 *   (-1 / 0 for a multi-reference slice or when no record is placed), and the slice tags of ContentDigests.ALL
 *   [cram/digest/ContentDigests.java:26-83]: for BD (bases) and SD (scores) the CRC32 of every record's array, summed as
 *   Java ints (Crc32Hasher + IntegerSumCombine), 4 bytes little endian.
 * The other four tags (B5, S5: SHA-512; B1, S1: SHA-1; ByteSumCombine) come out of MessageDigestHasher.asByteArray(), which
 * returns messageDigest.digest() -- of a digest that getValue() has just finished and thereby reset -- instead of the combined
 * value: in htsjdk 2.4.1 they are the hash of the EMPTY message whatever the records hold (tests/golden/streams_jvm.json
 * shows it: B1 = S1 = da39a3ee..., B5 = S5 = cf83e135...).  Constants, therefore, and no work.
 * CramCompressionRecord.getAlignmentEnd / calculateAlignmentBoundaries [CramCompressionRecord.java:160-194]: 0 for a record
 * with the unmapped flag; else start + span - 1 with span = readLength - (#'i') - len('I') - len('S') + len('D').  Reference
 * skips ('N') are NOT counted: that is the reference's arithmetic.
 * CramSerilization.convert then clamps a negative span to 0 [CramSerilization.java:184-186].
 * Pinned by tests/golden/streams_jvm.json ('slice', 'slice_tags').
 */
typedef struct {
    int32_t sequence_id, alignment_start, alignment_span;
    uint32_t n_records;
    uint64_t bases;
    uint32_t bd, sd;
} slice_header_t;

static uint32_t crc32_table[256];
static void crc32_init(void)
{
    if (crc32_table[1]) return;
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc32_table[i] = c;
    }
}
uint32_t rec_oracle_crc32(const uint8_t *p, int64_t n)      /* java.util.zip.CRC32 */
{
    crc32_init();
    uint32_t c = 0xFFFFFFFFu;
    for (int64_t i = 0; i < n; i++) c = crc32_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
    return ~c;
}

/* scores: record.qualityScores as addQualityScores / createCramRecord leave it, laid out like bases (0xFF = -1);
 * rflags bit0 = NULL_QUALS (empty array).  Features as rec_oracle_extract gives them (Q features, if merged in, are ignored). */
void rec_oracle_slice_header(int64_t n, const uint8_t *bases, const uint8_t *scores, const int64_t *read_off, const uint8_t *rflags,
                             const int32_t *ref_id, const int32_t *pos, const uint16_t *flag, const uint8_t *f_op,
                             const int32_t *f_val, const int32_t *f_aux, const int64_t *feat_off, slice_header_t *out)
{
    memset(out, 0, sizeof *out);
    out->n_records = (uint32_t)n;
    out->sequence_id = n ? (ref_id ? ref_id[0] : 0) : -1;            /* ref_id == NULL: every record on reference 0 */
    int32_t min_start = INT32_MAX, max_end = 0;
    for (int64_t r = 0; r < n; r++) {
        const int64_t lo = read_off[r], rlen = read_off[r + 1] - lo;
        out->bases += (uint64_t)rlen;
        out->bd += rec_oracle_crc32(bases + lo, rlen);
        out->sd += rec_oracle_crc32(scores + lo, (rflags && (rflags[r] & RF_NO_QUALS)) ? 0 : rlen);
        if (out->sequence_id == -2) continue;
        if (out->sequence_id != (ref_id ? ref_id[r] : 0)) { out->sequence_id = -2; continue; }
        if (pos[r] == 0) continue;
        int32_t end = 0;
        if (!(flag[r] & SAM_UNMAPPED)) {
            int32_t span = (int32_t)rlen;
            for (int64_t f = feat_off[r]; f < feat_off[r + 1]; f++) {
                if (f_op[f] == 'i') span -= 1;
                else if (f_op[f] == 'S') span -= f_aux[f];
                else if (f_op[f] == 'D') span += f_val[f];
            }
            end = pos[r] + span - 1;
        }
        if (pos[r] < min_start) min_start = pos[r];
        if (end > max_end) max_end = end;
    }
    if (out->sequence_id == -2 || min_start == INT32_MAX) { out->alignment_start = -1; out->alignment_span = 0; }
    else { out->alignment_start = min_start; out->alignment_span = max_end - min_start + 1; }
    if (out->alignment_span < 0) out->alignment_span = 0;
}
