/*
 * oracle/rans_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, single-threaded CPU restatement of htsjdk 2.4.1's rANS block codec
 * (htsjdk/samtools/cram/encoding/rans/*, bundled as bytecode in
 * /root/reference/lib/htsjdk-2.4.1.jar and /root/reference/cramtools-3.0.jar).
 * It is the checker for the CUDA path; nothing under cramtools_b200/ may call
 * it.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs load this file's shared object.
 *
 * Parity pin: tests/golden/rans_jvm_*.json holds (input, order, output)
 * triples produced by executing the reference jar's own bytecode with
 * oracle/minijvm.py (see that file); tests/test_oracle_rans.py checks this
 * restatement byte-for-byte against them.
 *
 * Citations "[X.java:NN]" are the original htsjdk source lines recorded in the
 * class files' LineNumberTable.
 *
 * Java semantics kept explicit: int = int32 wrap-around, long = int64,
 * (int)double truncates toward zero.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TF_SHIFT 12
#define TOTFREQ 4096
#define RANS_BYTE_L (1u << 23)

enum {
    RANS_OK = 0,
    RANS_ERR_INPUT_LENGTH = -1,   /* "Incorrect input length."          [RANS.java:37] */
    RANS_ERR_OUT_TOO_SMALL = -2,  /* "Output buffer too small ..."      [RANS.java:44] */
    RANS_ERR_UNKNOWN_ORDER = -3,  /* "Unknown rANS order: X"            [RANS.java:55] */
    RANS_ERR_CORRUPT = -4         /* Java would throw AIOOBE / BufferUnderflow */
};

typedef struct {
    uint32_t x_max, rcp_freq, bias, cmpl_freq, rcp_shift;
} enc_sym_t;

/* Encoding.RansEncSymbolInit [Encoding.java:21-43] */
static void enc_sym_init(enc_sym_t *s, uint32_t start, uint32_t freq)
{
    s->x_max = ((RANS_BYTE_L >> TF_SHIFT) << 8) * freq;
    s->cmpl_freq = (1u << TF_SHIFT) - freq;
    if (freq < 2) {
        s->rcp_freq = 0xFFFFFFFFu;
        s->rcp_shift = 0;
        s->bias = start + (1u << TF_SHIFT) - 1;
    } else {
        uint32_t shift = 0;
        while (freq > (1u << shift)) shift++;
        s->rcp_freq = (uint32_t)(((1ull << (shift + 31)) + freq - 1) / freq);
        s->rcp_shift = shift - 1;
        s->bias = start;
    }
    s->rcp_shift += 32;
}

/* byte sink that grows toward lower addresses == "emit then reverse whole buffer"
 * of E04/E14 + Utils.reverse [rans/Utils.java:7-33] */
typedef struct { uint8_t *base, *p; int overflow; } bsink_t;
static inline void emit(bsink_t *k, uint8_t b)
{
    if (k->p == k->base) { k->overflow = 1; return; }
    *--k->p = b;
}

/* Encoding.RansEncPutSymbol [Encoding.java:46-71] */
static inline uint32_t put_symbol(uint32_t x, bsink_t *k, const enc_sym_t *s)
{
    if (x >= s->x_max) {
        emit(k, (uint8_t)(x & 0xFF)); x >>= 8;
        if (x >= s->x_max) { emit(k, (uint8_t)(x & 0xFF)); x >>= 8; }
    }
    uint32_t q = (uint32_t)(((uint64_t)x * s->rcp_freq) >> s->rcp_shift);
    return x + s->bias + q * s->cmpl_freq;   /* int32 wrap-around */
}

static inline void emit_state_be_then_reversed(bsink_t *k, uint32_t r)
{   /* putInt big-endian in emission order: b3,b2,b1,b0; the sink reverses */
    emit(k, (uint8_t)(r >> 24)); emit(k, (uint8_t)(r >> 16));
    emit(k, (uint8_t)(r >> 8));  emit(k, (uint8_t)r);
}

/* one frequency row, shared by o0 and each o1 context
 * Frequencies.writeFrequencies_o0 [Frequencies.java:215-244], _o1 rows [:297-324] */
static size_t write_row(uint8_t *cp, const int *F)
{
    uint8_t *s = cp; int rle = 0;
    for (int j = 0; j < 256; j++) {
        if (!F[j]) continue;
        if (rle > 0) rle--;
        else {
            *cp++ = (uint8_t)j;
            if (j > 0 && F[j - 1] != 0) {
                int r = j + 1;
                while (r < 256 && F[r] != 0) r++;
                rle = r - (j + 1);
                *cp++ = (uint8_t)rle;
            }
        }
        if (F[j] < 128) *cp++ = (uint8_t)F[j];
        else { *cp++ = (uint8_t)(128 | (F[j] >> 8)); *cp++ = (uint8_t)(F[j] & 0xFF); }
    }
    *cp++ = 0;
    return (size_t)(cp - s);
}

/* Frequencies.calcFrequencies_o0 [Frequencies.java:109-145] */
void rans_oracle_freqs_o0(const uint8_t *in, uint32_t n, int *F)
{
    memset(F, 0, 256 * sizeof(int));
    for (uint32_t i = 0; i < n; i++) F[in[i]]++;
    int64_t T = n;
    uint64_t tr = (((uint64_t)TOTFREQ << 31) / (uint64_t)T) + ((1u << 30) / (uint64_t)T);
    int m = 0, M = 0, fsum = 0;
    for (int j = 0; j < 256; j++) {
        if (!F[j]) continue;
        if (m < F[j]) { m = F[j]; M = j; }
        if ((F[j] = (int)(((uint64_t)F[j] * tr) >> 31)) == 0) F[j] = 1;
        fsum += F[j];
    }
    fsum++;
    if (fsum < TOTFREQ) F[M] += TOTFREQ - fsum;
    else F[M] -= fsum - TOTFREQ;
}

/* Frequencies.calcFrequencies_o1 [Frequencies.java:149-193]; F is [256][256] */
void rans_oracle_freqs_o1(const uint8_t *in, uint32_t n, int *F, int *T)
{
    memset(F, 0, 256 * 256 * sizeof(int));
    memset(T, 0, 256 * sizeof(int));
    int last = 0;
    for (uint32_t i = 0; i < n; i++) { int c = in[i]; F[last * 256 + c]++; T[last]++; last = c; }
    uint32_t q = n >> 2;
    F[in[1 * q]]++; F[in[2 * q]]++; F[in[3 * q]]++; T[0] += 3;
    for (int i = 0; i < 256; i++) {
        if (T[i] == 0) continue;
        double p = ((double)TOTFREQ) / (double)T[i];
        int t2 = 0, m = 0, M = 0;
        int *Fi = F + i * 256;
        for (int j = 0; j < 256; j++) {
            if (!Fi[j]) continue;
            if (m < Fi[j]) { m = Fi[j]; M = j; }
            if ((Fi[j] = (int)((double)Fi[j] * p)) == 0) Fi[j] = 1;
            t2 += Fi[j];
        }
        t2++;
        if (t2 < TOTFREQ) Fi[M] += TOTFREQ - t2;
        else Fi[M] -= t2 - TOTFREQ;
    }
}

/* RANS.allocateIfNeeded [RANS.java:79-85]: (int)(1.05*n + 257*257*3 + 4) */
uint32_t rans_oracle_bound(uint32_t n)
{
    return (uint32_t)(int)(1.05 * (double)n + 257 * 257 * 3 + 4);
}

/* RANS.finalizeCompressed [RANS.java:132-142] */
static void write_prefix(uint8_t *out, int order, uint32_t csize, uint32_t rsize)
{
    out[0] = (uint8_t)order;
    for (int i = 0; i < 4; i++) out[1 + i] = (uint8_t)(csize >> (8 * i));
    for (int i = 0; i < 4; i++) out[5 + i] = (uint8_t)(rsize >> (8 * i));
}

/* o0: Frequencies.buildSyms_o0 [:197-211], E04.compress [E04.java:12-57] */
static int compress_o0(const uint8_t *in, uint32_t n, uint8_t *out, uint32_t cap, uint32_t *out_len)
{
    int F[256];
    enc_sym_t syms[256];
    rans_oracle_freqs_o0(in, n, F);
    uint32_t x = 0;
    for (int j = 0; j < 256; j++) if (F[j]) { enc_sym_init(&syms[j], x, (uint32_t)F[j]); x += (uint32_t)F[j]; }

    uint8_t tab[1024];
    size_t tlen = write_row(tab, F);
    if (9 + tlen > cap) return RANS_ERR_OUT_TOO_SMALL;

    bsink_t k = { out + 9 + tlen, out + cap, 0 };
    uint32_t r0 = RANS_BYTE_L, r1 = RANS_BYTE_L, r2 = RANS_BYTE_L, r3 = RANS_BYTE_L;
    uint32_t i = n & 3;
    switch (i) {                                    /* [E04.java:22-31] fall-through */
    case 3: r2 = put_symbol(r2, &k, &syms[in[n - (i - 2)]]); /* fallthrough */
    case 2: r1 = put_symbol(r1, &k, &syms[in[n - (i - 1)]]); /* fallthrough */
    case 1: r0 = put_symbol(r0, &k, &syms[in[n - (i - 0)]]); /* fallthrough */
    default: break;
    }
    for (i = n & ~3u; i > 0; i -= 4) {              /* [E04.java:35-44] */
        r3 = put_symbol(r3, &k, &syms[in[i - 1]]);
        r2 = put_symbol(r2, &k, &syms[in[i - 2]]);
        r1 = put_symbol(r1, &k, &syms[in[i - 3]]);
        r0 = put_symbol(r0, &k, &syms[in[i - 4]]);
    }
    emit_state_be_then_reversed(&k, r3); emit_state_be_then_reversed(&k, r2);
    emit_state_be_then_reversed(&k, r1); emit_state_be_then_reversed(&k, r0);
    if (k.overflow) return RANS_ERR_OUT_TOO_SMALL;
    size_t plen = (size_t)((out + cap) - k.p);
    memcpy(out + 9, tab, tlen);
    memmove(out + 9 + tlen, k.p, plen);
    write_prefix(out, 0, (uint32_t)(tlen + plen), n);
    *out_len = (uint32_t)(9 + tlen + plen);
    return RANS_OK;
}

/* o1: buildSyms_o1 [:248-265], writeFrequencies_o1 [:269-328], E14.compress [E14.java:13-86] */
static int compress_o1(const uint8_t *in, uint32_t n, uint8_t *out, uint32_t cap, uint32_t *out_len)
{
    int *F = (int *)malloc(256 * 256 * sizeof(int));
    int T[256];
    enc_sym_t *syms = (enc_sym_t *)malloc(256 * 256 * sizeof(enc_sym_t));
    uint8_t *tab = (uint8_t *)malloc(257 * 257 * 3 + 16);
    int rc = RANS_OK;
    rans_oracle_freqs_o1(in, n, F, T);

    size_t tlen = 0; int rle_i = 0;
    for (int i = 0; i < 256; i++) {
        if (T[i] == 0) continue;
        if (rle_i > 0) rle_i--;
        else {
            tab[tlen++] = (uint8_t)i;
            if (i > 0 && T[i - 1] != 0) {
                int r = i + 1;
                while (r < 256 && T[r] != 0) r++;
                rle_i = r - (i + 1);
                tab[tlen++] = (uint8_t)rle_i;
            }
        }
        const int *Fi = F + i * 256;
        tlen += write_row(tab + tlen, Fi);
        uint32_t x = 0;
        for (int j = 0; j < 256; j++) if (Fi[j]) { enc_sym_init(&syms[i * 256 + j], x, (uint32_t)Fi[j]); x += (uint32_t)Fi[j]; }
    }
    tab[tlen++] = 0;
    if (9 + tlen > cap) { rc = RANS_ERR_OUT_TOO_SMALL; goto done; }

    {
        bsink_t k = { out + 9 + tlen, out + cap, 0 };
        uint32_t r0 = RANS_BYTE_L, r1 = RANS_BYTE_L, r2 = RANS_BYTE_L, r3 = RANS_BYTE_L;
        int64_t q = n >> 2;
        int64_t i0 = 1 * q - 2, i1 = 2 * q - 2, i2 = 3 * q - 2, i3;
        uint8_t l0 = (i0 + 1 >= 0) ? in[i0 + 1] : 0;
        uint8_t l1 = (i1 + 1 >= 0) ? in[i1 + 1] : 0;
        uint8_t l2 = (i2 + 1 >= 0) ? in[i2 + 1] : 0;
        uint8_t l3 = in[n - 1];
        for (i3 = (int64_t)n - 2; i3 > 4 * q - 2 && i3 >= 0; i3--) {   /* [E14.java:45-49] */
            uint8_t c3 = in[i3];
            r3 = put_symbol(r3, &k, &syms[c3 * 256 + l3]);
            l3 = c3;
        }
        for (; i0 >= 0; i0--, i1--, i2--, i3--) {                       /* [E14.java:51-66] */
            uint8_t c0 = in[i0], c1 = in[i1], c2 = in[i2], c3 = in[i3];
            r3 = put_symbol(r3, &k, &syms[c3 * 256 + l3]);
            r2 = put_symbol(r2, &k, &syms[c2 * 256 + l2]);
            r1 = put_symbol(r1, &k, &syms[c1 * 256 + l1]);
            r0 = put_symbol(r0, &k, &syms[c0 * 256 + l0]);
            l0 = c0; l1 = c1; l2 = c2; l3 = c3;
        }
        r3 = put_symbol(r3, &k, &syms[0 * 256 + l3]);                   /* [E14.java:68-71] */
        r2 = put_symbol(r2, &k, &syms[0 * 256 + l2]);
        r1 = put_symbol(r1, &k, &syms[0 * 256 + l1]);
        r0 = put_symbol(r0, &k, &syms[0 * 256 + l0]);
        emit_state_be_then_reversed(&k, r3); emit_state_be_then_reversed(&k, r2);
        emit_state_be_then_reversed(&k, r1); emit_state_be_then_reversed(&k, r0);
        if (k.overflow) { rc = RANS_ERR_OUT_TOO_SMALL; goto done; }
        size_t plen = (size_t)((out + cap) - k.p);
        memcpy(out + 9, tab, tlen);
        memmove(out + 9 + tlen, k.p, plen);
        write_prefix(out, 1, (uint32_t)(tlen + plen), n);
        *out_len = (uint32_t)(9 + tlen + plen);
    }
done:
    free(F); free(syms); free(tab);
    return rc;
}

/* RANS.compress [RANS.java:60-73] */
int rans_oracle_compress(const uint8_t *in, uint32_t n, int order, uint8_t *out, uint32_t cap, uint32_t *out_len)
{
    if (n == 0) { *out_len = 0; return RANS_OK; }
    if (n < 4) order = 0;
    if (order == 0) return compress_o0(in, n, out, cap, out_len);
    if (order == 1) return compress_o1(in, n, out, cap, out_len);
    return RANS_ERR_UNKNOWN_ORDER;
}

/* ---------------------------------------------------------------- decode */

typedef struct { uint16_t F[256], C[256]; uint8_t R[TOTFREQ]; } dec_t;

typedef struct { const uint8_t *p, *end; int under; } bsrc_t;
static inline uint32_t getb(bsrc_t *s) { if (s->p >= s->end) { s->under = 1; return 0; } return *s->p++; }
static inline int peekb(bsrc_t *s) { if (s->p >= s->end) { s->under = 1; return -1; } return *s->p; }

/* Frequencies.readStats_o0 [Frequencies.java:13-46]; also one o1 row [:53-94] */
static int read_row(bsrc_t *s, dec_t *d, int o1)
{
    int rle = 0, x = 0, j = (int)getb(s);
    do {
        int f = (int)getb(s);
        if (f >= 128) { f &= ~128; f = ((f & 127) << 8) | (int)getb(s); }
        if (o1 && f == 0) f = TOTFREQ;                                  /* [:68-69] */
        if (x + f > TOTFREQ) return RANS_ERR_CORRUPT;                   /* Java: AIOOBE from Arrays.fill */
        d->F[j] = (uint16_t)f; d->C[j] = (uint16_t)x;
        memset(d->R + x, j, (size_t)f);
        x += f;
        if (rle == 0 && peekb(s) == j + 1) { j = (int)getb(s); rle = (int)getb(s); }
        else if (rle != 0) { rle--; j++; if (j > 255) return RANS_ERR_CORRUPT; }
        else j = (int)getb(s);
        if (s->under) return RANS_ERR_CORRUPT;
    } while (j != 0);
    return RANS_OK;
}

/* Decoding.RansDecRenorm [Decoding.java:84-90] */
static inline uint32_t renorm(uint32_t x, bsrc_t *s)
{
    if (x < RANS_BYTE_L) {
        do x = (x << 8) | getb(s); while (x < RANS_BYTE_L);
    }
    return x;
}
static inline uint32_t get_le32(bsrc_t *s)
{
    uint32_t v = getb(s); v |= getb(s) << 8; v |= getb(s) << 16; v |= getb(s) << 24; return v;
}

/* D04.uncompress [D04.java:9-85] */
static int uncompress_o0(bsrc_t *s, uint8_t *out, uint32_t n)
{
    dec_t *d = (dec_t *)calloc(1, sizeof(dec_t));
    int rc = read_row(s, d, 0);
    if (rc) { free(d); return rc; }
    uint32_t r[4];
    for (int k = 0; k < 4; k++) r[k] = get_le32(s);
    uint32_t out_end = n & ~3u;
    for (uint32_t i = 0; i < out_end; i += 4) {
        uint8_t c[4];
        for (int k = 0; k < 4; k++) c[k] = d->R[r[k] & (TOTFREQ - 1)];
        for (int k = 0; k < 4; k++) out[i + k] = c[k];
        for (int k = 0; k < 4; k++) r[k] = d->F[c[k]] * (r[k] >> TF_SHIFT) + (r[k] & (TOTFREQ - 1)) - d->C[c[k]];
        for (int k = 0; k < 4; k++) r[k] = renorm(r[k], s);
    }
    /* tail [D04.java:60-83]: lanes 0..(n&3)-1 supply one symbol each, no further state advance needed */
    for (uint32_t k = 0; k < (n & 3); k++) out[out_end + k] = d->R[r[k] & (TOTFREQ - 1)];
    rc = s->under ? RANS_ERR_CORRUPT : RANS_OK;
    free(d);
    return rc;
}

/* readStats_o1 [Frequencies.java:50-106], D14.uncompress [D14.java:9-70] */
static int uncompress_o1(bsrc_t *s, uint8_t *out, uint32_t n)
{
    dec_t *D = (dec_t *)calloc(256, sizeof(dec_t));
    int rc = RANS_OK;
    {
        int rle_i = 0, i = (int)getb(s);
        do {
            rc = read_row(s, &D[i], 1);
            if (rc) goto done;
            if (rle_i == 0 && peekb(s) == i + 1) { i = (int)getb(s); rle_i = (int)getb(s); }
            else if (rle_i != 0) { rle_i--; i++; if (i > 255) { rc = RANS_ERR_CORRUPT; goto done; } }
            else i = (int)getb(s);
            if (s->under) { rc = RANS_ERR_CORRUPT; goto done; }
        } while (i != 0);
    }
    {
        uint32_t r[4]; uint32_t l[4] = {0, 0, 0, 0};
        for (int k = 0; k < 4; k++) r[k] = get_le32(s);
        uint32_t q = n >> 2;
        uint32_t i4[4] = {0, q, 2 * q, 3 * q};
        for (; i4[0] < q; i4[0]++, i4[1]++, i4[2]++, i4[3]++) {
            uint8_t c[4];
            for (int k = 0; k < 4; k++) c[k] = D[l[k]].R[r[k] & (TOTFREQ - 1)];
            for (int k = 0; k < 4; k++) out[i4[k]] = c[k];
            for (int k = 0; k < 4; k++)
                r[k] = D[l[k]].F[c[k]] * (r[k] >> TF_SHIFT) + (r[k] & (TOTFREQ - 1)) - D[l[k]].C[c[k]];
            for (int k = 0; k < 4; k++) r[k] = renorm(r[k], s);
            for (int k = 0; k < 4; k++) l[k] = c[k];
        }
        for (; i4[3] < n; i4[3]++) {                                     /* [D14.java:62-68] */
            uint8_t c3 = D[l[3]].R[r[3] & (TOTFREQ - 1)];
            out[i4[3]] = c3;
            r[3] = D[l[3]].F[c3] * (r[3] >> TF_SHIFT) + (r[3] & (TOTFREQ - 1)) - D[l[3]].C[c3];
            r[3] = renorm(r[3], s);
            l[3] = c3;
        }
        if (s->under) rc = RANS_ERR_CORRUPT;
    }
done:
    free(D);
    return rc;
}

/* RANS.uncompress [RANS.java:29-55].  *out_len receives the raw size even on
 * RANS_ERR_OUT_TOO_SMALL so the caller can size its buffer. */
int rans_oracle_uncompress(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t cap, uint32_t *out_len)
{
    if (in_len == 0) { *out_len = 0; return RANS_OK; }
    if (in_len < 9) return RANS_ERR_CORRUPT;
    int order = in[0];
    uint32_t in_sz = in[1] | (in[2] << 8) | (in[3] << 16) | ((uint32_t)in[4] << 24);
    if (in_sz != in_len - 1 - 4 - 4) return RANS_ERR_INPUT_LENGTH;
    uint32_t out_sz = in[5] | (in[6] << 8) | (in[7] << 16) | ((uint32_t)in[8] << 24);
    *out_len = out_sz;
    if (cap < out_sz) return RANS_ERR_OUT_TOO_SMALL;
    bsrc_t s = { in + 9, in + in_len, 0 };
    if (order == 0) return uncompress_o0(&s, out, out_sz);
    if (order == 1) return uncompress_o1(&s, out, out_sz);
    return RANS_ERR_UNKNOWN_ORDER;
}

/* ---------------------------------------------------------------- multi-threaded driver
 * Runs the single-threaded codec above over independent blocks on `nthreads` host threads (the
 * reference's own concurrency model: one container / block per pool thread,
 * paralell/BamToCram_OBA_Function.java:51-94).  Used by bench.py for the CPU baseline only. */
#include <pthread.h>

typedef struct {
    int encode, nthreads, tid;
    uint32_t nblocks;
    const uint8_t *const *in; const uint32_t *in_len; const uint8_t *order;
    uint8_t *const *out; const uint32_t *out_cap; uint32_t *out_len; int *rc;
} many_t;

static void *many_worker(void *arg)
{
    many_t *m = (many_t *)arg;
    for (uint32_t i = (uint32_t)m->tid; i < m->nblocks; i += (uint32_t)m->nthreads) {
        if (m->encode) m->rc[i] = rans_oracle_compress(m->in[i], m->in_len[i], m->order[i], m->out[i], m->out_cap[i], &m->out_len[i]);
        else m->rc[i] = rans_oracle_uncompress(m->in[i], m->in_len[i], m->out[i], m->out_cap[i], &m->out_len[i]);
    }
    return NULL;
}

int rans_oracle_many(int encode, int nthreads, uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len,
                     const uint8_t *order, uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len, int *rc)
{
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    many_t a[256];
    for (int t = 0; t < nthreads; t++) {
        a[t] = (many_t){encode, nthreads, t, nblocks, in, in_len, order, out, out_cap, out_len, rc};
        pthread_create(&th[t], NULL, many_worker, &a[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    for (uint32_t i = 0; i < nblocks; i++) if (rc[i]) return rc[i];
    return 0;
}
