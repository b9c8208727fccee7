/*
 * cramgpu.h -- C ABI of libcramgpu.so, the B200 (sm_100a) CRAM slice codec.
 *
 * This is the drop-in boundary for the data-parallel hot path of cramtools'
 * `cram` / `bam` commands.  The reference (enasequence/cramtools 3.0 + htsjdk
 * 2.4.1) is pure Java and has no FFI today; each entry point below names the
 * Java method it stands in for ("replaces").  Citations of the form
 * [X.java:NN] are htsjdk 2.4.1 source lines as recorded in the class files
 * inside /root/reference/lib/htsjdk-2.4.1.jar; plain paths are under
 * /root/reference/.  INTEGRATION.md shows the JNI / Panama shim that binds
 * these symbols from the Java side.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++ / torch types.
 *  - every function returns 0 (CRAMGPU_OK) or a negative cramgpu_status;
 *    cramgpu_last_error(ctx) gives the message of the last failure on ctx,
 *    worded like the RuntimeException the reference throws.
 *  - a cramgpu_ctx owns one CUDA stream, device scratch and pinned staging.
 *    A ctx may be used by one thread at a time; use one ctx per calling
 *    thread (the reference's codecs are static and stateless and are called
 *    concurrently from pool threads: paralell/BamToCram_OBA_Function.java:51-94).
 *  - "host" entry points take caller-owned host memory, never retain it past
 *    the call and never modify inputs.  "dev" entry points take device
 *    pointers and enqueue on a caller stream; they are what bench.py times
 *    with inputs resident in HBM.
 *  - there is no CPU fallback: without a usable CUDA device cramgpu_init
 *    fails with CRAMGPU_ERR_CUDA.
 */
#ifndef CRAMGPU_H
#define CRAMGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cramgpu_ctx cramgpu_ctx;

typedef enum cramgpu_status {
    CRAMGPU_OK = 0,
    CRAMGPU_ERR_INPUT_LENGTH = -1,  /* "Incorrect input length."                 [RANS.java:37] */
    CRAMGPU_ERR_OUT_TOO_SMALL = -2, /* decoder: "Output buffer too small to fit N bytes." [RANS.java:44];
                                       coder: "Insufficient buffer size." [RANS.allocateIfNeeded, RANS.java:83] */
    CRAMGPU_ERR_UNKNOWN_ORDER = -3, /* "Unknown rANS order: X"                   [RANS.java:55] */
    CRAMGPU_ERR_CORRUPT = -4,       /* malformed table/payload; Java: ArrayIndexOutOfBounds / BufferUnderflow */
    CRAMGPU_ERR_CUDA = -5,          /* CUDA runtime failure or no device */
    CRAMGPU_ERR_ARG = -6,           /* bad argument (null pointer, misaligned device offset, ...) */
    CRAMGPU_ERR_UNSUPPORTED = -7    /* input the reference itself cannot process (documented per call) */
} cramgpu_status;

/* ------------------------------------------------------------------ context */

/* Create a context on CUDA device `device`.  No Java counterpart (the JVM codecs are static). */
int cramgpu_init(int device, cramgpu_ctx **ctx);
void cramgpu_destroy(cramgpu_ctx *ctx);
/* Message of the last failure on ctx (ctx == NULL: last failure of cramgpu_init on this thread). */
const char *cramgpu_last_error(const cramgpu_ctx *ctx);
/* Block until everything enqueued on the context's own stream has finished. */
int cramgpu_sync(cramgpu_ctx *ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches claim). */
uint64_t cramgpu_launch_count(const cramgpu_ctx *ctx);
/* Tuning knobs (no counterpart in the reference).  Returns CRAMGPU_ERR_ARG for an unknown name.
 *   "dec_wide_pool_blocks"  how many largest-size (201 KB) table images of order-1 blocks with more than 64 distinct bytes the
 *                           decoder's scratch holds per call (default 2048); blocks beyond it take a slower path */
int cramgpu_set_option(cramgpu_ctx *ctx, const char *name, uint64_t value);

/* Copy `bytes` from device memory (e.g. a result of a *_dev call) to host memory and wait for it.  `stream`:
 * the stream the producing work was enqueued on (NULL = the context's own).  Convenience for shims that own no
 * CUDA binding of their own; no Java counterpart. */
int cramgpu_copy_to_host(cramgpu_ctx *ctx, void *dst_host, const void *src_dev, uint64_t bytes, void *stream);

/*
 * Per-kernel device timing, the CUDA-event counterpart of the reference's System.nanoTime() stage
 * counters (Cram2Bam.java:172-178,291-294; Container.buildHeaderTime/buildSlicesTime).  While enabled,
 * every kernel the context launches is bracketed by events on the launching stream.
 * cramgpu_profile_read drains them: synchronise the stream first; names[i] / ms[i] receive up to `cap`
 * spans in launch order; returns the number of spans recorded since the previous read.
 */
int cramgpu_profile_enable(cramgpu_ctx *ctx, int on);
int cramgpu_profile_read(cramgpu_ctx *ctx, uint32_t cap, const char **names, float *ms);

/* ------------------------------------------------------------------ rANS block codec */

/* Worst-case compressed size the reference allocates for n raw bytes:
 * (int)(1.05*n + 257*257*3 + 4).  Replaces RANS.allocateIfNeeded [RANS.java:79-85]. */
uint32_t cramgpu_rans_bound(uint32_t n);

/*
 * Compress nblocks independent byte blocks, each exactly as
 *   ExternalCompression.rans(byte[] data, int order)          [ExternalCompression.java:89-109]
 *   -> RANS.compress(ByteBuffer in, ORDER order, ByteBuffer out) [RANS.java:60-73]
 * would: empty input -> empty output; n < 4 -> order 0 regardless of `order`;
 * output = order:u8 | compressedSize:u32le | rawSize:u32le | frequency table | payload,
 * byte-identical to htsjdk.  One call carries every rANS block of >= 1 slices
 * (the loop ContainerFactory.buildSlice runs over external blocks, [ContainerFactory.java:190-203]).
 *
 *   in[i], in_len[i]   caller-owned raw bytes of block i
 *   order[i]           0 or 1 (anything else: CRAMGPU_ERR_UNKNOWN_ORDER for that block)
 *   out[i], out_cap[i] caller-owned destination; cramgpu_rans_bound(in_len[i]) always suffices.  The reference refuses a
 *                      caller's buffer below that bound outright ("Insufficient buffer size."); here a smaller buffer is
 *                      accepted when the compressed block fits and fails with that message (CRAMGPU_ERR_OUT_TOO_SMALL,
 *                      nothing written) when it does not
 *   out_len[i]         bytes written
 *   status             optional per-block status array (may be NULL)
 * Returns the first non-OK per-block status, else CRAMGPU_OK.
 *
 * Where the bytes live: in[i] and out[i] may each point to pageable host memory (staged), to host memory the
 * caller page-locked (handed to the DMA engine as it is), or to DEVICE memory (unified addressing tells them
 * apart).  A raw block in device memory is coded where it is -- e.g. a series cramgpu_features_encode_resident
 * left on the device -- and a destination in device memory receives its bytes without crossing the bus; that is
 * how a host keeps intermediate series on the GPU between calls.  (Compressed INPUT blocks of the decoder are
 * host memory; blocks that already sit on the device go through cramgpu_rans_decode_dev.)  No counterpart in
 * the reference, whose blocks are always Java arrays.
 */
int cramgpu_rans_encode_batch(cramgpu_ctx *ctx, uint32_t nblocks,
                              const uint8_t *const *in, const uint32_t *in_len, const uint8_t *order,
                              uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len,
                              int32_t *status);

/*
 * Raw (uncompressed) size of each compressed block, read from its 9-byte prefix, so the
 * caller can allocate.  Host-only; validates compressedSize as RANS.uncompress does
 * ("Incorrect input length.", [RANS.java:33-37]).  Empty block -> 0.
 */
int cramgpu_rans_raw_sizes(uint32_t nblocks, const uint8_t *const *in, const uint32_t *in_len,
                           uint32_t *raw_len, int32_t *status);

/*
 * Decompress nblocks blocks, each as ExternalCompression.unrans(byte[]) [ExternalCompression.java:112-113]
 * -> RANS.uncompress(ByteBuffer in, ByteBuffer out) [RANS.java:29-55] would.  One call carries every
 * rANS block of >= 1 slices (Block.uncompress call site, [Block.java:228-262], driven from
 * ContainerParser.getRecords [ContainerParser.java:155-158]).
 */
int cramgpu_rans_decode_batch(cramgpu_ctx *ctx, uint32_t nblocks,
                              const uint8_t *const *in, const uint32_t *in_len,
                              uint8_t *const *out, const uint32_t *out_cap, uint32_t *out_len,
                              int32_t *status);

/*
 * Device-resident variants.  Same results, but data stay in HBM.
 *   d_in            device buffer holding all blocks; block i is d_in[in_off[i] .. +in_len[i]);
 *                   in_off[i] must be a multiple of 16.  d_in may be NULL, in which case in_off[i] are absolute
 *                   device addresses (blocks scattered over several buffers); every block must be followed by
 *                   at least 3 readable bytes (the coders fetch whole 4-byte words)
 *   in_off/in_len/order   HOST arrays (the Java host knows every block length)
 *   d_out           device buffer of out_cap bytes; compressed/decoded blocks are packed into it
 *                   back to back, each starting at a multiple of 16
 *   d_out_off       DEVICE array of nblocks+1 u64, written by the call: block i occupies
 *                   d_out[d_out_off[i] .. +d_out_len[i]); d_out_off[nblocks] = bytes used
 *   d_out_len       DEVICE array of nblocks u32, written by the call
 *   d_status        DEVICE array of nblocks i32, written by the call (cramgpu_status per block)
 *   stream          cudaStream_t to enqueue on (NULL = the context's own stream)
 * Work is enqueued asynchronously; results are valid once `stream` has been synchronised.  (The ENCODER synchronises `stream`
 * once near its start: the alphabets of the order-1 blocks are found on the device, and the frequency tables, coder cells and
 * work areas are sized by what the blocks need -- 1 GB for 16 384 blocks where the worst case would be 12.6 GB.  The decoder,
 * and everything the encoder enqueues after that point, is asynchronous.)
 * The context's scratch is reused by the next call on the same ctx, so calls on one ctx must be
 * stream-ordered (same stream, or synchronised in between).
 * For decode, out_cap must be >= sum over blocks of roundup16(raw size); sizes come from the prefixes,
 * which the call reads from the HOST copy `in_prefix` (9 bytes per block, row-major nblocks x 9)
 * so it can plan launches without a device round trip.
 */
int cramgpu_rans_encode_dev(cramgpu_ctx *ctx, uint32_t nblocks,
                            const uint8_t *d_in, const uint64_t *in_off, const uint32_t *in_len,
                            const uint8_t *order,
                            uint8_t *d_out, uint64_t out_cap, uint64_t *d_out_off, uint32_t *d_out_len,
                            int32_t *d_status, void *stream);

int cramgpu_rans_decode_dev(cramgpu_ctx *ctx, uint32_t nblocks,
                            const uint8_t *d_in, const uint64_t *in_off, const uint32_t *in_len,
                            const uint8_t *in_prefix,
                            uint8_t *d_out, uint64_t out_cap, uint64_t *d_out_off, uint32_t *d_out_len,
                            int32_t *d_status, void *stream);

/*
 * Order-0 rANS compression of ONE block without any allocation.  This entry point has no counterpart in the
 * reference (Java callers never manage device memory).  It exists for hosts that are not allowed to obtain
 * memory from the CUDA runtime while they run -- a JVM-side pool or a serving process may have handed all of HBM
 * to its own memory manager, so that a cudaMalloc fails or upsets its book-keeping, and latency-critical callers
 * cannot take the device-wide synchronisation that cudaMalloc / cudaFree bring.  The result is byte-identical to
 * cramgpu_rans_encode_dev with order 0.
 *
 *   d_in, n          device pointer (16-byte aligned, at least 3 readable bytes after the block) and raw size;
 *                    n may not exceed 1 MiB (1 048 576): larger inputs are refused with CRAMGPU_ERR_ARG and a
 *                    message from cramgpu_last_error(NULL)
 *   d_out, out_cap   device destination; cramgpu_rans_bound(n) always suffices
 *   d_out_len        DEVICE u32, receives the compressed size
 *   d_status         DEVICE i32, receives the cramgpu_status of the block
 *   stream           cudaStream_t to enqueue on (NULL = the legacy default stream); asynchronous
 *
 * All scratch is a file-scope device array of about 1.26 MiB that every GPU holds from the moment the library is
 * loaded until the process ends; no cudaMalloc, cudaFree or other allocating routine is called.  The first call on
 * a device resolves the array's device address and raises the coder kernel's shared-memory limit: issue that
 * first call before any stream capture starts.  Every call reuses the same array and finds the previous call's
 * data in it; the counters that have to start from zero are cleared on `stream` at the beginning of each call.
 * Calls whose GPU work can overlap (different streams without an ordering between them) would race on the array
 * and must be ordered by the caller.
 */
int cramgpu_rans0_encode_static(const uint8_t *d_in, uint32_t n, uint8_t *d_out, uint32_t out_cap,
                                uint32_t *d_out_len, int32_t *d_status, void *stream);


/* ------------------------------------------------------------------ record-feature work
 *
 * Encode side  (replaces the per-record loop of CramSerilization.convert,
 *               src/main/java/htsjdk/samtools/cram/CramSerilization.java:144-166):
 *   Sam2CramRecordFactory.createCramRecord / createVariations / addSubstitutionsAndMaskedBases
 *                                                       [Sam2CramRecordFactory.java:108-335]
 *   QualityScorePreservation.addQualityScores           lossy/QualityScorePreservation.java:276-301,332-489
 *   CompressionHeaderFactory.buildFrequencies, SubstitutionMatrix(long[][]), updateSubstitutionCodes
 *                                                       [CompressionHeaderFactory.java:171-209]
 *   Writer.write for the series fed by the feature work [Writer.java:180-258]
 * Decode side  (replaces CramNormalizer.restoreBases / restoreQualityScores,
 *               src/main/java/htsjdk/samtools/cram/build/CramNormalizer.java:208-244,287-434).
 *
 * Records travel as structure-of-arrays over one or more containers (a container = the <= 10 000 records
 * CRAMContainerStreamWriter flushes together, all on one reference sequence).  The Java shim flattens
 * List<SAMRecord> into this layout once per call.
 */

/* one parsed lossy-quality policy (lossy/PreservationPolicy + ReadCategory + BaseCategory + QualityScoreTreatment) */
typedef struct cramgpu_qs_policy {
    int32_t read_cat;      /* 0 none, 1 ALL (*), 2 UNPLACED (U), 3 LOWER_MAPPING_SCORE (m), 4 HIGHER_MAPPING_SCORE (M) */
    int32_t read_param;
    int32_t n_base_cats;   /* base categories in spec order (they overwrite one another exactly as in the reference) */
    int32_t base_cat[8];   /* 1 MATCH (R), 2 MISMATCH (N), 3 FLANKING_DELETION (D), 4 INSERTION (I); 5 LOWER_COVERAGE (X) and
                              6 PILEUP (P) with their parameter in bits 8..31 -- these need cramgpu_tracks_update_dev first */
    int32_t treatment;     /* 0 PRESERVE, 1 BIN (Illumina 8-binning), 2 DROP */
    int32_t treat_param;
} cramgpu_qs_policy;

/*
 * Replaces QualityScorePreservation.parsePolicies(String) (lossy/QualityScorePreservation.java:167-256):
 * "-"-separated policies, sorted DROP, BIN, PRESERVE (stable).  Host only.
 * Returns the number of policies written, or CRAMGPU_ERR_ARG for "Unknown read or base category".  The X<n> / P<n>
 * selectors (coverage below n; more than n mismatches piled up) read ReferenceTracks counters that carry state across
 * containers: see cramgpu_tracks_update_dev.
 */
int cramgpu_parse_lossy_spec(const char *spec, cramgpu_qs_policy *out, int cap);

/* series produced by cramgpu_features_encode; values of the index into stream_off / stream_len */
enum {
    CRAMGPU_S_FN = 0, /* number of read features, ITF8                       content id 15 */
    CRAMGPU_S_FC,     /* feature operator byte                                content id 3  */
    CRAMGPU_S_FP,     /* feature position delta, ITF8                         content id 2  */
    CRAMGPU_S_BS,     /* substitution code byte                               content id 16 */
    CRAMGPU_S_DL,     /* deletion length, ITF8                                content id 5  */
    CRAMGPU_S_BA,     /* bases: InsertBase features and unmapped reads        content id 6  */
    CRAMGPU_S_SC,     /* soft clip bases, TAB terminated                      content id 27 */
    CRAMGPU_S_QS,     /* quality scores: Q features and preserved arrays      content id 4  */
    CRAMGPU_S_RS,     /* reference skip length, ITF8                          content id 26 */
    CRAMGPU_S_HC,     /* hard clip length, ITF8                               content id 28 */
    CRAMGPU_S_PD,     /* padding length, ITF8                                 content id 29 */
    CRAMGPU_S_MQ,     /* mapping quality, ITF8                                content id 11 */
    CRAMGPU_S_COUNT
};

#define CRAMGPU_RF_NO_QUALS 1u   /* rflags bit: SAMRecord.NULL_QUALS ('*') */
#define CRAMGPU_RF_NO_BASES 2u   /* rflags bit: SAMRecord.NULL_SEQUENCE ('*'); read_off gives length 0 */

/* input records, structure of arrays.  cont_off, ref_off, ref_len are always HOST arrays. */
typedef struct cramgpu_reads {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;  /* [n_containers+1] container c holds reads cont_off[c] .. cont_off[c+1]-1 */
    const uint64_t *ref_off;   /* [n_containers] start of the container's reference sequence inside `ref` */
    const uint64_t *ref_len;   /* [n_containers] its length (bases beyond it read as 'N', [:306]) */
    const uint8_t *ref;        /* reference bases as ReferenceSource hands them out (upper-cased or not) */
    uint64_t ref_total;        /* bytes in `ref` */
    const uint8_t *bases;      /* SAMRecord.getReadBases(), concatenated */
    const uint8_t *quals;      /* SAMRecord.getBaseQualities(), same offsets as bases */
    const uint64_t *read_off;  /* [n_reads+1] */
    const uint32_t *cigar;     /* BAM-packed: length << 4 | op, ops MIDNSHP=X */
    const uint64_t *cigar_off; /* [n_reads+1] */
    const int32_t *pos;        /* SAMRecord.getAlignmentStart(), 1-based, 0 = none */
    const uint16_t *flag;      /* SAM flags (0x4 = read unmapped) */
    const uint8_t *mapq;
    const uint8_t *rflags;     /* CRAMGPU_RF_* per read; NULL = all 0 */
} cramgpu_reads;

/*
 * Result of feature extraction.  Every pointer refers to memory owned by the context (device memory for
 * cramgpu_features_encode_dev, host memory for cramgpu_features_encode) and stays valid until the next
 * record call on the same context.
 *   features: f_op (operator byte as in ReadFeature.getOperator()), f_pos (1-based position in read),
 *     'X': f_val = substitution code (after ranking), f_aux = read base | reference base << 8
 *     'i': f_val = base      'D','N','H','P': f_val = length      'Q': f_val = score
 *     'S': f_val = offset into payload, f_aux = length
 *   Q features are already merged in by position, after the other features of the same position, exactly
 *   as Collections.sort leaves them (QualityScorePreservation.java:288-299).
 */
typedef struct cramgpu_features {
    uint64_t n_feat;
    const uint8_t *f_op;
    const int32_t *f_pos, *f_val, *f_aux;
    const uint64_t *feat_off;     /* [n_reads+1] */
    const uint8_t *payload;
    uint64_t payload_len;
    const uint8_t *force_qs;      /* [n_reads] bit0 = CramCompressionRecord.isForcePreserveQualityScores(), bit1 = has a score array,
                                     bit7 = the reference throws for this record (see CRAMGPU_ERR_UNSUPPORTED below) */
    const uint8_t *matrix;        /* [n_containers][8]: 5 bytes of SubstitutionMatrix.getEncodedMatrix(), 3 pad */
    const uint8_t *streams;       /* all series bytes, SERIES-major: the FN regions of all containers, then FC, ...; every
                                     (container, series) region starts at a multiple of 16 */
    uint64_t streams_len;
    const uint64_t *stream_off;   /* [n_containers][CRAMGPU_S_COUNT] */
    const uint64_t *stream_len;   /* [n_containers][CRAMGPU_S_COUNT] */
    const uint64_t *read_qs_off;  /* [n_reads] offset in `streams` of the read's preserved score array, ~0 = none */
    const uint64_t *read_ba_off;  /* [n_reads] offset in `streams` of an unmapped read's bases, ~0 = none */
} cramgpu_features;

/*
 * Feature extraction + quality preservation + series packing for a batch of containers.
 *   n_policies < 0          lossless (CramSerilization.java:160-161: preserve iff quals != NULL_QUALS)
 *   n_policies >= 0         the parsed lossy model; 0 policies = drop all
 * Host variant: `reads` holds host pointers; results are host pointers.
 * Device variant: the big arrays of `reads` (ref, bases, quals, read_off, cigar, cigar_off, pos, flag, mapq,
 * rflags) are device pointers; results are device pointers, except that the small per-container tables
 * stream_off / stream_len and the scalars are also mirrored on the host (h_stream_off, h_stream_len) so that
 * the caller can hand the series straight to cramgpu_rans_encode_dev.  The call returns after the device
 * has finished (it needs the totals).
 * In the host variants the bulk arrays of `reads` (ref, bases, quals, cigar, pos, flag, mapq, rflags) may also be
 * device pointers -- e.g. a reference that stays resident across calls -- and are then used where they are;
 * read_off and cigar_off are read on the host and must be host memory.
 *
 * A BIN treatment ('*8', 'R8', ...) whose lookup meets a score of 110..255: the reference throws ArrayIndexOutOfBoundsException
 * out of Binning.Illumina_binning_matrix (110 entries, signed byte index; lossy/Binning.java:63-71,
 * QualityScorePreservation.java:258-266).  The call then returns CRAMGPU_ERR_UNSUPPORTED after filling `out` as usual:
 * bit 7 of force_qs marks the records concerned (their scores are binned as 40), the other records are unaffected.
 *
 * cramgpu_features_encode_resident: the host variant for callers that code the big series on the GPU right away.
 * Results stay in device memory (`dev_out`, device pointers, valid until the next record call on ctx) and come
 * back to the host (`out`) as usual, except that of `streams` only the series whose bit is set in `host_series`
 * (1u << CRAMGPU_S_x) are copied; `out->streams` keeps the device layout, the other series' bytes are left
 * unwritten.  Typical use: host_series = all but QS and BA, then cramgpu_rans_encode_batch with
 * in[i] = dev_out->streams + stream_off[c][QS].  With CRAMGPU_FEATURES_ON_DEVICE_ONLY or-ed into host_series the feature arrays
 * (f_op f_pos f_val f_aux feat_off payload read_qs_off read_ba_off) are not copied to the host either (`out` holds NULL for
 * them): a host that hands the series straight to the block layer never looks at them.  No counterpart in the reference.
 */
#define CRAMGPU_FEATURES_ON_DEVICE_ONLY 0x80000000u
int cramgpu_features_encode(cramgpu_ctx *ctx, const cramgpu_reads *reads, const cramgpu_qs_policy *policies,
                            int n_policies, cramgpu_features *out);
int cramgpu_features_encode_resident(cramgpu_ctx *ctx, const cramgpu_reads *reads, const cramgpu_qs_policy *policies,
                                     int n_policies, uint32_t host_series, cramgpu_features *out,
                                     cramgpu_features *dev_out);
int cramgpu_features_encode_dev(cramgpu_ctx *ctx, const cramgpu_reads *reads, const cramgpu_qs_policy *policies,
                                int n_policies, cramgpu_features *out, uint64_t *h_stream_off, uint64_t *h_stream_len,
                                void *stream);

/*
 * Series -> read features: the feature part of CramRecordReader.read
 * (htsjdk/samtools/cram/encoding/reader/CramRecordReader.java:59-..., mirror of Writer.write [Writer.java:180-258])
 * for the series listed above, so that decoded series can go to cramgpu_normalize_decode without a trip through
 * CramCompressionRecord objects.  Per mapped record: FN; per feature FC, FP (delta) and the
 * operator's value (X: BS, D: DL, N: RS, H: HC, P: PD, i: BA, Q: QS, S: SC bytes up to the TAB); MQ; then the score
 * array in QS when the record carries one.  Unmapped: readLength bytes of BA unless the bases are unknown, then the
 * score array.  What the host still parses (BF/CF/RL/AP...) arrives as read_off (read lengths) and rflags.
 */
typedef struct cramgpu_series {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    const uint8_t *streams;       /* decoded series bytes of all containers; device variant: NULL = stream_off holds absolute
                                     device addresses (series decoded into different buffers), results' offsets then too */
    uint64_t streams_len;
    const uint64_t *stream_off;   /* HOST [n_containers][CRAMGPU_S_COUNT] */
    const uint64_t *stream_len;   /* HOST [n_containers][CRAMGPU_S_COUNT] */
    const uint64_t *read_off;     /* [n_reads+1] cumulative readLength */
    const uint8_t *rflags;        /* [n_reads] CRAMGPU_NF_* */
} cramgpu_series;

/*
 * Result: feature arrays as cramgpu_features_encode returns them (f_aux of an X feature is 0: the bases are not in the
 * series), payload = soft-clip bytes, read_qs_off / read_ba_off = offsets into `streams`, plus the mapping qualities.
 * Not set: force_qs, matrix, stream_off / stream_len.  status[c] (HOST, optional) = CRAMGPU_ERR_CORRUPT when a series
 * of container c ends before its records do, CRAMGPU_ERR_UNSUPPORTED for an operator outside the list; the call
 * returns the first such status.  Host variant: streams / read_off / rflags may be host or device memory, results are
 * host memory owned by the context (mapq: caller's array, optional).  Device variant: those three are device
 * pointers, results are device memory owned by the context (*d_mapq receives the mapq array), valid until the next
 * unpack call on ctx (an encode result of the same context stays valid: it may be the input).
 */
int cramgpu_features_unpack(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *out, uint8_t *mapq,
                            int32_t *status);
int cramgpu_features_unpack_dev(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *out, uint8_t **d_mapq,
                                int32_t *status, void *stream);
/* cramgpu_features_unpack_resident: the host variant for a decode pipeline that never looks at the features itself.  in->streams
 * must be NULL and stream_off holds ABSOLUTE addresses: a series the host inflated (gzip) lies in host memory and is staged,
 * a series the block codec left on the device (QS, BA out of cramgpu_rans_decode_batch) is used where it is.  The features
 * stay on the device (dev_out: device pointers, read_qs_off / read_ba_off absolute; valid until the next unpack call on ctx)
 * and go to cramgpu_normalize_decode as they are -- with qs = ba = NULL there.  mapq (optional) and status come back to the host. */
int cramgpu_features_unpack_resident(cramgpu_ctx *ctx, const cramgpu_series *in, cramgpu_features *dev_out, uint8_t *mapq,
                                     int32_t *status);

/* input of the decode-side normaliser: records as CramRecordReader leaves them, structure of arrays */
typedef struct cramgpu_norm_in {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    const uint64_t *win_off;      /* HOST [n_containers] start of the container's reference region inside `ref` */
    const uint64_t *win_len;      /* HOST [n_containers] region.array.length */
    const int64_t *win_start0;    /* HOST [n_containers] refOffsetZeroBased = region.alignmentStart - 1 */
    const uint8_t *ref;           /* region bytes of all containers */
    uint64_t ref_total;
    const uint8_t *matrix;        /* [n_containers][8] substitution matrix bytes */
    const int32_t *pos;           /* [n_reads] alignmentStart */
    const uint64_t *read_off;     /* [n_reads+1] output offsets = cumulative readLength */
    const uint8_t *rflags;        /* [n_reads] bit0 forcePreserveQualityScores, bit1 segment unmapped, bit2 unknownBases */
    const uint64_t *feat_off;     /* [n_reads+1] */
    const uint8_t *f_op;
    const int32_t *f_pos, *f_val, *f_aux;
    const uint8_t *payload;
    const uint8_t *qs;            /* decoded QS series (or any buffer holding the preserved arrays) */
    const uint64_t *read_qs_off;  /* [n_reads] offset of the read's preserved array in qs; ~0 = none */
    const uint8_t *ba;            /* decoded BA series */
    const uint64_t *read_ba_off;  /* [n_reads] offset of an unmapped read's bases in ba; ~0 = none */
    uint64_t payload_len;         /* bytes in payload; an S / I feature that points outside it makes its record CRAMGPU_ERR_CORRUPT */
} cramgpu_norm_in;

#define CRAMGPU_NF_FORCE_QS 1u
#define CRAMGPU_NF_UNMAPPED 2u
#define CRAMGPU_NF_UNKNOWN_BASES 4u

/*
 * Base and quality restoration.  bases / quals receive read_off[n_reads] bytes; qual_star[r] = 1 where the
 * reference yields SAMRecord.NULL_QUALS; status[r] = CRAMGPU_ERR_CORRUPT where the reference would throw
 * ArrayIndexOutOfBoundsException (record outside its region, a feature position outside [1, readLength], a
 * soft clip or insertion that overruns the read or the payload); such a record's own bases / scores are then
 * undefined, every other record is unaffected.  Host variant takes/returns host memory; the
 * device variant takes device pointers for everything not marked HOST and enqueues on `stream`.
 * In the host variant ref, matrix, the feature arrays, payload, qs, ba and the outputs bases / quals may also be
 * device pointers (a QS series cramgpu_rans_decode_batch left on the device, say) and are then used in place;
 * the per-record tables (pos, read_off, rflags, feat_off, read_qs_off, read_ba_off) and qual_star / status are
 * host memory.
 */
int cramgpu_normalize_decode(cramgpu_ctx *ctx, const cramgpu_norm_in *in, uint8_t *bases, uint8_t *quals,
                             uint8_t *qual_star, int32_t *status);
int cramgpu_normalize_decode_dev(cramgpu_ctx *ctx, const cramgpu_norm_in *in, uint8_t *d_bases, uint8_t *d_quals,
                                 uint8_t *d_qual_star, int32_t *d_status, void *stream);


/* ------------------------------------------------------------------ the record fields in front of the read features
 *
 * Replaces the first half of Writer.write [Writer.java:139-172], driven per record from ContainerFactory.buildSlice
 * [ContainerFactory.java:179-184], and the flag part of Sam2CramRecordFactory.createCramRecord [:114-131,143-152]:
 * the series BF CF RI RL AP RG RN MF NS NP TS NF TL of every container of the batch, byte-identical to what the
 * reference's Writer produces (tests/golden/streams_jvm.json).  What stays on the Java host is what needs the whole
 * container at once: mate linking (CramSerilization.setMateInfo, :274-373 -- it decides `detached`, `mate downstream`
 * and recordsToNextFragment) and the tag dictionary.
 */
enum {
    CRAMGPU_HS_BF = 0, /* CRAM bit flags, ITF8                                   content id 0  */
    CRAMGPU_HS_CF,     /* compression bit flags, one byte                        content id 21 */
    CRAMGPU_HS_RI,     /* reference id, ITF8 (multi-reference slices only)       content id 25 */
    CRAMGPU_HS_RL,     /* read length, ITF8                                      content id 9  */
    CRAMGPU_HS_AP,     /* alignment start or its delta, ITF8                     content id 1  */
    CRAMGPU_HS_RG,     /* read group index, ITF8                                 content id 10 */
    CRAMGPU_HS_RN,     /* read name + TAB                                        content id 12 */
    CRAMGPU_HS_MF,     /* mate flags, one byte                                   content id 19 */
    CRAMGPU_HS_NS,     /* mate reference id, ITF8                                content id 20 */
    CRAMGPU_HS_NP,     /* mate alignment start, ITF8                             content id 13 */
    CRAMGPU_HS_TS,     /* template size, ITF8                                    content id 14 */
    CRAMGPU_HS_NF,     /* records to the next fragment, ITF8                     content id 8  */
    CRAMGPU_HS_TL,     /* tag id list index, ITF8                                content id 24 */
    CRAMGPU_HS_COUNT
};

typedef struct cramgpu_host_fields {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    const int32_t *slice_start;   /* HOST [n_containers] Slice.alignmentStart: AP of the first record is its start minus this (ap_delta only) */
    const uint8_t *multi_ref;     /* HOST [n_containers] 1 = Slice.MULTI_REFERENCE: RI is written; NULL = none */
    uint32_t ap_delta;            /* CompressionHeader.APDelta (coordinate-sorted input) */
    uint32_t names_included;      /* CompressionHeader.readNamesIncluded: names of all records, else of detached ones only */
    const uint16_t *flag;         /* [n_reads] SAM flags: BF and MF are derived from them */
    const uint8_t *cf;            /* [n_reads] CF as CramCompressionRecord.getCompressionFlags() returns it, or NULL: then composed from */
    const uint8_t *mate_cf;       /*   the mate-linking bits (2 detached, 4 mate downstream; NULL = every record detached), */
    const uint8_t *force_qs;      /*   bit 0 of cramgpu_features.force_qs (NULL = 0) and */
    const uint8_t *rflags;        /*   CRAMGPU_RF_NO_BASES of the reads' rflags (NULL = 0) */
    const int32_t *ref_id;        /* [n_reads] sequenceId (needed for multi-reference slices) */
    const int32_t *pos;           /* [n_reads] alignmentStart */
    const uint64_t *read_off;     /* [n_reads+1] cumulative read length */
    const int32_t *read_group;    /* [n_reads] readGroupID; NULL = -1 everywhere */
    const uint8_t *names;         /* read names, concatenated (no terminators); NULL = RN stays empty */
    const uint64_t *name_off;     /* [n_reads+1] */
    const int32_t *mate_ref_id;   /* [n_reads] mateSequenceID      ) read for detached records only (NULL: -1, 0, 0) */
    const int32_t *mate_pos;      /* [n_reads] mateAlignmentStart  )                                                 */
    const int32_t *tlen;          /* [n_reads] templateSize        )                                                 */
    const int32_t *nf;            /* [n_reads] recordsToNextFragment (records with their mate downstream) */
    const int32_t *tl;            /* [n_reads] tagIdsIndex; NULL = 0 */
} cramgpu_host_fields;

/* Result: all series bytes SERIES-major like cramgpu_features.streams (the BF regions of all containers, then CF, ...),
 * every (container, series) region at a multiple of 16.  Memory belongs to the context and stays valid until the next
 * call of the same kind on it. */
typedef struct cramgpu_host_series {
    const uint8_t *streams;
    uint64_t streams_len;
    const uint64_t *stream_off;   /* [n_containers][CRAMGPU_HS_COUNT] */
    const uint64_t *stream_len;   /* [n_containers][CRAMGPU_HS_COUNT] */
} cramgpu_host_series;

/* Device variant: the per-record arrays of `in` are device pointers, results are device pointers; the layout tables are
 * mirrored into h_stream_off / h_stream_len ([n_containers][CRAMGPU_HS_COUNT], optional) so that the caller can hand the
 * series to cramgpu_rans_encode_dev.  The call returns after the counting pass (it sizes the result buffer from the
 * totals); the packing kernel is still running on `stream` then.
 * Host variant: everything in host memory; results are host memory owned by the context. */
int cramgpu_host_series_pack_dev(cramgpu_ctx *ctx, const cramgpu_host_fields *in, cramgpu_host_series *out,
                                 uint64_t *h_stream_off, uint64_t *h_stream_len, void *stream);
int cramgpu_host_series_pack(cramgpu_ctx *ctx, const cramgpu_host_fields *in, cramgpu_host_series *out);
/* The host variant for callers that code the rANS series on the GPU right away (the counterpart of
 * cramgpu_features_encode_resident): the result stays in device memory (`dev_out`: device pointers, valid until the next
 * call of this kind on ctx) and only the series whose bit is set in `host_series` (1u << CRAMGPU_HS_x) are copied to the
 * host; `out->streams` keeps the device layout.  Per-record arrays that already are device pointers are used in place. */
int cramgpu_host_series_pack_resident(cramgpu_ctx *ctx, const cramgpu_host_fields *in, uint32_t host_series,
                                      cramgpu_host_series *out, cramgpu_host_series *dev_out);

/* ------------------------------------------------------------------ slice header fields and content digests
 *
 * What ContainerFactory.buildSlice derives from a slice's records before it writes a byte
 * [htsjdk/samtools/cram/build/ContainerFactory.java:138-176]: nofRecords, bases (sum of readLength), sequenceId (the first
 * record's; Slice.MULTI_REFERENCE = -2 once a record differs), alignmentStart / alignmentSpan (min start over the placed
 * records, max CramCompressionRecord.getAlignmentEnd() [CramCompressionRecord.java:160-194] -- span = readLength - #'i' -
 * len('S') + len('D'), reference skips NOT counted, 0 for a record with the unmapped flag; -1 / 0 for a multi-reference
 * slice or one without a placed record; a negative span becomes 0, CramSerilization.java:184-186) and the content digests
 * ContentDigests.ALL [cram/digest/ContentDigests.java:26-83]:
 *   BD / SD   CRC32 (java.util.zip.CRC32) of every record's readBases / qualityScores array, summed as Java ints
 *             (Crc32Hasher + IntegerSumCombine); the tag value is the sum, 4 bytes little endian
 *   B5 / S5 / B1 / S1   SHA-512 / SHA-1 "byte-wise sums": MessageDigestHasher.asByteArray() returns messageDigest.digest() of
 *             a digest that getValue() has just finished (and so reset), not the combined value: in htsjdk 2.4.1 these
 *             four tags are the digest of the EMPTY message whatever the records hold.  cramgpu_slice_sha_tags hands
 *             out those constants; there is nothing to compute.
 * One slice per container, as the writer builds them (CRAMContainerStreamWriter: one slice per container by default).
 *
 * Every per-record pointer may be host or device memory in the host variant (device memory is used where it is);
 * cont_off is always a HOST array.
 *   scores != NULL   record.qualityScores, laid out like bases (0xFF = -1): the qualities themselves when lossless
 *   scores == NULL   take them from a cramgpu_features_encode[_dev] result: the score array in `streams` at read_qs_off[r]
 *                    when force_qs[r] bit 0 is set, else 0xFF everywhere except at the record's 'Q' features
 *                    (f_pos, f_val); needs force_qs, read_qs_off, streams, f_pos.  (One case the features cannot express: a
 *                    record with the unmapped flag AND an alignment start under a per-base quality model keeps sparse
 *                    scores in memory that Writer.write never emits; pass `scores` when such records can occur.)
 *   rflags bit CRAMGPU_RF_NO_QUALS: qualityScores = SAMRecord.NULL_QUALS, an empty array (CRC32 = 0)
 */
typedef struct cramgpu_slice_in {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    const uint8_t *bases;         /* CramCompressionRecord.readBases, concatenated */
    const uint8_t *scores;        /* see above; may be NULL */
    const uint64_t *read_off;     /* [n_reads+1] */
    const uint8_t *rflags;        /* CRAMGPU_RF_* per record; NULL = all 0 */
    const int32_t *ref_id;        /* CramCompressionRecord.sequenceId (-1 = unmapped); NULL = all 0 */
    const int32_t *pos;           /* alignmentStart, 0 = none */
    const uint16_t *flag;         /* SAM flags */
    const uint8_t *f_op;          /* read features, as cramgpu_features */
    const int32_t *f_pos, *f_val, *f_aux;
    const uint64_t *feat_off;     /* [n_reads+1] */
    const uint8_t *force_qs;      /* scores == NULL only */
    const uint64_t *read_qs_off;  /* scores == NULL only */
    const uint8_t *streams;       /* scores == NULL only */
    /* slice reference MD5 (optional: ref == NULL leaves ref_md5 zero): the reference as in cramgpu_reads */
    const uint8_t *ref;           /* reference bases, raw (as ReferenceSource hands them out); host or device in the host variant */
    uint64_t ref_total;           /* bytes in `ref` */
    const uint64_t *ref_off;      /* HOST [n_containers] start of the slice's reference sequence inside `ref` */
    const uint64_t *ref_len;      /* HOST [n_containers] its length */
} cramgpu_slice_in;

typedef struct cramgpu_slice_header {
    int32_t sequence_id;          /* Slice.sequenceId */
    int32_t alignment_start;      /* Slice.alignmentStart */
    int32_t alignment_span;       /* Slice.alignmentSpan */
    uint32_t n_records;           /* Slice.nofRecords */
    uint64_t bases;               /* Slice.bases */
    uint32_t bd, sd;              /* tags BD and SD */
    uint8_t ref_md5[16];          /* Slice.refMD5 as CramSerilization.convert sets it (CramSerilization.java:184-197, Slice.setRefMD5
                                     [Slice.java:145-159]): MD5 of the reference bases [alignmentStart - 1, + min(span, length -
                                     alignmentStart + 1)); 16 zero bytes for a multi-reference or unplaced slice, or without `ref` */
} cramgpu_slice_header;

/* out: HOST array of n_containers headers.  _dev: every pointer but cont_off is device memory, d_out is a DEVICE array,
 * the call is asynchronous on `stream` (NULL = the context's stream). */
int cramgpu_slice_headers(cramgpu_ctx *ctx, const cramgpu_slice_in *in, cramgpu_slice_header *out);
int cramgpu_slice_headers_dev(cramgpu_ctx *ctx, const cramgpu_slice_in *in, cramgpu_slice_header *d_out, void *stream);
/* the four constant tags: sha1 = B1 = S1 (20 bytes), sha512 = B5 = S5 (64 bytes) */
void cramgpu_slice_sha_tags(uint8_t sha1[20], uint8_t sha512[64]);

/* ------------------------------------------------------------------ coverage / mismatch tracks: the lossy selectors X and P
 *
 * The reference keeps one ReferenceTracks per sequence (src/main/java/htsjdk/samtools/cram/ref/ReferenceTracks.java) across
 * the containers of a file: CramSerilization.convert [CramSerilization.java:112-131] moves its window to the container's span
 * (getSpans :60-86, ensureRange) and counts the container's records into it (updateTracks :225-257: coverage of every
 * reference position an alignment covers, mismatches of the bases of M / = / X operators, with that loop's indices as written)
 * before QualityScorePreservation.addQS asks coverageAt(pos) < n (X<n>, lossy/QualityScorePreservation.java:447-465) and
 * mismatchesAt(alignment start + i) > n (P<n>, :466-470) for every record.
 *
 * cramgpu_tracks_create      the tracks of one sequence; ref = its raw bases (HOST), window = 0 for the reference's 1 000 000.
 * cramgpu_tracks_update_dev  everything above for a batch of containers of that sequence (reads: DEVICE arrays as
 *     cramgpu_features_encode_dev takes them, cont_off HOST; records with pos = 0 are unplaced; ref / ref_off / ref_len unused):
 *     container c sees the counts of containers 0..c of this call and of all earlier calls.  It leaves one byte per base in the
 *     context; the NEXT cramgpu_features_encode* call on the context with the same reads and the same policies consumes them
 *     (without them a model with X or P fails with "The quality score lossy model requires reference tracks.", :121).  A model
 *     without X and P leaves the tracks untouched, as the reference does.  container_status (HOST [n_containers], may be NULL):
 *     CRAMGPU_ERR_ARG where the reference throws ("Cannot move the window backwards", "New position is beyond the reference"),
 *     CRAMGPU_ERR_UNSUPPORTED for a container that reaches beyond the window after the window was moved to its start (span
 *     above the window: the reference then moves the window between records -- not modelled) or that holds a record without
 *     bases; the first such container fails the call (nothing is left for the feature call) and the tracks stay unusable,
 *     as after the reference's exception.  One stream synchronisation inside.
 * cramgpu_tracks_position    ReferenceTracks.getWindowPosition() after the last call.
 */
typedef struct cramgpu_tracks cramgpu_tracks;
int cramgpu_tracks_create(cramgpu_ctx *ctx, const uint8_t *ref, uint64_t ref_len, uint32_t window, cramgpu_tracks **out);
void cramgpu_tracks_destroy(cramgpu_ctx *ctx, cramgpu_tracks *t);
int64_t cramgpu_tracks_position(const cramgpu_tracks *t);
int cramgpu_tracks_update_dev(cramgpu_ctx *ctx, cramgpu_tracks *t, const cramgpu_reads *reads, const cramgpu_qs_policy *pol, int n_pol,
                              int32_t *container_status, void *stream);

/* ------------------------------------------------------------------ mate linking
 *
 * CramSerilization.setMateInfo / detach (src/main/java/htsjdk/samtools/cram/CramSerilization.java:258-373): which records of a
 * coordinate-sorted container are written as linked mates -- CF bit "mate downstream" (value 4) on the first, the NF series =
 * recordsToNextFragment -- and which are detached (CF value 2: MF, NS, NP, TS written in full).  A multi-fragment record is
 * linked to the next record of the same name and class (primary / secondary) when exactly two such records are in the
 * container, their first/last-of-pair flags fit, both template sizes equal +-CramNormalizer.computeInsertSize
 * [CramNormalizer.java:455-470] and each record's mate fields describe the other [:318-349]; three or more records of one
 * name are all detached, as the reference's map walk leaves them; so is everything when the header is not coordinate sorted
 * [:175-179].  The results are the mate_cf / nf inputs of cramgpu_host_series_pack.
 * Record fields as Sam2CramRecordFactory.createCramRecord sets them: flag = SAM flags; mate_ref_id / mate_pos / tlen =
 * SAMRecord.getMateReferenceIndex / getMateAlignmentStart / getInferredInsertSize; the read features give
 * CramCompressionRecord.getAlignmentEnd [CramCompressionRecord.java:160-194] for reverse-strand records.
 * cont_off: HOST.  Host variant: every other pointer may be host or device memory; mate_cf / nf are HOST arrays [n_reads].
 */
typedef struct cramgpu_mate_in {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    uint32_t coordinate_sorted;   /* SAMFileHeader.getSortOrder() == coordinate */
    uint32_t reserved;
    const uint16_t *flag;
    const uint8_t *names;         /* read names, concatenated */
    const uint64_t *name_off;     /* [n_reads+1] */
    const int32_t *ref_id;        /* NULL = all 0 */
    const int32_t *pos, *mate_ref_id, *mate_pos, *tlen;
    const uint64_t *read_off;     /* [n_reads+1] */
    const uint8_t *f_op;          /* read features, as cramgpu_features */
    const int32_t *f_val, *f_aux;
    const uint64_t *feat_off;     /* [n_reads+1] */
} cramgpu_mate_in;
int cramgpu_mate_link(cramgpu_ctx *ctx, const cramgpu_mate_in *in, uint8_t *mate_cf, int32_t *nf);
int cramgpu_mate_link_dev(cramgpu_ctx *ctx, const cramgpu_mate_in *in, uint8_t *d_mate_cf, int32_t *d_nf, void *stream);

/* ------------------------------------------------------------------ decode: restore the pairing (inverse of cramgpu_mate_link)
 *
 * The pairing part of CramNormalizer.normalize (src/main/java/htsjdk/samtools/cram/build/CramNormalizer.java:67-102, 236-270) on
 * the records as the reader left them (cramgpu_features_unpack*): an unplaced record (ref_id -1) gets alignment start 0; a record
 * that is not multi-fragment (flag 0x1) or is detached (cf & 2) gets nf = -1 and no links; a linked record with a mate downstream
 * (cf & 4) points to record index + 1 + nf of its container (beyond the container: d_status[c] = CRAMGPU_ERR_CORRUPT, records.get()
 * throws); then along every chain that starts at a record nobody points to: setNextMate (mate start, mate sequence, and in the
 * record's mate flags -- the MF byte, CramCompressionRecord.mateFlags [:270-283] -- bit 2 mate unmapped and bit 1 mate reverse from
 * the next record; the last record takes them from the first) and the template size of the first and last record = +-computeInsertSize [:455-470] (getAlignmentEnd from the read features).  Detached records keep
 * the mate fields the reader gave them.  All arrays DEVICE (cont_off HOST), nf / pos / mate_* / tlen are updated in place;
 * asynchronous on `stream`.
 */
typedef struct cramgpu_mate_restore {
    uint64_t n_reads;
    uint32_t n_containers;
    const uint64_t *cont_off;     /* HOST [n_containers+1] */
    const uint16_t *flag;         /* SAM flags (BF series) */
    const uint8_t *cf;            /* compression flags (CF series) */
    uint8_t *mate_flags;          /* MF: 1 = mate on the reverse strand, 2 = mate unmapped */
    int32_t *nf;                  /* recordsToNextFragment (NF series; anything where cf & 4 is not set) */
    const int32_t *ref_id;        /* sequenceId; NULL = all 0 */
    int32_t *pos;                 /* alignmentStart */
    int32_t *mate_ref_id, *mate_pos, *tlen;
    const uint64_t *read_off;     /* [n_reads+1] read lengths */
    const uint64_t *feat_off;     /* [n_reads+1] read features, as cramgpu_features */
    const uint8_t *f_op; const int32_t *f_val, *f_aux;
} cramgpu_mate_restore;
int cramgpu_mate_restore_dev(cramgpu_ctx *ctx, const cramgpu_mate_restore *io, int32_t *d_status, void *stream);

/* ------------------------------------------------------------------ CIGAR rebuild and BAM records
 *
 * Device-resident entry points: every pointer is DEVICE memory; results are device memory owned by the context, valid until
 * the next call of the same name on it.  Each call synchronises `stream` once (the result buffers are sized from totals).
 *
 * cramgpu_cigar_rebuild_dev   Cram2SamRecordFactory.getCigar2 [Cram2SamRecordFactory.java:119-225]: the CIGAR of every record
 *     from its read features ('X' 'B' -> M 1, 'i' -> I 1, 'I' 'S' -> their f_aux bases, 'D' 'N' 'H' 'P' -> f_val; gaps are M;
 *     quality-score features are passed over), BAM-packed (length << 4 | op of MIDNSHP).  A record with the unmapped flag
 *     gets no elements (Cram2SamRecordFactory.create sets "*", [:68-69]).  Note that '=' / 'X' of the original come back as M.
 *
 * cramgpu_bam_unpack_dev      a stripe of raw BAM records (each with its block_size prefix, back to back, as
 *     paralell/Bam_OBA_Supplier.java:45-82 cuts them; rec_off[i] = byte offset of record i, rec_off[n] = stripe length)
 *     -> structure of arrays: pos / mate_pos 1-based, bases as upper-case ASCII of "=ACMGRSVTWYHKDBN", names without the NUL,
 *     rflags = CRAMGPU_RF_NO_QUALS when the first quality byte is 0xFF (BAMRecord: missing qualities), CRAMGPU_RF_NO_BASES when
 *     l_seq = 0.  The arrays feed cramgpu_features_encode_dev (cramgpu_reads), cramgpu_mate_link_dev, cramgpu_host_series_pack_dev
 *     and cramgpu_slice_headers_dev as they are.  status[i] = CRAMGPU_ERR_CORRUPT for a record whose lengths do not add up to
 *     its block_size (Bam_OBA_Supplier: "Invalid record length" below 32); such a record comes out empty.
 *
 * cramgpu_bam_pack_dev        the inverse, laid out as cramtools' flat-buffer BAM writer does
 *     (encoding/reader/BAMRecordView.java:70-241): block_size refID pos-1 l_read_name mapq bin n_cigar flag l_seq next_refID
 *     next_pos-1 tlen name NUL cigar seq (4 bit, high nybble first) qual tags; bin = SAMRecord.computeIndexingBin
 *     [SAMRecord.java:1524-1531] (`in->bin` is ignored); CRAMGPU_RF_NO_QUALS -> l_seq bytes 0xFF [BAMRecordView.java:176-180].
 *     A base outside "=ACMGRSVTWYHKDBN" (either case) and '.': status[i] = CRAMGPU_ERR_ARG (IllegalArgumentException,
 *     [:405-407]), the nybble is written as N.
 */
typedef struct cramgpu_bam_records {
    uint64_t n_reads;
    const int32_t *ref_id, *pos, *mate_ref_id, *mate_pos, *tlen;
    const uint16_t *flag, *bin;
    const uint8_t *mapq, *rflags;
    const uint8_t *names;   const uint64_t *name_off;    /* [n_reads+1] */
    const uint32_t *cigar;  const uint64_t *cigar_off;   /* [n_reads+1] */
    const uint8_t *bases, *quals; const uint64_t *read_off;   /* [n_reads+1] */
    const uint8_t *tags;    const uint64_t *tag_off;     /* [n_reads+1]; tags may be NULL on input: no tags */
    uint64_t names_len, cigar_len, bases_len, tags_len;
} cramgpu_bam_records;
int cramgpu_cigar_rebuild_dev(cramgpu_ctx *ctx, uint64_t n_reads, const uint16_t *d_flag, const uint64_t *d_read_off,
                              const uint64_t *d_feat_off, const uint8_t *d_f_op, const int32_t *d_f_pos, const int32_t *d_f_val,
                              const int32_t *d_f_aux, uint32_t **d_cigar, uint64_t **d_cigar_off, uint64_t *cigar_len, void *stream);
int cramgpu_bam_unpack_dev(cramgpu_ctx *ctx, const uint8_t *d_bam, const uint64_t *d_rec_off, uint64_t n_reads,
                           cramgpu_bam_records *out, int32_t *d_status, void *stream);
int cramgpu_bam_pack_dev(cramgpu_ctx *ctx, const cramgpu_bam_records *in, uint8_t **d_bam, uint64_t *bam_len, uint64_t **d_rec_off,
                         int32_t *d_status, void *stream);

/* ------------------------------------------------------------------ the CRAM v3 block layer (block part)
 *
 * Block.write(3, os) [htsjdk Block.java:272-294] for blocks whose (compressed) bytes are in device memory -- e.g. the output of
 * cramgpu_rans_encode_dev (d_comp, d_out_off, d_out_len there): method | content type | ITF8 content id | ITF8 compressed size |
 * ITF8 raw size | data | CRC32 of all of it, little endian [CRC32OutputStream.java:50-51].  method / content_type =
 * BlockCompressionMethod / BlockContentType ordinals (RAW 0 GZIP 1 BZIP2 2 LZMA 3 RANS 4; EXTERNAL 4, CORE 5, ...).
 * content_id, method, content_type: HOST arrays [nblocks]; the other inputs DEVICE arrays (d_comp NULL: d_comp_off holds absolute
 * addresses).  Result: the framed blocks back to back, every block at a multiple of 16, in device memory owned by ctx (valid
 * until the next call); d_out_off[i] (DEVICE, [nblocks+1]) = start of block i, its length = header + data + 4.  One stream
 * synchronisation inside.
 */
int cramgpu_blocks_frame_dev(cramgpu_ctx *ctx, uint32_t nblocks, const uint8_t *d_comp, const uint64_t *d_comp_off,
                             const uint32_t *d_comp_len, const uint32_t *d_raw_len, const int32_t *content_id, const uint8_t *method,
                             const uint8_t *content_type, uint8_t **d_out, uint64_t *out_len, uint64_t **d_out_off, void *stream);

/* ------------------------------------------------------------------ the CRAM v3 container (container part)
 *
 * ContainerIO.writeContainer [htsjdk ContainerIO.java:152-209] for containers of ONE slice (how cramtools builds them:
 * ContainerFactory's records per slice = records per container): the bytes of every container, ready to be appended to the file --
 *   container header   ContainerHeaderIO.writeContainerHeader(3, ..) [ContainerHeaderIO.java:69-84]: int32 size of the blocks,
 *                      ITF8 sequence id / start / span / records, LTF8 global record counter / bases, ITF8 block count,
 *                      landmarks (one: the slice starts behind the compression header block), CRC32 little endian.
 *                      Sequence id, start and span are the slice's (ContainerFactory.buildContainer [:93-94],
 *                      calculateAlignmentBoundaries [:108-123]); -1 / -1 / 0 for an unplaced slice.
 *   compression header block   RAW, content type COMPRESSION_HEADER, the bytes of CompressionHeader.toByteArray() the Java host
 *                      built (comp_header; the encoding maps are host business)
 *   slice header block SliceIO.createSliceHeaderBlockContent(3, ..) [SliceIO.java:76-107] serialised on the device from the
 *                      result of cramgpu_slice_headers_dev: ITF8 fields, LTF8 global record counter, ITF8 number of blocks,
 *                      the content ids, embedded reference id -1, the reference MD5, and with CRAMGPU_CONTAINER_TAGS the
 *                      content digest tags in the slice's order B1 S1 B5 S5 BD SD (BinaryTagCodec.writeTag of byte arrays)
 *   core block, external blocks   Block.write(3, ..) as cramgpu_blocks_frame_dev, in the order of the block table: the core
 *                      block first, then the external blocks in the order the host's Slice.external map iterates.
 * Blocks of container c = rows [block_first[c], block_first[c+1]) of the block table (the table is cramgpu_blocks_frame_dev's).
 * global_record_counter: Container.globalRecordCounter = Slice.globalRecordCounter (cramgpu_allgather_sizes returns it).
 * Result: the containers back to back without padding in device memory owned by ctx (shared with cramgpu_blocks_frame_dev:
 * valid until the next call of either); d_out_off (DEVICE, [n_containers+1]) = where container c starts -- added to the file
 * position of the first one it is Container.offset.  One stream synchronisation inside.
 */
#define CRAMGPU_CONTAINER_TAGS 1u
typedef struct cramgpu_container_in {
    uint32_t n_containers;
    uint32_t flags;                          /* CRAMGPU_CONTAINER_TAGS */
    const cramgpu_slice_header *d_slices;    /* DEVICE [n_containers] */
    const uint64_t *global_record_counter;   /* HOST [n_containers] */
    const uint8_t *comp_header;              /* HOST: CompressionHeader.toByteArray() of every container, concatenated */
    const uint64_t *comp_header_off;         /* HOST [n_containers+1] */
    const uint32_t *block_first;             /* HOST [n_containers+1] */
    uint32_t nblocks;
    const uint8_t *d_comp;                   /* DEVICE (or NULL: d_comp_off holds absolute addresses) */
    const uint64_t *d_comp_off;              /* DEVICE [nblocks] */
    const uint32_t *d_comp_len, *d_raw_len;  /* DEVICE [nblocks] */
    const int32_t *content_id;               /* HOST [nblocks] */
    const uint8_t *method, *content_type;    /* HOST [nblocks] */
} cramgpu_container_in;
int cramgpu_containers_assemble_dev(cramgpu_ctx *ctx, const cramgpu_container_in *in, uint8_t **d_out, uint64_t *out_len,
                                    uint64_t **d_out_off, void *stream);

/* ------------------------------------------------------------------ multi-GPU: where this rank's containers go
 *
 * Containers are independent and are dealt round-robin: the container with global ordinal g is converted by
 * rank g % world.  The writer's two running sums -- container.offset (CRAMContainerStreamWriter.java:229-230) and
 * ContainerFactory.globalRecordCounter [ContainerFactory.java:77,105] -- become an exclusive scan over the per-container
 * sizes of all ranks.  The library does the bookkeeping; the exchange is a callback, so that the same entry point works
 * over NCCL / gloo (cramtools_b200/sharding.py), MPI, or the sockets of a JVM cluster.
 *
 *   local[i]          sizes of this rank's i-th container (global ordinal i * world + rank)
 *   gather            all-gather of `bytes` bytes: recv receives world * bytes, rank r's contribution at r * bytes;
 *                     returns 0 on success.  Called twice (counts, then sizes); not called when world == 1.
 *   base_offset       file offset of the first container (after the CRAM header); base_record: its global record counter
 *   offset[i], record_counter[i]   results for this rank's containers;  total (optional): sums over all ranks
 */
typedef struct cramgpu_container_size {
    uint64_t compressed_bytes;    /* bytes ContainerIO.writeContainer produced for the container */
    uint64_t n_bases;             /* Container.bases */
    uint32_t n_records;           /* Container.nofRecords */
    uint32_t reserved;
} cramgpu_container_size;
typedef int (*cramgpu_allgather_fn)(void *user, const void *send, void *recv, uint64_t bytes);
int cramgpu_allgather_sizes(cramgpu_ctx *ctx, uint32_t n_local, const cramgpu_container_size *local, uint32_t rank, uint32_t world,
                            cramgpu_allgather_fn gather, void *user, uint64_t base_offset, uint64_t base_record,
                            uint64_t *offset, uint64_t *record_counter, cramgpu_container_size *total);

#ifdef __cplusplus
}
#endif
#endif /* CRAMGPU_H */
