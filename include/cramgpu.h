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
    CRAMGPU_ERR_OUT_TOO_SMALL = -2, /* "Output buffer too small to fit N bytes." [RANS.java:44] */
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
 *   out[i], out_cap[i] caller-owned destination; cramgpu_rans_bound(in_len[i]) always suffices
 *   out_len[i]         bytes written
 *   status             optional per-block status array (may be NULL)
 * Returns the first non-OK per-block status, else CRAMGPU_OK.
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
 *                   in_off[i] must be a multiple of 16
 *   in_off/in_len/order   HOST arrays (the Java host knows every block length)
 *   d_out           device buffer of out_cap bytes; compressed/decoded blocks are packed into it
 *                   back to back, each starting at a multiple of 16
 *   d_out_off       DEVICE array of nblocks+1 u64, written by the call: block i occupies
 *                   d_out[d_out_off[i] .. +d_out_len[i]); d_out_off[nblocks] = bytes used
 *   d_out_len       DEVICE array of nblocks u32, written by the call
 *   d_status        DEVICE array of nblocks i32, written by the call (cramgpu_status per block)
 *   stream          cudaStream_t to enqueue on (NULL = the context's own stream)
 * Work is enqueued asynchronously; results are valid once `stream` has been synchronised.
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

#ifdef __cplusplus
}
#endif
#endif /* CRAMGPU_H */
