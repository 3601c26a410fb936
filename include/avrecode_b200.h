/* avrecode_b200.h -- C ABI of the B200-native avrecode hot path (libavrecode_b200.so).
 *
 * Drop-in boundary for ddkang/avrecode's per-bin loop.  In the reference that loop lives behind
 * struct AVCodecHooks (ffmpeg/libavcodec/coding_hooks.h:27-57): libavcodec calls back into
 * recode.cpp once PER BIN (get / get_bypass / get_terminate, coding_hooks.h:39-48) and once per
 * residual block (begin_sub_mb / begin_coding_type / end_coding_type, :27-37).  A device cannot be
 * called per bin, so this ABI moves the boundary up one level: the host hands over whole CABAC
 * slice payloads plus the few slice-header facts the syntax walk needs, and the library runs
 *   CABAC decode -> H.264 slice_data state machine -> avrecode model -> range coder   (compress)
 * and its exact inverse (decompress) in sm_100a kernels.  Each entry point cites what it replaces.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 or a
 * negative AVR_ERR_* code and never throws; per-slice problems are reported in
 * avr_slice_result.status (the host then emits skip_coded + literal, exactly like
 * recode.cpp:1289-1296); a context is bound to one CUDA device and is single-caller, different
 * contexts may be used concurrently from different threads.  There is NO CPU fallback: without a
 * usable CUDA device avr_create fails with AVR_ERR_NO_DEVICE.
 */
#ifndef AVRECODE_B200_H
#define AVRECODE_B200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AVR_ABI_VERSION 2

/* slice_type % 5 (H.264 7.4.3); SP/SI are reported unsupported */
#define AVR_SLICE_P 0
#define AVR_SLICE_B 1
#define AVR_SLICE_I 2

/* avr_slice_desc.flags */
#define AVR_SLICE_TRANSFORM_8X8        0x01 /* pps.transform_8x8_mode          (h264_cabac.c:1930) */
#define AVR_SLICE_DIRECT_8X8_INFERENCE 0x02 /* sps.direct_8x8_inference_flag   (h264.h:1165-1177)  */
#define AVR_SLICE_NEW_PICTURE          0x04 /* frame_spec hook fires before this slice (h264_slice.c:568-573) */
#define AVR_SLICE_MBAFF                0x08 /* unsupported on device -> status AVR_SLICE_UNSUPPORTED */
#define AVR_SLICE_FIELD                0x10 /* unsupported on device */
#define AVR_SLICE_STOP_BYTE            0x40 /* the rbsp_stop_one_bit sits alone in the byte right after the payload
                                               (payload rule h264.c:1298-1315): the reference's bin decoder, which reads the
                                               whole NAL, sees 0x80 there and zero padding after it; without the flag every
                                               byte past the payload reads as zero.  Compress direction only. */
#define AVR_SLICE_SKIP                 0x80 /* a CABAC slice the host leaves as skip_coded + literal (recode.cpp:1289-1296).
                                               The reference disables its bin hooks for it (recode.cpp:1139-1145) but the
                                               model hooks still fire, so the FrameBuffer flip of frame_spec
                                               (recode.cpp:824-843) must still happen: the device applies that and nothing
                                               else (in_size = 0, result out_size = 0, status OK).  Only meaningful when
                                               slices share a model (file mode). */
#define AVR_SLICE_NEW_MODEL            0x20 /* chunk boundary: a fresh h264_model (recode.cpp:618-1059)
                                               -- estimators {1,1}, both FrameBuffers empty -- starts here.
                                               Reference-parity ("file") mode sets it on the first slice of a
                                               file only; chunk mode sets it on every slice. */

/* One CABAC slice payload.  Replaces the arguments of CABACCodingHooks.init_decoder
 * (coding_hooks.h:41: buf,size) plus the H264SliceContext/H264Context fields the reference's
 * walker reads through container_of (h264_slice.c:2357-2368, h264_cabac.c:1267-1286). */
typedef struct avr_slice_desc {
  uint64_t in_off;        /* compress: offset of the slice_data payload in `in`; decompress: offset of Block.cabac */
  uint32_t in_size;       /* bytes at in_off */
  uint32_t payload_size;  /* Recoded.Block.size == `size` of init_decoder (bytes of slice_data) */
  uint32_t first_mb;      /* first_mb_in_slice */
  uint16_t mb_width;      /* h->mb_width  */
  uint16_t mb_height;     /* h->mb_height */
  int32_t  frame_num;     /* h->frame_num, key of the FrameBuffer flip (recode.cpp:824-843) */
  uint8_t  slice_type;    /* AVR_SLICE_P/B/I (sl->slice_type_nos) */
  uint8_t  cabac_init_idc;
  int8_t   slice_qp;      /* sl->qscale at slice start */
  uint8_t  num_ref_idx_l0;/* sl->ref_count[0] */
  uint8_t  num_ref_idx_l1;/* sl->ref_count[1] */
  uint8_t  chroma_format_idc; /* 1 = 4:2:0 (the only format the kernels walk) */
  uint8_t  flags;         /* AVR_SLICE_* */
  uint8_t  n_streams;     /* bits 0-3 (AVR_STREAMS_MASK): 0 or 1: an ordinary slice.  2..AVR_MAX_STREAMS: a "wide" slice of the chunked container
                             mode -- macroblock row r is coded in range-coder stream (r - first row) mod n_streams, so
                             that the slice decompresses as a wavefront of n_streams rows.  Block.cabac of such a slice
                             is: u32 macroblocks | n_streams x u32 stream bytes | the streams (little endian).  Only
                             for slices that carry AVR_SLICE_NEW_MODEL and are followed by one (a model of their own);
                             the format is this library's own, the reference has no counterpart.
                             bits 4-5: zero bytes (0..2) that end the slice header right before the payload; bit 7
                             (AVR_SLICE_ESCAPED), decompress only: the slice_data in the file contains emulation-
                             prevention bytes -- the coded block describes the UNESCAPED payload and the re-encoder
                             inserts them again (H.264 7.4.1), payload_size being the escaped size.  Chunked mode only:
                             the reference leaves such slices as literals (recode.cpp:1289-1296).
                             bit 6 (AVR_SLICE_PRIORS): a model that starts with this slice starts from the context's prior
                             table (avr_set_priors) instead of the reference's {1, 1} everywhere.  Chunked mode only. */
} avr_slice_desc;
#define AVR_STREAMS_MASK   0x0F
#define AVR_ESC_ZEROS_SHIFT 4
#define AVR_SLICE_ESCAPED  0x80
#define AVR_SLICE_PRIORS   0x40
#define AVR_MAX_PICTURE_MBS 139264u /* Level 6.2 MaxFS (H.264 table A-1): larger pictures are reported unsupported */
#define AVR_DEFAULT_WIDE_STREAMS 2
#define AVR_MAX_STREAMS 8   /* what this build runs: every stream of a wide slice is worked on by a thread block of its own */

/* avr_slice_result.status */
#define AVR_SLICE_OK           0
#define AVR_SLICE_UNSUPPORTED  1 /* MBAFF/field/I_PCM/4:2:2/4:4:4/SP/SI: leave the slice as a literal */
#define AVR_SLICE_CORRUPT      2 /* syntax walk failed (overrun, illegal value) */
#define AVR_SLICE_OVERFLOW     3 /* output did not fit its slot */

typedef struct avr_slice_result {
  uint64_t out_off;   /* offset of this slice's output in the buffer returned by the call */
  uint32_t out_size;  /* compress: bytes of Block.cabac; decompress: bytes of rebuilt slice_data */
  int32_t  status;    /* AVR_SLICE_* */
  uint32_t n_mbs;     /* macroblocks walked */
  uint32_t reserved;
  uint64_t n_bins;    /* CABAC bins decoded (compress) / rebuilt (decompress) */
} avr_slice_result;

typedef struct avr_config {
  int32_t abi_version;       /* AVR_ABI_VERSION */
  int32_t device;            /* CUDA device ordinal */
  int32_t blocks_per_sm;     /* 0 = default: thread blocks per SM of the kernel variant that keeps the estimators in global memory (1..8) */
  int32_t wide_stream_bytes; /* chunked mode of the file calls: a slice of at least 2 x this many payload bytes becomes a wide slice.
                                -1 = default (32768), 0 = never.  Environment AVR_WIDE_STREAM_BYTES overrides (debugging aid). */
  int32_t resident;          /* where a block keeps its estimators: -1 = choose per batch (estimated makespan), 1 = shared memory
                                (2 blocks per SM, fastest per slice), 0 = global memory (more blocks per SM).  Environment
                                AVR_RESIDENT overrides (debugging aid). */
  int32_t wide_batch_limit;  /* wide slices are a latency device: the file calls only use them while a batch has at most this many
                                CABAC slices.  -1 = default (4 x the device's SM count), 0 = no limit. */
  int32_t wide_max_streams;  /* most streams the file calls give a wide slice: 2..AVR_MAX_STREAMS, 0 = default (AVR_DEFAULT_WIDE_STREAMS:
                                as measured, decompression of a wide slice is bound by the one warp that re-encodes the CABAC
                                stream from two parsers on, and every further stream costs ratio -- DESIGN.md "Wide slices") */
  int32_t reserved[1];
} avr_config;

/* fills *cfg with the defaults (abi_version set, device 0, everything else "choose for me") */
void avr_config_init(avr_config* cfg);

typedef struct avr_ctx avr_ctx;

#define AVR_OK                 0
#define AVR_ERR_NO_DEVICE     -1 /* no CUDA device / driver: there is no CPU fallback */
#define AVR_ERR_CUDA          -2
#define AVR_ERR_ARG           -3
#define AVR_ERR_NOMEM         -4
#define AVR_ERR_FORMAT        -5 /* container / bitstream cannot be parsed */
#define AVR_ERR_STATE         -6 /* staged call sequence violated */
#define AVR_ERR_PRIORS        -7 /* the container was written from a prior table this context does not hold (avr_set_priors) */
#define AVR_ERR_LIMIT         -8 /* avr_recode_noalloc only: the batch is beyond that entry point's fixed limits */

const char* avr_strerror(int code);
int  avr_abi_version(void);

/* Lifetime.  Replaces av_decoder<Driver> owning the AVCodecHooks table (recode.cpp:212-229). */
int  avr_create(const avr_config* cfg, avr_ctx** out);
void avr_destroy(avr_ctx* ctx);

/* One-shot calls with HOST buffers (H2D, kernel, D2H inside).
 * avr_compress replaces compressor::cabac_decoder (recode.cpp:1134-1268): for each slice the
 * payload's bins are decoded and re-coded with the model; result i is Block.cabac of slice i.
 * avr_decompress replaces decompressor::cabac_decoder (recode.cpp:1411-1520): for each slice
 * Block.cabac is decoded with the model and the original slice_data bytes are rebuilt by the
 * device CABAC encoder (cabac_code.h:16-86), trailing lone 0x80 already dropped (:1501-1505).
 * `*out` points into context-owned pinned memory valid until the next call on `ctx`. */
int  avr_compress(avr_ctx* ctx, const avr_slice_desc* descs, int n_slices,
                  const uint8_t* in, uint64_t in_bytes,
                  avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes);
int  avr_decompress(avr_ctx* ctx, const avr_slice_desc* descs, int n_slices,
                    const uint8_t* in, uint64_t in_bytes,
                    avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes);

/* Staged form of the same work, for pipelines that keep inputs resident in HBM:
 *   avr_stage   copies descs + input bytes to the device (H2D) and sizes the output slots;
 *   avr_launch  enqueues the kernels on the context's stream (direction 0 = compress, 1 = decompress);
 *               may be called repeatedly on the same staged input;
 *   avr_sync    waits for the stream;
 *   avr_collect copies results + output bytes back (D2H) and compacts them.
 * avr_last_kernel_ms returns the CUDA-event duration of the most recent avr_launch's hot kernel
 * measured on the context's own stream (valid after avr_sync). */
int  avr_stage(avr_ctx* ctx, const avr_slice_desc* descs, int n_slices, const uint8_t* in, uint64_t in_bytes);
int  avr_launch(avr_ctx* ctx, int direction);
int  avr_sync(avr_ctx* ctx);
int  avr_collect(avr_ctx* ctx, avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes);
/* Per-class bin counts of the last launch, 8 words per slice in batch order (after avr_sync / avr_collect): what the
 * reference prints as its per-CodingType bills (recode.cpp:635-655), as counts.  [0] macroblock-level syntax (mb_type,
 * modes, ref_idx, mvd, cbp, qp_delta ...), [1] coded_block_flag, [2] significance map (sig + last), [3] coefficient level
 * prefixes, [4] bypass bins of residuals (signs, escapes), [5] bits of the nnz side channel (model-only bins, recode.cpp:
 * 844-930; not part of n_bins), [6] intra prediction modes, [7] end_of_slice (one per macroblock).  [0]+[1]+[2]+[3]+[4]+[6]+[7] = n_bins. */
#define AVR_HIST_CLASSES 8
int  avr_collect_hist(avr_ctx* ctx, uint32_t* hist, int n_slices);
int  avr_last_kernel_ms(avr_ctx* ctx, float* ms);
int  avr_launch_count(avr_ctx* ctx, uint64_t* n_launches); /* kernels launched by this context so far */
int  avr_flush_l2(avr_ctx* ctx);                           /* bench hygiene: write > L2-size scratch */

/* ---- host side: slice index + recode.proto container (no GPU needed) --------------------------
 * avr_index_h264 replaces what the reference gets from running libavcodec over the file
 * (recode.cpp:114-134 + find_next_coded_block_and_emit_literal :1275-1297): it scans an Annex-B
 * H.264 byte stream -- or walks the AVC video track of an ISO base media file (mp4 / mov: avcC parameter
 * sets, stsz/stsc/stco sample table, length-prefixed NAL units; what the reference gets from libavformat,
 * recode.cpp:92, 118) -- parses SPS/PPS/slice headers and returns one entry per CABAC slice payload in
 * file order.  recodable == 0 marks payloads the reference would emit as skip_coded (NAL-escaped
 * payload, size < 8, CAVLC never listed). */
typedef struct avr_index_entry {
  avr_slice_desc desc;     /* in_off = payload offset in the file, in_size = payload_size */
  int32_t  recodable;      /* 1: the reference would recode it.  0: it would not; when the only reason is emulation-
                              prevention bytes inside the payload, desc.n_streams carries AVR_SLICE_ESCAPED and the
                              chunked mode of the file calls recodes it anyway */
  int32_t  nal_type;
} avr_index_entry;

int  avr_index_h264(const uint8_t* file, uint64_t file_bytes, avr_index_entry** entries, int* n_entries);
/* The n_streams the file calls give this slice in AVR_MODE_SLICE when the batch is small enough for wide slices
 * (avr_config.wide_batch_limit): one stream per quantum of payload, 2..avr_config.wide_max_streams, when it is at least two stream quanta (avr_config.wide_stream_bytes of
 * `ctx`; ctx == NULL: the default of 32 KiB each), else 0.  No reference counterpart. */
int  avr_slice_streams(const avr_slice_desc* desc);
int  avr_ctx_slice_streams(avr_ctx* ctx, const avr_slice_desc* desc, int n_batch_slices);

/* chunking mode of the container produced by avr_file_compress */
#define AVR_MODE_FILE   0  /* reference parity: one model for the whole file -- output is byte-identical
                              to `recode compress` of the reference (recode.cpp:1115-1125) */
#define AVR_MODE_SLICE  1  /* B200 format: model reset per slice (every slice is a chunk); recorded in
                              Recoded.Metadata.version = "avr-b200/slice" */

/* recode compress / decompress (recode.cpp:1102-1309, 1312-1591) on in-memory files.
 * Output buffers are malloc'ed by the library; release with avr_free. */
int  avr_file_compress(avr_ctx* ctx, const uint8_t* file, uint64_t file_bytes, int mode,
                       uint8_t** rec, uint64_t* rec_bytes);
int  avr_file_decompress(avr_ctx* ctx, const uint8_t* rec, uint64_t rec_bytes,
                         uint8_t** file, uint64_t* file_bytes);
/* The same for a corpus of independent files in ONE device batch (BASELINE config 5: many clips per GPU).
 * Files never share a model: in AVR_MODE_FILE the result for file i is byte-identical to avr_file_compress of that
 * file alone; in AVR_MODE_SLICE it is too as long as the batch stays within avr_config.wide_batch_limit (beyond it
 * long slices are no longer split into streams -- a throughput batch needs no latency device; every container
 * decompresses to the same file either way).  recs[i] / files[i] are malloc'ed (avr_free each). */
int  avr_files_compress(avr_ctx* ctx, const uint8_t* const* files, const uint64_t* file_bytes, int n_files, int mode,
                        uint8_t** recs, uint64_t* rec_bytes);
int  avr_files_decompress(avr_ctx* ctx, const uint8_t* const* recs, const uint64_t* rec_bytes, int n_files,
                          uint8_t** files, uint64_t* file_bytes);

/* ---- static priors (chunked mode) -------------------------------------------------------------------------------
 * In AVR_MODE_SLICE every slice starts a model of its own, which costs about half of what the reference's one model per
 * file saves (a slice is over before the estimators have learnt much).  A PRIOR TABLE gives every fresh model a head
 * start: AVR_PRIOR_ENTRIES estimators in the table layout of avr_core.cuh (CNT_*), each a uint16 `pos | (pos + neg) << 8`
 * with pos, neg >= 1 and pos + neg <= 0x50 (what recode.cpp:1030-1047 can reach).  tools/train_priors.py makes one from
 * sample streams; avrecode_b200/priors_v1.bin is the table that ships.  The container records the table's digest
 * (Metadata field 2) and the decompress calls refuse a container whose digest is not that of the context's table
 * (AVR_ERR_PRIORS).  AVR_MODE_FILE never uses priors: its output is the reference's.  No reference counterpart.
 * table == NULL clears.  avr_priors_digest: 0 = no table. */
#define AVR_PRIOR_ENTRIES 47892
int      avr_set_priors(avr_ctx* ctx, const uint16_t* table, uint32_t n_entries);
uint32_t avr_priors_digest(avr_ctx* ctx);
uint32_t avr_priors_digest_of(const uint16_t* table, uint32_t n_entries);

/* ---- several devices of one node behind one handle ----------------------------------------------------------------
 * The reference runs one file on one core (recode.cpp main); its unit of parallelism is the process.  Here the units are
 * the slices (AVR_MODE_SLICE) or the files (AVR_MODE_FILE: the slices of a file share one model) of a call: they are dealt
 * out over the devices longest first, every device gets only the bytes of its share and runs it on a host thread of its
 * own, and nothing is exchanged between devices (the path shards, it has no collective).  The containers
 * are byte-identical to what one device writes.  `cfg` as for avr_create (its `device` is ignored; NULL = defaults). */
typedef struct avr_multi avr_multi;
int  avr_multi_create(const avr_config* cfg, const int* devices, int n_devices, avr_multi** out);
void avr_multi_destroy(avr_multi* m);
int  avr_multi_devices(const avr_multi* m);
avr_ctx* avr_multi_ctx(avr_multi* m, int i);   /* the context of device i (owned by m) */
int  avr_multi_set_priors(avr_multi* m, const uint16_t* table, uint32_t n_entries);
int  avr_multi_files_compress(avr_multi* m, const uint8_t* const* files, const uint64_t* file_bytes, int n_files, int mode,
                              uint8_t** recs, uint64_t* rec_bytes);
int  avr_multi_files_decompress(avr_multi* m, const uint8_t* const* recs, const uint64_t* rec_bytes, int n_files,
                                uint8_t** files, uint64_t* file_bytes);
/* ---- the entry point that allocates nothing --------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * One batch of slices through the job kernel of avr_compress (direction 0) / avr_decompress (direction 1), with the
 * caller's DEVICE buffers for input and output and everything else the kernel needs (its descriptor / job tables, job
 * counter, results) in a fixed file-scope device array of AVR_NOALLOC_SCRATCH_BYTES, which every GPU that loads the
 * library holds until the process exits.  The call makes no cudaMalloc / cudaFree / cudaHostAlloc and no host
 * allocation; it copies the tables in, launches, copies `results` out and returns when `stream` has finished them.
 *
 *   descs, results   HOST arrays of n_slices (descs[i].in_off is an offset into d_in, as for avr_compress)
 *   d_in, d_out      DEVICE pointers on `device`; slice i's bytes are at d_out + results[i].out_off
 *   out_cap          bytes at d_out; *out_need (may be NULL) receives the bytes this batch needs -- the sum over the
 *                    slices of in_size + in_size / 8 + 256 (direction 0) or payload_size + 64 (direction 1), each
 *                    rounded up to 16.  d_out == NULL: only reports *out_need.
 *   stream           a cudaStream_t of `device` (NULL: the default stream)
 *
 * Limits (AVR_ERR_LIMIT beyond them): at most AVR_NOALLOC_MAX_SLICES slices per call; out_cap >= *out_need; pictures whose
 * neighbour row fits beside the estimators in shared memory (any Level 6.2 width does).  AVR_ERR_ARG: a slice that
 * does not start a model of its own (every slice but the first must carry AVR_SLICE_NEW_MODEL: chunked mode only), a
 * wide slice (n_streams >= 2), AVR_SLICE_PRIORS or AVR_SLICE_SKIP -- those need buffers this entry point does not have.
 * The output bytes and results are those of avr_compress / avr_decompress for the same descriptors.
 * All calls of a process share the scratch array (per device), which keeps the last call's contents; the library clears
 * what must start clean and SERIALISES the calls (a mutex held until the results are back), because two calls
 * overlapping on one GPU would race on it.  Do not expect concurrency from this entry point. */
#define AVR_NOALLOC_MAX_SLICES    4096
#define AVR_NOALLOC_SCRATCH_BYTES (512 * 1024)
int  avr_recode_noalloc(int device, int direction, const avr_slice_desc* descs, int n_slices, const uint8_t* d_in,
                        uint64_t in_bytes, uint8_t* d_out, uint64_t out_cap, uint64_t* out_need,
                        avr_slice_result* results, void* stream);

void avr_free(void* p);

/* ---- measurement: single-lane microbenchmarks of the per-bin primitives ------------------------
 * The path is a serial integer recurrence per slice (cabac_functions.h:113-132 is ~20 dependent operations per
 * decision), so its speed limit is "cycles per bin of one lane".  avr_probe MEASURES that floor on `device`: probe
 * `which` runs one lane over the engine code for `iters` iterations and returns clock cycles, bins done and a
 * checksum.  avr_probe_name(which) describes it (NULL past the last probe).  No reference counterpart. */
const char* avr_probe_name(int which);
int  avr_probe(int device, int which, uint32_t iters, uint64_t* cycles, uint64_t* bins, uint64_t* checksum);

#ifdef __cplusplus
}
#endif
#endif /* AVRECODE_B200_H */
