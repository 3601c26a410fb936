// avr_core.cuh -- the per-bin hot path of avrecode, written for one warp per chunk on sm_100a.
//
//   compress  : CABAC bin decode -> H.264 slice_data walk -> avrecode model -> 64-bit range ENcoder
//   decompress: 64-bit range DEcoder -> avrecode model -> slice_data walk -> CABAC bin ENcoder
//
// What it replaces in the reference (paths relative to /root/reference):
//   ffmpeg/libavcodec/cabac_functions.h:56-205, cabac.c:193-218      CabacIn   (bin decoder)
//   cabac_code.h:16-86 (on arithmetic_code<u32,u16,0x200>)            CabacOut  (bin encoder)
//   arithmetic_code.h:87-201 / 209-298  (arithmetic_code<u64,u8>)      RangeEnc / RangeDec
//   recode.cpp:618-1059 (h264_model), block.h, framebuffer.h          Model tables (see layout below)
//   ffmpeg/libavcodec/h264_cabac.c:1267-2480, h264_mvpred.h:356-834   Walker<Dir>
//   recode.cpp:1134-1268 / 1411-1520 (the per-slice glue)             Dir = Compress / Decompress
//
// This is a re-design, not a translation: there are no per-bin callbacks, no std::map, no symbol
// queue and no 64-bit division.  A residual block is handled as a unit (significance mask in a
// 64-bit register, one nnz side-channel code, one pass over a contiguous counter slab); ordinary
// bins touch ONE packed 32-bit word in shared memory that holds the CABAC state and both avrecode
// counters; the neighbour state the context selection needs lives in a per-warp row buffer.
//
// Execution model: one warp owns one job (a run of slices that share a model).  Lane 0 runs the
// serial recurrence; all 32 lanes cooperate on table initialisation, frame-buffer clears and the
// 16-byte-per-lane refills of the shared-memory input window.
//
// The file also compiles as plain C++ (AVR_HOST_EMULATION) so that the exact device logic can be
// debugged in the GPU-less development container by tests/; that build is never shipped.
#pragma once
#include <stdint.h>
#include "../../include/avrecode_b200.h"

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_DEV __device__ __forceinline__
#define AVR_DEV_NOINLINE __device__ __noinline__
#define AVR_CONST __constant__
#define AVR_TABLE_QUAL static __constant__ const
#else
#define AVR_DEV inline
#define AVR_DEV_NOINLINE inline
#define AVR_CONST static const
#define AVR_TABLE_QUAL static const
#endif
#include "avr_tables.h"

namespace avr {

// ------------------------------------------------------------------------------------------------
// Model table layout (global memory, one set per job).  Counters are {pos, neg} bytes packed in a
// u16 (pos | neg << 8); both start at 1 and never exceed 0x61 (recode.cpp:1030-1047).
//
//  sig[]  significance-map estimators, key (cat, nnz, observed, zz) of recode.cpp:798-800.
//         slab(cat, nnz) = contiguous [observed][zz] array -> one block touches one slab.
//  nz[]   nnz side-channel estimators, key of recode.cpp:884:
//         (bit i, so_far, prev>=2^i, left_bit{0,1,2}, above_bit{0,1,2}, cat, is_8x8)
//  nnz[]  FrameBuffer::meta.num_nonzeros (framebuffer.h, block.h): [2 frames][mb][NNZ_STRIDE]
// ------------------------------------------------------------------------------------------------
enum : uint32_t {
  SIG_ZZ = 16,                     // zz index width (4x4: position 0..14, 8x8: ctxIdxInc 0..14)
  SIG_SLAB_4 = 16 * SIG_ZZ,        // [observed 0..15][zz]
  SIG_SLAB_8 = 64 * SIG_ZZ,        // [observed 0..63][zz]
  SIG_CAT_4 = 17 * SIG_SLAB_4,     // nnz 0..16
  SIG_CAT_8 = 64 * SIG_SLAB_8,     // nnz 0..63 (64 is stored as 0, SURVEY hazard H9)
  SIG_OFF_CAT5 = 5 * SIG_CAT_4,
  SIG_ENTRIES = 5 * SIG_CAT_4 + SIG_CAT_8,
  NZ_PER_BIT = 32 * 2 * 3 * 3 * 12,  // so_far(<32) x prev x left x above x (cat*2 + is_8x8)
  NZ_ENTRIES = 6 * NZ_PER_BIT,
  NNZ_STRIDE = 52,                 // 51 block slots + pad
  MODEL_COUNTERS = SIG_ENTRIES + NZ_ENTRIES,
};

struct JobDesc {        // one warp's work: slices [first, first + count) share one model
  uint32_t first_slice;
  uint32_t n_slices;
  uint64_t model_off;   // byte offset of this job's counter tables in the workspace
  uint64_t nnz_off;     // byte offset of this job's [2][nmb][NNZ_STRIDE] frame buffers
  uint32_t nnz_frame_bytes;  // bytes of ONE frame buffer
  uint32_t reserved;
};

struct SliceIO {        // device-side view of one slice's buffers
  uint64_t in_off;      // into the input arena
  uint64_t out_off;     // into the output arena
  uint32_t in_size;
  uint32_t out_cap;
};

// ------------------------------------------------------------------------------------------------
// Small shared tables (per thread block): LPS ranges, LPS transitions, reciprocals for the
// (range / total) of recode.cpp:819, 8x8 ctxIdxInc tables.
// ------------------------------------------------------------------------------------------------
struct SharedTables {
  uint64_t recip[100];     // floor(2^64 / d), d = 2..99  (recip[0], recip[1] unused)
  uint8_t range_lps[256];
  uint8_t trans_lps[64];
  uint8_t sig8[64];
  uint8_t last8[64];
};

AVR_DEV uint64_t mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
  return __umul64hi(a, b);
#else
  return uint64_t((unsigned __int128)a * b >> 64);
#endif
}
AVR_DEV int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz(int(x));
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
AVR_DEV int popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __popcll(x);
#else
  return __builtin_popcountll(x);
#endif
}

// exact floor(n / d) for 2 <= d < 100 without a hardware divide: one high multiply + one fix-up
AVR_DEV uint64_t div_small(const SharedTables& t, uint64_t n, uint32_t d) {
  uint64_t q = mulhi64(n, t.recip[d]);
  uint64_t r = n - q * d;
  return q + (r >= d);
}

// ------------------------------------------------------------------------------------------------
// Input byte window.  Lane 0 reads slice bytes through a per-warp shared-memory window that all 32
// lanes refill with one 16-byte load each (512 B per refill).  Bytes past the end read as zero
// (the reference reads NAL trailer bytes there; they cannot influence a decoded bin).
// ------------------------------------------------------------------------------------------------
enum : uint32_t { WIN_BYTES = 1024, WIN_HALF = 512 };

struct ByteWindow {
  const uint8_t* base;   // slice bytes in global memory (any alignment)
  uint8_t* win;          // WIN_BYTES of shared memory, 16-byte aligned
  uint32_t size;         // valid bytes
  uint32_t tail;         // value of the byte at position `size` (0x80 with AVR_SLICE_STOP_BYTE, else 0); zero after it
  uint32_t skew;         // base & 15: ring positions are counted from the 16-byte aligned address below base
  uint32_t win_start;    // the ring holds aligned positions [win_start, win_start + WIN_BYTES); multiple of WIN_HALF
  AVR_DEV void attach(const uint8_t* src, uint32_t n, uint8_t* window, uint32_t tail_byte) {
    base = src; size = n; win = window; win_start = 0; tail = tail_byte;
    skew = uint32_t(reinterpret_cast<uintptr_t>(src) & 15);
  }
  AVR_DEV uint32_t byte(uint32_t pos) const {
    const uint32_t a = pos + skew;
    if (a - win_start < WIN_BYTES) return win[a & (WIN_BYTES - 1)];
    return pos < size ? base[pos] : (pos == size ? tail : 0);   // slow path: ran ahead of the window inside one macroblock
  }
  AVR_DEV uint32_t be16(uint32_t pos) const { return (byte(pos) << 8) | byte(pos + 1); }
};

// ------------------------------------------------------------------------------------------------
// CABAC bin decoder (H.264 9.3.3.2).  val = offset << nbits | lookahead, 8 <= nbits <= 23 between
// calls, so every compare is a 32-bit compare against range << nbits.  Renormalisation = clz.
// ------------------------------------------------------------------------------------------------
struct CabacIn {
  uint32_t val, range, pos;
  int nbits;
  AVR_DEV void init(const ByteWindow& w) {
    val = (w.be16(0) << 16) | w.be16(2);
    pos = 4; nbits = 23; range = 510;
  }
  AVR_DEV void refill(const ByteWindow& w) {
    if (nbits < 8) { val = (val << 16) | w.be16(pos); pos += 2; nbits += 16; }
  }
  // state = pStateIdx << 1 | valMPS
  AVR_DEV int decision(const ByteWindow& w, const SharedTables& t, uint32_t& state) {
    uint32_t p = state >> 1;
    uint32_t rlps = t.range_lps[p * 4 + ((range >> 6) & 3)];
    range -= rlps;
    uint32_t scaled = range << nbits;
    int bit = int(state & 1);
    if (val >= scaled) {            // LPS
      val -= scaled;
      range = rlps;
      bit ^= 1;
      state = (uint32_t(t.trans_lps[p]) << 1) | ((state & 1) ^ (p == 0));
    } else if (p < 62) state += 2;
    int shift = clz32(range) - 23;
    range <<= shift;
    nbits -= shift;
    refill(w);
    return bit;
  }
  AVR_DEV int bypass(const ByteWindow& w) {
    nbits -= 1;
    uint32_t scaled = range << nbits;
    int bit = val >= scaled;
    if (bit) val -= scaled;
    refill(w);
    return bit;
  }
  AVR_DEV int terminate(const ByteWindow& w) {
    range -= 2;
    uint32_t scaled = range << nbits;
    if (val >= scaled) return 1;
    if (range < 256) { range <<= 1; nbits -= 1; refill(w); }
    return 0;
  }
};

// ------------------------------------------------------------------------------------------------
// Byte sink with carry resolution: one cached byte + a count of pending 0xFF bytes is equivalent to
// the reference's vector of deferred digits (arithmetic_code.h:148-180).
// ------------------------------------------------------------------------------------------------
struct ByteSink {
  uint8_t* out;
  uint32_t n, cap, pending;
  int cache;          // -1 = nothing cached yet
  bool overflow;
  AVR_DEV void init(uint8_t* o, uint32_t c) { out = o; n = 0; cap = c; pending = 0; cache = -1; overflow = false; }
  AVR_DEV void raw(uint32_t b) { if (n < cap) out[n] = uint8_t(b); else overflow = true; n++; }
  AVR_DEV void push(uint32_t byte, uint32_t carry) {
    if (carry) {
      raw(uint32_t(cache + 1));
      for (; pending; pending--) raw(0x00);
      cache = int(byte);
    } else if (byte == 0xFF && cache >= 0) {
      pending++;
    } else {
      if (cache >= 0) raw(uint32_t(cache));
      for (; pending; pending--) raw(0xFF);
      cache = int(byte);
    }
  }
  AVR_DEV void flush() {
    if (cache >= 0) raw(uint32_t(cache));
    for (; pending; pending--) raw(0xFF);
    cache = -1;
  }
};

// ------------------------------------------------------------------------------------------------
// avrecode's 64-bit range coder, arithmetic_code<uint64_t, uint8_t> (arithmetic_code.h:31-299):
// fixed_one = 2^63, one byte per digit, renormalise only when range < 2^51, then up to >= 2^55.
// P(1) = (range / (pos + neg)) * pos   (recode.cpp:816-820).
// ------------------------------------------------------------------------------------------------
#define AVR_MIN_RANGE (uint64_t(1) << 51)
#define AVR_NORM_RANGE (uint64_t(1) << 55)

struct RangeEnc {
  uint64_t low, range;
  ByteSink sink;
  AVR_DEV void init(uint8_t* out, uint32_t cap) { low = 0; range = uint64_t(1) << 63; sink.init(out, cap); }
  AVR_DEV void shift_byte() {
    uint32_t carry = uint32_t(low >> 63);
    uint32_t digit = uint32_t(low >> 55) & 0xFF;
    low = (low & ((uint64_t(1) << 55) - 1)) << 8;
    range <<= 8;
    sink.push(digit, carry);
  }
  AVR_DEV void put(const SharedTables& t, int bit, uint32_t pos, uint32_t neg) {
    uint64_t r1 = div_small(t, range, pos + neg) * pos;
    uint64_t r0 = range - r1;
    if (bit) { low += r0; range = r1; } else range = r0;
    if (range < AVR_MIN_RANGE) { do shift_byte(); while (range < AVR_NORM_RANGE); }
  }
  AVR_DEV void finish() {   // arithmetic_code.h:128-144
    for (uint64_t stop = uint64_t(1) << 62; stop > 0; stop >>= 1) {
      uint64_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; shift_byte(); }
    sink.flush();
  }
};

struct RangeDec {
  uint64_t low, range;
  const uint8_t* in;
  uint32_t pos, size, next;
  AVR_DEV uint32_t rd() { uint32_t v = pos < size ? in[pos] : 0; pos++; return v; }
  AVR_DEV void consume() {  // arithmetic_code.h:259-275: digits are read one bit off the byte grid
    uint32_t b = rd();
    uint32_t d = ((next << 7) | (b >> 1)) & 0xFF;
    next = b;
    low = (low << 8) | d;
    range <<= 8;
  }
  AVR_DEV void init(const uint8_t* p, uint32_t n) {  // arithmetic_code.h:218-230
    in = p; size = n; pos = 0;
    next = rd();
    low = next >> 1;
    range = 128;
    for (int i = 0; i < 7; i++) consume();
  }
  AVR_DEV int get(const SharedTables& t, uint32_t pos_, uint32_t neg_) {
    uint64_t r1 = div_small(t, range, pos_ + neg_) * pos_;
    uint64_t r0 = range - r1;
    int bit = low >= r0;
    if (bit) { low -= r0; range = r1; } else range = r0;
    if (range < AVR_MIN_RANGE) { do consume(); while (range < AVR_NORM_RANGE); }
    return bit;
  }
};

// ------------------------------------------------------------------------------------------------
// CABAC bin encoder (H.264 9.3.4.2-9.3.4.5), byte-batched: `queue` counts bits beyond the 10-bit
// low register (starts at -9: the very first bit is dropped, "firstBitFlag").  Output is identical
// to cabac_code.h:16-86 for every legal bin sequence.
// ------------------------------------------------------------------------------------------------
struct CabacOut {
  uint32_t low, range;
  int queue;
  ByteSink sink;
  AVR_DEV void init(uint8_t* out, uint32_t cap) { low = 0; range = 510; queue = -9; sink.init(out, cap); }
  AVR_DEV void put_bytes() {
    while (queue >= 0) {
      uint32_t o = low >> (queue + 10);
      low &= (0x400u << queue) - 1;
      queue -= 8;
      sink.push(o & 0xFF, o >> 8);
    }
  }
  AVR_DEV void decision(const SharedTables& t, int bin, uint32_t& state) {
    uint32_t p = state >> 1;
    uint32_t rlps = t.range_lps[p * 4 + ((range >> 6) & 3)];
    range -= rlps;
    if (uint32_t(bin) != (state & 1)) {
      low += range;
      range = rlps;
      state = (uint32_t(t.trans_lps[p]) << 1) | ((state & 1) ^ (p == 0));
    } else if (p < 62) state += 2;
    int shift = clz32(range) - 23;
    range <<= shift; low <<= shift; queue += shift;
    put_bytes();
  }
  AVR_DEV void bypass(int bin) {
    low <<= 1;
    if (bin) low += range;
    queue += 1;
    put_bytes();
  }
  AVR_DEV void terminate(int bin) {
    range -= 2;
    if (bin) {
      low += range;
      // range = 2 -> RenormE shifts 7, then 3 bits: (low >> 9) & 1, ((low >> 7) & 3) | 1
      low = ((low << 7) | 0x80) << 3;
      queue += 10;
      put_bytes();
      if (queue > -8) { low <<= -queue; queue = 0; put_bytes(); }  // rbsp_alignment_zero_bits
      sink.flush();
      // "Omit trailing byte if it's only a stop bit" (recode.cpp:1501-1505)
      if (sink.n > 0 && sink.n <= sink.cap && sink.out[sink.n - 1] == 0x80) sink.n--;
    } else if (range < 256) {
      range <<= 1; low <<= 1; queue += 1;
      put_bytes();
    }
  }
};

// ------------------------------------------------------------------------------------------------
// Per-warp working set in shared memory.
// ------------------------------------------------------------------------------------------------
struct MbCtx {          // what later macroblocks need from a decoded one to select CABAC contexts
  uint16_t flags;       // MBF_*
  uint16_t cbp;         // cbp_table semantics: luma 0..3, chroma 4..5, 0x40<<c chroma DC, 0x100 luma DC
  uint8_t nz_bottom;    // coefficient flags of the bottom block row: luma x=0..3 -> bits 0..3, Cb x=0,1 -> 4,5, Cr -> 6,7
  uint8_t nz_right;     // right block column: luma y=0..3 -> bits 0..3, Cb y=0,1 -> 4,5, Cr -> 6,7
  uint8_t ref_pos;      // bit (list*4 + b8): ref_idx > 0
  uint8_t reserved;
  uint8_t mvd_bottom[4][2];  // list 0, bottom row x=0..3, {|mvdx|,|mvdy|} clipped to 70
  uint8_t mvd_right[4][2];   // list 0, right column y=0..3
};
enum : uint16_t {
  MBF_I16 = 1, MBF_IPCM = 2, MBF_INTRA = 4, MBF_SKIP = 8, MBF_DIRECT = 16, MBF_T8x8 = 32,
  MBF_CHROMA_PRED_NZ = 64, MBF_L0 = 128, MBF_L1 = 256,
};

struct WarpState {
  alignas(16) uint8_t window[WIN_BYTES];
  uint32_t ctx[1024];      // state | pos << 8 | neg << 16  per ctxIdx (KS_CTX estimators of recode.cpp:679-683)
  uint32_t bypass_cnt;     // pos | neg << 8 (recode.cpp:1183)
  uint32_t term_cnt;       // (recode.cpp:1190)
  uint32_t pad_[2];
};

AVR_DEV void est_update(uint32_t& pos, uint32_t& neg, int bit, uint32_t limit) {  // recode.cpp:1035-1045
  if (bit) pos++; else neg++;
  if (pos + neg > limit) { pos = (pos + 1) >> 1; neg = (neg + 1) >> 1; }
}

// ------------------------------------------------------------------------------------------------
// Model view of one job.
// ------------------------------------------------------------------------------------------------
struct ModelMem {
  uint16_t* sig;      // SIG_ENTRIES
  uint16_t* nz;       // NZ_ENTRIES
  uint8_t* nnz[2];    // frame buffers, [mb][NNZ_STRIDE]
  int cur;            // index of the current frame buffer (recode.cpp:624, 824-843)
};

AVR_DEV uint32_t sig_slab(int cat, uint32_t key_nnz) {
  return cat == 5 ? SIG_OFF_CAT5 + key_nnz * SIG_SLAB_8 : uint32_t(cat) * SIG_CAT_4 + key_nnz * SIG_SLAB_4;
}

// ctxIdx bases per block category 0..5 (h264_cabac.c:1589-1599, :1556)
AVR_TABLE_QUAL int16_t k_sig_base[6] = {105, 120, 134, 149, 152, 402};
AVR_TABLE_QUAL int16_t k_last_base[6] = {166, 181, 195, 210, 213, 417};
AVR_TABLE_QUAL int16_t k_abs_base[6] = {227, 237, 247, 257, 266, 426};
AVR_TABLE_QUAL int16_t k_cbf_base[6] = {85, 89, 93, 97, 101, 1012};

// block index <-> 4x4 raster position (h264.h:1004-1024 scan8 geometry)
AVR_DEV int blk_x(int n) { return (n & 1) + ((n >> 1) & 2); }
AVR_DEV int blk_y(int n) { return ((n >> 1) & 1) + ((n >> 2) & 2); }
AVR_DEV int blk_at(int x, int y) { return (x & 1) | ((y & 1) << 1) | ((x & 2) << 1) | ((y & 2) << 2); }

// ------------------------------------------------------------------------------------------------
// Direction policies: the three bin primitives + the residual-block unit.
// ------------------------------------------------------------------------------------------------
struct Compress {
  static constexpr bool kDecompress = false;
  ByteWindow in;
  CabacIn cabac;
  RangeEnc rc;
  AVR_DEV void begin(const uint8_t* src, uint32_t src_size, uint8_t* dst, uint32_t dst_cap, uint8_t* window, uint32_t tail_byte) {
    in.attach(src, src_size, window, tail_byte);
    rc.init(dst, dst_cap);
  }
  AVR_DEV void start_bins() { cabac.init(in); }
  AVR_DEV int bin(const SharedTables& t, uint32_t& w) {
    uint32_t state = w & 0xFF, pos = (w >> 8) & 0xFF, neg = w >> 16;
    int b = cabac.decision(in, t, state);
    rc.put(t, b, pos, neg);
    est_update(pos, neg, b, 0x60);
    w = state | (pos << 8) | (neg << 16);
    return b;
  }
  AVR_DEV int bypass(const SharedTables& t, uint32_t& cnt) {
    uint32_t pos = cnt & 0xFF, neg = cnt >> 8;
    int b = cabac.bypass(in);
    rc.put(t, b, pos, neg);
    est_update(pos, neg, b, 0x60);
    cnt = pos | (neg << 8);
    return b;
  }
  AVR_DEV int terminate(const SharedTables& t, uint32_t& cnt) {
    uint32_t pos = cnt & 0xFF, neg = cnt >> 8;
    int b = cabac.terminate(in);
    rc.put(t, b, pos, neg);
    est_update(pos, neg, b, 0x60);
    cnt = pos | (neg << 8);
    return b;
  }
  // model-coded bit: encode `bit` (compress) / decode and return it (decompress)
  AVR_DEV int model_bit(const SharedTables& t, int bit, uint32_t pos, uint32_t neg) { rc.put(t, bit, pos, neg); return bit; }
  AVR_DEV void finish() { rc.finish(); }
  AVR_DEV uint32_t out_bytes() const { return rc.sink.n; }
  AVR_DEV bool overflowed() const { return rc.sink.overflow; }
  AVR_DEV uint32_t in_pos() const { return cabac.pos; }
};

struct Decompress {
  static constexpr bool kDecompress = true;
  RangeDec rc;
  CabacOut cabac;
  AVR_DEV void begin(const uint8_t* src, uint32_t src_size, uint8_t* dst, uint32_t dst_cap, uint8_t*, uint32_t) {
    rc.init(src, src_size);
    cabac.init(dst, dst_cap);
  }
  AVR_DEV void start_bins() {}
  AVR_DEV int bin(const SharedTables& t, uint32_t& w) {
    uint32_t state = w & 0xFF, pos = (w >> 8) & 0xFF, neg = w >> 16;
    int b = rc.get(t, pos, neg);
    cabac.decision(t, b, state);
    est_update(pos, neg, b, 0x60);
    w = state | (pos << 8) | (neg << 16);
    return b;
  }
  AVR_DEV int bypass(const SharedTables& t, uint32_t& cnt) {
    uint32_t pos = cnt & 0xFF, neg = cnt >> 8;
    int b = rc.get(t, pos, neg);
    est_update(pos, neg, b, 0x60);
    cabac.bypass(b);
    cnt = pos | (neg << 8);
    return b;
  }
  AVR_DEV int terminate(const SharedTables& t, uint32_t& cnt) {
    uint32_t pos = cnt & 0xFF, neg = cnt >> 8;
    int b = rc.get(t, pos, neg);
    est_update(pos, neg, b, 0x60);
    cabac.terminate(b);
    cnt = pos | (neg << 8);
    return b;
  }
  AVR_DEV int model_bit(const SharedTables& t, int, uint32_t pos, uint32_t neg) { return rc.get(t, pos, neg); }
  AVR_DEV void finish() {}
  AVR_DEV uint32_t out_bytes() const { return cabac.sink.n; }
  AVR_DEV bool overflowed() const { return cabac.sink.overflow; }
  AVR_DEV uint32_t in_pos() const { return rc.pos; }
};

// ------------------------------------------------------------------------------------------------
// The slice_data walker.  One instance per slice, lane 0 only.
// ------------------------------------------------------------------------------------------------
enum WalkStatus { WALK_OK = 0, WALK_UNSUPPORTED = 1, WALK_CORRUPT = 2, WALK_OVERFLOW = 3 };

template <class Dir>
struct Walker {
  Dir dir;
  const SharedTables* tab;
  WarpState* ws;
  MbCtx* row;            // [mb_width] most recent macroblock of every column
  ModelMem mm;
  avr_slice_desc d;
  // running state
  int mbw, mbh, mb_x, mb_y, mb_xy;
  int last_qp_delta;
  MbCtx left, cur;       // left neighbour (previous macroblock), macroblock being walked
  bool left_ok, top_ok;
  uint32_t left_cbp, top_cbp;
  uint64_t n_bins;
  int status;
  bool model_is8x8;      // BlockMeta::is_8x8 of the current macroblock (recode.cpp:944)
  uint8_t cur_nz[48];    // coefficient presence of the current macroblock, by block index (cbf contexts)
  uint8_t cur_mvd[16][2];
  int8_t cur_ref[4];

  AVR_DEV int bin(int ctx) { n_bins++; return dir.bin(*tab, ws->ctx[ctx]); }
  AVR_DEV int bypass() { n_bins++; return dir.bypass(*tab, ws->bypass_cnt); }
  AVR_DEV int terminate() { n_bins++; return dir.terminate(*tab, ws->term_cnt); }
  AVR_DEV void fail(int s) { if (status == WALK_OK) status = s; }

  // ---- avrecode model side channel: nnz of a coded block (recode.cpp:844-930) -------------------
  AVR_DEV uint8_t* nnz_slot(int frame, int mbi, int blk) { return mm.nnz[frame] + size_t(mbi) * NNZ_STRIDE + blk; }

  // get_neighbor_sub_mb (recode.cpp:419-471): frame geometry only -- slice edges are ignored on purpose
  AVR_DEV bool model_neighbor(bool above, int max_coeff, int blk, uint32_t* value) {
    int mbi = mb_xy, nb = blk;
    if (blk >= 48) {
      if (above) { if (mb_y == 0) return false; mbi -= mbw; }
      else { if (mb_x == 0) return false; mbi -= 1; }
    } else {
      int plane = blk & 48, x = blk_x(blk & 15), y = blk_y(blk & 15);
      if (above) { if (--y < 0) { if (mb_y == 0) return false; mbi -= mbw; y = 3; } }
      else { if (--x < 0) { if (mb_x == 0) return false; mbi -= 1; x = 3; } }
      nb = plane | blk_at(x, y);
      if (max_coeff >= 32) nb &= ~3;
    }
    *value = *nnz_slot(mm.cur, mbi, nb);
    return true;
  }
  AVR_DEV uint32_t nz_index(int i, uint32_t so_far, uint32_t prev_ge, uint32_t left_bit, uint32_t above_bit, int cat) {
    uint32_t ctx = uint32_t(cat) * 2 + (model_is8x8 ? 1 : 0);
    return uint32_t(i) * NZ_PER_BIT + ((((so_far * 2 + prev_ge) * 3 + left_bit) * 3 + above_bit) * 12 + ctx);
  }
  // codes (compress) or reads (decompress) the block's nnz; returns the value the compressor stores
  AVR_DEV uint32_t code_nnz(int cat, int blk, int max_coeff, uint32_t nnz) {
    const int nbits = max_coeff > 16 ? 6 : max_coeff > 4 ? 4 : 2;
    uint32_t left_nz = 0, above_nz = 0;
    bool has_left = model_neighbor(false, max_coeff, blk, &left_nz);
    model_neighbor(true, max_coeff, blk, &above_nz);
    uint32_t prev_nz = *nnz_slot(mm.cur ^ 1, mb_xy, blk);
    uint32_t so_far = 0;
    for (int i = 0; i < nbits; i++) {
      uint32_t cur_bit = 1u << i;
      uint32_t lb = has_left ? (left_nz >= cur_bit) : 2;
      uint32_t ab = above_nz ? (above_nz >= cur_bit) : 2;   // sic: value test (recode.cpp:881)
      uint16_t* e = mm.nz + nz_index(i, so_far, prev_nz >= cur_bit, lb, ab, cat);
      uint32_t c = *e, pos = c & 0xFF, neg = c >> 8;
      const int b = dir.model_bit(*tab, int((nnz >> i) & 1), pos, neg);
      est_update(pos, neg, b, 0x60);
      *e = uint16_t(pos | (neg << 8));
      if (b) so_far |= cur_bit;
    }
    if (Dir::kDecompress) return so_far;
    return nnz;
  }

  // ---- one residual block (h264_cabac.c:1582-1780 + recode.cpp queueing 1160-1255 / 1435-1492) --
  AVR_DEV void sig_state(int ctx, int bit_or_neg1, int* out_bit) {
    // CABAC-side half of a significance bin: only the 8-bit state of the packed word changes
    uint32_t w = ws->ctx[ctx], st = w & 0xFF;
    if (Dir::kDecompress) dir_cabac_put(bit_or_neg1, st);
    else *out_bit = dir_cabac_get(st);
    ws->ctx[ctx] = (w & ~0xFFu) | st;
    n_bins++;
  }
  AVR_DEV int dir_cabac_get(uint32_t& st) { return cabac_get_impl(dir, st); }
  AVR_DEV void dir_cabac_put(int b, uint32_t& st) { cabac_put_impl(dir, b, st); }
  AVR_DEV int cabac_get_impl(Compress& c, uint32_t& st) { return c.cabac.decision(c.in, *tab, st); }
  AVR_DEV int cabac_get_impl(Decompress&, uint32_t&) { return 0; }
  AVR_DEV void cabac_put_impl(Compress&, int, uint32_t&) {}
  AVR_DEV void cabac_put_impl(Decompress& c, int b, uint32_t& st) { c.cabac.decision(*tab, b, st); }

  AVR_DEV int residual(int cat, int blk, int max_coeff) {
    const bool is8 = max_coeff == 64;
    const int coefs = max_coeff - 1;
    const int sig_base = k_sig_base[cat], last_base = k_last_base[cat], abs_base = k_abs_base[cat];
    if (is8) model_is8x8 = true;          // before the nnz bits, as the compressor does
    uint32_t nnz = 0;
    uint64_t mask = 0;
    if (!Dir::kDecompress) {
      // 1. pull the significance map out of the CABAC stream (h264_cabac.c:1667-1681)
      int last;
      for (last = 0; last < coefs; last++) {
        int s = 0;
        sig_state(sig_base + (is8 ? tab->sig8[last] : last), -1, &s);
        if (s) {
          mask |= uint64_t(1) << last;
          nnz++;
          int l = 0;
          sig_state(last_base + (is8 ? tab->last8[last] : last), -1, &l);
          if (l) { last = max_coeff; break; }
        }
      }
      if (last == max_coeff - 1) { mask |= uint64_t(1) << last; nnz++; }
    }
    // 2. nnz side channel
    uint32_t coded = code_nnz(cat, blk, max_coeff, nnz & 63);
    uint32_t nnz_true, key_nnz;
    if (Dir::kDecompress) {
      nnz_true = coded == 0 ? uint32_t(max_coeff) : coded;
      if (nnz_true > uint32_t(max_coeff)) { fail(WALK_CORRUPT); nnz_true = uint32_t(max_coeff); }
      key_nnz = is8 ? coded : nnz_true;
    } else { nnz_true = nnz; key_nnz = nnz & 63; }
    *nnz_slot(mm.cur, mb_xy, blk) = uint8_t(key_nnz);   // stored count: 64 -> 0, else exact
    // 3. significance bins through the (cat, nnz) counter slab; last_significant_coeff_flag is
    //    never coded, it is implied by nnz (recode.cpp:1073, 1437-1438)
    uint16_t* slab = mm.sig + sig_slab(cat, key_nnz);
    uint32_t observed = 0;
    int zz = 0;
    while (observed < nnz_true && zz < coefs) {
      const int inc = is8 ? tab->sig8[zz] : zz;
      uint16_t* e = slab + observed * SIG_ZZ + inc;
      uint32_t c = *e, pos = c & 0xFF, neg = c >> 8;
      const int s = dir.model_bit(*tab, int((mask >> zz) & 1), pos, neg);
      if (Dir::kDecompress) sig_state(sig_base + inc, s, nullptr);
      est_update(pos, neg, s, 0x50);
      *e = uint16_t(pos | (neg << 8));
      if (s) {
        observed++;
        if (Dir::kDecompress) sig_state(last_base + (is8 ? tab->last8[zz] : zz), observed == nnz_true, nullptr);
      }
      zz++;
    }
    // 4. coefficient levels (h264_cabac.c:1724-1765): ordinary context bins + bypass
    int node = 0;
    for (uint32_t c = nnz_true; c > 0; c--) {
      const int l1 = node < 4 ? node + 1 : 0;                     // coeff_abs_level1_ctx
      if (bin(abs_base + l1) == 0) {
        node = node < 3 ? node + 1 : node;                         // transition after |level| == 1
        bypass();
      } else {
        const int gctx = abs_base + (node < 4 ? 5 : (node + 2 < 9 ? node + 2 : 9));  // coeff_abs_levelgt1_ctx
        node = node < 4 ? 4 : (node < 7 ? node + 1 : 7);           // transition after |level| > 1
        int abs = 2;
        while (abs < 15 && bin(gctx)) abs++;
        if (abs >= 15) {
          int j = 0;
          while (bypass() && j < 30) j++;
          while (j--) bypass();
        }
        bypass();
      }
    }
    return int(nnz_true);
  }

  // ---- context helpers --------------------------------------------------------------------------
  AVR_DEV int nz_neighbor(int blk, bool above, bool intra) const {   // h264_mvpred.h:542-587 semantics
    int plane = blk >> 4, x = blk_x(blk & 15), y = blk_y(blk & 15);
    int w = plane ? 2 : 4;
    if (above) {
      if (y > 0) return cur_nz[(blk & 48) | blk_at(x, y - 1)];
      if (!top_ok) return intra ? 1 : 0;
      const MbCtx& t = row[mb_x];
      int bit = plane == 0 ? x : (plane == 1 ? 4 + x : 6 + x);
      return (t.nz_bottom >> bit) & 1;
    }
    if (x > 0) return cur_nz[(blk & 48) | blk_at(x - 1, y)];
    (void)w;
    if (!left_ok) return intra ? 1 : 0;
    int bit = plane == 0 ? y : (plane == 1 ? 4 + y : 6 + y);
    return (left.nz_right >> bit) & 1;
  }
  AVR_DEV int cbf_ctx(int cat, int blk, bool intra) const {   // h264_cabac.c:1550-1580
    int a, b;
    if (cat == 3) { int i = blk - 49; a = (left_cbp >> (6 + i)) & 1; b = (top_cbp >> (6 + i)) & 1; }
    else if (cat == 0) { a = (left_cbp >> 8) & 1; b = (top_cbp >> 8) & 1; }
    else { a = nz_neighbor(blk, false, intra); b = nz_neighbor(blk, true, intra); }
    return k_cbf_base[cat] + a + 2 * b;
  }
  AVR_DEV int mvd_sum(int n, int comp) const {   // h264_cabac.c:1539-1548
    int x = blk_x(n), y = blk_y(n), a = 0, b = 0;
    if (x > 0) a = cur_mvd[blk_at(x - 1, y)][comp];
    else if (left_ok && (left.flags & MBF_L0)) a = left.mvd_right[y][comp];
    if (y > 0) b = cur_mvd[blk_at(x, y - 1)][comp];
    else if (top_ok && (row[mb_x].flags & MBF_L0)) b = row[mb_x].mvd_bottom[x][comp];
    return a + b;
  }
  AVR_DEV int mvd_component(int ctxbase, int amvd) {   // h264_cabac.c:1502-1537; returns stored |mvd|
    if (!bin(ctxbase + (amvd > 2) + (amvd > 32))) return 0;
    int mvd = 1;
    ctxbase += 3;
    while (mvd < 9 && bin(ctxbase)) { if (mvd < 4) ctxbase++; mvd++; }
    if (mvd >= 9) {
      int k = 3;
      while (bypass()) { mvd += 1 << k; k++; if (k > 24) { fail(WALK_CORRUPT); break; } }
      while (k--) mvd += bypass() << k;
    }
    bypass();   // sign
    return mvd < 70 ? mvd : 70;
  }
  AVR_DEV void mvd_part(int n, int w, int h) {
    int s0 = mvd_component(40, mvd_sum(n, 0));
    int s1 = mvd_component(47, mvd_sum(n, 1));
    int x = blk_x(n), y = blk_y(n);
    for (int j = 0; j < h; j++)
      for (int i = 0; i < w; i++) { uint8_t* m = cur_mvd[blk_at(x + i, y + j)]; m[0] = uint8_t(s0); m[1] = uint8_t(s1); }
  }
  AVR_DEV int ref_pos_neighbor(int n, bool above) const {   // "refa > 0" of h264_cabac.c:1475-1490 (P slices)
    int x = blk_x(n), y = blk_y(n);
    if (above) {
      if (y > 0) return cur_ref[(x >> 1) + 2 * ((y - 1) >> 1)] > 0;
      if (!top_ok) return 0;
      return (row[mb_x].ref_pos >> (2 + (x >> 1))) & 1;
    }
    if (x > 0) return cur_ref[((x - 1) >> 1) + 2 * (y >> 1)] > 0;
    if (!left_ok) return 0;
    return (left.ref_pos >> (1 + 2 * (y >> 1))) & 1;
  }
  AVR_DEV int read_ref(int n) {   // h264_cabac.c:1473-1500 with the rc > 1 guard of :2232-2240
    int rc = d.num_ref_idx_l0;
    if (rc <= 1) return 0;
    int ctx = ref_pos_neighbor(n, false) + 2 * ref_pos_neighbor(n, true);
    int ref = 0;
    while (bin(54 + ctx)) {
      ref++;
      ctx = (ctx >> 2) + 4;
      if (ref >= 32) { fail(WALK_CORRUPT); break; }
    }
    if (ref >= rc) fail(WALK_CORRUPT);
    return ref;
  }

  AVR_DEV int intra_mb_type(int base, bool intra_slice) {   // h264_cabac.c:1300-1330
    if (intra_slice) {
      int ctx = 0;
      if (left_ok && (left.flags & (MBF_I16 | MBF_IPCM))) ctx++;
      if (top_ok && (row[mb_x].flags & (MBF_I16 | MBF_IPCM))) ctx++;
      if (!bin(base + ctx)) return 0;
      base += 2;
    } else if (!bin(base)) return 0;
    if (terminate()) return 25;
    int is = intra_slice ? 1 : 0;
    int t = 1 + 12 * bin(base + 1);
    if (bin(base + 2)) t += 4 + 4 * bin(base + 2 + is);
    t += 2 * bin(base + 3 + is);
    t += bin(base + 3 + 2 * is);
    return t;
  }

  AVR_DEV void luma_block4(int cat, int n, int max_coeff, bool intra) {
    if (bin(cbf_ctx(cat, n, intra))) { residual(cat, n, max_coeff); cur_nz[n] = 1; }
  }

  // ---- one macroblock (h264_cabac.c:1926-2480, frame macroblocks, 4:2:0) --------------------------
  AVR_DEV void macroblock() {
    const bool is_p = d.slice_type == AVR_SLICE_P;
    cur = MbCtx();
    model_is8x8 = false;
    left_ok = mb_x > 0 && uint32_t(mb_xy - 1) >= d.first_mb;
    top_ok = mb_y > 0 && mb_xy - mbw >= int(d.first_mb);
    if (is_p) {   // mb_skip_flag, h264_cabac.c:1332-1367
      int ctx = 0;
      if (left_ok && !(left.flags & MBF_SKIP)) ctx++;
      if (top_ok && !(row[mb_x].flags & MBF_SKIP)) ctx++;
      if (bin(11 + ctx)) {
        cur.flags = MBF_SKIP | MBF_L0;
        last_qp_delta = 0;
        return;
      }
    }
    for (int i = 0; i < 48; i++) cur_nz[i] = 0;
    int itype = -1, ptype = -1;
    if (is_p) {   // h264_cabac.c:2008-2023
      if (!bin(14)) {
        if (!bin(15)) ptype = 3 * bin(16);
        else ptype = 2 - bin(17);
      } else itype = intra_mb_type(17, false);
    } else itype = intra_mb_type(3, true);

    const bool intra = itype >= 0;
    bool i16 = false;
    int cbp = 0;
    if (intra) {
      if (itype == 25) { fail(WALK_UNSUPPORTED); return; }   // I_PCM
      if (itype > 0) {
        i16 = true;
        int k = itype - 1;
        int chroma = (k >> 2) % 3;
        cbp = (k >= 12 ? 15 : 0) + 16 * chroma;
      }
      cur.flags = uint16_t(MBF_INTRA | (i16 ? MBF_I16 : 0));
    } else cur.flags = MBF_L0;

    // neighbour cbp with libavcodec's "unavailable" defaults (h264_mvpred.h:589-603)
    const MbCtx& top = row[mb_x];
    top_cbp = top_ok ? top.cbp : (intra ? 0x7CF : 0x00F);
    left_cbp = left_ok ? ((left.cbp & 0x7F0) | (left.cbp & 2) | ((left.cbp & 8))) : (intra ? 0x7CF : 0x00F);
    const int nts = ((top_ok && (top.flags & MBF_T8x8)) ? 1 : 0) + ((left_ok && (left.flags & MBF_T8x8)) ? 1 : 0);
    bool t8_allowed = (d.flags & AVR_SLICE_TRANSFORM_8X8) != 0;
    bool t8 = false;

    if (intra) {
      if (!i16) {   // I_NxN: transform_size_8x8_flag + prediction modes (h264_cabac.c:2081-2099, 1369-1381)
        int nmodes = 16;
        if (t8_allowed && bin(399 + nts)) { t8 = true; nmodes = 4; }
        for (int i = 0; i < nmodes; i++)
          if (!bin(68)) { bin(69); bin(69); bin(69); }
      }
      // intra_chroma_pred_mode (h264_cabac.c:1383-1406)
      int ctx = 0;
      if (left_ok && (left.flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (top_ok && (top.flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (bin(64 + ctx)) {
        cur.flags |= MBF_CHROMA_PRED_NZ;
        if (bin(67)) bin(67);
      }
    } else {
      for (int i = 0; i < 16; i++) { cur_mvd[i][0] = 0; cur_mvd[i][1] = 0; }
      for (int i = 0; i < 4; i++) cur_ref[i] = 0;
      if (ptype == 3) {   // P_8x8 (h264_cabac.c:2114-2220)
        int sub[4];
        for (int i = 0; i < 4; i++) {   // decode_cabac_p_mb_sub_type :1445-1454
          if (bin(21)) sub[i] = 0;
          else if (!bin(22)) sub[i] = 1;
          else sub[i] = bin(23) ? 2 : 3;
        }
        for (int i = 0; i < 4; i++) cur_ref[i] = int8_t(read_ref(4 * i));
        for (int i = 0; i < 4; i++) if (sub[i] != 0) t8_allowed = false;   // get_dct8x8_allowed, h264.h:1165-1177
        for (int i = 0; i < 4; i++) {
          int n = 4 * i;
          if (sub[i] == 0) mvd_part(n, 2, 2);
          else if (sub[i] == 1) { mvd_part(n, 2, 1); mvd_part(n + 2, 2, 1); }
          else if (sub[i] == 2) { mvd_part(n, 1, 2); mvd_part(n + 1, 1, 2); }
          else for (int j = 0; j < 4; j++) mvd_part(n + j, 1, 1);
        }
      } else if (ptype == 0) {
        int r = read_ref(0);
        for (int i = 0; i < 4; i++) cur_ref[i] = int8_t(r);
        mvd_part(0, 4, 4);
      } else if (ptype == 1) {   // 16x8
        int r0 = read_ref(0); cur_ref[0] = cur_ref[1] = int8_t(r0);
        int r1 = read_ref(8); cur_ref[2] = cur_ref[3] = int8_t(r1);
        mvd_part(0, 4, 2);
        mvd_part(8, 4, 2);
      } else {                   // 8x16
        int r0 = read_ref(0); cur_ref[0] = cur_ref[2] = int8_t(r0);
        int r1 = read_ref(4); cur_ref[1] = cur_ref[3] = int8_t(r1);
        mvd_part(0, 2, 4);
        mvd_part(4, 2, 4);
      }
    }

    if (!i16) {   // coded_block_pattern (h264_cabac.c:1408-1443)
      uint32_t a = left_cbp, b = top_cbp;
      int c = 0;
      c += bin(73 + !(a & 2) + 2 * !(b & 4));
      c += bin(73 + !(c & 1) + 2 * !(b & 8)) << 1;
      c += bin(73 + !(a & 8) + 2 * !(c & 1)) << 2;
      c += bin(73 + !(c & 4) + 2 * !(c & 2)) << 3;
      int ca = (a >> 4) & 3, cb = (b >> 4) & 3;
      if (bin(77 + (ca > 0) + 2 * (cb > 0))) c |= (1 + bin(77 + 4 + (ca == 2) + 2 * (cb == 2))) << 4;
      cbp = c;
    }
    cur.cbp = uint16_t(cbp);
    if (t8_allowed && (cbp & 15) && !intra) t8 = bin(399 + nts) != 0;   // h264_cabac.c:2348-2350
    if (t8) cur.flags |= MBF_T8x8;

    if (cbp || i16) {
      // mb_qp_delta (h264_cabac.c:2388-2416)
      if (bin(60 + (last_qp_delta != 0))) {
        int val = 1, ctx = 2;
        while (bin(60 + ctx)) { ctx = 3; val++; if (val > 102) { fail(WALK_CORRUPT); break; } }
        last_qp_delta = val;   // only "!= 0" is ever read
      } else last_qp_delta = 0;
      // luma (h264_cabac.c:1874-1920)
      if (i16) {
        if (bin(cbf_ctx(0, 48, true))) { residual(0, 48, 16); cur.cbp |= 0x100; }
        if (cbp & 15) for (int i = 0; i < 16; i++) luma_block4(1, i, 15, true);
      } else {
        for (int i8 = 0; i8 < 4; i8++) {
          if (!(cbp & (1 << i8))) continue;
          if (t8) { residual(5, 4 * i8, 64); for (int k = 0; k < 4; k++) cur_nz[4 * i8 + k] = 1; }
          else for (int i4 = 0; i4 < 4; i4++) luma_block4(2, 4 * i8 + i4, 16, intra);
        }
      }
      if (cbp & 0x30)   // chroma DC (h264_cabac.c:2449-2453)
        for (int c = 0; c < 2; c++)
          if (bin(cbf_ctx(3, 49 + c, intra))) { residual(3, 49 + c, 4); cur.cbp |= uint16_t(0x40 << c); }
      if (cbp & 0x20)   // chroma AC (h264_cabac.c:2455-2463)
        for (int c = 0; c < 2; c++)
          for (int i = 0; i < 4; i++) luma_block4(4, 16 + 16 * c + i, 15, intra);
    } else last_qp_delta = 0;

    // pack what the neighbours will need
    uint32_t nb = 0, nr = 0;
    for (int x = 0; x < 4; x++) nb |= uint32_t(cur_nz[blk_at(x, 3)]) << x;
    for (int y = 0; y < 4; y++) nr |= uint32_t(cur_nz[blk_at(3, y)]) << y;
    nb |= uint32_t(cur_nz[18]) << 4 | uint32_t(cur_nz[19]) << 5 | uint32_t(cur_nz[34]) << 6 | uint32_t(cur_nz[35]) << 7;
    nr |= uint32_t(cur_nz[17]) << 4 | uint32_t(cur_nz[19]) << 5 | uint32_t(cur_nz[33]) << 6 | uint32_t(cur_nz[35]) << 7;
    cur.nz_bottom = uint8_t(nb);
    cur.nz_right = uint8_t(nr);
    if (!intra) {
      uint32_t rp = 0;
      for (int i = 0; i < 4; i++) rp |= uint32_t(cur_ref[i] > 0) << i;
      cur.ref_pos = uint8_t(rp);
      for (int i = 0; i < 4; i++) {
        cur.mvd_bottom[i][0] = cur_mvd[blk_at(i, 3)][0]; cur.mvd_bottom[i][1] = cur_mvd[blk_at(i, 3)][1];
        cur.mvd_right[i][0] = cur_mvd[blk_at(3, i)][0]; cur.mvd_right[i][1] = cur_mvd[blk_at(3, i)][1];
      }
    }
  }

  // ---- slice loop (h264_slice.c:2366-2443) -------------------------------------------------------
  // Runs macroblocks until the slice ends or `budget` of them are done; returns true when finished.
  AVR_DEV void start() {
    mbw = d.mb_width; mbh = d.mb_height;
    mb_xy = int(d.first_mb);
    mb_x = mb_xy % mbw; mb_y = mb_xy / mbw;
    last_qp_delta = 0;
    n_bins = 0;
    status = WALK_OK;
    n_mbs = 0;
    done = false;
    left = MbCtx();
    if (d.slice_type > AVR_SLICE_I || d.slice_type == AVR_SLICE_B || d.chroma_format_idc != 1 ||
        (d.flags & (AVR_SLICE_MBAFF | AVR_SLICE_FIELD)) || mbw < 1 || mbh < 1 || mb_xy >= mbw * mbh) {
      status = WALK_UNSUPPORTED; done = true; return;
    }
    dir.start_bins();
  }
  uint32_t n_mbs;
  bool done;
  AVR_DEV void step() {
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; return; }
    macroblock();
    n_mbs++;
    int eos = 0;
    if (status == WALK_OK) eos = terminate();
    row[mb_x] = cur;
    left = cur;
    mb_xy++;
    if (++mb_x >= mbw) { mb_x = 0; mb_y++; }
    if (status != WALK_OK) { done = true; return; }
    if (dir.overflowed()) { fail(WALK_OVERFLOW); done = true; return; }
    if (eos) { dir.finish(); done = true; return; }
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; }   // picture full without end_of_slice: reference asserts here
  }
};

}  // namespace avr
