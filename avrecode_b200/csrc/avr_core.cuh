// avr_core.cuh -- the per-bin hot path of avrecode for sm_100a: engines, tables and model layout.
//
//   compress  : CABAC bin decode -> H.264 slice_data walk -> avrecode model -> 64-bit range ENcoder
//   decompress: 64-bit range DEcoder -> avrecode model -> slice_data walk -> CABAC bin ENcoder
//
// What it replaces in the reference (paths relative to /root/reference):
//   ffmpeg/libavcodec/cabac_functions.h:56-205, cabac.c:193-218      CabacIn   (bin decoder)
//   cabac_code.h:16-86 (on arithmetic_code<u32,u16,0x200>)            CabacOut  (bin encoder)
//   arithmetic_code.h:87-201 / 209-298  (arithmetic_code<u64,u8>)      RangeEnc / RangeDec
//   recode.cpp:618-1059 (h264_model), block.h, framebuffer.h          counter tables (layout below)
//   ffmpeg/libavcodec/h264_cabac.c:1267-2480, h264_mvpred.h:356-834   Walker<Dir>   (avr_walk.cuh)
//   recode.cpp:1134-1268 / 1411-1520 (the per-slice glue)             parser / coder roles (avr_job.cuh)
//
// This is a re-design, not a translation: no per-bin callbacks, no std::map, no symbol vector and
// no 64-bit division.  A slice is worked on by a PAIR of warps of one thread block that form a
// two-stage pipeline over a shared-memory symbol ring:
//   compress  : parser warp = CABAC decode + syntax walk  ->  coder warp = counters + range encoder
//   decompress: parser warp = range decoder + counters + syntax walk -> coder warp = CABAC encoder
// so the two serial recurrences of each direction overlap.  All adaptive state of a slice -- CABAC
// context states, every avrecode estimator (compacted to 96 KB), the neighbour row -- lives in
// shared memory when the batch is small enough for <= 2 blocks per SM (the latency-bound case);
// big batches keep the estimators in global memory / L1 and run more blocks per SM instead.
//
// The file also compiles as plain C++ (AVR_HOST_EMULATION) so that the exact device logic can be
// debugged in the GPU-less development container by tests/; that build is never shipped.
#pragma once
#include <stdint.h>
#include "../../include/avrecode_b200.h"

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_DEV __device__ __forceinline__
#define AVR_DEV_NOINLINE static __device__ __noinline__
#define AVR_TABLE_QUAL static __constant__ const
#else
#define AVR_DEV inline
#define AVR_DEV_NOINLINE static inline
#define AVR_TABLE_QUAL static const
#endif
#include "avr_tables.h"

namespace avr {

// ------------------------------------------------------------------------------------------------
// Dense CABAC context numbering.  Frame macroblocks in 4:2:0 use only ctxIdx 0..276, 399..435 and
// 1012..1023 (h264_cabac.c:1589-1599); they are renumbered 0..325 so that the per-slice state fits
// a few KB of shared memory.  The model key of an ordinary bin is its ctxIdx (recode.cpp:679-683),
// so the renumbering is invisible in the output.
// ------------------------------------------------------------------------------------------------
enum : int { N_CTX = 326, N_CTX_PAD = 328 };
constexpr int dctx(int ctx_idx) { return ctx_idx < 277 ? ctx_idx : ctx_idx < 1012 ? ctx_idx - 399 + 277 : ctx_idx - 1012 + 314; }
AVR_DEV int ctx_idx_of_dense(int i) { return i < 277 ? i : i < 314 ? i - 277 + 399 : i - 314 + 1012; }

// ------------------------------------------------------------------------------------------------
// Estimator tables of one model (recode.cpp:1057-1058), compacted.  One estimator = {pos, pos + neg}
// bytes packed in a u16 (pos | total << 8): the coders need the total for the reciprocal and pos for the
// product, and an update is ONE add (bin + 0x100).  pos and neg start at 1 (EST_INIT) and the total never
// exceeds 0x61 (recode.cpp:1030-1047).
//
//   [CNT_CTX, +N_CTX_PAD)  ordinary bins, keyed by (dense) ctxIdx            (recode.cpp:679-683)
//   CNT_BYPASS, CNT_TERM   the adaptive bypass / terminate estimators         (recode.cpp:1183, 1190)
//   [CNT_NZ, +NZ_ENTRIES)  nnz side channel, key of recode.cpp:884:
//        group(cat, is_8x8) -> [bit i: so_far < 2^i][prev >= 2^i][left_bit 0..2][above_bit 0..2]
//   [CNT_SIG, +SIG_ENTRIES) significance map, key (cat, nnz, observed, zz) of recode.cpp:798-800:
//        only observed < nnz exists, so (nnz, observed) pairs are stored triangularly; one block
//        touches one contiguous [observed][16] slab.
// ------------------------------------------------------------------------------------------------
enum : uint32_t {
  CNT_CTX = 0,
  CNT_BYPASS = N_CTX_PAD,
  CNT_TERM = N_CTX_PAD + 1,
  CNT_NZ = N_CTX_PAD + 2,
  NZ_GROUP_2 = 3 * 18,      // 2 coded bits:  (2^2 - 1) so_far slots x 2 x 3 x 3
  NZ_GROUP_4 = 15 * 18,
  NZ_GROUP_6 = 63 * 18,
  NZ_OFF_CAT3 = 8 * NZ_GROUP_4,                 // cats 0,1,2,4 x is_8x8 first
  NZ_OFF_CAT5 = NZ_OFF_CAT3 + 2 * NZ_GROUP_2,
  NZ_ENTRIES = NZ_OFF_CAT5 + NZ_GROUP_6,        // 3402
  CNT_SIG = CNT_NZ + NZ_ENTRIES,
  SIG_ZZ = 16,
  SIG_CAT_4 = 136 * SIG_ZZ,                     // nnz 1..16, observed < nnz
  SIG_OFF_CAT5 = 5 * SIG_CAT_4,
  SIG_CAT_8 = 2080 * SIG_ZZ,                    // key 0 (= 64 coefficients): observed < 64; key n: observed < n
  SIG_ENTRIES = SIG_OFF_CAT5 + SIG_CAT_8,       // 44160
  MODEL_COUNTERS = CNT_SIG + SIG_ENTRIES,       // 47892 estimators = 95784 bytes
  MODEL_COUNTERS_PAD = MODEL_COUNTERS + 36,     // the coders prefetch the estimator one step past a block (<= +17): allocated, never used
  // Split placement (the kernel variant for throughput batches): the 8x8 significance slab -- 70 % of the table, used by
  // one block category only -- lives in the global workspace (L1 / L2, prefetched per block), the rest in shared
  // memory, so that four blocks fit an SM instead of two.  cnt8 points CNT8_SPLIT entries before that slab, so the
  // same indices work with either placement (resident: cnt8 == cnt).
  CNT8_SPLIT = CNT_SIG + SIG_OFF_CAT5,
  CNT_LOW_PAD = CNT8_SPLIT + 36,
  CNT8_ENTRIES = SIG_CAT_8 + 36,
  EST_INIT = 0x0201,                            // {pos = 1, neg = 1}
  NNZ_STRIDE = 52,                              // FrameBuffer meta: 51 block slots + pad per macroblock
};
static_assert(MODEL_COUNTERS % 2 == 0 && MODEL_COUNTERS_PAD % 8 == 0 && MODEL_COUNTERS_PAD < 65536, "counter index must fit 16 bits");

AVR_DEV uint32_t nz_group(int cat, bool is8) {
  if (cat == 5) return NZ_OFF_CAT5;
  if (cat == 3) return NZ_OFF_CAT3 + (is8 ? NZ_GROUP_2 : 0);
  const uint32_t c = cat == 4 ? 3u : uint32_t(cat);   // cats 0,1,2,4 -> 0..3
  return (c * 2 + (is8 ? 1 : 0)) * NZ_GROUP_4;
}
AVR_DEV uint32_t sig_slab(int cat, uint32_t key_nnz) {   // index of [observed = 0][zz = 0]
  if (cat == 5) return CNT_SIG + SIG_OFF_CAT5 + (key_nnz == 0 ? 0u : 64u + key_nnz * (key_nnz - 1) / 2) * SIG_ZZ;
  return CNT_SIG + uint32_t(cat) * SIG_CAT_4 + (key_nnz * (key_nnz - 1) / 2) * SIG_ZZ;
}

// ------------------------------------------------------------------------------------------------
// Small read-only tables in shared memory (per thread block).
// ------------------------------------------------------------------------------------------------
struct alignas(8) U2 { uint32_t x, y; };
struct alignas(16) U4 { uint32_t x, y, z, w; };   // two U2 side by side: one 16-byte load
struct alignas(16) DivMagic { uint64_t magic; uint32_t shift, pad; };   // one 16-byte load per division
struct SharedTables {
  DivMagic div[100];       // magic = ceil(2^(63+l) / d), l = ceil(log2 d), shift = l - 1: exact floor(n/d) for n <= 2^63
  alignas(16) U2 trans[256];   // [state * 2 + lps] -> successor entry; the MPS/LPS pair of a state is one 16-byte load
  uint8_t sig8[64];        // significant_coeff_flag ctxIdxInc, 8x8 frame blocks
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

AVR_DEV uint32_t clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return uint32_t(__clzll((long long)x));
#else
  return x ? uint32_t(__builtin_clzll(x)) : 64u;
#endif
}

// exact floor(n / d) for 2 <= d < 100, n <= 2^63 (Granlund-Montgomery round-up reciprocal)
AVR_DEV uint64_t div_magic(uint64_t n, const DivMagic& m) { return mulhi64(n, m.magic) >> m.shift; }
AVR_DEV uint64_t div_small(const SharedTables& t, uint64_t n, uint32_t d) { return div_magic(n, t.div[d]); }

AVR_DEV uint32_t rlps4_of(uint32_t p) {
  return uint32_t(avr_range_lps[p * 4]) | uint32_t(avr_range_lps[p * 4 + 1]) << 8 | uint32_t(avr_range_lps[p * 4 + 2]) << 16 |
         uint32_t(avr_range_lps[p * 4 + 3]) << 24;
}
// Context entry word x: bit 0 valMPS, bits 4..10 the state (2 * pStateIdx + valMPS) -- times 16, the byte offset of
// its row in the transition table --, bits 16..27 the renormalisation shift of an LPS in each of the four range
// classes (clz(rangeTabLPS) - 23, 1..7, three bits each), so that the bin decoder needs no count-leading-zeros
// on its critical path.
enum : uint32_t { CTX_STATE_MASK = 0x7F0, CTX_SHLPS_SHIFT = 16 };
AVR_DEV uint32_t ctx_word_x(uint32_t state) {
  uint32_t x = (state & 0x7F) << 4 | (state & 1);
  for (int q = 0; q < 4; q++) x |= uint32_t(clz32(avr_range_lps[(state >> 1) * 4 + q]) - 23) << (CTX_SHLPS_SHIFT + 3 * q);
  return x;
}
AVR_DEV void fill_shared_tables(SharedTables* t, int tid, int nthreads) {
  for (int d = tid; d < 100; d += nthreads) {
    uint64_t m = 0; int l = 0;
    if (d >= 2) {
      while ((1u << l) < uint32_t(d)) l++;
      // ceil(2^(63+l)/d) without 128-bit arithmetic: 2^(63+l) = (q * d + r) * 2^l with 2^63 = q * d + r
      const uint64_t q = (uint64_t(1) << 63) / uint64_t(d), r = (uint64_t(1) << 63) % uint64_t(d);
      const uint64_t num = r << l;
      m = (q << l) + num / uint64_t(d) + (num % uint64_t(d) ? 1 : 0);
    }
    t->div[d].magic = m;
    t->div[d].shift = uint32_t(l > 0 ? l - 1 : 0);
    t->div[d].pad = 0;
  }
  for (int i = tid; i < 256; i += nthreads) {
    const uint32_t s = uint32_t(i) >> 1, p = s >> 1, mps = s & 1;
    uint32_t ns;
    if (i & 1) ns = (uint32_t(avr_trans_lps[p]) << 1) | (mps ^ (p == 0 ? 1u : 0u));
    else ns = p < 62 ? s + 2 : s;
    t->trans[i].x = ctx_word_x(ns);
    t->trans[i].y = rlps4_of(ns >> 1);
  }
  for (int i = tid; i < 64; i += nthreads) {
    t->sig8[i] = i < 63 ? avr_sig8x8_inc[i] : 0;
    t->last8[i] = i < 63 ? avr_last8x8_inc[i] : 0;
  }
}

// ------------------------------------------------------------------------------------------------
// Input byte window.  Lane 0 reads the slice (compress) or the coded block (decompress) through a
// shared-memory ring that all 32 lanes of the parser warp refill with one aligned 16-byte load
// each (512 B per refill).  Past the payload a lone stop-bit byte may follow (AVR_SLICE_STOP_BYTE),
// then zeros -- exactly what the reference's decoder sees in libavcodec's padded NAL buffer.
// ------------------------------------------------------------------------------------------------
enum : uint32_t { WIN_BYTES = 2048, WIN_HALF = 1024, WIN_GUARD = 16 };

struct ByteWindow {
  const uint8_t* base;   // bytes in global memory (any alignment)
  uint8_t* win;          // WIN_BYTES of shared memory, 16-byte aligned
  uint32_t size;         // valid bytes
  uint32_t tail;         // value of the byte at position `size`; zero after it
  uint32_t skew;         // base & 15: ring positions are counted from the 16-byte aligned address below base
  uint32_t win_start;    // the ring holds aligned positions [win_start, win_start + WIN_BYTES); multiple of WIN_HALF
  AVR_DEV void attach(const uint8_t* src, uint32_t n, uint8_t* window, uint32_t tail_byte) {
    base = src; size = n; win = window; win_start = 0; tail = tail_byte;
    skew = uint32_t(reinterpret_cast<uintptr_t>(src) & 15);
  }
  // No range check on the hot path: between two refills the reader may run at most WIN_HALF - WIN_GUARD
  // bytes ahead (a conformant macroblock is <= 400 bytes: 3200 bits, H.264 A.3.1); the parser verifies
  // that after every macroblock (ran_past) and gives the slice up as unsupported otherwise.
  AVR_DEV uint32_t byte(uint32_t pos) const { return win[(pos + skew) & (WIN_BYTES - 1)]; }
  AVR_DEV bool ran_past(uint32_t pos) const { return pos + skew + WIN_GUARD > win_start + WIN_BYTES; }
  AVR_DEV uint32_t be16(uint32_t pos) const { return (byte(pos) << 8) | byte(pos + 1); }
  AVR_DEV uint32_t limit() const { return win_start + WIN_HALF - skew; }   // consume up to here, then refill
};

// ------------------------------------------------------------------------------------------------
// CABAC bin decoder (H.264 9.3.3.2).  val = offset << nbits | lookahead, 8 <= nbits <= 23 between
// calls, so every compare is a 32-bit compare against range << nbits.  Renormalisation = clz.
// A context is one 8-byte shared-memory entry {state, rangeTabLPS[state][0..3]}: one load gives
// everything the decision needs; the transition is one more 8-byte table load off the critical path.
// ------------------------------------------------------------------------------------------------
// rangeTabLPS[state][(range >> 6) & 3] out of the four bytes packed in a context entry: one byte permute
AVR_DEV uint32_t rlps_sel(uint32_t four, uint32_t q) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(four, 0u, q | 0x4440u);
#else
  return (four >> (q * 8)) & 0xFF;
#endif
}
AVR_DEV uint32_t rlps_byte(uint32_t four, uint32_t range) { return rlps_sel(four, (range >> 6) & 3u); }
// the same with t = range >> 6 (4..7 for a normalised range 256..510): as a PRMT selector 4 + q picks byte q of the
// SECOND operand and the zero nibbles above it pick byte 0 of the first (zero) -- no masking, no or
AVR_DEV uint32_t rlps_by_t(uint32_t four, uint32_t t) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(0u, four, t);
#else
  return (four >> ((t & 3) * 8)) & 0xFF;
#endif
}
AVR_DEV const U4& ctx_trans_row(const SharedTables& t, uint32_t x) {
  return *reinterpret_cast<const U4*>(reinterpret_cast<const char*>(t.trans) + (x & CTX_STATE_MASK));
}

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
  // decides one bin from a context entry `e` that was loaded earlier (the caller may have prefetched
  // it while the previous bin was being decoded) and writes the successor state to *ctx
  // core of a decision: consumes entry e, returns the bin and leaves the successor entry in `next`.
  // Both successors (MPS / LPS) are fetched with ONE 16-byte load whose address only needs the state,
  // so the transition is a select, not a load, once the bin is known.
  AVR_DEV int decide_core(const ByteWindow& w, const SharedTables& t, const U2 e, U2* next) {
    const U4 pair = ctx_trans_row(t, e.x);
    const uint32_t tq = range >> 6;                         // 4 + quantile
    const uint32_t rlps = rlps_by_t(e.y, tq);
    const uint32_t sh_lps = (e.x >> (3 * tq + (CTX_SHLPS_SHIFT - 12))) & 7;   // renormalisation shift if this bin is the LPS
    range -= rlps;
    const uint32_t sh_mps = (range >> 8) ^ 1;               // range in [128, 510]: one shift iff below 256
    const uint32_t scaled = range << nbits;
    const uint32_t lps = val >= scaled ? 1u : 0u;
    if (lps) { val -= scaled; range = rlps; }
    const int shift = int(lps ? sh_lps : sh_mps);
    next->x = lps ? pair.z : pair.x;
    next->y = lps ? pair.w : pair.y;
    range <<= shift;
    nbits -= shift;
    refill(w);
    return int((e.x & 1) ^ lps);
  }
  // decides one bin from a context entry `e` that was loaded earlier (the caller may have prefetched
  // it while the previous bin was being decoded) and writes the successor state to *ctx
  AVR_DEV int decide(const ByteWindow& w, const SharedTables& t, const U2 e, U2* ctx) {
    U2 n;
    const int b = decide_core(w, t, e, &n);
    *ctx = n;
    return b;
  }
  AVR_DEV int decision(const ByteWindow& w, const SharedTables& t, U2* ctx) { return decide(w, t, *ctx, ctx); }
  AVR_DEV int bypass(const ByteWindow& w) {
    nbits -= 1;
    const uint32_t scaled = range << nbits;
    const int bit = val >= scaled;
    if (bit) val -= scaled;
    refill(w);
    return bit;
  }
  AVR_DEV int terminate(const ByteWindow& w) {
    range -= 2;
    const uint32_t scaled = range << nbits;
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
  int esc;            // >= 0: emulation prevention (H.264 7.4.1) is applied on the way out, esc = zero bytes just written
  uint32_t overflow;
  AVR_DEV void init(uint8_t* o, uint32_t c, int esc_zeros = -1) { out = o; n = 0; cap = c; pending = 0; cache = -1; esc = esc_zeros; overflow = 0; }
  AVR_DEV void store(uint32_t b) { if (n < cap) out[n] = uint8_t(b); else overflow = 1; n++; }
  AVR_DEV void raw(uint32_t b) {
    if (esc >= 0) {   // chunked mode, escaped slices: 00 00 0x (x <= 3) becomes 00 00 03 0x
      if (esc >= 2 && b <= 3) { store(3); esc = 0; }
      esc = b == 0 ? esc + 1 : 0;
    }
    store(b);
  }
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
// The sink lives in shared memory and is only touched once per output BYTE (every 8-9 bins), through ONE out-of-line
// copy of its code: inlined at every bin site it was a third of the kernels' instructions (instruction cache).
// dc: bits 0..7 the byte, bit 8 the carry into the bytes before it.
AVR_DEV_NOINLINE void sink_push(ByteSink* s, uint32_t dc) { s->push(dc & 0xFF, dc >> 8); }

// ------------------------------------------------------------------------------------------------
// avrecode's 64-bit range coder, arithmetic_code<uint64_t, uint8_t> (arithmetic_code.h:31-299):
// fixed_one = 2^63, one byte per digit, renormalise only when range < 2^51, then up to >= 2^55.
// P(1) = (range / (pos + neg)) * pos   (recode.cpp:816-820).
// ------------------------------------------------------------------------------------------------
#define AVR_MIN_RANGE (uint64_t(1) << 51)
#define AVR_NORM_RANGE (uint64_t(1) << 55)

struct RangeEnc {
  uint64_t low, range;
  ByteSink* sink;
  AVR_DEV void init(ByteSink* s, uint8_t* out, uint32_t cap) { low = 0; range = uint64_t(1) << 63; sink = s; s->init(out, cap); }
  AVR_DEV void shift_byte() {
    const uint32_t dc = uint32_t(low >> 55);   // digit + carry
    low = (low & ((uint64_t(1) << 55) - 1)) << 8;
    range <<= 8;
    sink_push(sink, dc);
  }
  AVR_DEV void put(const DivMagic& mg, int bit, uint32_t pos) {   // mg = t.div[pos + neg]
    const uint64_t r1 = div_magic(range, mg) * pos;
    const uint64_t r0 = range - r1;
    if (bit) { low += r0; range = r1; } else range = r0;
    if (range < AVR_MIN_RANGE) { do shift_byte(); while (range < AVR_NORM_RANGE); }
  }
};
// arithmetic_code.h:128-144, once per slice
AVR_DEV_NOINLINE void range_enc_finish(ByteSink* sink, uint64_t low, uint64_t range) {
  for (uint64_t stop = uint64_t(1) << 62; stop > 0; stop >>= 1) {
    const uint64_t x = (low | stop) & ~(stop - 1);
    if (stop < range && low <= x && x < low + range) { low = x; break; }
  }
  while (low != 0) {
    const uint32_t dc = uint32_t(low >> 55);
    low = (low & ((uint64_t(1) << 55) - 1)) << 8;
    sink->push(dc & 0xFF, dc >> 8);
  }
  sink->flush();
}

// The decoder's read position: shared memory, touched once per input byte through ONE out-of-line copy.
struct DecHome {
  const uint8_t* win;   // the input window
  uint32_t skew, pos, next;
};
AVR_DEV_NOINLINE uint32_t dec_digit(DecHome* h) {  // arithmetic_code.h:259-275: digits are read one bit off the byte grid
  const uint32_t b = h->win[(h->pos + h->skew) & (WIN_BYTES - 1)];
  const uint32_t d = ((h->next << 7) | (b >> 1)) & 0xFF;
  h->pos++;
  h->next = b;
  return d;
}
struct RangeDec {
  uint64_t low, range;
  DecHome* home;
  AVR_DEV void consume() {
    low = (low << 8) | dec_digit(home);
    range <<= 8;
  }
  AVR_DEV void init(DecHome* h, const ByteWindow& w) {  // arithmetic_code.h:218-230
    home = h;
    h->win = w.win; h->skew = w.skew; h->pos = 1;
    h->next = w.byte(0);
    low = h->next >> 1;
    range = 128;
    for (int i = 0; i < 7; i++) consume();
  }
  AVR_DEV int get(const DivMagic& mg, uint32_t pos_) {   // mg = t.div[pos + neg]
    const uint64_t r1 = div_magic(range, mg) * pos_;
    const uint64_t r0 = range - r1;
    const int bit = low >= r0;
    if (bit) { low -= r0; range = r1; } else range = r0;
    if (range < AVR_MIN_RANGE) { do consume(); while (range < AVR_NORM_RANGE); }
    return bit;
  }
  AVR_DEV uint32_t pos() const { return home->pos; }
};

// ------------------------------------------------------------------------------------------------
// CABAC bin encoder (H.264 9.3.4.2-9.3.4.5), byte-batched: `queue` counts bits beyond the 10-bit
// low register (starts at -9: the very first bit is dropped, "firstBitFlag").  Output is identical
// to cabac_code.h:16-86 for every legal bin sequence.
// ------------------------------------------------------------------------------------------------
struct CabacOut {
  uint32_t low, range;
  int queue;
  ByteSink* sink;
  AVR_DEV void init(ByteSink* s, uint8_t* out, uint32_t cap, int esc_zeros = -1) { low = 0; range = 510; queue = -9; sink = s; s->init(out, cap, esc_zeros); }
  AVR_DEV void put_bytes() {
    while (queue >= 0) {
      const uint32_t o = low >> (queue + 10);
      low &= (0x400u << queue) - 1;
      queue -= 8;
      sink_push(sink, o & 0x1FF);
    }
  }
  // core of a decision, the mirror image of CabacIn::decide_core: consumes context entry e, leaves the successor
  // in `next`.  Both successors come with ONE 16-byte load addressed by the state alone, and the renormalisation
  // shift comes out of the entry (LPS) or of range's top bit (MPS): no count-leading-zeros, no dependent load.
  AVR_DEV void encode_core(const SharedTables& t, const U2 e, int bin, U2* next) {
    const U4 pair = ctx_trans_row(t, e.x);
    const uint32_t tq = range >> 6;                         // 4 + quantile
    const uint32_t rlps = rlps_by_t(e.y, tq);
    const uint32_t sh_lps = (e.x >> (3 * tq + (CTX_SHLPS_SHIFT - 12))) & 7;
    range -= rlps;
    const uint32_t sh_mps = (range >> 8) ^ 1;
    const uint32_t lps = uint32_t(bin) != (e.x & 1) ? 1u : 0u;
    if (lps) { low += range; range = rlps; }
    const int shift = int(lps ? sh_lps : sh_mps);
    next->x = lps ? pair.z : pair.x;
    next->y = lps ? pair.w : pair.y;
    range <<= shift; low <<= shift; queue += shift;
    put_bytes();
  }
  AVR_DEV void decision(const SharedTables& t, int bin, U2* ctx) {
    U2 n;
    encode_core(t, *ctx, bin, &n);
    *ctx = n;
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
      sink->flush();
      // "Omit trailing byte if it's only a stop bit" (recode.cpp:1501-1505)
      if (sink->n > 0 && sink->n <= sink->cap && sink->out[sink->n - 1] == 0x80) sink->n--;
    } else if (range < 256) {
      range <<= 1; low <<= 1; queue += 1;
      put_bytes();
    }
  }
};

// Estimator update (recode.cpp:1035-1045): count the bin, halve both counts (rounding up) past the limit.
AVR_DEV_NOINLINE uint32_t est_halve(uint32_t c) {   // out of line: rare, and inlined it is 7 predicated instructions on every bin
  uint32_t pos = c & 0xFF, neg = (c >> 8) - pos;
  pos = (pos + 1) >> 1; neg = (neg + 1) >> 1;
  return pos | (pos + neg) << 8;
}
AVR_DEV uint32_t est_next(uint32_t c, uint32_t bit, uint32_t limit) {
  uint32_t c2 = c + bit + 0x100u;
  if (c2 >= ((limit + 1u) << 8)) c2 = est_halve(c2);
  return c2;
}
// the rows of the 8x8 slab a block is about to walk (split placement: global memory; a shared-memory address: no-op)
AVR_DEV void est_prefetch(const uint16_t* p, uint32_t bytes) {
#if defined(__CUDA_ARCH__)
  const char* a = reinterpret_cast<const char*>(p);
#pragma unroll 1
  for (uint32_t o = 0; o < bytes; o += 128) asm volatile("prefetch.L1 [%0];" ::"l"(a + o));
#else
  (void)p; (void)bytes;
#endif
}
// limits: 0x50 while the bin is a significance-map bin, else 0x60 (recode.cpp:1041-1042)
enum : uint32_t { LIMIT_SIG = 0x50, LIMIT_ANY = 0x60 };

// (dense) ctxIdx bases per block category 0..5 (h264_cabac.c:1589-1599, :1556)
enum : int { CTX_SIG8 = 402 - 399 + 277, CTX_LAST8 = 417 - 399 + 277, CTX_ABS8 = 426 - 399 + 277, CTX_CBF8 = 1012 - 1012 + 314 };
AVR_TABLE_QUAL int16_t k_sig_base[6] = {105, 120, 134, 149, 152, CTX_SIG8};
AVR_TABLE_QUAL int16_t k_last_base[6] = {166, 181, 195, 210, 213, CTX_LAST8};
AVR_TABLE_QUAL int16_t k_abs_base[6] = {227, 237, 247, 257, 266, CTX_ABS8};
AVR_TABLE_QUAL int16_t k_cbf_base[6] = {85, 89, 93, 97, 101, CTX_CBF8};

// coefficients of a residual block by category (h264_cabac.c:1813-1818)
AVR_TABLE_QUAL uint8_t k_maxc[8] = {16, 15, 16, 4, 15, 64, 0, 0};

// ------------------------------------------------------------------------------------------------
// Symbol ring between the parser warp and the coder warp of a block (single producer, single
// consumer).  The unit is an ENTRY of 1..4 words; what crosses is syntax, not bins:
//
//   plain bin   bits 0..15 index (estimator index when compressing, dense ctx when decompressing),
//               bit 16 the bin, bits 20..21 the CABAC primitive (decompress)
//   SYM_LEVEL   one coefficient level: bits 0..8 first context, 9 its bin, 10..18 the "greater than one"
//               context, 19..22 how many 1-bins it coded, 23 whether a 0-bin ended that run, 24 whether
//               the sign follows, 25 the sign (bypass)
//   SYM_BLOCK   one residual block's significance map: bits 0..2 category, 8..14 coefficients (1..64);
//               compress adds bit 15 "the macroblock has an 8x8 block" (NZ key, recode.cpp:884), bit 16
//               "left neighbour exists" and a word {left, above, previous-picture count} for the nnz side
//               channel (recode.cpp:844-930); then the 64-bit significance mask, one word unless category 5.
//               The coder derives every bin from it: the compress coder the nnz bits and the model-coded
//               significance bins (last_significant_coeff_flag is never coded, recode.cpp:1073), the
//               decompress coder the CABAC significant_coeff_flag / last_significant_coeff_flag bins.
//   SYM_END     end of the slice
//
// The producer publishes its tail between entries only (every RING_PUBLISH words and at macroblock
// ends), so an entry the consumer can see the first word of is complete.
// ------------------------------------------------------------------------------------------------
enum : uint32_t {
  RING_ENTRIES = 512, RING_PUBLISH = 64, RING_BATCH = 128, ENTRY_MAX_WORDS = 4,
  SYM_BIT_SHIFT = 16, SYM_KIND_SHIFT = 20, SYM_DECISION = 0u << 20, SYM_BYPASS = 1u << 20, SYM_TERMINATE = 2u << 20,
  SYM_END = 1u << 31,
  // Wide slices only (one slice spread over several range-coder streams, macroblock row r in stream
  // r mod K): kind 3 entries hand a row over between the parser warps of a decompression --
  //   SP_ROW_END  a row is complete; bit 0 = "the last macroblock coded a non-zero mb_qp_delta"
  //   SP_QPD      first mb_qp_delta bin of a row: its CABAC context (60 + that bit) is only known to the
  //               re-encoder, which sees the rows in order; the bin value sits at SYM_BIT_SHIFT
  SYM_SPECIAL = 3u << 20, SP_ROW_END = 1u << 8, SP_QPD = 2u << 8,
  WIDE_MAX = AVR_MAX_STREAMS,   // most streams per wide slice (every stream is worked on by a thread block of its own)
  SYM_LEVEL = 1u << 30, LVL_BIT1 = 9, LVL_GCTX = 10, LVL_RUN = 19, LVL_TERM = 23, LVL_HAS_SIGN = 24, LVL_SIGN = 25,
  SYM_BLOCK = 1u << 29, BLK_NNZ = 8, BLK_MB8 = 15, BLK_HAS_LEFT = 16,
};
AVR_DEV uint32_t lvl_entry(int ctx1, int b1, int gctx, int run, int term, int has_sign, int sign) {
  return SYM_LEVEL | uint32_t(ctx1) | uint32_t(b1) << LVL_BIT1 | uint32_t(gctx) << LVL_GCTX | uint32_t(run) << LVL_RUN |
         uint32_t(term) << LVL_TERM | uint32_t(has_sign) << LVL_HAS_SIGN | uint32_t(sign) << LVL_SIGN;
}

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
// release/acquire at block scope is all the ring needs (entries before the published index)
// ... at block scope; rings whose two ends are in DIFFERENT thread blocks (wide slices: control words and entries in
// global memory) need device scope
AVR_DEV void ring_fence(bool far = false) {
  if (far) asm volatile("fence.acq_rel.gpu;" ::: "memory");
  else asm volatile("fence.acq_rel.cta;" ::: "memory");
}
// Waits of the wide-slice kernel are long (a coder warp idles while the rows of the other streams are parsed, a
// parser warp idles behind the row above), and a warp that polls flat out takes issue slots and shared-memory
// cycles from the warps of its block that do the serial work: poll a few times, then sleep between polls.
#define AVR_SYNCWARP_CORE() __syncwarp()
#define AVR_BCAST_CORE(v) __shfl_sync(0xFFFFFFFFu, (v), 0)
AVR_DEV void wide_backoff(uint32_t& polls) { if (++polls > 32) __nanosleep(polls > 256 ? 1000 : 100); }
// No wait of this library lasts seconds unless something is wrong (a role of a wide slice that never got a block, a
// protocol bug): rather than hang the device, give up the launch -- the host sees a CUDA error, not a dead GPU.
AVR_DEV void wait_watchdog(long long t0) { if (clock64() - t0 > (1ll << 34)) __trap(); }
struct RingCtl {
  uint32_t tail_pub, head_pub;
  uint32_t end_status, end_mbs, end_bins_lo, end_bins_hi;   // written by the parser before SYM_END
  uint32_t prod_tail, cons_head;   // wide slices: where a producer / consumer that moved to another ring left off
  uint32_t prod_wait, cons_wait;   // cycles / 256 the producer waited for room, the consumer for entries (which stage is the limit)
};
enum : uint32_t { REC_ENTRIES = 256 };
struct RingShared : RingCtl { uint32_t e[RING_ENTRIES]; };   // parser -> the warp behind it
struct RecRing : RingCtl { uint32_t e[REC_ENTRIES]; };       // compress: estimator warp -> range-coder warp
AVR_DEV void ring_reset(RingCtl* r) { r->tail_pub = 0; r->head_pub = 0; r->prod_tail = 0; r->cons_head = 0; r->prod_wait = 0; r->cons_wait = 0; }
// makes room for an entry of n words: publishes what is complete, waits for the consumer if the ring is full; returns
// the tail value up to which the producer may go on without looking again.  One out-of-line copy (instruction cache).
AVR_DEV_NOINLINE uint32_t ring_slow(RingCtl* r, uint32_t tail, uint32_t n, uint32_t cap, bool far) {
  ring_fence(far);
  *reinterpret_cast<volatile uint32_t*>(&r->tail_pub) = tail;
  uint32_t head_seen = *reinterpret_cast<volatile uint32_t*>(&r->head_pub);
  if (tail + n - head_seen > cap) {
    const long long t0 = clock64();
    do { head_seen = *reinterpret_cast<volatile uint32_t*>(&r->head_pub); wait_watchdog(t0); } while (tail + n - head_seen > cap);
    r->prod_wait += uint32_t((clock64() - t0) >> 8);
  }
  const uint32_t a = head_seen + cap, b = tail + RING_PUBLISH;
  return int32_t(a - b) < 0 ? a : b;
}
#else
}  // namespace avr
#include <vector>
namespace avr {
// host emulation: the parser runs a whole slice into an unbounded queue, then the coder drains it
inline void ring_fence(bool = false) {}
#define AVR_SYNCWARP_CORE() ((void)0)
#define AVR_BCAST_CORE(v) (v)
struct RingShared {
  std::vector<uint32_t> e;
  uint32_t end_status, end_mbs, end_bins_lo, end_bins_hi;
  size_t cons_head;
  uint32_t prod_wait = 0, cons_wait = 0;
};
inline void ring_reset(RingShared* r) { r->e.clear(); r->cons_head = 0; }
typedef RingShared RecRing;
typedef RingShared RingCtl;
#endif
// ------------------------------------------------------------------------------------------------
// Coder warp.  All lanes fetch a batch of ring words (a poll of the published tail; for a ring in global
// memory a coalesced copy into shared memory), lane 0 works through the entries of the batch: every bin of
// an entry is derived from the entry alone, so the estimators / contexts an entry will touch are known
// up front and are loaded before the serial recurrence (range coder / CABAC encoder) needs them.
// Tried and dropped (profiles/r2d_probes.txt): expanding the entries into one item per bin with all 32
// lanes and running one software-pipelined loop over the items -- 369 / 358 cycles per bin against
// 297 / 187 for deriving the bins in place: a warp instruction costs a lone warp the same 3-5 cycles
// whether 1 or 32 lanes execute it, so the expansion only pays for entries with many bins.
// ------------------------------------------------------------------------------------------------
struct WordSrc {       // words of a batch: the shared-memory ring itself (mask 511) or a staged copy (mask ~0)
  const uint32_t* e;
  uint32_t at, mask;
  AVR_DEV uint32_t take() { return e[at++ & mask]; }
};

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
// the consumer side of a ring as the coder warp (all lanes) sees it
struct RingReader {
  RingCtl* r;
  const uint32_t* ents;   // the ring's words when they are in shared memory
  uint32_t mask, head;
  bool patient;
  // patient: the other end is in another thread block (wide slices) -- long waits, device-scope fences
  AVR_DEV void init(RingShared* ring) { r = ring; ents = ring->e; mask = RING_ENTRIES - 1; head = 0; patient = false; }
  AVR_DEV void init_far(RingCtl* ring) { r = ring; ents = nullptr; mask = 0; head = 0; patient = true; }
  AVR_DEV void init_rec(RecRing* ring) { r = ring; ents = ring->e; mask = REC_ENTRIES - 1; head = 0; patient = false; }
  AVR_DEV uint32_t wait_avail() {   // all lanes poll the same word: a broadcast, no divergence
    uint32_t tail = *reinterpret_cast<volatile uint32_t*>(&r->tail_pub);
    if (head == tail) {
      const long long t0 = clock64();
      uint32_t polls = 0;
      while (head == (tail = *reinterpret_cast<volatile uint32_t*>(&r->tail_pub))) { if (patient) wide_backoff(polls); wait_watchdog(t0); }
      const uint32_t w = uint32_t((clock64() - t0) >> 8);
      if ((threadIdx.x & 31) == 0) r->cons_wait += w;
    }
    ring_fence(patient);
    return tail - head;   // every entry that starts inside this count is complete
  }
  AVR_DEV void release(uint32_t n, int lane) {
    head += n;
    __syncwarp();
    if (lane == 0) { ring_fence(patient); *reinterpret_cast<volatile uint32_t*>(&r->head_pub) = head; }
  }
};
#else
struct RingReader {
  RingShared* r;
  size_t head;
  void init(RingShared* ring) { r = ring; head = 0; }
  void init_far(RingShared* ring) { r = ring; head = 0; }
  void init_rec(RingShared* ring) { r = ring; head = 0; }
  uint32_t wait_avail() { return uint32_t(r->e.size() - head); }
  void release(uint32_t n, int) { head += n; }
};
#endif
// Brings the next batch of a ring into reach of lane 0: returns the words to work through (entries that START
// inside them are complete) and points `src` at them.  ALL lanes call this.
AVR_DEV uint32_t coder_fetch(RingReader& q, const uint32_t* far, uint32_t far_cap, uint32_t* stage, int lane, WordSrc* src) {
  const uint32_t avail = q.wait_avail();
  const uint32_t n = avail < RING_BATCH ? avail : uint32_t(RING_BATCH);
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
  if (far) {
    const uint32_t ns = avail < RING_BATCH + ENTRY_MAX_WORDS ? avail : uint32_t(RING_BATCH + ENTRY_MAX_WORDS);
    const volatile uint32_t* g = far;
    for (uint32_t i = uint32_t(lane); i < ns; i += 32) stage[i] = g[(q.head + i) & (far_cap - 1)];
    __syncwarp();
    src->e = stage; src->at = 0; src->mask = 0xFFFFFFFFu;
  } else { src->e = q.ents; src->at = q.head; src->mask = q.mask; }
#else
  (void)far; (void)far_cap; (void)lane;
  for (uint32_t i = 0; i < n + ENTRY_MAX_WORDS; i++) stage[i] = q.head + i < q.r->e.size() ? q.r->e[q.head + i] : uint32_t(SYM_END);
  src->e = stage; src->at = 0; src->mask = 0xFFFFFFFFu;
#endif
  return n;
}

// Compression has TWO recurrences behind the parser: the estimators (serial per estimator) and the 64-bit range
// coder (serial per slice, ~35 instructions of multi-word arithmetic per bin).  One warp doing both was the slowest
// stage of the direction (320-340 cycles per bin against 225 for the parser, profiles/r2e_*), so they are two warps
// joined by a second ring of one RECORD per bin: the estimator as it stood before the bin {pos, total} (bits 0..15),
// the bin (16), REC_FINISH = flush the range coder after this bin (end_of_slice = 1, recode.cpp:1092-1095); SYM_END.
enum : uint32_t { REC_BIT_SHIFT = 16, REC_FINISH = 1u << 17 };
struct RecWriter {     // producer side of the record ring, in registers (lane 0 of the estimator warp)
  RecRing* r;
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
  uint32_t tail, limit;
  AVR_DEV void init(RecRing* ring) { r = ring; tail = 0; limit = RING_PUBLISH; }
  AVR_DEV void put(uint32_t w) {
    if (int32_t(limit - tail) < 1) limit = ring_slow(r, tail, 1, REC_ENTRIES, false);
    r->e[tail & (REC_ENTRIES - 1)] = w;
    tail++;
  }
  AVR_DEV void publish() { ring_fence(); *reinterpret_cast<volatile uint32_t*>(&r->tail_pub) = tail; }
#else
  void init(RecRing* ring) { r = ring; }
  void put(uint32_t w) { r->e.push_back(w); }
  void publish() {}
#endif
};

struct CoderCompress {     // the estimator warp (recode.cpp:1030-1047 + the model keys of 676-815 as indices)
  RecWriter out;
  uint16_t* cnt8;          // estimators of the 8x8 significance slab (CNT8_SPLIT)
  AVR_DEV void begin(RecRing* records, uint16_t* counters8) { out.init(records); cnt8 = counters8; }
  // one bin: the estimator's state goes to the range-coder warp, its update stays here
  AVR_DEV void code(uint16_t* cnt, uint32_t idx, uint32_t bit, uint32_t limit, uint32_t flags = 0) {
    const uint32_t c = cnt[idx];
    out.put(c | bit << REC_BIT_SHIFT | flags);
    cnt[idx] = uint16_t(est_next(c, bit, limit));
  }
  // Significance bins of a 4x4-class block: the estimator index only ever grows inside a block (+1, or +17 after a
  // coefficient), so the next estimator is fetched while the current one is updated.
  AVR_DEV void sig_bins4(uint16_t* cnt, uint32_t slab, uint32_t mask, uint32_t nnz_true, uint32_t coefs) {
    uint32_t idx = slab, zz = 0, observed = 0;
    uint32_t c = cnt[idx];
#pragma unroll 1
    for (;;) {
      const uint32_t bit = (mask >> zz) & 1u;
      const uint32_t idx_n = idx + 1u + (bit << 4);
      const uint32_t c_n = cnt[idx_n];
      out.put(c | bit << REC_BIT_SHIFT);
      cnt[idx] = uint16_t(est_next(c, bit, LIMIT_SIG));
      observed += bit; zz++;
      if (observed >= nnz_true || zz >= coefs) break;
      idx = idx_n; c = c_n;
    }
  }
  // 8x8 blocks: several positions share an estimator (sig_coeff_flag_offset_8x8)
  AVR_DEV void sig_bins8(const SharedTables& t, uint16_t* cnt, uint32_t slab, uint64_t mask, uint32_t nnz_true) {
    uint32_t zz = 0, observed = 0;
    est_prefetch(cnt + slab, nnz_true * (SIG_ZZ * 2));
#pragma unroll 1
    while (observed < nnz_true && zz < 63) {
      const uint32_t bit = uint32_t(mask >> zz) & 1u;
      code(cnt, slab + observed * SIG_ZZ + t.sig8[zz], bit, LIMIT_SIG);
      observed += bit; zz++;
    }
  }
  // nnz of a block as 2 / 4 / 6 model-coded bits, least significant first (recode.cpp:858, 874-888)
  AVR_DEV void nnz_bits(uint16_t* cnt, uint32_t cat, uint32_t w0, uint32_t w1) {
    const uint32_t maxc = k_maxc[cat];
    const uint32_t nbits = maxc > 16 ? 6 : maxc > 4 ? 4 : 2;
    const uint32_t left_nz = w1 & 0xFF, above_nz = (w1 >> 8) & 0xFF, prev_nz = (w1 >> 16) & 0xFF;
    const bool has_left = (w0 >> BLK_HAS_LEFT) & 1;
    const uint32_t nnz = (w0 >> BLK_NNZ) & 63;   // 64 is coded as 0
    const uint32_t group = CNT_NZ + nz_group(int(cat), (w0 >> BLK_MB8) & 1);
#pragma unroll 1
    for (uint32_t i = 0; i < nbits; i++) {
      const uint32_t cur_bit = 1u << i;
      const uint32_t so_far = nnz & (cur_bit - 1);
      const uint32_t lb = has_left ? (left_nz >= cur_bit) : 2;
      const uint32_t ab = above_nz ? (above_nz >= cur_bit) : 2;   // sic: value test (recode.cpp:881)
      code(cnt, group + ((cur_bit - 1 + so_far) * 2 + (prev_nz >= cur_bit)) * 9 + lb * 3 + ab, (nnz >> i) & 1, LIMIT_ANY);
    }
  }
  // one entry; returns true on SYM_END
  AVR_DEV bool entry(WordSrc& s, const SharedTables& t, uint16_t* cnt) {
    const uint32_t e = s.take();
    if (e & (SYM_END | SYM_LEVEL | SYM_BLOCK)) {
      if (e & SYM_END) { out.put(SYM_END); return true; }
      if (e & SYM_BLOCK) {
        const uint32_t cat = e & 7, nnz_true = (e >> BLK_NNZ) & 0x7F;
        const uint32_t w1 = s.take();
        const uint32_t mlo = s.take();
        nnz_bits(cnt, cat, e, w1);
        const uint32_t slab = sig_slab(int(cat), nnz_true & 63);
        if (cat == 5) {
          const uint32_t mhi = s.take();
          sig_bins8(t, cnt8, slab, uint64_t(mlo) | uint64_t(mhi) << 32, nnz_true);
        } else sig_bins4(cnt, slab, mlo, nnz_true, uint32_t(k_maxc[cat]) - 1);
        return false;
      }
      // one coefficient level: first bin, run of "greater than one" bins, sign -- three distinct estimators
      const uint32_t b1 = (e >> LVL_BIT1) & 1, has_sign = (e >> LVL_HAS_SIGN) & 1;
      const uint32_t i1 = CNT_CTX + (e & 0x1FF), g = CNT_CTX + ((e >> LVL_GCTX) & 0x1FF);
      const uint32_t c1 = cnt[i1];
      uint32_t cg = cnt[g];
      const uint32_t cs = cnt[CNT_BYPASS];
      out.put(c1 | b1 << REC_BIT_SHIFT);
      cnt[i1] = uint16_t(est_next(c1, b1, LIMIT_ANY));
      if (b1) {
#pragma unroll 1
        for (uint32_t r = (e >> LVL_RUN) & 15; r > 0; r--) { out.put(cg | 1u << REC_BIT_SHIFT); cg = est_next(cg, 1, LIMIT_ANY); }
        if ((e >> LVL_TERM) & 1) { out.put(cg); cg = est_next(cg, 0, LIMIT_ANY); }
        cnt[g] = uint16_t(cg);
      }
      if (has_sign) {
        const uint32_t sg = (e >> LVL_SIGN) & 1;
        out.put(cs | sg << REC_BIT_SHIFT);
        cnt[CNT_BYPASS] = uint16_t(est_next(cs, sg, LIMIT_ANY));
      }
      return false;
    }
    const uint32_t idx = e & 0xFFFF, bit = (e >> SYM_BIT_SHIFT) & 1;
    code(cnt, idx, bit, LIMIT_ANY, (idx == CNT_TERM && bit) ? uint32_t(REC_FINISH) : 0u);
    return false;
  }
  // Drains a ring until SYM_END.  ALL lanes of the estimator warp call this.  far: nullptr = the ring's own shared-memory
  // entries; else (wide slices) the ring's entries are in global memory at `far`, staged through `stage`, so that this
  // warp may fall a good part of a macroblock row behind the parser and catch up while the parser feeds another stream.
  AVR_DEV void run(RingShared* r, RingCtl* far_ctl, const uint32_t* far, uint32_t far_cap, const SharedTables& t, uint16_t* cnt,
                   uint32_t* stage, int lane) {
    RingReader q;
    if (far_ctl) q.init_far(far_ctl); else q.init(r);
    for (;;) {
      WordSrc src;
      const uint32_t n = coder_fetch(q, far, far_cap, stage, lane, &src);
      uint32_t used = 0, done = 0;
      if (lane == 0) {
        const uint32_t at0 = src.at;
#pragma unroll 1
        while (!done && src.at - at0 < n) done = entry(src, t, cnt) ? 1u : 0u;
        used = src.at - at0;
        out.publish();
      }
      used = AVR_BCAST_CORE(used); done = AVR_BCAST_CORE(done);
      q.release(used, lane);
      if (done) break;
    }
  }
};

// the range-coder warp of a compression (recode.cpp:1074-1075 -> arithmetic_code.h:106-126): one record per bin
struct RangeWarp {
  RangeEnc rc;
  AVR_DEV void begin(ByteSink* sink, uint8_t* dst, uint32_t cap) { rc.init(sink, dst, cap); }
  AVR_DEV void run(RecRing* records, const SharedTables& t, uint32_t* stage, int lane) {
    RingReader q; q.init_rec(records);
    for (;;) {
      WordSrc src;
      const uint32_t n = coder_fetch(q, nullptr, 0, stage, lane, &src);
      uint32_t done = 0;
      if (lane == 0) {
        // records never change once written: record k + 2 and the reciprocal of record k + 1 are fetched while record k
        // goes through the coder, so that no load sits on the range recurrence
        uint32_t rec = src.take();
        uint32_t rec1 = n > 1 ? src.take() : uint32_t(SYM_END);
        DivMagic mg = t.div[(rec >> 8) & 0xFF];
#pragma unroll 2
        for (uint32_t k = 2;; k++) {
          if (rec & SYM_END) { done = 1; break; }
          const uint32_t rec2 = k < n ? src.take() : uint32_t(SYM_END);
          const DivMagic mg1 = t.div[(rec1 >> 8) & 0xFF];
          rc.put(mg, int((rec >> REC_BIT_SHIFT) & 1), rec & 0xFF);
          if (rec & REC_FINISH) range_enc_finish(rc.sink, rc.low, rc.range);
          if (k > n) break;
          rec = rec1; rec1 = rec2; mg = mg1;
        }
      }
      done = AVR_BCAST_CORE(done);
      q.release(n, lane);
      if (done) break;
    }
  }
  AVR_DEV uint32_t out_bytes() const { return rc.sink->n; }
  AVR_DEV bool overflowed() const { return rc.sink->overflow != 0; }
};

// Decompression behind the parser stays ONE warp (context states + CABAC arithmetic).  Tried and dropped
// (profiles/r2o_ab_decompress_split.txt): splitting it like the compressor's coder into a context warp and an arithmetic
// warp joined by the record ring.  The parser is the slow stage of this direction, so a lone slice gained nothing
// (409.6 -> 408.4 ms on the profile case), and with every SM full the third warp's instructions cost 21 % of the
// throughput (1728 slices: 999 -> 1209 ms).
struct CoderDecompress {   // CABAC re-encoder (cabac_code.h:16-86; recode.cpp:1443, 1455, 1466)
  CabacOut cabac;
  AVR_DEV void begin(ByteSink* sink, uint8_t* dst, uint32_t cap, int esc_zeros = -1) { cabac.init(sink, dst, cap, esc_zeros); }
  // significant_coeff_flag / last_significant_coeff_flag bins of a block out of its mask (h264_cabac.c:1667-1703);
  // the last coefficient of a full scan is inferred, not coded.  4x4-class blocks: position i uses contexts
  // sig_base + i and last_base + i, all distinct, so the entries of the next position are fetched ahead.
  AVR_DEV void sig_bins4(const SharedTables& t, U2* sig, U2* last, uint32_t mask, uint32_t nnz_true, uint32_t coefs) {
    uint32_t zz = 0, observed = 0;
    U2 es = sig[0], el = last[0];
#pragma unroll 1
    for (;;) {
      const uint32_t bit = (mask >> zz) & 1u;
      const U2 es_n = sig[zz + 1], el_n = last[zz + 1];   // one past the block at the end: loaded, never used
      U2 n;
      cabac.encode_core(t, es, int(bit), &n);
      sig[zz] = n;
      if (bit) {
        observed++;
        cabac.encode_core(t, el, observed == nnz_true, &n);
        last[zz] = n;
      }
      zz++;
      if (observed >= nnz_true || zz >= coefs) break;
      es = es_n; el = el_n;
    }
  }
  AVR_DEV void sig_bins8(const SharedTables& t, U2* sig, U2* last, uint64_t mask, uint32_t nnz_true) {
    uint32_t zz = 0, observed = 0;
#pragma unroll 1
    while (observed < nnz_true && zz < 63) {
      const uint32_t bit = uint32_t(mask >> zz) & 1u;
      cabac.decision(t, int(bit), &sig[t.sig8[zz]]);
      if (bit) { observed++; cabac.decision(t, observed == nnz_true, &last[t.last8[zz]]); }
      zz++;
    }
  }
  // everything but SYM_END and the wide-slice specials
  AVR_DEV void entry(WordSrc& s, const SharedTables& t, U2* cstate, uint32_t e) {
    if (e & (SYM_LEVEL | SYM_BLOCK)) {
      if (e & SYM_BLOCK) {
        const uint32_t cat = e & 7, nnz_true = (e >> BLK_NNZ) & 0x7F;
        const uint32_t mlo = s.take();
        U2* sig = cstate + k_sig_base[cat];
        U2* last = cstate + k_last_base[cat];
        if (cat == 5) {
          const uint32_t mhi = s.take();
          sig_bins8(t, sig, last, uint64_t(mlo) | uint64_t(mhi) << 32, nnz_true);
        } else sig_bins4(t, sig, last, mlo, nnz_true, uint32_t(k_maxc[cat]) - 1);
        return;
      }
      // one coefficient level: first bin, run of "greater than one" bins, sign
      const uint32_t g = (e >> LVL_GCTX) & 0x1FF;
      const int b1 = int((e >> LVL_BIT1) & 1);
      U2 eg = cstate[g];                      // both contexts are loaded up front (they differ); the run keeps its
      cabac.decision(t, b1, &cstate[e & 0x1FF]);   // context entry in registers and writes it back once
      if (b1) {
#pragma unroll 1
        for (uint32_t r = (e >> LVL_RUN) & 15; r > 0; r--) cabac.encode_core(t, eg, 1, &eg);
        if ((e >> LVL_TERM) & 1) cabac.encode_core(t, eg, 0, &eg);
        cstate[g] = eg;
      }
      if ((e >> LVL_HAS_SIGN) & 1) cabac.bypass(int((e >> LVL_SIGN) & 1));
      return;
    }
    const int bit = int((e >> SYM_BIT_SHIFT) & 1);
    const uint32_t kind = e & (3u << SYM_KIND_SHIFT);
    if (kind == SYM_DECISION) cabac.decision(t, bit, &cstate[e & 0xFFFF]);
    else if (kind == SYM_BYPASS) cabac.bypass(bit);
    else cabac.terminate(bit);
  }
  // Ordinary slices: drains the shared-memory ring until SYM_END.  ALL lanes call this.
  AVR_DEV void run(RingShared* r, const SharedTables& t, U2* cstate, uint32_t* stage, int lane) {
    RingReader q; q.init(r);
    for (;;) {
      WordSrc src;
      const uint32_t n = coder_fetch(q, nullptr, 0, stage, lane, &src);
      uint32_t used = 0, done = 0;
      if (lane == 0) {
        const uint32_t at0 = src.at;
#pragma unroll 1
        while (src.at - at0 < n) {
          const uint32_t e = src.take();
          if (e & SYM_END) { done = 1; break; }
          entry(src, t, cstate, e);
        }
        used = src.at - at0;
      }
      used = AVR_BCAST_CORE(used); done = AVR_BCAST_CORE(done);
      q.release(used, lane);
      if (done) break;
    }
  }
  // Wide slices: ring j (entries in global memory, far[j * far_cap ...]; indices in rings[j]) carries the
  // macroblock rows j, j + K, ... and every row ends in SP_ROW_END.  ALL lanes of the warp call this: they
  // stage a batch of words with coalesced loads, lane 0 encodes the entries in it.  Leaves on the
  // first SYM_END and then drains the other rings so that their parser warps can finish.
  AVR_DEV void run_wide(RingCtl* rings, uint32_t n_rings, const uint32_t* far, uint32_t far_cap, const SharedTables& t, U2* cstate,
                        uint32_t* stage, int lane) {
    uint32_t cur = 0, qpd = 0;
    bool draining = false;
    RingReader q; q.init_far(&rings[0]);
    for (;;) {
      WordSrc src;
      const uint32_t n = coder_fetch(q, far ? far + size_t(cur) * far_cap : nullptr, far_cap, stage, lane, &src);
      uint32_t used = 0, why = 0;   // why: 0 go on, 1 row end, 2 SYM_END
      if (lane == 0) {   // the serial part: plain single-lane code, nothing warp-wide per entry
        const uint32_t at0 = src.at;
#pragma unroll 1
        while (src.at - at0 < n) {
          const uint32_t e = src.take();
          if (e & SYM_END) { why = 2; break; }
          if (draining) { if (e & SYM_BLOCK) src.at += (e & 7) == 5 ? 2 : 1; continue; }   // skip a block's mask words
          if (!(e & (SYM_LEVEL | SYM_BLOCK)) && (e & (3u << SYM_KIND_SHIFT)) == SYM_SPECIAL) {
            if (e & SP_ROW_END) { qpd = e & 1; why = 1; break; }
            cabac.decision(t, int((e >> SYM_BIT_SHIFT) & 1), &cstate[60 + qpd]);
          } else entry(src, t, cstate, e);
        }
        used = src.at - at0;
      }
      used = AVR_BCAST_CORE(used); why = AVR_BCAST_CORE(why);
      q.release(used, lane);
      if (why == 0) continue;
      if (why == 2) { draining = true; q.r->cons_head = 0xFFFFFFFFu; }
      else q.r->cons_head = uint32_t(q.head);
      AVR_SYNCWARP_CORE();
      // next ring: the next row's, or (draining) the next one that has not delivered SYM_END yet
      uint32_t tries = 0;
      do { cur = cur + 1 == n_rings ? 0 : cur + 1; tries++; } while (draining && tries <= n_rings && rings[cur].cons_head == 0xFFFFFFFFu);
      if (draining && tries > n_rings) break;
      q.r = &rings[cur];
      q.head = uint32_t(rings[cur].cons_head);
    }
  }
  AVR_DEV uint32_t out_bytes() const { return cabac.sink->n; }
  AVR_DEV bool overflowed() const { return cabac.sink->overflow != 0; }
};

// ------------------------------------------------------------------------------------------------
// Where a parser warp keeps its engines BETWEEN residual blocks: shared memory.  Macroblock-level syntax
// (a tenth of the bins, but 57 call sites in the walker) goes through ONE out-of-line copy of each bin
// primitive that works on this state; Walker::residual() loads the state into registers (enter), runs
// its loops with everything inlined, and stores it back (leave).  With every primitive inlined at
// every site the decompression kernels were 230-250 KB of SASS and stalled one cycle per issued
// instruction on instruction fetch (profiles/r1k_ncu_full_summary.txt).
// ------------------------------------------------------------------------------------------------
// Per-class bin counts of a slice (the reference's per-CodingType bills, recode.cpp:635-655, as counts): the fast
// regions of the walker keep them in registers and add them to the home state when they leave.
enum : int { HIST_MB = 0, HIST_CBF = 1, HIST_SIG = 2, HIST_LEVEL = 3, HIST_BYPASS = 4, HIST_NNZ = 5, HIST_PRED = 6, HIST_TERM = 7 };
struct ResBins {
  uint32_t cbf, sig, level, bypass, nnz, pred, mb;   // mb: macroblock-level bins walked in register mode (HIST_MB is the remainder)
  AVR_DEV void clear() { cbf = sig = level = bypass = nnz = pred = mb = 0; }
  AVR_DEV uint32_t cabac_bins() const { return cbf + sig + level + bypass + pred + mb; }   // nnz bits are model-only bins
};
struct ParserHome {
  RingCtl* r;               // ring, producer side: its control words
  uint32_t* ents;           //   its entries: the block's shared-memory ring, or (wide slices) a much longer one in global memory
  uint32_t* far_base;       //   wide slices: ring j's entries live at far_base + j * cap, its control words at far_ctl[j]
  RingCtl* far_ctl;
  uint32_t cap, tail, limit;
  const SharedTables* tab;
  U2* cstate;
  uint16_t* cnt;
  const uint8_t* win;       // input window
  uint32_t skew;
  uint32_t val, range, pos; // compress: CabacIn
  int32_t nbits;
  uint64_t low, drange;     // decompress: RangeDec (its read position: dec)
  DecHome dec;
  uint32_t n_bins;          // bins walked so far (low word)
  uint32_t hist[8];         // per class, HIST_*: [HIST_MB] and [HIST_TERM] are filled in at the end of the slice
};
AVR_DEV void home_add_bins(ParserHome* h, const ResBins& b) {
  h->n_bins += b.cabac_bins();
  h->hist[HIST_CBF] += b.cbf; h->hist[HIST_SIG] += b.sig; h->hist[HIST_LEVEL] += b.level;
  h->hist[HIST_BYPASS] += b.bypass; h->hist[HIST_NNZ] += b.nnz; h->hist[HIST_PRED] += b.pred;
}

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
// ordinary slices: the block's own ring.  Wide slices: ring `index` of the slice's rings in global memory.
AVR_DEV void home_ring_init(ParserHome* h, RingShared* ring, RingCtl* far_ctl, uint32_t* far_entries, uint32_t far_cap, uint32_t index) {
  h->tail = 0; h->limit = RING_PUBLISH;
  h->far_base = far_entries; h->far_ctl = far_ctl;
  if (far_ctl) { h->r = &far_ctl[index]; h->ents = far_entries + size_t(index) * far_cap; h->cap = far_cap; }
  else { h->r = ring; h->ents = ring->e; h->cap = RING_ENTRIES; }
}
AVR_DEV void home_push(ParserHome* h, uint32_t e) {
  uint32_t tail = h->tail;
  if (int32_t(h->limit - tail) < 1) h->limit = ring_slow(h->r, tail, 1, h->cap, h->far_ctl != nullptr);
  h->ents[tail & (h->cap - 1)] = e;
  h->tail = tail + 1;
}
AVR_DEV void home_publish(ParserHome* h) {
  ring_fence(h->far_ctl != nullptr);
  *reinterpret_cast<volatile uint32_t*>(&h->r->tail_pub) = h->tail;
}
AVR_DEV void home_switch_to(ParserHome* h, uint32_t index) {   // wide compression: the next macroblock row goes to another ring
  home_publish(h);
  h->r->prod_tail = h->tail;
  RingCtl* next = &h->far_ctl[index];
  h->r = next;
  h->ents = h->far_base + size_t(index) * h->cap;
  h->tail = next->prod_tail;
  h->limit = ring_slow(next, h->tail, ENTRY_MAX_WORDS, h->cap, true);
}
#else
inline void home_ring_init(ParserHome* h, RingShared* ring, RingCtl* far_ctl, uint32_t*, uint32_t, uint32_t index) {
  h->far_ctl = far_ctl;
  h->r = far_ctl ? &far_ctl[index] : ring;
}
inline void home_push(ParserHome* h, uint32_t e) { h->r->e.push_back(e); }
inline void home_publish(ParserHome*) {}
inline void home_switch_to(ParserHome* h, uint32_t index) { h->r = &h->far_ctl[index]; }
#endif

// register copy of the ring's producer side, valid between enter() and leave()
struct RingWriter {
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
  uint32_t* ents;
  uint32_t cap, tail, limit;
  AVR_DEV void enter(const ParserHome* h) { ents = h->ents; cap = h->cap; tail = h->tail; limit = h->limit; }
  AVR_DEV void leave(ParserHome* h) const { h->tail = tail; h->limit = limit; }
  AVR_DEV void reserve(ParserHome* h, uint32_t n) { if (int32_t(limit - tail) < int32_t(n)) limit = ring_slow(h->r, tail, n, cap, h->far_ctl != nullptr); }
  AVR_DEV void put(ParserHome*, uint32_t w) { ents[tail & (cap - 1)] = w; tail++; }
#else
  void enter(const ParserHome*) {}
  void leave(ParserHome*) const {}
  void reserve(ParserHome*, uint32_t) {}
  void put(ParserHome* h, uint32_t w) { h->r->e.push_back(w); }
#endif
};

// ---- the out-of-line bin primitives of macroblock-level syntax --------------------------------------
// What a call costs is its chain of dependent memory operations, so: the caller passes the addresses it already has
// in registers (context entry / estimator, tables), and every field of the home state is read up front, before the
// first store (a store through one of these generic pointers could alias any of them as far as the compiler knows).
AVR_DEV void home_load(const ParserHome* h, CabacIn& c, ByteWindow& w) {
  c.val = h->val; c.range = h->range; c.pos = h->pos; c.nbits = h->nbits;
  w.win = const_cast<uint8_t*>(h->win); w.skew = h->skew;
}
AVR_DEV void home_store(ParserHome* h, const CabacIn& c) { h->val = c.val; h->range = c.range; h->pos = c.pos; h->nbits = c.nbits; }
struct HomeRing {   // the producer side of the ring, read in one go
  RingCtl* r; uint32_t* ents; uint32_t cap, tail, limit, n_bins; bool far;
  AVR_DEV void load(const ParserHome* h) { r = h->r; ents = h->ents; cap = h->cap; tail = h->tail; limit = h->limit; n_bins = h->n_bins; far = h->far_ctl != nullptr; }
  AVR_DEV void push_and_store(ParserHome* h, uint32_t e) {
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
    if (int32_t(limit - tail) < 1) { limit = ring_slow(r, tail, 1, cap, far); h->limit = limit; }
    ents[tail & (cap - 1)] = e;
    h->tail = tail + 1;
#else
    h->r->e.push_back(e);
#endif
    h->n_bins = n_bins + 1;
  }
};
// compress: CABAC decision on the context entry at `ctx`, modelled with estimator `model_ctx`
AVR_DEV_NOINLINE uint32_t slow_c_bin(ParserHome* h, U2* ctx, const SharedTables* tab, uint32_t model_ctx) {
  const U2 e = *ctx;
  CabacIn c; ByteWindow w; HomeRing q;
  home_load(h, c, w);
  q.load(h);
  U2 n;
  const int b = c.decide_core(w, *tab, e, &n);
  *ctx = n;
  home_store(h, c);
  q.push_and_store(h, uint32_t(CNT_CTX + model_ctx) | uint32_t(b) << SYM_BIT_SHIFT);
  return uint32_t(b);
}
AVR_DEV_NOINLINE uint32_t slow_c_bypass(ParserHome* h) {
  CabacIn c; ByteWindow w; HomeRing q;
  home_load(h, c, w);
  q.load(h);
  const int b = c.bypass(w);
  home_store(h, c);
  q.push_and_store(h, CNT_BYPASS | uint32_t(b) << SYM_BIT_SHIFT);
  return uint32_t(b);
}
AVR_DEV_NOINLINE uint32_t slow_c_terminate(ParserHome* h) {
  CabacIn c; ByteWindow w; HomeRing q;
  home_load(h, c, w);
  q.load(h);
  const int b = c.terminate(w);
  home_store(h, c);
  q.push_and_store(h, CNT_TERM | uint32_t(b) << SYM_BIT_SHIFT);
  return uint32_t(b);
}
// decompress: one model-coded bin with the estimator at `est`; `entry` (without the bin) goes to the CABAC re-encoder
AVR_DEV_NOINLINE uint32_t slow_d_bin(ParserHome* h, uint16_t* est, const SharedTables* tab, uint32_t entry) {
  const uint32_t c = *est;
  RangeDec rc; rc.low = h->low; rc.range = h->drange; rc.home = &h->dec;
  HomeRing q;
  q.load(h);
  const DivMagic mg = tab->div[c >> 8];
  const int b = rc.get(mg, c & 0xFF);
  *est = uint16_t(est_next(c, uint32_t(b), LIMIT_ANY));
  h->low = rc.low; h->drange = rc.range;
  q.push_and_store(h, entry | uint32_t(b) << SYM_BIT_SHIFT);
  return uint32_t(b);
}

// ------------------------------------------------------------------------------------------------
// Parser-warp direction policies (lane 0).  bin / bypass / terminate / bin_row_qpd / row_end work on
// the home state (macroblock-level syntax); everything else is register-resident and only valid
// between enter() and leave() (Walker::residual).
// ------------------------------------------------------------------------------------------------
struct Compress {
  static constexpr bool kDecompress = false;
  typedef CoderCompress Coder;
  typedef RangeWarp Tail;
  ByteWindow in;
  ParserHome* h;
  U2* cstate;        // parser owns the CABAC context states
  const SharedTables* tabp;
  CabacIn cabac;     // register copy
  RingWriter q;
  AVR_DEV void set_cnt8(uint16_t*) {}   // the parser of a compression touches no estimator
  AVR_DEV void begin(ParserHome* home, const uint8_t* src, uint32_t n, uint8_t* window, uint32_t tail_byte, RingShared* ring,
                     const SharedTables* tab, uint16_t* counters, U2* states, RingCtl* far_ctl = nullptr, uint32_t* far_entries = nullptr,
                     uint32_t far_cap = 0, uint32_t far_index = 0) {
    h = home;
    in.attach(src, n, window, tail_byte);
    cstate = states; tabp = tab;
    home_ring_init(h, ring, far_ctl, far_entries, far_cap, far_index);
    h->tab = tab; h->cstate = states; h->cnt = counters; h->win = window; h->skew = in.skew; h->n_bins = 0;
    for (int i = 0; i < 8; i++) h->hist[i] = 0;
  }
  AVR_DEV void start_bins(const SharedTables&) { CabacIn c; c.init(in); home_store(h, c); }
  AVR_DEV void finish_bins() {}
  // ---- home state ----
  AVR_DEV int bin(int ctx) { return int(slow_c_bin(h, &cstate[ctx], tabp, uint32_t(ctx))); }
  // wide slices: first mb_qp_delta bin of a macroblock row -- decoded with its true context, modelled with
  // the estimator of context 60 whatever the previous row ended on (the rows are decoded in parallel)
  AVR_DEV int bin_row_qpd(int true_ctx) { return int(slow_c_bin(h, &cstate[true_ctx], tabp, 60)); }
  AVR_DEV void row_end(int) {}
  AVR_DEV int bypass() { return int(slow_c_bypass(h)); }
  AVR_DEV int terminate() { return int(slow_c_terminate(h)); }
  AVR_DEV uint32_t in_pos() const { return h->pos; }
  // After end_of_slice = 1 the last bit in the arithmetic decoder's offset register is rbsp_stop_one_bit (H.264
  // 9.3.4.5 writes it as the low bit of the flush), so the bits taken so far minus that one are exactly the payload
  // -- the byte of a stop bit that sits alone being outside it (h264.c:1298-1315).  A slice that ends anywhere
  // else was mis-located or carries junk: it would decompress to another size and lose the whole file, so it is
  // reported corrupt and stays a literal (recode.cpp:1289-1296).
  AVR_DEV bool consumed_all() const { return (h->pos * 8u - uint32_t(h->nbits) + 6u) / 8u == in.size; }
  // ---- registers ----
  AVR_DEV void enter(const SharedTables&) { ByteWindow w; home_load(h, cabac, w); q.enter(h); }
  AVR_DEV void leave(const ResBins& bins) { home_store(h, cabac); q.leave(h); home_add_bins(h, bins); }
  AVR_DEV int fast_terminate(const SharedTables&) {   // end_of_slice / I_PCM bin in register mode
    const int b = cabac.terminate(in);
    q.reserve(h, 1); q.put(h, CNT_TERM | uint32_t(b) << SYM_BIT_SHIFT);
    return b;
  }
  AVR_DEV int fast_bin(const SharedTables& t, int ctx) {   // an ordinary bin in register mode
    const int b = cabac.decision(in, t, &cstate[ctx]);
    q.reserve(h, 1); q.put(h, uint32_t(CNT_CTX + ctx) | uint32_t(b) << SYM_BIT_SHIFT);
    return b;
  }
  // coefficient levels: CABAC side only, the whole level crosses the ring as one entry (lvl_push)
  typedef U2 BinEntry;
  AVR_DEV BinEntry bin_fetch(const SharedTables&, int ctx) const { return cstate[ctx]; }
  AVR_DEV void bin_store(const BinEntry& e, int ctx) { cstate[ctx] = e; }
  AVR_DEV int lvl_first(const SharedTables& t, const BinEntry e, int ctx) { return cabac.decide(in, t, e, &cstate[ctx]); }
  AVR_DEV int lvl_run(const SharedTables& t, BinEntry& e, int) {
    BinEntry n;
    const int b = cabac.decide_core(in, t, e, &n);
    e = n;
    return b;
  }
  AVR_DEV int lvl_sign(const SharedTables&) { return cabac.bypass(in); }
  AVR_DEV void lvl_push(uint32_t entry) { q.reserve(h, 1); q.put(h, entry); }
  AVR_DEV int lvl_bypass(const SharedTables&) {   // Exp-Golomb escape bins
    const int b = cabac.bypass(in);
    q.reserve(h, 1); q.put(h, CNT_BYPASS | uint32_t(b) << SYM_BIT_SHIFT);
    return b;
  }
  // significance map: CABAC side (the model side is derived from the block entry by the coder warp)
  AVR_DEV U2 sig_peek(int ctx) const { return cstate[ctx]; }
  AVR_DEV int sig_get(const SharedTables& t, int ctx) { return cabac.decision(in, t, &cstate[ctx]); }
  AVR_DEV int sig_get_pre(const SharedTables& t, const U2 e, int ctx) { return cabac.decide(in, t, e, &cstate[ctx]); }
  // nnz: count of this block; left / above / previous-picture counts of the NZ key; mb8: the macroblock has an 8x8 block
  AVR_DEV void block_push(int cat, uint32_t nnz, bool mb8, bool has_left, uint32_t left_nz, uint32_t above_nz, uint32_t prev_nz, uint64_t mask) {
    q.reserve(h, 4);
    q.put(h, SYM_BLOCK | uint32_t(cat) | nnz << BLK_NNZ | (mb8 ? 1u << BLK_MB8 : 0u) | (has_left ? 1u << BLK_HAS_LEFT : 0u));
    q.put(h, left_nz | above_nz << 8 | prev_nz << 16);
    q.put(h, uint32_t(mask));
    if (cat == 5) q.put(h, uint32_t(mask >> 32));
  }
};

struct Decompress {
  static constexpr bool kDecompress = true;
  typedef CoderDecompress Coder;
  ByteWindow in;
  ParserHome* h;
  uint16_t* cnt;     // parser owns the estimators
  uint16_t* cnt8;    // ... those of the 8x8 significance slab (CNT8_SPLIT)
  const SharedTables* tabp;
  RangeDec rc;       // register copy
  RingWriter q;
  // The bypass estimator codes every sign and escape bit: inside a residual block it is kept in registers
  // together with its reciprocal.
  uint32_t byp_c;
  DivMagic byp_mg;
  AVR_DEV void begin(ParserHome* home, const uint8_t* src, uint32_t n, uint8_t* window, uint32_t, RingShared* ring,
                     const SharedTables* tab, uint16_t* counters, U2* states, RingCtl* far_ctl = nullptr, uint32_t* far_entries = nullptr,
                     uint32_t far_cap = 0, uint32_t far_index = 0) {
    h = home;
    in.attach(src, n, window, 0);
    cnt = counters; cnt8 = counters; tabp = tab;
    home_ring_init(h, ring, far_ctl, far_entries, far_cap, far_index);
    h->tab = tab; h->cstate = states; h->cnt = counters; h->win = window; h->skew = in.skew; h->n_bins = 0;
    for (int i = 0; i < 8; i++) h->hist[i] = 0;
  }
  AVR_DEV void start_bins(const SharedTables&) {
    RangeDec d; d.init(&h->dec, in);
    h->low = d.low; h->drange = d.range;
  }
  AVR_DEV void finish_bins() {}
  // ---- home state ----
  AVR_DEV int bin(int ctx) { return int(slow_d_bin(h, &cnt[CNT_CTX + ctx], tabp, uint32_t(ctx) | SYM_DECISION)); }
  AVR_DEV int bin_row_qpd(int) { return int(slow_d_bin(h, &cnt[CNT_CTX + 60], tabp, SYM_SPECIAL | SP_QPD)); }   // see Compress::bin_row_qpd
  AVR_DEV void row_end(int qpd_nonzero) { home_push(h, SYM_SPECIAL | SP_ROW_END | uint32_t(qpd_nonzero)); }
  AVR_DEV int bypass() { return int(slow_d_bin(h, &cnt[CNT_BYPASS], tabp, SYM_BYPASS)); }
  AVR_DEV int terminate() { return int(slow_d_bin(h, &cnt[CNT_TERM], tabp, SYM_TERMINATE)); }
  AVR_DEV uint32_t in_pos() const { return h->dec.pos; }
  AVR_DEV bool consumed_all() const { return true; }   // the rebuilt size is checked by the host against Block.size
  // ---- registers ----
  AVR_DEV void enter(const SharedTables& t) {
    rc.low = h->low; rc.range = h->drange; rc.home = &h->dec;
    q.enter(h);
    byp_c = cnt[CNT_BYPASS];
    byp_mg = t.div[byp_c >> 8];
  }
  AVR_DEV void leave(const ResBins& bins) {
    h->low = rc.low; h->drange = rc.range;
    q.leave(h);
    cnt[CNT_BYPASS] = uint16_t(byp_c);
    home_add_bins(h, bins);
  }
  AVR_DEV int coded(const SharedTables& t, uint32_t idx, uint32_t limit) {
    const uint32_t c = cnt[idx];
    const int b = rc.get(t.div[c >> 8], c & 0xFF);
    cnt[idx] = uint16_t(est_next(c, uint32_t(b), limit));
    return b;
  }
  AVR_DEV int fast_terminate(const SharedTables& t) {   // end_of_slice / I_PCM bin in register mode
    const int b = coded(t, CNT_TERM, LIMIT_ANY);
    q.reserve(h, 1); q.put(h, uint32_t(b) << SYM_BIT_SHIFT | SYM_TERMINATE);
    return b;
  }
  AVR_DEV int fast_bin(const SharedTables& t, int ctx) {   // an ordinary bin in register mode
    const int b = coded(t, uint32_t(CNT_CTX + ctx), LIMIT_ANY);
    q.reserve(h, 1); q.put(h, uint32_t(ctx) | uint32_t(b) << SYM_BIT_SHIFT | SYM_DECISION);
    return b;
  }
  AVR_DEV int byp_bin(const SharedTables& t) {
    const int b = rc.get(byp_mg, byp_c & 0xFF);
    byp_c = est_next(byp_c, uint32_t(b), LIMIT_ANY);
    byp_mg = t.div[byp_c >> 8];
    return b;
  }
  // significance-map bins with both possible successors prefetched (see Walker::residual)
  struct SigEntry { uint32_t c; DivMagic mg; };
  AVR_DEV SigEntry sig_fetch(const SharedTables& t, uint32_t idx) const {
    SigEntry e; e.c = cnt[idx]; e.mg = t.div[e.c >> 8]; return e;
  }
  AVR_DEV int sig_decode(const SigEntry& e, uint32_t idx) {
    const int b = rc.get(e.mg, e.c & 0xFF);
    cnt[idx] = uint16_t(est_next(e.c, uint32_t(b), LIMIT_SIG));
    return b;
  }
  // an ordinary bin whose estimator + reciprocal were fetched ahead of time (no update of it in between)
  typedef SigEntry BinEntry;
  AVR_DEV BinEntry bin_fetch(const SharedTables& t, int ctx) const { return sig_fetch(t, uint32_t(CNT_CTX + ctx)); }
  AVR_DEV void bin_store(const BinEntry& e, int ctx) { cnt[CNT_CTX + ctx] = uint16_t(e.c); }
  // coefficient levels: no per-bin push; lvl_push sends the whole level at once
  AVR_DEV int lvl_first(const SharedTables&, const BinEntry& e, int ctx) {
    const int b = rc.get(e.mg, e.c & 0xFF);
    cnt[CNT_CTX + ctx] = uint16_t(est_next(e.c, uint32_t(b), LIMIT_ANY));
    return b;
  }
  // a run of bins on ONE context: estimator and reciprocal stay in registers; the next reciprocal is
  // div[total + 1], known before the bin is, unless the update halves the counts (rare: reloaded)
  AVR_DEV int lvl_run(const SharedTables& t, BinEntry& e, int) {
    const uint32_t tot = e.c >> 8;
    const DivMagic mg_next = t.div[tot + 1];
    const int b = rc.get(e.mg, e.c & 0xFF);
    e.c = est_next(e.c, uint32_t(b), LIMIT_ANY);
    e.mg = mg_next;
    if (tot + 1 > LIMIT_ANY) e.mg = t.div[e.c >> 8];
    return b;
  }
  AVR_DEV int lvl_sign(const SharedTables& t) { return byp_bin(t); }
  AVR_DEV void lvl_push(uint32_t entry) { q.reserve(h, 1); q.put(h, entry); }
  AVR_DEV int lvl_bypass(const SharedTables& t) {   // Exp-Golomb escape bins
    const int b = byp_bin(t);
    q.reserve(h, 1); q.put(h, uint32_t(b) << SYM_BIT_SHIFT | SYM_BYPASS);
    return b;
  }
  AVR_DEV void set_cnt8(uint16_t* p) { cnt8 = p; }
  AVR_DEV void sig8_prefetch(uint32_t slab, uint32_t rows) const { est_prefetch(cnt8 + slab, rows * (SIG_ZZ * 2)); }
  AVR_DEV int sig8_bit(const SharedTables& t, uint32_t idx) {   // a significance bin of an 8x8 block
    const uint32_t c = cnt8[idx];
    const int b = rc.get(t.div[c >> 8], c & 0xFF);
    cnt8[idx] = uint16_t(est_next(c, uint32_t(b), LIMIT_SIG));
    return b;
  }
  AVR_DEV int model_bit(const SharedTables& t, uint32_t idx) { return coded(t, idx, idx >= CNT_SIG ? uint32_t(LIMIT_SIG) : uint32_t(LIMIT_ANY)); }
  // the block's significance map for the CABAC re-encoder
  AVR_DEV void block_push(int cat, uint32_t nnz_true, uint64_t mask) {
    q.reserve(h, 3);
    q.put(h, SYM_BLOCK | uint32_t(cat) | nnz_true << BLK_NNZ);
    q.put(h, uint32_t(mask));
    if (cat == 5) q.put(h, uint32_t(mask >> 32));
  }
};

}  // namespace avr
