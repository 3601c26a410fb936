// avr_oracle.hpp -- CPU restatement of avrecode's CABAC decode -> model -> re-encode path.
//
// *** TEST INFRASTRUCTURE ONLY. ***  Nothing under avrecode_b200/ (the product) may include,
// link or call this.  It exists so tests/, bench.py's cpu_baseline leg and
// __graft_entry__.smoke() can check the CUDA path bit for bit.
//
// Parity status: PINNED.  This restatement is checked byte for byte against the unmodified
// reference binary (oracle/_ref/recode_ref*, built by oracle/build_ref.sh from /root/reference)
// on every stream in tests/golden/ and on freshly generated streams (tests/test_oracle_pin.py).
//
// Every function cites the reference file:line it follows (paths relative to /root/reference).
// Scope: frame macroblocks, 4:2:0, 8-bit, I/P/B slices, 4x4 and 8x8 transform.  MBAFF / field
// pictures / I_PCM / 4:2:2 / 4:4:4 are reported as AVR_O_UNSUPPORTED.
#pragma once
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#include <map>
#include <memory>
#include <tuple>
#include <vector>
#include <stdexcept>
#include <string>
#include "../avrecode_b200/csrc/avr_tables.h"
#include "../include/avrecode_b200.h"

namespace avr_oracle {

struct Unsupported : std::runtime_error { using std::runtime_error::runtime_error; };
struct Corrupt : std::runtime_error { using std::runtime_error::runtime_error; };

// ---------------------------------------------------------------------------------------------
// Generic binary arithmetic coder -- restates arithmetic_code.h:31-299.
// Fixed = fixed point word, DigitBytes = sizeof(CompressedDigit); output/input digits are bytes.
// ---------------------------------------------------------------------------------------------
template <typename Fixed, int DigitBytes, uint64_t MinRangeOverride>
struct ArithParams {
  static constexpr Fixed fixed_one = (Fixed(~Fixed(0)) >> 1) + 1;                 // arithmetic_code.h:54-55
  static constexpr Fixed digit_base = Fixed(1) << (8 * DigitBytes);               // :57
  static constexpr Fixed min_range =
      MinRangeOverride ? Fixed(MinRangeOverride) : (fixed_one / digit_base) / 16;  // :61-62
  static constexpr Fixed max_range = fixed_one;                                   // :64
};

template <typename Fixed, int DigitBytes, uint64_t MinRangeOverride = 0>
class ArithEncoder {
  typedef ArithParams<Fixed, DigitBytes, MinRangeOverride> P;

 public:
  explicit ArithEncoder(std::vector<uint8_t>* out, Fixed initial_range = P::fixed_one)
      : out_(out), low_(0), range_(initial_range) {}
  Fixed range() const { return range_; }

  // arithmetic_code.h:106-126.  range_of_1 = probability_of_1(range) is computed by the caller.
  void put(int symbol, Fixed range_of_1) {
    Fixed range_of_0 = range_ - range_of_1;
    if (symbol) { low_ += range_of_0; range_ = range_of_1; } else { range_ = range_of_0; }
    if (range_ < P::min_range) {
      if (range_ == 0) throw Corrupt("encoder: zero-probability symbol");
      while (range_ < P::max_range / P::digit_base) renorm_emit(DigitBytes);
    }
  }
  // arithmetic_code.h:128-144
  void finish() {
    for (Fixed stop = P::fixed_one >> 1; stop > 0; stop >>= 1) {
      Fixed x = (low_ | stop) & ~(stop - 1);
      if (stop < range_ && low_ <= x && x < low_ + range_) { low_ = x; break; }
    }
    while (low_ != 0) { range_ = 1; renorm_emit(1); }
    range_ = 0;
  }

 private:
  // arithmetic_code.h:147-180; nbytes = sizeof(Digit) of this call (CompressedDigit or OutputDigit)
  void renorm_emit(int nbytes) {
    const Fixed base = Fixed(1) << (8 * nbytes);
    const Fixed msd = P::fixed_one / base;
    if (low_ >= P::fixed_one) {  // carry into deferred digits, lowest first (:154-159)
      for (int i = int(overflow_.size()) - 1; i >= 0; i--) {
        overflow_[i].v = (overflow_[i].v + 1) & ((uint64_t(1) << (8 * DigitBytes)) - 1);
        if (overflow_[i].v != 0) break;
      }
      low_ -= P::fixed_one;
    }
    Fixed digit = (low_ / msd) & (base - 1);
    Fixed digit_hi = ((low_ + range_ - 1) / msd) & (base - 1);
    if (digit != digit_hi) {
      overflow_.push_back(Def{uint64_t(digit)});  // deferred: may still receive a carry (:167-169)
    } else {
      for (const Def& d : overflow_) emit(d.v, DigitBytes);  // overflow holds CompressedDigits (:171)
      overflow_.clear();
      emit(uint64_t(digit), nbytes);
    }
    low_ -= digit * msd;
    low_ *= base;
    range_ *= base;
  }
  void emit(uint64_t digit, int nbytes) {  // big-endian bytes, arithmetic_code.h:185-190
    for (int i = nbytes - 1; i >= 0; i--) out_->push_back(uint8_t(digit >> (8 * i)));
  }
  struct Def { uint64_t v; };
  std::vector<uint8_t>* out_;
  Fixed low_, range_;
  std::vector<Def> overflow_;
};

template <typename Fixed, int DigitBytes, uint64_t MinRangeOverride = 0>
class ArithDecoder {
  typedef ArithParams<Fixed, DigitBytes, MinRangeOverride> P;
  static constexpr uint64_t digit_alignment = uint64_t(Fixed(~Fixed(0)) / P::fixed_one) + 1;  // :251-252

 public:
  // arithmetic_code.h:218-230
  ArithDecoder(const uint8_t* in, const uint8_t* end) : in_(in), end_(end) {
    next_digit_ = aligned();
    low_ = Fixed(next_digit_ / digit_alignment);
    range_ = Fixed(P::digit_base / digit_alignment);
    while (range_ < P::fixed_one) renorm();
  }
  Fixed range() const { return range_; }
  // arithmetic_code.h:232-248
  int get(Fixed range_of_1) {
    Fixed range_of_0 = range_ - range_of_1;
    int symbol = low_ >= range_of_0;
    if (symbol) { low_ -= range_of_0; range_ = range_of_1; } else { range_ = range_of_0; }
    if (range_ < P::min_range)
      while (range_ < P::max_range / P::digit_base) renorm();
    return symbol;
  }

 private:
  void renorm() {  // :259-265
    uint64_t d = digit();
    low_ = low_ * P::digit_base + Fixed(d);
    range_ *= P::digit_base;
  }
  uint64_t digit() {  // :269-275 -- reads are bit-misaligned by log2(digit_alignment)
    uint64_t in_digit = aligned();
    uint64_t mask = (uint64_t(1) << (8 * DigitBytes)) - 1;
    uint64_t d = ((next_digit_ * (uint64_t(P::digit_base) / digit_alignment)) | (in_digit / digit_alignment)) & mask;
    next_digit_ = in_digit;
    return d;
  }
  uint64_t aligned() {  // :279-288 -- zero past the end
    uint64_t d = 0;
    for (int i = 0; i < DigitBytes; i++) {
      d <<= 8;
      if (in_ != end_) d |= *in_++;
    }
    return d;
  }
  const uint8_t *in_, *end_;
  uint64_t next_digit_;
  Fixed low_, range_;
};

typedef ArithEncoder<uint64_t, 1> RecodedEncoder;  // recode.cpp:315-316
typedef ArithDecoder<uint64_t, 1> RecodedDecoder;

// ---------------------------------------------------------------------------------------------
// H.264 CABAC *encoder* on top of the generic coder -- restates cabac_code.h:16-86.
// State byte = 2*pStateIdx + valMPS (ffmpeg convention, cabac.c:36-164).
// ---------------------------------------------------------------------------------------------
class CabacEncoder {
  typedef ArithEncoder<uint32_t, 2, 0x200> E;

 public:
  explicit CabacEncoder(std::vector<uint8_t>* out) : e_(out, (uint32_t(0x80000000u) / 0x200) * 0x1FE) {}  // :30
  void put(int symbol, uint8_t* state) {  // :33-49
    int s = *state;
    bool lps = symbol != (s & 1);
    uint32_t range = e_.range();
    int normalize = ilog2(range / 0x100);
    uint32_t range_approx = range >> (normalize - 1);
    uint32_t q = (range_approx & 0x180) >> 7;
    uint32_t rlps = uint32_t(avr_range_lps[(s >> 1) * 4 + q]) << normalize;
    e_.put(lps, rlps);
    *state = next_state(s, lps);
  }
  void put_bypass(int symbol) { e_.put(symbol, e_.range() / 2); }  // :52-54
  void put_terminate(int symbol) {                                    // :57-67
    int normalize = ilog2(e_.range() / 0x100);
    e_.put(symbol, uint32_t(2) << normalize);
    if (symbol) e_.finish();
  }
  // ff_h264_mlps_state restated in spec form (cabac.c:130-162): LPS -> transIdxLPS (MPS flips at
  // pStateIdx 0), MPS -> min(p+1, 62).
  static uint8_t next_state(int s, bool lps) {
    int p = s >> 1, m = s & 1;
    if (lps) { if (p == 0) m ^= 1; p = avr_trans_lps[p]; }
    else if (p < 62) p++;
    return uint8_t(2 * p + m);
  }

 private:
  static int ilog2(uint32_t x) { int i = 0; while (x >>= 1) i++; return i; }  // :70-79
  E e_;
};

// ---------------------------------------------------------------------------------------------
// H.264 CABAC bin *decoder* -- restates ffmpeg/libavcodec/cabac_functions.h:56-205 and
// cabac.c:193-218 (CABAC_BITS = 16 form).  Bytes past `size` read as zero (the reference reads
// the NAL's trailing bytes there; they can never influence a decoded bin).
// ---------------------------------------------------------------------------------------------
class CabacDecoder {
 public:
  void init(const uint8_t* buf, int size, int tail_byte = 0) {  // cabac.c:193-218, 3-byte ("unaligned") prime
    buf_ = buf; size_ = size; pos_ = 0; tail_ = tail_byte;
    low_ = byte() << 18;
    low_ += byte() << 10;
    low_ += (byte() << 2) + 2;
    range_ = 0x1FE;
    if ((range_ << 17) < low_) throw Corrupt("cabac init: offset out of range");
  }
  int get(uint8_t* state) {  // cabac_functions.h:113-132
    int s = *state;
    int rlps = avr_range_lps[(s >> 1) * 4 + ((range_ >> 6) & 3)];
    int bit;
    range_ -= rlps;
    if (low_ < (range_ << 17)) {
      bit = s & 1;
      *state = CabacEncoder::next_state(s, false);
    } else {
      low_ -= range_ << 17;
      range_ = rlps;
      bit = (s & 1) ^ 1;
      *state = CabacEncoder::next_state(s, true);
    }
    int shift = 0;
    while ((range_ << shift) < 0x100) shift++;  // norm_shift[range]
    range_ <<= shift;
    low_ <<= shift;
    if (!(low_ & 0xFFFF)) refill2();
    return bit;
  }
  int get_bypass() {  // cabac_functions.h:145-163
    low_ += low_;
    if (!(low_ & 0xFFFF)) refill();
    if (low_ < (range_ << 17)) return 0;
    low_ -= range_ << 17;
    return 1;
  }
  int get_terminate() {  // cabac_functions.h:192-204 (returns 0/1 like the hook, recode.cpp:1188-1189)
    range_ -= 2;
    if (low_ < (range_ << 17)) {
      int shift = range_ < 0x100;
      range_ <<= shift;
      low_ <<= shift;
      if (!(low_ & 0xFFFF)) refill();
      return 0;
    }
    return 1;
  }

 private:
  // past the payload the reference reads the rest of the NAL: a lone stop-bit byte 0x80 (when the
  // slice has one, AVR_SLICE_STOP_BYTE) and then libavcodec's zero padding (h264.c:299-302)
  int byte() { int v = pos_ < size_ ? buf_[pos_] : (pos_ == size_ ? tail_ : 0); pos_++; return v; }
  void refill() {  // cabac_functions.h:56-68
    int b0 = byte(), b1 = byte();
    low_ += (b0 << 9) + (b1 << 1);
    low_ -= 0xFFFF;
  }
  void refill2() {  // cabac_functions.h:81-104
    int i = 7 - norm_of((low_ ^ (low_ - 1)) >> 15);
    int b0 = byte(), b1 = byte();
    int x = -0xFFFF + ((b0 << 9) + (b1 << 1));
    low_ += x * (1 << i);
  }
  static int norm_of(int v) { int s = 0; while ((v << s) < 0x100) s++; return s; }  // norm_shift[]
  const uint8_t* buf_; int size_, pos_, tail_ = 0;
  int low_, range_;
};

// ---------------------------------------------------------------------------------------------
// avrecode's adaptive model -- restates recode.cpp:618-1059 (class h264_model) with the
// FrameBuffer/BlockMeta state of framebuffer.h:5-87 and block.h:4-23 reduced to what is read.
// ---------------------------------------------------------------------------------------------
enum CodingType { CT_UNKNOWN = 0, CT_UNREACHABLE, CT_SIGMAP, CT_EOB, CT_RESIDUALS, CT_NZ };  // coding_hooks.h:10-22

enum KeySpace { KS_CTX = 0, KS_BYPASS, KS_TERMINATE, KS_SIG, KS_EOB, KS_NZ0 /* +bit */ };
typedef std::tuple<int, int, int, int> ModelKey;  // (space, ctxIdx-or-0, a, b) -- pointer identity of recode.cpp:318

struct Estimator { int pos = 1, neg = 1; };  // recode.cpp:1057

struct MbMeta {  // BlockMeta (block.h:10-23): only num_nonzeros / is_8x8 / coded are ever used
  uint8_t num_nonzeros[51];
  uint8_t is_8x8;
};

struct FrameBuf {  // framebuffer.h:5-87
  int w = 0, h = 0, frame_num = 0;
  std::vector<MbMeta> meta;
  std::vector<uint8_t> sig;  // Block::residual reduced to the 0/1 flags, [mb][51*16]
  void init(int w_, int h_) { w = w_; h = h_; meta.assign(size_t(w) * h, MbMeta()); sig.assign(size_t(w) * h * 51 * 16, 0); bzero(); }
  void bzero() { memset(meta.data(), 0, meta.size() * sizeof(MbMeta)); memset(sig.data(), 0, sig.size()); }
  MbMeta& at(int x, int y) { return meta[size_t(x) + size_t(y) * w]; }
  uint8_t* sig_at(int x, int y) { return &sig[(size_t(x) + size_t(y) * w) * 51 * 16]; }
};

// recode.cpp:263-312 -- scan_8 / reverse_scan_8, recomputed from geometry rather than tabulated.
// Block index n: luma 0..15, Cb 16..31, Cr 32..47 (4:2:0 uses 16..19 / 32..35), DC 48..50.
inline void blk_xy(int n, int* x, int* y) { *x = (n & 1) + 2 * ((n >> 2) & 1); *y = ((n >> 1) & 1) + 2 * ((n >> 3) & 1); }
inline int blk_of(int x, int y) { return (x & 1) | ((y & 1) << 1) | ((x >> 1) << 2) | ((y >> 1) << 3); }

// recode.cpp:419-471 (get_neighbor_sub_mb): the block left of / above block `n` of MB (mx,my).
inline bool neighbor_sub_mb(bool above, int sub_mb_size, int mx, int my, int n, int* omx, int* omy, int* on) {
  *omx = mx; *omy = my; *on = n;
  if (n >= 48) {  // DC blocks: same slot of the neighbouring MB (:429-443)
    if (above) { if (my > 0) { *omy = my - 1; return true; } return false; }
    if (mx > 0) { *omx = mx - 1; return true; }
    return false;
  }
  int plane = n >> 4, x, y;
  blk_xy(n & 15, &x, &y);
  if (above) { if (--y < 0) { if (my == 0) return false; --my; y = 3; } }
  else       { if (--x < 0) { if (mx == 0) return false; --mx; x = 3; } }
  int nn = plane * 16 + blk_of(x, y);
  if (sub_mb_size >= 32) nn = (nn / 4) * 4;  // :463-466
  *omx = mx; *omy = my; *on = nn;
  return true;
}

inline FILE* trace_file() {  // debugging aid: AVR_ORACLE_TRACE=<path> logs every model-coded bin
  static FILE* f = getenv("AVR_ORACLE_TRACE") ? fopen(getenv("AVR_ORACLE_TRACE"), "w") : nullptr;
  return f;
}

// ---------------------------------------------------------------------------------------------
// Static priors (chunked mode of the B200 library; no reference counterpart -- the reference starts every estimator at
// {1, 1}, recode.cpp:1057).  A prior table is AVR_PRIOR_ENTRIES uint16 `pos | (pos + neg) << 8` in the table layout that
// include/avrecode_b200.h and the header comment of avr_core.cuh document.  This function restates that layout
// independently, from the model KEY of recode.cpp:676-815 / :884 to the table index:
//   [0, 328)        ordinary bins by dense ctxIdx (0..276 as is, 399..435 -> 277.., 1012..1023 -> 314..)
//   328, 329        bypass, terminate
//   330 + ...       nnz side channel: group(cat, is_8x8) + ((2^i - 1 + so_far) * 2 + prev_bit) * 9 + left_bit * 3 + above_bit
//                   groups: cats 0, 1, 2, 4 (in that order) x is_8x8 of 15 * 18 each, then cat 3 x is_8x8 of 3 * 18, then cat 5 of 63 * 18
//   3732 + ...      significance map: cat 0..4: cat * 136 * 16 + (nnz (nnz - 1) / 2 + observed) * 16 + zz
//                   cat 5 (after 5 * 136 * 16): key nnz 0 (= 64 coefficients): observed * 16 + zz; key nnz n: (64 + n (n - 1) / 2 + observed) * 16 + zz
// Returns -1 for keys that have no estimator in the table (KS_EOB: never coded).
// ---------------------------------------------------------------------------------------------
inline int prior_index(const ModelKey& k) {
  const int space = std::get<0>(k), ctx = std::get<1>(k), a = std::get<2>(k), b = std::get<3>(k);
  if (space == KS_CTX) {
    if (ctx >= 0 && ctx < 277) return ctx;
    if (ctx >= 399 && ctx <= 435) return ctx - 399 + 277;
    if (ctx >= 1012 && ctx <= 1023) return ctx - 1012 + 314;
    return -1;
  }
  if (space == KS_BYPASS) return 328;
  if (space == KS_TERMINATE) return 329;
  if (space == KS_SIG) {
    static const int cat_first_ctx[6] = {105, 120, 134, 149, 152, 402};   // the cat_lookup values of recode.cpp:696 for cats 0..5
    static const int cat_coefs[6] = {16, 15, 16, 4, 15, 64};
    int nnz = a >> 6;
    const int observed = a & 63, zz = (b >> 1) & 15, lookup = b >> 5;
    int cat = -1;
    for (int c = 0; c < 6; c++) if (cat_first_ctx[c] == lookup) cat = c;
    if (cat < 0) return -1;
    // A full block's count does not fit its 2 / 4 / 6 serialised bits and travels as 0: the DEcompressor
    // keys such a block's bins with nnz = 0 where the compressor keyed them with nnz = all -- one and the same estimator.
    // (8x8 blocks: the compressor's stored count wraps to 0 as well, so both sides say 0 and the table has a row for it.)
    if (nnz == 0 && cat != 5) nnz = cat_coefs[cat];
    if (observed >= (nnz ? nnz : 64)) return -1;
    const int base = 330 + 3402;
    if (cat == 5) return base + 5 * 136 * 16 + ((nnz == 0 ? 0 : 64 + nnz * (nnz - 1) / 2) + observed) * 16 + zz;
    if (nnz < 1 || nnz > 16) return -1;
    return base + cat * 136 * 16 + (nnz * (nnz - 1) / 2 + observed) * 16 + zz;
  }
  if (space >= KS_NZ0 && space < KS_NZ0 + 6) {
    const int i = space - KS_NZ0, so_far = a & 63, prev_bit = (a >> 6) & 1, lb = (a >> 7) % 3, ab = (a >> 7) / 3;
    const int cat = b >> 2, is8 = b & 1;
    int group;
    if (cat == 5) group = 8 * 15 * 18 + 2 * 3 * 18;
    else if (cat == 3) group = 8 * 15 * 18 + (is8 ? 3 * 18 : 0);
    else if (cat == 0 || cat == 1 || cat == 2 || cat == 4) group = ((cat == 4 ? 3 : cat) * 2 + is8) * 15 * 18;
    else return -1;
    return 330 + group + (((1 << i) - 1 + so_far) * 2 + prev_bit) * 9 + lb * 3 + ab;
  }
  return -1;
}

class Model {
 public:
  const uint16_t* priors = nullptr;   // a fresh estimator starts from priors[prior_index(key)] instead of {1, 1}
  Estimator& lookup(const ModelKey& k) {
    Estimators::iterator it = est->find(k);
    if (it == est->end()) {
      Estimator e;
      const int idx = priors ? prior_index(k) : -1;
      if (idx >= 0) { e.pos = priors[idx] & 0xFF; e.neg = (priors[idx] >> 8) - e.pos; }
      it = est->insert(std::make_pair(k, e)).first;
    }
    return it->second;
  }
  CodingType coding_type = CT_UNKNOWN;
  int mb_x = 0, mb_y = 0, blk = -1, zigzag = 0;  // mb_coord (recode.cpp:1050)
  int nonzeros_observed = 0, sub_mb_cat = -1, sub_mb_size = -1, sub_mb_is_dc = 0, sub_mb_chroma422 = 0;
  FrameBuf frames[2];
  int cur = 0;
  typedef std::map<ModelKey, Estimator> Estimators;
  Estimators* est = &estimators_;   // wide slices (not in the reference) switch between several sets; the geometry is shared
  void use_own_estimators() { est = &estimators_; }

  void reset_all() {  // a brand-new h264_model (chunk mode: one per slice)
    estimators_.clear(); frames[0] = FrameBuf(); frames[1] = FrameBuf(); cur = 0; coding_type = CT_UNKNOWN;
  }
  // recode.cpp:824-843
  void update_frame_spec(int frame_num, int mbw, int mbh) {
    FrameBuf& c = frames[cur];
    bool same = c.frame_num == frame_num && c.w != 0 && c.h != 0;  // framebuffer.h:40-42
    if (c.w != mbw || c.h != mbh || !same) {
      cur = !cur;
      if (frames[cur].w != mbw || frames[cur].h != mbh) {
        frames[cur].init(mbw, mbh);
        if (frames[!cur].w != mbw || frames[!cur].h != mbh) frames[!cur].init(mbw, mbh);
      } else {
        frames[cur].bzero();
      }
      frames[cur].frame_num = frame_num;
    }
  }
  MbMeta& meta() { return frames[cur].at(mb_x, mb_y); }

  // recode.cpp:676-815
  ModelKey key_for_ctx(int space, int ctx_idx) {
    switch (coding_type) {
      case CT_NZ: case CT_UNKNOWN: case CT_UNREACHABLE: case CT_RESIDUALS:
        return ModelKey(space, ctx_idx, 0, 0);
      case CT_SIGMAP: {
        static const int cat_lookup[14] = {105 + 0, 105 + 15, 105 + 29, 105 + 44, 105 + 47, 402, 484 + 0,
                                           484 + 15, 484 + 29, 660, 528 + 0, 528 + 15, 528 + 29, 718};  // :696
        static const uint8_t sig_coeff_offset_dc[7] = {0, 0, 1, 1, 2, 2, 2};
        int zz = zigzag;
        if (sub_mb_is_dc && sub_mb_chroma422) zz = sig_coeff_offset_dc[zigzag];
        else if (sub_mb_size > 32) zz = avr_sig8x8_inc[zigzag];  // frame table always (:704)
        int nnz = meta().num_nonzeros[blk];
        return ModelKey(KS_SIG, 0, 64 * nnz + nonzeros_observed, sub_mb_is_dc + zz * 2 + 16 * 2 * cat_lookup[sub_mb_cat]);
      }
      case CT_EOB: {
        int nnz = meta().num_nonzeros[blk];
        return ModelKey(KS_EOB, 0, nnz == nonzeros_observed, 0);  // :802-809
      }
    }
    throw Corrupt("model: bad coding type");
  }
  // recode.cpp:816-820 -- divide first, then multiply
  uint64_t probability(uint64_t range, const ModelKey& k) {
    Estimator& e = lookup(k);
    return (range / uint64_t(e.pos + e.neg)) * uint64_t(e.pos);
  }
  // recode.cpp:1030-1047
  void update(int symbol, const ModelKey& k) {
    Estimator& e = lookup(k);
    if (FILE* tf = trace_file())  // same columns as oracle/_ref/recode_ref_trace: type, key identity, ints, symbol, counters
      fprintf(tf, "%d %d %d %d %d %d %d\n", int(coding_type), std::get<0>(k) * 4096 + std::get<1>(k), std::get<2>(k), std::get<3>(k), symbol, e.pos, e.neg);
    if (symbol) e.pos++; else e.neg++;
    int limit = coding_type == CT_SIGMAP ? 0x50 : 0x60;
    if (e.pos + e.neg > limit) { e.pos = (e.pos + 1) / 2; e.neg = (e.neg + 1) / 2; }
    update_state_tracking(symbol);
  }
  // recode.cpp:980-1026
  void update_state_tracking(int symbol) {
    uint8_t* sig = frames[cur].sig_at(mb_x, mb_y);
    switch (coding_type) {
      case CT_NZ: break;
      case CT_SIGMAP:
        sig[blk * 16 + zigzag] = uint8_t(symbol);
        nonzeros_observed += symbol;
        if (zigzag + 1 == sub_mb_size) { coding_type = CT_UNREACHABLE; zigzag = 0; }
        else if (symbol) coding_type = CT_EOB;
        else {
          ++zigzag;
          if (zigzag + 1 == sub_mb_size) {  // trailing coefficient is implied (:995-1002)
            sig[blk * 16 + zigzag] = 1; ++nonzeros_observed; coding_type = CT_UNREACHABLE; zigzag = 0;
          }
        }
        break;
      case CT_EOB:
        if (symbol) { zigzag = 0; coding_type = CT_UNREACHABLE; }
        else if (zigzag + 2 == sub_mb_size) { sig[blk * 16 + zigzag + 1] = 1; coding_type = CT_UNREACHABLE; }
        else { coding_type = CT_SIGMAP; ++zigzag; }
        break;
      case CT_RESIDUALS: case CT_UNKNOWN: break;
      default: throw Corrupt("model: bin in unreachable state");
    }
  }
  // recode.cpp:951-974
  void begin_sigmap() {
    coding_type = CT_SIGMAP;
    meta().num_nonzeros[blk] = 0;
    nonzeros_observed = 0;
    zigzag = 0;
  }
  // recode.cpp:931-950
  void end_sigmap() {
    uint8_t* sig = frames[cur].sig_at(mb_x, mb_y);
    int n = 0;
    for (int i = 0; i < sub_mb_size; i++) n += sig[blk * 16 + i] != 0;
    MbMeta& m = meta();
    m.is_8x8 = m.is_8x8 || sub_mb_size > 32;
    m.num_nonzeros[blk] = uint8_t(n);
    coding_type = CT_UNKNOWN;
  }
  void reset_sig_tracking() { zigzag = 0; nonzeros_observed = 0; coding_type = CT_SIGMAP; }  // :975-979

  // recode.cpp:844-930 -- the num_nonzeros side channel.  put_or_get(key, &bit)
  template <class F>
  void finished_queueing(const F& put_or_get) {
    CodingType last = coding_type;
    coding_type = CT_NZ;
    MbMeta& m = meta();
    int bits[6];
    for (int i = 0; i < 6; i++) bits[i] = (m.num_nonzeros[blk] >> i) & 1;
    const int serialized_bits = sub_mb_size > 16 ? 6 : sub_mb_size > 4 ? 4 : 2;  // :858
    int nx, ny, nb;
    unsigned left_nz = 0, above_nz = 0, so_far = 0;
    bool has_left = neighbor_sub_mb(false, sub_mb_size, mb_x, mb_y, blk, &nx, &ny, &nb);
    if (has_left) left_nz = frames[cur].at(nx, ny).num_nonzeros[nb];
    bool has_above = neighbor_sub_mb(true, sub_mb_size, mb_x, mb_y, blk, &nx, &ny, &nb);
    if (has_above) above_nz = frames[cur].at(nx, ny).num_nonzeros[nb];
    unsigned prev_nz = frames[!cur].at(mb_x, mb_y).num_nonzeros[blk];
    for (int i = 0; i < serialized_bits; i++) {
      unsigned cur_bit = 1u << i;
      int left_bit = 2, above_bit = 2;
      if (has_left) left_bit = left_nz >= cur_bit;
      if (above_nz) above_bit = above_nz >= cur_bit;  // sic: tests the value, not has_above (:881)
      ModelKey k(KS_NZ0 + i, 0, int(so_far + 64 * (prev_nz >= cur_bit) + 128 * left_bit + 384 * above_bit),
                 m.is_8x8 + sub_mb_is_dc * 2 + sub_mb_chroma422 + sub_mb_cat * 4);
      put_or_get(k, &bits[i]);
      if (bits[i]) so_far |= cur_bit;
    }
    m.num_nonzeros[blk] = 0;
    for (int i = 0; i < 6; i++) m.num_nonzeros[blk] |= uint8_t(bits[i] << i);  // :921-924
    coding_type = last;
  }

 private:
  Estimators estimators_;
};


// Hints: the walker tags every bin request with what it MEANS (kind | arg << 8).  The real
// back ends ignore the tag; the synthetic stream generator (tools/h264gen.cpp) uses it to pick
// legal, statistically shaped values.  Not part of the reference.
enum HintKind {
  HK_NONE = 0, HK_SKIP, HK_MBTYPE, HK_I16_CBPL, HK_I16_CBPC, HK_I16_PRED_HI, HK_I16_PRED_LO, HK_PCM_TERM,
  HK_CHROMA_PRED, HK_I4_FLAG, HK_I4_REM, HK_SUB_TYPE, HK_REF, HK_MVD_NZ, HK_MVD_UNARY, HK_MVD_ESC_PREFIX,
  HK_MVD_ESC_SUFFIX, HK_MVD_SIGN, HK_QPD_NZ, HK_QPD_MORE, HK_CBP_LUMA, HK_CBP_CHROMA_NZ, HK_CBP_CHROMA_2,
  HK_T8X8, HK_CBF, HK_SIG, HK_LAST, HK_LVL_GT1, HK_LVL_MORE, HK_LVL_ESC_PREFIX, HK_LVL_ESC_SUFFIX, HK_SIGN, HK_EOS,
};
inline int mk_hint(int kind, int arg = 0) { return kind | (arg << 8); }

// ---------------------------------------------------------------------------------------------
// Bin I/O back ends.  The slice_data walker below is written once against this interface:
//   int  get(int ctxIdx); int bypass(); int terminate();
//   void mb_xy(x,y); void begin_block(cat,n,max_coeff,is_dc); void end_block();
//   void begin_sigmap(); void end_sigmap();
// ---------------------------------------------------------------------------------------------

// compressor::cabac_decoder -- recode.cpp:1134-1268 (+ h264_symbol::execute :1061-1100)
class CompressIO {
 public:
  CompressIO(Model* m, const uint8_t* payload, int size, std::vector<uint8_t>* out, int tail_byte = 0)
      : model_(m), enc_(out), out_(out) { dec_.init(payload, size, tail_byte); }
  uint8_t state[1024];
  uint64_t n_bins = 0;

  int get(int ctx, int = 0) { int s = dec_.get(&state[ctx]); symbol(s, KS_CTX, ctx); return s; }         // :1175-1179
  int bypass(int = 0) { int s = dec_.get_bypass(); symbol(s, KS_BYPASS, 0); return s; }                 // :1181-1185
  int terminate(int = 0) { int s = dec_.get_terminate(); symbol(s, KS_TERMINATE, 0); return s; }        // :1187-1192
  void mb_xy(int x, int y) { model_->mb_x = x; model_->mb_y = y; }                               // :170-174
  void begin_block(int cat, int n, int max_coeff, int is_dc) {                                   // :175-182
    model_->sub_mb_cat = cat; model_->blk = n; model_->sub_mb_size = max_coeff; model_->sub_mb_is_dc = is_dc;
    model_->sub_mb_chroma422 = 0;
  }
  void end_block() { model_->sub_mb_cat = -1; model_->blk = -1; model_->sub_mb_size = -1; model_->sub_mb_is_dc = 0; }
  void begin_sigmap() { model_->begin_sigmap(); queueing_ = true; }                              // :1194-1203
  void end_sigmap() {                                                                            // :1204-1229
    model_->end_sigmap();
    queueing_ = false;
    model_->finished_queueing([&](const ModelKey& k, int* bit) {
      enc_.put(*bit, model_->probability(enc_.range(), k));
      model_->update(*bit, k);
    });
    model_->reset_sig_tracking();
    for (const Queued& q : queue_) execute(q.symbol, q.space, q.ctx);
    queue_.clear();
    model_->coding_type = CT_UNKNOWN;
  }
  bool finished() const { return finished_; }

 private:
  struct Queued { int symbol, space, ctx; };
  void symbol(int s, int space, int ctx) {  // execute_symbol, recode.cpp:1160-1173
    n_bins++;
    if (queueing_ || !queue_.empty()) { queue_.push_back(Queued{s, space, ctx}); model_->update_state_tracking(s); }
    else execute(s, space, ctx);
  }
  void execute(int s, int space, int ctx) {  // h264_symbol::execute, recode.cpp:1068-1096
    ModelKey k = model_->key_for_ctx(space, ctx);
    if (model_->coding_type != CT_EOB) enc_.put(s, model_->probability(enc_.range(), k));
    model_->update(s, k);
    if (space == KS_TERMINATE && s) { enc_.finish(); finished_ = true; }
  }
  Model* model_;
  CabacDecoder dec_;
  RecodedEncoder enc_;
  std::vector<uint8_t>* out_;
  std::vector<Queued> queue_;
  bool queueing_ = false, finished_ = false;
};

// decompressor::cabac_decoder -- recode.cpp:1411-1520
class DecompressIO {
 public:
  DecompressIO(Model* m, const uint8_t* coded, int size, std::vector<uint8_t>* out)
      : model_(m), dec_(coded, coded + size), cabac_(out), out_(out) {}
  uint8_t state[1024];
  uint64_t n_bins = 0;

  int get(int ctx, int = 0) {  // :1435-1449
    n_bins++;
    ModelKey k = model_->key_for_ctx(KS_CTX, ctx);
    int s;
    if (model_->coding_type == CT_EOB) s = std::get<2>(k);
    else s = dec_.get(model_->probability(dec_.range(), k));
    cabac_.put(s, &state[ctx]);
    model_->update(s, k);
    return s;
  }
  int bypass(int = 0) {  // :1451-1460
    n_bins++;
    ModelKey k = model_->key_for_ctx(KS_BYPASS, 0);
    int s = dec_.get(model_->probability(dec_.range(), k));
    model_->update(s, k);
    cabac_.put_bypass(s);
    return s;
  }
  int terminate(int = 0) {  // :1462-1474, finish() :1501-1508
    n_bins++;
    ModelKey k = model_->key_for_ctx(KS_TERMINATE, 0);
    int s = dec_.get(model_->probability(dec_.range(), k));
    model_->update(s, k);
    cabac_.put_terminate(s);
    if (s) {
      if (!out_->empty() && out_->back() == 0x80) out_->pop_back();
      finished_ = true;
    }
    return s;
  }
  void mb_xy(int x, int y) { model_->mb_x = x; model_->mb_y = y; }
  void begin_block(int cat, int n, int max_coeff, int is_dc) {
    model_->sub_mb_cat = cat; model_->blk = n; model_->sub_mb_size = max_coeff; model_->sub_mb_is_dc = is_dc;
    model_->sub_mb_chroma422 = 0;
  }
  void end_block() { model_->sub_mb_cat = -1; model_->blk = -1; model_->sub_mb_size = -1; model_->sub_mb_is_dc = 0; }
  // mirror_compressor: the stock reference learns is_8x8 only at end_coding_type (recode.cpp:944),
  // i.e. AFTER its NZ bits, while the compressor learns it BEFORE (SURVEY hazard H8).  The format
  // is defined by the compressor, so by default we set it first (== oracle/_ref/recode_ref_h8).
  bool mirror_compressor = true;
  void begin_sigmap() {  // :1476-1492
    model_->begin_sigmap();
    if (mirror_compressor) { MbMeta& m = model_->meta(); m.is_8x8 = m.is_8x8 || model_->sub_mb_size > 32; }
    model_->finished_queueing([&](const ModelKey& k, int* bit) {
      *bit = dec_.get(model_->probability(dec_.range(), k));
      model_->update(*bit, k);
    });
    stored_nnz_ = model_->meta().num_nonzeros[model_->blk];
  }
  void end_sigmap() {  // :1493-1498
    model_->end_sigmap();
    // H9: the compressor keeps the 6-bit-truncated count (64 -> 0) for neighbours.
    if (mirror_compressor && model_->sub_mb_size > 32) model_->meta().num_nonzeros[model_->blk] &= 63;
  }
  bool finished() const { return finished_; }

 private:
  Model* model_;
  RecodedDecoder dec_;
  CabacEncoder cabac_;
  std::vector<uint8_t>* out_;
  int stored_nnz_ = 0;
  bool finished_ = false;
};

// ---------------------------------------------------------------------------------------------
// Wide slices -- avrecode-b200's own chunked-mode format, NO reference counterpart (include/avrecode_b200.h,
// avr_slice_desc.n_streams).  Restated here from the format's definition, with the per-bin primitives above:
//   * a slice is coded as K range-coder streams with K estimator sets; macroblock row r goes to stream
//     (r - first row of the slice) mod K.  Everything else of the model -- the nnz planes, the significance
//     tracking -- is the slice's, shared by the streams (a block's neighbours may sit in another stream's row);
//   * the first bin of mb_qp_delta of the FIRST macroblock of a row is modelled with the estimator of ctxIdx 60
//     whatever its CABAC context is (60 or 61: "the previous macroblock in decoding order had a non-zero
//     mb_qp_delta" is not known to a decoder that starts the row before the row above is finished);
//   * every stream ends with a terminate bin of value 1 that flushes its coder (arithmetic_code.h:128-144): the
//     slice's own end_of_slice_flag in the stream of the last row, one extra model-coded bin in each other;
//   * blob = u32 macroblocks | K x u32 stream bytes | the streams, little endian.
// ---------------------------------------------------------------------------------------------
inline void put_le32(std::vector<uint8_t>* o, uint32_t v) { for (int i = 0; i < 4; i++) o->push_back(uint8_t(v >> (8 * i))); }
inline uint32_t get_le32(const uint8_t* p) { return uint32_t(p[0]) | uint32_t(p[1]) << 8 | uint32_t(p[2]) << 16 | uint32_t(p[3]) << 24; }

class WideCompressIO {
 public:
  WideCompressIO(Model* m, int k, int first_row, const uint8_t* payload, int size, std::vector<uint8_t>* out, int tail_byte = 0)
      : model_(m), k_(k), first_row_(first_row), out_(out), ests_(size_t(k)), bytes_(size_t(k)) {
    dec_.init(payload, size, tail_byte);
    for (int j = 0; j < k; j++) encs_.emplace_back(new RecodedEncoder(&bytes_[size_t(j)]));
  }
  uint8_t state[1024];
  uint64_t n_bins = 0;
  int get(int ctx, int hint = 0) {
    int s = dec_.get(&state[ctx]);
    const bool row_qpd = (hint & 0xFF) == HK_QPD_NZ && model_->mb_x == 0;
    symbol(s, KS_CTX, row_qpd ? 60 : ctx);
    return s;
  }
  int bypass(int = 0) { int s = dec_.get_bypass(); symbol(s, KS_BYPASS, 0); return s; }
  int terminate(int = 0) {
    int s = dec_.get_terminate();
    symbol(s, KS_TERMINATE, 0);
    if (s) {   // the other streams end with a terminate bin of their own; then the blob
      const int last = cur_;
      for (int j = 0; j < k_; j++) if (j != last) { select(j); execute(1, KS_TERMINATE, 0); }
      finished_ = true;
    }
    return s;
  }
  void mb_xy(int x, int y) { model_->mb_x = x; model_->mb_y = y; mbs_++; select((y - first_row_) % k_); }
  void begin_block(int cat, int n, int max_coeff, int is_dc) {
    model_->sub_mb_cat = cat; model_->blk = n; model_->sub_mb_size = max_coeff; model_->sub_mb_is_dc = is_dc; model_->sub_mb_chroma422 = 0;
  }
  void end_block() { model_->sub_mb_cat = -1; model_->blk = -1; model_->sub_mb_size = -1; model_->sub_mb_is_dc = 0; }
  void begin_sigmap() { model_->begin_sigmap(); queueing_ = true; }
  void end_sigmap() {   // as CompressIO::end_sigmap, into the current stream
    model_->end_sigmap();
    queueing_ = false;
    model_->finished_queueing([&](const ModelKey& k, int* bit) {
      encs_[size_t(cur_)]->put(*bit, model_->probability(encs_[size_t(cur_)]->range(), k));
      model_->update(*bit, k);
    });
    model_->reset_sig_tracking();
    for (const Queued& q : queue_) execute(q.symbol, q.space, q.ctx);
    queue_.clear();
    model_->coding_type = CT_UNKNOWN;
  }
  bool finished() const { return finished_; }
  void write_blob() {
    put_le32(out_, uint32_t(mbs_));
    for (int j = 0; j < k_; j++) put_le32(out_, uint32_t(bytes_[size_t(j)].size()));
    for (int j = 0; j < k_; j++) out_->insert(out_->end(), bytes_[size_t(j)].begin(), bytes_[size_t(j)].end());
    model_->use_own_estimators();
  }

 private:
  struct Queued { int symbol, space, ctx; };
  void select(int j) { cur_ = j; model_->est = &ests_[size_t(j)]; }
  void symbol(int s, int space, int ctx) {
    n_bins++;
    if (queueing_ || !queue_.empty()) { queue_.push_back(Queued{s, space, ctx}); model_->update_state_tracking(s); }
    else execute(s, space, ctx);
  }
  void execute(int s, int space, int ctx) {
    RecodedEncoder& enc = *encs_[size_t(cur_)];
    ModelKey k = model_->key_for_ctx(space, ctx);
    if (model_->coding_type != CT_EOB) enc.put(s, model_->probability(enc.range(), k));
    model_->update(s, k);
    if (space == KS_TERMINATE && s) enc.finish();
  }
  Model* model_;
  int k_, first_row_, cur_ = 0, mbs_ = 0;
  CabacDecoder dec_;
  std::vector<uint8_t>* out_;
  std::vector<Model::Estimators> ests_;
  std::vector<std::vector<uint8_t>> bytes_;
  std::vector<std::unique_ptr<RecodedEncoder>> encs_;
  std::vector<Queued> queue_;
  bool queueing_ = false, finished_ = false;
};

// The inverse, serially: macroblock by macroblock in slice order, each row from its own stream.
class WideDecompressIO {
 public:
  WideDecompressIO(Model* m, int k, int first_row, const uint8_t* blob, int size, std::vector<uint8_t>* out)
      : model_(m), k_(k), first_row_(first_row), cabac_(out), out_(out), ests_(size_t(k)) {
    if (size < 4 + 4 * k) throw Corrupt("wide blob header");
    total_mbs_ = int(get_le32(blob));
    size_t off = size_t(4 + 4 * k);
    for (int j = 0; j < k; j++) {
      const size_t n = get_le32(blob + 4 + 4 * j);
      if (off + n > size_t(size)) throw Corrupt("wide blob sizes");
      decs_.emplace_back(new RecodedDecoder(blob + off, blob + off + n));
      off += n;
    }
    if (off != size_t(size)) throw Corrupt("wide blob sizes");
  }
  uint8_t state[1024];
  uint64_t n_bins = 0;
  int get(int ctx, int hint = 0) {
    n_bins++;
    const bool row_qpd = (hint & 0xFF) == HK_QPD_NZ && model_->mb_x == 0;
    ModelKey k = model_->key_for_ctx(KS_CTX, row_qpd ? 60 : ctx);
    int s;
    if (model_->coding_type == CT_EOB) s = std::get<2>(k);
    else s = dec().get(model_->probability(dec().range(), k));
    cabac_.put(s, &state[ctx]);
    model_->update(s, k);
    return s;
  }
  int bypass(int = 0) {
    n_bins++;
    ModelKey k = model_->key_for_ctx(KS_BYPASS, 0);
    int s = dec().get(model_->probability(dec().range(), k));
    model_->update(s, k);
    cabac_.put_bypass(s);
    return s;
  }
  int terminate(int = 0) {
    n_bins++;
    ModelKey k = model_->key_for_ctx(KS_TERMINATE, 0);
    int s = dec().get(model_->probability(dec().range(), k));
    model_->update(s, k);
    cabac_.put_terminate(s);
    if (s) {
      if (!out_->empty() && out_->back() == 0x80) out_->pop_back();
      finished_ = mbs_ == total_mbs_;
      model_->use_own_estimators();
    }
    return s;
  }
  void mb_xy(int x, int y) { model_->mb_x = x; model_->mb_y = y; mbs_++; cur_ = (y - first_row_) % k_; model_->est = &ests_[size_t(cur_)]; }
  void begin_block(int cat, int n, int max_coeff, int is_dc) {
    model_->sub_mb_cat = cat; model_->blk = n; model_->sub_mb_size = max_coeff; model_->sub_mb_is_dc = is_dc; model_->sub_mb_chroma422 = 0;
  }
  void end_block() { model_->sub_mb_cat = -1; model_->blk = -1; model_->sub_mb_size = -1; model_->sub_mb_is_dc = 0; }
  void begin_sigmap() {   // as DecompressIO::begin_sigmap with mirror_compressor
    model_->begin_sigmap();
    MbMeta& m = model_->meta(); m.is_8x8 = m.is_8x8 || model_->sub_mb_size > 32;
    model_->finished_queueing([&](const ModelKey& k, int* bit) {
      *bit = dec().get(model_->probability(dec().range(), k));
      model_->update(*bit, k);
    });
  }
  void end_sigmap() {
    model_->end_sigmap();
    if (model_->sub_mb_size > 32) model_->meta().num_nonzeros[model_->blk] &= 63;
  }
  bool finished() const { return finished_; }

 private:
  RecodedDecoder& dec() { return *decs_[size_t(cur_)]; }
  Model* model_;
  int k_, first_row_, cur_ = 0, mbs_ = 0, total_mbs_ = 0;
  CabacEncoder cabac_;
  std::vector<uint8_t>* out_;
  std::vector<Model::Estimators> ests_;
  std::vector<std::unique_ptr<RecodedDecoder>> decs_;
  bool finished_ = false;
};

// ---------------------------------------------------------------------------------------------
// Slice-level syntax state shared by all slices of one picture (ffmpeg keeps these per H264Context).
// ---------------------------------------------------------------------------------------------
enum MbFlags : uint32_t {
  MB_I4x4 = 1, MB_I16x16 = 2, MB_IPCM = 4, MB_16x16 = 8, MB_16x8 = 16, MB_8x16 = 32, MB_8x8 = 64,
  MB_DIRECT = 256, MB_SKIP = 2048, MB_P0L0 = 1 << 12, MB_P1L0 = 1 << 13, MB_P0L1 = 1 << 14, MB_P1L1 = 1 << 15,
  MB_8x8DCT = 1 << 24,
};
inline bool is_intra(uint32_t t) { return t & (MB_I4x4 | MB_I16x16 | MB_IPCM); }
inline bool uses_list(uint32_t t, int list) { return t & ((MB_P0L0 | MB_P1L0) << (2 * list)); }  // USES_LIST
inline bool is_dir(uint32_t t, int part, int list) { return t & (MB_P0L0 << (part + 2 * list)); }  // IS_DIR

struct MbSyntax {
  uint32_t type = 0;      // cur_pic.mb_type
  uint32_t slice_id = 0;  // slice_table (0 = not decoded in this picture)
  uint16_t cbp = 0;       // cbp_table (incl. DC bits 0x40<<c, 0x100)
  uint8_t chroma_pred = 0;
  uint8_t nnz[48] = {0};  // non_zero_count, by block index
  uint8_t mvd[2][16][2] = {{{0}}};
  int8_t ref[2][4] = {{0}};
  uint8_t direct8[4] = {0};
};

struct PictureSyntax {
  int mbw = 0, mbh = 0;
  uint32_t next_slice_id = 1;
  std::vector<MbSyntax> mbs;
  void start(int w, int h) { mbw = w; mbh = h; mbs.assign(size_t(w) * h, MbSyntax()); next_slice_id = 1; }
};

// ---------------------------------------------------------------------------------------------
// slice_data() walker -- restates ffmpeg/libavcodec/h264_cabac.c:1267-2480, h264_mvpred.h:356-834
// and the CABAC loop of h264_slice.c:2357-2443, keeping only the state that selects contexts.
// ---------------------------------------------------------------------------------------------
template <class IO>
class SliceWalker {
 public:
  SliceWalker(IO* io, PictureSyntax* pic, const avr_slice_desc& d) : io_(*io), pic_(*pic), d_(d) {}

  // h264_cabac.c:1267-1286
  static void init_states(uint8_t* state, const avr_slice_desc& d) {
    int set = d.slice_type == AVR_SLICE_I ? 0 : 1 + d.cabac_init_idc;
    int qp = d.slice_qp < 0 ? 0 : d.slice_qp > 51 ? 51 : d.slice_qp;
    for (int i = 0; i < 1024; i++) {
      uint16_t mn = avr_ctx_init_mn[set * 1024 + i];
      int m = int8_t(mn >> 8), n = int8_t(mn & 0xFF);
      int pre = 2 * (((m * qp) >> 4) + n) - 127;
      pre ^= pre >> 31;
      if (pre > 124) pre = 124 + (pre & 1);
      state[i] = uint8_t(pre);
    }
  }

  // h264_slice.c:2366-2443.  Returns number of macroblocks walked.
  int run() {
    if (d_.flags & (AVR_SLICE_MBAFF | AVR_SLICE_FIELD)) throw Unsupported("field/MBAFF");
    if (d_.chroma_format_idc != 1) throw Unsupported("chroma format");
    if (d_.slice_type > AVR_SLICE_I) throw Unsupported("SP/SI slice");
    init_states(io_.state, d_);
    slice_id_ = pic_.next_slice_id++;
    qp_ = d_.slice_qp;
    last_qp_delta_ = 0;
    int mbs = 0;
    mb_x_ = int(d_.first_mb % uint32_t(pic_.mbw));
    mb_y_ = int(d_.first_mb / uint32_t(pic_.mbw));
    for (;;) {
      if (mb_y_ >= pic_.mbh) throw Corrupt("slice runs past the picture");
      io_.mb_xy(mb_x_, mb_y_);
      macroblock();
      mbs++;
      int eos = io_.terminate(mk_hint(HK_EOS, mbs));
      if (++mb_x_ >= pic_.mbw) { mb_x_ = 0; ++mb_y_; }
      if (eos) break;
      if (mb_y_ >= pic_.mbh) break;  // h264_slice.c:2431 -- picture full, no terminate bin = 1 seen
    }
    return mbs;
  }
  bool left_available() const { return left_ != nullptr; }
  bool top_available() const { return top_ != nullptr; }
  int qp() const { return qp_; }

 private:
  IO& io_;
  PictureSyntax& pic_;
  const avr_slice_desc& d_;
  uint32_t slice_id_ = 0;
  int mb_x_ = 0, mb_y_ = 0, qp_ = 0, last_qp_delta_ = 0;

  MbSyntax* cur_ = nullptr;
  const MbSyntax *left_ = nullptr, *top_ = nullptr;  // null when unavailable (h264_mvpred.h:425-442)
  uint32_t left_type_ = 0, top_type_ = 0;
  int left_cbp_ = 0, top_cbp_ = 0;
  int cur_ref_[2][4];
  bool cur_direct_[4];

  const MbSyntax* avail(int x, int y) const {
    if (x < 0 || y < 0 || x >= pic_.mbw) return nullptr;
    const MbSyntax* m = &pic_.mbs[size_t(x) + size_t(y) * pic_.mbw];
    return m->slice_id == slice_id_ ? m : nullptr;
  }

  // --- neighbour fetch helpers (fill_decode_caches, h264_mvpred.h:445-798) -------------------
  // nnz of the block left of / above luma-or-chroma block n (cache semantics :542-587)
  int nnz_neighbor(int n, bool above, bool cur_intra) const {
    int plane = n >> 4, x, y;
    blk_xy(n & 15, &x, &y);
    int w = plane ? 2 : 4;  // 4:2:0 chroma is 2x2 (blocks 16..19 / 32..35)
    const MbSyntax* m = cur_;
    if (above) { if (--y < 0) { m = top_; y = w - 1; } }
    else       { if (--x < 0) { m = left_; x = w - 1; } }
    if (!m) return cur_intra ? 64 : 0;
    return m->nnz[plane * 16 + blk_of(x, y)];
  }
  int mvd_neighbor_sum(int list, int n, int comp) const {  // DECODE_CABAC_MB_MVD, h264_cabac.c:1539-1548
    int x, y;
    blk_xy(n, &x, &y);
    int a = 0, b = 0;
    if (x > 0) a = cur_->mvd[list][blk_of(x - 1, y)][comp];
    else if (left_ && uses_list(left_->type, list)) a = left_->mvd[list][blk_of(3, y)][comp];
    if (y > 0) b = cur_->mvd[list][blk_of(x, y - 1)][comp];
    else if (top_ && uses_list(top_->type, list)) b = top_->mvd[list][blk_of(x, 3)][comp];
    return a + b;
  }
  // ref_cache / direct_cache at the 8x8 left of / above 4x4 block n (h264_mvpred.h:618-662, 723-753)
  void ref_neighbor(int list, int n, bool above, int* ref, bool* direct) const {
    int x, y;
    blk_xy(n, &x, &y);
    const MbSyntax* m = cur_;
    if (above) { if (--y < 0) { m = top_; y = 3; } }
    else       { if (--x < 0) { m = left_; x = 3; } }
    *direct = false;
    if (!m) { *ref = -2; return; }  // PART_NOT_AVAILABLE
    int b8 = (x >> 1) + 2 * (y >> 1);
    if (m == cur_) { *ref = cur_ref_[list][b8]; *direct = cur_direct_[b8]; return; }
    if (m->type & MB_DIRECT) *direct = true;                         // :728-730, 740-741
    else if (m->type & MB_8x8) *direct = m->direct8[b8] != 0;        // :731-734, 742-743
    *ref = uses_list(m->type, list) ? m->ref[list][b8] : -1;         // LIST_NOT_USED
  }

  // --- syntax elements ------------------------------------------------------------------------
  int mb_skip_flag() {  // h264_cabac.c:1332-1367
    const MbSyntax* a = avail(mb_x_ - 1, mb_y_);
    const MbSyntax* b = avail(mb_x_, mb_y_ - 1);
    int ctx = 0;
    if (a && !(a->type & MB_SKIP)) ctx++;
    if (b && !(b->type & MB_SKIP)) ctx++;
    if (d_.slice_type == AVR_SLICE_B) ctx += 13;
    return io_.get(11 + ctx, mk_hint(HK_SKIP));
  }
  // which Intra16x16 prediction modes the available neighbours allow (h264.c:181-224), as a mask over
  // the bitstream's 2-bit code (0 = V, 1 = H, 2 = DC, 3 = plane).  Only a hint for the generator.
  int i16_pred_mask() const {
    bool l = left_ != nullptr, t = top_ != nullptr;
    return 4 | (t ? 1 : 0) | (l ? 2 : 0) | ((l && t && avail(mb_x_ - 1, mb_y_ - 1)) ? 8 : 0);
  }
  int intra_mb_type(int ctx_base, bool intra_slice) {  // h264_cabac.c:1300-1330
    int base = ctx_base;
    if (intra_slice) {
      int ctx = 0;
      if (left_type_ & (MB_I16x16 | MB_IPCM)) ctx++;
      if (top_type_ & (MB_I16x16 | MB_IPCM)) ctx++;
      if (io_.get(base + ctx, mk_hint(HK_MBTYPE, 100)) == 0) return 0;
      base += 2;
    } else if (io_.get(base, mk_hint(HK_MBTYPE, 100)) == 0) return 0;
    if (io_.terminate(mk_hint(HK_PCM_TERM))) return 25;
    int is = intra_slice ? 1 : 0;
    int t = 1;
    t += 12 * io_.get(base + 1, mk_hint(HK_I16_CBPL));
    if (io_.get(base + 2, mk_hint(HK_I16_CBPC, 0))) t += 4 + 4 * io_.get(base + 2 + is, mk_hint(HK_I16_CBPC, 1));
    int mask = i16_pred_mask();
    int hi = io_.get(base + 3 + is, mk_hint(HK_I16_PRED_HI, mask));
    t += 2 * hi;
    t += 1 * io_.get(base + 3 + 2 * is, mk_hint(HK_I16_PRED_LO, mask | (hi << 4)));
    return t;
  }
  int chroma_pred_mode() {  // h264_cabac.c:1383-1406 ; bitstream code 0 = DC, 1 = H, 2 = V, 3 = plane
    int ctx = 0;
    if (left_ && left_->chroma_pred != 0) ctx++;
    if (top_ && top_->chroma_pred != 0) ctx++;
    bool l = left_ != nullptr, t = top_ != nullptr;
    int mask = 1 | (l ? 2 : 0) | (t ? 4 : 0) | ((l && t && avail(mb_x_ - 1, mb_y_ - 1)) ? 8 : 0);
    if (io_.get(64 + ctx, mk_hint(HK_CHROMA_PRED, mask)) == 0) return 0;
    if (io_.get(64 + 3, mk_hint(HK_CHROMA_PRED, mask | (1 << 4))) == 0) return 1;
    return io_.get(64 + 3, mk_hint(HK_CHROMA_PRED, mask | (2 << 4))) == 0 ? 2 : 3;
  }
  int cbp_luma() {  // h264_cabac.c:1408-1424
    int a = left_cbp_, b = top_cbp_, cbp = 0, ctx;
    ctx = !(a & 0x02) + 2 * !(b & 0x04);   cbp += io_.get(73 + ctx, mk_hint(HK_CBP_LUMA, 0));
    ctx = !(cbp & 0x01) + 2 * !(b & 0x08); cbp += io_.get(73 + ctx, mk_hint(HK_CBP_LUMA, 1)) << 1;
    ctx = !(a & 0x08) + 2 * !(cbp & 0x01); cbp += io_.get(73 + ctx, mk_hint(HK_CBP_LUMA, 2)) << 2;
    ctx = !(cbp & 0x04) + 2 * !(cbp & 0x02); cbp += io_.get(73 + ctx, mk_hint(HK_CBP_LUMA, 3)) << 3;
    return cbp;
  }
  int cbp_chroma() {  // h264_cabac.c:1425-1443
    int a = (left_cbp_ >> 4) & 3, b = (top_cbp_ >> 4) & 3;
    int ctx = (a > 0) + 2 * (b > 0);
    if (io_.get(77 + ctx, mk_hint(HK_CBP_CHROMA_NZ)) == 0) return 0;
    ctx = 4 + (a == 2) + 2 * (b == 2);
    return 1 + io_.get(77 + ctx, mk_hint(HK_CBP_CHROMA_2));
  }
  int p_sub_mb_type() {  // h264_cabac.c:1445-1454
    if (io_.get(21, mk_hint(HK_SUB_TYPE, 0))) return 0;
    if (!io_.get(22, mk_hint(HK_SUB_TYPE, 1))) return 1;
    return io_.get(23, mk_hint(HK_SUB_TYPE, 2)) ? 2 : 3;
  }
  int b_sub_mb_type() {  // h264_cabac.c:1455-1471
    const int h = mk_hint(HK_SUB_TYPE, 8);
    if (!io_.get(36, h)) return 0;
    if (!io_.get(37, h)) return 1 + io_.get(39, h);
    int type = 3;
    if (io_.get(38, h)) {
      if (io_.get(39, h)) return 11 + io_.get(39, h);
      type += 4;
    }
    type += 2 * io_.get(39, h);
    type += io_.get(39, h);
    return type;
  }
  int ref_idx(int list, int n, int rc) {  // h264_cabac.c:1473-1500
    int refa, refb; bool da, db;
    ref_neighbor(list, n, false, &refa, &da);
    ref_neighbor(list, n, true, &refb, &db);
    int ctx = 0;
    if (d_.slice_type == AVR_SLICE_B) { if (refa > 0 && !da) ctx++; if (refb > 0 && !db) ctx += 2; }
    else { if (refa > 0) ctx++; if (refb > 0) ctx += 2; }
    int ref = 0;
    while (io_.get(54 + ctx, mk_hint(HK_REF, rc - 1 - ref))) {
      ref++;
      ctx = (ctx >> 2) + 4;
      if (ref >= 32) throw Corrupt("ref_idx overflow");
    }
    return ref;
  }
  int mvd_component(int ctxbase, int amvd, int* stored) {  // h264_cabac.c:1502-1537
    if (!io_.get(ctxbase + (amvd > 2) + (amvd > 32), mk_hint(HK_MVD_NZ))) { *stored = 0; return 0; }
    int mvd = 1;
    ctxbase += 3;
    while (mvd < 9 && io_.get(ctxbase, mk_hint(HK_MVD_UNARY, mvd))) { if (mvd < 4) ctxbase++; mvd++; }
    if (mvd >= 9) {
      int k = 3;
      while (io_.bypass(mk_hint(HK_MVD_ESC_PREFIX, k))) { mvd += 1 << k; k++; if (k > 24) throw Corrupt("mvd overflow"); }
      while (k--) mvd += io_.bypass(mk_hint(HK_MVD_ESC_SUFFIX)) << k;
      *stored = mvd < 70 ? mvd : 70;
    } else *stored = mvd;
    return io_.bypass(mk_hint(HK_MVD_SIGN)) ? -mvd : mvd;
  }
  void mvd_pair(int list, int n, int w, int h) {  // one partition's mvd, stored over w x h 4x4 blocks
    int a0 = mvd_neighbor_sum(list, n, 0), a1 = mvd_neighbor_sum(list, n, 1);
    int s0, s1;
    mvd_component(40, a0, &s0);
    mvd_component(47, a1, &s1);
    int x, y;
    blk_xy(n, &x, &y);
    for (int j = 0; j < h; j++)
      for (int i = 0; i < w; i++) {
        cur_->mvd[list][blk_of(x + i, y + j)][0] = uint8_t(s0);
        cur_->mvd[list][blk_of(x + i, y + j)][1] = uint8_t(s1);
      }
  }
  int read_ref(int list, int n) {  // "rc > 1 ? decode : 0" idiom, e.g. h264_cabac.c:2232-2240
    int rc = list ? d_.num_ref_idx_l1 : d_.num_ref_idx_l0;
    if (rc > 1) {
      int r = ref_idx(list, n, rc);
      if (r >= rc) throw Corrupt("reference index out of range");
      return r;
    }
    return 0;
  }

  int cbf_ctx(int cat, int n, bool is_dc, bool cur_intra) const {  // h264_cabac.c:1550-1580
    static const uint16_t base_ctx[14] = {85, 89, 93, 97, 101, 1012, 460, 464, 468, 1016, 472, 476, 480, 1020};
    int nza, nzb;
    if (is_dc) {
      if (cat == 3) { int i = n - 49; nza = (left_cbp_ >> (6 + i)) & 1; nzb = (top_cbp_ >> (6 + i)) & 1; }
      else { int i = n - 48; nza = left_cbp_ & (0x100 << i); nzb = top_cbp_ & (0x100 << i); }
    } else {
      nza = nnz_neighbor(n, false, cur_intra);
      nzb = nnz_neighbor(n, true, cur_intra);
    }
    return base_ctx[cat] + (nza > 0) + 2 * (nzb > 0);
  }
  int cbf(int cat, int n, bool is_dc, bool cur_intra) { return io_.get(cbf_ctx(cat, n, is_dc, cur_intra), mk_hint(HK_CBF, cat)); }

  // decode_cabac_residual_internal, h264_cabac.c:1582-1780 (frame MBs: MB_FIELD == 0)
  int residual_block(int cat, int n, int max_coeff, bool is_dc) {
    static const int sig_off[14] = {105 + 0, 105 + 15, 105 + 29, 105 + 44, 105 + 47, 402, 484 + 0, 484 + 15, 484 + 29, 660, 528 + 0, 528 + 15, 528 + 29, 718};
    static const int last_off[14] = {166 + 0, 166 + 15, 166 + 29, 166 + 44, 166 + 47, 417, 572 + 0, 572 + 15, 572 + 29, 690, 616 + 0, 616 + 15, 616 + 29, 748};
    static const int abs_off[14] = {227 + 0, 227 + 10, 227 + 20, 227 + 30, 227 + 39, 426, 952 + 0, 952 + 10, 952 + 20, 708, 982 + 0, 982 + 10, 982 + 20, 766};
    static const uint8_t level1_ctx[8] = {1, 2, 3, 4, 0, 0, 0, 0};
    static const uint8_t levelgt1_ctx[8] = {5, 5, 5, 5, 6, 7, 8, 9};
    static const uint8_t trans[2][8] = {{1, 2, 3, 3, 4, 5, 6, 7}, {4, 4, 4, 4, 5, 6, 7, 7}};
    io_.begin_block(cat, n, max_coeff, is_dc);
    io_.begin_sigmap();
    int coeff_count = 0, last;
    bool is8 = !is_dc && max_coeff == 64;
    int coefs = is8 ? 63 : max_coeff - 1;
    for (last = 0; last < coefs; last++) {  // DECODE_SIGNIFICANCE :1667-1681
      int so = is8 ? avr_sig8x8_inc[last] : last;
      if (io_.get(sig_off[cat] + so, mk_hint(HK_SIG, last | (max_coeff << 8) | (coeff_count << 16)))) {
        int lo = is8 ? avr_last8x8_inc[last] : last;
        coeff_count++;
        if (io_.get(last_off[cat] + lo, mk_hint(HK_LAST, last | (max_coeff << 8) | (coeff_count << 16)))) { last = max_coeff; break; }
      }
    }
    if (last == max_coeff - 1) coeff_count++;
    io_.end_sigmap();
    int node = 0;
    for (int c = coeff_count; c > 0; c--) {  // STORE_BLOCK :1724-1765
      if (io_.get(abs_off[cat] + level1_ctx[node], mk_hint(HK_LVL_GT1, c)) == 0) {
        node = trans[0][node];
        io_.bypass(mk_hint(HK_SIGN));
      } else {
        int ctx = abs_off[cat] + levelgt1_ctx[node];
        node = trans[1][node];
        int coeff_abs = 2;
        while (coeff_abs < 15 && io_.get(ctx, mk_hint(HK_LVL_MORE, coeff_abs))) coeff_abs++;
        if (coeff_abs >= 15) {
          int j = 0;
          while (io_.bypass(mk_hint(HK_LVL_ESC_PREFIX, j)) && j < 30) j++;
          while (j--) io_.bypass(mk_hint(HK_LVL_ESC_SUFFIX));
        }
        io_.bypass(mk_hint(HK_SIGN));
      }
    }
    io_.end_block();
    return coeff_count;
  }

  void set_nnz8x8(int n, int v) { for (int i = 0; i < 4; i++) cur_->nnz[n + i] = uint8_t(v); }

  // decode_cabac_luma_residual, h264_cabac.c:1874-1920 (p == 0) + cbf wrappers :1825-1872
  void luma_residual(uint32_t mb_type, int cbp) {
    bool intra = is_intra(mb_type);
    if (mb_type & MB_I16x16) {
      if (cbf(0, 48, true, intra)) {
        residual_block(0, 48, 16, true);
        cur_->cbp |= 0x100;  // :1710
      }
      if (cbp & 15) {
        for (int i = 0; i < 16; i++) cur_->nnz[i] = cbf(1, i, false, intra) ? uint8_t(residual_block(1, i, 15, false)) : 0;
      } else memset(cur_->nnz, 0, 16);
    } else {
      for (int i8 = 0; i8 < 4; i8++) {
        if (cbp & (1 << i8)) {
          if (mb_type & MB_8x8DCT) {
            set_nnz8x8(4 * i8, residual_block(5, 4 * i8, 64, false));  // no cbf for cat 5 unless 4:4:4 (:1863)
          } else {
            for (int i4 = 0; i4 < 4; i4++) {
              int n = 4 * i8 + i4;
              cur_->nnz[n] = cbf(2, n, false, intra) ? uint8_t(residual_block(2, n, 16, false)) : 0;
            }
          }
        } else set_nnz8x8(4 * i8, 0);
      }
    }
  }

  void intra4x4_pred_mode(bool can_rem) {  // h264_cabac.c:1369-1381 (value irrelevant for context selection)
    if (io_.get(68, mk_hint(HK_I4_FLAG, can_rem))) return;
    io_.get(69, mk_hint(HK_I4_REM)); io_.get(69, mk_hint(HK_I4_REM)); io_.get(69, mk_hint(HK_I4_REM));
  }

  // ff_h264_decode_mb_cabac, h264_cabac.c:1926-2480 (non-MBAFF)
  void macroblock() {
    cur_ = &pic_.mbs[size_t(mb_x_) + size_t(mb_y_) * pic_.mbw];
    uint32_t mb_type = 0;
    int partition_count = 0, cbp = 0;
    bool dct8x8_allowed = d_.flags & AVR_SLICE_TRANSFORM_8X8;
    const bool is_b = d_.slice_type == AVR_SLICE_B;

    if (d_.slice_type != AVR_SLICE_I) {
      if (mb_skip_flag()) {  // decode_mb_skip, h264_mvpred.h:803-834 + h264_cabac.c:1953-1959
        *cur_ = MbSyntax();
        cur_->slice_id = slice_id_;
        cur_->type = is_b ? (MB_DIRECT | MB_SKIP | MB_P0L0 | MB_P1L0 | MB_P0L1 | MB_P1L1)
                          : (MB_16x16 | MB_P0L0 | MB_P1L0 | MB_SKIP);
        last_qp_delta_ = 0;
        return;
      }
    }
    left_ = avail(mb_x_ - 1, mb_y_);
    top_ = avail(mb_x_, mb_y_ - 1);
    left_type_ = left_ ? left_->type : 0;
    top_type_ = top_ ? top_->type : 0;
    *cur_ = MbSyntax();
    bool intra_path = false;
    int itype = 0;
    if (is_b) {  // :1973-2007
      int ctx = 0;
      if (left_ && !(left_type_ & MB_DIRECT)) ctx++;
      if (top_ && !(top_type_ & MB_DIRECT)) ctx++;
      int t = 0;
      const int h = mk_hint(HK_MBTYPE, 2);
      if (!io_.get(27 + ctx, h)) t = 0;
      else if (!io_.get(27 + 3, h)) t = 1 + io_.get(27 + 5, h);
      else {
        int bits = io_.get(27 + 4, h) << 3;
        bits += io_.get(27 + 5, h) << 2;
        bits += io_.get(27 + 5, h) << 1;
        bits += io_.get(27 + 5, h);
        if (bits < 8) t = bits + 3;
        else if (bits == 13) { itype = intra_mb_type(32, false); intra_path = true; }
        else if (bits == 14) t = 11;
        else if (bits == 15) t = 22;
        else { bits = (bits << 1) + io_.get(27 + 5, h); t = bits - 4; }
      }
      if (!intra_path) b_type_info(t, &mb_type, &partition_count);
    } else if (d_.slice_type == AVR_SLICE_P) {  // :2008-2023
      if (io_.get(14, mk_hint(HK_MBTYPE, 10)) == 0) {
        int t;
        if (io_.get(15, mk_hint(HK_MBTYPE, 11)) == 0) t = 3 * io_.get(16, mk_hint(HK_MBTYPE, 12));
        else t = 2 - io_.get(17, mk_hint(HK_MBTYPE, 13));
        static const uint32_t ptype[4] = {MB_16x16 | MB_P0L0, MB_16x8 | MB_P0L0 | MB_P1L0, MB_8x16 | MB_P0L0 | MB_P1L0,
                                          MB_8x8 | MB_P0L0 | MB_P1L0};
        static const int pcount[4] = {1, 2, 2, 4};
        mb_type = ptype[t]; partition_count = pcount[t];
      } else { itype = intra_mb_type(17, false); intra_path = true; }
    } else { itype = intra_mb_type(3, true); intra_path = true; }

    if (intra_path) {  // decode_intra_mb :2029-2034 ; i_mb_type_info h264data.h:79-106
      partition_count = 0;
      if (itype == 0) mb_type = MB_I4x4;
      else if (itype == 25) mb_type = MB_IPCM;
      else {
        mb_type = MB_I16x16;
        int k = itype - 1;
        int chroma = (k >> 2) % 3, luma = k >= 12;
        cbp = (luma ? 15 : 0) + 16 * chroma;
      }
    }
    cur_->slice_id = slice_id_;
    if (mb_type & MB_IPCM) throw Unsupported("I_PCM");  // SURVEY hazard H3: reference cannot handle it either

    // fill_decode_caches bits that feed contexts (h264_mvpred.h:589-603)
    bool intra = is_intra(mb_type);
    top_cbp_ = top_ ? top_->cbp : (intra ? 0x7CF : 0x00F);
    left_cbp_ = left_ ? ((left_->cbp & 0x7F0) | (left_->cbp & 2) | (((left_->cbp >> 2) & 2) << 2)) : (intra ? 0x7CF : 0x00F);
    int neighbor_transform_size = ((top_type_ & MB_8x8DCT) ? 1 : 0) + ((left_type_ & MB_8x8DCT) ? 1 : 0);  // :797
    for (int l = 0; l < 2; l++) for (int i = 0; i < 4; i++) cur_ref_[l][i] = -1;
    for (int i = 0; i < 4; i++) cur_direct_[i] = false;
    int list_count = is_b ? 2 : 1;

    if (intra) {
      if (mb_type & MB_I4x4) {  // :2081-2099
        bool can_rem = left_ && top_;
        if (dct8x8_allowed && io_.get(399 + neighbor_transform_size, mk_hint(HK_T8X8, 1))) {
          mb_type |= MB_8x8DCT;
          for (int i = 0; i < 4; i++) intra4x4_pred_mode(can_rem);
        } else {
          for (int i = 0; i < 16; i++) intra4x4_pred_mode(can_rem);
        }
      }
      cur_->chroma_pred = uint8_t(chroma_pred_mode());  // :2104-2110
    } else if (partition_count == 4) {  // :2114-2220
      uint32_t sub[4]; int subcount[4];
      if (is_b) {
        for (int i = 0; i < 4; i++) b_sub_type_info(b_sub_mb_type(), &sub[i], &subcount[i]);
        for (int i = 0; i < 4; i++) { cur_direct_[i] = (sub[i] & MB_DIRECT) != 0; cur_->direct8[i] = cur_direct_[i]; }
      } else {
        static const uint32_t pst[4] = {MB_16x16 | MB_P0L0, MB_16x8 | MB_P0L0, MB_8x16 | MB_P0L0, MB_8x8 | MB_P0L0};
        static const int psc[4] = {1, 2, 2, 4};
        for (int i = 0; i < 4; i++) { int t = p_sub_mb_type(); sub[i] = pst[t]; subcount[i] = psc[t]; }
      }
      for (int list = 0; list < list_count; list++)
        for (int i = 0; i < 4; i++) {
          if (sub[i] & MB_DIRECT) continue;  // ref_cache keeps what direct prediction put there; never read for ctx
          if (is_dir(sub[i], 0, list)) cur_ref_[list][i] = read_ref(list, 4 * i);
          else cur_ref_[list][i] = -1;
        }
      if (dct8x8_allowed) {  // get_dct8x8_allowed, h264.h:1165-1177
        for (int i = 0; i < 4; i++) {
          if (sub[i] & (MB_16x8 | MB_8x16 | MB_8x8)) dct8x8_allowed = false;
          if (!(d_.flags & AVR_SLICE_DIRECT_8X8_INFERENCE) && (sub[i] & MB_DIRECT)) dct8x8_allowed = false;
        }
      }
      for (int list = 0; list < list_count; list++)
        for (int i = 0; i < 4; i++) {
          if ((sub[i] & MB_DIRECT) || !is_dir(sub[i], 0, list)) continue;  // mvd stays 0
          int bw = (sub[i] & (MB_16x16 | MB_16x8)) ? 2 : 1;
          for (int j = 0; j < subcount[i]; j++) {
            int n = 4 * i + bw * j;
            if (sub[i] & MB_16x16) mvd_pair(list, n, 2, 2);
            else if (sub[i] & MB_16x8) mvd_pair(list, n, 2, 1);
            else if (sub[i] & MB_8x16) mvd_pair(list, n, 1, 2);
            else mvd_pair(list, n, 1, 1);
          }
        }
      for (int list = 0; list < 2; list++) for (int i = 0; i < 4; i++) cur_->ref[list][i] = int8_t(cur_ref_[list][i]);
    } else if (mb_type & MB_DIRECT) {  // :2221-2225
      dct8x8_allowed = dct8x8_allowed && (d_.flags & AVR_SLICE_DIRECT_8X8_INFERENCE);
      for (int i = 0; i < 4; i++) cur_direct_[i] = true;
    } else {  // 16x16 / 16x8 / 8x16 :2226-2327
      if (mb_type & MB_16x16) {
        for (int list = 0; list < list_count; list++)
          if (is_dir(mb_type, 0, list)) { int r = read_ref(list, 0); for (int i = 0; i < 4; i++) cur_ref_[list][i] = r; }
        for (int list = 0; list < list_count; list++)
          if (is_dir(mb_type, 0, list)) mvd_pair(list, 0, 4, 4);
      } else if (mb_type & MB_16x8) {
        for (int list = 0; list < list_count; list++)
          for (int i = 0; i < 2; i++)
            if (is_dir(mb_type, i, list)) { int r = read_ref(list, 8 * i); cur_ref_[list][2 * i] = cur_ref_[list][2 * i + 1] = r; }
        for (int list = 0; list < list_count; list++)
          for (int i = 0; i < 2; i++)
            if (is_dir(mb_type, i, list)) mvd_pair(list, 8 * i, 4, 2);
      } else {
        for (int list = 0; list < list_count; list++)
          for (int i = 0; i < 2; i++)
            if (is_dir(mb_type, i, list)) { int r = read_ref(list, 4 * i); cur_ref_[list][i] = cur_ref_[list][i + 2] = r; }
        for (int list = 0; list < list_count; list++)
          for (int i = 0; i < 2; i++)
            if (is_dir(mb_type, i, list)) mvd_pair(list, 4 * i, 2, 4);
      }
      for (int list = 0; list < 2; list++) for (int i = 0; i < 4; i++) cur_->ref[list][i] = int8_t(cur_ref_[list][i]);
    }

    if (!(mb_type & MB_I16x16)) {  // :2335-2339
      cbp = cbp_luma();
      cbp |= cbp_chroma() << 4;
    }
    cur_->cbp = uint16_t(cbp);  // :2346
    if (dct8x8_allowed && (cbp & 15) && !intra)  // :2348-2350
      if (io_.get(399 + neighbor_transform_size, mk_hint(HK_T8X8, 0))) mb_type |= MB_8x8DCT;
    cur_->type = mb_type;  // :2374

    if (cbp || (mb_type & MB_I16x16)) {  // :2376-2468
      if (io_.get(60 + (last_qp_delta_ != 0), mk_hint(HK_QPD_NZ))) {  // mb_qp_delta :2388-2416
        int val = 1, ctx = 2;
        while (io_.get(60 + ctx, mk_hint(HK_QPD_MORE, val))) { ctx = 3; val++; if (val > 2 * 51) throw Corrupt("mb_qp_delta overflow"); }
        val = (val & 1) ? (val + 1) >> 1 : -((val + 1) >> 1);
        last_qp_delta_ = val;
        qp_ += val;
        if (unsigned(qp_) > 51u) { if (qp_ < 0) qp_ += 52; else qp_ -= 52; }
      } else last_qp_delta_ = 0;
      luma_residual(mb_type, cbp);
      if (cbp & 0x30) {  // chroma DC :2449-2453
        for (int c = 0; c < 2; c++)
          if (cbf(3, 49 + c, true, intra)) { residual_block(3, 49 + c, 4, true); cur_->cbp |= uint16_t(0x40 << c); }
      }
      if (cbp & 0x20) {  // chroma AC :2455-2463
        for (int c = 0; c < 2; c++)
          for (int i = 0; i < 4; i++) {
            int n = 16 + 16 * c + i;
            cur_->nnz[n] = cbf(4, n, false, intra) ? uint8_t(residual_block(4, n, 15, false)) : 0;
          }
      }
    } else {
      last_qp_delta_ = 0;  // :2473
    }
  }

  // b_mb_type_info / b_sub_mb_type_info, h264data.h:127-170, rebuilt from the spec's partition table
  static void b_type_info(int t, uint32_t* type, int* count) {
    // pred codes per partition: 1 = L0, 2 = L1, 3 = Bi
    static const uint8_t shape[23] = {0, 1, 1, 1, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 4};
    static const uint8_t p0[23] = {0, 1, 2, 3, 1, 1, 2, 2, 1, 1, 2, 2, 1, 1, 2, 2, 3, 3, 3, 3, 3, 3, 3};
    static const uint8_t p1[23] = {0, 0, 0, 0, 1, 1, 2, 2, 2, 2, 1, 1, 3, 3, 3, 3, 1, 1, 2, 2, 3, 3, 3};
    uint32_t ty = 0;
    switch (shape[t]) {
      case 0: *type = MB_DIRECT | MB_P0L0 | MB_P1L0 | MB_P0L1 | MB_P1L1; *count = 1; return;
      case 1: ty = MB_16x16; *count = 1; break;
      case 2: ty = MB_16x8; *count = 2; break;
      case 3: ty = MB_8x16; *count = 2; break;
      default: *type = MB_8x8 | MB_P0L0 | MB_P1L0 | MB_P0L1 | MB_P1L1; *count = 4; return;
    }
    if (p0[t] & 1) ty |= MB_P0L0;
    if (p0[t] & 2) ty |= MB_P0L1;
    if (p1[t] & 1) ty |= MB_P1L0;
    if (p1[t] & 2) ty |= MB_P1L1;
    *type = ty;
  }
  static void b_sub_type_info(int t, uint32_t* type, int* count) {
    static const uint8_t shape[13] = {0, 1, 1, 1, 2, 3, 2, 3, 2, 3, 4, 4, 4};
    static const uint8_t pred[13] = {0, 1, 2, 3, 1, 1, 2, 2, 3, 3, 1, 2, 3};
    static const uint32_t sh[5] = {MB_DIRECT, MB_16x16, MB_16x8, MB_8x16, MB_8x8};
    static const int cnt[5] = {1, 1, 2, 2, 4};
    uint32_t ty = sh[shape[t]];
    if (pred[t] & 1) ty |= MB_P0L0 | (shape[t] >= 2 ? MB_P1L0 : 0);
    if (pred[t] & 2) ty |= MB_P0L1 | (shape[t] >= 2 ? MB_P1L1 : 0);
    *type = ty; *count = cnt[shape[t]];
  }
};

}  // namespace avr_oracle
