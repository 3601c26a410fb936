// h264gen -- synthetic CABAC H.264 (Annex-B) stream generator.  TEST INFRASTRUCTURE, not product.
//
// There is no sample video and no H.264 encoder on the box, so the parity
// and bench inputs are synthesised at the SYNTAX level: legal SPS/PPS/slice headers followed by
// random-but-legal slice_data.  The slice_data is produced by running the oracle's slice walker
// (oracle/avr_oracle.hpp, SliceWalker) "in reverse": every time the walker asks for a bin, a
// seeded policy picks its value (shaped by the hint the walker attaches) and a textbook CABAC
// encoder (H.264 9.3.4.2-9.3.4.5, same algorithm as ffmpeg/libavcodec/cabac.c:255-332) writes it.
// Pixel content is irrelevant (libavcodec reconstructs garbage happily); what matters is legal
// syntax and controllable statistics.  Every generated stream used by the tests is first checked
// against the reference binary itself (`recode roundtrip`), see tests/make_golden.py.
//
// Hazards of the reference that the generator avoids: frame_num changes on
// every picture (H2), no I_PCM (H3); only picture 0 is IDR.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <string>
#include <vector>
#include "../oracle/avr_oracle.hpp"

using namespace avr_oracle;

struct Rng {  // splitmix64 / xorshift*
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567) { next(); next(); }
  uint64_t next() { s ^= s >> 12; s ^= s << 25; s ^= s >> 27; return s * 0x2545F4914F6CDD1Dull; }
  double uni() { return double(next() >> 11) * (1.0 / 9007199254740992.0); }
  bool bern(double p) { return uni() < p; }
  int below(int n) { return int(next() % uint64_t(n)); }
};

struct BitWriter {
  std::vector<uint8_t> bytes;
  int nbits = 0;
  void put(int bit) {
    if ((nbits & 7) == 0) bytes.push_back(0);
    if (bit) bytes.back() |= uint8_t(0x80 >> (nbits & 7));
    nbits++;
  }
  void u(int n, uint32_t v) { for (int i = n - 1; i >= 0; i--) put((v >> i) & 1); }
  void ue(uint32_t v) { uint32_t x = v + 1; int len = 0; while ((x >> len) > 1) len++; for (int i = 0; i < len; i++) put(0); u(len + 1, x); }
  void se(int v) { ue(v > 0 ? uint32_t(2 * v - 1) : uint32_t(-2 * v)); }
  bool aligned() const { return (nbits & 7) == 0; }
  void trailing() { put(1); while (!aligned()) put(0); }
};

// CABAC arithmetic encoder, H.264 9.3.4.2 (EncodeDecision), 9.3.4.3 (RenormE/PutBit), 9.3.4.4 (bypass),
// 9.3.4.5 (terminate + flush).  State byte = 2*pStateIdx + valMPS.
struct CabacWriter {
  BitWriter* bw;
  uint32_t low = 0, range = 510;
  int outstanding = 0;
  bool first = true;
  void put_bit(int b) {
    if (first) first = false; else bw->put(b);
    while (outstanding > 0) { bw->put(1 - b); outstanding--; }
  }
  void renorm() {
    while (range < 256) {
      if (low < 256) put_bit(0);
      else if (low >= 512) { low -= 512; put_bit(1); }
      else { low -= 256; outstanding++; }
      range <<= 1; low <<= 1;
    }
  }
  void decision(uint8_t* state, int bin) {
    int p = *state >> 1, m = *state & 1;
    uint32_t rlps = avr_range_lps[p * 4 + ((range >> 6) & 3)];
    range -= rlps;
    if (bin != m) { low += range; range = rlps; if (p == 0) m ^= 1; p = avr_trans_lps[p]; }
    else if (p < 62) p++;
    *state = uint8_t(2 * p + m);
    renorm();
  }
  void bypass(int bin) {
    low <<= 1;
    if (bin) low += range;
    if (low >= 1024) { put_bit(1); low -= 1024; }
    else if (low < 512) put_bit(0);
    else { low -= 512; outstanding++; }
  }
  void terminate(int bin) {
    range -= 2;
    if (bin) {
      low += range;
      range = 2;
      renorm();
      put_bit((low >> 9) & 1);
      bw->u(2, ((low >> 7) & 3) | 1);  // final 1 = rbsp_stop_one_bit
    } else renorm();
  }
};

struct GenConfig {
  int mbw = 80, mbh = 45, frames = 30, slices = 1;
  bool high = false;       // High profile (transform_8x8_mode = 1)
  double p_t8x8 = 0.5;     // P(transform_size_8x8_flag) when allowed
  int intra_period = 0;    // 0 = only the first picture is intra; 1 = all intra; n = I every n pictures
  int qp = 26;
  double density = 1.0;    // scales residual density
  double target_bytes_per_frame = 0;  // > 0: adapt density picture by picture
  double p_skip = 0.35, p_intra_in_p = 0.08;
  double p_dense = 0.0;    // P(4x4 block forced fully significant) -- exercises SURVEY hazard H1
  int num_refs = 2;
  double p_bpic = 0.0;     // P(a non-intra picture is coded as B slices); B pictures stay reference pictures
  bool ragged_slices = false;  // slice boundaries at arbitrary macroblocks instead of row starts
  bool rich_headers = false;   // slice headers with ref_pic_list_modification, pred_weight_table and memory management
                               // control operations 1, 2, 4, 6 -- syntax the slice indexer has to get past exactly
  uint64_t seed = 1;
};

// Bin policy + CABAC writer behind the walker's IO interface.
class GeneratorIO {
 public:
  GeneratorIO(const GenConfig& c, Rng* rng, BitWriter* bw, int planned_mbs, const std::vector<float>* act,
              int mbw, double density)
      : cfg_(c), rng_(*rng), planned_mbs_(planned_mbs), act_(act), mbw_(mbw), density_(density) { cw_.bw = bw; }
  uint8_t state[1024];
  uint64_t n_bins = 0;

  int get(int ctx, int hint = 0) { int b = choose(hint); cw_.decision(&state[ctx], b); n_bins++; return b; }
  int bypass(int hint = 0) { int b = choose(hint); cw_.bypass(b); n_bins++; return b; }
  int terminate(int hint = 0) { int b = choose(hint); cw_.terminate(b); n_bins++; return b; }
  void mb_xy(int x, int y) { a_ = (*act_)[size_t(x) + size_t(y) * mbw_]; }
  void begin_block(int cat, int, int max_coeff, int) { cat_ = cat; dense_ = max_coeff <= 16 && rng_.bern(cfg_.p_dense); }
  void end_block() {}
  void begin_sigmap() {}
  void end_sigmap() {}

 private:
  int choose(int hint) {
    int kind = hint & 0xFF, arg = hint >> 8;
    double a = a_ * density_;  // local activity in roughly [0, 2]
    switch (kind) {
      case HK_SKIP: return rng_.bern(clamp01(cfg_.p_skip * (1.6 - a)));
      case HK_MBTYPE:
        if (arg == 100) return rng_.bern(0.3);        // I: 0 = I_NxN, 1 = I16x16
        if (arg == 10) return rng_.bern(cfg_.p_intra_in_p);  // P: 1 = intra
        if (arg == 11) return rng_.bern(0.35);
        if (arg == 12) return rng_.bern(0.4);          // 16x16 vs 8x8
        if (arg == 13) return rng_.bern(0.5);
        return rng_.bern(0.5);                         // B mb_type tree (uniform)
      case HK_I16_CBPL: return rng_.bern(clamp01(0.5 * a));
      case HK_I16_CBPC: return rng_.bern(arg ? 0.5 : clamp01(0.5 * a));
      case HK_I16_PRED_HI: {
        int mask = arg & 15, opts[4], n = 0;
        for (int i = 0; i < 4; i++) if (mask & (1 << i)) opts[n++] = i;
        pick_ = opts[rng_.below(n)];
        return pick_ >> 1;
      }
      case HK_I16_PRED_LO: return pick_ & 1;
      case HK_PCM_TERM: return 0;
      case HK_CHROMA_PRED: {
        int stage = arg >> 4, mask = arg & 15;
        if (stage == 0) {
          int opts[4], n = 0;
          for (int i = 0; i < 4; i++) if (mask & (1 << i)) opts[n++] = i;
          pick_ = rng_.bern(0.5) ? 0 : opts[rng_.below(n)];
          return pick_ != 0;
        }
        if (stage == 1) return pick_ >= 2;
        return pick_ == 3;
      }
      case HK_I4_FLAG: return arg ? rng_.bern(0.7) : 1;
      case HK_I4_REM: return rng_.bern(0.5);
      case HK_SUB_TYPE: return rng_.bern(arg == 0 ? 0.6 : 0.5);
      case HK_REF: return arg > 0 ? rng_.bern(0.3) : 0;
      case HK_MVD_NZ: return rng_.bern(0.55);
      case HK_MVD_UNARY: return rng_.bern(0.62);
      case HK_MVD_ESC_PREFIX: return arg < 6 ? rng_.bern(0.3) : 0;
      case HK_MVD_ESC_SUFFIX: return rng_.bern(0.5);
      case HK_MVD_SIGN: return rng_.bern(0.5);
      case HK_QPD_NZ: return rng_.bern(0.08);
      case HK_QPD_MORE: return arg < 6 ? rng_.bern(0.4) : 0;
      case HK_CBP_LUMA: return rng_.bern(clamp01(0.15 + 0.55 * a));
      case HK_CBP_CHROMA_NZ: return rng_.bern(clamp01(0.1 + 0.45 * a));
      case HK_CBP_CHROMA_2: return rng_.bern(0.5);
      case HK_T8X8: return rng_.bern(cfg_.p_t8x8);
      case HK_CBF: return rng_.bern(clamp01(0.25 + 0.5 * a));
      case HK_SIG: {
        if (dense_) return 1;
        int pos = arg & 0xFF, maxc = (arg >> 8) & 0xFF;
        double rel = double(pos) / double(maxc);
        return rng_.bern(clamp01((0.2 + 0.65 * a) * exp(-2.2 * rel)));
      }
      case HK_LAST: {
        if (dense_) return 0;
        int pos = arg & 0xFF, maxc = (arg >> 8) & 0xFF;
        double rel = double(pos) / double(maxc);
        return rng_.bern(clamp01(0.12 + 0.5 * rel + 0.25 * (1.0 - clamp01(a))));
      }
      case HK_LVL_GT1: return rng_.bern(clamp01(0.15 + 0.3 * a));
      case HK_LVL_MORE: return rng_.bern(clamp01(0.35 + 0.25 * a));
      case HK_LVL_ESC_PREFIX: return arg < 5 ? rng_.bern(0.5) : 0;
      case HK_LVL_ESC_SUFFIX: return rng_.bern(0.5);
      case HK_SIGN: return rng_.bern(0.5);
      case HK_EOS: return arg >= planned_mbs_;
      default: return rng_.bern(0.5);
    }
  }
  static double clamp01(double v) { return v < 0.02 ? 0.02 : v > 0.98 ? 0.98 : v; }
  const GenConfig& cfg_;
  Rng& rng_;
  CabacWriter cw_;
  int planned_mbs_;
  const std::vector<float>* act_;
  int mbw_;
  double density_;
  double a_ = 1.0;
  int cat_ = 0, pick_ = 0;
  bool dense_ = false;
};

static void write_nal(std::vector<uint8_t>* out, int ref_idc, int type, const std::vector<uint8_t>& rbsp, bool long_sc) {
  if (long_sc) out->push_back(0);
  out->push_back(0); out->push_back(0); out->push_back(1);
  out->push_back(uint8_t((ref_idc << 5) | type));
  int zeros = 0;
  for (uint8_t b : rbsp) {  // emulation prevention, H.264 7.4.1
    if (zeros >= 2 && b <= 3) { out->push_back(3); zeros = 0; }
    out->push_back(b);
    zeros = b == 0 ? zeros + 1 : 0;
  }
}

static std::vector<uint8_t> make_sps(const GenConfig& c) {
  BitWriter b;
  b.u(8, c.high ? 100 : 77); b.u(8, 0); b.u(8, 51);
  b.ue(0);  // sps id
  if (c.high) { b.ue(1); b.ue(0); b.ue(0); b.u(1, 0); b.u(1, 0); }  // chroma 4:2:0, 8 bit, no qpprime bypass, no scaling lists
  b.ue(0);            // log2_max_frame_num_minus4 -> frame_num is u(4)
  b.ue(2);            // pic_order_cnt_type 2
  b.ue(uint32_t(c.num_refs));  // max_num_ref_frames
  b.u(1, 0);          // gaps
  b.ue(uint32_t(c.mbw - 1)); b.ue(uint32_t(c.mbh - 1));
  b.u(1, 1);          // frame_mbs_only
  b.u(1, 1);          // direct_8x8_inference
  b.u(1, 0);          // cropping
  b.u(1, 0);          // vui
  b.trailing();
  return b.bytes;
}
static std::vector<uint8_t> make_pps(const GenConfig& c) {
  BitWriter b;
  b.ue(0); b.ue(0);
  b.u(1, 1);   // entropy_coding_mode_flag = CABAC
  b.u(1, 0);   // bottom_field_pic_order_in_frame_present
  b.ue(0);     // slice groups
  b.ue(0); b.ue(0);  // num_ref_idx_default_active_minus1
  b.u(1, c.rich_headers ? 1 : 0); b.u(2, c.rich_headers ? 1 : 0);  // weighted_pred_flag / weighted_bipred_idc (1 = explicit)
  b.se(0); b.se(0); b.se(0);  // pic_init_qp-26, qs, chroma_qp_index_offset
  b.u(1, 1);   // deblocking_filter_control_present
  b.u(1, 0);   // constrained_intra_pred
  b.u(1, 0);   // redundant_pic_cnt_present
  if (c.high) { b.u(1, 1); b.u(1, 0); b.se(0); }  // transform_8x8_mode, no scaling matrix, second chroma qp offset
  b.trailing();
  return b.bytes;
}

struct GenStats { uint64_t bins = 0, payload_bytes = 0, slices = 0, mbs = 0; };

static std::vector<uint8_t> generate(const GenConfig& c, GenStats* st) {
  Rng rng(c.seed);
  std::vector<uint8_t> out;
  write_nal(&out, 3, 7, make_sps(c), true);
  write_nal(&out, 3, 8, make_pps(c), true);
  const int nmb = c.mbw * c.mbh;
  std::vector<float> act(size_t(nmb), 1.0f), prev_act(size_t(nmb), 1.0f);
  double density = c.density;
  PictureSyntax pic;
  for (int f = 0; f < c.frames; f++) {
    bool intra = f == 0 || c.intra_period == 1 || (c.intra_period > 1 && f % c.intra_period == 0);
    bool idr = f == 0;
    const bool bpic = !intra && f >= 2 && c.num_refs >= 2 && c.p_bpic > 0 && rng.bern(c.p_bpic);
    // smooth activity field, correlated in space and time, so that avrecode's priors have something to learn
    for (int y = 0; y < c.mbh; y++)
      for (int x = 0; x < c.mbw; x++) {
        double l = x ? act[size_t(y) * c.mbw + x - 1] : 1.0, t = y ? act[size_t(y - 1) * c.mbw + x] : 1.0;
        double p = prev_act[size_t(y) * c.mbw + x];
        double v = 0.36 * l + 0.30 * t + 0.22 * p + 0.12 * (2.6 * rng.uni());
        act[size_t(y) * c.mbw + x] = float(v < 0.05 ? 0.05 : v > 2.0 ? 2.0 : v);
      }
    pic.start(c.mbw, c.mbh);
    // slice boundaries
    std::vector<int> starts;
    for (int s = 0; s < c.slices; s++) {
      int first;
      if (c.ragged_slices && s > 0) first = int((int64_t(nmb) * s) / c.slices) + rng.below(c.mbw / 2 + 1);
      else first = ((c.mbh * s) / c.slices) * c.mbw;
      if (!starts.empty() && first <= starts.back()) first = starts.back() + 1;
      if (first < nmb) starts.push_back(first);
    }
    size_t frame_bytes0 = out.size();
    for (size_t s = 0; s < starts.size(); s++) {
      int first = starts[s], end = s + 1 < starts.size() ? starts[s + 1] : nmb;
      BitWriter b;
      int refs_avail = f < c.num_refs ? f : c.num_refs;
      int nref = intra ? 0 : (refs_avail < 1 ? 1 : refs_avail);
      b.ue(uint32_t(first));
      b.ue(bpic ? 6 : intra ? 7 : 5);    // slice_type (all slices of the picture alike)
      b.ue(0);                // pps id
      b.u(4, uint32_t(f & 15));  // frame_num
      if (idr) b.ue(0);       // idr_pic_id
      if (bpic) b.u(1, 1);    // direct_spatial_mv_pred_flag
      if (!intra) {
        b.u(1, 1); b.ue(uint32_t(nref - 1));  // num_ref_idx_active_override
        if (bpic) b.ue(uint32_t(nref - 1));   // num_ref_idx_l1_active_minus1
        if (!c.rich_headers) {
          b.u(1, 0);            // ref_pic_list_modification_flag_l0
          if (bpic) b.u(1, 0);  // ref_pic_list_modification_flag_l1
        } else {
          // H.264 7.3.3.1: "previous picture first" spelled out (idc 0, abs_diff_pic_num_minus1 0, end) -- a no-op reordering
          for (int l = 0; l < (bpic ? 2 : 1); l++) { b.u(1, 1); b.ue(0); b.ue(0); b.ue(3); }
          // pred_weight_table (7.3.3.2): denominators, then per reference optional luma / chroma weights
          b.ue(uint32_t(rng.below(8)));
          b.ue(uint32_t(rng.below(8)));
          for (int l = 0; l < (bpic ? 2 : 1); l++)
            for (int r = 0; r < nref; r++) {
              const int lf = rng.bern(0.5);
              b.u(1, uint32_t(lf));
              if (lf) { b.se(rng.below(255) - 127); b.se(rng.below(255) - 127); }
              const int cf = rng.bern(0.5);
              b.u(1, uint32_t(cf));
              if (cf) for (int k = 0; k < 4; k++) b.se(rng.below(255) - 127);
            }
        }
      }
      if (idr) { b.u(1, 0); b.u(1, 0); }  // no_output_of_prior_pics, long_term_reference_flag
      else if (!c.rich_headers || intra) b.u(1, 0);   // adaptive_ref_pic_marking_mode_flag
      else {
        // dec_ref_pic_marking (7.3.3.3) with operations 4 + 6 (this picture becomes long-term reference 0), later 2
        // (long-term picture 0 is dropped) and 1 (the oldest short-term picture is dropped).  The pictures still
        // reference whatever libavcodec keeps: only the SYNTAX matters to a stream that is never displayed.
        b.u(1, 1);
        const int phase = f % 4;
        if (phase == 1) { b.ue(4); b.ue(2); b.ue(6); b.ue(0); }
        else if (phase == 3) { b.ue(2); b.ue(0); if (f > c.num_refs + 1 && c.num_refs >= 2) { b.ue(1); b.ue(uint32_t(c.num_refs - 1)); } }
        b.ue(0);
      }
      int cabac_init_idc = intra ? 0 : rng.below(3);
      if (!intra) b.ue(uint32_t(cabac_init_idc));
      int qp = c.qp + rng.below(5) - 2;
      b.se(qp - 26);
      b.ue(1);                // disable_deblocking_filter_idc = 1
      while (!b.aligned()) b.put(1);  // cabac_alignment_one_bit
      size_t hdr_bytes = b.bytes.size();

      avr_slice_desc d;
      memset(&d, 0, sizeof d);
      d.first_mb = uint32_t(first); d.mb_width = uint16_t(c.mbw); d.mb_height = uint16_t(c.mbh);
      d.slice_type = bpic ? AVR_SLICE_B : intra ? AVR_SLICE_I : AVR_SLICE_P;
      d.cabac_init_idc = uint8_t(cabac_init_idc); d.slice_qp = int8_t(qp);
      d.num_ref_idx_l0 = uint8_t(nref); d.num_ref_idx_l1 = uint8_t(bpic ? nref : 0); d.chroma_format_idc = 1;
      d.flags = uint8_t((c.high ? AVR_SLICE_TRANSFORM_8X8 : 0) | AVR_SLICE_DIRECT_8X8_INFERENCE);
      GeneratorIO io(c, &rng, &b, end - first, &act, c.mbw, density);
      SliceWalker<GeneratorIO> w(&io, &pic, d);
      int mbs = w.run();
      while (!b.aligned()) b.put(0);  // rbsp_alignment_zero_bit (stop bit was written by terminate)
      st->bins += io.n_bins; st->mbs += uint64_t(mbs); st->slices++;
      st->payload_bytes += b.bytes.size() - hdr_bytes;
      write_nal(&out, idr ? 3 : 2, idr ? 5 : 1, b.bytes, s == 0);
    }
    if (c.target_bytes_per_frame > 0) {
      double actual = double(out.size() - frame_bytes0);
      double r = c.target_bytes_per_frame / (actual + 1.0);
      r = r < 0.5 ? 0.5 : r > 2.0 ? 2.0 : r;
      density *= pow(r, 0.6);
      if (density < 0.02) density = 0.02;
      if (density > 3.0) density = 3.0;
    }
    prev_act.swap(act);
  }
  return out;
}

static void usage() {
  fprintf(stderr,
          "usage: h264gen -o out.h264 [--width W --height H | --mbw N --mbh N] [--frames N] [--slices N]\n"
          "       [--high] [--p8x8 P] [--intra-period N] [--qp Q] [--density D] [--kbps K --fps F]\n"
          "       [--pskip P] [--pintra P] [--pdense P] [--refs N] [--bpics P] [--ragged] [--rich-headers] [--seed S]\n");
}

int main(int argc, char** argv) {
  GenConfig c;
  std::string outp;
  double kbps = 0, fps = 30;
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i];
    auto next = [&]() -> const char* { if (i + 1 >= argc) { usage(); exit(2); } return argv[++i]; };
    if (a == "-o") outp = next();
    else if (a == "--width") c.mbw = (atoi(next()) + 15) / 16;
    else if (a == "--height") c.mbh = (atoi(next()) + 15) / 16;
    else if (a == "--mbw") c.mbw = atoi(next());
    else if (a == "--mbh") c.mbh = atoi(next());
    else if (a == "--frames") c.frames = atoi(next());
    else if (a == "--slices") c.slices = atoi(next());
    else if (a == "--high") c.high = true;
    else if (a == "--p8x8") c.p_t8x8 = atof(next());
    else if (a == "--intra-period") c.intra_period = atoi(next());
    else if (a == "--qp") c.qp = atoi(next());
    else if (a == "--density") c.density = atof(next());
    else if (a == "--kbps") kbps = atof(next());
    else if (a == "--fps") fps = atof(next());
    else if (a == "--pskip") c.p_skip = atof(next());
    else if (a == "--pintra") c.p_intra_in_p = atof(next());
    else if (a == "--pdense") c.p_dense = atof(next());
    else if (a == "--refs") c.num_refs = atoi(next());
    else if (a == "--bpics") c.p_bpic = atof(next());
    else if (a == "--ragged") c.ragged_slices = true;
    else if (a == "--rich-headers") c.rich_headers = true;
    else if (a == "--seed") c.seed = strtoull(next(), nullptr, 10);
    else { usage(); return 2; }
  }
  if (outp.empty() || c.mbw < 1 || c.mbh < 1 || c.frames < 1 || c.slices < 1) { usage(); return 2; }
  if (kbps > 0) c.target_bytes_per_frame = kbps * 1000.0 / 8.0 / fps;
  GenStats st;
  std::vector<uint8_t> data;
  try { data = generate(c, &st); }
  catch (const std::exception& e) { fprintf(stderr, "h264gen: %s\n", e.what()); return 1; }
  FILE* f = fopen(outp.c_str(), "wb");
  if (!f) { perror("h264gen"); return 1; }
  fwrite(data.data(), 1, data.size(), f);
  fclose(f);
  fprintf(stderr, "h264gen: %zu bytes, %llu slices, %llu mbs, %llu bins, %llu payload bytes (%.2f bins/byte)\n", data.size(),
          (unsigned long long)st.slices, (unsigned long long)st.mbs, (unsigned long long)st.bins,
          (unsigned long long)st.payload_bytes, double(st.bins) / double(st.payload_bytes ? st.payload_bytes : 1));
  return 0;
}
