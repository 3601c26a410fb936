// h264_index.cpp -- host-side slice indexer (the "genuinely serial and light" part of the path).
//
// Scans an Annex-B H.264 byte stream, parses SPS / PPS / slice headers and returns, for every
// CABAC slice, where its slice_data payload lies in the file plus the handful of header facts the
// device syntax walk needs.  This replaces what the reference obtains by running all of libavcodec
// over the file: the (buf,size) handed to init_decoder (ffmpeg/libavcodec/h264_slice.c:2357-2364),
// the frame_spec hook (h264_slice.c:568-573), and the memmem-based payload location of
// recode.cpp:1275-1297.  No macroblock is decoded here.
//
// Payload rule (h264.c:1298-1315 get_bit_length + h264_slice.c:2359-2364): the payload starts at the
// first byte-aligned position after the slice header and ends with the byte that holds the last bit
// before rbsp_stop_one_bit; trailing zero bytes and a stop bit alone in its byte are outside it.
// A payload that contains an emulation-prevention byte is reported recodable = 0 (the reference's
// memmem cannot find it, recode.cpp:1289-1296), as is one shorter than 8 bytes (recode.cpp:27,1279).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../include/avrecode_b200.h"

namespace {

struct Sps {
  bool valid = false;
  int chroma_format_idc = 1, separate_colour_plane = 0, bit_depth_luma = 8;
  int log2_max_frame_num = 4, poc_type = 0, log2_max_poc_lsb = 4, delta_pic_order_always_zero = 0;
  int mb_width = 0, mb_height_units = 0, frame_mbs_only = 1, mbaff = 0, direct_8x8_inference = 0;
};
struct Pps {
  bool valid = false;
  int sps_id = 0, cabac = 0, bottom_field_pic_order = 0, slice_groups = 1;
  int ref_count[2] = {1, 1};
  int weighted_pred = 0, weighted_bipred_idc = 0, init_qp = 26, deblocking_control = 0;
  int redundant_pic_cnt_present = 0, transform_8x8_mode = 0;
};

// Bit reader over the RAW NAL bytes that drops emulation-prevention bytes on the fly (H.264 7.4.1)
// and remembers where it is in the raw stream.
class NalBits {
 public:
  NalBits(const uint8_t* p, size_t n) : p_(p), n_(n) {}
  bool ok() const { return !err_; }
  int bit() {
    if (bitpos_ == 0) {
      if (!fetch()) { err_ = true; return 0; }
    }
    int b = (cur_ >> (7 - bitpos_)) & 1;
    bitpos_ = (bitpos_ + 1) & 7;
    return b;
  }
  uint32_t u(int n) { uint32_t v = 0; while (n--) v = (v << 1) | uint32_t(bit()); return v; }
  uint32_t ue() {
    int z = 0;
    while (!bit()) { if (++z > 32 || err_) { err_ = true; return 0; } }
    return z ? ((1u << z) - 1 + u(z)) : 0;
  }
  int32_t se() { uint32_t k = ue(); return (k & 1) ? int32_t((k + 1) >> 1) : -int32_t(k >> 1); }
  bool aligned() const { return bitpos_ == 0; }
  void align() { bitpos_ = 0; }
  // raw offset of the next unread RBSP byte (skipping an emulation byte that sits right there)
  size_t raw_pos_next() {
    if (pos_ < n_ && zeros_ >= 2 && p_[pos_] == 3) { pos_++; zeros_ = 0; }
    return pos_;
  }
  int zeros() const { return zeros_; }
  // more_rbsp_data(): anything but the trailing stop bit pattern left?
  bool more_rbsp_data() {
    size_t last = n_;
    while (last > pos_ && p_[last - 1] == 0) last--;
    if (bitpos_ == 0) {
      if (pos_ >= last) return false;
      if (pos_ + 1 == last) return p_[pos_] != 0x80;
      return true;
    }
    if (pos_ < last) return true;
    uint8_t rest = uint8_t(cur_ << bitpos_);
    return rest != 0x80 && rest != 0;
  }

 private:
  bool fetch() {
    if (pos_ < n_ && zeros_ >= 2 && p_[pos_] == 3) { pos_++; zeros_ = 0; }
    if (pos_ >= n_) return false;
    cur_ = p_[pos_++];
    zeros_ = cur_ == 0 ? zeros_ + 1 : 0;
    return true;
  }
  const uint8_t* p_; size_t n_, pos_ = 0;
  int bitpos_ = 0, zeros_ = 0;
  uint8_t cur_ = 0;
  bool err_ = false;
};

void skip_scaling_list(NalBits& b, int size) {
  int last = 8, next = 8;
  for (int j = 0; j < size; j++) {
    if (next != 0) next = (last + b.se() + 256) % 256;
    last = next == 0 ? last : next;
  }
}

void parse_sps(NalBits& b, Sps* table) {
  int profile = int(b.u(8)); b.u(8); b.u(8);
  uint32_t id = b.ue();
  if (id >= 32 || !b.ok()) return;
  Sps s;
  if (profile == 100 || profile == 110 || profile == 122 || profile == 244 || profile == 44 || profile == 83 ||
      profile == 86 || profile == 118 || profile == 128 || profile == 138 || profile == 139 || profile == 134 || profile == 135) {
    s.chroma_format_idc = int(b.ue());
    if (s.chroma_format_idc == 3) s.separate_colour_plane = int(b.u(1));
    s.bit_depth_luma = int(b.ue()) + 8;
    b.ue(); b.u(1);
    if (b.u(1)) {
      int lists = s.chroma_format_idc != 3 ? 8 : 12;
      for (int i = 0; i < lists; i++) if (b.u(1)) skip_scaling_list(b, i < 6 ? 16 : 64);
    }
  }
  s.log2_max_frame_num = int(b.ue()) + 4;
  s.poc_type = int(b.ue());
  if (s.poc_type == 0) s.log2_max_poc_lsb = int(b.ue()) + 4;
  else if (s.poc_type == 1) {
    s.delta_pic_order_always_zero = int(b.u(1));
    b.se(); b.se();
    uint32_t n = b.ue();
    if (n > 255) return;
    for (uint32_t i = 0; i < n; i++) b.se();
  }
  b.ue(); b.u(1);
  s.mb_width = int(b.ue()) + 1;
  s.mb_height_units = int(b.ue()) + 1;
  s.frame_mbs_only = int(b.u(1));
  if (!s.frame_mbs_only) s.mbaff = int(b.u(1));
  s.direct_8x8_inference = int(b.u(1));
  if (!b.ok()) return;
  s.valid = true;
  table[id] = s;
}

void parse_pps(NalBits& b, Pps* table, const Sps* sps) {
  uint32_t id = b.ue();
  if (id >= 256) return;
  Pps p;
  p.sps_id = int(b.ue());
  if (p.sps_id >= 32) return;
  p.cabac = int(b.u(1));
  p.bottom_field_pic_order = int(b.u(1));
  p.slice_groups = int(b.ue()) + 1;
  if (p.slice_groups > 1) { p.valid = true; table[id] = p; return; }  // FMO: slices will be marked unsupported
  p.ref_count[0] = int(b.ue()) + 1;
  p.ref_count[1] = int(b.ue()) + 1;
  p.weighted_pred = int(b.u(1));
  p.weighted_bipred_idc = int(b.u(2));
  p.init_qp = int(b.se()) + 26;
  b.se(); b.se();
  p.deblocking_control = int(b.u(1));
  b.u(1);
  p.redundant_pic_cnt_present = int(b.u(1));
  if (!b.ok()) return;
  if (b.more_rbsp_data()) {
    p.transform_8x8_mode = int(b.u(1));
    if (b.u(1)) {  // pic_scaling_matrix_present_flag
      int cf = sps[p.sps_id].valid ? sps[p.sps_id].chroma_format_idc : 1;
      int lists = 6 + (cf != 3 ? 2 : 6) * p.transform_8x8_mode;
      for (int i = 0; i < lists; i++) if (b.u(1)) skip_scaling_list(b, i < 6 ? 16 : 64);
    }
    b.se();
  }
  p.valid = true;
  table[id] = p;
}

// One indexing pass over a file: SPS / PPS tables, access-unit tracking, the entries found so far.
// Fed NAL by NAL -- by the Annex-B start-code scanner or by the ISO-BMFF (mp4 / mov) sample walker below.
struct Indexer {
  const uint8_t* file;
  std::vector<avr_index_entry> out;
  std::vector<Sps> sps;
  std::vector<Pps> pps;
  bool frame_start_found = false, boundary_nal_seen = true;
  uint32_t last_first_mb = 0;
  explicit Indexer(const uint8_t* f) : file(f), sps(32), pps(256) {}
  void parameter_set(const uint8_t* nal, size_t n);   // a NAL that is not in the file proper (avcC box)
  void nal(size_t nal_start, size_t nal_end);         // raw NAL bytes [nal_start, nal_end) of the file, header byte first
};

void Indexer::parameter_set(const uint8_t* nal, size_t n) {
  if (n < 2) return;
  const int type = nal[0] & 31;
  NalBits b(nal + 1, n - 1);
  if (type == 7) parse_sps(b, sps.data());
  else if (type == 8) parse_pps(b, pps.data(), sps.data());
}

void Indexer::nal(size_t nal_start, size_t nal_end) {
  if (nal_end <= nal_start) return;
  int hdr = file[nal_start];
  int ref_idc = (hdr >> 5) & 3, type = hdr & 31;
  const uint8_t* body = file + nal_start + 1;
  size_t body_n = nal_end - nal_start - 1;
  if (type == 7) { NalBits b(body, body_n); parse_sps(b, sps.data()); boundary_nal_seen = true; }
  else if (type == 8) { NalBits b(body, body_n); parse_pps(b, pps.data(), sps.data()); boundary_nal_seen = true; }
  else if (type == 6 || type == 9) boundary_nal_seen = true;
  else if (type == 1 || type == 5) {
    NalBits b(body, body_n);
    uint32_t first_mb = b.ue();
    uint32_t slice_type = b.ue();
    uint32_t pps_id = b.ue();
    const Pps* pp = pps_id < 256 && pps[pps_id].valid ? &pps[pps_id] : nullptr;
    const Sps* sp = pp && sps[pp->sps_id].valid ? &sps[pp->sps_id] : nullptr;
    // access-unit split as libavcodec's parser does it (h264_parser.c:86-118)
    bool new_picture = !frame_start_found || boundary_nal_seen || first_mb <= last_first_mb;
    frame_start_found = true; boundary_nal_seen = false; last_first_mb = first_mb;
    if (!(pp && sp && pp->cabac && b.ok() && slice_type < 10)) return;
    bool unsupported = pp->slice_groups > 1;
    int st = int(slice_type % 5);
    if (sp->separate_colour_plane) b.u(2);
    int frame_num = int(b.u(sp->log2_max_frame_num));
    int field_pic = 0;
    if (!sp->frame_mbs_only) { field_pic = int(b.u(1)); if (field_pic) b.u(1); }
    if (type == 5) b.ue();
    if (sp->poc_type == 0) { b.u(sp->log2_max_poc_lsb); if (pp->bottom_field_pic_order && !field_pic) b.se(); }
    if (sp->poc_type == 1 && !sp->delta_pic_order_always_zero) { b.se(); if (pp->bottom_field_pic_order && !field_pic) b.se(); }
    int redundant = 0;
    if (pp->redundant_pic_cnt_present) redundant = int(b.ue());
    if (st == 1) b.u(1);  // direct_spatial_mv_pred_flag
    int rc0 = pp->ref_count[0], rc1 = pp->ref_count[1];
    if (st == 0 || st == 1 || st == 3) {
      if (b.u(1)) { rc0 = int(b.ue()) + 1; if (st == 1) rc1 = int(b.ue()) + 1; }
    }
    if (st != 1) rc1 = 0;
    if (st == 2 || st == 4) rc0 = 0;
    if (st != 2 && st != 4) {  // ref_pic_list_modification (H.264 7.3.3.1): ops 0..2 carry one ue, 3 ends the list
      for (int l = 0; l < (st == 1 ? 2 : 1); l++)
        if (b.u(1)) {
          for (int guard = 0; guard < 70 && b.ok(); guard++) { uint32_t op = b.ue(); if (op == 3) break; b.ue(); }
        }
    }
    if ((pp->weighted_pred && (st == 0 || st == 3)) || (pp->weighted_bipred_idc == 1 && st == 1)) {  // pred_weight_table
      b.ue();
      int chroma = sp->chroma_format_idc != 0 && !sp->separate_colour_plane;
      if (chroma) b.ue();
      for (int l = 0; l < (st == 1 ? 2 : 1); l++) {
        int cnt = l ? rc1 : rc0;
        for (int r = 0; r < cnt && b.ok(); r++) {
          if (b.u(1)) { b.se(); b.se(); }
          if (chroma && b.u(1)) { b.se(); b.se(); b.se(); b.se(); }
        }
      }
    }
    if (ref_idc != 0) {  // dec_ref_pic_marking (H.264 7.3.3.3; libavcodec h264_refs.c ff_h264_decode_ref_pic_marking)
      if (type == 5) { b.u(1); b.u(1); }
      else if (b.u(1)) {
        for (int guard = 0; guard < 70 && b.ok(); guard++) {
          uint32_t op = b.ue();
          if (op == 0) break;
          if (op == 1 || op == 3) b.ue();             // difference_of_pic_nums_minus1
          if (op == 2) b.ue();                        // long_term_pic_num
          if (op == 3 || op == 6) b.ue();             // long_term_frame_idx
          if (op == 4) b.ue();                        // max_long_term_frame_idx_plus1
        }
      }
    }
    int cabac_init_idc = 0;
    if (st != 2 && st != 4) cabac_init_idc = int(b.ue());
    int qp = pp->init_qp + b.se();
    if (st == 3) b.u(1);
    if (st == 3 || st == 4) b.se();
    if (pp->deblocking_control) {
      uint32_t idc = b.ue();
      if (idc != 1) { b.se(); b.se(); }
    }
    if (!(b.ok() && cabac_init_idc <= 2 && redundant == 0)) return;
    b.align();  // cabac_alignment_one_bit
    size_t pay_raw = b.raw_pos_next();  // offset within body
    // end of payload: strip trailing zero bytes, then the stop bit
    size_t end = body_n;
    while (end > pay_raw && body[end - 1] == 0) end--;
    bool escaped = false;
    {   // an emulation-prevention byte: 0x03 after two zero bytes (the zeros may end the slice header)
      const int z = b.zeros();
      size_t k = pay_raw;
      while (k < end) {
        const void* hit = memchr(body + k, 3, end - k);
        if (!hit) break;
        k = size_t(static_cast<const uint8_t*>(hit) - body);
        const size_t back = k - pay_raw;   // payload bytes before the 0x03
        const bool z1 = back >= 1 ? body[k - 1] == 0 : z >= 1;
        const bool z2 = back >= 2 ? body[k - 2] == 0 : (back == 1 ? z >= 1 : z >= 2);
        if (z1 && z2) { escaped = true; break; }
        k++;
      }
    }
    if (end <= pay_raw) return;
    uint8_t lastb = body[end - 1];
    size_t size = (lastb == 0x80) ? end - 1 - pay_raw : end - pay_raw;
    if (size == 0 || size >= (size_t(1) << 31)) return;
    avr_index_entry e;
    memset(&e, 0, sizeof e);
    e.desc.in_off = uint64_t(nal_start + 1 + pay_raw);
    e.desc.in_size = uint32_t(size);
    e.desc.payload_size = uint32_t(size);
    e.desc.first_mb = first_mb;
    const int mbh = sp->mb_height_units * (2 - sp->frame_mbs_only);
    // pictures beyond Level 6.2 (139 264 macroblocks, H.264 table A-1) or wider than the descriptor holds are not walked
    const bool too_big = sp->mb_width > 0xFFFF || mbh > 0xFFFF || uint64_t(sp->mb_width) * uint64_t(mbh) > AVR_MAX_PICTURE_MBS;
    e.desc.mb_width = uint16_t(too_big ? 0 : sp->mb_width);
    e.desc.mb_height = uint16_t(too_big ? 0 : mbh);
    e.desc.frame_num = frame_num;
    e.desc.slice_type = uint8_t(st == 3 ? 0 : st == 4 ? 2 : st);
    e.desc.cabac_init_idc = uint8_t(cabac_init_idc);
    e.desc.slice_qp = int8_t(qp);
    e.desc.num_ref_idx_l0 = uint8_t(rc0 > 255 ? 255 : rc0);
    e.desc.num_ref_idx_l1 = uint8_t(rc1 > 255 ? 255 : rc1);
    e.desc.chroma_format_idc = uint8_t(sp->separate_colour_plane ? 0 : sp->chroma_format_idc);
    uint8_t fl = 0;
    if (pp->transform_8x8_mode) fl |= AVR_SLICE_TRANSFORM_8X8;
    if (sp->direct_8x8_inference) fl |= AVR_SLICE_DIRECT_8X8_INFERENCE;
    if (new_picture) fl |= AVR_SLICE_NEW_PICTURE;
    if (lastb == 0x80) fl |= AVR_SLICE_STOP_BYTE;
    if (field_pic) fl |= AVR_SLICE_FIELD;
    if (sp->mbaff && !field_pic) fl |= AVR_SLICE_MBAFF;
    if (unsupported || st >= 3 || sp->bit_depth_luma != 8 || too_big) fl |= AVR_SLICE_FIELD;  // routed to "unsupported"
    e.desc.flags = fl;
    e.recodable = (!escaped && size >= 8) ? 1 : 0;
    {   // what a re-escaping decoder needs: the zero bytes that end the header, and whether the payload is escaped
      const int z = b.zeros();
      e.desc.n_streams = uint8_t((z > 2 ? 2 : z) << AVR_ESC_ZEROS_SHIFT);
      if (escaped && size >= 8) e.desc.n_streams |= AVR_SLICE_ESCAPED;
    }
    e.nal_type = type;
    out.push_back(e);
  }
}

// ---- ISO base media files (mp4 / mov): the reference demuxes these with libavformat (recode.cpp:92, 118; its
// README's example is a GoPro .MP4).  What the slice index needs from the container is light: SPS / PPS and the NAL
// length size from the avcC box, and where the samples of that track lie (stsz / stsc / stco|co64); every sample is
// a run of length-prefixed NAL units (ISO/IEC 14496-15 5.2.3), emulation-prevention bytes included. ------------------
uint32_t be32(const uint8_t* p) { return uint32_t(p[0]) << 24 | uint32_t(p[1]) << 16 | uint32_t(p[2]) << 8 | p[3]; }
uint64_t be64(const uint8_t* p) { return uint64_t(be32(p)) << 32 | be32(p + 4); }

struct Box { uint32_t type; size_t body, end; };   // body .. end = payload of the box
bool next_box(const uint8_t* f, size_t pos, size_t limit, Box* b) {
  if (limit - pos < 8 || pos > limit) return false;
  uint64_t size = be32(f + pos);
  b->type = be32(f + pos + 4);
  size_t hdr = 8;
  if (size == 1) { if (limit - pos < 16) return false; size = be64(f + pos + 8); hdr = 16; }
  else if (size == 0) size = limit - pos;
  if (size < hdr || size > limit - pos) return false;
  b->body = pos + hdr; b->end = pos + size_t(size);
  return true;
}
constexpr uint32_t fourcc(char a, char b, char c, char d) { return uint32_t(uint8_t(a)) << 24 | uint32_t(uint8_t(b)) << 16 | uint32_t(uint8_t(c)) << 8 | uint8_t(d); }
bool find_box(const uint8_t* f, size_t pos, size_t limit, uint32_t type, Box* out) {
  Box b;
  while (next_box(f, pos, limit, &b)) { if (b.type == type) { *out = b; return true; } pos = b.end; }
  return false;
}

struct AvcTrack {
  int nal_len = 0;
  std::vector<std::pair<size_t, size_t>> psets;      // parameter sets inside avcC: (offset, size) in the file
  std::vector<std::pair<uint64_t, uint32_t>> samples; // (file offset, size) in decoding order
};

bool read_avc_track(const uint8_t* f, size_t n, const Box& trak, AvcTrack* t) {
  Box mdia, minf, stbl, stsd, b;
  if (!find_box(f, trak.body, trak.end, fourcc('m','d','i','a'), &mdia) || !find_box(f, mdia.body, mdia.end, fourcc('m','i','n','f'), &minf) ||
      !find_box(f, minf.body, minf.end, fourcc('s','t','b','l'), &stbl) || !find_box(f, stbl.body, stbl.end, fourcc('s','t','s','d'), &stsd))
    return false;
  if (stsd.end - stsd.body < 8) return false;
  Box entry;
  if (!next_box(f, stsd.body + 8, stsd.end, &entry) || (entry.type != fourcc('a','v','c','1') && entry.type != fourcc('a','v','c','3'))) return false;
  Box avcc;
  if (entry.end - entry.body < 78 || !find_box(f, entry.body + 78, entry.end, fourcc('a','v','c','C'), &avcc)) return false;   // VisualSampleEntry is 78 bytes
  size_t p = avcc.body;
  if (avcc.end - p < 7) return false;
  t->nal_len = (f[p + 4] & 3) + 1;
  int n_sps = f[p + 5] & 31;
  p += 6;
  for (int pass = 0; pass < 2; pass++) {
    int cnt = pass == 0 ? n_sps : (p < avcc.end ? f[p++] : 0);
    for (int i = 0; i < cnt; i++) {
      if (avcc.end - p < 2) return false;
      size_t len = size_t(f[p]) << 8 | f[p + 1];
      p += 2;
      if (len > avcc.end - p) return false;
      t->psets.push_back({p, len});
      p += len;
    }
  }
  // sample table: sizes, samples per chunk, chunk offsets
  Box stsz, stsc, stco;
  if (!find_box(f, stbl.body, stbl.end, fourcc('s','t','s','z'), &stsz) || !find_box(f, stbl.body, stbl.end, fourcc('s','t','s','c'), &stsc)) return false;
  bool co64 = false;
  if (!find_box(f, stbl.body, stbl.end, fourcc('s','t','c','o'), &stco)) {
    if (!find_box(f, stbl.body, stbl.end, fourcc('c','o','6','4'), &stco)) return false;
    co64 = true;
  }
  if (stsz.end - stsz.body < 12 || stsc.end - stsc.body < 8 || stco.end - stco.body < 8) return false;
  const uint32_t fixed = be32(f + stsz.body + 4), n_samples = be32(f + stsz.body + 8);
  if (!fixed && (stsz.end - stsz.body - 12) / 4 < n_samples) return false;
  const uint32_t n_stsc = be32(f + stsc.body + 4), n_chunks = be32(f + stco.body + 4);
  if ((stsc.end - stsc.body - 8) / 12 < n_stsc || (stco.end - stco.body - 8) / (co64 ? 8 : 4) < n_chunks) return false;
  (void)b;
  uint32_t s = 0, run = 0;
  for (uint32_t c = 0; c < n_chunks && s < n_samples; c++) {
    while (run + 1 < n_stsc && be32(f + stsc.body + 8 + 12 * (run + 1)) <= c + 1) run++;
    const uint32_t per = n_stsc ? be32(f + stsc.body + 8 + 12 * run + 4) : 0;
    uint64_t off = co64 ? be64(f + stco.body + 8 + 8 * size_t(c)) : be32(f + stco.body + 8 + 4 * size_t(c));
    for (uint32_t k = 0; k < per && s < n_samples; k++, s++) {
      const uint32_t sz = fixed ? fixed : be32(f + stsz.body + 12 + 4 * size_t(s));
      if (off > n || sz > n - off) return false;
      t->samples.push_back({off, sz});
      off += sz;
    }
  }
  return true;
}

// true: `file` is an ISO base media file with an AVC video track, and its slices are now in ix.out
bool index_isobmff(Indexer& ix, size_t n) {
  const uint8_t* f = ix.file;
  Box b;
  if (!next_box(f, 0, n, &b)) return false;
  if (b.type != fourcc('f','t','y','p') && b.type != fourcc('m','o','o','v') && b.type != fourcc('m','d','a','t') &&
      b.type != fourcc('f','r','e','e') && b.type != fourcc('s','k','i','p') && b.type != fourcc('w','i','d','e')) return false;
  Box moov;
  if (!find_box(f, 0, n, fourcc('m','o','o','v'), &moov)) return false;
  size_t pos = moov.body;
  Box trak;
  while (next_box(f, pos, moov.end, &trak)) {
    pos = trak.end;
    if (trak.type != fourcc('t','r','a','k')) continue;
    AvcTrack t;
    if (!read_avc_track(f, n, trak, &t)) continue;
    for (auto& ps : t.psets) ix.parameter_set(f + ps.first, ps.second);
    for (auto& sm : t.samples) {
      ix.boundary_nal_seen = true;   // a sample is an access unit
      size_t p = size_t(sm.first), e = p + sm.second;
      while (e - p >= size_t(t.nal_len)) {
        uint32_t len = 0;
        for (int k = 0; k < t.nal_len; k++) len = len << 8 | f[p + k];
        p += t.nal_len;
        if (len > e - p) break;
        ix.nal(p, p + len);
        p += len;
      }
    }
    return true;   // the first AVC track, like the reference (recode.cpp:119-121 takes the first video stream)
  }
  return false;
}

}  // namespace

extern "C" int avr_index_h264(const uint8_t* file, uint64_t file_bytes, avr_index_entry** entries, int* n_entries) {
  if (!file || !entries || !n_entries) return AVR_ERR_ARG;
  Indexer ix(file);
  size_t n = size_t(file_bytes), i = 0;
  if (!index_isobmff(ix, n)) {
    // Annex-B elementary stream: 00 00 01 by memchr on the 0x01 (SIMD in libc; 0x01 is rare in entropy-coded data),
    // then two bytes back
    auto find_sc = [&](size_t from) -> size_t {
      size_t k = from + 2;
      while (k < n) {
        const void* hit = memchr(file + k, 1, n - k);
        if (!hit) break;
        k = size_t(static_cast<const uint8_t*>(hit) - file);
        if (file[k - 1] == 0 && file[k - 2] == 0) return k - 2;
        k++;
      }
      return n;
    };
    i = find_sc(0);
    while (i < n) {
      size_t nal_start = i + 3;
      size_t next = find_sc(nal_start);   // raw NAL bytes [nal_start, next); trailing zeros belong to the next start code
      if (nal_start >= n) break;
      ix.nal(nal_start, next);
      i = next;
    }
  }
  std::vector<avr_index_entry>& out = ix.out;
  *n_entries = int(out.size());
  *entries = nullptr;
  if (!out.empty()) {
    *entries = static_cast<avr_index_entry*>(malloc(out.size() * sizeof(avr_index_entry)));
    if (!*entries) return AVR_ERR_NOMEM;
    memcpy(*entries, out.data(), out.size() * sizeof(avr_index_entry));
  }
  return AVR_OK;
}
