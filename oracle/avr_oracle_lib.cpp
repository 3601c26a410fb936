// avr_oracle_lib.cpp -- C entry points over oracle/avr_oracle.hpp for ctypes (tests/, bench.py's
// cpu_baseline leg, __graft_entry__.smoke()).  TEST INFRASTRUCTURE ONLY -- never linked into
// libavrecode_b200.so.  Semantics mirror avr_compress / avr_decompress of include/avrecode_b200.h
// so that results can be compared slot by slot with the CUDA path.
#include <stdlib.h>
#include <string.h>
#include <memory>
#include "avr_oracle.hpp"

using namespace avr_oracle;

namespace {
const uint16_t* g_priors = nullptr;   // avr_oracle_set_priors: the table of slices that carry AVR_SLICE_PRIORS

struct Session {  // one h264_model + the picture-level syntax tables, carried across slices until NEW_MODEL
  Model model;
  PictureSyntax pic;
};

inline void finish_wide(WideCompressIO* io) { io->write_blob(); }
inline void finish_wide(WideDecompressIO*) {}

template <class IO, class WideIO, class Make, class MakeWide>
int run_all(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
            uint8_t** out, uint64_t* out_bytes, Make make_io, MakeWide make_wide) {
  std::vector<uint8_t> all;
  std::unique_ptr<Session> s;
  for (int i = 0; i < n; i++) {
    const avr_slice_desc& d = descs[i];
    avr_slice_result& r = res[i];
    memset(&r, 0, sizeof r);
    r.out_off = all.size();
    if (d.in_off + d.in_size > in_bytes) { r.status = AVR_SLICE_CORRUPT; continue; }
    if (!s || (d.flags & AVR_SLICE_NEW_MODEL)) {
      s.reset(new Session());
      if (d.n_streams & AVR_SLICE_PRIORS) {
        if (!g_priors) { r.status = AVR_SLICE_CORRUPT; s.reset(); continue; }
        s->model.priors = g_priors;
      }
    }
    if ((d.flags & AVR_SLICE_NEW_PICTURE) || s->pic.mbw != d.mb_width || s->pic.mbh != d.mb_height || s->model.frames[0].w == 0) {
      s->model.update_frame_spec(d.frame_num, d.mb_width, d.mb_height);  // frame_spec hook, h264_slice.c:568-573
      s->pic.start(d.mb_width, d.mb_height);
    }
    if (d.flags & AVR_SLICE_SKIP) continue;  // model hooks only (recode.cpp:1139-1145): the frame flip above
    std::vector<uint8_t> bytes;
    const int k_streams = d.n_streams & AVR_STREAMS_MASK;
    try {
      if (k_streams >= 2) {   // a wide slice of the chunked format: K range-coder streams, row r in stream (r - first row) mod K
        if (!(d.flags & AVR_SLICE_NEW_MODEL) || d.mb_width == 0) throw Corrupt("wide slice without a model of its own");
        WideIO io = make_wide(&s->model, k_streams, int(d.first_mb / d.mb_width), in + d.in_off, int(d.in_size), &bytes,
                              (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80 : 0);
        SliceWalker<WideIO> w(&io, &s->pic, d);
        r.n_mbs = uint32_t(w.run());
        r.n_bins = io.n_bins;
        r.status = io.finished() ? AVR_SLICE_OK : AVR_SLICE_CORRUPT;
        finish_wide(&io);
      } else {
        IO io = make_io(&s->model, in + d.in_off, int(d.in_size), &bytes, (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80 : 0);
        SliceWalker<IO> w(&io, &s->pic, d);
        r.n_mbs = uint32_t(w.run());
        r.n_bins = io.n_bins;
        r.status = io.finished() ? AVR_SLICE_OK : AVR_SLICE_CORRUPT;
      }
    } catch (const Unsupported&) { r.status = AVR_SLICE_UNSUPPORTED; }
    catch (const std::exception& ex) { if (getenv("AVR_ORACLE_DEBUG")) fprintf(stderr, "oracle: slice %d: %s\n", i, ex.what()); r.status = AVR_SLICE_CORRUPT; }
    if (r.status == AVR_SLICE_OK) { r.out_size = uint32_t(bytes.size()); all.insert(all.end(), bytes.begin(), bytes.end()); }
  }
  *out_bytes = all.size();
  *out = static_cast<uint8_t*>(malloc(all.size() ? all.size() : 1));
  if (!*out) return AVR_ERR_NOMEM;
  if (!all.empty()) memcpy(*out, all.data(), all.size());
  return AVR_OK;
}
}  // namespace

extern "C" {

// the caller keeps the table (AVR_PRIOR_ENTRIES entries) alive; null: none
void avr_oracle_set_priors(const uint16_t* table) { g_priors = table; }

int avr_oracle_compress(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                        uint8_t** out, uint64_t* out_bytes) {
  return run_all<CompressIO, WideCompressIO>(descs, n, in, in_bytes, res, out, out_bytes,
      [](Model* m, const uint8_t* p, int sz, std::vector<uint8_t>* o, int tail) { return CompressIO(m, p, sz, o, tail); },
      [](Model* m, int k, int row, const uint8_t* p, int sz, std::vector<uint8_t>* o, int tail) { return WideCompressIO(m, k, row, p, sz, o, tail); });
}

// mirror_compressor = 1: the format as the reference COMPRESSOR defines it (SURVEY hazards H8/H9);
// 0: bug-compatible with the stock reference decompressor.
int avr_oracle_decompress(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                          uint8_t** out, uint64_t* out_bytes, int mirror_compressor) {
  return run_all<DecompressIO, WideDecompressIO>(descs, n, in, in_bytes, res, out, out_bytes,
      [mirror_compressor](Model* m, const uint8_t* p, int sz, std::vector<uint8_t>* o, int) {
        DecompressIO io(m, p, sz, o);
        io.mirror_compressor = mirror_compressor != 0;
        return io;
      },
      [](Model* m, int k, int row, const uint8_t* p, int sz, std::vector<uint8_t>* o, int) { return WideDecompressIO(m, k, row, p, sz, o); });
}

void avr_oracle_free(void* p) { free(p); }

// Known-answer helpers for the reference's own unit test (test/arithmetic_code.cpp:48-112):
// n bits with p = 1/2 through arithmetic_code<uint64_t, uint16_t> -> compressed size.
uint64_t avr_oracle_kat_half(const uint8_t* bits, uint64_t n, uint8_t* decoded) {
  std::vector<uint8_t> out;
  ArithEncoder<uint64_t, 2> enc(&out);
  for (uint64_t i = 0; i < n; i++) enc.put(bits[i], enc.range() / 2);
  enc.finish();
  ArithDecoder<uint64_t, 2> dec(out.data(), out.data() + out.size());
  for (uint64_t i = 0; i < n; i++) decoded[i] = uint8_t(dec.get(dec.range() / 2));
  return out.size();
}

// The disabled regression vector of test/arithmetic_code.cpp:14-47 (cabac::encoder vs a CABAC decoder).
// Encodes (states[i], bits[i]) pairs, then put_terminate(0), then `tail` MPS bins on a zero state, then
// terminate(1); decodes with the restated ffmpeg decoder.  Returns 0 when every bin round-trips.
int avr_oracle_kat_cabac(const uint8_t* states, const uint8_t* bits, int n, int tail, uint8_t* stream, int* stream_len) {
  std::vector<uint8_t> out;
  {
    CabacEncoder e(&out);
    std::vector<uint8_t> st(states, states + n);
    for (int i = 0; i < n; i++) e.put(bits[i], &st[i]);
    e.put_terminate(0);
    uint8_t z = 0;
    for (int i = 0; i < tail; i++) e.put(0, &z);
    e.put_terminate(1);
  }
  if (int(out.size()) > *stream_len) return -1;
  memcpy(stream, out.data(), out.size());
  *stream_len = int(out.size());
  CabacDecoder d;
  d.init(out.data(), int(out.size()));
  std::vector<uint8_t> st(states, states + n);
  for (int i = 0; i < n; i++) if (d.get(&st[i]) != bits[i]) return 1 + i;
  if (d.get_terminate() != 0) return 1000;
  uint8_t z = 0;
  for (int i = 0; i < tail; i++) if (d.get(&z) != 0) return 2000 + i;
  if (d.get_terminate() != 1) return 3000;
  return 0;
}

}  // extern "C"
