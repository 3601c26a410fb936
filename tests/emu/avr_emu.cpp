// avr_emu.cpp -- HOST EMULATION of the device code, for debugging in the GPU-less dev container.
// TEST INFRASTRUCTURE ONLY: compiles avrecode_b200/csrc/avr_core.cuh + avr_job.cuh as plain C++
// (one emulated lane) behind the same call shape as the oracle's C entry points, so that the exact
// device logic can be diffed against the oracle and the reference on the CPU.  It is never linked
// into libavrecode_b200.so and is not a fallback: the product fails loudly without a CUDA device.
#define AVR_HOST_EMULATION 1
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../avrecode_b200/csrc/avr_job.cuh"

using namespace avr;

static thread_local std::vector<uint32_t> g_hist;   // per-class bin counts of the last call (of this thread)
static thread_local const uint16_t* g_priors = nullptr;   // prior table of the next call (avr_emu_set_priors / the file shim)
static thread_local std::vector<uint16_t> g_last_counters;   // estimators as the last ordinary job of the last call left them

template <class Dir>
static int run(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
               uint8_t** out, uint64_t* out_bytes) {
  std::vector<SliceIO> ios(n);
  uint64_t out_total = 0;
  int max_mbw = 1; uint32_t max_frame = 0;
  for (int i = 0; i < n; i++) {
    ios[i].in_off = descs[i].in_off; ios[i].in_size = descs[i].in_size;
    ios[i].out_off = out_total;
    ios[i].out_cap = Dir::kDecompress ? descs[i].payload_size + 64 : descs[i].in_size + descs[i].in_size / 8 + 256;
    if (!Dir::kDecompress && (descs[i].n_streams & AVR_STREAMS_MASK) >= 2)
      ios[i].out_cap = 64 + (descs[i].n_streams & AVR_STREAMS_MASK) * ((descs[i].in_size + descs[i].in_size / 8 + 512 + 15) & ~15u);   // any one stream may get it all
    out_total += ios[i].out_cap;
    if (descs[i].mb_width > max_mbw) max_mbw = descs[i].mb_width;
    uint32_t fb = uint32_t(descs[i].mb_width) * descs[i].mb_height * NNZ_STRIDE;
    if (fb > max_frame) max_frame = fb;
    if (descs[i].in_off + descs[i].in_size > in_bytes) return AVR_ERR_ARG;
  }
  std::vector<uint8_t> arena(out_total + 16);
  std::vector<uint16_t> counters(MODEL_COUNTERS_PAD);
  max_frame = (max_frame + 15u) & ~15u;
  std::vector<uint8_t> planes(2 * size_t(max_frame) + 64);
  BlockShared* sh = new BlockShared();
  fill_shared_tables(&sh->tab, 0, 1);
  std::vector<MbCtx> row(max_mbw);
  g_hist.assign(size_t(n) * 8, 0);
  JobArgs a{descs, ios.data(), in, arena.data(), res, g_hist.data(), g_priors};
  JobMem m;
  m.sh = sh; m.row = row.data(); m.cnt = counters.data(); m.cnt8 = m.cnt;
  m.nnz0 = planes.data(); m.nnz1 = planes.data() + max_frame; m.nnz_frame_bytes = max_frame; m.max_mbw = uint32_t(max_mbw);
  // wide slices (n_streams >= 2): one set of block state per role, the shared control block / rows / rings on the heap
  std::vector<BlockShared*> wsh;
  std::vector<std::vector<uint16_t>> wcnt(WIDE_MAX + 1, std::vector<uint16_t>(MODEL_COUNTERS_PAD));
  std::vector<std::vector<MbCtx>> wrow(WIDE_MAX + 1, std::vector<MbCtx>(max_mbw));
  std::vector<JobMem> wmems(WIDE_MAX + 1);
  for (uint32_t r = 0; r <= WIDE_MAX; r++) {
    wsh.push_back(new BlockShared());
    fill_shared_tables(&wsh[r]->tab, 0, 1);
    wmems[r] = m;
    wmems[r].sh = wsh[r]; wmems[r].row = wrow[r].data(); wmems[r].cnt = wcnt[r].data(); wmems[r].cnt8 = wmems[r].cnt;
  }
  std::vector<MbCtx> wrows(size_t(WIDE_MAX) * max_mbw);
  int i = 0;
  while (i < n) {
    if ((descs[i].n_streams & AVR_STREAMS_MASK) >= 2) {
      if (!(descs[i].flags & AVR_SLICE_NEW_MODEL) || (i + 1 < n && !(descs[i + 1].flags & AVR_SLICE_NEW_MODEL))) return AVR_ERR_ARG;
      WideCtl* ctl = new WideCtl();
      memset(wrows.data(), 0, wrows.size() * sizeof(MbCtx));
      WideMem wm;
      wm.ctl = ctl; wm.rows = wrows.data(); wm.far = nullptr; wm.far_cap = 0;
      emu_wide_slice<Dir>(a, uint32_t(i), wmems.data(), wm);
      delete ctl;
      i++;
      continue;
    }
    int j = i + 1;
    while (j < n && !(descs[j].flags & AVR_SLICE_NEW_MODEL)) j++;
    JobDesc job;
    job.first_slice = uint32_t(i); job.n_slices = uint32_t(j - i); job.weight = 0; job.wide = 0;
    run_job<Dir>(a, job, m, nullptr, 0, 0);
    i = j;
  }
  g_last_counters.assign(counters.begin(), counters.begin() + MODEL_COUNTERS);
  delete sh;
  for (BlockShared* p : wsh) delete p;
  // compact
  std::vector<uint8_t> all;
  for (int k = 0; k < n; k++) {
    uint64_t off = res[k].out_off;
    res[k].out_off = all.size();
    if (res[k].status == AVR_SLICE_OK) all.insert(all.end(), arena.begin() + off, arena.begin() + off + res[k].out_size);
    else res[k].out_size = 0;
  }
  *out_bytes = all.size();
  *out = static_cast<uint8_t*>(malloc(all.size() ? all.size() : 1));
  if (!all.empty()) memcpy(*out, all.data(), all.size());
  return AVR_OK;
}

extern "C" {
int avr_emu_compress(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                     uint8_t** out, uint64_t* out_bytes) {
  return run<Compress>(descs, n, in, in_bytes, res, out, out_bytes);
}
int avr_emu_decompress(const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                       uint8_t** out, uint64_t* out_bytes) {
  return run<Decompress>(descs, n, in, in_bytes, res, out, out_bytes);
}
void avr_emu_free(void* p) { free(p); }
// the table slices with AVR_SLICE_PRIORS start from (the caller keeps it alive); null: none
void avr_emu_set_priors(const uint16_t* table) { g_priors = table; }
// the estimators (avr_core.cuh layout, AVR_PRIOR_ENTRIES of them) as the last ordinary job left them: tools/train_priors.py
int avr_emu_last_counters(uint16_t* dst, int n) {
  if (n != int(MODEL_COUNTERS) || g_last_counters.size() != size_t(n)) return -1;
  memcpy(dst, g_last_counters.data(), size_t(n) * 2);
  return 0;
}
int avr_emu_last_hist(uint32_t* dst, int n_slices) {
  if (size_t(n_slices) * 8 != g_hist.size()) return -1;
  memcpy(dst, g_hist.data(), g_hist.size() * 4);
  return 0;
}
}
