// avr_job.cuh -- what one warp does with one job: cooperative set-up, lane-0 serial walk,
// cooperative window refills.  Shared by the CUDA kernels (avr_kernels.cu) and by the host
// emulation used for debugging in the GPU-less container (tests/emu/avr_emu.cpp, never shipped).
//
// Lane abstraction: on the device `lane` = threadIdx.x & 31 and AVR_LANES = 32; under
// AVR_HOST_EMULATION there is a single lane, so every "for (i = lane; i < n; i += AVR_LANES)" loop
// degenerates to a plain loop and the sync / broadcast primitives are no-ops.
#pragma once
#include "avr_core.cuh"

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_LANES 32
#define AVR_SYNCWARP() __syncwarp()
#define AVR_BCAST(v) __shfl_sync(0xFFFFFFFFu, (v), 0)
#else
#define AVR_LANES 1
#define AVR_SYNCWARP() ((void)0)
#define AVR_BCAST(v) (v)
#endif

namespace avr {

AVR_DEV void fill_shared_tables(SharedTables* t, int tid, int nthreads) {
  for (int i = tid; i < 100; i += nthreads) t->recip[i] = i >= 2 ? ~uint64_t(0) / uint64_t(i) : 0;  // floor((2^64-1)/d) == floor(2^64/d) unless d | 2^64
  for (int i = tid; i < 256; i += nthreads) t->range_lps[i] = avr_range_lps[i];
  for (int i = tid; i < 64; i += nthreads) {
    t->trans_lps[i] = avr_trans_lps[i];
    t->sig8[i] = i < 63 ? avr_sig8x8_inc[i] : 0;
    t->last8[i] = i < 63 ? avr_last8x8_inc[i] : 0;
  }
}

// CABAC context initialisation (H.264 9.3.1.1; h264_cabac.c:1267-1286): 1024 states per slice.
// The avrecode counters packed in the same word survive unless `reset_counters`.
AVR_DEV void init_contexts(WarpState* ws, const avr_slice_desc& d, bool reset_counters, int lane) {
  const int set = d.slice_type == AVR_SLICE_I ? 0 : 1 + d.cabac_init_idc;
  const int qp = d.slice_qp < 0 ? 0 : (d.slice_qp > 51 ? 51 : d.slice_qp);
  for (int i = lane; i < 1024; i += AVR_LANES) {
    uint32_t mn = avr_ctx_init_mn[set * 1024 + i];
    int m = int(int8_t(mn >> 8)), n = int(int8_t(mn & 0xFF));
    int pre = ((m * qp) >> 4) + n;
    pre = pre < 1 ? 1 : (pre > 126 ? 126 : pre);
    // pStateIdx / valMPS from preCtxState; same numbers as libavcodec's 2*(..)-127 folding
    uint32_t state = pre <= 63 ? uint32_t((63 - pre) << 1) : uint32_t(((pre - 64) << 1) | 1);
    uint32_t w = ws->ctx[i];
    ws->ctx[i] = reset_counters ? (state | (1u << 8) | (1u << 16)) : ((w & ~0xFFu) | state);
  }
  if (lane == 0 && reset_counters) { ws->bypass_cnt = 1 | (1 << 8); ws->term_cnt = 1 | (1 << 8); }
}

AVR_DEV void fill_u32(uint32_t* p, uint64_t n_words, uint32_t v, int lane) {
  for (uint64_t i = uint64_t(lane); i < n_words; i += AVR_LANES) p[i] = v;
}

// loads aligned positions [from, from + WIN_HALF) into their ring slots, one aligned 16-byte load per
// lane; bytes past the end of the slice read as zero, the <= 15 bytes before its start are don't-care
AVR_DEV void window_load_half(const ByteWindow& w, uint32_t from, int lane) {
  const uint8_t* abase = w.base - w.skew;
  const uint32_t aend = w.size + w.skew;
  for (uint32_t o = uint32_t(lane) * 16; o < WIN_HALF; o += AVR_LANES * 16) {
    uint32_t apos = from + o;
    uint8_t* dst = w.win + (apos & (WIN_BYTES - 1));
#if defined(__CUDA_ARCH__)
    if (apos + 16 <= aend) {
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(abase + apos);
      continue;
    }
#endif
    for (uint32_t k = 0; k < 16; k++)
      dst[k] = uint8_t((apos + k >= w.skew && apos + k < aend) ? abase[apos + k] : (apos + k == aend ? w.tail : 0));
  }
}

// Frame-buffer bookkeeping of h264_model::update_frame_spec (recode.cpp:824-843).
struct FrameTags {
  int w[2], h[2], num[2], cur;
};

struct JobArgs {
  const avr_slice_desc* descs;
  const SliceIO* ios;
  const uint8_t* in_arena;
  uint8_t* out_arena;
  avr_slice_result* results;
  uint8_t* workspace;
};

AVR_DEV ByteWindow& window_of(Compress& c) { return c.in; }
AVR_DEV uint32_t window_limit(const Compress& c) { return c.in.win_start + WIN_HALF - c.in.skew; }
AVR_DEV uint32_t window_limit(const Decompress&) { return 0xFFFFFFFFu; }

// Runs one job on the calling warp.
template <class Dir>
AVR_DEV_NOINLINE void run_job(const JobArgs& a, const JobDesc& job, const SharedTables* tab, WarpState* ws, MbCtx* row, int lane) {
  Walker<Dir> wk;
  wk.tab = tab;
  wk.ws = ws;
  wk.row = row;
  wk.mm.sig = reinterpret_cast<uint16_t*>(a.workspace + job.model_off);
  wk.mm.nz = wk.mm.sig + SIG_ENTRIES;
  wk.mm.nnz[0] = a.workspace + job.nnz_off;
  wk.mm.nnz[1] = a.workspace + job.nnz_off + job.nnz_frame_bytes;
  wk.mm.cur = 0;
  FrameTags ft;
  ft.w[0] = ft.w[1] = ft.h[0] = ft.h[1] = 0; ft.num[0] = ft.num[1] = 0; ft.cur = 0;
  bool ctx_counters_stale = false;
  bool poisoned = false;   // a failed slice leaves the shared model undefined: fail the rest of the job

  for (uint32_t s = job.first_slice; s < job.first_slice + job.n_slices; s++) {
    const avr_slice_desc d = a.descs[s];
    const SliceIO io = a.ios[s];
    const bool fresh = s == job.first_slice || (d.flags & AVR_SLICE_NEW_MODEL);
    const bool skip = (d.flags & AVR_SLICE_SKIP) != 0;   // frame_spec bookkeeping only
    const uint32_t frame_bytes = uint32_t(d.mb_width) * d.mb_height * NNZ_STRIDE;
    bool usable = !poisoned && frame_bytes <= job.nnz_frame_bytes && d.mb_width > 0 && d.mb_height > 0;
    if (usable) {
      if (fresh) {
        // a brand-new h264_model: every estimator {1,1} (recode.cpp:1057), no frame buffers yet
        fill_u32(reinterpret_cast<uint32_t*>(wk.mm.sig), MODEL_COUNTERS / 2, 0x01010101u, lane);
        ft.w[0] = ft.w[1] = ft.h[0] = ft.h[1] = 0; ft.cur = 0;
      }
      if ((d.flags & AVR_SLICE_NEW_PICTURE) || fresh || ft.w[ft.cur] != d.mb_width || ft.h[ft.cur] != d.mb_height) {
        // update_frame_spec (recode.cpp:824-843): flip + clear unless same frame_num and size
        bool same = ft.num[ft.cur] == d.frame_num && ft.w[ft.cur] != 0 && ft.h[ft.cur] != 0;
        if (ft.w[ft.cur] != d.mb_width || ft.h[ft.cur] != d.mb_height || !same) {
          ft.cur ^= 1;
          const int c = ft.cur;
          if (ft.w[c] != d.mb_width || ft.h[c] != d.mb_height) {
            ft.w[c] = d.mb_width; ft.h[c] = d.mb_height;
            fill_u32(reinterpret_cast<uint32_t*>(wk.mm.nnz[c]), frame_bytes / 4, 0, lane);
            if (ft.w[c ^ 1] != d.mb_width || ft.h[c ^ 1] != d.mb_height) {
              ft.w[c ^ 1] = d.mb_width; ft.h[c ^ 1] = d.mb_height;
              fill_u32(reinterpret_cast<uint32_t*>(wk.mm.nnz[c ^ 1]), frame_bytes / 4, 0, lane);
            }
          } else {
            fill_u32(reinterpret_cast<uint32_t*>(wk.mm.nnz[c]), frame_bytes / 4, 0, lane);
          }
          ft.num[c] = d.frame_num;
        }
      }
      wk.mm.cur = ft.cur;
      if (fresh && skip) ctx_counters_stale = true;   // the ctxIdx estimators are reset by the next walked slice
    }
    if (usable && !skip) {
      init_contexts(ws, d, fresh || ctx_counters_stale, lane);
      ctx_counters_stale = false;
      wk.d = d;
      wk.dir.begin(a.in_arena + io.in_off, io.in_size, a.out_arena + io.out_off, io.out_cap, ws->window,
                   (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80u : 0u);
      if constexpr (!Dir::kDecompress) {
        ByteWindow& bw = window_of(wk.dir);
        window_load_half(bw, 0, lane);
        window_load_half(bw, WIN_HALF, lane);
      }
    }
    AVR_SYNCWARP();
    int done = 1;
    if (usable && !skip) {
      if (lane == 0) wk.start();
      for (;;) {
        uint32_t in_pos = 0;
        if (lane == 0) {
          const uint32_t limit = window_limit(wk.dir);
          while (!wk.done && wk.dir.in_pos() < limit) wk.step();
          done = wk.done;
          in_pos = wk.dir.in_pos();
        }
        done = AVR_BCAST(done);
        if (done) break;
        in_pos = AVR_BCAST(in_pos);
        if constexpr (!Dir::kDecompress) {
          ByteWindow& bw = window_of(wk.dir);
          AVR_SYNCWARP();
          while (in_pos + bw.skew >= bw.win_start + WIN_HALF) {   // the consumed half is free: load the next 512 bytes into it
            window_load_half(bw, bw.win_start + WIN_BYTES, lane);
            bw.win_start += WIN_HALF;
          }
          AVR_SYNCWARP();
        }
      }
    }
    if (lane == 0) {
      avr_slice_result r;
      r.out_off = io.out_off;
      r.reserved = 0;
      if (!usable) { r.status = poisoned ? AVR_SLICE_CORRUPT : AVR_SLICE_UNSUPPORTED; r.out_size = 0; r.n_mbs = 0; r.n_bins = 0; }
      else if (skip) { r.status = AVR_SLICE_OK; r.out_size = 0; r.n_mbs = 0; r.n_bins = 0; }
      else {
        r.status = wk.status;
        r.out_size = wk.status == WALK_OK ? wk.dir.out_bytes() : 0;
        r.n_mbs = wk.n_mbs;
        r.n_bins = wk.n_bins;
      }
      a.results[s] = r;
      if (r.status != AVR_SLICE_OK) poisoned = true;
    }
    poisoned = AVR_BCAST(int(poisoned)) != 0;
    AVR_SYNCWARP();
  }
}

}  // namespace avr
