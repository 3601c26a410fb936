// avr_job.cuh -- what one thread block (parser warp + coder warp) does with one job: cooperative
// set-up, the two lane-0 serial roles joined by the symbol ring, cooperative window refills.
// Shared by the CUDA kernels (avr_lib.cu) and by the host emulation used for debugging in the
// GPU-less container (tests/emu/avr_emu.cpp, never shipped).
//
// Thread abstraction: on the device a block has AVR_THREADS = 64 threads (warp 0 = parser, warp 1 =
// coder); under AVR_HOST_EMULATION there is one thread that plays the parser for a whole slice into
// an unbounded queue and then the coder, so every "for (i = tid; i < n; i += nthreads)" loop
// degenerates to a plain loop and the sync / broadcast primitives are no-ops.
#pragma once
#include "avr_walk.cuh"

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_THREADS 64
#define AVR_LANES 32
#define AVR_SYNCWARP() __syncwarp()
#define AVR_SYNCBLOCK() __syncthreads()
#define AVR_BCAST(v) __shfl_sync(0xFFFFFFFFu, (v), 0)
#else
#define AVR_THREADS 1
#define AVR_LANES 1
#define AVR_SYNCWARP() ((void)0)
#define AVR_SYNCBLOCK() ((void)0)
#define AVR_BCAST(v) (v)
#endif

namespace avr {

struct JobDesc {        // one block's work: slices [first, first + count) share one model
  uint32_t first_slice;
  uint32_t n_slices;
  uint32_t weight;      // host-side ordering key (input bytes)
  uint32_t reserved;
};

struct SliceIO {        // device-side view of one slice's buffers
  uint64_t in_off;      // into the input arena
  uint64_t out_off;     // into the output arena
  uint32_t in_size;
  uint32_t out_cap;
};

// Fixed part of a block's shared memory; the neighbour row and (when resident) the estimators follow.
struct BlockShared {
  SharedTables tab;
  RingShared ring;
  alignas(16) uint8_t window[WIN_BYTES];
  U2 cstate[N_CTX_PAD];     // CABAC context states {state, rangeTabLPS row}
  WalkShared walk;          // left macroblock + per-macroblock scratch of the walker
  uint32_t job;             // job handed out by the atomic counter
  uint32_t final_status;    // status of the slice just finished (coder writes, everyone reads)
  uint32_t pad_[2];
};

struct JobArgs {
  const avr_slice_desc* descs;
  const SliceIO* ios;
  const uint8_t* in_arena;
  uint8_t* out_arena;
  avr_slice_result* results;
};

struct JobMem {             // where this block keeps its state
  BlockShared* sh;
  MbCtx* row;               // [max_mbw], shared memory
  uint16_t* cnt;            // MODEL_COUNTERS estimators: shared memory (resident) or global workspace
  uint8_t* nnz0;            // two FrameBuffer nnz planes, global workspace (scalars: the struct must stay in registers)
  uint8_t* nnz1;
  uint32_t nnz_frame_bytes; // capacity of one plane
};

// CABAC context initialisation (H.264 9.3.1.1; h264_cabac.c:1267-1286) for the contexts in use.
AVR_DEV void init_contexts(U2* cstate, const avr_slice_desc& d, int tid, int nthreads) {
  const int set = d.slice_type == AVR_SLICE_I ? 0 : 1 + d.cabac_init_idc;
  const int qp = d.slice_qp < 0 ? 0 : (d.slice_qp > 51 ? 51 : d.slice_qp);
  for (int i = tid; i < N_CTX; i += nthreads) {
    const uint32_t mn = avr_ctx_init_mn[set * 1024 + ctx_idx_of_dense(i)];
    const int m = int(int8_t(mn >> 8)), n = int(int8_t(mn & 0xFF));
    int pre = ((m * qp) >> 4) + n;
    pre = pre < 1 ? 1 : (pre > 126 ? 126 : pre);
    // pStateIdx / valMPS from preCtxState; same numbers as libavcodec's 2*(..)-127 folding
    const uint32_t state = pre <= 63 ? uint32_t((63 - pre) << 1) : uint32_t(((pre - 64) << 1) | 1);
    cstate[i].x = ctx_word_x(state);
    cstate[i].y = rlps4_of(state >> 1);
  }
}

AVR_DEV void fill_u32(uint32_t* p, uint32_t n_words, uint32_t v, int tid, int nthreads) {
  for (uint32_t i = uint32_t(tid); i < n_words; i += uint32_t(nthreads)) p[i] = v;
}

// loads aligned positions [from, from + WIN_HALF) into their ring slots, one aligned 16-byte load per
// lane; past the end of the slice: the tail byte, then zeros; the <= 15 bytes before its start are don't-care
AVR_DEV void window_load_half(const ByteWindow& w, uint32_t from, int lane) {
  const uint8_t* abase = w.base - w.skew;
  const uint32_t aend = w.size + w.skew;
  for (uint32_t o = uint32_t(lane) * 16; o < WIN_HALF; o += AVR_LANES * 16) {
    const uint32_t apos = from + o;
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

// Frame-buffer bookkeeping of h264_model::update_frame_spec (recode.cpp:824-843).  Scalars only (no
// run-time indexed arrays) so that it lives in registers.
struct FrameTags {
  int w0, w1, h0, h1, num0, num1, cur;
  AVR_DEV int w(int i) const { return i ? w1 : w0; }
  AVR_DEV int h(int i) const { return i ? h1 : h0; }
  AVR_DEV int num(int i) const { return i ? num1 : num0; }
  AVR_DEV void set_size(int i, int ww, int hh) { if (i) { w1 = ww; h1 = hh; } else { w0 = ww; h0 = hh; } }
  AVR_DEV void set_num(int i, int n) { if (i) num1 = n; else num0 = n; }
  AVR_DEV void clear() { w0 = w1 = h0 = h1 = 0; cur = 0; }
};

// ---- parser warp: lane 0 walks the slice, all 32 lanes keep its input window filled ---------------
template <class Dir>
AVR_DEV void parser_role(const JobArgs& a, const JobMem& m, const avr_slice_desc& d, const SliceIO& io, int cur_frame, bool use_planes, int lane) {
  Walker<Dir> wk;
  wk.tab = &m.sh->tab;
  wk.row = m.row;
  wk.ws = &m.sh->walk;
  wk.use_planes = use_planes;
  wk.nnz_cur = cur_frame ? m.nnz1 : m.nnz0; wk.nnz_prev = cur_frame ? m.nnz0 : m.nnz1;
  wk.dir.begin(a.in_arena + io.in_off, io.in_size, m.sh->window, (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80u : 0u, &m.sh->ring,
               m.cnt, m.sh->cstate);
  window_load_half(wk.dir.in, 0, lane);
  window_load_half(wk.dir.in, WIN_HALF, lane);
  AVR_SYNCWARP();
  if (lane == 0) wk.start(d);
  for (;;) {
    int done = 1;
    uint32_t in_pos = 0;
    if (lane == 0) {
      const uint32_t limit = wk.dir.in.limit();
      while (!wk.done && wk.dir.in_pos() < limit) {
        wk.step();
        if (wk.dir.in.ran_past(wk.dir.in_pos())) { wk.fail(WALK_UNSUPPORTED); wk.done = true; }   // macroblock > window
      }
      done = wk.done;
      in_pos = wk.dir.in_pos();
    }
    done = AVR_BCAST(done);
    if (done) break;
    in_pos = AVR_BCAST(in_pos);
    ByteWindow& bw = wk.dir.in;
    AVR_SYNCWARP();
    while (in_pos + bw.skew >= bw.win_start + WIN_HALF) {   // the consumed half is free: load the next 512 bytes into it
      window_load_half(bw, bw.win_start + WIN_BYTES, lane);
      bw.win_start += WIN_HALF;
    }
    AVR_SYNCWARP();
  }
  if (lane == 0) {
    wk.dir.finish_bins();
    RingShared* r = &m.sh->ring;
    r->end_status = uint32_t(wk.status);
    r->end_mbs = wk.n_mbs;
    r->end_bins_lo = wk.n_bins_lo;
    r->end_bins_hi = wk.n_bins_hi;
    wk.dir.q.push(SYM_END);
    wk.dir.q.publish();
  }
  AVR_SYNCWARP();
}

// ---- coder warp: lane 0 drains the ring and owns the output bytes and the slice result -------------
template <class Dir>
AVR_DEV void coder_role(const JobArgs& a, const JobMem& m, const SliceIO& io, uint32_t s, int lane) {
  if (lane == 0) {
    typename Dir::Coder c;
    c.begin(a.out_arena + io.out_off, io.out_cap);
    Consumer q;
    q.init(&m.sh->ring);
    c.run(q, m.sh->tab, m.cnt, m.sh->cstate);
    const volatile RingShared* r = &m.sh->ring;
    avr_slice_result res;
    res.out_off = io.out_off;
    res.status = int32_t(r->end_status);
    if (res.status == AVR_SLICE_OK && c.overflowed()) res.status = AVR_SLICE_OVERFLOW;
    res.out_size = res.status == AVR_SLICE_OK ? c.out_bytes() : 0;
    res.n_mbs = r->end_mbs;
    res.reserved = 0;
    res.n_bins = uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
    a.results[s] = res;
    m.sh->final_status = uint32_t(res.status);
  }
  AVR_SYNCWARP();
}

// Runs one job on the calling block.
template <class Dir>
AVR_DEV void run_job(const JobArgs& a, const JobDesc& job, const JobMem& m, int tid) {
  const int nthreads = AVR_THREADS;
  const int lane = tid & (AVR_LANES - 1);
  FrameTags ft;
  ft.clear(); ft.num0 = ft.num1 = 0;
  bool poisoned = false;   // a failed slice leaves the shared model undefined: fail the rest of the job
  // A job of one slice is a chunk whose model dies with it: the previous picture is empty and the
  // neighbour counts live in the shared-memory row, so the global nnz planes are not touched at all.
  const bool use_planes = job.n_slices > 1;

  for (uint32_t s = job.first_slice; s < job.first_slice + job.n_slices; s++) {
    const avr_slice_desc d = a.descs[s];
    const SliceIO io = a.ios[s];
    const bool fresh = s == job.first_slice || (d.flags & AVR_SLICE_NEW_MODEL);
    const bool skip = (d.flags & AVR_SLICE_SKIP) != 0;   // frame_spec bookkeeping only
    const uint32_t frame_bytes = uint32_t(d.mb_width) * d.mb_height * NNZ_STRIDE;
    const bool usable = !poisoned && (!use_planes || frame_bytes <= m.nnz_frame_bytes) && d.mb_width > 0 && d.mb_height > 0;
    if (usable) {
      if (fresh) {
        // a brand-new h264_model: every estimator {1,1} (recode.cpp:1057), no frame buffers yet
        fill_u32(reinterpret_cast<uint32_t*>(m.cnt), MODEL_COUNTERS / 2, 0x01010101u, tid, nthreads);
        ft.clear();
      }
      const int mw = d.mb_width, mh = d.mb_height;
      if ((d.flags & AVR_SLICE_NEW_PICTURE) || fresh || ft.w(ft.cur) != mw || ft.h(ft.cur) != mh) {
        // update_frame_spec (recode.cpp:824-843): flip + clear unless same frame_num and size
        const bool same = ft.num(ft.cur) == d.frame_num && ft.w(ft.cur) != 0 && ft.h(ft.cur) != 0;
        if (ft.w(ft.cur) != mw || ft.h(ft.cur) != mh || !same) {
          ft.cur ^= 1;
          const int c = ft.cur;
          if (use_planes) fill_u32(reinterpret_cast<uint32_t*>(c ? m.nnz1 : m.nnz0), frame_bytes / 4, 0, tid, nthreads);
          if (ft.w(c) != mw || ft.h(c) != mh) {
            ft.set_size(c, mw, mh);
            if (ft.w(c ^ 1) != mw || ft.h(c ^ 1) != mh) {
              ft.set_size(c ^ 1, mw, mh);
              if (use_planes) fill_u32(reinterpret_cast<uint32_t*>(c ? m.nnz0 : m.nnz1), frame_bytes / 4, 0, tid, nthreads);
            }
          }
          ft.set_num(c, d.frame_num);
        }
      }
    }
    const bool walk = usable && !skip;
    if (walk) {
      init_contexts(m.sh->cstate, d, tid, nthreads);
      if (!use_planes) fill_u32(reinterpret_cast<uint32_t*>(m.row), uint32_t(d.mb_width) * MBCTX_WORDS, 0, tid, nthreads);
      if (tid == 0) ring_reset(&m.sh->ring);
    }
    AVR_SYNCBLOCK();
    if (walk) {
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
      if (tid < 32) parser_role<Dir>(a, m, d, io, ft.cur, use_planes, lane);
      else coder_role<Dir>(a, m, io, s, lane);
#else
      parser_role<Dir>(a, m, d, io, ft.cur, use_planes, lane);
      coder_role<Dir>(a, m, io, s, lane);
#endif
    } else if (tid == 0) {
      avr_slice_result r;
      r.out_off = io.out_off; r.out_size = 0; r.n_mbs = 0; r.reserved = 0; r.n_bins = 0;
      r.status = usable ? AVR_SLICE_OK : (poisoned ? AVR_SLICE_CORRUPT : AVR_SLICE_UNSUPPORTED);
      a.results[s] = r;
      m.sh->final_status = uint32_t(r.status);
    }
    AVR_SYNCBLOCK();
    if (m.sh->final_status != AVR_SLICE_OK) poisoned = true;
  }
}

}  // namespace avr
