// avr_job.cuh -- what one thread block (parser warp + coder warp) does with one job: cooperative
// set-up, the two lane-0 serial roles joined by the symbol ring, cooperative window refills.
// Shared by the CUDA kernels (avr_lib.cu) and by the host emulation used for debugging in the
// GPU-less container (tests/emu/avr_emu.cpp, never shipped).
//
// Thread abstraction: on the device a block has 64 threads when decompressing (warp 0 = parser, warp 1 = CABAC
// re-encoder) and 96 when compressing (parser, estimator warp, range-coder warp); under AVR_HOST_EMULATION there is one thread that plays the parser for a whole slice into
// an unbounded queue and then the coder, so every "for (i = tid; i < n; i += nthreads)" loop
// degenerates to a plain loop and the sync / broadcast primitives are no-ops.
#pragma once
#include "avr_walk.cuh"

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_THREADS_OF(Dir) ((Dir::kDecompress) ? 64 : 96)   // compress: parser + estimator warp + range-coder warp
#define AVR_HOST_STAGE(x) nullptr   // the device reads a shared-memory ring in place
#define AVR_LANES 32
#define AVR_SYNCWARP() __syncwarp()
#define AVR_SYNCBLOCK() __syncthreads()
#define AVR_BCAST(v) __shfl_sync(0xFFFFFFFFu, (v), 0)
#else
#define AVR_THREADS_OF(Dir) 1
#define AVR_HOST_STAGE(x) (x)
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
#if !defined(__CUDACC__) || defined(AVR_HOST_EMULATION)
  uint32_t stage[RING_BATCH + 2 * ENTRY_MAX_WORDS];   // host emulation: the coder's view of a batch (the device reads the ring in place)
#endif
  RecRing records;          // compress: estimator warp -> range-coder warp, one word per bin
#if !defined(__CUDACC__) || defined(AVR_HOST_EMULATION)
  uint32_t stage2[RING_BATCH + 2 * ENTRY_MAX_WORDS];  // host emulation: the range-coder's view of a batch
#endif
  ByteSink sink;            // output bytes of the last warp (touched once per byte, out of line)
  ParserHome phome;         // the parser's engines between residual blocks
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
  uint32_t* hist;           // [n_slices][8] per-class bin counts (HIST_*), or null
};

struct JobMem {             // where this block keeps its state
  BlockShared* sh;
  MbCtx* row;               // [max_mbw], shared memory
  uint16_t* cnt;            // MODEL_COUNTERS estimators: shared memory (resident) or global workspace
  uint8_t* nnz0;            // two FrameBuffer nnz planes, global workspace (scalars: the struct must stay in registers)
  uint8_t* nnz1;
  uint32_t nnz_frame_bytes; // capacity of one plane
  uint32_t max_mbw;         // capacity of the neighbour row
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
struct Lane {               // what one parser warp works with
  const SharedTables* tab;
  RingShared* ring;
  uint8_t* window;
  U2* cstate;
  WalkShared* walk;
  MbCtx* row;
  uint16_t* cnt;
  uint32_t* far;            // wide decompression: this warp's symbol ring in global memory (else null)
  uint32_t far_cap;
  ParserHome* home;
};

// stream / total_mbs: wide decompression only (the caller has set wk.wide_k and what goes with it)
template <class Dir>
AVR_DEV void parser_begin(Walker<Dir>& wk, const Lane& ln, const uint8_t* src, uint32_t n, const avr_slice_desc& d,
                          uint8_t* nnz_cur, const uint8_t* nnz_prev, bool use_planes, int lane, uint32_t stream, uint32_t total_mbs) {
  wk.tab = ln.tab;
  wk.row = ln.row;
  wk.row_top = ln.row;
  wk.ws = ln.walk;
  wk.use_planes = use_planes;
  wk.nnz_cur = nnz_cur; wk.nnz_prev = nnz_prev;
  wk.dir.begin(ln.home, src, n, ln.window, (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80u : 0u, ln.ring, ln.tab, ln.cnt, ln.cstate, ln.far, ln.far_cap);
  window_load_half(wk.dir.in, 0, lane);
  window_load_half(wk.dir.in, WIN_HALF, lane);
  AVR_SYNCWARP();
  if (lane == 0) wk.start(d, stream, total_mbs);
}

// Runs until the slice (or this warp's share of a wide slice) is done.  Host emulation of a wide
// decompression also returns when the walker has to wait for the row above (wk.blocked).
template <class Dir>
AVR_DEV void parser_run(Walker<Dir>& wk, int lane) {
  for (;;) {
    int done = 1;
    uint32_t in_pos = 0;
    if (lane == 0) {
      const uint32_t limit = wk.dir.in.limit();
      while (!wk.done && wk.dir.in_pos() < limit) {
        wk.step();
        if (wk.blocked) break;
        if (wk.dir.in.ran_past(wk.dir.in_pos())) { wk.fail(WALK_UNSUPPORTED); wk.done = true; }   // macroblock > window
      }
      done = wk.done || wk.blocked;
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
}

AVR_DEV void ring_end_record(RingShared* r, int status, uint32_t mbs, uint32_t bins_lo, uint32_t bins_hi) {
  r->end_status = uint32_t(status);
  r->end_mbs = mbs;
  r->end_bins_lo = bins_lo;
  r->end_bins_hi = bins_hi;
}

template <class Dir>
AVR_DEV void parser_end(Walker<Dir>& wk, int lane) {
  if (lane == 0) {
    wk.dir.finish_bins();
    if (wk.wide_k && Dir::kDecompress) {   // nobody may wait for this warp any more
      if (wk.status != WALK_OK) *wk.abort_flag = 1;
      ring_fence();
      *wk.prog_mine = 0xFFFFFFFFu;
    }
    ring_end_record(wk.dir.h->r, wk.status, wk.n_mbs, wk.n_bins_lo(), wk.n_bins_hi);
    home_push(wk.dir.h, SYM_END);
    home_publish(wk.dir.h);
    if (wk.wide_k && !Dir::kDecompress) {
      // Wide compression: every stream ends in a terminate bin, which flushes its range coder -- the
      // slice's own end_of_slice bin in the stream of the last row, one extra bin in each of the others.
      RingShared* first = wk.rings;
      ring_end_record(first, wk.status, wk.n_mbs, wk.n_bins_lo(), wk.n_bins_hi);
#pragma unroll 1
      for (uint32_t j = 0; j < wk.wide_k; j++) {
        if (j == wk.cur_ring) continue;
        home_switch_to(wk.dir.h, &wk.rings[j], j);
        if (wk.status == WALK_OK) home_push(wk.dir.h, CNT_TERM | 1u << SYM_BIT_SHIFT);
        home_push(wk.dir.h, SYM_END);
        home_publish(wk.dir.h);
      }
    }
  }
  AVR_SYNCWARP();
}

template <class Dir>
AVR_DEV void parser_role(const JobArgs& a, const JobMem& m, const avr_slice_desc& d, const SliceIO& io, int cur_frame, bool use_planes, int lane) {
  Walker<Dir> wk;
  wk.wide_k = 0;
  wk.top_copy = false;
  Lane ln;
  ln.tab = &m.sh->tab; ln.ring = &m.sh->ring; ln.window = m.sh->window; ln.cstate = m.sh->cstate;
  ln.walk = &m.sh->walk; ln.row = m.row; ln.cnt = m.cnt; ln.far = nullptr; ln.far_cap = 0; ln.home = &m.sh->phome;
  parser_begin<Dir>(wk, ln, a.in_arena + io.in_off, io.in_size, d, cur_frame ? m.nnz1 : m.nnz0, cur_frame ? m.nnz0 : m.nnz1,
                    use_planes, lane, 0, 0);
  parser_run<Dir>(wk, lane);
  parser_end<Dir>(wk, lane);
}

// ---- the warps behind the parser ---------------------------------------------------------------------------------
// which stage waited for which: four bytes in units of 2^22 cycles (2.1 ms at 1965 MHz) -- parser for room in its ring,
// the warp behind it for entries, estimator warp for room in the record ring, range-coder warp for records
AVR_DEV uint32_t wait_bytes(const volatile RingCtl* a, const volatile RingCtl* b) {
  uint32_t v = 0;
  const uint32_t w[4] = {a->prod_wait, a->cons_wait, b ? b->prod_wait : 0u, b ? b->cons_wait : 0u};
  for (int i = 0; i < 4; i++) { const uint32_t u = w[i] >> 14; v |= (u < 255 ? u : 255u) << (8 * i); }
  return v;
}
// per-class bin counts of slice s out of the parser homes that walked it
AVR_DEV void write_hist(const JobArgs& a, uint32_t s, const ParserHome* homes, uint32_t n_homes, uint32_t n_mbs, uint64_t n_bins) {
  if (!a.hist) return;
  uint32_t h[8];
  for (int i = 0; i < 8; i++) { h[i] = 0; for (uint32_t j = 0; j < n_homes; j++) h[i] += homes[j].hist[i]; }
  h[HIST_TERM] = n_mbs;   // one end_of_slice bin per macroblock
  h[HIST_MB] = uint32_t(n_bins) - (h[HIST_CBF] + h[HIST_SIG] + h[HIST_LEVEL] + h[HIST_BYPASS] + h[HIST_PRED] + h[HIST_TERM]);
  for (int i = 0; i < 8; i++) a.hist[size_t(s) * 8 + i] = h[i];
}
AVR_DEV void write_result(const JobArgs& a, const JobMem& m, const SliceIO& io, uint32_t s, uint32_t out_bytes, bool overflowed, bool two_rings) {
  const volatile RingShared* r = &m.sh->ring;
  avr_slice_result res;
  res.out_off = io.out_off;
  res.status = int32_t(r->end_status);
  if (res.status == AVR_SLICE_OK && overflowed) res.status = AVR_SLICE_OVERFLOW;
  res.out_size = res.status == AVR_SLICE_OK ? out_bytes : 0;
  res.n_mbs = r->end_mbs;
  res.reserved = wait_bytes(r, two_rings ? &m.sh->records : nullptr);
  res.n_bins = uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
  a.results[s] = res;
  write_hist(a, s, &m.sh->phome, 1, res.n_mbs, res.n_bins);
  m.sh->final_status = uint32_t(res.status);
}
// decompress: the CABAC re-encoder, which owns the output bytes and the slice result; compress: the estimator warp
template <class Dir>
AVR_DEV void coder_role(const JobArgs& a, const JobMem& m, const SliceIO& io, uint32_t s, int lane) {
  typename Dir::Coder c;
  if constexpr (Dir::kDecompress) {
    const uint8_t lay = a.descs[s].n_streams;   // put emulation-prevention bytes back (escaped slices, chunked mode)
    c.begin(&m.sh->sink, a.out_arena + io.out_off, io.out_cap, (lay & AVR_SLICE_ESCAPED) ? int((lay >> AVR_ESC_ZEROS_SHIFT) & 3) : -1);
    c.run(&m.sh->ring, m.sh->tab, m.sh->cstate, AVR_HOST_STAGE(m.sh->stage), lane);
    if (lane == 0) write_result(a, m, io, s, c.out_bytes(), c.overflowed(), false);
  } else {
    c.begin(&m.sh->records);
    c.run(&m.sh->ring, nullptr, 0, false, m.sh->tab, m.cnt, AVR_HOST_STAGE(m.sh->stage), lane);
  }
  AVR_SYNCWARP();
}
// compress: the range-coder warp, which owns the output bytes and the slice result
AVR_DEV void range_role(const JobArgs& a, const JobMem& m, const SliceIO& io, uint32_t s, int lane) {
  RangeWarp c;
  c.begin(&m.sh->sink, a.out_arena + io.out_off, io.out_cap);
  c.run(&m.sh->records, false, m.sh->tab, AVR_HOST_STAGE(m.sh->stage2), lane);
  if (lane == 0) write_result(a, m, io, s, c.out_bytes(), c.overflowed(), true);
  AVR_SYNCWARP();
}

// Runs one job on the calling block.
template <class Dir>
AVR_DEV void run_job(const JobArgs& a, const JobDesc& job, const JobMem& m, int tid) {
  const int nthreads = AVR_THREADS_OF(Dir);
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
    const uint64_t picture_mbs = uint64_t(d.mb_width) * d.mb_height;   // Level 6.2 at most: frame_bytes cannot wrap
    const uint32_t frame_bytes = picture_mbs <= AVR_MAX_PICTURE_MBS ? uint32_t(picture_mbs) * NNZ_STRIDE : 0xFFFFFFFFu;
    const bool usable = !poisoned && picture_mbs > 0 && picture_mbs <= AVR_MAX_PICTURE_MBS && d.mb_width <= m.max_mbw &&
                        (!use_planes || frame_bytes <= m.nnz_frame_bytes);
    if (usable) {
      if (fresh) {
        // a brand-new h264_model: every estimator {1,1} (recode.cpp:1057), no frame buffers yet
        fill_u32(reinterpret_cast<uint32_t*>(m.cnt), MODEL_COUNTERS_PAD / 2, EST_INIT * 0x10001u, tid, nthreads);
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
      if (tid == 0) { ring_reset(&m.sh->ring); ring_reset(&m.sh->records); }
    }
    AVR_SYNCBLOCK();
    if (walk) {
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
      if (tid < 32) parser_role<Dir>(a, m, d, io, ft.cur, use_planes, lane);
      else if (tid < 64) coder_role<Dir>(a, m, io, s, lane);
      else range_role(a, m, io, s, lane);
#else
      parser_role<Dir>(a, m, d, io, ft.cur, use_planes, lane);
      coder_role<Dir>(a, m, io, s, lane);
      if (!Dir::kDecompress) range_role(a, m, io, s, lane);
#endif
    } else if (tid == 0) {
      avr_slice_result r;
      r.out_off = io.out_off; r.out_size = 0; r.n_mbs = 0; r.reserved = 0; r.n_bins = 0;
      r.status = usable ? AVR_SLICE_OK : (poisoned ? AVR_SLICE_CORRUPT : AVR_SLICE_UNSUPPORTED);
      a.results[s] = r;
      if (a.hist) for (int i = 0; i < 8; i++) a.hist[size_t(s) * 8 + i] = 0;
      m.sh->final_status = uint32_t(r.status);
    }
    AVR_SYNCBLOCK();
    if (m.sh->final_status != AVR_SLICE_OK) poisoned = true;
  }
}

// ------------------------------------------------------------------------------------------------
// Wide slices (container mode "avr-b200/slice" only; avr_slice_desc.n_streams = K >= 2).
//
// A slice that is much longer than the rest of its batch sets the time of the whole batch, because one
// slice is one serial CABAC stream.  The CABAC side cannot be split, but OUR side of the format can: the
// compressor writes macroblock row r into range-coder stream (r - first row) mod K, each stream with its
// own estimators.  Decompression then runs K parser warps as a wavefront -- the contexts of a macroblock
// need its left and top neighbours only, so row r at column x waits for row r - 1 to be through column
// x -- and one warp re-encodes the CABAC stream from the K rings in row order.  The single piece of
// state that crosses a row boundary in decoding order, "previous macroblock had a non-zero mb_qp_delta",
// is resolved by the re-encoder (SP_QPD / SP_ROW_END) and left out of the model (context 60).
//
// Blob layout (Block.cabac of a wide slice): u32 macroblocks | K x u32 stream bytes | streams.
// Block = 1 + WIDE_MAX warps.  compress: warp 0 parser, warps 1..K coders;  decompress: warp 0 CABAC
// re-encoder, warps 1..K parsers.  Estimators: one set per stream, all in shared memory -- which is what
// limits K to WIDE_MAX = 2 (2 x 96 KB); the neighbour rows and the long symbol rings are in the global workspace.
// ------------------------------------------------------------------------------------------------
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
#define AVR_WIDE_THREADS_OF(Dir) ((Dir::kDecompress) ? (WIDE_MAX + 1) * 32 : (2 * WIDE_MAX + 1) * 32)
#else
#define AVR_WIDE_THREADS_OF(Dir) 1
#endif

struct WideShared {
  SharedTables tab;
  RingShared ring[WIDE_MAX];
  alignas(16) uint8_t window[WIDE_MAX][WIN_BYTES];
  U2 cstate[N_CTX_PAD];
  WalkShared walk[WIDE_MAX];
  uint32_t prog[WIDE_MAX];       // decompress: wavefront progress of each parser warp
  uint32_t abort_flag;
  uint32_t out_size[WIDE_MAX];   // compress: bytes of each stream;  decompress: [0] = bytes of the slice
  uint32_t overflow;
  uint32_t job;
  uint32_t stage[WIDE_MAX][144]; // the batch of ring words a coder warp is working on (decompress: [0] only)
  RecRing records[WIDE_MAX];     // compress: estimator warp j -> range-coder warp j
  ByteSink sink[WIDE_MAX];       // compress: one per range-coder warp; decompress: [0]
  ParserHome phome[WIDE_MAX];    // one per parser warp (compress: [0])
};

struct WideMem {
  WideShared* sh;
  MbCtx* rows;      // [WIDE_MAX][max_mbw] (compress uses one): shared memory when they fit beside the estimators, else
  bool rows_global; //   the global workspace -- slower: every publication fence then waits for the row stores
  uint16_t* cnt;    // [WIDE_MAX][MODEL_COUNTERS_PAD], shared memory
  uint32_t* far;    // [WIDE_MAX][far_cap] symbol rings of a decompression, global workspace: a parser warp may run a
  uint32_t far_cap; //   whole macroblock row ahead of the re-encoder, far more than a shared-memory ring holds
  uint32_t max_mbw;
};

AVR_DEV uint32_t load_le32(const uint8_t* p) { return uint32_t(p[0]) | uint32_t(p[1]) << 8 | uint32_t(p[2]) << 16 | uint32_t(p[3]) << 24; }
AVR_DEV void store_le32(uint8_t* p, uint32_t v) { p[0] = uint8_t(v); p[1] = uint8_t(v >> 8); p[2] = uint8_t(v >> 16); p[3] = uint8_t(v >> 24); }

template <class Dir>
AVR_DEV void wide_fail(const JobArgs& a, uint32_t s, const SliceIO& io, int status, int tid) {
  if (tid == 0) {
    avr_slice_result r;
    r.out_off = io.out_off; r.out_size = 0; r.n_mbs = 0; r.reserved = 0; r.n_bins = 0; r.status = status;
    a.results[s] = r;
  }
  AVR_SYNCBLOCK();
}

template <class Dir>
AVR_DEV void wide_lane(Lane& ln, const WideMem& m, uint32_t j, uint32_t mbw) {
  WideShared* sh = m.sh;
  ln.tab = &sh->tab; ln.ring = &sh->ring[j]; ln.window = sh->window[j]; ln.cstate = sh->cstate;
  ln.walk = &sh->walk[j]; ln.row = m.rows + size_t(j) * mbw; ln.cnt = m.cnt + size_t(j) * MODEL_COUNTERS_PAD;
  ln.far = m.far ? m.far + size_t(j) * m.far_cap : nullptr;   // compress: ring 0, the parser moves on from there
  ln.far_cap = m.far ? m.far_cap : 0;
  ln.home = &sh->phome[j];
}

template <class Dir>
AVR_DEV void wide_walker(Walker<Dir>& wk, const WideMem& m, uint32_t K, uint32_t j) {
  WideShared* sh = m.sh;
  wk.wide_k = K;
  wk.top_copy = m.rows_global;
  wk.rings = sh->ring;
  wk.rows_base = m.rows;
  wk.prog_mine = &sh->prog[j];
  wk.prog_above = &sh->prog[j == 0 ? K - 1 : j - 1];
  wk.abort_flag = &sh->abort_flag;
}

template <class Dir>
AVR_DEV void run_wide_job(const JobArgs& a, uint32_t s, const WideMem& m, int tid) {
  const int nthreads = AVR_WIDE_THREADS_OF(Dir);
  const int lane = tid & (AVR_LANES - 1);
  WideShared* sh = m.sh;
  const avr_slice_desc d = a.descs[s];
  const SliceIO io = a.ios[s];
  const uint32_t K = d.n_streams & AVR_STREAMS_MASK;
  const int esc_zeros = (Dir::kDecompress && (d.n_streams & AVR_SLICE_ESCAPED)) ? int((d.n_streams >> AVR_ESC_ZEROS_SHIFT) & 3) : -1;
  (void)esc_zeros;
  const uint32_t mbw = d.mb_width;
  const uint32_t hdr = 4 + 4 * K;
  if (K < 2 || K > WIDE_MAX || mbw < 1 || d.mb_height < 1 || mbw > m.max_mbw) { wide_fail<Dir>(a, s, io, AVR_SLICE_UNSUPPORTED, tid); return; }
  const uint8_t* blob = a.in_arena + io.in_off;
  uint8_t* out = a.out_arena + io.out_off;
  uint32_t total_mbs = 0;
  uint32_t subcap = 0;
  if (Dir::kDecompress) {
    bool ok = io.in_size >= hdr;
    uint64_t sum = hdr;
    if (ok) {
      total_mbs = load_le32(blob);
      for (uint32_t j = 0; j < K; j++) sum += load_le32(blob + 4 + 4 * j);
      const uint32_t mbs_left = uint32_t(d.mb_width) * d.mb_height - (d.first_mb < uint32_t(d.mb_width) * d.mb_height ? d.first_mb : 0);
      ok = sum == io.in_size && total_mbs >= 1 && total_mbs <= mbs_left && d.first_mb < uint32_t(d.mb_width) * d.mb_height;
    }
    if (!ok) { wide_fail<Dir>(a, s, io, AVR_SLICE_CORRUPT, tid); return; }
  } else {
    if (io.out_cap < hdr + 64 * K) { wide_fail<Dir>(a, s, io, AVR_SLICE_OVERFLOW, tid); return; }
    subcap = ((io.out_cap - hdr) / K) & ~15u;
  }
  init_contexts(sh->cstate, d, tid, nthreads);
  fill_u32(reinterpret_cast<uint32_t*>(m.rows), (Dir::kDecompress ? K : 1u) * mbw * MBCTX_WORDS, 0, tid, nthreads);   // compress walks in place: one row
  fill_u32(reinterpret_cast<uint32_t*>(m.cnt), K * (MODEL_COUNTERS_PAD / 2), EST_INIT * 0x10001u, tid, nthreads);
  for (uint32_t j = uint32_t(tid); j < K; j += uint32_t(nthreads)) { ring_reset(&sh->ring[j]); ring_reset(&sh->records[j]); sh->prog[j] = d.first_mb; sh->out_size[j] = 0; }
  if (tid == 0) { sh->abort_flag = 0; sh->overflow = 0; }
  AVR_SYNCBLOCK();

#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
  const uint32_t warp = uint32_t(tid) >> 5;
  if constexpr (!Dir::kDecompress) {
    if (warp == 0) {
      Walker<Dir> wk;
      wide_walker<Dir>(wk, m, K, 0);
      Lane ln;
      wide_lane<Dir>(ln, m, 0, mbw);
      parser_begin<Dir>(wk, ln, blob, io.in_size, d, nullptr, nullptr, false, lane, 0, 0);
      parser_run<Dir>(wk, lane);
      parser_end<Dir>(wk, lane);
    } else if (warp <= K) {   // estimator warp of stream j
      const uint32_t j = warp - 1;
      typename Dir::Coder c;
      c.begin(&sh->records[j]);
      c.run(&sh->ring[j], m.far + size_t(j) * m.far_cap, m.far_cap, true, sh->tab, m.cnt + size_t(j) * MODEL_COUNTERS_PAD, sh->stage[j], lane);
      AVR_SYNCWARP();   // the other lanes must not run ahead into the block barrier
    } else if (warp <= 2 * K) {   // range-coder warp of stream j
      const uint32_t j = warp - 1 - K;
      RangeWarp c;
      c.begin(&sh->sink[j], out + hdr + size_t(j) * subcap, subcap);
      c.run(&sh->records[j], true, sh->tab, nullptr, lane);
      if (lane == 0) {
        sh->out_size[j] = c.out_bytes();
        if (c.overflowed()) sh->overflow = 1;
      }
      AVR_SYNCWARP();
    }
  } else {
    if (warp == 0) {
      typename Dir::Coder c;
      c.begin(&sh->sink[0], out, io.out_cap, esc_zeros);
      c.run_wide(sh->ring, K, m.far, m.far_cap, sh->tab, sh->cstate, sh->stage[0], lane);
      if (lane == 0) {
        sh->out_size[0] = c.out_bytes();
        if (c.overflowed()) sh->overflow = 1;
      }
      AVR_SYNCWARP();
    } else if (warp <= K) {
      const uint32_t j = warp - 1;
      Walker<Dir> wk;
      wide_walker<Dir>(wk, m, K, j);
      Lane ln;
      wide_lane<Dir>(ln, m, j, mbw);
      uint32_t off = hdr;
      for (uint32_t i = 0; i < j; i++) off += load_le32(blob + 4 + 4 * i);
      parser_begin<Dir>(wk, ln, blob + off, load_le32(blob + 4 + 4 * j), d, nullptr, nullptr, false, lane, j, total_mbs);
      parser_run<Dir>(wk, lane);
      parser_end<Dir>(wk, lane);
    }
  }
#else
  if constexpr (!Dir::kDecompress) {
    Walker<Dir> wk;
    wide_walker<Dir>(wk, m, K, 0);
    Lane ln;
    wide_lane<Dir>(ln, m, 0, mbw);
    parser_begin<Dir>(wk, ln, blob, io.in_size, d, nullptr, nullptr, false, lane, 0, 0);
    parser_run<Dir>(wk, lane);
    parser_end<Dir>(wk, lane);
    for (uint32_t j = 0; j < K; j++) {
      typename Dir::Coder c;
      c.begin(&sh->records[j]);
      c.run(&sh->ring[j], nullptr, 0, true, sh->tab, m.cnt + size_t(j) * MODEL_COUNTERS_PAD, sh->stage[j], lane);
      RangeWarp rw;
      rw.begin(&sh->sink[j], out + hdr + size_t(j) * subcap, subcap);
      rw.run(&sh->records[j], true, sh->tab, sh->stage[j], lane);
      sh->out_size[j] = rw.out_bytes();
      if (rw.overflowed()) sh->overflow = 1;
    }
  } else {
    Walker<Dir>* wk = new Walker<Dir>[WIDE_MAX];
    uint32_t off = hdr;
    for (uint32_t j = 0; j < K; j++) {
      wide_walker<Dir>(wk[j], m, K, j);
      Lane ln;
      wide_lane<Dir>(ln, m, j, mbw);
      const uint32_t sz = load_le32(blob + 4 + 4 * j);
      parser_begin<Dir>(wk[j], ln, blob + off, sz, d, nullptr, nullptr, false, lane, j, total_mbs);
      off += sz;
    }
    for (bool busy = true; busy;) {   // round robin: a walker runs until it has to wait for the row above
      busy = false;
      for (uint32_t j = 0; j < K; j++)
        if (!wk[j].done) { parser_run<Dir>(wk[j], lane); busy = true; }
    }
    for (uint32_t j = 0; j < K; j++) parser_end<Dir>(wk[j], lane);
    delete[] wk;
    typename Dir::Coder c;
    c.begin(&sh->sink[0], out, io.out_cap, esc_zeros);
    c.run_wide(sh->ring, K, m.far, m.far_cap, sh->tab, sh->cstate, sh->stage[0], lane);
    sh->out_size[0] = c.out_bytes();
    if (c.overflowed()) sh->overflow = 1;
  }
#endif
  AVR_SYNCBLOCK();

  // result (and, compressing, the blob: header + the K streams moved up against each other)
  int status;
  uint32_t mbs = 0, out_size = 0;
  uint64_t bins = 0;
  if (!Dir::kDecompress) {
    const volatile RingShared* r = &sh->ring[0];
    status = int(r->end_status);
    mbs = r->end_mbs;
    bins = uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
    if (status == AVR_SLICE_OK && sh->overflow) status = AVR_SLICE_OVERFLOW;
    if (status == AVR_SLICE_OK) {
      uint32_t dst = hdr + sh->out_size[0];
#pragma unroll 1
      for (uint32_t j = 1; j < K; j++) {
        const uint8_t* src = out + hdr + size_t(j) * subcap;
        const uint32_t n = sh->out_size[j];
#pragma unroll 1
        for (uint32_t o = 0; o < n; o += uint32_t(nthreads)) {   // dst <= src: chunk by chunk, read before write
          const uint32_t i = o + uint32_t(tid);
          const uint8_t v = i < n ? src[i] : uint8_t(0);
          AVR_SYNCBLOCK();
          if (i < n) out[dst + i] = v;
          AVR_SYNCBLOCK();
        }
        dst += n;
      }
      out_size = dst;
      if (tid == 0) {
        store_le32(out, mbs);
        for (uint32_t j = 0; j < K; j++) store_le32(out + 4 + 4 * j, sh->out_size[j]);
      }
    }
  } else {
    status = AVR_SLICE_OK;
    for (uint32_t j = 0; j < K; j++) {
      const volatile RingShared* r = &sh->ring[j];
      if (status == AVR_SLICE_OK && r->end_status != AVR_SLICE_OK) status = int(r->end_status);
      mbs += r->end_mbs;
      bins += uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
    }
    if (status == AVR_SLICE_OK && mbs != total_mbs) status = AVR_SLICE_CORRUPT;
    if (status == AVR_SLICE_OK && sh->overflow) status = AVR_SLICE_OVERFLOW;
    out_size = sh->out_size[0];
  }
  if (tid == 0) {
    avr_slice_result res;
    res.out_off = io.out_off;
    res.status = status;
    res.out_size = status == AVR_SLICE_OK ? out_size : 0;
    res.n_mbs = mbs;
    res.reserved = 0;
    res.n_bins = bins;
    a.results[s] = res;
    write_hist(a, s, sh->phome, Dir::kDecompress ? K : 1u, mbs, bins);
  }
  AVR_SYNCBLOCK();
}

}  // namespace avr
