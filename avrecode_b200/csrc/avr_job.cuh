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
#define AVR_THREADS_OF(Dir) ((Dir::kDecompress) ? 64 : 96)   // parser warp + coder warp (+ compress: range-coder warp)
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
  uint32_t wide;        // 0: an ordinary job.  Else a role of the wide slice first_slice: bits 0..7 role + 1, bits 8.. its workspace slot
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
  uint32_t stage[RING_BATCH + 2 * ENTRY_MAX_WORDS];   // a batch of a wide slice's ring (global memory) staged for lane 0; host emulation: of any ring
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
  const uint16_t* priors;   // [MODEL_COUNTERS] what a fresh model starts from when its slice says AVR_SLICE_PRIORS, or null
};

struct JobMem {             // where this block keeps its state
  BlockShared* sh;
  MbCtx* row;               // [max_mbw], shared memory
  uint16_t* cnt;            // MODEL_COUNTERS estimators: shared memory (resident: all of them; split: the first CNT_LOW_PAD)
  uint16_t* cnt8;           // the 8x8 significance slab is at cnt8 + CNT8_SPLIT: cnt8 == cnt (resident) or in the global workspace (split)
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

// n words of estimators from the prior table (or {1, 1} each), the padding behind them {1, 1}
AVR_DEV void init_estimators(uint32_t* dst, const uint32_t* prior, uint32_t n, uint32_t n_pad, int tid, int nthreads) {
  for (uint32_t i = uint32_t(tid); i < n_pad; i += uint32_t(nthreads)) dst[i] = (prior && i < n) ? prior[i] : EST_INIT * 0x10001u;
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
  uint16_t* cnt8;
  RingCtl* far_ctl;         // wide slices: the slice's rings in global memory (control words / entries), else null
  uint32_t* far;
  uint32_t far_cap, far_index;
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
  wk.dir.begin(ln.home, src, n, ln.window, (d.flags & AVR_SLICE_STOP_BYTE) ? 0x80u : 0u, ln.ring, ln.tab, ln.cnt, ln.cstate, ln.far_ctl, ln.far,
               ln.far_cap, ln.far_index);
  wk.dir.set_cnt8(ln.cnt8);
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

AVR_DEV void ring_end_record(RingCtl* r, int status, uint32_t mbs, uint32_t bins_lo, uint32_t bins_hi) {
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
      ring_end_record(&wk.dir.h->far_ctl[0], wk.status, wk.n_mbs, wk.n_bins_lo(), wk.n_bins_hi);
#pragma unroll 1
      for (uint32_t j = 0; j < wk.wide_k; j++) {
        if (j == wk.cur_ring) continue;
        home_switch_to(wk.dir.h, j);
        if (wk.status == WALK_OK) home_push(wk.dir.h, CNT_TERM | 1u << SYM_BIT_SHIFT);
        home_push(wk.dir.h, SYM_END);
        home_publish(wk.dir.h);
      }
    }
  }
  AVR_SYNCWARP();
}

// ------------------------------------------------------------------------------------------------
// Wide slices (container mode "avr-b200/slice" only; avr_slice_desc.n_streams = K >= 2).
//
// A slice that is much longer than the rest of its batch sets the time of the whole batch, because one
// slice is one serial CABAC stream.  The CABAC side cannot be split, but OUR side of the format can: the
// compressor writes macroblock row r into range-coder stream (r - first row) mod K, each stream with its
// own estimators.  Decompression then runs K parsers as a wavefront -- the contexts of a macroblock
// need its left and top neighbours only, so row r at column x waits for row r - 1 to be through column
// x -- and one warp re-encodes the CABAC stream from the K rings in row order.  The single piece of
// state that crosses a row boundary in decoding order, "previous macroblock had a non-zero mb_qp_delta",
// is resolved by the re-encoder (SP_QPD / SP_ROW_END) and left out of the model (context 60).
//
// Blob layout (Block.cabac of a wide slice): u32 macroblocks | K x u32 stream bytes | streams.
//
// Every stream is worked on by a THREAD BLOCK of its own (its estimators fill half an SM's shared memory): a wide
// slice is K + 1 ROLES that the persistent blocks of the job kernel take from the job queue like any other job --
//   compress    role 0: the parser (CABAC decode + walk, feeds K rings)   role 1 + j: estimator + range-coder warps of stream j
//   decompress  role 0: the CABAC re-encoder (drains K rings in row order) role 1 + j: the parser of stream j
// The roles talk through global memory (WideCtl: ring control words, wavefront progress; the rings' entries; the
// neighbour rows) with device-scope fences.  No deadlock: the roles of a slice are consecutive in the queue with the
// one the others wait for last in line... first in line, jobs are handed out in order, and a block keeps its job until
// it is done -- so when any role of a slice is running, all roles of every slice before it in the queue are, and the
// first unfinished slice of the queue always has all of its roles running (the grid has at least K + 1 blocks).
// ------------------------------------------------------------------------------------------------
struct WideCtl {            // one per wide slice, global memory, zeroed by the host before the launch
  RingCtl ring[WIDE_MAX];   // ring j carries the entries of stream j (its words: `far`)
  uint32_t prog[WIDE_MAX];  // decompress: macroblocks of the picture (raster order) parser j has got past
  uint32_t out_size[WIDE_MAX];
  uint32_t hist[WIDE_MAX][8];   // per-class bin counts of each parser (HIST_*)
  uint32_t abort_flag, overflow, roles_done, pad;
};

struct WideMem {            // where a wide slice keeps what its roles share (global memory)
  WideCtl* ctl;
  MbCtx* rows;              // decompress: [K][mb_width], row r in buffer r mod K
  uint32_t* far;            // [K][far_cap] ring entries
  uint32_t far_cap;
};

AVR_DEV uint32_t load_le32(const uint8_t* p) { return uint32_t(p[0]) | uint32_t(p[1]) << 8 | uint32_t(p[2]) << 16 | uint32_t(p[3]) << 24; }
AVR_DEV void store_le32(uint8_t* p, uint32_t v) { p[0] = uint8_t(v); p[1] = uint8_t(v >> 8); p[2] = uint8_t(v >> 16); p[3] = uint8_t(v >> 24); }

// what every role works out for itself from the descriptor (and, decompressing, the blob header)
struct WideGeom {
  uint32_t K, hdr, total_mbs, subcap;
  int status;   // AVR_SLICE_OK or why the slice cannot be done (every role comes to the same verdict)
};
AVR_DEV WideGeom wide_geometry(const avr_slice_desc& d, const SliceIO& io, const uint8_t* blob, bool decompress, uint32_t max_mbw) {
  WideGeom g;
  g.K = d.n_streams & AVR_STREAMS_MASK; g.hdr = 4 + 4 * g.K; g.total_mbs = 0; g.subcap = 0; g.status = AVR_SLICE_OK;
  const uint64_t picture_mbs = uint64_t(d.mb_width) * d.mb_height;
  if (g.K < 2 || g.K > WIDE_MAX || picture_mbs < 1 || picture_mbs > AVR_MAX_PICTURE_MBS || d.mb_width > max_mbw) { g.status = AVR_SLICE_UNSUPPORTED; return g; }
  if (decompress) {
    bool ok = io.in_size >= g.hdr && d.first_mb < picture_mbs;
    uint64_t sum = g.hdr;
    if (ok) {
      g.total_mbs = load_le32(blob);
      for (uint32_t j = 0; j < g.K; j++) sum += load_le32(blob + 4 + 4 * j);
      ok = sum == io.in_size && g.total_mbs >= 1 && g.total_mbs <= picture_mbs - d.first_mb;
    }
    if (!ok) g.status = AVR_SLICE_CORRUPT;
  } else {
    if (io.out_cap < g.hdr + 64 * g.K) g.status = AVR_SLICE_OVERFLOW;
    else g.subcap = ((io.out_cap - g.hdr) / g.K) & ~15u;
  }
  return g;
}

template <class Dir>
AVR_DEV void wide_walker(Walker<Dir>& wk, const WideMem& w, uint32_t K, uint32_t j) {
  wk.wide_k = K;
  wk.top_copy = Dir::kDecompress;   // the row above was written by another thread block
  wk.rows_base = w.rows;
  wk.prog_mine = &w.ctl->prog[j];
  wk.prog_above = &w.ctl->prog[j == 0 ? K - 1 : j - 1];
  wk.abort_flag = &w.ctl->abort_flag;
}

AVR_DEV void wide_result(const JobArgs& a, uint32_t s, const SliceIO& io, int status, uint32_t out_size, uint32_t mbs, uint64_t bins) {
  avr_slice_result res;
  res.out_off = io.out_off;
  res.status = status;
  res.out_size = status == AVR_SLICE_OK ? out_size : 0;
  res.n_mbs = mbs;
  res.reserved = 0;
  res.n_bins = bins;
  a.results[s] = res;
}

// How the calling block takes part in slice s.  An ordinary slice: all of it, through the block's own rings.  A wide
// slice: ONE role of it (see above) -- the stages that are not this role's stay idle, the rings / rows / progress
// counters are the slice's in global memory.  One code path (one inlined copy of the walker and of each coder) serves both.
struct Unit {
  bool parse, code;          // which stages run in this block
  uint32_t K, j;             // wide: streams of the slice, the stream this role works on (0 = ordinary)
  WideMem w;                 // wide: w.ctl != null
  const uint8_t* src;        // parser input (payload / coded bytes of the stream)
  uint32_t src_size;
  uint8_t* dst;              // output bytes of the last stage
  uint32_t dst_cap, total_mbs;
  int esc_zeros;
};

template <class Dir>
AVR_DEV void parser_role(const JobMem& m, const Unit& u, const avr_slice_desc& d, int cur_frame, bool use_planes, int lane) {
  Walker<Dir> wk;
  wk.wide_k = 0;
  wk.top_copy = false;
  Lane ln;
  ln.tab = &m.sh->tab; ln.ring = &m.sh->ring; ln.window = m.sh->window; ln.cstate = m.sh->cstate;
  ln.walk = &m.sh->walk; ln.row = m.row; ln.cnt = m.cnt; ln.cnt8 = m.cnt8; ln.far_ctl = nullptr; ln.far = nullptr; ln.far_cap = 0; ln.far_index = 0; ln.home = &m.sh->phome;
  if (u.w.ctl) {
    wide_walker<Dir>(wk, u.w, u.K, u.j);
    ln.far_ctl = u.w.ctl->ring; ln.far = u.w.far; ln.far_cap = u.w.far_cap; ln.far_index = u.j;
    if (Dir::kDecompress) ln.row = u.w.rows + size_t(u.j) * d.mb_width;   // (compress walks in place: the block's own row)
  }
  parser_begin<Dir>(wk, ln, u.src, u.src_size, d, cur_frame ? m.nnz1 : m.nnz0, cur_frame ? m.nnz0 : m.nnz1, use_planes, lane, u.j, u.total_mbs);
  parser_run<Dir>(wk, lane);
  if (u.w.ctl && lane == 0) for (int i = 0; i < 8; i++) u.w.ctl->hist[u.j][i] = m.sh->phome.hist[i];   // before the end of the ring is published
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
AVR_DEV void write_hist_rows(const JobArgs& a, uint32_t s, const volatile uint32_t (*rows)[8], uint32_t n_rows, uint32_t n_mbs, uint64_t n_bins) {
  if (!a.hist) return;
  uint32_t h[8];
  for (int i = 0; i < 8; i++) { h[i] = 0; for (uint32_t j = 0; j < n_rows; j++) h[i] += rows[j][i]; }
  h[HIST_TERM] = n_mbs;
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
// the warp behind the parser -- decompress: the CABAC re-encoder, which owns the output bytes and (ordinary slices) the
// slice result; compress: the estimator warp, which feeds the record ring
template <class Dir>
AVR_DEV void coder_role(const JobArgs& a, const JobMem& m, const Unit& u, const SliceIO& io, uint32_t s, int lane) {
  typename Dir::Coder c;
  if constexpr (Dir::kDecompress) {
    c.begin(&m.sh->sink, u.dst, u.dst_cap, u.esc_zeros);
    if (u.w.ctl) {
      c.run_wide(u.w.ctl->ring, u.K, u.w.far, u.w.far_cap, m.sh->tab, m.sh->cstate, m.sh->stage, lane);
      if (lane == 0) { u.w.ctl->out_size[0] = c.out_bytes(); if (c.overflowed()) u.w.ctl->overflow = 1; }
    } else {
      c.run(&m.sh->ring, m.sh->tab, m.sh->cstate, AVR_HOST_STAGE(m.sh->stage), lane);
      if (lane == 0) write_result(a, m, io, s, c.out_bytes(), c.overflowed(), false);
    }
  } else {
    c.begin(&m.sh->records, m.cnt8);
    c.run(&m.sh->ring, u.w.ctl ? &u.w.ctl->ring[u.j] : nullptr, u.w.ctl ? u.w.far + size_t(u.j) * u.w.far_cap : nullptr, u.w.far_cap,
          m.sh->tab, m.cnt, m.sh->stage, lane);
  }
  AVR_SYNCWARP();
}
// compress: the range-coder warp, which owns the output bytes and (ordinary slices) the slice result
AVR_DEV void range_role(const JobArgs& a, const JobMem& m, const Unit& u, const SliceIO& io, uint32_t s, int lane) {
  RangeWarp c;
  c.begin(&m.sh->sink, u.dst, u.dst_cap);
  c.run(&m.sh->records, m.sh->tab, AVR_HOST_STAGE(m.sh->stage2), lane);
  if (lane == 0) {
    if (u.w.ctl) { u.w.ctl->out_size[u.j] = c.out_bytes(); if (c.overflowed()) u.w.ctl->overflow = 1; }
    else write_result(a, m, io, s, c.out_bytes(), c.overflowed(), true);
  }
  AVR_SYNCWARP();
}

// What remains to be done for a wide slice once this block's role is through: compressing, the LAST role to finish moves
// the K streams up against each other and writes the blob header and the result; decompressing, role 0 (the re-encoder,
// which ends last by construction) writes the result.
template <class Dir>
AVR_DEV void wide_finish(const JobArgs& a, const JobMem& m, const Unit& u, const WideGeom& g, const SliceIO& io, uint32_t s, uint32_t role, int tid) {
  const int nthreads = AVR_THREADS_OF(Dir);
  WideCtl* ctl = u.w.ctl;
  const uint32_t K = u.K;
  uint8_t* out = a.out_arena + io.out_off;
  if constexpr (!Dir::kDecompress) {
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
    if (tid == 0) { __threadfence(); m.sh->job = atomicAdd(&ctl->roles_done, 1u); }
    AVR_SYNCBLOCK();
    const bool last = m.sh->job == K;
    AVR_SYNCBLOCK();
    if (!last) return;
    __threadfence();
#else
    (void)role;
    if (++ctl->roles_done != K + 1) return;
#endif
    const volatile RingCtl* r0 = &ctl->ring[0];
    int status = int(r0->end_status);
    const uint32_t mbs = r0->end_mbs;
    const uint64_t bins = uint64_t(r0->end_bins_lo) | uint64_t(r0->end_bins_hi) << 32;
    if (status == AVR_SLICE_OK && *reinterpret_cast<volatile uint32_t*>(&ctl->overflow)) status = AVR_SLICE_OVERFLOW;
    uint32_t out_size = 0;
    if (status == AVR_SLICE_OK) {
      uint32_t dst = g.hdr + *reinterpret_cast<volatile uint32_t*>(&ctl->out_size[0]);
#pragma unroll 1
      for (uint32_t j = 1; j < K; j++) {
        const uint8_t* src = out + g.hdr + size_t(j) * g.subcap;
        const uint32_t n = *reinterpret_cast<volatile uint32_t*>(&ctl->out_size[j]);
#pragma unroll 1
        for (uint32_t o = 0; o < n; o += uint32_t(nthreads)) {   // dst <= src: chunk by chunk, read before write
          const uint32_t i = o + uint32_t(tid);
          const uint8_t v = i < n ? reinterpret_cast<const volatile uint8_t*>(src)[i] : uint8_t(0);
          AVR_SYNCBLOCK();
          if (i < n) out[dst + i] = v;
          AVR_SYNCBLOCK();
        }
        dst += n;
      }
      out_size = dst;
      if (tid == 0) {
        store_le32(out, mbs);
        for (uint32_t j = 0; j < K; j++) store_le32(out + 4 + 4 * j, *reinterpret_cast<volatile uint32_t*>(&ctl->out_size[j]));
      }
    }
    if (tid == 0) {
      wide_result(a, s, io, status, out_size, mbs, bins);
      write_hist_rows(a, s, ctl->hist, 1, mbs, bins);
    }
  } else {
    if (role != 0 || tid != 0) return;
    int status = AVR_SLICE_OK;
    uint32_t mbs = 0;
    uint64_t bins = 0;
    for (uint32_t j = 0; j < K; j++) {
      const volatile RingCtl* r = &ctl->ring[j];
      if (status == AVR_SLICE_OK && r->end_status != AVR_SLICE_OK) status = int(r->end_status);
      mbs += r->end_mbs;
      bins += uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
    }
    if (status == AVR_SLICE_OK && mbs != g.total_mbs) status = AVR_SLICE_CORRUPT;
    if (status == AVR_SLICE_OK && ctl->overflow) status = AVR_SLICE_OVERFLOW;
    wide_result(a, s, io, status, ctl->out_size[0], mbs, bins);
    write_hist_rows(a, s, ctl->hist, K, mbs, bins);
  }
}

// Runs one job on the calling block: the slices of a job share one model (one slice in chunked mode, a file in
// reference-parity mode); wide != null: the job is role `role` of the wide slice job.first_slice.
template <class Dir>
AVR_DEV void run_job(const JobArgs& a, const JobDesc& job, const JobMem& m, const WideMem* wide, uint32_t role, int tid) {
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
    Unit u;
    u.parse = u.code = true; u.K = 0; u.j = 0; u.w.ctl = nullptr; u.w.rows = nullptr; u.w.far = nullptr; u.w.far_cap = 0;
    u.src = a.in_arena + io.in_off; u.src_size = io.in_size;
    u.dst = a.out_arena + io.out_off; u.dst_cap = io.out_cap; u.total_mbs = 0;
    u.esc_zeros = (Dir::kDecompress && (d.n_streams & AVR_SLICE_ESCAPED)) ? int((d.n_streams >> AVR_ESC_ZEROS_SHIFT) & 3) : -1;
    WideGeom g;
    g.status = AVR_SLICE_OK; g.K = 0; g.hdr = 0; g.total_mbs = 0; g.subcap = 0;
    if (wide) {
      g = wide_geometry(d, io, u.src, Dir::kDecompress, m.max_mbw);
      if (g.status != AVR_SLICE_OK || role > g.K) {   // every role comes to the same verdict: role 0 reports, the others leave
        if (g.status != AVR_SLICE_OK && role == 0 && tid == 0) {
          wide_result(a, s, io, g.status, 0, 0, 0);
          if (a.hist) for (int i = 0; i < 8; i++) a.hist[size_t(s) * 8 + i] = 0;
        }
        AVR_SYNCBLOCK();
        return;
      }
      u.w = *wide; u.K = g.K; u.total_mbs = g.total_mbs;
      // compress: role 0 parses, role 1 + j codes stream j.  decompress: role 0 re-encodes, role 1 + j parses stream j.
      u.parse = Dir::kDecompress ? role >= 1 : role == 0;
      u.code = !u.parse;
      u.j = role >= 1 ? role - 1 : 0;
      if (Dir::kDecompress) {
        if (u.parse) {
          uint32_t off = g.hdr;
          for (uint32_t i = 0; i < u.j; i++) off += load_le32(u.src + 4 + 4 * i);
          u.src_size = load_le32(u.src + 4 + 4 * u.j);
          u.src += off;
        }
      } else if (u.code) { u.dst += g.hdr + size_t(u.j) * g.subcap; u.dst_cap = g.subcap; }
    }
    const bool fresh = s == job.first_slice || (d.flags & AVR_SLICE_NEW_MODEL);
    const bool skip = (d.flags & AVR_SLICE_SKIP) != 0;   // frame_spec bookkeeping only
    const uint64_t picture_mbs = uint64_t(d.mb_width) * d.mb_height;   // Level 6.2 at most: frame_bytes cannot wrap
    const uint32_t frame_bytes = picture_mbs <= AVR_MAX_PICTURE_MBS ? uint32_t(picture_mbs) * NNZ_STRIDE : 0xFFFFFFFFu;
    const bool usable = !poisoned && picture_mbs > 0 && picture_mbs <= AVR_MAX_PICTURE_MBS && d.mb_width <= m.max_mbw &&
                        (!use_planes || frame_bytes <= m.nnz_frame_bytes);
    if (usable) {
      if (fresh) {
        // a brand-new h264_model: every estimator {1,1} (recode.cpp:1057), no frame buffers yet
        // ... or, chunked mode with a prior table, a copy of that table (avr_set_priors)
        const uint32_t* pri = (a.priors && (d.n_streams & AVR_SLICE_PRIORS)) ? reinterpret_cast<const uint32_t*>(a.priors) : nullptr;
        if (m.cnt8 == m.cnt) init_estimators(reinterpret_cast<uint32_t*>(m.cnt), pri, MODEL_COUNTERS / 2, MODEL_COUNTERS_PAD / 2, tid, nthreads);
        else {
          init_estimators(reinterpret_cast<uint32_t*>(m.cnt), pri, CNT8_SPLIT / 2, CNT_LOW_PAD / 2, tid, nthreads);
          init_estimators(reinterpret_cast<uint32_t*>(m.cnt8 + CNT8_SPLIT), pri ? pri + CNT8_SPLIT / 2 : nullptr, SIG_CAT_8 / 2, CNT8_ENTRIES / 2, tid, nthreads);
        }
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
      if (tid < 32) { if (u.parse) parser_role<Dir>(m, u, d, ft.cur, use_planes, lane); }
      else if (tid < 64) { if (u.code) coder_role<Dir>(a, m, u, io, s, lane); }
      else if (u.code) range_role(a, m, u, io, s, lane);
#else
      if (u.parse) parser_role<Dir>(m, u, d, ft.cur, use_planes, lane);
      if (u.code) coder_role<Dir>(a, m, u, io, s, lane);
      if (u.code && !Dir::kDecompress) range_role(a, m, u, io, s, lane);
#endif
    } else if (tid == 0 && (!wide || role == 0)) {
      avr_slice_result r;
      r.out_off = io.out_off; r.out_size = 0; r.n_mbs = 0; r.reserved = 0; r.n_bins = 0;
      r.status = usable ? AVR_SLICE_OK : (poisoned ? AVR_SLICE_CORRUPT : AVR_SLICE_UNSUPPORTED);
      a.results[s] = r;
      if (a.hist) for (int i = 0; i < 8; i++) a.hist[size_t(s) * 8 + i] = 0;
      m.sh->final_status = uint32_t(r.status);
    }
    AVR_SYNCBLOCK();
    if (wide) { if (walk) wide_finish<Dir>(a, m, u, g, io, s, role, tid); AVR_SYNCBLOCK(); return; }
    if (m.sh->final_status != AVR_SLICE_OK) poisoned = true;
  }
}

#if !defined(__CUDACC__) || defined(AVR_HOST_EMULATION)
// Host emulation of a wide slice: the roles one after the other (the decompression parsers round robin: a walker
// runs until it has to wait for the row above), on one set of "block" state per role.
template <class Dir>
void emu_wide_slice(const JobArgs& a, uint32_t s, JobMem* mems, const WideMem& w) {
  const avr_slice_desc d = a.descs[s];
  const SliceIO io = a.ios[s];
  const uint8_t* blob = a.in_arena + io.in_off;
  const WideGeom g = wide_geometry(d, io, blob, Dir::kDecompress, mems[0].max_mbw);
  if (g.status != AVR_SLICE_OK || !Dir::kDecompress) {
    JobDesc job;
    job.first_slice = s; job.n_slices = 1; job.weight = 0; job.wide = 1;
    for (uint32_t role = 0; role <= (g.status == AVR_SLICE_OK ? g.K : 0u); role++) run_job<Dir>(a, job, mems[role], &w, role, 0);
    return;
  }
  if constexpr (Dir::kDecompress) {
    const uint32_t K = g.K, mbw = d.mb_width;
    Walker<Dir>* wk = new Walker<Dir>[WIDE_MAX];
    uint32_t off = g.hdr;
    for (uint32_t j = 0; j < K; j++) {
      const JobMem& m = mems[1 + j];
      init_estimators(reinterpret_cast<uint32_t*>(m.cnt), (a.priors && (d.n_streams & AVR_SLICE_PRIORS)) ? reinterpret_cast<const uint32_t*>(a.priors) : nullptr,
                      MODEL_COUNTERS / 2, MODEL_COUNTERS_PAD / 2, 0, 1);
      wide_walker<Dir>(wk[j], w, K, j);
      Lane ln;
      ln.tab = &m.sh->tab; ln.ring = &m.sh->ring; ln.window = m.sh->window; ln.cstate = m.sh->cstate; ln.walk = &m.sh->walk;
      ln.row = w.rows + size_t(j) * mbw; ln.cnt = m.cnt; ln.cnt8 = m.cnt8; ln.home = &m.sh->phome;
      ln.far_ctl = w.ctl->ring; ln.far = nullptr; ln.far_cap = 0; ln.far_index = j;
      const uint32_t sz = load_le32(blob + 4 + 4 * j);
      parser_begin<Dir>(wk[j], ln, blob + off, sz, d, nullptr, nullptr, false, 0, j, g.total_mbs);
      off += sz;
    }
    for (bool busy = true; busy;) {
      busy = false;
      for (uint32_t j = 0; j < K; j++)
        if (!wk[j].done) { parser_run<Dir>(wk[j], 0); busy = true; }
    }
    for (uint32_t j = 0; j < K; j++) parser_end<Dir>(wk[j], 0);
    delete[] wk;
    BlockShared* sh = mems[0].sh;
    init_contexts(sh->cstate, d, 0, 1);
    CoderDecompress c;
    c.begin(&sh->sink, a.out_arena + io.out_off, io.out_cap, (d.n_streams & AVR_SLICE_ESCAPED) ? int((d.n_streams >> AVR_ESC_ZEROS_SHIFT) & 3) : -1);
    c.run_wide(w.ctl->ring, K, nullptr, 0, sh->tab, sh->cstate, sh->stage, 0);
    int status = AVR_SLICE_OK;
    uint32_t mbs = 0;
    uint64_t bins = 0;
    for (uint32_t j = 0; j < K; j++) {
      const RingCtl* r = &w.ctl->ring[j];
      if (status == AVR_SLICE_OK && r->end_status != AVR_SLICE_OK) status = int(r->end_status);
      mbs += r->end_mbs;
      bins += uint64_t(r->end_bins_lo) | uint64_t(r->end_bins_hi) << 32;
    }
    if (status == AVR_SLICE_OK && mbs != g.total_mbs) status = AVR_SLICE_CORRUPT;
    if (status == AVR_SLICE_OK && c.overflowed()) status = AVR_SLICE_OVERFLOW;
    for (uint32_t j = 0; j < K; j++) for (int i = 0; i < 8; i++) w.ctl->hist[j][i] = mems[1 + j].sh->phome.hist[i];
    wide_result(a, s, io, status, c.out_bytes(), mbs, bins);
    write_hist_rows(a, s, w.ctl->hist, K, mbs, bins);
  }
}
#endif

}  // namespace avr
