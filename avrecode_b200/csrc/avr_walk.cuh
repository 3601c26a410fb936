// avr_walk.cuh -- the H.264 slice_data context-selection state machine (parser warp, lane 0).
//
// Restates, for frame macroblocks in 4:2:0 / 8 bit / I, P and B slices, what libavcodec does between
// two per-bin hook calls: ffmpeg/libavcodec/h264_cabac.c:1926-2480 (ff_h264_decode_mb_cabac and its
// syntax-element decoders), :1550-1780 (coded_block_flag contexts, residual blocks), the neighbour
// plumbing of h264_mvpred.h:356-834 reduced to the syntax-level facts context selection reads, and
// avrecode's significance-map side channel (recode.cpp:844-930, 980-1026, 1160-1255, 1435-1492).
// A residual block is handled as a unit: significance mask in a 64-bit register, one nnz code, one
// contiguous estimator slab.  Pixel reconstruction, MV prediction and dequantisation -- the bulk of
// the reference's CPU time -- do not exist here because parsing never needs them.
//
// Register discipline: the walker holds SCALARS only.  Everything that is indexed at run time lives
// in shared memory (WalkShared, the neighbour row) or in bit masks, so that the whole walker --
// arithmetic-decoder state included -- stays in registers (0 bytes of stack frame); and every
// syntax element that repeats (residual blocks, partitions, reference indices) is walked by ONE
// loop with ONE inlined copy of its bin decoders, which keeps the kernel small enough for the
// instruction cache.
#pragma once
#include "avr_core.cuh"

namespace avr {

struct alignas(4) MbCtx {          // what later macroblocks need from a decoded one to select contexts / model keys
  uint16_t flags;       // MBF_*
  uint16_t cbp;         // cbp_table semantics: luma 0..3, chroma 4..5, 0x40<<c chroma DC, 0x100 luma DC
  uint8_t nz_bottom;    // coefficient flags of the bottom block row: luma x=0..3 -> bits 0..3, Cb x=0,1 -> 4,5, Cr -> 6,7
  uint8_t nz_right;     // right block column: luma y=0..3 -> bits 0..3, Cb y=0,1 -> 4,5, Cr -> 6,7
  uint8_t ref_pos;      // bit (4 * list + b8): ref_idx > 0
  uint8_t dir8;         // bit b8: that 8x8 is direct-predicted (0xF for B_Skip / B_Direct_16x16), h264_mvpred.h:723-753
  uint8_t mvd_bottom[4][2];  // list 0, bottom row x=0..3, {|mvdx|,|mvdy|} clipped to 70
  uint8_t mvd_right[4][2];   // list 0, right column y=0..3
  uint8_t nnz[20];           // avrecode model: stored nnz of the blocks a neighbour can ask for (k_nnz_slot)
  uint8_t mvd1_bottom[4][2]; // list 1 (B slices)
  uint8_t mvd1_right[4][2];
};
static_assert(sizeof(MbCtx) == 60, "MbCtx layout");
enum : int { MBCTX_WORDS = 15, MBCTX_NNZ_WORD = 6 };

// block index -> slot in MbCtx::nnz.  A left/above neighbour lookup (recode.cpp:419-471) can only land
// on the right column / bottom row of the neighbouring macroblock, on the first block of an 8x8
// quadrant (the "& ~3" rounding for 8x8 blocks), or on a DC slot.
AVR_TABLE_QUAL uint8_t k_nnz_slot[52] = {
  19, 19, 19, 19, 0, 1, 19, 2, 3, 19, 4, 5, 6, 7, 8, 9,            // luma 4, 5, 7, 8, 10..15
  19, 10, 11, 12, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19,   // Cb 17, 18, 19
  19, 13, 14, 15, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19,   // Cr 33, 34, 35
  16, 17, 18, 19};                                                    // DC 48, 49, 50
enum : uint16_t {
  MBF_I16 = 1, MBF_IPCM = 2, MBF_INTRA = 4, MBF_SKIP = 8, MBF_DIRECT = 16, MBF_T8x8 = 32,
  MBF_CHROMA_PRED_NZ = 64, MBF_L0 = 128, MBF_L1 = 256,
};

struct alignas(8) WalkShared {     // per-block scratch of the walker (shared memory on the device)
  MbCtx left;           // the previous macroblock
  MbCtx top;            // wide slices: copy of the macroblock above (their rows are in global memory)
  uint8_t mvd[2][16][2];  // |mvd| of the macroblock being walked, [list][4x4 block index][x, y]
  uint8_t nnz_mb[52];   // avrecode model: stored nnz of the macroblock being walked, by block index
};

enum : int { CTX_T8X8 = dctx(399) };
// B macroblock types 0..22 (h264data.h:127-151 b_mb_type_info): partition shape (0 direct, 1 16x16, 2 16x8,
// 3 8x16, 4 8x8) and the prediction lists of partition 0 / 1 (bit 0 = list 0, bit 1 = list 1)
AVR_TABLE_QUAL uint8_t k_b_shape[23] = {0, 1, 1, 1, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 2, 3, 4};
AVR_TABLE_QUAL uint8_t k_b_p0[23] = {0, 1, 2, 3, 1, 1, 2, 2, 1, 1, 2, 2, 1, 1, 2, 2, 3, 3, 3, 3, 3, 3, 3};
AVR_TABLE_QUAL uint8_t k_b_p1[23] = {0, 0, 0, 0, 1, 1, 2, 2, 2, 2, 1, 1, 3, 3, 3, 3, 1, 1, 2, 2, 3, 3, 3};
// B sub-macroblock types 0..12 (h264data.h:153-170): shape (0 direct, 1 8x8, 2 8x4, 3 4x8, 4 4x4) and lists
AVR_TABLE_QUAL uint8_t k_bsub_shape[13] = {0, 1, 1, 1, 2, 3, 2, 3, 2, 3, 4, 4, 4};
AVR_TABLE_QUAL uint8_t k_bsub_pred[13] = {0, 1, 2, 3, 1, 1, 2, 2, 3, 3, 1, 2, 3};

// block index <-> 4x4 raster position (h264.h:1004-1024 scan8 geometry)
AVR_DEV int blk_x(int n) { return (n & 1) + ((n >> 1) & 2); }
AVR_DEV int blk_y(int n) { return ((n >> 1) & 1) + ((n >> 2) & 2); }
AVR_DEV int blk_at(int x, int y) { return (x & 1) | ((y & 1) << 1) | ((x & 2) << 1) | ((y & 2) << 2); }

enum WalkStatus { WALK_OK = 0, WALK_UNSUPPORTED = 1, WALK_CORRUPT = 2, WALK_OVERFLOW = 3 };

// ------------------------------------------------------------------------------------------------
// The slice_data walker.  One instance per slice, lane 0 only.
// ------------------------------------------------------------------------------------------------
template <class Dir>
struct Walker {
  Dir dir;
  const SharedTables* tab;
  MbCtx* row;            // [mb_width] most recent macroblock of every column (shared memory)
  const MbCtx* top_mb;   // the macroblock above the current one (row_top[mb_x], or its copy in ws->top)
  MbCtx* row_top;        // the row above: the same array, except in a wide decompression (one buffer per parser warp)
  // Wide slices: one slice over K range-coder streams, macroblock row r in stream (r - first row) mod K, so that
  // K parser warps can decompress it as a wavefront -- row r at column x needs row r - 1 through column x only.
  uint32_t wide_k;       // 0: ordinary slice
  bool top_copy;         // the rows are in global memory, not in shared memory
  uint32_t cur_ring;     // compress: ring (= stream) of the current row
  MbCtx* rows_base;      // decompress: K row buffers, row r in buffer r mod K
  int first_row, last_row;
  volatile uint32_t* prog_mine;          // decompress: macroblocks of the picture (raster order) this warp has got past
  const volatile uint32_t* prog_above;   // ... and the warp that decodes the row above
  volatile uint32_t* abort_flag;
  uint32_t above_seen;
  bool blocked;          // host emulation only: the row above is not far enough yet
  WalkShared* ws;
  uint8_t* nnz_cur;      // FrameBuffer meta.num_nonzeros of the current picture, [mb][NNZ_STRIDE] (recode.cpp:624, 824-843)
  const uint8_t* nnz_prev;  // ... of the previous picture
  bool use_planes;       // false: the model lives and dies with this slice (chunk mode) -- the previous picture is
                         // empty and every neighbour count comes from the shared-memory row, no global traffic
  // slice constants
  uint32_t first_mb;
  int mbw, mbh, num_ref_l0, num_ref_l1;
  bool is_p, is_b, t8_mode, d8inf;
  // running state
  int mb_x, mb_y, mb_xy;
  int last_qp_delta;
  uint32_t n_bins_hi, n_mbs;
  int status;
  bool done;
  // per macroblock
  bool left_ok, top_ok, model_is8x8;
  uint32_t left_cbp, top_cbp, left_flags, top_flags, left_nzr, top_nzb, left_refpos, top_refpos, left_dir8, top_dir8;
  uint32_t cur_flags, cur_cbp, cur_refpos, cur_dir8;
  uint64_t nzmask;       // coefficient presence of the current macroblock, bit = block index (cbf contexts)

  // macroblock-level syntax: out-of-line primitives on the shared-memory copy of the engines (they count the bins)
  AVR_DEV int bin(int ctx) { return dir.bin(ctx); }   // ctx = dense ctxIdx
  AVR_DEV int bypass() { return dir.bypass(); }
  AVR_DEV int terminate() { return dir.terminate(); }
  AVR_DEV void fail(int s) { if (status == WALK_OK) status = s; }
  AVR_DEV uint32_t n_bins_lo() const { return dir.h->n_bins; }
  AVR_DEV uint64_t n_bins() const { return uint64_t(n_bins_lo()) | uint64_t(n_bins_hi) << 32; }

  // ---- avrecode model side channel: nnz of a coded block (recode.cpp:844-930) -------------------
  // get_neighbor_sub_mb (recode.cpp:419-471): frame geometry only -- slice edges are ignored on purpose
  AVR_DEV bool model_neighbor(bool above, int max_coeff, int blk, uint32_t* value) const {
    int where = 0, nb = blk;   // where: 0 = this macroblock, 1 = left, 2 = above
    if (blk >= 48) {
      if (above) { if (mb_y == 0) return false; where = 2; }
      else { if (mb_x == 0) return false; where = 1; }
    } else {
      int plane = blk & 48, x = blk_x(blk & 15), y = blk_y(blk & 15);
      if (above) { if (--y < 0) { if (mb_y == 0) return false; where = 2; y = 3; } }
      else { if (--x < 0) { if (mb_x == 0) return false; where = 1; x = 3; } }
      nb = plane | blk_at(x, y);
      if (max_coeff >= 32) nb &= ~3;
    }
    if (use_planes) {
      const int mbi = where == 0 ? mb_xy : (where == 1 ? mb_xy - 1 : mb_xy - mbw);
      *value = nnz_cur[size_t(mbi) * NNZ_STRIDE + nb];
    } else if (where == 0) *value = ws->nnz_mb[nb];
    else *value = (where == 1 ? ws->left.nnz : top_mb->nnz)[k_nnz_slot[nb]];
    return true;
  }
  // decompress: reads the block's nnz, 2 / 4 / 6 model-coded bits, least significant first (recode.cpp:858, 874-888)
  AVR_DEV uint32_t read_nnz(int cat, int blk, int max_coeff) {
    const int nbits = max_coeff > 16 ? 6 : max_coeff > 4 ? 4 : 2;
    uint32_t left_nz = 0, above_nz = 0;
    const bool has_left = model_neighbor(false, max_coeff, blk, &left_nz);
    model_neighbor(true, max_coeff, blk, &above_nz);
    const uint32_t prev_nz = use_planes ? nnz_prev[size_t(mb_xy) * NNZ_STRIDE + blk] : 0u;
    const uint32_t group = CNT_NZ + nz_group(cat, model_is8x8);
    uint32_t so_far = 0;
#pragma unroll 1
    for (int i = 0; i < nbits; i++) {
      const uint32_t cur_bit = 1u << i;
      const uint32_t lb = has_left ? (left_nz >= cur_bit) : 2;
      const uint32_t ab = above_nz ? (above_nz >= cur_bit) : 2;   // sic: value test (recode.cpp:881)
      const uint32_t idx = group + ((cur_bit - 1 + so_far) * 2 + (prev_nz >= cur_bit)) * 9 + lb * 3 + ab;
      if (dir.model_bit(*tab, idx)) so_far |= cur_bit;
    }
    return so_far;
  }

  // ---- one residual block (h264_cabac.c:1582-1780 + recode.cpp queueing 1160-1255 / 1435-1492) --
  // The significance map crosses to the coder warp as ONE entry (category, count, mask): the compress
  // coder derives the nnz bits and the model-coded significance bins from it, the decompress coder the
  // CABAC sig / last bins.  What stays here is what is serial by nature: pulling the map out of the
  // CABAC stream (compress) or out of the range coder (decompress).
  // Runs between dir.enter() and dir.leave(): adds the bins of the block to `rb`.
  AVR_DEV void residual(int cat, int blk, int max_coeff, ResBins& rb) {
    const bool is8 = max_coeff == 64;
    const int coefs = max_coeff - 1;
    const int abs_base = k_abs_base[cat];
    if (is8) model_is8x8 = true;          // before the nnz bits, as the compressor does
    uint32_t nnz_true, key_nnz;
    rb.nnz += max_coeff > 16 ? 6 : max_coeff > 4 ? 4 : 2;
    if constexpr (!Dir::kDecompress) {
      // 1. pull the significance map out of the CABAC stream (h264_cabac.c:1667-1681)
      const int sig_base = k_sig_base[cat], last_base = k_last_base[cat];
      uint64_t mask = 0;
      uint32_t nnz = 0;
      int last;
      bool ended = false;   // by last_significant_coeff_flag = 1
      if (!is8) {
        // 4x4-class blocks: the context of position i is base + i, all distinct, so the two contexts
        // that can come next (last[i], sig[i+1]) are loaded while sig[i] is still being decided
        uint32_t m32 = 0;
        U2 e_sig = dir.sig_peek(sig_base);
#pragma unroll 1
        for (last = 0; last < coefs; last++) {
          const U2 e_last = dir.sig_peek(last_base + last);
          const U2 e_next = dir.sig_peek(sig_base + last + 1);   // one past the block at the end: loaded, never used
          if (dir.sig_get_pre(*tab, e_sig, sig_base + last)) {
            m32 |= 1u << last;
            nnz++;
            if (dir.sig_get_pre(*tab, e_last, last_base + last)) { ended = true; last++; break; }
          }
          e_sig = e_next;
        }
        rb.sig += uint32_t(last) + nnz;   // one sig bin per position scanned, one last bin per coefficient found
        mask = m32;
      } else {
#pragma unroll 1
        for (last = 0; last < coefs; last++) {
          if (dir.sig_get(*tab, sig_base + tab->sig8[last])) {
            mask |= uint64_t(1) << last;
            nnz++;
            if (dir.sig_get(*tab, last_base + tab->last8[last])) { ended = true; last++; break; }
          }
        }
        rb.sig += uint32_t(last) + nnz;
      }
      // a scan that never met last_significant_coeff_flag = 1 ends in an inferred coefficient at the last position
      if (!ended) { mask |= uint64_t(1) << coefs; nnz++; }
      nnz_true = nnz; key_nnz = nnz & 63;   // stored count: 64 -> 0, else exact
      // 2. nnz side channel + 3. significance bins: both derived by the coder warp from this entry
      uint32_t left_nz = 0, above_nz = 0;
      const bool has_left = model_neighbor(false, max_coeff, blk, &left_nz);
      model_neighbor(true, max_coeff, blk, &above_nz);
      const uint32_t prev_nz = use_planes ? nnz_prev[size_t(mb_xy) * NNZ_STRIDE + blk] : 0u;
      dir.block_push(cat, nnz, model_is8x8, has_left, left_nz, above_nz, prev_nz, mask);
    } else {
      // 2. nnz side channel
      const uint32_t coded = read_nnz(cat, blk, max_coeff);
      nnz_true = coded == 0 ? uint32_t(max_coeff) : coded;
      if (nnz_true > uint32_t(max_coeff)) { fail(WALK_CORRUPT); nnz_true = uint32_t(max_coeff); }
      key_nnz = is8 ? coded : nnz_true;
      // 3. significance bins through the (cat, nnz) counter slab; last_significant_coeff_flag is
      //    never coded, it is implied by nnz (recode.cpp:1073, 1437-1438)
      const uint32_t slab = sig_slab(cat, key_nnz);
      uint32_t observed = 0;
      int zz = 0;
      uint64_t mask = 0;
      if (!is8) {
        // 4x4-class blocks: the estimator (and its reciprocal) of BOTH possible successors -- next position
        // with the same / with one more coefficient seen -- is fetched while the current bin is being
        // range-decoded; the three indices are always distinct, so no forwarding
        uint32_t m32 = 0;
        typename Dir::SigEntry e_cur = dir.sig_fetch(*tab, slab);
#pragma unroll 1
        while (observed < nnz_true && zz < coefs) {
          const uint32_t idx = slab + observed * SIG_ZZ + uint32_t(zz);
          const typename Dir::SigEntry e0 = dir.sig_fetch(*tab, idx + 1);
          const typename Dir::SigEntry e1 = dir.sig_fetch(*tab, idx + 1 + SIG_ZZ);
          const int s = dir.sig_decode(e_cur, idx);
          m32 |= uint32_t(s) << zz;
          observed += uint32_t(s);
          e_cur = s ? e1 : e0;
          zz++;
        }
        mask = m32;
      } else {
        dir.sig8_prefetch(slab, nnz_true);
#pragma unroll 1
        while (observed < nnz_true && zz < coefs) {
          const int s = dir.sig8_bit(*tab, slab + observed * SIG_ZZ + uint32_t(tab->sig8[zz]));
          mask |= uint64_t(s) << zz;
          observed += uint32_t(s);
          zz++;
        }
      }
      rb.sig += uint32_t(zz) + observed;   // the CABAC bins the re-encoder derives from the entry
      dir.block_push(cat, nnz_true, mask);
    }
    ws->nnz_mb[blk] = uint8_t(key_nnz);
    if (use_planes) nnz_cur[size_t(mb_xy) * NNZ_STRIDE + blk] = uint8_t(key_nnz);
    // 4. coefficient levels (h264_cabac.c:1724-1765): ordinary context bins + bypass
    // The first context of the NEXT coefficient is known as soon as this one's magnitude class is, so
    // it is fetched before the sign bin and has arrived when the sign is done.
    int node = 0;
    typename Dir::BinEntry e1 = dir.bin_fetch(*tab, abs_base + 1);
#pragma unroll 1
    for (uint32_t c = nnz_true; c > 0; c--) {
      const int ctx1 = abs_base + (node < 4 ? node + 1 : 0);      // coeff_abs_level1_ctx
      int gctx = abs_base + 5, run = 0, term = 0;
      bool escape = false;
      rb.level++;
      const int b1 = dir.lvl_first(*tab, e1, ctx1);
      if (b1 == 0) {
        node = node < 3 ? node + 1 : node;                         // transition after |level| == 1
      } else {
        gctx = abs_base + (node < 4 ? 5 : (node + 2 < 9 ? node + 2 : 9));  // coeff_abs_levelgt1_ctx
        node = node < 4 ? 4 : (node < 7 ? node + 1 : 7);           // transition after |level| > 1
        typename Dir::BinEntry g = dir.bin_fetch(*tab, gctx);      // a run on one context: state kept in registers
#pragma unroll 1
        while (run < 13) { rb.level++; if (!dir.lvl_run(*tab, g, gctx)) { term = 1; break; } run++; }
        dir.bin_store(g, gctx);
        escape = !term;                                             // |level| >= 15: Exp-Golomb suffix follows
      }
      e1 = dir.bin_fetch(*tab, abs_base + (node < 4 ? node + 1 : 0));
      if (!escape) {
        rb.bypass++;
        const int sg = dir.lvl_sign(*tab);
        dir.lvl_push(lvl_entry(ctx1, b1, gctx, run, term, 1, sg));   // the whole level as one ring entry
      } else {
        dir.lvl_push(lvl_entry(ctx1, b1, gctx, run, term, 0, 0));
        int j = 0;
#pragma unroll 1
        while (dir.lvl_bypass(*tab) && j < 30) j++;
        rb.bypass += uint32_t(2 * j + 2);   // prefix (j ones and the zero -- or 30 ones), j suffix bits, sign
        if (j >= 30) rb.bypass--;
#pragma unroll 1
        while (j--) dir.lvl_bypass(*tab);
        dir.lvl_bypass(*tab);   // sign
      }
    }
  }

  // ---- context helpers --------------------------------------------------------------------------
  AVR_DEV int nz_neighbor(int blk, bool above, bool intra) const {   // h264_mvpred.h:542-587 semantics
    const int plane = blk >> 4, x = blk_x(blk & 15), y = blk_y(blk & 15);
    if (above) {
      if (y > 0) return int((nzmask >> ((blk & 48) | blk_at(x, y - 1))) & 1);
      if (!top_ok) return intra ? 1 : 0;
      const int bit = plane == 0 ? x : (plane == 1 ? 4 + x : 6 + x);
      return int((top_nzb >> bit) & 1);
    }
    if (x > 0) return int((nzmask >> ((blk & 48) | blk_at(x - 1, y))) & 1);
    if (!left_ok) return intra ? 1 : 0;
    const int bit = plane == 0 ? y : (plane == 1 ? 4 + y : 6 + y);
    return int((left_nzr >> bit) & 1);
  }
  AVR_DEV int cbf_ctx(int cat, int blk, bool intra) const {   // h264_cabac.c:1550-1580
    int a, b;
    if (cat == 3) { const int i = blk - 49; a = int((left_cbp >> (6 + i)) & 1); b = int((top_cbp >> (6 + i)) & 1); }
    else if (cat == 0) { a = int((left_cbp >> 8) & 1); b = int((top_cbp >> 8) & 1); }
    else { a = nz_neighbor(blk, false, intra); b = nz_neighbor(blk, true, intra); }
    return k_cbf_base[cat] + a + 2 * b;
  }
  AVR_DEV int mvd_sum(int list, int n, int comp) const {   // h264_cabac.c:1539-1548
    const int x = blk_x(n), y = blk_y(n);
    const uint32_t uses = list ? MBF_L1 : MBF_L0;
    int a = 0, b = 0;
    if (x > 0) a = ws->mvd[list][blk_at(x - 1, y)][comp];
    else if (left_ok && (left_flags & uses)) a = list ? ws->left.mvd1_right[y][comp] : ws->left.mvd_right[y][comp];
    if (y > 0) b = ws->mvd[list][blk_at(x, y - 1)][comp];
    else if (top_ok && (top_flags & uses)) b = list ? top_mb->mvd1_bottom[x][comp] : top_mb->mvd_bottom[x][comp];
    return a + b;
  }
  // reference indices and motion vector differences run in register mode (between dir.enter() and dir.leave()): in
  // P and B slices they are the bulk of the macroblock-level bins, a quarter of all bins (profiles/r2k_bin_classes.txt)
  AVR_DEV int mvd_component(int ctxbase, int amvd, ResBins& rb) {   // h264_cabac.c:1502-1537; returns stored |mvd|
    rb.mb++;
    if (!dir.fast_bin(*tab, ctxbase + (amvd > 2) + (amvd > 32))) return 0;
    int mvd = 1;
    ctxbase += 3;
#pragma unroll 1
    for (;;) {
      if (mvd >= 9) break;
      rb.mb++;
      if (!dir.fast_bin(*tab, ctxbase)) break;
      if (mvd < 4) ctxbase++;
      mvd++;
    }
    if (mvd >= 9) {
      int k = 3;
#pragma unroll 1
      for (;;) {   // Exp-Golomb prefix, then k suffix bits: one inlined bypass site
        rb.mb++;
        if (!dir.lvl_bypass(*tab)) break;
        mvd += 1 << k; k++;
        if (k > 24) { fail(WALK_CORRUPT); break; }
      }
#pragma unroll 1
      while (k--) { rb.mb++; mvd += dir.lvl_bypass(*tab) << k; }
    }
    rb.mb++;
    dir.lvl_bypass(*tab);   // sign
    return mvd < 70 ? mvd : 70;
  }
  AVR_DEV void mvd_part(int list, int n, int w, int h, ResBins& rb) {
    int s[2];
#pragma unroll 1
    for (int comp = 0; comp < 2; comp++) {
      const int v = mvd_component(40 + 7 * comp, mvd_sum(list, n, comp), rb);
      if (comp == 0) s[0] = v; else s[1] = v;
    }
    const int x = blk_x(n), y = blk_y(n);
    for (int j = 0; j < h; j++)
      for (int i = 0; i < w; i++) { uint8_t* m = ws->mvd[list][blk_at(x + i, y + j)]; m[0] = uint8_t(s[0]); m[1] = uint8_t(s[1]); }
  }
  // "refa > 0 [&& !direct]" of h264_cabac.c:1475-1490: the 8x8 left of / above 4x4 block n
  AVR_DEV int ref_pos_neighbor(int list, int n, bool above) const {
    const int x = blk_x(n), y = blk_y(n);
    uint32_t refpos, dir8;
    int b8;
    if (above) {
      if (y > 0) { refpos = cur_refpos; dir8 = cur_dir8; b8 = (x >> 1) + 2 * ((y - 1) >> 1); }
      else { if (!top_ok) return 0; refpos = top_refpos; dir8 = top_dir8; b8 = 2 + (x >> 1); }
    } else {
      if (x > 0) { refpos = cur_refpos; dir8 = cur_dir8; b8 = ((x - 1) >> 1) + 2 * (y >> 1); }
      else { if (!left_ok) return 0; refpos = left_refpos; dir8 = left_dir8; b8 = 1 + 2 * (y >> 1); }
    }
    if (is_b && ((dir8 >> b8) & 1)) return 0;
    return int((refpos >> (4 * list + b8)) & 1);
  }
  AVR_DEV int read_ref(int list, int n, ResBins& rb) {   // h264_cabac.c:1473-1500 with the rc > 1 guard of :2232-2240
    const int rc = list ? num_ref_l1 : num_ref_l0;
    if (rc <= 1) return 0;
    int ctx = ref_pos_neighbor(list, n, false) + 2 * ref_pos_neighbor(list, n, true);
    int ref = 0;
#pragma unroll 1
    for (;;) {
      rb.mb++;
      if (!dir.fast_bin(*tab, 54 + ctx)) break;
      ref++;
      ctx = (ctx >> 2) + 4;
      if (ref >= 32) { fail(WALK_CORRUPT); break; }
    }
    if (ref >= rc) fail(WALK_CORRUPT);
    return ref;
  }

  AVR_DEV int intra_mb_type(int base, bool intra_slice) {   // h264_cabac.c:1300-1330
    int first = base;
    if (intra_slice) {
      if (left_ok && (left_flags & (MBF_I16 | MBF_IPCM))) first++;
      if (top_ok && (top_flags & (MBF_I16 | MBF_IPCM))) first++;
      base += 2;
    }
    if (!bin(first)) return 0;
    if (terminate()) return 25;
    const int is = intra_slice ? 1 : 0;
    int t = 1 + 12 * bin(base + 1);
    if (bin(base + 2)) t += 4 + 4 * bin(base + 2 + is);
    t += 2 * bin(base + 3 + is);
    t += bin(base + 3 + 2 * is);
    return t;
  }

  // ---- one macroblock (h264_cabac.c:1926-2480, frame macroblocks, 4:2:0) --------------------------
  AVR_DEV void macroblock() {
    cur_flags = 0; cur_cbp = 0; cur_refpos = 0; cur_dir8 = 0;
    nzmask = 0;
    model_is8x8 = false;
    for (int i = 0; i < 13; i++) reinterpret_cast<uint32_t*>(ws->nnz_mb)[i] = 0;
    left_ok = mb_x > 0 && uint32_t(mb_xy - 1) >= first_mb;
    top_ok = mb_y > 0 && mb_xy - mbw >= int(first_mb);
    {   // neighbour facts into registers
      const MbCtx* l = &ws->left;
      if (top_copy) {   // rows in global memory (a wide slice of a very wide picture): one burst into shared memory per macroblock
        const volatile uint32_t* g = reinterpret_cast<const volatile uint32_t*>(&row_top[mb_x]);   // another block wrote it: not through L1
        uint32_t w[MBCTX_WORDS];
        for (int i = 0; i < MBCTX_WORDS; i++) w[i] = g[i];
        for (int i = 0; i < MBCTX_WORDS; i++) reinterpret_cast<uint32_t*>(&ws->top)[i] = w[i];
        top_mb = &ws->top;
      } else top_mb = &row_top[mb_x];
      const MbCtx* t = top_mb;
      left_flags = l->flags; left_cbp = l->cbp; left_nzr = l->nz_right; left_refpos = l->ref_pos; left_dir8 = l->dir8;
      top_flags = t->flags; top_cbp = t->cbp; top_nzb = t->nz_bottom; top_refpos = t->ref_pos; top_dir8 = t->dir8;
    }
    const bool inter_slice = is_p || is_b;
    if (inter_slice) {   // mb_skip_flag, h264_cabac.c:1332-1367
      int ctx = is_b ? 13 : 0;
      if (left_ok && !(left_flags & MBF_SKIP)) ctx++;
      if (top_ok && !(top_flags & MBF_SKIP)) ctx++;
      if (bin(11 + ctx)) {   // decode_mb_skip (h264_mvpred.h:803-834): P_Skip, or B_Skip = direct with both lists
        cur_flags = is_b ? uint32_t(MBF_SKIP | MBF_DIRECT | MBF_L0 | MBF_L1) : uint32_t(MBF_SKIP | MBF_L0);
        cur_dir8 = is_b ? 0xFu : 0u;
        last_qp_delta = 0;
        for (int i = 0; i < 16; i++) reinterpret_cast<uint32_t*>(ws->mvd)[i] = 0;
        return;
      }
    }
    // macroblock type.  Inter macroblocks are reduced to: shape (0 direct 16x16, 1 16x16, 2 16x8, 3 8x16,
    // 4 8x8) and the prediction lists of partition 0 / 1 (bit 0 = list 0, bit 1 = list 1).
    int itype = -1, shape = -1;
    uint32_t pred0 = 1, pred1 = 1;
    bool want_intra = true;
    if (is_b) {   // h264_cabac.c:1973-2007
      int ctx = 0;
      if (left_ok && !(left_flags & MBF_DIRECT)) ctx++;
      if (top_ok && !(top_flags & MBF_DIRECT)) ctx++;
      int t = -1;
      if (!bin(27 + ctx)) t = 0;
      else if (!bin(27 + 3)) t = 1 + bin(27 + 5);
      else {
        int bits = bin(27 + 4) << 3;
        bits += bin(27 + 5) << 2;
        bits += bin(27 + 5) << 1;
        bits += bin(27 + 5);
        if (bits < 8) t = bits + 3;
        else if (bits == 13) t = -1;
        else if (bits == 14) t = 11;
        else if (bits == 15) t = 22;
        else { bits = (bits << 1) + bin(27 + 5); t = bits - 4; }
      }
      if (t >= 0) {
        want_intra = false;
        if (t > 22) { fail(WALK_CORRUPT); return; }
        shape = k_b_shape[t]; pred0 = k_b_p0[t]; pred1 = k_b_p1[t];
      }
    } else if (is_p && !bin(14)) {   // h264_cabac.c:2008-2023
      want_intra = false;
      int ptype;
      if (!bin(15)) ptype = 3 * bin(16);
      else ptype = 2 - bin(17);
      shape = ptype + 1;
    }
    if (want_intra) itype = intra_mb_type(is_b ? 32 : is_p ? 17 : 3, !inter_slice);

    const bool intra = itype >= 0;
    bool i16 = false;
    int cbp = 0;
    if (intra) {
      if (itype == 25) { fail(WALK_UNSUPPORTED); return; }   // I_PCM
      if (itype > 0) {
        i16 = true;
        const int k = itype - 1;
        const int chroma = (k >> 2) % 3;
        cbp = (k >= 12 ? 15 : 0) + 16 * chroma;
      }
      cur_flags = uint32_t(MBF_INTRA | (i16 ? MBF_I16 : 0));
    }

    // neighbour cbp with libavcodec's "unavailable" defaults (h264_mvpred.h:589-603)
    top_cbp = top_ok ? top_cbp : (intra ? 0x7CFu : 0x00Fu);
    left_cbp = left_ok ? ((left_cbp & 0x7F0) | (left_cbp & 2) | (left_cbp & 8)) : (intra ? 0x7CFu : 0x00Fu);
    const int nts = ((top_ok && (top_flags & MBF_T8x8)) ? 1 : 0) + ((left_ok && (left_flags & MBF_T8x8)) ? 1 : 0);
    bool t8_allowed = t8_mode;
    bool t8 = false;

    if (intra) {
      if (!i16) {   // I_NxN: transform_size_8x8_flag + prediction modes (h264_cabac.c:2081-2099, 1369-1381)
        int nmodes = 16;
        if (t8_allowed && bin(CTX_T8X8 + nts)) { t8 = true; nmodes = 4; }
        // up to 64 bins on two contexts: register mode, two inlined sites
        dir.enter(*tab);
        ResBins bins; bins.clear();
#pragma unroll 1
        for (int i = 0; i < nmodes; i++) {
          bins.pred++;
          if (!dir.fast_bin(*tab, 68)) {
            bins.pred += 3;
#pragma unroll 1
            for (int r = 0; r < 3; r++) dir.fast_bin(*tab, 69);
          }
        }
        dir.leave(bins);
      }
      // intra_chroma_pred_mode (h264_cabac.c:1383-1406)
      int ctx = 0;
      if (left_ok && (left_flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (top_ok && (top_flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (bin(64 + ctx)) {
        cur_flags |= MBF_CHROMA_PRED_NZ;
        if (bin(67)) bin(67);
      }
    } else {
      for (int i = 0; i < 16; i++) reinterpret_cast<uint32_t*>(ws->mvd)[i] = 0;
      const int lists = is_b ? 2 : 1;
      if (shape == 0) {   // B_Direct_16x16 (h264_cabac.c:2221-2225): nothing coded, everything inferred
        cur_flags = MBF_DIRECT | MBF_L0 | MBF_L1;
        cur_dir8 = 0xF;
        t8_allowed = t8_allowed && d8inf;
      } else {
        uint32_t subs = 0;      // 8x8 only, 2 bits per quadrant: 0 = 8x8, 1 = 8x4, 2 = 4x8, 3 = 4x4
        uint32_t subpred = 0;   // 8x8 only, 2 bits per quadrant: prediction lists (0 for a direct quadrant)
        if (shape == 4) {
#pragma unroll 1
          for (int i = 0; i < 4; i++) {
            int sub, pr = 1;
            if (is_b) {   // decode_cabac_b_mb_sub_type, h264_cabac.c:1455-1471
              int t;
              if (!bin(36)) t = 0;
              else if (!bin(37)) t = 1 + bin(39);
              else {
                t = 3;
                bool done13 = false;
                if (bin(38)) {
                  if (bin(39)) { t = 11 + bin(39); done13 = true; }
                  else t += 4;
                }
                if (!done13) { t += 2 * bin(39); t += bin(39); }
              }
              const int sh = k_bsub_shape[t];
              pr = k_bsub_pred[t];
              if (sh == 0) { cur_dir8 |= 1u << i; sub = 0; pr = 0; if (!d8inf) t8_allowed = false; }
              else sub = sh - 1;
            } else {      // decode_cabac_p_mb_sub_type, h264_cabac.c:1445-1454
              if (bin(21)) sub = 0;
              else if (!bin(22)) sub = 1;
              else sub = bin(23) ? 2 : 3;
            }
            subs |= uint32_t(sub) << (2 * i);
            subpred |= uint32_t(pr) << (2 * i);
          }
          if (subs) t8_allowed = false;   // get_dct8x8_allowed, h264.h:1165-1177
          cur_flags = is_b ? uint32_t(MBF_L0 | MBF_L1) : uint32_t(MBF_L0);   // B_8x8 counts as using both lists
        } else {
          const uint32_t any = pred0 | (shape == 1 ? 0u : pred1);
          cur_flags = ((any & 1) ? uint32_t(MBF_L0) : 0u) | ((any & 2) ? uint32_t(MBF_L1) : 0u);
        }
        // partitions: 16x16 one, 16x8 / 8x16 two, 8x8 four quadrants (h264_cabac.c:2114-2330); for every
        // list all reference indices first, then for every list all motion vector differences
        const int nq = shape == 4 ? 4 : (shape == 1 ? 1 : 2);
        dir.enter(*tab);
        ResBins bins; bins.clear();
#pragma unroll 1
        for (int list = 0; list < lists; list++) {
#pragma unroll 1
          for (int q = 0; q < nq; q++) {
            const uint32_t pr = shape == 4 ? ((subpred >> (2 * q)) & 3) : (q ? pred1 : pred0);
            if (!((pr >> list) & 1)) continue;
            const int n = shape == 2 ? 8 * q : (shape == 1 ? 0 : 4 * q);
            const uint32_t covers = shape == 1 ? 0xFu : shape == 2 ? (3u << (2 * q)) : shape == 3 ? (5u << q) : (1u << q);
            if (read_ref(list, n, bins) > 0) cur_refpos |= covers << (4 * list);
          }
        }
#pragma unroll 1
        for (int list = 0; list < lists; list++) {
#pragma unroll 1
          for (int q = 0; q < nq; q++) {
            const uint32_t pr = shape == 4 ? ((subpred >> (2 * q)) & 3) : (q ? pred1 : pred0);
            if (!((pr >> list) & 1)) continue;
            int base, w, h, cnt = 1, stride = 0;
            if (shape == 4) {
              const int sb = int((subs >> (2 * q)) & 3);
              base = 4 * q; w = sb < 2 ? 2 : 1; h = (sb == 0 || sb == 2) ? 2 : 1;
              cnt = sb == 0 ? 1 : (sb == 3 ? 4 : 2); stride = sb == 1 ? 2 : 1;
            } else if (shape == 1) { base = 0; w = 4; h = 4; }
            else if (shape == 2) { base = 8 * q; w = 4; h = 2; }
            else { base = 4 * q; w = 2; h = 4; }
#pragma unroll 1
            for (int j = 0; j < cnt; j++) mvd_part(list, base + j * stride, w, h, bins);
          }
        }
        dir.leave(bins);
      }
    }

    // Everything from here to the end of the macroblock runs in register mode: coded_block_pattern, the
    // transform-size flag, mb_qp_delta (six inlined bin sites) and the residual blocks.
    // First bin of mb_qp_delta: in a wide slice the first macroblock of a row cannot see how the previous row ended, so
    // that bin crosses the ring as a special entry -- on the home state, before the region, and only once it is known
    // that the macroblock has one (I16: no coded_block_pattern in between; else after it, which costs such a
    // macroblock a second enter / leave pair).
    const bool row_qpd = wide_k && mb_x == 0;
    const int qctx = 60 + (last_qp_delta != 0);
    int q1 = -1;
    if (row_qpd && i16) q1 = dir.bin_row_qpd(qctx);
    dir.enter(*tab);
    ResBins bins; bins.clear();
    if (!i16) {   // coded_block_pattern (h264_cabac.c:1408-1443)
      const uint32_t a = left_cbp, b = top_cbp;
      uint32_t c = 0;
#pragma unroll 1
      for (int i = 0; i < 4; i++) {   // luma: the 8x8 to the left / above, inside the macroblock or in the neighbour
        const uint32_t lf = (i & 1) ? (c >> (i - 1)) & 1 : (a >> (1 + i)) & 1;
        const uint32_t tp = (i & 2) ? (c >> (i - 2)) & 1 : (b >> (2 + i)) & 1;
        bins.mb++;
        c |= uint32_t(dir.fast_bin(*tab, 73 + int(!lf) + 2 * int(!tp))) << i;
      }
      const int ca = int((a >> 4) & 3), cb = int((b >> 4) & 3);
      bins.mb++;
      if (dir.fast_bin(*tab, 77 + (ca > 0) + 2 * (cb > 0))) { bins.mb++; c |= uint32_t(1 + dir.fast_bin(*tab, 77 + 4 + (ca == 2) + 2 * (cb == 2))) << 4; }
      cbp = int(c);
    }
    cur_cbp = uint32_t(cbp);
    if (t8_allowed && (cbp & 15) && !intra) { bins.mb++; t8 = dir.fast_bin(*tab, CTX_T8X8 + nts) != 0; }   // h264_cabac.c:2348-2350
    if (t8) cur_flags |= MBF_T8x8;

    if (cbp || i16) {
      // mb_qp_delta (h264_cabac.c:2388-2416)
      if (row_qpd && !i16) { dir.leave(bins); bins.clear(); q1 = dir.bin_row_qpd(qctx); dir.enter(*tab); }
      if (q1 < 0) { bins.mb++; q1 = dir.fast_bin(*tab, qctx); }
      if (q1) {
        int val = 1, ctx = 2;
#pragma unroll 1
        for (;;) {
          bins.mb++;
          if (!dir.fast_bin(*tab, 60 + ctx)) break;
          ctx = 3; val++;
          if (val > 102) { fail(WALK_CORRUPT); break; }
        }
        last_qp_delta = val;   // only "!= 0" is ever read
      } else last_qp_delta = 0;
      // residual blocks in coding order (h264_cabac.c:1874-1920, 2449-2463), one loop, one call site:
      //   k = 0        Intra16x16 DC          cat 0, block 48
      //   k = 1..16    luma 4x4 blocks 0..15  cat 1 (I16 AC) / 2 (4x4) / 5 (8x8, first block of a quadrant only)
      //   k = 17,18    chroma DC Cb, Cr       cat 3, blocks 49, 50
      //   k = 19..26   chroma AC              cat 4, blocks 16..19, 32..35
      // coded_block_flag bins: one inlined site
#pragma unroll 1
      for (int k = i16 ? 0 : 1; k < 27; k++) {
        int cat, blk, maxc;
        bool has_cbf = true;
        if (k == 0) { cat = 0; blk = 48; maxc = 16; }
        else if (k <= 16) {
          blk = k - 1;
          if (!(cbp & (1 << (blk >> 2)))) { k = (blk | 3) + 1; continue; }   // quadrant not coded: on to the next one
          if (i16) { cat = 1; maxc = 15; }
          else if (t8) { if (blk & 3) continue; cat = 5; maxc = 64; has_cbf = false; }
          else { cat = 2; maxc = 16; }
        } else if (k <= 18) {
          if (!(cbp & 0x30)) break;
          cat = 3; blk = 49 + (k - 17); maxc = 4;
        } else {
          if (!(cbp & 0x20)) break;
          const int j = k - 19;
          cat = 4; blk = 16 + 16 * (j >> 2) + (j & 3); maxc = 15;
        }
        if (has_cbf) { bins.cbf++; if (!dir.fast_bin(*tab, cbf_ctx(cat, blk, intra))) continue; }
        residual(cat, blk, maxc, bins);
        if (cat == 0) cur_cbp |= 0x100;
        else if (cat == 3) cur_cbp |= uint32_t(0x40) << (blk - 49);
        else if (cat == 5) nzmask |= uint64_t(0xF) << blk;
        else nzmask |= uint64_t(1) << blk;
      }
    } else last_qp_delta = 0;
    dir.leave(bins);
  }

  // what the neighbours will need from this macroblock -> row[mb_x] and ws->left
  AVR_DEV void store_mb() {
    const uint64_t m = nzmask;
    // bottom row: luma blocks 10, 11, 14, 15; Cb 18, 19; Cr 34, 35.  right column: luma 5, 7, 13, 15; Cb 17, 19; Cr 33, 35
    const uint32_t nb = uint32_t((m >> 10) & 3) | uint32_t((m >> 14) & 3) << 2 | uint32_t((m >> 18) & 3) << 4 | uint32_t((m >> 34) & 3) << 6;
    const uint32_t nr = uint32_t((m >> 5) & 1) | uint32_t((m >> 7) & 1) << 1 | uint32_t((m >> 13) & 1) << 2 | uint32_t((m >> 15) & 1) << 3 |
                        uint32_t((m >> 17) & 1) << 4 | uint32_t((m >> 19) & 1) << 5 | uint32_t((m >> 33) & 1) << 6 | uint32_t((m >> 35) & 1) << 7;
    uint32_t w[MBCTX_WORDS];
    {   // the 19 model counts a neighbour can ask for (k_nnz_slot order)
      const uint8_t* z = ws->nnz_mb;
      w[6] = uint32_t(z[4]) | uint32_t(z[5]) << 8 | uint32_t(z[7]) << 16 | uint32_t(z[8]) << 24;
      w[7] = uint32_t(z[10]) | uint32_t(z[11]) << 8 | uint32_t(z[12]) << 16 | uint32_t(z[13]) << 24;
      w[8] = uint32_t(z[14]) | uint32_t(z[15]) << 8 | uint32_t(z[17]) << 16 | uint32_t(z[18]) << 24;
      w[9] = uint32_t(z[19]) | uint32_t(z[33]) << 8 | uint32_t(z[34]) << 16 | uint32_t(z[35]) << 24;
      w[10] = uint32_t(z[48]) | uint32_t(z[49]) << 8 | uint32_t(z[50]) << 16;
    }
    w[0] = (cur_flags & 0xFFFF) | (cur_cbp & 0xFFFF) << 16;
    w[1] = nb | nr << 8 | (cur_refpos & 0xFF) << 16 | (cur_dir8 & 0xF) << 24;
    if (cur_flags & MBF_INTRA) { w[2] = w[3] = w[4] = w[5] = 0; w[11] = w[12] = w[13] = w[14] = 0; }
    else {
      const uint16_t* mv = reinterpret_cast<const uint16_t*>(ws->mvd);   // one u16 = {|mvdx|, |mvdy|} of a 4x4 block; list 1 at +16
      w[2] = uint32_t(mv[10]) | uint32_t(mv[11]) << 16;   // bottom row x = 0..3: blocks 10, 11, 14, 15
      w[3] = uint32_t(mv[14]) | uint32_t(mv[15]) << 16;
      w[4] = uint32_t(mv[5]) | uint32_t(mv[7]) << 16;     // right column y = 0..3: blocks 5, 7, 13, 15
      w[5] = uint32_t(mv[13]) | uint32_t(mv[15]) << 16;
      w[11] = uint32_t(mv[16 + 10]) | uint32_t(mv[16 + 11]) << 16;
      w[12] = uint32_t(mv[16 + 14]) | uint32_t(mv[16 + 15]) << 16;
      w[13] = uint32_t(mv[16 + 5]) | uint32_t(mv[16 + 7]) << 16;
      w[14] = uint32_t(mv[16 + 13]) | uint32_t(mv[16 + 15]) << 16;
    }
    uint32_t* r = reinterpret_cast<uint32_t*>(&row[mb_x]);
    uint32_t* l = reinterpret_cast<uint32_t*>(&ws->left);
    for (int i = 0; i < MBCTX_WORDS; i++) { r[i] = w[i]; l[i] = w[i]; }
  }

  // ---- slice loop (h264_slice.c:2366-2443) -------------------------------------------------------
  // wide decompression: may macroblock (mb_x, mb_y) start, i.e. is the row above done through column mb_x?
  AVR_DEV bool top_ready() {
    if (mb_y <= first_row) return true;
    const uint32_t need = uint32_t((mb_y - 1) * mbw + mb_x + 1);
    if (above_seen >= need) return true;
#if defined(__CUDACC__) && !defined(AVR_HOST_EMULATION)
    const long long t0 = clock64();
    for (uint32_t polls = 0;;) {
      above_seen = *prog_above;
      if (*abort_flag) { fail(WALK_CORRUPT); done = true; return false; }
      if (above_seen >= need) break;
      wide_backoff(polls);
      wait_watchdog(t0);
    }
    ring_fence(true);
    return true;
#else
    above_seen = *prog_above;
    if (above_seen >= need) return true;
    if (*abort_flag) { fail(WALK_CORRUPT); done = true; return false; }
    blocked = true;
    return false;
#endif
  }
  // stream: wide decompression only -- which of the K streams this walker decodes; total_mbs from the blob header
  AVR_DEV void start(const avr_slice_desc& d, uint32_t stream = 0, uint32_t total_mbs = 0) {
    mbw = d.mb_width; mbh = d.mb_height;
    first_mb = d.first_mb;
    is_p = d.slice_type == AVR_SLICE_P;
    is_b = d.slice_type == AVR_SLICE_B;
    d8inf = (d.flags & AVR_SLICE_DIRECT_8X8_INFERENCE) != 0;
    t8_mode = (d.flags & AVR_SLICE_TRANSFORM_8X8) != 0;
    num_ref_l0 = d.num_ref_idx_l0;
    num_ref_l1 = d.num_ref_idx_l1;
    mb_xy = int(d.first_mb);
    mb_x = mb_xy % (mbw > 0 ? mbw : 1); mb_y = mb_xy / (mbw > 0 ? mbw : 1);
    last_qp_delta = 0;
    n_bins_hi = 0;
    status = WALK_OK;
    n_mbs = 0;
    done = false;
    blocked = false;
    first_row = mb_y;
    cur_ring = 0;
    above_seen = 0;
    if (Dir::kDecompress && wide_k) {
      last_row = int((d.first_mb + (total_mbs ? total_mbs : 1u) - 1) / uint32_t(mbw > 0 ? mbw : 1));
      if (stream > 0) { mb_x = 0; mb_y = first_row + int(stream); mb_xy = mb_y * mbw; }
      row = rows_base + (uint32_t(mb_y) % wide_k) * uint32_t(mbw);
      row_top = rows_base + ((uint32_t(mb_y) + wide_k - 1) % wide_k) * uint32_t(mbw);
      if (mb_y > last_row) { done = true; return; }   // more streams than rows: nothing in this one
    }
    for (int i = 0; i < MBCTX_WORDS; i++) reinterpret_cast<uint32_t*>(&ws->left)[i] = 0;
    if (d.slice_type > AVR_SLICE_I || d.chroma_format_idc != 1 ||
        (d.flags & (AVR_SLICE_MBAFF | AVR_SLICE_FIELD)) || mbw < 1 || mbh < 1 || mb_xy >= mbw * mbh) {
      status = WALK_UNSUPPORTED; done = true; return;
    }
    dir.start_bins(*tab);
  }
  AVR_DEV void step() {
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; return; }
    if (Dir::kDecompress && wide_k) { blocked = false; if (!top_ready()) return; }
    const uint32_t bins_before = n_bins_lo();
    macroblock();
    n_mbs++;
    int eos = 0;
    if (status == WALK_OK) eos = terminate();
    if (eos && !dir.consumed_all()) fail(WALK_CORRUPT);
    if (n_bins_lo() < bins_before) n_bins_hi++;
    store_mb();
    mb_xy++;
    bool wrapped = false;
    if (++mb_x >= mbw) { mb_x = 0; mb_y++; wrapped = true; }
    if (Dir::kDecompress && wide_k) { ring_fence(true); *prog_mine = uint32_t(mb_xy); }   // the row below may use this macroblock
    home_publish(dir.h);
    if (status != WALK_OK) { done = true; return; }
    if (eos) { done = true; return; }
    if (wide_k && wrapped) {
      if (!Dir::kDecompress) {   // next row -> next stream
        cur_ring = cur_ring + 1 == wide_k ? 0 : cur_ring + 1;
        home_switch_to(dir.h, cur_ring);
      } else {                   // on to this warp's next row, K - 1 rows further down
        dir.row_end(last_qp_delta != 0);
        mb_y += int(wide_k) - 1;
        mb_xy += (int(wide_k) - 1) * mbw;
        if (mb_y > last_row) { done = true; return; }
        row = rows_base + (uint32_t(mb_y) % wide_k) * uint32_t(mbw);
        row_top = rows_base + (uint32_t(mb_y - 1) % wide_k) * uint32_t(mbw);
      }
    }
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; }   // picture full without end_of_slice: reference asserts here
  }
};

}  // namespace avr
