// avr_walk.cuh -- the H.264 slice_data context-selection state machine (parser warp, lane 0).
//
// Restates, for frame macroblocks in 4:2:0 / 8 bit / I and P slices, what libavcodec does between
// two per-bin hook calls: ffmpeg/libavcodec/h264_cabac.c:1926-2480 (ff_h264_decode_mb_cabac and its
// syntax-element decoders), :1550-1780 (coded_block_flag contexts, residual blocks), the neighbour
// plumbing of h264_mvpred.h:356-834 reduced to the syntax-level facts context selection reads, and
// avrecode's significance-map side channel (recode.cpp:844-930, 980-1026, 1160-1255, 1435-1492).
// A residual block is handled as a unit: significance mask in a 64-bit register, one nnz code, one
// contiguous estimator slab.  Pixel reconstruction, MV prediction and dequantisation -- the bulk of
// the reference's CPU time -- do not exist here because parsing never needs them.
#pragma once
#include "avr_core.cuh"

namespace avr {

struct MbCtx {          // what later macroblocks need from a decoded one to select CABAC contexts
  uint16_t flags;       // MBF_*
  uint16_t cbp;         // cbp_table semantics: luma 0..3, chroma 4..5, 0x40<<c chroma DC, 0x100 luma DC
  uint8_t nz_bottom;    // coefficient flags of the bottom block row: luma x=0..3 -> bits 0..3, Cb x=0,1 -> 4,5, Cr -> 6,7
  uint8_t nz_right;     // right block column: luma y=0..3 -> bits 0..3, Cb y=0,1 -> 4,5, Cr -> 6,7
  uint8_t ref_pos;      // bit (list*4 + b8): ref_idx > 0
  uint8_t reserved;
  uint8_t mvd_bottom[4][2];  // list 0, bottom row x=0..3, {|mvdx|,|mvdy|} clipped to 70
  uint8_t mvd_right[4][2];   // list 0, right column y=0..3
};
enum : uint16_t {
  MBF_I16 = 1, MBF_IPCM = 2, MBF_INTRA = 4, MBF_SKIP = 8, MBF_DIRECT = 16, MBF_T8x8 = 32,
  MBF_CHROMA_PRED_NZ = 64, MBF_L0 = 128, MBF_L1 = 256,
};

// ------------------------------------------------------------------------------------------------
// Model view of one job.
// ------------------------------------------------------------------------------------------------
struct ModelMem {
  uint8_t* nnz[2];    // FrameBuffer meta.num_nonzeros of the two pictures (global memory), [mb][NNZ_STRIDE]
  int cur;            // index of the current frame buffer (recode.cpp:624, 824-843)
};

// (dense) ctxIdx bases per block category 0..5 (h264_cabac.c:1589-1599, :1556)
AVR_TABLE_QUAL int16_t k_sig_base[6] = {105, 120, 134, 149, 152, dctx(402)};
AVR_TABLE_QUAL int16_t k_last_base[6] = {166, 181, 195, 210, 213, dctx(417)};
AVR_TABLE_QUAL int16_t k_abs_base[6] = {227, 237, 247, 257, 266, dctx(426)};
AVR_TABLE_QUAL int16_t k_cbf_base[6] = {85, 89, 93, 97, 101, dctx(1012)};
enum : int { CTX_T8X8 = dctx(399) };

// block index <-> 4x4 raster position (h264.h:1004-1024 scan8 geometry)
AVR_DEV int blk_x(int n) { return (n & 1) + ((n >> 1) & 2); }
AVR_DEV int blk_y(int n) { return ((n >> 1) & 1) + ((n >> 2) & 2); }
AVR_DEV int blk_at(int x, int y) { return (x & 1) | ((y & 1) << 1) | ((x & 2) << 1) | ((y & 2) << 2); }

// ------------------------------------------------------------------------------------------------
// The slice_data walker.  One instance per slice, lane 0 only.
// ------------------------------------------------------------------------------------------------
enum WalkStatus { WALK_OK = 0, WALK_UNSUPPORTED = 1, WALK_CORRUPT = 2, WALK_OVERFLOW = 3 };

template <class Dir>
struct Walker {
  Dir dir;
  const SharedTables* tab;
  MbCtx* row;            // [mb_width] most recent macroblock of every column
  ModelMem mm;
  avr_slice_desc d;
  // running state
  int mbw, mbh, mb_x, mb_y, mb_xy;
  int last_qp_delta;
  MbCtx left, cur;       // left neighbour (previous macroblock), macroblock being walked
  bool left_ok, top_ok;
  uint32_t left_cbp, top_cbp;
  uint64_t n_bins;
  int status;
  bool model_is8x8;      // BlockMeta::is_8x8 of the current macroblock (recode.cpp:944)
  uint8_t cur_nz[48];    // coefficient presence of the current macroblock, by block index (cbf contexts)
  uint8_t cur_mvd[16][2];
  int8_t cur_ref[4];

  AVR_DEV int bin(int ctx) { n_bins++; return dir.bin(*tab, ctx); }   // ctx = dense ctxIdx
  AVR_DEV int bypass() { n_bins++; return dir.bypass(*tab); }
  AVR_DEV int terminate() { n_bins++; return dir.terminate(*tab); }
  AVR_DEV void fail(int s) { if (status == WALK_OK) status = s; }

  // ---- avrecode model side channel: nnz of a coded block (recode.cpp:844-930) -------------------
  AVR_DEV uint8_t* nnz_slot(int frame, int mbi, int blk) { return mm.nnz[frame] + size_t(mbi) * NNZ_STRIDE + blk; }

  // get_neighbor_sub_mb (recode.cpp:419-471): frame geometry only -- slice edges are ignored on purpose
  AVR_DEV bool model_neighbor(bool above, int max_coeff, int blk, uint32_t* value) {
    int mbi = mb_xy, nb = blk;
    if (blk >= 48) {
      if (above) { if (mb_y == 0) return false; mbi -= mbw; }
      else { if (mb_x == 0) return false; mbi -= 1; }
    } else {
      int plane = blk & 48, x = blk_x(blk & 15), y = blk_y(blk & 15);
      if (above) { if (--y < 0) { if (mb_y == 0) return false; mbi -= mbw; y = 3; } }
      else { if (--x < 0) { if (mb_x == 0) return false; mbi -= 1; x = 3; } }
      nb = plane | blk_at(x, y);
      if (max_coeff >= 32) nb &= ~3;
    }
    *value = *nnz_slot(mm.cur, mbi, nb);
    return true;
  }
  // codes (compress) or reads (decompress) the block's nnz; returns the value the compressor stores
  AVR_DEV uint32_t code_nnz(int cat, int blk, int max_coeff, uint32_t nnz) {
    const int nbits = max_coeff > 16 ? 6 : max_coeff > 4 ? 4 : 2;
    uint32_t left_nz = 0, above_nz = 0;
    const bool has_left = model_neighbor(false, max_coeff, blk, &left_nz);
    model_neighbor(true, max_coeff, blk, &above_nz);
    const uint32_t prev_nz = *nnz_slot(mm.cur ^ 1, mb_xy, blk);
    const uint32_t group = CNT_NZ + nz_group(cat, model_is8x8);
    uint32_t so_far = 0;
    for (int i = 0; i < nbits; i++) {
      const uint32_t cur_bit = 1u << i;
      const uint32_t lb = has_left ? (left_nz >= cur_bit) : 2;
      const uint32_t ab = above_nz ? (above_nz >= cur_bit) : 2;   // sic: value test (recode.cpp:881)
      const uint32_t idx = group + ((cur_bit - 1 + so_far) * 2 + (prev_nz >= cur_bit)) * 9 + lb * 3 + ab;
      if (dir.model_bit(*tab, idx, int((nnz >> i) & 1))) so_far |= cur_bit;
    }
    if (Dir::kDecompress) return so_far;
    return nnz;
  }

  // ---- one residual block (h264_cabac.c:1582-1780 + recode.cpp queueing 1160-1255 / 1435-1492) --
  AVR_DEV int residual(int cat, int blk, int max_coeff) {
    const bool is8 = max_coeff == 64;
    const int coefs = max_coeff - 1;
    const int sig_base = k_sig_base[cat], last_base = k_last_base[cat], abs_base = k_abs_base[cat];
    if (is8) model_is8x8 = true;          // before the nnz bits, as the compressor does
    uint32_t nnz = 0;
    uint64_t mask = 0;
    if (!Dir::kDecompress) {
      // 1. pull the significance map out of the CABAC stream (h264_cabac.c:1667-1681)
      int last;
      for (last = 0; last < coefs; last++) {
        n_bins++;
        if (dir.sig_get(*tab, sig_base + (is8 ? tab->sig8[last] : last))) {
          mask |= uint64_t(1) << last;
          nnz++;
          n_bins++;
          if (dir.sig_get(*tab, last_base + (is8 ? tab->last8[last] : last))) { last = max_coeff; break; }
        }
      }
      if (last == max_coeff - 1) { mask |= uint64_t(1) << last; nnz++; }
    }
    // 2. nnz side channel
    uint32_t coded = code_nnz(cat, blk, max_coeff, nnz & 63);
    uint32_t nnz_true, key_nnz;
    if (Dir::kDecompress) {
      nnz_true = coded == 0 ? uint32_t(max_coeff) : coded;
      if (nnz_true > uint32_t(max_coeff)) { fail(WALK_CORRUPT); nnz_true = uint32_t(max_coeff); }
      key_nnz = is8 ? coded : nnz_true;
    } else { nnz_true = nnz; key_nnz = nnz & 63; }
    *nnz_slot(mm.cur, mb_xy, blk) = uint8_t(key_nnz);   // stored count: 64 -> 0, else exact
    // 3. significance bins through the (cat, nnz) counter slab; last_significant_coeff_flag is
    //    never coded, it is implied by nnz (recode.cpp:1073, 1437-1438)
    const uint32_t slab = sig_slab(cat, key_nnz);
    uint32_t observed = 0;
    int zz = 0;
    while (observed < nnz_true && zz < coefs) {
      const int inc = is8 ? tab->sig8[zz] : zz;
      const int s = dir.model_bit(*tab, slab + observed * SIG_ZZ + uint32_t(inc), int((mask >> zz) & 1));
      if (Dir::kDecompress) { n_bins++; dir.sig_put(sig_base + inc, s); }
      if (s) {
        observed++;
        if (Dir::kDecompress) { n_bins++; dir.sig_put(last_base + (is8 ? tab->last8[zz] : zz), observed == nnz_true); }
      }
      zz++;
    }
    // 4. coefficient levels (h264_cabac.c:1724-1765): ordinary context bins + bypass
    int node = 0;
    for (uint32_t c = nnz_true; c > 0; c--) {
      const int l1 = node < 4 ? node + 1 : 0;                     // coeff_abs_level1_ctx
      if (bin(abs_base + l1) == 0) {
        node = node < 3 ? node + 1 : node;                         // transition after |level| == 1
        bypass();
      } else {
        const int gctx = abs_base + (node < 4 ? 5 : (node + 2 < 9 ? node + 2 : 9));  // coeff_abs_levelgt1_ctx
        node = node < 4 ? 4 : (node < 7 ? node + 1 : 7);           // transition after |level| > 1
        int abs = 2;
        while (abs < 15 && bin(gctx)) abs++;
        if (abs >= 15) {
          int j = 0;
          while (bypass() && j < 30) j++;
          while (j--) bypass();
        }
        bypass();
      }
    }
    return int(nnz_true);
  }

  // ---- context helpers --------------------------------------------------------------------------
  AVR_DEV int nz_neighbor(int blk, bool above, bool intra) const {   // h264_mvpred.h:542-587 semantics
    int plane = blk >> 4, x = blk_x(blk & 15), y = blk_y(blk & 15);
    int w = plane ? 2 : 4;
    if (above) {
      if (y > 0) return cur_nz[(blk & 48) | blk_at(x, y - 1)];
      if (!top_ok) return intra ? 1 : 0;
      const MbCtx& t = row[mb_x];
      int bit = plane == 0 ? x : (plane == 1 ? 4 + x : 6 + x);
      return (t.nz_bottom >> bit) & 1;
    }
    if (x > 0) return cur_nz[(blk & 48) | blk_at(x - 1, y)];
    (void)w;
    if (!left_ok) return intra ? 1 : 0;
    int bit = plane == 0 ? y : (plane == 1 ? 4 + y : 6 + y);
    return (left.nz_right >> bit) & 1;
  }
  AVR_DEV int cbf_ctx(int cat, int blk, bool intra) const {   // h264_cabac.c:1550-1580
    int a, b;
    if (cat == 3) { int i = blk - 49; a = (left_cbp >> (6 + i)) & 1; b = (top_cbp >> (6 + i)) & 1; }
    else if (cat == 0) { a = (left_cbp >> 8) & 1; b = (top_cbp >> 8) & 1; }
    else { a = nz_neighbor(blk, false, intra); b = nz_neighbor(blk, true, intra); }
    return k_cbf_base[cat] + a + 2 * b;
  }
  AVR_DEV int mvd_sum(int n, int comp) const {   // h264_cabac.c:1539-1548
    int x = blk_x(n), y = blk_y(n), a = 0, b = 0;
    if (x > 0) a = cur_mvd[blk_at(x - 1, y)][comp];
    else if (left_ok && (left.flags & MBF_L0)) a = left.mvd_right[y][comp];
    if (y > 0) b = cur_mvd[blk_at(x, y - 1)][comp];
    else if (top_ok && (row[mb_x].flags & MBF_L0)) b = row[mb_x].mvd_bottom[x][comp];
    return a + b;
  }
  AVR_DEV int mvd_component(int ctxbase, int amvd) {   // h264_cabac.c:1502-1537; returns stored |mvd|
    if (!bin(ctxbase + (amvd > 2) + (amvd > 32))) return 0;
    int mvd = 1;
    ctxbase += 3;
    while (mvd < 9 && bin(ctxbase)) { if (mvd < 4) ctxbase++; mvd++; }
    if (mvd >= 9) {
      int k = 3;
      while (bypass()) { mvd += 1 << k; k++; if (k > 24) { fail(WALK_CORRUPT); break; } }
      while (k--) mvd += bypass() << k;
    }
    bypass();   // sign
    return mvd < 70 ? mvd : 70;
  }
  AVR_DEV void mvd_part(int n, int w, int h) {
    int s0 = mvd_component(40, mvd_sum(n, 0));
    int s1 = mvd_component(47, mvd_sum(n, 1));
    int x = blk_x(n), y = blk_y(n);
    for (int j = 0; j < h; j++)
      for (int i = 0; i < w; i++) { uint8_t* m = cur_mvd[blk_at(x + i, y + j)]; m[0] = uint8_t(s0); m[1] = uint8_t(s1); }
  }
  AVR_DEV int ref_pos_neighbor(int n, bool above) const {   // "refa > 0" of h264_cabac.c:1475-1490 (P slices)
    int x = blk_x(n), y = blk_y(n);
    if (above) {
      if (y > 0) return cur_ref[(x >> 1) + 2 * ((y - 1) >> 1)] > 0;
      if (!top_ok) return 0;
      return (row[mb_x].ref_pos >> (2 + (x >> 1))) & 1;
    }
    if (x > 0) return cur_ref[((x - 1) >> 1) + 2 * (y >> 1)] > 0;
    if (!left_ok) return 0;
    return (left.ref_pos >> (1 + 2 * (y >> 1))) & 1;
  }
  AVR_DEV int read_ref(int n) {   // h264_cabac.c:1473-1500 with the rc > 1 guard of :2232-2240
    int rc = d.num_ref_idx_l0;
    if (rc <= 1) return 0;
    int ctx = ref_pos_neighbor(n, false) + 2 * ref_pos_neighbor(n, true);
    int ref = 0;
    while (bin(54 + ctx)) {
      ref++;
      ctx = (ctx >> 2) + 4;
      if (ref >= 32) { fail(WALK_CORRUPT); break; }
    }
    if (ref >= rc) fail(WALK_CORRUPT);
    return ref;
  }

  AVR_DEV int intra_mb_type(int base, bool intra_slice) {   // h264_cabac.c:1300-1330
    if (intra_slice) {
      int ctx = 0;
      if (left_ok && (left.flags & (MBF_I16 | MBF_IPCM))) ctx++;
      if (top_ok && (row[mb_x].flags & (MBF_I16 | MBF_IPCM))) ctx++;
      if (!bin(base + ctx)) return 0;
      base += 2;
    } else if (!bin(base)) return 0;
    if (terminate()) return 25;
    int is = intra_slice ? 1 : 0;
    int t = 1 + 12 * bin(base + 1);
    if (bin(base + 2)) t += 4 + 4 * bin(base + 2 + is);
    t += 2 * bin(base + 3 + is);
    t += bin(base + 3 + 2 * is);
    return t;
  }

  AVR_DEV void luma_block4(int cat, int n, int max_coeff, bool intra) {
    if (bin(cbf_ctx(cat, n, intra))) { residual(cat, n, max_coeff); cur_nz[n] = 1; }
  }

  // ---- one macroblock (h264_cabac.c:1926-2480, frame macroblocks, 4:2:0) --------------------------
  AVR_DEV void macroblock() {
    const bool is_p = d.slice_type == AVR_SLICE_P;
    cur = MbCtx();
    model_is8x8 = false;
    left_ok = mb_x > 0 && uint32_t(mb_xy - 1) >= d.first_mb;
    top_ok = mb_y > 0 && mb_xy - mbw >= int(d.first_mb);
    if (is_p) {   // mb_skip_flag, h264_cabac.c:1332-1367
      int ctx = 0;
      if (left_ok && !(left.flags & MBF_SKIP)) ctx++;
      if (top_ok && !(row[mb_x].flags & MBF_SKIP)) ctx++;
      if (bin(11 + ctx)) {
        cur.flags = MBF_SKIP | MBF_L0;
        last_qp_delta = 0;
        return;
      }
    }
    for (int i = 0; i < 48; i++) cur_nz[i] = 0;
    int itype = -1, ptype = -1;
    if (is_p) {   // h264_cabac.c:2008-2023
      if (!bin(14)) {
        if (!bin(15)) ptype = 3 * bin(16);
        else ptype = 2 - bin(17);
      } else itype = intra_mb_type(17, false);
    } else itype = intra_mb_type(3, true);

    const bool intra = itype >= 0;
    bool i16 = false;
    int cbp = 0;
    if (intra) {
      if (itype == 25) { fail(WALK_UNSUPPORTED); return; }   // I_PCM
      if (itype > 0) {
        i16 = true;
        int k = itype - 1;
        int chroma = (k >> 2) % 3;
        cbp = (k >= 12 ? 15 : 0) + 16 * chroma;
      }
      cur.flags = uint16_t(MBF_INTRA | (i16 ? MBF_I16 : 0));
    } else cur.flags = MBF_L0;

    // neighbour cbp with libavcodec's "unavailable" defaults (h264_mvpred.h:589-603)
    const MbCtx& top = row[mb_x];
    top_cbp = top_ok ? top.cbp : (intra ? 0x7CF : 0x00F);
    left_cbp = left_ok ? ((left.cbp & 0x7F0) | (left.cbp & 2) | ((left.cbp & 8))) : (intra ? 0x7CF : 0x00F);
    const int nts = ((top_ok && (top.flags & MBF_T8x8)) ? 1 : 0) + ((left_ok && (left.flags & MBF_T8x8)) ? 1 : 0);
    bool t8_allowed = (d.flags & AVR_SLICE_TRANSFORM_8X8) != 0;
    bool t8 = false;

    if (intra) {
      if (!i16) {   // I_NxN: transform_size_8x8_flag + prediction modes (h264_cabac.c:2081-2099, 1369-1381)
        int nmodes = 16;
        if (t8_allowed && bin(CTX_T8X8 + nts)) { t8 = true; nmodes = 4; }
        for (int i = 0; i < nmodes; i++)
          if (!bin(68)) { bin(69); bin(69); bin(69); }
      }
      // intra_chroma_pred_mode (h264_cabac.c:1383-1406)
      int ctx = 0;
      if (left_ok && (left.flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (top_ok && (top.flags & MBF_CHROMA_PRED_NZ)) ctx++;
      if (bin(64 + ctx)) {
        cur.flags |= MBF_CHROMA_PRED_NZ;
        if (bin(67)) bin(67);
      }
    } else {
      for (int i = 0; i < 16; i++) { cur_mvd[i][0] = 0; cur_mvd[i][1] = 0; }
      for (int i = 0; i < 4; i++) cur_ref[i] = 0;
      if (ptype == 3) {   // P_8x8 (h264_cabac.c:2114-2220)
        int sub[4];
        for (int i = 0; i < 4; i++) {   // decode_cabac_p_mb_sub_type :1445-1454
          if (bin(21)) sub[i] = 0;
          else if (!bin(22)) sub[i] = 1;
          else sub[i] = bin(23) ? 2 : 3;
        }
        for (int i = 0; i < 4; i++) cur_ref[i] = int8_t(read_ref(4 * i));
        for (int i = 0; i < 4; i++) if (sub[i] != 0) t8_allowed = false;   // get_dct8x8_allowed, h264.h:1165-1177
        for (int i = 0; i < 4; i++) {
          int n = 4 * i;
          if (sub[i] == 0) mvd_part(n, 2, 2);
          else if (sub[i] == 1) { mvd_part(n, 2, 1); mvd_part(n + 2, 2, 1); }
          else if (sub[i] == 2) { mvd_part(n, 1, 2); mvd_part(n + 1, 1, 2); }
          else for (int j = 0; j < 4; j++) mvd_part(n + j, 1, 1);
        }
      } else if (ptype == 0) {
        int r = read_ref(0);
        for (int i = 0; i < 4; i++) cur_ref[i] = int8_t(r);
        mvd_part(0, 4, 4);
      } else if (ptype == 1) {   // 16x8
        int r0 = read_ref(0); cur_ref[0] = cur_ref[1] = int8_t(r0);
        int r1 = read_ref(8); cur_ref[2] = cur_ref[3] = int8_t(r1);
        mvd_part(0, 4, 2);
        mvd_part(8, 4, 2);
      } else {                   // 8x16
        int r0 = read_ref(0); cur_ref[0] = cur_ref[2] = int8_t(r0);
        int r1 = read_ref(4); cur_ref[1] = cur_ref[3] = int8_t(r1);
        mvd_part(0, 2, 4);
        mvd_part(4, 2, 4);
      }
    }

    if (!i16) {   // coded_block_pattern (h264_cabac.c:1408-1443)
      uint32_t a = left_cbp, b = top_cbp;
      int c = 0;
      c += bin(73 + !(a & 2) + 2 * !(b & 4));
      c += bin(73 + !(c & 1) + 2 * !(b & 8)) << 1;
      c += bin(73 + !(a & 8) + 2 * !(c & 1)) << 2;
      c += bin(73 + !(c & 4) + 2 * !(c & 2)) << 3;
      int ca = (a >> 4) & 3, cb = (b >> 4) & 3;
      if (bin(77 + (ca > 0) + 2 * (cb > 0))) c |= (1 + bin(77 + 4 + (ca == 2) + 2 * (cb == 2))) << 4;
      cbp = c;
    }
    cur.cbp = uint16_t(cbp);
    if (t8_allowed && (cbp & 15) && !intra) t8 = bin(CTX_T8X8 + nts) != 0;   // h264_cabac.c:2348-2350
    if (t8) cur.flags |= MBF_T8x8;

    if (cbp || i16) {
      // mb_qp_delta (h264_cabac.c:2388-2416)
      if (bin(60 + (last_qp_delta != 0))) {
        int val = 1, ctx = 2;
        while (bin(60 + ctx)) { ctx = 3; val++; if (val > 102) { fail(WALK_CORRUPT); break; } }
        last_qp_delta = val;   // only "!= 0" is ever read
      } else last_qp_delta = 0;
      // luma (h264_cabac.c:1874-1920)
      if (i16) {
        if (bin(cbf_ctx(0, 48, true))) { residual(0, 48, 16); cur.cbp |= 0x100; }
        if (cbp & 15) for (int i = 0; i < 16; i++) luma_block4(1, i, 15, true);
      } else {
        for (int i8 = 0; i8 < 4; i8++) {
          if (!(cbp & (1 << i8))) continue;
          if (t8) { residual(5, 4 * i8, 64); for (int k = 0; k < 4; k++) cur_nz[4 * i8 + k] = 1; }
          else for (int i4 = 0; i4 < 4; i4++) luma_block4(2, 4 * i8 + i4, 16, intra);
        }
      }
      if (cbp & 0x30)   // chroma DC (h264_cabac.c:2449-2453)
        for (int c = 0; c < 2; c++)
          if (bin(cbf_ctx(3, 49 + c, intra))) { residual(3, 49 + c, 4); cur.cbp |= uint16_t(0x40 << c); }
      if (cbp & 0x20)   // chroma AC (h264_cabac.c:2455-2463)
        for (int c = 0; c < 2; c++)
          for (int i = 0; i < 4; i++) luma_block4(4, 16 + 16 * c + i, 15, intra);
    } else last_qp_delta = 0;

    // pack what the neighbours will need
    uint32_t nb = 0, nr = 0;
    for (int x = 0; x < 4; x++) nb |= uint32_t(cur_nz[blk_at(x, 3)]) << x;
    for (int y = 0; y < 4; y++) nr |= uint32_t(cur_nz[blk_at(3, y)]) << y;
    nb |= uint32_t(cur_nz[18]) << 4 | uint32_t(cur_nz[19]) << 5 | uint32_t(cur_nz[34]) << 6 | uint32_t(cur_nz[35]) << 7;
    nr |= uint32_t(cur_nz[17]) << 4 | uint32_t(cur_nz[19]) << 5 | uint32_t(cur_nz[33]) << 6 | uint32_t(cur_nz[35]) << 7;
    cur.nz_bottom = uint8_t(nb);
    cur.nz_right = uint8_t(nr);
    if (!intra) {
      uint32_t rp = 0;
      for (int i = 0; i < 4; i++) rp |= uint32_t(cur_ref[i] > 0) << i;
      cur.ref_pos = uint8_t(rp);
      for (int i = 0; i < 4; i++) {
        cur.mvd_bottom[i][0] = cur_mvd[blk_at(i, 3)][0]; cur.mvd_bottom[i][1] = cur_mvd[blk_at(i, 3)][1];
        cur.mvd_right[i][0] = cur_mvd[blk_at(3, i)][0]; cur.mvd_right[i][1] = cur_mvd[blk_at(3, i)][1];
      }
    }
  }

  // ---- slice loop (h264_slice.c:2366-2443) -------------------------------------------------------
  // Runs macroblocks until the slice ends or `budget` of them are done; returns true when finished.
  AVR_DEV void start() {
    mbw = d.mb_width; mbh = d.mb_height;
    mb_xy = int(d.first_mb);
    mb_x = mb_xy % mbw; mb_y = mb_xy / mbw;
    last_qp_delta = 0;
    n_bins = 0;
    status = WALK_OK;
    n_mbs = 0;
    done = false;
    left = MbCtx();
    if (d.slice_type > AVR_SLICE_I || d.slice_type == AVR_SLICE_B || d.chroma_format_idc != 1 ||
        (d.flags & (AVR_SLICE_MBAFF | AVR_SLICE_FIELD)) || mbw < 1 || mbh < 1 || mb_xy >= mbw * mbh) {
      status = WALK_UNSUPPORTED; done = true; return;
    }
    dir.start_bins();
  }
  uint32_t n_mbs;
  bool done;
  AVR_DEV void step() {
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; return; }
    macroblock();
    n_mbs++;
    int eos = 0;
    if (status == WALK_OK) eos = terminate();
    row[mb_x] = cur;
    left = cur;
    mb_xy++;
    if (++mb_x >= mbw) { mb_x = 0; mb_y++; }
    dir.q.publish();
    if (status != WALK_OK) { done = true; return; }
    if (eos) { done = true; return; }
    if (mb_y >= mbh) { fail(WALK_CORRUPT); done = true; }   // picture full without end_of_slice: reference asserts here
  }
};

}  // namespace avr
