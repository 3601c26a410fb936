// avr_probe.cu -- single-lane microbenchmarks of the per-bin primitives (C ABI: avr_probe).
//
// The hot path is a serial integer recurrence per slice, so its speed limit is "cycles per bin of one lane".
// These probes MEASURE that floor on the device instead of asserting it (bench.py reports them as
// roofline.bounds.latency.floor_cycles_per_bin): each runs ONE lane of ONE warp of ONE block over the exact
// engine code of avr_core.cuh in a loop and returns clock64() cycles per iteration.  Inputs are pseudo-random
// bytes / entries in shared memory: the engines do not care whether the bytes are a legal slice.
#include <cuda_runtime.h>
#include <stdint.h>
#include "avr_job.cuh"

using namespace avr;

namespace {

struct ProbeShared {
  SharedTables tab;
  RingShared ring;
  alignas(16) uint8_t window[WIN_BYTES];
  U2 cstate[N_CTX_PAD];
  uint16_t cnt[10240];
  uint8_t out[8192];
  RecRing records;
  ByteSink sink;
  ParserHome home;
};

struct ProbeArgs {
  uint32_t iters;
  uint64_t* result;   // [0] cycles, [1] iterations (bins) done, [2] checksum (keeps the work alive)
};

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

__device__ void probe_setup(ProbeShared* sh, int tid, int nthreads) {
  fill_shared_tables(&sh->tab, tid, nthreads);
  uint32_t s = 12345u + uint32_t(tid) * 977u;
  for (int i = tid; i < int(WIN_BYTES); i += nthreads) sh->window[i] = uint8_t(lcg(s));
  for (int i = tid; i < N_CTX_PAD; i += nthreads) {
    const uint32_t state = uint32_t(i * 7) % 100u;   // a spread of probabilities
    sh->cstate[i].x = ctx_word_x(state);
    sh->cstate[i].y = rlps4_of(state >> 1);
  }
  for (int i = tid; i < 10240; i += nthreads) sh->cnt[i] = EST_INIT;
  if (tid == 0) ring_reset(&sh->ring);
}

// 0: a chain of dependent integer instructions (the price of one instruction when nothing overlaps)
__device__ void probe_alu(ProbeShared*, const ProbeArgs& a, uint64_t* r) {
  uint32_t x = a.iters | 1u, y = 0x9E3779B9u;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
#pragma unroll
    for (int k = 0; k < 16; k++) { x = (x ^ y) + 0x3C6EF372u; x = (x << 1) | (x >> 31); }
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = uint64_t(a.iters) * 32; r[2] = x;
}

// 1: CABAC decision on one context (H.264 9.3.3.2.1): the recurrence alone
__device__ void probe_decide(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  ByteWindow w; w.attach(sh->window, 0x7FFFFFFFu, sh->window, 0);
  CabacIn c; c.init(w);
  U2 ctx = sh->cstate[5];
  uint32_t acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) { U2 n; acc += uint32_t(c.decide_core(w, sh->tab, ctx, &n)); ctx = n; }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = a.iters; r[2] = acc + c.pos;
}

// 9: the 64-bit range-coder recurrence alone (RangeEnc::put, arithmetic_code.h:106-126): reciprocals, positions and bins
//    in registers, nothing loaded inside the loop -- what one bin costs the range-coder warp at the very least
__device__ void probe_range_put(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  RangeEnc rc; rc.init(&sh->sink, sh->out, sizeof sh->out);
  const DivMagic m0 = sh->tab.div[7], m1 = sh->tab.div[23], m2 = sh->tab.div[64], m3 = sh->tab.div[90];
  const uint32_t bits = a.iters * 2654435761u;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i += 4) {
    rc.put(m0, int((bits >> (i & 15)) & 1), 3);
    rc.put(m1, int((bits >> ((i + 5) & 15)) & 1), 11);
    rc.put(m2, int((bits >> ((i + 9) & 15)) & 1), 40);
    rc.put(m3, int((bits >> ((i + 13) & 15)) & 1), 77);
    if (sh->sink.n > 2000) sh->sink.n = 0;
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = a.iters; r[2] = rc.low ^ rc.range;
}

// 2: the compress-side significance loop of a 4x4 block (distinct contexts, branch on the bin, last flag)
__device__ void probe_sigloop_c(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  Compress d;
  d.begin(&sh->home, sh->window, 0x7FFFFFFFu, sh->window, 0, &sh->ring, &sh->tab, sh->cnt, sh->cstate);
  d.start_bins(sh->tab);
  d.enter(sh->tab);
  uint64_t bins = 0, acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    const int sig_base = 134, last_base = 195, coefs = 15;
    uint64_t mask = 0;
    int last;
    U2 e_sig = d.sig_peek(sig_base);
#pragma unroll 1
    for (last = 0; last < coefs; last++) {
      const U2 e_last = d.sig_peek(last_base + last);
      const U2 e_next = d.sig_peek(sig_base + last + 1);
      bins++;
      if (d.sig_get_pre(sh->tab, e_sig, sig_base + last)) {
        mask |= uint64_t(1) << last;
        bins++;
        if (d.sig_get_pre(sh->tab, e_last, last_base + last)) break;
      }
      e_sig = e_next;
    }
    acc += mask;
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = bins; r[2] = acc + d.cabac.pos;
}

// 3: decompress-side bin: estimator load, exact division, range decode, update, store
__device__ void probe_coded(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  Decompress d;
  d.begin(&sh->home, sh->window, 0x7FFFFFFFu, sh->window, 0, &sh->ring, &sh->tab, sh->cnt, sh->cstate);
  d.start_bins(sh->tab);
  d.enter(sh->tab);
  uint32_t acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) acc += uint32_t(d.coded(sh->tab, (i * 5u) & 63u, LIMIT_ANY));
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = a.iters; r[2] = acc + d.rc.pos() + uint32_t(d.rc.low);
}

// 4: the decompress-side significance loop (both successors prefetched), without the ring push
__device__ void probe_sigloop_d(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  Decompress d;
  d.begin(&sh->home, sh->window, 0x7FFFFFFFu, sh->window, 0, &sh->ring, &sh->tab, sh->cnt, sh->cstate);
  d.start_bins(sh->tab);
  d.enter(sh->tab);
  uint64_t bins = 0, acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    const uint32_t slab = 64, nnz_true = 5, coefs = 15;
    uint32_t observed = 0, zz = 0;
    Decompress::SigEntry e_cur = d.sig_fetch(sh->tab, slab);
#pragma unroll 1
    while (observed < nnz_true && zz < coefs) {
      const uint32_t idx = slab + observed * SIG_ZZ + zz;
      const Decompress::SigEntry e0 = d.sig_fetch(sh->tab, idx + 1);
      const Decompress::SigEntry e1 = d.sig_fetch(sh->tab, idx + 1 + SIG_ZZ);
      const int s = d.sig_decode(e_cur, idx);
      bins++;
      if (s) { observed++; e_cur = e1; acc += zz; } else e_cur = e0;
      zz++;
    }
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = bins; r[2] = acc + d.rc.pos() + uint32_t(d.rc.low);
}

// Fills the ring with a residual-heavy mix of entries (4x4 blocks with their levels, a few plain bins), ends it with
// SYM_END and returns how many bins the coder warp will derive from it.
__device__ uint32_t fill_ring(ProbeShared* sh, bool decompress) {
  uint32_t s = 777u, n = 0, bins = 0;
  uint32_t* e = sh->ring.e;
  while (n + 24 < RING_ENTRIES) {
    // one block of category 2 (luma 4x4): a few coefficients in the first positions
    const uint32_t nnz = 1 + lcg(s) % 5;
    uint32_t mask = 0, placed = 0, zz = 0;
    while (placed < nnz && zz < 15) { if (lcg(s) % 3 == 0 || 15 - zz == nnz - placed) { mask |= 1u << zz; placed++; } zz++; }
    const uint32_t scanned = zz;   // positions up to and including the last coefficient
    if (decompress) {
      e[n++] = SYM_BLOCK | 2u | nnz << BLK_NNZ;
      e[n++] = mask;
      bins += scanned + nnz;                 // sig + last bins
    } else {
      e[n++] = SYM_BLOCK | 2u | nnz << BLK_NNZ | 1u << BLK_HAS_LEFT;
      e[n++] = (lcg(s) % 6) | (lcg(s) % 6) << 8;
      e[n++] = mask;
      bins += 4 + scanned;                   // nnz bits + model-coded sig bins
    }
    for (uint32_t c = 0; c < nnz; c++) {     // levels: mostly 1, sometimes a short run
      const uint32_t b1 = lcg(s) % 4 == 0, run = b1 ? lcg(s) % 3 : 0;
      e[n++] = lvl_entry(227 + 1 + int(c % 4), int(b1), 227 + 5, int(run), int(b1), 1, int(lcg(s) & 1));
      bins += 1 + (b1 ? run + 1 : 0) + 1;
    }
    e[n++] = (decompress ? 85u | SYM_DECISION : CNT_CTX + 85u) | (lcg(s) & 1) << SYM_BIT_SHIFT;   // a coded_block_flag
    bins += 1;
  }
  e[n++] = SYM_END;
  sh->ring.tail_pub = n;
  sh->ring.head_pub = 0;
  return bins;
}

// 5: the two warps behind the compress parser as a pipeline: estimator warp (warp 0) -> record ring -> range-coder
//    warp (warp 1), over a full ring of block / level entries
__device__ void probe_coder_c(ProbeShared* sh, const ProbeArgs& a, uint64_t* r, int tid) {
  const int lane = tid & 31;
  uint32_t per = 0;
  if (tid == 0) per = fill_ring(sh, false);
  __syncthreads();
  uint64_t bins = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    if (tid == 0) ring_reset(&sh->records);
    __syncthreads();
    if (tid < 32) {
      CoderCompress c;
      c.begin(&sh->records, sh->cnt);
      c.run(&sh->ring, nullptr, nullptr, 0, sh->tab, sh->cnt, nullptr, lane);
    } else {
      RangeWarp w;
      w.begin(&sh->sink, sh->out, sizeof sh->out);
      w.run(&sh->records, sh->tab, nullptr, lane);
    }
    __syncthreads();
    bins += per;
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = bins; r[2] = sh->sink.n;
}

// 8: the range-coder warp alone over a full record ring (RangeWarp::run)
__device__ void probe_range(ProbeShared* sh, const ProbeArgs& a, uint64_t* r, int lane) {
  if (lane == 0) {
    uint32_t s = 4242u;
    for (uint32_t i = 0; i + 1 < REC_ENTRIES; i++) {
      const uint32_t tot = 2 + lcg(s) % 90, pos = 1 + lcg(s) % (tot - 1);
      sh->records.e[i] = pos | tot << 8 | (lcg(s) % 3 == 0 ? 1u : 0u) << REC_BIT_SHIFT;
    }
    sh->records.e[REC_ENTRIES - 1] = SYM_END;
    sh->records.tail_pub = REC_ENTRIES; sh->records.head_pub = 0;
  }
  __syncwarp();
  uint64_t bins = 0;
  RangeWarp w;
  w.begin(&sh->sink, sh->out, sizeof sh->out);
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    w.run(&sh->records, sh->tab, nullptr, lane);
    if (lane == 0) sh->sink.n = 0;
    bins += REC_ENTRIES - 1;
    __syncwarp();
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = bins; r[2] = w.rc.low;
}

// 6: the decompress coder warp (lane 0 runs the CABAC re-encoder: context states + arithmetic) over a full ring
__device__ void probe_coder_d(ProbeShared* sh, const ProbeArgs& a, uint64_t* r, int lane) {
  uint32_t per = 0;
  if (lane == 0) per = fill_ring(sh, true);
  per = __shfl_sync(0xFFFFFFFFu, per, 0);
  uint64_t bins = 0;
  CoderDecompress c;
  c.begin(&sh->sink, sh->out, sizeof sh->out);
  __syncwarp();
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    c.run(&sh->ring, sh->tab, sh->cstate, nullptr, lane);
    if (lane == 0) sh->sink.n = 0;
    bins += per;
    __syncwarp();
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = bins; r[2] = c.cabac.low;
}

// 7: one macroblock-level bin through the out-of-line primitive (compress: CABAC decision on the shared-memory copy
//    of the engine + plain ring entry)
__device__ void probe_push(ProbeShared* sh, const ProbeArgs& a, uint64_t* r) {
  Compress d;
  d.begin(&sh->home, sh->window, 0x7FFFFFFFu, sh->window, 0, &sh->ring, &sh->tab, sh->cnt, sh->cstate);
  d.start_bins(sh->tab);
  uint32_t acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (uint32_t i = 0; i < a.iters; i++) {
    *reinterpret_cast<volatile uint32_t*>(&sh->ring.head_pub) = sh->home.tail;   // a consumer that always keeps up
    acc += uint32_t(d.bin(int(11 + (i & 7))));
  }
  const long long t1 = clock64();
  r[0] = uint64_t(t1 - t0); r[1] = a.iters; r[2] = acc;
}

template <int WHICH>
__global__ void __launch_bounds__(64, 1) avr_probe_kernel(ProbeArgs a) {
  __shared__ ProbeShared sh;
  probe_setup(&sh, threadIdx.x, int(blockDim.x));
  __syncthreads();
  if (WHICH == 5 || WHICH == 6 || WHICH == 8) {   // warp-wide roles
    uint64_t r[3] = {0, 0, 0};
    if (WHICH == 5) probe_coder_c(&sh, a, r, int(threadIdx.x));
    if (WHICH == 6) probe_coder_d(&sh, a, r, int(threadIdx.x));
    if (WHICH == 8) probe_range(&sh, a, r, int(threadIdx.x));
    if (threadIdx.x == 0) { a.result[0] = r[0]; a.result[1] = r[1]; a.result[2] = r[2]; }
  } else if (threadIdx.x == 0) {
    uint64_t r[3] = {0, 0, 0};
    if (WHICH == 0) probe_alu(&sh, a, r);
    if (WHICH == 1) probe_decide(&sh, a, r);
    if (WHICH == 2) probe_sigloop_c(&sh, a, r);
    if (WHICH == 3) probe_coded(&sh, a, r);
    if (WHICH == 4) probe_sigloop_d(&sh, a, r);
    if (WHICH == 7) probe_push(&sh, a, r);
    if (WHICH == 9) probe_range_put(&sh, a, r);
    a.result[0] = r[0]; a.result[1] = r[1]; a.result[2] = r[2];
  }
}

}  // namespace

extern "C" {

const char* avr_probe_name(int which) {
  static const char* const names[] = {
    "dependent integer instruction (ALU chain)",
    "CABAC decision, one context (CabacIn::decide_core)",
    "compress significance loop, 4x4 block (CABAC decode of sig/last bins)",
    "decompress bin: estimator + exact division + RangeDec::get + update (Decompress::coded)",
    "decompress significance loop, 4x4 block (both successors prefetched)",
    "compress, the two warps behind the parser as a pipeline: estimator warp -> records -> range-coder warp, over block / level entries",
    "decompress coder warp: CABAC re-encoder over a ring of block / level entries (CoderDecompress::run)",
    "macroblock-level bin, compress: out-of-line CABAC decision on the shared-memory engine copy + ring entry",
    "compress range-coder warp alone: reciprocal + RangeEnc::put over a ring of records (RangeWarp::run)",
    "64-bit range-coder recurrence alone, operands in registers (RangeEnc::put)",
  };
  return which >= 0 && which < int(sizeof names / sizeof names[0]) ? names[which] : nullptr;
}

// Runs probe `which` for `iters` iterations on `device`; returns cycles, bins (iterations) and a checksum.
int avr_probe(int device, int which, uint32_t iters, uint64_t* cycles, uint64_t* bins, uint64_t* checksum) {
  if (!cycles || !bins || !avr_probe_name(which)) return AVR_ERR_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return AVR_ERR_NO_DEVICE; }
  if (device < 0 || device >= ndev || cudaSetDevice(device) != cudaSuccess) return AVR_ERR_ARG;
  uint64_t* d = nullptr;
  if (cudaMalloc(reinterpret_cast<void**>(&d), 3 * sizeof(uint64_t)) != cudaSuccess) return AVR_ERR_NOMEM;
  ProbeArgs a; a.result = d;
  // the warp-wide probes work through a whole ring (~1500 bins) per iteration, the single-lane ones through one bin
  a.iters = (which == 5 || which == 6 || which == 8) ? iters / 64 + 1 : iters;
  for (int rep = 0; rep < 2; rep++) {   // the first run also loads the kernel
    switch (which) {
      case 0: avr_probe_kernel<0><<<1, 32>>>(a); break;
      case 1: avr_probe_kernel<1><<<1, 32>>>(a); break;
      case 2: avr_probe_kernel<2><<<1, 32>>>(a); break;
      case 3: avr_probe_kernel<3><<<1, 32>>>(a); break;
      case 4: avr_probe_kernel<4><<<1, 32>>>(a); break;
      case 5: avr_probe_kernel<5><<<1, 64>>>(a); break;
      case 6: avr_probe_kernel<6><<<1, 32>>>(a); break;
      case 7: avr_probe_kernel<7><<<1, 32>>>(a); break;
      case 8: avr_probe_kernel<8><<<1, 32>>>(a); break;
      default: avr_probe_kernel<9><<<1, 32>>>(a); break;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { cudaGetLastError(); cudaFree(d); return AVR_ERR_CUDA; }
  }
  uint64_t h[3] = {0, 0, 0};
  const cudaError_t e = cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return AVR_ERR_CUDA;
  *cycles = h[0]; *bins = h[1];
  if (checksum) *checksum = h[2];
  return AVR_OK;
}

}  // extern "C"
