// avr_lib.cu -- sm_100a kernels + the C ABI of include/avrecode_b200.h (device side).
//
// One persistent warp per job slot; jobs are handed out through an atomic counter in descending
// size order.  All algorithmic code is in avr_core.cuh / avr_job.cuh.  There is no CPU path in
// this library: without a CUDA device every entry point that needs one fails with
// AVR_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>
#include <vector>
#include "avr_job.cuh"

using namespace avr;

namespace {

struct KernelArgs {
  JobArgs a;
  const JobDesc* jobs;
  uint32_t n_jobs;
  uint32_t* counter;        // next job to hand out
  uint8_t* workspace;       // per-block slots: [estimators unless resident] [nnz plane 0] [nnz plane 1]
  uint64_t work_stride;     // bytes of workspace per resident block
  uint32_t nnz_frame_bytes; // bytes of one nnz plane inside a slot
  uint32_t max_mbw;         // neighbour-row entries per block
  uint8_t* wide_ws;         // per wide slice: WideCtl | neighbour rows [WIDE_MAX][wide_mbw] | ring entries [WIDE_MAX][far_cap]
  uint64_t wide_stride;
  uint32_t wide_mbw, far_cap;
};

__host__ __device__ inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
__host__ __device__ inline size_t wide_ctl_bytes() { return align_up(sizeof(WideCtl), 256); }
__host__ __device__ inline size_t wide_rows_bytes(uint32_t mbw) { return align_up(size_t(WIDE_MAX) * mbw * sizeof(MbCtx), 256); }
size_t row_bytes(uint32_t max_mbw) { return align_up(size_t(max_mbw) * sizeof(MbCtx), 16); }
size_t counters_bytes() { return align_up(size_t(CNT8_ENTRIES) * 2, 256); }   // split placement: the 8x8 slab of a block
size_t smem_bytes(uint32_t max_mbw, bool resident) {
  return sizeof(BlockShared) + row_bytes(max_mbw) + (resident ? size_t(MODEL_COUNTERS_PAD) : size_t(CNT_LOW_PAD)) * 2;
}

// Block = parser warp + coder warp.  Shared memory: BlockShared | MbCtx row[max_mbw] | estimators (RESIDENT).
// Persistent blocks take jobs from an atomic counter (longest first).
template <class Dir, bool RESIDENT>
__global__ void __launch_bounds__(AVR_THREADS_OF(Dir), RESIDENT ? 2 : 4) avr_job_kernel(KernelArgs k) {
  extern __shared__ __align__(16) uint8_t smem[];
  JobMem m;
  m.sh = reinterpret_cast<BlockShared*>(smem);
  m.row = reinterpret_cast<MbCtx*>(smem + sizeof(BlockShared));
  uint8_t* slot = k.workspace + uint64_t(blockIdx.x) * k.work_stride;
  const size_t rb = (size_t(k.max_mbw) * sizeof(MbCtx) + 15) / 16 * 16;
  m.cnt = reinterpret_cast<uint16_t*>(smem + sizeof(BlockShared) + rb);
  if (RESIDENT) m.cnt8 = m.cnt;
  else { m.cnt8 = reinterpret_cast<uint16_t*>(slot) - CNT8_SPLIT; slot += (size_t(CNT8_ENTRIES) * 2 + 255) / 256 * 256; }
  m.nnz0 = slot;
  m.nnz1 = slot + k.nnz_frame_bytes;
  m.nnz_frame_bytes = k.nnz_frame_bytes;
  m.max_mbw = k.max_mbw;
  const int tid = threadIdx.x;
  fill_shared_tables(&m.sh->tab, tid, AVR_THREADS_OF(Dir));
  __syncthreads();
  for (;;) {
    if (tid == 0) m.sh->job = atomicAdd(k.counter, 1u);
    __syncthreads();
    const uint32_t j = m.sh->job;
    __syncthreads();
    if (j >= k.n_jobs) break;
    const JobDesc job = k.jobs[j];
    WideMem w;
    w.ctl = nullptr;
    if (job.wide) {   // one role of a wide slice (avr_job.cuh, "Wide slices")
      uint8_t* ws = k.wide_ws + uint64_t(job.wide >> 8) * k.wide_stride;
      w.ctl = reinterpret_cast<WideCtl*>(ws);
      w.rows = reinterpret_cast<MbCtx*>(ws + wide_ctl_bytes());
      w.far = reinterpret_cast<uint32_t*>(ws + wide_ctl_bytes() + wide_rows_bytes(k.wide_mbw));
      w.far_cap = k.far_cap;
    }
    run_job<Dir>(k.a, job, m, job.wide ? &w : nullptr, job.wide ? (job.wide & 0xFF) - 1 : 0u, tid);
  }
}

enum : uint32_t { WIDE_FAR_CAP = 1u << 16 };   // words per ring of a wide slice (global memory)

__global__ void avr_fill_kernel(uint4* p, size_t n, uint32_t v) {
  size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
  size_t stride = size_t(gridDim.x) * blockDim.x;
  uint4 x = make_uint4(v, v, v, v);
  for (; i < n; i += stride) p[i] = x;
}

}  // namespace

struct avr_ctx {
  int device = 0;
  int sm_count = 0;
  int max_smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // device buffers, grown on demand
  uint8_t* d_in = nullptr; size_t d_in_cap = 0;
  uint8_t* d_out = nullptr; size_t d_out_cap = 0;
  uint8_t* d_work = nullptr; size_t d_work_cap = 0;
  uint8_t* d_wide = nullptr; size_t d_wide_cap = 0;   // workspace of the wide-slice kernel
  uint8_t* d_meta = nullptr; size_t d_meta_cap = 0;   // descs | ios | jobs | results | counter
  uint8_t* d_flush = nullptr; size_t flush_bytes = 0;
  uint16_t* d_priors = nullptr; uint32_t priors_digest = 0;   // avr_set_priors
  // pinned host buffers
  uint8_t* h_out = nullptr; size_t h_out_cap = 0;
  uint8_t* h_meta = nullptr; size_t h_meta_cap = 0;
  // staged batch
  int n_slices = 0, n_jobs = 0;
  bool staged = false, launched = false;
  int direction = 0;
  std::vector<avr_slice_desc> descs;
  std::vector<SliceIO> ios[2];       // per direction
  std::vector<JobDesc> jobs;
  std::vector<uint32_t> wide;        // slices with n_streams >= 2, longest first
  int grid_wide = 0;
  uint32_t max_mbw_wide = 1;
  uint64_t wide_stride = 0;
  size_t off_wide = 0;
  uint64_t out_total[2] = {0, 0};
  size_t in_bytes = 0;
  uint32_t max_mbw = 1, max_frame_bytes = 0;
  int grid[2] = {0, 0};                   // per direction (0 compress, 1 decompress)
  bool resident[2] = {true, true};
  int force_resident = -1;   // avr_config.resident / AVR_RESIDENT env: 0 / 1, -1 = per batch
  int wide_stream_bytes = 32768;   // avr_config.wide_stream_bytes / AVR_WIDE_STREAM_BYTES env
  int wide_batch_limit = -1;       // avr_config.wide_batch_limit (-1: 4 x SMs, 0: none)
  int wide_max_streams = 0;        // avr_config.wide_max_streams (0: default)
  // meta offsets
  size_t off_descs = 0, off_ios = 0, off_jobs = 0, off_res = 0, off_hist = 0, off_counter = 0;
  uint64_t work_stride = 0;
  int max_smem_sm = 0;
  float last_ms = 0.f;
  uint64_t launches = 0;
  int cfg_warps = 0;
};

namespace {

bool ck(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  fprintf(stderr, "avrecode_b200: CUDA error in %s: %s\n", what, cudaGetErrorString(e));
  return false;
}
#define CK(x) do { if (!ck((x), #x)) return AVR_ERR_CUDA; } while (0)

int grow_dev(uint8_t** p, size_t* cap, size_t need) {
  if (need <= *cap) return AVR_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  size_t want = need + need / 4 + 4096;
  if (!ck(cudaMalloc(reinterpret_cast<void**>(p), want), "cudaMalloc")) return AVR_ERR_NOMEM;
  *cap = want;
  return AVR_OK;
}
int grow_pinned(uint8_t** p, size_t* cap, size_t need) {
  if (need <= *cap) return AVR_OK;
  if (*p) cudaFreeHost(*p);
  *p = nullptr; *cap = 0;
  size_t want = need + need / 4 + 4096;
  if (!ck(cudaMallocHost(reinterpret_cast<void**>(p), want), "cudaMallocHost")) return AVR_ERR_NOMEM;
  *cap = want;
  return AVR_OK;
}
}  // namespace

extern "C" {

const char* avr_strerror(int code) {
  switch (code) {
    case AVR_OK: return "ok";
    case AVR_ERR_NO_DEVICE: return "no usable CUDA device (this library has no CPU fallback)";
    case AVR_ERR_CUDA: return "CUDA runtime error";
    case AVR_ERR_ARG: return "invalid argument";
    case AVR_ERR_NOMEM: return "out of memory";
    case AVR_ERR_FORMAT: return "malformed container or bitstream";
    case AVR_ERR_STATE: return "call sequence violated (stage -> launch -> sync -> collect)";
    case AVR_ERR_PRIORS: return "the container or slice asks for a prior table this context does not hold (avr_set_priors)";
    case AVR_ERR_LIMIT: return "batch beyond the fixed limits of avr_recode_noalloc (AVR_NOALLOC_MAX_SLICES slices, the caller's output capacity, the picture width): use avr_compress / avr_decompress";
    default: return "unknown error";
  }
}

int avr_abi_version(void) { return AVR_ABI_VERSION; }

void avr_config_init(avr_config* cfg) {
  if (!cfg) return;
  memset(cfg, 0, sizeof *cfg);
  cfg->abi_version = AVR_ABI_VERSION;
  cfg->wide_stream_bytes = -1; cfg->resident = -1; cfg->wide_batch_limit = -1; cfg->wide_max_streams = 0;
}

int avr_create(const avr_config* cfg, avr_ctx** out) {
  if (!out) return AVR_ERR_ARG;
  *out = nullptr;
  if (cfg && cfg->abi_version != AVR_ABI_VERSION) return AVR_ERR_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return AVR_ERR_NO_DEVICE; }
  int dev = cfg ? cfg->device : 0;
  if (dev < 0 || dev >= ndev) return AVR_ERR_ARG;
  avr_ctx* c = new (std::nothrow) avr_ctx();
  if (!c) return AVR_ERR_NOMEM;
  c->device = dev;
  c->cfg_warps = cfg ? cfg->blocks_per_sm : 0;
  if (cfg) {
    c->force_resident = cfg->resident < 0 ? -1 : (cfg->resident ? 1 : 0);
    if (cfg->wide_stream_bytes >= 0) c->wide_stream_bytes = cfg->wide_stream_bytes;
    c->wide_batch_limit = cfg->wide_batch_limit;
    c->wide_max_streams = cfg->wide_max_streams;
  }
  if (const char* e = getenv("AVR_WIDE_STREAM_BYTES")) c->wide_stream_bytes = int(strtoul(e, nullptr, 10));
  if (!ck(cudaSetDevice(dev), "cudaSetDevice")) { delete c; return AVR_ERR_NO_DEVICE; }
  cudaDeviceProp prop;
  if (!ck(cudaGetDeviceProperties(&prop, dev), "cudaGetDeviceProperties")) { delete c; return AVR_ERR_CUDA; }
  c->sm_count = prop.multiProcessorCount;
  c->max_smem_optin = int(prop.sharedMemPerBlockOptin);
  c->max_smem_sm = int(prop.sharedMemPerMultiprocessor);
  if (const char* e = getenv("AVR_RESIDENT")) c->force_resident = atoi(e) ? 1 : 0;
  if (!ck(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
      !ck(cudaEventCreate(&c->ev0), "cudaEventCreate") || !ck(cudaEventCreate(&c->ev1), "cudaEventCreate")) {
    avr_destroy(c);
    return AVR_ERR_CUDA;
  }
  *out = c;
  return AVR_OK;
}

void avr_destroy(avr_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->d_in) cudaFree(c->d_in);
  if (c->d_out) cudaFree(c->d_out);
  if (c->d_work) cudaFree(c->d_work);
  if (c->d_wide) cudaFree(c->d_wide);
  if (c->d_meta) cudaFree(c->d_meta);
  if (c->d_flush) cudaFree(c->d_flush);
  if (c->d_priors) cudaFree(c->d_priors);
  if (c->h_out) cudaFreeHost(c->h_out);
  if (c->h_meta) cudaFreeHost(c->h_meta);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

// FNV-1a over the table bytes; never 0 (0 = "no table")
uint32_t avr_priors_digest_of(const uint16_t* table, uint32_t n_entries) {
  if (!table) return 0;
  uint32_t h = 2166136261u;
  for (uint32_t i = 0; i < n_entries; i++) { h = (h ^ (table[i] & 0xFF)) * 16777619u; h = (h ^ (table[i] >> 8)) * 16777619u; }
  return h ? h : 1u;
}
int avr_set_priors(avr_ctx* c, const uint16_t* table, uint32_t n_entries) {
  if (!c) return AVR_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  if (!table) {
    if (c->d_priors) cudaFree(c->d_priors);
    c->d_priors = nullptr; c->priors_digest = 0;
    return AVR_OK;
  }
  if (n_entries != AVR_PRIOR_ENTRIES) return AVR_ERR_ARG;
  static_assert(AVR_PRIOR_ENTRIES == MODEL_COUNTERS, "the header's table size is the model's");
  for (uint32_t i = 0; i < n_entries; i++) {   // what recode.cpp:1030-1047 can reach (the coders' reciprocal table ends at 99)
    const uint32_t pos = table[i] & 0xFF, tot = table[i] >> 8;
    if (pos < 1 || tot <= pos || tot > LIMIT_SIG) return AVR_ERR_ARG;
  }
  if (!c->d_priors && cudaMalloc(reinterpret_cast<void**>(&c->d_priors), size_t(MODEL_COUNTERS) * 2) != cudaSuccess) { cudaGetLastError(); c->d_priors = nullptr; return AVR_ERR_NOMEM; }
  CK(cudaMemcpy(c->d_priors, table, size_t(MODEL_COUNTERS) * 2, cudaMemcpyHostToDevice));
  c->priors_digest = avr_priors_digest_of(table, n_entries);
  return AVR_OK;
}
uint32_t avr_priors_digest(avr_ctx* c) { return c ? c->priors_digest : 0; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Staged pipeline: stage (H2D) -> launch (kernels) -> sync -> collect (D2H).
// ------------------------------------------------------------------------------------------------
extern "C" {

int avr_stage(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes) {
  if (!c || n < 0 || (n > 0 && (!descs || !in))) return AVR_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));   // previous batch may still be reading the pinned metadata
  c->staged = false; c->launched = false;
  c->n_slices = n; c->in_bytes = size_t(in_bytes);
  c->descs.assign(descs, descs + n);
  c->jobs.clear();
  c->wide.clear();
  c->max_mbw = 1; c->max_frame_bytes = 0; c->max_mbw_wide = 1;
  for (int dir = 0; dir < 2; dir++) { c->ios[dir].resize(n); c->out_total[dir] = 0; }
  for (int i = 0; i < n; i++) {
    const avr_slice_desc& d = descs[i];
    const uint32_t n_str = d.n_streams & AVR_STREAMS_MASK;
    if (d.in_off + d.in_size > in_bytes) return AVR_ERR_ARG;
    if ((d.n_streams & AVR_SLICE_PRIORS) && !c->d_priors) return AVR_ERR_PRIORS;   // the slice asks for a table this context does not have
    for (int dir = 0; dir < 2; dir++) {
      SliceIO& io = c->ios[dir][i];
      io.in_off = d.in_off; io.in_size = d.in_size;
      io.out_off = c->out_total[dir];
      io.out_cap = dir == 0 ? d.in_size + d.in_size / 8 + 256 : d.payload_size + 64;
      // wide slice: header + n_streams slots, any one of which may end up with (nearly) all of the slice
      if (dir == 0 && n_str >= 2) io.out_cap = 64 + n_str * uint32_t(align_up(size_t(d.in_size) + d.in_size / 8 + 512, 16));
      c->out_total[dir] += align_up(io.out_cap, 16);
    }
    if (n_str >= 2) {
      // a wide slice is a chunk of its own (its K streams have K models of their own)
      const bool own_model = (i == 0 || (d.flags & AVR_SLICE_NEW_MODEL)) && (i + 1 == n || (descs[i + 1].flags & AVR_SLICE_NEW_MODEL));
      if (n_str > AVR_MAX_STREAMS || !own_model || (d.flags & AVR_SLICE_SKIP)) return AVR_ERR_ARG;
      c->wide.push_back(uint32_t(i));
      if (d.mb_width <= 1024) c->max_mbw_wide = std::max<uint32_t>(c->max_mbw_wide, d.mb_width);
      continue;
    }
    // pictures beyond Level 6.2 are reported unsupported by the device (run_job) and size nothing here
    const bool sane = uint64_t(d.mb_width) * d.mb_height <= AVR_MAX_PICTURE_MBS && d.mb_width <= 1024;
    if (sane) {
      c->max_mbw = std::max<uint32_t>(c->max_mbw, d.mb_width);
      c->max_frame_bytes = std::max<uint32_t>(c->max_frame_bytes, uint32_t(d.mb_width) * d.mb_height * NNZ_STRIDE);
    }
    if (i == 0 || (d.flags & AVR_SLICE_NEW_MODEL)) {
      JobDesc j; memset(&j, 0, sizeof j);
      j.first_slice = uint32_t(i); j.n_slices = 0;
      c->jobs.push_back(j);
    }
    c->jobs.back().n_slices++;
    c->jobs.back().weight += d.in_size > 0x7FFFFFFFu - c->jobs.back().weight ? 0 : d.in_size;
  }
  // longest jobs first (LPT order for the dynamic hand-out)
  std::stable_sort(c->jobs.begin(), c->jobs.end(), [](const JobDesc& a, const JobDesc& b) { return a.weight > b.weight; });
  // Wide slices go to the FRONT of the queue, longest first, each as K + 1 consecutive role jobs with role 0 -- the one
  // the others feed or are fed by -- first (see avr_job.cuh for why that order cannot deadlock).
  std::stable_sort(c->wide.begin(), c->wide.end(), [&](uint32_t a, uint32_t b) { return descs[a].in_size > descs[b].in_size; });
  uint32_t max_roles = 0;
  if (!c->wide.empty()) {
    std::vector<JobDesc> front;
    for (size_t slot = 0; slot < c->wide.size(); slot++) {
      const uint32_t i = c->wide[slot];
      const uint32_t K = descs[i].n_streams & AVR_STREAMS_MASK;
      max_roles = std::max(max_roles, K + 1);
      for (uint32_t role = 0; role <= K; role++) {
        JobDesc j; memset(&j, 0, sizeof j);
        j.first_slice = i; j.n_slices = 1; j.weight = descs[i].in_size / (role == 0 ? 1 : K);
        j.wide = (role + 1) | uint32_t(slot) << 8;
        front.push_back(j);
      }
    }
    c->jobs.insert(c->jobs.begin(), front.begin(), front.end());
  }
  c->n_jobs = int(c->jobs.size());
  c->grid_wide = int(c->wide.size());
  c->wide_stride = align_up(wide_ctl_bytes() + wide_rows_bytes(c->max_mbw_wide) + size_t(WIDE_MAX) * WIDE_FAR_CAP * 4, 256);
  c->max_mbw = std::max(c->max_mbw, c->max_mbw_wide);   // a role works in an ordinary block: its row and window
  c->max_frame_bytes = uint32_t(align_up(c->max_frame_bytes, 16));

  // Placement of the estimators: all of them in shared memory (2 blocks per SM), or split -- the 8x8 significance
  // slab (70 % of the table) in the global workspace behind L1 / L2, the rest in shared memory, 4 blocks per SM.
  const size_t smem_res = smem_bytes(c->max_mbw, true), smem_glb = smem_bytes(c->max_mbw, false);
  const size_t smem_sm = size_t(c->max_smem_sm);
  int res_per_sm = smem_res + 1024 <= size_t(c->max_smem_optin) ? int(smem_sm / (smem_res + 1024)) : 0;
  res_per_sm = std::min(res_per_sm, 2);
  int glb_per_sm = std::max(1, std::min(c->cfg_warps > 0 ? c->cfg_warps : 4, int(smem_sm / (smem_glb + 1024))));
  glb_per_sm = std::min(glb_per_sm, 4);
  if (smem_glb + 1024 > size_t(c->max_smem_optin)) return AVR_ERR_ARG;   // picture too wide for the neighbour row
  // Measured with every SM full (profiles/r2p_throughput_placement.txt, 1728 slices of 32 corpus clips; compress /
  // decompress): resident 771 / 985 ms, split 825 / 871 ms.  Twice the blocks make a block 2.14x / 1.77x slower: a loss
  // for compression (three warps per block), a 13 % gain for decompression (two) -- when the batch has the slices to
  // fill them for the whole launch.  So per direction: compression stays resident whenever that fits, decompression goes
  // split when the batch is throughput-bound.  (The factors are averages over a launch with full SMs: the longest jobs
  // start first and finish among fewer and fewer neighbours, so "1.77 x the longest job" overestimates the split launch.)
  // The split placement is also the one for pictures whose neighbour row leaves the estimators no room, and
  // avr_config.resident forces either.
  uint64_t total_w = 0, longest_w = 0;
  for (const JobDesc& j : c->jobs) { total_w += j.weight; longest_w = std::max<uint64_t>(longest_w, j.weight); }
  const bool fits = res_per_sm > 0;
  for (int dir = 0; dir < 2; dir++) {
    bool resident = fits;
    if (fits && dir == 1 && c->wide.empty()) {
      // throughput-bound with 2 blocks per SM (the slots stay full for longer than the longest job lasts)?  Then 4 slower
      // blocks per SM finish sooner; a latency-bound batch -- a file with a few long slices -- stays resident
      resident = total_w / uint64_t(res_per_sm * c->sm_count) < longest_w;
    }
    if (c->force_resident == 0) resident = false;
    if (c->force_resident == 1 && fits) resident = true;
    c->resident[dir] = resident;
    c->grid[dir] = std::max(1, std::min(c->n_jobs, c->sm_count * (resident ? res_per_sm : glb_per_sm)));
    if (uint32_t(c->grid[dir]) < max_roles) return AVR_ERR_ARG;   // (a device with fewer than 9 resident blocks: not a B200)
  }
  const bool all_resident = c->resident[0] && c->resident[1];
  const int max_grid = std::max(c->grid[0], c->grid[1]);

  // device memory
  // the nnz planes are only needed by jobs whose model spans several slices (file mode)
  bool any_multi = false;
  for (const JobDesc& j : c->jobs) any_multi = any_multi || j.n_slices > 1;
  if (!any_multi) c->max_frame_bytes = 0;
  const uint64_t work_stride = align_up((all_resident ? 0 : counters_bytes()) + 2 * size_t(c->max_frame_bytes) + 16, 256);
  c->work_stride = work_stride;
  int rc;
  if ((rc = grow_dev(&c->d_in, &c->d_in_cap, size_t(in_bytes) + 64))) return rc;
  if ((rc = grow_dev(&c->d_out, &c->d_out_cap, size_t(std::max(c->out_total[0], c->out_total[1])) + 64))) return rc;
  if ((rc = grow_dev(&c->d_work, &c->d_work_cap, size_t(work_stride) * max_grid + 256))) return rc;
  if (c->grid_wide && (rc = grow_dev(&c->d_wide, &c->d_wide_cap, size_t(c->wide_stride) * c->grid_wide + 256))) return rc;
  c->off_descs = 0;
  c->off_ios = align_up(c->off_descs + sizeof(avr_slice_desc) * n, 256);
  c->off_jobs = align_up(c->off_ios + 2 * sizeof(SliceIO) * n, 256);
  c->off_wide = align_up(c->off_jobs + sizeof(JobDesc) * c->jobs.size(), 256);
  c->off_res = align_up(c->off_wide + sizeof(uint32_t) * c->wide.size(), 256);
  c->off_hist = align_up(c->off_res + sizeof(avr_slice_result) * n, 256);
  c->off_counter = align_up(c->off_hist + 8 * sizeof(uint32_t) * n, 256);
  size_t meta_bytes = c->off_counter + 256;
  if ((rc = grow_dev(&c->d_meta, &c->d_meta_cap, meta_bytes))) return rc;
  if ((rc = grow_pinned(&c->h_meta, &c->h_meta_cap, meta_bytes))) return rc;
  if (n > 0) {
    memcpy(c->h_meta + c->off_descs, descs, sizeof(avr_slice_desc) * n);
    memcpy(c->h_meta + c->off_ios, c->ios[0].data(), sizeof(SliceIO) * n);
    memcpy(c->h_meta + c->off_ios + sizeof(SliceIO) * n, c->ios[1].data(), sizeof(SliceIO) * n);
    memcpy(c->h_meta + c->off_jobs, c->jobs.data(), sizeof(JobDesc) * c->jobs.size());
    memcpy(c->h_meta + c->off_wide, c->wide.data(), sizeof(uint32_t) * c->wide.size());
    CK(cudaMemcpyAsync(c->d_meta, c->h_meta, c->off_res, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_in, in, size_t(in_bytes), cudaMemcpyHostToDevice, c->stream));
  }
  c->staged = true;
  return AVR_OK;
}

int avr_launch(avr_ctx* c, int direction) {
  if (!c || (direction != 0 && direction != 1)) return AVR_ERR_ARG;
  if (!c->staged) return AVR_ERR_STATE;
  CK(cudaSetDevice(c->device));
  c->direction = direction;
  c->launched = true;
  if (c->n_slices == 0) return AVR_OK;
  const int n = c->n_slices;
  KernelArgs k;
  k.a.descs = reinterpret_cast<const avr_slice_desc*>(c->d_meta + c->off_descs);
  k.a.ios = reinterpret_cast<const SliceIO*>(c->d_meta + c->off_ios) + (direction ? n : 0);
  k.a.in_arena = c->d_in;
  k.a.out_arena = c->d_out;
  k.a.results = reinterpret_cast<avr_slice_result*>(c->d_meta + c->off_res);
  k.a.hist = reinterpret_cast<uint32_t*>(c->d_meta + c->off_hist);
  k.a.priors = c->d_priors;
  k.jobs = reinterpret_cast<const JobDesc*>(c->d_meta + c->off_jobs);
  k.n_jobs = uint32_t(c->n_jobs);
  k.counter = reinterpret_cast<uint32_t*>(c->d_meta + c->off_counter);
  k.workspace = c->d_work;
  k.work_stride = c->work_stride;
  k.nnz_frame_bytes = c->max_frame_bytes;
  k.max_mbw = c->max_mbw;
  const size_t smem = smem_bytes(c->max_mbw, c->resident[direction]);
  if (int(smem) > c->max_smem_optin) return AVR_ERR_ARG;
  CK(cudaMemsetAsync(k.counter, 0, 2 * sizeof(uint32_t), c->stream));
  k.wide_ws = c->d_wide;
  k.wide_stride = c->wide_stride;
  k.wide_mbw = c->max_mbw_wide;
  k.far_cap = WIDE_FAR_CAP;
  // the control block of every wide slice starts out zero (ring indices, wavefront progress, role count), and so do
  // its neighbour rows: a macroblock of the slice's first row, or left of its first macroblock in the second, reads
  // "nothing there" (all zero) from the row above, exactly as an ordinary slice does from its freshly cleared row
  if (c->grid_wide)
    CK(cudaMemset2DAsync(c->d_wide, size_t(c->wide_stride), 0, wide_ctl_bytes() + wide_rows_bytes(c->max_mbw_wide), size_t(c->grid_wide), c->stream));
  cudaLaunchConfig_t lc;
  memset(&lc, 0, sizeof lc);
  lc.gridDim = dim3(unsigned(c->grid[direction])); lc.blockDim = dim3(direction == 0 ? AVR_THREADS_OF(Compress) : AVR_THREADS_OF(Decompress)); lc.dynamicSmemBytes = smem; lc.stream = c->stream;
  lc.attrs = nullptr; lc.numAttrs = 0;
  CK(cudaEventRecord(c->ev0, c->stream));
#define AVR_LAUNCH(DIR, RES)                                                                                              \
  do {                                                                                                                    \
    CK(cudaFuncSetAttribute(avr_job_kernel<DIR, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));          \
    CK(cudaLaunchKernelEx(&lc, avr_job_kernel<DIR, RES>, k));                                                             \
  } while (0)
  if (c->n_jobs > 0) {
    if (direction == 0) { if (c->resident[0]) AVR_LAUNCH(Compress, true); else AVR_LAUNCH(Compress, false); }
    else { if (c->resident[1]) AVR_LAUNCH(Decompress, true); else AVR_LAUNCH(Decompress, false); }
    c->launches++;
  }
#undef AVR_LAUNCH
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev1, c->stream));
  return AVR_OK;
}

int avr_sync(avr_ctx* c) {
  if (!c) return AVR_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  if (c->launched && c->n_slices > 0) CK(cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1));
  return AVR_OK;
}

int avr_collect(avr_ctx* c, avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes) {
  if (!c || !out || !out_bytes) return AVR_ERR_ARG;
  if (!c->launched) return AVR_ERR_STATE;
  if (!results && c->n_slices > 0) return AVR_ERR_ARG;   // an empty batch has no results to return
  CK(cudaSetDevice(c->device));
  const int n = c->n_slices;
  const size_t total = size_t(c->out_total[c->direction]);
  int rc;
  if ((rc = grow_pinned(&c->h_out, &c->h_out_cap, total + 64))) return rc;
  if (n > 0) {
    CK(cudaMemcpyAsync(c->h_meta + c->off_res, c->d_meta + c->off_res, sizeof(avr_slice_result) * n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(c->h_out, c->d_out, total, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    memcpy(results, c->h_meta + c->off_res, sizeof(avr_slice_result) * n);
  }
  *out = c->h_out;
  *out_bytes = total;
  return AVR_OK;
}

// what the host file layer (avr_file.cpp) needs to know about the context
int avr_ctx_wide_stream_bytes(avr_ctx* c) { return c ? c->wide_stream_bytes : 32768; }
int avr_ctx_wide_max_streams(avr_ctx* c) { return !c || c->wide_max_streams < 2 ? AVR_DEFAULT_WIDE_STREAMS : c->wide_max_streams > AVR_MAX_STREAMS ? AVR_MAX_STREAMS : c->wide_max_streams; }
int avr_ctx_wide_batch_limit(avr_ctx* c) { return !c ? 0 : c->wide_batch_limit < 0 ? 4 * c->sm_count : c->wide_batch_limit; }

int avr_collect_hist(avr_ctx* c, uint32_t* hist, int n_slices) {
  if (!c || !hist || n_slices != c->n_slices) return AVR_ERR_ARG;
  if (!c->launched) return AVR_ERR_STATE;
  if (n_slices == 0) return AVR_OK;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaMemcpy(hist, c->d_meta + c->off_hist, size_t(8) * sizeof(uint32_t) * n_slices, cudaMemcpyDeviceToHost));
  return AVR_OK;
}

int avr_last_kernel_ms(avr_ctx* c, float* ms) {
  if (!c || !ms) return AVR_ERR_ARG;
  *ms = c->last_ms;
  return AVR_OK;
}

int avr_launch_count(avr_ctx* c, uint64_t* n) {
  if (!c || !n) return AVR_ERR_ARG;
  *n = c->launches;
  return AVR_OK;
}

int avr_flush_l2(avr_ctx* c) {
  if (!c) return AVR_ERR_ARG;
  CK(cudaSetDevice(c->device));
  if (!c->d_flush) {
    c->flush_bytes = size_t(256) << 20;   // > 126 MB L2
    if (!ck(cudaMalloc(reinterpret_cast<void**>(&c->d_flush), c->flush_bytes), "cudaMalloc flush")) return AVR_ERR_NOMEM;
  }
  avr_fill_kernel<<<c->sm_count * 8, 256, 0, c->stream>>>(reinterpret_cast<uint4*>(c->d_flush), c->flush_bytes / 16, uint32_t(c->launches));
  CK(cudaGetLastError());
  return AVR_OK;
}

static int run_oneshot(avr_ctx* c, int direction, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes,
                       avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes) {
  int rc;
  if ((rc = avr_stage(c, descs, n, in, in_bytes))) return rc;
  if ((rc = avr_launch(c, direction))) return rc;
  if ((rc = avr_sync(c))) return rc;
  return avr_collect(c, results, out, out_bytes);
}

int avr_compress(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes,
                 avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes) {
  return run_oneshot(c, 0, descs, n, in, in_bytes, results, out, out_bytes);
}

int avr_decompress(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes,
                   avr_slice_result* results, const uint8_t** out, uint64_t* out_bytes) {
  return run_oneshot(c, 1, descs, n, in, in_bytes, results, out, out_bytes);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// The entry point that allocates nothing (include/avrecode_b200.h: avr_recode_noalloc).  No counterpart in the
// reference; added for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).
//
// It launches the same persistent job kernel as avr_launch (estimators resident in shared memory, every slice a job of
// its own), but everything that kernel needs besides the caller's input and output -- the descriptor / buffer / job
// tables, the job counter and the per-slice results -- lives in ONE file-scope device array instead of cudaMalloc'ed
// buffers, and the host-side staging copy of those tables in a file-scope host array.  Every GPU that loads the
// library holds AVR_NOALLOC_SCRATCH_BYTES of device memory for it until the process exits.
//
// The array is shared by every call and keeps the last call's contents: the job counter and the results are cleared
// at the start of each call, the tables are overwritten.  Two calls that overlapped on one GPU would race on it, so
// the library does not let them: a mutex is held from the first copy until the results are back on the host.
// ------------------------------------------------------------------------------------------------
#include <mutex>

namespace {

static __device__ __align__(256) uint8_t g_noalloc_scratch[AVR_NOALLOC_SCRATCH_BYTES];
static uint8_t g_noalloc_host[AVR_NOALLOC_SCRATCH_BYTES];   // host-side staging of the tables (no allocation per call either)
static std::mutex g_noalloc_mutex;

// layout of the scratch: descs | ios | jobs | results | counter
constexpr size_t NA_OFF_DESCS = 0;
constexpr size_t NA_OFF_IOS = NA_OFF_DESCS + sizeof(avr_slice_desc) * AVR_NOALLOC_MAX_SLICES;
constexpr size_t NA_OFF_JOBS = NA_OFF_IOS + sizeof(SliceIO) * AVR_NOALLOC_MAX_SLICES;
constexpr size_t NA_OFF_RES = NA_OFF_JOBS + sizeof(JobDesc) * AVR_NOALLOC_MAX_SLICES;
constexpr size_t NA_OFF_COUNTER = NA_OFF_RES + sizeof(avr_slice_result) * AVR_NOALLOC_MAX_SLICES;
static_assert(NA_OFF_IOS % 16 == 0 && NA_OFF_JOBS % 16 == 0 && NA_OFF_RES % 16 == 0 && NA_OFF_COUNTER % 16 == 0, "table alignment");
static_assert(NA_OFF_COUNTER + 256 <= AVR_NOALLOC_SCRATCH_BYTES, "the tables of AVR_NOALLOC_MAX_SLICES slices fit the scratch");

struct NoAllocDevice {       // per device, filled in once with that device current
  bool ready = false;
  uint8_t* scratch = nullptr;   // DEVICE address of g_noalloc_scratch on this device (the name itself is a host-side stand-in)
  int sm_count = 0, max_smem_optin = 0, max_smem_sm = 0;
};
enum : int { NA_MAX_DEVICES = 64 };
NoAllocDevice g_noalloc_dev[NA_MAX_DEVICES];

int noalloc_setup(int device, NoAllocDevice** out) {
  if (device < 0 || device >= NA_MAX_DEVICES) return AVR_ERR_ARG;
  NoAllocDevice& d = g_noalloc_dev[device];
  if (!d.ready) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return AVR_ERR_CUDA;
    CK(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
    CK(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    CK(cudaDeviceGetAttribute(&d.max_smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device));
    d.scratch = static_cast<uint8_t*>(p);
    d.ready = true;
  }
  *out = &d;
  return AVR_OK;
}

}  // namespace

extern "C" int avr_recode_noalloc(int device, int direction, const avr_slice_desc* descs, int n_slices, const uint8_t* d_in,
                                  uint64_t in_bytes, uint8_t* d_out, uint64_t out_cap, uint64_t* out_need,
                                  avr_slice_result* results, void* stream) {
  if (out_need) *out_need = 0;
  if ((direction != 0 && direction != 1) || n_slices < 0 || (n_slices > 0 && !descs)) return AVR_ERR_ARG;
  if (n_slices > AVR_NOALLOC_MAX_SLICES) return AVR_ERR_LIMIT;   // beyond the static scratch: use avr_compress / avr_decompress
  const int n = n_slices;
  // what the batch needs of the caller's output buffer: the slots of avr_stage, in order
  uint64_t need = 0;
  uint32_t max_mbw = 1;
  for (int i = 0; i < n; i++) {
    const avr_slice_desc& d = descs[i];
    // every slice a model of its own, no wide slices, no bookkeeping-only slices: the static scratch holds no
    // neighbour-count planes and no wide-slice rings
    if (i > 0 && !(d.flags & AVR_SLICE_NEW_MODEL)) return AVR_ERR_ARG;
    if ((d.n_streams & AVR_STREAMS_MASK) >= 2 || (d.n_streams & AVR_SLICE_PRIORS) || (d.flags & AVR_SLICE_SKIP)) return AVR_ERR_ARG;
    if (d.in_off + d.in_size > in_bytes) return AVR_ERR_ARG;
    const uint32_t cap = direction == 0 ? d.in_size + d.in_size / 8 + 256 : d.payload_size + 64;
    need += align_up(cap, 16);
    if (uint64_t(d.mb_width) * d.mb_height <= AVR_MAX_PICTURE_MBS && d.mb_width <= 1024) max_mbw = std::max<uint32_t>(max_mbw, d.mb_width);
  }
  if (out_need) *out_need = need;
  if (!d_out) return n == 0 || out_need ? AVR_OK : AVR_ERR_ARG;   // a query: how many bytes of output the batch needs
  if (n == 0) return AVR_OK;
  if (!d_in || !results) return AVR_ERR_ARG;
  if (out_cap < need) return AVR_ERR_LIMIT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return AVR_ERR_NO_DEVICE; }
  if (device < 0 || device >= ndev) return AVR_ERR_ARG;

  std::lock_guard<std::mutex> lock(g_noalloc_mutex);   // the scratch is shared: overlapping calls would race on it
  CK(cudaSetDevice(device));
  NoAllocDevice* dev = nullptr;
  int rc = noalloc_setup(device, &dev);
  if (rc) return rc;
  const size_t smem = smem_bytes(max_mbw, true);
  if (smem + 1024 > size_t(dev->max_smem_optin)) return AVR_ERR_LIMIT;   // picture too wide for the resident variant
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  // the tables, built in the static host array
  avr_slice_desc* h_descs = reinterpret_cast<avr_slice_desc*>(g_noalloc_host + NA_OFF_DESCS);
  SliceIO* h_ios = reinterpret_cast<SliceIO*>(g_noalloc_host + NA_OFF_IOS);
  JobDesc* h_jobs = reinterpret_cast<JobDesc*>(g_noalloc_host + NA_OFF_JOBS);
  uint64_t off = 0;
  for (int i = 0; i < n; i++) {
    const avr_slice_desc& d = descs[i];
    h_descs[i] = d;
    h_ios[i].in_off = d.in_off; h_ios[i].in_size = d.in_size;
    h_ios[i].out_off = off;
    h_ios[i].out_cap = direction == 0 ? d.in_size + d.in_size / 8 + 256 : d.payload_size + 64;
    off += align_up(h_ios[i].out_cap, 16);
    memset(&h_jobs[i], 0, sizeof(JobDesc));
    h_jobs[i].first_slice = uint32_t(i); h_jobs[i].n_slices = 1; h_jobs[i].weight = d.in_size;
  }
  // longest first; std::sort works in place (no temporary buffer), the tie-break keeps the order deterministic
  std::sort(h_jobs, h_jobs + n, [](const JobDesc& a, const JobDesc& b) { return a.weight != b.weight ? a.weight > b.weight : a.first_slice < b.first_slice; });

  uint8_t* s = dev->scratch;
  CK(cudaMemcpyAsync(s + NA_OFF_DESCS, h_descs, sizeof(avr_slice_desc) * n, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s + NA_OFF_IOS, h_ios, sizeof(SliceIO) * n, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s + NA_OFF_JOBS, h_jobs, sizeof(JobDesc) * n, cudaMemcpyHostToDevice, st));
  // what must not survive from the last call: the job counter (zero), the results (0xFF: "not written")
  CK(cudaMemsetAsync(s + NA_OFF_COUNTER, 0, 256, st));
  CK(cudaMemsetAsync(s + NA_OFF_RES, 0xFF, sizeof(avr_slice_result) * n, st));

  KernelArgs k;
  memset(&k, 0, sizeof k);
  k.a.descs = reinterpret_cast<const avr_slice_desc*>(s + NA_OFF_DESCS);
  k.a.ios = reinterpret_cast<const SliceIO*>(s + NA_OFF_IOS);
  k.a.in_arena = d_in;
  k.a.out_arena = d_out;
  k.a.results = reinterpret_cast<avr_slice_result*>(s + NA_OFF_RES);
  k.a.hist = nullptr;
  k.jobs = reinterpret_cast<const JobDesc*>(s + NA_OFF_JOBS);
  k.n_jobs = uint32_t(n);
  k.counter = reinterpret_cast<uint32_t*>(s + NA_OFF_COUNTER);
  k.workspace = s;            // jobs of one slice keep their neighbour counts in shared memory: never dereferenced
  k.work_stride = 0;
  k.nnz_frame_bytes = 0;
  k.max_mbw = max_mbw;
  k.wide_ws = nullptr; k.wide_stride = 0; k.wide_mbw = 1; k.far_cap = WIDE_FAR_CAP;
  const int per_sm = std::max(1, std::min(2, int(size_t(dev->max_smem_sm) / (smem + 1024))));
  cudaLaunchConfig_t lc;
  memset(&lc, 0, sizeof lc);
  lc.gridDim = dim3(unsigned(std::max(1, std::min(n, dev->sm_count * per_sm))));
  lc.blockDim = dim3(direction == 0 ? AVR_THREADS_OF(Compress) : AVR_THREADS_OF(Decompress));
  lc.dynamicSmemBytes = smem; lc.stream = st;
  if (direction == 0) {
    CK(cudaFuncSetAttribute(avr_job_kernel<Compress, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CK(cudaLaunchKernelEx(&lc, avr_job_kernel<Compress, true>, k));
  } else {
    CK(cudaFuncSetAttribute(avr_job_kernel<Decompress, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CK(cudaLaunchKernelEx(&lc, avr_job_kernel<Decompress, true>, k));
  }
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(results, s + NA_OFF_RES, sizeof(avr_slice_result) * n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return AVR_OK;
}
