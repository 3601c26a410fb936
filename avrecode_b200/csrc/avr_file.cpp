// avr_file.cpp -- host side of `recode compress | decompress`: recode.proto container codec and the
// drivers that turn a file into slice batches for the device and back.
//
// Mirrors, on the host, compressor::run / find_next_coded_block_and_emit_literal
// (recode.cpp:1115-1125, 1275-1297) and decompressor::run / read_packet / recognize_coded_block
// (recode.cpp:1338-1409, 1546-1573).  The container is the reference's Recoded message
// (recode.proto:1-20) written in field-number order like libprotobuf does, so AVR_MODE_FILE output
// is byte-identical to the reference's.  libprotobuf is not used (it is not on the box).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <stdio.h>
#include <string>
#include <thread>
#include <new>
#include <vector>
#include "../../include/avrecode_b200.h"

namespace {

// AVR_FILE_TIMING=1: where a file call spends its time, to stderr (debugging aid)
struct Stopwatch {
  bool on = getenv("AVR_FILE_TIMING") != nullptr;
  std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
  void lap(const char* what) {
    if (!on) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "avr_file: %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

const char kSliceModeTag[] = "avr-b200/slice";

void put_varint(std::string& o, uint64_t v) {
  while (v >= 0x80) { o.push_back(char((v & 0x7F) | 0x80)); v >>= 7; }
  o.push_back(char(v));
}
void put_len(std::string& o, int field, const uint8_t* p, size_t n) {
  put_varint(o, (uint64_t(field) << 3) | 2);
  put_varint(o, n);
  o.append(reinterpret_cast<const char*>(p), n);
}
void put_var(std::string& o, int field, uint64_t v) { put_varint(o, (uint64_t(field) << 3) | 0); put_varint(o, v); }

bool get_varint(const uint8_t*& p, const uint8_t* e, uint64_t* v) {
  uint64_t r = 0;
  for (int sh = 0; p < e && sh < 64; sh += 7) {
    uint8_t b = *p++;
    r |= uint64_t(b & 0x7F) << sh;
    if (!(b & 0x80)) { *v = r; return true; }
  }
  return false;
}

struct Block {   // Recoded.Block (recode.proto:9-18)
  bool has_size = false, has_literal = false, has_skip = false, has_cabac = false, has_parity = false, has_last = false;
  int64_t size = 0;
  uint64_t streams = 0;   // field 7, chunked mode only: range-coder streams of a wide slice (avr_slice_desc.n_streams)
  uint64_t escaped = 0;   // field 8, chunked mode only: the payload in the file has emulation-prevention bytes, cabac
                          //   codes the unescaped slice_data (AVR_SLICE_ESCAPED)
  bool skip = false, parity = false;
  const uint8_t* literal = nullptr; size_t literal_n = 0;
  const uint8_t* cabac = nullptr; size_t cabac_n = 0;
  const uint8_t* last = nullptr; size_t last_n = 0;
};

bool parse_block(const uint8_t* p, const uint8_t* e, Block* b) {
  while (p < e) {
    uint64_t key;
    if (!get_varint(p, e, &key)) return false;
    int field = int(key >> 3), wt = int(key & 7);
    if (wt == 0) {
      uint64_t v;
      if (!get_varint(p, e, &v)) return false;
      if (field == 1) { b->size = int64_t(v); b->has_size = true; }
      else if (field == 3) { b->skip = v != 0; b->has_skip = true; }
      else if (field == 5) { b->parity = v != 0; b->has_parity = true; }
      else if (field == 7) b->streams = v;
      else if (field == 8) b->escaped = v;
    } else if (wt == 2) {
      uint64_t n;
      if (!get_varint(p, e, &n) || n > uint64_t(e - p)) return false;
      if (field == 2) { b->literal = p; b->literal_n = size_t(n); b->has_literal = true; }
      else if (field == 4) { b->cabac = p; b->cabac_n = size_t(n); b->has_cabac = true; }
      else if (field == 6) { b->last = p; b->last_n = size_t(n); b->has_last = true; }
      p += n;
    } else if (wt == 1) { if (e - p < 8) return false; p += 8; }
    else if (wt == 5) { if (e - p < 4) return false; p += 4; }
    else return false;
  }
  return true;
}

bool parse_container(const uint8_t* p, size_t n, std::vector<Block>* blocks, std::string* version, uint64_t* priors) {
  const uint8_t* e = p + n;
  while (p < e) {
    uint64_t key;
    if (!get_varint(p, e, &key)) return false;
    int field = int(key >> 3), wt = int(key & 7);
    if (wt == 2) {
      uint64_t len;
      if (!get_varint(p, e, &len) || len > uint64_t(e - p)) return false;
      if (field == 2) { Block b; if (!parse_block(p, p + len, &b)) return false; blocks->push_back(b); }
      else if (field == 1) {   // Metadata: version = field 1
        const uint8_t* q = p; const uint8_t* qe = p + len;
        while (q < qe) {
          uint64_t k2;
          if (!get_varint(q, qe, &k2)) return false;
          if ((k2 & 7) == 2) {
            uint64_t l2;
            if (!get_varint(q, qe, &l2) || l2 > uint64_t(qe - q)) return false;
            if ((k2 >> 3) == 1) version->assign(reinterpret_cast<const char*>(q), size_t(l2));
            q += l2;
          } else if ((k2 & 7) == 0) { uint64_t v; if (!get_varint(q, qe, &v)) return false; if ((k2 >> 3) == 2) *priors = v; }   // Metadata field 2: digest of the prior table
          else return false;
        }
      }
      p += len;
    } else if (wt == 0) { uint64_t v; if (!get_varint(p, e, &v)) return false; }
    else if (wt == 1) { if (e - p < 8) return false; p += 8; }
    else if (wt == 5) { if (e - p < 4) return false; p += 4; }
    else return false;
  }
  return true;
}

// Chunked mode: a slice whose payload is at least two of these becomes a wide slice, one stream per this many
// bytes (avr_config.wide_stream_bytes; 0 = never).  Long slices set the time of a small batch -- see avr_job.cuh;
// a batch with more slices than avr_config.wide_batch_limit is a throughput job and gets none.
extern "C" int avr_ctx_wide_stream_bytes(avr_ctx* c);
extern "C" int avr_ctx_wide_batch_limit(avr_ctx* c);
extern "C" int avr_ctx_wide_max_streams(avr_ctx* c);
uint32_t wide_stream_bytes(avr_ctx* ctx, size_t n_batch_slices) {
  const int limit = ctx ? avr_ctx_wide_batch_limit(ctx) : 0;
  if (limit > 0 && n_batch_slices > size_t(limit)) return 0;
  if (ctx) return uint32_t(avr_ctx_wide_stream_bytes(ctx));
  if (const char* e = getenv("AVR_WIDE_STREAM_BYTES")) return uint32_t(strtoul(e, nullptr, 10));
  return 32768;
}
// what the device walker rejects whatever the bytes say (avr_walk.cuh Walker::start): known from the descriptor alone
bool statically_unsupported(const avr_slice_desc& d) {
  return d.slice_type > AVR_SLICE_I || d.chroma_format_idc != 1 || (d.flags & (AVR_SLICE_MBAFF | AVR_SLICE_FIELD)) ||
         d.mb_width < 1 || d.mb_height < 1 || uint64_t(d.mb_width) * d.mb_height > AVR_MAX_PICTURE_MBS ||
         d.first_mb >= uint32_t(d.mb_width) * d.mb_height;
}
uint8_t streams_for(const avr_slice_desc& d, uint32_t per_stream, uint32_t max_streams) {
  if (per_stream == 0 || d.mb_width < 1 || d.mb_width > 256) return 0;
  uint32_t k = d.payload_size / per_stream;
  const uint32_t rows = d.first_mb / d.mb_width < d.mb_height ? d.mb_height - d.first_mb / d.mb_width : 1;
  if (k > rows) k = rows;
  if (k > max_streams) k = max_streams;
  return k >= 2 ? uint8_t(k) : uint8_t(0);
}

// H.264 7.4.1 on a slice_data payload; z = zero bytes that end the slice header (0..2)
void unescape_payload(const uint8_t* p, size_t n, int z, std::vector<uint8_t>* out) {
  for (size_t i = 0; i < n; i++) {
    if (z >= 2 && p[i] == 3) { z = 0; continue; }
    out->push_back(p[i]);
    z = p[i] == 0 ? z + 1 : 0;
  }
}
bool escapes_back_to(const uint8_t* raw, size_t raw_n, int z, const uint8_t* p, size_t n) {   // what the device will write
  size_t k = 0;
  for (size_t i = 0; i < raw_n; i++) {
    if (z >= 2 && raw[i] <= 3) { if (k >= n || p[k++] != 3) return false; z = 0; }
    if (k >= n || p[k++] != raw[i]) return false;
    z = raw[i] == 0 ? z + 1 : 0;
  }
  return k == n;
}

void emit_block(std::string& out, const std::string& body) {
  put_varint(out, (2u << 3) | 2);
  put_varint(out, body.size());
  out.append(body);
}

// a byte vector whose resize() does not zero-fill (gigabytes of arena are overwritten right away)
template <class T>
struct no_init_alloc : std::allocator<T> {
  template <class U> struct rebind { typedef no_init_alloc<U> other; };
  template <class U> void construct(U* p) { ::new (static_cast<void*>(p)) U; }
  template <class U, class A0, class... A> void construct(U* p, A0&& a0, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A0>(a0), std::forward<A>(a)...); }
};
typedef std::vector<uint8_t, no_init_alloc<uint8_t>> RawBytes;

// Files of a call are independent on the host side too: index, container parse and assembly run one task per file.
template <class F>
void parallel_for(int n, F f) {
  const unsigned hw = std::thread::hardware_concurrency();
  const int workers = std::min(n, int(std::min(hw ? hw : 1u, 16u)));
  if (workers <= 1) { for (int i = 0; i < n; i++) f(i); return; }
  std::atomic<int> next(0);
  std::vector<std::thread> threads;
  for (int t = 0; t < workers; t++)
    threads.emplace_back([&]() { for (int i = next++; i < n; i = next++) f(i); });
  for (std::thread& t : threads) t.join();
}

// The devices a file call runs on: one context (avr_file_* / avr_files_*) or the contexts of an avr_multi.
struct DevSet { avr_ctx* const* ctx; int n; };

// One batch of slices over the devices of `ds`.  A UNIT is a run of slices that share a model -- from one
// AVR_SLICE_NEW_MODEL to the next: a slice in chunked mode, a file in the reference's mode -- and units are independent,
// so they are dealt out over the devices longest first (by input bytes) with no exchange between devices: every device
// gets its own compact arena (only the bytes of its slices travel) and runs its share on a host thread of its own.
// out_at[i]: where slice i's output bytes are (in the buffer of the context that produced them, valid until that
// context's next call).
int run_batch(const DevSet& ds, int direction, const std::vector<avr_slice_desc>& descs, const uint8_t* arena, uint64_t arena_bytes,
              std::vector<avr_slice_result>* res, std::vector<const uint8_t*>* out_at) {
  const size_t n = descs.size();
  res->assign(n, avr_slice_result());
  out_at->assign(n, nullptr);
  auto call = direction == 0 ? avr_compress : avr_decompress;
  if (ds.n == 1) {
    const uint8_t* out = nullptr;
    uint64_t out_bytes = 0;
    const int rc = call(ds.ctx[0], descs.data(), int(n), arena ? arena : reinterpret_cast<const uint8_t*>(""), arena_bytes, res->data(), &out, &out_bytes);
    if (rc) return rc;
    for (size_t i = 0; i < n; i++) (*out_at)[i] = out + (*res)[i].out_off;
    return AVR_OK;
  }
  struct Unit { size_t first, count; uint64_t weight; };
  std::vector<Unit> units;
  for (size_t i = 0; i < n; i++) {
    if (units.empty() || (descs[i].flags & AVR_SLICE_NEW_MODEL)) units.push_back(Unit{i, 0, 0});
    units.back().count++;
    units.back().weight += descs[i].in_size + 64;
  }
  std::vector<size_t> order(units.size());
  for (size_t u = 0; u < units.size(); u++) order[u] = u;
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return units[a].weight > units[b].weight; });
  struct Share { std::vector<size_t> idx; std::vector<avr_slice_desc> descs; std::vector<uint8_t> arena; std::vector<avr_slice_result> res;
                 const uint8_t* out = nullptr; uint64_t out_bytes = 0, load = 0; int rc = AVR_OK; };
  std::vector<Share> shares(size_t(ds.n));
  for (size_t u : order) {   // longest first onto the least loaded device
    size_t best = 0;
    for (size_t k = 1; k < shares.size(); k++) if (shares[k].load < shares[best].load) best = k;
    shares[best].load += units[u].weight;
    for (size_t i = units[u].first; i < units[u].first + units[u].count; i++) shares[best].idx.push_back(i);
  }
  for (Share& sh : shares) {
    std::sort(sh.idx.begin(), sh.idx.end());   // units stay contiguous and in file order
    for (size_t i : sh.idx) {
      avr_slice_desc d = descs[i];
      const size_t at = (sh.arena.size() + 15) & ~size_t(15);
      sh.arena.resize(at);
      if (d.in_size) {
        if (uint64_t(d.in_off) + d.in_size > arena_bytes) return AVR_ERR_ARG;
        sh.arena.insert(sh.arena.end(), arena + d.in_off, arena + d.in_off + d.in_size);
      }
      d.in_off = at;
      sh.descs.push_back(d);
    }
    sh.arena.resize(sh.arena.size() + 16);
    sh.res.assign(sh.descs.size(), avr_slice_result());
  }
  std::vector<std::thread> threads;
  for (int k = 0; k < ds.n; k++) {
    threads.emplace_back([&, k]() {
      Share& sh = shares[size_t(k)];
      if (sh.descs.empty()) return;
      sh.rc = call(ds.ctx[k], sh.descs.data(), int(sh.descs.size()), sh.arena.data(), sh.arena.size(), sh.res.data(), &sh.out, &sh.out_bytes);
    });
  }
  for (std::thread& t : threads) t.join();
  for (const Share& sh : shares) {
    if (sh.rc) return sh.rc;
    for (size_t k = 0; k < sh.idx.size(); k++) { (*res)[sh.idx[k]] = sh.res[k]; (*out_at)[sh.idx[k]] = sh.out + sh.res[k].out_off; }
  }
  return AVR_OK;
}

int to_malloc(const std::string& s, uint8_t** p, uint64_t* n) {
  *p = static_cast<uint8_t*>(malloc(s.size() ? s.size() : 1));
  if (!*p) return AVR_ERR_NOMEM;
  memcpy(*p, s.data(), s.size());
  *n = s.size();
  return AVR_OK;
}

}  // namespace

extern "C" {

void avr_free(void* p) { free(p); }

int avr_slice_streams(const avr_slice_desc* desc) { return desc ? streams_for(*desc, wide_stream_bytes(nullptr, 0), AVR_DEFAULT_WIDE_STREAMS) : 0; }
int avr_ctx_slice_streams(avr_ctx* ctx, const avr_slice_desc* desc, int n_batch_slices) {
  return desc ? streams_for(*desc, wide_stream_bytes(ctx, size_t(n_batch_slices < 0 ? 0 : n_batch_slices)), uint32_t(avr_ctx_wide_max_streams(ctx))) : 0;
}

// Shared driver of avr_file_compress / avr_files_compress: all CABAC slices of all files go to the
// device as ONE batch (files are independent, so in either mode every file starts a fresh model).
static int compress_many(const DevSet& ds, const uint8_t* const* files, const uint64_t* sizes, int n_files, int mode,
                         uint8_t** recs, uint64_t* rec_sizes) {
  avr_ctx* ctx = ds.ctx[0];   // the policy (wide slices) is that of the first context
  struct Plan {
    avr_index_entry* ents = nullptr;
    int n_ents = 0;
    std::vector<char> coded;      // 1 = goes to the device; 2 = goes there unescaped (chunked mode, escaped payload)
    std::vector<uint64_t> raw_off; // coded == 2: offset of the unescaped copy in the arena
    std::vector<uint32_t> raw_size;
    std::vector<int> map;         // batch position -> entry, for this file
    size_t first = 0;             // first batch position of this file
    uint64_t base = 0;            // offset of the file in the arena
  };
  std::vector<Plan> plans;
  plans.resize(size_t(n_files));
  auto cleanup = [&]() { for (Plan& p : plans) free(p.ents); };
  Stopwatch sw;
  // one contiguous arena (a single file is used in place)
  RawBytes arena_store;
  const uint8_t* arena = n_files == 1 ? files[0] : nullptr;
  uint64_t arena_bytes = n_files == 1 ? sizes[0] : 0;
  if (n_files != 1) {
    uint64_t total = 0;
    for (int f = 0; f < n_files; f++) { plans[f].base = total; total += (sizes[f] + 15) & ~uint64_t(15); }
    arena_store.resize(size_t(total) + 16);
    memset(arena_store.data() + total, 0, 16);
    arena = arena_store.data(); arena_bytes = total;
  }
  std::vector<int> rcs(size_t(n_files), AVR_OK);
  parallel_for(n_files, [&](int f) {   // one task per file: its copy into the arena and its slice index
    Plan& p = plans[f];
    if (n_files != 1) {
      memcpy(arena_store.data() + p.base, files[f], size_t(sizes[f]));
      memset(arena_store.data() + p.base + sizes[f], 0, size_t(((sizes[f] + 15) & ~uint64_t(15)) - sizes[f]));   // the pad up to the next file
    }
    rcs[f] = avr_index_h264(files[f], sizes[f], &p.ents, &p.n_ents);
    if (rcs[f]) return;
    p.coded.assign(size_t(p.n_ents), 0);
    p.raw_off.assign(size_t(p.n_ents), 0);
    p.raw_size.assign(size_t(p.n_ents), 0);
    // slices the device would only hand back as unsupported stay literals from the start (in file mode every failed
    // attempt costs a whole pass: the slices of a file share one model)
    for (int i = 0; i < p.n_ents; i++) p.coded[i] = (p.ents[i].recodable && !statically_unsupported(p.ents[i].desc)) ? 1 : 0;
  });
  for (int f = 0; f < n_files; f++) if (rcs[f]) { cleanup(); return rcs[f]; }
  // Chunked mode recodes slices whose payload contains emulation-prevention bytes too (the reference leaves
  // them as literals): the device gets an unescaped copy, appended to the arena, and puts the bytes back when
  // it decompresses.  Only if escaping the copy gives the file's bytes again.
  std::vector<uint8_t> raws;
  if (mode == AVR_MODE_SLICE) {
    for (int f = 0; f < n_files; f++) {
      Plan& p = plans[f];
      for (int i = 0; i < p.n_ents; i++) {
        const avr_slice_desc& d = p.ents[i].desc;
        if (p.coded[i] || !(d.n_streams & AVR_SLICE_ESCAPED) || statically_unsupported(d)) continue;
        const int z = (d.n_streams >> AVR_ESC_ZEROS_SHIFT) & 3;
        const size_t at = (raws.size() + 15) & ~size_t(15);
        raws.resize(at);
        unescape_payload(files[f] + d.in_off, d.payload_size, z, &raws);
        if (raws.size() - at >= 8 && escapes_back_to(raws.data() + at, raws.size() - at, z, files[f] + d.in_off, d.payload_size)) {
          p.coded[i] = 2; p.raw_off[i] = at; p.raw_size[i] = uint32_t(raws.size() - at);
        } else raws.resize(at);
      }
    }
    if (!raws.empty()) {   // one arena: the files, then the unescaped copies
      uint64_t total = 0;
      if (n_files == 1) {   // the file was going to be used in place: one copy, no zero fill first
        plans[0].base = 0;
        total = (sizes[0] + 15) & ~uint64_t(15);
        arena_store.reserve(size_t(total) + raws.size() + 16);
        arena_store.insert(arena_store.end(), files[0], files[0] + sizes[0]);
        arena_store.resize(size_t(total));
        memset(arena_store.data() + sizes[0], 0, size_t(total - sizes[0]));
      } else {
        total = arena_bytes;
        arena_store.resize(size_t(total));
      }
      arena_store.insert(arena_store.end(), raws.begin(), raws.end());
      arena_store.resize(arena_store.size() + 16);
      memset(arena_store.data() + arena_store.size() - 16, 0, 16);
      for (Plan& p : plans) for (int i = 0; i < p.n_ents; i++) if (p.coded[i] == 2) p.raw_off[i] += total;
      arena = arena_store.data(); arena_bytes = total + raws.size();
    }
  }
  sw.lap("compress: arena + index");
  size_t n_batch = 0;
  for (const Plan& p : plans) for (int i = 0; i < p.n_ents; i++) n_batch += p.coded[i] ? 1 : 0;
  const uint32_t per_stream = mode == AVR_MODE_SLICE ? wide_stream_bytes(ctx, n_batch) : 0;
  // chunked mode: fresh models start from the context's prior table, if it has one (every device the same one)
  const uint32_t priors = mode == AVR_MODE_SLICE ? avr_priors_digest(ctx) : 0;
  for (int k = 1; k < ds.n && mode == AVR_MODE_SLICE; k++) if (avr_priors_digest(ds.ctx[k]) != priors) { cleanup(); return AVR_ERR_PRIORS; }
  std::vector<avr_slice_desc> descs;
  std::vector<avr_slice_result> res;
  std::vector<const uint8_t*> out_at;
  // In file mode the slices of a file share one model, so a slice the device rejects poisons everything
  // after it; drop it (it becomes skip_coded + literal, recode.cpp:1289-1296) and run again.
  for (int attempt = 0; attempt < 64; attempt++) {
    descs.clear();
    for (Plan& p : plans) {
      p.map.clear();
      p.first = descs.size();
      for (int i = 0; i < p.n_ents; i++) {
        if (!p.coded[i] && mode == AVR_MODE_SLICE) continue;
        avr_slice_desc d = p.ents[i].desc;
        d.in_off += p.base;
        if (!p.coded[i]) {   // stays skip_coded + literal, but its frame_spec still flips the shared model's buffers
          d.flags |= AVR_SLICE_SKIP;
          d.in_off = 0; d.in_size = 0;
        }
        if (p.map.empty() || mode == AVR_MODE_SLICE) d.flags |= AVR_SLICE_NEW_MODEL;
        if (p.coded[i] == 2) { d.in_off = p.raw_off[i]; d.in_size = p.raw_size[i]; }
        d.n_streams = p.coded[i] ? streams_for(d, per_stream, uint32_t(avr_ctx_wide_max_streams(ctx))) : 0;
        if (priors && p.coded[i]) d.n_streams |= AVR_SLICE_PRIORS;
        descs.push_back(d);
        p.map.push_back(i);
      }
    }
    int rc = run_batch(ds, 0, descs, arena, arena_bytes, &res, &out_at);
    if (rc) { cleanup(); return rc; }
    // slice mode: a failed slice becomes a literal, done.  File mode: the first slice of a file that fails poisons the
    // model of everything behind it (those come back CORRUPT whatever they are): drop it, and with it every slice the
    // device called UNSUPPORTED on its own account (the walker checks that before it touches the model), and run again
    bool retry = false;
    for (Plan& p : plans) {
      bool first_failure = true;
      for (size_t k = 0; k < p.map.size(); k++) {
        const int st = res[p.first + k].status;
        if (st == AVR_SLICE_OK || !p.coded[p.map[k]]) continue;
        if (mode == AVR_MODE_SLICE) { p.coded[p.map[k]] = 0; continue; }
        if (first_failure || st == AVR_SLICE_UNSUPPORTED) { p.coded[p.map[k]] = 0; retry = true; }
        first_failure = false;
      }
    }
    if (!retry) break;
  }
  sw.lap("compress: device batch");
  parallel_for(n_files, [&](int f) {   // one task per file: its container
    const Plan& p = plans[f];
    const uint8_t* file = files[f];
    std::string o;
    o.reserve(size_t(sizes[f]) + size_t(sizes[f]) / 16 + 64 * size_t(p.n_ents) + 256);
    if (mode == AVR_MODE_SLICE) {
      std::string meta;
      put_len(meta, 1, reinterpret_cast<const uint8_t*>(kSliceModeTag), sizeof(kSliceModeTag) - 1);
      if (priors) put_var(meta, 2, priors);
      put_len(o, 1, reinterpret_cast<const uint8_t*>(meta.data()), meta.size());
    }
    uint64_t prev = 0;
    size_t k = 0;
    for (int i = 0; i < p.n_ents; i++) {
      const avr_slice_desc& d = p.ents[i].desc;
      const bool on_device = k < p.map.size() && p.map[k] == i;
      const avr_slice_result* r = on_device ? &res[p.first + k] : nullptr;
      const bool ok = on_device && p.coded[i] && r->status == AVR_SLICE_OK;
      std::string body;   // field-number order inside the block: size(1) literal(2) skip_coded(3) cabac(4) parity(5) last(6)
      if (ok) {
        std::string lit;
        put_len(lit, 2, file + prev, size_t(d.in_off - prev));
        emit_block(o, lit);
        put_var(body, 1, d.payload_size);
        put_len(body, 4, out_at[p.first + k], r->out_size);
        put_var(body, 5, d.payload_size & 1);
        if (d.payload_size > 1) put_len(body, 6, file + d.in_off + d.payload_size - 1, 1);
        if ((descs[p.first + k].n_streams & AVR_STREAMS_MASK) >= 2) put_var(body, 7, descs[p.first + k].n_streams & AVR_STREAMS_MASK);
        if (p.coded[i] == 2) put_var(body, 8, 1);
        prev = d.in_off + d.payload_size;
      } else {   // {size, skip_coded}: the bytes stay in the surrounding literal
        put_var(body, 1, d.payload_size);
        put_var(body, 3, 1);
      }
      emit_block(o, body);
      if (on_device) k++;
    }
    std::string lit;
    put_len(lit, 2, file + prev, size_t(sizes[f] - prev));
    emit_block(o, lit);
    rcs[f] = to_malloc(o, &recs[f], &rec_sizes[f]);
  });
  sw.lap("compress: containers");
  cleanup();
  int rc_all = AVR_OK;
  for (int f = 0; f < n_files; f++) if (rcs[f]) rc_all = rcs[f];
  return rc_all;
}

int avr_file_compress(avr_ctx* ctx, const uint8_t* file, uint64_t file_bytes, int mode, uint8_t** rec, uint64_t* rec_bytes) {
  if (!ctx || !file || !rec || !rec_bytes || (mode != AVR_MODE_FILE && mode != AVR_MODE_SLICE)) return AVR_ERR_ARG;
  const DevSet ds{&ctx, 1};
  return compress_many(ds, &file, &file_bytes, 1, mode, rec, rec_bytes);
}

int avr_files_compress(avr_ctx* ctx, const uint8_t* const* files, const uint64_t* file_bytes, int n_files, int mode,
                       uint8_t** recs, uint64_t* rec_bytes) {
  if (!ctx || n_files < 0 || (n_files > 0 && (!files || !file_bytes || !recs || !rec_bytes)) ||
      (mode != AVR_MODE_FILE && mode != AVR_MODE_SLICE)) return AVR_ERR_ARG;
  for (int f = 0; f < n_files; f++) { if (!files[f]) return AVR_ERR_ARG; recs[f] = nullptr; rec_bytes[f] = 0; }
  if (n_files == 0) return AVR_OK;
  const DevSet ds{&ctx, 1};
  int rc = compress_many(ds, files, file_bytes, n_files, mode, recs, rec_bytes);
  if (rc) for (int f = 0; f < n_files; f++) { free(recs[f]); recs[f] = nullptr; rec_bytes[f] = 0; }
  return rc;
}

// Shared driver of avr_file_decompress / avr_files_decompress: every Block.cabac of every container in one batch.
static int decompress_many(const DevSet& ds, const uint8_t* const* recs, const uint64_t* rec_sizes, int n_files,
                           uint8_t** files, uint64_t* file_sizes) {
  struct CodedRef { size_t block; uint64_t off; bool is_cabac; };
  struct Plan {
    std::vector<Block> blocks;
    std::string stream;               // the file being rebuilt, surrogate payloads first
    std::vector<CodedRef> coded;
    std::vector<size_t> which;        // batch position (relative to first) -> coded index
    size_t first = 0, count = 0;
    std::vector<avr_slice_desc> descs; // this file's share of the batch, in_off relative to its own coded bytes
    uint64_t in_bytes = 0, in_base = 0;
    int rc = AVR_OK;
  };
  std::vector<Plan> plans;
  plans.resize(size_t(n_files));
  std::vector<avr_slice_desc> descs;
  Stopwatch sw;
  std::string in;   // all Block.cabac bytes back to back
  // Phase A, one task per container (files are independent): parse, rebuild the stream with surrogate payloads, index it,
  // pair the entries with the coded blocks.  Descriptors carry offsets into the file's own run of coded bytes.
  parallel_for(n_files, [&](int f) {
    Plan& p = plans[f];
    p.rc = [&]() -> int {
      std::string version;
      uint64_t priors = 0;
      if (!parse_container(recs[f], size_t(rec_sizes[f]), &p.blocks, &version, &priors)) return AVR_ERR_FORMAT;
      const bool slice_mode = version == kSliceModeTag;
      if (!version.empty() && !slice_mode) return AVR_ERR_FORMAT;
      if (priors) {   // written from a prior table: every device must hold that very table
        if (!slice_mode) return AVR_ERR_FORMAT;
        for (int k = 0; k < ds.n; k++) if (avr_priors_digest(ds.ctx[k]) != priors) return AVR_ERR_PRIORS;
      }
      p.stream.reserve(size_t(rec_sizes[f]) + size_t(rec_sizes[f]) / 8 + 4096);
      // 1. rebuild the stream with surrogate payloads (recode.cpp:1359-1409, 1537-1544) so the slice
      //    headers can be indexed exactly as on the compress side
      for (size_t i = 0; i < p.blocks.size(); i++) {
        const Block& b = p.blocks[i];
        if (int(b.has_literal) + int(b.has_cabac) + int(b.has_skip) != 1) return AVR_ERR_FORMAT;   // recode.cpp:1364-1366
        if (b.has_literal) p.stream.append(reinterpret_cast<const char*>(b.literal), b.literal_n);
        else if (b.has_cabac) {
          // :1377-1379, 1538-1540; and no honest block is 4096x its coded size (an estimator never goes below 1/97:
          // a bin costs > 0.01 bit), so a few bytes of container cannot demand gigabytes of output
          if (!b.has_size || b.size < 8 || b.size > (int64_t(1) << 31) || uint64_t(b.size) > (uint64_t(b.cabac_n) + 16) * 4096) return AVR_ERR_FORMAT;
          p.coded.push_back(CodedRef{i, p.stream.size(), true});
          p.stream.append(size_t(b.size), 'X');
        } else {
          if (!b.has_size) return AVR_ERR_FORMAT;
          p.coded.push_back(CodedRef{i, 0, false});
        }
      }
      avr_index_entry* ents = nullptr;
      int n_ents = 0;
      int rc = avr_index_h264(reinterpret_cast<const uint8_t*>(p.stream.data()), p.stream.size(), &ents, &n_ents);
      if (rc) return rc;
      struct Free { avr_index_entry* e; ~Free() { free(e); } } guard{ents};
      // 2. pair index entries with coded blocks in order (recognize_coded_block, recode.cpp:1546-1573)
      if (size_t(n_ents) != p.coded.size()) return AVR_ERR_FORMAT;
      for (int i = 0; i < n_ents; i++) {
        const Block& b = p.blocks[p.coded[i].block];
        if (uint64_t(b.size) != ents[i].desc.payload_size) return AVR_ERR_FORMAT;   // "Invalid surrogate block size"
        if (!p.coded[i].is_cabac && slice_mode) continue;
        if (p.coded[i].is_cabac && ents[i].desc.in_off != p.coded[i].off) return AVR_ERR_FORMAT;
        avr_slice_desc d = ents[i].desc;
        d.in_off = p.in_bytes;
        d.in_size = p.coded[i].is_cabac ? uint32_t(b.cabac_n) : 0;
        if (!p.coded[i].is_cabac) d.flags |= AVR_SLICE_SKIP;
        if (p.which.empty() || slice_mode) d.flags |= AVR_SLICE_NEW_MODEL;
        if (b.streams) {
          if (!slice_mode || !p.coded[i].is_cabac || b.streams < 2 || b.streams > AVR_MAX_STREAMS) return AVR_ERR_FORMAT;
          d.n_streams = uint8_t((d.n_streams & ~AVR_STREAMS_MASK) | b.streams);
        }
        if (b.escaped) {
          if (!slice_mode || !p.coded[i].is_cabac) return AVR_ERR_FORMAT;
          d.n_streams |= AVR_SLICE_ESCAPED;
        }
        if (priors && p.coded[i].is_cabac) d.n_streams |= AVR_SLICE_PRIORS;
        p.in_bytes += d.in_size;
        p.descs.push_back(d);
        p.which.push_back(size_t(i));
      }
      p.count = p.which.size();
      return AVR_OK;
    }();
  });
  for (const Plan& p : plans) if (p.rc) return p.rc;
  // Phase B: the batch -- every file's coded bytes back to back (copied in parallel), descriptors rebased
  uint64_t in_total = 0;
  for (Plan& p : plans) { p.in_base = in_total; in_total += p.in_bytes; p.first = descs.size(); for (avr_slice_desc d : p.descs) { d.in_off += p.in_base; descs.push_back(d); } }
  in.resize(size_t(in_total));
  parallel_for(n_files, [&](int f) {
    const Plan& p = plans[f];
    for (size_t k = 0; k < p.descs.size(); k++)
      if (p.descs[k].in_size) memcpy(&in[size_t(p.in_base + p.descs[k].in_off)], p.blocks[p.coded[p.which[k]].block].cabac, p.descs[k].in_size);
  });
  sw.lap("decompress: parse + index");
  std::vector<avr_slice_result> res;
  std::vector<const uint8_t*> out_at;
  if (in.empty()) in.push_back(0);
  int rc = run_batch(ds, 1, descs, reinterpret_cast<const uint8_t*>(in.data()), in.size(), &res, &out_at);
  if (rc) return rc;
  sw.lap("decompress: device batch");
  // 3. splice the rebuilt payloads in, with the x264 padding fix-up of recode.cpp:1345-1354, and hand the files over
  std::vector<int> rcs(size_t(n_files), AVR_OK);
  parallel_for(n_files, [&](int f) {
    Plan& p = plans[f];
    for (size_t k = 0; k < p.count; k++) {
      const avr_slice_result& r = res[p.first + k];
      if (r.status != AVR_SLICE_OK) { rcs[f] = AVR_ERR_FORMAT; return; }   // "Not all blocks were decoded."
      if (descs[p.first + k].flags & AVR_SLICE_SKIP) continue;
      const CodedRef& cr = p.coded[p.which[k]];
      const Block& b = p.blocks[cr.block];
      size_t n = r.out_size;
      const bool fix = b.has_parity && b.has_last && b.last_n > 0;
      if (fix && int(b.parity) != int(n & 1)) n++;              // the byte x264 padded with comes back (recode.cpp:1347-1354)
      if (n != size_t(b.size)) { rcs[f] = AVR_ERR_FORMAT; return; }
      memcpy(&p.stream[cr.off], out_at[p.first + k], std::min<size_t>(n, r.out_size));
      if (fix && n > 0) p.stream[cr.off + n - 1] = char(b.last[0]);
    }
    rcs[f] = to_malloc(p.stream, &files[f], &file_sizes[f]);
  });
  sw.lap("decompress: splice + copy out");
  int rc_all = AVR_OK;
  for (int f = 0; f < n_files; f++) if (rcs[f]) rc_all = rcs[f];
  if (rc_all) return rc_all;
  return rc_all;
}

int avr_file_decompress(avr_ctx* ctx, const uint8_t* rec, uint64_t rec_bytes, uint8_t** file, uint64_t* file_bytes) {
  if (!ctx || !rec || !file || !file_bytes) return AVR_ERR_ARG;
  *file = nullptr; *file_bytes = 0;
  const DevSet ds{&ctx, 1};
  return decompress_many(ds, &rec, &rec_bytes, 1, file, file_bytes);
}

int avr_files_decompress(avr_ctx* ctx, const uint8_t* const* recs, const uint64_t* rec_bytes, int n_files,
                         uint8_t** files, uint64_t* file_bytes) {
  if (!ctx || n_files < 0 || (n_files > 0 && (!recs || !rec_bytes || !files || !file_bytes))) return AVR_ERR_ARG;
  for (int f = 0; f < n_files; f++) { if (!recs[f]) return AVR_ERR_ARG; files[f] = nullptr; file_bytes[f] = 0; }
  if (n_files == 0) return AVR_OK;
  const DevSet ds{&ctx, 1};
  int rc = decompress_many(ds, recs, rec_bytes, n_files, files, file_bytes);
  if (rc) for (int f = 0; f < n_files; f++) { free(files[f]); files[f] = nullptr; file_bytes[f] = 0; }
  return rc;
}

// ---- several devices behind one handle ---------------------------------------------------------------
struct avr_multi { std::vector<avr_ctx*> ctx; };

int avr_multi_create(const avr_config* cfg, const int* devices, int n_devices, avr_multi** out) {
  if (!out) return AVR_ERR_ARG;
  *out = nullptr;
  if (n_devices < 1 || n_devices > 64 || !devices) return AVR_ERR_ARG;
  for (int i = 0; i < n_devices; i++) for (int j = 0; j < i; j++) if (devices[i] == devices[j]) return AVR_ERR_ARG;
  avr_multi* m = new (std::nothrow) avr_multi();
  if (!m) return AVR_ERR_NOMEM;
  for (int i = 0; i < n_devices; i++) {
    avr_config c;
    if (cfg) c = *cfg; else avr_config_init(&c);
    c.device = devices[i];
    avr_ctx* x = nullptr;
    const int rc = avr_create(&c, &x);
    if (rc) { avr_multi_destroy(m); return rc; }
    m->ctx.push_back(x);
  }
  *out = m;
  return AVR_OK;
}
void avr_multi_destroy(avr_multi* m) {
  if (!m) return;
  for (avr_ctx* c : m->ctx) avr_destroy(c);
  delete m;
}
int avr_multi_devices(const avr_multi* m) { return m ? int(m->ctx.size()) : 0; }
avr_ctx* avr_multi_ctx(avr_multi* m, int i) { return m && i >= 0 && size_t(i) < m->ctx.size() ? m->ctx[size_t(i)] : nullptr; }

int avr_multi_set_priors(avr_multi* m, const uint16_t* table, uint32_t n_entries) {
  if (!m) return AVR_ERR_ARG;
  for (avr_ctx* c : m->ctx) { const int rc = avr_set_priors(c, table, n_entries); if (rc) return rc; }
  return AVR_OK;
}
int avr_multi_files_compress(avr_multi* m, const uint8_t* const* files, const uint64_t* file_bytes, int n_files, int mode,
                             uint8_t** recs, uint64_t* rec_bytes) {
  if (!m || m->ctx.empty() || n_files < 0 || (n_files > 0 && (!files || !file_bytes || !recs || !rec_bytes)) ||
      (mode != AVR_MODE_FILE && mode != AVR_MODE_SLICE)) return AVR_ERR_ARG;
  for (int f = 0; f < n_files; f++) { if (!files[f]) return AVR_ERR_ARG; recs[f] = nullptr; rec_bytes[f] = 0; }
  if (n_files == 0) return AVR_OK;
  const DevSet ds{m->ctx.data(), int(m->ctx.size())};
  int rc = compress_many(ds, files, file_bytes, n_files, mode, recs, rec_bytes);
  if (rc) for (int f = 0; f < n_files; f++) { free(recs[f]); recs[f] = nullptr; rec_bytes[f] = 0; }
  return rc;
}
int avr_multi_files_decompress(avr_multi* m, const uint8_t* const* recs, const uint64_t* rec_bytes, int n_files,
                               uint8_t** files, uint64_t* file_bytes) {
  if (!m || m->ctx.empty() || n_files < 0 || (n_files > 0 && (!recs || !rec_bytes || !files || !file_bytes))) return AVR_ERR_ARG;
  for (int f = 0; f < n_files; f++) { if (!recs[f]) return AVR_ERR_ARG; files[f] = nullptr; file_bytes[f] = 0; }
  if (n_files == 0) return AVR_OK;
  const DevSet ds{m->ctx.data(), int(m->ctx.size())};
  int rc = decompress_many(ds, recs, rec_bytes, n_files, files, file_bytes);
  if (rc) for (int f = 0; f < n_files; f++) { free(files[f]); files[f] = nullptr; file_bytes[f] = 0; }
  return rc;
}

}  // extern "C"
