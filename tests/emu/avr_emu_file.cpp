// avr_emu_file.cpp -- TEST INFRASTRUCTURE ONLY: the product's host file layer (avrecode_b200/csrc/avr_file.cpp and
// h264_index.cpp, compiled unchanged) linked against the HOST EMULATION of the device code instead of the CUDA
// library, so that the container codec, the retry policy and the multi-device sharding of avr_multi_* can be tested in
// the GPU-less dev container (-m "not gpu").  "Devices" here are just emulation contexts; any ordinal is accepted.
// Never part of libavrecode_b200.so.
#include "avr_emu.cpp"

struct avr_ctx {
  avr_config cfg;
  std::vector<uint8_t> out;
  std::vector<uint16_t> priors;
  int calls = 0;
};

template <class Dir>
static int emu_call(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                    const uint8_t** out, uint64_t* out_bytes) {
  if (!c || n < 0) return AVR_ERR_ARG;
  for (int i = 0; i < n; i++) {   // what avr_stage refuses (avr_lib.cu)
    const uint32_t k = descs[i].n_streams & AVR_STREAMS_MASK;
    if (k > AVR_MAX_STREAMS) return AVR_ERR_ARG;
    if ((descs[i].n_streams & AVR_SLICE_PRIORS) && c->priors.empty()) return AVR_ERR_PRIORS;
  }
  g_priors = c->priors.empty() ? nullptr : c->priors.data();
  uint8_t* o = nullptr;
  uint64_t ob = 0;
  const int rc = run<Dir>(descs, n, in, in_bytes, res, &o, &ob);
  if (rc) return rc;
  c->out.assign(o, o + ob);
  c->out.resize(c->out.size() + 16);
  free(o);
  c->calls++;
  *out = c->out.data();
  *out_bytes = ob;
  return AVR_OK;
}

extern "C" {
int avr_abi_version(void) { return AVR_ABI_VERSION; }
void avr_config_init(avr_config* cfg) {
  memset(cfg, 0, sizeof *cfg);
  cfg->abi_version = AVR_ABI_VERSION;
  cfg->wide_stream_bytes = -1; cfg->resident = -1; cfg->wide_batch_limit = -1;
}
int avr_create(const avr_config* cfg, avr_ctx** out) {
  if (!out) return AVR_ERR_ARG;
  avr_ctx* c = new avr_ctx();
  if (cfg) c->cfg = *cfg; else avr_config_init(&c->cfg);
  *out = c;
  return AVR_OK;
}
void avr_destroy(avr_ctx* c) { delete c; }
int avr_emu_ctx_calls(avr_ctx* c) { return c ? c->calls : -1; }
int avr_compress(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                 const uint8_t** out, uint64_t* out_bytes) {
  return emu_call<Compress>(c, descs, n, in, in_bytes, res, out, out_bytes);
}
int avr_decompress(avr_ctx* c, const avr_slice_desc* descs, int n, const uint8_t* in, uint64_t in_bytes, avr_slice_result* res,
                   const uint8_t** out, uint64_t* out_bytes) {
  return emu_call<Decompress>(c, descs, n, in, in_bytes, res, out, out_bytes);
}
uint32_t avr_priors_digest_of(const uint16_t* table, uint32_t n_entries) {   // as avr_lib.cu
  if (!table) return 0;
  uint32_t h = 2166136261u;
  for (uint32_t i = 0; i < n_entries; i++) { h = (h ^ (table[i] & 0xFF)) * 16777619u; h = (h ^ (table[i] >> 8)) * 16777619u; }
  return h ? h : 1u;
}
int avr_set_priors(avr_ctx* c, const uint16_t* table, uint32_t n_entries) {
  if (!c) return AVR_ERR_ARG;
  if (!table) { c->priors.clear(); return AVR_OK; }
  if (n_entries != AVR_PRIOR_ENTRIES) return AVR_ERR_ARG;
  for (uint32_t i = 0; i < n_entries; i++) {
    const uint32_t pos = table[i] & 0xFF, tot = table[i] >> 8;
    if (pos < 1 || tot <= pos || tot > LIMIT_SIG) return AVR_ERR_ARG;
  }
  c->priors.assign(table, table + n_entries);
  return AVR_OK;
}
uint32_t avr_priors_digest(avr_ctx* c) { return c && !c->priors.empty() ? avr_priors_digest_of(c->priors.data(), uint32_t(c->priors.size())) : 0; }
int avr_ctx_wide_stream_bytes(avr_ctx* c) { return !c || c->cfg.wide_stream_bytes < 0 ? 32768 : c->cfg.wide_stream_bytes; }
int avr_ctx_wide_batch_limit(avr_ctx* c) { return !c ? 0 : c->cfg.wide_batch_limit < 0 ? 4 * 148 : c->cfg.wide_batch_limit; }
int avr_ctx_wide_max_streams(avr_ctx* c) {
  return !c || c->cfg.wide_max_streams < 2 ? AVR_DEFAULT_WIDE_STREAMS : c->cfg.wide_max_streams > AVR_MAX_STREAMS ? AVR_MAX_STREAMS : c->cfg.wide_max_streams;
}
}
