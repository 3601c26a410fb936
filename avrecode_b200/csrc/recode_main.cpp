// recode -- drop-in CLI: recode [compress|decompress|roundtrip] <input> [output]
// Same verbs, argument order, exit codes and roundtrip report as the reference (recode.cpp:1594-1659).
// Options come from the environment so the argv surface stays identical:
//   AVR_MODE=slice (default; B200 chunked format) | file (reference-parity bytes)
//   AVR_DEVICE=<cuda ordinal>
//   AVR_DEVICES=<ordinal>,<ordinal>,...   several devices: the slices (AVR_MODE=slice) of the file are dealt out over them
//   AVR_PRIORS=<path> | none   the prior table of chunked mode (default: priors_v1.bin beside this executable, if it is there)
#include <stdio.h>
#include <unistd.h>
#include <stdlib.h>
#include <string.h>
#include <fstream>
#include <iostream>
#include <iterator>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <vector>
#include "../../include/avrecode_b200.h"

namespace {

std::vector<uint8_t> read_file(const std::string& name) {
  std::ifstream f(name, std::ios::binary);
  if (!f) throw std::invalid_argument("Failed to open file: " + name);
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

void check(int rc, const char* what) {
  if (rc != AVR_OK) throw std::runtime_error(std::string(what) + ": " + avr_strerror(rc));
}

struct Ctx {
  avr_multi* m = nullptr;
  Ctx() {
    avr_config cfg;
    avr_config_init(&cfg);
    std::vector<int> devices;
    if (const char* list = getenv("AVR_DEVICES")) {
      for (const char* p = list; *p;) {
        char* end = nullptr;
        const long v = strtol(p, &end, 10);
        if (end == p) throw std::invalid_argument(std::string("AVR_DEVICES: not a list of device ordinals: ") + list);
        devices.push_back(int(v));
        p = *end == ',' ? end + 1 : end;
        if (*end && *end != ',') throw std::invalid_argument(std::string("AVR_DEVICES: not a list of device ordinals: ") + list);
      }
    }
    if (devices.empty()) devices.push_back(getenv("AVR_DEVICE") ? atoi(getenv("AVR_DEVICE")) : 0);
    check(avr_multi_create(&cfg, devices.data(), int(devices.size()), &m), "avr_multi_create");
    load_priors();
  }
  // chunked mode starts fresh models from a prior table; a container records which one (AVR_ERR_PRIORS on a mismatch)
  void load_priors() {
    const char* env = getenv("AVR_PRIORS");
    if (env && strcmp(env, "none") == 0) return;
    std::string path = env ? env : "";
    if (path.empty()) {
      char self[4096];
      const ssize_t n = readlink("/proc/self/exe", self, sizeof self - 1);
      if (n <= 0) return;
      self[n] = 0;
      path = self;
      path = path.substr(0, path.rfind('/') + 1) + "priors_v1.bin";
    }
    std::ifstream f(path, std::ios::binary);
    if (!f) { if (env) throw std::invalid_argument("AVR_PRIORS: cannot open " + path); return; }
    std::vector<char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (raw.size() != size_t(AVR_PRIOR_ENTRIES) * 2) throw std::invalid_argument("not a prior table: " + path);
    check(avr_multi_set_priors(m, reinterpret_cast<const uint16_t*>(raw.data()), AVR_PRIOR_ENTRIES), "avr_multi_set_priors");
  }
  ~Ctx() { avr_multi_destroy(m); }
};

int mode_from_env() {
  const char* m = getenv("AVR_MODE");
  if (m && strcmp(m, "file") == 0) return AVR_MODE_FILE;
  return AVR_MODE_SLICE;
}

std::string compress(Ctx& ctx, const std::vector<uint8_t>& in) {
  uint8_t* rec = nullptr; uint64_t n = 0;
  const uint8_t* file = in.data();
  const uint64_t size = in.size();
  check(avr_multi_files_compress(ctx.m, &file, &size, 1, mode_from_env(), &rec, &n), "compress");
  std::string s(reinterpret_cast<char*>(rec), size_t(n));
  avr_free(rec);
  return s;
}
std::string decompress(Ctx& ctx, const uint8_t* p, size_t n_in) {
  uint8_t* out = nullptr; uint64_t n = 0;
  const uint64_t size = n_in;
  check(avr_multi_files_decompress(ctx.m, &p, &size, 1, &out, &n), "decompress");
  std::string s(reinterpret_cast<char*>(out), size_t(n));
  avr_free(out);
  return s;
}

// sums literal + cabac payload bytes of a container (for the "protobuf overhead" line, recode.cpp:1608-1614)
size_t payload_bytes(const std::string& rec) {
  size_t total = 0;
  const uint8_t* p = reinterpret_cast<const uint8_t*>(rec.data());
  const uint8_t* e = p + rec.size();
  auto varint = [&](const uint8_t*& q, const uint8_t* end) { uint64_t v = 0; int sh = 0; while (q < end) { uint8_t b = *q++; v |= uint64_t(b & 0x7F) << sh; if (!(b & 0x80)) break; sh += 7; } return v; };
  while (p < e) {
    uint64_t key = varint(p, e);
    if ((key & 7) != 2) { varint(p, e); continue; }
    uint64_t len = varint(p, e);
    if ((key >> 3) == 2) {
      const uint8_t* q = p; const uint8_t* qe = p + len;
      while (q < qe) {
        uint64_t k2 = varint(q, qe);
        if ((k2 & 7) == 2) { uint64_t l2 = varint(q, qe); if ((k2 >> 3) == 2 || (k2 >> 3) == 4) total += l2; q += l2; }
        else varint(q, qe);
      }
    }
    p += len;
  }
  return total;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 3 || argc > 4) {
    std::cerr << "Usage: " << argv[0] << " [compress|decompress|roundtrip] <input> [output]" << std::endl;
    return 1;
  }
  std::string command = argv[1], input = argv[2];
  std::ofstream out_file;
  if (argc > 3) out_file.open(argv[3], std::ios::binary);
  try {
    if (command == "compress") {
      Ctx ctx;
      std::string rec = compress(ctx, read_file(input));
      (out_file.is_open() ? static_cast<std::ostream&>(out_file) : std::cout) << rec;
    } else if (command == "decompress") {
      Ctx ctx;
      std::vector<uint8_t> in = read_file(input);
      std::string out = decompress(ctx, in.data(), in.size());
      (out_file.is_open() ? static_cast<std::ostream&>(out_file) : std::cout) << out;
    } else if (command == "roundtrip") {
      Ctx ctx;
      std::vector<uint8_t> in = read_file(input);
      std::string rec = compress(ctx, in);
      std::string back = decompress(ctx, reinterpret_cast<const uint8_t*>(rec.data()), rec.size());
      if (back.size() == in.size() && memcmp(back.data(), in.data(), in.size()) == 0) {
        if (out_file.is_open()) out_file << rec;
        double ratio = rec.size() * 1.0 / in.size();
        double overhead = (rec.size() - payload_bytes(rec)) * 1.0 / rec.size();
        std::cout << "Compress-decompress roundtrip succeeded:" << std::endl;
        std::cout << " compression ratio: " << ratio * 100. << "%" << std::endl;
        std::cout << " protobuf overhead: " << overhead * 100. << "%" << std::endl;
        return 0;
      }
      std::cerr << "Compress-decompress roundtrip failed." << std::endl;
      return 1;
    } else {
      throw std::invalid_argument("Unknown command: " + command);
    }
  } catch (const std::exception& e) {
    std::cerr << "Exception (" << typeid(e).name() << "): " << e.what() << std::endl;
    return 1;
  }
  return 0;
}
