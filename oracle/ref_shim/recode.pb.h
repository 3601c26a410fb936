// Minimal stand-in for the protoc-generated recode.pb.h (protoc / libprotobuf are not
// installed in this image).  Written from the schema in recode.proto:1-20 and the public
// proto2 wire format; used ONLY to link the unmodified reference recode.cpp into
// oracle/_ref/ (test infrastructure -- never part of the product).
//
// Serialisation order follows libprotobuf: fields in field-number order, repeated
// messages in insertion order, unset optional fields omitted.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>
#include <functional>

namespace avr_pb_shim {
inline void put_varint(std::string& o, uint64_t v) {
  while (v >= 0x80) { o.push_back(char((v & 0x7F) | 0x80)); v >>= 7; }
  o.push_back(char(v));
}
inline bool get_varint(const uint8_t*& p, const uint8_t* e, uint64_t* v) {
  uint64_t r = 0; int sh = 0;
  while (p < e && sh < 64) {
    uint8_t b = *p++;
    r |= uint64_t(b & 0x7F) << sh;
    if (!(b & 0x80)) { *v = r; return true; }
    sh += 7;
  }
  return false;
}
inline void put_bytes(std::string& o, int field, const std::string& s) {
  put_varint(o, (uint64_t(field) << 3) | 2);
  put_varint(o, s.size());
  o.append(s);
}
}  // namespace avr_pb_shim

class Recoded {
 public:
  class Block {
   public:
    bool has_size() const { return has_size_; }
    int64_t size() const { return size_; }
    void set_size(int64_t v) { size_ = v; has_size_ = true; }

    bool has_literal() const { return has_literal_; }
    const std::string& literal() const { return literal_; }
    void set_literal(const void* p, size_t n) { literal_.assign((const char*)p, n); has_literal_ = true; }
    void set_literal(const std::string& s) { literal_ = s; has_literal_ = true; }

    bool has_skip_coded() const { return has_skip_coded_; }
    bool skip_coded() const { return skip_coded_; }
    void set_skip_coded(bool v) { skip_coded_ = v; has_skip_coded_ = true; }

    bool has_cabac() const { return has_cabac_; }
    const std::string& cabac() const { return cabac_; }
    void set_cabac(const void* p, size_t n) { cabac_.assign((const char*)p, n); has_cabac_ = true; }

    bool has_length_parity() const { return has_length_parity_; }
    bool length_parity() const { return length_parity_; }
    void set_length_parity(bool v) { length_parity_ = v; has_length_parity_ = true; }

    bool has_last_byte() const { return has_last_byte_; }
    const std::string& last_byte() const { return last_byte_; }
    void set_last_byte(const void* p, size_t n) { last_byte_.assign((const char*)p, n); has_last_byte_ = true; }

    void AppendTo(std::string& o) const {
      using namespace avr_pb_shim;
      if (has_size_) { put_varint(o, (1 << 3) | 0); put_varint(o, uint64_t(size_)); }
      if (has_literal_) put_bytes(o, 2, literal_);
      if (has_skip_coded_) { put_varint(o, (3 << 3) | 0); put_varint(o, skip_coded_ ? 1 : 0); }
      if (has_cabac_) put_bytes(o, 4, cabac_);
      if (has_length_parity_) { put_varint(o, (5 << 3) | 0); put_varint(o, length_parity_ ? 1 : 0); }
      if (has_last_byte_) put_bytes(o, 6, last_byte_);
    }
    bool Parse(const uint8_t* p, const uint8_t* e) {
      using namespace avr_pb_shim;
      while (p < e) {
        uint64_t key; if (!get_varint(p, e, &key)) return false;
        int field = int(key >> 3), wt = int(key & 7);
        if (wt == 0) {
          uint64_t v; if (!get_varint(p, e, &v)) return false;
          if (field == 1) set_size(int64_t(v));
          else if (field == 3) set_skip_coded(v != 0);
          else if (field == 5) set_length_parity(v != 0);
        } else if (wt == 2) {
          uint64_t n; if (!get_varint(p, e, &n) || n > uint64_t(e - p)) return false;
          if (field == 2) set_literal(p, n);
          else if (field == 4) set_cabac(p, n);
          else if (field == 6) set_last_byte(p, n);
          p += n;
        } else if (wt == 1) { if (e - p < 8) return false; p += 8; }
        else if (wt == 5) { if (e - p < 4) return false; p += 4; }
        else return false;
      }
      return true;
    }

   private:
    int64_t size_ = 0; bool has_size_ = false;
    std::string literal_; bool has_literal_ = false;
    bool skip_coded_ = false; bool has_skip_coded_ = false;
    std::string cabac_; bool has_cabac_ = false;
    bool length_parity_ = false; bool has_length_parity_ = false;
    std::string last_byte_; bool has_last_byte_ = false;
  };

  ~Recoded() { for (Block* b : blocks_) delete b; }
  Recoded() = default;
  Recoded(const Recoded&) = delete;
  Recoded& operator=(const Recoded&) = delete;

  Block* add_block() { blocks_.push_back(new Block()); return blocks_.back(); }
  int block_size() const { return int(blocks_.size()); }
  const Block& block(int i) const { return *blocks_[i]; }

  // range-for support: `for (const auto& block : proto.block())`
  struct BlockRange {
    const std::vector<Block*>* v;
    struct It {
      std::vector<Block*>::const_iterator it;
      const Block& operator*() const { return **it; }
      It& operator++() { ++it; return *this; }
      bool operator!=(const It& o) const { return it != o.it; }
    };
    It begin() const { return It{v->begin()}; }
    It end() const { return It{v->end()}; }
  };
  BlockRange block() const { return BlockRange{&blocks_}; }

  std::string SerializeAsString() const {
    std::string o;
    if (has_metadata_) avr_pb_shim::put_bytes(o, 1, metadata_raw_);
    for (const Block* b : blocks_) {
      std::string body; b->AppendTo(body);
      avr_pb_shim::put_bytes(o, 2, body);
    }
    return o;
  }
  bool ParseFromArray(const void* data, int size) {
    using namespace avr_pb_shim;
    for (Block* b : blocks_) delete b;
    blocks_.clear(); has_metadata_ = false; metadata_raw_.clear();
    const uint8_t* p = (const uint8_t*)data; const uint8_t* e = p + size;
    while (p < e) {
      uint64_t key; if (!get_varint(p, e, &key)) return false;
      int field = int(key >> 3), wt = int(key & 7);
      if (wt == 2) {
        uint64_t n; if (!get_varint(p, e, &n) || n > uint64_t(e - p)) return false;
        if (field == 1) { metadata_raw_.assign((const char*)p, n); has_metadata_ = true; }
        else if (field == 2) { if (!add_block()->Parse(p, p + n)) return false; }
        p += n;
      } else if (wt == 0) { uint64_t v; if (!get_varint(p, e, &v)) return false; }
      else if (wt == 1) { if (e - p < 8) return false; p += 8; }
      else if (wt == 5) { if (e - p < 4) return false; p += 4; }
      else return false;
    }
    return true;
  }
  bool ParseFromString(const std::string& s) { return ParseFromArray(s.data(), int(s.size())); }

 private:
  std::vector<Block*> blocks_;
  std::string metadata_raw_; bool has_metadata_ = false;
};
