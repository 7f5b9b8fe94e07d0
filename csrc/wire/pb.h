// pb.h — minimal proto3 wire codec (no protobuf runtime on the box).
//
// Encodes / decodes exactly the 19 messages of the reference's two .proto files
// (/root/reference/proto/coordinator.proto:12-59, proto/parameter_server.proto:13-77) so the
// C++ binaries stay wire-compatible with any gRPC peer built from those files.
// proto3 rules implemented: zero-valued scalars are omitted, repeated scalars are written
// packed, decoders accept packed and unpacked repeated scalars and skip unknown fields,
// negative int32 are 10-byte varints.  Golden bytes: SURVEY.md App. A, tests/test_wire.py.
#pragma once

#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

namespace psb {
namespace pb {

enum WireType { kVarint = 0, kFixed64 = 1, kLen = 2, kFixed32 = 5 };

inline size_t varint_size(uint64_t v) {
  size_t n = 1;
  while (v >= 0x80) {
    v >>= 7;
    ++n;
  }
  return n;
}

class Writer {
 public:
  std::string buf;

  void varint(uint64_t v) {
    while (v >= 0x80) {
      buf.push_back(static_cast<char>((v & 0x7f) | 0x80));
      v >>= 7;
    }
    buf.push_back(static_cast<char>(v));
  }
  void tag(int field, int wt) { varint((static_cast<uint64_t>(field) << 3) | wt); }
  void i32(int field, int32_t v) {
    if (v == 0) return;
    tag(field, kVarint);
    varint(static_cast<uint64_t>(static_cast<int64_t>(v)));  // sign-extended: 10 bytes if < 0
  }
  void i64(int field, int64_t v) {
    if (v == 0) return;
    tag(field, kVarint);
    varint(static_cast<uint64_t>(v));
  }
  void boolean(int field, bool v) {
    if (!v) return;
    tag(field, kVarint);
    buf.push_back(1);
  }
  void str(int field, std::string_view s) {
    if (s.empty()) return;
    tag(field, kLen);
    varint(s.size());
    buf.append(s.data(), s.size());
  }
  // embedded message: always written, even when empty (presence of a repeated element)
  void msg(int field, const std::string& body) {
    tag(field, kLen);
    varint(body.size());
    buf.append(body);
  }
  void packed_i32(int field, const std::vector<int32_t>& v) {
    if (v.empty()) return;
    Writer inner;
    for (int32_t x : v) inner.varint(static_cast<uint64_t>(static_cast<int64_t>(x)));
    tag(field, kLen);
    varint(inner.buf.size());
    buf.append(inner.buf);
  }
  void f32(int field, float v) {
    if (v == 0.0f) return;
    tag(field, kFixed32);
    buf.append(reinterpret_cast<const char*>(&v), 4);
  }
  void packed_sfixed64(int field, const int64_t* data, size_t n) {
    if (n == 0) return;
    tag(field, kLen);
    varint(n * 8);
    buf.append(reinterpret_cast<const char*>(data), n * 8);  // little-endian host
  }
  void packed_f32(int field, const float* data, size_t n) {
    if (n == 0) return;
    tag(field, kLen);
    varint(n * 4);
    buf.append(reinterpret_cast<const char*>(data), n * 4);  // little-endian host
  }
};

class Reader {
 public:
  Reader(const void* data, size_t n)
      : p_(static_cast<const uint8_t*>(data)), end_(p_ + n), ok_(true) {}
  explicit Reader(std::string_view s) : Reader(s.data(), s.size()) {}

  bool ok() const { return ok_; }
  bool next(int& field, int& wt) {
    if (!ok_ || p_ >= end_) return false;
    uint64_t t = varint();
    if (!ok_) return false;
    field = static_cast<int>(t >> 3);
    wt = static_cast<int>(t & 7);
    if (field == 0) ok_ = false;
    return ok_;
  }
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 70; shift += 7) {
      if (p_ >= end_) {
        ok_ = false;
        return 0;
      }
      uint8_t b = *p_++;
      v |= static_cast<uint64_t>(b & 0x7f) << shift;
      if (!(b & 0x80)) return v;
    }
    ok_ = false;
    return 0;
  }
  std::string_view bytes() {
    uint64_t n = varint();
    if (!ok_ || n > static_cast<uint64_t>(end_ - p_)) {
      ok_ = false;
      return {};
    }
    std::string_view s(reinterpret_cast<const char*>(p_), n);
    p_ += n;
    return s;
  }
  uint32_t fixed32() {
    if (end_ - p_ < 4) {
      ok_ = false;
      return 0;
    }
    uint32_t v;
    memcpy(&v, p_, 4);
    p_ += 4;
    return v;
  }
  uint64_t fixed64() {
    if (end_ - p_ < 8) {
      ok_ = false;
      return 0;
    }
    uint64_t v;
    memcpy(&v, p_, 8);
    p_ += 8;
    return v;
  }
  void skip(int wt) {
    switch (wt) {
      case kVarint: varint(); break;
      case kFixed64:
        if (end_ - p_ < 8) ok_ = false; else p_ += 8;
        break;
      case kLen: bytes(); break;
      case kFixed32:
        if (end_ - p_ < 4) ok_ = false; else p_ += 4;
        break;
      default: ok_ = false;
    }
  }

 private:
  const uint8_t* p_;
  const uint8_t* end_;
  bool ok_;
};

// ------------------------------------------------------------------ coordinator.proto
struct WorkerInfo {
  int32_t worker_id = 0;
  std::string address;
  int32_t port = 0;
  std::string hostname;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.str(2, address), w.i32(3, port), w.str(4, hostname);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kLen) address = std::string(r.bytes());
      else if (f == 3 && wt == kVarint) port = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kLen) hostname = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct RegisterResponse {
  bool success = false;
  std::string message;
  std::string parameter_server_address;
  int32_t total_workers = 0;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.str(2, message), w.str(3, parameter_server_address);
    w.i32(4, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kLen) message = std::string(r.bytes());
      else if (f == 3 && wt == kLen) parameter_server_address = std::string(r.bytes());
      else if (f == 4 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

enum WorkerStatus { IDLE = 0, TRAINING = 1, CHECKPOINTING = 2, ERROR = 3 };

struct HeartbeatRequest {
  int32_t worker_id = 0;
  int32_t status = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, status);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) status = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct HeartbeatResponse {
  bool success = false;
  int64_t timestamp = 0;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.i64(2, timestamp);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kVarint) timestamp = static_cast<int64_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct Empty {  // ListWorkersRequest, GetPSAddressRequest
  std::string encode() const { return {}; }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) r.skip(wt);
    return r.ok();
  }
};
using ListWorkersRequest = Empty;
using GetPSAddressRequest = Empty;

struct ListWorkersResponse {
  std::vector<WorkerInfo> workers;
  int32_t total_workers = 0;
  std::string encode() const {
    Writer w;
    for (const auto& x : workers) w.msg(1, x.encode());
    w.i32(2, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) {
        WorkerInfo wi;
        if (!wi.decode(r.bytes())) return false;
        workers.push_back(std::move(wi));
      } else if (f == 2 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct GetPSAddressResponse {
  std::string address;
  int32_t port = 0;
  std::string encode() const {
    Writer w;
    w.str(1, address), w.i32(2, port);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) address = std::string(r.bytes());
      else if (f == 2 && wt == kVarint) port = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

}  // namespace pb
}  // namespace psb

#include "pb_ps.h"
