// pb_ext.h — messages of proto/psb_ext.proto (package psb, service CoordinatorExt).
// No counterpart in the reference: these carry the north-star additions.
#pragma once

#include "pb.h"

namespace psb {
namespace pb {

struct ShardInfo {
  int32_t shard_id = 0;
  std::string address;
  int32_t device = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, shard_id), w.str(2, address), w.i32(3, device);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) shard_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kLen) address = std::string(r.bytes());
      else if (f == 3 && wt == kVarint) device = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct ShardMapResponse {
  std::vector<ShardInfo> shards;
  int32_t total_workers = 0;
  std::string encode() const {
    Writer w;
    for (const auto& s : shards) w.msg(1, s.encode());
    w.i32(2, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) {
        ShardInfo x;
        if (!x.decode(r.bytes())) return false;
        shards.push_back(std::move(x));
      } else if (f == 2 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct BarrierRequest {
  std::string name;
  int32_t worker_id = 0, generation = 0, expected = 0;
  int64_t timeout_ms = 0;
  std::string encode() const {
    Writer w;
    w.str(1, name), w.i32(2, worker_id), w.i32(3, generation), w.i32(4, expected);
    w.i64(5, timeout_ms);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) name = std::string(r.bytes());
      else if (f == 2 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 3 && wt == kVarint) generation = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kVarint) expected = static_cast<int32_t>(r.varint());
      else if (f == 5 && wt == kVarint) timeout_ms = static_cast<int64_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct BarrierResponse {
  bool released = false;
  int32_t arrived = 0;
  std::string encode() const {
    Writer w;
    w.boolean(1, released), w.i32(2, arrived);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) released = r.varint() != 0;
      else if (f == 2 && wt == kVarint) arrived = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct ClockRequest {
  int32_t worker_id = 0, clock = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, clock);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) clock = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct ClockResponse {
  int32_t min_clock = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, min_clock);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) min_clock = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct HandleBlob {
  std::string key;
  int32_t rank = 0;
  std::string blob;
  std::string encode() const {
    Writer w;
    w.str(1, key), w.i32(2, rank), w.str(3, blob);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) key = std::string(r.bytes());
      else if (f == 2 && wt == kVarint) rank = static_cast<int32_t>(r.varint());
      else if (f == 3 && wt == kLen) blob = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct HandleQuery {
  std::string key;
  std::string encode() const {
    Writer w;
    w.str(1, key);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) key = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct HandleList {
  std::vector<HandleBlob> handles;
  std::string encode() const {
    Writer w;
    for (const auto& h : handles) w.msg(1, h.encode());
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) {
        HandleBlob h;
        if (!h.decode(r.bytes())) return false;
        handles.push_back(std::move(h));
      } else r.skip(wt);
    }
    return r.ok();
  }
};

struct Ack {
  bool success = false;
  std::string encode() const {
    Writer w;
    w.boolean(1, success);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else r.skip(wt);
    }
    return r.ok();
  }
};

// ---- sparse embedding tables (service psb.ParameterServerExt; no counterpart in the reference,
//      whose Tensor is dense: proto/parameter_server.proto:19-24)
namespace detail {
// packed or unpacked `repeated float` / `repeated sfixed64` (proto3 decoders accept both)
inline bool read_f32s(Reader& r, int wt, std::vector<float>& out) {
  if (wt == kLen) {
    std::string_view b = r.bytes();
    if (!r.ok() || b.size() % 4) return false;
    const size_t at = out.size();
    out.resize(at + b.size() / 4);
    if (!b.empty()) memcpy(out.data() + at, b.data(), b.size());
    return true;
  }
  if (wt != kFixed32) return false;
  uint32_t u = r.fixed32();
  float f;
  memcpy(&f, &u, 4);
  out.push_back(f);
  return r.ok();
}
inline bool read_i64s(Reader& r, int wt, std::vector<int64_t>& out) {
  if (wt == kLen) {
    std::string_view b = r.bytes();
    if (!r.ok() || b.size() % 8) return false;
    const size_t at = out.size();
    out.resize(at + b.size() / 8);
    if (!b.empty()) memcpy(out.data() + at, b.data(), b.size());
    return true;
  }
  if (wt != kFixed64) return false;
  out.push_back(static_cast<int64_t>(r.fixed64()));
  return r.ok();
}
}  // namespace detail

struct TableSpec {
  std::string name;
  int64_t rows = 0;
  int32_t dim = 0;
  float init_std = 0.0f;
  int64_t seed = 0;
  std::string encode() const {
    Writer w;
    w.str(1, name), w.i64(2, rows), w.i32(3, dim), w.f32(4, init_std), w.i64(5, seed);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) name = std::string(r.bytes());
      else if (f == 2 && wt == kVarint) rows = static_cast<int64_t>(r.varint());
      else if (f == 3 && wt == kVarint) dim = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kFixed32) {
        uint32_t u = r.fixed32();
        memcpy(&init_std, &u, 4);
      } else if (f == 5 && wt == kVarint) seed = static_cast<int64_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct SparseRows {
  std::string name;
  int32_t dim = 0;
  std::vector<int64_t> indices;
  std::vector<float> values;
  std::string encode() const {
    Writer w;
    w.str(1, name), w.i32(2, dim);
    w.packed_sfixed64(3, indices.data(), indices.size());
    w.packed_f32(4, values.data(), values.size());
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) name = std::string(r.bytes());
      else if (f == 2 && wt == kVarint) dim = static_cast<int32_t>(r.varint());
      else if (f == 3 && (wt == kLen || wt == kFixed64)) {
        if (!detail::read_i64s(r, wt, indices)) return false;
      } else if (f == 4 && (wt == kLen || wt == kFixed32)) {
        if (!detail::read_f32s(r, wt, values)) return false;
      } else r.skip(wt);
    }
    return r.ok();
  }
};

struct SparsePush {
  int32_t worker_id = 0, iteration = 0;
  std::vector<SparseRows> tables;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, iteration);
    for (const auto& t : tables) w.msg(3, t.encode());
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 3 && wt == kLen) {
        SparseRows t;
        if (!t.decode(r.bytes())) return false;
        tables.push_back(std::move(t));
      } else r.skip(wt);
    }
    return r.ok();
  }
};

struct SparsePushResponse {
  bool success = false;
  std::string message;
  bool aggregation_complete = false;
  int32_t workers_received = 0, total_workers = 0;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.str(2, message), w.boolean(3, aggregation_complete);
    w.i32(4, workers_received), w.i32(5, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kLen) message = std::string(r.bytes());
      else if (f == 3 && wt == kVarint) aggregation_complete = r.varint() != 0;
      else if (f == 4 && wt == kVarint) workers_received = static_cast<int32_t>(r.varint());
      else if (f == 5 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct RowsRequest {
  int32_t worker_id = 0, iteration = 0;
  std::string name;
  std::vector<int64_t> indices;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, iteration), w.str(3, name);
    w.packed_sfixed64(4, indices.data(), indices.size());
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 3 && wt == kLen) name = std::string(r.bytes());
      else if (f == 4 && (wt == kLen || wt == kFixed64)) {
        if (!detail::read_i64s(r, wt, indices)) return false;
      } else r.skip(wt);
    }
    return r.ok();
  }
};

struct RowsResponse {
  bool ready = false;
  int32_t dim = 0;
  std::vector<float> values;
  std::string message;
  std::string encode() const {
    Writer w;
    w.boolean(1, ready), w.i32(2, dim);
    w.packed_f32(3, values.data(), values.size());
    w.str(4, message);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) ready = r.varint() != 0;
      else if (f == 2 && wt == kVarint) dim = static_cast<int32_t>(r.varint());
      else if (f == 3 && (wt == kLen || wt == kFixed32)) {
        if (!detail::read_f32s(r, wt, values)) return false;
      } else if (f == 4 && wt == kLen) message = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

}  // namespace pb
}  // namespace psb
