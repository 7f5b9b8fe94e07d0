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

}  // namespace pb
}  // namespace psb
