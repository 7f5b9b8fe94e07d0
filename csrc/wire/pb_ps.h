// pb_ps.h — messages of /root/reference/proto/parameter_server.proto:13-77 (included by pb.h).
#pragma once

#include "pb.h"

namespace psb {
namespace pb {

struct Tensor {
  std::string name;
  std::vector<int32_t> shape;
  std::vector<float> data;
  int32_t dtype = 0;  // 0 = float32, 1 = float64 (payload is float32 either way)

  // Exact number of bytes encode_into() appends: lets a message with many large tensors be
  // written into one buffer reserved once (no per-tensor staging copy, no growth copies).
  size_t encoded_size() const {
    size_t n = 0;
    if (!name.empty()) n += 1 + varint_size(name.size()) + name.size();
    if (!shape.empty()) {
      size_t packed = 0;
      for (int32_t x : shape) packed += varint_size(static_cast<uint64_t>(static_cast<int64_t>(x)));
      n += 1 + varint_size(packed) + packed;
    }
    if (!data.empty()) n += 1 + varint_size(data.size() * 4) + data.size() * 4;
    if (dtype != 0) n += 1 + varint_size(static_cast<uint64_t>(static_cast<int64_t>(dtype)));
    return n;
  }
  void encode_into(Writer& w) const {
    w.str(1, name);
    w.packed_i32(2, shape);
    w.packed_f32(3, data.data(), data.size());
    w.i32(4, dtype);
  }
  std::string encode() const {
    Writer w;
    w.buf.reserve(data.size() * 4 + name.size() + 32);
    encode_into(w);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) {
        name = std::string(r.bytes());
      } else if (f == 2 && wt == kLen) {
        Reader in(r.bytes());
        while (in.ok()) {
          uint64_t v = in.varint();
          if (!in.ok()) break;
          shape.push_back(static_cast<int32_t>(v));
        }
      } else if (f == 2 && wt == kVarint) {
        shape.push_back(static_cast<int32_t>(r.varint()));
      } else if (f == 3 && wt == kLen) {
        std::string_view b = r.bytes();
        if (b.size() % 4) return false;
        size_t old = data.size();
        data.resize(old + b.size() / 4);
        memcpy(data.data() + old, b.data(), b.size());
      } else if (f == 3 && wt == kFixed32) {
        uint32_t u = r.fixed32();
        float x;
        memcpy(&x, &u, 4);
        data.push_back(x);
      } else if (f == 4 && wt == kVarint) {
        dtype = static_cast<int32_t>(r.varint());
      } else {
        r.skip(wt);
      }
    }
    return r.ok();
  }
};

inline size_t tensors_encoded_size(const std::vector<Tensor>& ts) {
  size_t total = 0;
  for (const auto& t : ts) {
    const size_t n = t.encoded_size();
    total += 1 + varint_size(n) + n;  // field numbers here are < 16: one tag byte
  }
  return total;
}

inline void encode_tensors(Writer& w, int field, const std::vector<Tensor>& ts) {
  w.buf.reserve(w.buf.size() + tensors_encoded_size(ts) + 32);
  for (const auto& t : ts) {
    w.tag(field, kLen);  // embedded message: always written, even when empty
    w.varint(t.encoded_size());
    t.encode_into(w);
  }
}

struct GradientUpdate {
  int32_t worker_id = 0;
  int32_t iteration = 0;
  std::vector<Tensor> gradients;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, iteration);
    encode_tensors(w, 3, gradients);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 3 && wt == kLen) {
        Tensor t;
        if (!t.decode(r.bytes())) return false;
        gradients.push_back(std::move(t));
      } else r.skip(wt);
    }
    return r.ok();
  }
};

struct PushResponse {
  bool success = false;
  std::string message;
  int32_t iteration = 0;
  bool aggregation_complete = false;
  int32_t workers_received = 0;
  int32_t total_workers = 0;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.str(2, message), w.i32(3, iteration);
    w.boolean(4, aggregation_complete), w.i32(5, workers_received), w.i32(6, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kLen) message = std::string(r.bytes());
      else if (f == 3 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kVarint) aggregation_complete = r.varint() != 0;
      else if (f == 5 && wt == kVarint) workers_received = static_cast<int32_t>(r.varint());
      else if (f == 6 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct PullRequest {
  int32_t worker_id = 0;
  int32_t iteration = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, worker_id), w.i32(2, iteration);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) worker_id = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct ParameterUpdate {
  int32_t iteration = 0;
  std::vector<Tensor> parameters;
  bool ready = false;
  std::string encode() const {
    Writer w;
    w.i32(1, iteration);
    encode_tensors(w, 2, parameters);
    w.boolean(3, ready);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kLen) {
        Tensor t;
        if (!t.decode(r.bytes())) return false;
        parameters.push_back(std::move(t));
      } else if (f == 3 && wt == kVarint) ready = r.varint() != 0;
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct SyncStatusRequest {
  int32_t iteration = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, iteration);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct SyncStatusResponse {
  int32_t iteration = 0;
  bool ready = false;
  int32_t workers_received = 0;
  int32_t total_workers = 0;
  std::string encode() const {
    Writer w;
    w.i32(1, iteration), w.boolean(2, ready), w.i32(3, workers_received), w.i32(4, total_workers);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) iteration = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kVarint) ready = r.varint() != 0;
      else if (f == 3 && wt == kVarint) workers_received = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kVarint) total_workers = static_cast<int32_t>(r.varint());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct SaveCheckpointRequest {
  int32_t epoch = 0;
  std::string path;
  std::string encode() const {
    Writer w;
    w.i32(1, epoch), w.str(2, path);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) epoch = static_cast<int32_t>(r.varint());
      else if (f == 2 && wt == kLen) path = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct SaveCheckpointResponse {
  bool success = false;
  std::string message;
  std::string checkpoint_path;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.str(2, message), w.str(3, checkpoint_path);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kLen) message = std::string(r.bytes());
      else if (f == 3 && wt == kLen) checkpoint_path = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct LoadCheckpointRequest {
  std::string path;
  std::string encode() const {
    Writer w;
    w.str(1, path);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kLen) path = std::string(r.bytes());
      else r.skip(wt);
    }
    return r.ok();
  }
};

struct LoadCheckpointResponse {
  bool success = false;
  std::string message;
  int32_t epoch = 0;
  std::vector<Tensor> parameters;
  std::string encode() const {
    Writer w;
    w.boolean(1, success), w.str(2, message), w.i32(3, epoch);
    encode_tensors(w, 4, parameters);
    return std::move(w.buf);
  }
  bool decode(std::string_view s) {
    Reader r(s);
    int f, wt;
    while (r.next(f, wt)) {
      if (f == 1 && wt == kVarint) success = r.varint() != 0;
      else if (f == 2 && wt == kLen) message = std::string(r.bytes());
      else if (f == 3 && wt == kVarint) epoch = static_cast<int32_t>(r.varint());
      else if (f == 4 && wt == kLen) {
        Tensor t;
        if (!t.decode(r.bytes())) return false;
        parameters.push_back(std::move(t));
      } else r.skip(wt);
    }
    return r.ok();
  }
};

}  // namespace pb
}  // namespace psb
