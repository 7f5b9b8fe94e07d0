// client.cpp — libpsb_control.so: native coordinator client (C ABI in include/psb_control.h).
//
// Counterpart of the coordinator-facing half of the reference worker
// (/root/reference/src/worker.cpp:108-238): discovery, registration with back-off, a heartbeat
// thread — here also re-registration after an eviction, a membership watch that raises the
// data plane's abort word when a peer disappears, and the extension RPCs (shard map, barriers,
// blob exchange) through which the GPU ranks of a job bootstrap their symmetric heap.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/psb_control.h"
#include "../wire/h2.h"
#include "../wire/pb.h"
#include "../wire/pb_ext.h"

using psb::grpc_lite::Channel;
using psb::grpc_lite::Status;
namespace pb = psb::pb;

namespace {

thread_local char g_err[512] = "";
void set_err(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int fail(int code, const char* what, const Status& s) {
  set_err("%s: grpc status %d %s", what, s.code, s.message.c_str());
  return code;
}

const char* kCoord = "/coordinator.Coordinator/";
const char* kExt = "/psb.CoordinatorExt/";

}  // namespace

struct PsbCoord {
  std::string address;
  pb::WorkerInfo self;
  int heartbeat_ms = 5000;
  std::unique_ptr<Channel> ch;     // calls of the owning thread(s), serialised by the channel
  std::unique_ptr<Channel> hb_ch;  // the heartbeat thread's own connection
  std::string ps_address;
  std::thread hb;
  std::mutex hb_mu;
  std::condition_variable hb_cv;
  bool running = true;
  std::atomic<int32_t> status{pb::IDLE};
  std::atomic<int32_t> live{-1};
  std::atomic<int32_t> evictions{0};
  std::atomic<int32_t> expected_world{0};
  std::atomic<volatile uint32_t*> abort_word{nullptr};
  std::atomic<bool> peer_lost{false};
  std::mutex seq_mu;
  std::map<std::string, int32_t> barrier_gen;
  std::map<std::string, int64_t> gather_seq;

  bool do_register(Channel& on, int timeout_ms, std::string* ps_out) {
    std::string resp;
    Status s = on.call(std::string(kCoord) + "RegisterWorker", self.encode(), resp, timeout_ms);
    pb::RegisterResponse r;
    if (!s.ok() || !r.decode(resp) || !r.success) return false;
    if (ps_out && !r.parameter_server_address.empty()) *ps_out = r.parameter_server_address;
    return true;
  }

  void raise_abort() {
    peer_lost.store(true);
    volatile uint32_t* w = abort_word.load();
    if (w) *w = 1u;
  }

  void heartbeat_once() {
    pb::HeartbeatRequest req;
    req.worker_id = self.worker_id;
    req.status = status.load();
    std::string resp;
    Status s = hb_ch->call(std::string(kCoord) + "Heartbeat", req.encode(), resp, 3000);
    pb::HeartbeatResponse r;
    if (s.ok() && r.decode(resp) && !r.success) {
      // the coordinator no longer knows this worker: it was evicted (or the coordinator
      // restarted).  The reference ignores the reply (src/worker.cpp:227-228); here the worker
      // registers again and the data plane is told, because the peers have already seen the
      // world shrink.
      evictions.fetch_add(1);
      do_register(*hb_ch, 3000, nullptr);
      if (expected_world.load() > 0) raise_abort();
    }
    std::string lresp;
    s = hb_ch->call(std::string(kCoord) + "ListWorkers", std::string(), lresp, 3000);
    pb::ListWorkersResponse lw;
    if (s.ok() && lw.decode(lresp)) {
      const int32_t n = static_cast<int32_t>(lw.workers.size());
      live.store(n);
      const int32_t want = expected_world.load();
      if (want > 0 && n < want) raise_abort();
    }
  }

  void heartbeat_loop() {
    std::unique_lock<std::mutex> lk(hb_mu);
    while (running) {
      lk.unlock();
      heartbeat_once();
      lk.lock();
      if (!running) break;
      hb_cv.wait_for(lk, std::chrono::milliseconds(heartbeat_ms), [this] { return !running; });
    }
  }
};

extern "C" {

const char* psbc_last_error(void) { return g_err; }

PsbCoord* psbc_connect(const char* coordinator_address, int32_t worker_id, const char* self_address,
                       int32_t self_port, const char* hostname, int32_t heartbeat_ms,
                       int32_t connect_timeout_ms) {
  if (!coordinator_address || !*coordinator_address) {
    set_err("psbc_connect: empty coordinator address");
    return nullptr;
  }
  std::unique_ptr<PsbCoord> c(new PsbCoord);
  c->address = coordinator_address;
  c->self.worker_id = worker_id;
  c->self.address = self_address ? self_address : "";
  c->self.port = self_port;
  c->self.hostname = hostname ? hostname : "";
  c->heartbeat_ms = heartbeat_ms > 0 ? heartbeat_ms : 5000;
  c->ch.reset(new Channel(c->address));
  c->hb_ch.reset(new Channel(c->address));

  const auto deadline = std::chrono::steady_clock::now() +
                        std::chrono::milliseconds(connect_timeout_ms > 0 ? connect_timeout_ms : 10000);
  int backoff = 50;
  bool ok = false;
  std::string host;
  int port = 0;
  while (true) {
    // discovery, then registration; the register reply may carry a dialable host:port or the
    // host part only (reference coordinator) — in that case keep the discovered port
    std::string resp;
    Status s = c->ch->call(std::string(kCoord) + "GetParameterServerAddress", std::string(), resp, 3000);
    pb::GetPSAddressResponse ps;
    if (s.ok() && ps.decode(resp)) {
      host = ps.address, port = ps.port;
      std::string h2;
      int p2 = 0;
      if (psb::grpc_lite::split_host_port(host, h2, p2)) host = h2, port = port ? port : p2;
      std::string from_reg;
      if (c->do_register(*c->ch, 3000, &from_reg)) {
        std::string h3;
        int p3 = 0;
        if (!from_reg.empty()) {
          if (psb::grpc_lite::split_host_port(from_reg, h3, p3)) host = h3, port = p3;
          else host = from_reg;
        }
        ok = true;
        break;
      }
    }
    if (std::chrono::steady_clock::now() >= deadline) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(backoff));
    backoff = backoff < 800 ? backoff * 2 : 800;
  }
  if (!ok) {
    set_err("psbc_connect: coordinator %s did not answer within %d ms", coordinator_address,
            connect_timeout_ms);
    return nullptr;
  }
  c->ps_address = host + ":" + std::to_string(port);
  PsbCoord* raw = c.release();
  raw->hb = std::thread(&PsbCoord::heartbeat_loop, raw);
  return raw;
}

void psbc_close(PsbCoord* c) {
  if (!c) return;
  {
    std::lock_guard<std::mutex> lk(c->hb_mu);
    c->running = false;
  }
  c->hb_cv.notify_all();
  if (c->hb.joinable()) c->hb.join();
  delete c;
}

const char* psbc_ps_address(PsbCoord* c) { return c ? c->ps_address.c_str() : ""; }

int psbc_set_status(PsbCoord* c, int32_t status) {
  if (!c) return PSBC_EINVAL;
  c->status.store(status);
  return 0;
}

int psbc_list_workers(PsbCoord* c, int32_t* ids, int32_t cap, int32_t* count) {
  if (!c || !count) return PSBC_EINVAL;
  std::string resp;
  Status s = c->ch->call(std::string(kCoord) + "ListWorkers", std::string(), resp, 5000);
  pb::ListWorkersResponse lw;
  if (!s.ok() || !lw.decode(resp)) return fail(PSBC_EUNAVAILABLE, "ListWorkers", s);
  *count = static_cast<int32_t>(lw.workers.size());
  for (int32_t i = 0; i < *count && i < cap && ids; ++i) ids[i] = lw.workers[i].worker_id;
  return 0;
}

int psbc_live_workers(PsbCoord* c) { return c ? c->live.load() : -1; }
int psbc_evictions(PsbCoord* c) { return c ? c->evictions.load() : 0; }

int psbc_watch(PsbCoord* c, int32_t expected_world, volatile uint32_t* abort_word) {
  if (!c || expected_world < 0) return PSBC_EINVAL;
  c->abort_word.store(abort_word);
  c->expected_world.store(expected_world);
  if (expected_world == 0) c->peer_lost.store(false);
  return 0;
}

int psbc_peer_lost(PsbCoord* c) { return c && c->peer_lost.load() ? 1 : 0; }

int psbc_register_shard(PsbCoord* c, int32_t shard_id, const char* address, int32_t device) {
  if (!c) return PSBC_EINVAL;
  pb::ShardInfo in;
  in.shard_id = shard_id;
  in.address = address ? address : "";
  in.device = device + 1;  // wire convention: GPU ordinal + 1, 0 = host shard
  std::string resp;
  Status s = c->ch->call(std::string(kExt) + "RegisterShard", in.encode(), resp, 5000);
  pb::Ack ack;
  if (!s.ok() || !ack.decode(resp) || !ack.success) return fail(PSBC_EUNAVAILABLE, "RegisterShard", s);
  return 0;
}

int psbc_shard_map(PsbCoord* c, int32_t* shard_ids, int32_t* devices, int32_t cap, char* addr_buf,
                   size_t addr_cap, int32_t* count, int32_t* total_workers) {
  if (!c || !count) return PSBC_EINVAL;
  std::string resp;
  Status s = c->ch->call(std::string(kExt) + "GetShardMap", std::string(), resp, 5000);
  pb::ShardMapResponse m;
  if (!s.ok() || !m.decode(resp)) return fail(PSBC_EUNAVAILABLE, "GetShardMap", s);
  *count = static_cast<int32_t>(m.shards.size());
  if (total_workers) *total_workers = m.total_workers;
  size_t used = 0;
  for (int32_t i = 0; i < *count && i < cap; ++i) {
    if (shard_ids) shard_ids[i] = m.shards[i].shard_id;
    if (devices) devices[i] = m.shards[i].device - 1;
    if (addr_buf) {
      const std::string& a = m.shards[i].address;
      if (used + a.size() + 1 > addr_cap) {
        set_err("psbc_shard_map: address buffer of %zu bytes too small", addr_cap);
        return PSBC_ERANGE;
      }
      memcpy(addr_buf + used, a.c_str(), a.size() + 1);
      used += a.size() + 1;
    }
  }
  return 0;
}

int psbc_barrier(PsbCoord* c, const char* name, int32_t expected, int32_t timeout_ms) {
  if (!c || !name) return PSBC_EINVAL;
  int32_t gen;
  {
    std::lock_guard<std::mutex> lk(c->seq_mu);
    gen = c->barrier_gen[name]++;
  }
  if (timeout_ms <= 0) timeout_ms = 300000;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int transport_failures = 0;
  while (true) {
    const auto left = std::chrono::duration_cast<std::chrono::milliseconds>(
                          deadline - std::chrono::steady_clock::now()).count();
    if (left <= 0) break;
    pb::BarrierRequest in;
    in.name = name;
    in.worker_id = c->self.worker_id;
    in.generation = gen;
    in.expected = expected;
    in.timeout_ms = left < 2000 ? left : 2000;  // wait server-side in slices, re-arrive after each
    std::string resp;
    Status s = c->ch->call(std::string(kExt) + "Barrier", in.encode(), resp, (int)in.timeout_ms + 5000);
    pb::BarrierResponse out;
    if (!s.ok() || !out.decode(resp)) {
      if (++transport_failures > 5) return fail(PSBC_EUNAVAILABLE, "Barrier", s);
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
      continue;
    }
    if (out.released) return 0;
  }
  set_err("psbc_barrier(%s): not released within %d ms", name, timeout_ms);
  return PSBC_ETIMEOUT;
}

int psbc_put(PsbCoord* c, const char* key, const void* blob, size_t n) {
  if (!c || !key || (n && !blob)) return PSBC_EINVAL;
  pb::HandleBlob in;
  in.key = key;
  in.rank = c->self.worker_id;
  in.blob.assign(static_cast<const char*>(blob), n);
  std::string resp;
  Status s = c->ch->call(std::string(kExt) + "PutHandle", in.encode(), resp, 10000);
  pb::Ack ack;
  if (!s.ok() || !ack.decode(resp) || !ack.success) return fail(PSBC_EUNAVAILABLE, "PutHandle", s);
  return 0;
}

static int get_all(PsbCoord* c, const std::string& key, pb::HandleList& out) {
  pb::HandleQuery q;
  q.key = key;
  std::string resp;
  Status s = c->ch->call(std::string(kExt) + "GetHandles", q.encode(), resp, 10000);
  if (!s.ok() || !out.decode(resp)) return fail(PSBC_EUNAVAILABLE, "GetHandles", s);
  return 0;
}

int psbc_get(PsbCoord* c, const char* key, int32_t rank, void* out, size_t cap, size_t* n) {
  if (!c || !key || !n) return PSBC_EINVAL;
  pb::HandleList l;
  if (int rc = get_all(c, key, l)) return rc;
  *n = 0;
  for (const auto& h : l.handles) {
    if (h.rank != rank) continue;
    *n = h.blob.size();
    if (h.blob.size() > cap) {
      set_err("psbc_get: blob of %zu bytes exceeds the buffer (%zu)", h.blob.size(), cap);
      return PSBC_ERANGE;
    }
    if (out) memcpy(out, h.blob.data(), h.blob.size());
  }
  return 0;
}

int psbc_allgather(PsbCoord* c, const char* key, const void* blob, size_t n, int32_t world,
                   void* out, int32_t timeout_ms) {
  if (!c || !key || !blob || !out || n == 0 || world < 1) return PSBC_EINVAL;
  int64_t seq;
  {
    std::lock_guard<std::mutex> lk(c->seq_mu);
    seq = c->gather_seq[key]++;
  }
  const std::string k = std::string(key) + "#" + std::to_string(seq);
  if (int rc = psbc_put(c, k.c_str(), blob, n)) return rc;
  // every rank has put its piece when the barrier opens — and has finished reading the
  // previous round of this key, whose entry can therefore be dropped
  if (int rc = psbc_barrier(c, (std::string("ag:") + key).c_str(), world, timeout_ms)) return rc;
  if (seq > 0) psbc_put(c, (std::string(key) + "#" + std::to_string(seq - 1)).c_str(), nullptr, 0);
  pb::HandleList l;
  if (int rc = get_all(c, k, l)) return rc;
  std::vector<bool> have(world, false);
  for (const auto& h : l.handles) {
    if (h.rank < 0 || h.rank >= world) continue;
    if (h.blob.size() != n) {
      set_err("psbc_allgather(%s): rank %d contributed %zu bytes, expected %zu", key, h.rank,
              h.blob.size(), n);
      return PSBC_EINVAL;
    }
    memcpy(static_cast<char*>(out) + (size_t)h.rank * n, h.blob.data(), n);
    have[h.rank] = true;
  }
  for (int32_t r = 0; r < world; ++r)
    if (!have[r]) {
      set_err("psbc_allgather(%s): contribution of rank %d missing", key, r);
      return PSBC_ETIMEOUT;
    }
  return 0;
}

int psbc_update_clock(PsbCoord* c, int32_t clock, int32_t* min_clock) {
  if (!c) return PSBC_EINVAL;
  pb::ClockRequest in;
  in.worker_id = c->self.worker_id;
  in.clock = clock;
  std::string resp;
  Status s = c->ch->call(std::string(kExt) + "UpdateClock", in.encode(), resp, 5000);
  pb::ClockResponse out;
  if (!s.ok() || !out.decode(resp)) return fail(PSBC_EUNAVAILABLE, "UpdateClock", s);
  if (min_clock) *min_clock = out.min_clock;
  return 0;
}

}  // extern "C"
