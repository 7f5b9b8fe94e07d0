// h2.h — minimal HTTP/2 connection (RFC 7540) + gRPC unary transport on top of it.
//
// Wire-compatible stand-in for the gRPC C++ library the reference links
// (/root/reference/CMakeLists.txt:20, src/*_service.cpp, src/worker.cpp:141-314) and which is
// not installable on the target box.  Scope: cleartext h2 with prior knowledge (what
// grpc::InsecureChannelCredentials / InsecureServerCredentials speak), unary calls, identity
// encoding, full flow control in both directions, PING / SETTINGS / GOAWAY / RST_STREAM.
// Unlike the reference (4 MiB default receive limit never raised — SURVEY D7) message size is
// bounded only by `max_message_bytes`.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <string_view>
#include <thread>
#include <vector>

#include "hpack.h"

namespace psb {
namespace h2 {

enum FrameType : uint8_t {
  DATA = 0, HEADERS = 1, PRIORITY = 2, RST_STREAM = 3, SETTINGS = 4, PUSH_PROMISE = 5, PING = 6,
  GOAWAY = 7, WINDOW_UPDATE = 8, CONTINUATION = 9
};
enum Flags : uint8_t { END_STREAM = 0x1, ACK = 0x1, END_HEADERS = 0x4, PADDED = 0x8, PRIO = 0x20 };

// Limits a server advertises / enforces: any peer on the port could otherwise make it spawn
// handler threads or buffer header fragments without bound.
constexpr uint32_t kMaxConcurrentStreams = 128;  // SETTINGS_MAX_CONCURRENT_STREAMS
constexpr size_t kMaxHeaderBlock = 64 * 1024;    // HEADERS + CONTINUATION fragments of one block

struct Stream {
  uint32_t id = 0;
  std::vector<hpack::Header> headers;   // request headers (server) / response headers (client)
  std::vector<hpack::Header> trailers;  // client side: trailing HEADERS
  std::string body;
  bool headers_done = false;
  bool remote_closed = false;
  bool reset = false;
  int64_t send_window = 65535;
  int64_t recv_unacked = 0;  // DATA bytes received and not yet returned by a WINDOW_UPDATE
};

// One HTTP/2 connection over a connected TCP socket.
// Threading: exactly one thread calls pump() (the reader).  In `threaded` mode other threads may
// call send_headers / send_data / erase concurrently (the server runs one handler thread per
// request so RPCs multiplexed on one connection do not serialise); a sender blocked on flow
// control then waits for the reader to process WINDOW_UPDATE.  In non-threaded mode (client
// channel) the single owner thread pumps frames itself while blocked.
class Conn {
 public:
  Conn(int fd, bool is_server, bool threaded = false);
  ~Conn();
  bool handshake();  // preface + SETTINGS exchange start
  // Reads and processes exactly one frame.  Returns false when the connection is finished.
  bool pump();
  // Copy of a stream that became complete (remote END_STREAM) since the last call; false if none.
  bool take_completed(Stream& out);
  void erase(uint32_t id);
  bool send_headers(uint32_t sid, const std::vector<hpack::Header>& h, bool end_stream);
  // Sends `data` as DATA frames, honouring flow control (pumps incoming frames while blocked).
  bool send_data(uint32_t sid, const std::string& data, bool end_stream);
  bool send_data(uint32_t sid, const char* data, size_t size, bool end_stream);
  // One unary message in ONE socket write when it is small and the flow-control windows allow:
  // [HEADERS(hs)] + DATA(data) + [HEADERS(trailers), END_STREAM]; `hs` may be null (headers
  // already sent), `trailers` null means END_STREAM goes on the last DATA frame.  Three writes
  // become one segment — a unary RPC is latency-bound by syscalls and wake-ups, not by bytes.
  // Falls back to send_headers / send_data for large bodies.
  // `grpc_prefix`: put the 5-byte gRPC message header (flag 0 + big-endian length) in front of
  // the body here, so that callers never build a framed copy of a large message; `tail` is a
  // second piece of the same body (e.g. a cached encoding shared between responses).
  bool send_message(uint32_t sid, const std::vector<hpack::Header>* hs, const std::string& data,
                    const std::vector<hpack::Header>* trailers, bool grpc_prefix = false,
                    const std::string* tail = nullptr);
  bool send_rst(uint32_t sid, uint32_t code);
  bool send_goaway(uint32_t code);
  uint32_t next_stream_id() { uint32_t s = next_sid_; next_sid_ += 2; return s; }
  void open_stream(uint32_t sid);
  bool alive() const { return alive_.load(); }
  void mark_dead();
  int fd() const { return fd_; }
  size_t max_message_bytes = (size_t)1 << 31;

 private:
  bool read_exact(void* buf, size_t n);
  bool write_all(const void* buf, size_t n);
  bool write_frame(uint8_t type, uint8_t flags, uint32_t sid, const void* payload, size_t n);
  bool write_frame_locked(uint8_t type, uint8_t flags, uint32_t sid, const void* payload,
                          size_t n);  // write_mu_ held
  bool fail();
  void append_frame(std::string& out, uint8_t type, uint8_t flags, uint32_t sid, const char* payload,
                    size_t n);
  bool on_headers_block(uint32_t sid, uint8_t flags);
  bool on_data(uint32_t sid, uint8_t flags, size_t len);
  bool on_settings(const std::string& p, uint8_t flags);
  Stream* find(uint32_t id);  // state_mu_ held

  int fd_;
  bool server_;
  bool threaded_;
  std::atomic<bool> alive_{true};
  bool goaway_ = false;
  std::mutex state_mu_;   // streams_, windows, completed_ (lock order: state_mu_ -> write_mu_)
  std::mutex write_mu_;   // socket writes; a header block is written under one hold
  std::condition_variable window_cv_;
  hpack::Decoder dec_;
  hpack::Encoder enc_;
  std::map<uint32_t, Stream> streams_;
  std::vector<uint32_t> completed_;
  std::string hdr_block_;
  uint32_t hdr_sid_ = 0;
  uint8_t hdr_flags_ = 0;
  int64_t conn_send_window_ = 65535;
  int64_t peer_initial_window_ = 65535;
  uint32_t peer_max_frame_ = 16384;
  uint32_t next_sid_ = 1;
  uint32_t last_peer_sid_ = 0;
  int64_t conn_recv_unacked_ = 0;  // reader thread only
  // reader-side buffer: one recv() usually brings every frame of a small response
  std::vector<uint8_t> rbuf_;
  size_t rpos_ = 0, rlen_ = 0;
};

}  // namespace h2

namespace grpc_lite {

// gRPC status codes used here
enum Code { OK = 0, UNKNOWN = 2, DEADLINE_EXCEEDED = 4, UNIMPLEMENTED = 12, INTERNAL = 13,
            UNAVAILABLE = 14 };

struct Status {
  int code = OK;
  std::string message;
  bool ok() const { return code == OK; }
};

// What a handler returns: the encoded response message, optionally continued by a shared,
// immutable piece (`shared_tail`) that is sent straight from that buffer — protobuf fields may
// arrive in any order, so "small fields, then the big cached repeated field" is one message.
struct Reply {
  std::string body;
  std::shared_ptr<const std::string> shared_tail;
  Reply& operator=(std::string encoded) {
    body = std::move(encoded);
    return *this;
  }
};

// `request` is a view into the connection's receive buffer (valid for the call).
using Handler = std::function<Status(std::string_view request, Reply& response)>;

class HandlerPool;  // grpc_lite.cpp

class Server {
 public:
  Server();
  ~Server();
  void add(const std::string& path, Handler h);  // path = "/package.Service/Method"
  // Binds and listens ("host:port"; port 0 picks a free one).  Returns the bound port or -1.
  int listen(const std::string& address);
  void serve_forever();  // blocks until shutdown()
  void start();          // serve on a background thread
  void shutdown();
  int port() const { return port_; }

 private:
  void accept_loop();
  void serve_conn(int fd);
  std::map<std::string, Handler> handlers_;
  std::atomic<int> listen_fd_{-1};
  int port_ = -1;
  std::atomic<bool> running_{false};
  std::thread accept_thread_;
  std::mutex conns_mu_;
  std::vector<int> conn_fds_;
  std::atomic<int> active_{0};
  std::shared_ptr<HandlerPool> pool_;
};

// Client channel: one persistent connection, one unary call at a time, reconnects on failure.
class Channel {
 public:
  explicit Channel(const std::string& target);  // "host:port" (host-only => port 443 like gRPC)
  ~Channel();
  Status call(const std::string& path, const std::string& request, std::string& response,
              int timeout_ms = 30000);

 private:
  bool connect_locked(int timeout_ms);
  void close_locked();
  std::string host_;
  int port_;
  std::mutex mu_;
  std::unique_ptr<h2::Conn> conn_;
  int socket_timeout_ms_ = -1;
};

// helpers shared by binaries
bool split_host_port(const std::string& s, std::string& host, int& port);

}  // namespace grpc_lite
}  // namespace psb
