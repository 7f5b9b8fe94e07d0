// grpc_lite.cpp — gRPC unary server / client over h2::Conn (see h2.h).
//
// Protocol (gRPC over HTTP/2): request = HEADERS(:method POST, :path /pkg.Svc/Method,
// content-type application/grpc, te trailers) + DATA(1-byte compressed flag, 4-byte big-endian
// length, message) with END_STREAM; response = HEADERS(:status 200, content-type) + DATA +
// trailing HEADERS(grpc-status, grpc-message) with END_STREAM.
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <chrono>
#include <cstring>
#include <deque>

#include "h2.h"

namespace psb {
namespace grpc_lite {

namespace {

// Locates the single message of a unary body (a view into `body`).  False on framing errors.
bool unframe_message(const std::string& body, std::string_view& msg, std::string& err) {
  if (body.size() < 5) {
    err = "missing gRPC message frame";
    return false;
  }
  if (body[0] != 0) {
    err = "compressed messages are not supported (grpc-encoding must be identity)";
    return false;
  }
  const uint8_t* p = reinterpret_cast<const uint8_t*>(body.data());
  uint32_t n = (uint32_t(p[1]) << 24) | (uint32_t(p[2]) << 16) | (uint32_t(p[3]) << 8) | p[4];
  if (body.size() != 5 + static_cast<size_t>(n)) {
    err = "gRPC message length does not match the body";
    return false;
  }
  msg = std::string_view(body.data() + 5, n);
  return true;
}

const std::string* find_header(const std::vector<hpack::Header>& hs, const char* name) {
  for (const auto& h : hs)
    if (h.first == name) return &h.second;
  return nullptr;
}

std::string percent_encode(const std::string& s) {
  static const char* hex = "0123456789ABCDEF";
  std::string out;
  for (unsigned char c : s) {
    if (c >= 0x20 && c <= 0x7e && c != '%') out.push_back(static_cast<char>(c));
    else {
      out.push_back('%');
      out.push_back(hex[c >> 4]);
      out.push_back(hex[c & 15]);
    }
  }
  return out;
}

void set_timeouts(int fd, int ms) {
  timeval tv{ms / 1000, (ms % 1000) * 1000};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
}

}  // namespace

bool split_host_port(const std::string& s, std::string& host, int& port) {
  size_t c = s.rfind(':');
  if (c == std::string::npos) {
    host = s;
    return false;
  }
  host = s.substr(0, c);
  try {
    port = std::stoi(s.substr(c + 1));
  } catch (...) {
    return false;
  }
  return true;
}

// ------------------------------------------------------------------------------ HandlerPool
// Request handlers run on pooled threads: waking an idle thread costs a futex, creating one per
// RPC costs a clone + stack mapping + teardown.  Handlers may block for long (Barrier, the
// long-polled status queries), so the pool grows whenever no thread is idle — it is bounded by
// what the connections admit (kMaxConcurrentStreams each), not by a fixed size that could
// deadlock a barrier — and idle threads retire after a few seconds.
class HandlerPool : public std::enable_shared_from_this<HandlerPool> {
 public:
  void submit(std::function<void()> fn) {
    std::unique_lock<std::mutex> lk(mu_);
    q_.push_back(std::move(fn));
    // One thread per queued task: a task may block until a LATER task has run (two Barrier
    // calls of the same barrier), so it must never wait behind another one.  An idle thread
    // that was notified is still counted in idle_ until it wakes up — hence the comparison
    // with the queue length instead of `idle_ > 0`.
    if (q_.size() <= static_cast<size_t>(idle_)) {
      lk.unlock();
      cv_.notify_one();
      return;
    }
    ++threads_;
    lk.unlock();
    auto self = shared_from_this();
    std::thread([self] { self->run(); }).detach();
  }
  int threads() {
    std::lock_guard<std::mutex> lk(mu_);
    return threads_;
  }

 private:
  void run() {
    std::unique_lock<std::mutex> lk(mu_);
    while (true) {
      if (q_.empty()) {
        ++idle_;
        const bool got = cv_.wait_for(lk, std::chrono::seconds(5), [&] { return !q_.empty(); });
        --idle_;
        if (!got) break;
      }
      std::function<void()> fn = std::move(q_.front());
      q_.pop_front();
      lk.unlock();
      fn();
      fn = nullptr;
      lk.lock();
    }
    --threads_;
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::function<void()>> q_;
  int idle_ = 0, threads_ = 0;
};

// ------------------------------------------------------------------------------ Server
Server::Server() : pool_(std::make_shared<HandlerPool>()) {
  // grpc.health.v1.Health/Check — what load balancers, `grpc_health_probe` and Kubernetes gRPC
  // probes call.  Every server built on this class answers SERVING (HealthCheckResponse
  // {status = 1}) for any service name while it is accepting requests.
  add("/grpc.health.v1.Health/Check", [](std::string_view, Reply& reply) {
    reply.body.assign("\x08\x01", 2);
    return Status{};
  });
}
Server::~Server() { shutdown(); }

void Server::add(const std::string& path, Handler h) { handlers_[path] = std::move(h); }

int Server::listen(const std::string& address) {
  std::string host;
  int port = 0;
  if (!split_host_port(address, host, port)) return -1;
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  hints.ai_flags = AI_PASSIVE;
  const char* node = (host.empty() || host == "0.0.0.0" || host == "[::]") ? nullptr : host.c_str();
  if (getaddrinfo(node, std::to_string(port).c_str(), &hints, &res) != 0 || !res) return -1;
  int fd = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
  if (fd < 0) {
    freeaddrinfo(res);
    return -1;
  }
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  if (::bind(fd, res->ai_addr, res->ai_addrlen) != 0 || ::listen(fd, 128) != 0) {
    ::close(fd);
    freeaddrinfo(res);
    return -1;
  }
  freeaddrinfo(res);
  sockaddr_in sa{};
  socklen_t sl = sizeof(sa);
  getsockname(fd, reinterpret_cast<sockaddr*>(&sa), &sl);
  port_ = ntohs(sa.sin_port);
  listen_fd_ = fd;
  return port_;
}

void Server::start() {
  running_ = true;
  accept_thread_ = std::thread(&Server::accept_loop, this);
}

void Server::serve_forever() {
  running_ = true;
  accept_loop();
}

void Server::shutdown() {
  bool was = running_.exchange(false);
  const int lfd = listen_fd_.load();
  if (lfd >= 0) ::shutdown(lfd, SHUT_RDWR);  // wakes poll(); the fd is closed after the join
  {
    std::lock_guard<std::mutex> lk(conns_mu_);
    for (int fd : conn_fds_) ::shutdown(fd, SHUT_RDWR);
  }
  if (accept_thread_.joinable()) accept_thread_.join();
  if (lfd >= 0 && listen_fd_.exchange(-1) == lfd) ::close(lfd);
  if (was)
    for (int i = 0; i < 500 && active_.load() > 0; ++i)
      std::this_thread::sleep_for(std::chrono::milliseconds(10));
}

void Server::accept_loop() {
  while (running_) {
    const int lfd = listen_fd_.load();
    if (lfd < 0) break;
    pollfd pfd{lfd, POLLIN, 0};
    int pr = ::poll(&pfd, 1, 200);
    if (pr <= 0) continue;
    int fd = ::accept(lfd, nullptr, nullptr);
    if (fd < 0) {
      if (!running_) break;
      continue;
    }
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    {
      std::lock_guard<std::mutex> lk(conns_mu_);
      conn_fds_.push_back(fd);
    }
    active_++;
    std::thread([this, fd] {
      serve_conn(fd);
      {
        std::lock_guard<std::mutex> lk(conns_mu_);
        for (auto it = conn_fds_.begin(); it != conn_fds_.end(); ++it)
          if (*it == fd) {
            conn_fds_.erase(it);
            break;
          }
      }
      active_--;
    }).detach();
  }
}

void Server::serve_conn(int fd) {
  // The reader (this thread) only pumps frames; every completed request runs on its own
  // pooled handler thread so RPCs multiplexed on one connection (gRPC shares connections
  // between channels of a process) never wait for each other — e.g. two Barrier calls.
  auto conn = std::make_shared<h2::Conn>(fd, /*is_server=*/true, /*threaded=*/true);
  if (!conn->handshake()) return;
  auto inflight = std::make_shared<std::atomic<int>>(0);
  while (running_ && conn->pump()) {
    h2::Stream req;
    while (conn->take_completed(req)) {
      if (req.reset) {
        conn->erase(req.id);
        continue;
      }
      if (inflight->load() >= (int)h2::kMaxConcurrentStreams) {  // above what we advertised
        conn->send_rst(req.id, 7 /*REFUSED_STREAM*/);
        conn->erase(req.id);
        req = h2::Stream();
        continue;
      }
      auto st = std::make_shared<h2::Stream>(std::move(req));
      req = h2::Stream();
      (*inflight)++;
      pool_->submit([this, conn, st, inflight] {
        const uint32_t sid = st->id;
        const std::string* path = find_header(st->headers, ":path");
        const std::string* ctype = find_header(st->headers, "content-type");
        Status status;
        Reply response;
        if (!path || !ctype || ctype->rfind("application/grpc", 0) != 0) {
          status = {UNKNOWN, "not a gRPC request"};
        } else {
          auto it = handlers_.find(*path);
          std::string_view request;
          std::string err;
          if (it == handlers_.end()) status = {UNIMPLEMENTED, "unknown method " + *path};
          else if (!unframe_message(st->body, request, err)) status = {INTERNAL, err};
          else {
            try {
              status = it->second(request, response);
            } catch (const std::exception& e) {
              status = {INTERNAL, e.what()};
            }
          }
        }
        std::vector<hpack::Header> trailers = {{"grpc-status", std::to_string(status.code)}};
        if (!status.message.empty())
          trailers.push_back({"grpc-message", percent_encode(status.message)});
        bool ok;
        if (status.ok()) {  // HEADERS + DATA + trailers leave in one write when small
          const std::vector<hpack::Header> hs = {{":status", "200"},
                                                 {"content-type", "application/grpc"}};
          ok = conn->send_message(sid, &hs, response.body, &trailers, /*grpc_prefix=*/true,
                                  response.shared_tail.get());
        } else {  // trailers-only response
          std::vector<hpack::Header> hs = {{":status", "200"},
                                           {"content-type", "application/grpc"}};
          hs.insert(hs.end(), trailers.begin(), trailers.end());
          ok = conn->send_headers(sid, hs, true);
        }
        conn->erase(sid);
        if (!ok) conn->mark_dead();
        (*inflight)--;
      });
    }
  }
  conn->mark_dead();
  ::shutdown(fd, SHUT_RDWR);  // wakes handler threads blocked in send()
  for (int i = 0; i < 1000 && inflight->load() > 0; ++i)
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
}

// ------------------------------------------------------------------------------ Channel
Channel::Channel(const std::string& target) : port_(443) {
  // A port-less target dials 443 exactly like gRPC does; callers repair
  // host-only addresses before they get here.
  if (!split_host_port(target, host_, port_)) {
    host_ = target;
    port_ = 443;
  }
}

Channel::~Channel() {
  std::lock_guard<std::mutex> lk(mu_);
  close_locked();
}

void Channel::close_locked() { conn_.reset(); }

bool Channel::connect_locked(int timeout_ms) {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(host_.c_str(), std::to_string(port_).c_str(), &hints, &res) != 0 || !res)
    return false;
  int fd = -1;
  for (addrinfo* a = res; a; a = a->ai_next) {
    fd = ::socket(a->ai_family, a->ai_socktype, a->ai_protocol);
    if (fd < 0) continue;
    set_timeouts(fd, timeout_ms);
    if (::connect(fd, a->ai_addr, a->ai_addrlen) == 0) break;
    ::close(fd);
    fd = -1;
  }
  freeaddrinfo(res);
  if (fd < 0) return false;
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  conn_.reset(new h2::Conn(fd, /*is_server=*/false));
  if (!conn_->handshake()) {
    conn_.reset();
    return false;
  }
  return true;
}

Status Channel::call(const std::string& path, const std::string& request, std::string& response,
                     int timeout_ms) {
  std::lock_guard<std::mutex> lk(mu_);
  for (int attempt = 0; attempt < 2; ++attempt) {
    const bool fresh = !conn_ || !conn_->alive();
    if (fresh && !connect_locked(timeout_ms))
      return {UNAVAILABLE, "failed to connect to " + host_ + ":" + std::to_string(port_)};
    if (fresh || timeout_ms != socket_timeout_ms_) {  // two setsockopt calls saved per RPC
      set_timeouts(conn_->fd(), timeout_ms);
      socket_timeout_ms_ = timeout_ms;
    }
    h2::Conn& c = *conn_;
    const uint32_t sid = c.next_stream_id();
    c.open_stream(sid);
    std::vector<hpack::Header> hs = {
        {":method", "POST"}, {":scheme", "http"}, {":path", path},
        {":authority", host_ + ":" + std::to_string(port_)},
        {"content-type", "application/grpc"}, {"te", "trailers"},
        {"user-agent", "psb-grpc-lite/0.1"}};
    bool sent = c.send_message(sid, &hs, request, nullptr, /*grpc_prefix=*/true);
    bool done = false;
    h2::Stream got, tmp;
    h2::Stream* st = &got;
    while (sent && !done && c.pump()) {
      while (c.take_completed(tmp)) {
        if (tmp.id == sid) {
          got = std::move(tmp);
          done = true;
        } else {
          c.erase(tmp.id);
        }
        tmp = h2::Stream();
      }
    }
    if (!done) {
      close_locked();
      if (!fresh && attempt == 0) continue;  // stale pooled connection: retry once on a new one
      return {UNAVAILABLE, "connection to " + host_ + ":" + std::to_string(port_) + " lost"};
    }
    Status status;
    if (st->reset) {
      status = {UNAVAILABLE, "stream reset by peer"};
    } else {
      const std::vector<hpack::Header>& tr = st->trailers.empty() ? st->headers : st->trailers;
      const std::string* gs = find_header(tr, "grpc-status");
      const std::string* gm = find_header(tr, "grpc-message");
      if (!gs) status = {UNKNOWN, "response without grpc-status"};
      else status = {std::atoi(gs->c_str()), gm ? *gm : std::string()};
      if (status.ok()) {
        std::string err;
        std::string_view msg;
        if (!unframe_message(st->body, msg, err)) {
          status = {INTERNAL, err};
        } else {  // hand the receive buffer over; dropping the 5-byte header is a warm memmove
          response = std::move(st->body);
          response.erase(0, 5);
        }
      }
    }
    c.erase(sid);
    return status;
  }
  return {UNAVAILABLE, "unreachable"};
}

}  // namespace grpc_lite
}  // namespace psb
