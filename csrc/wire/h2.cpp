// h2.cpp — HTTP/2 connection state machine (see h2.h).
#include "h2.h"

#include <errno.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cstring>

namespace psb {
namespace h2 {

namespace {
const char kPreface[] = "PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n";
constexpr size_t kPrefaceLen = 24;
constexpr uint32_t kOurWindow = 0x7fffffff;  // we never want to be the flow-control bottleneck
constexpr uint32_t kOurMaxFrame = 1 << 20;
// Both receive windows are opened to 2^31-1 at the handshake; they are replenished lazily, once
// this much has been consumed, instead of after every DATA frame (two extra frames per RPC in
// each direction for no benefit at that window size).
constexpr int64_t kWindowUpdateThreshold = 1 << 26;
constexpr size_t kReadBuf = 64 * 1024;
constexpr size_t kCoalesceMax = 256 * 1024;

void put32(std::string& s, uint32_t v) {
  char b[4] = {static_cast<char>(v >> 24), static_cast<char>(v >> 16), static_cast<char>(v >> 8),
               static_cast<char>(v)};
  s.append(b, 4);
}
uint32_t get32(const uint8_t* p) {
  return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3];
}
}  // namespace

Conn::Conn(int fd, bool is_server, bool threaded)
    : fd_(fd), server_(is_server), threaded_(threaded) {
  next_sid_ = is_server ? 2 : 1;
  rbuf_.resize(kReadBuf);
}

void Conn::mark_dead() {
  alive_ = false;
  window_cv_.notify_all();
}

Conn::~Conn() {
  if (fd_ >= 0) ::close(fd_);
}

bool Conn::read_exact(void* buf, size_t n) {
  uint8_t* p = static_cast<uint8_t*>(buf);
  while (n > 0) {
    if (rpos_ < rlen_) {  // serve from the buffer first
      size_t k = rlen_ - rpos_;
      if (k > n) k = n;
      memcpy(p, rbuf_.data() + rpos_, k);
      rpos_ += k;
      p += k;
      n -= k;
      continue;
    }
    // large remainder: straight into the caller's memory; otherwise refill the buffer
    const bool direct = n >= rbuf_.size();
    ssize_t r = direct ? ::recv(fd_, p, n, 0) : ::recv(fd_, rbuf_.data(), rbuf_.size(), 0);
    if (r == 0) { mark_dead(); return false; }
    if (r < 0) {
      if (errno == EINTR) continue;
      mark_dead();  // includes SO_RCVTIMEO expiry
      return false;
    }
    if (direct) {
      p += r;
      n -= static_cast<size_t>(r);
    } else {
      rpos_ = 0;
      rlen_ = static_cast<size_t>(r);
    }
  }
  return true;
}

bool Conn::write_all(const void* buf, size_t n) {
  const uint8_t* p = static_cast<const uint8_t*>(buf);
  while (n > 0) {
    ssize_t r = ::send(fd_, p, n, MSG_NOSIGNAL);
    if (r < 0) {
      if (errno == EINTR) continue;
      mark_dead();
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

bool Conn::write_frame(uint8_t type, uint8_t flags, uint32_t sid, const void* payload, size_t n) {
  std::lock_guard<std::mutex> wl(write_mu_);
  return write_frame_locked(type, flags, sid, payload, n);
}

bool Conn::write_frame_locked(uint8_t type, uint8_t flags, uint32_t sid, const void* payload,
                              size_t n) {
  uint8_t h[9] = {static_cast<uint8_t>(n >> 16), static_cast<uint8_t>(n >> 8),
                  static_cast<uint8_t>(n), type, flags,
                  static_cast<uint8_t>((sid >> 24) & 0x7f), static_cast<uint8_t>(sid >> 16),
                  static_cast<uint8_t>(sid >> 8), static_cast<uint8_t>(sid)};
  if (n <= 1024) {  // one syscall for small frames
    uint8_t buf[9 + 1024];
    memcpy(buf, h, 9);
    if (n) memcpy(buf + 9, payload, n);
    return write_all(buf, 9 + n);
  }
  return write_all(h, 9) && write_all(payload, n);
}

bool Conn::handshake() {
  if (server_) {
    char pre[kPrefaceLen];
    if (!read_exact(pre, kPrefaceLen) || memcmp(pre, kPreface, kPrefaceLen) != 0) {
      mark_dead();
      return false;
    }
  } else {
    if (!write_all(kPreface, kPrefaceLen)) return false;
  }
  std::string s;
  auto setting = [&](uint16_t id, uint32_t v) {
    s.push_back(static_cast<char>(id >> 8));
    s.push_back(static_cast<char>(id));
    put32(s, v);
  };
  if (!server_) setting(2, 0);  // ENABLE_PUSH = 0
  else setting(3, kMaxConcurrentStreams);  // MAX_CONCURRENT_STREAMS: bounds the handler threads
  setting(4, kOurWindow);       // INITIAL_WINDOW_SIZE
  setting(5, kOurMaxFrame);     // MAX_FRAME_SIZE
  if (!write_frame(SETTINGS, 0, 0, s.data(), s.size())) return false;
  std::string wu;
  put32(wu, kOurWindow - 65535);  // open the connection-level receive window as well
  return write_frame(WINDOW_UPDATE, 0, 0, wu.data(), wu.size());
}

void Conn::open_stream(uint32_t sid) {
  std::lock_guard<std::mutex> lk(state_mu_);
  Stream& st = streams_[sid];
  st.id = sid;
  st.send_window = peer_initial_window_;
}

Stream* Conn::find(uint32_t id) {
  auto it = streams_.find(id);
  return it == streams_.end() ? nullptr : &it->second;
}

void Conn::erase(uint32_t id) {
  std::lock_guard<std::mutex> lk(state_mu_);
  streams_.erase(id);
}

bool Conn::take_completed(Stream& out) {
  std::lock_guard<std::mutex> lk(state_mu_);
  while (!completed_.empty()) {
    uint32_t id = completed_.front();
    completed_.erase(completed_.begin());
    if (Stream* s = find(id)) {
      out.id = s->id;
      out.headers = std::move(s->headers);
      out.trailers = std::move(s->trailers);
      out.body = std::move(s->body);
      out.reset = s->reset;
      out.remote_closed = s->remote_closed;
      return true;
    }
  }
  return false;
}

bool Conn::fail() {
  mark_dead();
  return false;
}

// state_mu_ held by the caller (pump)
bool Conn::on_settings(const std::string& p, uint8_t flags) {
  if (flags & ACK) return true;
  if (p.size() % 6) return fail();
  const uint8_t* d = reinterpret_cast<const uint8_t*>(p.data());
  for (size_t i = 0; i + 6 <= p.size(); i += 6) {
    uint16_t id = static_cast<uint16_t>((d[i] << 8) | d[i + 1]);
    uint32_t v = get32(d + i + 2);
    if (id == 4) {  // INITIAL_WINDOW_SIZE: adjusts every open stream by the delta
      int64_t delta = static_cast<int64_t>(v) - peer_initial_window_;
      peer_initial_window_ = v;
      for (auto& kv : streams_) kv.second.send_window += delta;
      window_cv_.notify_all();
    } else if (id == 5) {
      if (v >= 16384 && v <= 16777215) peer_max_frame_ = v;
    }
    // HEADER_TABLE_SIZE (1): our encoder never indexes, nothing to do.
  }
  return write_frame(SETTINGS, ACK, 0, nullptr, 0);
}

// state_mu_ held by the caller (pump)
bool Conn::on_headers_block(uint32_t sid, uint8_t flags) {
  std::vector<hpack::Header> hs;
  if (!dec_.decode(hdr_block_, hs)) {
    send_goaway(9 /*COMPRESSION_ERROR*/);
    return fail();
  }
  hdr_block_.clear();
  Stream* st = find(sid);
  if (!st) {
    if (!server_) return true;  // response for a stream we already dropped
    st = &streams_[sid];
    st->id = sid;
    st->send_window = peer_initial_window_;
    last_peer_sid_ = sid;
  }
  if (!st->headers_done) {
    st->headers = std::move(hs);
    st->headers_done = true;
  } else {
    st->trailers = std::move(hs);
  }
  if (flags & END_STREAM) {
    st->remote_closed = true;
    completed_.push_back(sid);
  }
  return true;
}

// An unpadded DATA frame: the payload is received straight into the stream's body — no
// intermediate frame buffer, and the body is sized once from the gRPC message header at its
// start, so a 400 MB tensor message is written exactly once on its way in.  Only the reader
// thread touches the body of a stream that is still open (take_completed moves it out after
// END_STREAM), so the socket read happens without state_mu_.
bool Conn::on_data(uint32_t sid, uint8_t flags, size_t len) {
  Stream* st = nullptr;
  {
    std::lock_guard<std::mutex> lk(state_mu_);
    if (hdr_sid_ != 0) {  // header block must be contiguous
      send_goaway(1);
      return fail();
    }
    st = find(sid);
    if (st && st->reset) st = nullptr;
    if (st && st->body.size() + len > max_message_bytes + 5) {
      send_rst(sid, 11 /*ENHANCE_YOUR_CALM*/);
      st->reset = true;
      st = nullptr;
    }
  }
  if (st) {
    std::string& body = st->body;
    const size_t old = body.size();
    body.resize(old + len);
    if (len && !read_exact(&body[old], len)) return false;
    if (old < 5 && body.size() >= 5) {
      const uint8_t* b = reinterpret_cast<const uint8_t*>(body.data());
      const size_t want = 5 + static_cast<size_t>(get32(b + 1));
      if (want > body.size() && want <= max_message_bytes + 5) body.reserve(want);
    }
  } else if (len) {  // nobody wants it: drain
    char scratch[16384];
    for (size_t left = len; left > 0;) {
      const size_t k = left < sizeof(scratch) ? left : sizeof(scratch);
      if (!read_exact(scratch, k)) return false;
      left -= k;
    }
  }
  uint32_t stream_inc = 0, conn_inc = 0;
  {
    std::lock_guard<std::mutex> lk(state_mu_);
    st = find(sid);  // may have been erased meanwhile only if it was not ours to fill
    if (st && !(flags & END_STREAM)) {
      st->recv_unacked += static_cast<int64_t>(len);
      if (st->recv_unacked >= kWindowUpdateThreshold) {
        stream_inc = static_cast<uint32_t>(st->recv_unacked);
        st->recv_unacked = 0;
      }
    }
    if (st && (flags & END_STREAM)) {
      st->remote_closed = true;
      completed_.push_back(sid);
    }
    conn_recv_unacked_ += static_cast<int64_t>(len);
    if (conn_recv_unacked_ >= kWindowUpdateThreshold) {
      conn_inc = static_cast<uint32_t>(conn_recv_unacked_);
      conn_recv_unacked_ = 0;
    }
  }
  // The window updates go out WITHOUT state_mu_: a peer that stops reading can then stall this
  // reader, but not the handler threads that need the lock to send their responses.
  std::string wu;
  if (conn_inc) {
    put32(wu, conn_inc);
    if (!write_frame(WINDOW_UPDATE, 0, 0, wu.data(), 4)) return false;
  }
  if (stream_inc) {
    wu.clear();
    put32(wu, stream_inc);
    if (!write_frame(WINDOW_UPDATE, 0, sid, wu.data(), 4)) return false;
  }
  return true;
}

bool Conn::pump() {
  if (!alive_) return false;
  uint8_t h[9];
  if (!read_exact(h, 9)) return false;
  const size_t len = (size_t(h[0]) << 16) | (size_t(h[1]) << 8) | h[2];
  const uint8_t type = h[3], flags = h[4];
  const uint32_t sid = get32(h + 5) & 0x7fffffffu;
  if (len > kOurMaxFrame + 1024) {
    send_goaway(6 /*FRAME_SIZE_ERROR*/);
    return fail();
  }
  if (type == DATA && !(flags & PADDED)) return on_data(sid, flags, len);
  std::string p(len, '\0');
  if (len && !read_exact(&p[0], len)) return false;
  const uint8_t* d = reinterpret_cast<const uint8_t*>(p.data());

  std::unique_lock<std::mutex> lk(state_mu_);
  if (hdr_sid_ != 0 && type != CONTINUATION) {  // header block must be contiguous
    send_goaway(1);
    return fail();
  }
  switch (type) {
    case DATA: {
      size_t off = 0, pad = 0;
      if (flags & PADDED) {
        if (len < 1) return fail();
        pad = d[0];
        off = 1;
      }
      if (off + pad > len) return fail();
      const size_t n = len - off - pad;
      Stream* st = find(sid);
      if (st && !st->reset) {
        if (st->body.size() + n > max_message_bytes + 5) {
          send_rst(sid, 11 /*ENHANCE_YOUR_CALM*/);
          st->reset = true;
        } else {
          st->body.append(p, off, n);
        }
      }
      // Replenish the windows by what was consumed — lazily (kWindowUpdateThreshold).
      uint32_t stream_inc = 0, conn_inc = 0;
      if (st && !(flags & END_STREAM)) {
        st->recv_unacked += static_cast<int64_t>(len);
        if (st->recv_unacked >= kWindowUpdateThreshold) {
          stream_inc = static_cast<uint32_t>(st->recv_unacked);
          st->recv_unacked = 0;
        }
      }
      if (st && (flags & END_STREAM)) {
        st->remote_closed = true;
        completed_.push_back(sid);
      }
      conn_recv_unacked_ += static_cast<int64_t>(len);
      if (conn_recv_unacked_ >= kWindowUpdateThreshold) {
        conn_inc = static_cast<uint32_t>(conn_recv_unacked_);
        conn_recv_unacked_ = 0;
      }
      if (conn_inc || stream_inc) {
        // The socket write happens WITHOUT state_mu_: a peer that stops reading can then stall
        // this reader, but not the handler threads that need the lock to send their responses.
        lk.unlock();
        std::string wu;
        if (conn_inc) {
          put32(wu, conn_inc);
          if (!write_frame(WINDOW_UPDATE, 0, 0, wu.data(), 4)) return false;
        }
        if (stream_inc) {
          wu.clear();
          put32(wu, stream_inc);
          if (!write_frame(WINDOW_UPDATE, 0, sid, wu.data(), 4)) return false;
        }
      }
      break;
    }
    case HEADERS: {
      size_t off = 0, pad = 0;
      if (flags & PADDED) {
        if (len < 1) return fail();
        pad = d[0];
        off = 1;
      }
      if (flags & PRIO) off += 5;
      if (off + pad > len) return fail();
      hdr_block_.assign(p, off, len - off - pad);
      if (hdr_block_.size() > kMaxHeaderBlock) {
        send_goaway(11 /*ENHANCE_YOUR_CALM*/);
        return fail();
      }
      if (flags & END_HEADERS) return on_headers_block(sid, flags);
      hdr_sid_ = sid;
      hdr_flags_ = flags;
      break;
    }
    case CONTINUATION: {
      if (hdr_sid_ != sid) return fail();
      hdr_block_.append(p);
      if (hdr_block_.size() > kMaxHeaderBlock) {  // CONTINUATION frames are not a memory reservoir
        send_goaway(11 /*ENHANCE_YOUR_CALM*/);
        return fail();
      }
      if (flags & END_HEADERS) {
        uint8_t f = hdr_flags_;
        hdr_sid_ = 0;
        return on_headers_block(sid, f);
      }
      break;
    }
    case SETTINGS: return on_settings(p, flags);
    case PING:
      if (!(flags & ACK)) return write_frame(PING, ACK, 0, p.data(), p.size());
      break;
    case WINDOW_UPDATE: {
      if (len != 4) return fail();
      const uint32_t inc = get32(d) & 0x7fffffffu;
      if (sid == 0) conn_send_window_ += inc;
      else if (Stream* st = find(sid)) st->send_window += inc;
      window_cv_.notify_all();
      break;
    }
    case RST_STREAM:
      if (Stream* st = find(sid)) {
        st->reset = true;
        st->remote_closed = true;
        completed_.push_back(sid);
        window_cv_.notify_all();
      }
      break;
    case GOAWAY: goaway_ = true; break;
    case PRIORITY:
    case PUSH_PROMISE:
    default: break;  // ignored (unknown frame types MUST be ignored)
  }
  return true;
}

bool Conn::send_headers(uint32_t sid, const std::vector<hpack::Header>& hs, bool end_stream) {
  std::string block;
  enc_.encode(hs, block);  // stateless encoder (no dynamic table): safe from any thread
  uint32_t max_frame;
  {
    std::lock_guard<std::mutex> lk(state_mu_);
    max_frame = peer_max_frame_;
  }
  std::lock_guard<std::mutex> wl(write_mu_);  // HEADERS + CONTINUATION must not be interleaved
  size_t off = 0;
  bool first = true;
  do {
    size_t n = block.size() - off;
    if (n > max_frame) n = max_frame;
    const bool last = off + n == block.size();
    uint8_t flags = last ? END_HEADERS : 0;
    if (first && end_stream) flags |= END_STREAM;
    if (!write_frame_locked(first ? HEADERS : CONTINUATION, flags, sid, block.data() + off, n))
      return false;
    off += n;
    first = false;
  } while (off < block.size());
  return true;
}

bool Conn::send_data(uint32_t sid, const std::string& data, bool end_stream) {
  return send_data(sid, data.data(), data.size(), end_stream);
}

bool Conn::send_data(uint32_t sid, const char* data, size_t size, bool end_stream) {
  size_t off = 0;
  if (size == 0) return write_frame(DATA, end_stream ? END_STREAM : 0, sid, nullptr, 0);
  while (off < size) {
    size_t n = 0;
    {
      std::unique_lock<std::mutex> lk(state_mu_);
      while (true) {
        if (!alive_) return false;
        Stream* st = find(sid);
        if (!st || st->reset) return false;
        int64_t win = conn_send_window_ < st->send_window ? conn_send_window_ : st->send_window;
        if (win > 0) {
          n = size - off;
          if (n > static_cast<size_t>(win)) n = static_cast<size_t>(win);
          if (n > peer_max_frame_) n = peer_max_frame_;
          conn_send_window_ -= static_cast<int64_t>(n);
          st->send_window -= static_cast<int64_t>(n);
          break;
        }
        // blocked on flow control until a WINDOW_UPDATE is processed
        if (threaded_) {
          window_cv_.wait(lk);
        } else {
          lk.unlock();
          if (!pump()) return false;
          lk.lock();
        }
      }
    }
    const bool last = off + n == size;
    if (!write_frame(DATA, (last && end_stream) ? END_STREAM : 0, sid, data + off, n))
      return false;
    off += n;
  }
  return true;
}

void Conn::append_frame(std::string& out, uint8_t type, uint8_t flags, uint32_t sid,
                        const char* payload, size_t n) {
  const char h[9] = {static_cast<char>(n >> 16), static_cast<char>(n >> 8), static_cast<char>(n),
                     static_cast<char>(type), static_cast<char>(flags),
                     static_cast<char>((sid >> 24) & 0x7f), static_cast<char>(sid >> 16),
                     static_cast<char>(sid >> 8), static_cast<char>(sid)};
  out.append(h, 9);
  if (n) out.append(payload, n);
}

bool Conn::send_message(uint32_t sid, const std::vector<hpack::Header>* hs, const std::string& data,
                        const std::vector<hpack::Header>* trailers, bool grpc_prefix,
                        const std::string* tail) {
  std::string hblock, tblock;
  if (hs) enc_.encode(*hs, hblock);
  if (trailers) enc_.encode(*trailers, tblock);
  const size_t tail_n = tail ? tail->size() : 0;
  const size_t msg_n = data.size() + tail_n;
  char prefix[5] = {0, static_cast<char>(msg_n >> 24), static_cast<char>(msg_n >> 16),
                    static_cast<char>(msg_n >> 8), static_cast<char>(msg_n)};
  const size_t prefix_n = grpc_prefix ? 5 : 0;
  const size_t body_n = prefix_n + msg_n;
  uint32_t max_frame = 0;
  bool coalesce = body_n <= kCoalesceMax;
  if (coalesce) {
    std::lock_guard<std::mutex> lk(state_mu_);
    max_frame = peer_max_frame_;
    Stream* st = find(sid);
    const int64_t need = static_cast<int64_t>(body_n);
    coalesce = st && !st->reset && hblock.size() <= max_frame && tblock.size() <= max_frame &&
               conn_send_window_ >= need && st->send_window >= need;
    if (coalesce) {  // reserve the whole body now
      conn_send_window_ -= need;
      st->send_window -= need;
    }
  }
  if (!coalesce) {
    // large body: frames go out straight from the caller's buffers, piece by piece
    if (hs && !send_headers(sid, *hs, false)) return false;
    const bool end = trailers == nullptr;
    if (prefix_n && !send_data(sid, prefix, prefix_n, end && msg_n == 0)) return false;
    if ((!data.empty() || (!prefix_n && !tail_n)) &&
        !send_data(sid, data.data(), data.size(), end && tail_n == 0))
      return false;
    if (tail_n && !send_data(sid, tail->data(), tail_n, end)) return false;
    return trailers ? send_headers(sid, *trailers, true) : true;
  }
  std::string out;
  out.reserve(hblock.size() + body_n + tblock.size() + 64);
  if (hs) append_frame(out, HEADERS, END_HEADERS, sid, hblock.data(), hblock.size());
  std::string body;
  const std::string* src = &data;
  if (prefix_n || tail_n) {  // small: one contiguous body is simplest
    body.reserve(body_n);
    body.append(prefix, prefix_n);
    body.append(data);
    if (tail_n) body.append(*tail);
    src = &body;
  }
  size_t off = 0;
  do {
    size_t n = src->size() - off;
    if (n > max_frame) n = max_frame;
    const bool last = off + n == src->size();
    append_frame(out, DATA, (last && !trailers) ? END_STREAM : 0, sid, src->data() + off, n);
    off += n;
  } while (off < src->size());
  if (trailers)
    append_frame(out, HEADERS, END_HEADERS | END_STREAM, sid, tblock.data(), tblock.size());
  std::lock_guard<std::mutex> wl(write_mu_);
  return write_all(out.data(), out.size());
}

bool Conn::send_rst(uint32_t sid, uint32_t code) {
  std::string p;
  put32(p, code);
  return write_frame(RST_STREAM, 0, sid, p.data(), 4);
}

bool Conn::send_goaway(uint32_t code) {
  std::string p;
  put32(p, last_peer_sid_);
  put32(p, code);
  return write_frame(GOAWAY, 0, 0, p.data(), p.size());
}

}  // namespace h2
}  // namespace psb
