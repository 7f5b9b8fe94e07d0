// hpack.cpp — see hpack.h.
#include "hpack.h"

#include <mutex>

#include "hpack_huffman_table.h"

namespace psb {
namespace hpack {

namespace {

const Header kStatic[61] = {
    {":authority", ""}, {":method", "GET"}, {":method", "POST"}, {":path", "/"},
    {":path", "/index.html"}, {":scheme", "http"}, {":scheme", "https"}, {":status", "200"},
    {":status", "204"}, {":status", "206"}, {":status", "304"}, {":status", "400"},
    {":status", "404"}, {":status", "500"}, {"accept-charset", ""},
    {"accept-encoding", "gzip, deflate"}, {"accept-language", ""}, {"accept-ranges", ""},
    {"accept", ""}, {"access-control-allow-origin", ""}, {"age", ""}, {"allow", ""},
    {"authorization", ""}, {"cache-control", ""}, {"content-disposition", ""},
    {"content-encoding", ""}, {"content-language", ""}, {"content-length", ""},
    {"content-location", ""}, {"content-range", ""}, {"content-type", ""}, {"cookie", ""},
    {"date", ""}, {"etag", ""}, {"expect", ""}, {"expires", ""}, {"from", ""}, {"host", ""},
    {"if-match", ""}, {"if-modified-since", ""}, {"if-none-match", ""}, {"if-range", ""},
    {"if-unmodified-since", ""}, {"last-modified", ""}, {"link", ""}, {"location", ""},
    {"max-forwards", ""}, {"proxy-authenticate", ""}, {"proxy-authorization", ""}, {"range", ""},
    {"referer", ""}, {"refresh", ""}, {"retry-after", ""}, {"server", ""}, {"set-cookie", ""},
    {"strict-transport-security", ""}, {"transfer-encoding", ""}, {"user-agent", ""},
    {"vary", ""}, {"via", ""}, {"www-authenticate", ""}};

// Huffman decoding tree, built once from the code table.
struct Node {
  int32_t child[2] = {-1, -1};
  int32_t sym = -1;
};
std::vector<Node> g_tree;
std::once_flag g_tree_once;

void build_tree() {
  g_tree.emplace_back();
  for (int s = 0; s < 257; ++s) {
    uint32_t code = kHuffTable[s].code;
    int bits = kHuffTable[s].bits;
    int cur = 0;
    for (int b = bits - 1; b >= 0; --b) {
      int bit = (code >> b) & 1;
      if (g_tree[cur].child[bit] < 0) {
        g_tree[cur].child[bit] = static_cast<int32_t>(g_tree.size());
        g_tree.emplace_back();
      }
      cur = g_tree[cur].child[bit];
    }
    g_tree[cur].sym = s;
  }
}

// prefix-coded integer (RFC 7541 §5.1)
bool read_int(const uint8_t*& p, const uint8_t* end, int prefix_bits, uint64_t& out) {
  if (p >= end) return false;
  const uint32_t mask = (1u << prefix_bits) - 1;
  out = *p++ & mask;
  if (out < mask) return true;
  int shift = 0;
  while (p < end) {
    uint8_t b = *p++;
    out += static_cast<uint64_t>(b & 0x7f) << shift;
    shift += 7;
    if (!(b & 0x80)) return true;
    if (shift > 56) return false;
  }
  return false;
}

bool read_string(const uint8_t*& p, const uint8_t* end, std::string& out) {
  if (p >= end) return false;
  const bool huff = (*p & 0x80) != 0;
  uint64_t len;
  if (!read_int(p, end, 7, len)) return false;
  if (len > static_cast<uint64_t>(end - p)) return false;
  std::string_view raw(reinterpret_cast<const char*>(p), len);
  p += len;
  if (!huff) {
    out.assign(raw);
    return true;
  }
  return huffman_decode(raw, out);
}

void write_int(std::string& out, uint8_t first_byte_flags, int prefix_bits, uint64_t v) {
  const uint32_t mask = (1u << prefix_bits) - 1;
  if (v < mask) {
    out.push_back(static_cast<char>(first_byte_flags | v));
    return;
  }
  out.push_back(static_cast<char>(first_byte_flags | mask));
  v -= mask;
  while (v >= 128) {
    out.push_back(static_cast<char>((v & 0x7f) | 0x80));
    v >>= 7;
  }
  out.push_back(static_cast<char>(v));
}

}  // namespace

bool huffman_decode(std::string_view in, std::string& out) {
  std::call_once(g_tree_once, build_tree);
  out.clear();
  int cur = 0;
  int pad_bits = 0;
  bool pad_all_ones = true;
  for (unsigned char c : in) {
    for (int b = 7; b >= 0; --b) {
      int bit = (c >> b) & 1;
      cur = g_tree[cur].child[bit];
      if (cur < 0) return false;
      ++pad_bits;
      pad_all_ones = pad_all_ones && bit;
      if (g_tree[cur].sym >= 0) {
        if (g_tree[cur].sym == 256) return false;  // EOS inside a string is an error
        out.push_back(static_cast<char>(g_tree[cur].sym));
        cur = 0;
        pad_bits = 0;
        pad_all_ones = true;
      }
    }
  }
  // leftover bits must be a prefix of EOS (all ones) and shorter than one byte
  return pad_bits < 8 && pad_all_ones;
}

bool Decoder::lookup(uint64_t index, Header& h) const {
  if (index == 0) return false;
  if (index <= 61) {
    h = kStatic[index - 1];
    return true;
  }
  index -= 62;
  if (index >= dyn_.size()) return false;
  h = dyn_[index];
  return true;
}

void Decoder::evict() {
  while (size_ > max_size_ && !dyn_.empty()) {
    size_ -= dyn_.back().first.size() + dyn_.back().second.size() + 32;
    dyn_.pop_back();
  }
}

void Decoder::insert(const Header& h) {
  size_t need = h.first.size() + h.second.size() + 32;
  if (need > max_size_) {  // larger than the table: empties it
    dyn_.clear();
    size_ = 0;
    return;
  }
  dyn_.push_front(h);
  size_ += need;
  evict();
}

bool Decoder::decode(std::string_view block, std::vector<Header>& out) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(block.data());
  const uint8_t* end = p + block.size();
  while (p < end) {
    const uint8_t b = *p;
    if (b & 0x80) {  // indexed field
      uint64_t idx;
      Header h;
      if (!read_int(p, end, 7, idx) || !lookup(idx, h)) return false;
      out.push_back(std::move(h));
    } else if ((b & 0xe0) == 0x20) {  // dynamic table size update
      uint64_t sz;
      if (!read_int(p, end, 5, sz) || sz > limit_) return false;
      max_size_ = sz;
      evict();
    } else {
      const bool incremental = (b & 0xc0) == 0x40;
      const int prefix = incremental ? 6 : 4;  // 0000xxxx without indexing, 0001xxxx never indexed
      uint64_t idx;
      if (!read_int(p, end, prefix, idx)) return false;
      Header h;
      if (idx != 0) {
        Header named;
        if (!lookup(idx, named)) return false;
        h.first = named.first;
      } else if (!read_string(p, end, h.first)) {
        return false;
      }
      if (!read_string(p, end, h.second)) return false;
      if (incremental) insert(h);
      out.push_back(std::move(h));
    }
  }
  return true;
}

void Encoder::encode(const std::vector<Header>& headers, std::string& out) {
  for (const auto& h : headers) {
    int exact = 0, name_only = 0;
    for (int i = 0; i < 61; ++i) {
      if (kStatic[i].first == h.first) {
        if (!name_only) name_only = i + 1;
        if (kStatic[i].second == h.second && !h.second.empty()) {
          exact = i + 1;
          break;
        }
      }
    }
    if (exact) {
      write_int(out, 0x80, 7, exact);
      continue;
    }
    write_int(out, 0x00, 4, name_only);  // literal without indexing
    if (!name_only) {
      write_int(out, 0x00, 7, h.first.size());
      out.append(h.first);
    }
    write_int(out, 0x00, 7, h.second.size());
    out.append(h.second);
  }
}

}  // namespace hpack
}  // namespace psb
