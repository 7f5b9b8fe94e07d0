// hpack.h — HPACK (RFC 7541) header compression for the gRPC-lite transport.
//
// Decoder: complete (static + dynamic table, all literal forms, table-size updates,
// Huffman strings) because real gRPC C-core peers use all of it.  Encoder: deliberately
// trivial — literals without indexing, indexed names from the static table, no Huffman —
// which every compliant peer accepts.  The reference links the gRPC C++ library instead
// (/root/reference/CMakeLists.txt:20); that library does not exist on the target box.
#pragma once

#include <cstdint>
#include <deque>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace psb {
namespace hpack {

using Header = std::pair<std::string, std::string>;

bool huffman_decode(std::string_view in, std::string& out);

class Decoder {
 public:
  explicit Decoder(size_t max_table = 4096) : max_size_(max_table), limit_(max_table) {}
  // Decodes one complete header block; returns false on a malformed block.
  bool decode(std::string_view block, std::vector<Header>& out);
  void set_limit(size_t n) { limit_ = n; }  // our SETTINGS_HEADER_TABLE_SIZE

 private:
  bool lookup(uint64_t index, Header& h) const;
  void insert(const Header& h);
  void evict();
  std::deque<Header> dyn_;
  size_t size_ = 0;
  size_t max_size_;
  size_t limit_;
};

class Encoder {
 public:
  void encode(const std::vector<Header>& headers, std::string& out);
};

}  // namespace hpack
}  // namespace psb
