// metafile.hpp -- reader of x-search v1 meta files (C++20, header-only).
//
// The v1 preprocessor wrote, next to a data file, a table of chunks with a line mapping; the classes
// that read it are absent from the reference snapshot, the format is pinned by the fixtures
// test/files/*.meta and by the field names printed in metafile_cat.cpp:39-49:
//
//   u32 compression_type            (1 = none, 2 = zstd, 3 = lz4), little endian
//   per chunk:  u64 original_offset, u64 actual_offset, u64 original_size, u64 actual_size,
//               u64 n_map,  n_map x { u64 globalByteOffset, u64 globalLineIndex }
//
// The GPU path uses the chunk table as the reference-chunk boundary list (every chunk starts at a
// line start and is >= 16 MiB extended through the next '\n'); compressed files are rejected.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace xs {

enum class CompressionType : uint32_t { UNKNOWN = 0, NONE = 1, ZSTD = 2, LZ4 = 3 };

struct ByteToNewLineMapping {
  uint64_t globalByteOffset;
  uint64_t globalLineIndex;
};

struct ChunkMetaData {
  uint64_t original_offset;
  uint64_t actual_offset;
  uint64_t original_size;
  uint64_t actual_size;
  std::vector<ByteToNewLineMapping> line_mapping_data;
};

class MetaFile {
 public:
  explicit MetaFile(const std::string& path) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("xs::MetaFile: cannot open '" + path + "'");
    std::vector<unsigned char> raw;
    unsigned char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) raw.insert(raw.end(), buf, buf + got);
    std::fclose(f);
    parse(raw, path);
  }

  CompressionType compression_type() const { return _compression; }
  const std::vector<ChunkMetaData>& chunks() const { return _chunks; }

  // exclusive end offsets of the chunks in the ORIGINAL (uncompressed) byte stream
  std::vector<uint64_t> chunk_bounds() const {
    std::vector<uint64_t> b;
    b.reserve(_chunks.size());
    for (const auto& c : _chunks) b.push_back(c.original_offset + c.original_size);
    return b;
  }
  uint64_t original_size() const {
    return _chunks.empty() ? 0 : _chunks.back().original_offset + _chunks.back().original_size;
  }

 private:
  void parse(const std::vector<unsigned char>& raw, const std::string& path) {
    size_t pos = 0;
    auto need = [&](size_t n) {
      if (pos + n > raw.size()) throw std::runtime_error("xs::MetaFile: '" + path + "' is truncated");
    };
    auto u64 = [&]() {
      need(8);
      uint64_t v;
      std::memcpy(&v, raw.data() + pos, 8);
      pos += 8;
      return v;
    };
    need(4);
    uint32_t ct;
    std::memcpy(&ct, raw.data(), 4);
    pos = 4;
    if (ct < 1 || ct > 3) throw std::runtime_error("xs::MetaFile: '" + path + "' has an unknown compression type");
    _compression = static_cast<CompressionType>(ct);
    uint64_t expect = 0;
    while (pos < raw.size()) {
      ChunkMetaData c;
      c.original_offset = u64();
      c.actual_offset = u64();
      c.original_size = u64();
      c.actual_size = u64();
      const uint64_t n_map = u64();
      if (n_map > (raw.size() - pos) / 16) throw std::runtime_error("xs::MetaFile: '" + path + "' is truncated");
      c.line_mapping_data.resize(n_map);
      for (auto& m : c.line_mapping_data) {
        m.globalByteOffset = u64();
        m.globalLineIndex = u64();
      }
      if (c.original_offset != expect || c.original_size == 0)
        throw std::runtime_error("xs::MetaFile: '" + path + "' has a gap in its chunk table");
      expect = c.original_offset + c.original_size;
      _chunks.push_back(std::move(c));
    }
  }

  CompressionType _compression = CompressionType::UNKNOWN;
  std::vector<ChunkMetaData> _chunks;
};

}  // namespace xs
