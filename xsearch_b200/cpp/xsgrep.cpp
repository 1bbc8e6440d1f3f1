// xsgrep -- grep-like command line tool on top of xs::extern_search (the shape of the reference's
// example/grep.cpp:66-80 and checkit.cpp).  Also the driver of the C++ API tests (tests/test_cpp_api.py).
//
//   xsgrep [--tag lines|count|count_lines|match_byte_offsets|line_byte_offsets|line_indices] [--meta FILE]
//          [-j N] [--readers N] [--reader pread|mmap] [--min-chunk BYTES] [--segment-bytes BYTES] [--device D]
//          [--devices D0,D1,...] [--live]
//          [-i] [-E | -F]
//          PATTERN FILE
//   -E / -F read PATTERN as a regular expression / as plain text; default: the reference's rule (utils.h:17-24).
//   A regular expression is searched by scanning for its required literals on the GPU and running the expression
//   over the candidate lines on -j host threads (xsearch_b200/regex_prefilter.hpp).
//   xsgrep --required-literals PATTERN   print how PATTERN would be read and the literals the GPU would scan for
//   xsgrep --selftest-host          host-only checks of the result collector (no GPU needed)
//   xsgrep --metafile-cat FILE      print a v1 meta file (the reference's metafile_cat.cpp:23-49)
#include <xsearch_b200/xsearch.hpp>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>

namespace {

struct Args {
  std::string tag = "lines", meta, pattern, file;
  int threads = 1, readers = 1;
  bool live = false;
  xs::Options opt;
};

template <class Tag>
int run(const Args& a) {
  auto searcher = xs::extern_search<Tag>(a.pattern, a.file, a.meta, a.threads, a.readers, a.opt);
  auto* res = searcher->getResult();
  if constexpr (Tag::is_count) {
    uint64_t last = 0, yielded = 0;
    if (a.live) {
      for (auto v : *res) {  // running totals; the last one is the total
        last = v;
        ++yielded;
      }
      searcher->join();
      if (last != res->size()) {
        std::fprintf(stderr, "running total %llu != size %zu\n", (unsigned long long)last, res->size());
        return 3;
      }
      std::fprintf(stderr, "yielded %llu running totals\n", (unsigned long long)yielded);
    } else {
      searcher->join();
    }
    std::cout << res->size() << '\n';
  } else {
    std::string out;
    auto emit = [&out](const auto& v) {
      if constexpr (std::is_same_v<typename Tag::value_type, std::string>) out += v;
      else out += std::to_string(v);
      out += '\n';
      if (out.size() > (1u << 20)) {
        std::fwrite(out.data(), 1, out.size(), stdout);
        out.clear();
      }
    };
    if (a.live) {
      for (auto const& v : *res) emit(v);
      searcher->join();
    } else {
      searcher->join();
      for (auto const& v : res->copyResultSafe()) emit(v);
    }
    std::fwrite(out.data(), 1, out.size(), stdout);
  }
  if (!searcher->ok()) {
    std::fprintf(stderr, "xsgrep: %s\n", searcher->error().c_str());
    return 2;
  }
  const xsg_timings t = searcher->timings();
  std::fprintf(stderr, "xsgrep: %llu bytes, pipeline %.1f ms (%.2f GB/s), device scans %.2f ms, %u kernel launches, %zu device(s), pinned alloc %.1f ms\n",
               (unsigned long long)searcher->bytes(), t.h2d_ms, t.h2d_ms > 0 ? searcher->bytes() / t.h2d_ms / 1e6 : 0.0,
               t.scan_ms, t.kernel_launches, searcher->num_devices(), searcher->pinned_alloc_ms());
  return 0;
}

int selftest_host() {
  // blocking iterator: a producer thread adds parts with pauses, the consumer sees all of them in order
  xs::Result<uint64_t> r(false);
  std::thread prod([&] {
    for (uint64_t p = 0; p < 5; ++p) {
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
      r.add({p * 3, p * 3 + 1, p * 3 + 2});
    }
    r.close();
  });
  uint64_t expect = 0;
  for (auto v : r) {
    if (v != expect) return 1;
    ++expect;
  }
  prod.join();
  if (expect != 15 || r.size() != 15 || !r.is_closed() || r.add({1})) return 1;
  if (r.copyResultSafe().size() != 15 || r.at(7) != 7) return 1;
  // parts of different sizes (also empty ones): indexing crosses part edges, every begin() starts at the front
  xs::Result<uint64_t> q(false);
  q.add({});
  q.add({0});
  q.add({1, 2, 3});
  q.add({});
  q.add({4, 5});
  if (q.size() != 6) return 1;
  for (uint64_t i = 0; i < 6; ++i)
    if (q.at(i) != i || q[i] != i) return 1;
  try {
    q.at(6);
    return 1;
  } catch (const std::out_of_range&) {
  }
  q.close();
  for (int pass = 0; pass < 2; ++pass) {
    uint64_t n = 0;
    for (auto it = q.begin(); it != q.end(); ++it, ++n)
      if (*it != n) return 1;
    if (n != 6) return 1;
  }
  if (q.copyResultSafe() != std::vector<uint64_t>{0, 1, 2, 3, 4, 5}) return 1;
  // count tags: elements are running totals, size() is the total
  xs::Result<uint64_t> c(true);
  c.add({4});
  c.add({9});
  c.close();
  uint64_t last = 0;
  for (auto v : c) last = v;
  if (last != 9 || c.size() != 9) return 1;
  xs::Result<std::string> s(false);
  s.close();
  for (auto const& v : s) (void)v, std::abort();
  // argument errors are exceptions, before any thread starts
  try {
    xs::extern_search<xs::count>("a\nb", "/nonexistent");
    return 1;
  } catch (const std::invalid_argument&) {
  }
  try {
    xs::extern_search<xs::count>("abc", "/nonexistent/file");
    return 1;
  } catch (const std::runtime_error&) {
  }
  // regex prefilter: the literals every match must contain, and the per-line verification
  {
    using V = std::vector<std::string>;
    if (xs::required_literals("She[r ]lock") != V{"lock"}) return 1;
    if (xs::required_literals("(a[n|m]t)") != V{"a"}) return 1;
    if (xs::required_literals("foo(bar|bazqux)+x?") != V{"foo"}) return 1;
    if (xs::required_literals("(?:alpha|beta)[0-9]") != V{"alpha", "beta"}) return 1;
    if (xs::required_literals("a{2}bc\\.d*") != V{"bc."}) return 1;
    for (const char* bad : {"a*", "[ab]+", "x|.", "(?=a)b", "(a)\\1", "a\\nb"}) {
      try {
        xs::required_literals(bad);
        return 1;
      } catch (const std::invalid_argument&) {
      }
    }
    if (!xs::use_str_as_regex("She[r ]lock") || xs::use_str_as_regex("Sherlock") || xs::use_str_as_regex("a(b")) return 1;
    const xs::LineVerifier v("a[nm]t", false), vi("SHE[r ]lock", true);
    std::vector<uint64_t> pos;
    v.match_positions("an ant and an amt, anti", &pos);
    if (pos != std::vector<uint64_t>{3, 14, 19} || v.matches("and") || !vi.matches("a sherLOCK")) return 1;
    try {
      xs::Options o;
      o.pattern_mode = xs::Options::PatternMode::REGEX;
      xs::extern_search<xs::count>("[ab]+", "/nonexistent", "", 1, 1, o);
      return 1;
    } catch (const std::invalid_argument&) {
    }
  }
  std::puts("selftest-host ok");
  return 0;
}

int show_required_literals(const std::string& pattern) {
  std::printf("regex: %d\n", xs::use_str_as_regex(pattern) ? 1 : 0);
  try {
    for (const std::string& lit : xs::required_literals(pattern)) std::printf("literal: %s\n", lit.c_str());
  } catch (const std::invalid_argument& e) {
    std::fprintf(stderr, "xsgrep: %s\n", e.what());
    return 1;
  }
  return 0;
}

int metafile_cat(const std::string& path) {
  xs::MetaFile mf(path);
  std::printf("compression_type: %u\nchunks: %zu\noriginal_size: %llu\n", (unsigned)mf.compression_type(),
              mf.chunks().size(), (unsigned long long)mf.original_size());
  for (const auto& c : mf.chunks())
    std::printf("original_offset=%llu actual_offset=%llu original_size=%llu actual_size=%llu mappings=%zu first=(%llu,%llu)\n",
                (unsigned long long)c.original_offset, (unsigned long long)c.actual_offset,
                (unsigned long long)c.original_size, (unsigned long long)c.actual_size, c.line_mapping_data.size(),
                (unsigned long long)(c.line_mapping_data.empty() ? 0 : c.line_mapping_data[0].globalByteOffset),
                (unsigned long long)(c.line_mapping_data.empty() ? 0 : c.line_mapping_data[0].globalLineIndex));
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  Args a;
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    const std::string s = argv[i];
    auto next = [&]() -> std::string {
      if (i + 1 >= argc) {
        std::fprintf(stderr, "xsgrep: %s needs a value\n", s.c_str());
        std::exit(64);
      }
      return argv[++i];
    };
    if (s == "--selftest-host") return selftest_host();
    else if (s == "--metafile-cat") return metafile_cat(next());
    else if (s == "--required-literals") return show_required_literals(next());
    else if (s == "-E" || s == "--regex") a.opt.pattern_mode = xs::Options::PatternMode::REGEX;
    else if (s == "-F" || s == "--fixed-strings") a.opt.pattern_mode = xs::Options::PatternMode::LITERAL;
    else if (s == "--tag") a.tag = next();
    else if (s == "--meta") a.meta = next();
    else if (s == "-j") a.threads = std::atoi(next().c_str());
    else if (s == "--readers") a.readers = std::atoi(next().c_str());
    else if (s == "--min-chunk") a.opt.min_chunk = std::strtoull(next().c_str(), nullptr, 10);
    else if (s == "--segment-bytes") a.opt.segment_bytes = std::strtoull(next().c_str(), nullptr, 10);
    else if (s == "--device") a.opt.device = std::atoi(next().c_str());
    else if (s == "--devices") {  // one Searcher over several devices: --devices 0,1,2,3
      a.opt.devices.clear();
      const std::string list = next();
      for (size_t p = 0; p < list.size();) {
        const size_t q = list.find(',', p);
        a.opt.devices.push_back(std::atoi(list.substr(p, q == std::string::npos ? q : q - p).c_str()));
        if (q == std::string::npos) break;
        p = q + 1;
      }
    }
    else if (s == "--reader") {
      const std::string m = next();
      if (m != "pread" && m != "mmap") {
        std::fprintf(stderr, "xsgrep: --reader pread|mmap\n");
        return 64;
      }
      a.opt.reader_mode = m == "mmap" ? xs::Options::ReaderMode::MMAP : xs::Options::ReaderMode::PREAD;
    } else if (s == "--live") a.live = true;
    else if (s == "-i" || s == "--ignore-case") a.opt.ignore_case = true;
    else pos.push_back(s);
  }
  if (pos.size() != 2) {
    std::fprintf(stderr, "usage: xsgrep [--tag T] [--meta FILE] [-j N] [--readers N] [--live] PATTERN FILE\n");
    return 64;
  }
  a.pattern = pos[0];
  a.file = pos[1];
  try {
    if (a.tag == "lines") return run<xs::lines>(a);
    if (a.tag == "count" || a.tag == "count_matches") return run<xs::count>(a);
    if (a.tag == "count_lines") return run<xs::count_lines>(a);
    if (a.tag == "match_byte_offsets") return run<xs::match_byte_offsets>(a);
    if (a.tag == "line_byte_offsets") return run<xs::line_byte_offsets>(a);
    if (a.tag == "line_indices") return run<xs::line_indices>(a);
    std::fprintf(stderr, "xsgrep: unknown tag %s\n", a.tag.c_str());
    return 64;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "xsgrep: %s\n", e.what());
    return 1;
  }
}
