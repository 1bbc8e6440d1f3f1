// ref_shim.cpp -- thin harness around the UNMODIFIED reference sources.
//
// TEST INFRASTRUCTURE ONLY.  Compiled together with
//   /root/reference/src/string_search/simd_search.cpp
// and including /root/reference/include/xsearch/string_search/search_wrappers.h where
// they lie (oracle/Makefile) into oracle/_ref/libxsref.so.  No reference source is copied
// into this repository.  The shim supplies only what the reference snapshot lacks
// (SURVEY.md 0.1/0.3): v1 chunking is done by the caller (bounds array), chunks are copied
// into xs::strtype like the reference's FileReader does (tasks/readers.h:39-48), per-chunk
// results are globalised (+chunk base) and concatenated in file order, and line indices
// are derived from byte_offsets_line by counting '\n' in one pass per chunk (std::count).
#include <xsearch/string_search/search_wrappers.h>
#include <xsearch/string_search/simd_search.h>
#include <xsearch/types.h>
#include <xsearch/utils/string_utils.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define XSR_API extern "C" __attribute__((visibility("default")))

namespace {

enum Kind { K_COUNT = 0, K_MATCH_OFF, K_LINE_OFF, K_LINE_IDX, K_LINES };

struct Part {
  std::vector<uint64_t> a, b;
  uint64_t count = 0, newlines = 0;
};

// '\n' in d[from, to): ONE vectorised pass (std::count).  v1 answered xs::line_indices from the meta file's
// line-index mapping plus a short newline count (metafile_cat.cpp:47-49); that mapping is absent from the snapshot, so
// the harness counts.  Round 1 counted with one simd::findNextNewLine call per newline, which cost the timed
// baseline more than the search itself.
uint64_t count_newlines(const xs::strtype& d, uint64_t from, uint64_t to) {
  return static_cast<uint64_t>(std::count(d.data() + from, d.data() + to, '\n'));
}

void search_chunk(const char* p, uint64_t n, const std::string& pat, int kind, int skip_to_nl, Part* out) {
  // the reference searches its own copy of the chunk (reader stage, tasks/readers.h:39-48); the
  // buffer is reused per thread so that the timed baseline does not pay page faults per chunk
  thread_local xs::strtype data;
  data.assign(p, p + n);
  switch (kind) {
    case K_COUNT:
      out->count = xs::search::count(data, pat, skip_to_nl != 0);
      break;
    case K_MATCH_OFF:
      out->a = xs::search::byte_offsets_match(data, pat, false);
      break;
    case K_LINE_OFF:
      out->a = xs::search::byte_offsets_line(data, pat);
      break;
    case K_LINE_IDX: {
      out->a = xs::search::byte_offsets_line(data, pat);
      uint64_t seen = 0, pos = 0;
      for (auto& v : out->a) {
        seen += count_newlines(data, pos, v);
        pos = v;
        v = seen;
      }
      out->newlines = seen + count_newlines(data, pos, n);
      break;
    }
    case K_LINES: {
      // line() returns strings; recover (start, len) from byte_offsets_line, which visits the
      // same lines in the same order; line() drops a final unterminated line (:199-201).
      std::vector<std::string> ls = xs::search::line(data, pat);
      std::vector<uint64_t> st = xs::search::byte_offsets_line(data, pat);
      for (size_t i = 0; i < ls.size(); ++i) {
        out->a.push_back(st[i]);
        out->b.push_back(ls[i].size());
        // sanity: the string really is the bytes at [start, start+len)
        if (std::memcmp(ls[i].data(), p + st[i], ls[i].size()) != 0) out->b.back() = UINT64_MAX;
      }
      break;
    }
  }
}

uint64_t run(const char* buf, const char* pat, uint32_t m, const uint64_t* bounds, uint64_t nb, int kind,
             int skip_to_nl, uint64_t* out_a, uint64_t* out_b, uint64_t cap, int nthreads) {
  if (m == 0 || nb == 0) return 0;
  const std::string pattern(pat, pat + m);
  std::vector<Part> parts(nb);
  std::atomic<uint64_t> next{0};
  auto work = [&]() {
    for (;;) {
      uint64_t c = next.fetch_add(1);
      if (c >= nb) break;
      uint64_t b0 = c ? bounds[c - 1] : 0;
      search_chunk(buf + b0, bounds[c] - b0, pattern, kind, skip_to_nl, &parts[c]);
    }
  };
  if (nthreads <= 1) {
    work();
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work);
    for (auto& t : th) t.join();
  }
  uint64_t total = 0, nl_base = 0;
  for (uint64_t c = 0; c < nb; ++c) {
    const Part& p = parts[c];
    const uint64_t base = c ? bounds[c - 1] : 0;
    if (kind == K_COUNT) {
      total += p.count;
      continue;
    }
    for (size_t i = 0; i < p.a.size(); ++i, ++total) {
      if (total >= cap) continue;
      out_a[total] = p.a[i] + (kind == K_LINE_IDX ? nl_base : base);
      if (kind == K_LINES) out_b[total] = p.b[i];
    }
    nl_base += p.newlines;
  }
  return total;
}

}  // namespace

XSR_API int64_t xsr_find_next(const char* pat, size_t m, const char* buf, size_t n, size_t shift) {
  return xs::search::simd::findNext(pat, m, buf, n, shift);
}
// the reference's own toLower (src/utils/string_utils.cpp:11-33), the v1 ignore_case processor
XSR_API void xsr_to_lower(char* buf, uint64_t n) { xs::utils::str::simd::toLower(buf, n); }
XSR_API int64_t xsr_find_next_newline(const char* buf, size_t n, size_t shift) {
  return xs::search::simd::findNextNewLine(buf, n, shift);
}
// defined in simd_search.cpp:305-336 but declared under other names in the header
namespace xs::search::simd {
uint64_t countMatchingLines(const char* pattern, size_t pattern_len, const char* str, size_t str_len);
uint64_t countMatches(const char* pattern, size_t pattern_len, const char* str, size_t str_len);
}  // namespace xs::search::simd
XSR_API uint64_t xsr_count_matches(const char* pat, size_t m, const char* buf, size_t n) {
  return xs::search::simd::countMatches(pat, m, buf, n);
}
XSR_API uint64_t xsr_count_matching_lines(const char* pat, size_t m, const char* buf, size_t n) {
  return xs::search::simd::countMatchingLines(pat, m, buf, n);
}
XSR_API uint64_t xsr_count(const char* buf, uint64_t, const char* pat, uint32_t m, const uint64_t* bounds,
                           uint64_t nb, int skip_to_nl, int nthreads) {
  return run(buf, pat, m, bounds, nb, K_COUNT, skip_to_nl, nullptr, nullptr, 0, nthreads);
}
XSR_API uint64_t xsr_match_byte_offsets(const char* buf, uint64_t, const char* pat, uint32_t m,
                                        const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap,
                                        int nthreads) {
  return run(buf, pat, m, bounds, nb, K_MATCH_OFF, 0, out, nullptr, cap, nthreads);
}
XSR_API uint64_t xsr_line_byte_offsets(const char* buf, uint64_t, const char* pat, uint32_t m,
                                       const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap,
                                       int nthreads) {
  return run(buf, pat, m, bounds, nb, K_LINE_OFF, 1, out, nullptr, cap, nthreads);
}
XSR_API uint64_t xsr_line_indices(const char* buf, uint64_t, const char* pat, uint32_t m, const uint64_t* bounds,
                                  uint64_t nb, uint64_t* out, uint64_t cap, int nthreads) {
  return run(buf, pat, m, bounds, nb, K_LINE_IDX, 1, out, nullptr, cap, nthreads);
}
XSR_API uint64_t xsr_lines(const char* buf, uint64_t, const char* pat, uint32_t m, const uint64_t* bounds,
                           uint64_t nb, uint64_t* starts, uint64_t* lens, uint64_t cap, int nthreads) {
  return run(buf, pat, m, bounds, nb, K_LINES, 1, starts, lens, cap, nthreads);
}
