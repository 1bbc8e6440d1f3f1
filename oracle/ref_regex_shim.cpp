// ref_regex_shim.cpp -- thin harness around the reference's UNMODIFIED regex wrappers and its vendored re2.
//
// TEST INFRASTRUCTURE ONLY.  Compiled (oracle/Makefile, target _ref/libxsref_re.so) together with
//   /root/reference/src/string_search/simd_search.cpp
//   /root/reference/third_party/re2/{re2,util}/*.cc   (the library sources, no tests / benchmarks)
// and including /root/reference/include/xsearch/string_search/search_wrappers.h where they lie.  No reference
// source is copied into this repository.  It is the checker of the regex prefilter (SURVEY.md 8(f) rank 4,
// include/xsearch_b200/regex_prefilter.hpp): xs::search::regex::{count, byte_offsets_match, byte_offsets_line}
// (search_wrappers.h:63-88,209-271) are called per chunk, chunk results are globalised (+chunk base) and
// concatenated in file order -- exactly what ref_shim.cpp does for the literal wrappers.
#include <xsearch/string_search/search_wrappers.h>
#include <xsearch/types.h>

#include <cstdint>
#include <string>
#include <vector>

#define XSR_API extern "C" __attribute__((visibility("default")))

namespace {

enum Kind { K_COUNT = 0, K_COUNT_LINES, K_MATCH_OFF, K_LINE_OFF };

uint64_t run(const char* buf, const char* pattern, int ignore_case, const uint64_t* bounds, uint64_t nb, int kind,
             uint64_t* out, uint64_t cap) {
  re2::RE2::Options opt;
  opt.set_case_sensitive(!ignore_case);
  opt.set_log_errors(false);
  // The wrappers read "the match" from the FIRST capture group (RE2::PartialMatch(input, pattern, &match)), so their
  // callers wrap the whole expression in one: benchmarks/search_algorithms/count_matches.cpp:22,
  // count_matches_icase.cpp:22-25 (which is also where set_case_sensitive(false) comes from) and the test patterns
  // "(a[n|m]t)" (search_wrappersTest.cpp:78-103).
  const re2::RE2 re("(" + std::string(pattern) + ")", opt);
  if (!re.ok()) return UINT64_MAX;
  uint64_t total = 0;
  xs::strtype data;
  for (uint64_t c = 0; c < nb; ++c) {
    const uint64_t b0 = c ? bounds[c - 1] : 0;
    data.assign(buf + b0, buf + bounds[c]);
    if (kind == K_COUNT || kind == K_COUNT_LINES) {
      total += xs::search::regex::count(data, re, kind == K_COUNT_LINES);
      continue;
    }
    const std::vector<uint64_t> v = kind == K_MATCH_OFF ? xs::search::regex::byte_offsets_match(data, re, false)
                                                        : xs::search::regex::byte_offsets_line(data, re);
    for (uint64_t x : v) {
      if (total < cap) out[total] = x + b0;
      ++total;
    }
  }
  return total;
}

}  // namespace

// UINT64_MAX: re2 rejected the pattern
XSR_API uint64_t xsr_re_count(const char* buf, uint64_t, const char* pattern, int ignore_case, const uint64_t* bounds,
                              uint64_t nb, int skip_to_nl) {
  return run(buf, pattern, ignore_case, bounds, nb, skip_to_nl ? K_COUNT_LINES : K_COUNT, nullptr, 0);
}
XSR_API uint64_t xsr_re_match_byte_offsets(const char* buf, uint64_t, const char* pattern, int ignore_case,
                                           const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap) {
  return run(buf, pattern, ignore_case, bounds, nb, K_MATCH_OFF, out, cap);
}
XSR_API uint64_t xsr_re_line_byte_offsets(const char* buf, uint64_t, const char* pattern, int ignore_case,
                                          const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap) {
  return run(buf, pattern, ignore_case, bounds, nb, K_LINE_OFF, out, cap);
}
