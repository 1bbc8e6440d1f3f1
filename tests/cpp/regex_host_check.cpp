// regex_host_check -- TEST DRIVER for the host stage of the regex prefilter (include/xsearch_b200/regex_prefilter.hpp).
// Not part of the product: it hands EVERY line of a file that holds a required literal (found with memmem) to
// xs::LineVerifier, i.e. it replaces the GPU literal scan by a trivial host loop so that the literal extraction and the
// VM / DFA verification (regex_vm.hpp) can be pinned against the reference's re2 wrappers on a machine without a GPU
// (tests/test_regex.py).  Prints what `xsgrep -E --tag TAG` prints.
#include <xsearch_b200/regex_prefilter.hpp>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <iterator>
#include <string>

static bool holds(std::string_view line, const std::string& lit, bool icase) {
  if (!icase) return line.find(lit) != std::string_view::npos;
  auto lower = [](unsigned char c) { return (c >= 'A' && c <= 'Z') ? char(c + 32) : char(c); };
  std::string a(line), b(lit);
  for (auto& c : a) c = lower(c);
  for (auto& c : b) c = lower(c);
  return a.find(b) != std::string::npos;
}

int main(int argc, char** argv) {
  if (argc < 4) return 64;
  const std::string tag = argv[1], pattern = argv[2];
  const bool icase = argc > 4 && std::string(argv[4]) == "-i";
  std::ifstream f(argv[3], std::ios::binary);
  const std::string data((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  std::vector<std::string> lits;
  try {
    lits = xs::required_literals(pattern);
  } catch (const std::invalid_argument& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 1;
  }
  const xs::LineVerifier v(pattern, icase);
  const auto cache = v.make_cache();  // the lazy DFA the Searcher uses for the line kinds
  uint64_t count = 0, count_lines = 0, index = 0;
  std::string out;
  for (size_t pos = 0; pos < data.size(); ++index) {
    const size_t nl = data.find('\n', pos);
    const size_t end = nl == std::string::npos ? data.size() : nl;
    const std::string_view line(data.data() + pos, end - pos);
    bool candidate = false;
    for (const auto& l : lits) candidate = candidate || holds(line, l, icase);
    std::vector<uint64_t> at;
    if (candidate) v.match_positions(line, &at);
    if (xs::LineVerifier::matches(line, *cache) != v.matches(line)) {
      std::fprintf(stderr, "line %llu: the DFA and the VM disagree\n", (unsigned long long)index);
      return 3;
    }
    if (candidate && at.empty() == v.matches(line)) {
      std::fprintf(stderr, "line %llu: match_positions and matches disagree\n", (unsigned long long)index);
      return 3;
    }
    if (!candidate && v.matches(line)) {  // the prefilter would have lost this line
      std::fprintf(stderr, "line %llu matches but holds no required literal\n", (unsigned long long)index);
      return 2;
    }
    if (!at.empty()) {
      count += at.size();
      ++count_lines;
      if (tag == "match_byte_offsets")
        for (uint64_t p : at) out += std::to_string(pos + p) + "\n";
      else if (tag == "line_byte_offsets")
        out += std::to_string(pos) + "\n";
      else if (tag == "line_indices")
        out += std::to_string(index) + "\n";
      else if (tag == "lines" && nl != std::string::npos)
        out += std::string(line) + "\n";
    }
    if (nl == std::string::npos) break;
    pos = nl + 1;
  }
  if (tag == "count") out = std::to_string(count) + "\n";
  if (tag == "count_lines") out = std::to_string(count_lines) + "\n";
  std::fwrite(out.data(), 1, out.size(), stdout);
  return 0;
}
