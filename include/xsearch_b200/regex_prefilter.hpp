// regex_prefilter.hpp -- regular expressions on top of the GPU literal scan.
//
// The reference dispatches a pattern to its re2 wrappers when xs::utils::use_str_as_regex says so
// (include/xsearch/utils/utils.h:17-24) and then runs re2 over every byte of every chunk
// (include/xsearch/string_search/search_wrappers.h:63-88,209-271).  Regular expressions stay off the GPU
// (north_star); what the GPU does for them is the expensive part, touching every byte:
//
//   1. required_literals(pattern): a set L of literals such that EVERY match of the expression contains at
//      least one member of L (the classic grep prefilter).  No such set -> the pattern is rejected.
//   2. one GPU stream per member of L reports the lines holding it (line start + 0-based line index);
//   3. the host runs the expression over those candidate lines only (LineVerifier, `num_threads` threads).
//
// A line that holds a match holds a required literal, so the candidate lines are a superset of the matching
// lines and step 3 makes the result exact.  Results are per line, like grep: an expression that could match
// across a '\n' ("a[^b]c", "a\sb") is searched line by line (documented difference to the reference, whose re2
// call sees whole chunks).
//
// The expression syntax is what re2 and the ECMAScript grammar of <regex> read the same way: literals, escapes,
// classes, '.', groups "(...)" / "(?:...)", alternation, the quantifiers * + ? {n,m} and their lazy forms, the
// anchors ^ $ \b.  Look-around, back-references and re2's "(?flags)" are rejected.  Matching follows re2's rules
// (regex_vm.hpp); quantifiers bind to the last BYTE of a non-ASCII literal (re2: to the whole UTF-8 character).
#pragma once

#include <algorithm>
#include <cstdint>
#include <memory>
#include <regex>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "regex_vm.hpp"

namespace xs {

// The reference's dispatch rule (utils.h:17-24): a pattern is a regular expression when it compiles as one
// and does not match itself; anything that does not compile is searched as plain text.
inline bool use_str_as_regex(const std::string& pattern) {
  try {
    const std::regex probe("^" + pattern + "$");
    return !std::regex_match(pattern, probe);
  } catch (const std::regex_error&) {
    return false;
  }
}

namespace regex_detail {

constexpr size_t kMaxLiterals = 8;   // one GPU stream each
constexpr size_t kMaxLiteralLen = 64;  // XSG_MAX_PATTERN

struct Info {
  bool exact = false;              // the node matches exactly one string: s
  std::string s;
  bool has_req = false;            // every match of the node contains one of req
  std::vector<std::string> req;
};

inline size_t min_len(const std::vector<std::string>& v) {
  size_t m = SIZE_MAX;
  for (const auto& x : v) m = std::min(m, x.size());
  return v.empty() ? 0 : m;
}

// a better prefilter: longer shortest literal first, then fewer literals
inline bool better(const std::vector<std::string>& a, const std::vector<std::string>& b) {
  const size_t la = std::min(min_len(a), kMaxLiteralLen), lb = std::min(min_len(b), kMaxLiteralLen);
  if (la != lb) return la > lb;
  return a.size() < b.size();
}

class Parser {
 public:
  explicit Parser(const std::string& p) : _p(p) {}

  Info parse() {
    Info r = alternation();
    if (_i != _p.size()) bad("unbalanced ')'");
    return r;
  }

 private:
  [[noreturn]] void bad(const std::string& why) const {
    throw std::invalid_argument("xs::extern_search: regular expression not usable with the GPU prefilter: " + why);
  }
  bool more() const { return _i < _p.size(); }
  char peek() const { return _p[_i]; }

  Info alternation() {
    std::vector<Info> alts;
    alts.push_back(concatenation());
    while (more() && peek() == '|') {
      ++_i;
      alts.push_back(concatenation());
    }
    if (alts.size() == 1) return alts[0];
    Info r;  // a match of the alternation is a match of one branch: the union of the branches' sets
    r.has_req = true;
    for (const Info& a : alts) {
      if (!a.has_req) {
        r.has_req = false;
        break;
      }
      for (const auto& s : a.req)
        if (std::find(r.req.begin(), r.req.end(), s) == r.req.end()) r.req.push_back(s);
    }
    if (!r.has_req || r.req.size() > kMaxLiterals) {
      r.has_req = false;
      r.req.clear();
    }
    return r;
  }

  Info concatenation() {
    std::vector<std::vector<std::string>> cands;
    std::string run;  // literal formed by adjacent exact children
    bool all_exact = true;
    while (more() && peek() != '|' && peek() != ')') {
      Info a = repetition();
      if (a.exact) {
        run += a.s;
        continue;
      }
      all_exact = false;
      if (!run.empty()) cands.push_back({run});
      run.clear();
      if (a.has_req) cands.push_back(a.req);
    }
    Info r;
    if (all_exact) {
      r.exact = true;
      r.s = run;
      r.has_req = !run.empty();
      if (r.has_req) r.req = {run};
      return r;
    }
    if (!run.empty()) cands.push_back({run});
    for (auto& c : cands)
      if (!r.has_req || better(c, r.req)) {
        r.has_req = true;
        r.req = c;
      }
    return r;
  }

  Info repetition() {
    Info a = atom();
    for (;;) {
      if (!more()) break;
      const char c = peek();
      unsigned lo = 1;
      bool one = false;  // exactly once
      if (c == '*' || c == '?') {
        lo = 0;
        ++_i;
      } else if (c == '+') {
        ++_i;
      } else if (c == '{' && counted(&lo, &one)) {
        // _i advanced by counted()
      } else {
        break;
      }
      if (more() && peek() == '?') ++_i;  // lazy form: same language
      if (lo == 0) {
        a = Info();
      } else if (!one) {
        a.exact = false;  // still contains one copy of everything the atom requires
        if (!a.has_req && !a.s.empty()) {
          a.has_req = true;
          a.req = {a.s};
        }
        a.s.clear();
      }
    }
    return a;
  }

  // "{n}", "{n,}", "{n,m}" at _i; anything else is not a quantifier
  bool counted(unsigned* lo, bool* one) {
    size_t j = _i + 1;
    unsigned n = 0, m = 0;
    size_t d = 0;
    while (j < _p.size() && _p[j] >= '0' && _p[j] <= '9' && d < 6) n = n * 10 + unsigned(_p[j++] - '0'), ++d;
    if (d == 0) return false;
    bool has_m = false, open = false;
    if (j < _p.size() && _p[j] == ',') {
      ++j;
      size_t e = 0;
      while (j < _p.size() && _p[j] >= '0' && _p[j] <= '9' && e < 6) m = m * 10 + unsigned(_p[j++] - '0'), ++e;
      has_m = e > 0;
      open = e == 0;
    }
    if (j >= _p.size() || _p[j] != '}') return false;
    _i = j + 1;
    *lo = n;
    *one = n == 1 && !open && (!has_m || m == 1);
    return true;
  }

  static Info literal(char c) {
    Info r;
    r.exact = true;
    r.s.assign(1, c);
    r.has_req = true;
    r.req = {r.s};
    return r;
  }

  Info atom() {
    const char c = _p[_i++];
    switch (c) {
      case '(': {
        if (more() && peek() == '?') {
          if (_i + 1 < _p.size() && _p[_i + 1] == ':')
            _i += 2;
          else
            bad("look-around and \"(?flags)\" groups are not supported");
        }
        Info r = alternation();
        if (!more() || peek() != ')') bad("missing ')'");
        ++_i;
        return r;
      }
      case '[':
        skip_class();
        return Info();
      case '.':
      case '^':
      case '$':
        return Info();
      case '\\':
        return escape();
      case '*':
      case '+':
      case '?':
        bad("quantifier without an operand");
      default:
        if (c == '\n') bad("a pattern must not contain a newline");
        return literal(c);
    }
  }

  void skip_class() {
    if (more() && peek() == '^') ++_i;
    if (more() && peek() == ']') ++_i;  // a leading ']' is a member
    while (more() && peek() != ']') {
      if (peek() == '\\' && _i + 1 < _p.size()) {
        _i += 2;
      } else if (peek() == '[' && _i + 1 < _p.size() && (_p[_i + 1] == ':' || _p[_i + 1] == '.' || _p[_i + 1] == '=')) {
        const char k = _p[_i + 1];  // "[:alpha:]" and friends
        size_t j = _i + 2;
        while (j + 1 < _p.size() && !(_p[j] == k && _p[j + 1] == ']')) ++j;
        _i = std::min(j + 2, _p.size());
      } else {
        ++_i;
      }
    }
    if (!more()) bad("missing ']'");
    ++_i;
  }

  static bool is_hex(char c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F'); }

  Info escape() {
    if (!more()) bad("trailing backslash");
    const char c = _p[_i++];
    switch (c) {
      case 'd': case 'D': case 'w': case 'W': case 's': case 'S': case 'b': case 'B':
        return Info();
      case 'n':
        bad("a pattern must not contain a newline");
      case 't': return literal('\t');
      case 'r': return literal('\r');
      case 'f': return literal('\f');
      case 'v': return literal('\v');
      case 'x': {  // \xHH: a byte we do not fold into the literal
        size_t k = 0;
        while (k < 2 && more() && is_hex(peek())) ++_i, ++k;
        if (k != 2) bad("\\x needs two hex digits");
        return Info();
      }
      case 'u': {
        size_t k = 0;
        while (k < 4 && more() && is_hex(peek())) ++_i, ++k;
        if (k != 4) bad("\\u needs four hex digits");
        return Info();
      }
      case 'c':
        if (more()) ++_i;
        return Info();
      default:
        if (c >= '1' && c <= '9') bad("back-references are not supported");
        if (c == '0') return Info();
        if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) bad(std::string("escape \\") + c + " is not supported");
        return literal(c);  // escaped punctuation
    }
  }

  const std::string& _p;
  size_t _i = 0;
};

}  // namespace regex_detail

// The prefilter of `pattern`: every match contains at least one of the returned literals (1..64 bytes each,
// at most 8 of them).  Throws std::invalid_argument when the expression has no such set (e.g. "a*", "[ab]+",
// "x|.") or uses syntax outside the supported subset.
inline std::vector<std::string> required_literals(const std::string& pattern) {
  regex_detail::Parser parser(pattern);
  regex_detail::Info info = parser.parse();
  if (!info.has_req || info.req.empty() || regex_detail::min_len(info.req) == 0)
    throw std::invalid_argument(
        "xs::extern_search: regular expression not usable with the GPU prefilter: no literal is part of every match");
  std::vector<std::string> out;
  for (std::string s : info.req) {
    if (s.size() > regex_detail::kMaxLiteralLen) s.resize(regex_detail::kMaxLiteralLen);  // a prefix is still required
    if (std::find(out.begin(), out.end(), s) == out.end()) out.push_back(std::move(s));
  }
  // a literal that contains another one is redundant: the shorter one already selects its lines
  std::vector<std::string> keep;
  for (size_t i = 0; i < out.size(); ++i) {
    bool redundant = false;
    for (size_t j = 0; j < out.size() && !redundant; ++j)
      redundant = j != i && out[i].find(out[j]) != std::string::npos && (out[i].size() > out[j].size() || j < i);
    if (!redundant) keep.push_back(out[i]);
  }
  return keep;
}

// Step 3: the expression over one candidate line (the bytes between two '\n', the '\n' excluded), on the Pike VM of
// regex_vm.hpp (re2's matching rules; <regex> is only used for the reference's dispatch rule above).
class LineVerifier {
 public:
  LineVerifier(const std::string& pattern, bool ignore_case) : _prog(pattern, ignore_case) {}

  bool matches(std::string_view line) const {
    regex_vm::Span m;
    return _prog.search(line.data(), line.size(), 0, &m);
  }

  // The same answer from a lazily built DFA (regex_vm::Dfa), several times faster per byte.  The cache is not
  // thread-safe: every verification thread owns one and passes it in.
  using Cache = regex_vm::Dfa;
  std::unique_ptr<Cache> make_cache() const { return std::make_unique<Cache>(_prog); }
  static bool matches(std::string_view line, Cache& cache) { return cache.matches(line.data(), line.size()); }

  // start of every match, left to right, non-overlapping: the loop of search_wrappers.h:63-88 within one line.
  // Like that loop (input.remove_prefix(shift)), the search for the next match sees the text BEHIND the previous
  // match only: "^" and "\b" treat that position as a start of text ("\bfoo" finds two matches in "foofoo").
  void match_positions(std::string_view line, std::vector<uint64_t>* out) const {
    regex_vm::Span m;
    for (size_t from = 0; from < line.size() && _prog.search(line.data() + from, line.size() - from, 0, &m);) {
      if (m.end > m.begin) out->push_back(static_cast<uint64_t>(from + m.begin));
      from += m.end > m.begin ? m.end : m.end + 1;  // an expression with a required literal never matches empty
    }
  }

 private:
  regex_vm::Program _prog;
};

}  // namespace xs
