// regex_vm.hpp -- the host-side matcher of the regex prefilter (regex_prefilter.hpp).
//
// The reference hands regular expressions to re2 (include/xsearch/string_search/search_wrappers.h:63-88,209-271).
// re2 is not part of this repository, and <regex> costs about a microsecond per candidate line, so the candidate
// lines the GPU reports are verified by this small engine instead: the expression is compiled to a program for
// a Pike VM (one pass over the line, all alternatives in lock step, no backtracking: time O(line x program)), with
// re2's matching rules for the syntax both read the same way --
//   * leftmost match, and among the matches starting there the one a backtracking engine would find first
//     (alternatives left to right, greedy quantifiers longest first, lazy ones shortest first);
//   * '.' and negated classes match one UTF-8 encoded character (any ASCII byte but '\n', or a 2..4-byte sequence);
//   * \d \w \s \b and their negations are ASCII, ignore_case folds ASCII letters only;
//   * '^' and '$' anchor the candidate line (see regex_prefilter.hpp for why results are per line).
// Supported: literals, escapes (\t \r \f \v \0 \xHH, escaped punctuation), classes with ranges / escapes /
// [:posix:] names, '.', groups "(...)" and "(?:...)", alternation, * + ? {n} {n,} {n,m} and their lazy forms,
// ^ $ \b \B.  Rejected with std::invalid_argument: look-around, back-references, "(?flags)", \p{..}, \u, \c, other
// letter escapes, non-ASCII class members, repetition counts above 1000.
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace xs::regex_vm {

struct ByteSet {
  uint64_t w[4] = {0, 0, 0, 0};
  void add(unsigned c) { w[c >> 6] |= 1ull << (c & 63); }
  void add_range(unsigned lo, unsigned hi) {
    for (unsigned c = lo; c <= hi; ++c) add(c);
  }
  bool has(unsigned c) const { return (w[c >> 6] >> (c & 63)) & 1; }
  void merge(const ByteSet& o) {
    for (int i = 0; i < 4; ++i) w[i] |= o.w[i];
  }
  void fold_case() {  // ignore_case: both spellings of every ASCII letter
    for (unsigned c = 'a'; c <= 'z'; ++c)
      if (has(c) || has(c - 32)) add(c), add(c - 32);
  }
};

inline bool is_word(unsigned char c) {
  return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_';
}

// ---- syntax tree -------------------------------------------------------------------------------------------
struct Node {
  enum Kind { EMPTY, SET, ANYCHAR, CAT, ALT, REPEAT, BOL, EOL, WORDB, NWORDB } kind = EMPTY;
  ByteSet set;           // SET: the bytes; ANYCHAR: its ASCII part (multi-byte UTF-8 sequences are added on emit)
  std::vector<std::unique_ptr<Node>> sub;
  int lo = 0, hi = 0;    // REPEAT: hi < 0 = unbounded
  bool greedy = true;
};
using NodePtr = std::unique_ptr<Node>;

class TreeParser {
 public:
  TreeParser(const std::string& p, bool ignore_case) : _p(p), _icase(ignore_case) {}

  NodePtr parse() {
    NodePtr r = alternation();
    if (_i != _p.size()) bad("unbalanced ')'");
    return r;
  }

 private:
  [[noreturn]] void bad(const std::string& why) const {
    throw std::invalid_argument("xs::extern_search: regular expression not supported: " + why);
  }
  bool more() const { return _i < _p.size(); }
  char peek() const { return _p[_i]; }
  static NodePtr make(Node::Kind k) {
    auto n = std::make_unique<Node>();
    n->kind = k;
    return n;
  }
  NodePtr byte_node(unsigned char c) const {
    NodePtr n = make(Node::SET);
    n->set.add(c);
    if (_icase) n->set.fold_case();
    return n;
  }

  NodePtr alternation() {
    NodePtr first = concatenation();
    if (!more() || peek() != '|') return first;
    NodePtr alt = make(Node::ALT);
    alt->sub.push_back(std::move(first));
    while (more() && peek() == '|') {
      ++_i;
      alt->sub.push_back(concatenation());
    }
    return alt;
  }

  NodePtr concatenation() {
    NodePtr cat = make(Node::CAT);
    while (more() && peek() != '|' && peek() != ')') cat->sub.push_back(repetition());
    return cat;
  }

  NodePtr repetition() {
    NodePtr a = atom();
    while (more()) {
      const char c = peek();
      int lo = 0, hi = -1;
      if (c == '*') {
        ++_i;
      } else if (c == '+') {
        lo = 1;
        ++_i;
      } else if (c == '?') {
        hi = 1;
        ++_i;
      } else if (c == '{' && counted(&lo, &hi)) {
        // _i advanced by counted()
      } else {
        break;
      }
      NodePtr r = make(Node::REPEAT);
      r->lo = lo;
      r->hi = hi;
      if (more() && peek() == '?') {
        r->greedy = false;
        ++_i;
      }
      r->sub.push_back(std::move(a));
      a = std::move(r);
    }
    return a;
  }

  // "{n}", "{n,}", "{n,m}" at _i (the same reading as regex_detail::Parser::counted); anything else is a literal '{'
  bool counted(int* lo, int* hi) {
    size_t j = _i + 1;
    int n = 0, m = 0;
    size_t d = 0;
    while (j < _p.size() && _p[j] >= '0' && _p[j] <= '9' && d < 6) n = n * 10 + (_p[j++] - '0'), ++d;
    if (d == 0) return false;
    bool has_m = false, open = false;
    if (j < _p.size() && _p[j] == ',') {
      ++j;
      size_t e = 0;
      while (j < _p.size() && _p[j] >= '0' && _p[j] <= '9' && e < 6) m = m * 10 + (_p[j++] - '0'), ++e;
      has_m = e > 0;
      open = e == 0;
    }
    if (j >= _p.size() || _p[j] != '}') return false;
    _i = j + 1;
    *lo = n;
    *hi = open ? -1 : (has_m ? m : n);
    if (n > 1000 || m > 1000) bad("repetition count above 1000");
    if (*hi >= 0 && *hi < *lo) bad("bad repetition range");
    return true;
  }

  NodePtr atom() {
    const char c = _p[_i++];
    switch (c) {
      case '(': {
        if (more() && peek() == '?') {
          if (_i + 1 < _p.size() && _p[_i + 1] == ':')
            _i += 2;
          else
            bad("look-around and \"(?flags)\" groups");
        }
        NodePtr r = alternation();
        if (!more() || peek() != ')') bad("missing ')'");
        ++_i;
        return r;
      }
      case '[':
        return char_class();
      case '.': {
        NodePtr n = make(Node::ANYCHAR);
        n->set.add_range(0, 127);
        n->set.w[0] &= ~(1ull << '\n');
        return n;
      }
      case '^':
        return make(Node::BOL);
      case '$':
        return make(Node::EOL);
      case '\\':
        return escape();
      case '*':
      case '+':
      case '?':
        bad("quantifier without an operand");
      default:
        if (c == '\n') bad("a pattern must not contain a newline");
        return byte_node(static_cast<unsigned char>(c));
    }
  }

  static bool is_hex(char c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F'); }
  static unsigned hex_val(char c) { return c <= '9' ? unsigned(c - '0') : unsigned((c | 32) - 'a' + 10); }

  // \d \w \s and negations into *out; false when c is none of them
  static bool perl_class(char c, ByteSet* out) {
    ByteSet s;
    switch (c | 32) {
      case 'd': s.add_range('0', '9'); break;
      case 'w': s.add_range('0', '9'), s.add_range('a', 'z'), s.add_range('A', 'Z'), s.add('_'); break;
      case 's': s.add(' '), s.add('\t'), s.add('\n'), s.add('\f'), s.add('\r'); break;  // re2's \s has no \v
      default: return false;
    }
    if (c < 'a') {  // upper case: the complement within ASCII (non-ASCII characters are handled by the caller)
      for (int i = 0; i < 2; ++i) s.w[i] = ~s.w[i];
    }
    out->merge(s);
    return true;
  }

  // one escaped byte (after the backslash) shared by atoms and class members; returns -1 for a class escape
  int simple_escape(char c) {
    switch (c) {
      case 't': return '\t';
      case 'r': return '\r';
      case 'f': return '\f';
      case 'v': return '\v';
      case 'a': return '\a';
      case '0': return 0;
      case 'x': {
        if (_i + 1 < _p.size() && is_hex(_p[_i]) && is_hex(_p[_i + 1])) {
          const int v = int(hex_val(_p[_i]) * 16 + hex_val(_p[_i + 1]));
          _i += 2;
          return v;
        }
        bad("\\x needs two hex digits");
      }
      default:
        if (c >= '1' && c <= '9') bad("back-references");
        if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) bad(std::string("escape \\") + c);
        return static_cast<unsigned char>(c);  // escaped punctuation
    }
  }

  NodePtr escape() {
    if (!more()) bad("trailing backslash");
    const char c = _p[_i++];
    if (c == 'b') return make(Node::WORDB);
    if (c == 'B') return make(Node::NWORDB);
    if (c == 'n') bad("a pattern must not contain a newline");
    ByteSet s;
    if (perl_class(c, &s)) {
      const bool negated = c < 'a';
      NodePtr n = make(negated ? Node::ANYCHAR : Node::SET);  // \D \W \S also match any non-ASCII character
      n->set = s;
      n->set.w[0] &= ~(1ull << '\n');
      return n;
    }
    if (c == 'a') bad("escape \\a");  // the literal extractor does not read it either
    const int v = simple_escape(c);
    if (v >= 128) bad("\\x above 7f");
    return byte_node(static_cast<unsigned char>(v));
  }

  // "[...]" after the '[' (the same scan for the closing bracket as regex_detail::Parser::skip_class)
  NodePtr char_class() {
    struct Item {
      int v = -1;         // a single member (0..127), or -1: `set`
      bool dash = false;  // an unescaped '-'
      ByteSet set;
    };
    bool negated = false, nonascii = false;  // nonascii: \D \W \S inside the class also match every non-ASCII character
    if (more() && peek() == '^') negated = true, ++_i;
    std::vector<Item> items;
    bool first = true;
    while (more() && (peek() != ']' || first)) {
      first = false;
      Item it;
      if (peek() == '\\' && _i + 1 < _p.size()) {
        const char c = _p[_i + 1];
        _i += 2;
        if (perl_class(c, &it.set)) {
          nonascii = nonascii || c < 'a';
        } else if (c == 'n') {
          it.v = '\n';
        } else if (c == 'b') {
          it.v = '\b';
        } else {
          it.v = simple_escape(c);
        }
      } else if (peek() == '[' && _i + 1 < _p.size() && (_p[_i + 1] == ':' || _p[_i + 1] == '.' || _p[_i + 1] == '=')) {
        const char k = _p[_i + 1];
        size_t j = _i + 2;
        while (j + 1 < _p.size() && !(_p[j] == k && _p[j + 1] == ']')) ++j;
        if (j + 1 >= _p.size()) bad("missing ']'");
        if (k != ':') bad("collating elements in a class");
        posix_class(_p.substr(_i + 2, j - (_i + 2)), &it.set);
        _i = j + 2;
      } else {
        it.v = static_cast<unsigned char>(_p[_i++]);
        it.dash = it.v == '-';
      }
      if (it.v >= 128) bad("non-ASCII class members");
      items.push_back(it);
    }
    if (!more()) bad("missing ']'");
    ++_i;
    ByteSet s;
    for (size_t k = 0; k < items.size(); ++k) {
      const Item& it = items[k];
      if (it.v < 0) {
        s.merge(it.set);
      } else if (k + 2 < items.size() && items[k + 1].dash && items[k + 2].v >= 0) {  // "a-z"
        if (items[k + 2].v < it.v) bad("bad class range");
        s.add_range(unsigned(it.v), unsigned(items[k + 2].v));
        k += 2;
      } else {
        s.add(unsigned(it.v));  // includes a '-' that is first, last or next to a class escape
      }
    }
    if (_icase) s.fold_case();
    bool wide = nonascii;  // does the class match non-ASCII characters?
    if (negated) {
      for (int i = 0; i < 2; ++i) s.w[i] = ~s.w[i];
      wide = !nonascii;
    }
    s.w[2] = s.w[3] = 0;
    s.w[0] &= ~(1ull << '\n');  // a line holds no newline; keeps "[^a]" line-local like grep
    NodePtr n = make(wide ? Node::ANYCHAR : Node::SET);
    n->set = s;
    return n;
  }

  void posix_class(const std::string& name, ByteSet* s) {
    if (name == "alpha") s->add_range('a', 'z'), s->add_range('A', 'Z');
    else if (name == "digit") s->add_range('0', '9');
    else if (name == "alnum") s->add_range('a', 'z'), s->add_range('A', 'Z'), s->add_range('0', '9');
    else if (name == "upper") s->add_range('A', 'Z');
    else if (name == "lower") s->add_range('a', 'z');
    else if (name == "space") s->add(' '), s->add_range('\t', '\r');
    else if (name == "blank") s->add(' '), s->add('\t');
    else if (name == "punct") s->add_range('!', '/'), s->add_range(':', '@'), s->add_range('[', '`'), s->add_range('{', '~');
    else if (name == "xdigit") s->add_range('0', '9'), s->add_range('a', 'f'), s->add_range('A', 'F');
    else if (name == "word") s->add_range('a', 'z'), s->add_range('A', 'Z'), s->add_range('0', '9'), s->add('_');
    else if (name == "print") s->add_range(' ', '~');
    else if (name == "graph") s->add_range('!', '~');
    else if (name == "cntrl") s->add_range(0, 31), s->add(127);
    else bad("class name [:" + name + ":]");
  }

  const std::string& _p;
  const bool _icase;
  size_t _i = 0;
};

// ---- program -----------------------------------------------------------------------------------------------
struct Inst {
  enum Op : uint8_t { BYTE, SET, SPLIT, JMP, BOL, EOL, WORDB, NWORDB, MATCH } op = MATCH;
  uint8_t byte = 0;    // BYTE
  uint32_t x = 0;      // SET: index into sets; SPLIT: preferred target; JMP: target
  uint32_t y = 0;      // SPLIT: the other target
};

struct Span {
  size_t begin = 0, end = 0;
};

class Program {
 public:
  Program(const std::string& pattern, bool ignore_case) {
    TreeParser parser(pattern, ignore_case);
    const NodePtr tree = parser.parse();
    emit(*tree);
    push(Inst::MATCH);
    find_first_bytes();
  }

  size_t size() const { return _prog.size(); }
  // true when the program holds ^ $ \b \B: what follows a position then depends on its neighbours (no DFA)
  bool has_assertions() const {
    for (const Inst& i : _prog)
      if (i.op == Inst::BOL || i.op == Inst::EOL || i.op == Inst::WORDB || i.op == Inst::NWORDB) return true;
    return false;
  }

  // Leftmost match starting at or behind `from` in the line [s, s+n); assertions see the whole line.
  bool search(const char* s, size_t n, size_t from, Span* out) const;

 private:
  static constexpr size_t kMaxInsts = 20000;

  uint32_t push(Inst::Op op, uint32_t x = 0, uint32_t y = 0, uint8_t byte = 0) {
    if (_prog.size() >= kMaxInsts)
      throw std::invalid_argument("xs::extern_search: regular expression not supported: program too large");
    Inst i;
    i.op = op, i.x = x, i.y = y, i.byte = byte;
    _prog.push_back(i);
    return static_cast<uint32_t>(_prog.size() - 1);
  }
  uint32_t here() const { return static_cast<uint32_t>(_prog.size()); }

  void emit_set(const ByteSet& set) {
    int count = 0, last = 0;
    for (int c = 0; c < 256; ++c)
      if (set.has(unsigned(c))) ++count, last = c;
    if (count == 1) {
      push(Inst::BYTE, 0, 0, static_cast<uint8_t>(last));
      return;
    }
    for (size_t k = 0; k < _sets.size(); ++k)
      if (std::memcmp(_sets[k].w, set.w, sizeof set.w) == 0) {
        push(Inst::SET, static_cast<uint32_t>(k));
        return;
      }
    _sets.push_back(set);
    push(Inst::SET, static_cast<uint32_t>(_sets.size() - 1));
  }

  // one character: an ASCII byte of `ascii`, or any 2..4-byte UTF-8 sequence
  void emit_anychar(const ByteSet& ascii) {
    ByteSet lead2, lead3, lead4, cont;
    lead2.add_range(0xC2, 0xDF), lead3.add_range(0xE0, 0xEF), lead4.add_range(0xF0, 0xF4), cont.add_range(0x80, 0xBF);
    //   split A, L2 ; A: set ascii ; jmp END ; L2: split B, L3 ; B: lead2 cont ; jmp END ; L3: split C, D ; ...
    const uint32_t s1 = push(Inst::SPLIT);
    _prog[s1].x = here();
    emit_set(ascii);
    const uint32_t j1 = push(Inst::JMP);
    const uint32_t s2 = push(Inst::SPLIT);
    _prog[s1].y = s2;
    _prog[s2].x = here();
    emit_set(lead2), emit_set(cont);
    const uint32_t j2 = push(Inst::JMP);
    const uint32_t s3 = push(Inst::SPLIT);
    _prog[s2].y = s3;
    _prog[s3].x = here();
    emit_set(lead3), emit_set(cont), emit_set(cont);
    const uint32_t j3 = push(Inst::JMP);
    _prog[s3].y = here();
    emit_set(lead4), emit_set(cont), emit_set(cont), emit_set(cont);
    _prog[j1].x = _prog[j2].x = _prog[j3].x = here();
  }

  void emit(const Node& n) {
    switch (n.kind) {
      case Node::EMPTY:
        break;
      case Node::SET:
        emit_set(n.set);
        break;
      case Node::ANYCHAR:
        emit_anychar(n.set);
        break;
      case Node::BOL: push(Inst::BOL); break;
      case Node::EOL: push(Inst::EOL); break;
      case Node::WORDB: push(Inst::WORDB); break;
      case Node::NWORDB: push(Inst::NWORDB); break;
      case Node::CAT:
        for (const auto& c : n.sub) emit(*c);
        break;
      case Node::ALT: {
        std::vector<uint32_t> jumps;
        for (size_t k = 0; k < n.sub.size(); ++k) {
          if (k + 1 < n.sub.size()) {
            const uint32_t sp = push(Inst::SPLIT);
            _prog[sp].x = here();
            emit(*n.sub[k]);
            jumps.push_back(push(Inst::JMP));
            _prog[sp].y = here();
          } else {
            emit(*n.sub[k]);
          }
        }
        for (uint32_t j : jumps) _prog[j].x = here();
        break;
      }
      case Node::REPEAT: {
        const Node& body = *n.sub[0];
        for (int k = 0; k < n.lo; ++k) emit(body);
        if (n.hi < 0) {  // L: split body, out ; body ; jmp L
          const uint32_t sp = push(Inst::SPLIT);
          const uint32_t b = here();
          emit(body);
          push(Inst::JMP, sp);
          set_split(sp, b, here(), n.greedy);
        } else {  // (body (body ...)?)? : every split leaves to the common end
          std::vector<uint32_t> splits;
          for (int k = n.lo; k < n.hi; ++k) {
            splits.push_back(push(Inst::SPLIT));
            emit(body);
          }
          for (uint32_t sp : splits) set_split(sp, sp + 1, here(), n.greedy);
        }
        break;
      }
    }
  }

  void set_split(uint32_t sp, uint32_t body, uint32_t out, bool greedy) {
    _prog[sp].x = greedy ? body : out;
    _prog[sp].y = greedy ? out : body;
  }

  // bytes a match can start with, when the start of the program holds no assertion and cannot match empty
  void find_first_bytes() {
    std::vector<uint32_t> stack{0};
    std::vector<uint8_t> seen(_prog.size(), 0);
    ByteSet first;
    bool usable = true;
    while (!stack.empty() && usable) {
      const uint32_t pc = stack.back();
      stack.pop_back();
      if (seen[pc]) continue;
      seen[pc] = 1;
      const Inst& i = _prog[pc];
      switch (i.op) {
        case Inst::BYTE: first.add(i.byte); break;
        case Inst::SET: first.merge(_sets[i.x]); break;
        case Inst::JMP: stack.push_back(i.x); break;
        case Inst::SPLIT: stack.push_back(i.y), stack.push_back(i.x); break;
        default: usable = false; break;  // assertion or MATCH
      }
    }
    _has_first = usable;
    if (usable)
      for (int c = 0; c < 256; ++c) _first[c] = first.has(unsigned(c));
  }

  friend class Dfa;
  std::vector<Inst> _prog;
  std::vector<ByteSet> _sets;
  bool _has_first = false;
  uint8_t _first[256] = {};
};

// ---- Pike VM -----------------------------------------------------------------------------------------------
namespace detail {

struct ThreadList {  // the program counters waiting at one text position, in priority order, each at most once
  std::vector<uint32_t> pc, start, where;  // where[pc] = index into pc[] when pc is a member
  uint32_t n = 0;
  void reset(size_t prog_size) {
    if (pc.size() < prog_size) pc.resize(prog_size), start.resize(prog_size), where.resize(prog_size, 0);
    n = 0;
  }
  bool has(uint32_t p) const { return where[p] < n && pc[where[p]] == p; }
  void add(uint32_t p, uint32_t s) {
    where[p] = n;
    pc[n] = p;
    start[n] = s;
    ++n;
  }
};

struct Scratch {
  ThreadList a, b;
  std::vector<uint32_t> stack;
};

inline Scratch& scratch() {
  thread_local Scratch s;
  return s;
}

}  // namespace detail

inline bool Program::search(const char* text, size_t n, size_t from, Span* out) const {
  const unsigned char* s = reinterpret_cast<const unsigned char*>(text);
  detail::Scratch& sc = detail::scratch();
  detail::ThreadList* clist = &sc.a;
  detail::ThreadList* nlist = &sc.b;
  clist->reset(_prog.size());
  nlist->reset(_prog.size());
  std::vector<uint32_t>& stack = sc.stack;

  // follow jumps, splits and assertions from pc at text position pos; consuming instructions (and MATCH) join l.
  // ThreadList marks every visited pc (also the non-consuming ones), which both keeps priorities -- the first
  // path to reach an instruction wins -- and ends empty loops like "(a*)*".
  auto add_thread = [&](detail::ThreadList* l, uint32_t pc0, uint32_t start, size_t pos) {
    stack.clear();
    stack.push_back(pc0);
    while (!stack.empty()) {
      const uint32_t pc = stack.back();
      stack.pop_back();
      if (l->has(pc)) continue;
      l->add(pc, start);
      const Inst& i = _prog[pc];
      switch (i.op) {
        case Inst::JMP:
          stack.push_back(i.x);
          break;
        case Inst::SPLIT:
          stack.push_back(i.y);
          stack.push_back(i.x);
          break;
        case Inst::BOL:
          if (pos == 0) stack.push_back(pc + 1);
          break;
        case Inst::EOL:
          if (pos == n) stack.push_back(pc + 1);
          break;
        case Inst::WORDB:
        case Inst::NWORDB: {
          const bool before = pos > 0 && is_word(s[pos - 1]), after = pos < n && is_word(s[pos]);
          if ((before != after) == (i.op == Inst::WORDB)) stack.push_back(pc + 1);
          break;
        }
        default:
          break;  // BYTE, SET, MATCH wait in the list
      }
    }
  };

  bool matched = false;
  for (size_t pos = from;; ++pos) {
    if (clist->n == 0) {
      if (matched) break;
      if (_has_first) {  // nothing in flight: skip to the next byte a match can start with
        while (pos < n && !_first[s[pos]]) ++pos;
        if (pos >= n) break;
      }
    }
    if (!matched) add_thread(clist, 0, static_cast<uint32_t>(pos), pos);  // a later start has the lowest priority
    nlist->n = 0;
    for (uint32_t t = 0; t < clist->n; ++t) {
      const uint32_t pc = clist->pc[t];
      const Inst& i = _prog[pc];
      if (i.op == Inst::MATCH) {
        out->begin = clist->start[t];
        out->end = pos;
        matched = true;
        break;  // threads behind this one have lower priority
      }
      if (pos < n && ((i.op == Inst::BYTE && s[pos] == i.byte) || (i.op == Inst::SET && _sets[i.x].has(s[pos]))))
        add_thread(nlist, pc + 1, clist->start[t], pos + 1);
    }
    std::swap(clist, nlist);
    if (pos >= n) break;
  }
  return matched;
}

// ---- lazy DFA: does the line hold a match at all? --------------------------------------------------------------
// For the line kinds (count_lines, line_byte_offsets, line_indices, lines) only the existence of a match counts, and
// existence does not depend on priorities: the set of program counters alive after each byte is a DFA state.  States
// and transitions are built on first use (the classic lazy subset construction, as in re2's DFA), one table row of
// 256 entries per state, at most kMaxStates states (then the cache is flushed).  Not thread-safe: one Dfa per
// verification thread; the Program is shared.
//
// Assertions (^ $ \b \B) look at the neighbours of a position, so a program that holds some runs a second form of
// the same construction (as re2's DFA does with its "needs flags" states): closure() STOPS at an assertion and keeps
// it in the state as pending; the state also carries what lies before the position (FLAG_START: nothing, FLAG_WORD:
// a word byte).  A transition on byte c first resolves the pending assertions with c as the byte behind the position
// (expand()), then consumes c; the end of the line resolves them once more with "nothing behind".  States without
// pending assertions do not depend on the flags and are stored with flags 0, so a program whose start holds no
// assertion keeps ONE start state and the first-byte skip.
class Dfa {
 public:
  explicit Dfa(const Program& p) : _p(p), _assert(p.has_assertions()) { reset(); }

  bool matches(const char* text, size_t n) {
    const unsigned char* s = reinterpret_cast<const unsigned char*>(text);
    if (_assert) return matches_assert(s, n);
    int32_t st = _start;
    if (_accept[size_t(st)]) return true;
    for (size_t i = 0; i < n; ++i) {
      if (st == _start && _p._has_first) {  // nothing in flight: skip to the next byte a match can start with
        while (i < n && !_p._first[s[i]]) ++i;
        if (i >= n) break;
      }
      int32_t nx = _next[size_t(st) * 256 + s[i]];  // the accept flag travels in the table entry: one load per byte
      if (nx < 0) nx = step(st, s[i]);
      if (nx & kAcceptBit) return true;
      st = nx;
    }
    return false;
  }

  size_t states() const { return _sets.size(); }

 private:
  static constexpr size_t kMaxStates = 2048;
  static constexpr int32_t kAcceptBit = 1 << 30;
  static constexpr uint8_t FLAG_WORD = 1, FLAG_START = 2;

  void reset() {
    _sets.clear();
    _accept.clear();
    _flags.clear();
    _eot.clear();
    _next.clear();
    _index.clear();
    _mark.assign(_p._prog.size(), 0);
    _epoch = 0;
    std::vector<uint32_t> set;
    bool acc = false;
    begin_set();
    closure(0, &set, &acc);
    _start = intern(std::move(set), acc, FLAG_START);
  }

  static bool is_assertion(const Inst& i) {
    return i.op == Inst::BOL || i.op == Inst::EOL || i.op == Inst::WORDB || i.op == Inst::NWORDB;
  }

  // ---- programs with assertions ----------------------------------------------------------------------------
  bool matches_assert(const unsigned char* s, size_t n) {
    int32_t st = _start;
    if (_accept[size_t(st)]) return true;
    for (size_t i = 0; i < n; ++i) {
      if (st == _start && _p._has_first) {  // the start state holds no assertion then: flags 0, one state
        while (i < n && !_p._first[s[i]]) ++i;
        if (i >= n) break;
      }
      int32_t nx = _next[size_t(st) * 256 + s[i]];
      if (nx < 0) nx = step_assert(st, s[i]);
      if (nx & kAcceptBit) return true;
      st = nx;
    }
    if (_eot[size_t(st)] < 0) {  // pending assertions against the end of the line
      _cons.clear();
      _eot[size_t(st)] = expand(st, /*eot=*/true, 0) ? 1 : 0;
    }
    return _eot[size_t(st)] != 0;
  }

  // Resolves the pending assertions of state st at its position, `c` being the byte behind it (eot: none).  The
  // consuming instructions alive there are left in _cons; true when MATCH is reached.
  bool expand(int32_t st, bool eot, unsigned char c) {
    const uint8_t flags = _flags[size_t(st)];
    const bool before = (flags & FLAG_WORD) != 0, after = !eot && is_word(c);
    bool acc = false;
    begin_set();
    _stack.clear();
    const std::vector<uint32_t>& set = _sets[size_t(st)];
    for (size_t k = set.size(); k-- > 0;) _stack.push_back(set[k]);
    while (!_stack.empty()) {
      const uint32_t pc = _stack.back();
      _stack.pop_back();
      if (_mark[pc] == _epoch) continue;
      _mark[pc] = _epoch;
      const Inst& i = _p._prog[pc];
      switch (i.op) {
        case Inst::JMP: _stack.push_back(i.x); break;
        case Inst::SPLIT: _stack.push_back(i.y), _stack.push_back(i.x); break;
        case Inst::MATCH: acc = true; break;
        case Inst::BOL:
          if (flags & FLAG_START) _stack.push_back(pc + 1);
          break;
        case Inst::EOL:
          if (eot) _stack.push_back(pc + 1);
          break;
        case Inst::WORDB:
        case Inst::NWORDB:
          if ((before != after) == (i.op == Inst::WORDB)) _stack.push_back(pc + 1);
          break;
        default: _cons.push_back(pc); break;  // BYTE, SET
      }
    }
    return acc;
  }

  int32_t step_assert(int32_t st, unsigned char c) {
    _cons.clear();
    if (expand(st, /*eot=*/false, c)) {  // a match ends at this position: the search is over
      _next[size_t(st) * 256 + c] = kAcceptBit;
      return kAcceptBit;
    }
    std::vector<uint32_t> set;
    bool acc = false;
    begin_set();
    for (uint32_t pc : _cons) {
      const Inst& i = _p._prog[pc];
      if ((i.op == Inst::BYTE && i.byte == c) || (i.op == Inst::SET && _p._sets[i.x].has(c))) closure(pc + 1, &set, &acc);
    }
    closure(0, &set, &acc);  // a match may start at the next byte as well
    if (_sets.size() >= kMaxStates) {
      std::vector<uint32_t> cur = _sets[size_t(st)];
      const bool cur_acc = _accept[size_t(st)] != 0;
      const uint8_t cur_flags = _flags[size_t(st)];
      reset();
      st = intern(std::move(cur), cur_acc, cur_flags);
    }
    int32_t nx = intern(std::move(set), acc, is_word(c) ? FLAG_WORD : 0);
    if (acc) nx |= kAcceptBit;
    _next[size_t(st) * 256 + c] = nx;
    return nx;
  }

  void begin_set() {
    if (++_epoch == 0) {  // wrapped: clear the marks
      std::fill(_mark.begin(), _mark.end(), 0);
      _epoch = 1;
    }
  }

  // consuming instructions reachable from pc without reading a byte; *acc when MATCH is among them
  void closure(uint32_t pc0, std::vector<uint32_t>* set, bool* acc) {
    _stack.clear();
    _stack.push_back(pc0);
    while (!_stack.empty()) {
      const uint32_t pc = _stack.back();
      _stack.pop_back();
      if (_mark[pc] == _epoch) continue;
      _mark[pc] = _epoch;
      const Inst& i = _p._prog[pc];
      switch (i.op) {
        case Inst::JMP: _stack.push_back(i.x); break;
        case Inst::SPLIT: _stack.push_back(i.y), _stack.push_back(i.x); break;
        case Inst::MATCH: *acc = true; break;
        default: set->push_back(pc); break;  // BYTE, SET; an assertion stays pending in the state (see expand)
      }
    }
  }

  int32_t intern(std::vector<uint32_t>&& set, bool acc, uint8_t flags = 0) {
    std::sort(set.begin(), set.end());
    bool pending = false;
    for (uint32_t pc : set) pending = pending || is_assertion(_p._prog[pc]);
    if (!pending) flags = 0;  // nothing in the state looks at its neighbours
    std::string key(reinterpret_cast<const char*>(set.data()), set.size() * sizeof(uint32_t));
    key.push_back(acc ? 1 : 0);
    key.push_back(static_cast<char>(flags));
    auto it = _index.find(key);
    if (it != _index.end()) return it->second;
    const int32_t id = static_cast<int32_t>(_sets.size());
    _sets.push_back(std::move(set));
    _accept.push_back(acc ? 1 : 0);
    _flags.push_back(flags);
    _eot.push_back(-1);
    _next.resize(_next.size() + 256, -1);
    _index.emplace(std::move(key), id);
    return id;
  }

  int32_t step(int32_t st, unsigned char c) {
    std::vector<uint32_t> set;
    bool acc = false;
    begin_set();
    for (uint32_t pc : _sets[size_t(st)]) {
      const Inst& i = _p._prog[pc];
      if ((i.op == Inst::BYTE && i.byte == c) || (i.op == Inst::SET && _p._sets[i.x].has(c))) closure(pc + 1, &set, &acc);
    }
    closure(0, &set, &acc);  // a match may start at the next byte as well
    if (_sets.size() >= kMaxStates) {  // cache full: start over (the current transition is rebuilt below)
      std::vector<uint32_t> cur = _sets[size_t(st)];
      const bool cur_acc = _accept[size_t(st)] != 0;
      reset();
      st = intern(std::move(cur), cur_acc);
    }
    int32_t nx = intern(std::move(set), acc);
    if (acc) nx |= kAcceptBit;  // the search ends there, so the flagged value is never used as a state
    _next[size_t(st) * 256 + c] = nx;
    return nx;
  }

  const Program& _p;
  const bool _assert;  // the program holds assertions: states carry flags, transitions resolve them (expand)
  int32_t _start = 0;
  std::vector<std::vector<uint32_t>> _sets;  // state -> its program counters (sorted)
  std::vector<uint8_t> _accept;
  std::vector<uint8_t> _flags;               // assertion programs: FLAG_* of the position the state stands at
  std::vector<int8_t> _eot;                  // assertion programs: does the state match at the end of the line (-1 = unknown)
  std::vector<uint32_t> _cons;               // scratch of expand()
  std::vector<int32_t> _next;                // state x byte -> state, -1 = not built yet
  std::unordered_map<std::string, int32_t> _index;
  std::vector<uint32_t> _mark, _stack;
  uint32_t _epoch = 0;
};

}  // namespace xs::regex_vm
