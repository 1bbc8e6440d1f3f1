// xsg_testdouble.cpp -- TEST DOUBLE of the part of the C ABI (include/xsg.h) that include/xsearch_b200/xsearch.hpp
// calls.  TEST INFRASTRUCTURE ONLY: it is built into tests/cpp/_build/ by tests/test_host_pipeline.py, linked into a
// second copy of the xsgrep driver there, and nothing of the product links or loads it (the product's libxsg.so has
// no CPU path and fails loudly without a CUDA device).
//
// Purpose: the HOST logic of xs::extern_search -- reader threads, segment forming with the v1 chunk rule, carries
// across segment edges, buffer recycling, the live result iterator, the regex host stage, the failure path -- can be
// exercised, differentially against the CPU oracle and under ThreadSanitizer, in a container without a GPU.  The
// double answers every segment with the oracle (oracle/xs_oracle.c, compiled in) and models the one property of the
// real stream interface that the host layer can get wrong:
//
//   * a pushed segment is read LATE: xsg_stream_push(k) only remembers the host pointer; the bytes are looked at
//     inside push(k+1) / xsg_stream_close, which is when the real library's scan of segment k waits for its H2D copy
//     (xsearch_b200/csrc/xsg_scan_api.inc, stream_scan_pending).  A checksum taken at push time is compared then: a
//     host layer that recycles a pinned buffer too early fails with XSG_ERR_STATE instead of passing by luck.
//
// The argument rules of xsg_stream_open / xsg_stream_push are the real library's.  Fault injection for the failure
// path: XSG_DOUBLE_FAIL_PUSH=k makes the k-th push (0-based) of every stream return XSG_ERR_CUDA.
// XSG_DOUBLE_NOSCAN=1 answers every segment with "no results" without reading it: what is left is the host pipeline
// alone (readers, carries, pushes), whose throughput can then be timed on any machine.
#include <xsg.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../oracle/xs_oracle.h"

namespace {

thread_local std::string g_error = "no error";
void set_error(const std::string& s) { g_error = s; }

constexpr uint32_t kAllKinds =
    XSG_COUNT | XSG_COUNT_LINES | XSG_MATCH_BYTE_OFFSETS | XSG_LINE_BYTE_OFFSETS | XSG_LINE_INDICES | XSG_LINES;

struct Part {
  uint64_t count = 0, count_lines = 0, newlines = 0;
  std::vector<uint64_t> match_offsets, line_offsets, line_indices, line_starts, line_lengths;
};

uint64_t checksum(const uint8_t* p, uint64_t n) {  // order-sensitive, cheap enough for test-sized segments
  uint64_t h = 0x9e3779b97f4a7c15ull ^ n;
  uint64_t i = 0;
  for (; i + 8 <= n; i += 8) {
    uint64_t w;
    std::memcpy(&w, p + i, 8);
    h = (h ^ w) * 0x100000001b3ull;
    h ^= h >> 29;
  }
  for (; i < n; ++i) h = (h ^ p[i]) * 0x100000001b3ull;
  return h;
}

}  // namespace

struct xsg_ctx {
  int device = 0;
};

struct xsg_result {
  uint32_t kinds = 0;
  float pipeline_ms = 0;  // open -> close, reported as xsg_timings.h2d_ms like the real stream does
  std::vector<Part> parts;
};

struct xsg_stream {
  std::string pat;  // lowered when icase
  uint32_t kinds = 0;
  bool icase = false;
  uint64_t goff_base = 0, lidx = 0, bytes_pushed = 0, pushed = 0;
  bool local_lidx = false;  // XSG_SCAN_LOCAL_LINE_INDEX
  bool last_terminated = true;
  long fail_push = -1;
  bool noscan = false;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  xsg_result* res = nullptr;
  // the pending segment (pushed, not scanned)
  bool pending = false;
  const uint8_t* pend_ptr = nullptr;
  uint64_t pend_n = 0, pend_off = 0, pend_sum = 0;
  std::vector<uint64_t> pend_bounds;
};

namespace {

int scan_pending(xsg_stream* st) {
  const uint8_t* raw = st->pend_ptr;
  const uint64_t n = st->pend_n;
  if (st->noscan) {
    st->res->parts.emplace_back();
    st->pending = false;
    return XSG_OK;
  }
  if (checksum(raw, n) != st->pend_sum) {
    set_error("test double: the host buffer of segment " + std::to_string(st->pushed - 1) +
              " changed between its push and its scan (recycled before the next push returned)");
    return XSG_ERR_STATE;
  }
  std::vector<char> lowered;
  const char* buf = reinterpret_cast<const char*>(raw);
  if (st->icase) {
    lowered.assign(buf, buf + n);
    xso_to_lower(lowered.data(), n);
    buf = lowered.data();
  }
  const char* pat = st->pat.data();
  const uint32_t m = static_cast<uint32_t>(st->pat.size());
  const uint64_t* b = st->pend_bounds.data();
  const uint64_t nb = st->pend_bounds.size();
  const uint64_t goff = st->goff_base + st->pend_off;
  st->res->parts.emplace_back();
  Part& p = st->res->parts.back();
  for (uint64_t i = 0; i < n; ++i) p.newlines += raw[i] == '\n';
  if (st->kinds & XSG_COUNT) p.count = xso_count(buf, n, pat, m, b, nb, 0, 1);
  if (st->kinds & XSG_COUNT_LINES) p.count_lines = xso_count(buf, n, pat, m, b, nb, 1, 1);
  auto fill = [&](std::vector<uint64_t>& v, auto fn, uint64_t add) {
    const uint64_t total = fn(buf, n, pat, m, b, nb, static_cast<uint64_t*>(nullptr), uint64_t(0), 1);
    v.resize(total);
    if (total) fn(buf, n, pat, m, b, nb, v.data(), total, 1);
    for (uint64_t& x : v) x += add;
  };
  if (st->kinds & XSG_MATCH_BYTE_OFFSETS) fill(p.match_offsets, xso_match_byte_offsets, goff);
  if (st->kinds & XSG_LINE_BYTE_OFFSETS) fill(p.line_offsets, xso_line_byte_offsets, goff);
  if (st->kinds & XSG_LINE_INDICES) fill(p.line_indices, xso_line_indices, st->lidx);
  if (st->kinds & XSG_LINES) {
    const uint64_t total = xso_lines(buf, n, pat, m, b, nb, nullptr, nullptr, 0, 1);
    p.line_starts.resize(total);
    p.line_lengths.resize(total);
    if (total) xso_lines(buf, n, pat, m, b, nb, p.line_starts.data(), p.line_lengths.data(), total, 1);
    for (uint64_t& x : p.line_starts) x += goff;
  }
  if (!st->local_lidx) st->lidx += p.newlines;
  st->pending = false;
  return XSG_OK;
}

const std::vector<uint64_t>* part_vector(const Part& p, uint32_t kind) {
  switch (kind) {
    case XSG_MATCH_BYTE_OFFSETS: return &p.match_offsets;
    case XSG_LINE_BYTE_OFFSETS: return &p.line_offsets;
    case XSG_LINE_INDICES: return &p.line_indices;
    case XSG_LINES: return &p.line_starts;
    default: return nullptr;
  }
}

int copy_out(const std::vector<uint64_t>& v, uint64_t* dst, uint64_t cap) {
  if (!dst && cap) return XSG_ERR_ARG;
  const uint64_t k = cap < v.size() ? cap : v.size();
  if (k) std::memcpy(dst, v.data(), k * sizeof(uint64_t));
  return XSG_OK;
}

}  // namespace

extern "C" {

const char* xsg_last_error(void) { return g_error.c_str(); }

int xsg_init(int device, xsg_ctx** out) {
  if (!out) return XSG_ERR_ARG;
  *out = new (std::nothrow) xsg_ctx();
  if (!*out) return XSG_ERR_NOMEM;
  (*out)->device = device;
  return XSG_OK;
}
int xsg_init_multi(const int* devices, int n, xsg_ctx** out) {  // any number of pretend devices
  if (!devices || !out || n <= 0) return XSG_ERR_ARG;
  for (int i = 0; i < n; ++i)
    if (xsg_init(devices[i], &out[i]) != XSG_OK) return XSG_ERR_NOMEM;
  return XSG_OK;
}
void xsg_destroy(xsg_ctx* ctx) { delete ctx; }

// plain heap memory stands in for cudaHostAlloc; freed buffers are poisoned so that a late reader sees garbage
int xsg_host_alloc(uint64_t n, void** out) {
  if (!out) return XSG_ERR_ARG;
  uint64_t* p = static_cast<uint64_t*>(std::malloc(n + 16));
  if (!p) {
    set_error("test double: out of memory");
    return XSG_ERR_NOMEM;
  }
  p[0] = n;
  *out = p + 2;
  return XSG_OK;
}
void xsg_host_free(void* q) {
  if (!q) return;
  uint64_t* p = static_cast<uint64_t*>(q) - 2;
  std::memset(q, 0xdb, p[0]);
  std::free(p);
}

int xsg_stream_open(xsg_ctx* ctx, const uint8_t* pattern, uint32_t m, uint32_t kinds, const xsg_scan_opts* opts,
                    xsg_stream** out) {
  if (!ctx || !out) return XSG_ERR_ARG;
  *out = nullptr;
  if (kinds == 0 || (kinds & ~kAllKinds)) {
    set_error("kinds must be a non-empty OR of XSG_* result kinds");
    return XSG_ERR_ARG;
  }
  if (!pattern || m == 0 || m > XSG_MAX_PATTERN) {
    set_error("pattern length must be 1..64");
    return XSG_ERR_ARG;
  }
  if (std::memchr(pattern, '\n', m)) {
    set_error("the pattern must not contain '\\n'");
    return XSG_ERR_ARG;
  }
  xsg_stream* st = new xsg_stream();
  st->pat.assign(reinterpret_cast<const char*>(pattern), m);
  st->icase = opts && (opts->flags & XSG_SCAN_IGNORE_CASE);
  if (st->icase) xso_to_lower(st->pat.data(), m);
  st->kinds = kinds;
  st->goff_base = opts ? opts->global_offset_base : 0;
  st->lidx = opts ? opts->line_index_base : 0;
  st->local_lidx = opts && (opts->flags & XSG_SCAN_LOCAL_LINE_INDEX);
  if (const char* f = std::getenv("XSG_DOUBLE_FAIL_PUSH")) st->fail_push = std::atol(f);
  st->noscan = std::getenv("XSG_DOUBLE_NOSCAN") != nullptr;
  st->res = new xsg_result();
  st->res->kinds = kinds;
  *out = st;
  return XSG_OK;
}

int xsg_stream_push(xsg_stream* st, const void* host, uint64_t n, const uint64_t* bounds, uint64_t nb, uint64_t* ticket) {
  return xsg_stream_push_at(st, host, n, bounds, nb, st ? st->bytes_pushed : 0, ticket);
}

int xsg_stream_push_at(xsg_stream* st, const void* host, uint64_t n, const uint64_t* bounds, uint64_t nb,
                       uint64_t stream_offset, uint64_t* ticket) {
  if (!st || !host || n == 0 || !bounds || nb == 0) {
    set_error("xsg_stream_push: a segment needs bytes and at least one chunk bound");
    return XSG_ERR_ARG;
  }
  const uint8_t* hbuf = static_cast<const uint8_t*>(host);
  if (!st->last_terminated) {
    set_error("xsg_stream_push: only the last segment of a stream may end without a newline");
    return XSG_ERR_BOUNDS;
  }
  uint64_t prev = 0;
  for (uint64_t i = 0; i < nb; ++i) {
    const uint64_t e = bounds[i];
    if (!(e > prev && e <= n && (i + 1 == nb ? e == n : hbuf[e - 1] == '\n'))) {
      set_error("chunk bounds must be increasing, end at the segment size, and every bound but the last must follow a '\\n'");
      return XSG_ERR_BOUNDS;
    }
    prev = e;
  }
  if (st->fail_push >= 0 && static_cast<uint64_t>(st->fail_push) == st->pushed) {
    set_error("test double: injected failure of push " + std::to_string(st->pushed));
    return XSG_ERR_CUDA;
  }
  if (st->pending) {
    const int rc = scan_pending(st);
    if (rc) return rc;
  }
  st->pend_ptr = hbuf;
  st->pend_n = n;
  st->pend_off = stream_offset;
  st->pend_sum = st->noscan ? 0 : checksum(hbuf, n);
  st->pend_bounds.assign(bounds, bounds + nb);
  st->pending = true;
  st->last_terminated = hbuf[n - 1] == '\n';
  if (ticket) *ticket = st->pushed;
  st->pushed += 1;
  st->bytes_pushed = stream_offset + n;
  return XSG_OK;
}

const xsg_result* xsg_stream_result(const xsg_stream* st) { return st ? st->res : nullptr; }

int xsg_stream_close(xsg_stream* st, xsg_result** out) {
  if (!st || !out) return XSG_ERR_ARG;
  *out = nullptr;
  if (st->pending) {
    const int rc = scan_pending(st);
    if (rc) {
      delete st->res;
      delete st;
      return rc;
    }
  }
  st->res->pipeline_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - st->t0).count();
  *out = st->res;
  delete st;
  return XSG_OK;
}

void xsg_stream_abort(xsg_stream* st) {
  if (!st) return;
  delete st->res;
  delete st;
}

uint64_t xsg_result_num_parts(const xsg_result* r) { return r ? r->parts.size() : 0; }

uint64_t xsg_result_part_count(const xsg_result* r, uint64_t part, uint32_t kind) {
  if (!r || part >= r->parts.size() || !(r->kinds & kind)) return 0;
  const Part& p = r->parts[part];
  if (kind == XSG_COUNT) return p.count;
  if (kind == XSG_COUNT_LINES) return p.count_lines;
  const std::vector<uint64_t>* v = part_vector(p, kind);
  return v ? v->size() : 0;
}

int xsg_result_part_copy(const xsg_result* r, uint64_t part, uint32_t kind, uint64_t* dst, uint64_t cap) {
  if (!r || part >= r->parts.size() || !(r->kinds & kind)) return XSG_ERR_ARG;
  const std::vector<uint64_t>* v = part_vector(r->parts[part], kind);
  return v ? copy_out(*v, dst, cap) : XSG_ERR_ARG;
}

int xsg_result_part_copy_line_lengths(const xsg_result* r, uint64_t part, uint64_t* dst, uint64_t cap) {
  if (!r || part >= r->parts.size() || !(r->kinds & XSG_LINES)) return XSG_ERR_ARG;
  return copy_out(r->parts[part].line_lengths, dst, cap);
}

uint64_t xsg_result_part_newlines(const xsg_result* r, uint64_t part) {
  return r && part < r->parts.size() ? r->parts[part].newlines : 0;
}

int xsg_result_timings(const xsg_result* r, xsg_timings* t) {
  if (!r || !t) return XSG_ERR_ARG;
  std::memset(t, 0, sizeof *t);
  t->h2d_ms = r->pipeline_ms;
  return XSG_OK;
}

void xsg_result_free(xsg_result* r) { delete r; }

}  // extern "C"
