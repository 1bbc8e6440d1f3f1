/*
 * xs_oracle.c -- CPU restatement of x-search's chunked literal scan-and-match.
 * TEST INFRASTRUCTURE ONLY (see xs_oracle.h).  Parity status: PINNED against the
 * reference's KATs and against oracle/_ref/libxsref.so (the reference's own sources).
 *
 * Build: gcc -O2 -fPIC -shared -pthread -o libxsoracle.so xs_oracle.c   (oracle/Makefile)
 */
#define _GNU_SOURCE
#include "xs_oracle.h"

#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <string.h>

/* ---- L0: matcher model ------------------------------------------------------------------- */

/* reference: simd_search.cpp:58-78 scalar_strstr.  On a partial match of k bytes the
 * reference advances to shift + k + 1 (line 75: shift = str_index), which can skip a real
 * occurrence; reproduced deliberately. */
static int64_t scalar_tail(const char* b, size_t r, const char* p, size_t m) {
  size_t s = 0;
  while (s < r) {
    if (r - s < m) return -1;
    size_t k = 0;
    while (k < m && b[s + k] == p[k]) k++;
    if (k == m) return (int64_t)s;
    s = s + k + 1;
  }
  return -1;
}

int64_t xso_find_next(const char* pat, size_t m, const char* buf, size_t n, size_t shift) {
  if (shift > n || m == 0) return -1; /* simd_search.cpp:290-292 */
  size_t L = n - shift;
  const char* str = buf + shift;
  if (m == 1) { /* simd_search.cpp:164-166 -> strchr :116-144 (exact incl. tail :107-114) */
    const char* r = (const char*)memchr(str, (unsigned char)pat[0], L);
    return r ? (int64_t)(r - buf) : -1;
  }
  /* AVX2 loop runs while remaining >= 32 + m (:169,179), 32 bytes per iteration (:199-201);
   * every candidate in a processed block is verified with memcmp (:192) => exact. */
  size_t K = (L >= 32 + m) ? (L - m) / 32 : 0;
  size_t T = 32 * K;
  if (T > 0) {
    const char* r = (const char*)memmem(str, T + m - 1, pat, m);
    if (r) return (int64_t)(r - buf);
  }
  int64_t t = scalar_tail(str + T, L - T, pat, m); /* :171 / :203 */
  return t < 0 ? -1 : (int64_t)(shift + T) + t;
}

void xso_to_lower(char* buf, uint64_t n) {
  /* string_utils.cpp:14-25: signed compares data > 'A'-1 and 'Z'+1 > data, then += 'a'-'A'; the scalar
   * tail (:28-32) calls std::tolower, which in the "C" locale maps exactly the same 26 bytes */
  for (uint64_t i = 0; i < n; i++) {
    signed char c = (signed char)buf[i];
    if (c > 'A' - 1 && 'Z' + 1 > c) buf[i] = (char)(c + ('a' - 'A'));
  }
}

int64_t xso_find_next_newline(const char* buf, size_t n, size_t shift) {
  if (shift > n) return -1; /* simd_search.cpp:298-300 */
  const char* r = (const char*)memchr(buf + shift, '\n', n - shift);
  return r ? (int64_t)(r - buf) : -1;
}

uint64_t xso_count_matches(const char* pat, size_t m, const char* buf, size_t n) {
  uint64_t c = 0; /* simd_search.cpp:324-336 */
  size_t shift = 0;
  for (;;) {
    int64_t q = xso_find_next(pat, m, buf, n, shift);
    if (q < 0) break;
    c++;
    shift = (size_t)q + m;
  }
  return c;
}

uint64_t xso_count_matching_lines(const char* pat, size_t m, const char* buf, size_t n) {
  uint64_t c = 0; /* simd_search.cpp:305-322 */
  size_t shift = 0;
  for (;;) {
    int64_t q = xso_find_next(pat, m, buf, n, shift);
    if (q < 0) break;
    c++;
    shift = (size_t)q + m;
    q = xso_find_next_newline(buf, n, shift);
    if (q < 0) break;
    shift = (size_t)q + 1;
  }
  return c;
}

/* ---- chunking ---------------------------------------------------------------------------- */

uint64_t xso_chunk_bounds(const char* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap) {
  uint64_t nb = 0, start = 0;
  if (min_chunk == 0) min_chunk = 1;
  while (start < n) {
    uint64_t end;
    if (n - start <= min_chunk) {
      end = n;
    } else {
      /* >= min_chunk bytes, then through the next '\n' */
      const char* r = (const char*)memchr(buf + start + min_chunk - 1, '\n', n - (start + min_chunk - 1));
      end = r ? (uint64_t)(r - buf) + 1 : n;
    }
    if (bounds && nb < cap) bounds[nb] = end;
    nb++;
    start = end;
  }
  return nb;
}

/* ---- L1: per-chunk wrappers, restated ---------------------------------------------------- */

enum { K_COUNT = 0, K_MATCH_OFF, K_LINE_OFF, K_LINE_IDX, K_LINES };

typedef struct {
  uint64_t* a; /* primary values (chunk-local unless noted) */
  uint64_t* b; /* lens for K_LINES */
  uint64_t n, cap;
  uint64_t count;    /* for K_COUNT */
  uint64_t newlines; /* '\n' in the chunk (K_LINE_IDX) */
} part_t;

static void part_push(part_t* p, uint64_t a, uint64_t b, int want_b) {
  if (p->n == p->cap) {
    p->cap = p->cap ? p->cap * 2 : 64;
    p->a = (uint64_t*)realloc(p->a, p->cap * sizeof(uint64_t));
    if (want_b) p->b = (uint64_t*)realloc(p->b, p->cap * sizeof(uint64_t));
  }
  p->a[p->n] = a;
  if (want_b) p->b[p->n] = b;
  p->n++;
}

/* reference: search_wrappers.h:111-123 previous_new_line_offset_relative_to_match, folded
 * into "start of the line containing q". */
static uint64_t line_start(const char* d, uint64_t q) {
  const char* r = (const char*)memrchr(d, '\n', q); /* pattern holds no '\n', so d[q] != '\n' */
  return r ? (uint64_t)(r - d) + 1 : 0;
}

static uint64_t count_nl(const char* d, uint64_t from, uint64_t to) {
  uint64_t c = 0;
  const char* p = d + from;
  const char* e = d + to;
  while (p < e) {
    const char* r = (const char*)memchr(p, '\n', (size_t)(e - p));
    if (!r) break;
    c++;
    p = r + 1;
  }
  return c;
}

/* One chunk, one result kind.  Mirrors the loops at search_wrappers.h:36-50 (_byte_offsets),
 * :168-183 (count) and :192-205 (line). */
static void search_chunk(const char* d, uint64_t n, const char* pat, uint32_t m, int kind, int skip_to_nl,
                         part_t* out) {
  size_t shift = 0;
  uint64_t nl_seen = 0, nl_pos = 0; /* incremental newline count for K_LINE_IDX */
  while (shift < n) {
    int64_t q = xso_find_next(pat, m, d, n, shift);
    if (q < 0) break;
    if (kind == K_LINES) {
      uint64_t ls = line_start(d, (uint64_t)q);
      shift = (size_t)q + m;
      int64_t e = xso_find_next_newline(d, n, shift);
      if (e < 0) break; /* :199-201: unterminated last line is dropped */
      shift = (size_t)e + 1;
      part_push(out, ls, (uint64_t)e - ls, 1);
      continue;
    }
    if (kind == K_COUNT) {
      out->count++;
    } else if (kind == K_MATCH_OFF) {
      part_push(out, (uint64_t)q, 0, 0);
    } else {
      uint64_t ls = line_start(d, (uint64_t)q);
      if (kind == K_LINE_OFF) {
        part_push(out, ls, 0, 0);
      } else {
        nl_seen += count_nl(d, nl_pos, ls);
        nl_pos = ls;
        part_push(out, nl_seen, 0, 0);
      }
    }
    shift = (size_t)q + m;
    if (skip_to_nl) {
      int64_t e = xso_find_next_newline(d, n, shift);
      if (e < 0) break;
      shift = (size_t)e + 1;
    }
  }
  if (kind == K_LINE_IDX) out->newlines = nl_seen + count_nl(d, nl_pos, n);
}

/* ---- driver: chunks -> (threads) -> globalised, file-ordered result ------------------------ */

typedef struct {
  const char* buf;
  const char* pat;
  uint32_t m;
  const uint64_t* bounds;
  uint64_t nb;
  int kind, skip_to_nl;
  part_t* parts;
  atomic_ullong next;
} job_t;

static void* worker(void* arg) {
  job_t* j = (job_t*)arg;
  for (;;) {
    uint64_t c = atomic_fetch_add(&j->next, 1);
    if (c >= j->nb) break;
    uint64_t b0 = c ? j->bounds[c - 1] : 0, b1 = j->bounds[c];
    search_chunk(j->buf + b0, b1 - b0, j->pat, j->m, j->kind, j->skip_to_nl, &j->parts[c]);
  }
  return NULL;
}

static uint64_t run(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds, uint64_t nb,
                    int kind, int skip_to_nl, uint64_t* out_a, uint64_t* out_b, uint64_t cap, int nthreads) {
  (void)n;
  if (m == 0 || nb == 0) return 0;
  job_t j;
  j.buf = buf; j.pat = pat; j.m = m; j.bounds = bounds; j.nb = nb;
  j.kind = kind; j.skip_to_nl = skip_to_nl;
  j.parts = (part_t*)calloc(nb, sizeof(part_t));
  atomic_init(&j.next, 0);
  if (nthreads <= 1) {
    worker(&j);
  } else {
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, worker, &j);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th);
  }
  uint64_t total = 0, nl_base = 0;
  for (uint64_t c = 0; c < nb; c++) {
    part_t* p = &j.parts[c];
    uint64_t base = c ? bounds[c - 1] : 0;
    if (kind == K_COUNT) {
      total += p->count;
    } else {
      for (uint64_t i = 0; i < p->n; i++, total++) {
        if (total >= cap) continue;
        if (kind == K_LINE_IDX) out_a[total] = p->a[i] + nl_base;
        else out_a[total] = p->a[i] + base;
        if (kind == K_LINES) out_b[total] = p->b[i];
      }
      nl_base += p->newlines;
    }
    free(p->a);
    free(p->b);
  }
  free(j.parts);
  return total;
}

uint64_t xso_count(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds, uint64_t nb,
                   int skip_to_nl, int nthreads) {
  return run(buf, n, pat, m, bounds, nb, K_COUNT, skip_to_nl, NULL, NULL, 0, nthreads);
}
uint64_t xso_match_byte_offsets(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds,
                                uint64_t nb, uint64_t* out, uint64_t cap, int nthreads) {
  return run(buf, n, pat, m, bounds, nb, K_MATCH_OFF, 0, out, NULL, cap, nthreads);
}
uint64_t xso_line_byte_offsets(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds,
                               uint64_t nb, uint64_t* out, uint64_t cap, int nthreads) {
  return run(buf, n, pat, m, bounds, nb, K_LINE_OFF, 1, out, NULL, cap, nthreads);
}
uint64_t xso_line_indices(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds,
                          uint64_t nb, uint64_t* out, uint64_t cap, int nthreads) {
  return run(buf, n, pat, m, bounds, nb, K_LINE_IDX, 1, out, NULL, cap, nthreads);
}
uint64_t xso_lines(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds, uint64_t nb,
                   uint64_t* starts, uint64_t* lens, uint64_t cap, int nthreads) {
  return run(buf, n, pat, m, bounds, nb, K_LINES, 1, starts, lens, cap, nthreads);
}
