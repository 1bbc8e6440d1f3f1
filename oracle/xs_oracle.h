/*
 * xs_oracle.h -- CPU restatement of x-search's chunked literal scan-and-match.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and only as the checker or the timed CPU baseline.
 *
 * Parity status: PINNED.  This restatement is checked (tests/test_oracle.py)
 *   (1) against the reference's own known-answer tests
 *       (test/src/string_search/simd_searchTest.cpp:63-98,
 *        test/src/string_search/search_wrappersTest.cpp:26,39,52), and
 *   (2) differentially against the reference's unmodified sources compiled into
 *       oracle/_ref/libxsref.so (oracle/Makefile, oracle/ref_shim.cpp).
 *
 * All functions work on one in-memory corpus buf[0..n) split into "reference
 * chunks" by a bounds array (bounds[i] = exclusive end offset of chunk i,
 * bounds[nb-1] == n).  Each chunk is searched exactly as the reference searches
 * one DataChunk; results are globalised (offset + chunk base) and concatenated
 * in file order.
 */
#ifndef XS_ORACLE_H
#define XS_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* reference: src/string_search/simd_search.cpp:289-295 (findNext) incl. the
 * AVX2 main loop :162-204 and the scalar tail :58-78. Returns offset or -1. */
int64_t xso_find_next(const char* pat, size_t m, const char* buf, size_t n, size_t shift);
/* reference: simd_search.cpp:297-303 (findNextNewLine) / :116-144 (strchr). */
int64_t xso_find_next_newline(const char* buf, size_t n, size_t shift);
/* reference: simd_search.cpp:324-336 (countMatches) / :305-322 (countMatchingLines) */
uint64_t xso_count_matches(const char* pat, size_t m, const char* buf, size_t n);
uint64_t xso_count_matching_lines(const char* pat, size_t m, const char* buf, size_t n);

/* reference: src/utils/string_utils.cpp:11-33 (simd::toLower), the v1 "tolower" chunk processor behind
 * ignore_case: ASCII 'A'..'Z' -> 'a'..'z' in place, every other byte (incl. >= 0x80) unchanged. */
void xso_to_lower(char* buf, uint64_t n);

/* v1 chunking rule decoded from test/files/sample.meta: each
 * chunk is >= min_chunk bytes and extended through the next '\n' (or to n).
 * Writes exclusive end offsets; returns the number of chunks (0 for n == 0).
 * If bounds == NULL only counts. */
uint64_t xso_chunk_bounds(const char* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap);

/* result kinds (README.md:72-81 tags). Each returns the TOTAL number of results
 * and writes at most cap of them. nthreads > 1 distributes chunks over threads
 * (used only by the timed CPU baseline). */
/* xs::count (skip_to_nl = 0) / xs::count_lines (skip_to_nl = 1): search_wrappers.h:163-185 */
uint64_t xso_count(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds,
                   uint64_t nb, int skip_to_nl, int nthreads);
/* xs::match_byte_offsets: search_wrappers.h:136-139 (skip_to_nl=false) */
uint64_t xso_match_byte_offsets(const char* buf, uint64_t n, const char* pat, uint32_t m,
                                const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap,
                                int nthreads);
/* xs::line_byte_offsets: search_wrappers.h:149-154 (+ :111-123) */
uint64_t xso_line_byte_offsets(const char* buf, uint64_t n, const char* pat, uint32_t m,
                               const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap,
                               int nthreads);
/* xs::line_indices: 0-based number of '\n' before the matching line
 * (v1 golden vectors test/src/xsearchTest.cpp:95-101). */
uint64_t xso_line_indices(const char* buf, uint64_t n, const char* pat, uint32_t m,
                          const uint64_t* bounds, uint64_t nb, uint64_t* out, uint64_t cap,
                          int nthreads);
/* xs::lines: search_wrappers.h:187-207; emitted as (start, length) pairs, '\n'
 * excluded, an unterminated last line is dropped. */
uint64_t xso_lines(const char* buf, uint64_t n, const char* pat, uint32_t m, const uint64_t* bounds,
                   uint64_t nb, uint64_t* starts, uint64_t* lens, uint64_t cap, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
