/*
 * xsg.h -- C ABI of the B200 scan-and-match library (libxsg.so).
 *
 * This is the drop-in boundary: plain C types, opaque handles, 0 / negative status codes.
 * Everything below replaces, for the literal (non-regex, case-sensitive) path, what the
 * reference does on CPU threads:
 *
 *   xsg_scan / xsg_search_host
 *        <- the per-chunk matchers   src/string_search/simd_search.cpp:116-204,289-336
 *        <- the chunk wrappers       include/xsearch/string_search/search_wrappers.h:29-52,111-207
 *        <- the searcher functors    include/xsearch/tasks/searchers.h:38-92
 *        <- the worker loop          include/xsearch/Searcher.h:100-120  (chunk -> search -> result)
 *   xsg_result_*
 *        <- the result collector     include/xsearch/ResultTypes.h:31-130
 *   xsg_corpus_chunk / xsg_chunk_bounds_host
 *        <- the v1 chunker (>= 16 MiB, extended through the next '\n'); evidence
 *           test/files/sample.meta + metafile_cat.cpp:39-49
 *   xsg_search_host
 *        <- the reader stage         include/xsearch/tasks/readers.h:39-48 (pinned, double-buffered H2D)
 *
 * The C++ surface users see (xs::extern_search<Tag>, Searcher::join(), getResult()) lives in
 * include/xsearch_b200/xsearch.hpp and only calls the functions declared here.
 *
 * Semantics are bit-exact with the reference, including the scalar-tail quirk of
 * simd::strstr in the last 31+m bytes of every reference chunk (simd_search.cpp:58-78,169-171,203).
 * Requirements: 1 <= m <= XSG_MAX_PATTERN, the pattern holds no '\n', and every reference-chunk
 * boundary except the last lies directly behind a '\n' (the v1 chunker guarantees it).
 *
 * Threading: a context may be driven by one host thread at a time.
 */
#ifndef XSG_H
#define XSG_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XSG_MAX_PATTERN 64u
#define XSG_DEFAULT_MIN_CHUNK 16777216ull /* v1 chunk size, SURVEY.md 3.5 */

/* result kinds == the reference's tag types (README.md:72-81) */
enum {
  XSG_COUNT = 1u << 0,              /* xs::count            : greedy non-overlapping matches       */
  XSG_COUNT_LINES = 1u << 1,        /* xs::count_lines      : lines holding >= 1 match             */
  XSG_MATCH_BYTE_OFFSETS = 1u << 2, /* xs::match_byte_offsets : global offset of every match       */
  XSG_LINE_BYTE_OFFSETS = 1u << 3,  /* xs::line_byte_offsets  : global offset of matching lines    */
  XSG_LINE_INDICES = 1u << 4,       /* xs::line_indices       : 0-based index of matching lines    */
  XSG_LINES = 1u << 5               /* xs::lines : (start, length) of matching lines, '\n' excluded */
};

/* status codes */
enum {
  XSG_OK = 0,
  XSG_ERR_CUDA = -1,     /* CUDA runtime failure, see xsg_last_error() */
  XSG_ERR_ARG = -2,      /* bad argument (m == 0, m > 64, '\n' in pattern, unknown kind ...) */
  XSG_ERR_BOUNDS = -3,   /* chunk bounds not increasing / not newline-aligned / not ending at n */
  XSG_ERR_NOMEM = -4,    /* device or host allocation failed */
  XSG_ERR_STATE = -5     /* call order (e.g. scan before bounds were set) */
};

typedef struct xsg_ctx xsg_ctx;       /* one device + stream + scratch */
typedef struct xsg_corpus xsg_corpus; /* device-resident bytes + reference-chunk bounds */
typedef struct xsg_result xsg_result; /* device-resident result vectors of one scan */
typedef struct xsg_stream xsg_stream; /* a streamed search: segments are pushed, result parts appear */

typedef struct xsg_scan_opts {
  uint64_t global_offset_base; /* added to every byte offset (shard / segment base)           */
  uint64_t line_index_base;    /* added to every line index ('\n' count before this corpus)   */
  uint64_t occ_capacity;       /* 0 = automatic; max raw occurrences buffered on the device   */
  uint32_t flags;              /* OR of XSG_SCAN_* flags                                      */
  uint32_t reserved;
} xsg_scan_opts;
/* ignore_case of the reference's one-call API: v1 ran a "tolower" processor over every chunk
 * (simd::toLower, src/utils/string_utils.cpp:11-33: ASCII A-Z only) and searched the lowered pattern in
 * the lowered bytes.  Here the fold happens in registers while the bytes are scanned. */
#define XSG_SCAN_IGNORE_CASE 1u
/* xsg_scan only: the scan's last kernel stores the scan's summary into every rank's table of the exchange attached
 * to the context (xsg_exchange_create) -- the publish step of the multi-GPU exchange without a launch of its own;
 * replaces the xsg_exchange_push() call of that round */
#define XSG_SCAN_PUBLISH 2u
/* streams only: the line indices of every part count from the part's own first byte (+ line_index_base) instead of
 * from the first byte of the stream -- for callers that deal the segments of one file to several streams/devices
 * and add each part's base themselves (the '\n' count of the earlier parts, xsg_result_part_newlines) */
#define XSG_SCAN_LOCAL_LINE_INDEX 4u

typedef struct xsg_timings {
  float scan_ms;        /* all device work of the scan (CUDA events on the library's stream) */
  float main_kernel_ms; /* the byte-scanning kernel alone                                     */
  float post_ms;        /* list-domain kernels (tail emulation, line selection, compaction)   */
  float h2d_ms;         /* xsg_search_host only: wall time of the streamed pipeline           */
  uint32_t kernel_launches; /* kernels launched by the library for this scan                  */
  uint32_t reserved;
} xsg_timings;

/* parameters of the deterministic synthetic corpus (block-wise pure function of seed, block) */
typedef struct xsg_synth_params {
  uint64_t seed;
  uint32_t needle_len;           /* 0 = no needle */
  uint8_t needle[XSG_MAX_PATTERN];
  uint32_t needle_rate_u32;      /* per line: inject when rand32 < rate (0 = never) */
} xsg_synth_params;
#define XSG_SYNTH_BLOCK 4096u /* every synthetic block is this long and ends with '\n' */

const char* xsg_last_error(void);
int xsg_version(void);

/* ---- context -------------------------------------------------------------------------- */
int xsg_init(int device, xsg_ctx** out);
/* one host process, n devices: out[0..n) = one context per device, peer access enabled between them where the
 * hardware allows (NVLink / NVSwitch); nothing is left behind on failure.  Each context is destroyed on its own. */
int xsg_init_multi(const int* devices, int n, xsg_ctx** out);
void xsg_destroy(xsg_ctx* ctx);
int xsg_device_info(xsg_ctx* ctx, int* sm_count, uint64_t* total_mem, char* name, size_t name_cap);
/* run all device work of this context on a caller-owned CUDA stream (a cudaStream_t passed as
 * void*; NULL = back to the context's own stream).  Lets the caller bracket scans with its own
 * events and order them with its own work (e.g. an NCCL exchange of shard counts). */
int xsg_set_stream(xsg_ctx* ctx, void* cuda_stream);

/* pinned host memory for xsg_search_host / xsg_corpus_upload sources (cudaHostAlloc) */
int xsg_host_alloc(uint64_t n, void** out);
void xsg_host_free(void* p);

/* ---- corpus (device-resident) ----------------------------------------------------------- */
int xsg_corpus_alloc(xsg_ctx* ctx, uint64_t n, xsg_corpus** out);
/* borrow caller-owned device memory (16-byte aligned, readable up to the next 16-byte boundary) */
int xsg_corpus_wrap(xsg_ctx* ctx, const void* dev_ptr, uint64_t n, xsg_corpus** out);
void xsg_corpus_free(xsg_corpus* c);
uint64_t xsg_corpus_size(const xsg_corpus* c);
const void* xsg_corpus_device_ptr(const xsg_corpus* c);
int xsg_corpus_upload(xsg_corpus* c, const void* host, uint64_t offset, uint64_t n);
int xsg_corpus_download(const xsg_corpus* c, void* host, uint64_t offset, uint64_t n);
/* fill [0, size) with synthetic blocks first_block, first_block+1, ... (size % XSG_SYNTH_BLOCK == 0) */
int xsg_corpus_fill_synth(xsg_corpus* c, uint64_t first_block, const xsg_synth_params* p);
/* reference chunks: either computed on the device with the v1 rule, or supplied */
int xsg_corpus_chunk(xsg_corpus* c, uint64_t min_chunk);
int xsg_corpus_set_bounds(xsg_corpus* c, const uint64_t* bounds, uint64_t nb);
uint64_t xsg_corpus_num_chunks(const xsg_corpus* c);
int xsg_corpus_get_bounds(const xsg_corpus* c, uint64_t* bounds, uint64_t cap);

/* ---- host-side helpers (same arithmetic as the device paths; no GPU needed) --------------- */
void xsg_synth_host(void* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params* p);
uint64_t xsg_chunk_bounds_host(const void* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap);

/* ---- scan ------------------------------------------------------------------------------- */
/* device-resident corpus; kinds = OR of XSG_* result kinds; opts may be NULL */
int xsg_scan(xsg_ctx* ctx, const xsg_corpus* c, const uint8_t* pattern, uint32_t m, uint32_t kinds,
             const xsg_scan_opts* opts, xsg_result** out);
/* host buffer (pinned or pageable), streamed through double-buffered device segments of whole
 * reference chunks; bounds may be NULL (v1 rule with min_chunk, 0 = default) */
int xsg_search_host(xsg_ctx* ctx, const void* host, uint64_t n, const uint64_t* bounds, uint64_t nb,
                    uint64_t min_chunk, const uint8_t* pattern, uint32_t m, uint32_t kinds,
                    const xsg_scan_opts* opts, xsg_result** out);
/* the same search on n_ctx devices driven by ONE host thread (the reference's num_threads workers over the chunks
 * of one file, include/xsearch/Searcher.h:141-145, README.md:72): the segments are dealt round-robin to the devices,
 * every device copies and scans its own, and ONE result comes back with its parts in file order, byte offsets and
 * line indices global (a part's line-index base is the '\n' count of the parts before it, known on the host one
 * segment later: no collective).  The result belongs to ctxs[0]; all contexts must outlive it. */
int xsg_search_host_multi(xsg_ctx* const* ctxs, int n_ctx, const void* host, uint64_t n, const uint64_t* bounds,
                          uint64_t nb, uint64_t min_chunk, const uint8_t* pattern, uint32_t m, uint32_t kinds,
                          const xsg_scan_opts* opts, xsg_result** out);

/* ---- streamed search (the reference's reader -> searcher -> result loop, Searcher.h:100-120,
 *      tasks/readers.h:39-48, as a push interface) -------------------------------------------------
 * A segment is a contiguous run of whole reference chunks: n host bytes (pinned memory for full
 * speed) and the exclusive chunk ends RELATIVE to the segment start (bounds[nb-1] == n; every bound
 * but the very last of the stream follows a '\n').  xsg_stream_push enqueues the H2D copy of the
 * segment on the copy stream, then scans the PREVIOUS segment while that copy runs, and returns a
 * ticket.  The host buffer may be reused once xsg_stream_wait_copied(ticket) has returned.  Results
 * of finished segments are visible as parts of xsg_stream_result() (borrowed; parts only grow) --
 * this is what the live iterator of the C++ layer reads.  xsg_stream_close scans the last segment,
 * hands the complete result to the caller (who frees it) and destroys the stream. */
int xsg_stream_open(xsg_ctx* ctx, const uint8_t* pattern, uint32_t m, uint32_t kinds, const xsg_scan_opts* opts,
                    xsg_stream** out);
int xsg_stream_push(xsg_stream* st, const void* host, uint64_t n, const uint64_t* bounds, uint64_t nb,
                    uint64_t* ticket);
/* the same with the segment's offset in the stream given by the caller (xsg_stream_push continues behind the previous
 * segment): segments of one file dealt to several streams keep their global byte offsets */
int xsg_stream_push_at(xsg_stream* st, const void* host, uint64_t n, const uint64_t* bounds, uint64_t nb,
                       uint64_t stream_offset, uint64_t* ticket);
int xsg_stream_wait_copied(xsg_stream* st, uint64_t ticket);
const xsg_result* xsg_stream_result(const xsg_stream* st);
int xsg_stream_close(xsg_stream* st, xsg_result** out);
void xsg_stream_abort(xsg_stream* st); /* destroy without finishing */

/* ---- results ---------------------------------------------------------------------------- */
uint64_t xsg_result_count(const xsg_result* r, uint32_t kind);
/* copy up to cap values of one vector kind to host memory (XSG_LINES: line starts) */
int xsg_result_copy(const xsg_result* r, uint32_t kind, uint64_t* dst, uint64_t cap);
int xsg_result_copy_line_lengths(const xsg_result* r, uint64_t* dst, uint64_t cap);
/* part-wise access (one part per scanned segment, in file order) */
uint64_t xsg_result_num_parts(const xsg_result* r);
uint64_t xsg_result_part_count(const xsg_result* r, uint64_t part, uint32_t kind);
int xsg_result_part_copy(const xsg_result* r, uint64_t part, uint32_t kind, uint64_t* dst, uint64_t cap);
int xsg_result_part_copy_line_lengths(const xsg_result* r, uint64_t part, uint64_t* dst, uint64_t cap);
/* '\n' in the bytes of one part (valid when XSG_LINE_INDICES was asked for): lets a host stage that adds lines of
 * its own -- the regex prefilter of xsearch_b200/regex_prefilter.hpp -- number them like the device does */
uint64_t xsg_result_part_newlines(const xsg_result* r, uint64_t part);
/* device pointer of a result vector (stays valid until xsg_result_free) */
const uint64_t* xsg_result_device_ptr(const xsg_result* r, uint32_t kind);
uint64_t xsg_result_newlines(const xsg_result* r);        /* '\n' in the scanned bytes (line kinds) */
uint64_t xsg_result_raw_occurrences(const xsg_result* r); /* all true occurrences seen            */
int xsg_result_timings(const xsg_result* r, xsg_timings* t);
/* multi-GPU fix-up: add offset_delta to every byte offset and line_index_delta to
 * every line index of a finished result, on the device -- the global line-number base of a shard
 * is only known after the exchange of all shards' newline counts. */
int xsg_result_rebase(xsg_result* r, uint64_t offset_delta, uint64_t line_index_delta);
/* The same fix-up without a host round trip.  xsg_scan_summary_device() points at 4 u64 in device
 * memory, {newlines, matches, matching lines, lines} of the LAST scan of this context (written in stream
 * order at the end of xsg_scan; valid until the next scan).  After an all-gather of every rank's summary
 * into table[world][stride] (NCCL, on the context's stream), xsg_result_rebase_from_table adds
 * sum(table[r][0] for r < rank) to every line index, reading the table on the device. */
const uint64_t* xsg_scan_summary_device(xsg_ctx* ctx);
int xsg_result_rebase_from_table(xsg_result* r, const uint64_t* table_dev, uint32_t stride, uint32_t rank);
void xsg_result_free(xsg_result* r);

/* ---- shard exchange over NVLink peer memory (no counterpart in the single-host reference) ------
 * A rank is one context: a process with one GPU (handles travel between the processes by any transport -- a process
 * group, MPI, a file -- and xsg_exchange_connect maps the peers' tables with cudaIpcOpenMemHandle), or one of several
 * contexts of one process (xsg_init_multi + xsg_exchange_connect_local: peer access, no handles).  A context has at
 * most one exchange.  Per round, every rank
 *   1. publishes the summary {newlines, matches, matching lines, lines} of its shard into the table of EVERY rank --
 *      plain stores over NVLink, then a system-scope release of the round number -- either from inside the scan
 *      (xsg_scan with XSG_SCAN_PUBLISH: the scan's last kernel does it, no extra launch) or with xsg_exchange_push()
 *      for a finished result (any number of parts, e.g. a streamed one: its totals are known on the host);
 *   2. waits for the round: xsg_result_rebase_exchanged() launches the fix-up kernel, which spins on the device until
 *      all ranks have published this round and adds the newline counts of the lower ranks to the result's line
 *      indices; xsg_exchange_totals() waits and copies the gathered table [world][4] to the host.
 * No host round trip and no collective-library call is involved.  Publishing a round before the previous one was
 * waited for is XSG_ERR_STATE (the two slot parities would be overrun).  A wait that times out (60 s: a peer died)
 * adds nothing and raises a flag that xsg_exchange_totals, xsg_exchange_check and every xsg_result_copy* of a
 * rebased result report as XSG_ERR_STATE (and clear).  The same fix-up through an NCCL all-gather:
 * xsg_scan_summary_device + xsg_result_rebase_from_table. */
typedef struct xsg_exchange xsg_exchange;
#define XSG_EXCHANGE_HANDLE_BYTES 64
/* handle_out may be NULL when the exchange is only connected locally */
int xsg_exchange_create(xsg_ctx* ctx, uint32_t world, uint32_t rank, xsg_exchange** out, void* handle_out);
int xsg_exchange_connect(xsg_exchange* x, const void* handles /* world x XSG_EXCHANGE_HANDLE_BYTES */);
int xsg_exchange_connect_local(xsg_exchange* const* all /* [n], all[r] = rank r */, uint32_t n);
int xsg_exchange_push(xsg_exchange* x, const xsg_result* r);
int xsg_result_rebase_exchanged(xsg_result* r, xsg_exchange* x);
int xsg_exchange_totals(xsg_exchange* x, uint64_t* table_host /* world x 4 */);
int xsg_exchange_check(xsg_exchange* x); /* synchronizes the context's stream; XSG_ERR_STATE after a timed-out wait */
void xsg_exchange_destroy(xsg_exchange* x);

/* ---- xs::count without any allocation --------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The list-free count path of xsg_scan(kinds = XSG_COUNT) with its scratch memory -- one u32 per 4 KiB of text, the
 * scan's constants and scalars -- in a fixed __device__ array inside libxsg.so (1 MiB + 512 B, held by every GPU that
 * loads the library until the process exits) instead of the context's cudaMalloc'ed buffers.  The call allocates
 * nothing, needs no xsg_ctx, launches its kernels on `cuda_stream` of `device` (which it makes current) and returns
 * without synchronizing; in stream order d_out[0] = the reference's xs::count over the chunks d_bounds[0..nb)
 * (exclusive chunk ends on the device, as for xsg_corpus_set_bounds; NULL = the whole text is one chunk) and
 * d_out[1] = OR of XSG_STATIC_* flags (0 = d_out[0] is valid).  d_text: device memory, 16-byte aligned, readable up to
 * the next 16-byte boundary.  flags: 0 or XSG_SCAN_IGNORE_CASE.
 * Limits: n <= XSG_STATIC_SCRATCH_MAX_TEXT, else XSG_ERR_ARG (split the text at chunk bounds and add the counts, or
 * use xsg_scan); a self-overlapping pattern of more than 4 bytes is XSG_ERR_ARG (it needs the occurrence list).
 * Every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) race. */
#define XSG_STATIC_SCRATCH_MAX_TEXT (1ull << 30)
#define XSG_STATIC_OVERLAP 1u    /* occurrences of a self-overlapping pattern overlap in this text: use xsg_scan */
#define XSG_STATIC_BAD_BOUNDS 2u /* d_bounds not increasing / not newline-aligned / not ending at n             */
int xsg_count_static_scratch(int device, void* cuda_stream, const void* d_text, uint64_t n, const uint64_t* d_bounds,
                             uint64_t nb, const uint8_t* pattern, uint32_t m, uint32_t flags, uint64_t* d_out);

#ifdef __cplusplus
}
#endif
#endif
