// xsg_internal.h -- host-side objects behind the opaque handles of include/xsg.h.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "list_kernels.cuh"
#include "xsg_common.cuh"

namespace xsg {

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);
// validated, preprocessed pattern (xsg_scan.cu: filter bytes, borders, fold masks)
int build_pattern(const uint8_t* p, uint32_t m, bool icase, Pattern* out);

#define XSG_CUDA(expr)                                                    \
  do {                                                                    \
    cudaError_t e__ = (expr);                                             \
    if (e__ != cudaSuccess) return xsg::cuda_fail(e__, #expr, __FILE__, __LINE__); \
  } while (0)

// grow-only device scratch buffer
struct Scratch {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);  // returns XSG_OK / XSG_ERR_NOMEM; contents are NOT preserved
  void release();
  template <class T> T* as() const { return static_cast<T*>(p); }
};

}  // namespace xsg

struct xsg_exchange;

struct xsg_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;       // compute (own_stream unless xsg_set_stream replaced it)
  cudaStream_t own_stream = nullptr;
  cudaStream_t copy_stream = nullptr;  // H2D of streamed segments
  cudaStream_t aux_stream = nullptr;   // side work of a scan that does not depend on the occurrence list (tail replay)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaEvent_t ev[6] = {};
  cudaMemPool_t pool = nullptr;        // stream-ordered pool of this context: result vectors
  xsg::Scratch gran_count, gran_prefix, nl_granule, nl_prefix, scan_tmp, occ, sel, desc2, selq, gran_mask, gran_first;
  xsg::Scratch stage[2];               // streamed segments (double buffer)
  xsg::Scratch seg_bounds;
  xsg::Scratch stage_bounds[2];        // chunk bounds of the streamed segments
  cudaEvent_t stage_copied[2] = {}, stage_free[2] = {};
  xsg::DevScalars* d_scalars = nullptr;
  unsigned long long* d_summary = nullptr;  // {newlines, matches, matching lines, lines} of the last scan
  xsg::DevScalars* h_scalars = nullptr;  // pinned and mapped: the scan's kernels store into it (d_h_scalars)
  xsg::DevScalars* d_h_scalars = nullptr;  // h_scalars as the device addresses it
  uint32_t* d_konst = nullptr;           // 12 words: the scan kernel's constants in global memory (see ScanArgs::konst)
  uint32_t* h_konst = nullptr;           // pinned staging of the same (= what d_konst holds once konst_valid)
  bool konst_valid = false;
  double direct_density = 0;             // occurrences per byte the last direct-mode search saw (sizes the next result vector)
  uint32_t epoch = 0;                    // tag of the look-back descriptors of the next launch (xsg::lookback)
  xsg_exchange* exchange = nullptr;      // the shard exchange attached to this context (xsg_exchange_create), if any
  unsigned int* h_ex_bad = nullptr;      // mapped pinned flag: a wait of the exchange timed out (set by kernels)
  unsigned int* d_ex_bad = nullptr;      // the same word as the device sees it
};

struct xsg_exchange {
  xsg_ctx* ctx = nullptr;
  uint32_t world = 0, rank = 0;
  unsigned long long round = 0;         // last round published by this rank
  unsigned long long waited = 0;        // last round this rank waited for (rebase / totals)
  xsg::ExSlot* local = nullptr;         // slots[2][world], written by the peers
  xsg::ExPeers peers = {};
  std::vector<void*> opened;            // IPC mappings to close
  unsigned long long* d_out = nullptr;  // [world * 4] gathered table for xsg_exchange_totals
  unsigned long long* h_out = nullptr;  // pinned [world * 4]
  bool connected = false;
};

struct xsg_corpus {
  xsg_ctx* ctx = nullptr;
  uint8_t* dev = nullptr;
  uint64_t n = 0;
  bool owned = false;
  uint64_t* d_bounds = nullptr;
  std::vector<uint64_t> h_bounds;
};

struct xsg_part {  // results of one device-resident scan (one streamed segment)
  xsg_ctx* ctx = nullptr;  // the context (device, stream, pool) its vectors live in
  uint64_t* match_off = nullptr;
  uint64_t* line_off = nullptr;
  uint64_t* line_idx = nullptr;
  uint64_t* line_start = nullptr;
  uint64_t* line_len = nullptr;
  uint64_t n_match = 0;  // entries in match_off / value of xs::count
  uint64_t n_line = 0;   // entries in line_off / line_idx / value of xs::count_lines
  uint64_t n_lines = 0;  // entries in line_start / line_len (xs::lines; may be n_line - 1)
  uint64_t newlines = 0; // '\n' in the bytes of this part (counted when line indices are asked for)
};

namespace xsg {
int exchange_begin_round(xsg_ctx* ctx, ExPublish* pub);  // xsg_exchange.cu
int exchange_check_flag(xsg_ctx* ctx);
struct ScanStats {
  uint64_t newlines = 0, raw = 0;
  float main_ms = 0, post_ms = 0, scan_ms = 0;
  uint32_t launches = 0;
};
}  // namespace xsg

struct xsg_result {
  xsg_ctx* ctx = nullptr;
  uint32_t kinds = 0;
  std::vector<xsg_part> parts;
  uint64_t count = 0, count_lines = 0, n_lines = 0;
  uint64_t newlines = 0, raw_occurrences = 0;
  xsg_timings timings = {};
  bool exchanged = false;  // line indices were fixed up by xsg_result_rebase_exchanged: copies report a timed-out wait
  // lazily concatenated device copies for xsg_result_device_ptr on multi-part results
  uint64_t* concat[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

struct xsg_stream {
  xsg_ctx* ctx = nullptr;
  xsg::Pattern pat;
  uint32_t kinds = 0;
  uint64_t goff_base = 0, lidx = 0, occ_cap = 0;
  bool local_lidx = false;    // XSG_SCAN_LOCAL_LINE_INDEX: every part numbers its lines from its own first byte
  xsg_result* res = nullptr;
  xsg::ScanStats st;
  uint64_t pushed = 0;        // segments pushed so far (ticket of the next one)
  uint64_t bytes_pushed = 0;  // global offset of the next segment
  bool pending = false;       // segment pushed - 1 is copied (or being copied) but not scanned yet
  uint64_t pend_n = 0, pend_off = 0, pend_nb = 0;
  std::vector<uint64_t> h_bounds[2];
  bool last_terminated = true;  // the previous segment ended with a newline
  double t0 = 0;
};
