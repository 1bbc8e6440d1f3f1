// xsg_scan.cu -- the scan pipeline behind xsg_scan / xsg_search_host and the result accessors.
//
// Stages (the reference's chunk -> search -> result loop, Searcher.h:100-120, as stream-ordered
// kernel launches):
//   A  count / emit     : every byte once -> granule counters (+ '\n' per 4 KiB granule); positions of
//                         the occurrences from the granules that hold any
//   B  select / zone    : reference semantics on the occurrence list (greedy, per-line, chunk tail)
//   C  compact          : stable compaction + formatting of the requested result vectors
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "xsg_internal.h"

using namespace xsg;

namespace {

constexpr uint32_t kMatchKinds = XSG_COUNT | XSG_MATCH_BYTE_OFFSETS;
constexpr uint32_t kLineKinds = XSG_COUNT_LINES | XSG_LINE_BYTE_OFFSETS | XSG_LINE_INDICES | XSG_LINES;
constexpr uint32_t kAllKinds = kMatchKinds | kLineKinds;

// rough byte frequencies of English text; only used to pick the two filter bytes of a long pattern
uint32_t byte_weight(uint8_t c) {
  static const uint8_t lower[26] = {65, 12, 23, 35, 100, 18, 16, 50, 57, 1, 6, 33, 20,
                                    57, 62, 15, 1, 50, 52, 75, 23, 8, 19, 1, 16, 1};
  if (c == ' ') return 180;
  if (c >= 'a' && c <= 'z') return lower[c - 'a'];
  if (c >= 'A' && c <= 'Z') return lower[c - 'A'] / 15 + 1;
  if (c == '\n') return 30;
  if (c == '.' || c == ',') return 10;
  if (c >= '0' && c <= '9') return 3;
  return 1;
}

int make_pattern(const uint8_t* p_in, uint32_t m, bool icase, Pattern* out) {
  uint8_t lowered[XSG_MAX_PATTERN];
  const uint8_t* p = p_in;
  if (p_in && icase && m >= 1 && m <= XSG_MAX_PATTERN) {  // the reference lowers the pattern as well
    for (uint32_t i = 0; i < m; ++i) lowered[i] = fold_byte(p_in[i]);
    p = lowered;
  }
  if (!p || m == 0 || m > XSG_MAX_PATTERN) {
    set_error("pattern length must be 1..64");
    return XSG_ERR_ARG;
  }
  for (uint32_t i = 0; i < m; ++i)
    if (p[i] == '\n') {
      set_error("the pattern must not contain '\\n'");
      return XSG_ERR_ARG;
    }
  Pattern pat;
  std::memset(&pat, 0, sizeof pat);
  std::memcpy(pat.bytes, p, m);
  pat.m = m;
  pat.icase = icase ? 1 : 0;
  pat.c_nl = 0x0a0a0a0au;
  pat.c_7f = 0x7f7f7f7fu;
  pat.one = 1u;
  pat.c_fe = 0xfefefeffu;
  auto rep = [](uint8_t c) { return 0x01010101u * c; };
  if (m <= 4) {
    pat.nsimd = m;
    for (uint32_t i = 0; i < m; ++i) pat.rep[i] = rep(p[i]);
  } else {
    pat.nsimd = 2;
    uint32_t best = 0, best_w = ~0u;
    for (uint32_t a = 0; a + 4 < m; ++a) {
      const uint32_t w = byte_weight(p[a]) * byte_weight(p[a + 4]);
      if (w < best_w) { best_w = w; best = a; }
    }
    pat.anchor = best;
    // expected filter candidates per 512-byte row ~ 512 * best_w / 1e6 (the weights are per-mille frequencies):
    // from about 0.1 on, rejecting them in registers outweighs the ~3 % the extra live registers cost
    pat.precheck = best_w >= 200 ? 1u : 0u;
    pat.rep[0] = rep(p[best]);
    pat.rep[1] = rep(p[best + 4]);
    pat.word_a = (uint32_t)p[best] | ((uint32_t)p[best + 1] << 8) | ((uint32_t)p[best + 2] << 16) | ((uint32_t)p[best + 3] << 24);
  }
  if (icase) {  // the fold is done by OR-ing 0x20 into the text where the (lowered) pattern byte is a letter
    auto letter = [](uint8_t c) { return c >= 'a' && c <= 'z'; };
    if (m <= 4) {
      for (uint32_t i = 0; i < m; ++i) pat.orm[i] = letter(p[i]) ? 0x20202020u : 0u;
    } else {
      pat.orm[0] = letter(p[pat.anchor]) ? 0x20202020u : 0u;
      pat.orm[1] = letter(p[pat.anchor + 4]) ? 0x20202020u : 0u;
      for (uint32_t i = 0; i < 4; ++i) pat.word_or |= letter(p[pat.anchor + i]) ? 0x20u << (8 * i) : 0u;
    }
  }
  // bordered <=> the KMP failure value of the whole pattern is > 0
  uint32_t fail[XSG_MAX_PATTERN + 1];
  fail[0] = 0;
  if (m > 1) {
    fail[1] = 0;
    uint32_t k = 0;
    for (uint32_t i = 1; i < m; ++i) {
      while (k > 0 && p[i] != p[k]) k = fail[k];
      if (p[i] == p[k]) ++k;
      fail[i + 1] = k;
    }
    pat.bordered = fail[m] > 0;
    for (uint32_t b = fail[m]; b > 0; b = fail[b]) pat.periods |= 1ull << (m - b);
    if (pat.bordered) pat.ovl_shift = 8 * (m - fail[m]);  // the longest border gives the smallest period
  }
  *out = pat;
  return XSG_OK;
}

void free_part(xsg_ctx* ctx, xsg_part& p) {
  if (p.ctx) ctx = p.ctx;  // a result of a multi-device search holds parts of several contexts
  cudaSetDevice(ctx->device);
  uint64_t** arrs[] = {&p.match_off, &p.line_off, &p.line_idx, &p.line_start, &p.line_len};
  for (uint64_t** a : arrs) {
    if (*a) cudaFreeAsync(*a, ctx->stream);
    *a = nullptr;
  }
}

int alloc_out(xsg_ctx* ctx, uint64_t** p, uint64_t entries) {
  *p = nullptr;
  XSG_CUDA(cudaMallocFromPoolAsync(reinterpret_cast<void**>(p), std::max<uint64_t>(entries, 1) * sizeof(uint64_t), ctx->pool, ctx->stream));
  return XSG_OK;
}

// Waits for an event by polling for the first 20 ms (a scan of a device-resident shard takes a few ms, and a
// blocking wait adds its wake-up latency to every scan), then blocks.
cudaError_t wait_event(cudaEvent_t ev) {
  const auto t0 = std::chrono::steady_clock::now();
  for (;;) {
    const cudaError_t e = cudaEventQuery(ev);
    if (e != cudaErrorNotReady) return e;
    if (std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(20)) return cudaEventSynchronize(ev);
  }
}

// Look-back descriptor tables (scan_tmp, desc2) are never cleared between launches: every launch tags its
// descriptors with a fresh 14-bit epoch.  Cleared here: a table that has just been (re)allocated, and both
// tables when the epoch wraps.
int desc_table(xsg_ctx* ctx, Scratch& t, uint64_t entries, unsigned long long** out) {
  const void* before = t.p;
  if (int rc = t.reserve(entries * sizeof(unsigned long long))) return rc;
  if (t.p != before) XSG_CUDA(cudaMemsetAsync(t.p, 0, t.cap, ctx->stream));
  *out = t.as<unsigned long long>();
  return XSG_OK;
}

int next_epoch(xsg_ctx* ctx, uint32_t* out) {
  if (++ctx->epoch > kEpochMask) {
    ctx->epoch = 1;
    if (ctx->scan_tmp.p) XSG_CUDA(cudaMemsetAsync(ctx->scan_tmp.p, 0, ctx->scan_tmp.cap, ctx->stream));
    if (ctx->desc2.p) XSG_CUDA(cudaMemsetAsync(ctx->desc2.p, 0, ctx->desc2.cap, ctx->stream));
  }
  *out = ctx->epoch;
  return XSG_OK;
}

// The scan's last kernel (summary_kernel) has stored the scalars into the mapped host mirror; ev[3], recorded behind it,
// says when they are there.  (Was: a D2H copy + an event of its own -- one more trip through a copy engine per scan.)
int read_scalars(xsg_ctx* ctx) {
  XSG_CUDA(wait_event(ctx->ev[3]));
  return XSG_OK;
}

// One device-resident scan of buf[0..n) with reference chunks d_bounds[0..nb).
int scan_device(xsg_ctx* ctx, const uint8_t* buf, uint64_t n, const uint64_t* d_bounds, uint64_t nb,
                const Pattern& pat, uint32_t kinds, uint64_t goff, uint64_t lidx_base, uint64_t occ_cap_opt,
                xsg_part* part, ScanStats* st, const ExPublish* pub = nullptr) {
  cudaStream_t s = ctx->stream;
  *part = xsg_part();
  part->ctx = ctx;
  XSG_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, sizeof(DevScalars), s));
  if (n == 0 || nb == 0) {
    launch_summary(ctx->d_scalars, 0, 0, ctx->d_summary, pub, s);  // all zero (and published, if an exchange wants it)
    XSG_CUDA(cudaGetLastError());
    XSG_CUDA(cudaEventRecord(ctx->ev[3], s));
    XSG_CUDA(cudaStreamSynchronize(s));
    return XSG_OK;
  }
  const bool want_match = (kinds & kMatchKinds) != 0, want_line = (kinds & kLineKinds) != 0;
  const bool want_idx = (kinds & XSG_LINE_INDICES) != 0;
  // xs::count alone needs no occurrence list when no two occurrences overlap: always true for a border-free
  // pattern; for a bordered one pass 1 checks it on the fly (OVL variant) and the list path below is the fall-back
  // (bordered patterns of five bytes and more go straight to the list path: their occurrences are sparse, the
  // list costs next to nothing, and the filter path has no cheap whole-row overlap test)
  bool list_path = kinds != XSG_COUNT || (pat.bordered && pat.m > 4);
  const uint64_t n_tiles = div_up(n, kTileBytes);
  const uint64_t n_gran = n_tiles * kWarps;
  int rc;
  if ((rc = ctx->gran_count.reserve(n_gran * sizeof(uint32_t)))) return rc;
  ScanArgs a{};
  a.buf = buf;
  a.n = n;
  a.n_tiles = n_tiles;
  a.gran_count = ctx->gran_count.as<uint32_t>();
  DevScalars* ds = ctx->d_scalars;
  {  // the scan kernel's constants, also in global memory (ScanArgs::konst)
    uint32_t k[12];
    for (int i = 0; i < 4; ++i) k[i] = pat.rep[i], k[8 + i] = pat.orm[i];
    k[4] = pat.c_nl, k[5] = pat.c_7f, k[6] = pat.one, k[7] = pat.c_fe;
    if (!ctx->konst_valid || std::memcmp(k, ctx->h_konst, sizeof k) != 0) {  // same pattern as last time: already there
      std::memcpy(ctx->h_konst, k, sizeof k);
      XSG_CUDA(cudaMemcpyAsync(ctx->d_konst, ctx->h_konst, sizeof k, cudaMemcpyHostToDevice, s));
      ctx->konst_valid = true;
    }
    a.konst = reinterpret_cast<const uint4*>(ctx->d_konst);
  }

  if (!list_path) {
    a.overlap = pat.bordered ? &ds->overlap : nullptr;
    XSG_CUDA(cudaEventRecord(ctx->ev[1], s));
    XSG_CUDA(launch_count(pat, a, false, false, s));
    XSG_CUDA(cudaEventRecord(ctx->ev[2], s));
    launch_reduce_u32(a.gran_count, n_gran, &ds->occ_total, s);
    launch_zone_count(buf, n, d_bounds, nb, pat, a.gran_count, &ds->zone_sub, &ds->zone_add, s);
    XSG_CUDA(cudaGetLastError());
    launch_summary(ds, 1, 0, ctx->d_summary, pub, s, ctx->d_h_scalars);
    st->launches += 3 + (pat.m > 1 ? 1 : 0);
    XSG_CUDA(cudaEventRecord(ctx->ev[3], s));
    if ((rc = read_scalars(ctx))) return rc;
    const DevScalars& h = *ctx->h_scalars;
    if (h.overlap) {  // overlapping occurrences exist: the greedy selection needs the list
      list_path = true;
      a.overlap = nullptr;
      float aborted = 0;  // the aborted pass belongs to the scan's time (ev[1] is recorded again below)
      XSG_CUDA(cudaEventElapsedTime(&aborted, ctx->ev[1], ctx->ev[3]));
      st->scan_ms += aborted;
      XSG_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, sizeof(DevScalars), s));
    } else {
      st->raw += h.occ_total;
      part->n_match = h.occ_total - h.zone_sub + h.zone_add;
    }
  }
  if (list_path) {
    if ((rc = ctx->gran_prefix.reserve((n_gran + 2) * sizeof(uint64_t)))) return rc;
    unsigned long long* scan_desc = nullptr;
    if ((rc = desc_table(ctx, ctx->scan_tmp, 2 * (div_up(n_gran, 2048) + 1), &scan_desc))) return rc;
    if (want_idx) {
      if ((rc = ctx->nl_granule.reserve(n_gran * sizeof(uint32_t)))) return rc;
      if ((rc = ctx->nl_prefix.reserve((n_gran + 2) * sizeof(uint64_t)))) return rc;
      a.nl_granule = ctx->nl_granule.as<uint32_t>();
    }
    if ((rc = ctx->gran_mask.reserve(n_gran * 32 * sizeof(uint4)))) return rc;
    a.gran_mask = ctx->gran_mask.as<uint4>();
    if ((rc = ctx->gran_first.reserve(n_gran * sizeof(uint16_t)))) return rc;
    a.gran_first = ctx->gran_first.as<uint16_t>();
    a.gran_prefix = ctx->gran_prefix.as<uint64_t>();
    // pass 1: every byte once -> granule counters; then their exclusive prefix
    XSG_CUDA(cudaEventRecord(ctx->ev[1], s));
    XSG_CUDA(launch_count(pat, a, want_idx, true, s));
    XSG_CUDA(cudaEventRecord(ctx->ev[2], s));
    // one set of launches scans both tables: granule occurrence counts and granule newline counts
    uint32_t epoch;
    if ((rc = next_epoch(ctx, &epoch))) return rc;
    // direct mode: for xs::match_byte_offsets the occurrence list IS the result unless a chunk tail of the
    // reference drops one (rare) or -- bordered patterns only -- two occurrences overlap somewhere (checked on
    // the list).  So pass 2 writes global offsets straight into the result vector and the list kernels read
    // it with a bias; the select + compact fall-back below covers the exceptions.
    const bool direct = (kinds & XSG_MATCH_BYTE_OFFSETS) != 0;
    // Direct mode, border-free pattern: what the reference's chunk tails would not report can be counted WITHOUT the
    // list (zone_count_kernel: the text and the granule counters are all it reads), so that serial, latency-bound
    // replay -- one warp per reference chunk -- runs on a side stream next to the prefix sums and pass 2 instead of
    // behind them.
    const bool zone_aside = direct && !pat.bordered && pat.m > 1 && nb > 0;
    if (zone_aside) {
      XSG_CUDA(cudaEventRecord(ctx->ev_fork, s));
      XSG_CUDA(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0));
      launch_zone_count(buf, n, d_bounds, nb, pat, a.gran_count, &ds->zone_sub, &ds->zone_add, ctx->aux_stream);
      XSG_CUDA(cudaGetLastError());
      XSG_CUDA(cudaEventRecord(ctx->ev_join, ctx->aux_stream));
      st->launches += 1;
    }
    launch_exscan2_u32(a.gran_count, ctx->gran_prefix.as<uint64_t>(), &ds->occ_total, a.nl_granule,
                       want_idx ? ctx->nl_prefix.as<uint64_t>() : nullptr, &ds->nl_total, n_gran, scan_desc,
                       &ds->ticket[0], epoch, s, &ctx->d_h_scalars->occ_total);
    XSG_CUDA(cudaGetLastError());
    st->launches += 2;
    // pass 2: positions of the occurrences, expanded from the masks of the granules that hold any.
    // The host needs their number (allocation sizes, grids): the prefix-sum kernel has stored it into the mapped host
    // mirror, the event says when (no copy-engine operation between the prefix sums and pass 2).
    XSG_CUDA(cudaEventRecord(ctx->ev[4], s));
    uint64_t k = 0;
    uint64_t* occ_list = nullptr;
    if (direct) {
      // The result vector is sized by the exact count, which costs a host round trip with the GPU idle between pass 1
      // and pass 2.  A context that has searched before (the next segment of a streamed file, the same corpus again)
      // sizes it from the density the last search saw, 3 % on top, and starts pass 2 at once; if the guess was too
      // small -- a list that does not fit is simply not written past its capacity -- pass 2 runs again with the
      // exact size, as it would have anyway.
      a.occ_add = goff;
      const uint64_t guess = ctx->direct_density > 0 ? (uint64_t)(ctx->direct_density * 1.03 * (double)n) + 4096 : 0;
      bool emitted = false;
      if (guess) {
        if ((rc = alloc_out(ctx, &part->match_off, guess))) return rc;
        a.occ = part->match_off;
        a.occ_cap = guess;
        XSG_CUDA(launch_emit(pat, a, guess > 4 * n_gran, s));
        st->launches += 1;
        emitted = true;
      }
      XSG_CUDA(wait_event(ctx->ev[4]));
      k = ctx->h_scalars->occ_total;
      ctx->direct_density = (double)k / (double)n;
      if (emitted && k > guess) {
        XSG_CUDA(cudaFreeAsync(part->match_off, s));
        part->match_off = nullptr;
        emitted = false;
      }
      if (!emitted) {
        if ((rc = alloc_out(ctx, &part->match_off, k))) return rc;
        a.occ = part->match_off;
        a.occ_cap = k;
        if (k) {
          XSG_CUDA(launch_emit(pat, a, k > 4 * n_gran, s));
          st->launches += 1;
        }
      }
      occ_list = part->match_off;
      if (k) {
        if (pat.bordered) {
          launch_overlap_check(occ_list, k, pat.m, &ds->overlap, s);
          st->launches += 1;
        }
      }
    } else {
      uint64_t cap = occ_cap_opt ? occ_cap_opt : std::max<uint64_t>(1u << 20, n / 40);
      if (ctx->occ.cap / sizeof(uint64_t) > cap && !occ_cap_opt) cap = ctx->occ.cap / sizeof(uint64_t);
      if ((rc = ctx->occ.reserve(cap * sizeof(uint64_t)))) return rc;
      a.occ = ctx->occ.as<uint64_t>();
      a.occ_cap = cap;
      XSG_CUDA(launch_emit(pat, a, false, s));  // the count travels to the host while this runs
      st->launches += 1;
      XSG_CUDA(wait_event(ctx->ev[4]));
      k = ctx->h_scalars->occ_total;
      if (k > cap) {  // rare: denser than the estimate -> emit again with the exact capacity
        XSG_CUDA(cudaStreamSynchronize(s));
        if ((rc = ctx->occ.reserve(k * sizeof(uint64_t)))) return rc;
        a.occ = ctx->occ.as<uint64_t>();
        a.occ_cap = k;
        XSG_CUDA(launch_emit(pat, a, false, s));
        st->launches += 1;
      }
      occ_list = ctx->occ.as<uint64_t>();
    }
    st->raw += k;
    if (k > 0) {
      if ((rc = ctx->sel.reserve(k))) return rc;
      const uint64_t n_ct = div_up(k, 2048);
      unsigned long long* compact_desc = nullptr;
      if ((rc = desc_table(ctx, ctx->desc2, n_ct, &compact_desc))) return rc;
      ListArgs l{buf, n, d_bounds, nb, occ_list, k, ctx->sel.as<uint8_t>(), ctx->gran_prefix.as<uint64_t>(), n_gran,
                 direct ? goff : 0, &ds->drops};
      for (int fam = 0; fam < 2; ++fam) {
        if (fam == 0 ? !want_match : !want_line) continue;
        const int mode = fam == 1 ? kSelLine : (pat.bordered ? kSelMatchBordered : kSelMatch);
        if (fam == 0 && direct) {  // only count what the reference's chunk tails would not report
          if (!zone_aside) {
            ListArgs ld = l;
            ld.sel = nullptr;
            launch_zone(ld, pat, kSelMatch, s);  // bordered: meaningful only if nothing overlaps (else fall-back)
            st->launches += pat.m > 1 ? 1 : 0;
          }
          continue;
        }
        st->launches += launch_select_zone(l, pat, mode, s);
        CompactArgs c{};
        c.l = l;
        c.goff = goff;
        c.lidx_base = lidx_base;
        c.nl_prefix = ctx->nl_prefix.as<uint64_t>();
        c.nl_halves = ctx->nl_granule.as<uint32_t>();
        if (fam == 0) {
          if (kinds & XSG_MATCH_BYTE_OFFSETS) {
            if ((rc = alloc_out(ctx, &part->match_off, k))) return rc;
            c.out_match = part->match_off;
          }
        } else {
          if (kinds & XSG_LINE_BYTE_OFFSETS) {
            if ((rc = alloc_out(ctx, &part->line_off, k))) return rc;
            c.out_line_off = part->line_off;
          }
          if (kinds & XSG_LINE_INDICES) {
            if ((rc = alloc_out(ctx, &part->line_idx, k))) return rc;
            c.out_line_idx = part->line_idx;
          }
          if (kinds & XSG_LINES) {
            if ((rc = alloc_out(ctx, &part->line_start, k))) return rc;
            if ((rc = alloc_out(ctx, &part->line_len, k))) return rc;
            c.out_line_start = part->line_start;
            c.out_line_len = part->line_len;
          }
          if (c.out_line_off || c.out_line_idx || c.out_line_start) {
            if ((rc = ctx->selq.reserve(k * sizeof(uint64_t)))) return rc;
            c.selq = ctx->selq.as<uint64_t>();
          }
        }
        c.desc = compact_desc;
        if ((rc = next_epoch(ctx, &c.epoch))) return rc;
        c.ticket = &ds->ticket[2 + fam];
        c.total = &ds->sel_total[fam];
        c.unterminated = &ds->unterminated;
        launch_compact(c, pat, s);
        XSG_CUDA(cudaGetLastError());
        st->launches += c.selq ? 2 : 1;
      }
    }
    if (zone_aside) XSG_CUDA(cudaStreamWaitEvent(s, ctx->ev_join, 0));
    launch_summary(ds, zone_aside ? 3 : (direct ? 2 : 0), (kinds & XSG_LINES) ? 1 : 0, ctx->d_summary, pub, s, ctx->d_h_scalars);
    st->launches += 1;
    XSG_CUDA(cudaEventRecord(ctx->ev[3], s));
    if ((rc = read_scalars(ctx))) return rc;
    if (direct && k > 0 &&
        (ctx->h_scalars->drops || ctx->h_scalars->overlap || (zone_aside && ctx->h_scalars->zone_sub != ctx->h_scalars->zone_add))) {
      // a chunk tail of the reference skipped occurrences, or occurrences overlap (greedy selection needed):
      // select + compact the list into a new vector
      const int fb_mode = pat.bordered ? kSelMatchBordered : kSelMatch;
      ListArgs l{buf, n, d_bounds, nb, occ_list, k, ctx->sel.as<uint8_t>(), ctx->gran_prefix.as<uint64_t>(), n_gran,
                 goff, &ds->drops};
      launch_select(l, pat, fb_mode, s);
      launch_zone(l, pat, fb_mode, s);
      CompactArgs c{};
      c.l = l;
      c.goff = goff;
      uint64_t* fixed = nullptr;
      if ((rc = alloc_out(ctx, &fixed, k))) return rc;
      c.out_match = fixed;
      c.desc = ctx->desc2.as<unsigned long long>();  // sized above (k > 0)
      if ((rc = next_epoch(ctx, &c.epoch))) return rc;
      c.ticket = &ds->ticket[2];
      c.total = &ds->sel_total[0];
      c.unterminated = &ds->unterminated;
      launch_compact(c, pat, s);
      XSG_CUDA(cudaGetLastError());
      launch_summary(ds, 0, (kinds & XSG_LINES) ? 1 : 0, ctx->d_summary, pub, s, ctx->d_h_scalars);
      st->launches += 4;
      XSG_CUDA(cudaFreeAsync(part->match_off, s));
      part->match_off = fixed;
      XSG_CUDA(cudaEventRecord(ctx->ev[3], s));
      if ((rc = read_scalars(ctx))) return rc;
    } else if (direct) {
      ctx->h_scalars->sel_total[0] = k;
    }
    const DevScalars& h = *ctx->h_scalars;
    part->n_match = h.sel_total[0];
    part->n_line = h.sel_total[1];
    part->n_lines = (kinds & XSG_LINES) ? h.sel_total[1] - (h.unterminated ? 1 : 0) : 0;
    part->newlines = h.nl_total;
    st->newlines += h.nl_total;
  }
  // scan = pass 1 + everything behind it (two event queries: each costs the host about 3 us between two searches; the
  // few microseconds in front of pass 1 -- clearing 96 bytes of scalars -- are not in it)
  float main_ms = 0, post_ms = 0;
  XSG_CUDA(cudaEventElapsedTime(&main_ms, ctx->ev[1], ctx->ev[2]));
  XSG_CUDA(cudaEventElapsedTime(&post_ms, ctx->ev[2], ctx->ev[3]));
  st->main_ms += main_ms;
  st->post_ms += post_ms;
  st->scan_ms += main_ms + post_ms;
  return XSG_OK;
}

void finish_result(xsg_result* r, const ScanStats& st) {
  for (const xsg_part& p : r->parts) {
    r->count += p.n_match;
    r->count_lines += p.n_line;
    r->n_lines += p.n_lines;
  }
  r->newlines = st.newlines;
  r->raw_occurrences = st.raw;
  r->timings.scan_ms = st.scan_ms;
  r->timings.main_kernel_ms = st.main_ms;
  r->timings.post_ms = st.post_ms;
  r->timings.kernel_launches = st.launches;
}

}  // namespace

namespace xsg {
// make_pattern for the other translation units (xsg_static_scratch.cu)
int build_pattern(const uint8_t* p, uint32_t m, bool icase, Pattern* out) { return make_pattern(p, m, icase, out); }
}  // namespace xsg

#include "xsg_scan_api.inc"
