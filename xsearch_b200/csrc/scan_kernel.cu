// scan_kernel.cu -- the byte-scanning kernels (sm_100a).
//
// Pass 1 (count_kernel) reads every byte once and
//   * finds every true occurrence of the literal (exact SIMD-in-word compare for m <= 4, a two-byte
//     SIMD filter + verification for m >= 5), storing one counter per 4 KiB granule, and
//   * counts '\n' per 4 KiB granule (line-index accounting).
// Pass 2 (emit_kernel) re-reads only the granules that hold occurrences and writes their positions,
// in file order, at the exclusive prefix of the granule counters (stream compaction without any
// inter-CTA dependency; a decoupled look-back variant stalled on descriptor latency under full HBM
// load).
//
// Replaces the inner loops of the reference: simd::strstr / strchr (simd_search.cpp:116-204) driven
// by the findNext / findNextNewLine chains of search_wrappers.h:36-50,168-183,192-205.  The
// sequential "next match -> skip -> next" chain is reformulated data-parallel: these kernels report
// ALL true occurrences; greedy non-overlap, first-hit-per-line and the reference's chunk-tail quirk
// are resolved on the (much smaller) occurrence list by list_kernels.cu.
//
// Layout: a warp owns a 4 KiB granule (8 rows x 512 B), lane l holds bytes [16l,16l+16) of each row
// => every load is one fully coalesced 16-byte ld.global.nc per thread, 8 independent loads in
// flight per thread; the pattern-length halo is one extra 32-bit word per row obtained with a warp
// shuffle (m <= 4 needs <= 3 halo bytes; the m >= 5 filter tests pattern bytes a and a+4).
#include "xsg_common.cuh"

namespace xsg {

namespace {

// keep only the lowest `valid` byte lanes (valid may be <= 0 or >= 4)
__device__ __forceinline__ uint32_t keep_lanes(uint32_t w, long long valid) {
  if (valid >= 4) return w;
  if (valid <= 0) return 0u;
  return w & ((1u << (8 * (int)valid)) - 1u);
}

// full compare of the pattern at corpus position q (slow path of the m >= 5 filter)
__device__ __forceinline__ bool verify_at(const uint8_t* __restrict__ buf, uint64_t q, const Pattern& pat) {
  if (pat.icase) {
    for (uint32_t k = 0; k < pat.m; ++k)
      if (fold_byte(buf[q + k]) != pat.bytes[k]) return false;
    return true;
  }
  for (uint32_t k = 0; k < pat.m; ++k)
    if (buf[q + k] != pat.bytes[k]) return false;
  return true;
}

// acc + 128 * (number of bytes of w that differ from '\n').  Exact per-lane test in three logic/add
// operations: bit 7 of ((w ^ c) & 0x7f) + 0x7f is set iff the low seven bits differ from c; bit 7 of
// w itself differs from c iff it is set (c = 0x0a has it clear), so the xor is not needed for the
// final OR.  The four 0x80 flags are summed by one dp4a instead of popc + add.
__device__ __forceinline__ uint32_t not_newline_x128(uint32_t w, uint32_t acc, uint32_t c_nl, uint32_t c_7f, uint32_t one,
                                                     uint32_t c_01) {
  const uint32_t b = add_on_fma_pipe((w ^ c_nl) & c_7f, one, c_7f);
  return __dp4a((b | w) & k80, c_01, acc);
}

struct LaneResult {
  uint32_t mask[kRows / 2];  // EMIT: 16 occurrence bits per row, two rows per register
  uint32_t cnt;              // occurrences owned by this lane
  uint32_t nonnlh[2];        // per half of the granule (four rows): 128 x bytes that are not '\n' (padding included)
  uint32_t ovl;              // OVL: non-zero when an occurrence of this lane overlaps a later one
};

// Loads one warp granule: 8 independent coalesced 16-byte loads per thread (+ the halo word in lane 0).
__device__ __forceinline__ void load_granule(const uint8_t* __restrict__ buf, uint64_t n, uint64_t gran_off, int lane,
                                             uint4 (&v)[kRows], uint32_t& ext) {
  const uint64_t lane_off = gran_off + (uint64_t)lane * 16;
  const bool partial = gran_off + kGranuleBytes + 16 > n;
  ext = 0;  // lane 0: first word behind the granule (halo of the last row)
  if (!partial) {
#pragma unroll
    for (int u = 0; u < kRows; ++u) v[u] = ld_stream_v4(reinterpret_cast<const uint4*>(buf + lane_off + u * kRowBytes));
    if (lane == 0) ext = ld_stream_u32(reinterpret_cast<const uint32_t*>(buf + gran_off + kGranuleBytes));
  } else {
#pragma unroll
    for (int u = 0; u < kRows; ++u) {
      const uint64_t pos = lane_off + u * kRowBytes;
      uint4 t = make_uint4(0, 0, 0, 0);
      if (pos < n) {
        t = ld_stream_v4(reinterpret_cast<const uint4*>(buf + pos));
        const long long valid = (long long)(n - pos);
        t.x = keep_lanes(t.x, valid);
        t.y = keep_lanes(t.y, valid - 4);
        t.z = keep_lanes(t.z, valid - 8);
        t.w = keep_lanes(t.w, valid - 12);
      }
      v[u] = t;
    }
    if (lane == 0) {
      const uint64_t pos = gran_off + kGranuleBytes;
      if (pos < n) ext = keep_lanes(ld_stream_u32(reinterpret_cast<const uint32_t*>(buf + pos)), (long long)(n - pos));
    }
  }
}

// Slow path of the filter variants for one 16-byte lane row: the candidates (zero byte lanes of the filter words,
// recomputed here from the text words) are verified against the corpus; returns the 16 occurrence bits.
template <int NS, bool ICASE>
__device__ __forceinline__ uint32_t verify_row(const Pattern& pat, const uint8_t* __restrict__ buf, long long last_q,
                                               uint64_t pos0, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3,
                                               uint32_t w4, uint32_t rep0, uint32_t rep1, uint32_t o0, uint32_t o1) {
  const uint32_t zj[4] = {((w0 | o0) ^ rep0) | ((w1 | o1) ^ rep1), ((w1 | o0) ^ rep0) | ((w2 | o1) ^ rep1),
                          ((w2 | o0) ^ rep0) | ((w3 | o1) ^ rep1), ((w3 | o0) ^ rep0) | ((w4 | o1) ^ rep1)};
  const uint32_t wj[5] = {w0, w1, w2, w3, w4};
  uint32_t m16 = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t word = zero_lanes(zj[j]);
    while (word) {
      const int bit = __ffs(word) - 1;  // 7, 15, 23 or 31
      word &= word - 1;
      // NS == 6: the two filter bytes matched; the four bytes from the first one on are still in
      // registers, so almost every false candidate dies here without touching memory.  (Keeping the
      // text words alive costs the streaming path registers, hence a variant of its own, chosen by
      // the host when the filter bytes are frequent ones: Pattern::precheck.)
      if (NS == 6 && (__funnelshift_r(wj[j], wj[j + 1], bit - 7) | (ICASE ? pat.word_or : 0u)) != pat.word_a) continue;
      const int byte = 4 * j + (bit >> 3);
      const long long q = (long long)(pos0 + byte) - (long long)pat.anchor;
      if (q >= 0 && q <= last_q && verify_at(buf, (uint64_t)q, pat)) m16 |= 1u << byte;
    }
  }
  return m16;
}

// Phase 1 for one warp granule: classify, (verify), count.  NS: 1..4 = exact compare of m = NS
// consecutive bytes; 5 = filter on pattern bytes (a, a+4) + verification; 6 = 5 + register pre-check.
// OVL (bordered patterns, xs::count without a list): additionally report whether ANY two occurrences overlap.
// Two occurrences q < q' overlap only at a distance q' - q that is a period of the pattern (pat.periods).
#ifndef XSG_PAIR_ROWS
#define XSG_PAIR_ROWS 1
#endif
constexpr bool kPairRows = XSG_PAIR_ROWS != 0;

template <int NS, bool EMIT, bool NL, bool ICASE, bool OVL, bool STAGED = false>
__device__ __forceinline__ void classify_granule(const Pattern& pat, const uint8_t* __restrict__ buf, uint64_t n,
                                                 uint64_t gran_off, int lane, const uint4 (&v)[kRows], uint32_t ext,
                                                 const uint4* __restrict__ konst, LaneResult& r, uint32_t srow = 0) {
  const uint64_t lane_off = gran_off + (uint64_t)lane * 16;
  const bool partial = gran_off + kGranuleBytes + 16 > n;

  r.cnt = 0;
  r.ovl = 0;
#pragma unroll
  for (int i = 0; i < kRows / 2; ++i) r.mask[i] = 0;
  r.nonnlh[0] = r.nonnlh[1] = 0;
  const long long last_q = (long long)n - (long long)pat.m;  // last valid occurrence start
  // The filter variants take their constants from global memory so that they stay in
  // registers (ScanArgs::konst: saves nine constant-bank reads per row); the exact-compare variants are bound by
  // the ALU pipe either way and run slightly better with constant-bank operands.
  constexpr bool kFromGlobal = NS >= 5;
  const uint4 ka = kFromGlobal ? __ldg(konst) : make_uint4(pat.rep[0], pat.rep[1], pat.rep[2], pat.rep[3]);
  const uint4 kb = kFromGlobal ? __ldg(konst + 1) : make_uint4(pat.c_nl, pat.c_7f, pat.one, pat.c_fe);
  const uint4 kc = !ICASE ? make_uint4(0, 0, 0, 0)
                          : (kFromGlobal ? __ldg(konst + 2) : make_uint4(pat.orm[0], pat.orm[1], pat.orm[2], pat.orm[3]));
  const uint32_t k_nl = kb.x, k_7f = kb.y, k_one = kb.z, k_fe = kb.w;
  const uint32_t k_01 = kFromGlobal ? k_one * k01 : k01;
  const uint32_t k_rep[4] = {ka.x, ka.y, ka.z, ka.w}, k_orm[4] = {kc.x, kc.y, kc.z, kc.w};

  uint32_t pend = 0;  // kPairRows: candidate flags of the even row of the current pair
#pragma unroll
  for (int u = 0; u < kRows; ++u) {
    // ignore case costs no folding pass: for a (lowered) pattern byte c that is a letter, tolower(b) == c
    // <=> (b | 0x20) == c, and for any other c <=> b == c.  The per-byte OR mask (pat.orm[k]: 0x20 in every
    // lane if pattern byte k is a letter, else 0) rides along in the LOP3 that does the compare.
    const uint32_t w0 = v[u].x, w1 = v[u].y, w2 = v[u].z, w3 = v[u].w;
    // halo word: lane l takes word 0 of lane l+1; lane 31 takes word 0 of the next row (held by lane 0).
    // STAGED (the tile sits in shared memory, count_kernel_staged): it is simply the next word there.
    uint32_t w4;
    if (STAGED) {
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w4) : "r"(srow + u * kRowBytes + 16));
    } else {
      uint32_t send = w0;
      if (lane == 0) send = (u < kRows - 1) ? v[(u < kRows - 1) ? u + 1 : u].x : ext;
      w4 = __shfl_sync(kFull, send, (lane + 1) & 31);
    }
    const uint64_t pos0 = lane_off + (uint64_t)u * kRowBytes;

    if (NL) {  // 128 x (bytes that are not '\n'), accumulated on the multiply-add pipe
      uint32_t& acc = r.nonnlh[u >> 2];
      acc = not_newline_x128(w0, acc, k_nl, k_7f, k_one, k_01);
      acc = not_newline_x128(w1, acc, k_nl, k_7f, k_one, k_01);
      acc = not_newline_x128(w2, acc, k_nl, k_7f, k_one, k_01);
      acc = not_newline_x128(w3, acc, k_nl, k_7f, k_one, k_01);
    }

    uint32_t f0 = 0, f1 = 0, f2 = 0, f3 = 0;
    if (NS <= 4) {
      const uint32_t o0 = k_orm[0];
      uint32_t z0 = (w0 | o0) ^ k_rep[0], z1 = (w1 | o0) ^ k_rep[0], z2 = (w2 | o0) ^ k_rep[0],
               z3 = (w3 | o0) ^ k_rep[0];
#pragma unroll
      for (int k = 1; k < NS; ++k) {  // byte k of the window: data shifted down by k bytes
        const uint32_t ok = k_orm[k];
        const uint32_t s0 = __funnelshift_r(w0, w1, 8 * k), s1 = __funnelshift_r(w1, w2, 8 * k);
        const uint32_t s2 = __funnelshift_r(w2, w3, 8 * k), s3 = __funnelshift_r(w3, w4, 8 * k);
        z0 |= (s0 | ok) ^ k_rep[k];
        z1 |= (s1 | ok) ^ k_rep[k];
        z2 |= (s2 | ok) ^ k_rep[k];
        z3 |= (s3 | ok) ^ k_rep[k];
      }
      // exact zero test per byte lane, its addition issued on the multiply-add pipe (see add_on_fma_pipe)
      f0 = ~(add_on_fma_pipe(z0 & k_7f, k_one, k_7f) | z0) & k80;
      f1 = ~(add_on_fma_pipe(z1 & k_7f, k_one, k_7f) | z1) & k80;
      f2 = ~(add_on_fma_pipe(z2 & k_7f, k_one, k_7f) | z2) & k80;
      f3 = ~(add_on_fma_pipe(z3 & k_7f, k_one, k_7f) | z3) & k80;
      if (partial) {  // zero padding / bytes behind n must not match
        const long long lim = last_q + 1 - (long long)pos0;
        f0 = keep_lanes(f0, lim);
        f1 = keep_lanes(f1, lim - 4);
        f2 = keep_lanes(f2, lim - 8);
        f3 = keep_lanes(f3, lim - 12);
      }
      // The logic ops above saturate the ALU pipe, so everything that follows is written for the multiply-add pipe.
      // A flag is 0x80 in its byte lane: dp4a with the weights (1, 2, 4, 8) / (16, 32, 64, 128) sums the flags of two
      // words into 0x80 x (their eight flags as a byte) -- four IDP per row instead of shifts, multiplies and byte
      // permutes on the ALU pipe -- and dp4a with the weights (1, 1, 1, 1) counts them (0x80 x count), no popc.
      uint32_t row16 = 0;  // the 16 occurrence flags of this lane's 16 bytes, bit i <=> byte i
      auto pack_row = [&]() {
        const uint32_t lo = __dp4a(f1, 0x80402010u, __dp4a(f0, 0x08040201u, 0u));  // bits 7..14: words 0, 1
        const uint32_t hi = __dp4a(f3, 0x80402010u, __dp4a(f2, 0x08040201u, 0u));  // bits 7..14: words 2, 3
        row16 = add_on_fma_pipe(hi, 2u * k_one, lo >> 7);
      };
      if (EMIT) pack_row();
      const uint32_t cnt_before = r.cnt;
      if (!EMIT) {  // counted from the masks otherwise
        r.cnt = __dp4a(f0, k_01, r.cnt);
        r.cnt = __dp4a(f1, k_01, r.cnt);
        r.cnt = __dp4a(f2, k_01, r.cnt);
        r.cnt = __dp4a(f3, k_01, r.cnt);
      }
      // OVL: an overlapping pair needs two occurrences within m - 1 <= 3 bytes: either both in this lane's 16 bytes,
      // or the first one in its last three bytes.  Rows where no lane qualifies (three out of four for a word like
      // "that") skip the test altogether.
      if (OVL && __any_sync(kFull, r.cnt - cnt_before >= 256u || (f3 & 0x80808000u) != 0)) {
        // Two occurrences overlap iff they lie p < m bytes apart, p the smallest period (for m <= 4 every period is
        // a multiple of the smallest one, and occurrences at q and q + kp imply one at q + p).  With the flags of
        // this lane's 16 bytes in the low half of a word and those of the next lane's in the high half, that is one
        // shift and one AND.  The bytes behind lane 31 belong to another row or granule: checked byte-wise.
        pack_row();
        uint32_t nxt = __shfl_sync(kFull, row16, (lane + 1) & 31);
        if (lane == 31) nxt = 0;
        const uint32_t m32 = add_on_fma_pipe(nxt, 65536u * k_one, row16);
        r.ovl |= m32 & (m32 >> (pat.ovl_shift >> 3)) & 0xffffu;
        if (lane == 31 && (row16 >> (16 - (NS - 1))) != 0) {
          for (int j = 16 - (NS - 1); j < 16; ++j) {
            if (!((row16 >> j) & 1)) continue;
            for (int d = 16 - j; d < NS; ++d) {
              const long long q2 = (long long)(pos0 + j + d);
              if (((pat.periods >> d) & 1) && q2 <= last_q && verify_at(buf, (uint64_t)q2, pat)) r.ovl |= 1;
            }
          }
        }
      }
      if (EMIT) {  // two rows per mask register: even rows in the low half
        if (u & 1) r.mask[u >> 1] = add_on_fma_pipe(row16, 65536u * k_one, r.mask[u >> 1]);
        else r.mask[u >> 1] = row16;
      }
    } else {
      // filter: pattern byte a at this position AND pattern byte a+4 one word further
      const uint32_t o0 = k_orm[0], o1 = k_orm[1];
      const uint32_t z0 = ((w0 | o0) ^ k_rep[0]) | ((w1 | o1) ^ k_rep[1]);
      const uint32_t z1 = ((w1 | o0) ^ k_rep[0]) | ((w2 | o1) ^ k_rep[1]);
      const uint32_t z2 = ((w2 | o0) ^ k_rep[0]) | ((w3 | o1) ^ k_rep[1]);
      const uint32_t z3 = ((w3 | o0) ^ k_rep[0]) | ((w4 | o1) ^ k_rep[1]);
      // (z - 0x01010101) & ~z per word, the subtraction as IMAD (see add_on_fma_pipe)
      const uint32_t c_fe = k_fe;
      const uint32_t any = ((add_on_fma_pipe(z0, k_one, c_fe) & ~z0) | (add_on_fma_pipe(z1, k_one, c_fe) & ~z1) |
                            (add_on_fma_pipe(z2, k_one, c_fe) & ~z2) | (add_on_fma_pipe(z3, k_one, c_fe) & ~z3)) & k80;
      // One vote per PAIR of rows: the even row only remembers its candidate flags.  Its text words stay in
      // their registers one row longer (that does not raise the peak: all eight rows are live at row 0).
      constexpr bool kPair = kPairRows && !ICASE;  // measured: -0.8 % without ignore case, +1..4 % with it
      if (kPair && !(u & 1)) {
        pend = any;
      } else if (__any_sync(kFull, ((kPair ? pend : 0u) | any) != 0)) {
        if (kPair) {
          const int up = u > 0 ? u - 1 : 0;
          uint32_t sendp = v[up].x;
          if (lane == 0) sendp = w0;  // lane 31 of the previous row continues in lane 0 of this one
          const uint32_t w4p = __shfl_sync(kFull, sendp, (lane + 1) & 31);
          if (pend != 0) {
            const uint32_t m16 = verify_row<NS, ICASE>(pat, buf, last_q, pos0 - kRowBytes, v[up].x, v[up].y, v[up].z,
                                                      v[up].w, w4p, k_rep[0], k_rep[1], o0, o1);
            r.cnt += __popc(m16);
            if (EMIT) r.mask[up >> 1] |= m16 << (16 * (up & 1));
          }
        }
        if (any != 0) {
          const uint32_t m16 = verify_row<NS, ICASE>(pat, buf, last_q, pos0, w0, w1, w2, w3, w4, k_rep[0], k_rep[1], o0, o1);
          r.cnt += __popc(m16);
          if (EMIT) r.mask[u >> 1] |= m16 << (16 * (u & 1));
        }
      }
    }
  }
  if (NS <= 4) {
    if (EMIT) r.cnt = __popc(r.mask[0]) + __popc(r.mask[1]) + __popc(r.mask[2]) + __popc(r.mask[3]);
    else r.cnt >>= 7;  // accumulated as 0x80 x count
  }
}

// ---- pass 1: every byte once; per-granule occurrence counts (+ newline table, + occurrence masks) ----
// No inter-warp communication at all: each warp owns its granule and writes its own counters.  With
// MASK the 512-byte occurrence bitmask of a granule is stored when (and only when) the granule holds
// an occurrence, so pass 2 never touches the corpus again.
template <int NS, bool NL, bool MASK, bool ICASE, bool OVL, bool STAGED = false>
__device__ __forceinline__ bool count_tile(const Pattern& pat, const ScanArgs& a, uint64_t tile, int lane, int warp,
                                           uint32_t stile = 0) {
  const uint64_t g = tile * kWarps + warp;
  const uint64_t gran_off = g * (uint64_t)kGranuleBytes;
  uint4 v[kRows];
  uint32_t ext = 0;
  const uint32_t srow = stile + (uint32_t)warp * kGranuleBytes + (uint32_t)lane * 16;  // this lane's 16 bytes of row 0
  if (STAGED) {  // bytes behind n were zeroed in shared memory: no masking here
#pragma unroll
    for (int u = 0; u < kRows; ++u)
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "r"(srow + u * kRowBytes));
  } else {
    load_granule(a.buf, a.n, gran_off, lane, v, ext);
  }
  if (OVL) {
    // give up once any CTA has seen an overlap (see count_kernel).  The flag is read behind the granule
    // loads, so its latency hides under theirs; the loads of this one tile are all an aborted pass wastes.
    unsigned int seen;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.overlap) : "memory");
    if (seen != 0) return true;
  }
  LaneResult r;
  classify_granule<NS, MASK, NL, ICASE, OVL, STAGED>(pat, a.buf, a.n, gran_off, lane, v, ext, a.konst, r, srow);
  if (OVL && r.ovl != 0) *a.overlap = 1u;
  if (NL) {
    // '\n' per half of the granule, both counts in ONE warp reduction (a lane holds 64 bytes of each half)
    constexpr uint32_t kLaneHalf = kGranuleBytes / 2 / 32;
    const uint32_t mine = (kLaneHalf - (r.nonnlh[0] >> 7)) | ((kLaneHalf - (r.nonnlh[1] >> 7)) << 16);
    const uint32_t both = warp_sum(mine);
    if (lane == 0) a.nl_granule[g] = both;
  }
  const uint32_t wcnt = warp_sum(r.cnt);
  if (lane == 0) a.gran_count[g] = wcnt;
  if (MASK && wcnt == 1) {
    // the usual hit granule of a sparse search: its one position goes into a 2-byte table and pass 2 needs no
    // mask (512 bytes written here and read there for one set bit)
    if (r.cnt != 0) {
      uint32_t bit = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (r.mask[j] != 0) bit = 32 * j + __ffs(r.mask[j]) - 1;  // bit 16 u + b: row u, byte b of this lane
      a.gran_first[g] = (uint16_t)((bit >> 4) * (kGranuleBytes / kRows) + lane * 16 + (bit & 15));
    }
  } else if (MASK && wcnt != 0) {
    // stored as [row][lane] halfwords (eight 64-byte warp stores): 16 consecutive bytes of that layout are
    // the mask bits of 128 consecutive corpus bytes, which is what one lane of pass 2 expands
    uint16_t* __restrict__ mp = reinterpret_cast<uint16_t*>(a.gran_mask + g * 32) + lane;
#pragma unroll
    for (int u = 0; u < kRows; ++u) mp[u * 32] = (uint16_t)(r.mask[u >> 1] >> (16 * (u & 1)));
  }
  return false;
}

template <int NS, bool NL, bool MASK, bool ICASE, bool OVL = false>
__global__ void __launch_bounds__(kThreads, (NS >= 5 || OVL) ? 4 : 3)
count_kernel(const __grid_constant__ Pattern pat, const __grid_constant__ ScanArgs a) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  pdl_trigger();  // the kernel behind this one (prefix sums / reduction) waits for this grid at its top
  if (!OVL) {  // one tile per CTA
    count_tile<NS, NL, MASK, ICASE, OVL>(pat, a, blockIdx.x, lane, warp);
    return;
  }
  // OVL variant: a few CTAs per SM stride over the tiles.  Once any CTA has seen overlapping occurrences the
  // host will redo the search on the list path, so the rest of this pass is wasted work and wasted HBM
  // traffic: every CTA gives up at its next tile (with a strided grid that really ends the pass; a
  // tile-per-CTA grid would still have to drain a million empty CTAs).
  for (uint64_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x)
    if (count_tile<NS, NL, MASK, ICASE, OVL>(pat, a, tile, lane, warp)) return;
}

// ---- EXPERIMENT (north_star: "shared-memory (or TMA) staging of chunk tiles plus a pattern-length halo"): the CTA's
// 32 KiB tile + 16 halo bytes arrive in shared memory by ONE bulk async copy (cp.async.bulk, completion on an mbarrier;
// SASS: UBLKCP), the warps then read their rows with ld.shared.v4 and the halo word with ld.shared.u32 instead of a
// shuffle + select; bytes behind n are zeroed in shared memory.  Built for the exact-compare variants (m <= 4), which
// are bound by the ALU pipe: the question is whether taking the 64-bit address arithmetic, the ragged-tile masking and
// the halo select off that pipe pays for the extra trip through shared memory.  XSG_STAGED=1 selects it; the number
// is in DESIGN.md 8.
template <int NS, bool NL, bool MASK, bool ICASE>
__global__ void __launch_bounds__(kThreads, 3)
count_kernel_staged(const __grid_constant__ Pattern pat, const __grid_constant__ ScanArgs a) {
  extern __shared__ __align__(128) uint8_t s_tile[];  // kTileBytes + 16 halo bytes, then the mbarrier
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  pdl_trigger();
  const uint32_t stile = (uint32_t)__cvta_generic_to_shared(s_tile);
  const uint32_t bar = stile + kTileBytes + 16;
  const uint64_t off = (uint64_t)blockIdx.x * kTileBytes;
  const uint64_t readable = (a.n + 15) & ~15ull;  // the corpus is readable up to the next 16-byte boundary
  const uint32_t bytes = (uint32_t)(readable - off < kTileBytes + 16 ? readable - off : kTileBytes + 16);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(stile), "l"(a.buf + off), "r"(bytes), "r"(bar) : "memory");
  }
  // what the copy does not fill (ragged last tile) and the bytes behind n: zero
  const uint64_t valid = a.n - off < kTileBytes + 16 ? a.n - off : kTileBytes + 16;
  if (bytes < kTileBytes + 16)
    for (uint32_t i = bytes + threadIdx.x; i < kTileBytes + 16; i += kThreads) s_tile[i] = 0;
  uint32_t done = 0;
  while (!done)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar) : "memory");
  if (valid < bytes) {  // readable but behind n (at most 15 bytes)
    __syncthreads();
    if (threadIdx.x < bytes - valid) s_tile[valid + threadIdx.x] = 0;
  }
  if (bytes < kTileBytes + 16 || valid < bytes) __syncthreads();
  count_tile<NS, NL, MASK, ICASE, false, true>(pat, a, blockIdx.x, lane, warp, stile);
}

// ---- pass 2: ordered emission from the stored masks.  The output position of a granule is the
// exclusive prefix of the granule counts, so there is no inter-CTA dependency.  A warp inspects 32
// granule counters at a time and expands the masks of the non-empty ones (512 bytes each).
//
// Pass 1 stores a mask as [row][lane] halfwords.  Here lane L loads bytes [16 L, 16 L + 16) of it: row L/4,
// lanes 8(L%4)..8(L%4)+7, i.e. the 128 mask bits of the CONTIGUOUS byte range [128 L, 128 L + 128) of the
// granule.  File order is then lane order, one warp scan of the per-lane popcounts yields every slot.
struct GranMask {
  uint32_t w[4];
};

__device__ __forceinline__ void load_gran_mask(const ScanArgs& a, uint64_t g, int lane, GranMask& m) {
  const uint4 t = __ldg(a.gran_mask + g * 32 + lane);
  m.w[0] = t.x;
  m.w[1] = t.y;
  m.w[2] = t.z;
  m.w[3] = t.w;
}

// st.shared.u16 [addr + OFF], val  -- executed only if cond != 0 (a predicated store, no branch)
template <int OFF>
__device__ __forceinline__ void sts16_if(uint32_t addr, uint32_t val, uint32_t cond) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p st.shared.u16 [%0+%3], %1;\n\t}"
      :: "r"(addr), "h"((unsigned short)val), "r"(cond), "n"(OFF) : "memory");
}
// st.shared.v2.u32 [addr], {x, y}  -- executed only if cond != 0
__device__ __forceinline__ void sts64_if(uint32_t addr, uint32_t x, uint32_t y, uint32_t cond) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}"
      :: "r"(addr), "r"(x), "r"(y), "r"(cond) : "memory");
}

constexpr uint32_t kStageCap = 1024;   // hits staged per warp (2 KiB); denser granules (>= 1 hit per 4 bytes) go direct
constexpr uint32_t kStageMin = 16;     // below this many hits staging costs more than it saves
constexpr uint32_t kQueueCap = 128;    // mask words of one granule

// Expansion of one hit granule, dense case.  A lane holds the 128 mask bits of 128 consecutive corpus bytes: on
// English text a 32-bit word of them holds 0.6 hits, so walking "the k-th hit of my j-th word" in lock step leaves
// four lanes out of five idle in every step (round 1: twelve predicated chains per lane, 3.3 warp instructions per
// emitted offset).  Instead the NON-EMPTY words are queued in shared memory first -- every lane appends its own at the
// warp-scanned position, each entry carrying its first output slot -- and the lanes then take one queued word each:
// every lane of a round has at least one hit to place, and a second one in a quarter of the cases.
__device__ __forceinline__ void expand_queued(const GranMask& cur, uint32_t hits_excl, uint32_t nz_excl, uint32_t qn,
                                              int lane, uint2* __restrict__ queue, uint16_t* __restrict__ stage) {
  const uint32_t qaddr = (uint32_t)__cvta_generic_to_shared(queue);
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage);
  uint32_t q = qaddr + 8 * nz_excl;
  uint32_t tag = (hits_excl << 16) | (uint32_t)(lane * 128);  // {first output slot, position of bit 0} of word 0
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t w = cur.w[k];
    sts64_if(q, w, tag, w);
    q += w ? 8u : 0u;
    tag += ((uint32_t)__popc(w) << 16) + 32u;
  }
  __syncwarp();
#pragma unroll 1
  for (uint32_t e = lane; e < qn; e += 32) {
    const uint2 it = queue[e];
    uint32_t w = it.x;
    const uint32_t base = it.y & 0xffffu;
    const uint32_t dst = saddr + 2 * (it.y >> 16);
    sts16_if<0>(dst, base + __ffs(w) - 1, 1u);  // a queued word is not empty
    w &= w - 1;
    sts16_if<2>(dst, base + __ffs(w) - 1, w);
    w &= w - 1;
    uint32_t d = dst + 4;
#pragma unroll 1
    while (w) {  // three and more hits in four bytes of text: rare
      sts16_if<0>(d, base + __ffs(w) - 1, 1u);
      w &= w - 1;
      d += 2;
    }
  }
  __syncwarp();
}

// The dense unit, one granule per round (the variant of emit_kernel that keeps its shared memory small, below).
__device__ __forceinline__ void emit_dense_unit1(const Pattern& pat, const ScanArgs& a, uint64_t g0, int lane,
                                                uint2* __restrict__ queue, uint16_t* __restrict__ stage) {
  const uint32_t qaddr = (uint32_t)__cvta_generic_to_shared(queue);
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage);
  const uint4* __restrict__ mp = a.gran_mask + g0 * 32 + lane;
  unsigned long long base = a.gran_prefix[g0];
  uint64_t pos0 = g0 * (uint64_t)kGranuleBytes - pat.anchor + a.occ_add;
  uint64_t* __restrict__ outp = a.occ + base + lane;  // this lane's first output slot of the current granule
  uint4 nxt = __ldg(mp);
#pragma unroll 1
  for (int i = 0; i < 32; ++i) {
    const uint4 cur = nxt;
    mp += 32;
    if (i < 31) nxt = __ldg(mp);
    const uint32_t c0 = __popc(cur.x), c1 = __popc(cur.y), c2 = __popc(cur.z), c3 = __popc(cur.w);
    const uint32_t n0 = min(c0, 1u), n1 = min(c1, 1u), n2 = min(c2, 1u), n3 = min(c3, 1u);
    // one warp scan for both: hits (low half) and non-empty words (high half) before this lane
    const uint32_t mine = (c0 + c1 + c2 + c3) + ((n0 + n1 + n2 + n3) << 16);
    const uint32_t inc = warp_inclusive_scan(mine);
    const uint32_t all = __shfl_sync(kFull, inc, 31);
    const uint32_t total = all & 0xffffu, qn = all >> 16;
    const uint32_t excl = inc - mine;
    if (base + total <= a.occ_cap) {  // (only an estimate-sized list of the non-direct path can be too short)
      if (total <= kStageCap) {
        // queue the non-empty words: {mask word, first output slot x 2 (a byte offset into the stage) : position of bit 0}
        uint32_t q = qaddr + 8 * (excl >> 16);
        uint32_t tag = ((excl & 0xffffu) << 17) | (uint32_t)(lane * 128);
        sts64_if(q, cur.x, tag, cur.x);
        q += 8 * n0;
        tag += (c0 << 17) + 32u;
        sts64_if(q, cur.y, tag, cur.y);
        q += 8 * n1;
        tag += (c1 << 17) + 32u;
        sts64_if(q, cur.z, tag, cur.z);
        q += 8 * n2;
        tag += (c2 << 17) + 32u;
        sts64_if(q, cur.w, tag, cur.w);
        __syncwarp();
#pragma unroll 1
        for (uint32_t e = lane; e < qn; e += 32) {  // one queued word per lane: at least one hit to place
          uint32_t w, t;
          asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w), "=r"(t) : "r"(qaddr + 8 * e));
          const uint32_t bpos = t & 0xffffu;
          const uint32_t dst = saddr + (t >> 16);
          sts16_if<0>(dst, bpos + __ffs(w) - 1, 1u);
          w &= w - 1;
          sts16_if<2>(dst, bpos + __ffs(w) - 1, w);
          w &= w - 1;
          uint32_t d = dst + 4;
#pragma unroll 1
          while (w) {  // three and more hits in four bytes of text: rare
            sts16_if<0>(d, bpos + __ffs(w) - 1, 1u);
            w &= w - 1;
            d += 2;
          }
        }
        __syncwarp();
        // coalesced write-out: 32 consecutive offsets per store instruction
        uint64_t* __restrict__ o = outp;
#pragma unroll 1
        for (uint32_t j = lane; j < total; j += 32) {
          uint32_t v;
          asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(saddr + 2 * j));
          *o = pos0 + v;
          o += 32;
        }
        __syncwarp();  // stage and queue are rewritten for the next granule
      } else {  // denser than one hit per four bytes: every lane stores its own positions
        uint64_t* __restrict__ o = outp - lane + (excl & 0xffffu);
        const uint64_t p = pos0 + (uint32_t)(lane * 128);
        const uint32_t ws[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t w = ws[k];
#pragma unroll 1
          while (w) {
            const int b = __ffs(w) - 1;
            w &= w - 1;
            *o++ = p + (uint32_t)(32 * k + b);
          }
        }
      }
    }
    base += total;
    outp += total;
    pos0 += kGranuleBytes;
  }
}

// A unit whose 32 granules ALL hold two or more occurrences (a frequent short pattern): no search for the next hit
// granule, no per-granule base from a shuffle -- the output of the unit is one contiguous run, so the base just
// advances by every round's total -- and 32-bit shared-memory addresses and pointer increments throughout: this
// kernel is bound by instruction issue (ncu: 91 % of the issue slots), every instruction per granule is time.
// The granules are taken TWO per round: one packed scan serves both (hits of A in the low half, of B in the high
// half; the same for the non-empty words), and the fixed cost of a round -- loop control, capacity check, the warp
// barriers around queue and stage, pointer updates -- is paid once per 8 KiB of text instead of once per 4 KiB.
__device__ __forceinline__ void enqueue_words(const uint4& m, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t n0, uint32_t n1,
                                              uint32_t n2, uint32_t q, uint32_t tag) {
  // {mask word, first output slot x 2 (a byte offset into the stage) : position of bit 0}, non-empty words only
  sts64_if(q, m.x, tag, m.x);
  q += 8 * n0;
  tag += (c0 << 17) + 32u;
  sts64_if(q, m.y, tag, m.y);
  q += 8 * n1;
  tag += (c1 << 17) + 32u;
  sts64_if(q, m.z, tag, m.z);
  q += 8 * n2;
  tag += (c2 << 17) + 32u;
  sts64_if(q, m.w, tag, m.w);
}

__device__ __forceinline__ void store_own_positions(const uint4& m, uint64_t* __restrict__ o, uint64_t p) {
  const uint32_t ws[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t w = ws[k];
#pragma unroll 1
    while (w) {
      const int b = __ffs(w) - 1;
      w &= w - 1;
      *o++ = p + (uint32_t)(32 * k + b);
    }
  }
}

__device__ __forceinline__ void emit_dense_unit(const Pattern& pat, const ScanArgs& a, uint64_t g0, int lane,
                                                uint2* __restrict__ queue, uint16_t* __restrict__ stage) {
  const uint32_t qaddr = (uint32_t)__cvta_generic_to_shared(queue);
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(stage);
  const uint4* __restrict__ mp = a.gran_mask + g0 * 32 + lane;
  unsigned long long base = a.gran_prefix[g0];
  uint64_t pos0 = g0 * (uint64_t)kGranuleBytes - pat.anchor + a.occ_add;
  uint64_t* __restrict__ outp = a.occ + base + lane;  // this lane's first output slot of the current round
  uint4 nxa = __ldg(mp), nxb = __ldg(mp + 32);
#pragma unroll 1
  for (int i = 0; i < 32; i += 2) {
    const uint4 ma = nxa, mb = nxb;
    mp += 64;
    if (i < 30) {
      nxa = __ldg(mp);
      nxb = __ldg(mp + 32);
    }
    const uint32_t a0 = __popc(ma.x), a1 = __popc(ma.y), a2 = __popc(ma.z), a3 = __popc(ma.w);
    const uint32_t b0 = __popc(mb.x), b1 = __popc(mb.y), b2 = __popc(mb.z), b3 = __popc(mb.w);
    const uint32_t na0 = min(a0, 1u), na1 = min(a1, 1u), na2 = min(a2, 1u), na3 = min(a3, 1u);
    const uint32_t nb0 = min(b0, 1u), nb1 = min(b1, 1u), nb2 = min(b2, 1u), nb3 = min(b3, 1u);
    const uint32_t hits = (a0 + a1 + a2 + a3) + ((b0 + b1 + b2 + b3) << 16);      // granule A low, B high
    const uint32_t nz = (na0 + na1 + na2 + na3) + ((nb0 + nb1 + nb2 + nb3) << 16);
    const uint32_t inc_h = warp_inclusive_scan(hits), inc_n = warp_inclusive_scan(nz);
    const uint32_t all_h = __shfl_sync(kFull, inc_h, 31), all_n = __shfl_sync(kFull, inc_n, 31);
    const uint32_t tot_a = all_h & 0xffffu, total = tot_a + (all_h >> 16);
    const uint32_t qn_a = all_n & 0xffffu, qn = qn_a + (all_n >> 16);
    const uint32_t exc_h = inc_h - hits, exc_n = inc_n - nz;
    const uint32_t slot_a = exc_h & 0xffffu, slot_b = tot_a + (exc_h >> 16);
    if (base + total <= a.occ_cap) {  // (only an estimate-sized list of the non-direct path can be too short)
      if (total <= 2 * kStageCap) {
        enqueue_words(ma, a0, a1, a2, na0, na1, na2, qaddr + 8 * (exc_n & 0xffffu), (slot_a << 17) | (uint32_t)(lane * 128));
        enqueue_words(mb, b0, b1, b2, nb0, nb1, nb2, qaddr + 8 * (qn_a + (exc_n >> 16)),
                      (slot_b << 17) | (uint32_t)(kGranuleBytes + lane * 128));
        __syncwarp();
#pragma unroll 1
        for (uint32_t e = lane; e < qn; e += 32) {  // one queued word per lane: at least one hit to place
          uint32_t w, t;
          asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w), "=r"(t) : "r"(qaddr + 8 * e));
          const uint32_t bpos = t & 0xffffu;
          const uint32_t dst = saddr + (t >> 16);
          sts16_if<0>(dst, bpos + __ffs(w) - 1, 1u);
          w &= w - 1;
          sts16_if<2>(dst, bpos + __ffs(w) - 1, w);
          w &= w - 1;
          uint32_t d = dst + 4;
#pragma unroll 1
          while (w) {  // three and more hits in four bytes of text: rare
            sts16_if<0>(d, bpos + __ffs(w) - 1, 1u);
            w &= w - 1;
            d += 2;
          }
        }
        __syncwarp();
        // coalesced write-out: 32 consecutive offsets per store instruction
        uint64_t* __restrict__ o = outp;
#pragma unroll 1
        for (uint32_t j = lane; j < total; j += 32) {
          uint32_t v;
          asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(saddr + 2 * j));
          *o = pos0 + v;
          o += 32;
        }
        __syncwarp();  // stage and queue are rewritten in the next round
      } else {  // denser than one hit per four bytes: every lane stores its own positions
        store_own_positions(ma, outp - lane + slot_a, pos0 + (uint32_t)(lane * 128));
        store_own_positions(mb, outp - lane + slot_b, pos0 + kGranuleBytes + (uint32_t)(lane * 128));
      }
    }
    base += total;
    outp += total;
    pos0 += 2 * kGranuleBytes;
  }
}

// PAIRS: the dense loop takes two granules per round and the kernel carries 48 KiB of shared memory (4 CTAs per SM) --
// right when most granules hold several occurrences, which only direct mode knows at launch time.  Otherwise: 24 KiB,
// 6 CTAs per SM -- a sparse search is a few waves of latency-bound warps and wants them resident (35 -> 25 us on the
// headline workload).
template <bool PAIRS>
__global__ void __launch_bounds__(kThreads, PAIRS ? 4 : 6)
emit_kernel(const __grid_constant__ Pattern pat, const __grid_constant__ ScanArgs a) {
  pdl_wait();
  pdl_trigger();
  __shared__ uint16_t s_stage[kWarps][(PAIRS ? 2 : 1) * kStageCap];
  __shared__ uint2 s_queue[kWarps][(PAIRS ? 2 : 1) * kQueueCap];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint64_t n_gran = a.n_tiles * kWarps;
  // one warp per unit of 32 granules: the serial part of a warp is the handful of hit granules of its unit.  The sparse
  // variant takes kUnits units per warp and loads their counters -- then the slots and positions of their single-hit
  // granules -- side by side: with one 4-byte load per lane in flight the kernel was bound by memory latency.
  constexpr int kUnits = PAIRS ? 1 : 4;
  const uint64_t gw = ((uint64_t)blockIdx.x * kWarps + warp) * (32 * kUnits);
  if (gw >= n_gran) return;
  uint32_t cu[kUnits];
#pragma unroll
  for (int u = 0; u < kUnits; ++u) {
    const uint64_t gi = gw + 32 * u + lane;
    cu[u] = gi < n_gran ? a.gran_count[gi] : 0;
  }
  {  // no mask for a single occurrence: pass 1 left its position in gran_first
    unsigned long long slot[kUnits];
    uint32_t first[kUnits];
#pragma unroll
    for (int u = 0; u < kUnits; ++u) {
      const uint64_t gi = gw + 32 * u + lane;
      slot[u] = cu[u] == 1 ? a.gran_prefix[gi] : ~0ull;
      first[u] = cu[u] == 1 ? a.gran_first[gi] : 0u;
    }
#pragma unroll
    for (int u = 0; u < kUnits; ++u) {
      const uint64_t gi = gw + 32 * u + lane;
      if (cu[u] == 1 && slot[u] < a.occ_cap) a.occ[slot[u]] = gi * (uint64_t)kGranuleBytes - pat.anchor + a.occ_add + first[u];
    }
  }
#pragma unroll 1
  for (int u = 0; u < kUnits; ++u) {
  const uint64_t g0 = gw + 32 * u;
  const uint64_t gi = g0 + lane;
  uint32_t c = cu[0];
#pragma unroll
  for (int v = 1; v < kUnits; ++v) c = u == v ? cu[v] : c;
  unsigned todo = __ballot_sync(kFull, c > 1);
  if (todo == 0) continue;
  if (todo == kFull) {  // every granule of the unit holds several occurrences: the dense case, as a straight loop
    if (PAIRS) emit_dense_unit(pat, a, g0, lane, s_queue[warp], s_stage[warp]);
    else emit_dense_unit1(pat, a, g0, lane, s_queue[warp], s_stage[warp]);
    continue;
  }
  // lane i fetches the output base of granule g0 + i (one coalesced load, handed out by shuffle below).  The
  // masks are expanded one granule after the other with the next one in flight; to make that next load an L2
  // hit rather than a DRAM access, the masks of the hit granules four ahead are pulled into L2 (only four:
  // prefetching a warp's whole unit put 155 MB of masks in flight, more than L2 holds, and DRAM read them twice)
  const uint32_t lt = (1u << lane) - 1u;
  unsigned long long my_base = 0;
  if (c > 1) {
    my_base = a.gran_prefix[gi];
    if (__popc(todo & lt) < 4) {
      const char* mp = reinterpret_cast<const char*>(a.gran_mask + gi * 32);
#pragma unroll
      for (int i = 0; i < 4; ++i) asm volatile("prefetch.global.L2 [%0];" ::"l"(mp + 128 * i));
    }
  }
  GranMask next{};
  load_gran_mask(a, g0 + (__ffs(todo) - 1), lane, next);
  uint16_t* __restrict__ stage = s_stage[warp];
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const GranMask cur = next;
    const unsigned long long base = __shfl_sync(kFull, my_base, src);
    if (todo) {
      load_gran_mask(a, g0 + (__ffs(todo) - 1), lane, next);
      unsigned t = todo & (todo - 1);
      t &= t - 1;
      t &= t - 1;  // the hit granule four behind the current one
      if (t != 0 && lane < 4)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(a.gran_mask + (g0 + (__ffs(t) - 1)) * 32) + 128 * lane));
    }
    // one warp scan for both: hits (low half) and non-empty words (high half) before this lane
    const uint32_t cnt = __popc(cur.w[0]) + __popc(cur.w[1]) + __popc(cur.w[2]) + __popc(cur.w[3]);
    const uint32_t nzc = (cur.w[0] != 0) + (cur.w[1] != 0) + (cur.w[2] != 0) + (cur.w[3] != 0);
    const uint32_t mine = cnt | (nzc << 16);
    const uint32_t inc = warp_inclusive_scan(mine);
    const uint32_t all = __shfl_sync(kFull, inc, 31);
    const uint32_t total = all & 0xffffu;
    // one capacity check per granule (the list is sized from the exact total; it only overflows on the
    // estimate-based first attempt of the non-direct path)
    if (base + total > a.occ_cap) continue;
    const uint32_t excl = inc - mine;
    uint64_t* __restrict__ out = a.occ + base;
    const uint64_t pos0 = (g0 + src) * (uint64_t)kGranuleBytes - pat.anchor + a.occ_add;
    if (total < kStageMin || total > kStageCap) {
      uint64_t* __restrict__ o = out + (excl & 0xffffu);
      const uint64_t p = pos0 + (uint32_t)(lane * 128);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t w = cur.w[k];
#pragma unroll 1
        while (w) {  // a word rarely holds more than two hits: an unrolled loop is all overhead
          const int b = __ffs(w) - 1;
          w &= w - 1;
          *o++ = p + (uint32_t)(32 * k + b);
        }
      }
    } else {
      expand_queued(cur, excl & 0xffffu, excl >> 16, all >> 16, lane, s_queue[warp], stage);
#pragma unroll 1
      for (uint32_t j = lane; j < total; j += 32) out[j] = pos0 + stage[j];
      __syncwarp();  // stage and queue are rewritten for the next hit granule
    }
  }
  }  // units
}

inline bool staged_enabled() {
  static const bool on = std::getenv("XSG_STAGED") != nullptr;
  return on;
}

template <int NS, bool ICASE>
cudaError_t launch_count_ns(const Pattern& pat, const ScanArgs& a, bool newlines, bool masks, cudaStream_t s) {
  const dim3 grid((unsigned)a.n_tiles), block(kThreads);
  if (a.overlap) {  // list-free xs::count of a bordered pattern of <= 4 bytes (no masks, no newline table)
    const dim3 strided((unsigned)(a.n_tiles < 148u * 4 * 4 ? a.n_tiles : 148u * 4 * 4));
    if (NS <= 4) count_kernel<(NS <= 4 ? NS : 1), false, false, ICASE, true><<<strided, block, 0, s>>>(pat, a);
    else return cudaErrorInvalidValue;
  } else if (NS <= 4 && staged_enabled()) {
    constexpr int N4 = NS <= 4 ? NS : 1;
    const size_t smem = kTileBytes + 16 + 16;
    if (masks && newlines) count_kernel_staged<N4, true, true, ICASE><<<grid, block, smem, s>>>(pat, a);
    else if (masks) count_kernel_staged<N4, false, true, ICASE><<<grid, block, smem, s>>>(pat, a);
    else if (newlines) count_kernel_staged<N4, true, false, ICASE><<<grid, block, smem, s>>>(pat, a);
    else count_kernel_staged<N4, false, false, ICASE><<<grid, block, smem, s>>>(pat, a);
  } else if (masks) {
    if (newlines) count_kernel<NS, true, true, ICASE><<<grid, block, 0, s>>>(pat, a);
    else count_kernel<NS, false, true, ICASE><<<grid, block, 0, s>>>(pat, a);
  } else {
    if (newlines) count_kernel<NS, true, false, ICASE><<<grid, block, 0, s>>>(pat, a);
    else count_kernel<NS, false, false, ICASE><<<grid, block, 0, s>>>(pat, a);
  }
  return cudaGetLastError();
}

template <bool ICASE>
cudaError_t launch_count_ic(const Pattern& pat, const ScanArgs& a, bool newlines, bool masks, cudaStream_t s) {
  switch (pat.m) {
    case 1: return launch_count_ns<1, ICASE>(pat, a, newlines, masks, s);
    case 2: return launch_count_ns<2, ICASE>(pat, a, newlines, masks, s);
    case 3: return launch_count_ns<3, ICASE>(pat, a, newlines, masks, s);
    case 4: return launch_count_ns<4, ICASE>(pat, a, newlines, masks, s);
    default:
      return pat.precheck ? launch_count_ns<6, ICASE>(pat, a, newlines, masks, s)
                          : launch_count_ns<5, ICASE>(pat, a, newlines, masks, s);
  }
}

}  // namespace

cudaError_t launch_count(const Pattern& pat, const ScanArgs& a, bool newlines, bool masks, cudaStream_t s) {
  if (a.n_tiles == 0) return cudaSuccess;
  return pat.icase ? launch_count_ic<true>(pat, a, newlines, masks, s) : launch_count_ic<false>(pat, a, newlines, masks, s);
}

cudaError_t launch_emit(const Pattern& pat, const ScanArgs& a, bool dense, cudaStream_t s) {
  if (a.n_tiles == 0) return cudaSuccess;
  const uint64_t units = div_up(a.n_tiles * kWarps, 32);  // units of 32 granules: one per warp (dense), four per warp (sparse)
  const dim3 block(kThreads);
  if (dense) launch_dependent(emit_kernel<true>, dim3((unsigned)div_up(units, kWarps)), block, s, pat, a);
  else launch_dependent(emit_kernel<false>, dim3((unsigned)div_up(units, 4 * kWarps)), block, s, pat, a);
  return cudaGetLastError();
}

}  // namespace xsg
