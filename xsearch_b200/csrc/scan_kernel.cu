// scan_kernel.cu -- the byte-scanning kernel (sm_100a): one pass over the corpus that
//   * finds every true occurrence of the literal (exact SIMD-in-word compare for m <= 4, a two-byte
//     SIMD filter + verification for m >= 5),
//   * counts '\n' per 4 KiB granule (line-index accounting), and
//   * either counts occurrences per tile or emits them, in file order, through a single-pass
//     decoupled look-back scan (stream compaction).
//
// Replaces the inner loops of the reference: simd::strstr / strchr (simd_search.cpp:116-204) driven
// by the findNext / findNextNewLine chains of search_wrappers.h:36-50,168-183,192-205.  The
// sequential "next match -> skip -> next" chain is reformulated data-parallel: this kernel reports
// ALL true occurrences; greedy non-overlap, first-hit-per-line and the reference's chunk-tail quirk
// are resolved on the (much smaller) occurrence list by list_kernels.cu.
//
// Layout: a warp owns a 4 KiB granule (8 rows x 512 B), lane l holds bytes [16l,16l+16) of each row
// => every load is one fully coalesced 16-byte ld.global.nc per thread, 8 independent loads in
// flight per thread; the pattern-length halo is one extra 32-bit word per row obtained with a warp
// shuffle (m <= 4 needs <= 3 halo bytes; the m >= 5 filter tests pattern bytes a and a+4).
//
// Emit mode is a persistent, cooperatively launched kernel: CTA b handles tiles b, b+grid, ... so the
// predecessors of a tile are in flight in the neighbouring CTAs; the per-lane occurrence masks are kept
// as 16 bits per row (4 registers per thread) across the look-back, and tiles without occurrences
// publish their aggregate and skip the look-back altogether.
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
__device__ __noinline__ bool verify_at(const uint8_t* __restrict__ buf, uint64_t q, const Pattern& pat) {
  for (uint32_t k = 0; k < pat.m; ++k)
    if (buf[q + k] != pat.bytes[k]) return false;
  return true;
}

// flags at bits 7/15/23/31 of f -> 4 bits in lane order
__device__ __forceinline__ uint32_t nibble_of(uint32_t f) { return (f * 0x00204081u) >> 28; }

struct LaneResult {
  uint32_t mask[kRows / 2];  // EMIT: 16 occurrence bits per row, two rows per register
  uint32_t cnt;              // occurrences owned by this lane
  uint32_t nonnl;            // bytes that are not '\n' (zero padding included)
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

// Phase 1 for one warp granule: classify, (verify), count.  NS: 1..4 = exact compare of m = NS
// consecutive bytes; 5 = filter on pattern bytes (a, a+4) + verification.
template <int NS, bool EMIT, bool NL>
__device__ __forceinline__ void classify_granule(const Pattern& pat, const uint8_t* __restrict__ buf, uint64_t n,
                                                 uint64_t gran_off, int lane, const uint4 (&v)[kRows], uint32_t ext,
                                                 LaneResult& r) {
  const uint64_t lane_off = gran_off + (uint64_t)lane * 16;
  const bool partial = gran_off + kGranuleBytes + 16 > n;

  r.cnt = 0;
  r.nonnl = 0;
#pragma unroll
  for (int i = 0; i < kRows / 2; ++i) r.mask[i] = 0;
  const long long last_q = (long long)n - (long long)pat.m;  // last valid occurrence start

#pragma unroll
  for (int u = 0; u < kRows; ++u) {
    const uint32_t w0 = v[u].x, w1 = v[u].y, w2 = v[u].z, w3 = v[u].w;
    // halo word: lane l takes word 0 of lane l+1; lane 31 takes word 0 of the next row (held by lane 0)
    uint32_t send = w0;
    if (lane == 0) send = (u < kRows - 1) ? v[(u < kRows - 1) ? u + 1 : u].x : ext;
    const uint32_t w4 = __shfl_sync(kFull, send, (lane + 1) & 31);
    const uint64_t pos0 = lane_off + (uint64_t)u * kRowBytes;

    if (NL) {
      r.nonnl += __popc(nonzero_lanes(w0 ^ 0x0a0a0a0au)) + __popc(nonzero_lanes(w1 ^ 0x0a0a0a0au)) +
                 __popc(nonzero_lanes(w2 ^ 0x0a0a0a0au)) + __popc(nonzero_lanes(w3 ^ 0x0a0a0a0au));
    }

    uint32_t f0 = 0, f1 = 0, f2 = 0, f3 = 0;
    if (NS <= 4) {
      uint32_t z0 = w0 ^ pat.rep[0], z1 = w1 ^ pat.rep[0], z2 = w2 ^ pat.rep[0], z3 = w3 ^ pat.rep[0];
#pragma unroll
      for (int k = 1; k < NS; ++k) {  // byte k of the window: data shifted down by k bytes
        z0 |= __funnelshift_r(w0, w1, 8 * k) ^ pat.rep[k];
        z1 |= __funnelshift_r(w1, w2, 8 * k) ^ pat.rep[k];
        z2 |= __funnelshift_r(w2, w3, 8 * k) ^ pat.rep[k];
        z3 |= __funnelshift_r(w3, w4, 8 * k) ^ pat.rep[k];
      }
      f0 = zero_lanes(z0);
      f1 = zero_lanes(z1);
      f2 = zero_lanes(z2);
      f3 = zero_lanes(z3);
      if (partial) {  // zero padding / bytes behind n must not match
        const long long lim = last_q + 1 - (long long)pos0;
        f0 = keep_lanes(f0, lim);
        f1 = keep_lanes(f1, lim - 4);
        f2 = keep_lanes(f2, lim - 8);
        f3 = keep_lanes(f3, lim - 12);
      }
      r.cnt += __popc(f0) + __popc(f1) + __popc(f2) + __popc(f3);
      if (EMIT && (f0 | f1 | f2 | f3)) {
        const uint32_t m16 = nibble_of(f0) | (nibble_of(f1) << 4) | (nibble_of(f2) << 8) | (nibble_of(f3) << 12);
        r.mask[u >> 1] |= m16 << (16 * (u & 1));
      }
    } else {
      // filter: pattern byte a at this position AND pattern byte a+4 one word further
      const uint32_t z0 = (w0 ^ pat.rep[0]) | (w1 ^ pat.rep[1]);
      const uint32_t z1 = (w1 ^ pat.rep[0]) | (w2 ^ pat.rep[1]);
      const uint32_t z2 = (w2 ^ pat.rep[0]) | (w3 ^ pat.rep[1]);
      const uint32_t z3 = (w3 ^ pat.rep[0]) | (w4 ^ pat.rep[1]);
      const uint32_t any = (haszero_filter(z0) | haszero_filter(z1) | haszero_filter(z2) | haszero_filter(z3)) & k80;
      if (__any_sync(kFull, any != 0)) {
        if (any != 0) {
          const uint32_t e[4] = {zero_lanes(z0), zero_lanes(z1), zero_lanes(z2), zero_lanes(z3)};
          uint32_t m16 = 0;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t word = e[j];
            while (word) {
              const int bit = __ffs(word) - 1;  // 7, 15, 23 or 31
              word &= word - 1;
              const int byte = 4 * j + (bit >> 3);
              const long long q = (long long)(pos0 + byte) - (long long)pat.anchor;
              if (q >= 0 && q <= last_q && verify_at(buf, (uint64_t)q, pat)) m16 |= 1u << byte;
            }
          }
          r.cnt += __popc(m16);
          if (EMIT) r.mask[u >> 1] |= m16 << (16 * (u & 1));
        }
      }
    }
  }
}

// ---- count mode: one tile per CTA, per-tile occurrence count (+ newline table) -------------------
template <int NS, bool NL>
__global__ void __launch_bounds__(kThreads, 3)
count_kernel(const __grid_constant__ Pattern pat, const __grid_constant__ ScanArgs a) {
  __shared__ uint32_t s_warp_cnt[kWarps];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint64_t tile = blockIdx.x;
  const uint64_t gran_off = tile * (uint64_t)kTileBytes + (uint64_t)warp * kGranuleBytes;
  uint4 v[kRows];
  uint32_t ext;
  load_granule(a.buf, a.n, gran_off, lane, v, ext);
  LaneResult r;
  classify_granule<NS, false, NL>(pat, a.buf, a.n, gran_off, lane, v, ext, r);
  if (NL) {
    const uint32_t s = warp_sum(r.nonnl);
    if (lane == 0) a.nl_granule[tile * kWarps + warp] = kGranuleBytes - s;
  }
  const uint32_t wcnt = warp_sum(r.cnt);
  if (lane == 0) s_warp_cnt[warp] = wcnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += s_warp_cnt[w];
    a.tile_count[tile] = t;
  }
}

// ---- emit mode: persistent CTAs, ordered stream compaction through decoupled look-back ------------
template <int NS, bool NL>
__global__ void __launch_bounds__(kThreads, 3)
emit_kernel(const __grid_constant__ Pattern pat, const __grid_constant__ ScanArgs a) {
  __shared__ uint32_t s_warp_cnt2[2][kWarps];  // double-buffered: no barrier at the end of a tile
  __shared__ unsigned long long s_tile_base;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;

  // static strided tile assignment: tile t-1 is processed by the neighbouring CTA in the same round,
  // so predecessors are always in flight (the grid is launched cooperatively => co-resident)
  // Software pipeline: the loads of the next tile are issued before this tile's barrier / look-back /
  // emission, so HBM stays busy while the CTA waits on its predecessors.
  int round = 0;
  uint4 v[kRows];
  uint32_t ext = 0;
  if (blockIdx.x < a.n_tiles)
    load_granule(a.buf, a.n, blockIdx.x * (uint64_t)kTileBytes + (uint64_t)warp * kGranuleBytes, lane, v, ext);
  for (uint64_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, round ^= 1) {
    uint32_t* s_warp_cnt = s_warp_cnt2[round];
    const uint64_t tile_off = tile * (uint64_t)kTileBytes;
    LaneResult r;
    classify_granule<NS, true, NL>(pat, a.buf, a.n, tile_off + (uint64_t)warp * kGranuleBytes, lane, v, ext, r);
    if (tile + gridDim.x < a.n_tiles)
      load_granule(a.buf, a.n, (tile + gridDim.x) * (uint64_t)kTileBytes + (uint64_t)warp * kGranuleBytes, lane, v, ext);
    if (NL) {
      const uint32_t s = warp_sum(r.nonnl);
      if (lane == 0) a.nl_granule[tile * kWarps + warp] = kGranuleBytes - s;
    }
    const uint32_t wcnt = warp_sum(r.cnt);
    if (lane == 0) s_warp_cnt[warp] = wcnt;
    __syncthreads();
    uint32_t total = 0, before = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const uint32_t c = s_warp_cnt[w];
      total += c;
      if (w < warp) before += c;
    }
    if (total == 0) {
      // nothing to place: publish the aggregate so later tiles can look past this one
      if (threadIdx.x == 0) {
        st_relaxed_u64(&a.desc[tile], kAggregate);
        a.tile_count[tile] = 0;
      }
    } else {
      if (warp == 0) {
        const unsigned long long excl = lookback(a.desc, tile, total, lane);
        if (lane == 0) {
          s_tile_base = excl;
          a.tile_count[tile] = total;
          atomicAdd(a.occ_total, (unsigned long long)total);
        }
      }
      __syncthreads();
      if (wcnt != 0) {  // warp-uniform
        unsigned long long base = s_tile_base + before;
        const uint64_t lane_off = tile_off + (uint64_t)warp * kGranuleBytes + (uint64_t)lane * 16 - pat.anchor;
#pragma unroll
        for (int u = 0; u < kRows; ++u) {
          uint32_t m16 = (r.mask[u >> 1] >> (16 * (u & 1))) & 0xffffu;
          const uint32_t c = __popc(m16);
          if (!__any_sync(kFull, c != 0)) continue;
          uint32_t inc = c;  // inclusive warp scan of the per-lane counts of this row
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, inc, o);
            if (lane >= o) inc += t;
          }
          unsigned long long slot = base + (inc - c);
          base += __shfl_sync(kFull, inc, 31);
          const uint64_t pos0 = lane_off + (uint64_t)u * kRowBytes;
          while (m16) {
            const int b = __ffs(m16) - 1;
            m16 &= m16 - 1;
            if (slot < a.occ_cap) a.occ[slot] = pos0 + b;
            ++slot;
          }
        }
      }
    }
  }
}

template <int NS, bool NL>
cudaError_t launch_emit(const Pattern& pat, const ScanArgs& a, int sm_count, cudaStream_t s) {
  // persistent grid sized to what is co-resident; the cooperative launch makes that a checked
  // guarantee (it fails instead of dead-locking the look-back if the grid could not be resident)
  static int per_sm = 0;
  if (per_sm == 0) {
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, emit_kernel<NS, NL>, kThreads, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    if (per_sm > 4) per_sm = 4;
  }
  const uint64_t want = (uint64_t)sm_count * per_sm;
  const dim3 grid((unsigned)(a.n_tiles < want ? a.n_tiles : want)), block(kThreads);
  void* args[] = {const_cast<Pattern*>(&pat), const_cast<ScanArgs*>(&a)};
  return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(emit_kernel<NS, NL>), grid, block, args, 0, s);
}

template <int NS>
cudaError_t launch_ns(const Pattern& pat, const ScanArgs& a, bool emit, bool newlines, int sm_count, cudaStream_t s) {
  if (emit) return newlines ? launch_emit<NS, true>(pat, a, sm_count, s) : launch_emit<NS, false>(pat, a, sm_count, s);
  const dim3 grid((unsigned)a.n_tiles), block(kThreads);
  if (newlines) count_kernel<NS, true><<<grid, block, 0, s>>>(pat, a);
  else count_kernel<NS, false><<<grid, block, 0, s>>>(pat, a);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_scan(const Pattern& pat, const ScanArgs& a, bool emit, bool newlines, int sm_count,
                        cudaStream_t s) {
  if (a.n_tiles == 0) return cudaSuccess;
  switch (pat.m) {
    case 1: return launch_ns<1>(pat, a, emit, newlines, sm_count, s);
    case 2: return launch_ns<2>(pat, a, emit, newlines, sm_count, s);
    case 3: return launch_ns<3>(pat, a, emit, newlines, sm_count, s);
    case 4: return launch_ns<4>(pat, a, emit, newlines, sm_count, s);
    default: return launch_ns<5>(pat, a, emit, newlines, sm_count, s);
  }
}

}  // namespace xsg
