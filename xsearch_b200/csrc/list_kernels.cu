// list_kernels.cu -- everything around the byte scan that works on chunks or on the occurrence
// list instead of on every byte: synthetic corpus fill, v1 chunker, newline prefix table, and the
// exact restatement of the reference's per-chunk result semantics (search_wrappers.h:36-50,
// 111-123,149-154,163-207) including the scalar-tail quirk of simd::strstr
// (simd_search.cpp:58-78,169-171,203) -- see SURVEY.md 3.3 for the model.
#include <cstdlib>
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>

#include "corpus_gen.h"
#include "list_kernels.cuh"

namespace cg = cooperative_groups;

namespace xsg {

namespace {

__constant__ xs_vocab_t d_voc = XS_VOCAB_INIT;

// ---- synthetic corpus ------------------------------------------------------------------------
constexpr int kSynthThreads = 32;  // blocks per CTA; 32 x 4 KiB = 128 KiB of shared memory

__global__ void __launch_bounds__(kSynthThreads) synth_kernel(uint8_t* dst, uint64_t first_block, uint64_t n_blocks,
                                                              const __grid_constant__ xsg_synth_params p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const uint64_t blk0 = (uint64_t)blockIdx.x * kSynthThreads;
  const uint64_t b = blk0 + threadIdx.x;
  if (b < n_blocks) xs_synth_block(smem + threadIdx.x * XSG_SYNTH_BLOCK, first_block + b, p, d_voc);
  __syncthreads();
  const uint64_t nb_here = min((uint64_t)kSynthThreads, n_blocks - blk0);
  const uint32_t n16 = (uint32_t)(nb_here * XSG_SYNTH_BLOCK / 16);
  uint4* out = reinterpret_cast<uint4*>(dst + blk0 * XSG_SYNTH_BLOCK);
  const uint4* in = reinterpret_cast<const uint4*>(smem);
  for (uint32_t i = threadIdx.x; i < n16; i += kSynthThreads) out[i] = in[i];
}

// ---- v1 chunker: >= min_chunk bytes, extended through the next '\n'  -----------
__global__ void chunker_kernel(const uint8_t* __restrict__ buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds,
                               uint64_t cap, unsigned long long* nb_out) {
  const int lane = threadIdx.x;
  uint64_t start = 0, nb = 0;
  while (start < n) {
    uint64_t end = n;
    if (n - start > min_chunk) {
      uint64_t pos = start + min_chunk - 1;
      while (pos < n) {
        const uint64_t p = pos + lane;
        const bool hit = p < n && buf[p] == '\n';
        const unsigned b = __ballot_sync(kFull, hit);
        if (b) {
          end = pos + (__ffs(b) - 1) + 1;
          break;
        }
        pos += 32;
      }
    }
    if (lane == 0 && nb < cap) bounds[nb] = end;
    ++nb;
    start = end;
  }
  if (lane == 0) *nb_out = nb;
}

__global__ void check_bounds_kernel(const uint8_t* __restrict__ buf, uint64_t n, const uint64_t* __restrict__ bounds,
                                    uint64_t nb, unsigned int* bad) {
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nb) return;
  const uint64_t e = bounds[c];
  const uint64_t b = c ? bounds[c - 1] : 0;
  bool ok = e > b && e <= n;
  if (ok && c + 1 < nb) ok = buf[e - 1] == '\n';
  if (ok && c + 1 == nb) ok = e == n;
  if (!ok) atomicOr(bad, 1u);
}

// ---- reductions / scans over the per-tile and per-granule tables -------------------------------
__global__ void reduce_u32_kernel(const uint32_t* __restrict__ in, uint64_t n, unsigned long long* out) {
  pdl_wait();
  pdl_trigger();
  unsigned long long acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    acc += in[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
  __shared__ unsigned long long s[32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (unsigned w = 0; w < blockDim.x / 32; ++w) t += s[w];
    if (t) atomicAdd(out, t);
  }
}

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048

// block-wide exclusive scan of one value per thread (256 threads); returns exclusive prefix, total in *total
__device__ __forceinline__ uint32_t block_exscan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t s_w[kScanThreads / 32];
  __shared__ uint32_t s_total;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(kFull, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_w[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kScanThreads / 32 ? s_w[lane] : 0;
    uint32_t winc = w;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      const uint32_t t = __shfl_up_sync(kFull, winc, o);
      if (lane >= o) winc += t;
    }
    if (lane < kScanThreads / 32) s_w[lane] = winc - w;
    if (lane == kScanThreads / 32 - 1) s_total = winc;
  }
  __syncthreads();
  const uint32_t r = s_w[warp] + inc - v;
  *total = s_total;
  __syncthreads();
  return r;
}

// The scans below handle up to two u32 tables of the same length in one set of launches (blockIdx.y
// selects the table): granule occurrence counts and granule newline counts.
struct Scan2 {
  const uint32_t* in[2];
  uint64_t* out[2];
  unsigned long long* total[2];
  unsigned long long* host_total;  // mapped host memory: the total of table 0 goes there as well (or nullptr)
};

// Single pass (decoupled look-back between the tiles of a table): a tile of 8192 counters is read once, its
// exclusive prefix over the earlier tiles comes from their descriptors, and the 64-bit prefixes are written.
// Large tiles: the tables are small (2.4 M counters for 10 GB), so what this kernel costs is the serial
// look-back chain, and that is as long as the number of tiles.
constexpr int kExItems = 32;
constexpr int kExTile = kScanThreads * kExItems;

__global__ void __launch_bounds__(kScanThreads, 5) exscan_kernel(const __grid_constant__ Scan2 a, uint64_t n, uint64_t nblk,
                                                              unsigned long long* desc_all, unsigned int* tickets,
                                                              uint32_t epoch) {
  __shared__ unsigned int s_tile;
  __shared__ unsigned long long s_base;
  pdl_wait();
  pdl_trigger();
  const uint32_t* __restrict__ in = a.in[blockIdx.y];
  uint64_t* __restrict__ out = a.out[blockIdx.y];
  if (threadIdx.x == 0) s_tile = atomicAdd(&tickets[blockIdx.y], 1u);
  __syncthreads();
  const uint64_t tile = s_tile;
  const uint64_t base = tile * kExTile + (uint64_t)threadIdx.x * kExItems;
  uint32_t x[kExItems];
  uint32_t v = 0;
  const bool full = base + kExItems <= n;
  if (full) {
#pragma unroll
    for (int i = 0; i < kExItems / 8; ++i)  // 32 bytes per load, 64 per store: a thread's items are contiguous, its
      asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"  // neighbours' are 128 / 256 bytes away
                   : "=r"(x[8 * i]), "=r"(x[8 * i + 1]), "=r"(x[8 * i + 2]), "=r"(x[8 * i + 3]), "=r"(x[8 * i + 4]),
                     "=r"(x[8 * i + 5]), "=r"(x[8 * i + 6]), "=r"(x[8 * i + 7])
                   : "l"(in + base + 8 * i));
  } else {
#pragma unroll
    for (int i = 0; i < kExItems; ++i) x[i] = base + i < n ? in[base + i] : 0;
  }
  if (blockIdx.y == 1) {  // the newline table: two 16-bit counts per granule (halves), scanned as their sum
#pragma unroll
    for (int i = 0; i < kExItems; ++i) x[i] = (x[i] & 0xffffu) + (x[i] >> 16);
  }
#pragma unroll
  for (int i = 0; i < kExItems; ++i) v += x[i];
  uint32_t total;
  const uint32_t local = block_exscan(v, &total);
  if (threadIdx.x < 32) {
    const unsigned long long excl = lookback(desc_all + (uint64_t)blockIdx.y * nblk, tile, total, threadIdx.x, epoch);
    if (threadIdx.x == 0) {
      s_base = excl;
      if (tile + 1 == nblk) {
        *a.total[blockIdx.y] = excl + total;
        if (blockIdx.y == 0 && a.host_total) *a.host_total = excl + total;  // the host reads it behind an event: no copy
      }
    }
  }
  __syncthreads();
  uint64_t run = s_base + local;
  if (full) {
    uint64_t* o = out + base;
#pragma unroll
    for (int i = 0; i < kExItems / 4; ++i) {
      const uint64_t r0 = run, r1 = r0 + x[4 * i], r2 = r1 + x[4 * i + 1], r3 = r2 + x[4 * i + 2];
      run = r3 + x[4 * i + 3];
      asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(o + 4 * i), "l"(r0), "l"(r1), "l"(r2), "l"(r3) : "memory");
    }
    if (base + kExItems == n) out[n] = run;
  } else {
#pragma unroll
    for (int i = 0; i < kExItems; ++i) {
      if (base + i < n) out[base + i] = run;
      run += x[i];
    }
    if (base < n) out[n] = run;  // only the thread that holds the ragged end gets here
  }
}

__global__ void add_u64_kernel(uint64_t* v, uint64_t n, uint64_t delta) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    v[i] += delta;
}

__global__ void add_u64_from_table_kernel(uint64_t* v, uint64_t n, const uint64_t* __restrict__ table, uint32_t stride,
                                          uint32_t rank) {
  uint64_t delta = 0;
  for (uint32_t r = 0; r < rank; ++r) delta += table[(uint64_t)r * stride];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    v[i] += delta;
}

__global__ void __launch_bounds__(256) overlap_check_kernel(const uint64_t* __restrict__ occ, uint64_t k, uint32_t m,
                                                            unsigned int* flag) {
  pdl_wait();
  pdl_trigger();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  bool seen = false;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < k; i += stride)
    seen |= occ[i + 1] - occ[i] < m;
  if (seen) *flag = 1u;
}

// Thread 0 writes the summary of the scan; with an exchange attached (pub.world != 0) thread p < world also stores it
// into rank p's table over peer memory and releases the round number -- the push of the shard exchange costs no
// launch of its own.  A pass that will be redone (the list-free count saw overlapping occurrences; the direct path
// saw tail drops or overlaps) publishes nothing: the summary of the fall-back pass does.
__global__ void __launch_bounds__(kMaxWorld) summary_kernel(const DevScalars* __restrict__ d, int mode, int want_lines,
                                                           unsigned long long* out, const __grid_constant__ ExPublish pub,
                                                           DevScalars* mirror) {
  pdl_wait();
  pdl_trigger();
  if (mirror && threadIdx.x == 0) *mirror = *d;  // mapped host memory: the host reads the scalars behind the scan's last event
  unsigned long long v[4];
  v[0] = d->nl_total;
  v[1] = (mode == 1 || mode == 3) ? d->occ_total - d->zone_sub + d->zone_add
                                  : (mode == 2 ? d->occ_total - d->drops : d->sel_total[0]);
  v[2] = d->sel_total[1];
  v[3] = want_lines ? d->sel_total[1] - (d->unterminated ? 1 : 0) : 0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) out[i] = v[i];
  }
  const bool redo = (mode == 1 && d->overlap != 0) || (mode == 2 && (d->drops != 0 || d->overlap != 0)) ||
                    (mode == 3 && d->zone_sub != d->zone_add);
  if (threadIdx.x < pub.world && !redo) {
    ExSlot* slot = pub.peers.table[threadIdx.x] + (pub.round & 1) * pub.world + pub.rank;
#pragma unroll
    for (int i = 0; i < 4; ++i) slot->v[i] = v[i];
    st_release_sys(&slot->seq, pub.round);  // orders the four stores above before the round number, system-wide
  }
}

}  // namespace

void launch_overlap_check(const uint64_t* occ, uint64_t k, uint32_t m, unsigned int* flag, cudaStream_t s) {
  if (k < 2 || m < 2) return;
  const uint64_t want = div_up(k, 256);
  launch_dependent(overlap_check_kernel, dim3((unsigned)(want < 148 * 16 ? want : 148 * 16)), dim3(256), s, occ, k, m, flag);
}

void launch_summary(const void* dev_scalars, int mode, int want_lines, unsigned long long* summary, const ExPublish* pub,
                    cudaStream_t s, void* host_mirror) {
  static const ExPublish none = {};
  const ExPublish& p = pub ? *pub : none;
  launch_dependent(summary_kernel, dim3(1), dim3(p.world ? kMaxWorld : 1), s, static_cast<const DevScalars*>(dev_scalars), mode,
                   want_lines, summary, p, static_cast<DevScalars*>(host_mirror));
}

void launch_add_u64_from_table(uint64_t* v, uint64_t n, const uint64_t* table, uint32_t stride, uint32_t rank,
                               cudaStream_t s) {
  if (n == 0 || rank == 0) return;
  const uint64_t want = div_up(n, 256);
  add_u64_from_table_kernel<<<(unsigned)(want < 4736 ? want : 4736), 256, 0, s>>>(v, n, table, stride, rank);
}

void launch_add_u64(uint64_t* v, uint64_t n, uint64_t delta, cudaStream_t s) {
  if (n == 0 || delta == 0) return;
  const uint64_t want = div_up(n, 256);
  add_u64_kernel<<<(unsigned)(want < 4736 ? want : 4736), 256, 0, s>>>(v, n, delta);
}

void launch_synth(uint8_t* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params& p, cudaStream_t s) {
  if (n_blocks == 0) return;
  // (the attribute belongs to the function ON THE CURRENT DEVICE: a process that drives several devices sets it on each)
  const int smem = kSynthThreads * XSG_SYNTH_BLOCK;
  cudaFuncSetAttribute(synth_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  // grid.x is limited to 2^31-1 CTAs of 32 blocks (128 KiB) each: far beyond 180 GB
  synth_kernel<<<(unsigned)div_up(n_blocks, kSynthThreads), kSynthThreads, smem, s>>>(dst, first_block, n_blocks, p);
}

void launch_chunker(const uint8_t* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap,
                    unsigned long long* nb_out, cudaStream_t s) {
  chunker_kernel<<<1, 32, 0, s>>>(buf, n, min_chunk, bounds, cap, nb_out);
}

void launch_check_bounds(const uint8_t* buf, uint64_t n, const uint64_t* bounds, uint64_t nb, unsigned int* bad_out,
                         cudaStream_t s) {
  if (nb == 0) return;
  check_bounds_kernel<<<(unsigned)div_up(nb, 256), 256, 0, s>>>(buf, n, bounds, nb, bad_out);
}

void launch_reduce_u32(const uint32_t* in, uint64_t n, unsigned long long* out, cudaStream_t s) {
  if (n == 0) return;
  const unsigned grid = (unsigned)(div_up(n, 256 * 8) < 1184 ? div_up(n, 256 * 8) : 1184);
  launch_dependent(reduce_u32_kernel, dim3(grid), dim3(256), s, in, n, out);
}

void launch_exscan2_u32(const uint32_t* in_a, uint64_t* out_a, unsigned long long* total_a, const uint32_t* in_b,
                        uint64_t* out_b, unsigned long long* total_b, uint64_t n, unsigned long long* desc,
                        unsigned int* tickets, uint32_t epoch, cudaStream_t s, unsigned long long* host_total_a) {
  if (n == 0) return;
  const uint64_t nblk = div_up(n, kExTile);
  Scan2 a{};
  a.host_total = host_total_a;
  a.in[0] = in_a; a.out[0] = out_a; a.total[0] = total_a;
  a.in[1] = in_b; a.out[1] = out_b; a.total[1] = total_b;
  const unsigned tables = in_b ? 2 : 1;
  launch_dependent(exscan_kernel, dim3((unsigned)nblk, tables), dim3(kScanThreads), s, a, n, nblk, desc, tickets, epoch);
}

}  // namespace xsg

#include "list_select.inc"
