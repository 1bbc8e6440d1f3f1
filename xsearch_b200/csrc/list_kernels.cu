// list_kernels.cu -- everything around the byte scan that works on chunks or on the occurrence
// list instead of on every byte: synthetic corpus fill, v1 chunker, newline prefix table, and the
// exact restatement of the reference's per-chunk result semantics (search_wrappers.h:36-50,
// 111-123,149-154,163-207) including the scalar-tail quirk of simd::strstr
// (simd_search.cpp:58-78,169-171,203) -- see SURVEY.md 3.3 for the model.
#include "corpus_gen.h"
#include "list_kernels.cuh"

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

__global__ void __launch_bounds__(kScanThreads) exscan_sums_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                                   uint64_t* tmp) {
  const uint64_t base = (uint64_t)blockIdx.x * kScanTile + (uint64_t)threadIdx.x * kScanItems;
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) v += in[base + i];
  uint32_t total;
  block_exscan(v, &total);
  if (threadIdx.x == 0) tmp[blockIdx.x] = total;
}

// single block: exclusive scan of tmp[0..nblk) in place, total -> *total
__global__ void __launch_bounds__(1024) exscan_top_kernel(uint64_t* tmp, uint64_t nblk, unsigned long long* total) {
  __shared__ unsigned long long s_w[32];
  __shared__ unsigned long long s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint64_t base = 0; base < nblk; base += 1024) {
    const uint64_t i = base + threadIdx.x;
    const unsigned long long v = i < nblk ? tmp[i] : 0;
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(kFull, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const unsigned long long w = s_w[lane];
      unsigned long long winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(kFull, winc, o);
        if (lane >= o) winc += t;
      }
      s_w[lane] = winc - w;
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    const unsigned long long ex = carry + s_w[warp] + inc - v;
    if (i < nblk) tmp[i] = ex;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = ex + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = s_carry;
}

__global__ void __launch_bounds__(kScanThreads) exscan_apply_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                                    const uint64_t* __restrict__ tmp, uint64_t* out) {
  const uint64_t base = (uint64_t)blockIdx.x * kScanTile + (uint64_t)threadIdx.x * kScanItems;
  uint32_t x[kScanItems];
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    x[i] = base + i < n ? in[base + i] : 0;
    v += x[i];
  }
  uint32_t total;
  uint64_t run = tmp[blockIdx.x] + block_exscan(v, &total);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = run;
    run += x[i];
  }
}

__global__ void add_u64_kernel(uint64_t* v, uint64_t n, uint64_t delta) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    v[i] += delta;
}

}  // namespace

void launch_add_u64(uint64_t* v, uint64_t n, uint64_t delta, cudaStream_t s) {
  if (n == 0 || delta == 0) return;
  const uint64_t want = div_up(n, 256);
  add_u64_kernel<<<(unsigned)(want < 4736 ? want : 4736), 256, 0, s>>>(v, n, delta);
}

void launch_synth(uint8_t* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params& p, cudaStream_t s) {
  if (n_blocks == 0) return;
  static bool attr_set = false;
  const int smem = kSynthThreads * XSG_SYNTH_BLOCK;
  if (!attr_set) {
    cudaFuncSetAttribute(synth_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    attr_set = true;
  }
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
  reduce_u32_kernel<<<grid, 256, 0, s>>>(in, n, out);
}

void launch_exscan_u32(const uint32_t* in, uint64_t n, uint64_t* out, uint64_t* tmp, unsigned long long* total,
                       cudaStream_t s) {
  if (n == 0) return;
  const uint64_t nblk = div_up(n, kScanTile);
  exscan_sums_kernel<<<(unsigned)nblk, kScanThreads, 0, s>>>(in, n, tmp);
  exscan_top_kernel<<<1, 1024, 0, s>>>(tmp, nblk, total);
  exscan_apply_kernel<<<(unsigned)nblk, kScanThreads, 0, s>>>(in, n, tmp, out);
}

}  // namespace xsg

#include "list_select.inc"
