// xsg_exchange.cu -- the shard exchange of a multi-GPU search over NVLink peer memory.
//
// One process per GPU, every rank scans its own contiguous shard.  Byte offsets carry the shard base from the
// start; line indices need the number of '\n' in all lower shards, which is only known once those shards have
// been scanned.  The reference has no counterpart (it runs on one host); SURVEY.md 8(e) asks for an all-gather
// or exclusive scan of the per-shard counts.  Here that exchange is two small kernels and no library call:
//
//   push    (after the scan, same stream)  one thread per peer writes this rank's 4-word scan summary
//           {newlines, matches, matching lines, lines} into THAT peer's table -- plain stores over NVLink into
//           memory mapped with cudaIpcOpenMemHandle -- then a system-scope release of the round number;
//   rebase  the fix-up kernel acquires the round number of every rank in its own table (spinning while a shard
//           is still being scanned), sums the newline counts of the lower ranks and adds them to the line indices.
//
// Two parities of slots: a rank can only be one round ahead of the slowest one (every rebase waits for all ranks),
// so round s + 2 may overwrite the slots of round s.
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "xsg_internal.h"

namespace xsg {

struct ExSlot {
  unsigned long long v[4];
  unsigned long long seq;
  unsigned long long pad[3];
};
static_assert(sizeof(ExSlot) == 64, "one slot per 64-byte line");

constexpr int kMaxWorld = 64;
constexpr long long kSpinLimit = 60ll * 2000000000ll;  // ~60 s of clock64 ticks: then give up and flag it

struct ExPeers {
  ExSlot* table[kMaxWorld];  // [rank] -> that rank's slots[2][world]
};

namespace {

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void exchange_push_kernel(const __grid_constant__ ExPeers peers, uint32_t world, uint32_t rank,
                                     unsigned long long round, const unsigned long long* __restrict__ summary) {
  const uint32_t p = threadIdx.x;
  if (p >= world) return;
  ExSlot* slot = peers.table[p] + (round & 1) * world + rank;
#pragma unroll
  for (int i = 0; i < 4; ++i) slot->v[i] = summary[i];
  st_release_sys(&slot->seq, round);  // orders the four stores above before the round number, system-wide
}

// waits until every rank has published `round` in this rank's table; *bad is set on time-out
__device__ __forceinline__ void wait_round(const ExSlot* slots, uint32_t world, unsigned long long round,
                                           unsigned int* bad) {
  const long long t0 = clock64();
  for (uint32_t r = threadIdx.x; r < world; r += blockDim.x) {
    while (ld_acquire_sys(&slots[r].seq) < round) {
      if (clock64() - t0 > kSpinLimit) {
        *bad = 1u;
        break;
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256) exchange_rebase_kernel(uint64_t* v, uint64_t n, const ExSlot* local, uint32_t world,
                                                              uint32_t rank, unsigned long long round, unsigned int* bad) {
  const ExSlot* slots = local + (round & 1) * world;
  wait_round(slots, world, round, bad);
  uint64_t delta = 0;
  for (uint32_t r = 0; r < rank; ++r) delta += slots[r].v[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    v[i] += delta;
}

__global__ void exchange_gather_kernel(const ExSlot* local, uint32_t world, unsigned long long round,
                                       unsigned long long* out, unsigned int* bad) {
  const ExSlot* slots = local + (round & 1) * world;
  wait_round(slots, world, round, bad);
  for (uint32_t i = threadIdx.x; i < world * 4; i += blockDim.x) out[i] = slots[i / 4].v[i % 4];
}

}  // namespace
}  // namespace xsg

using namespace xsg;

struct xsg_exchange {
  xsg_ctx* ctx = nullptr;
  uint32_t world = 0, rank = 0;
  unsigned long long round = 0;
  ExSlot* local = nullptr;            // slots[2][world], exported to the peers
  ExPeers peers = {};
  std::vector<void*> opened;          // peer mappings to close
  unsigned long long* d_out = nullptr;  // [world * 4] gathered table for xsg_exchange_totals
  unsigned int* d_bad = nullptr;
  unsigned long long* h_out = nullptr;  // pinned [world * 4 + 1]
  bool connected = false;
};

extern "C" {

int xsg_exchange_create(xsg_ctx* ctx, uint32_t world, uint32_t rank, xsg_exchange** out, void* handle_out) {
  if (!ctx || !out || !handle_out || world == 0 || world > (uint32_t)kMaxWorld || rank >= world) return XSG_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == XSG_EXCHANGE_HANDLE_BYTES, "handle size");
  *out = nullptr;
  XSG_CUDA(cudaSetDevice(ctx->device));
  xsg_exchange* x = new (std::nothrow) xsg_exchange();
  if (!x) return XSG_ERR_NOMEM;
  x->ctx = ctx;
  x->world = world;
  x->rank = rank;
  const size_t bytes = 2 * (size_t)world * sizeof(ExSlot);
  cudaError_t e = cudaMalloc(&x->local, bytes);
  if (e == cudaSuccess) e = cudaMemset(x->local, 0, bytes);
  if (e == cudaSuccess) e = cudaMalloc(&x->d_out, (size_t)world * 4 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&x->d_bad, sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaMemset(x->d_bad, 0, sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaHostAlloc(&x->h_out, ((size_t)world * 4 + 1) * sizeof(unsigned long long), cudaHostAllocDefault);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->local);
  if (e != cudaSuccess) {
    xsg_exchange_destroy(x);
    return cuda_fail(e, "xsg_exchange_create", __FILE__, __LINE__);
  }
  std::memcpy(handle_out, &h, sizeof h);
  *out = x;
  return XSG_OK;
}

int xsg_exchange_connect(xsg_exchange* x, const void* handles) {
  if (!x || !handles) return XSG_ERR_ARG;
  if (x->connected) return XSG_ERR_STATE;
  XSG_CUDA(cudaSetDevice(x->ctx->device));
  const cudaIpcMemHandle_t* hs = static_cast<const cudaIpcMemHandle_t*>(handles);
  for (uint32_t r = 0; r < x->world; ++r) {
    if (r == x->rank) {
      x->peers.table[r] = x->local;
      continue;
    }
    void* p = nullptr;
    XSG_CUDA(cudaIpcOpenMemHandle(&p, hs[r], cudaIpcMemLazyEnablePeerAccess));
    x->opened.push_back(p);
    x->peers.table[r] = static_cast<ExSlot*>(p);
  }
  x->connected = true;
  return XSG_OK;
}

int xsg_exchange_push(xsg_exchange* x) {
  if (!x) return XSG_ERR_ARG;
  if (!x->connected) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  ++x->round;
  exchange_push_kernel<<<1, kMaxWorld, 0, ctx->stream>>>(x->peers, x->world, x->rank, x->round, ctx->d_summary);
  XSG_CUDA(cudaGetLastError());
  return XSG_OK;
}

int xsg_result_rebase_exchanged(xsg_result* r, xsg_exchange* x) {
  if (!r || !x || r->ctx != x->ctx) return XSG_ERR_ARG;
  if (!x->connected || x->round == 0) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  bool any = false;
  for (xsg_part& p : r->parts) {
    if (!p.line_idx || p.n_line == 0) continue;
    const uint64_t want = div_up(p.n_line, 256);
    exchange_rebase_kernel<<<(unsigned)(want < 1184 ? want : 1184), 256, 0, ctx->stream>>>(
        p.line_idx, p.n_line, x->local, x->world, x->rank, x->round, x->d_bad);
    any = true;
  }
  if (!any)  // nothing to fix up, but the lockstep of the rounds still needs every rank to wait for every rank
    exchange_rebase_kernel<<<1, 64, 0, ctx->stream>>>(nullptr, 0, x->local, x->world, x->rank, x->round, x->d_bad);
  XSG_CUDA(cudaGetLastError());
  return XSG_OK;
}

int xsg_exchange_totals(xsg_exchange* x, uint64_t* table_host) {
  if (!x || !table_host) return XSG_ERR_ARG;
  if (!x->connected || x->round == 0) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  exchange_gather_kernel<<<1, 256, 0, ctx->stream>>>(x->local, x->world, x->round, x->d_out, x->d_bad);
  XSG_CUDA(cudaGetLastError());
  XSG_CUDA(cudaMemcpyAsync(x->h_out, x->d_out, (size_t)x->world * 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                           ctx->stream));
  XSG_CUDA(cudaMemcpyAsync(x->h_out + (size_t)x->world * 4, x->d_bad, sizeof(unsigned int), cudaMemcpyDeviceToHost,
                           ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(ctx->stream));
  if ((unsigned int)x->h_out[(size_t)x->world * 4] != 0) {
    set_error("xsg_exchange: timed out waiting for a peer's shard summary");
    return XSG_ERR_STATE;
  }
  std::memcpy(table_host, x->h_out, (size_t)x->world * 4 * sizeof(unsigned long long));
  return XSG_OK;
}

void xsg_exchange_destroy(xsg_exchange* x) {
  if (!x) return;
  cudaSetDevice(x->ctx->device);
  cudaStreamSynchronize(x->ctx->stream);
  for (void* p : x->opened) cudaIpcCloseMemHandle(p);
  if (x->local) cudaFree(x->local);
  if (x->d_out) cudaFree(x->d_out);
  if (x->d_bad) cudaFree(x->d_bad);
  if (x->h_out) cudaFreeHost(x->h_out);
  delete x;
}

}  // extern "C"
