// xsg_exchange.cu -- the shard exchange of a multi-GPU search over NVLink peer memory.
//
// Every rank (a process with one GPU, or one of several contexts of one process) scans its own contiguous shard.
// Byte offsets carry the shard base from the start; line indices need the number of '\n' in all lower shards,
// which is only known once those shards have been scanned.  The reference has no counterpart (it runs on one host);
// SURVEY.md 8(e) asks for an all-gather or exclusive scan of the per-shard counts.  Here that exchange is no library
// call and at most one small kernel:
//
//   publish  the scan's own summary kernel (list_kernels.cu: summary_kernel, XSG_SCAN_PUBLISH) stores the 4-word
//            summary {newlines, matches, matching lines, lines} into the table of EVERY rank -- plain stores over
//            NVLink into memory mapped with cudaIpcOpenMemHandle (other processes) or reached through peer access
//            (contexts of this process) -- then a system-scope release of the round number.  xsg_exchange_push()
//            does the same for a finished result (e.g. a streamed, multi-part one) with one launch of its own.
//   rebase   the fix-up kernel acquires the round number of every rank in its own table (spinning while a shard
//            is still being scanned), sums the newline counts of the lower ranks and adds them to the line indices.
//
// Two parities of slots: a rank can only be one round ahead of the slowest one, because every round must be waited
// for (rebase or totals) before the next one is published -- enforced here on the host (XSG_ERR_STATE otherwise) --
// so round s + 2 may overwrite the slots of round s.
// A wait that times out (a peer died) sets the context's mapped error flag and adds nothing; the flag is reported
// by xsg_exchange_totals / xsg_exchange_check and by every copy of a rebased result, and cleared once reported.
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "xsg_internal.h"

namespace xsg {

constexpr long long kSpinLimit = 60ll * 2000000000ll;  // ~60 s of clock64 ticks: then give up and flag it

namespace {

__global__ void __launch_bounds__(kMaxWorld) exchange_push_kernel(const __grid_constant__ ExPublish pub, unsigned long long v0,
                                                                  unsigned long long v1, unsigned long long v2,
                                                                  unsigned long long v3) {
  pdl_wait();
  pdl_trigger();
  const uint32_t p = threadIdx.x;
  if (p >= pub.world) return;
  ExSlot* slot = pub.peers.table[p] + (pub.round & 1) * pub.world + pub.rank;
  slot->v[0] = v0;
  slot->v[1] = v1;
  slot->v[2] = v2;
  slot->v[3] = v3;
  st_release_sys(&slot->seq, pub.round);  // orders the four stores above before the round number, system-wide
}

// waits until every rank has published `round` in this rank's table; returns false (and sets *bad) on time-out
__device__ __forceinline__ bool wait_round(const ExSlot* slots, uint32_t world, unsigned long long round,
                                           volatile unsigned int* bad) {
  __shared__ unsigned int s_bad;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (uint32_t r = threadIdx.x; r < world; r += blockDim.x) {
    while (ld_acquire_sys(&slots[r].seq) < round) {
      if (clock64() - t0 > kSpinLimit) {
        *bad = 1u;
        s_bad = 1u;
        break;
      }
    }
  }
  __syncthreads();
  return s_bad == 0;
}

__global__ void __launch_bounds__(256) exchange_rebase_kernel(uint64_t* v, uint64_t n, const ExSlot* local, uint32_t world,
                                                              uint32_t rank, unsigned long long round,
                                                              volatile unsigned int* bad) {
  pdl_wait();  // v was written by the kernels of the scan (the chain ends in summary_kernel, which publishes)
  pdl_trigger();
  const ExSlot* slots = local + (round & 1) * world;
  if (!wait_round(slots, world, round, bad)) return;  // a peer never published: add nothing, the flag reports it
  uint64_t delta = 0;
  for (uint32_t r = 0; r < rank; ++r) delta += slots[r].v[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    v[i] += delta;
}

__global__ void exchange_gather_kernel(const ExSlot* local, uint32_t world, unsigned long long round,
                                       unsigned long long* out, volatile unsigned int* bad) {
  pdl_wait();
  pdl_trigger();
  const ExSlot* slots = local + (round & 1) * world;
  if (!wait_round(slots, world, round, bad)) return;
  for (uint32_t i = threadIdx.x; i < world * 4; i += blockDim.x) out[i] = slots[i / 4].v[i % 4];
}

}  // namespace

// the next round of the exchange attached to ctx, as the argument of summary_kernel (xsg_scan with XSG_SCAN_PUBLISH)
int exchange_begin_round(xsg_ctx* ctx, ExPublish* pub) {
  xsg_exchange* x = ctx->exchange;
  if (!x || !x->connected) {
    set_error("XSG_SCAN_PUBLISH: no connected exchange is attached to this context");
    return XSG_ERR_STATE;
  }
  if (x->round != x->waited) {
    set_error("xsg_exchange: the previous round was not waited for (xsg_result_rebase_exchanged / xsg_exchange_totals)");
    return XSG_ERR_STATE;
  }
  ++x->round;
  pub->peers = x->peers;
  pub->world = x->world;
  pub->rank = x->rank;
  pub->round = x->round;
  return XSG_OK;
}

// reports (and clears) a timed-out wait; call with the context's stream synchronized
int exchange_check_flag(xsg_ctx* ctx) {
  if (ctx->h_ex_bad && *ctx->h_ex_bad) {
    *ctx->h_ex_bad = 0;
    set_error("xsg_exchange: timed out waiting for a peer's shard summary; the rebased line indices are not valid");
    return XSG_ERR_STATE;
  }
  return XSG_OK;
}

}  // namespace xsg

using namespace xsg;

extern "C" {

int xsg_exchange_create(xsg_ctx* ctx, uint32_t world, uint32_t rank, xsg_exchange** out, void* handle_out) {
  if (!ctx || !out || world == 0 || world > (uint32_t)kMaxWorld || rank >= world) return XSG_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == XSG_EXCHANGE_HANDLE_BYTES, "handle size");
  *out = nullptr;
  if (ctx->exchange) {
    set_error("xsg_exchange_create: the context already has an exchange");
    return XSG_ERR_STATE;
  }
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
  if (e == cudaSuccess) e = cudaHostAlloc(&x->h_out, (size_t)world * 4 * sizeof(unsigned long long), cudaHostAllocDefault);
  if (e == cudaSuccess && handle_out) {  // handle_out == NULL: in-process use only (xsg_exchange_connect_local)
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, x->local);
    if (e == cudaSuccess) std::memcpy(handle_out, &h, sizeof h);
  }
  if (e != cudaSuccess) {
    xsg_exchange_destroy(x);
    return cuda_fail(e, "xsg_exchange_create", __FILE__, __LINE__);
  }
  ctx->exchange = x;
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

// All ranks live in THIS process (one context per device, xsg_init_multi): the tables are reached through peer
// access, no IPC handle is involved.  all[r] must be the exchange of rank r.
int xsg_exchange_connect_local(xsg_exchange* const* all, uint32_t n) {
  if (!all || n == 0) return XSG_ERR_ARG;
  for (uint32_t r = 0; r < n; ++r)
    if (!all[r] || all[r]->world != n || all[r]->rank != r || all[r]->connected) {
      set_error("xsg_exchange_connect_local: all[r] must be the unconnected exchange of rank r of a world of n");
      return XSG_ERR_ARG;
    }
  for (uint32_t a = 0; a < n; ++a) {
    XSG_CUDA(cudaSetDevice(all[a]->ctx->device));
    for (uint32_t b = 0; b < n; ++b) {
      const int da = all[a]->ctx->device, db = all[b]->ctx->device;
      if (da != db) {
        int can = 0;
        XSG_CUDA(cudaDeviceCanAccessPeer(&can, da, db));
        if (!can) {
          set_error("xsg_exchange_connect_local: the devices cannot reach each other's memory");
          return XSG_ERR_CUDA;
        }
        const cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        cudaGetLastError();
      }
      all[a]->peers.table[b] = all[b]->local;
    }
  }
  for (uint32_t r = 0; r < n; ++r) all[r]->connected = true;
  return XSG_OK;
}

int xsg_exchange_push(xsg_exchange* x, const xsg_result* r) {
  if (!x || !r || r->ctx != x->ctx) return XSG_ERR_ARG;
  if (!x->connected) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  ExPublish pub;
  int rc = exchange_begin_round(ctx, &pub);
  if (rc) return rc;
  // the totals of a finished result are known on the host (every scan ends with a read-back), whatever number of
  // parts it has: they travel as kernel arguments
  launch_dependent(exchange_push_kernel, dim3(1), dim3(kMaxWorld), ctx->stream, pub, (unsigned long long)r->newlines,
                   (unsigned long long)r->count, (unsigned long long)r->count_lines, (unsigned long long)r->n_lines);
  XSG_CUDA(cudaGetLastError());
  return XSG_OK;
}

int xsg_result_rebase_exchanged(xsg_result* r, xsg_exchange* x) {
  if (!r || !x || r->ctx != x->ctx) return XSG_ERR_ARG;
  if (!x->connected || x->round == 0) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  for (const xsg_part& p : r->parts)
    if (p.ctx && p.ctx != ctx) {
      set_error("xsg_result_rebase_exchanged: the result holds parts of other contexts");
      return XSG_ERR_ARG;
    }
  bool any = false;
  for (xsg_part& p : r->parts) {
    if (!p.line_idx || p.n_line == 0) continue;
    const uint64_t want = div_up(p.n_line, 256);
    launch_dependent(exchange_rebase_kernel, dim3((unsigned)(want < 1184 ? want : 1184)), dim3(256), ctx->stream, p.line_idx,
                     p.n_line, (const ExSlot*)x->local, x->world, x->rank, x->round, (volatile unsigned int*)ctx->d_ex_bad);
    any = true;
  }
  if (!any)  // nothing to fix up, but the lockstep of the rounds still needs every rank to wait for every rank
    launch_dependent(exchange_rebase_kernel, dim3(1), dim3(64), ctx->stream, (uint64_t*)nullptr, (uint64_t)0,
                     (const ExSlot*)x->local, x->world, x->rank, x->round, (volatile unsigned int*)ctx->d_ex_bad);
  XSG_CUDA(cudaGetLastError());
  x->waited = x->round;
  r->exchanged = true;
  return XSG_OK;
}

int xsg_exchange_totals(xsg_exchange* x, uint64_t* table_host) {
  if (!x || !table_host) return XSG_ERR_ARG;
  if (!x->connected || x->round == 0) return XSG_ERR_STATE;
  xsg_ctx* ctx = x->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  launch_dependent(exchange_gather_kernel, dim3(1), dim3(256), ctx->stream, (const ExSlot*)x->local, x->world, x->round,
                   x->d_out, (volatile unsigned int*)ctx->d_ex_bad);
  XSG_CUDA(cudaGetLastError());
  XSG_CUDA(cudaMemcpyAsync(x->h_out, x->d_out, (size_t)x->world * 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                           ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(ctx->stream));
  x->waited = x->round;
  if (int rc = exchange_check_flag(ctx)) return rc;
  std::memcpy(table_host, x->h_out, (size_t)x->world * 4 * sizeof(unsigned long long));
  return XSG_OK;
}

int xsg_exchange_check(xsg_exchange* x) {
  if (!x) return XSG_ERR_ARG;
  XSG_CUDA(cudaSetDevice(x->ctx->device));
  XSG_CUDA(cudaStreamSynchronize(x->ctx->stream));
  return exchange_check_flag(x->ctx);
}

void xsg_exchange_destroy(xsg_exchange* x) {
  if (!x) return;
  cudaSetDevice(x->ctx->device);
  cudaStreamSynchronize(x->ctx->stream);
  for (void* p : x->opened) cudaIpcCloseMemHandle(p);
  if (x->local) cudaFree(x->local);
  if (x->d_out) cudaFree(x->d_out);
  if (x->h_out) cudaFreeHost(x->h_out);
  if (x->ctx->exchange == x) x->ctx->exchange = nullptr;
  delete x;
}

}  // extern "C"
