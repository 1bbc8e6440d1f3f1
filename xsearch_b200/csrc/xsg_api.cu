// xsg_api.cu -- implementation of the C ABI declared in include/xsg.h: contexts, corpora,
// chunking and the host-side helpers.  The scan pipeline itself lives in xsg_scan.cu.
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <algorithm>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "corpus_gen.h"
#include "xsg_internal.h"

namespace xsg {

static thread_local std::string g_error;

void set_error(const std::string& msg) { g_error = msg; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  std::snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_error = buf;
  cudaGetLastError();  // clear the sticky-free error state
  return e == cudaErrorMemoryAllocation ? XSG_ERR_NOMEM : XSG_ERR_CUDA;
}

int Scratch::reserve(size_t bytes) {
  if (bytes <= cap) return XSG_OK;
  release();
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    want = bytes;
    e = cudaMalloc(&p, want);
  }
  if (e != cudaSuccess) {
    p = nullptr;
    return cuda_fail(e, "cudaMalloc(scratch)", __FILE__, __LINE__);
  }
  cap = want;
  return XSG_OK;
}

void Scratch::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}

}  // namespace xsg

using namespace xsg;

extern "C" {

const char* xsg_last_error(void) { return g_error.c_str(); }
int xsg_version(void) { return 1; }

// ---- context ----------------------------------------------------------------------------------
// everything xsg_init allocates; any failure leaves a partially built context that xsg_destroy takes apart
static int init_ctx(xsg_ctx* c, int device) {
  c->device = device;
  XSG_CUDA(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
  XSG_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  XSG_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  XSG_CUDA(cudaStreamCreateWithFlags(&c->aux_stream, cudaStreamNonBlocking));
  XSG_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  XSG_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  for (auto& ev : c->ev) XSG_CUDA(cudaEventCreate(&ev));
  for (int i = 0; i < 2; ++i) {
    XSG_CUDA(cudaEventCreateWithFlags(&c->stage_copied[i], cudaEventDisableTiming));
    XSG_CUDA(cudaEventCreateWithFlags(&c->stage_free[i], cudaEventDisableTiming));
  }
  XSG_CUDA(cudaMalloc(&c->d_scalars, sizeof(DevScalars)));
  XSG_CUDA(cudaMalloc(&c->d_summary, 4 * sizeof(unsigned long long)));
  XSG_CUDA(cudaMemset(c->d_summary, 0, 4 * sizeof(unsigned long long)));
  XSG_CUDA(cudaHostAlloc(&c->h_scalars, sizeof(DevScalars), cudaHostAllocMapped));
  XSG_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->d_h_scalars), c->h_scalars, 0));
  XSG_CUDA(cudaMalloc(&c->d_konst, 12 * sizeof(uint32_t)));
  XSG_CUDA(cudaHostAlloc(&c->h_konst, 12 * sizeof(uint32_t), cudaHostAllocDefault));
  XSG_CUDA(cudaHostAlloc(&c->h_ex_bad, sizeof(unsigned int), cudaHostAllocMapped));
  *c->h_ex_bad = 0;
  XSG_CUDA(cudaHostGetDevicePointer(&c->d_ex_bad, c->h_ex_bad, 0));
  // result vectors come from a stream-ordered pool of the context's OWN (freed memory stays in it instead of going
  // back to the driver); the device's default pool -- shared with every other cudaMallocAsync user of the process --
  // is left as it is
  cudaMemPoolProps props = {};
  props.allocType = cudaMemAllocationTypePinned;
  props.handleTypes = cudaMemHandleTypeNone;
  props.location.type = cudaMemLocationTypeDevice;
  props.location.id = device;
  XSG_CUDA(cudaMemPoolCreate(&c->pool, &props));
  unsigned long long thr = ~0ull;
  XSG_CUDA(cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &thr));
  return XSG_OK;
}

int xsg_init(int device, xsg_ctx** out) {
  if (!out) return XSG_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    // No CPU fallback on purpose: the product path is the CUDA path.
    set_error("xsg_init: no CUDA device available (the B200 library has no CPU fallback)");
    cudaGetLastError();
    return XSG_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) {
    set_error("xsg_init: device index out of range");
    return XSG_ERR_ARG;
  }
  XSG_CUDA(cudaSetDevice(device));
  xsg_ctx* c = new (std::nothrow) xsg_ctx();
  if (!c) return XSG_ERR_NOMEM;
  const int rc = init_ctx(c, device);
  if (rc != XSG_OK) {
    const std::string why = g_error;  // xsg_destroy's own CUDA calls must not overwrite the reason
    c->device = device;
    xsg_destroy(c);
    cudaGetLastError();
    g_error = why;
    return rc;
  }
  *out = c;
  return XSG_OK;
}

// One host process, several devices (xsg_init(const int* devs, int n, ...)): n contexts, one per
// device, with peer access enabled between all pairs that support it (for xsg_exchange_connect_local).  On failure
// nothing is left behind.
int xsg_init_multi(const int* devices, int n, xsg_ctx** out) {
  if (!devices || !out || n <= 0) return XSG_ERR_ARG;
  for (int i = 0; i < n; ++i) out[i] = nullptr;
  for (int i = 0; i < n; ++i) {
    for (int j = 0; j < i; ++j)
      if (devices[i] == devices[j]) {
        set_error("xsg_init_multi: a device is listed twice");
        for (int k = 0; k < i; ++k) xsg_destroy(out[k]), out[k] = nullptr;
        return XSG_ERR_ARG;
      }
    const int rc = xsg_init(devices[i], &out[i]);
    if (rc != XSG_OK) {
      const std::string why = g_error;
      for (int k = 0; k < i; ++k) xsg_destroy(out[k]), out[k] = nullptr;
      g_error = why;
      return rc;
    }
  }
  for (int i = 0; i < n; ++i) {
    cudaSetDevice(devices[i]);
    for (int j = 0; j < n; ++j) {
      int can = 0;
      if (i == j || cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) != cudaSuccess || !can) continue;
      cudaDeviceEnablePeerAccess(devices[j], 0);  // "already enabled" is fine
      cudaGetLastError();
    }
  }
  return XSG_OK;
}

void xsg_destroy(xsg_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
  if (c->aux_stream) cudaStreamSynchronize(c->aux_stream);
  Scratch* all[] = {&c->gran_count, &c->gran_prefix, &c->nl_granule, &c->nl_prefix, &c->scan_tmp, &c->occ,
                    &c->sel, &c->desc2, &c->selq, &c->gran_mask, &c->gran_first, &c->stage[0], &c->stage[1], &c->seg_bounds, &c->stage_bounds[0],
                    &c->stage_bounds[1]};
  for (Scratch* s : all) s->release();
  for (auto& ev : c->ev)
    if (ev) cudaEventDestroy(ev);
  for (int i = 0; i < 2; ++i) {
    if (c->stage_copied[i]) cudaEventDestroy(c->stage_copied[i]);
    if (c->stage_free[i]) cudaEventDestroy(c->stage_free[i]);
  }
  if (c->d_scalars) cudaFree(c->d_scalars);
  if (c->d_summary) cudaFree(c->d_summary);
  if (c->h_scalars) cudaFreeHost(c->h_scalars);
  if (c->d_konst) cudaFree(c->d_konst);
  if (c->h_konst) cudaFreeHost(c->h_konst);
  if (c->h_ex_bad) cudaFreeHost(c->h_ex_bad);
  if (c->pool) cudaMemPoolDestroy(c->pool);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  cudaGetLastError();
  delete c;
}

int xsg_device_info(xsg_ctx* c, int* sm_count, uint64_t* total_mem, char* name, size_t name_cap) {
  if (!c) return XSG_ERR_ARG;
  cudaDeviceProp prop;
  XSG_CUDA(cudaGetDeviceProperties(&prop, c->device));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (total_mem) *total_mem = prop.totalGlobalMem;
  if (name && name_cap) std::snprintf(name, name_cap, "%s", prop.name);
  return XSG_OK;
}

int xsg_set_stream(xsg_ctx* c, void* cuda_stream) {
  if (!c) return XSG_ERR_ARG;
  XSG_CUDA(cudaSetDevice(c->device));
  XSG_CUDA(cudaStreamSynchronize(c->stream));
  c->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : c->own_stream;
  return XSG_OK;
}

// NUMA node of the current CUDA device (sysfs), or -1
static int device_numa_node() {
  int dev = 0;
  char bus[32];
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  for (char* c = bus; *c; ++c) *c = (char)std::tolower((unsigned char)*c);
  char path[128];
  std::snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
  std::FILE* f = std::fopen(path, "r");
  if (!f) return -1;
  int node = -1;
  if (std::fscanf(f, "%d", &node) != 1) node = -1;
  std::fclose(f);
  return node;
}

// Pinned memory on the NUMA node the current device hangs off: on a two-socket box a buffer that lands
// on the other socket halves the H2D bandwidth of a streamed search.  set_mempolicy is called directly
// (no libnuma in the image); failures are ignored, the allocation then simply follows the default policy.
//
// Large buffers are NOT taken from cudaHostAlloc: it faults and pins 4 KiB pages on one thread (measured on the B200
// boxes of this pool: 162 ms per 256 MiB, 0.6 s per GiB -- a one-shot file search spent more time page-locking its
// segment buffers than reading the file).  Instead: 2 MiB-aligned anonymous memory, huge pages advised, first touch on
// eight threads (8 ms per 256 MiB), then cudaHostRegister (5 ms: 128 huge pages instead of 65 536 small ones).
// XSG_HOST_ALLOC=cuda forces the old path.
namespace {
std::mutex g_host_mu;
std::unordered_map<void*, size_t> g_host_registered;  // blocks from the fast path: base -> rounded size

bool bind_to_node(int node) {
#if defined(__linux__) && defined(SYS_set_mempolicy)
  if (node >= 0 && node < 1024) {
    unsigned long mask[16] = {0};
    mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
    return syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, mask, 1025ul) == 0;
  }
#endif
  (void)node;
  return false;
}
void unbind() {
#if defined(__linux__) && defined(SYS_set_mempolicy)
  syscall(SYS_set_mempolicy, 0 /* MPOL_DEFAULT */, nullptr, 0ul);
#endif
}
}  // namespace

int xsg_host_alloc(uint64_t n, void** out) {
  if (!out) return XSG_ERR_ARG;
  *out = nullptr;
  const int node = std::getenv("XSG_NO_NUMA") ? -1 : device_numa_node();
  constexpr size_t kHuge = size_t{2} << 20;
  const char* mode = std::getenv("XSG_HOST_ALLOC");
  if (n >= 4 * kHuge && !(mode && std::strcmp(mode, "cuda") == 0)) {
    const size_t rounded = (size_t)((n + kHuge - 1) / kHuge * kHuge);
    void* p = std::aligned_alloc(kHuge, rounded);
    if (p) {
      ::madvise(p, rounded, MADV_HUGEPAGE);  // advisory: small pages work as well, only slower
      const size_t nthreads = std::min<size_t>(8, std::max<size_t>(1, std::thread::hardware_concurrency()));
      const size_t per = (rounded / kHuge + nthreads - 1) / nthreads * kHuge;
      std::vector<std::thread> th;
      for (size_t t = 0; t < nthreads; ++t)
        th.emplace_back([=] {
          const size_t lo = t * per, hi = std::min(rounded, lo + per);
          if (lo >= hi) return;
          const bool bound = bind_to_node(node);  // the policy is per thread: the thread that faults decides the node
          for (size_t o = lo; o < hi; o += 4096) static_cast<volatile char*>(p)[o] = 0;
          if (bound) unbind();
        });
      for (auto& t : th) t.join();
      const cudaError_t e = cudaHostRegister(p, rounded, cudaHostRegisterPortable);
      if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lock(g_host_mu);
        g_host_registered[p] = rounded;
        *out = p;
        return XSG_OK;
      }
      cudaGetLastError();  // fall through to cudaHostAlloc
      std::free(p);
    }
  }
  const bool bound = bind_to_node(node);
  const cudaError_t e = cudaHostAlloc(out, n ? n : 1, cudaHostAllocPortable);
  if (bound) unbind();
  if (e != cudaSuccess) return cuda_fail(e, "cudaHostAlloc", __FILE__, __LINE__);
  return XSG_OK;
}

void xsg_host_free(void* p) {
  if (!p) return;
  {
    std::lock_guard<std::mutex> lock(g_host_mu);
    auto it = g_host_registered.find(p);
    if (it != g_host_registered.end()) {
      g_host_registered.erase(it);
      cudaHostUnregister(p);
      cudaGetLastError();
      std::free(p);
      return;
    }
  }
  cudaFreeHost(p);
}

// ---- corpus -------------------------------------------------------------------------------------
int xsg_corpus_alloc(xsg_ctx* ctx, uint64_t n, xsg_corpus** out) {
  if (!ctx || !out) return XSG_ERR_ARG;
  *out = nullptr;
  XSG_CUDA(cudaSetDevice(ctx->device));
  xsg_corpus* c = new (std::nothrow) xsg_corpus();
  if (!c) return XSG_ERR_NOMEM;
  c->ctx = ctx;
  c->n = n;
  c->owned = true;
  const size_t padded = (size_t)(div_up(n, 256) * 256 + 256);  // 16-byte loads never leave the allocation
  cudaError_t e = cudaMalloc(&c->dev, padded);
  if (e != cudaSuccess) {
    delete c;
    return cuda_fail(e, "cudaMalloc(corpus)", __FILE__, __LINE__);
  }
  *out = c;
  return XSG_OK;
}

int xsg_corpus_wrap(xsg_ctx* ctx, const void* dev_ptr, uint64_t n, xsg_corpus** out) {
  if (!ctx || !out || (!dev_ptr && n)) return XSG_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(dev_ptr) % 16 != 0) {
    set_error("xsg_corpus_wrap: device pointer must be 16-byte aligned");
    return XSG_ERR_ARG;
  }
  xsg_corpus* c = new (std::nothrow) xsg_corpus();
  if (!c) return XSG_ERR_NOMEM;
  c->ctx = ctx;
  c->dev = const_cast<uint8_t*>(static_cast<const uint8_t*>(dev_ptr));
  c->n = n;
  c->owned = false;
  *out = c;
  return XSG_OK;
}

void xsg_corpus_free(xsg_corpus* c) {
  if (!c) return;
  cudaSetDevice(c->ctx->device);
  cudaStreamSynchronize(c->ctx->stream);
  if (c->owned && c->dev) cudaFree(c->dev);
  if (c->d_bounds) cudaFree(c->d_bounds);
  delete c;
}

uint64_t xsg_corpus_size(const xsg_corpus* c) { return c ? c->n : 0; }
const void* xsg_corpus_device_ptr(const xsg_corpus* c) { return c ? c->dev : nullptr; }

int xsg_corpus_upload(xsg_corpus* c, const void* host, uint64_t offset, uint64_t n) {
  if (!c || (!host && n) || offset + n > c->n) return XSG_ERR_ARG;
  XSG_CUDA(cudaSetDevice(c->ctx->device));
  XSG_CUDA(cudaMemcpyAsync(c->dev + offset, host, n, cudaMemcpyHostToDevice, c->ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(c->ctx->stream));
  return XSG_OK;
}

int xsg_corpus_download(const xsg_corpus* c, void* host, uint64_t offset, uint64_t n) {
  if (!c || (!host && n) || offset + n > c->n) return XSG_ERR_ARG;
  XSG_CUDA(cudaSetDevice(c->ctx->device));
  XSG_CUDA(cudaMemcpyAsync(host, c->dev + offset, n, cudaMemcpyDeviceToHost, c->ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(c->ctx->stream));
  return XSG_OK;
}

int xsg_corpus_fill_synth(xsg_corpus* c, uint64_t first_block, const xsg_synth_params* p) {
  if (!c || !p || c->n % XSG_SYNTH_BLOCK != 0 || p->needle_len > XSG_MAX_PATTERN) return XSG_ERR_ARG;
  XSG_CUDA(cudaSetDevice(c->ctx->device));
  launch_synth(c->dev, first_block, c->n / XSG_SYNTH_BLOCK, *p, c->ctx->stream);
  XSG_CUDA(cudaGetLastError());
  XSG_CUDA(cudaStreamSynchronize(c->ctx->stream));
  return XSG_OK;
}

static int store_bounds(xsg_corpus* c, const uint64_t* d_src, uint64_t nb) {
  if (c->d_bounds) cudaFree(c->d_bounds);
  c->d_bounds = nullptr;
  c->h_bounds.assign(nb, 0);
  if (nb == 0) return XSG_OK;
  XSG_CUDA(cudaMalloc(&c->d_bounds, nb * sizeof(uint64_t)));
  XSG_CUDA(cudaMemcpyAsync(c->d_bounds, d_src, nb * sizeof(uint64_t), cudaMemcpyDeviceToDevice, c->ctx->stream));
  XSG_CUDA(cudaMemcpyAsync(c->h_bounds.data(), d_src, nb * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(c->ctx->stream));
  return XSG_OK;
}

int xsg_corpus_chunk(xsg_corpus* c, uint64_t min_chunk) {
  if (!c) return XSG_ERR_ARG;
  if (min_chunk == 0) min_chunk = XSG_DEFAULT_MIN_CHUNK;
  xsg_ctx* ctx = c->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  const uint64_t cap = c->n / min_chunk + 2;
  int rc = ctx->seg_bounds.reserve(cap * sizeof(uint64_t));
  if (rc) return rc;
  XSG_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, sizeof(DevScalars), ctx->stream));
  launch_chunker(c->dev, c->n, min_chunk, ctx->seg_bounds.as<uint64_t>(), cap, &ctx->d_scalars->nb_out, ctx->stream);
  XSG_CUDA(cudaGetLastError());
  XSG_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, sizeof(DevScalars), cudaMemcpyDeviceToHost, ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(ctx->stream));
  const uint64_t nb = ctx->h_scalars->nb_out;
  if (nb > cap) {
    set_error("xsg_corpus_chunk: internal bound on the number of chunks exceeded");
    return XSG_ERR_STATE;
  }
  return store_bounds(c, ctx->seg_bounds.as<uint64_t>(), nb);
}

int xsg_corpus_set_bounds(xsg_corpus* c, const uint64_t* bounds, uint64_t nb) {
  if (!c || (!bounds && nb)) return XSG_ERR_ARG;
  if ((nb == 0) != (c->n == 0)) {
    set_error("xsg_corpus_set_bounds: bounds must cover the corpus");
    return XSG_ERR_BOUNDS;
  }
  xsg_ctx* ctx = c->ctx;
  XSG_CUDA(cudaSetDevice(ctx->device));
  if (nb == 0) return store_bounds(c, nullptr, 0);
  int rc = ctx->seg_bounds.reserve(nb * sizeof(uint64_t));
  if (rc) return rc;
  uint64_t* d = ctx->seg_bounds.as<uint64_t>();
  XSG_CUDA(cudaMemcpyAsync(d, bounds, nb * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
  XSG_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, sizeof(DevScalars), ctx->stream));
  launch_check_bounds(c->dev, c->n, d, nb, &ctx->d_scalars->bad_bounds, ctx->stream);
  XSG_CUDA(cudaGetLastError());
  XSG_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, sizeof(DevScalars), cudaMemcpyDeviceToHost, ctx->stream));
  XSG_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->h_scalars->bad_bounds) {
    set_error("chunk bounds must be increasing, end at the corpus size, and every bound but the last must follow a '\\n'");
    return XSG_ERR_BOUNDS;
  }
  return store_bounds(c, d, nb);
}

uint64_t xsg_corpus_num_chunks(const xsg_corpus* c) { return c ? c->h_bounds.size() : 0; }

int xsg_corpus_get_bounds(const xsg_corpus* c, uint64_t* bounds, uint64_t cap) {
  if (!c || (!bounds && cap)) return XSG_ERR_ARG;
  const uint64_t k = c->h_bounds.size() < cap ? c->h_bounds.size() : cap;
  if (k) std::memcpy(bounds, c->h_bounds.data(), k * sizeof(uint64_t));
  return XSG_OK;
}

// ---- host-side helpers ------------------------------------------------------------------------
void xsg_synth_host(void* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params* p) {
  static const xs_vocab_t voc = XS_VOCAB_INIT;
  uint8_t* out = static_cast<uint8_t*>(dst);
  for (uint64_t b = 0; b < n_blocks; ++b) xs_synth_block(out + b * XSG_SYNTH_BLOCK, first_block + b, *p, voc);
}

uint64_t xsg_chunk_bounds_host(const void* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap) {
  const uint8_t* b = static_cast<const uint8_t*>(buf);
  if (min_chunk == 0) min_chunk = XSG_DEFAULT_MIN_CHUNK;
  uint64_t nb = 0, start = 0;
  while (start < n) {
    uint64_t end = n;
    if (n - start > min_chunk) {
      const void* r = std::memchr(b + start + min_chunk - 1, '\n', n - (start + min_chunk - 1));
      if (r) end = (uint64_t)(static_cast<const uint8_t*>(r) - b) + 1;
    }
    if (bounds && nb < cap) bounds[nb] = end;
    ++nb;
    start = end;
  }
  return nb;
}

}  // extern "C"
