// xsg_static_scratch.cu -- xsg_count_static_scratch(): xs::count of a device-resident text without a single
// allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).  No counterpart in the reference.
//
// The list-free count path of xsg_scan (xsg_scan.cu, `!list_path`: count_kernel -> reduce -> zone_count -> summary)
// needs scratch memory: one u32 per 4 KiB granule of text, the scan's 12 constant words and its device scalars.
// xsg_scan takes them from the context (cudaMalloc'ed, grow-only); here they live in ONE file-scope __device__ array
// of fixed size, which every GPU that loads libxsg.so holds until the process exits (1 MiB + 512 B).  That bounds
// the text of one call to XSG_STATIC_SCRATCH_MAX_TEXT bytes (1 GiB).
//
// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, taken once per device with that device current (scratch_of()).  Device code of this
// file names the array directly.
//
// Every call shares the array, and the array keeps the last call's contents: the set-up kernel of a call clears what
// must start at zero (the scalars) and rewrites the constants; the granule counters are all rewritten by pass 1.
// Calls that overlap on the GPU (two streams, two threads, two graphs in flight on one device) race on it.
#include <mutex>

#include "xsg_internal.h"

using namespace xsg;

namespace {

constexpr uint64_t kMaxText = XSG_STATIC_SCRATCH_MAX_TEXT;
constexpr uint64_t kMaxGranules = kMaxText / kTileBytes * kWarps;  // n_tiles * kWarps of the largest text
static_assert(kMaxText % kTileBytes == 0, "the maximum is a whole number of tiles");

struct StaticScratch {
  uint32_t gran_count[kMaxGranules];  // ScanArgs::gran_count
  uint32_t konst[12];                 // ScanArgs::konst (16-byte aligned: the table above is a multiple of 16)
  uint32_t pad[4];
  unsigned long long whole_text;      // the one chunk bound used when the caller passes none (= n)
  DevScalars scal;
};

static __device__ __align__(256) unsigned char g_static_scratch[(sizeof(StaticScratch) + 255) / 256 * 256];

struct KonstWords {
  uint32_t w[12];
};

// first kernel of a call: the array still holds the previous call's scalars and constants
__global__ void __launch_bounds__(64) static_setup_kernel(const KonstWords k, const unsigned long long n) {
  StaticScratch* sc = reinterpret_cast<StaticScratch*>(g_static_scratch);
  if (threadIdx.x < 12) sc->konst[threadIdx.x] = k.w[threadIdx.x];
  uint32_t* z = reinterpret_cast<uint32_t*>(&sc->scal);
  for (uint32_t i = threadIdx.x; i < sizeof(DevScalars) / sizeof(uint32_t); i += blockDim.x) z[i] = 0;
  if (threadIdx.x == 0) sc->whole_text = n;
}

// last kernel of a call: out[0] = xs::count, out[1] = flags
__global__ void __launch_bounds__(32) static_finish_kernel(unsigned long long* out) {
  pdl_wait();
  const StaticScratch* sc = reinterpret_cast<const StaticScratch*>(g_static_scratch);
  if (threadIdx.x == 0) {
    const DevScalars& d = sc->scal;
    out[0] = d.occ_total - d.zone_sub + d.zone_add;
    out[1] = (d.overlap ? (unsigned long long)XSG_STATIC_OVERLAP : 0ull) |
             (d.bad_bounds ? (unsigned long long)XSG_STATIC_BAD_BOUNDS : 0ull);
  }
}

constexpr int kMaxDevices = 64;

// device address of the array on `device` (which must be current): asked for once per device
int scratch_of(int device, StaticScratch** out) {
  static std::mutex mu;
  static StaticScratch* known[kMaxDevices] = {};
  std::lock_guard<std::mutex> lock(mu);
  if (!known[device]) {
    void* p = nullptr;
    XSG_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    known[device] = static_cast<StaticScratch*>(p);
  }
  *out = known[device];
  return XSG_OK;
}

}  // namespace

extern "C" int xsg_count_static_scratch(int device, void* cuda_stream, const void* d_text, uint64_t n,
                                        const uint64_t* d_bounds, uint64_t nb, const uint8_t* pattern, uint32_t m,
                                        uint32_t flags, uint64_t* d_out) {
  if (!d_out || (!d_text && n) || (d_bounds && nb == 0 && n) || device < 0 || device >= kMaxDevices) {
    set_error("xsg_count_static_scratch: bad argument (null text / output, bounds without a count, device out of range)");
    return XSG_ERR_ARG;
  }
  if (n > kMaxText) {
    set_error("xsg_count_static_scratch: text of " + std::to_string(n) + " bytes exceeds the static scratch's maximum of " +
              std::to_string(kMaxText) + " bytes (XSG_STATIC_SCRATCH_MAX_TEXT); split it or use xsg_scan");
    return XSG_ERR_ARG;
  }
  if (reinterpret_cast<uintptr_t>(d_text) % 16 != 0) {
    set_error("xsg_count_static_scratch: device pointer must be 16-byte aligned");
    return XSG_ERR_ARG;
  }
  if (flags & ~XSG_SCAN_IGNORE_CASE) {
    set_error("xsg_count_static_scratch: unknown flag");
    return XSG_ERR_ARG;
  }
  Pattern pat;
  if (int rc = build_pattern(pattern, m, (flags & XSG_SCAN_IGNORE_CASE) != 0, &pat)) return rc;
  if (pat.bordered && pat.m > 4) {  // xsg_scan sends these down the list path (xsg_scan.cu, scan_device)
    set_error("xsg_count_static_scratch: a self-overlapping pattern of more than 4 bytes needs the occurrence list; use xsg_scan");
    return XSG_ERR_ARG;
  }
  XSG_CUDA(cudaSetDevice(device));
  StaticScratch* sc = nullptr;
  if (int rc = scratch_of(device, &sc)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);

  KonstWords k;
  for (int i = 0; i < 4; ++i) k.w[i] = pat.rep[i], k.w[8 + i] = pat.orm[i];
  k.w[4] = pat.c_nl, k.w[5] = pat.c_7f, k.w[6] = pat.one, k.w[7] = pat.c_fe;
  static_setup_kernel<<<1, 64, 0, s>>>(k, n);
  XSG_CUDA(cudaGetLastError());
  if (n > 0) {
    const uint8_t* buf = static_cast<const uint8_t*>(d_text);
    const uint64_t* bounds = d_bounds ? d_bounds : reinterpret_cast<const uint64_t*>(&sc->whole_text);
    if (!d_bounds) nb = 1;
    else launch_check_bounds(buf, n, bounds, nb, &sc->scal.bad_bounds, s);
    ScanArgs a{};
    a.buf = buf;
    a.n = n;
    a.n_tiles = div_up(n, kTileBytes);
    a.gran_count = sc->gran_count;
    a.konst = reinterpret_cast<const uint4*>(sc->konst);
    a.overlap = pat.bordered ? &sc->scal.overlap : nullptr;
    XSG_CUDA(launch_count(pat, a, false, false, s));
    launch_reduce_u32(a.gran_count, a.n_tiles * kWarps, &sc->scal.occ_total, s);
    launch_zone_count(buf, n, bounds, nb, pat, a.gran_count, &sc->scal.zone_sub, &sc->scal.zone_add, s);
    XSG_CUDA(cudaGetLastError());
  }
  XSG_CUDA(launch_dependent(static_finish_kernel, dim3(1), dim3(32), s, reinterpret_cast<unsigned long long*>(d_out)));
  return XSG_OK;
}
