// xsg_common.cuh -- shared declarations of the sm_100a scan-and-match kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>
#include <utility>

#include "../../include/xsg.h"

namespace xsg {

// ---- geometry of the byte-scanning kernel -------------------------------------------------
// A warp owns a contiguous "granule" of ROWS rows x 512 B; lane l holds bytes [16l, 16l+16) of
// every row (one coalesced 16-byte load per row).  A CTA (tile) is WARPS consecutive granules.
constexpr int kRows = 8;                          // rows per warp granule
constexpr int kWarps = 8;                         // warps per CTA
constexpr int kThreads = kWarps * 32;
constexpr uint32_t kRowBytes = 512;               // 32 lanes x 16 B
constexpr uint32_t kGranuleBytes = kRows * kRowBytes;       // 4 KiB: granularity of the newline table
constexpr uint32_t kTileBytes = kWarps * kGranuleBytes;     // 32 KiB per CTA

// ---- pattern descriptor handed to every kernel (lives in the constant bank) ----------------
struct Pattern {
  uint8_t bytes[XSG_MAX_PATTERN];
  uint32_t m;        // pattern length
  uint32_t nsimd;    // bytes tested SIMD-in-word: m for m <= 4 (exact), 2 for m >= 5 (filter)
  uint32_t anchor;   // m >= 5: pattern index a of the first filter byte (second is a + 4)
  uint32_t rep[4];   // the SIMD-tested bytes replicated into all four lanes of a word
  uint32_t bordered; // pattern overlaps itself (has a proper border) -> greedy resolution needed
  // 0x0a0a0a0a and 0x7f7f7f7f as run-time values: read from the constant bank they can BOTH be operands of
  // one LOP3 ((w ^ c_nl) & c_7f); as literals the compiler spends one instruction per immediate
  uint32_t c_nl, c_7f;
  uint32_t icase;    // ignore case: bytes[] is already lowered, corpus bytes are folded (ASCII A-Z) before every compare
  // the value 1 as a run-time multiplier: a + c written as mad.lo(a, one, c) is issued as IMAD on the
  // multiply-add pipe, which idles while the logic ops of the scan saturate the ALU pipe
  uint32_t one;
  uint32_t c_fe;     // 0xfefefeff = -0x01010101 (addend of the has-zero-byte filter)
  // bit d set <=> d is a period of the pattern, 1 <= d < m (a border of length m - d exists): two occurrences can
  // only overlap at such a distance
  uint64_t periods;
  uint32_t word_a;   // m >= 5: pattern bytes [anchor, anchor + 4) as a little-endian word (register pre-check of a candidate)
  uint32_t ovl_shift; // m <= 4, bordered: 8 x the smallest period (byte-lane shift of the overlap test in pass 1)
  uint32_t orm[4];   // ignore case: 0x20202020 where SIMD-tested pattern byte k is a letter (see classify_granule), else 0
  uint32_t word_or;  // ignore case: 0x20 in the lanes of word_a that hold letters
  uint32_t precheck; // m >= 5: the chosen filter bytes are frequent ones -> use the variant that pre-checks candidates in registers
};

// ---- scan kernel launch arguments ---------------------------------------------------------
struct ScanArgs {
  const uint8_t* buf;        // corpus bytes, 16-byte aligned
  uint64_t n;                // corpus size in bytes
  uint64_t n_tiles;          // CTAs of pass 1; n_tiles * kWarps granules
  uint32_t* gran_count;      // [granules] occurrences owned by each 4 KiB granule (ownership by anchor byte)
  uint32_t* nl_granule;      // [granules] '\n' per granule (newline mode) or nullptr, as TWO 16-bit counts: low half =
                             // first 2 KiB of the granule, high half = second 2 KiB (the line-index stage rescans at
                             // most 1 KiB; one packed warp reduction yields both)
  uint4* gran_mask;          // [granules x 32] occurrence bitmask (16 bits per lane and row) of every granule
                             // that holds two or more occurrences
  uint16_t* gran_first;      // [granules] position in the granule of the occurrence of a granule that holds ONE
                             // with gran_count != 0, written by pass 1 and read by pass 2; nullptr = none
  const uint64_t* gran_prefix;  // [granules] exclusive prefix of gran_count (pass 2)
  uint64_t* occ;             // raw occurrence list (pass 2)
  uint64_t occ_cap;
  uint64_t occ_add;          // added to every emitted position (direct mode: the global offset base)
  unsigned int* overlap;     // OVL variant of pass 1: set to 1 when two occurrences overlap anywhere
  // {rep[0..3]}, {c_nl, c_7f, one, c_fe}, {orm[0..3]} in GLOBAL memory: values loaded from there stay in
  // registers, whereas ptxas re-reads kernel parameters from the constant bank in every row of the scan (nine
  // LDC/LDCU per 16-byte load; it sees through shuffles and empty asm, so a real load it is)
  const uint4* konst;
};

// launchers implemented in scan_kernel.cu
cudaError_t launch_count(const Pattern& pat, const ScanArgs& a, bool newlines, bool masks, cudaStream_t s);  // pat.icase selects the folding variant
// dense: most granules hold several occurrences (known in direct mode): the variant with two granules per round
cudaError_t launch_emit(const Pattern& pat, const ScanArgs& a, bool dense, cudaStream_t s);

// ---- list-domain (post) kernels, list_kernels.cu -------------------------------------------
struct ListArgs {
  const uint8_t* buf;
  uint64_t n;
  const uint64_t* bounds;  // device: exclusive chunk ends
  uint64_t nb;
  const uint64_t* occ;     // sorted raw occurrences
  uint64_t k;              // number of raw occurrences
  uint8_t* sel;            // [k] selection flags (output of select / zone kernels)
  const uint64_t* gran_prefix;  // [n_gran + 1] exclusive prefix of the granule counts (brackets searches in occ) or nullptr
  uint64_t n_gran;
  uint64_t bias;                // the list holds position + bias (direct mode: the global offset base)
  unsigned long long* drops;    // direct mode (sel == nullptr): occurrences the reference's chunk tails do not report
};

// device scalars of one scan (zeroed at the start, read back through a pinned mirror)
struct DevScalars {
  unsigned long long nl_total;
  unsigned long long occ_total;
  unsigned long long zone_sub, zone_add;
  unsigned long long sel_total[2];  // [0] match family, [1] line family
  unsigned long long nb_out;
  unsigned long long drops;
  unsigned int unterminated;
  unsigned int bad_bounds;
  unsigned int ticket[5];  // tile counters: [0], [1] the two prefix-sum tables, [2], [3] compaction (match / line family)
  unsigned int overlap;    // count_kernel<.., OVL>: some pair of occurrences overlaps (bordered patterns)
};
// ---- shard exchange over peer memory (xsg_exchange.cu) ----------------------------------------------
// One 64-byte slot per (parity, source rank) in every rank's table; a source rank stores its 4-word scan
// summary, then releases the round number system-wide.
struct ExSlot {
  unsigned long long v[4];  // {newlines, matches, matching lines, lines}
  unsigned long long seq;   // round number, written last (st.release.sys)
  unsigned long long pad[3];
};
static_assert(sizeof(ExSlot) == 64, "one slot per 64-byte line");
constexpr int kMaxWorld = 64;
struct ExPeers {
  ExSlot* table[kMaxWorld];  // [rank] -> that rank's slots[2][world] (own memory, IPC mapping or peer mapping)
};
// what summary_kernel needs to publish a scan's summary itself (world == 0: nothing to publish)
struct ExPublish {
  ExPeers peers;
  uint32_t world, rank;
  unsigned long long round;
};

// small helpers usable from host code

inline uint64_t div_up(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// ---- device helpers -------------------------------------------------------------------------
#if defined(__CUDACC__)

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// streaming 16-byte load: read-only path, do not allocate in L1 (each byte is read once)
__device__ __forceinline__ uint4 ld_stream_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

// ASCII upper -> lower of one byte / of the four byte lanes of a word (simd::toLower, string_utils.cpp:11-33)
__host__ __device__ __forceinline__ uint8_t fold_byte(uint8_t c) { return (uint8_t)(c - 'A') < 26u ? (uint8_t)(c | 0x20) : c; }
__device__ __forceinline__ uint32_t fold_word(uint32_t w) {
  const uint32_t t = w & 0x7f7f7f7fu;
  const uint32_t ge = t + 0x3f3f3f3fu;  // bit 7 of a lane: low seven bits >= 'A'
  const uint32_t gt = t + 0x25252525u;  // bit 7 of a lane: low seven bits >  'Z'
  return w | (((ge & ~gt & ~w) & 0x80808080u) >> 2);
}

constexpr uint32_t k7F = 0x7f7f7f7fu;
constexpr uint32_t k80 = 0x80808080u;
constexpr uint32_t k01 = 0x01010101u;

// exact per-lane zero test: 0x80 in every byte lane of z that is zero (3 instructions)
__device__ __forceinline__ uint32_t zero_lanes(uint32_t z) {
  uint32_t b = (z & k7F) + k7F;
  return ~(b | z) & k80;
}
// exact per-lane non-zero test: 0x80 in every byte lane of z that is NOT zero
__device__ __forceinline__ uint32_t nonzero_lanes(uint32_t z) {
  uint32_t b = (z & k7F) + k7F;
  return (b | z) & k80;
}
// a + c on the multiply-add pipe (IMAD) instead of the ALU pipe (IADD3 / VIADD); one == 1 at run time
__device__ __forceinline__ uint32_t add_on_fma_pipe(uint32_t a, uint32_t one, uint32_t c) {
  uint32_t d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(c));
  return d;
}
// word-level "has a zero byte" predicate (non-zero result <=> some lane is zero); lanes above the
// lowest zero lane may carry false flags, so use it only as a filter (2 instructions)
__device__ __forceinline__ uint32_t haszero_filter(uint32_t z) { return (z - k01) & ~z; }

// flags at bits 7/15/23/31 of f -> 4 bits in lane order
__device__ __forceinline__ uint32_t nibble16(uint32_t f) { return (f * 0x00204081u) >> 28; }

constexpr unsigned kFull = 0xffffffffu;
constexpr unsigned long long kStatusShift = 62;
constexpr unsigned long long kAggregate = 1ull << kStatusShift;
constexpr unsigned long long kPrefix = 2ull << kStatusShift;
constexpr unsigned long long kValueMask = (1ull << kStatusShift) - 1;

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) { return __reduce_add_sync(kFull, v); }

// inclusive warp scan, two instructions per step: the shuffle itself says (predicate) whether a lane had a source
__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
    asm volatile(
        "{\n\t.reg .u32 t;\n\t.reg .pred p;\n\tshfl.sync.up.b32 t|p, %0, %1, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t}"
        : "+r"(v) : "r"(o));
  return v;
}


// Decoupled look-back (Merrill & Garland): returns the exclusive prefix of `aggregate` over all
// tiles before `tile`; called by warp 0 only.  Tiles are handed out by a ticket counter, so every
// predecessor tile is owned by a CTA that is already running => forward progress.
//
// A descriptor is {status:2, epoch:14, value:48}.  It only counts if it carries the epoch of the launch
// that reads it, so a table needs no clearing between launches: what an earlier launch left behind reads
// as "not published yet".  (The host clears the tables when the 14-bit epoch wraps and when a table is
// reallocated; epoch 0 is never used.)
constexpr unsigned kEpochShift = 48;
constexpr unsigned kEpochMask = 0x3fffu;
constexpr unsigned long long kDescValueMask = (1ull << kEpochShift) - 1;

__device__ __forceinline__ unsigned long long lookback(unsigned long long* desc, uint64_t tile,
                                                       unsigned long long aggregate, int lane, uint32_t epoch) {
  const unsigned long long tag = (unsigned long long)epoch << kEpochShift;
  if (tile == 0) {
    if (lane == 0) st_relaxed_u64(&desc[0], kPrefix | tag | aggregate);
    return 0;
  }
  if (lane == 0) st_relaxed_u64(&desc[tile], kAggregate | tag | aggregate);
  unsigned long long excl = 0;
  long long idx = (long long)tile - 1;
  while (true) {
    const long long my = idx - lane;
    unsigned long long d;
    unsigned pm, need;
    do {
      d = my >= 0 ? ld_relaxed_u64(&desc[my]) : (kPrefix | tag);  // virtual tile -1: prefix 0
      unsigned st = (unsigned)(d >> kStatusShift);
      if (((unsigned)(d >> kEpochShift) & kEpochMask) != epoch) st = 0;  // left over from an earlier launch
      pm = __ballot_sync(kFull, st == 2);
      const unsigned inv = __ballot_sync(kFull, st == 0);
      // only the tiles between this one and the nearest inclusive prefix have to be published
      need = pm ? (inv & ((1u << (__ffs(pm) - 1)) - 1u)) : inv;
    } while (need);
    unsigned long long val = d & kDescValueMask;
    if (pm && lane > __ffs(pm) - 1) val = 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(kFull, val, o);
    excl += val;
    if (pm) break;
    idx -= 32;
  }
  if (lane == 0) st_relaxed_u64(&desc[tile], kPrefix | tag | (excl + aggregate));
  return excl;
}

// ---- programmatic dependent launch ----------------------------------------------------------------
// The kernels of one search form chains on one stream (count -> prefix sums; select -> compact -> format ->
// summary; count -> reduce -> zone count -> summary), each a few microseconds long, so the launch latency
// between them is a visible part of the step.  A kernel launched with launch_dependent() may be scheduled
// while its predecessor is still running: the predecessor says pdl_trigger() at its top ("my CTAs have all
// started, get the next grid ready"), the dependent says pdl_wait() at ITS top, which returns once the
// predecessor grid has completed and its writes are visible.  Nothing is read or written before pdl_wait().
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
  static const bool on = std::getenv("XSG_NO_PDL") == nullptr;
  return on;
}

template <class... KArgs, class... Args>
inline cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

#endif  // __CUDACC__

}  // namespace xsg
