// list_kernels.cuh -- launchers of the small kernels around the byte scan (list_kernels.cu).
#pragma once
#include "xsg_common.cuh"

namespace xsg {

enum SelectMode : int {
  kSelMatch = 0,          // xs::count / xs::match_byte_offsets, border-free pattern: every occurrence
  kSelMatchBordered = 1,  // same tags, self-overlapping pattern: greedy leftmost non-overlapping
  kSelLine = 2            // xs::count_lines / line_*: first occurrence of every line
};

// synthetic corpus: blocks [first_block, first_block + n_blocks) -> dst
void launch_synth(uint8_t* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params& p, cudaStream_t s);

// v1 chunker on the device: writes bounds (capacity cap) and *nb_out; one warp, sequential chunks
void launch_chunker(const uint8_t* buf, uint64_t n, uint64_t min_chunk, uint64_t* bounds, uint64_t cap,
                    unsigned long long* nb_out, cudaStream_t s);
// *bad_out != 0 if bounds are not increasing / newline-aligned / do not end at n
void launch_check_bounds(const uint8_t* buf, uint64_t n, const uint64_t* bounds, uint64_t nb, unsigned int* bad_out,
                         cudaStream_t s);

// sum of a u32 array into *out (out must be zero before)
void launch_reduce_u32(const uint32_t* in, uint64_t n, unsigned long long* out, cudaStream_t s);
// exclusive scans u32 -> u64 of one or two tables of n entries in ONE launch (out[i] = sum in[0..i),
// out[n] = *total = sum); in_b may be nullptr.  Table B holds two 16-bit counts per entry (ScanArgs::nl_granule): their sum
// is what is scanned.  desc holds 2 * div_up(n, 2048) look-back descriptors (tagged
// with `epoch`, see lookback()), tickets two zeroed counters.  The tables must be 16-byte aligned.
void launch_exscan2_u32(const uint32_t* in_a, uint64_t* out_a, unsigned long long* total_a, const uint32_t* in_b,
                        uint64_t* out_b, unsigned long long* total_b, uint64_t n, unsigned long long* desc,
                        unsigned int* tickets, uint32_t epoch, cudaStream_t s, unsigned long long* host_total_a = nullptr);

// *flag = 1 if two consecutive entries of the sorted occurrence list are closer than m (overlapping occurrences)
void launch_overlap_check(const uint64_t* occ, uint64_t k, uint32_t m, unsigned int* flag, cudaStream_t s);
// selection flags of the non-zone part of every chunk
void launch_select(const ListArgs& a, const Pattern& pat, int mode, cudaStream_t s);
// per-chunk: greedy resolution (bordered) and emulation of the reference's chunk tail; sets sel[]
void launch_zone(const ListArgs& a, const Pattern& pat, int mode, cudaStream_t s);
// both; one launch where the zone replay does not depend on select's flags.  Returns the number of launches.
int launch_select_zone(const ListArgs& a, const Pattern& pat, int mode, cudaStream_t s);
// count-only variant without an occurrence list: adds to sub (true occurrences inside the tail zones)
// and add (occurrences the reference reports there)
void launch_zone_count(const uint8_t* buf, uint64_t n, const uint64_t* bounds, uint64_t nb, const Pattern& pat,
                       const uint32_t* gran_count, unsigned long long* sub, unsigned long long* add, cudaStream_t s);

struct CompactArgs {
  ListArgs l;
  uint64_t goff;               // global offset base
  uint64_t lidx_base;          // line index base
  const uint64_t* nl_prefix;   // [granules] exclusive '\n' prefix (line indices) or nullptr
  const uint32_t* nl_halves;   // [granules] '\n' per half granule, 2 x u16 packed (ScanArgs::nl_granule)
  uint64_t* out_match;         // any of these may be nullptr
  uint64_t* out_line_off;
  uint64_t* out_line_idx;
  uint64_t* out_line_start;    // xs::lines: (start, len)
  uint64_t* out_line_len;
  uint64_t* selq;              // [k] scratch: compacted positions of the selected occurrences (line kinds)
  unsigned long long* desc;    // [div_up(k, 2048)] look-back descriptors (tagged with epoch, see lookback())
  uint32_t epoch;
  unsigned int* ticket;        // zeroed
  unsigned long long* total;   // number of selected entries
  unsigned int* unterminated;  // set when the last selected line has no '\n' (xs::lines drops it)
};
void launch_compact(const CompactArgs& a, const Pattern& pat, cudaStream_t s);

// v[i] += delta (multi-GPU rebase of offsets / line indices)
void launch_add_u64(uint64_t* v, uint64_t n, uint64_t delta, cudaStream_t s);
// v[i] += sum(table[r * stride] for r < rank), the table read on the device
void launch_add_u64_from_table(uint64_t* v, uint64_t n, const uint64_t* table, uint32_t stride, uint32_t rank,
                               cudaStream_t s);
// summary[0..4) = {newlines, matches, matching lines, lines} from the scan's device scalars
// mode 0: list path, 1: count-only path (occ_total - zone_sub + zone_add), 2: direct path (occ_total - drops),
// 3: direct path whose tail replay ran list-free on the side stream (zone_sub / zone_add; a difference means redo)
// pub (may be nullptr): the kernel also stores the summary into every rank's exchange table and releases the round
// (xsg_exchange.cu) -- unless a fall-back pass will follow (overlap seen / tail drops), whose summary publishes instead
// host_mirror (may be nullptr): DevScalars in mapped host memory that receives a copy of the scalars
void launch_summary(const void* dev_scalars, int mode, int want_lines, unsigned long long* summary, const ExPublish* pub,
                    cudaStream_t s, void* host_mirror = nullptr);

}  // namespace xsg
