// corpus_host.cpp -- the synthetic corpus on the host cores WITHOUT the CUDA library.
//
// TEST / BENCH INFRASTRUCTURE ONLY (like everything under oracle/).  bench.py --impl reference must time the
// reference's CPU code in a process that maps nothing of the product (task 7): this file compiles the
// block generator of xsearch_b200/csrc/corpus_gen.h -- a pure function of (seed, block), shared with the device fill
// so that both arms see the same bytes -- into oracle/libxscorpus.so.  No matcher code lives here.
#include <cstdint>
#include <thread>
#include <vector>

#include "../xsearch_b200/csrc/corpus_gen.h"

extern "C" __attribute__((visibility("default")))
void xsc_synth(void* dst, uint64_t first_block, uint64_t n_blocks, const xsg_synth_params* p, int nthreads) {
  static const xs_vocab_t voc = XS_VOCAB_INIT;
  uint8_t* out = static_cast<uint8_t*>(dst);
  if (nthreads < 1) nthreads = 1;
  const uint64_t per = (n_blocks + (uint64_t)nthreads - 1) / (uint64_t)nthreads;
  auto work = [&](uint64_t b0, uint64_t b1) {
    for (uint64_t b = b0; b < b1; ++b) xs_synth_block(out + b * XSG_SYNTH_BLOCK, first_block + b, *p, voc);
  };
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) {
    const uint64_t b0 = (uint64_t)t * per, b1 = b0 + per < n_blocks ? b0 + per : n_blocks;
    if (b0 < b1) th.emplace_back(work, b0, b1);
  }
  for (auto& t : th) t.join();
}

// '\n' in buf[0..n): the independent newline count bench.py's parity block compares line-index bases with
extern "C" __attribute__((visibility("default")))
uint64_t xsc_count_newlines(const void* buf, uint64_t n, int nthreads) {
  const uint8_t* b = static_cast<const uint8_t*>(buf);
  if (nthreads < 1) nthreads = 1;
  std::vector<uint64_t> part((size_t)nthreads, 0);
  const uint64_t per = (n + (uint64_t)nthreads - 1) / (uint64_t)nthreads;
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t)
    th.emplace_back([&, t] {
      const uint64_t lo = (uint64_t)t * per, hi = lo + per < n ? lo + per : n;
      uint64_t c = 0;
      for (uint64_t i = lo; i < hi; ++i) c += b[i] == '\n';
      part[(size_t)t] = c;
    });
  for (auto& t : th) t.join();
  uint64_t total = 0;
  for (uint64_t c : part) total += c;
  return total;
}
