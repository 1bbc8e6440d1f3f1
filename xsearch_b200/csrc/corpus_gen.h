// corpus_gen.h -- deterministic synthetic English-like corpus, one pure function per 4 KiB block.
//
// The same code runs on the host (xsg_synth_host: feeds the CPU oracle / reference) and on the
// device (xsg_corpus_fill_synth: fills HBM directly, so a 400 GB corpus never crosses PCIe).
// Shape follows SURVEY.md 8(d): words from a fixed 256-word vocabulary with a Zipf-like skew,
// single spaces, lines of 1..12 words (~30 B/line like the reference's sample),
// every block ends with '\n', and with probability rate/2^32 per line one word is replaced by
// the needle.
#pragma once
#include <stdint.h>

#include "../../include/xsg.h"
#include "vocab.inc"

#if defined(__CUDACC__)
#define XS_HD __host__ __device__ __forceinline__
#else
#define XS_HD inline
#endif

struct xs_vocab_t {
  char chars[XS_VOCAB_WORDS * XS_VOCAB_SLOT + 1];
  uint8_t lens[XS_VOCAB_WORDS];
};

#define XS_VOCAB_INIT {XS_VOCAB_CHARS, {XS_VOCAB_LENS}}

// the splitmix64 finaliser (a bijection of 64-bit words)
XS_HD uint64_t xs_mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// One independent stream per block: start AND increment are hashes of (seed, block), so no two blocks walk the
// same sequence.  (Round 1 seeded block b with seed * C + b * phi and stepped by the same phi: draw j of block b
// was draw j - 1 of block b + 1, every block its neighbour shifted by one word.)
struct xs_rng {
  uint64_t s, inc;
};
XS_HD xs_rng xs_rng_for_block(uint64_t seed, uint64_t blk) {
  xs_rng g;
  const uint64_t h = xs_mix64(seed * 0xD1342543DE82EF95ull + 0x1234567ull) ^ xs_mix64(blk + 0x9E3779B97F4A7C15ull);
  g.s = xs_mix64(h);
  g.inc = xs_mix64(h ^ 0xDA942042E4DD58B5ull) | 1ull;
  return g;
}
XS_HD uint64_t xs_next(xs_rng& g) {
  g.s += g.inc;
  return xs_mix64(g.s);
}

// Writes exactly XSG_SYNTH_BLOCK bytes of block `blk` to out (any byte-addressable memory).
XS_HD void xs_synth_block(uint8_t* out, uint64_t blk, const xsg_synth_params& p, const xs_vocab_t& voc) {
  xs_rng s = xs_rng_for_block(p.seed, blk);
  uint32_t pos = 0;
  uint32_t words_in_line = 0, line_target = 1, inject_word = 0xffffffffu;
  const uint32_t reserve = 16 + (p.needle_len > 12 ? p.needle_len : 12);
  while (XSG_SYNTH_BLOCK - pos >= reserve) {
    uint64_t r = xs_next(s);
    if (words_in_line == 0) {
      line_target = 1 + (uint32_t)(r & 0xffff) % 12;
      bool inject = p.needle_len > 0 && (uint32_t)(r >> 32) < p.needle_rate_u32;
      inject_word = inject ? (uint32_t)((r >> 16) & 0xffff) % line_target : 0xffffffffu;
      r = xs_next(s);
    }
    if (words_in_line == inject_word) {
      for (uint32_t k = 0; k < p.needle_len; ++k) out[pos++] = p.needle[k];
    } else {
      uint32_t idx = ((uint32_t)(r & 0xff) * (uint32_t)((r >> 8) & 0xff)) >> 8;
      uint32_t len = voc.lens[idx];
      for (uint32_t k = 0; k < len; ++k) out[pos++] = (uint8_t)voc.chars[idx * XS_VOCAB_SLOT + k];
    }
    ++words_in_line;
    if (words_in_line == line_target) {
      out[pos++] = '\n';
      words_in_line = 0;
    } else {
      out[pos++] = ' ';
    }
  }
  // pad the rest of the block as one filler line so the block ends with '\n'
  while (pos < XSG_SYNTH_BLOCK - 1) out[pos++] = '.';
  out[pos] = '\n';
}
