// xsearch.hpp -- the reference's one-call API on top of the B200 C ABI (C++20, header-only).
//
//   auto s = xs::extern_search<xs::lines>(pattern, file, meta_file, num_threads, max_num_readers);
//   for (auto const& line : *s->getResult()) ...      // live: blocks until more results or closed
//   s->join();  s->getResult()->size();  s->getResult()->copyResultSafe();
//
// Shapes kept from the reference (file:line in /root/reference):
//   * extern_search<Tag>(pattern, file, meta_file, num_threads, max_num_readers)     README.md:72, checkit.cpp:4
//     and the overloads (pattern, file, ignore_case, num_threads)                    README.md:37, example/grep.cpp:69-77
//     / (pattern, file, meta_file, ignore_case, num_threads, max_num_readers)        test/src/xsearchTest.cpp:1227-1228
//   * tags count (alias count_matches), count_lines, match_byte_offsets, line_byte_offsets,
//     line_indices, lines                                                             README.md:72-81
//   * Searcher: work starts in the constructor, join() idempotent, destructor joins,
//     non-copyable / non-movable                                                      include/xsearch/Searcher.h:41-59
//   * Result: add / close / size / begin / end with an iterator whose operator!= blocks until more data
//     or close                                                                        include/xsearch/ResultTypes.h:31-130
//     count tags yield running totals, the last one is the total                      test/src/xsearchTest.cpp:735-739
//
// What runs underneath is NOT the reference's thread pool: max_num_readers reader threads pread the file
// into pinned segments (all of them 8 MiB slices of the same segment, so segments complete in file order at the
// speed of all readers together), one pusher thread hands them to libxsg.so (xsg_stream_push: H2D copy overlapped
// with the scan of the previous segment on the GPU) and publishes the result parts in file order.
// There is no CPU search path: without a CUDA device extern_search throws.
//
// Regular expressions (the reference sends a pattern to re2 when utils.h:17-24 says it is one): the GPU scans for
// the literals every match must contain, the host runs the expression over the candidate lines only
// (regex_prefilter.hpp).  Options::pattern_mode forces either reading of the pattern.
#pragma once

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#include "../xsg.h"
#include "metafile.hpp"
#include "regex_prefilter.hpp"

namespace xs {

// ---- tags ------------------------------------------------------------------------------------------
struct count {
  using value_type = uint64_t;
  static constexpr uint32_t kind = XSG_COUNT;
  static constexpr bool is_count = true;
};
using count_matches = count;  // spelling used by the reference's tests (xsearchTest.cpp:344)
struct count_lines {
  using value_type = uint64_t;
  static constexpr uint32_t kind = XSG_COUNT_LINES;
  static constexpr bool is_count = true;
};
struct match_byte_offsets {
  using value_type = uint64_t;
  static constexpr uint32_t kind = XSG_MATCH_BYTE_OFFSETS;
  static constexpr bool is_count = false;
};
struct line_byte_offsets {
  using value_type = uint64_t;
  static constexpr uint32_t kind = XSG_LINE_BYTE_OFFSETS;
  static constexpr bool is_count = false;
};
struct line_indices {
  using value_type = uint64_t;
  static constexpr uint32_t kind = XSG_LINE_INDICES;
  static constexpr bool is_count = false;
};
struct lines {
  using value_type = std::string;
  static constexpr uint32_t kind = XSG_LINES;
  static constexpr bool is_count = false;
};

// knobs the reference does not have (device selection, chunking); defaults reproduce v1 chunking
struct Options {
  int device = 0;
  // more than one entry: ONE Searcher over several devices of this host (the reference's num_threads workers over the
  // chunks of one file, Searcher.h:141-145): segments are dealt round-robin to the devices, results are published in
  // file order, line indices get their base on the host (the '\n' count of the earlier segments).  Literal patterns
  // only; a regular expression runs on devices[0].  Empty: {device}.
  std::vector<int> devices;
  uint64_t min_chunk = XSG_DEFAULT_MIN_CHUNK;   // v1: chunks of >= 16 MiB, extended through the next '\n'
  // bytes read / copied / scanned per pipeline step.  64 MiB: the first H2D copy starts after 2 ms of reading and
  // the pinned buffers are small (81 MB each); measured on an 8 GiB file, 16 readers: 41.3 GB/s against 39.8 (128 MiB),
  // 37.1 (256 MiB), 31 (512 MiB) -- repo:profiles/r2_file_segment_sweep.txt
  uint64_t segment_bytes = 64ull << 20;
  bool ignore_case = false;                     // set by the ignore_case overloads of extern_search
  // AUTO = the reference's rule (xs::use_str_as_regex, utils.h:17-24); LITERAL / REGEX force one reading
  enum class PatternMode { AUTO, LITERAL, REGEX };
  PatternMode pattern_mode = PatternMode::AUTO;
  // how the reader threads get the bytes into the pinned segments: PREAD (default; a shrinking file is a reported
  // read error) or MMAP (the file is mapped once and the readers memcpy out of the mapping: no system call per
  // slice, 1.2-1.5x faster per thread out of a tmpfs page cache; a file truncated WHILE it is searched
  // raises SIGBUS, as with any mapping)
  enum class ReaderMode { PREAD, MMAP };
  ReaderMode reader_mode = ReaderMode::PREAD;
};

// ---- result collector ------------------------------------------------------------------------------
namespace detail {

// Allocator of the result parts.  (1) vector<T, PartAllocator<T>>(n) default-initialises: no zero-fill pass over a
// buffer that the D2H copy overwrites anyway (what the reference's UninitializedAllocator.h is for on its chunk
// buffers).  (2) Large parts are 2 MiB-aligned and advised MADV_HUGEPAGE: a dense result is gigabytes of fresh
// memory, and its first touch is what the host then spends its time on (1 GiB: 0.79 s with 4 KiB faults, 0.34 s
// with huge pages).
template <class T>
struct PartAllocator {
  using value_type = T;
  static constexpr size_t kHuge = size_t{2} << 20;
  PartAllocator() = default;
  template <class U>
  PartAllocator(const PartAllocator<U>&) noexcept {}
  T* allocate(size_t n) {
    if (n > static_cast<size_t>(-1) / sizeof(T) - kHuge) throw std::bad_alloc();
    const size_t bytes = n * sizeof(T);
    void* p;
    if (bytes >= 4 * kHuge) {
      const size_t rounded = (bytes + kHuge - 1) / kHuge * kHuge;
      p = std::aligned_alloc(kHuge, rounded);
      if (p) ::madvise(p, rounded, MADV_HUGEPAGE);  // advisory: failure leaves ordinary pages
    } else {
      p = std::malloc(bytes ? bytes : 1);
    }
    if (!p) throw std::bad_alloc();
    return static_cast<T*>(p);
  }
  void deallocate(T* p, size_t) noexcept { std::free(p); }
  template <class U>
  void construct(U* p) noexcept(std::is_nothrow_default_constructible_v<U>) {
    ::new (static_cast<void*>(p)) U;
  }
  template <class U, class A0, class... A>
  void construct(U* p, A0&& a0, A&&... a) {
    ::new (static_cast<void*>(p)) U(std::forward<A0>(a0), std::forward<A>(a)...);
  }
  template <class U>
  bool operator==(const PartAllocator<U>&) const noexcept {
    return true;
  }
  template <class U>
  bool operator!=(const PartAllocator<U>&) const noexcept {
    return false;
  }
};

}  // namespace detail

// Ordered collection of result elements with the reference's blocking-iterator behaviour.
// For count tags the elements are running totals and size() is the total.
// Elements are kept as the PARTS they arrive in (one per scanned segment, moved in, never touched again): adding a
// part is O(1) whatever its size, and the iterator takes the lock once per part, not per element -- a dense search
// delivers 10^8 offsets per 10 GB, and a lock round trip per element would cost seconds where the scan costs
// milliseconds.
template <class T>
class Result {
 public:
  using part_type = std::vector<T, detail::PartAllocator<T>>;  // what the producer hands over, one per segment

  class iterator {
   public:
    explicit iterator(Result* r) : _r(r) {}
    iterator& operator++() {
      ++_k;
      return *this;
    }
    // valid after operator!= returned true (what a range-for does); blocks like operator!= otherwise
    const T& operator*() const {
      if (_k >= _n) fetch();
      return _cur[_k];
    }
    // blocks until the next element exists or the result is closed (ResultTypes.h:48-60)
    bool operator!=(const iterator&) const { return _k < _n || fetch(); }
    bool operator==(const iterator& o) const { return !(*this != o); }

   private:
    bool fetch() const {  // next part (parts are never empty); false: closed and nothing left
      std::unique_lock lock(_r->_m);
      _r->_cv.wait(lock, [&] { return _next < _r->_parts.size() || _r->_closed; });
      if (_next >= _r->_parts.size()) return false;
      const part_type& p = _r->_parts[_next++];
      _cur = p.data();
      _n = p.size();
      _k = 0;
      return true;
    }
    Result* _r;
    mutable const T* _cur = nullptr;  // the part being walked: immutable once added, read without the lock
    mutable size_t _n = 0, _k = 0, _next = 0;
  };

  explicit Result(bool counting) : _counting(counting) {}
  Result(const Result&) = delete;
  Result& operator=(const Result&) = delete;

  // producer side ----------------------------------------------------------------------------------
  bool add(part_type&& part) {
    std::unique_lock lock(_m);
    if (_closed) return false;
    if (part.empty()) return true;
    _starts.push_back(_count);
    _count += part.size();
    _parts.push_back(std::move(part));
    _cv.notify_all();
    return true;
  }
  void close() {
    {
      std::unique_lock lock(_m);
      _closed = true;
    }
    _cv.notify_all();
  }

  // consumer side ----------------------------------------------------------------------------------
  // count tags: the count so far (the total once closed); vector tags: number of elements so far
  size_t size() const {
    std::unique_lock lock(_m);
    if constexpr (std::is_same_v<T, uint64_t>) {
      if (_counting) return _parts.empty() ? 0 : static_cast<size_t>(_parts.back().back());
    }
    return _count;
  }
  bool empty() const { return size() == 0; }
  bool is_closed() const {
    std::unique_lock lock(_m);
    return _closed;
  }
  void wait_closed() const {
    std::unique_lock lock(_m);
    _cv.wait(lock, [&] { return _closed; });
  }
  std::vector<T> copyResultSafe() const {
    std::unique_lock lock(_m);
    std::vector<T> out;
    out.reserve(_count);
    for (const part_type& p : _parts) out.insert(out.end(), p.begin(), p.end());
    return out;
  }
  T at(size_t i) const {
    std::unique_lock lock(_m);
    if (i >= _count) throw std::out_of_range("xs::Result::at");
    const size_t part = static_cast<size_t>(std::upper_bound(_starts.begin(), _starts.end(), i) - _starts.begin()) - 1;
    return _parts[part][i - _starts[part]];
  }
  T operator[](size_t i) const { return at(i); }
  iterator begin() { return iterator(this); }
  iterator end() { return iterator(this); }

 private:
  const bool _counting;
  mutable std::mutex _m;
  mutable std::condition_variable _cv;
  std::deque<part_type> _parts;  // deque: a part handed to an iterator stays where it is when later parts arrive
  std::vector<size_t> _starts;        // index of the first element of every part
  size_t _count = 0;
  bool _closed = false;
};

// ---- searcher ----------------------------------------------------------------------------------------
namespace detail {

inline std::string last_error(const char* what) { return std::string(what) + ": " + xsg_last_error(); }

struct Pinned {  // one pinned host buffer (cudaHostAlloc through the C ABI)
  uint8_t* p = nullptr;
  size_t cap = 0;
  Pinned() = default;
  Pinned(const Pinned&) = delete;
  Pinned& operator=(const Pinned&) = delete;
  ~Pinned() { release(); }
  bool alloc(size_t n) {
    release();
    void* q = nullptr;
    if (xsg_host_alloc(n, &q) != XSG_OK) return false;
    p = static_cast<uint8_t*>(q);
    cap = n;
    return true;
  }
  void release() {
    if (p) xsg_host_free(p);
    p = nullptr;
    cap = 0;
  }
};

struct Block {  // one unit of reading: a byte range of the file
  uint64_t off = 0, len = 0;
  std::vector<uint64_t> rel_bounds;  // meta-file mode: chunk ends relative to off
};

}  // namespace detail

template <class Tag>
class Searcher {
 public:
  using value_type = typename Tag::value_type;
  using ResultT = Result<value_type>;

  Searcher(std::string pattern, const std::string& file_path, const std::string& meta_file_path, int num_threads,
           int max_num_readers, Options opt = {})
      : _pattern(std::move(pattern)), _opt(opt), _result(Tag::is_count) {
    _regex_mode = _opt.pattern_mode == Options::PatternMode::REGEX ||
                  (_opt.pattern_mode == Options::PatternMode::AUTO && use_str_as_regex(_pattern));
    std::vector<std::string> literals{_pattern};
    if (_regex_mode) {
      if (_pattern.find('\n') != std::string::npos)
        throw std::invalid_argument("xs::extern_search: a pattern must not contain '\\n'");
      literals = required_literals(_pattern);  // throws std::invalid_argument when there is nothing to scan for
      _verifier = std::make_unique<LineVerifier>(_pattern, _opt.ignore_case);  // std::invalid_argument: unsupported syntax
    } else if (_pattern.empty() || _pattern.size() > XSG_MAX_PATTERN || _pattern.find('\n') != std::string::npos) {
      throw std::invalid_argument("xs::extern_search: the GPU literal path needs a pattern of 1..64 bytes without '\\n'");
    }
    _num_threads = num_threads < 1 ? 1 : num_threads;  // accepted for API compatibility; the search itself runs on the GPU
    _num_readers = max_num_readers < 1 ? 1 : max_num_readers;
    if (_opt.min_chunk == 0) _opt.min_chunk = XSG_DEFAULT_MIN_CHUNK;
    if (_opt.segment_bytes < 2 * _opt.min_chunk) _opt.segment_bytes = 2 * _opt.min_chunk;
    _fd = ::open(file_path.c_str(), O_RDONLY);
    if (_fd < 0) throw std::runtime_error("xs::extern_search: cannot open '" + file_path + "'");
    struct stat sb;
    if (::fstat(_fd, &sb) != 0 || !S_ISREG(sb.st_mode)) {
      ::close(_fd);
      throw std::runtime_error("xs::extern_search: '" + file_path + "' is not a regular file");
    }
    _size = static_cast<uint64_t>(sb.st_size);
    try {
      if (_opt.reader_mode == Options::ReaderMode::MMAP && _size > 0) {
        void* m = ::mmap(nullptr, static_cast<size_t>(_size), PROT_READ, MAP_SHARED, _fd, 0);
        if (m == MAP_FAILED) throw std::runtime_error("xs::extern_search: cannot map '" + file_path + "'");
        _map = static_cast<const uint8_t*>(m);
      }
      plan(meta_file_path);
      xsg_scan_opts so{};
      so.flags = _opt.ignore_case ? XSG_SCAN_IGNORE_CASE : 0u;  // ASCII fold, like v1's tolower processor
      // literal: one lane per device, each with one stream of the asked kind.  regex: one lane (devices[0]) with one
      // stream per required literal (each with a context of its own: the staging buffers belong to the context),
      // reporting candidate lines with their indices
      std::vector<int> devs = _opt.devices.empty() ? std::vector<int>{_opt.device} : _opt.devices;
      if (_regex_mode) devs.resize(1);
      const uint32_t kinds = _regex_mode ? (XSG_LINE_BYTE_OFFSETS | XSG_LINE_INDICES) : Tag::kind;
      if (devs.size() > 1) so.flags |= XSG_SCAN_LOCAL_LINE_INDEX;  // bases are added in harvest(), in file order
      _lanes.resize(devs.size());
      if (devs.size() > 1) {
        std::vector<xsg_ctx*> cs(devs.size(), nullptr);
        if (xsg_init_multi(devs.data(), static_cast<int>(devs.size()), cs.data()) != XSG_OK)
          throw std::runtime_error(detail::last_error("xs::extern_search: xsg_init_multi"));
        for (size_t d = 0; d < devs.size(); ++d) _lanes[d].ctxs.push_back(cs[d]);
      }
      for (size_t d = 0; d < devs.size(); ++d) {
        for (size_t k = 0; k < literals.size(); ++k) {
          const std::string& lit = literals[k];
          if (_lanes[d].ctxs.size() <= k) {
            xsg_ctx* ctx = nullptr;
            if (xsg_init(devs[d], &ctx) != XSG_OK) throw std::runtime_error(detail::last_error("xs::extern_search: xsg_init"));
            _lanes[d].ctxs.push_back(ctx);
          }
          xsg_stream* st = nullptr;
          if (xsg_stream_open(_lanes[d].ctxs[k], reinterpret_cast<const uint8_t*>(lit.data()), static_cast<uint32_t>(lit.size()),
                              kinds, &so, &st) != XSG_OK)
            throw std::runtime_error(detail::last_error("xs::extern_search: xsg_stream_open"));
          _lanes[d].streams.push_back(st);
        }
      }
    } catch (...) {
      destroy_lanes();
      if (_map) ::munmap(const_cast<uint8_t*>(_map), static_cast<size_t>(_size));
      ::close(_fd);
      throw;
    }
    _running.store(true);
    _pusher = std::thread(&Searcher::run_pusher, this);
  }

  ~Searcher() {
    join();
    _bufs.clear();
    destroy_lanes();
    if (_map) ::munmap(const_cast<uint8_t*>(_map), static_cast<size_t>(_size));
    if (_fd >= 0) ::close(_fd);
  }

  Searcher(Searcher&&) = delete;
  Searcher(const Searcher&) = delete;
  Searcher& operator=(Searcher&&) = delete;
  Searcher& operator=(const Searcher&) = delete;

  // block until the search has finished (idempotent; Searcher.h:52-59)
  void join() {
    std::unique_lock lock(_join_m);
    if (_pusher.joinable()) _pusher.join();
    _running.store(false);
  }
  [[nodiscard]] bool running() const { return _running.load(); }
  [[nodiscard]] size_t num_devices() const { return _lanes.size(); }
  // host time spent page-locking the segment buffers (part of the pipeline time; valid after join())
  [[nodiscard]] double pinned_alloc_ms() const { return _pinned_alloc_ms; }
  ResultT* getResult() { return &_result; }
  // nothing throws after extern_search returned: a failure closes the result early and is reported here
  [[nodiscard]] bool ok() const { return !_failed.load(); }
  std::string error() const {
    std::unique_lock lock(_m);
    return _error;
  }
  // device time of all scans / wall time of the streamed pipeline (valid after join())
  xsg_timings timings() const { return _timings; }
  uint64_t bytes() const { return _size; }

 private:
  // one device: its context(s) and stream(s) (literal: one of each; regex: one per required literal)
  struct Lane {
    std::vector<xsg_ctx*> ctxs;
    std::vector<xsg_stream*> streams;
    uint64_t parts_done = 0;  // parts of this lane's result that have been harvested
  };

  void destroy_lanes() {
    for (Lane& l : _lanes) {
      for (xsg_stream* st : l.streams)
        if (st) xsg_stream_abort(st);
      for (xsg_ctx* ctx : l.ctxs)
        if (ctx) xsg_destroy(ctx);
    }
    _lanes.clear();
  }

  // ---- planning: which byte ranges are read as one block --------------------------------------------
  void plan(const std::string& meta_file_path) {
    if (!meta_file_path.empty()) {
      MetaFile mf(meta_file_path);
      if (mf.compression_type() != CompressionType::NONE)
        throw std::runtime_error("xs::extern_search: compressed inputs are outside the GPU path (decompress first)");
      if (mf.original_size() != _size) throw std::runtime_error("xs::extern_search: the meta file does not describe this file");
      _meta_mode = true;
      const std::vector<uint64_t> b = mf.chunk_bounds();
      for (size_t c0 = 0; c0 < b.size();) {
        const uint64_t b0 = c0 ? b[c0 - 1] : 0;
        size_t c1 = c0 + 1;
        while (c1 < b.size() && b[c1] - b0 <= _opt.segment_bytes) ++c1;
        detail::Block blk;
        blk.off = b0;
        blk.len = b[c1 - 1] - b0;
        for (size_t c = c0; c < c1; ++c) blk.rel_bounds.push_back(b[c] - b0);
        _max_block = std::max<uint64_t>(_max_block, blk.len);
        _blocks.push_back(std::move(blk));
        c0 = c1;
      }
      _headroom = 0;
    } else {
      for (uint64_t off = 0; off < _size; off += _opt.segment_bytes) {
        detail::Block blk;
        blk.off = off;
        blk.len = std::min<uint64_t>(_opt.segment_bytes, _size - off);
        _max_block = std::max<uint64_t>(_max_block, blk.len);
        _blocks.push_back(std::move(blk));
      }
      // the unfinished last chunk of a block (< min_chunk + one line) is carried into the next buffer
      _headroom = _opt.min_chunk + (1u << 20);
    }
  }

  void fail(const std::string& msg) {
    {
      std::unique_lock lock(_m);
      if (_error.empty()) _error = msg;
    }
    _failed.store(true);
    _cv.notify_all();
  }

  // ---- reader threads: ALL of them pread slices of the SAME block (then the next one) ------------------------
  // One pread is a kernel memcpy out of the page cache, a few GB/s per thread.  With a block per reader the
  // first block -- and so the first H2D copy -- is ready only after one thread has copied all of it, and blocks
  // complete out of order; with slices the blocks complete in order at the speed of all readers together, and
  // the number of pinned buffers does not grow with the number of readers.
  static constexpr uint64_t kSliceBytes = 8ull << 20;

  void run_reader() {
    for (;;) {
      int buf;
      size_t i;
      uint64_t lo, hi;
      {
        std::unique_lock lock(_m);
        _cv.wait(lock, [&] { return _cur_buf >= 0 || !_free.empty() || _failed.load() || _next_block >= _blocks.size(); });
        if (_failed.load() || _next_block >= _blocks.size()) return;
        if (_cur_buf < 0) {  // open the next block: it gets its buffer before any later block can
          _cur_buf = _free.back();
          _free.pop_back();
          _next_slice = 0;
          _slices_left[_next_block] = (_blocks[_next_block].len + kSliceBytes - 1) / kSliceBytes;
        }
        i = _next_block;
        buf = _cur_buf;
        const detail::Block& b = _blocks[i];
        lo = _next_slice * kSliceBytes;
        hi = std::min<uint64_t>(b.len, lo + kSliceBytes);
        if (hi >= b.len) {  // that was its last slice: the next reader opens the next block
          ++_next_block;
          _cur_buf = -1;
        } else {
          ++_next_slice;
        }
      }
      const detail::Block& b = _blocks[i];
      uint8_t* dst = _bufs[buf]->p + _headroom;
      uint64_t done = lo;
      if (_map) {  // ReaderMode::MMAP: a copy out of the mapping, page faults instead of system calls
        std::memcpy(dst + lo, _map + b.off + lo, static_cast<size_t>(hi - lo));
        done = hi;
      }
      while (done < hi) {
        const ssize_t r = ::pread(_fd, dst + done, static_cast<size_t>(hi - done), static_cast<off_t>(b.off + done));
        if (r <= 0) {
          fail("xs::extern_search: read error (file truncated while searching?)");
          return;
        }
        done += static_cast<uint64_t>(r);
      }
      bool complete;
      {
        std::unique_lock lock(_m);
        complete = --_slices_left[i] == 0;
        if (complete) _ready[i] = buf;
      }
      _cv.notify_all();
    }
  }

  void give_back(int buf) {
    if (buf < 0) return;
    {
      std::unique_lock lock(_m);
      _free.push_back(buf);
    }
    _cv.notify_all();
  }

  // a pushed segment whose results have not been collected yet
  struct Inflight {
    bool valid = false;
    int buf = -1;                           // pool buffer holding its bytes
    std::unique_ptr<detail::Pinned> extra;  // or a one-off buffer (a carry longer than the headroom)
    const uint8_t* ptr = nullptr;           // first byte of the segment on the host
    uint64_t goff = 0;                      // its global file offset
    uint64_t len = 0;                       // bytes of it that were pushed (whole chunks)
    size_t lane = 0;                        // the device lane it was pushed to
  };

  // ---- results of one finished part -> the result collector ------------------------------------------
  bool harvest(const xsg_result* res, uint64_t part, const Inflight& seg) {
    const uint64_t n = xsg_result_part_count(res, part, Tag::kind);
    if constexpr (Tag::is_count) {
      _running_total += n;
      _result.add(typename ResultT::part_type{_running_total});
      _published = true;
    } else if constexpr (std::is_same_v<value_type, uint64_t>) {
      typename ResultT::part_type v(n);  // default-initialised: the copy below writes every element
      if (n && xsg_result_part_copy(res, part, Tag::kind, v.data(), n) != XSG_OK) return false;
      if constexpr (Tag::kind == XSG_LINE_INDICES) {
        // several devices: every part numbered its lines from its own first byte; parts are harvested in file
        // order, so the base -- the '\n' count of everything before this segment -- is known here
        if (_lanes.size() > 1) {
          for (uint64_t& x : v) x += _newlines_seen;
          _newlines_seen += xsg_result_part_newlines(res, part);
        }
      }
      _result.add(std::move(v));
    } else {
      std::vector<uint64_t> st(n), ln(n);
      if (n && (xsg_result_part_copy(res, part, XSG_LINES, st.data(), n) != XSG_OK ||
                xsg_result_part_copy_line_lengths(res, part, ln.data(), n) != XSG_OK))
        return false;
      // one std::string per matching line (as in the reference, search_wrappers.h:187-207) is one allocation per
      // line: the work num_threads is for on this path
      typename ResultT::part_type out(n);
      const char* base = reinterpret_cast<const char*>(seg.ptr);
      auto build = [&](uint64_t lo, uint64_t hi) {
        for (uint64_t i = lo; i < hi; ++i) out[i].assign(base + (st[i] - seg.goff), static_cast<size_t>(ln[i]));
      };
      const uint64_t nt = std::min<uint64_t>(static_cast<uint64_t>(_num_threads), n / 16384 + 1);
      if (nt <= 1) {
        build(0, n);
      } else {
        std::vector<std::thread> pool;
        for (uint64_t t = 0; t < nt; ++t) pool.emplace_back(build, n * t / nt, n * (t + 1) / nt);
        for (auto& t : pool) t.join();
      }
      _result.add(std::move(out));
    }
    return true;
  }

  // ---- regex mode: candidate lines of one finished part -> verified results --------------------------------
  // Candidates are the lines the literal streams report, plus -- unconditionally -- the lines that reach into the
  // last kZoneBytes of the segment: the literal scan reproduces the reference matcher's chunk-tail behaviour
  // (DESIGN.md 4.3), which may skip an occurrence among the last 31+m positions of a chunk, and a regular
  // expression (re2 in the reference) has no such tail.  Segments are pushed as ONE chunk in this mode, so that
  // zone exists once per segment.
  static constexpr uint64_t kZoneBytes = 32 + 2 * XSG_MAX_PATTERN;

  struct Candidate {
    uint64_t start, index;
  };

  bool harvest_regex(const std::vector<const xsg_result*>& rs, uint64_t part, const Inflight& seg) {
    std::vector<Candidate> cand;
    std::vector<size_t> runs;  // cand is a concatenation of sorted runs (one per literal stream + the tail zone)
    std::vector<uint64_t> off, idx;
    for (const xsg_result* r : rs) {
      runs.push_back(cand.size());
      const uint64_t n = xsg_result_part_count(r, part, XSG_LINE_BYTE_OFFSETS);
      off.resize(n);
      idx.resize(n);
      if (n && (xsg_result_part_copy(r, part, XSG_LINE_BYTE_OFFSETS, off.data(), n) != XSG_OK ||
                xsg_result_part_copy(r, part, XSG_LINE_INDICES, idx.data(), n) != XSG_OK))
        return false;
      for (uint64_t i = 0; i < n; ++i) cand.push_back({off[i], idx[i]});
    }
    const uint8_t* base = seg.ptr;
    const uint64_t end = seg.len;
    _newlines_seen += xsg_result_part_newlines(rs[0], part);  // '\n' in the file up to the end of this segment
    {
      uint64_t s0 = end > kZoneBytes ? end - kZoneBytes : 0;
      while (s0 > 0 && base[s0 - 1] != '\n') --s0;  // start of the line holding the first zone byte
      uint64_t index = _newlines_seen - static_cast<uint64_t>(std::count(base + s0, base + end, uint8_t('\n')));
      runs.push_back(cand.size());
      for (uint64_t pos = s0; pos < end; ++index) {
        cand.push_back({seg.goff + pos, index});
        const void* nl = std::memchr(base + pos, '\n', static_cast<size_t>(end - pos));
        if (!nl) break;
        pos = static_cast<uint64_t>(static_cast<const uint8_t*>(nl) - base) + 1;
      }
    }
    // merging the runs is linear in the candidates per run (10^7 lines per GiB for a frequent literal), a sort is not
    const auto by_start = [](const Candidate& a, const Candidate& b) { return a.start < b.start; };
    runs.push_back(cand.size());
    for (size_t k = 1; k + 1 < runs.size(); ++k)  // [0, runs[k]) is sorted; take in the run [runs[k], runs[k + 1])
      std::inplace_merge(cand.begin(), cand.begin() + static_cast<std::ptrdiff_t>(runs[k]),
                         cand.begin() + static_cast<std::ptrdiff_t>(runs[k + 1]), by_start);
    cand.erase(std::unique(cand.begin(), cand.end(), [](const Candidate& a, const Candidate& b) { return a.start == b.start; }),
               cand.end());

    // verification on up to num_threads host threads (what the reference's worker threads are left with)
    constexpr bool kMatchLevel = Tag::kind == XSG_COUNT || Tag::kind == XSG_MATCH_BYTE_OFFSETS;
    const size_t n = cand.size();
    std::vector<uint64_t> line_len(n);
    std::vector<uint8_t> keep(n, 0), terminated(n, 0);
    std::vector<std::vector<uint64_t>> pos(kMatchLevel ? n : 0);
    auto work = [&](size_t thread, size_t lo, size_t hi) {
      LineVerifier::Cache& cache = *_caches[thread];
      for (size_t i = lo; i < hi; ++i) {
        const uint64_t rel = cand[i].start - seg.goff;
        const void* nl = std::memchr(base + rel, '\n', static_cast<size_t>(end - rel));
        terminated[i] = nl != nullptr;
        line_len[i] = nl ? static_cast<uint64_t>(static_cast<const uint8_t*>(nl) - (base + rel)) : end - rel;
        const std::string_view line(reinterpret_cast<const char*>(base + rel), static_cast<size_t>(line_len[i]));
        if constexpr (kMatchLevel) {
          _verifier->match_positions(line, &pos[i]);
          keep[i] = !pos[i].empty();
        } else {
          keep[i] = LineVerifier::matches(line, cache);
        }
      }
    };
    const size_t nt = std::min<size_t>(static_cast<size_t>(_num_threads), n / 64 + 1);
    while (_caches.size() < nt) _caches.push_back(_verifier->make_cache());  // one lazy DFA per verification thread
    if (nt <= 1) {
      work(0, 0, n);
    } else {
      std::vector<std::thread> pool;
      for (size_t t = 0; t < nt; ++t) pool.emplace_back(work, t, n * t / nt, n * (t + 1) / nt);
      for (auto& t : pool) t.join();
    }

    if constexpr (Tag::is_count) {
      for (size_t i = 0; i < n; ++i) _running_total += kMatchLevel ? pos[i].size() : keep[i];
      _result.add(typename ResultT::part_type{_running_total});
      _published = true;
    } else if constexpr (std::is_same_v<value_type, uint64_t>) {
      typename ResultT::part_type v;
      for (size_t i = 0; i < n; ++i) {
        if (!keep[i]) continue;
        if constexpr (Tag::kind == XSG_MATCH_BYTE_OFFSETS) {
          for (uint64_t p : pos[i]) v.push_back(cand[i].start + p);
        } else if constexpr (Tag::kind == XSG_LINE_INDICES) {
          v.push_back(cand[i].index);
        } else {
          v.push_back(cand[i].start);
        }
      }
      _result.add(std::move(v));
    } else {
      typename ResultT::part_type out;
      for (size_t i = 0; i < n; ++i)  // like the literal xs::lines: an unterminated last line is not a line
        if (keep[i] && terminated[i])
          out.emplace_back(reinterpret_cast<const char*>(base + (cand[i].start - seg.goff)), static_cast<size_t>(line_len[i]));
      _result.add(std::move(out));
    }
    return true;
  }

  // results of every part of the segment's lane that has finished since the last call (literal: one stream; regex:
  // all streams of the lane)
  template <class GetResult>
  bool harvest_finished(GetResult&& result_of, const Inflight& seg) {
    Lane& lane = _lanes[seg.lane];
    std::vector<const xsg_result*> rs;
    for (size_t k = 0; k < lane.streams.size(); ++k) rs.push_back(result_of(k));
    while (lane.parts_done < xsg_result_num_parts(rs[0])) {
      const bool ok = _regex_mode ? harvest_regex(rs, lane.parts_done, seg) : harvest(rs[0], lane.parts_done, seg);
      if (!ok) return false;
      ++lane.parts_done;
    }
    return true;
  }

  void release(Inflight& s) {
    give_back(s.buf);
    s = Inflight();
  }

  // ---- pusher thread: forms segments of whole reference chunks and drives the GPU -------------------
  void run_pusher() {
    try {
      pusher_body();
    } catch (const std::exception& e) {
      fail(e.what());
    }
    // wake and collect the readers, then close the result (the reference closes when the last worker ends)
    {
      std::unique_lock lock(_m);
      _next_block = _blocks.size();
    }
    _cv.notify_all();
    for (auto& t : _readers)
      if (t.joinable()) t.join();
    _result.close();
  }

  void pusher_body() {
    // per device lane one segment being copied / scanned and one whose results are collected, plus the block(s) being
    // read (two around a block edge)
    const size_t nlanes = _lanes.size();
    const size_t nbuf = nlanes + 3;
    // The buffers are page-locked one after the other on a thread of its own: the readers start on the first one
    // while the others are still being pinned (page-locking costs as much per byte as reading from the page cache).
    const size_t want = _blocks.empty() ? 0 : std::min(nbuf, _blocks.size() + 1);
    _bufs.reserve(want);  // readers index _bufs without the lock: it must never reallocate
    std::thread allocator([this, want] {
      const auto t_alloc = std::chrono::steady_clock::now();
      for (size_t i = 0; i < want && !_failed.load(); ++i) {
        auto b = std::make_unique<detail::Pinned>();
        if (!b->alloc(static_cast<size_t>(_headroom + _max_block + 64))) {
          fail(detail::last_error("pinned buffer"));
          break;
        }
        {
          std::unique_lock lock(_m);
          _bufs.push_back(std::move(b));
          _free.push_back(static_cast<int>(i));
        }
        _cv.notify_all();
      }
      _pinned_alloc_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_alloc).count();
    });
    struct Joiner {  // whatever way this function is left: the allocator thread is collected first
      std::thread& t;
      ~Joiner() {
        if (t.joinable()) t.join();
      }
    } join_allocator{allocator};
    _ready.assign(_blocks.size(), -1);
    _slices_left.assign(_blocks.size(), 0);
    for (int t = 0; t < _num_readers; ++t) _readers.emplace_back(&Searcher::run_reader, this);

    std::deque<Inflight> pending;  // pushed, results not collected yet: file order, at most one per lane + the newest
    Inflight carry_holder;         // holds the bytes of `carry` when they belong to a segment that was not pushed
    const uint8_t* carry = nullptr;
    uint64_t carry_len = 0;
    uint64_t n_pushed = 0;
    std::vector<uint64_t> bounds;

    for (size_t i = 0; i < _blocks.size(); ++i) {
      int buf;
      {
        std::unique_lock lock(_m);
        _cv.wait(lock, [&] { return _ready[i] >= 0 || _failed.load(); });
        if (_failed.load()) throw std::runtime_error(_error);
        buf = _ready[i];
      }
      const detail::Block& b = _blocks[i];
      Inflight cur;
      cur.valid = true;
      cur.buf = buf;
      uint64_t seg_len;
      if (carry_len <= _headroom) {
        cur.ptr = _bufs[buf]->p + _headroom - carry_len;
        if (carry_len) std::memcpy(const_cast<uint8_t*>(cur.ptr), carry, carry_len);
      } else {  // a line (or unfinished chunk) longer than the headroom: one-off buffer
        cur.extra = std::make_unique<detail::Pinned>();
        if (!cur.extra->alloc(static_cast<size_t>(carry_len + b.len + 64))) throw std::runtime_error(detail::last_error("pinned buffer"));
        std::memcpy(cur.extra->p, carry, carry_len);
        std::memcpy(cur.extra->p + carry_len, _bufs[buf]->p + _headroom, b.len);
        cur.ptr = cur.extra->p;
        give_back(buf);
        cur.buf = -1;
      }
      seg_len = carry_len + b.len;
      cur.goff = b.off - carry_len;
      if (carry_holder.valid) release(carry_holder);  // its bytes have been copied

      // chunk bounds of this segment
      uint64_t consumed;
      const bool eof = i + 1 == _blocks.size();
      if (_meta_mode) {
        bounds = b.rel_bounds;
        consumed = seg_len;
      } else {
        bounds.clear();
        uint64_t pos = 0;
        while (seg_len - pos >= _opt.min_chunk) {  // v1 rule: >= min_chunk bytes, then through the next '\n'
          const uint64_t from = pos + _opt.min_chunk - 1;
          const void* nl = std::memchr(cur.ptr + from, '\n', static_cast<size_t>(seg_len - from));
          if (!nl) break;
          pos = static_cast<uint64_t>(static_cast<const uint8_t*>(nl) - cur.ptr) + 1;
          bounds.push_back(pos);
        }
        if (eof && pos < seg_len) {
          pos = seg_len;
          bounds.push_back(pos);
        }
        consumed = pos;
      }
      carry = cur.ptr + consumed;
      carry_len = seg_len - consumed;
      cur.len = consumed;
      if (bounds.empty()) {  // nothing complete yet: everything is carried on
        carry_holder = std::move(cur);
        continue;
      }
      if (_regex_mode) bounds.assign(1, consumed);  // re2 has no chunk-tail behaviour to reproduce (see harvest_regex)
      cur.lane = static_cast<size_t>(n_pushed++ % nlanes);
      Lane& lane = _lanes[cur.lane];
      for (xsg_stream* st : lane.streams)
        if (xsg_stream_push_at(st, cur.ptr, consumed, bounds.data(), bounds.size(), cur.goff, nullptr) != XSG_OK)
          throw std::runtime_error(detail::last_error("xsg_stream_push"));
      // this lane's previous segment was scanned inside the push: it is the oldest one pending (segments go to the
      // lanes in turn) -- publish its results, free its buffer
      pending.push_back(std::move(cur));
      if (pending.size() > nlanes) {
        Inflight& done = pending.front();
        if (!harvest_finished([&](size_t k) { return xsg_stream_result(_lanes[done.lane].streams[k]); }, done))
          throw std::runtime_error(detail::last_error("result copy"));
        release(done);
        pending.pop_front();
      }
    }
    // the last segment of every lane is scanned when its streams are closed: in file order
    std::string close_error;
    bool ok = true;
    while (!pending.empty()) {
      Inflight& seg = pending.front();
      Lane& lane = _lanes[seg.lane];
      std::vector<xsg_result*> fin(lane.streams.size(), nullptr);
      bool closed = true;
      for (size_t k = 0; k < lane.streams.size(); ++k) {
        xsg_stream* st = lane.streams[k];
        lane.streams[k] = nullptr;  // close destroys it whatever happens
        if (st) closed = xsg_stream_close(st, &fin[k]) == XSG_OK && closed;
        else closed = false;
      }
      if (!closed && close_error.empty()) close_error = detail::last_error("xsg_stream_close");
      if (closed && ok && close_error.empty())
        ok = harvest_finished([&](size_t k) { return static_cast<const xsg_result*>(fin[k]); }, seg);
      if (fin[0]) {
        xsg_timings t{};
        xsg_result_timings(fin[0], &t);
        _timings.scan_ms += t.scan_ms;
        _timings.main_kernel_ms += t.main_kernel_ms;
        _timings.post_ms += t.post_ms;
        _timings.kernel_launches += t.kernel_launches;
        _timings.h2d_ms = std::max(_timings.h2d_ms, t.h2d_ms);
      }
      for (xsg_result* r : fin)
        if (r) xsg_result_free(r);
      release(seg);
      pending.pop_front();
    }
    if (carry_holder.valid) release(carry_holder);
    if (!close_error.empty()) throw std::runtime_error(close_error);
    if (!ok) throw std::runtime_error(detail::last_error("result copy"));
    if constexpr (Tag::is_count) {
      if (!_published) _result.add(typename ResultT::part_type{0});  // empty input: the only running total is 0
    }
  }

  const std::string _pattern;
  Options _opt;
  int _num_threads = 1, _num_readers = 1;
  int _fd = -1;
  const uint8_t* _map = nullptr;  // ReaderMode::MMAP: the whole file
  uint64_t _size = 0;
  bool _meta_mode = false;
  std::vector<detail::Block> _blocks;
  uint64_t _max_block = 0, _headroom = 0;

  std::vector<Lane> _lanes;           // one per device (regex: one lane)
  xsg_timings _timings{};
  double _pinned_alloc_ms = 0;
  bool _regex_mode = false;
  std::unique_ptr<LineVerifier> _verifier;
  std::vector<std::unique_ptr<LineVerifier::Cache>> _caches;  // regex mode: per verification thread, kept across parts
  uint64_t _newlines_seen = 0;  // regex mode / several devices: '\n' in all harvested segments

  mutable std::mutex _m;  // guards _free, _ready, _next_block, _next_slice, _cur_buf, _slices_left, _error
  std::condition_variable _cv;
  std::vector<std::unique_ptr<detail::Pinned>> _bufs;
  std::vector<int> _free;
  std::vector<int> _ready;
  size_t _next_block = 0;           // block whose slices are being handed out
  uint64_t _next_slice = 0;         // next slice of it
  int _cur_buf = -1;                // its buffer (-1: not opened yet)
  std::vector<uint64_t> _slices_left;
  std::string _error;
  std::atomic<bool> _failed{false};

  std::mutex _join_m;
  std::thread _pusher;
  std::vector<std::thread> _readers;
  std::atomic<bool> _running{false};

  ResultT _result;
  uint64_t _running_total = 0;
  bool _published = false;
};

// ---- the one-call API -------------------------------------------------------------------------------------
template <class Tag>
std::shared_ptr<Searcher<Tag>> extern_search(const std::string& pattern, const std::string& file_path,
                                             const std::string& meta_file_path, int num_threads = 1,
                                             int max_num_readers = 1, Options opt = {}) {
  return std::make_shared<Searcher<Tag>>(pattern, file_path, meta_file_path, num_threads, max_num_readers, opt);
}

// a string literal / char* meta path must not decay to the bool of the ignore_case overload
template <class Tag>
std::shared_ptr<Searcher<Tag>> extern_search(const std::string& pattern, const std::string& file_path,
                                             const char* meta_file_path, int num_threads = 1, int max_num_readers = 1,
                                             Options opt = {}) {
  return std::make_shared<Searcher<Tag>>(pattern, file_path, std::string(meta_file_path ? meta_file_path : ""),
                                         num_threads, max_num_readers, opt);
}

template <class Tag>
std::shared_ptr<Searcher<Tag>> extern_search(const std::string& pattern, const std::string& file_path,
                                             const std::string& meta_file_path, bool ignore_case, int num_threads,
                                             int max_num_readers, Options opt = {}) {
  opt.ignore_case = ignore_case;
  return std::make_shared<Searcher<Tag>>(pattern, file_path, meta_file_path, num_threads, max_num_readers, opt);
}

template <class Tag>
std::shared_ptr<Searcher<Tag>> extern_search(const std::string& pattern, const std::string& file_path, bool ignore_case,
                                             int num_threads = 1, Options opt = {}) {
  opt.ignore_case = ignore_case;
  return std::make_shared<Searcher<Tag>>(pattern, file_path, std::string(), num_threads, 1, opt);
}

template <class Tag>
std::shared_ptr<Searcher<Tag>> extern_search(const std::string& pattern, const std::string& file_path) {
  return std::make_shared<Searcher<Tag>>(pattern, file_path, std::string(), 1, 1, Options{});
}

}  // namespace xs
