// hostreg_probe.cu -- how fast can the bytes of a FILE reach the device without a CPU copying them?
// (DESIGN.md 10 item 5 / VERDICT r1 task 6.)  Measures, on the box it runs on:
//   A  H2D from cudaHostAlloc memory                              (the ceiling: what xsg_search_host reaches)
//   B  cudaHostRegister(ReadOnly) over slices of a MAP_SHARED mapping of a tmpfs file: cost of register /
//      unregister per slice and the H2D rate out of the registered page cache
//   C  cudaMemcpyAsync straight from the unregistered mapping (the driver stages through its own pinned buffer)
//   D  pread / O_DIRECT pread into pinned memory with T threads  (the copy the readers do today)
// Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o hostreg_probe hostreg_probe.cu -lpthread
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}
#define CK(x)                                                                       \
  do {                                                                              \
    cudaError_t e = (x);                                                            \
    if (e != cudaSuccess) {                                                         \
      std::printf("FAIL %s: %s\n", #x, cudaGetErrorString(e));                      \
      cudaGetLastError();                                                           \
    }                                                                               \
  } while (0)

int main(int argc, char** argv) {
  const size_t gib = argc > 1 ? std::atol(argv[1]) : 4;
  const char* path = argc > 2 ? argv[2] : "/dev/shm/xsg_hostreg_probe.bin";
  const char* dpath = argc > 3 ? argv[3] : "/tmp/xsg_hostreg_probe_direct.bin";
  const size_t N = gib << 30, S = 256ull << 20;
  uint8_t* dev = nullptr;
  CK(cudaMalloc(&dev, 2 * S));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));

  // the file: written once (tmpfs = page cache)
  {
    int fd = ::open(path, O_CREAT | O_TRUNC | O_WRONLY, 0644);
    std::vector<uint8_t> blk(64 << 20);
    for (size_t i = 0; i < blk.size(); ++i) blk[i] = (uint8_t)(i * 2654435761u >> 24);
    for (size_t o = 0; o < N; o += blk.size()) if (::write(fd, blk.data(), blk.size()) < 0) return 1;
    ::close(fd);
  }

  // A: pinned
  {
    uint8_t* h = nullptr;
    CK(cudaHostAlloc(&h, N, cudaHostAllocDefault));
    std::memset(h, 1, N);
    for (int rep = 0; rep < 2; ++rep) {
      const double t0 = now();
      for (size_t o = 0, i = 0; o < N; o += S, ++i) CK(cudaMemcpyAsync(dev + (i & 1) * S, h + o, S, cudaMemcpyHostToDevice, st));
      CK(cudaStreamSynchronize(st));
      std::printf("A pinned H2D                : %.2f GB/s\n", N / (now() - t0) / 1e9);
    }
    // D: pread into pinned with T threads
    int fd = ::open(path, O_RDONLY);
    for (int T : {1, 4, 8, 16}) {
      const double t0 = now();
      std::vector<std::thread> th;
      for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] {
          for (size_t o = (size_t)t * (8 << 20); o < N; o += (size_t)T * (8 << 20)) {
            size_t done = 0;
            while (done < (8u << 20)) {
              ssize_t r = ::pread(fd, h + o + done, (8 << 20) - done, o + done);
              if (r <= 0) break;
              done += r;
            }
          }
        });
      for (auto& x : th) x.join();
      std::printf("D pread tmpfs -> pinned T=%2d : %.2f GB/s\n", T, N / (now() - t0) / 1e9);
    }
    ::close(fd);
    // O_DIRECT from a disk-backed file, if the filesystem allows it
    {
      int wfd = ::open(dpath, O_CREAT | O_TRUNC | O_WRONLY, 0644);
      const size_t DN = N < (2ull << 30) ? N : (2ull << 30);
      bool ok = wfd >= 0;
      for (size_t o = 0; ok && o < DN; o += 64 << 20) ok = ::write(wfd, h + o, 64 << 20) == (64 << 20);
      if (wfd >= 0) { ::fsync(wfd); ::close(wfd); }
      int dfd = ok ? ::open(dpath, O_RDONLY | O_DIRECT) : -1;
      if (dfd < 0) {
        std::printf("D O_DIRECT                  : not available on %s (%s)\n", dpath, std::strerror(errno));
      } else {
        for (int T : {1, 4, 8, 16}) {
          const double t0 = now();
          std::vector<std::thread> th;
          bool fail = false;
          for (int t = 0; t < T; ++t)
            th.emplace_back([&, t] {
              for (size_t o = (size_t)t * (8 << 20); o < DN; o += (size_t)T * (8 << 20)) {
                size_t done = 0;
                while (done < (8u << 20)) {
                  ssize_t r = ::pread(dfd, h + o + done, (8 << 20) - done, o + done);
                  if (r <= 0) { fail = true; break; }
                  done += r;
                }
              }
            });
          for (auto& x : th) x.join();
          std::printf("D O_DIRECT disk -> pinned T=%2d: %.2f GB/s%s\n", T, DN / (now() - t0) / 1e9, fail ? " (READ FAILED)" : "");
        }
        ::close(dfd);
      }
      ::unlink(dpath);
    }
    CK(cudaFreeHost(h));
  }

  // E: what page-locking costs (the fixed cost of a one-shot file search)
  for (size_t mb : {64, 256}) {
    void* h = nullptr;
    double t0 = now();
    CK(cudaHostAlloc(&h, mb << 20, cudaHostAllocDefault));
    const double t_alloc = now() - t0;
    t0 = now();
    CK(cudaFreeHost(h));
    const double t_free = now() - t0;
    // anonymous memory, huge pages advised, first touch by 8 threads, then registered
    const size_t bytes = mb << 20;
    t0 = now();
    uint8_t* m = (uint8_t*)::aligned_alloc(2 << 20, bytes);
    ::madvise(m, bytes, MADV_HUGEPAGE);
    {
      std::vector<std::thread> th;
      for (int t = 0; t < 8; ++t) th.emplace_back([&, t] { std::memset(m + bytes / 8 * t, 0, bytes / 8); });
      for (auto& x : th) x.join();
    }
    const double t_touch = now() - t0;
    t0 = now();
    CK(cudaHostRegister(m, bytes, cudaHostRegisterDefault));
    const double t_reg = now() - t0;
    t0 = now();
    CK(cudaMemcpyAsync(dev, m, bytes < 2 * S ? bytes : 2 * S, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    const double t_cp = now() - t0;
    t0 = now();
    CK(cudaHostUnregister(m));
    const double t_unreg = now() - t0;
    ::free(m);
    std::printf("E %3zu MiB: cudaHostAlloc %.1f ms, cudaFreeHost %.1f ms | THP alloc+touch(8 thr) %.1f ms, cudaHostRegister %.1f ms, H2D %.2f GB/s, unregister %.1f ms\n",
                mb, t_alloc * 1e3, t_free * 1e3, t_touch * 1e3, t_reg * 1e3, (bytes < 2 * S ? bytes : 2 * S) / t_cp / 1e9, t_unreg * 1e3);
  }

  // B0: other ways to map the file before registering it
  {
    int attr = -1;
    cudaDeviceGetAttribute(&attr, cudaDevAttrHostRegisterReadOnlySupported, 0);
    std::printf("B0 cudaDevAttrHostRegisterReadOnlySupported = %d\n", attr);
    int attr2 = -1;
    cudaDeviceGetAttribute(&attr2, cudaDevAttrPageableMemoryAccess, 0);
    std::printf("B0 cudaDevAttrPageableMemoryAccess = %d\n", attr2);
    struct V { const char* name; int oflag; int prot; int mflag; };
    const V vs[] = {{"O_RDWR  PROT_RW MAP_SHARED ", O_RDWR, PROT_READ | PROT_WRITE, MAP_SHARED},
                    {"O_RDONLY PROT_RW MAP_PRIVATE", O_RDONLY, PROT_READ | PROT_WRITE, MAP_PRIVATE},
                    {"O_RDONLY PROT_R  MAP_PRIVATE", O_RDONLY, PROT_READ, MAP_PRIVATE}};
    for (const V& v : vs) {
      int fd = ::open(path, v.oflag);
      uint8_t* m = (uint8_t*)::mmap(nullptr, N, v.prot, v.mflag, fd, 0);
      if (m == MAP_FAILED) { std::printf("B0 %s: mmap failed\n", v.name); ::close(fd); continue; }
      for (unsigned flag : {(unsigned)cudaHostRegisterDefault, (unsigned)cudaHostRegisterReadOnly}) {
        double t0 = now();
        cudaError_t e = cudaHostRegister(m, S, flag);
        const double t_reg = now() - t0;
        if (e != cudaSuccess) {
          std::printf("B0 %s flag=%u: cudaHostRegister -> %s\n", v.name, flag, cudaGetErrorString(e));
          cudaGetLastError();
          continue;
        }
        t0 = now();
        CK(cudaMemcpyAsync(dev, m, S, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        const double t_cp = now() - t0;
        t0 = now();
        CK(cudaHostUnregister(m));
        std::printf("B0 %s flag=%u: register 256 MiB %.1f ms (%.2f GB/s), H2D %.2f GB/s, unregister %.1f ms\n", v.name, flag,
                    t_reg * 1e3, S / t_reg / 1e9, S / t_cp / 1e9, (now() - t0) * 1e3);
      }
      ::munmap(m, N);
      ::close(fd);
    }
  }

  // B: register slices of the mapping
  if (argc > 4) {
    int fd = ::open(path, O_RDONLY);
    for (int populate = 0; populate < 2; ++populate) {
      uint8_t* m = (uint8_t*)::mmap(nullptr, N, PROT_READ, MAP_SHARED | (populate ? MAP_POPULATE : 0), fd, 0);
      if (m == MAP_FAILED) { std::printf("mmap failed\n"); return 1; }
      for (size_t slice : {S, 4 * S}) {
        double t_reg = 0, t_cp = 0, t_unreg = 0;
        bool ok = true;
        const double t_all = now();
        for (size_t o = 0; o < N && ok; o += slice) {
          double t0 = now();
          cudaError_t e = cudaHostRegister(m + o, slice, cudaHostRegisterReadOnly);
          if (e != cudaSuccess) {
            cudaGetLastError();
            e = cudaHostRegister(m + o, slice, cudaHostRegisterDefault);
            if (e != cudaSuccess) {
              std::printf("B cudaHostRegister failed: %s\n", cudaGetErrorString(e));
              cudaGetLastError();
              ok = false;
              break;
            }
          }
          t_reg += now() - t0;
          t0 = now();
          for (size_t q = 0; q < slice; q += S) CK(cudaMemcpyAsync(dev, m + o + q, S, cudaMemcpyHostToDevice, st));
          CK(cudaStreamSynchronize(st));
          t_cp += now() - t0;
          t0 = now();
          CK(cudaHostUnregister(m + o));
          t_unreg += now() - t0;
        }
        if (ok)
          std::printf("B populate=%d slice=%4zu MiB : register %.3f s (%.2f GB/s), H2D %.2f GB/s, unregister %.3f s, serial total %.2f GB/s\n",
                      populate, slice >> 20, t_reg, N / t_reg / 1e9, N / t_cp / 1e9, t_unreg, N / (now() - t_all) / 1e9);
      }
      // pipelined: a helper thread registers slice i+1 while slice i is copied
      {
        const size_t ns = N / S;
        std::vector<int> ready(ns, 0);
        const double t0 = now();
        std::thread reg([&] {
          for (size_t i = 0; i < ns; ++i) {
            if (cudaHostRegister(m + i * S, S, cudaHostRegisterReadOnly) != cudaSuccess) { cudaGetLastError(); cudaHostRegister(m + i * S, S, cudaHostRegisterDefault); }
            __atomic_store_n(&ready[i], 1, __ATOMIC_RELEASE);
          }
        });
        for (size_t i = 0; i < ns; ++i) {
          while (!__atomic_load_n(&ready[i], __ATOMIC_ACQUIRE)) {}
          CK(cudaMemcpyAsync(dev + (i & 1) * S, m + i * S, S, cudaMemcpyHostToDevice, st));
        }
        CK(cudaStreamSynchronize(st));
        const double t1 = now();
        reg.join();
        for (size_t i = 0; i < ns; ++i) CK(cudaHostUnregister(m + i * S));
        std::printf("B populate=%d pipelined (1 register thread): %.2f GB/s to last copy, %.2f GB/s incl. unregister\n", populate,
                    N / (t1 - t0) / 1e9, N / (now() - t0) / 1e9);
      }
      // C: unregistered mapping
      if (!populate) {
        const double t0 = now();
        for (size_t o = 0, i = 0; o < N; o += S, ++i) CK(cudaMemcpyAsync(dev + (i & 1) * S, m + o, S, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        std::printf("C pageable mapping H2D      : %.2f GB/s\n", N / (now() - t0) / 1e9);
      }
      ::munmap(m, N);
    }
    ::close(fd);
  }
  ::unlink(path);
  return 0;
}
