// scripts/dev/pcie_probe.cu -- measurement aid, not product code.
// What bounds the host->host number of b200jpeg_decode_batch is the device->host copy of the RGB (1.59 GB per C2 batch).
// This probe separates the box from the allocation: the same traffic (D2H of `rgb_mb` MB in `chunk_mb` pieces on one
// stream, H2D of `scan_mb` MB on a second stream at the same time) for every kind of pinned host memory, on 1..N GPUs
// driven by one host thread each (the shape of b200jpeg_multi_decode_batch).
//   nvcc -O2 -std=c++17 -o build/pcie_probe scripts/dev/pcie_probe.cu && build/pcie_probe [max_gpus] [rgb_mb] [chunk_mb]
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <thread>
#include <vector>
#include <atomic>

enum Mode { kDefault = 0, kWriteCombined = 1, kThpRegister = 2, kHugeTlbRegister = 3, kMallocRegister = 4, kModes = 5 };
static const char *kNames[kModes] = {"cudaHostAlloc(default)", "cudaHostAlloc(writeCombined)", "mmap+MADV_HUGEPAGE+cudaHostRegister",
                                     "mmap(MAP_HUGETLB)+cudaHostRegister", "malloc(4K pages)+cudaHostRegister"};

struct HostBuf { void *p = nullptr; size_t n = 0; Mode m = kDefault; bool ok = false; };

static HostBuf host_alloc(Mode m, size_t n) {
  HostBuf b; b.n = n; b.m = m;
  const size_t huge = (size_t)2 << 20;
  const size_t len = (n + huge - 1) / huge * huge;
  switch (m) {
    case kDefault: b.ok = cudaHostAlloc(&b.p, n, cudaHostAllocDefault) == cudaSuccess; break;
    case kWriteCombined: b.ok = cudaHostAlloc(&b.p, n, cudaHostAllocWriteCombined) == cudaSuccess; break;
    case kThpRegister: {
      void *p = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
      if (p == MAP_FAILED) break;
      madvise(p, len, MADV_HUGEPAGE);
      memset(p, 1, len);
      b.p = p; b.n = len;
      b.ok = cudaHostRegister(p, len, cudaHostRegisterDefault) == cudaSuccess;
      if (!b.ok) munmap(p, len);
      break;
    }
    case kHugeTlbRegister: {
      void *p = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_HUGETLB, -1, 0);
      if (p == MAP_FAILED) break;
      memset(p, 1, len);
      b.p = p; b.n = len;
      b.ok = cudaHostRegister(p, len, cudaHostRegisterDefault) == cudaSuccess;
      if (!b.ok) munmap(p, len);
      break;
    }
    case kMallocRegister: {
      void *p = nullptr;
      if (posix_memalign(&p, 4096, len)) break;
      madvise(p, len, MADV_NOHUGEPAGE);
      memset(p, 1, len);
      b.p = p; b.n = len;
      b.ok = cudaHostRegister(p, len, cudaHostRegisterDefault) == cudaSuccess;
      if (!b.ok) free(p);
      break;
    }
    default: break;
  }
  cudaGetLastError();
  return b;
}
static void host_free(HostBuf &b) {
  if (!b.ok) return;
  if (b.m == kDefault || b.m == kWriteCombined) cudaFreeHost(b.p);
  else { cudaHostUnregister(b.p); if (b.m == kMallocRegister) free(b.p); else munmap(b.p, b.n); }
  b.ok = false;
}

int main(int argc, char **argv) {
  int ndev = 0;
  cudaGetDeviceCount(&ndev);
  const int max_gpus = argc > 1 ? atoi(argv[1]) : ndev;
  const size_t rgb = (size_t)(argc > 2 ? atoi(argv[2]) : 1593) << 20, chunk = (size_t)(argc > 3 ? atoi(argv[3]) : 100) << 20;
  const size_t scan = (size_t)153 << 20;
  for (int n = 1; n <= max_gpus && n <= ndev; n *= 2) {
    for (int m = 0; m < kModes; ++m) {
      std::vector<double> ms((size_t)n, 0.0);
      std::vector<int> okv((size_t)n, 0);
      std::atomic<int> ready(0);
      auto work = [&](int k) {
        cudaSetDevice(k);
        void *d_rgb = nullptr, *d_scan = nullptr;
        cudaMalloc(&d_rgb, rgb); cudaMalloc(&d_scan, scan);
        cudaMemset(d_rgb, 3, rgb);
        HostBuf h_rgb = host_alloc((Mode)m, rgb), h_scan = host_alloc(kDefault, scan);
        cudaStream_t s0, s1;
        cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking);
        okv[(size_t)k] = h_rgb.ok && h_scan.ok && d_rgb && d_scan;
        double best = 1e30;
        for (int rep = 0; rep < 4; ++rep) {
          ready.fetch_add(1);                                    // every thread arrives once per repetition: start together
          while (ready.load() < (rep + 1) * n) std::this_thread::yield();
          if (!okv[(size_t)k]) continue;
          const auto t0 = std::chrono::steady_clock::now();
          for (size_t off = 0, so = 0; off < rgb; off += chunk, so += scan / ((rgb + chunk - 1) / chunk)) {
            const size_t len = rgb - off < chunk ? rgb - off : chunk, sl = scan / ((rgb + chunk - 1) / chunk);
            cudaMemcpyAsync((char *)d_scan + so, (char *)h_scan.p + so, sl, cudaMemcpyHostToDevice, s1);
            cudaMemcpyAsync((char *)h_rgb.p + off, (char *)d_rgb + off, len, cudaMemcpyDeviceToHost, s0);
          }
          cudaStreamSynchronize(s0); cudaStreamSynchronize(s1);
          const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
          if (rep > 0 && t < best) best = t;
        }
        ms[(size_t)k] = best;
        host_free(h_rgb); host_free(h_scan);
        cudaFree(d_rgb); cudaFree(d_scan);
        cudaStreamDestroy(s0); cudaStreamDestroy(s1);
      };
      std::vector<std::thread> pool;
      for (int k = 1; k < n; ++k) pool.emplace_back(work, k);
      work(0);
      for (auto &t : pool) t.join();
      double worst = 0; bool ok = true;
      for (int k = 0; k < n; ++k) { if (ms[(size_t)k] > worst) worst = ms[(size_t)k]; ok = ok && okv[(size_t)k]; }
      if (!ok) { printf("{\"gpus\":%d,\"mode\":\"%s\",\"available\":false}\n", n, kNames[m]); continue; }
      printf("{\"gpus\":%d,\"mode\":\"%s\",\"ms\":%.2f,\"d2h_gbs_per_gpu\":%.1f,\"d2h_gbs_total\":%.1f,\"chunk_mb\":%zu}\n", n, kNames[m], worst,
             rgb / worst / 1e6, n * (double)rgb / worst / 1e6, chunk >> 20);
      fflush(stdout);
    }
  }
  return 0;
}
