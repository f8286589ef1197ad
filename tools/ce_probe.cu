// ce_probe.cu -- what the copy engines give an all-to-all of 128 MiB blocks on this box, alone and under a kernel that
// saturates HBM.  One process, G devices, peer access on.  Prints one JSON line per mode.  Measurement tool for the
// design of hj_xchg_* (DESIGN.md section 5); not part of the product.
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/_bin/ce_probe tools/ce_probe.cu
//   tools/_bin/ce_probe <G> <block MiB> <rounds>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e_ = (x);                                                            \
    if (e_ != cudaSuccess) {                                                         \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));     \
      exit(1);                                                                       \
    }                                                                                \
  } while (0)

__global__ void hbm_load_kernel(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n, int reps) {
  for (int r = 0; r < reps; r++)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
      uint4 v = a[i];
      v.x += r;
      b[i] = v;
    }
}

__global__ void sm_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

int main(int argc, char** argv) {
  const int G = argc > 1 ? atoi(argv[1]) : 4;
  const size_t blk = (size_t)(argc > 2 ? atoi(argv[2]) : 128) << 20;
  const int rounds = argc > 3 ? atoi(argv[3]) : 8;
  int nd = 0;
  CK(cudaGetDeviceCount(&nd));
  if (nd < G) { fprintf(stderr, "need %d devices, have %d\n", G, nd); return 1; }
  std::vector<std::vector<char*>> src(G, std::vector<char*>(G)), dst(G, std::vector<char*>(G));
  std::vector<std::vector<cudaStream_t>> st(G, std::vector<cudaStream_t>(G));
  std::vector<cudaStream_t> one(G), ks(G);
  std::vector<uint4*> la(G), lb(G);
  const size_t load_bytes = (size_t)2 << 30;
  for (int d = 0; d < G; d++) {
    CK(cudaSetDevice(d));
    for (int p = 0; p < G; p++)
      if (p != d) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
        (void)cudaGetLastError();
      }
    for (int p = 0; p < G; p++) {
      CK(cudaMalloc(&src[d][p], blk));   // on d, to be sent to p
      CK(cudaMalloc(&dst[d][p], blk));   // on d, receives from p
      CK(cudaMemset(src[d][p], d * 16 + p, blk));
      CK(cudaStreamCreateWithFlags(&st[d][p], cudaStreamNonBlocking));
    }
    CK(cudaStreamCreateWithFlags(&one[d], cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&ks[d], cudaStreamNonBlocking));
    CK(cudaMalloc(&la[d], load_bytes));
    CK(cudaMalloc(&lb[d], load_bytes));
    CK(cudaMemset(la[d], 1, load_bytes));
  }
  auto sync_all = [&] {
    for (int d = 0; d < G; d++) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
  };
  // mode: 0 push ring, per-peer streams; 1 push index order; 2 push ring, one stream; 3 pull ring, per-peer streams;
  //       4 SM copy kernel (48 CTAs per peer stream... one kernel per peer, 8 CTAs each), ring
  const char* names[] = {"push_ring_streams", "push_index_streams", "push_ring_onestream", "pull_ring_streams",
                         "sm_copy_ring"};
  for (int load = 0; load < 2; load++)
    for (int mode = 0; mode < 5; mode++) {
      double best = 1e30;
      for (int rep = 0; rep < 3; rep++) {
        sync_all();
        auto t0 = std::chrono::steady_clock::now();
        if (load)
          for (int d = 0; d < G; d++) {
            CK(cudaSetDevice(d));
            hbm_load_kernel<<<148 * 4, 256, 0, ks[d]>>>(la[d], lb[d], load_bytes / 16, 100);
          }
        for (int r = 0; r < rounds; r++)
          for (int k = 1; k < G; k++)
            for (int d = 0; d < G; d++) {
              CK(cudaSetDevice(d));
              int p = (mode == 1) ? (k <= d ? k - 1 : k) : (d + k) % G;
              if (mode == 0 || mode == 1)
                CK(cudaMemcpyPeerAsync(dst[p][d], p, src[d][p], d, blk, st[d][p]));
              else if (mode == 2)
                CK(cudaMemcpyPeerAsync(dst[p][d], p, src[d][p], d, blk, one[d]));
              else if (mode == 3)   // d pulls from p
                CK(cudaMemcpyPeerAsync(dst[d][p], d, src[p][d], p, blk, st[d][p]));
              else
                sm_copy_kernel<<<8, 512, 0, st[d][p]>>>((const uint4*)src[d][p], (uint4*)dst[p][d], blk / 16);
            }
        // wait for the copies only (the load kernel runs on)
        for (int d = 0; d < G; d++) {
          CK(cudaSetDevice(d));
          for (int p = 0; p < G; p++) CK(cudaStreamSynchronize(st[d][p]));
          CK(cudaStreamSynchronize(one[d]));
        }
        double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (s < best) best = s;
        if (load) {
          // stop the load kernels: no flag, they just run out; wait for them
          sync_all();
        }
      }
      double gb = (double)blk * (G - 1) * rounds / 1e9;
      printf("{\"mode\": \"%s\", \"hbm_load\": %d, \"gpus\": %d, \"block_mib\": %zu, \"rounds\": %d, \"ms\": %.3f, "
             "\"egress_gbs_per_gpu\": %.1f}\n",
             names[mode], load, G, blk >> 20, rounds, best * 1e3, gb / best);
      fflush(stdout);
    }
  return 0;
}
