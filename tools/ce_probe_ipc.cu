// ce_probe_ipc.cu -- the copy engines' all-to-all rate the way hj_xchg_* drives them: ONE PROCESS PER GPU, the
// receive windows mapped through cudaIpcOpenMemHandle, peers in ring order.  Companion of ce_probe.cu (one process,
// cudaDeviceEnablePeerAccess).  Measurement tool, not part of the product.
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/_bin/ce_probe_ipc tools/ce_probe_ipc.cu
//   tools/_bin/ce_probe_ipc <G> <block MiB> <rounds>
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>

static int g_rank = -1;
#define CK(x)                                                                                    \
  do {                                                                                           \
    cudaError_t e_ = (x);                                                                        \
    if (e_ != cudaSuccess) {                                                                     \
      fprintf(stderr, "rank %d %s:%d %s\n", g_rank, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      _exit(1);                                                                                  \
    }                                                                                            \
  } while (0)

struct Shared {
  std::atomic<int> arrived[256];
  cudaIpcMemHandle_t handle[16];
  float ms[16][16];
};

static void barrier(Shared* sh, int G, int& phase) {
  sh->arrived[phase].fetch_add(1);
  while (sh->arrived[phase].load() < G) usleep(50);
  phase++;
}

__global__ void hbm_load_kernel(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n, int reps) {
  for (int r = 0; r < reps; r++)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
      uint4 v = a[i];
      v.x += r;
      b[i] = v;
    }
}

static int child(Shared* sh, int G, int rank, size_t blk, int rounds) {
  g_rank = rank;
  int phase = 0;
  CK(cudaSetDevice(rank));
  char *window, *send;
  CK(cudaMalloc(&window, blk * G));  // block s of the window receives from rank s
  CK(cudaMalloc(&send, blk * G));
  CK(cudaMemset(send, rank + 1, blk * G));
  CK(cudaIpcGetMemHandle(&sh->handle[rank], window));
  uint4 *la, *lb;
  const size_t load_bytes = (size_t)2 << 30;
  CK(cudaMalloc(&la, load_bytes));
  CK(cudaMalloc(&lb, load_bytes));
  CK(cudaMemset(la, 1, load_bytes));
  CK(cudaDeviceSynchronize());
  barrier(sh, G, phase);
  char* peer[16];
  for (int p = 0; p < G; p++) {
    peer[p] = window;
    if (p != rank) CK(cudaIpcOpenMemHandle((void**)&peer[p], sh->handle[p], cudaIpcMemLazyEnablePeerAccess));
  }
  cudaStream_t cs, ks, ps[16];
  cudaEvent_t e0, e1, pe[16];
  CK(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&ks, cudaStreamNonBlocking));
  for (int p = 0; p < G; p++) {
    CK(cudaStreamCreateWithFlags(&ps[p], cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&pe[p], cudaEventDisableTiming));
  }
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  // 0: cudaMemcpyAsync D2D, one stream; 1: cudaMemcpyPeerAsync, one stream; 2: cudaMemcpyAsync, a stream per peer;
  // 3: mode 0, each block as two halves on two streams
  const char* names[] = {"ipc_memcpy_onestream", "ipc_memcpypeer_onestream", "ipc_memcpy_peerstreams",
                         "ipc_memcpy_halves"};
  for (int load = 0; load < 2; load++)
    for (int mode = 0; mode < 4; mode++) {
      float best = 1e30f;
      for (int rep = 0; rep < 3; rep++) {
        CK(cudaDeviceSynchronize());
        barrier(sh, G, phase);
        if (load) hbm_load_kernel<<<148 * 4, 256, 0, ks>>>(la, lb, load_bytes / 16, 30);
        CK(cudaEventRecord(e0, cs));
        for (int p = 0; p < G; p++) CK(cudaStreamWaitEvent(ps[p], e0, 0));
        for (int r = 0; r < rounds; r++)
          for (int k = 1; k < G; k++) {
            const int p = (rank + k) % G;
            char* dst = peer[p] + blk * rank;
            const char* src = send + blk * p;
            if (mode == 0) CK(cudaMemcpyAsync(dst, src, blk, cudaMemcpyDeviceToDevice, cs));
            else if (mode == 1) CK(cudaMemcpyPeerAsync(dst, p, src, rank, blk, cs));
            else if (mode == 2) CK(cudaMemcpyAsync(dst, src, blk, cudaMemcpyDeviceToDevice, ps[p]));
            else {
              CK(cudaMemcpyAsync(dst, src, blk / 2, cudaMemcpyDeviceToDevice, cs));
              CK(cudaMemcpyAsync(dst + blk / 2, src + blk / 2, blk / 2, cudaMemcpyDeviceToDevice, ps[rank]));
            }
          }
        for (int p = 0; p < G; p++) {
          CK(cudaEventRecord(pe[p], ps[p]));
          CK(cudaStreamWaitEvent(cs, pe[p], 0));
        }
        CK(cudaEventRecord(e1, cs));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
        CK(cudaDeviceSynchronize());
      }
      sh->ms[load * 4 + mode][rank] = best;
      barrier(sh, G, phase);
      if (rank == 0) {
        float worst = 0.f;
        for (int r = 0; r < G; r++) worst = sh->ms[load * 4 + mode][r] > worst ? sh->ms[load * 4 + mode][r] : worst;
        const double gb = (double)blk * (G - 1) * rounds / 1e9;
        printf("{\"mode\": \"%s\", \"hbm_load\": %d, \"gpus\": %d, \"block_mib\": %zu, \"rounds\": %d, "
               "\"ms_max_rank\": %.3f, \"egress_gbs_per_gpu\": %.1f}\n",
               names[mode], load, G, blk >> 20, rounds, worst, gb / (worst * 1e-3));
        fflush(stdout);
      }
    }
  barrier(sh, G, phase);
  for (int p = 0; p < G; p++)
    if (p != rank) cudaIpcCloseMemHandle(peer[p]);
  return 0;
}

int main(int argc, char** argv) {
  const int G = argc > 1 ? atoi(argv[1]) : 4;
  const size_t blk = (size_t)(argc > 2 ? atoi(argv[2]) : 128) << 20;
  const int rounds = argc > 3 ? atoi(argv[3]) : 8;
  if (G < 2 || G > 16) return 1;
  Shared* sh = (Shared*)mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  if (sh == MAP_FAILED) return 1;
  memset((void*)sh, 0, sizeof(Shared));
  pid_t pids[16];
  for (int r = 0; r < G; r++) {
    pids[r] = fork();
    if (pids[r] == 0) _exit(child(sh, G, r, blk, rounds));
  }
  int bad = 0;
  for (int r = 0; r < G; r++) {
    int st = 0;
    waitpid(pids[r], &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st)) bad = 1;
  }
  return bad;
}
