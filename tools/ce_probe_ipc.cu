// ce_probe_ipc.cu -- the copy engines' all-to-all rate the way hj_xchg_* drives them: ONE PROCESS PER GPU, the
// receive windows mapped through cudaIpcOpenMemHandle, peers in ring order.  Companion of ce_probe.cu (one process,
// cudaDeviceEnablePeerAccess).  Measurement tool, not part of the product.
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/_bin/ce_probe_ipc tools/ce_probe_ipc.cu
//   tools/_bin/ce_probe_ipc <G> <block MiB> <rounds>
#include <cuda.h>
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
  std::atomic<int> arrived[512];
  cudaIpcMemHandle_t handle[16];
  float ms[32][16];
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
  CK(cudaMalloc(&window, blk * G + 65536));  // block s of the window receives from rank s; then 8192 token words
  CK(cudaMemset(window + blk * G, 0, 65536));
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
  // PROBE_DUMMY_STREAMS=n: n streams created (and used once) before the copy stream, as a host program with its own
  // compute / staging streams would have; PROBE_MODES=bitmask of the modes to run
  const int ndummy = getenv("PROBE_DUMMY_STREAMS") ? atoi(getenv("PROBE_DUMMY_STREAMS")) : 0;
  const unsigned mode_mask = getenv("PROBE_MODES") ? (unsigned)strtoul(getenv("PROBE_MODES"), nullptr, 0) : ~0u;
  for (int i = 0; i < ndummy; i++) {
    cudaStream_t d;
    CK(cudaStreamCreateWithFlags(&d, cudaStreamNonBlocking));
    CK(cudaMemcpyAsync(lb, la, 1 << 20, cudaMemcpyDeviceToDevice, d));
    hbm_load_kernel<<<1, 32, 0, d>>>(la, lb, 32, 1);
    CK(cudaStreamSynchronize(d));
  }
  CK(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&ks, cudaStreamNonBlocking));
  for (int p = 0; p < G; p++) {
    CK(cudaStreamCreateWithFlags(&ps[p], cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&pe[p], cudaEventDisableTiming));
  }
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  // 0: cudaMemcpyAsync D2D, one stream; 1: cudaMemcpyPeerAsync, one stream; 2: cudaMemcpyAsync, a stream per peer;
  // 3: mode 0, each block as two halves on two streams; 4: mode 0 + two 8-byte copies after every block (fill levels
  // and flag of hj_xchg); 5: mode 0, rank r starts r * 0.4 ms late; 6: mode 2, rank r starts r * 0.4 ms late;
  // 7: mode 2 + the two 8-byte copies; 8: mode 5 (skewed) with the ring kept in lock step by tokens: before a rank
  // sends round k+1 (target rank+k+1) it waits (cuStreamWaitValue64, no SM involved) until its successor rank+1 --
  // the previous sender to that target -- has finished its round k
  const char* names[] = {"ipc_memcpy_onestream", "ipc_memcpypeer_onestream", "ipc_memcpy_peerstreams", "ipc_memcpy_halves",
                         "ipc_onestream_tiny", "ipc_onestream_skewed", "ipc_peerstreams_skewed", "ipc_peerstreams_tiny",
                         "ipc_onestream_skewed_tokens"};
  constexpr int kModes = 9;
  typedef CUresult (*WaitFn)(CUstream, CUdeviceptr, cuuint64_t, unsigned int);
  WaitFn wait_value = nullptr;
  {
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuStreamWaitValue64", (void**)&wait_value, cudaEnableDefault, &qr));
    if (qr != cudaDriverEntryPointSuccess || !wait_value) { fprintf(stderr, "no cuStreamWaitValue64\n"); _exit(1); }
  }
  unsigned long long* gen_word;  // device word holding the current generation (source of the token copies)
  CK(cudaMalloc(&gen_word, 256));
  unsigned long long gen = 0;
  unsigned long long* my_tok = (unsigned long long*)(window + blk * G);
  const int succ = (rank + 1) % G, pred = (rank + G - 1) % G;
  cudaEvent_t first[16];
  for (int p = 0; p < G; p++) CK(cudaEventCreate(&first[p]));
  for (int load = 0; load < 2; load++)
    for (int mode = 0; mode < kModes; mode++) {
      if (!((mode_mask >> mode) & 1u)) continue;
      float best = 1e30f, first_ms[16] = {0};
      for (int rep = 0; rep < 3; rep++) {
        CK(cudaDeviceSynchronize());
        barrier(sh, G, phase);
        gen++;
        CK(cudaMemcpy(gen_word, &gen, 8, cudaMemcpyHostToDevice));
        barrier(sh, G, phase);
        if (mode == 5 || mode == 6 || mode == 8) usleep(400 * rank);
        if (load) hbm_load_kernel<<<148 * 4, 256, 0, ks>>>(la, lb, load_bytes / 16, 30);
        CK(cudaEventRecord(e0, cs));
        for (int p = 0; p < G; p++) CK(cudaStreamWaitEvent(ps[p], e0, 0));
        const bool per_peer = mode == 2 || mode == 6 || mode == 7;
        for (int r = 0; r < rounds; r++)
          for (int k = 1; k < G; k++) {
            const int p = (rank + k) % G;
            char* dst = peer[p] + blk * rank;
            const char* src = send + blk * p;
            cudaStream_t s = per_peer ? ps[p] : cs;
            if (mode == 8 && k >= 2) {
              CUresult cr = wait_value((CUstream)cs, (CUdeviceptr)(my_tok + r * G + (k - 1)), gen, CU_STREAM_WAIT_VALUE_GEQ);
              if (cr != CUDA_SUCCESS) { fprintf(stderr, "rank %d cuStreamWaitValue64 -> %d\n", rank, (int)cr); _exit(1); }
            }
            if (mode == 1) CK(cudaMemcpyPeerAsync(dst, p, src, rank, blk, cs));
            else if (mode == 3) {
              CK(cudaMemcpyAsync(dst, src, blk / 2, cudaMemcpyDeviceToDevice, cs));
              CK(cudaMemcpyAsync(dst + blk / 2, src + blk / 2, blk / 2, cudaMemcpyDeviceToDevice, ps[rank]));
            } else CK(cudaMemcpyAsync(dst, src, blk, cudaMemcpyDeviceToDevice, s));
            if (mode == 4 || mode == 7) {
              CK(cudaMemcpyAsync(dst + 1024, src + 1024, 1024, cudaMemcpyDeviceToDevice, s));
              CK(cudaMemcpyAsync(dst, src, 8, cudaMemcpyDeviceToDevice, s));
            }
            if (mode == 8 && k <= G - 2)  // tell the predecessor: my round k is done, its round k+1 has the target to itself
              CK(cudaMemcpyAsync((unsigned long long*)(peer[pred] + blk * G) + r * G + k, gen_word, 8, cudaMemcpyDeviceToDevice, s));
            if (r == 0) CK(cudaEventRecord(first[p], s));
          }
        for (int p = 0; p < G; p++) {
          CK(cudaEventRecord(pe[p], ps[p]));
          CK(cudaStreamWaitEvent(cs, pe[p], 0));
        }
        CK(cudaEventRecord(e1, cs));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) {
          best = ms;
          for (int k = 1; k < G; k++) CK(cudaEventElapsedTime(&first_ms[k], e0, first[(rank + k) % G]));
        }
        CK(cudaDeviceSynchronize());
      }
      sh->ms[load * kModes + mode][rank] = best;
      barrier(sh, G, phase);
      if (rank == 0) {
        float worst = 0.f;
        for (int r = 0; r < G; r++) worst = sh->ms[load * kModes + mode][r] > worst ? sh->ms[load * kModes + mode][r] : worst;
        const double gb = (double)blk * (G - 1) * rounds / 1e9;
        printf("{\"mode\": \"%s\", \"hbm_load\": %d, \"gpus\": %d, \"block_mib\": %zu, \"rounds\": %d, "
               "\"ms_max_rank\": %.3f, \"egress_gbs_per_gpu\": %.1f, \"rank0_first_round_done_ms\": [",
               names[mode], load, G, blk >> 20, rounds, worst, gb / (worst * 1e-3));
        for (int k = 1; k < G; k++) printf("%s%.2f", k > 1 ? ", " : "", first_ms[k]);
        printf("]}\n");
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
  size_t blk = (size_t)(argc > 2 ? atoi(argv[2]) : 128) << 20;
  if (getenv("PROBE_BLOCK_BYTES")) blk = strtoull(getenv("PROBE_BLOCK_BYTES"), nullptr, 0);  // e.g. a 256-byte aligned odd size
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
