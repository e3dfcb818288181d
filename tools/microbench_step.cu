// Where does k_step spend its time?  Launches the real kernel on dummy buffers with parts switched off.
#include "../ddp_practice_b200/csrc/step_kernel.cuh"
#include <cstdio>
using namespace dstep;
template <typename F> float time_graph(F enqueue, int nodes, int reps) {
  cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  cudaGraph_t g; cudaGraphExec_t e;
  cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < nodes; ++i) enqueue(s);
  cudaStreamEndCapture(s, &g); cudaGraphInstantiate(&e, g, 0);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 20; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(a, s);
  for (int i = 0; i < reps; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(b, s); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  cudaError_t err = cudaGetLastError(); if (err != cudaSuccess) printf("  (error: %s)\n", cudaGetErrorString(err));
  return ms * 1e3f / reps / nodes;
}
int main() {
  float *params, *grads, *gs, *gpart, *gred, *gdirect, *nll, *bnbuf; long long* nbt; DevState* ds; void* tiles;
  cudaMalloc(&params, N_PAD * 4); cudaMalloc(&grads, N_PAD * 4); cudaMalloc(&gs, N_PAD * 4);
  cudaMalloc(&gpart, (size_t)160 * N_PAD * 4); cudaMalloc(&gred, N_PAD * 4); cudaMalloc(&gdirect, N_PAD * 4);
  cudaMalloc(&nll, 4096 * 4); cudaMalloc(&bnbuf, 96 * 4); cudaMalloc(&nbt, 16); cudaMalloc(&ds, sizeof(DevState)); cudaMalloc(&tiles, 51200);
  cudaMemset(params, 0, N_PAD * 4); cudaMemset(gpart, 0, (size_t)160 * N_PAD * 4); cudaMemset(gred, 0, N_PAD * 4);
  cudaMemset(gdirect, 0, N_PAD * 4); cudaMemset(nll, 0, 4096 * 4); cudaMemset(ds, 0, sizeof(DevState));
  DevState h; memset(&h, 0, sizeof h); h.scale = 32768.f; for (int t = 0; t < 10; ++t) h.grp[t] = {1e-4f, 0.f, 0.f, 0.f, 0}; h.timeout_cycles = 1LL << 40; cudaMemcpy(ds, &h, sizeof h, cudaMemcpyHostToDevice);
  CommDev comm; memset(&comm, 0, sizeof comm); comm.world = 1;
  struct V { const char* name; int r1, r2w, r2b, rfc; bool red, tl; } vs[] = {
      {"as in the step (rows 32/1/8/4, tiles)", 32, 32, 8, 4, true, true},
      {"no tiles", 32, 32, 8, 4, true, false},
      {"conv2w from 32 rows in-kernel", 32, 32, 8, 4, false, true},
      {"all rows = 1", 1, 1, 1, 1, true, true},
      {"all rows = 0", 0, 0, 0, 0, true, false},
  };
  for (auto& v : vs) {
    StepArgs a{params, grads, gs, gpart, v.red ? gred : nullptr, gdirect, nll, nullptr, bnbuf, nbt, ds, comm, 32,
               v.r1, v.r2w, v.r2b, v.rfc, 1, 1, 2.f, 0.5f, 2000, nullptr, 0, nullptr, 0, nullptr, 0};
    float us = time_graph([&](cudaStream_t s) { k_step<__half><<<STEP_CTAS, STEP_THREADS, 0, s>>>(a); }, 8, 500);
    printf("%-42s : %.2f us\n", v.name, us);
  }
  return 0;
}
