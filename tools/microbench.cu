// Launch-latency floors on the box: per-node cost of a chain of tiny dependent kernels inside a CUDA graph,
// for plain launches, 8-CTA cluster launches, and programmatic dependent launch (PDL) edges.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/microbench tools/microbench.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__global__ void k_plain(int* p) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(p, 1); }
__global__ void __cluster_dims__(8, 1, 1) k_cluster(int* p) {
  cg::this_cluster().sync();
  if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(p, 1);
}
__global__ void k_pdl(int* p) {
  cudaGridDependencySynchronize();
  if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(p, 1);
  cudaTriggerProgrammaticLaunchCompletion();
}

template <typename F> float time_graph(F enqueue, int nodes, int reps) {
  cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  cudaGraph_t g; cudaGraphExec_t e;
  cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < nodes; ++i) enqueue(s);
  cudaStreamEndCapture(s, &g);
  cudaGraphInstantiate(&e, g, 0);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 20; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(a, s);
  for (int i = 0; i < reps; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(b, s); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) printf("  (error: %s)\n", cudaGetErrorString(err));
  return ms * 1e3f / reps / nodes;
}

int main() {
  int* d; cudaMalloc(&d, 4); cudaMemset(d, 0, 4);
  const int nodes = 13, reps = 2000;
  for (int grid : {8, 32, 148, 592}) for (int block : {128, 1024}) {
    float us = time_graph([&](cudaStream_t s) { k_plain<<<grid, block, 0, s>>>(d); }, nodes, reps);
    printf("plain   grid %4d block %4d : %.2f us per node\n", grid, block, us);
  }
  for (int block : {128, 1024}) {
    float us = time_graph([&](cudaStream_t s) { k_cluster<<<8, block, 0, s>>>(d); }, nodes, reps);
    printf("cluster grid    8 block %4d : %.2f us per node\n", block, us);
  }
  for (int grid : {32, 148}) {
    float us = time_graph([&](cudaStream_t s) {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = s;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1; cfg.attrs = at; cfg.numAttrs = 1;
      cudaLaunchKernelEx(&cfg, k_pdl, d);
    }, nodes, reps);
    printf("pdl     grid %4d block  256 : %.2f us per node\n", grid, us);
  }
  // one graph launch with a single node: fixed cost of the graph launch itself
  float us1 = time_graph([&](cudaStream_t s) { k_plain<<<32, 256, 0, s>>>(d); }, 1, reps);
  printf("single-node graph            : %.2f us per launch\n", us1);
  return 0;
}
