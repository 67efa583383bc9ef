// What can this host deliver? Pinned-memory H2D (and D2H) copy rate on N GPUs AT ONCE — the denominator of every end-to-end
// figure of bench.py (the LLR ingest is PCIe-bound by construction). One thread per GPU, 256 MB pinned buffers, cudaMemcpyAsync in
// a loop, all threads released together; prints per-GPU and aggregate GB/s.
//   nvcc -O2 -o tools/h2d_peak tools/h2d_peak.cu -lpthread && tools/h2d_peak [ngpus] [MB] [reps]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <pthread.h>

struct Arg { int dev; size_t bytes; int reps; pthread_barrier_t *bar; float ms_h2d, ms_d2h, ms_both; };

static void *worker(void *p)
{
  Arg *a = (Arg *)p;
  cudaSetDevice(a->dev);
  void *h = nullptr, *h2 = nullptr, *d = nullptr, *d2 = nullptr;
  cudaStream_t s1, s2;
  cudaEvent_t e0, e1;
  if (cudaHostAlloc(&h, a->bytes, cudaHostAllocDefault) != cudaSuccess || cudaHostAlloc(&h2, a->bytes, cudaHostAllocDefault) != cudaSuccess ||
      cudaMalloc(&d, a->bytes) != cudaSuccess || cudaMalloc(&d2, a->bytes) != cudaSuccess) { fprintf(stderr, "device %d: allocation failed\n", a->dev); exit(1); }
  cudaStreamCreate(&s1); cudaStreamCreate(&s2); cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int pass = 0; pass < 3; pass++) {
    cudaMemcpyAsync(d, h, a->bytes, cudaMemcpyHostToDevice, s1);            // warm
    cudaDeviceSynchronize();
    pthread_barrier_wait(a->bar);
    cudaEventRecord(e0, s1);
    if (pass == 2) cudaStreamWaitEvent(s2, e0, 0);
    for (int r = 0; r < a->reps; r++) {
      if (pass != 1) cudaMemcpyAsync(d, h, a->bytes, cudaMemcpyHostToDevice, s1);
      if (pass == 1) cudaMemcpyAsync(h2, d2, a->bytes, cudaMemcpyDeviceToHost, s1);
      if (pass == 2) cudaMemcpyAsync(h2, d2, a->bytes, cudaMemcpyDeviceToHost, s2);
    }
    if (pass == 2) { cudaEvent_t e2; cudaEventCreate(&e2); cudaEventRecord(e2, s2); cudaStreamWaitEvent(s1, e2, 0); }
    cudaEventRecord(e1, s1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    (pass == 0 ? a->ms_h2d : pass == 1 ? a->ms_d2h : a->ms_both) = ms;
    pthread_barrier_wait(a->bar);
  }
  return nullptr;
}

int main(int argc, char **argv)
{
  int n = 0;
  cudaGetDeviceCount(&n);
  const int ng = argc > 1 ? atoi(argv[1]) : n;
  const size_t bytes = (size_t)(argc > 2 ? atoi(argv[2]) : 256) << 20;
  const int reps = argc > 3 ? atoi(argv[3]) : 20;
  if (ng < 1 || ng > n) { fprintf(stderr, "%d GPUs asked, %d present\n", ng, n); return 1; }
  pthread_barrier_t bar;
  pthread_barrier_init(&bar, nullptr, ng);
  pthread_t th[64]; Arg a[64];
  for (int i = 0; i < ng; i++) { a[i] = Arg{i, bytes, reps, &bar, 0, 0, 0}; pthread_create(&th[i], nullptr, worker, &a[i]); }
  for (int i = 0; i < ng; i++) pthread_join(th[i], nullptr);
  double th2d = 0, td2h = 0, tbo = 0;
  printf("%d GPU(s) at once, %zu MB pinned buffers x %d copies each\n", ng, bytes >> 20, reps);
  for (int i = 0; i < ng; i++) {
    const double g = (double)bytes * reps / 1e6;
    printf("  gpu %d: H2D %.1f GB/s   D2H %.1f GB/s   both directions at once: %.1f + %.1f GB/s\n", i, g / a[i].ms_h2d, g / a[i].ms_d2h, g / a[i].ms_both, g / a[i].ms_both);
    th2d += g / a[i].ms_h2d; td2h += g / a[i].ms_d2h; tbo += g / a[i].ms_both;
  }
  printf("aggregate: H2D %.1f GB/s   D2H %.1f GB/s   bidirectional %.1f GB/s each way\n", th2d, td2h, tbo);
  return 0;
}
