/*
 * fec_noalloc.cu — fec_cuda_rm_rx_noalloc: the standalone rate de-matcher (fec_cuda_rm_rx) for callers that must not
 * allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the
 * reference.
 *
 * Every byte of device scratch the call needs — the staged LLRs, the caller's soft buffer in the reference layout, the
 * device-layout slot, the per-K rank table, the launch descriptor and the max|soft| word — lives in ONE file-scope
 * __device__ array of FEC_NOALLOC_SCRATCH_BYTES (every GPU that loads the library holds it until exit). In host code the
 * array's name is only a host-side stand-in for the symbol: its device address is looked up once per device with
 * cudaGetSymbolAddress (that device current) and that pointer is what the kernels get. The entry point calls no
 * cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine; the kernels are the ones fec_cuda_rm_rx launches.
 *
 * All calls on one device share the array and it keeps the previous call's contents: the word that must start at zero is
 * cleared on the stream before the launch, the slot is rebuilt from the caller's soft buffer, the rank table is rewritten
 * when K changes. Calls are serialised on the host by a mutex and synchronous; work enqueued on the array by any other
 * route that overlaps a call on the GPU would race.
 */
#include "../../include/srslte_fec_cuda.h"
#include "fec_kernels.cuh"
#include <pthread.h>
#include <stdio.h>
#include <string.h>

#define NA_ALIGN(x)        (((x) + 255u) & ~(size_t)255u)
#define NA_OFF_IN          0
#define NA_OFF_REF         (NA_OFF_IN + NA_ALIGN((size_t)FEC_CUDA_NOALLOC_MAX_E * 2 + 16))
#define NA_OFF_SOFT        (NA_OFF_REF + NA_ALIGN((size_t)18600 * 2))
#define NA_OFF_RANK        (NA_OFF_SOFT + NA_ALIGN((size_t)FEC_CUDA_SOFT_SLOT * 2))
#define NA_OFF_DESC        (NA_OFF_RANK + NA_ALIGN((size_t)FEC_CUDA_SOFT_SLOT * 2))
#define NA_OFF_MAX         (NA_OFF_DESC + NA_ALIGN(sizeof(FecRmDesc)))
#define NA_SCRATCH_BYTES   (NA_OFF_MAX + 256)
#define NA_MAX_DEV         64

static __device__ __align__(256) unsigned char fec_noalloc_scratch[NA_SCRATCH_BYTES];

static pthread_mutex_t g_na_mu = PTHREAD_MUTEX_INITIALIZER;
static unsigned char  *g_na_ptr[NA_MAX_DEV];     /* device address of the array on each device, looked up on first use */
static int             g_na_K[NA_MAX_DEV];       /* code-block size and redundancy version (K * 4 + rv) whose table the array holds */
static uint32_t        g_na_off[NA_MAX_DEV][4];
/* host-side staging of the rank table and the descriptor (plain static memory, copied with cudaMemcpyAsync) */
static uint16_t        g_na_rank[FEC_CUDA_SOFT_SLOT], g_na_rank0[FEC_CUDA_SOFT_SLOT];
static FecRmDesc       g_na_desc;

static int na_fail(int code, const char *what, cudaError_t e)
{
  fprintf(stderr, "fec_cuda_rm_rx_noalloc: %s%s%s\n", what, e != cudaSuccess ? ": " : "", e != cudaSuccess ? cudaGetErrorString(e) : "");
  pthread_mutex_unlock(&g_na_mu);
  return code;
}
#define NA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return na_fail(FEC_CUDA_ENODEV, #call, e_); } while (0)

extern "C" int fec_cuda_rm_rx_noalloc(int device, const int16_t *in, uint32_t E, uint32_t cb_idx, uint32_t rv,
                                      int sb_layout, int16_t *soft)
{
  static_assert(NA_SCRATCH_BYTES <= FEC_CUDA_NOALLOC_SCRATCH_BYTES, "header states the scratch size");
  if (!in || !soft || rv > 3 || fec_k_size(cb_idx) < 0 || device < 0 || device >= NA_MAX_DEV) return FEC_CUDA_EINVAL;
  if (E > FEC_CUDA_NOALLOC_MAX_E) {
    fprintf(stderr, "fec_cuda_rm_rx_noalloc: E = %u LLRs is beyond the static scratch (FEC_CUDA_NOALLOC_MAX_E = %u); "
                    "use fec_cuda_rm_rx\n", E, (unsigned)FEC_CUDA_NOALLOC_MAX_E);
    return FEC_CUDA_EINVAL;
  }
  fec_geom_t g;
  if (fec_geom((uint32_t)fec_k_size(cb_idx), &g)) return FEC_CUDA_EINVAL;
  /* the extent of the caller's array the reference touches (rm_turbo.c:250-264,411-420) */
  const size_t next = (sb_layout && g.P) ? 3 * ((size_t)g.K + 32) + 12 : 3 * (size_t)g.K + 12;

  pthread_mutex_lock(&g_na_mu);
  NA(cudaSetDevice(device));
  if (!g_na_ptr[device]) {
    /* set-up, once per device, with that device current: the device address of the static array */
    unsigned char *p = NULL;
    NA(cudaGetSymbolAddress((void **)&p, fec_noalloc_scratch));
    if (!p) return na_fail(FEC_CUDA_ENODEV, "cudaGetSymbolAddress returned a null address", cudaSuccess);
    g_na_ptr[device] = p; g_na_K[device] = 0;
  }
  unsigned char *sc = g_na_ptr[device];
  int16_t  *d_in = (int16_t *)(sc + NA_OFF_IN), *d_ref = (int16_t *)(sc + NA_OFF_REF), *d_soft = (int16_t *)(sc + NA_OFF_SOFT);
  uint16_t *d_rank = (uint16_t *)(sc + NA_OFF_RANK);
  FecRmDesc *d_desc = (FecRmDesc *)(sc + NA_OFF_DESC);
  uint32_t *d_max = (uint32_t *)(sc + NA_OFF_MAX);
  cudaStream_t st = cudaStreamPerThread;

  if (g_na_K[device] != g.K * 4 + (int)rv) {
    g_na_K[device] = 0;
    fec_gen_rm_rank(&g, g_na_rank0, g_na_off[device]);
    fec_rm_rank_rebase(&g, g_na_rank0, g_na_off[device][rv], g_na_rank);        /* the table of this rv, as fec_cuda_rm_rx uses it */
    NA(cudaMemcpyAsync(d_rank, g_na_rank, (3 * (size_t)g.SS + 16) * 2, cudaMemcpyHostToDevice, st));
    NA(cudaStreamSynchronize(st));       /* g_na_rank is reusable, the table is in place */
    g_na_K[device] = g.K * 4 + (int)rv;
  }
  memset(&g_na_desc, 0, sizeof(g_na_desc));
  g_na_desc.e = d_in; g_na_desc.soft = d_soft; g_na_desc.rank = d_rank; g_na_desc.E = E; g_na_desc.N = 3 * (uint32_t)g.K + 12;
  g_na_desc.off = g_na_off[device][rv]; g_na_desc.n_out = 3 * (uint32_t)g.SS + 16; g_na_desc.new_data = 0; g_na_desc.slot = 0;

  if (E) NA(cudaMemcpyAsync(d_in, in, (size_t)E * 2, cudaMemcpyHostToDevice, st));
  NA(cudaMemcpyAsync(d_ref, soft, next * 2, cudaMemcpyHostToDevice, st));
  NA(cudaMemcpyAsync(d_desc, &g_na_desc, sizeof(FecRmDesc), cudaMemcpyHostToDevice, st));
  NA(cudaMemsetAsync(d_max, 0, 256, st));                       /* atomicMax target: must start at zero */
  /* reference layout -> device layout (zeroes the slot first), combine, and back; entries of the caller's array that no
   * coded bit maps to are never written (keep = 1), as in the reference */
  NA(fec_launch_ref_to_soft(d_ref, d_soft, g, sb_layout, st));
  NA(fec_launch_rm(d_desc, d_max, 1, g_na_desc.n_out, st, 0));
  NA(fec_launch_soft_to_ref(d_soft, d_ref, g, sb_layout, 1, st));
  NA(cudaMemcpyAsync(soft, d_ref, next * 2, cudaMemcpyDeviceToHost, st));
  NA(cudaStreamSynchronize(st));
  pthread_mutex_unlock(&g_na_mu);
  return 0;
}
