/*
 * fec_cuda_api.cu — host side of libsrslte_fec_cuda: context, HARQ soft-buffer pool, lazy per-K tables,
 * and the batched decode entry points declared in include/srslte_fec_cuda.h. C-style code, C ABI.
 *
 * What the reference does per transport block on one core (decode_tb -> decode_tb_cb, sch.c:307-498)
 * becomes, per BATCH of transport blocks: 1 rate-dematching launch, 1 decode launch per distinct K,
 * 1 assembly/CRC24A launch, all on the context's stream. There is no CPU compute path.
 */
#include "../../include/srslte_fec_cuda.h"
#include "fec_kernels.cuh"
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <pthread.h>

#define CB_BYTES 768     /* 6144/8: per-CB decoded byte slot */

typedef struct {
  int        ready;
  fec_geom_t g;
  uint32_t  *d_step;
  uint16_t  *d_rank;
  uint16_t  *d_ranki[4];      /* the rank table re-based to the start of each redundancy version (16-bit rate de-matcher), built on first use */
  uint16_t  *d_ranki8[4];     /* the same with byte offsets into an int8 tile (8-bit rate de-matcher) */
  uint32_t   off[4];
  uint32_t   slots;
  uint32_t  *d_pair_addr;     /* packed kernel: decoder-2 address words */
  uint2     *d_pair_crc[2];   /* packed kernel: CRC syndrome words, CRC24A / CRC24B */
  uint16_t  *d_txsrc;         /* transmit side: circular-buffer position -> d-stream bit (built on the first encode of this K) */
} ktab_t;

typedef struct {
  uint32_t first_id, n;
  int16_t *d_soft;    /* n x 13 x FEC_CUDA_SOFT_SLOT */
  uint8_t *d_cache;   /* n x 13 x CB_BYTES */
  uint8_t *cb_crc;    /* host: n x 13 */
  struct fec_cuda_ctx *owner;   /* context whose destruction releases the chunk; NULL: pooled single buffers, process lifetime */
} sbchunk_t;

/* HARQ soft buffers belong to the DEVICE, not to a context: the reference allocates them in one thread (MAC) and decodes
 * into them from others (PHY workers, one srslte_sch_t each: SURVEY 8b "Threading"), so every context of a device must be
 * able to decode into any of them. One pool per device; ids are unique per device. Bulk allocations (n > 1) are owned by
 * the allocating context and die with it; single buffers (n == 1) come from pooled chunks of SB_POOL_N, are recycled
 * through a free list (fec_cuda_softbuffer_free) and live as long as the process. */
#define FEC_MAX_DEV 64
#define SB_POOL_N   32
typedef struct { int16_t *soft0; uint8_t *cache0; uint8_t *crc; } sbref_t;
typedef struct {
  pthread_rwlock_t lock;
  sbchunk_t *chunks; uint32_t nchunks, next_id;
  uint32_t  *free_ids; uint32_t nfree, free_cap;
} sbpool_t;
static sbpool_t g_pool[FEC_MAX_DEV];
static pthread_once_t g_pool_once = PTHREAD_ONCE_INIT;
static void pool_init_all(void) { for (int i = 0; i < FEC_MAX_DEV; i++) pthread_rwlock_init(&g_pool[i].lock, NULL); }

typedef struct { void *h, *d; size_t cap; } stage_t;   /* pinned host + device mirror, grown on demand */
#define FEC_NPIPE 8
#ifndef FEC_DIAG_SKIP_RM
#define FEC_DIAG_SKIP_RM 0
#endif
#ifndef FEC_DIAG_SKIP_ASM
#define FEC_DIAG_SKIP_ASM 0
#endif
/* timing diagnosis only (tools/experiments, never in the shipped build): a kernel is left out after the 12th call, the buffers
 * keep what the previous identical step left */
#define FEC_DIAG_ON(skip) (!(skip) || ctx->call_seq <= 12)
typedef struct fec_pipe fec_pipe_t;

struct fec_cuda_ctx {
  int           device, sm_count, grid_sms;   /* grid_sms: CTAs of a pipelined decoder launch (FEC_CUDA_GRID_DIV, experiment) */
  cudaStream_t  stream;
  char          err[256];
  uint64_t      launches;
  uint32_t      max_tbs;
  ktab_t        ktab[FEC_NOF_K];
  ktab_t        ktab8[FEC_NOF_K];   /* the same per-K tables in the geometry of the 8-bit LLR path (P = 32 for K > 2048) */
  ktab_t        ktabm[3][FEC_NOF_K];/* ... and in a forced 16-bit class (srslte_tdec_init_manual): non-windowed, 8 and 16 sub-blocks */
  int           force_p;            /* -1: the AUTO mapping; 0 / 8 / 16: that class for every code block size */
  int           llr8;               /* srslte_sch_t.llr_is_8bit: e_bits are int8, int8 combining and decoders (sch.c:344-348,362-364) */
  uint32_t     *d_r24a, *d_r24b;
  sbpool_t     *pool;               /* the soft-buffer pool of this context's device */
  stage_t       st_e, st_data, st_cbd, st_rmd, st_tbd, st_status, st_ret, st_exd;
  stage_t       st_txin, st_txout, st_txtb, st_txcb;   /* transmit side: payloads, packed output, descriptors */
  uint8_t      *d_txcrc; size_t txcrc_cap;
  uint8_t      *d_cbbytes; size_t cbbytes_cap;
  int16_t      *d_tmpsoft; size_t tmpsoft_cap;
  unsigned int *d_queue;   /* per K: work-queue counters of the packed and the exact kernel (2 x FEC_NOF_K), FEC_NOF_K hand-over counts to the
                              exact kernel, then work-queue counters and hand-over counts of the certificate kernel (FEC_NQUEUE x FEC_NOF_K) */
  uint32_t     *d_softmax; size_t softmax_cap;   /* per code block of the call: max |soft-buffer entry| */
  uint32_t     *d_exact;   size_t exact_cap;     /* per code block of the call: hand-over lists of the packed kernels + continuation words */
  uint32_t     *d_econt;   size_t econt_cap;     /* extrinsic arrays of code blocks handed over in mid-decode */
  int           force_exact;
  int16_t      *d_dbg; size_t dbg_cap; uint32_t *dbg_slot; uint32_t dbg_n;   /* test hook: LLR dump */
  float         phase_ms[3];   /* device time of rm | decode | assemble summed over the sub-batches of the last decode_tbs call */
  int           ev_valid;
  cudaStream_t  s_copy, s_out, s_rm, s_comp[2];
  int           ramp;           /* FEC_CUDA_RAMP=1: small first sub-batches (measured slower on B200, off by default) */
  fec_pipe_t   *pipes;
  uint32_t      sub_tbs, sub_tbs_dev;
  size_t        hw[16];         /* high-water marks of the per-pipe staging buffers */
  uint64_t      call_seq;       /* tickets of fec_cuda_submit_tbs */
  uint32_t      pipe_next;      /* next slot of the pipe ring (runs on across calls) */
  int           trace;          /* FEC_CUDA_TRACE=1: print the device timeline of every sub-batch to stderr */
  uint32_t     *d_x2masks, *d_x1states;   /* front end: scrambling-sequence jump tables (fec_frontend.cu) */
  float         fe_ms;          /* device time of the front-end kernels of the last call */
  uint64_t      path_cnt[2];    /* last call: code blocks handed to the certificate kernel / to the exact kernel */
  int           eager_c_ttl;    /* the same for the exact kernel behind both */
  int           eager_ttl;      /* > 0: recent sub-batches handed blocks to the certificate kernel: launch it right behind the first
                                   kernel instead of after the host has seen the hand-over counts (pipe_finish) */
  unsigned int *h_queue;        /* pinned mirror of d_queue (fec_cuda_decode_cbs) */
  cudaEvent_t   ev_call;        /* start of the current decode_tbs call (copy stream) */
  double        t_call;         /* host clock at the start of the call */
};

static int pipes_create(fec_cuda_ctx_t *ctx);
static int pipe_finish(fec_cuda_ctx_t *ctx, fec_pipe_t *pp);
static void pipes_drain(fec_cuda_ctx_t *ctx);
static double now_ms(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6; }
static void pipes_destroy(fec_cuda_ctx_t *ctx);

static int fail(fec_cuda_ctx_t *c, int code, const char *fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  if (c) vsnprintf(c->err, sizeof(c->err), fmt, ap);
  else vfprintf(stderr, fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return fail(ctx, FEC_CUDA_ENODEV, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)

static int stage_reserve(fec_cuda_ctx_t *ctx, stage_t *s, size_t bytes)
{
  if (bytes <= s->cap) return 0;
  size_t cap = bytes + bytes / 4 + 4096;
  if (s->h) cudaFreeHost(s->h);
  if (s->d) cudaFree(s->d);
  s->h = s->d = NULL; s->cap = 0;
  CU(cudaHostAlloc(&s->h, cap, cudaHostAllocDefault));
  CU(cudaMalloc(&s->d, cap));
  s->cap = cap;
  return 0;
}
static void stage_free(stage_t *s) { if (s->h) cudaFreeHost(s->h); if (s->d) cudaFree(s->d); memset(s, 0, sizeof(*s)); }

static int dev_reserve(fec_cuda_ctx_t *ctx, void **p, size_t *cap, size_t bytes)
{
  if (bytes <= *cap) return 0;
  if (*p) cudaFree(*p);
  *p = NULL; *cap = 0;
  size_t c = bytes + bytes / 4 + 4096;
  CU(cudaMalloc(p, c));
  *cap = c;
  return 0;
}

extern "C" int fec_cuda_cb_index(uint32_t K) { return fec_k_index(K); }
extern "C" int fec_cuda_cb_size(uint32_t idx) { return fec_k_size(idx); }
extern "C" int fec_cuda_subblocks(uint32_t K) { return fec_k_subblocks(K); }
extern "C" int fec_cuda_cbsegm(uint32_t tbs, uint32_t out9[9]) { return fec_segm(tbs, out9); }

extern "C" int fec_cuda_create(fec_cuda_ctx_t **out, int device, uint32_t max_tbs_per_call)
{
  if (!out) return FEC_CUDA_EINVAL;
  *out = NULL;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device || device < 0)
    return fail(NULL, FEC_CUDA_ENODEV, "libsrslte_fec_cuda: no usable CUDA device %d (there is no CPU fallback)\n", device);
  fec_cuda_ctx_t *ctx = (fec_cuda_ctx_t *)calloc(1, sizeof(*ctx));
  if (!ctx) return FEC_CUDA_ERROR;
  if (device >= FEC_MAX_DEV) { free(ctx); return FEC_CUDA_EINVAL; }
  pthread_once(&g_pool_once, pool_init_all);
  ctx->pool = &g_pool[device];
  ctx->device = device; ctx->max_tbs = max_tbs_per_call ? max_tbs_per_call : 1;
  ctx->force_p = -1;
  cudaDeviceProp prop;
  if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) { free(ctx); return FEC_CUDA_ENODEV; }
  if (prop.major < 10) { fprintf(stderr, "libsrslte_fec_cuda: built for sm_100a only, device is sm_%d%d\n", prop.major, prop.minor); free(ctx); return FEC_CUDA_ENODEV; }
  ctx->sm_count = prop.multiProcessorCount;
  { const char *gd = getenv("FEC_CUDA_GRID_DIV"); const int dv = gd ? atoi(gd) : 1; ctx->grid_sms = ctx->sm_count / (dv > 0 ? dv : 1); if (ctx->grid_sms < 1) ctx->grid_sms = 1; }
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { free(ctx); return FEC_CUDA_ENODEV; }
  cudaError_t e = fec_kernels_init();
  if (e != cudaSuccess) { fprintf(stderr, "libsrslte_fec_cuda: kernel image unusable: %s\n", cudaGetErrorString(e)); free(ctx); return FEC_CUDA_ENODEV; }
  /* CRC power tables (replace gen_crc_table, crc.c:38-47) */
  uint32_t *R = (uint32_t *)malloc(sizeof(uint32_t) * FEC_R24A_LEN);
  cudaMalloc(&ctx->d_r24a, sizeof(uint32_t) * FEC_R24A_LEN);
  cudaMalloc(&ctx->d_r24b, sizeof(uint32_t) * FEC_MAX_K);
  fec_gen_crc_powers(0x1864CFBu, R, FEC_R24A_LEN);
  cudaMemcpy(ctx->d_r24a, R, sizeof(uint32_t) * FEC_R24A_LEN, cudaMemcpyHostToDevice);
  fec_gen_crc_powers(0x1800063u, R, FEC_MAX_K);
  cudaMemcpy(ctx->d_r24b, R, sizeof(uint32_t) * FEC_MAX_K, cudaMemcpyHostToDevice);
  free(R);
  cudaMalloc(&ctx->d_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K);
  cudaMallocHost(&ctx->h_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K);
  {
    uint32_t *m = (uint32_t *)malloc(sizeof(uint32_t) * FEC_FE_NBLK * 33);
    fec_gen_gold_tables(m, m + FEC_FE_NBLK * 32);
    cudaMalloc(&ctx->d_x2masks, sizeof(uint32_t) * FEC_FE_NBLK * 32);
    cudaMalloc(&ctx->d_x1states, sizeof(uint32_t) * FEC_FE_NBLK);
    cudaMemcpy(ctx->d_x2masks, m, sizeof(uint32_t) * FEC_FE_NBLK * 32, cudaMemcpyHostToDevice);
    cudaMemcpy(ctx->d_x1states, m + FEC_FE_NBLK * 32, sizeof(uint32_t) * FEC_FE_NBLK, cudaMemcpyHostToDevice);
    free(m);
  }
  { const char *fe = getenv("FEC_CUDA_FORCE_EXACT"); ctx->force_exact = fe && fe[0] == '1'; }
  cudaStreamSynchronize(cudaStreamLegacy);       /* the table copies above are in place, not just staged */
  if (pipes_create(ctx)) { free(ctx); return FEC_CUDA_ENODEV; }
  if (cudaGetLastError() != cudaSuccess) { free(ctx); return FEC_CUDA_ENODEV; }
  *out = ctx;
  return FEC_CUDA_SUCCESS;
}

extern "C" void fec_cuda_destroy(fec_cuda_ctx_t *ctx)
{
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (int i = 0; i < 5 * FEC_NOF_K; i++) {
    ktab_t *kt = i < FEC_NOF_K ? &ctx->ktab[i] : i < 2 * FEC_NOF_K ? &ctx->ktab8[i - FEC_NOF_K] : &ctx->ktabm[i / FEC_NOF_K - 2][i % FEC_NOF_K];
    if (!kt->ready) continue;
    cudaFree(kt->d_step); cudaFree(kt->d_rank); cudaFree(kt->d_pair_addr); cudaFree(kt->d_txsrc);
    for (int r = 0; r < 4; r++) { cudaFree(kt->d_ranki[r]); cudaFree(kt->d_ranki8[r]); }
    cudaFree(kt->d_pair_crc[0]); cudaFree(kt->d_pair_crc[1]);
  }
  cudaFree(ctx->d_softmax); cudaFree(ctx->d_exact); cudaFree(ctx->d_econt); cudaFreeHost(ctx->h_queue);
  pipes_drain(ctx);                              /* nothing in flight may touch the chunks */
  {
    sbpool_t *pl = ctx->pool;
    pthread_rwlock_wrlock(&pl->lock);
    uint32_t w = 0;
    for (uint32_t i = 0; i < pl->nchunks; i++) {
      if (pl->chunks[i].owner == ctx) { cudaFree(pl->chunks[i].d_soft); cudaFree(pl->chunks[i].d_cache); free(pl->chunks[i].cb_crc); }
      else pl->chunks[w++] = pl->chunks[i];
    }
    pl->nchunks = w;
    pthread_rwlock_unlock(&pl->lock);
  }
  stage_free(&ctx->st_e); stage_free(&ctx->st_data); stage_free(&ctx->st_cbd); stage_free(&ctx->st_rmd);
  stage_free(&ctx->st_tbd); stage_free(&ctx->st_status); stage_free(&ctx->st_ret); stage_free(&ctx->st_exd);
  stage_free(&ctx->st_txin); stage_free(&ctx->st_txout); stage_free(&ctx->st_txtb); stage_free(&ctx->st_txcb); cudaFree(ctx->d_txcrc);
  cudaFree(ctx->d_cbbytes); cudaFree(ctx->d_tmpsoft); cudaFree(ctx->d_r24a); cudaFree(ctx->d_r24b); cudaFree(ctx->d_queue);
  cudaFree(ctx->d_x2masks); cudaFree(ctx->d_x1states);
  pipes_destroy(ctx);
  cudaStreamDestroy(ctx->stream);
  free(ctx);
}

/* srslte_sch_t.llr_is_8bit (sch.h:60): from here on e_bits are int8 and the 8-bit rate de-matcher and decoders run. Soft
 * buffers written in one mode cannot be continued in the other (as in the reference): reset them when switching. */
extern "C" int fec_cuda_set_llr_8bit(fec_cuda_ctx_t *ctx, int on)
{
  if (!ctx) return FEC_CUDA_EINVAL;
  ctx->llr8 = on ? 1 : 0;
  return 0;
}

/* srslte_tdec_init_manual (turbodecoder.c:153-190): one 16-bit arithmetic class for every code block size instead of the
 * AUTO mapping — P = 0 the non-windowed decoder (SRSLTE_TDEC_SSE), 8 (SRSLTE_TDEC_SSE_WINDOW), 16 (SRSLTE_TDEC_AVX_WINDOW);
 * -1 returns to AUTO. Forced classes run in the exact kernels. Applies to fec_cuda_decode_cbs (the srslte_tdec_* level). */
extern "C" int fec_cuda_set_decoder_class(fec_cuda_ctx_t *ctx, int P)
{
  if (!ctx || !(P == -1 || P == 0 || P == 8 || P == 16)) return FEC_CUDA_EINVAL;
  ctx->force_p = P;
  return 0;
}

static ktab_t *ktab_of(fec_cuda_ctx_t *ctx, int idx)
{
  if (ctx->llr8) return &ctx->ktab8[idx];
  if (ctx->force_p >= 0) return &ctx->ktabm[ctx->force_p == 0 ? 0 : ctx->force_p == 8 ? 1 : 2][idx];
  return &ctx->ktab[idx];
}

extern "C" const char *fec_cuda_last_error(fec_cuda_ctx_t *ctx) { return ctx ? ctx->err : "null context"; }
extern "C" void *fec_cuda_stream(fec_cuda_ctx_t *ctx) { return ctx ? (void *)ctx->stream : NULL; }
extern "C" uint64_t fec_cuda_launch_count(fec_cuda_ctx_t *ctx) { return ctx ? ctx->launches : 0; }

/* pinned host memory for LLR / output buffers (what srslte_vec_malloc hands out in the drop-in build) */
extern "C" void *fec_cuda_host_alloc(size_t bytes)
{
  void *p = NULL;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return NULL; }
  return p;
}
extern "C" void fec_cuda_host_free(void *p) { if (p) cudaFreeHost(p); }

static void path_count_add(fec_cuda_ctx_t *ctx, const unsigned int *q)
{
  for (int k = 0; k < FEC_NOF_K; k++) { ctx->path_cnt[0] += q[4 * FEC_NOF_K + k]; ctx->path_cnt[1] += q[2 * FEC_NOF_K + k]; }
}

extern "C" int fec_cuda_last_path_counts(fec_cuda_ctx_t *ctx, uint64_t out2[2])
{
  if (!ctx || !out2) return FEC_CUDA_EINVAL;
  out2[0] = ctx->path_cnt[0]; out2[1] = ctx->path_cnt[1];
  return 0;
}

extern "C" int fec_cuda_last_phase_ms(fec_cuda_ctx_t *ctx, float out3[3])
{
  if (!ctx || !out3 || !ctx->ev_valid) return FEC_CUDA_EINVAL;
  for (int i = 0; i < 3; i++) out3[i] = ctx->phase_ms[i];
  return 0;
}

static int ktab_get(fec_cuda_ctx_t *ctx, uint32_t K, ktab_t **out)
{
  int idx = fec_k_index(K);
  if (idx < 0 || fec_k_size(idx) != (int)K) return fail(ctx, FEC_CUDA_EINVAL, "K=%u is not an LTE code block size", K);
  ktab_t *t = ktab_of(ctx, idx);
  if (!t->ready) {
    if (ctx->llr8) fec_geom_p(K, fec_k_subblocks8(K), &t->g);
    else if (ctx->force_p >= 0) {
      /* a windowed class needs whole sub-blocks that cover the 40-step warm-up of turbodecoder_win.h:321-384 */
      if (ctx->force_p && (K % ctx->force_p || K / ctx->force_p <= 40)) return fail(ctx, FEC_CUDA_EINVAL, "K=%u cannot run in the forced class of %d sub-blocks", K, ctx->force_p);
      fec_geom_p(K, ctx->force_p, &t->g);
    } else fec_geom(K, &t->g);
    const size_t nstep = 2 * (size_t)t->g.SS, nrank = 3 * (size_t)t->g.SS + 16;
    uint32_t *step = (uint32_t *)malloc(nstep * 4);
    uint16_t *rank = (uint16_t *)malloc(nrank * 2);
    fec_gen_step_table(&t->g, step);
    fec_gen_rm_rank(&t->g, rank, t->off);
    CU(cudaMalloc(&t->d_step, nstep * 4));
    CU(cudaMalloc(&t->d_rank, nrank * 2));
    CU(cudaMemcpy(t->d_step, step, nstep * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(t->d_rank, rank, nrank * 2, cudaMemcpyHostToDevice));
    free(step); free(rank);
    if (t->g.P) {
      const size_t rows = (size_t)(t->g.P / 2) * t->g.Lp;
      uint32_t *pa = (uint32_t *)malloc(rows * 4), *pc = (uint32_t *)malloc(rows * 16);
      fec_gen_pair_addr_table(&t->g, pa);
      CU(cudaMalloc(&t->d_pair_addr, rows * 4));
      CU(cudaMemcpy(t->d_pair_addr, pa, rows * 4, cudaMemcpyHostToDevice));
      for (int c = 0; c < 2; c++) {
        fec_gen_pair_crc_table(&t->g, c ? 0x1800063u : 0x1864CFBu, pc);
        CU(cudaMalloc(&t->d_pair_crc[c], rows * 16));
        CU(cudaMemcpy(t->d_pair_crc[c], pc, rows * 16, cudaMemcpyHostToDevice));
      }
      free(pa); free(pc);
    }
    t->slots = t->g.P ? fec_win_pick_slots(&t->g, ctx->llr8 && t->g.P >= 16) : 0;
    /* cudaMemcpy from pageable memory returns when the data is STAGED, not when it is in place, and the kernels that read
     * these tables run on non-blocking streams, which the legacy stream does not order: wait for the DMA itself */
    CU(cudaStreamSynchronize(cudaStreamLegacy));
    t->ready = 1;
  }
  *out = t;
  return 0;
}

/* table the rate de-matcher gathers through for redundancy version rv: in 8-bit mode the rank table itself (its kernel
 * rotates while staging), else the table re-based to the first LLR of that rv (bulk-copy staging, fec_rm_rx_kernel) */
static int ktab_rank(fec_cuda_ctx_t *ctx, ktab_t *t, uint32_t rv, const uint16_t **out)
{
  uint16_t **slot = ctx->llr8 ? &t->d_ranki8[rv] : &t->d_ranki[rv];
  if (!*slot) {
    const size_t n = 3 * (size_t)t->g.SS + 16;
    uint16_t *rank = (uint16_t *)malloc(n * 2), *ri = (uint16_t *)malloc(n * 2);
    uint32_t off[4];
    if (!rank || !ri) { free(rank); free(ri); return fail(ctx, FEC_CUDA_ERROR, "out of host memory"); }
    fec_gen_rm_rank(&t->g, rank, off);
    fec_rm_rank_rebase(&t->g, rank, off[rv], ri);
    if (ctx->llr8) for (size_t i = 0; i < n; i++) ri[i] >>= 1;              /* one byte per LLR in the tile */
    cudaError_t e = cudaMalloc(slot, n * 2);
    if (e == cudaSuccess) e = cudaMemcpy(*slot, ri, n * 2, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);      /* in place before a non-blocking stream reads it (see ktab_get) */
    free(rank); free(ri);
    if (e != cudaSuccess) { cudaFree(*slot); *slot = NULL; return fail(ctx, FEC_CUDA_ENODEV, "rank table: %s", cudaGetErrorString(e)); }
  }
  *out = *slot;
  return 0;
}

/* ---------------------------------------------------------------- soft buffers ---- */
/* new chunk of n buffers; the pool's write lock is held by the caller */
static int pool_add_chunk(fec_cuda_ctx_t *ctx, sbpool_t *pl, uint32_t n, fec_cuda_ctx_t *owner, uint32_t *first_id)
{
  sbchunk_t c;
  memset(&c, 0, sizeof(c));
  c.first_id = pl->next_id; c.n = n; c.owner = owner;
  CU(cudaMalloc(&c.d_soft, (size_t)n * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t)));
  if (cudaMalloc(&c.d_cache, (size_t)n * FEC_CUDA_MAX_CB * CB_BYTES) != cudaSuccess) { cudaFree(c.d_soft); return fail(ctx, FEC_CUDA_ENODEV, "soft buffers: out of device memory"); }
  c.cb_crc = (uint8_t *)calloc((size_t)n * FEC_CUDA_MAX_CB, 1);
  sbchunk_t *nc = (sbchunk_t *)realloc(pl->chunks, sizeof(sbchunk_t) * (pl->nchunks + 1));
  if (!c.cb_crc || !nc) { cudaFree(c.d_soft); cudaFree(c.d_cache); free(c.cb_crc); if (nc) pl->chunks = nc; return fail(ctx, FEC_CUDA_ERROR, "soft buffers: out of host memory"); }
  pl->chunks = nc;
  CU(cudaMemsetAsync(c.d_soft, 0, (size_t)n * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t), ctx->stream));
  CU(cudaMemsetAsync(c.d_cache, 0, (size_t)n * FEC_CUDA_MAX_CB * CB_BYTES, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));      /* decode calls run on several streams */
  pl->chunks[pl->nchunks++] = c;
  pl->next_id += n;
  *first_id = c.first_id;
  return 0;
}

/* id -> device slot 0, decoded-CB cache and host cb_crc flags of that buffer (copied out under the read lock: the chunk
 * array may be reallocated by another thread's allocation, the memory it points to stays until the owner is destroyed) */
static int sb_lookup(fec_cuda_ctx_t *ctx, uint32_t id, sbref_t *out)
{
  sbpool_t *pl = ctx->pool;
  int found = 0;
  pthread_rwlock_rdlock(&pl->lock);
  for (uint32_t i = 0; i < pl->nchunks; i++) {
    const sbchunk_t *c = &pl->chunks[i];
    if (id >= c->first_id && id - c->first_id < c->n) {
      const size_t l = id - c->first_id;
      out->soft0 = c->d_soft + l * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT;
      out->cache0 = c->d_cache + l * FEC_CUDA_MAX_CB * CB_BYTES;
      out->crc = c->cb_crc + l * FEC_CUDA_MAX_CB;
      found = 1;
      break;
    }
  }
  pthread_rwlock_unlock(&pl->lock);
  return found;
}

static int sb_zero(fec_cuda_ctx_t *ctx, const sbref_t *r)
{
  memset(r->crc, 0, FEC_CUDA_MAX_CB);
  CU(cudaMemsetAsync(r->soft0, 0, (size_t)FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t), ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int fec_cuda_softbuffer_alloc(fec_cuda_ctx_t *ctx, uint32_t n, uint32_t *first_id)
{
  if (!ctx || !n || !first_id) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  sbpool_t *pl = ctx->pool;
  int rc = 0;
  pthread_rwlock_wrlock(&pl->lock);
  if (n > 1) rc = pool_add_chunk(ctx, pl, n, ctx, first_id);
  else {
    if (pl->nfree == 0) {
      /* a pooled chunk: one id for the caller, the rest on the free list (softbuffer.c:46-100 allocates per UE on attach;
       * one cudaMalloc pair per SB_POOL_N of them instead of per buffer) */
      uint32_t f0 = 0;
      rc = pool_add_chunk(ctx, pl, SB_POOL_N, NULL, &f0);
      if (!rc && pl->free_cap < pl->nfree + SB_POOL_N) {
        uint32_t *nf = (uint32_t *)realloc(pl->free_ids, sizeof(uint32_t) * (pl->nfree + SB_POOL_N + pl->free_cap));
        if (nf) { pl->free_ids = nf; pl->free_cap = pl->nfree + SB_POOL_N + pl->free_cap; } else rc = FEC_CUDA_ERROR;
      }
      for (uint32_t i = 0; !rc && i < SB_POOL_N; i++) pl->free_ids[pl->nfree++] = f0 + SB_POOL_N - 1 - i;
    }
    if (!rc) *first_id = pl->free_ids[--pl->nfree];
  }
  pthread_rwlock_unlock(&pl->lock);
  return rc;
}

/* srslte_softbuffer_rx_free (softbuffer.c:102-126): a single buffer (n == 1 allocation) goes back to the pool, zeroed for
 * its next user; buffers of a bulk allocation are released with their context, freeing one of them is a no-op */
extern "C" int fec_cuda_softbuffer_free(fec_cuda_ctx_t *ctx, uint32_t id)
{
  sbref_t r;
  if (!ctx || !sb_lookup(ctx, id, &r)) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  sbpool_t *pl = ctx->pool;
  int pooled = 0, rc = 0;
  pthread_rwlock_wrlock(&pl->lock);
  for (uint32_t i = 0; i < pl->nchunks; i++)
    if (id >= pl->chunks[i].first_id && id - pl->chunks[i].first_id < pl->chunks[i].n) pooled = pl->chunks[i].owner == NULL;
  for (uint32_t i = 0; pooled && i < pl->nfree; i++) if (pl->free_ids[i] == id) pooled = 0;       /* double free */
  if (pooled) {
    if (pl->nfree == pl->free_cap) {
      uint32_t *nf = (uint32_t *)realloc(pl->free_ids, sizeof(uint32_t) * (pl->free_cap + SB_POOL_N));
      if (nf) { pl->free_ids = nf; pl->free_cap += SB_POOL_N; } else rc = FEC_CUDA_ERROR;
    }
    if (!rc) rc = sb_zero(ctx, &r);
    if (!rc) pl->free_ids[pl->nfree++] = id;
  }
  pthread_rwlock_unlock(&pl->lock);
  return rc;
}

extern "C" int fec_cuda_softbuffer_reset(fec_cuda_ctx_t *ctx, uint32_t id)
{
  sbref_t r;
  if (!ctx || !sb_lookup(ctx, id, &r)) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  return sb_zero(ctx, &r);
}

extern "C" int fec_cuda_softbuffer_read_ref_layout(fec_cuda_ctx_t *ctx, uint32_t id, uint32_t cb, uint32_t K, int16_t *out)
{
  sbref_t r;
  ktab_t *t;
  if (!ctx || !sb_lookup(ctx, id, &r) || cb >= FEC_CUDA_MAX_CB || !out) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  if (ktab_get(ctx, K, &t)) return FEC_CUDA_EINVAL;
  if (dev_reserve(ctx, (void **)&ctx->d_tmpsoft, &ctx->tmpsoft_cap, 2 * FEC_CUDA_SOFT_SLOT * sizeof(int16_t))) return FEC_CUDA_ENODEV;
  CU(fec_launch_soft_to_ref(r.soft0 + (size_t)cb * FEC_CUDA_SOFT_SLOT, ctx->d_tmpsoft, t->g, 1, 0, ctx->stream));
  ctx->launches++;
  CU(cudaMemcpyAsync(out, ctx->d_tmpsoft, 18600 * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static void fill_decode_args(fec_cuda_ctx_t *ctx, int k, FecDecodeArgs *a, const FecCbDesc *cbd_dev, FecCbStatus *status_dev,
                             unsigned int *queue3, const uint32_t *softmax, uint32_t *exact, uint32_t *trk, uint32_t *econt, uint32_t off, uint32_t cnt, int16_t *dbg)
{
  const ktab_t *t = ktab_of(ctx, k);
  memset(a, 0, sizeof(*a));
  a->llr8 = ctx->llr8 && t->g.P >= 16;     /* classes 8 / 0 of the 8-bit path run the 16-bit decoders on the widened values */
  a->g = t->g; a->step_tab = t->d_step; a->crcpow[0] = ctx->d_r24a; a->crcpow[1] = ctx->d_r24b;
  a->cbs = cbd_dev + off; a->status = status_dev;
  a->queue = queue3 + 2 * k; a->ncb = cnt; a->slots = t->slots; a->dbg_llr = dbg;
  a->pair_addr = t->d_pair_addr; a->pair_crc[0] = t->d_pair_crc[0]; a->pair_crc[1] = t->d_pair_crc[1];
  a->soft_max = softmax; a->status_off = off;
  if (t->g.P && !ctx->force_exact && ctx->force_p < 0 && !dbg && !a->llr8) {
    a->exact_list = exact + off; a->exact_count = queue3 + 2 * FEC_NOF_K + k;
    a->trk_list = trk + off; a->trk_queue = queue3 + 3 * FEC_NOF_K + k; a->trk_count = queue3 + 4 * FEC_NOF_K + k;
    /* hand-over in mid-decode: continuation words behind the two lists, extrinsic scratch slots shared by the K groups of the
     * call in proportion to their size */
    a->trk_cont = trk + (trk - exact) + off;
    a->econt = econt ? econt + (size_t)off * FEC_ECONT_WORDS : NULL;
    a->econt_cap = econt ? cnt : 0;
  }
}

/* ---------------------------------------------------------------- batched transport-block decode ---- */
/* A call is cut into sub-batches that flow through a ring of FEC_NPIPE "pipes" (own pinned staging, device
 * scratch and events each): while the GPU works on sub-batch i (H2D of its LLRs on the copy stream, kernels on
 * one of two compute streams, D2H on the output stream) the host builds the descriptors of sub-batch i+1 and
 * unpacks the results of sub-batch i-1. The call returns when every sub-batch is done. */
typedef struct {
  uint32_t seg[9]; sbref_t sb; int valid, pinned; uint32_t ndec; size_t data_off, e_off; int kidx[2];
  /* front end (symbols in): offsets into the pipe's symbol / list / control-information staging */
  int fe_valid, sym_pinned; size_t sym_off, csi_off, lst_off, uci_off; uint32_t nack_u, nri_u, nuci, last_ri_pos;
} tbplan_t;

struct fec_pipe {
  stage_t       st_e, st_data, st_cbd, st_rmd, st_tbd, st_status, st_ret;
  stage_t       st_fed, st_lst, st_uci;            /* front end: descriptors, position lists, control-information LLRs */
  int16_t      *d_llr; size_t llr_cap;             /* front end: LLRs produced on the device (decode_tb's e_bits / g_bits) */
  float        *d_csimax; size_t csimax_cap;
  const fec_cuda_sym_t *syms;
  cudaEvent_t   ev_fe;
  uint8_t      *d_cbbytes; size_t cbbytes_cap;
  uint32_t     *d_softmax; size_t softmax_cap;
  uint32_t     *d_exact;   size_t exact_cap;
  uint32_t     *d_econt;   size_t econt_cap;
  unsigned int *d_queue, *h_queue;
  struct { int k; uint32_t off, cnt; } *grp; int ngrp, stages;   /* decode launches of this sub-batch and the stages already enqueued */
  cudaEvent_t   ev_in, ev_k, ev_done, ev[4];
  tbplan_t     *plan; size_t plan_cap;
  fec_cuda_tb_t *tbs; uint32_t n, ncb, ntb; int dev_io, busy;
  size_t        data_total;
  uint64_t      call_id;        /* ticket of the call this sub-batch belongs to */
  double        t_submit;       /* host time (ms since call start) when the sub-batch had been submitted */
};

static int pipe_init(fec_cuda_ctx_t *ctx, fec_pipe_t *p)
{
  memset(p, 0, sizeof(*p));
  CU(cudaMalloc(&p->d_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K));
  CU(cudaMallocHost(&p->h_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K));
  p->grp = (decltype(p->grp))malloc(sizeof(*p->grp) * FEC_NOF_K);
  if (!p->grp) return -1;
  CU(cudaEventCreate(&p->ev_in));
  CU(cudaEventCreateWithFlags(&p->ev_k, cudaEventDisableTiming));
  CU(cudaEventCreate(&p->ev_done));
  CU(cudaEventCreate(&p->ev_fe));
  for (int i = 0; i < 4; i++) CU(cudaEventCreate(&p->ev[i]));
  return 0;
}

static void pipe_free(fec_pipe_t *p)
{
  stage_free(&p->st_e); stage_free(&p->st_data); stage_free(&p->st_cbd); stage_free(&p->st_rmd);
  stage_free(&p->st_tbd); stage_free(&p->st_status); stage_free(&p->st_ret);
  stage_free(&p->st_fed); stage_free(&p->st_lst); stage_free(&p->st_uci);
  cudaFree(p->d_llr); cudaFree(p->d_csimax);
  if (p->ev_fe) cudaEventDestroy(p->ev_fe);
  cudaFree(p->d_cbbytes); cudaFree(p->d_softmax); cudaFree(p->d_exact); cudaFree(p->d_econt); cudaFree(p->d_queue); cudaFreeHost(p->h_queue);
  if (p->ev_in) { cudaEventDestroy(p->ev_in); cudaEventDestroy(p->ev_k); cudaEventDestroy(p->ev_done); for (int i = 0; i < 4; i++) cudaEventDestroy(p->ev[i]); }
  free(p->plan); free(p->grp);
  memset(p, 0, sizeof(*p));
}

/* staging and scratch of one pipe for the high-water marks need[] (see pipe_submit): entries that are 0 reserve nothing */
static int pipe_reserve(fec_cuda_ctx_t *ctx, fec_pipe_t *pp, const size_t *need)
{
  if (stage_reserve(ctx, &pp->st_cbd, need[0]) || stage_reserve(ctx, &pp->st_rmd, need[1]) ||
      stage_reserve(ctx, &pp->st_tbd, need[2]) || stage_reserve(ctx, &pp->st_status, need[3]) ||
      stage_reserve(ctx, &pp->st_ret, need[4]) ||
      dev_reserve(ctx, (void **)&pp->d_cbbytes, &pp->cbbytes_cap, need[5]) ||
      dev_reserve(ctx, (void **)&pp->d_softmax, &pp->softmax_cap, need[6]) ||
      dev_reserve(ctx, (void **)&pp->d_exact, &pp->exact_cap, 3 * need[6]) ||
      dev_reserve(ctx, (void **)&pp->d_econt, &pp->econt_cap, need[14]) ||
      stage_reserve(ctx, &pp->st_e, need[7]) || stage_reserve(ctx, &pp->st_data, need[8])) return -1;
  if (stage_reserve(ctx, &pp->st_fed, need[9]) || dev_reserve(ctx, (void **)&pp->d_llr, &pp->llr_cap, need[10]) ||
      stage_reserve(ctx, &pp->st_lst, need[11]) || stage_reserve(ctx, &pp->st_uci, need[12]) ||
      dev_reserve(ctx, (void **)&pp->d_csimax, &pp->csimax_cap, need[13])) return -1;
  return 0;
}

/* build descriptors for tbs[0..n), enqueue copies + kernels; returns without waiting */
/* ascending, duplicates removed; returns the new length */
static int cmp_u32(const void *a, const void *b) { const uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b; return x < y ? -1 : x > y; }
static uint32_t sort_unique(uint32_t *a, uint32_t n)
{
  if (!n) return 0;
  qsort(a, n, sizeof(uint32_t), cmp_u32);
  uint32_t m = 1;
  for (uint32_t i = 1; i < n; i++) if (a[i] != a[m - 1]) a[m++] = a[i];
  return m;
}

/* two PUSCH allocations with the same control-information layout (a batch is mostly runs of them: the positions depend on the
 * grant's size and the number of HARQ-ACK / RI bits only): the checks and the sorted position lists of the first serve the
 * second. Sorting 48 positions three times per transport block made the uplink leg of the bench host-bound (1.07 us per
 * transport block against 0.75 us of GPU time) */
static int fe_same_uci(const fec_cuda_sym_t *a, const fec_cuda_sym_t *b)
{
  if (a->mod_bits != b->mod_bits || a->nof_re != b->nof_re || a->nof_symb != b->nof_symb || a->nof_cqi_bits != b->nof_cqi_bits ||
      a->nof_ack_pos != b->nof_ack_pos || a->nof_ri_pos != b->nof_ri_pos || !a->nof_symb) return 0;
  if (a->nof_ack_pos && a->ack_pos != b->ack_pos && (!a->ack_pos || !b->ack_pos || memcmp(a->ack_pos, b->ack_pos, sizeof(uint32_t) * a->nof_ack_pos))) return 0;
  if (a->nof_ri_pos && a->ri_pos != b->ri_pos && (!a->ri_pos || !b->ri_pos || memcmp(a->ri_pos, b->ri_pos, sizeof(uint32_t) * a->nof_ri_pos))) return 0;
  return 1;
}

/* front-end argument checks for one transport block given as equalised symbols. prev / prev_nri: the last uplink allocation
 * that passed and its number of distinct RI positions (NULL: none) */
static int fe_check(const fec_cuda_tb_t *t, const fec_cuda_sym_t *sy, const fec_cuda_sym_t *prev = NULL, uint32_t *prev_nri = NULL)
{
  const uint32_t mod = sy->mod_bits;
  if (!sy->symbols || !sy->nof_re || !(mod == 1 || mod == 2 || mod == 4 || mod == 6)) return 0;
  if (((uintptr_t)sy->symbols & 15) || (sy->csi && ((uintptr_t)sy->csi & 3))) return 0;
  const uint64_t nb = (uint64_t)sy->nof_re * mod;
  if (nb > (uint64_t)FEC_FE_NBLK * 256) return 0;                       /* MAX_SEQ_LEN, sequence.c:38 */
  if (sy->nof_symb == 0) {                                              /* DL: pdsch.c:778-834 */
    if (sy->nof_ack_pos || sy->nof_ri_pos || sy->nof_cqi_bits) return 0;
    if (sy->csi && mod == 1) return 0;
    return t->tbs == 0 || t->nof_e_bits == nb;
  }
  /* UL: pusch.c:617-657 */
  if (sy->csi || mod == 1 || sy->nof_re % sy->nof_symb || sy->nof_symb > 14) return 0;      /* 12 / 10 columns (normal / extended CP), one less with SRS */
  if ((sy->nof_ack_pos && !sy->ack_pos) || (sy->nof_ri_pos && !sy->ri_pos)) return 0;
  if ((sy->nof_ack_pos || sy->nof_ri_pos || sy->nof_cqi_bits) && !sy->uci_out) return 0;
  uint32_t nri_u = 0;
  if (prev && prev_nri && fe_same_uci(sy, prev)) nri_u = *prev_nri;       /* positions already checked against the same nb */
  else {
    for (uint32_t i = 0; i < sy->nof_ack_pos; i++) if (sy->ack_pos[i] >= nb) return 0;
    for (uint32_t i = 0; i < sy->nof_ri_pos; i++) if (sy->ri_pos[i] >= nb) return 0;
    if (sy->nof_cqi_bits > nb) return 0;
    /* the de-interleaved stream is shorter than the q stream by the RI positions taken out of it (sch.c:860-881): decode_tb
     * must not be asked to read behind its end (the pipelined path does not zero the LLR scratch) */
    if (sy->nof_ri_pos) {
      uint32_t *tmp = (uint32_t *)malloc(sizeof(uint32_t) * sy->nof_ri_pos);
      if (!tmp) return 0;
      memcpy(tmp, sy->ri_pos, sizeof(uint32_t) * sy->nof_ri_pos);
      nri_u = sort_unique(tmp, sy->nof_ri_pos);
      free(tmp);
    }
    if (prev_nri) *prev_nri = nri_u;
  }
  if (t->tbs == 0) return 1;
  return (uint64_t)t->nof_e_bits + sy->nof_cqi_bits + nri_u <= nb;
}

/* UL position lists of one transport block: 2*(nack+nri) words at lst (host) / dl (its device mirror):
 * [ACK, de-interleaved order, ascending | RI, same | the caller's q positions, ACK then RI] */
static void fe_build_lists(const fec_cuda_sym_t *sy, FecFeDesc *f, uint32_t *lst, const uint32_t *dl)
{
  const uint32_t mod = sy->mod_bits, nuci = sy->nof_ack_pos + sy->nof_ri_pos;
  uint32_t *raw = lst + nuci, last = 0;
  for (uint32_t k = 0; k < nuci; k++) {
    /* q position -> index after the channel de-interleaver without RI removal (ulsch_interleave_gen, sch.c:550-568) */
    const uint32_t q = k < sy->nof_ack_pos ? sy->ack_pos[k] : sy->ri_pos[k - sy->nof_ack_pos];
    const uint32_t sidx = q / mod, col = sidx / f->rows, row = sidx - col * f->rows;
    lst[k] = (row * f->cols + col) * mod + (q - sidx * mod);
    raw[k] = q;
    if (k >= sy->nof_ack_pos && q > last) last = q;
  }
  f->nack = sort_unique(lst, sy->nof_ack_pos);
  f->nri = sort_unique(lst + sy->nof_ack_pos, sy->nof_ri_pos);
  f->last_ri_pos = last;
  f->ack_ord = dl; f->ri_ord = dl + sy->nof_ack_pos; f->uci_pos = dl + nuci;
}

static int pipe_submit(fec_cuda_ctx_t *ctx, fec_pipe_t *pp, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags,
                       cudaStream_t sc)
{
  const int dev_io = (flags & FEC_CUDA_F_DEVICE_IO) != 0;
  const size_t esz = ctx->llr8 ? 1 : 2;        /* bytes per LLR of e_bits */
  if (pp->plan_cap < n) { free(pp->plan); pp->plan = (tbplan_t *)malloc(sizeof(tbplan_t) * n); pp->plan_cap = n; }
  tbplan_t *plan = pp->plan;
  memset(plan, 0, sizeof(tbplan_t) * n);
  uint32_t grp_cnt[FEC_NOF_K], grp_off[FEC_NOF_K], grp_fill[FEC_NOF_K];
  memset(grp_cnt, 0, sizeof(grp_cnt));
  uint32_t ncb = 0, ntb = 0, nfe = 0, max_re = 0;
  int any_csi = 0, any_uci = 0, fe_kinds = 0;
  size_t e_total = 0, data_total = 0, sym_total = 0, lst_total = 0, uci_total = 0;
  pp->tbs = tbs; pp->syms = syms; pp->n = n; pp->dev_io = dev_io; pp->ncb = pp->ntb = 0; pp->busy = 0;

  /* pass 1: argument checks of decode_tb (sch.c:444-457) + segmentation + HARQ state */
  const fec_cuda_sym_t *chk_prev = NULL; uint32_t chk_nri = 0;
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
    t->ret = FEC_CUDA_EINVAL; t->avg_its = 0;
    memset(t->cb_crc, 0, sizeof(t->cb_crc)); memset(t->cb_its, 0, sizeof(t->cb_its));
    const fec_cuda_sym_t *sy = syms ? &syms[i] : NULL;
    if ((!sy && !t->e_bits) || (!t->data && t->tbs) || t->rv > 3 || t->Qm == 0) continue;
    if (sy) {
      uint32_t nri_now = chk_nri;
      if (!fe_check(t, sy, chk_prev, &nri_now) || (ctx->llr8 && sy->nof_symb)) continue;     /* 8-bit LLR mode: the downlink front end only */
      if (sy->nof_symb) { chk_prev = sy; chk_nri = nri_now; }
    }
    if (fec_segm(t->tbs, p->seg)) continue;
    const int decode = !(t->tbs == 0 || p->seg[1] == 0);
    if (decode) {
      if (p->seg[0] || p->seg[1] > FEC_CUDA_MAX_CB) continue;            /* filler bits / too many CBs */
      if (!sb_lookup(ctx, t->softbuffer, &p->sb)) continue;
    }
    if (sy) {
      /* the front end runs for every transport block whose symbols are usable, also for control-only PUSCH (tbs == 0) */
      p->fe_valid = 1; nfe++;
      fe_kinds |= sy->nof_symb ? 2 : 1;
      if (sy->nof_re > max_re) max_re = sy->nof_re;
      p->sym_off = sym_total; sym_total += ((size_t)sy->nof_re * 8 + 15) & ~(size_t)15;
      if (sy->csi) { any_csi = 1; p->csi_off = sym_total; sym_total += ((size_t)sy->nof_re * 4 + 15) & ~(size_t)15; }
      p->lst_off = lst_total; lst_total += 2 * ((size_t)sy->nof_ack_pos + sy->nof_ri_pos);
      p->nuci = sy->nof_ack_pos + sy->nof_ri_pos;
      p->uci_off = uci_total; uci_total += ((size_t)p->nuci + sy->nof_cqi_bits + 7) & ~(size_t)7;
      if (p->nuci + sy->nof_cqi_bits) any_uci = 1;
      p->e_off = e_total; e_total += ((size_t)sy->nof_re * sy->mod_bits + 15) & ~(size_t)15;    /* 16-byte lines also for int8 LLRs */
    }
    if (!decode) { t->ret = 0; continue; }
    uint8_t *crc = p->sb.crc;
    if (t->new_data) memset(crc, 0, FEC_CUDA_MAX_CB);
    p->kidx[0] = (int)p->seg[4]; p->kidx[1] = (int)p->seg[5];
    p->valid = 1; p->ndec = 0;
    for (uint32_t cb = 0; cb < p->seg[1]; cb++)
      if (!crc[cb]) { grp_cnt[p->kidx[cb < p->seg[6] ? 0 : 1]]++; p->ndec++; }
    ncb += p->ndec; ntb++;
    if (!sy) { p->e_off = e_total; e_total += ((size_t)t->nof_e_bits + 7) & ~(size_t)7; }
    p->data_off = data_total; data_total += ((size_t)t->tbs / 8 + 3 + CB_BYTES + 15) & ~(size_t)15;
  }
  if (ntb == 0 && nfe == 0) return 0;
  uint32_t acc = 0;
  for (int k = 0; k < FEC_NOF_K; k++) { grp_off[k] = acc; grp_fill[k] = 0; acc += grp_cnt[k]; }

  /* every pipe of the ring is sized for the largest sub-batch seen so far: (re)allocation of pinned / device staging
   * synchronises the device, so it must stop after the first calls */
  size_t need[16] = { sizeof(FecCbDesc) * (ncb + 1), sizeof(FecRmDesc) * (ncb + 1), sizeof(FecTbDesc) * (ntb + 1), sizeof(FecCbStatus) * (ncb + 1),
                      sizeof(int32_t) * (ntb + 1), (size_t)(ncb + 1) * CB_BYTES, (size_t)(ncb + 1) * 4,
                      dev_io ? 0 : (syms ? sym_total : e_total * esz), dev_io ? 0 : data_total,
                      /* front end */ sizeof(FecFeDesc) * nfe, syms ? e_total * esz : 0, lst_total * 4, (dev_io || !syms) ? 0 : uci_total * 2, sizeof(float) * nfe, 0, 0 };
  need[14] = (size_t)(ncb + 1) * FEC_ECONT_WORDS * 4;
  int grew = 0;
  for (int i = 0; i < 16; i++) { if (need[i] > ctx->hw[i]) { ctx->hw[i] = need[i]; grew = 1; } need[i] = ctx->hw[i]; }
  if (pipe_reserve(ctx, pp, need)) return FEC_CUDA_ENODEV;
  /* a new high-water mark: the idle pipes of the ring grow now too, so that a caller with one sub-batch per call does not
   * meet an allocation (a device synchronisation) in each of its first FEC_NPIPE calls */
  if (grew)
    for (int i = 0; i < FEC_NPIPE; i++)
      if (&ctx->pipes[i] != pp && !ctx->pipes[i].busy && pipe_reserve(ctx, &ctx->pipes[i], need)) return FEC_CUDA_ENODEV;

  FecCbDesc *cbd = (FecCbDesc *)pp->st_cbd.h;
  FecRmDesc *rmd = (FecRmDesc *)pp->st_rmd.h;
  FecTbDesc *tbd = (FecTbDesc *)pp->st_tbd.h;
  uint32_t max_n_out = 0, max_tile = 0, itb = 0;

  /* pass 1b: front-end descriptors (symbols in): staging of symbols / CSI, position lists in de-interleaved order */
  FecFeDesc *fed = (FecFeDesc *)pp->st_fed.h;
  const fec_cuda_sym_t *bl_prev = NULL; const FecFeDesc *bl_f = NULL; const uint32_t *bl_lst = NULL;
  for (uint32_t i = 0, ife = 0; syms && i < n; i++) {
    tbplan_t *p = &plan[i];
    if (!p->fe_valid) continue;
    const fec_cuda_sym_t *sy = &syms[i];
    FecFeDesc *f = &fed[ife++];
    memset(f, 0, sizeof(*f));
    const uint32_t mod = sy->mod_bits;
    if (dev_io) { f->sym = (const float2 *)sy->symbols; f->csi = sy->csi; f->uci_out = sy->uci_out; }
    else {
      f->sym = (const float2 *)((char *)pp->st_e.d + p->sym_off);
      f->csi = sy->csi ? (const float *)((char *)pp->st_e.d + p->csi_off) : NULL;
      f->uci_out = (p->nuci + sy->nof_cqi_bits) ? (int16_t *)pp->st_uci.d + p->uci_off : NULL;
      cudaPointerAttributes pa;
      p->sym_pinned = (cudaPointerGetAttributes(&pa, sy->symbols) == cudaSuccess && pa.type == cudaMemoryTypeHost);
      cudaGetLastError();
      if (!p->sym_pinned) memcpy((char *)pp->st_e.h + p->sym_off, sy->symbols, (size_t)sy->nof_re * 8);
      if (sy->csi) memcpy((char *)pp->st_e.h + p->csi_off, sy->csi, (size_t)sy->nof_re * 4);
    }
    f->e = (int16_t *)((char *)pp->d_llr + p->e_off * esz);     /* int8 entries in 8-bit LLR mode */
    f->nof_re = sy->nof_re; f->mod = mod; f->c_init = sy->c_init & 0x7FFFFFFFu;
    f->cols = sy->nof_symb; f->rows = sy->nof_symb ? sy->nof_re / sy->nof_symb : 0;
    f->n_cqi = sy->nof_cqi_bits; f->nuci = p->nuci;
    if (p->nuci) {
      uint32_t *lst = (uint32_t *)pp->st_lst.h + p->lst_off;
      const uint32_t *dl = (const uint32_t *)pp->st_lst.d + p->lst_off;
      if (bl_prev && fe_same_uci(sy, bl_prev)) {
        /* same layout as the previous allocation: its sorted lists are this one's */
        memcpy(lst, bl_lst, sizeof(uint32_t) * 2 * p->nuci);
        f->nack = bl_f->nack; f->nri = bl_f->nri; f->last_ri_pos = bl_f->last_ri_pos;
        f->ack_ord = dl; f->ri_ord = dl + sy->nof_ack_pos; f->uci_pos = dl + p->nuci;
      } else fe_build_lists(sy, f, lst, dl);
      bl_prev = sy; bl_f = f; bl_lst = lst;
    }
  }

  /* pass 2: descriptors. E / rp per code block exactly as decode_tb_cb derives them (sch.c:328-342) */
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
    if (!p->valid) continue;
    const uint32_t C = p->seg[1], C1 = p->seg[6];
    uint8_t *crc = p->sb.crc;
    int16_t *soft0 = p->sb.soft0;
    uint8_t *cache0 = p->sb.cache0;
    const char *e_dev = syms ? (const char *)pp->d_llr + (p->e_off + syms[i].nof_cqi_bits) * esz   /* g_bits + Q'_cqi*Qm, sch.c:963-973 */
                             : dev_io ? (const char *)t->e_bits : (const char *)pp->st_e.d + p->e_off * esz;
    FecTbDesc *td = &tbd[itb];
    memset(td, 0, sizeof(*td));
    td->data = dev_io ? t->data : (uint8_t *)pp->st_data.d + p->data_off;
    td->C = C; td->tbs = t->tbs; td->all_ok = 1; td->ret = (int32_t *)pp->st_ret.d + itb;
    const uint32_t Gp = t->nof_e_bits / t->Qm, gamma = Gp % C, n_e = t->Qm * (Gp / C);
    ktab_t *kts[2] = {NULL, NULL};
    const uint16_t *rks[2] = {NULL, NULL};
    for (uint32_t cb = 0; cb < C; cb++) {
      const int which = cb < C1 ? 0 : 1;
      const uint32_t K = which ? p->seg[3] : p->seg[2];
      td->K[cb] = K; td->rlen[cb] = C == 1 ? K : K - 24; td->cache[cb] = cache0 + (size_t)cb * CB_BYTES;
      if (crc[cb]) { td->st_idx[cb] = -1; td->cb_bytes[cb] = td->cache[cb]; continue; }
      if (!kts[which] && (ktab_get(ctx, K, &kts[which]) || ktab_rank(ctx, kts[which], t->rv, &rks[which]))) return FEC_CUDA_EINVAL;
      ktab_t *kt = kts[which];
      uint32_t rp = cb * n_e, E = n_e;
      if (cb > C - gamma) { E = n_e + t->Qm; rp = (C - gamma) * n_e + (cb - (C - gamma)) * E; }   /* sch.c:339 off-by-one kept */
      const uint32_t slot = grp_off[kt->g.idx] + grp_fill[kt->g.idx]++;
      FecRmDesc *r = &rmd[slot];
      r->e = (const int16_t *)(e_dev + (size_t)rp * esz); r->soft = soft0 + (size_t)cb * FEC_CUDA_SOFT_SLOT; r->rank = rks[which];
      r->E = E; r->N = 3 * K + 12; r->off = kt->off[t->rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = t->new_data ? 1 : 0; r->slot = slot;
      if (r->n_out > max_n_out) max_n_out = r->n_out;
      if ((E < r->N ? E : r->N) + 1 > max_tile) max_tile = (E < r->N ? E : r->N) + 1;
      FecCbDesc *d = &cbd[slot];
      d->soft = r->soft; d->out = pp->d_cbbytes + (size_t)slot * CB_BYTES; d->crc_sel = C > 1 ? 1 : 0;
      d->max_its = t->max_its ? t->max_its : 1; d->flags = (flags & FEC_CUDA_F_NO_EARLY_STOP) ? FEC_F_NO_EARLY : 0; d->status = slot;
      td->st_idx[cb] = (int32_t)slot; td->cb_bytes[cb] = d->out;
    }
    if (!dev_io && !syms) {
      /* pinned caller memory (fec_cuda_host_alloc / cudaHostRegister) is DMA'd straight from where it lies */
      cudaPointerAttributes pa;
      p->pinned = (cudaPointerGetAttributes(&pa, t->e_bits) == cudaSuccess && pa.type == cudaMemoryTypeHost);
      cudaGetLastError();
      if (!p->pinned) memcpy((char *)pp->st_e.h + p->e_off * esz, t->e_bits, (size_t)t->nof_e_bits * esz);
    }
    itb++;
  }

  /* ---- copy stream: LLRs (one copy per run of TBs contiguous on both sides) + descriptors ---- */
  cudaStream_t sin = ctx->s_copy, sout = ctx->s_out;
  if (syms) {
    if (!dev_io) {
      /* symbols (+ CSI): pinned caller memory goes by DMA from where it lies, the rest through the pinned staging; runs of
       * transport blocks that are contiguous on both sides travel as one copy */
      uint32_t i = 0;
      while (i < n) {
        if (!plan[i].fe_valid) { i++; continue; }
        if (!plan[i].sym_pinned) {
          uint32_t j = i;
          size_t lo = plan[i].sym_off, hi = lo;
          for (; j < n && (!plan[j].fe_valid || !plan[j].sym_pinned); j++)
            if (plan[j].fe_valid) hi = (syms[j].csi ? plan[j].csi_off + (size_t)syms[j].nof_re * 4 : plan[j].sym_off + (size_t)syms[j].nof_re * 8);
          CU(cudaMemcpyAsync((char *)pp->st_e.d + lo, (char *)pp->st_e.h + lo, hi - lo, cudaMemcpyHostToDevice, sin));
          i = j;
        } else {
          const char *src0 = (const char *)syms[i].symbols;
          const size_t dst0 = plan[i].sym_off;
          size_t len = (size_t)syms[i].nof_re * 8;
          uint32_t j = i + 1;
          if (!syms[i].csi)
            for (; j < n; j++) {
              if (!plan[j].fe_valid || !plan[j].sym_pinned || syms[j].csi || plan[j].sym_off != dst0 + len ||
                  (const char *)syms[j].symbols != src0 + len || (len & 15)) break;
              len += (size_t)syms[j].nof_re * 8;
            }
          CU(cudaMemcpyAsync((char *)pp->st_e.d + dst0, src0, len, cudaMemcpyHostToDevice, sin));
          if (syms[i].csi) CU(cudaMemcpyAsync((char *)pp->st_e.d + plan[i].csi_off, (char *)pp->st_e.h + plan[i].csi_off,
                                              (size_t)syms[i].nof_re * 4, cudaMemcpyHostToDevice, sin));
          i = j;
        }
      }
    }
    if (lst_total) CU(cudaMemcpyAsync(pp->st_lst.d, pp->st_lst.h, lst_total * 4, cudaMemcpyHostToDevice, sin));
    CU(cudaMemcpyAsync(pp->st_fed.d, fed, sizeof(FecFeDesc) * nfe, cudaMemcpyHostToDevice, sin));
  } else if (!dev_io) {
    uint32_t i = 0;
    while (i < n) {
      if (!plan[i].valid) { i++; continue; }
      const int pinned = plan[i].pinned;
      const char *src0 = pinned ? (const char *)tbs[i].e_bits : (const char *)pp->st_e.h + plan[i].e_off * esz;
      const size_t dst0 = plan[i].e_off * esz;
      size_t len = (size_t)tbs[i].nof_e_bits * esz;
      uint32_t j = i + 1;
      for (; j < n; j++) {
        if (!plan[j].valid || plan[j].pinned != pinned || plan[j].e_off * esz != dst0 + len) break;
        if (pinned && (const char *)tbs[j].e_bits != src0 + len) break;
        len += (size_t)tbs[j].nof_e_bits * esz;
      }
      CU(cudaMemcpyAsync((char *)pp->st_e.d + dst0, src0, len, cudaMemcpyHostToDevice, sin));
      i = j;
    }
  }
  if (ncb) CU(cudaMemcpyAsync(pp->st_cbd.d, cbd, sizeof(FecCbDesc) * ncb, cudaMemcpyHostToDevice, sin));
  if (ncb) CU(cudaMemcpyAsync(pp->st_rmd.d, rmd, sizeof(FecRmDesc) * ncb, cudaMemcpyHostToDevice, sin));
  if (ntb) CU(cudaMemcpyAsync(pp->st_tbd.d, tbd, sizeof(FecTbDesc) * ntb, cudaMemcpyHostToDevice, sin));
  CU(cudaEventRecord(pp->ev_in, sin));

  /* ---- compute stream: rate de-matching, turbo decode per K group, TB assembly + CRC24A ---- */
  /* rate de-matching has its own stream: it depends on nothing but its inputs, so it must not queue up behind the
   * decode kernel of an earlier sub-batch (which can hold a compute stream for the length of its slowest code block) */
  cudaStream_t srm = (flags & FEC_CUDA_F_SERIAL) ? sc : ctx->s_rm;
  CU(cudaMemsetAsync(pp->d_queue, 0, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K, srm));
  if (ncb) CU(cudaMemsetAsync(pp->d_softmax, 0, (size_t)ncb * 4, srm));
  /* status entries of code blocks that are handed over stay unwritten until the deferred pass: they must not carry the
   * verdicts of the pipe's previous sub-batch into the first assembly (which refreshes the decoded-CB cache) */
  if (ncb) CU(cudaMemsetAsync(pp->st_status.d, 0, sizeof(FecCbStatus) * ncb, srm));
  CU(cudaStreamWaitEvent(srm, pp->ev_in, 0));
  CU(cudaEventRecord(pp->ev_fe, srm));
  if (nfe) {
    /* symbols -> LLRs in decode_tb order: soft demodulation, descrambling, CSI weights / UL channel de-interleaver */
    CU(fec_launch_frontend((const FecFeDesc *)pp->st_fed.d, pp->d_csimax, ctx->d_x2masks, ctx->d_x1states, nfe, max_re, any_csi, any_uci, fe_kinds, srm, ctx->llr8));
    ctx->launches += ctx->llr8 ? 1 + (any_csi ? 1 : 0) : (fe_kinds == 3 ? 2 : 1) + (any_csi ? 1 : 0) + (any_uci ? 1 : 0);
  }
  CU(cudaEventRecord(pp->ev[0], srm));
  if (ncb && FEC_DIAG_ON(FEC_DIAG_SKIP_RM)) { CU(fec_launch_rm((const FecRmDesc *)pp->st_rmd.d, pp->d_softmax, ncb, max_n_out, srm, ctx->llr8, max_tile)); ctx->launches++; }
  CU(cudaEventRecord(pp->ev[1], srm));
  if (srm != sc) CU(cudaStreamWaitEvent(sc, pp->ev[1], 0));
  /* the certificate kernel follows the first kernel at once only while recent sub-batches needed it; otherwise pipe_finish
   * launches it when it sees hand-over counts. Measured on the bench workload (nothing handed over, tools/variants.sh A/B
   * on one box): first + exact kernel 1.69 ms per step, first + certificate + exact 1.78 ms, first alone 1.72-1.78 ms — a
   * third persistent full-shared-memory grid between the sub-batches of the two compute streams costs 4 % even when it
   * finds nothing to do. Round 2 (CTA timeline, tools/experiments/timeline.py): the empty exact-kernel launch is no better — its
   * 148 CTAs ask for a whole SM's shared memory each, so on a busy GPU they run only when the decoder of the OTHER stream
   * starts to drain, and the next decoder launch of this stream, queued behind them and the assembly, came 0.2-0.45 ms late
   * to fill that tail. Both follow-up kernels are now launched only while recent sub-batches needed them */
  pp->stages = FEC_ST_A | (ctx->eager_ttl > 0 ? FEC_ST_B : 0) | (ctx->eager_c_ttl > 0 ? FEC_ST_C : 0);
  pp->ngrp = 0;
  for (int k = 0; k < FEC_NOF_K; k++) {
    if (!grp_cnt[k]) continue;
    FecDecodeArgs a;
    fill_decode_args(ctx, k, &a, (const FecCbDesc *)pp->st_cbd.d, (FecCbStatus *)pp->st_status.d, pp->d_queue, pp->d_softmax, pp->d_exact, pp->d_exact + pp->exact_cap / 12, pp->d_econt,
                     grp_off[k], grp_cnt[k], NULL);
    CU(fec_launch_decode(a, ctx->grid_sms, sc, pp->stages));
    ctx->launches += a.exact_list ? 1 + ((pp->stages & FEC_ST_B) != 0) + ((pp->stages & FEC_ST_C) != 0) : 1;
    if (a.exact_list) { pp->grp[pp->ngrp].k = k; pp->grp[pp->ngrp].off = grp_off[k]; pp->grp[pp->ngrp].cnt = grp_cnt[k]; pp->ngrp++; }
  }
  CU(cudaEventRecord(pp->ev[2], sc));
  if (ntb && FEC_DIAG_ON(FEC_DIAG_SKIP_ASM)) { CU(fec_launch_tb_assemble((const FecTbDesc *)pp->st_tbd.d, (const FecCbStatus *)pp->st_status.d, ctx->d_r24a, ntb, sc)); ctx->launches++; }
  CU(cudaEventRecord(pp->ev[3], sc));
  CU(cudaEventRecord(pp->ev_k, sc));

  /* ---- output stream: results back ---- */
  CU(cudaStreamWaitEvent(sout, pp->ev_k, 0));
  if (ncb) CU(cudaMemcpyAsync(pp->st_status.h, pp->st_status.d, sizeof(FecCbStatus) * ncb, cudaMemcpyDeviceToHost, sout));
  if (ntb) CU(cudaMemcpyAsync(pp->st_ret.h, pp->st_ret.d, sizeof(int32_t) * ntb, cudaMemcpyDeviceToHost, sout));
  if (ncb) CU(cudaMemcpyAsync(pp->h_queue, pp->d_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K, cudaMemcpyDeviceToHost, sout));
  if (!dev_io && data_total) CU(cudaMemcpyAsync(pp->st_data.h, pp->st_data.d, data_total, cudaMemcpyDeviceToHost, sout));
  if (!dev_io && syms && uci_total) CU(cudaMemcpyAsync(pp->st_uci.h, pp->st_uci.d, uci_total * 2, cudaMemcpyDeviceToHost, sout));
  CU(cudaEventRecord(pp->ev_done, sout));
  pp->ncb = ncb; pp->ntb = ntb; pp->data_total = data_total; pp->busy = 1;
  pp->t_submit = now_ms() - ctx->t_call;
  return 0;
}

/* wait for a submitted sub-batch and unpack its results + HARQ state (softbuffer->cb_crc, sch.c:384,404-408;
 * nof_iterations sch.c:420) */
static int pipe_finish(fec_cuda_ctx_t *ctx, fec_pipe_t *pp)
{
  if (!pp->busy) return 0;
  pp->busy = 0;
  CU(cudaEventSynchronize(pp->ev_done));
  for (int i = 0; i < 3; i++) { float ms = 0; if (cudaEventElapsedTime(&ms, pp->ev[i], pp->ev[i + 1]) == cudaSuccess) ctx->phase_ms[i] += ms; }
  { float ms = 0; if (cudaEventElapsedTime(&ms, pp->ev_fe, pp->ev[0]) == cudaSuccess) ctx->fe_ms += ms; }
  if (pp->ncb && pp->ngrp) {
    /* hand-over counts of the kernels launched so far: run the stages that were left out and turn out to be needed */
    uint64_t ntrk = 0, nexact = 0;
    for (int g = 0; g < pp->ngrp; g++) { ntrk += pp->h_queue[4 * FEC_NOF_K + pp->grp[g].k]; nexact += pp->h_queue[2 * FEC_NOF_K + pp->grp[g].k]; }
    const int did_b = (pp->stages & FEC_ST_B) != 0, did_c = (pp->stages & FEC_ST_C) != 0;
    const int need_b = ntrk && !did_b, need_c = need_b || (nexact && !did_c);
    if (need_c) {
      /* certificate kernel now (if it was left out and is needed), then the exact kernel for what neither packed variant
       * covers. If the exact kernel already ran in the first pass, that run is complete: its work-queue counter and its list
       * start again from zero; if it was left out, its list holds the first kernel's hand-overs and the certificate kernel
       * appends to it */
      cudaStream_t sc = ctx->s_comp[0];
      if (did_c) ctx->path_cnt[1] += nexact;
      CU(cudaEventRecord(pp->ev[0], sc));                  /* the events of the first pass have been read above */
      for (int g = 0; g < pp->ngrp; g++) {
        const int k = pp->grp[g].k;
        if (did_c) {
          CU(cudaMemsetAsync(pp->d_queue + 2 * k + 1, 0, sizeof(unsigned int), sc));
          CU(cudaMemsetAsync(pp->d_queue + 2 * FEC_NOF_K + k, 0, sizeof(unsigned int), sc));
        }
        FecDecodeArgs a;
        fill_decode_args(ctx, k, &a, (const FecCbDesc *)pp->st_cbd.d, (FecCbStatus *)pp->st_status.d, pp->d_queue, pp->d_softmax, pp->d_exact,
                         pp->d_exact + pp->exact_cap / 12, pp->d_econt, pp->grp[g].off, pp->grp[g].cnt, NULL);
        CU(fec_launch_decode(a, ctx->sm_count, sc, (need_b ? FEC_ST_B : 0) | FEC_ST_C));
        ctx->launches += 1 + need_b;
      }
      CU(cudaEventRecord(pp->ev[1], sc));
      if (pp->ntb) { CU(fec_launch_tb_assemble((const FecTbDesc *)pp->st_tbd.d, (const FecCbStatus *)pp->st_status.d, ctx->d_r24a, pp->ntb, sc)); ctx->launches++; }
      CU(cudaEventRecord(pp->ev[2], sc));
      CU(cudaMemcpyAsync(pp->st_status.h, pp->st_status.d, sizeof(FecCbStatus) * pp->ncb, cudaMemcpyDeviceToHost, sc));
      if (pp->ntb) CU(cudaMemcpyAsync(pp->st_ret.h, pp->st_ret.d, sizeof(int32_t) * pp->ntb, cudaMemcpyDeviceToHost, sc));
      CU(cudaMemcpyAsync(pp->h_queue, pp->d_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K, cudaMemcpyDeviceToHost, sc));
      if (!pp->dev_io && pp->data_total) CU(cudaMemcpyAsync(pp->st_data.h, pp->st_data.d, pp->data_total, cudaMemcpyDeviceToHost, sc));
      CU(cudaStreamSynchronize(sc));
      /* the deferred certificate / exact kernels and the second assembly belong to the phases of this call too */
      { float ms = 0; if (cudaEventElapsedTime(&ms, pp->ev[0], pp->ev[1]) == cudaSuccess) ctx->phase_ms[1] += ms;
        if (cudaEventElapsedTime(&ms, pp->ev[1], pp->ev[2]) == cudaSuccess) ctx->phase_ms[2] += ms; }
    }
    /* launch policy of the next sub-batches */
    if (ntrk) ctx->eager_ttl = 16; else if (ctx->eager_ttl > 0) ctx->eager_ttl--;
    if (nexact) ctx->eager_c_ttl = 16; else if (ctx->eager_c_ttl > 0) ctx->eager_c_ttl--;
  }
  if (pp->ncb) path_count_add(ctx, pp->h_queue);
  if (pp->syms && !pp->dev_io)
    for (uint32_t i = 0; i < pp->n; i++) {
      const tbplan_t *p = &pp->plan[i];
      const uint32_t nout = p->fe_valid ? p->nuci + pp->syms[i].nof_cqi_bits : 0;
      if (nout) memcpy(pp->syms[i].uci_out, (int16_t *)pp->st_uci.h + p->uci_off, (size_t)nout * 2);
    }
  if (ctx->trace) {
    float t[6] = {0, 0, 0, 0, 0, 0};
    cudaEventElapsedTime(&t[0], ctx->ev_call, pp->ev_in);
    for (int i = 0; i < 4; i++) cudaEventElapsedTime(&t[1 + i], ctx->ev_call, pp->ev[i]);
    cudaEventElapsedTime(&t[5], ctx->ev_call, pp->ev_done);
    fprintf(stderr, "[fec trace] sub-batch %u TBs %u CBs: submitted@%.3f h2d_done %.3f | rm %.3f-%.3f | dec -%.3f | asm -%.3f | d2h_done %.3f | host now %.3f ms\n",
            pp->n, pp->ncb, pp->t_submit, t[0], t[1], t[2], t[3], t[4], t[5], now_ms() - ctx->t_call);
  }
  const FecCbStatus *stt = (const FecCbStatus *)pp->st_status.h;
  const FecTbDesc *tbd = (const FecTbDesc *)pp->st_tbd.h;
  uint32_t itb = 0;
  for (uint32_t i = 0; i < pp->n; i++) {
    fec_cuda_tb_t *t = &pp->tbs[i];
    tbplan_t *p = &pp->plan[i];
    if (!p->valid) continue;
    const FecTbDesc *td = &tbd[itb];
    uint8_t *crc = p->sb.crc;
    uint32_t sum = 0;
    for (uint32_t cb = 0; cb < p->seg[1]; cb++) {
      if (td->st_idx[cb] >= 0) {
        const FecCbStatus *s = &stt[td->st_idx[cb]];
        t->cb_its[cb] = (uint8_t)s->its; sum += s->its;
        if (s->crc_ok) crc[cb] = 1;
      }
      t->cb_crc[cb] = crc[cb];
    }
    t->avg_its = sum / p->seg[1];
    t->ret = ((const int32_t *)pp->st_ret.h)[itb];
    if (!pp->dev_io) memcpy(t->data, (uint8_t *)pp->st_data.h + p->data_off, (size_t)t->tbs / 8 + 3);
    itb++;
  }
  return 0;
}

static int pipes_create(fec_cuda_ctx_t *ctx)
{
  ctx->pipes = (fec_pipe_t *)calloc(FEC_NPIPE, sizeof(fec_pipe_t));
  if (!ctx->pipes) return -1;
  for (int i = 0; i < FEC_NPIPE; i++) if (pipe_init(ctx, &ctx->pipes[i])) return -1;
  if (cudaStreamCreateWithFlags(&ctx->s_copy, cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->s_comp[1], cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->s_rm, cudaStreamNonBlocking) != cudaSuccess) return -1;
  ctx->s_comp[0] = ctx->stream;
  const char *sb = getenv("FEC_CUDA_SUB_TBS");
  ctx->sub_tbs = sb ? (uint32_t)atoi(sb) : 296;
  if (ctx->sub_tbs == 0) ctx->sub_tbs = 296;
  sb = getenv("FEC_CUDA_SUB_TBS_DEV");
  ctx->sub_tbs_dev = sb ? (uint32_t)atoi(sb) : 0;          /* 0: two sub-batches per call, at least 592 transport blocks each */
  sb = getenv("FEC_CUDA_RAMP");
  ctx->ramp = sb ? atoi(sb) : 0;
  sb = getenv("FEC_CUDA_TRACE");
  ctx->trace = sb ? atoi(sb) : 0;
  CU(cudaEventCreate(&ctx->ev_call));
  return 0;
}

static void pipes_drain(fec_cuda_ctx_t *ctx)
{
  for (int i = 0; i < FEC_NPIPE && ctx->pipes; i++) if (ctx->pipes[i].busy) pipe_finish(ctx, &ctx->pipes[i]);
}

static void pipes_destroy(fec_cuda_ctx_t *ctx)
{
  if (!ctx->pipes) return;
  for (int i = 0; i < FEC_NPIPE; i++) pipe_free(&ctx->pipes[i]);
  free(ctx->pipes); ctx->pipes = NULL;
  if (ctx->s_copy) cudaStreamDestroy(ctx->s_copy);
  if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
  if (ctx->s_comp[1]) cudaStreamDestroy(ctx->s_comp[1]);
  if (ctx->s_rm) cudaStreamDestroy(ctx->s_rm);
}

static int submit_common(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags, uint64_t *ticket)
{
  if (!ctx || (!tbs && n)) return FEC_CUDA_EINVAL;
  const uint64_t id = ++ctx->call_seq;
  if (ticket) *ticket = id;
  if (n == 0) return 0;
  cudaSetDevice(ctx->device);
  ctx->phase_ms[0] = ctx->phase_ms[1] = ctx->phase_ms[2] = 0; ctx->fe_ms = 0; ctx->ev_valid = 1;
  ctx->path_cnt[0] = ctx->path_cnt[1] = 0;
  ctx->t_call = now_ms();
  if (ctx->trace) CU(cudaEventRecord(ctx->ev_call, ctx->s_copy));
  /* host buffers: small sub-batches so that the PCIe copy of one overlaps the kernels of the previous (the copy is
   * the longer of the two); device buffers: nothing to overlap but host-side descriptor work, so few, large ones */
  /* device buffers: a decoder launch ends with a tail in which each CTA waits for the slowest of its 16 code blocks (about
   * 0.2 ms of SM time whatever the launch size; the next launch fills the SMs as they are released, but the slots of a
   * draining CTA idle), so launches are large: two sub-batches per call of at least 1184 transport blocks each — two rather
   * than one so that a single blocking call still overlaps the rate de-matching of its second half with the decoder of the
   * first. Measured, headline shape, Gbit/s with two calls in flight at 1184 / 4736 transport blocks per call: sub-batches
   * of 296: 55.6 / -; 592: 61.5 / 63.6; 1184: 64.5 / 67.5; 2368: - / 69.0; 4736: - / 68.9 */
  /* host buffers: sub-batches are sized by what crosses the link (FEC_CUDA_SUB_TBS transport blocks of the headline shape,
   * 90 000 int16 LLRs each = 53 MB by default): int8 LLRs are half the bytes per transport block, 64QAM symbols two thirds */
  uint32_t sub = ctx->sub_tbs;
  if (!(flags & FEC_CUDA_F_DEVICE_IO)) {
    const uint64_t per_tb = syms ? (uint64_t)syms[0].nof_re * (syms[0].csi ? 12 : 8) : (uint64_t)tbs[0].nof_e_bits * (ctx->llr8 ? 1 : 2);
    if (per_tb) {
      const uint64_t want = (uint64_t)ctx->sub_tbs * 180000u / per_tb;
      sub = (uint32_t)(want < 64 ? 64 : want > 8192 ? 8192 : want);
    }
  }
  if (flags & FEC_CUDA_F_DEVICE_IO) { sub = ctx->sub_tbs_dev ? ctx->sub_tbs_dev : (n + 1) / 2; if (!ctx->sub_tbs_dev && sub < 1184) sub = 1184; }
  int rc = 0;
  uint32_t s = 0, off = 0;
  while (off < n && !rc) {
    /* ramp: the first sub-batches are small so that the GPU starts after a short descriptor build and a short rate
     * de-matching launch; everything after that is hidden behind the kernels of the previous sub-batch */
    uint32_t cnt = (flags & FEC_CUDA_F_SERIAL) ? n : !ctx->ramp ? sub : (s == 0 ? sub / 4 : s == 1 ? sub / 2 : sub);
    if (cnt == 0) cnt = 1;
    if (n - off < cnt + cnt / 2) cnt = n - off;          /* no tiny last sub-batch */
    const uint32_t slot = ctx->pipe_next++;
    fec_pipe_t *pp = &ctx->pipes[slot % FEC_NPIPE];
    rc = pipe_finish(ctx, pp);                           /* ring full: the oldest sub-batch (maybe of an earlier call) completes first */
    if (!rc) { pp->call_id = id; rc = pipe_submit(ctx, pp, tbs + off, syms ? syms + off : NULL, cnt, flags, ctx->s_comp[slot & 1]); }
    off += cnt; s++;
  }
  return rc;
}

extern "C" int fec_cuda_submit_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags, uint64_t *ticket)
{
  return submit_common(ctx, tbs, NULL, n, flags, ticket);
}

/* same pipeline fed with equalised symbols: the front-end kernel produces decode_tb's LLRs on the device */
extern "C" int fec_cuda_submit_syms(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags, uint64_t *ticket)
{
  if (!syms && n) return FEC_CUDA_EINVAL;
  return submit_common(ctx, tbs, syms, n, flags, ticket);
}

extern "C" int fec_cuda_decode_syms(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags)
{
  uint64_t ticket = 0;
  const int rc = fec_cuda_submit_syms(ctx, tbs, syms, n, flags, &ticket);
  const int r2 = fec_cuda_wait(ctx, ticket);
  return rc ? rc : r2;
}

extern "C" int fec_cuda_last_frontend_ms(fec_cuda_ctx_t *ctx, float *ms)
{
  if (!ctx || !ms || !ctx->ev_valid) return FEC_CUDA_EINVAL;
  *ms = ctx->fe_ms;
  return 0;
}

extern "C" int fec_cuda_wait(fec_cuda_ctx_t *ctx, uint64_t ticket)
{
  if (!ctx) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  int rc = 0;
  for (uint32_t i = 0; i < FEC_NPIPE; i++) {             /* oldest first */
    fec_pipe_t *pp = &ctx->pipes[(ctx->pipe_next + i) % FEC_NPIPE];
    if (pp->busy && pp->call_id <= ticket) { const int r2 = pipe_finish(ctx, pp); if (!rc) rc = r2; }
  }
  return rc;
}

extern "C" int fec_cuda_decode_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags)
{
  uint64_t ticket = 0;
  const int rc = fec_cuda_submit_tbs(ctx, tbs, n, flags, &ticket);
  const int r2 = fec_cuda_wait(ctx, ticket);
  return rc ? rc : r2;
}

/* ---------------------------------------------------------------- bare code blocks (srslte_tdec_* level) ---- */
#define CUF(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { free(plan); \
  return fail(ctx, FEC_CUDA_ENODEV, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
#define DBG_LLR ctx->d_dbg
extern "C" int fec_cuda_decode_cbs(fec_cuda_ctx_t *ctx, fec_cuda_cb_t *cbs, uint32_t n, uint32_t flags)
{
  if (!ctx || (!cbs && n)) return FEC_CUDA_EINVAL;
  if (n == 0) return 0;
  cudaSetDevice(ctx->device);
  uint32_t grp_cnt[FEC_NOF_K], grp_off[FEC_NOF_K], grp_fill[FEC_NOF_K];
  memset(grp_cnt, 0, sizeof(grp_cnt));
  size_t llr_total = 0;
  for (uint32_t i = 0; i < n; i++) {
    ktab_t *kt;
    if (!cbs[i].llr || !cbs[i].out || ktab_get(ctx, cbs[i].K, &kt)) return FEC_CUDA_EINVAL;
    if (ctx->llr8) {                           /* the reference's 8-bit entry points take int8_t: anything wider is a caller error */
      const int16_t *v = cbs[i].llr;
      for (uint32_t t = 0; t < 3 * cbs[i].K + 12; t++)
        if (v[t] < -128 || v[t] > 127) return fail(ctx, FEC_CUDA_EINVAL, "8-bit LLR mode: code block %u has an LLR outside int8", i);
    }
    grp_cnt[kt->g.idx]++;
    llr_total += ((size_t)3 * cbs[i].K + 12 + 7) & ~(size_t)7;
  }
  uint32_t acc = 0;
  for (int k = 0; k < FEC_NOF_K; k++) { grp_off[k] = acc; grp_fill[k] = 0; acc += grp_cnt[k]; }
  void *plan = NULL;
  if (stage_reserve(ctx, &ctx->st_e, llr_total * 2) || stage_reserve(ctx, &ctx->st_cbd, sizeof(FecCbDesc) * n) ||
      stage_reserve(ctx, &ctx->st_exd, sizeof(FecExtractDesc) * n) || stage_reserve(ctx, &ctx->st_status, sizeof(FecCbStatus) * n) ||
      stage_reserve(ctx, &ctx->st_data, (size_t)n * CB_BYTES) ||
      dev_reserve(ctx, (void **)&ctx->d_softmax, &ctx->softmax_cap, (size_t)(n + 1) * 4) ||
      dev_reserve(ctx, (void **)&ctx->d_exact, &ctx->exact_cap, (size_t)(n + 1) * 12) ||
      dev_reserve(ctx, (void **)&ctx->d_econt, &ctx->econt_cap, (size_t)(n + 1) * FEC_ECONT_WORDS * 4) ||
      dev_reserve(ctx, (void **)&ctx->d_tmpsoft, &ctx->tmpsoft_cap, (size_t)n * FEC_CUDA_SOFT_SLOT * sizeof(int16_t))) return FEC_CUDA_ENODEV;
  if (ctx->dbg_cap && ctx->dbg_cap < (size_t)n * FEC_MAX_K * 2) { cudaFree(ctx->d_dbg); ctx->d_dbg = NULL; ctx->dbg_cap = 0; }
  FecCbDesc *cbd = (FecCbDesc *)ctx->st_cbd.h;
  FecExtractDesc *exd = (FecExtractDesc *)ctx->st_exd.h;
  uint32_t *slot_of = (uint32_t *)malloc(sizeof(uint32_t) * n);
  size_t eo = 0;
  for (uint32_t i = 0; i < n; i++) {
    ktab_t *kt;
    ktab_get(ctx, cbs[i].K, &kt);
    const uint32_t slot = grp_off[kt->g.idx] + grp_fill[kt->g.idx]++;
    slot_of[i] = slot;
    const size_t len = (size_t)3 * cbs[i].K + 12;
    memcpy((int16_t *)ctx->st_e.h + eo, cbs[i].llr, len * 2);
    exd[slot].in = (const int16_t *)ctx->st_e.d + eo; exd[slot].soft = ctx->d_tmpsoft + (size_t)slot * FEC_CUDA_SOFT_SLOT; exd[slot].g = kt->g; exd[slot].slot = slot;
    eo += (len + 7) & ~(size_t)7;
    cbd[slot].soft = exd[slot].soft; cbd[slot].out = (uint8_t *)ctx->st_data.d + (size_t)slot * CB_BYTES;
    cbd[slot].crc_sel = cbs[i].crc_type ? 1 : 0; cbd[slot].max_its = cbs[i].max_its ? cbs[i].max_its : 1;
    cbd[slot].flags = (flags & FEC_CUDA_F_NO_EARLY_STOP) ? FEC_F_NO_EARLY : 0; cbd[slot].status = slot;
  }
  plan = slot_of;
  cudaStream_t st = ctx->stream;
  CUF(cudaMemcpyAsync(ctx->st_e.d, ctx->st_e.h, llr_total * 2, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_cbd.d, cbd, sizeof(FecCbDesc) * n, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_exd.d, exd, sizeof(FecExtractDesc) * n, cudaMemcpyHostToDevice, st));
  CUF(cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K, st));
  CUF(cudaMemsetAsync(ctx->d_softmax, 0, (size_t)n * 4, st));
  for (uint32_t o = 0; o < n; o += 32768) {
    CUF(fec_launch_extract((const FecExtractDesc *)ctx->st_exd.d + o, ctx->d_softmax, n - o < 32768 ? n - o : 32768, st));
    ctx->launches++;
  }
  for (int k = 0; k < FEC_NOF_K; k++) {
    if (!grp_cnt[k]) continue;
    FecDecodeArgs a;
    fill_decode_args(ctx, k, &a, (const FecCbDesc *)ctx->st_cbd.d, (FecCbStatus *)ctx->st_status.d, ctx->d_queue, ctx->d_softmax, ctx->d_exact, ctx->d_exact + ctx->exact_cap / 12, ctx->d_econt,
                     grp_off[k], grp_cnt[k], DBG_LLR);
    CUF(fec_launch_decode(a, ctx->sm_count, st, FEC_ST_ALL));
    ctx->launches += a.exact_list ? 3 : 1;
  }
  CUF(cudaMemcpyAsync(ctx->st_status.h, ctx->st_status.d, sizeof(FecCbStatus) * n, cudaMemcpyDeviceToHost, st));
  CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, (size_t)n * CB_BYTES, cudaMemcpyDeviceToHost, st));
  CUF(cudaMemcpyAsync(ctx->h_queue, ctx->d_queue, sizeof(unsigned int) * FEC_NQUEUE * FEC_NOF_K, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  ctx->path_cnt[0] = ctx->path_cnt[1] = 0;
  path_count_add(ctx, ctx->h_queue);
  for (uint32_t i = 0; i < n; i++) {
    const FecCbStatus *s = (const FecCbStatus *)ctx->st_status.h + slot_of[i];
    cbs[i].its = s->its; cbs[i].crc_ok = s->crc_ok;
    memcpy(cbs[i].out, (uint8_t *)ctx->st_data.h + (size_t)slot_of[i] * CB_BYTES, cbs[i].K / 8);
  }
  free(ctx->dbg_slot); ctx->dbg_slot = slot_of; ctx->dbg_n = n;
  return 0;
}

/* ---------------------------------------------------------------- transmit side ---- */
/* encode_tb_off for n transport blocks (sch.c:188-296): CRC-24A, segmentation, CRC-24B, turbo encoding, rate matching.
 * Synchronous. Stateless: a retransmission is encoded from the data again (the reference reads the circular buffer it
 * kept in srslte_softbuffer_tx_t; the bits are the same). */
extern "C" int fec_cuda_encode_tbs(fec_cuda_ctx_t *ctx, fec_cuda_txtb_t *tbs, uint32_t n, uint32_t flags)
{
  if (!ctx || (!tbs && n)) return FEC_CUDA_EINVAL;
  if (n == 0) return 0;
  cudaSetDevice(ctx->device);
  const int dev_io = (flags & FEC_CUDA_F_DEVICE_IO) != 0;
  uint32_t ntb = 0, ncb = 0;
  size_t in_total = 0, out_total = 0;
  uint32_t (*seg)[9] = (uint32_t (*)[9])malloc(sizeof(uint32_t) * 9 * n);
  size_t *offs = (size_t *)malloc(sizeof(size_t) * 2 * n);
  if (!seg || !offs) { free(seg); free(offs); return FEC_CUDA_ERROR; }
  void *plan = seg;
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_txtb_t *t = &tbs[i];
    t->ret = FEC_CUDA_EINVAL;
    if (!t->data || !t->e_bits || t->rv > 3 || t->Qm == 0 || (t->tbs & 7) || fec_segm(t->tbs, seg[i])) continue;
    if (dev_io && ((uintptr_t)t->e_bits & 3)) continue;
    if (seg[i][0] || seg[i][1] > FEC_CUDA_MAX_CB) { t->ret = FEC_CUDA_ERROR; continue; }      /* filler bits / too many CBs: sch.c:202-210 */
    t->ret = 1;                                                                                   /* marks "planned" until the end */
    if (seg[i][1] == 0) continue;
    offs[2 * i] = in_total; in_total += ((size_t)t->tbs / 8 + 15) & ~(size_t)15;
    offs[2 * i + 1] = out_total; out_total += (((size_t)t->nof_e_bits + 7) / 8 + 15) & ~(size_t)15;
    ntb++; ncb += seg[i][1];
  }
  int rc = 0;
  if (ntb) {
    if (stage_reserve(ctx, &ctx->st_txtb, sizeof(FecTxTbDesc) * ntb) || stage_reserve(ctx, &ctx->st_txcb, sizeof(FecTxCbDesc) * ncb) ||
        dev_reserve(ctx, (void **)&ctx->d_txcrc, &ctx->txcrc_cap, (size_t)ntb * 4) ||
        (!dev_io && (stage_reserve(ctx, &ctx->st_txin, in_total) || stage_reserve(ctx, &ctx->st_txout, out_total)))) { free(plan); free(offs); return FEC_CUDA_ENODEV; }
    FecTxTbDesc *tbd = (FecTxTbDesc *)ctx->st_txtb.h;
    FecTxCbDesc *cbd = (FecTxCbDesc *)ctx->st_txcb.h;
    cudaStream_t st = ctx->stream;
    uint32_t itb = 0, icb = 0;
    for (uint32_t i = 0; i < n && !rc; i++) {
      fec_cuda_txtb_t *t = &tbs[i];
      if (t->ret != 1 || seg[i][1] == 0) continue;
      const uint32_t C = seg[i][1], C2 = seg[i][7];
      const uint8_t *data_dev = dev_io ? t->data : (const uint8_t *)ctx->st_txin.d + offs[2 * i];
      uint8_t *e_dev = dev_io ? t->e_bits : (uint8_t *)ctx->st_txout.d + offs[2 * i + 1];
      if (!dev_io) memcpy((uint8_t *)ctx->st_txin.h + offs[2 * i], t->data, t->tbs / 8);
      tbd[itb].data = data_dev; tbd[itb].tbs = t->tbs; tbd[itb].crc3 = ctx->d_txcrc + 4 * (size_t)itb;
      tbd[itb].e_out = (uint32_t *)e_dev; tbd[itb].e_words = (t->nof_e_bits + 31) / 32;
      const uint32_t Gp = t->nof_e_bits / t->Qm, gamma = Gp % C;
      uint32_t rp = 0, wp = 0;
      for (uint32_t cb = 0; cb < C; cb++) {
        const uint32_t K = cb < C2 ? seg[i][3] : seg[i][2], rlen = C > 1 ? K - 24 : K;
        const uint32_t n_e = (cb + gamma + 1 <= C) ? t->Qm * (Gp / C) : t->Qm * ((Gp + C - 1) / C);         /* sch.c:241-245 */
        ktab_t *kt;
        if (ktab_get(ctx, K, &kt)) { rc = FEC_CUDA_EINVAL; break; }
        if (!kt->d_txsrc) {
          const size_t N = 3 * (size_t)K + 12;
          uint16_t *src = (uint16_t *)malloc(N * 2);
          fec_gen_rm_txsrc(&kt->g, 32 * FEC_TX_W, src);
          if (cudaMalloc(&kt->d_txsrc, N * 2) != cudaSuccess || cudaMemcpy(kt->d_txsrc, src, N * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
              cudaStreamSynchronize(cudaStreamLegacy) != cudaSuccess) rc = FEC_CUDA_ENODEV;
          free(src);
          if (rc) break;
        }
        FecTxCbDesc *d = &cbd[icb++];
        memset(d, 0, sizeof(*d));
        const int last = cb == C - 1;
        d->src = data_dev + rp / 8; d->tbcrc = tbd[itb].crc3; d->e_out = e_dev; d->txsrc = kt->d_txsrc;
        d->nsrc = rlen / 8 - (last ? 3 : 0); d->ntb = last ? 3 : 0; d->cbcrc = C > 1;
        d->K = K; d->f1 = (uint32_t)kt->g.f1; d->f2 = (uint32_t)kt->g.f2;
        d->wp = wp; d->E = n_e; d->off = kt->off[t->rv]; d->N = 3 * K + 12;
        rp += rlen; wp += n_e;
      }
      itb++;
    }
    if (!rc) {
      cudaError_t e = cudaSuccess;
      if (!dev_io) {
        e = cudaMemcpyAsync(ctx->st_txin.d, ctx->st_txin.h, in_total, cudaMemcpyHostToDevice, st);
      }
      if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->st_txtb.d, tbd, sizeof(FecTxTbDesc) * ntb, cudaMemcpyHostToDevice, st);
      if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->st_txcb.d, cbd, sizeof(FecTxCbDesc) * icb, cudaMemcpyHostToDevice, st);
      if (e == cudaSuccess) e = fec_launch_tx((const FecTxTbDesc *)ctx->st_txtb.d, ntb, (const FecTxCbDesc *)ctx->st_txcb.d, icb, ctx->d_r24a, ctx->d_r24b, st);
      ctx->launches += 2;
      if (e == cudaSuccess && !dev_io) e = cudaMemcpyAsync(ctx->st_txout.h, ctx->st_txout.d, out_total, cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) rc = fail(ctx, FEC_CUDA_ERROR, "encode: %s", cudaGetErrorString(e));
    }
  }
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_txtb_t *t = &tbs[i];
    if (t->ret != 1) continue;
    if (rc) { t->ret = FEC_CUDA_ERROR; continue; }
    if (seg[i][1] == 0) { if (!dev_io) memset(t->e_bits, 0, ((size_t)t->nof_e_bits + 7) / 8); t->ret = 0; continue; }
    if (!dev_io) memcpy(t->e_bits, (uint8_t *)ctx->st_txout.h + offs[2 * i + 1], ((size_t)t->nof_e_bits + 7) / 8);
    t->ret = 0;
  }
  free(plan); free(offs);
  return rc;
}

/* ---------------------------------------------------------------- standalone rate de-matching ---- */
extern "C" int fec_cuda_rm_rx(fec_cuda_ctx_t *ctx, const int16_t *in, uint32_t E, uint32_t cb_idx, uint32_t rv,
                              int sb_layout, int16_t *soft)
{
  if (!ctx || !in || !soft || rv > 3 || fec_k_size(cb_idx) < 0) return FEC_CUDA_EINVAL;   /* rm_turbo.c:395,426-429 */
  cudaSetDevice(ctx->device);
  ktab_t *kt;
  const uint16_t *rk;
  if (ktab_get(ctx, (uint32_t)fec_k_size(cb_idx), &kt) || ktab_rank(ctx, kt, rv, &rk)) return FEC_CUDA_EINVAL;
  void *plan = NULL;
  if (stage_reserve(ctx, &ctx->st_e, ((size_t)E + 8) * 2) || stage_reserve(ctx, &ctx->st_rmd, sizeof(FecRmDesc)) ||
      stage_reserve(ctx, &ctx->st_data, 18600 * 2) ||
      dev_reserve(ctx, (void **)&ctx->d_softmax, &ctx->softmax_cap, 8) ||
      dev_reserve(ctx, (void **)&ctx->d_tmpsoft, &ctx->tmpsoft_cap, 2 * FEC_CUDA_SOFT_SLOT * sizeof(int16_t))) return FEC_CUDA_ENODEV;
  /* 8-bit LLR mode (srslte_rm_turbo_rx_lut_8bit): in and soft are int8 arrays; the soft buffer is widened for the device
   * layout and narrowed on the way back (the int8 kernel keeps sign-extended int8 values in int16 slots) */
  const size_t esz = ctx->llr8 ? 1 : 2;
  /* only the extent of the caller's array that the layout uses is read and written back: the reference touches nothing
   * else (rm_turbo.c:250-264,411-420), and its own tests pass arrays of exactly 3K+12 entries (rm_turbo_test.c:52,186) */
  const size_t next = (sb_layout && kt->g.P) ? 3 * ((size_t)kt->g.K + 32) + 12 : 3 * (size_t)kt->g.K + 12;
  memcpy(ctx->st_e.h, in, (size_t)E * esz);
  if (ctx->llr8) for (size_t i = 0; i < next; i++) ((int16_t *)ctx->st_data.h)[i] = ((const int8_t *)soft)[i];
  else memcpy(ctx->st_data.h, soft, next * 2);
  FecRmDesc *r = (FecRmDesc *)ctx->st_rmd.h;
  r->e = (const int16_t *)ctx->st_e.d; r->soft = ctx->d_tmpsoft; r->rank = rk; r->E = E; r->N = 3 * kt->g.K + 12;
  r->off = kt->off[rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = 0; r->slot = 0;
  cudaStream_t st = ctx->stream;
  /* the caller's soft buffer lives in host memory in the reference layout: bring it to the device layout, combine
   * there (wrapping int16 adds, rm_turbo.c:411-420), and hand the same layout back */
  CUF(cudaMemcpyAsync(ctx->st_e.d, ctx->st_e.h, (size_t)E * esz, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_data.d, ctx->st_data.h, next * 2, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_rmd.d, r, sizeof(FecRmDesc), cudaMemcpyHostToDevice, st));
  CUF(cudaMemsetAsync(ctx->d_softmax, 0, 8, st));
  CUF(fec_launch_ref_to_soft((const int16_t *)ctx->st_data.d, ctx->d_tmpsoft, kt->g, sb_layout, st));
  CUF(fec_launch_rm((const FecRmDesc *)ctx->st_rmd.d, ctx->d_softmax, 1, r->n_out, st, ctx->llr8));
  CUF(fec_launch_soft_to_ref(ctx->d_tmpsoft, (int16_t *)ctx->st_data.d, kt->g, sb_layout, 1, st));
  ctx->launches += 3;
  CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, next * 2, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  if (ctx->llr8) for (size_t i = 0; i < next; i++) ((int8_t *)soft)[i] = (int8_t)((const int16_t *)ctx->st_data.h)[i];
  else memcpy(soft, ctx->st_data.h, next * 2);
  return 0;
}

/* ---------------------------------------------------------------- standalone front end ---- */
extern "C" uint32_t fec_cuda_cinit_pdsch(uint32_t rnti, uint32_t codeword, uint32_t nslot, uint32_t cell_id)
{
  return (rnti << 14) + (codeword << 13) + ((nslot / 2) << 9) + cell_id;
}
extern "C" uint32_t fec_cuda_cinit_pusch(uint32_t rnti, uint32_t nslot, uint32_t cell_id)
{
  return (rnti << 14) + ((nslot / 2) << 9) + cell_id;
}

extern "C" int fec_cuda_frontend(fec_cuda_ctx_t *ctx, const fec_cuda_sym_t *sy, int16_t *llr_out)
{
  if (!ctx || !sy || !llr_out) return FEC_CUDA_EINVAL;
  fec_cuda_tb_t none;
  memset(&none, 0, sizeof(none));
  if (!fe_check(&none, sy) || (ctx->llr8 && sy->nof_symb)) return FEC_CUDA_EINVAL;       /* 8-bit LLR mode: downlink only */
  const size_t esz = ctx->llr8 ? 1 : 2;               /* llr_out points to int8 LLRs in 8-bit mode */
  cudaSetDevice(ctx->device);
  const uint32_t mod = sy->mod_bits, nb = sy->nof_re * mod, nuci = sy->nof_ack_pos + sy->nof_ri_pos;
  const size_t symb = ((size_t)sy->nof_re * 8 + 15) & ~(size_t)15, csib = sy->csi ? (((size_t)sy->nof_re * 4 + 15) & ~(size_t)15) : 0;
  const size_t lstb = ((size_t)2 * nuci * 4 + 15) & ~(size_t)15, ucib = (((size_t)nuci + sy->nof_cqi_bits) * 2 + 15) & ~(size_t)15;
  const size_t outb = ((size_t)nb * esz + 15) & ~(size_t)15;
  void *plan = NULL;
  /* one staging block: [symbols | csi | lists | descriptor] in, [LLRs | control-information LLRs | csi max] out */
  if (stage_reserve(ctx, &ctx->st_e, symb + csib + lstb + sizeof(FecFeDesc)) || stage_reserve(ctx, &ctx->st_data, outb + ucib + 16)) return FEC_CUDA_ENODEV;
  char *h = (char *)ctx->st_e.h, *d = (char *)ctx->st_e.d, *od = (char *)ctx->st_data.d;
  memcpy(h, sy->symbols, (size_t)sy->nof_re * 8);
  if (sy->csi) memcpy(h + symb, sy->csi, (size_t)sy->nof_re * 4);
  FecFeDesc *f = (FecFeDesc *)(h + symb + csib + lstb);
  memset(f, 0, sizeof(*f));
  f->sym = (const float2 *)d; f->csi = sy->csi ? (const float *)(d + symb) : NULL;
  f->e = (int16_t *)od; f->uci_out = (nuci + sy->nof_cqi_bits) ? (int16_t *)(od + outb) : NULL;
  f->nof_re = sy->nof_re; f->mod = mod; f->c_init = sy->c_init & 0x7FFFFFFFu;
  f->cols = sy->nof_symb; f->rows = sy->nof_symb ? sy->nof_re / sy->nof_symb : 0;
  f->n_cqi = sy->nof_cqi_bits; f->nuci = nuci;
  if (nuci) fe_build_lists(sy, f, (uint32_t *)(h + symb + csib), (const uint32_t *)(d + symb + csib));
  cudaStream_t st = ctx->stream;
  CUF(cudaMemcpyAsync(d, h, symb + csib + lstb + sizeof(FecFeDesc), cudaMemcpyHostToDevice, st));
  CUF(cudaMemsetAsync(od, 0, outb, st));       /* UL: the positions behind the last de-interleaved LLR are not written */
  CUF(fec_launch_frontend((const FecFeDesc *)(d + symb + csib + lstb), (float *)(od + outb + ucib), ctx->d_x2masks, ctx->d_x1states,
                          1, sy->nof_re, sy->csi != NULL, f->uci_out != NULL, sy->nof_symb ? 2 : 1, st, ctx->llr8));
  ctx->launches += 1 + (sy->csi ? 1 : 0) + (f->uci_out ? 1 : 0);
  CUF(cudaMemcpyAsync(ctx->st_data.h, od, outb + ucib, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  memcpy(llr_out, ctx->st_data.h, (size_t)nb * esz);
  if (f->uci_out) memcpy(sy->uci_out, (char *)ctx->st_data.h + outb, ((size_t)nuci + sy->nof_cqi_bits) * 2);
  return 0;
}

/* test hook: enable the LLR dump for later fec_cuda_decode_cbs calls (capacity n code blocks), and read it
 * back: out[i*6144 + x] = LLR of bit x of code block i after its last half-iteration. */
extern "C" int fec_cuda_debug_llr_enable(fec_cuda_ctx_t *ctx, uint32_t n)
{
  cudaSetDevice(ctx->device);
  if (ctx->d_dbg) cudaFree(ctx->d_dbg);
  ctx->d_dbg = NULL; ctx->dbg_cap = 0;
  if (!n) return 0;
  CU(cudaMalloc(&ctx->d_dbg, (size_t)n * FEC_MAX_K * 2));
  ctx->dbg_cap = (size_t)n * FEC_MAX_K * 2;
  return 0;
}
extern "C" int fec_cuda_debug_llr_read(fec_cuda_ctx_t *ctx, int16_t *out, uint32_t n)
{
  if (!ctx->d_dbg || n > ctx->dbg_n) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  for (uint32_t i = 0; i < n; i++)
    CU(cudaMemcpy(out + (size_t)i * FEC_MAX_K, ctx->d_dbg + (size_t)ctx->dbg_slot[i] * FEC_MAX_K, FEC_MAX_K * 2, cudaMemcpyDeviceToHost));
  return 0;
}
