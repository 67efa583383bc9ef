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

#define CB_BYTES 768     /* 6144/8: per-CB decoded byte slot */

typedef struct {
  int        ready;
  fec_geom_t g;
  uint32_t  *d_step;
  uint16_t  *d_rank;
  uint32_t   off[4];
  uint32_t   slots;
  uint32_t  *d_pair_addr;     /* packed kernel: decoder-2 address words */
  uint2     *d_pair_crc[2];   /* packed kernel: CRC syndrome words, CRC24A / CRC24B */
} ktab_t;

typedef struct {
  uint32_t first_id, n;
  int16_t *d_soft;    /* n x 13 x FEC_CUDA_SOFT_SLOT */
  uint8_t *d_cache;   /* n x 13 x CB_BYTES */
  uint8_t *cb_crc;    /* host: n x 13 */
} sbchunk_t;

typedef struct { void *h, *d; size_t cap; } stage_t;   /* pinned host + device mirror, grown on demand */
#define FEC_NPIPE 8
typedef struct fec_pipe fec_pipe_t;

struct fec_cuda_ctx {
  int           device, sm_count;
  cudaStream_t  stream;
  char          err[256];
  uint64_t      launches;
  uint32_t      max_tbs;
  ktab_t        ktab[FEC_NOF_K];
  uint32_t     *d_r24a, *d_r24b;
  sbchunk_t    *chunks;
  uint32_t      nchunks, next_id;
  stage_t       st_e, st_data, st_cbd, st_rmd, st_tbd, st_status, st_ret, st_exd;
  uint8_t      *d_cbbytes; size_t cbbytes_cap;
  int16_t      *d_tmpsoft; size_t tmpsoft_cap;
  unsigned int *d_queue;   /* per K: work-queue counters of the packed and the exact kernel (2 x FEC_NOF_K), then FEC_NOF_K hand-over counts */
  uint32_t     *d_softmax; size_t softmax_cap;   /* per code block of the call: max |soft-buffer entry| */
  uint32_t     *d_exact;   size_t exact_cap;     /* per code block of the call: hand-over list of the packed kernel */
  int           force_exact;
  int16_t      *d_dbg; size_t dbg_cap; uint32_t *dbg_slot; uint32_t dbg_n;   /* test hook: LLR dump */
  float         phase_ms[3];   /* device time of rm | decode | assemble summed over the sub-batches of the last decode_tbs call */
  int           ev_valid;
  cudaStream_t  s_copy, s_out, s_rm, s_comp[2];
  int           ramp;           /* FEC_CUDA_RAMP=1: small first sub-batches (measured slower on B200, off by default) */
  fec_pipe_t   *pipes;
  uint32_t      sub_tbs, sub_tbs_dev;
  size_t        hw[9];          /* high-water marks of the per-pipe staging buffers */
  uint64_t      call_seq;       /* tickets of fec_cuda_submit_tbs */
  uint32_t      pipe_next;      /* next slot of the pipe ring (runs on across calls) */
  int           trace;          /* FEC_CUDA_TRACE=1: print the device timeline of every sub-batch to stderr */
  cudaEvent_t   ev_call;        /* start of the current decode_tbs call (copy stream) */
  double        t_call;         /* host clock at the start of the call */
};

static int pipes_create(fec_cuda_ctx_t *ctx);
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
  ctx->device = device; ctx->max_tbs = max_tbs_per_call ? max_tbs_per_call : 1;
  cudaDeviceProp prop;
  if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) { free(ctx); return FEC_CUDA_ENODEV; }
  if (prop.major < 10) { fprintf(stderr, "libsrslte_fec_cuda: built for sm_100a only, device is sm_%d%d\n", prop.major, prop.minor); free(ctx); return FEC_CUDA_ENODEV; }
  ctx->sm_count = prop.multiProcessorCount;
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
  cudaMalloc(&ctx->d_queue, sizeof(unsigned int) * 3 * FEC_NOF_K);
  { const char *fe = getenv("FEC_CUDA_FORCE_EXACT"); ctx->force_exact = fe && fe[0] == '1'; }
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
  for (int i = 0; i < FEC_NOF_K; i++) if (ctx->ktab[i].ready) {
    cudaFree(ctx->ktab[i].d_step); cudaFree(ctx->ktab[i].d_rank); cudaFree(ctx->ktab[i].d_pair_addr);
    cudaFree(ctx->ktab[i].d_pair_crc[0]); cudaFree(ctx->ktab[i].d_pair_crc[1]);
  }
  cudaFree(ctx->d_softmax); cudaFree(ctx->d_exact);
  for (uint32_t i = 0; i < ctx->nchunks; i++) { cudaFree(ctx->chunks[i].d_soft); cudaFree(ctx->chunks[i].d_cache); free(ctx->chunks[i].cb_crc); }
  free(ctx->chunks);
  stage_free(&ctx->st_e); stage_free(&ctx->st_data); stage_free(&ctx->st_cbd); stage_free(&ctx->st_rmd);
  stage_free(&ctx->st_tbd); stage_free(&ctx->st_status); stage_free(&ctx->st_ret); stage_free(&ctx->st_exd);
  cudaFree(ctx->d_cbbytes); cudaFree(ctx->d_tmpsoft); cudaFree(ctx->d_r24a); cudaFree(ctx->d_r24b); cudaFree(ctx->d_queue);
  pipes_destroy(ctx);
  cudaStreamDestroy(ctx->stream);
  free(ctx);
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
  ktab_t *t = &ctx->ktab[idx];
  if (!t->ready) {
    fec_geom(K, &t->g);
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
    t->slots = t->g.P ? fec_win_pick_slots(&t->g) : 0;
    t->ready = 1;
  }
  *out = t;
  return 0;
}

/* ---------------------------------------------------------------- soft buffers ---- */
extern "C" int fec_cuda_softbuffer_alloc(fec_cuda_ctx_t *ctx, uint32_t n, uint32_t *first_id)
{
  if (!ctx || !n || !first_id) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  sbchunk_t c;
  memset(&c, 0, sizeof(c));
  c.first_id = ctx->next_id; c.n = n;
  CU(cudaMalloc(&c.d_soft, (size_t)n * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t)));
  CU(cudaMalloc(&c.d_cache, (size_t)n * FEC_CUDA_MAX_CB * CB_BYTES));
  CU(cudaMemsetAsync(c.d_soft, 0, (size_t)n * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t), ctx->stream));
  CU(cudaMemsetAsync(c.d_cache, 0, (size_t)n * FEC_CUDA_MAX_CB * CB_BYTES, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));      /* decode calls run on several streams */
  c.cb_crc = (uint8_t *)calloc((size_t)n * FEC_CUDA_MAX_CB, 1);
  ctx->chunks = (sbchunk_t *)realloc(ctx->chunks, sizeof(sbchunk_t) * (ctx->nchunks + 1));
  ctx->chunks[ctx->nchunks++] = c;
  ctx->next_id += n;
  *first_id = c.first_id;
  return 0;
}

static sbchunk_t *sb_find(fec_cuda_ctx_t *ctx, uint32_t id, uint32_t *local)
{
  for (uint32_t i = 0; i < ctx->nchunks; i++)
    if (id >= ctx->chunks[i].first_id && id < ctx->chunks[i].first_id + ctx->chunks[i].n) {
      *local = id - ctx->chunks[i].first_id;
      return &ctx->chunks[i];
    }
  return NULL;
}

extern "C" int fec_cuda_softbuffer_reset(fec_cuda_ctx_t *ctx, uint32_t id)
{
  uint32_t l;
  sbchunk_t *c = ctx ? sb_find(ctx, id, &l) : NULL;
  if (!c) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  memset(c->cb_crc + (size_t)l * FEC_CUDA_MAX_CB, 0, FEC_CUDA_MAX_CB);
  CU(cudaMemsetAsync(c->d_soft + (size_t)l * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT, 0,
                     (size_t)FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT * sizeof(int16_t), ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int fec_cuda_softbuffer_read_ref_layout(fec_cuda_ctx_t *ctx, uint32_t id, uint32_t cb, uint32_t K, int16_t *out)
{
  uint32_t l;
  sbchunk_t *c = ctx ? sb_find(ctx, id, &l) : NULL;
  ktab_t *t;
  if (!c || cb >= FEC_CUDA_MAX_CB || !out) return FEC_CUDA_EINVAL;
  cudaSetDevice(ctx->device);
  if (ktab_get(ctx, K, &t)) return FEC_CUDA_EINVAL;
  if (dev_reserve(ctx, (void **)&ctx->d_tmpsoft, &ctx->tmpsoft_cap, 2 * FEC_CUDA_SOFT_SLOT * sizeof(int16_t))) return FEC_CUDA_ENODEV;
  CU(fec_launch_soft_to_ref(c->d_soft + ((size_t)l * FEC_CUDA_MAX_CB + cb) * FEC_CUDA_SOFT_SLOT, ctx->d_tmpsoft, t->g, 1, 0, ctx->stream));
  ctx->launches++;
  CU(cudaMemcpyAsync(out, ctx->d_tmpsoft, 18600 * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static void fill_decode_args(fec_cuda_ctx_t *ctx, int k, FecDecodeArgs *a, const FecCbDesc *cbd_dev, FecCbStatus *status_dev,
                             unsigned int *queue3, const uint32_t *softmax, uint32_t *exact, uint32_t off, uint32_t cnt, int16_t *dbg)
{
  const ktab_t *t = &ctx->ktab[k];
  memset(a, 0, sizeof(*a));
  a->g = t->g; a->step_tab = t->d_step; a->crcpow[0] = ctx->d_r24a; a->crcpow[1] = ctx->d_r24b;
  a->cbs = cbd_dev + off; a->status = status_dev;
  a->queue = queue3 + 2 * k; a->ncb = cnt; a->slots = t->slots; a->dbg_llr = dbg;
  a->pair_addr = t->d_pair_addr; a->pair_crc[0] = t->d_pair_crc[0]; a->pair_crc[1] = t->d_pair_crc[1];
  a->soft_max = softmax; a->status_off = off;
  if (t->g.P && !ctx->force_exact && !dbg) { a->exact_list = exact + off; a->exact_count = queue3 + 2 * FEC_NOF_K + k; }
}

/* ---------------------------------------------------------------- batched transport-block decode ---- */
/* A call is cut into sub-batches that flow through a ring of FEC_NPIPE "pipes" (own pinned staging, device
 * scratch and events each): while the GPU works on sub-batch i (H2D of its LLRs on the copy stream, kernels on
 * one of two compute streams, D2H on the output stream) the host builds the descriptors of sub-batch i+1 and
 * unpacks the results of sub-batch i-1. The call returns when every sub-batch is done. */
typedef struct { uint32_t seg[9]; sbchunk_t *chunk; uint32_t local; int valid, pinned; uint32_t ndec; size_t data_off, e_off; int kidx[2]; } tbplan_t;

struct fec_pipe {
  stage_t       st_e, st_data, st_cbd, st_rmd, st_tbd, st_status, st_ret;
  uint8_t      *d_cbbytes; size_t cbbytes_cap;
  uint32_t     *d_softmax; size_t softmax_cap;
  uint32_t     *d_exact;   size_t exact_cap;
  unsigned int *d_queue;
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
  CU(cudaMalloc(&p->d_queue, sizeof(unsigned int) * 3 * FEC_NOF_K));
  CU(cudaEventCreate(&p->ev_in));
  CU(cudaEventCreateWithFlags(&p->ev_k, cudaEventDisableTiming));
  CU(cudaEventCreate(&p->ev_done));
  for (int i = 0; i < 4; i++) CU(cudaEventCreate(&p->ev[i]));
  return 0;
}

static void pipe_free(fec_pipe_t *p)
{
  stage_free(&p->st_e); stage_free(&p->st_data); stage_free(&p->st_cbd); stage_free(&p->st_rmd);
  stage_free(&p->st_tbd); stage_free(&p->st_status); stage_free(&p->st_ret);
  cudaFree(p->d_cbbytes); cudaFree(p->d_softmax); cudaFree(p->d_exact); cudaFree(p->d_queue);
  if (p->ev_in) { cudaEventDestroy(p->ev_in); cudaEventDestroy(p->ev_k); cudaEventDestroy(p->ev_done); for (int i = 0; i < 4; i++) cudaEventDestroy(p->ev[i]); }
  free(p->plan);
  memset(p, 0, sizeof(*p));
}

/* build descriptors for tbs[0..n), enqueue copies + kernels; returns without waiting */
static int pipe_submit(fec_cuda_ctx_t *ctx, fec_pipe_t *pp, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags, cudaStream_t sc)
{
  const int dev_io = (flags & FEC_CUDA_F_DEVICE_IO) != 0;
  if (pp->plan_cap < n) { free(pp->plan); pp->plan = (tbplan_t *)malloc(sizeof(tbplan_t) * n); pp->plan_cap = n; }
  tbplan_t *plan = pp->plan;
  memset(plan, 0, sizeof(tbplan_t) * n);
  uint32_t grp_cnt[FEC_NOF_K], grp_off[FEC_NOF_K], grp_fill[FEC_NOF_K];
  memset(grp_cnt, 0, sizeof(grp_cnt));
  uint32_t ncb = 0, ntb = 0;
  size_t e_total = 0, data_total = 0;
  pp->tbs = tbs; pp->n = n; pp->dev_io = dev_io; pp->ncb = pp->ntb = 0; pp->busy = 0;

  /* pass 1: argument checks of decode_tb (sch.c:444-457) + segmentation + HARQ state */
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
    t->ret = FEC_CUDA_EINVAL; t->avg_its = 0;
    memset(t->cb_crc, 0, sizeof(t->cb_crc)); memset(t->cb_its, 0, sizeof(t->cb_its));
    if (!t->e_bits || !t->data || t->rv > 3 || t->Qm == 0) continue;
    if (fec_segm(t->tbs, p->seg)) continue;
    if (t->tbs == 0 || p->seg[1] == 0) { t->ret = 0; continue; }
    if (p->seg[0] || p->seg[1] > FEC_CUDA_MAX_CB) continue;            /* filler bits / too many CBs */
    p->chunk = sb_find(ctx, t->softbuffer, &p->local);
    if (!p->chunk) continue;
    uint8_t *crc = p->chunk->cb_crc + (size_t)p->local * FEC_CUDA_MAX_CB;
    if (t->new_data) memset(crc, 0, FEC_CUDA_MAX_CB);
    p->kidx[0] = (int)p->seg[4]; p->kidx[1] = (int)p->seg[5];
    p->valid = 1; p->ndec = 0;
    for (uint32_t cb = 0; cb < p->seg[1]; cb++)
      if (!crc[cb]) { grp_cnt[p->kidx[cb < p->seg[6] ? 0 : 1]]++; p->ndec++; }
    ncb += p->ndec; ntb++;
    p->e_off = e_total; e_total += ((size_t)t->nof_e_bits + 7) & ~(size_t)7;
    p->data_off = data_total; data_total += ((size_t)t->tbs / 8 + 3 + CB_BYTES + 15) & ~(size_t)15;
  }
  if (ntb == 0) return 0;
  uint32_t acc = 0;
  for (int k = 0; k < FEC_NOF_K; k++) { grp_off[k] = acc; grp_fill[k] = 0; acc += grp_cnt[k]; }

  /* every pipe of the ring is sized for the largest sub-batch seen so far: (re)allocation of pinned / device staging
   * synchronises the device, so it must stop after the first calls */
  size_t need[9] = { sizeof(FecCbDesc) * (ncb + 1), sizeof(FecRmDesc) * (ncb + 1), sizeof(FecTbDesc) * ntb, sizeof(FecCbStatus) * (ncb + 1),
                     sizeof(int32_t) * ntb, (size_t)(ncb + 1) * CB_BYTES, (size_t)(ncb + 1) * 4, dev_io ? 0 : e_total * 2, dev_io ? 0 : data_total };
  for (int i = 0; i < 9; i++) { if (need[i] > ctx->hw[i]) ctx->hw[i] = need[i]; need[i] = ctx->hw[i]; }
  if (stage_reserve(ctx, &pp->st_cbd, need[0]) || stage_reserve(ctx, &pp->st_rmd, need[1]) ||
      stage_reserve(ctx, &pp->st_tbd, need[2]) || stage_reserve(ctx, &pp->st_status, need[3]) ||
      stage_reserve(ctx, &pp->st_ret, need[4]) ||
      dev_reserve(ctx, (void **)&pp->d_cbbytes, &pp->cbbytes_cap, need[5]) ||
      dev_reserve(ctx, (void **)&pp->d_softmax, &pp->softmax_cap, need[6]) ||
      dev_reserve(ctx, (void **)&pp->d_exact, &pp->exact_cap, need[6]) ||
      (!dev_io && (stage_reserve(ctx, &pp->st_e, need[7]) || stage_reserve(ctx, &pp->st_data, need[8])))) return FEC_CUDA_ENODEV;

  FecCbDesc *cbd = (FecCbDesc *)pp->st_cbd.h;
  FecRmDesc *rmd = (FecRmDesc *)pp->st_rmd.h;
  FecTbDesc *tbd = (FecTbDesc *)pp->st_tbd.h;
  uint32_t max_n_out = 0, itb = 0;

  /* pass 2: descriptors. E / rp per code block exactly as decode_tb_cb derives them (sch.c:328-342) */
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
    if (!p->valid) continue;
    const uint32_t C = p->seg[1], C1 = p->seg[6];
    uint8_t *crc = p->chunk->cb_crc + (size_t)p->local * FEC_CUDA_MAX_CB;
    int16_t *soft0 = p->chunk->d_soft + (size_t)p->local * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT;
    uint8_t *cache0 = p->chunk->d_cache + (size_t)p->local * FEC_CUDA_MAX_CB * CB_BYTES;
    const int16_t *e_dev = dev_io ? t->e_bits : (const int16_t *)pp->st_e.d + p->e_off;
    FecTbDesc *td = &tbd[itb];
    memset(td, 0, sizeof(*td));
    td->data = dev_io ? t->data : (uint8_t *)pp->st_data.d + p->data_off;
    td->C = C; td->tbs = t->tbs; td->all_ok = 1; td->ret = (int32_t *)pp->st_ret.d + itb;
    const uint32_t Gp = t->nof_e_bits / t->Qm, gamma = Gp % C, n_e = t->Qm * (Gp / C);
    ktab_t *kts[2] = {NULL, NULL};
    for (uint32_t cb = 0; cb < C; cb++) {
      const int which = cb < C1 ? 0 : 1;
      const uint32_t K = which ? p->seg[3] : p->seg[2];
      td->K[cb] = K; td->rlen[cb] = C == 1 ? K : K - 24; td->cache[cb] = cache0 + (size_t)cb * CB_BYTES;
      if (crc[cb]) { td->st_idx[cb] = -1; td->cb_bytes[cb] = td->cache[cb]; continue; }
      if (!kts[which] && ktab_get(ctx, K, &kts[which])) return FEC_CUDA_EINVAL;
      ktab_t *kt = kts[which];
      uint32_t rp = cb * n_e, E = n_e;
      if (cb > C - gamma) { E = n_e + t->Qm; rp = (C - gamma) * n_e + (cb - (C - gamma)) * E; }   /* sch.c:339 off-by-one kept */
      const uint32_t slot = grp_off[kt->g.idx] + grp_fill[kt->g.idx]++;
      FecRmDesc *r = &rmd[slot];
      r->e = e_dev + rp; r->soft = soft0 + (size_t)cb * FEC_CUDA_SOFT_SLOT; r->rank = kt->d_rank;
      r->E = E; r->N = 3 * K + 12; r->off = kt->off[t->rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = t->new_data ? 1 : 0; r->slot = slot;
      if (r->n_out > max_n_out) max_n_out = r->n_out;
      FecCbDesc *d = &cbd[slot];
      d->soft = r->soft; d->out = pp->d_cbbytes + (size_t)slot * CB_BYTES; d->crc_sel = C > 1 ? 1 : 0;
      d->max_its = t->max_its ? t->max_its : 1; d->flags = (flags & FEC_CUDA_F_NO_EARLY_STOP) ? FEC_F_NO_EARLY : 0; d->status = slot;
      td->st_idx[cb] = (int32_t)slot; td->cb_bytes[cb] = d->out;
    }
    if (!dev_io) {
      /* pinned caller memory (fec_cuda_host_alloc / cudaHostRegister) is DMA'd straight from where it lies */
      cudaPointerAttributes pa;
      p->pinned = (cudaPointerGetAttributes(&pa, t->e_bits) == cudaSuccess && pa.type == cudaMemoryTypeHost);
      cudaGetLastError();
      if (!p->pinned) memcpy((int16_t *)pp->st_e.h + p->e_off, t->e_bits, (size_t)t->nof_e_bits * 2);
    }
    itb++;
  }

  /* ---- copy stream: LLRs (one copy per run of TBs contiguous on both sides) + descriptors ---- */
  cudaStream_t sin = ctx->s_copy, sout = ctx->s_out;
  if (!dev_io) {
    uint32_t i = 0;
    while (i < n) {
      if (!plan[i].valid) { i++; continue; }
      const int pinned = plan[i].pinned;
      const char *src0 = pinned ? (const char *)tbs[i].e_bits : (const char *)((int16_t *)pp->st_e.h + plan[i].e_off);
      const size_t dst0 = plan[i].e_off * 2;
      size_t len = (size_t)tbs[i].nof_e_bits * 2;
      uint32_t j = i + 1;
      for (; j < n; j++) {
        if (!plan[j].valid || plan[j].pinned != pinned || plan[j].e_off * 2 != dst0 + len) break;
        if (pinned && (const char *)tbs[j].e_bits != src0 + len) break;
        len += (size_t)tbs[j].nof_e_bits * 2;
      }
      CU(cudaMemcpyAsync((char *)pp->st_e.d + dst0, src0, len, cudaMemcpyHostToDevice, sin));
      i = j;
    }
  }
  CU(cudaMemcpyAsync(pp->st_cbd.d, cbd, sizeof(FecCbDesc) * ncb, cudaMemcpyHostToDevice, sin));
  CU(cudaMemcpyAsync(pp->st_rmd.d, rmd, sizeof(FecRmDesc) * ncb, cudaMemcpyHostToDevice, sin));
  CU(cudaMemcpyAsync(pp->st_tbd.d, tbd, sizeof(FecTbDesc) * ntb, cudaMemcpyHostToDevice, sin));
  CU(cudaEventRecord(pp->ev_in, sin));

  /* ---- compute stream: rate de-matching, turbo decode per K group, TB assembly + CRC24A ---- */
  /* rate de-matching has its own stream: it depends on nothing but its inputs, so it must not queue up behind the
   * decode kernel of an earlier sub-batch (which can hold a compute stream for the length of its slowest code block) */
  cudaStream_t srm = (flags & FEC_CUDA_F_SERIAL) ? sc : ctx->s_rm;
  CU(cudaMemsetAsync(pp->d_queue, 0, sizeof(unsigned int) * 3 * FEC_NOF_K, srm));
  CU(cudaMemsetAsync(pp->d_softmax, 0, (size_t)ncb * 4, srm));
  CU(cudaStreamWaitEvent(srm, pp->ev_in, 0));
  CU(cudaEventRecord(pp->ev[0], srm));
  CU(fec_launch_rm((const FecRmDesc *)pp->st_rmd.d, pp->d_softmax, ncb, max_n_out, srm));
  ctx->launches++;
  CU(cudaEventRecord(pp->ev[1], srm));
  if (srm != sc) CU(cudaStreamWaitEvent(sc, pp->ev[1], 0));
  for (int k = 0; k < FEC_NOF_K; k++) {
    if (!grp_cnt[k]) continue;
    FecDecodeArgs a;
    fill_decode_args(ctx, k, &a, (const FecCbDesc *)pp->st_cbd.d, (FecCbStatus *)pp->st_status.d, pp->d_queue, pp->d_softmax, pp->d_exact,
                     grp_off[k], grp_cnt[k], NULL);
    CU(fec_launch_decode(a, ctx->sm_count, sc));
    ctx->launches += a.exact_list ? 2 : 1;
  }
  CU(cudaEventRecord(pp->ev[2], sc));
  CU(fec_launch_tb_assemble((const FecTbDesc *)pp->st_tbd.d, (const FecCbStatus *)pp->st_status.d, ctx->d_r24a, ntb, sc));
  ctx->launches++;
  CU(cudaEventRecord(pp->ev[3], sc));
  CU(cudaEventRecord(pp->ev_k, sc));

  /* ---- output stream: results back ---- */
  CU(cudaStreamWaitEvent(sout, pp->ev_k, 0));
  if (ncb) CU(cudaMemcpyAsync(pp->st_status.h, pp->st_status.d, sizeof(FecCbStatus) * ncb, cudaMemcpyDeviceToHost, sout));
  CU(cudaMemcpyAsync(pp->st_ret.h, pp->st_ret.d, sizeof(int32_t) * ntb, cudaMemcpyDeviceToHost, sout));
  if (!dev_io) CU(cudaMemcpyAsync(pp->st_data.h, pp->st_data.d, data_total, cudaMemcpyDeviceToHost, sout));
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
    uint8_t *crc = p->chunk->cb_crc + (size_t)p->local * FEC_CUDA_MAX_CB;
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
  ctx->sub_tbs_dev = sb ? (uint32_t)atoi(sb) : 592;
  if (ctx->sub_tbs_dev == 0) ctx->sub_tbs_dev = 592;
  sb = getenv("FEC_CUDA_RAMP");
  ctx->ramp = sb ? atoi(sb) : 0;
  sb = getenv("FEC_CUDA_TRACE");
  ctx->trace = sb ? atoi(sb) : 0;
  CU(cudaEventCreate(&ctx->ev_call));
  return 0;
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

extern "C" int fec_cuda_submit_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags, uint64_t *ticket)
{
  if (!ctx || (!tbs && n)) return FEC_CUDA_EINVAL;
  const uint64_t id = ++ctx->call_seq;
  if (ticket) *ticket = id;
  if (n == 0) return 0;
  cudaSetDevice(ctx->device);
  ctx->phase_ms[0] = ctx->phase_ms[1] = ctx->phase_ms[2] = 0; ctx->ev_valid = 1;
  ctx->t_call = now_ms();
  if (ctx->trace) CU(cudaEventRecord(ctx->ev_call, ctx->s_copy));
  /* host buffers: small sub-batches so that the PCIe copy of one overlaps the kernels of the previous (the copy is
   * the longer of the two); device buffers: nothing to overlap but host-side descriptor work, so few, large ones */
  const uint32_t sub = (flags & FEC_CUDA_F_DEVICE_IO) ? ctx->sub_tbs_dev : ctx->sub_tbs;
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
    if (!rc) { pp->call_id = id; rc = pipe_submit(ctx, pp, tbs + off, cnt, flags, ctx->s_comp[slot & 1]); }
    off += cnt; s++;
  }
  return rc;
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
      dev_reserve(ctx, (void **)&ctx->d_exact, &ctx->exact_cap, (size_t)(n + 1) * 4) ||
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
  CUF(cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned int) * 3 * FEC_NOF_K, st));
  CUF(cudaMemsetAsync(ctx->d_softmax, 0, (size_t)n * 4, st));
  for (uint32_t o = 0; o < n; o += 32768) {
    CUF(fec_launch_extract((const FecExtractDesc *)ctx->st_exd.d + o, ctx->d_softmax, n - o < 32768 ? n - o : 32768, st));
    ctx->launches++;
  }
  for (int k = 0; k < FEC_NOF_K; k++) {
    if (!grp_cnt[k]) continue;
    FecDecodeArgs a;
    fill_decode_args(ctx, k, &a, (const FecCbDesc *)ctx->st_cbd.d, (FecCbStatus *)ctx->st_status.d, ctx->d_queue, ctx->d_softmax, ctx->d_exact,
                     grp_off[k], grp_cnt[k], DBG_LLR);
    CUF(fec_launch_decode(a, ctx->sm_count, st));
    ctx->launches += a.exact_list ? 2 : 1;
  }
  CUF(cudaMemcpyAsync(ctx->st_status.h, ctx->st_status.d, sizeof(FecCbStatus) * n, cudaMemcpyDeviceToHost, st));
  CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, (size_t)n * CB_BYTES, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  for (uint32_t i = 0; i < n; i++) {
    const FecCbStatus *s = (const FecCbStatus *)ctx->st_status.h + slot_of[i];
    cbs[i].its = s->its; cbs[i].crc_ok = s->crc_ok;
    memcpy(cbs[i].out, (uint8_t *)ctx->st_data.h + (size_t)slot_of[i] * CB_BYTES, cbs[i].K / 8);
  }
  free(ctx->dbg_slot); ctx->dbg_slot = slot_of; ctx->dbg_n = n;
  return 0;
}

/* ---------------------------------------------------------------- standalone rate de-matching ---- */
extern "C" int fec_cuda_rm_rx(fec_cuda_ctx_t *ctx, const int16_t *in, uint32_t E, uint32_t cb_idx, uint32_t rv,
                              int sb_layout, int16_t *soft)
{
  if (!ctx || !in || !soft || rv > 3 || fec_k_size(cb_idx) < 0) return FEC_CUDA_EINVAL;   /* rm_turbo.c:395,426-429 */
  cudaSetDevice(ctx->device);
  ktab_t *kt;
  if (ktab_get(ctx, (uint32_t)fec_k_size(cb_idx), &kt)) return FEC_CUDA_EINVAL;
  void *plan = NULL;
  if (stage_reserve(ctx, &ctx->st_e, ((size_t)E + 8) * 2) || stage_reserve(ctx, &ctx->st_rmd, sizeof(FecRmDesc)) ||
      stage_reserve(ctx, &ctx->st_data, 18600 * 2) ||
      dev_reserve(ctx, (void **)&ctx->d_softmax, &ctx->softmax_cap, 8) ||
      dev_reserve(ctx, (void **)&ctx->d_tmpsoft, &ctx->tmpsoft_cap, 2 * FEC_CUDA_SOFT_SLOT * sizeof(int16_t))) return FEC_CUDA_ENODEV;
  memcpy(ctx->st_e.h, in, (size_t)E * 2);
  memcpy(ctx->st_data.h, soft, 18600 * 2);
  FecRmDesc *r = (FecRmDesc *)ctx->st_rmd.h;
  r->e = (const int16_t *)ctx->st_e.d; r->soft = ctx->d_tmpsoft; r->rank = kt->d_rank; r->E = E; r->N = 3 * kt->g.K + 12;
  r->off = kt->off[rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = 0; r->slot = 0;
  cudaStream_t st = ctx->stream;
  /* the caller's soft buffer lives in host memory in the reference layout: bring it to the device layout, combine
   * there (wrapping int16 adds, rm_turbo.c:411-420), and hand the same layout back */
  CUF(cudaMemcpyAsync(ctx->st_e.d, ctx->st_e.h, (size_t)E * 2, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_data.d, ctx->st_data.h, 18600 * 2, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_rmd.d, r, sizeof(FecRmDesc), cudaMemcpyHostToDevice, st));
  CUF(cudaMemsetAsync(ctx->d_softmax, 0, 8, st));
  CUF(fec_launch_ref_to_soft((const int16_t *)ctx->st_data.d, ctx->d_tmpsoft, kt->g, sb_layout, st));
  CUF(fec_launch_rm((const FecRmDesc *)ctx->st_rmd.d, ctx->d_softmax, 1, r->n_out, st));
  CUF(fec_launch_soft_to_ref(ctx->d_tmpsoft, (int16_t *)ctx->st_data.d, kt->g, sb_layout, 1, st));
  ctx->launches += 3;
  CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, 18600 * 2, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  memcpy(soft, ctx->st_data.h, 18600 * 2);
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
