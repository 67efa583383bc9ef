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
  cudaEvent_t   ev[4];     /* phase boundaries of the last decode_tbs call: rm | decode | assemble */
  int           ev_valid;
};

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
  for (int i = 0; i < 4; i++) cudaEventCreate(&ctx->ev[i]);
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
  for (int i = 0; i < 4; i++) cudaEventDestroy(ctx->ev[i]);
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
  for (int i = 0; i < 3; i++)
    if (cudaEventElapsedTime(&out3[i], ctx->ev[i], ctx->ev[i + 1]) != cudaSuccess) return FEC_CUDA_ERROR;
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
  CU(fec_launch_soft_to_ref(c->d_soft + ((size_t)l * FEC_CUDA_MAX_CB + cb) * FEC_CUDA_SOFT_SLOT, ctx->d_tmpsoft, t->g, 1, ctx->stream));
  ctx->launches++;
  CU(cudaMemcpyAsync(out, ctx->d_tmpsoft, 18600 * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static void fill_decode_args(fec_cuda_ctx_t *ctx, int k, FecDecodeArgs *a, uint32_t off, uint32_t cnt, int16_t *dbg)
{
  const ktab_t *t = &ctx->ktab[k];
  memset(a, 0, sizeof(*a));
  a->g = t->g; a->step_tab = t->d_step; a->crcpow[0] = ctx->d_r24a; a->crcpow[1] = ctx->d_r24b;
  a->cbs = (const FecCbDesc *)ctx->st_cbd.d + off; a->status = (FecCbStatus *)ctx->st_status.d;
  a->queue = ctx->d_queue + 2 * k; a->ncb = cnt; a->slots = t->slots; a->dbg_llr = dbg;
  a->pair_addr = t->d_pair_addr; a->pair_crc[0] = t->d_pair_crc[0]; a->pair_crc[1] = t->d_pair_crc[1];
  a->soft_max = ctx->d_softmax;
  if (t->g.P && !ctx->force_exact && !dbg) { a->exact_list = ctx->d_exact + off; a->exact_count = ctx->d_queue + 2 * FEC_NOF_K + k; }
}

/* ---------------------------------------------------------------- batched transport-block decode ---- */
typedef struct { uint32_t seg[9]; sbchunk_t *chunk; uint32_t local; int valid, pinned; uint32_t first_cb; uint32_t ndec; size_t data_off, e_off; } tbplan_t;

#define CUF(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { free(plan); \
  return fail(ctx, FEC_CUDA_ENODEV, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)

#define DBG_LLR NULL
extern "C" int fec_cuda_decode_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags)
{
  if (!ctx || (!tbs && n)) return FEC_CUDA_EINVAL;
  if (n == 0) return 0;
  cudaSetDevice(ctx->device);
  const int dev_io = (flags & FEC_CUDA_F_DEVICE_IO) != 0;
  tbplan_t *plan = (tbplan_t *)calloc(n, sizeof(tbplan_t));
  uint32_t grp_cnt[FEC_NOF_K], grp_off[FEC_NOF_K], grp_fill[FEC_NOF_K];
  memset(grp_cnt, 0, sizeof(grp_cnt));
  uint32_t ncb = 0, ntb = 0;
  size_t e_total = 0, data_total = 0;

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
    p->valid = 1; p->first_cb = ncb; p->ndec = 0;
    for (uint32_t cb = 0; cb < p->seg[1]; cb++)
      if (!crc[cb]) { uint32_t K = cb < p->seg[6] ? p->seg[2] : p->seg[3]; grp_cnt[fec_k_index(K)]++; p->ndec++; }
    ncb += p->ndec; ntb++;
    p->e_off = e_total; e_total += ((size_t)t->nof_e_bits + 7) & ~(size_t)7;
    p->data_off = data_total; data_total += ((size_t)t->tbs / 8 + 3 + CB_BYTES + 15) & ~(size_t)15;
  }
  if (ntb == 0) { free(plan); return 0; }
  uint32_t acc = 0;
  for (int k = 0; k < FEC_NOF_K; k++) { grp_off[k] = acc; grp_fill[k] = 0; acc += grp_cnt[k]; }

  int rc = 0;
  if (stage_reserve(ctx, &ctx->st_cbd, sizeof(FecCbDesc) * (ncb + 1)) || stage_reserve(ctx, &ctx->st_rmd, sizeof(FecRmDesc) * (ncb + 1)) ||
      stage_reserve(ctx, &ctx->st_tbd, sizeof(FecTbDesc) * ntb) || stage_reserve(ctx, &ctx->st_status, sizeof(FecCbStatus) * (ncb + 1)) ||
      stage_reserve(ctx, &ctx->st_ret, sizeof(int32_t) * ntb) ||
      dev_reserve(ctx, (void **)&ctx->d_cbbytes, &ctx->cbbytes_cap, (size_t)(ncb + 1) * CB_BYTES) ||
      dev_reserve(ctx, (void **)&ctx->d_softmax, &ctx->softmax_cap, (size_t)(ncb + 1) * 4) ||
      dev_reserve(ctx, (void **)&ctx->d_exact, &ctx->exact_cap, (size_t)(ncb + 1) * 4) ||
      (!dev_io && (stage_reserve(ctx, &ctx->st_e, e_total * 2) || stage_reserve(ctx, &ctx->st_data, data_total)))) { free(plan); return FEC_CUDA_ENODEV; }

  FecCbDesc *cbd = (FecCbDesc *)ctx->st_cbd.h;
  FecRmDesc *rmd = (FecRmDesc *)ctx->st_rmd.h;
  FecTbDesc *tbd = (FecTbDesc *)ctx->st_tbd.h;
  uint32_t max_n_out = 0, itb = 0;

  /* pass 2: descriptors. E / rp per code block exactly as decode_tb_cb derives them (sch.c:328-342) */
  for (uint32_t i = 0; i < n && !rc; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
    if (!p->valid) continue;
    const uint32_t C = p->seg[1], C1 = p->seg[6];
    uint8_t *crc = p->chunk->cb_crc + (size_t)p->local * FEC_CUDA_MAX_CB;
    int16_t *soft0 = p->chunk->d_soft + (size_t)p->local * FEC_CUDA_MAX_CB * FEC_CUDA_SOFT_SLOT;
    uint8_t *cache0 = p->chunk->d_cache + (size_t)p->local * FEC_CUDA_MAX_CB * CB_BYTES;
    const int16_t *e_dev = dev_io ? t->e_bits : (const int16_t *)ctx->st_e.d + p->e_off;
    FecTbDesc *td = &tbd[itb];
    memset(td, 0, sizeof(*td));
    td->data = dev_io ? t->data : (uint8_t *)ctx->st_data.d + p->data_off;
    td->C = C; td->tbs = t->tbs; td->all_ok = 1; td->ret = (int32_t *)ctx->st_ret.d + itb;
    const uint32_t Gp = t->nof_e_bits / t->Qm, gamma = Gp % C, n_e = t->Qm * (Gp / C);
    for (uint32_t cb = 0; cb < C; cb++) {
      const uint32_t K = cb < C1 ? p->seg[2] : p->seg[3];
      td->K[cb] = K; td->rlen[cb] = C == 1 ? K : K - 24; td->cache[cb] = cache0 + (size_t)cb * CB_BYTES;
      if (crc[cb]) { td->st_idx[cb] = -1; td->cb_bytes[cb] = td->cache[cb]; continue; }
      ktab_t *kt;
      if (ktab_get(ctx, K, &kt)) { rc = FEC_CUDA_EINVAL; break; }
      uint32_t rp = cb * n_e, E = n_e;
      if (cb > C - gamma) { E = n_e + t->Qm; rp = (C - gamma) * n_e + (cb - (C - gamma)) * E; }   /* sch.c:339 off-by-one kept */
      const uint32_t slot = grp_off[kt->g.idx] + grp_fill[kt->g.idx]++;
      FecRmDesc *r = &rmd[slot];
      r->e = e_dev + rp; r->soft = soft0 + (size_t)cb * FEC_CUDA_SOFT_SLOT; r->rank = kt->d_rank;
      r->E = E; r->N = 3 * K + 12; r->off = kt->off[t->rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = t->new_data ? 1 : 0; r->slot = slot;
      if (r->n_out > max_n_out) max_n_out = r->n_out;
      FecCbDesc *d = &cbd[slot];
      d->soft = r->soft; d->out = ctx->d_cbbytes + (size_t)slot * CB_BYTES; d->crc_sel = C > 1 ? 1 : 0;
      d->max_its = t->max_its ? t->max_its : 1; d->flags = (flags & FEC_CUDA_F_NO_EARLY_STOP) ? FEC_F_NO_EARLY : 0; d->status = slot;
      td->st_idx[cb] = (int32_t)slot; td->cb_bytes[cb] = d->out;
    }
    if (!dev_io) {
      /* pinned caller memory (fec_cuda_host_alloc / cudaHostRegister) is DMA'd straight from where it lies */
      cudaPointerAttributes pa;
      p->pinned = (cudaPointerGetAttributes(&pa, t->e_bits) == cudaSuccess && pa.type == cudaMemoryTypeHost);
      cudaGetLastError();
      if (!p->pinned) memcpy((int16_t *)ctx->st_e.h + p->e_off, t->e_bits, (size_t)t->nof_e_bits * 2);
    }
    itb++;
  }
  if (rc) { free(plan); return rc; }

  cudaStream_t st = ctx->stream;
  if (!dev_io) {
    /* one H2D copy per run of transport blocks that is contiguous on both sides */
    uint32_t i = 0;
    while (i < n) {
      if (!plan[i].valid) { i++; continue; }
      const int pinned = plan[i].pinned;
      const char *src0 = pinned ? (const char *)tbs[i].e_bits : (const char *)((int16_t *)ctx->st_e.h + plan[i].e_off);
      const size_t dst0 = plan[i].e_off * 2;
      size_t len = (size_t)tbs[i].nof_e_bits * 2;
      uint32_t j = i + 1;
      for (; j < n; j++) {
        if (!plan[j].valid || plan[j].pinned != pinned || plan[j].e_off * 2 != dst0 + len) break;
        if (pinned && (const char *)tbs[j].e_bits != src0 + len) break;
        if (!pinned && (((size_t)tbs[j - 1].nof_e_bits) & 7)) break;
        len += (size_t)tbs[j].nof_e_bits * 2;
      }
      CUF(cudaMemcpyAsync((char *)ctx->st_e.d + dst0, src0, len, cudaMemcpyHostToDevice, st));
      i = j;
    }
  }
  CUF(cudaMemcpyAsync(ctx->st_cbd.d, cbd, sizeof(FecCbDesc) * ncb, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_rmd.d, rmd, sizeof(FecRmDesc) * ncb, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_tbd.d, tbd, sizeof(FecTbDesc) * ntb, cudaMemcpyHostToDevice, st));
  CUF(cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned int) * 3 * FEC_NOF_K, st));
  CUF(cudaMemsetAsync(ctx->d_softmax, 0, (size_t)ncb * 4, st));
  CUF(cudaEventRecord(ctx->ev[0], st));
  for (uint32_t o = 0; o < ncb; o += 32768) {                       /* grid.y limit */
    const uint32_t m = ncb - o < 32768 ? ncb - o : 32768;
    CUF(fec_launch_rm((const FecRmDesc *)ctx->st_rmd.d + o, ctx->d_softmax, m, max_n_out, st));
    ctx->launches++;
  }
  CUF(cudaEventRecord(ctx->ev[1], st));
  for (int k = 0; k < FEC_NOF_K; k++) {
    if (!grp_cnt[k]) continue;
    FecDecodeArgs a;
    fill_decode_args(ctx, k, &a, grp_off[k], grp_cnt[k], DBG_LLR);
    CUF(fec_launch_decode(a, ctx->sm_count, st));
    ctx->launches += a.exact_list ? 2 : 1;
  }
  CUF(cudaEventRecord(ctx->ev[2], st));
  CUF(fec_launch_tb_assemble((const FecTbDesc *)ctx->st_tbd.d, (const FecCbStatus *)ctx->st_status.d, ctx->d_r24a, ntb, st));
  ctx->launches++;
  CUF(cudaEventRecord(ctx->ev[3], st));
  ctx->ev_valid = 1;
  if (ncb) CUF(cudaMemcpyAsync(ctx->st_status.h, ctx->st_status.d, sizeof(FecCbStatus) * ncb, cudaMemcpyDeviceToHost, st));
  CUF(cudaMemcpyAsync(ctx->st_ret.h, ctx->st_ret.d, sizeof(int32_t) * ntb, cudaMemcpyDeviceToHost, st));
  if (!dev_io) CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, data_total, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));

  /* pass 3: results + HARQ state (softbuffer->cb_crc, sch.c:384,404-408; nof_iterations sch.c:420) */
  const FecCbStatus *stt = (const FecCbStatus *)ctx->st_status.h;
  itb = 0;
  for (uint32_t i = 0; i < n; i++) {
    fec_cuda_tb_t *t = &tbs[i];
    tbplan_t *p = &plan[i];
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
    t->ret = ((const int32_t *)ctx->st_ret.h)[itb];
    if (!dev_io) memcpy(t->data, (uint8_t *)ctx->st_data.h + p->data_off, (size_t)t->tbs / 8 + 3);
    itb++;
  }
  free(plan);
  return 0;
}

/* ---------------------------------------------------------------- bare code blocks (srslte_tdec_* level) ---- */
#undef DBG_LLR
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
    fill_decode_args(ctx, k, &a, grp_off[k], grp_cnt[k], DBG_LLR);
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
  FecRmDesc *r = (FecRmDesc *)ctx->st_rmd.h;
  r->e = (const int16_t *)ctx->st_e.d; r->soft = ctx->d_tmpsoft; r->rank = kt->d_rank; r->E = E; r->N = 3 * kt->g.K + 12;
  r->off = kt->off[rv]; r->n_out = 3 * kt->g.SS + 16; r->new_data = 1; r->slot = 0;
  cudaStream_t st = ctx->stream;
  CUF(cudaMemcpyAsync(ctx->st_e.d, ctx->st_e.h, (size_t)E * 2, cudaMemcpyHostToDevice, st));
  CUF(cudaMemcpyAsync(ctx->st_rmd.d, r, sizeof(FecRmDesc), cudaMemcpyHostToDevice, st));
  CUF(fec_launch_rm((const FecRmDesc *)ctx->st_rmd.d, ctx->d_softmax, 1, r->n_out, st));
  CUF(fec_launch_soft_to_ref(ctx->d_tmpsoft, (int16_t *)ctx->st_data.d, kt->g, sb_layout, st));
  ctx->launches += 2;
  CUF(cudaMemcpyAsync(ctx->st_data.h, ctx->st_data.d, 18600 * 2, cudaMemcpyDeviceToHost, st));
  CUF(cudaStreamSynchronize(st));
  /* the caller's soft buffer lives in host memory: add the de-rate-matched block into it (wrapping int16) */
  const int16_t *d = (const int16_t *)ctx->st_data.h;
  for (int i = 0; i < 18600; i++) soft[i] = (int16_t)(soft[i] + d[i]);
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
