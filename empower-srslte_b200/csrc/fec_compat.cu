/*
 * fec_compat.cu — the reference's own C symbols for the decode path as batch-of-one wrappers over the batched
 * API (include/srslte_fec_compat.h).  Host code only; every computation goes through fec_cuda_* (device).
 *
 * Threading (SURVEY 8b "Threading"): the reference has no locks on this path and expects one object per worker thread,
 * with several workers decoding at once (the PDSCH co-worker decodes TB0 while the caller decodes TB1, pdsch.c:837-864).
 * Here every calling THREAD gets its own decode context (streams, staging, per-K tables) on first use, so calls from
 * different threads run concurrently and a 16-bit and an 8-bit object never share mode state; soft buffers live in the
 * device-wide pool of fec_cuda_api.cu and can be decoded into from any thread. Device: FEC_CUDA_DEVICE (default 0).
 */
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "../../include/srslte_fec_cuda.h"
#include "../../include/srslte_fec_compat.h"
#include "../../include/srslte_fec_abi.h"
#include "fec_tables.h"

static_assert(sizeof(srslte_tdec_t) == FEC_ABI_SIZEOF_TDEC, "srslte_tdec_t view must keep the reference size (SURVEY Appendix B)");
static_assert(offsetof(srslte_tdec_t, n_iter) == FEC_ABI_TDEC_N_ITER, "n_iter offset");
static_assert(sizeof(srslte_softbuffer_rx_t) == FEC_ABI_SIZEOF_SOFTBUFFER_RX, "srslte_softbuffer_rx_t view must keep the reference size");
static_assert(offsetof(srslte_softbuffer_rx_t, buffer_f) == FEC_ABI_SOFTBUFFER_RX_BUFFER_F && offsetof(srslte_softbuffer_rx_t, cb_crc) == FEC_ABI_SOFTBUFFER_RX_CB_CRC &&
              offsetof(srslte_softbuffer_rx_t, tb_crc) == FEC_ABI_SOFTBUFFER_RX_TB_CRC, "soft buffer field offsets");

#define COMPAT_MAGIC 0xFEC0B200u
#define SRSLTE_SUCCESS 0
#define SRSLTE_ERROR (-1)
#define SRSLTE_ERROR_INVALID_INPUTS (-2)

static int compat_device(void) { const char *e = getenv("FEC_CUDA_DEVICE"); return e ? atoi(e) : 0; }

/* one decode context per calling thread, destroyed when the thread ends */
struct CompatThread {
  fec_cuda_ctx_t *ctx = NULL;
  ~CompatThread() { if (ctx) fec_cuda_destroy(ctx); }
};
static thread_local CompatThread tl_compat;

static fec_cuda_ctx_t *compat_ctx(void)
{
  if (!tl_compat.ctx) {
    fec_cuda_ctx_t *c = NULL;
    if (fec_cuda_create(&c, compat_device(), 64) != 0) {
      fprintf(stderr, "libsrslte_fec_cuda: no usable CUDA device (there is no CPU fallback)\n");
      return NULL;
    }
    tl_compat.ctx = c;
  }
  return tl_compat.ctx;
}

/* soft buffers outlive the thread that allocates them (srsENB's MAC allocates per UE, PHY workers decode): they are taken
 * from the pooled, process-lifetime part of the device pool through one home context that is never destroyed */
static std::mutex g_home_mu;
static fec_cuda_ctx_t *g_home = NULL;
static fec_cuda_ctx_t *compat_home(void)
{
  std::lock_guard<std::mutex> lk(g_home_mu);
  if (!g_home && fec_cuda_create(&g_home, compat_device(), 1) != 0) {
    fprintf(stderr, "libsrslte_fec_cuda: no usable CUDA device (there is no CPU fallback)\n");
    g_home = NULL;
  }
  return g_home;
}

/* ================================================================= turbo decoder (turbodecoder.c:133-533) ==== */
extern "C" int srslte_tdec_init_manual(srslte_tdec_t *h, uint32_t max_long_cb, srslte_tdec_impl_type_t dec_type)
{
  if (!h) return SRSLTE_ERROR_INVALID_INPUTS;
  memset(h, 0, sizeof(*h));                                     /* turbodecoder.c:156 */
  /* manual modes pick ONE arithmetic class for every K (turbodecoder.c:160-190): the non-windowed decoder, 8 or 16
   * sub-blocks in 16 bits. The generic decoder (a different algorithm: full-length recursions, turbodecoder_gen.c) and the
   * manual 8-bit classes are not provided. */
  int cls;
  switch (dec_type) {
  case SRSLTE_TDEC_AUTO:       cls = -1; break;
  case SRSLTE_TDEC_SSE:        cls = 0;  break;
  case SRSLTE_TDEC_SSE_WINDOW: cls = 8;  break;
  case SRSLTE_TDEC_AVX_WINDOW: cls = 16; break;
  default:
    fprintf(stderr, "Error decoder %d not supported\n", (int)dec_type);               /* turbodecoder.c:191-193 */
    return SRSLTE_ERROR;
  }
  if (max_long_cb > FEC_CUDA_MAX_K) return SRSLTE_ERROR;
  if (!compat_ctx()) return SRSLTE_ERROR;
  h->fec_input = (int16_t *)malloc(sizeof(int16_t) * (3 * FEC_CUDA_MAX_K + 12));
  if (!h->fec_input) return SRSLTE_ERROR;
  h->max_long_cb = max_long_cb;
  h->fec_manual_type = (uint32_t)(cls + 1);                     /* 0 = AUTO, else forced class + 1 */
  h->fec_magic = COMPAT_MAGIC;
  return SRSLTE_SUCCESS;
}

extern "C" int srslte_tdec_init(srslte_tdec_t *h, uint32_t max_long_cb) { return srslte_tdec_init_manual(h, max_long_cb, SRSLTE_TDEC_AUTO); }

extern "C" void srslte_tdec_free(srslte_tdec_t *h)
{
  if (!h) return;
  if (h->fec_magic == COMPAT_MAGIC && h->fec_input) free(h->fec_input);
  memset(h, 0, sizeof(*h));                                     /* turbodecoder.c:350 */
}

extern "C" void srslte_tdec_force_not_sb(srslte_tdec_t *h) { if (h) h->fec_force_not_sb = 1; }            /* turbodecoder.c:353-355 */

extern "C" int srslte_tdec_new_cb(srslte_tdec_t *h, uint32_t long_cb)
{
  if (!h || h->fec_magic != COMPAT_MAGIC) return SRSLTE_ERROR;
  if (long_cb > h->max_long_cb) {                               /* turbodecoder.c:495-499 */
    fprintf(stderr, "TDEC was initialized for max_long_cb=%u\n", h->max_long_cb);
    return SRSLTE_ERROR;
  }
  const int idx = fec_k_index(long_cb);
  if (idx < 0 || (uint32_t)fec_k_size((uint32_t)idx) != long_cb) {   /* turbodecoder.c:501-505: not an LTE block size */
    fprintf(stderr, "Invalid CB length %u\n", long_cb);
    return SRSLTE_ERROR;
  }
  if (h->fec_manual_type > 1) {
    /* a forced windowed class needs whole sub-blocks that cover its 40-step warm-ups (turbodecoder_win.h:321-384); the
     * reference does not check and reads outside its arrays for the sizes refused here */
    const uint32_t P = h->fec_manual_type - 1;
    if (long_cb % P || long_cb / P <= 40) { fprintf(stderr, "CB length %u cannot run in %u sub-blocks\n", long_cb, P); return SRSLTE_ERROR; }
  }
  h->n_iter = 0;
  h->fec_K = long_cb;
  return SRSLTE_SUCCESS;
}

extern "C" int srslte_tdec_get_nof_iterations(srslte_tdec_t *h) { return h ? (int)h->n_iter : 0; }

extern "C" uint32_t srslte_tdec_autoimp_get_subblocks(uint32_t nof_bits) { return (uint32_t)fec_k_subblocks(nof_bits); }

/* sub-block count of the 8-bit decoders, AVX2 build (turbodecoder.c:392-406) */
extern "C" uint32_t srslte_tdec_autoimp_get_subblocks_8bit(uint32_t nof_bits) { return (uint32_t)fec_k_subblocks8(nof_bits); }

/* the caller's input is in the sub-block layout of srslte_rm_turbo_rx_lut_(…, true) unless force_not_sb was set
 * or the block size has no windowed decoder (turbodecoder_iter.h:48,61-70,91-103): bring it to the natural layout */
static void compat_take_input(srslte_tdec_t *h, const int16_t *in)
{
  const uint32_t K = h->fec_K;
  fec_geom_t g;
  if (h->fec_manual_type) fec_geom_p(K, (int)h->fec_manual_type - 1, &g); else fec_geom(K, &g);
  if (g.P == 0 || h->fec_force_not_sb) { memcpy(h->fec_input, in, sizeof(int16_t) * (3 * K + 12)); return; }
  for (uint32_t t = 0; t < 3 * K + 12; t++) h->fec_input[t] = in[fec_soft_offset_ref(&g, (int)t, 1)];
}

/* Runs the code block from its LLRs for (n_iter + 1) half-iterations: the device decoder is deterministic, so this
 * IS the state the reference reaches by its (n_iter+1)-th incremental call (turbodecoder_iter.h:73-147). */
static int compat_run(srslte_tdec_t *h, uint8_t *output, uint32_t halfits, int llr8 = 0)
{
  fec_cuda_cb_t cb;
  memset(&cb, 0, sizeof(cb));
  cb.K = h->fec_K; cb.max_its = halfits; cb.crc_type = 1; cb.llr = h->fec_input; cb.out = output;
  fec_cuda_ctx_t *ctx = compat_ctx();                              /* this thread's context: nobody else switches its mode */
  if (!ctx) return SRSLTE_ERROR;
  fec_cuda_set_llr_8bit(ctx, llr8);
  fec_cuda_set_decoder_class(ctx, llr8 ? -1 : (int)h->fec_manual_type - 1);
  const int r = fec_cuda_decode_cbs(ctx, &cb, 1, FEC_CUDA_F_NO_EARLY_STOP);
  fec_cuda_set_llr_8bit(ctx, 0);
  fec_cuda_set_decoder_class(ctx, -1);
  if (r) fprintf(stderr, "libsrslte_fec_cuda: %s\n", fec_cuda_last_error(ctx));
  return r;
}

extern "C" void srslte_tdec_iteration(srslte_tdec_t *h, int16_t *input, uint8_t *output)
{
  if (!h || h->fec_magic != COMPAT_MAGIC || !h->fec_K || !input || !output) return;
  if (h->n_iter == 0) compat_take_input(h, input);
  if (compat_run(h, output, h->n_iter + 1) == 0) h->n_iter++;
}

extern "C" int srslte_tdec_run_all(srslte_tdec_t *h, int16_t *input, uint8_t *output, uint32_t nof_iterations, uint32_t long_cb)
{
  if (srslte_tdec_new_cb(h, long_cb)) return SRSLTE_ERROR;       /* turbodecoder.c:510-533 */
  if (!input || !output) return SRSLTE_ERROR;
  compat_take_input(h, input);
  if (nof_iterations == 0) return SRSLTE_SUCCESS;
  if (compat_run(h, output, nof_iterations)) return SRSLTE_ERROR;
  h->n_iter = nof_iterations;
  return SRSLTE_SUCCESS;
}

/* 8-bit entry points (turbodecoder.c:408-452,536-563): the int8 input is in the sub-block layout of
 * srslte_rm_turbo_rx_lut_8bit for the 8-bit class of K (32 / 16 / 8 sub-blocks) unless force_not_sb was set or K <= 400 */
static void compat_take_input8(srslte_tdec_t *h, const int8_t *in)
{
  const uint32_t K = h->fec_K;
  fec_geom_t g;
  fec_geom_p(K, fec_k_subblocks8(K), &g);
  const int sb = !(g.P == 0 || h->fec_force_not_sb);
  for (uint32_t t = 0; t < 3 * K + 12; t++) h->fec_input[t] = in[sb ? fec_soft_offset_ref(&g, (int)t, 1) : (int)t];
}

extern "C" void srslte_tdec_iteration_8bit(srslte_tdec_t *h, int8_t *input, uint8_t *output)
{
  if (!h || h->fec_magic != COMPAT_MAGIC || !h->fec_K || !input || !output) return;
  if (h->n_iter == 0) compat_take_input8(h, input);
  if (compat_run(h, output, h->n_iter + 1, 1) == 0) h->n_iter++;
}

extern "C" int srslte_tdec_run_all_8bit(srslte_tdec_t *h, int8_t *input, uint8_t *output, uint32_t nof_iterations, uint32_t long_cb)
{
  if (srslte_tdec_new_cb(h, long_cb)) return SRSLTE_ERROR;       /* turbodecoder.c:545-563 */
  if (!input || !output) return SRSLTE_ERROR;
  compat_take_input8(h, input);
  if (nof_iterations == 0) return SRSLTE_SUCCESS;
  if (compat_run(h, output, nof_iterations, 1)) return SRSLTE_ERROR;
  h->n_iter = nof_iterations;
  return SRSLTE_SUCCESS;
}

/* ================================================================= rate de-matching (rm_turbo.c:267-430) ==== */
extern "C" void srslte_rm_turbo_gentables(void) { }
extern "C" void srslte_rm_turbo_free_tables(void) { }

extern "C" int srslte_rm_turbo_rx_lut_(int16_t *input, int16_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx, bool enable_input_tdec)
{
  if (!input || !output || rv_idx >= 4 || cb_idx >= FEC_NOF_K) {  /* rm_turbo.c:395,426-429 */
    fprintf(stderr, "Invalid inputs rv_idx=%d, cb_idx=%d\n", rv_idx, cb_idx);
    return SRSLTE_ERROR_INVALID_INPUTS;
  }
  fec_cuda_ctx_t *ctx = compat_ctx();
  if (!ctx) return SRSLTE_ERROR;
  const int r = fec_cuda_rm_rx(ctx, input, in_len, cb_idx, rv_idx, enable_input_tdec ? 1 : 0, output);
  return r == 0 ? SRSLTE_SUCCESS : (r == FEC_CUDA_EINVAL ? SRSLTE_ERROR_INVALID_INPUTS : SRSLTE_ERROR);
}

/* rm_turbo.c:432-469: int8 LLRs into an int8 soft buffer (18600 entries as the reference sizes it), wrapping int8 adds, in the
 * sub-block layout of the 8-bit decoder class of this K */
extern "C" int srslte_rm_turbo_rx_lut_8bit(int8_t *input, int8_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx)
{
  if (!input || !output || rv_idx >= 4 || cb_idx >= FEC_NOF_K) {
    fprintf(stderr, "Invalid inputs rv_idx=%d, cb_idx=%d\n", rv_idx, cb_idx);
    return SRSLTE_ERROR_INVALID_INPUTS;
  }
  fec_cuda_ctx_t *ctx = compat_ctx();
  if (!ctx) return SRSLTE_ERROR;
  fec_cuda_set_llr_8bit(ctx, 1);
  const int r = fec_cuda_rm_rx(ctx, (const int16_t *)input, in_len, cb_idx, rv_idx, 1, (int16_t *)output);
  fec_cuda_set_llr_8bit(ctx, 0);
  return r == 0 ? SRSLTE_SUCCESS : (r == FEC_CUDA_EINVAL ? SRSLTE_ERROR_INVALID_INPUTS : SRSLTE_ERROR);
}

extern "C" int srslte_rm_turbo_rx_lut(int16_t *input, int16_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx)
{
  return srslte_rm_turbo_rx_lut_(input, output, in_len, cb_idx, rv_idx, false);          /* rm_turbo.c:378-381 */
}

/* ================================================================= soft buffers (softbuffer.c:46-151) ==== */
typedef struct { uint32_t magic, id, pending_new; } compat_sb_t;

/* TBS of I_TBS = 26 for 1..110 PRB (36.213 table 7.1.7.2.1-1), divided as softbuffer.c:52-54: max_cb = TBS/6120 + 1 */
static const uint32_t tbs_itbs26[110] = {
  712, 1480, 2216, 2984, 3752, 4392, 5160, 5992, 6712, 7480, 8248, 8760, 9528, 10296, 11064, 11832, 12576, 13536, 14112, 14688, 15264,
  16416, 16992, 17568, 18336, 19080, 19848, 20616, 21384, 22152, 22920, 23688, 24496, 25456, 25456, 26416, 27376, 28336, 29296, 29296,
  30576, 30576, 31704, 32856, 32856, 34008, 35160, 35160, 36696, 36696, 37888, 37888, 39232, 40576, 40576, 40576, 42368, 42368, 43816,
  43816, 45352, 45352, 46888, 46888, 48936, 48936, 48936, 51024, 51024, 52752, 52752, 52752, 55056, 55056, 55056, 55056, 57336, 57336,
  57336, 59256, 59256, 59256, 61664, 61664, 61664, 63776, 63776, 63776, 66592, 66592, 66592, 68808, 68808, 68808, 71112, 71112, 71112,
  73712, 73712, 75376, 75376, 75376, 75376, 75376, 75376, 75376, 75376, 75376, 75376, 75376 };

static compat_sb_t *compat_sb(srslte_softbuffer_rx_t *q)
{
  compat_sb_t *s = q ? (compat_sb_t *)q->buffer_f : NULL;
  return (s && s->magic == COMPAT_MAGIC) ? s : NULL;
}

extern "C" int srslte_softbuffer_rx_init(srslte_softbuffer_rx_t *q, uint32_t nof_prb)
{
  if (!q) return SRSLTE_ERROR_INVALID_INPUTS;
  memset(q, 0, sizeof(*q));                                      /* softbuffer.c:50 */
  if (nof_prb < 1 || nof_prb > 110) return SRSLTE_ERROR;         /* srslte_ra_tbs_from_idx fails (softbuffer.c:52-53) */
  fec_cuda_ctx_t *ctx = compat_home();
  if (!ctx) return SRSLTE_ERROR;
  compat_sb_t *s = (compat_sb_t *)calloc(1, sizeof(*s));
  q->max_cb = tbs_itbs26[nof_prb - 1] / (6144 - 24) + 1;
  q->cb_crc = (bool *)calloc(q->max_cb, sizeof(bool));
  int arc = -1;
  if (s && q->cb_crc) { std::lock_guard<std::mutex> lk(g_home_mu); arc = fec_cuda_softbuffer_alloc(ctx, 1, &s->id); }
  if (arc) { free(s); free(q->cb_crc); memset(q, 0, sizeof(*q)); return SRSLTE_ERROR; }
  s->magic = COMPAT_MAGIC; s->pending_new = 1;
  q->buffer_f = (int16_t **)s;
  return SRSLTE_SUCCESS;
}

extern "C" void srslte_softbuffer_rx_free(srslte_softbuffer_rx_t *q)
{
  if (!q) return;
  /* the device buffer goes back to the pool's free list (zeroed for its next user) */
  if (compat_sb(q)) {
    fec_cuda_ctx_t *ctx = compat_home();
    if (ctx) { std::lock_guard<std::mutex> lk(g_home_mu); fec_cuda_softbuffer_free(ctx, compat_sb(q)->id); }
    compat_sb(q)->magic = 0;
    free(q->buffer_f);
  }
  if (q->cb_crc) free(q->cb_crc);
  memset(q, 0, sizeof(*q));
}

/* the reference zeroes the first nof_cb CB buffers and ALL cb_crc flags (softbuffer.c:138-151).  Zeroing is deferred:
 * the next decode combines with new_data = 1, i.e. the rate de-matcher writes instead of accumulating. */
extern "C" void srslte_softbuffer_rx_reset_cb(srslte_softbuffer_rx_t *q, uint32_t)
{
  compat_sb_t *s = compat_sb(q);
  if (s) s->pending_new = 1;
  if (q && q->cb_crc) memset(q->cb_crc, 0, sizeof(bool) * q->max_cb);
}
extern "C" void srslte_softbuffer_rx_reset_tbs(srslte_softbuffer_rx_t *q, uint32_t tbs) { srslte_softbuffer_rx_reset_cb(q, (tbs + 24) / (6144 - 24) + 1); }
extern "C" void srslte_softbuffer_rx_reset(srslte_softbuffer_rx_t *q) { if (q) srslte_softbuffer_rx_reset_cb(q, q->max_cb); }

/* ================================================================= decode_tb (sch.c:437-486) ==== */
static int compat_decode_tb(srslte_softbuffer_rx_t *softbuffer, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                            const void *e_bits, int llr8, uint8_t *data, uint32_t max_iterations, uint32_t *nof_iterations)
{
  compat_sb_t *s = compat_sb(softbuffer);
  if (!s || !data || !e_bits) return SRSLTE_ERROR_INVALID_INPUTS;          /* sch.c:444-448 */
  uint32_t seg[9];
  if (fec_segm(tbs, seg)) return SRSLTE_ERROR_INVALID_INPUTS;
  if (tbs == 0 || seg[1] == 0) return SRSLTE_SUCCESS;                          /* sch.c:450-452 */
  if (seg[0]) { fprintf(stderr, "Error filler bits are not supported. Use standard TBS\n"); return SRSLTE_ERROR_INVALID_INPUTS; }
  if (seg[1] > softbuffer->max_cb) {
    fprintf(stderr, "Error number of CB (%d) exceeds soft buffer size (%d CBs)\n", seg[1], softbuffer->max_cb);
    return SRSLTE_ERROR_INVALID_INPUTS;
  }
  fec_cuda_ctx_t *ctx = compat_ctx();
  if (!ctx) return SRSLTE_ERROR;
  fec_cuda_tb_t t;
  memset(&t, 0, sizeof(t));
  t.tbs = tbs; t.Qm = Qm; t.rv = rv; t.nof_e_bits = nof_e_bits; t.max_its = max_iterations; t.softbuffer = s->id;
  t.new_data = s->pending_new; t.e_bits = (const int16_t *)e_bits; t.data = data;
  fec_cuda_set_llr_8bit(ctx, llr8);                                            /* q->llr_is_8bit (sch.c:344-348,362-364) */
  const int rc = fec_cuda_decode_tbs(ctx, &t, 1, 0);
  fec_cuda_set_llr_8bit(ctx, 0);
  if (rc) { fprintf(stderr, "libsrslte_fec_cuda: %s\n", fec_cuda_last_error(ctx)); return SRSLTE_ERROR; }
  /* the deferred reset is spent only when the rate de-matcher ran with it; a transport block the library refused (-2)
   * wrote nothing to the device buffer, so the next valid decode still has to start from zero */
  if (t.ret != FEC_CUDA_EINVAL) s->pending_new = 0;
  else return SRSLTE_ERROR_INVALID_INPUTS;
  for (uint32_t c = 0; c < seg[1] && c < softbuffer->max_cb; c++) softbuffer->cb_crc[c] = t.cb_crc[c] != 0;
  bool all = true;
  for (uint32_t c = 0; c < seg[1]; c++) all = all && t.cb_crc[c];
  softbuffer->tb_crc = all;                                                    /* sch.c:404-408 */
  if (nof_iterations) *nof_iterations = t.avg_its;                             /* sch.c:420 */
  return t.ret;
}

extern "C" int srslte_sch_decode_tb_cuda(srslte_softbuffer_rx_t *softbuffer, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                                         int16_t *e_bits, uint8_t *data, uint32_t max_iterations, uint32_t *nof_iterations)
{
  return compat_decode_tb(softbuffer, tbs, Qm, rv, nof_e_bits, e_bits, 0, data, max_iterations, nof_iterations);
}

/* the same for q->llr_is_8bit: int8 LLRs. A soft buffer must be reset between uses in the two modes (as in the reference,
 * whose int8 and int16 views of buffer_f are incompatible). */
extern "C" int srslte_sch_decode_tb_cuda_8bit(srslte_softbuffer_rx_t *softbuffer, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                                              int8_t *e_bits, uint8_t *data, uint32_t max_iterations, uint32_t *nof_iterations)
{
  return compat_decode_tb(softbuffer, tbs, Qm, rv, nof_e_bits, e_bits, 1, data, max_iterations, nof_iterations);
}

/* ================================================================= srslte_dlsch_decode[2] (sch.c:488-517) ==== */
/* The reference's own DL-SCH entry points on the reference's own structs. Only the fields decode_tb reads are touched,
 * through the offsets of include/srslte_fec_abi.h (asserted against the real headers by the ABI checker that is compiled with the link tests):
 * q->max_iterations / nof_iterations / llr_is_8bit, cfg->cb_segm[tb].tbs, cfg->grant.Qm[tb] / tb_en[], cfg->rv[tb],
 * cfg->nbits[tb].nof_bits, cfg->nof_layers. */
template <typename T> static T &abi_at(void *base, size_t off) { return *(T *)((char *)base + off); }

extern "C" int srslte_dlsch_decode2(void *q, void *cfg, srslte_softbuffer_rx_t *softbuffer, int16_t *e_bits, uint8_t *data, int tb_idx)
{
  if (!q || !cfg || tb_idx < 0 || tb_idx >= FEC_ABI_MAX_CODEWORDS) return SRSLTE_ERROR_INVALID_INPUTS;
  const uint32_t ntb = (abi_at<bool>(cfg, FEC_ABI_PDSCH_CFG_GRANT + FEC_ABI_DL_GRANT_TB_EN) ? 1u : 0u) +
                       (abi_at<bool>(cfg, FEC_ABI_PDSCH_CFG_GRANT + FEC_ABI_DL_GRANT_TB_EN + 1) ? 1u : 0u);       /* SRSLTE_RA_DL_GRANT_NOF_TB */
  const uint32_t Nl = abi_at<uint32_t>(cfg, FEC_ABI_PDSCH_CFG_NOF_LAYERS) != ntb ? 2 : 1;                         /* sch.c:508-512 */
  const uint32_t tbs = abi_at<uint32_t>(cfg, FEC_ABI_PDSCH_CFG_CB_SEGM + (size_t)tb_idx * FEC_ABI_SIZEOF_CBSEGM + FEC_ABI_CBSEGM_TBS);
  const uint32_t Qm = abi_at<uint32_t>(cfg, FEC_ABI_PDSCH_CFG_GRANT + FEC_ABI_DL_GRANT_QM + 4 * (size_t)tb_idx) * Nl;
  const uint32_t rv = abi_at<uint32_t>(cfg, FEC_ABI_PDSCH_CFG_RV + 4 * (size_t)tb_idx);
  const uint32_t nbits = abi_at<uint32_t>(cfg, FEC_ABI_PDSCH_CFG_NBITS + (size_t)tb_idx * FEC_ABI_SIZEOF_RA_NBITS + FEC_ABI_RA_NBITS_NOF_BITS);
  return compat_decode_tb(softbuffer, tbs, Qm, rv, nbits, e_bits, abi_at<bool>(q, FEC_ABI_SCH_LLR_IS_8BIT) ? 1 : 0, data,
                          abi_at<uint32_t>(q, FEC_ABI_SCH_MAX_ITERATIONS), &abi_at<uint32_t>(q, FEC_ABI_SCH_NOF_ITERATIONS));
}

extern "C" int srslte_dlsch_decode(void *q, void *cfg, srslte_softbuffer_rx_t *softbuffer, int16_t *e_bits, uint8_t *data)
{
  return srslte_dlsch_decode2(q, cfg, softbuffer, e_bits, data, 0);
}

/* ================================================================= encode_tb_off (sch.c:188-296) ==== */
extern "C" int srslte_sch_encode_tb_cuda(uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, uint8_t *data, uint8_t *e_bits)
{
  if (!e_bits || !data) return SRSLTE_ERROR_INVALID_INPUTS;                    /* sch.c:196-199 */
  fec_cuda_ctx_t *ctx = compat_ctx();
  if (!ctx) return SRSLTE_ERROR;
  fec_cuda_txtb_t t;
  memset(&t, 0, sizeof(t));
  t.tbs = tbs; t.Qm = Qm; t.rv = rv; t.nof_e_bits = nof_e_bits; t.data = data; t.e_bits = e_bits;
  if (fec_cuda_encode_tbs(ctx, &t, 1, 0)) { fprintf(stderr, "libsrslte_fec_cuda: %s\n", fec_cuda_last_error(ctx)); return SRSLTE_ERROR; }
  if (t.ret == SRSLTE_ERROR) fprintf(stderr, "Error filler bits are not supported. Use standard TBS\n");
  return t.ret;
}
