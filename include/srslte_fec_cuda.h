/*
 * srslte_fec_cuda.h — C ABI of libsrslte_fec_cuda.so: the B200 (sm_100a) transport-block decode path
 * behind srsLTE's FEC / SCH decode API. Plain C, plain pointers and sizes; no torch / C++ types.
 *
 * Two layers:
 *  (1) the batched API (fec_cuda_*): the ADDITION a GPU needs — many transport blocks (subframes x HARQ
 *      processes x cells) per call. It replaces the per-TB loop the reference runs in
 *      decode_tb()/decode_tb_cb() (lib/src/phy/phch/sch.c:307-498).
 *  (2) the reference's own symbols (srslte_*): declared in srslte_fec_compat.h, batch-of-one wrappers over (1).
 *
 * Error codes follow lib/include/srslte/config.h:64-66: 0 success, -1 error (CRC fail), -2 invalid inputs.
 * There is NO CPU fallback: every entry point that computes returns FEC_CUDA_ENODEV (-100) when no
 * CUDA device / kernel image is usable.
 */
#ifndef SRSLTE_FEC_CUDA_H
#define SRSLTE_FEC_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEC_CUDA_SUCCESS          0
#define FEC_CUDA_ERROR           -1
#define FEC_CUDA_EINVAL          -2
#define FEC_CUDA_ENODEV          -100

#define FEC_CUDA_MAX_CB           13       /* SRSLTE_MAX_CODEBLOCKS as reachable from tbs_tables */
#define FEC_CUDA_MAX_K            6144
#define FEC_CUDA_SOFT_SLOT        18464    /* int16 per code-block soft-buffer slot on the device (>= reference 18600 role) */

/* flags for fec_cuda_decode_tbs */
#define FEC_CUDA_F_DEVICE_IO      1u       /* e_bits / data are DEVICE pointers; nothing bulk crosses PCIe */
#define FEC_CUDA_F_NO_EARLY_STOP  2u       /* run exactly max_its half-iterations (srslte_tdec_run_all semantics) */
#define FEC_CUDA_F_SERIAL         4u       /* do not cut the call into pipelined sub-batches: one launch per kernel, so that
                                              fec_cuda_last_phase_ms reports un-overlapped kernel durations (profiling) */

typedef struct fec_cuda_ctx fec_cuda_ctx_t;

/* One transport block of a batch. Mirrors the arguments of decode_tb() (sch.c:437-441) plus the HARQ
 * soft-buffer identity (srslte_softbuffer_rx_t, softbuffer.h:42-48) which lives on the device. */
typedef struct {
  /* in */
  uint32_t       tbs;          /* transport block size in bits (cb_segm->tbs) */
  uint32_t       Qm;           /* modulation order x layers, as decode_tb receives it */
  uint32_t       rv;           /* redundancy version 0..3 */
  uint32_t       nof_e_bits;   /* number of LLRs in e_bits (cfg->nbits[].nof_bits) */
  uint32_t       max_its;      /* max HALF-iterations per code block (srslte_sch_set_max_noi) */
  uint32_t       softbuffer;   /* device soft-buffer id from fec_cuda_softbuffer_alloc */
  uint32_t       new_data;     /* 1: srslte_softbuffer_rx_reset_tbs() semantics before combining */
  uint32_t       reserved;
  const int16_t *e_bits;       /* LLRs, host (default) or device (FEC_CUDA_F_DEVICE_IO) */
  uint8_t       *data;         /* out: tbs/8 + 3 bytes are written (TB + CRC24A); needs tbs/8+3+768 bytes of room */
  /* out */
  int32_t        ret;          /* 0 TB OK, -1 CRC fail, -2 invalid (decode_tb return value) */
  uint32_t       avg_its;      /* sum of half-iterations / C, integer division (sch.c:420, srslte_sch_last_noi) */
  uint8_t        cb_crc[FEC_CUDA_MAX_CB];   /* per-CB CRC state after this call (softbuffer->cb_crc) */
  uint8_t        cb_its[FEC_CUDA_MAX_CB];   /* half-iterations run per CB in this call (0 = skipped, already OK) */
  uint8_t        pad[6];
} fec_cuda_tb_t;

/* One bare code block for the srslte_tdec_* level (turbodecoder.h:102-140): natural-layout input
 * [s p p'] x K + 12 tail, as srslte_tdec_run_all / turbodecoder_test feed it. */
typedef struct {
  uint32_t       K;            /* code block size, must be one of the 188 LTE sizes */
  uint32_t       max_its;      /* half-iterations (cap, or exact count with FEC_CUDA_F_NO_EARLY_STOP) */
  uint32_t       crc_type;     /* 0 CRC24A, 1 CRC24B — only used for early stop */
  uint32_t       reserved;
  const int16_t *llr;          /* 3K+12 int16 */
  uint8_t       *out;          /* K/8 bytes */
  uint32_t       its;          /* out: half-iterations run */
  uint32_t       crc_ok;       /* out: 1 if the CRC over the K decoded bits is zero */
} fec_cuda_cb_t;

/* ---- lifecycle ------------------------------------------------------------------------------------ */
/* device: CUDA ordinal. max_tbs_per_call bounds the staging buffers (pinned host + device).
 * Environment read at creation (tuning and diagnosis, none is needed): FEC_CUDA_SUB_TBS = transport blocks per pipelined
 * sub-batch of a call with host buffers (default 296; twice that for int8 LLRs: sub-batches are sized by what crosses the
 * link); FEC_CUDA_SUB_TBS_DEV = the same with FEC_CUDA_F_DEVICE_IO (default: two sub-batches per call, at least 1184 each —
 * decoder launches as large as possible); FEC_CUDA_TRACE=1 prints the device timeline of every sub-batch to
 * stderr; FEC_CUDA_FORCE_EXACT=1 sends every code block to the exact 32-bit kernels. */
int  fec_cuda_create(fec_cuda_ctx_t **ctx, int device, uint32_t max_tbs_per_call);
void fec_cuda_destroy(fec_cuda_ctx_t *ctx);
const char *fec_cuda_last_error(fec_cuda_ctx_t *ctx);
/* the CUDA stream (cudaStream_t) all kernels of this context are launched on — for event timing */
void *fec_cuda_stream(fec_cuda_ctx_t *ctx);
/* number of kernels launched by this context so far */
uint64_t fec_cuda_launch_count(fec_cuda_ctx_t *ctx);
/* device time (CUDA events on the compute streams) of the three phases of the LAST fec_cuda_decode_tbs call, summed over
 * its sub-batches: out3 = { rate de-matching, turbo decode (all K groups), TB assembly + CRC24A } in milliseconds.
 * Sub-batches overlap on two streams; use FEC_CUDA_F_SERIAL for un-overlapped kernel durations. */
int  fec_cuda_last_phase_ms(fec_cuda_ctx_t *ctx, float out3[3]);

/* which kernels the code blocks of the LAST decode call took: out2 = { handed from the a-priori-range packed kernel to its
 * run-time-certificate variant, handed on to the exact 32-bit kernel }; everything else stayed in the first kernel */
int  fec_cuda_last_path_counts(fec_cuda_ctx_t *ctx, uint64_t out2[2]);

/* 8-bit LLR mode, the counterpart of srslte_sch_t.llr_is_8bit (sch.h:60, sch.c:344-348,362-364; set by
 * srslte_pdsch_init_ue / srslte_pusch_init_enb from the expert.*_8bit_decoder options). While on, fec_cuda_tb_t.e_bits point
 * to int8 LLRs (half the ingest), HARQ combining wraps at int8 (srslte_rm_turbo_rx_lut_8bit, rm_turbo.c:432-469) and code
 * blocks decode as srslte_tdec_iteration_8bit does (turbodecoder.c:392-452): K > 2048 in the 32-sub-block int8 decoder
 * (avx8), 800 < K <= 2048 with K mod 16 == 0 in the 16-sub-block one (sse8), every other K in the 16-bit decoder of its
 * class on the widened values. fec_cuda_cb_t.input stays int16 (values within int8). The symbol front end
 * (fec_cuda_decode_syms / fec_cuda_frontend) then runs the reference's 8-bit chain — srslte_demod_soft_demodulate_b
 * (demod_soft.c:458-478), srslte_scrambling_sb_offset (scrambling.c:53-56), the int8 branch of csi_correction
 * (pdsch.c:707-714) — for DOWNLINK transport blocks; uplink ones (nof_symb != 0) are refused in this mode. Soft buffers
 * must be reset when the mode changes. */
int  fec_cuda_set_llr_8bit(fec_cuda_ctx_t *ctx, int on);

/* srslte_tdec_init_manual (turbodecoder.c:153-190): force ONE 16-bit arithmetic class on every code block size instead of
 * the AUTO mapping of turbodecoder.c:364-376 — P = 0 the non-windowed decoder (SRSLTE_TDEC_SSE), 8 (SRSLTE_TDEC_SSE_WINDOW),
 * 16 (SRSLTE_TDEC_AVX_WINDOW); -1 returns to AUTO. Applies to fec_cuda_decode_cbs (the srslte_tdec_* level, what
 * turbodecoder_test -d N exercises); forced classes run in the exact kernels. A windowed class needs K mod P == 0 and
 * K / P > 40 (its warm-up length; the reference itself corrupts its heap at K / P = 40), else the code block is refused with FEC_CUDA_EINVAL. */
int  fec_cuda_set_decoder_class(fec_cuda_ctx_t *ctx, int P);

/* Pinned (page-locked) host memory for e_bits / data buffers: replaces srslte_vec_malloc
 * (lib/src/phy/utils/vector.c:125-132) for the buffers handed to the decode calls. Buffers obtained here (or
 * registered with cudaHostRegister) are DMA'd directly; any other host pointer is staged through a copy. */
void *fec_cuda_host_alloc(size_t bytes);
void  fec_cuda_host_free(void *p);

/* ---- HARQ soft buffers (replace srslte_softbuffer_rx_init/free/reset, softbuffer.c:46-151) ---------- */
/* allocates n TB soft buffers (13 code-block slots each) in HBM; ids are first_id .. first_id+n-1 */
int  fec_cuda_softbuffer_alloc(fec_cuda_ctx_t *ctx, uint32_t n, uint32_t *first_id);
int  fec_cuda_softbuffer_reset(fec_cuda_ctx_t *ctx, uint32_t id);           /* srslte_softbuffer_rx_reset */
/* srslte_softbuffer_rx_free (softbuffer.c:102-126). Soft buffers belong to the DEVICE, not to the context: any context of
 * the device can decode into any id (the reference allocates them in the MAC thread and decodes from PHY worker threads,
 * one srslte_sch_t each). A single buffer (n == 1 allocation) returns to a free list and lives as long as the process;
 * buffers of a bulk allocation (n > 1) are released when the allocating context is destroyed — freeing one is a no-op. */
int  fec_cuda_softbuffer_free(fec_cuda_ctx_t *ctx, uint32_t id);
/* debugging / parity: copy one CB slot out in the REFERENCE soft-buffer layout (18600 int16, rm_turbo.c:250-264) */
int  fec_cuda_softbuffer_read_ref_layout(fec_cuda_ctx_t *ctx, uint32_t id, uint32_t cb, uint32_t K, int16_t *out18600);

/* ---- the hot path ------------------------------------------------------------------------------------ */
/* rate de-matching + HARQ combine -> turbo decode with CRC early stop -> TB assembly + CRC24A, for n TBs.
 * Synchronous: returns when all outputs are in place. Per-TB results in tbs[i].ret; the call itself
 * returns 0 unless the batch could not be run at all. */
int  fec_cuda_decode_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags);

/* The same call split in two, for a caller that has the next batch ready before the previous one is done (a PHY worker
 * per subframe, as srsENB's phch_worker pool): submit enqueues the batch and returns; wait(ticket) blocks until that batch
 * and every earlier one is complete and their outputs are in place. fec_cuda_decode_tbs == submit + wait. Up to two
 * batches in flight overlap one batch's ragged tail (the few code blocks that run all their half-iterations) with the
 * next batch's bulk. tbs[] and every buffer it points to must stay valid until the wait; two batches in flight must
 * not share soft buffers. One submitting thread per context. */
int  fec_cuda_submit_tbs(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, uint32_t n, uint32_t flags, uint64_t *ticket);
int  fec_cuda_wait(fec_cuda_ctx_t *ctx, uint64_t ticket);

/* ---- front end: transport blocks given as equalised symbols (SURVEY 8f rows f1 + f2) ------------------------------
 * Replaces, per codeword, the calls between the equaliser and decode_tb:
 *   DL  srslte_pdsch_codeword_decode (pdsch.c:778-834): srslte_demod_soft_demodulate_s (demod_soft.c:437-456),
 *       srslte_scrambling_s_offset (scrambling.c:49-52) with the sequence of srslte_sequence_pdsch (sequences.c:64-66,
 *       sequence.c:54-86), csi_correction (pdsch.c:675-776), srslte_dlsch_decode2;
 *   UL  srslte_pusch_decode from the demodulator on (pusch.c:617-657): demodulation, HARQ-ACK positions zeroed
 *       (sch.c:913-916), descrambling, ulsch_deinterleave (sch.c:860-881), decode_tb on g_bits + Q'_cqi*Qm.
 * One fused kernel writes the int16 LLRs to HBM; the host ships 8 bytes per resource element instead of 2*Qm. The
 * LLRs are bit-identical with the reference's AVX2 build (rounding of its SIMD bodies and scalar tails, the CSI weights
 * of its SSE loop, g[0] overwritten by the last RI position). In 8-bit LLR mode (fec_cuda_set_llr_8bit) the downlink
 * chain is the reference's int8 one and the LLRs are int8. */
typedef struct {
  const void     *symbols;      /* nof_re complex floats (cf_t), 16-byte aligned: q->d[cw] (pdsch.c:797) / q->d (pusch.c:620);
                                   host, or device with FEC_CUDA_F_DEVICE_IO */
  const float    *csi;          /* DL: NULL (csi_enabled = false) or nof_re weights q->csi[cw]; same memory space as symbols */
  uint32_t        nof_re;       /* cfg->nbits.nof_re */
  uint32_t        mod_bits;     /* bits per symbol of grant.mcs.mod: 1, 2, 4 or 6 (tb.Qm stays "Qm x Nl" as decode_tb gets it) */
  uint32_t        c_init;       /* scrambling seed: fec_cuda_cinit_pdsch / fec_cuda_cinit_pusch */
  uint32_t        nof_symb;     /* 0: downlink. Uplink: cfg->nbits.nof_symb = columns of the channel interleaver */
  /* uplink control information multiplexed with the data (HOST arrays in either mode): q-bit positions as
   * srslte_uci_decode_ack_ri leaves them in ack_ri_bits[].position (uci.c:499-546) */
  const uint32_t *ack_pos;      /* zeroed before descrambling */
  const uint32_t *ri_pos;       /* taken out of the de-interleaved stream */
  uint32_t        nof_ack_pos, nof_ri_pos;
  uint32_t        nof_cqi_bits; /* Q'_cqi * Qm: LLRs at the head of g that belong to the CQI (sch.c:951-962); decode_tb starts after them */
  uint32_t        reserved;
  int16_t        *uci_out;      /* out (needed if any of the three counts is non-zero): raw q LLRs (before descrambling) at
                                   ack_pos[..] then ri_pos[..] — the input of srslte_uci_decode_ack_ri — then the first
                                   nof_cqi_bits LLRs of g — the input of srslte_uci_decode_cqi_pusch */
} fec_cuda_sym_t;

/* tbs[i] as for fec_cuda_decode_tbs except that e_bits is ignored: DL needs nof_e_bits == nof_re * mod_bits, UL
 * nof_e_bits = G * Qm (sch.c:965-969). tbs[i].tbs == 0 runs the front end only (control-only PUSCH). */
int  fec_cuda_decode_syms(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags);
int  fec_cuda_submit_syms(fec_cuda_ctx_t *ctx, fec_cuda_tb_t *tbs, const fec_cuda_sym_t *syms, uint32_t n, uint32_t flags,
                          uint64_t *ticket);
/* device time of the front-end kernels of the last call (ms) */
int  fec_cuda_last_frontend_ms(fec_cuda_ctx_t *ctx, float *ms);
/* front end alone: LLRs as decode_tb would receive them (e_bits for DL, g_bits for UL), nof_re*mod_bits int16 (int8 in
 * 8-bit LLR mode: llr_out then points to int8_t), host pointers; drop-in body of srslte_demod_soft_demodulate_s + srslte_scrambling_s_offset (+ csi / de-interleaver) */
int  fec_cuda_frontend(fec_cuda_ctx_t *ctx, const fec_cuda_sym_t *sym, int16_t *llr_out);
uint32_t fec_cuda_cinit_pdsch(uint32_t rnti, uint32_t codeword, uint32_t nslot, uint32_t cell_id);   /* sequences.c:64-66 */
uint32_t fec_cuda_cinit_pusch(uint32_t rnti, uint32_t nslot, uint32_t cell_id);                       /* sequences.c:71-73 */

/* ---- transmit side -------------------------------------------------------------------------------
 * The data part of srslte_dlsch_encode / srslte_ulsch_encode, i.e. encode_tb_off (sch.c:188-296): CRC-24A over the
 * transport block, segmentation, CRC-24B per code block, turbo encoding (srslte_tcod_encode_lut, turbocoder.c:196-330)
 * and rate matching (srslte_rm_turbo_tx_lut, rm_turbo.c:332-376) for a batch of transport blocks. Stateless: a
 * retransmission (rv != 0) is encoded from the data again instead of from srslte_softbuffer_tx_t; the bits are the same. */
typedef struct {
  uint32_t       tbs;          /* transport block size in bits (multiple of 8) */
  uint32_t       Qm;           /* modulation order x layers, as encode_tb_off receives it */
  uint32_t       rv;           /* redundancy version 0..3 */
  uint32_t       nof_e_bits;   /* cfg->nbits[].nof_bits */
  const uint8_t *data;         /* tbs/8 bytes; host, or device with FEC_CUDA_F_DEVICE_IO */
  uint8_t       *e_bits;       /* out: (nof_e_bits + 7) / 8 bytes, packed MSB first as srslte_dlsch_encode leaves them; a device
                                  pointer must be 4-byte aligned and have room up to the next multiple of 4 bytes */
  int32_t        ret;          /* 0, -1 (filler bits / too many code blocks, sch.c:202-210), -2 invalid */
  uint32_t       reserved;
} fec_cuda_txtb_t;
int  fec_cuda_encode_tbs(fec_cuda_ctx_t *ctx, fec_cuda_txtb_t *tbs, uint32_t n, uint32_t flags);

/* bare turbo decoding of n code blocks from natural-layout LLRs (host pointers) */
int  fec_cuda_decode_cbs(fec_cuda_ctx_t *ctx, fec_cuda_cb_t *cbs, uint32_t n, uint32_t flags);

/* standalone rate de-matching into a HOST soft buffer in the reference layout — drop-in body of
 * srslte_rm_turbo_rx_lut_ (rm_turbo.c:378-430); computed on the device. In 8-bit LLR mode (fec_cuda_set_llr_8bit) `in` and
 * `soft` point to int8 arrays: the body of srslte_rm_turbo_rx_lut_8bit (rm_turbo.c:432-469). Of `soft` only the extent
 * the layout uses is read and written: 3*(K+32)+12 entries with sb_layout and K > 400, else 3K+12 (what the reference's own
 * rm_turbo_test.c:52,186 allocates). */
int  fec_cuda_rm_rx(fec_cuda_ctx_t *ctx, const int16_t *in, uint32_t E, uint32_t cb_idx, uint32_t rv,
                    int sb_layout, int16_t *soft);

/* The same operation for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it). NO COUNTERPART IN THE REFERENCE. Takes a device
 * ordinal instead of a context and allocates nothing, on the host or on the device: all device scratch (staged LLRs,
 * the soft buffer in both layouts, the per-K rank table, the descriptor) is one static __device__ array of at most
 * FEC_CUDA_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit. Host pointers, 16-bit LLRs
 * only; of `soft` only the extent the layout uses is read and written (sb_layout and K > 400: 3*(K+32)+12 entries, else
 * 3K+12). E > FEC_CUDA_NOALLOC_MAX_E is rejected with FEC_CUDA_EINVAL and a message on stderr. Synchronous; calls are
 * serialised by a mutex because they all share the array (it keeps the last call's contents; what must start at zero is
 * cleared per call) — GPU work that overlaps a call on the same array by any other route would race. */
#define FEC_CUDA_NOALLOC_MAX_E          32768u
#define FEC_CUDA_NOALLOC_SCRATCH_BYTES  (176u * 1024u)
int  fec_cuda_rm_rx_noalloc(int device, const int16_t *in, uint32_t E, uint32_t cb_idx, uint32_t rv,
                            int sb_layout, int16_t *soft);

/* ---- test hooks (LLR dump of the last half-iteration of fec_cuda_decode_cbs; not a product feature) ---- */
int  fec_cuda_debug_llr_enable(fec_cuda_ctx_t *ctx, uint32_t n_code_blocks);
int  fec_cuda_debug_llr_read(fec_cuda_ctx_t *ctx, int16_t *out /* n x 6144 */, uint32_t n);

/* ---- host-side helpers (pure integer bookkeeping, no device) ---------------------------------------- */
int  fec_cuda_cb_index(uint32_t K);               /* srslte_cbsegm_cbindex   cbsegm.c:116-127 */
int  fec_cuda_cb_size(uint32_t idx);              /* srslte_cbsegm_cbsize    cbsegm.c:134-141 */
int  fec_cuda_subblocks(uint32_t K);              /* srslte_tdec_autoimp_get_subblocks turbodecoder.c:364-376 */
int  fec_cuda_cbsegm(uint32_t tbs, uint32_t out9[9]);   /* srslte_cbsegm: {F,C,K1,K2,K1_idx,K2_idx,C1,C2,tbs} */

#ifdef __cplusplus
}
#endif
#endif
