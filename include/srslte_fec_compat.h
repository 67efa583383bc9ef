/*
 * srslte_fec_compat.h — the reference's OWN symbols for the decode path, exported by libsrslte_fec_cuda.so as
 * batch-of-one wrappers over the batched API of srslte_fec_cuda.h.  Linking this library in place of
 * lib/src/phy/fec/{turbodecoder,turbodecoder_sse,turbodecoder_gen}.o, the receive half of rm_turbo.o and the rx
 * functions of softbuffer.o makes the reference's callers and tests run on the GPU unchanged.
 *
 * The structs below are ABI VIEWS of the reference structs: same size, same offsets for every field a caller
 * outside the replaced objects touches (probed with sizeof/offsetof on the reference headers);
 * everything else is private storage of this library.  A maintainer building inside the srsLTE tree includes the
 * reference headers instead (the prototypes are identical) and keeps the static asserts of csrc/fec_compat.cu.
 *
 * 16-bit and 8-bit LLR paths: srslte_tdec_iteration_8bit / srslte_tdec_run_all_8bit and
 * srslte_sch_decode_tb_cuda_8bit, srslte_rm_turbo_rx_lut_8bit.
 * No CPU fallback: without a usable CUDA device every init returns SRSLTE_ERROR and every decode returns -1.
 */
#ifndef SRSLTE_FEC_COMPAT_H
#define SRSLTE_FEC_COMPAT_H

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- turbo decoder object: lib/include/srslte/phy/fec/turbodecoder.h:68-100 (18264 bytes, caller-allocated) ---- */
typedef enum { SRSLTE_TDEC_AUTO = 0, SRSLTE_TDEC_GENERIC, SRSLTE_TDEC_SSE, SRSLTE_TDEC_SSE_WINDOW, SRSLTE_TDEC_AVX_WINDOW,
               SRSLTE_TDEC_SSE8_WINDOW, SRSLTE_TDEC_AVX8_WINDOW } srslte_tdec_impl_type_t;      /* turbodecoder.h:56-64 */

typedef struct {
  uint32_t max_long_cb;          /* @0, as in the reference */
  uint32_t fec_magic;            /* private from here on */
  void    *fec_reserved;         /* (decode contexts are per calling thread, not per object) */
  int16_t *fec_input;            /* private copy of the code block's LLRs, natural layout (taken at n_iter == 0) */
  uint32_t fec_K;                /* current code block size */
  uint32_t fec_force_not_sb;
  uint32_t fec_manual_type;      /* 0: AUTO, else forced sub-block class + 1 (srslte_tdec_init_manual) */
  uint8_t  fec_private[18256 - 36];
  uint32_t n_iter;               /* @18256, as in the reference */
  uint32_t fec_pad;
} srslte_tdec_t;

/* turbodecoder.h:102-140 */
int      srslte_tdec_init(srslte_tdec_t *h, uint32_t max_long_cb);
int      srslte_tdec_init_manual(srslte_tdec_t *h, uint32_t max_long_cb, srslte_tdec_impl_type_t dec_type);
void     srslte_tdec_free(srslte_tdec_t *h);
void     srslte_tdec_force_not_sb(srslte_tdec_t *h);
int      srslte_tdec_new_cb(srslte_tdec_t *h, uint32_t long_cb);
int      srslte_tdec_get_nof_iterations(srslte_tdec_t *h);
uint32_t srslte_tdec_autoimp_get_subblocks(uint32_t nof_bits);
uint32_t srslte_tdec_autoimp_get_subblocks_8bit(uint32_t nof_bits);
void     srslte_tdec_iteration(srslte_tdec_t *h, int16_t *input, uint8_t *output);
int      srslte_tdec_run_all(srslte_tdec_t *h, int16_t *input, uint8_t *output, uint32_t nof_iterations, uint32_t long_cb);
void     srslte_tdec_iteration_8bit(srslte_tdec_t *h, int8_t *input, uint8_t *output);                       /* turbodecoder.c:536-543 */
int      srslte_tdec_run_all_8bit(srslte_tdec_t *h, int8_t *input, uint8_t *output, uint32_t nof_iterations, uint32_t long_cb); /* turbodecoder.c:545-563 */

/* ---- rate de-matching: lib/include/srslte/phy/fec/rm_turbo.h:51-94 ---- */
void srslte_rm_turbo_gentables(void);                     /* tables are built per K on first use: nothing to do */
void srslte_rm_turbo_free_tables(void);
int  srslte_rm_turbo_rx_lut(int16_t *input, int16_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx);
int  srslte_rm_turbo_rx_lut_(int16_t *input, int16_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx, bool enable_input_tdec);
int  srslte_rm_turbo_rx_lut_8bit(int8_t *input, int8_t *output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx);   /* rm_turbo.c:432-469 */

/* ---- HARQ soft buffer: lib/include/srslte/phy/fec/softbuffer.h:42-48,57-75 (40 bytes, caller-allocated) ---- */
typedef struct {
  uint32_t  max_cb;              /* @0 */
  int16_t **buffer_f;            /* @8  reference: per-CB host buffers.  Here: opaque handle of the device soft buffer */
  uint8_t **data;                /* @16 reference: cached CB payloads.    Here: unused (the cache lives on the device) */
  bool     *cb_crc;              /* @24 per-CB CRC state, host array of max_cb flags, kept up to date by the decode calls */
  bool      tb_crc;              /* @32 */
} srslte_softbuffer_rx_t;

int  srslte_softbuffer_rx_init(srslte_softbuffer_rx_t *q, uint32_t nof_prb);
void srslte_softbuffer_rx_free(srslte_softbuffer_rx_t *q);
void srslte_softbuffer_rx_reset(srslte_softbuffer_rx_t *q);
void srslte_softbuffer_rx_reset_tbs(srslte_softbuffer_rx_t *q, uint32_t tbs);
void srslte_softbuffer_rx_reset_cb(srslte_softbuffer_rx_t *q, uint32_t nof_cb);

/* ---- decode_tb (lib/src/phy/phch/sch.c:437-486): the body the patched sch.c calls, see INTEGRATION.md ----
 * Arguments are those of decode_tb with the structs flattened: cb_segm->tbs, Qm (already multiplied by the layer
 * count as srslte_dlsch_decode2 does, sch.c:508-516), rv, nof_e_bits, e_bits, data; max_iterations is
 * q->max_iterations, *nof_iterations receives q->nof_iterations (sch.c:420).  Returns 0 / -1 / -2 like decode_tb. */
int  srslte_sch_decode_tb_cuda(srslte_softbuffer_rx_t *softbuffer, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                               int16_t *e_bits, uint8_t *data, uint32_t max_iterations, uint32_t *nof_iterations);

/* decode_tb with q->llr_is_8bit (sch.c:344-348,362-364): e_bits are int8 */
int  srslte_sch_decode_tb_cuda_8bit(srslte_softbuffer_rx_t *softbuffer, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                                    int8_t *e_bits, uint8_t *data, uint32_t max_iterations, uint32_t *nof_iterations);

/* ---- srslte_dlsch_decode / srslte_dlsch_decode2 (sch.h:108-121, sch.c:488-517): the reference's DL-SCH entry points on the
 * reference's own caller-allocated structs — q is a srslte_sch_t *, cfg a srslte_pdsch_cfg_t * (declared void * here because
 * this header does not carry those types; same ABI). Only the fields decode_tb reads are touched, through the offsets of
 * srslte_fec_abi.h. A build that takes these from the library compiles them out of sch.c (oracle/patch_sch.py shows how);
 * srslte_ulsch_decode / srslte_ulsch_uci_decode stay in sch.c (they de-interleave and decode UCI first, sch.c:883-985) and
 * reach the library through decode_tb -> srslte_sch_decode_tb_cuda. */
int  srslte_dlsch_decode(void *q, void *cfg, srslte_softbuffer_rx_t *softbuffer, int16_t *e_bits, uint8_t *data);
int  srslte_dlsch_decode2(void *q, void *cfg, srslte_softbuffer_rx_t *softbuffer, int16_t *e_bits, uint8_t *data, int tb_idx);

/* ---- encode_tb_off (sch.c:188-296): the body a patched sch.c calls on the transmit side. data = NULL (the reference
 * re-reads its tx soft buffer for a retransmission) is not supported: pass the transport block again. w_offset = 0. */
int  srslte_sch_encode_tb_cuda(uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, uint8_t *data, uint8_t *e_bits);

#ifdef __cplusplus
}
#endif
#endif
