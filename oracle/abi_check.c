/*
 * abi_check.c — TEST INFRASTRUCTURE. Compiled (never run) by `make -C oracle linktest` with the reference's own headers on
 * the include path: every constant of include/srslte_fec_abi.h — the sizes and offsets libsrslte_fec_cuda relies on when it
 * is handed the reference's caller-allocated structs — against sizeof / offsetof of the real types.
 */
#include <stddef.h>
#include "srslte/srslte.h"
#include "../include/srslte_fec_abi.h"

#define SZ(type, c)      _Static_assert(sizeof(type) == (c), "sizeof(" #type ") != " #c)
#define OF(type, f, c)   _Static_assert(offsetof(type, f) == (c), "offsetof(" #type ", " #f ") != " #c)

_Static_assert(SRSLTE_MAX_CODEWORDS == FEC_ABI_MAX_CODEWORDS, "SRSLTE_MAX_CODEWORDS");
SZ(srslte_tdec_t, FEC_ABI_SIZEOF_TDEC);
OF(srslte_tdec_t, max_long_cb, FEC_ABI_TDEC_MAX_LONG_CB);
OF(srslte_tdec_t, n_iter, FEC_ABI_TDEC_N_ITER);
SZ(srslte_softbuffer_rx_t, FEC_ABI_SIZEOF_SOFTBUFFER_RX);
OF(srslte_softbuffer_rx_t, max_cb, FEC_ABI_SOFTBUFFER_RX_MAX_CB);
OF(srslte_softbuffer_rx_t, buffer_f, FEC_ABI_SOFTBUFFER_RX_BUFFER_F);
OF(srslte_softbuffer_rx_t, data, FEC_ABI_SOFTBUFFER_RX_DATA);
OF(srslte_softbuffer_rx_t, cb_crc, FEC_ABI_SOFTBUFFER_RX_CB_CRC);
OF(srslte_softbuffer_rx_t, tb_crc, FEC_ABI_SOFTBUFFER_RX_TB_CRC);
SZ(srslte_sch_t, FEC_ABI_SIZEOF_SCH);
OF(srslte_sch_t, max_iterations, FEC_ABI_SCH_MAX_ITERATIONS);
OF(srslte_sch_t, nof_iterations, FEC_ABI_SCH_NOF_ITERATIONS);
OF(srslte_sch_t, llr_is_8bit, FEC_ABI_SCH_LLR_IS_8BIT);
OF(srslte_sch_t, decoder, FEC_ABI_SCH_DECODER);
SZ(srslte_cbsegm_t, FEC_ABI_SIZEOF_CBSEGM);
OF(srslte_cbsegm_t, F, FEC_ABI_CBSEGM_F);
OF(srslte_cbsegm_t, C, FEC_ABI_CBSEGM_C);
OF(srslte_cbsegm_t, tbs, FEC_ABI_CBSEGM_TBS);
SZ(srslte_pdsch_cfg_t, FEC_ABI_SIZEOF_PDSCH_CFG);
OF(srslte_pdsch_cfg_t, cb_segm, FEC_ABI_PDSCH_CFG_CB_SEGM);
OF(srslte_pdsch_cfg_t, grant, FEC_ABI_PDSCH_CFG_GRANT);
OF(srslte_pdsch_cfg_t, nbits, FEC_ABI_PDSCH_CFG_NBITS);
OF(srslte_pdsch_cfg_t, rv, FEC_ABI_PDSCH_CFG_RV);
OF(srslte_pdsch_cfg_t, nof_layers, FEC_ABI_PDSCH_CFG_NOF_LAYERS);
SZ(srslte_ra_dl_grant_t, FEC_ABI_SIZEOF_RA_DL_GRANT);
OF(srslte_ra_dl_grant_t, Qm, FEC_ABI_DL_GRANT_QM);
OF(srslte_ra_dl_grant_t, tb_en, FEC_ABI_DL_GRANT_TB_EN);
SZ(srslte_ra_nbits_t, FEC_ABI_SIZEOF_RA_NBITS);
OF(srslte_ra_nbits_t, nof_bits, FEC_ABI_RA_NBITS_NOF_BITS);
/* element types the library reads through those offsets */
_Static_assert(sizeof(((srslte_ra_dl_grant_t *)0)->Qm[0]) == 4 && sizeof(((srslte_pdsch_cfg_t *)0)->rv[0]) == 4 &&
               sizeof(((srslte_ra_dl_grant_t *)0)->tb_en[0]) == 1 && sizeof(((srslte_sch_t *)0)->llr_is_8bit) == 1, "field widths");
