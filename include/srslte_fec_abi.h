/*
 * srslte_fec_abi.h — sizes and field offsets of the reference structs that libsrslte_fec_cuda touches by address
 * (x86-64, srsLTE 18.09, AVX2 defines). Constants only, no typedefs, so that this header can be
 * included next to the reference's own headers: oracle/abi_check.c does exactly that and _Static_asserts every line against
 * sizeof / offsetof of the real types (built by `make -C oracle linktest` wherever /root/reference is mounted).
 */
#ifndef SRSLTE_FEC_ABI_H
#define SRSLTE_FEC_ABI_H

#define FEC_ABI_MAX_CODEWORDS            2        /* SRSLTE_MAX_CODEWORDS, phy_common.h */

/* srslte_tdec_t, lib/include/srslte/phy/fec/turbodecoder.h:68-100 */
#define FEC_ABI_SIZEOF_TDEC              18264
#define FEC_ABI_TDEC_MAX_LONG_CB         0
#define FEC_ABI_TDEC_N_ITER              18256

/* srslte_softbuffer_rx_t, lib/include/srslte/phy/fec/softbuffer.h:42-48 */
#define FEC_ABI_SIZEOF_SOFTBUFFER_RX     40
#define FEC_ABI_SOFTBUFFER_RX_MAX_CB     0
#define FEC_ABI_SOFTBUFFER_RX_BUFFER_F   8
#define FEC_ABI_SOFTBUFFER_RX_DATA       16
#define FEC_ABI_SOFTBUFFER_RX_CB_CRC     24
#define FEC_ABI_SOFTBUFFER_RX_TB_CRC     32

/* srslte_sch_t, lib/include/srslte/phy/phch/sch.h:57-81 */
#define FEC_ABI_SIZEOF_SCH               57648
#define FEC_ABI_SCH_MAX_ITERATIONS       0
#define FEC_ABI_SCH_NOF_ITERATIONS       4
#define FEC_ABI_SCH_LLR_IS_8BIT          8
#define FEC_ABI_SCH_DECODER              27728

/* srslte_cbsegm_t, lib/include/srslte/phy/fec/cbsegm.h:38-48 */
#define FEC_ABI_SIZEOF_CBSEGM            36
#define FEC_ABI_CBSEGM_F                 0
#define FEC_ABI_CBSEGM_C                 4
#define FEC_ABI_CBSEGM_TBS               32

/* srslte_pdsch_cfg_t, lib/include/srslte/phy/phch/pdsch_cfg.h */
#define FEC_ABI_SIZEOF_PDSCH_CFG         404
#define FEC_ABI_PDSCH_CFG_CB_SEGM        0
#define FEC_ABI_PDSCH_CFG_GRANT          72
#define FEC_ABI_PDSCH_CFG_NBITS          344
#define FEC_ABI_PDSCH_CFG_RV             376
#define FEC_ABI_PDSCH_CFG_NOF_LAYERS     388

/* srslte_ra_dl_grant_t and srslte_ra_nbits_t, lib/include/srslte/phy/phch/ra.h */
#define FEC_ABI_SIZEOF_RA_DL_GRANT       272
#define FEC_ABI_DL_GRANT_QM              224
#define FEC_ABI_DL_GRANT_TB_EN           260
#define FEC_ABI_SIZEOF_RA_NBITS          16
#define FEC_ABI_RA_NBITS_NOF_BITS        8

#endif
