/*
 * fec_kernels.cuh — device-side structures and kernel launch prototypes of libsrslte_fec_cuda.
 * sm_100a only. See DESIGN.md for the data layout and the roofline of each kernel.
 */
#ifndef FEC_KERNELS_CUH
#define FEC_KERNELS_CUH
#include <cuda_runtime.h>
#include <stdint.h>
#include "fec_tables.h"

#define FEC_WIN_THREADS   128     /* MAP lanes per CTA in the windowed kernel = slots x P */
#define FEC_WIN_CHUNK     16      /* beta checkpoint spacing S (trellis steps) */
#define FEC_F_NO_EARLY    1u
#define FEC_PK_WARPS      4       /* packed kernel: warps per CTA (= tensor-memory lane quarters) */
#define FEC_PK_GMAX       2000u   /* packed kernel: largest max|x| + max|y| for which int16 saturation is excluded a priori */
#define FEC_NQUEUE        5       /* device counters per K: packed queue, exact queue, exact count, certificate queue, certificate count */
#define FEC_ECONT_WORDS   3200    /* scratch words per handed-over code block: >= (P/2) * Lw + 1 for every K (3105 at K = 6144) */
#define FEC_PK_GMAX_TRK   4500u   /* ... and the largest for which the run-time range certificate can still hold (10000 + 5G <= 32767) */

/* one code block of a decode launch */
struct FecCbDesc {
  const int16_t *soft;     /* device-layout soft buffer slot (3 streams + 12 tail) */
  uint8_t       *out;      /* K/8 decoded bytes */
  uint32_t       crc_sel;  /* 0: CRC24A powers, 1: CRC24B powers */
  uint32_t       max_its;  /* half-iteration cap */
  uint32_t       flags;    /* FEC_F_NO_EARLY */
  uint32_t       status;   /* index into the status array */
};

struct FecCbStatus { uint32_t its; uint32_t crc_ok; };

/* everything a decode launch needs for one code-block size K */
struct FecDecodeArgs {
  fec_geom_t        g;
  const uint32_t   *step_tab;   /* [2][NL*Lp] (nat<<16 | eaddr) */
  const uint32_t   *crcpow[2];  /* R24A, R24B */
  const FecCbDesc  *cbs;
  FecCbStatus      *status;
  unsigned int     *queue;      /* work-queue counter, zeroed before launch */
  uint32_t          ncb;
  uint32_t          slots;      /* code blocks in flight per CTA */
  int16_t          *dbg_llr;    /* optional [ncb][6144]: LLRs of the last half-iteration, natural order (tests only) */
  /* packed fast path (fec_tdec_packed.cu) */
  const uint32_t   *pair_addr;  /* decoder-2 address words, (P/2) x Lp */
  const uint2      *pair_crc[2];/* CRC syndrome words per polynomial, [2 parities][(P/2) x Lp] */
  uint32_t          status_off; /* status / soft_max index of cbs[0] (cbs[q].status == status_off + q) */
  const uint32_t   *soft_max;   /* per status slot: max |soft-buffer entry| of the code block (rate de-matching / extract kernel) */
  uint32_t         *exact_list; /* out: code blocks (indices into cbs) outside the proven range -> exact kernel */
  unsigned int     *exact_count;
  uint32_t         *trk_list;   /* out: code blocks beyond the a-priori range -> certificate variant of the packed kernel */
  unsigned int     *trk_count;
  unsigned int     *trk_queue;  /* work-queue counter of the certificate kernel */
  uint32_t         *trk_cont;   /* per trk_list entry: 0 = decode from the soft buffer, else half-iterations done | G << 8 and ... */
  uint32_t         *econt;      /* ... the extrinsic array of the block in scratch slot [entry index] (Es words each) */
  uint32_t          econt_cap;  /* scratch slots */
  int               llr8;       /* 1: int8 decoders of the 8-bit LLR path (P = 16 / 32, exact kernel only) */
  /* exact kernels: optional indirect work list (NULL = all ncb code blocks) */
  const uint32_t   *list;
  const unsigned int *list_count;
};

/* one code block of a rate-dematching launch */
#define FEC_RM_THREADS 256
struct FecRmDesc {
  const int16_t  *e;        /* E transmitted LLRs of this CB */
  int16_t        *soft;     /* device-layout slot, read-modify-write */
  const uint16_t *rank;     /* per-K table in device layout: soft offset -> 2 x entry of the NULL-free circular buffer (byte offset into the tile; 2N: none) */
  uint32_t        E;
  uint32_t        N;        /* 3K+12 */
  uint32_t        off;      /* start rank for this rv */
  uint32_t        n_out;    /* 3*SS+16 (multiple of 8) */
  uint32_t        new_data; /* 1: overwrite (fresh HARQ process), 0: accumulate */
  uint32_t        slot;     /* status slot: soft_max[slot] receives max |soft-buffer entry| */
};

/* natural-layout [s p p']xK+12 input -> device-layout slot (srslte_tdec_* level entry) */
struct FecExtractDesc { const int16_t *in; int16_t *soft; fec_geom_t g; uint32_t slot; };

/* one transport block of the assembly/CRC24A launch */
struct FecTbDesc {
  uint8_t        *data;        /* out: device TB buffer */
  const uint8_t  *cb_bytes[13];/* decoded bytes per CB (fresh decode or cache) */
  uint32_t        C, rlen[13], K[13];
  uint32_t        tbs;
  uint32_t        all_ok;      /* host-known part of AND(cb_crc): CBs skipped because already OK */
  uint32_t        status0;     /* index of CB0's status entry, or per-CB status index array below */
  int32_t         st_idx[13];  /* status index per CB, -1 if CB not decoded in this call (was OK already) */
  uint8_t        *cache[13];   /* softbuffer->data[cb]: decoded-CB cache, updated when a CB newly passes */
  int32_t        *ret;         /* out: 0 / -1 */
};

/* one transport block of the front-end launch (fec_frontend.cu): equalised symbols -> int16 LLRs in decode_tb order */
#define FEC_FE_NBLK 512          /* jump-table entries: 512 x 256 = 131072 scrambling bits (MAX_SEQ_LEN, sequence.c:38) */
struct FecFeDesc {
  const float2   *sym;       /* nof_re equalised symbols, 16-byte aligned */
  const float    *csi;       /* DL: NULL or nof_re weights */
  int16_t        *e;         /* out: nof_re*mod LLRs (DL: e_bits; UL: g_bits) */
  const uint32_t *ack_ord;   /* UL: de-interleaved bit indices of the HARQ-ACK positions, ascending */
  const uint32_t *ri_ord;    /* UL: de-interleaved bit indices of the RI positions, ascending */
  const uint32_t *uci_pos;   /* UL: q positions to hand back raw (ACK then RI, caller's order) */
  int16_t        *uci_out;   /* UL: out, nuci raw LLRs then the first n_cqi LLRs of g */
  uint32_t        nof_re, mod, c_init;
  uint32_t        cols, rows;          /* UL: channel interleaver matrix (cols = N_pusch_symbs); DL: cols = 0 */
  uint32_t        nack, nri, nuci, n_cqi;
  uint32_t        last_ri_pos;         /* UL: largest RI q position (value that ends up in g[0]) */
  uint32_t        pad;
};
cudaError_t fec_launch_frontend(const FecFeDesc *descs, float *csi_max, const uint32_t *x2_masks, const uint32_t *x1_states,
                                uint32_t ntb, uint32_t max_re, int any_csi, int any_uci, int kinds /* 1: DL present, 2: UL present */, cudaStream_t st,
                                int llr8 = 0 /* 1: int8 LLRs out (srslte_demod_soft_demodulate_b chain, downlink only) */);
extern "C" void fec_gen_gold_tables(uint32_t *x2_masks, uint32_t *x1_states);

size_t fec_win_smem_bytes(const fec_geom_t *g, uint32_t slots, int llr8);
uint32_t fec_win_pick_slots(const fec_geom_t *g, int llr8);
/* ---- transmit side (fec_tx.cu) ---- */
#define FEC_TX_W 194              /* words per d-stream in shared memory: K + 4 <= 6148 bits */
struct FecTxTbDesc { const uint8_t *data; uint32_t tbs; uint8_t *crc3; uint32_t *e_out; uint32_t e_words; };   /* e_out: zeroed by the CRC kernel */
struct FecTxCbDesc {
  const uint8_t  *src;      /* payload bytes of this code block inside the transport block (device) */
  const uint8_t  *tbcrc;    /* the 3 CRC-24A bytes of the transport block (follow the payload of its last code block) */
  uint8_t        *e_out;    /* packed output of the transport block, 4-byte aligned, zeroed */
  const uint16_t *txsrc;    /* per K: NULL-free circular-buffer position -> bit index stream * 32 * FEC_TX_W + i */
  uint32_t        nsrc, ntb;/* payload bytes from src, CRC-24A bytes behind them (0 or 3) */
  uint32_t        cbcrc;    /* 1: CRC-24B over the first K - 24 bits goes into the last 24 */
  uint32_t        K, f1, f2;
  uint32_t        wp, E;    /* output bit offset inside e_out and number of bits (encode_tb_off: wp, n_e) */
  uint32_t        off, N;   /* start position for this rv and length 3K + 12 of the NULL-free circular buffer */
};
cudaError_t fec_launch_tx(const FecTxTbDesc *tbd, uint32_t ntb, const FecTxCbDesc *cbd, uint32_t ncb,
                          const uint32_t *r24a, const uint32_t *r24b, cudaStream_t st);

/* stages of a windowed decode launch (classes P = 8, 16 with hand-over lists): A the a-priori-range packed kernel, B its
 * run-time-certificate variant on what A handed over, C the exact kernel on what A and B handed over. A caller may leave B
 * and C out and launch them later, once it has seen non-zero hand-over counts (fec_cuda_api.cu does, so that the common
 * case is one launch); C must come after B. Ignored by the single-launch paths (P = 0, forced exact, LLR dump). */
#define FEC_ST_A 1
#define FEC_ST_B 2
#define FEC_ST_C 4
#define FEC_ST_ALL 7
cudaError_t fec_launch_decode(const FecDecodeArgs &a, int sm_count, cudaStream_t st, int stages);
cudaError_t fec_launch_decode_packed(const FecDecodeArgs &a, int sm_count, cudaStream_t st, int stages);
size_t fec_pk_smem_bytes(const fec_geom_t *g);
cudaError_t fec_pk_init(void);
cudaError_t fec_launch_rm(const FecRmDesc *descs, uint32_t *soft_max, uint32_t n, uint32_t max_n_out, cudaStream_t st, int llr8 = 0, uint32_t max_tile = 0);
cudaError_t fec_launch_extract(const FecExtractDesc *descs, uint32_t *soft_max, uint32_t n, cudaStream_t st);
cudaError_t fec_launch_tb_assemble(const FecTbDesc *descs, const FecCbStatus *status, const uint32_t *r24a,
                                   uint32_t n, cudaStream_t st);
cudaError_t fec_launch_soft_to_ref(const int16_t *soft_dev, int16_t *ref_dev, fec_geom_t g, int sb_layout, int keep, cudaStream_t st);
cudaError_t fec_launch_ref_to_soft(const int16_t *ref_dev, int16_t *soft_dev, fec_geom_t g, int sb_layout, cudaStream_t st);
cudaError_t fec_kernels_init(void);
#endif
