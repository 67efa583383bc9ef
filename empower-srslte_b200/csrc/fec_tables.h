/*
 * fec_tables.h — host-side integer bookkeeping of the decode path (plain C, no device code):
 * code-block sizes and segmentation, decoder class per K, device data-layout geometry, and the
 * generators of the per-K lookup tables the kernels read from HBM.
 */
#ifndef FEC_TABLES_H
#define FEC_TABLES_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEC_NOF_K        188
#define FEC_MAX_K        6144
#define FEC_R24A_LEN     (13 * 6144 + 64)   /* CRC power table long enough for the largest transport block */

/* Device layout of one code block ("geometry"). The reference's sub-block-interleaved soft buffer
 * (rm_turbo.c:250-264: element k of sub-block d at k*P+d) becomes, on the device, PAIR rows: sub-blocks
 * (lanes) 2j and 2j+1 share 32-bit words (lane 2j | lane 2j+1 << 16), so that one thread walks two sub-blocks
 * with packed s16x2 arithmetic; the words of the P/2 pairs are interleaved in groups of 4 steps, so that the
 * 16-byte loads of the P/2 threads of a code block fall into ONE 128-byte line (coalesced):
 *   stream s (0 syst, 1 parity, 2 parity'), lane d, step k  ->  word W = ((k>>2)*(P/2) + (d>>1))*4 + (k&3),
 *   int16 index s*SS + 2*W + (d&1)
 *   (non-windowed class P=0: one lane, index s*SS + k);  12 tail LLRs at 3*SS.
 * Exact (32-bit) kernels keep the extrinsic array E in shared memory with lane stride LpE (LpE/2 odd:
 * conflict-free gathers); the packed kernel keeps E in pair rows of Lw words (Lw mod 8 == 4). */
typedef struct {
  int K;      /* code block size */
  int idx;    /* index in the 188-entry table */
  int P;      /* decoder class as the reference's AUTO mode picks it: 0 (K<=400), 8, 16 */
  int NL;     /* lanes = P, or 1 for the non-windowed class */
  int L;      /* K / NL */
  int Lp;     /* lane stride in HBM, multiple of 8 */
  int LpE;    /* lane stride of E in shared memory (exact kernels) */
  int Lw;     /* pair-row stride of E in words (packed kernel), Lw >= L, Lw mod 8 == 4 */
  int SS;     /* stream stride NL*Lp */
  int f1, f2; /* QPP parameters */
} fec_geom_t;

int  fec_k_index(uint32_t K);            /* first table index with K_i >= K, -1 if none (cbsegm.c:116-127) */
int  fec_k_size(uint32_t idx);           /* K_i or -1 */
int  fec_k_subblocks(uint32_t K);        /* 16 / 8 / 0 (turbodecoder.c:364-376, AVX2 build) */
int  fec_segm(uint32_t tbs, uint32_t out9[9]);   /* cbsegm.c:58-108 */
int  fec_k_subblocks8(uint32_t K);       /* 32 / 16 / 8 / 0: class of the 8-bit LLR path (turbodecoder.c:392-406, AVX2 build) */
int  fec_geom(uint32_t K, fec_geom_t *g);        /* -1 if K is not an LTE size */
int  fec_geom_p(uint32_t K, int P, fec_geom_t *g);   /* the same with the sub-block count given (8-bit path: P up to 32) */

/* step table, 2 parities x NL x Lp entries of (natural_position << 16 | E_address); pads = 0 */
void fec_gen_step_table(const fec_geom_t *g, uint32_t *tab);
/* packed kernel: address table of the INTERLEAVED decoder, (P/2) x Lp words: (E halfword index of lane 2j+1) << 16
 * | (E halfword index of lane 2j) at step k, E halfword index of natural bit (d',k') = 2*((d'>>1)*Lw + k') + (d'&1) */
void fec_gen_pair_addr_table(const fec_geom_t *g, uint32_t *tab);
/* packed kernel: CRC syndrome words per pair-step, [2 parities][(P/2) x Lp] uint2 = {R of lane 2j's bit, R of lane 2j+1's bit},
 * pair j, step k at uint2 index ((k>>1)*(P/2) + j)*2 + (k&1) (two steps = one 16-byte load, P/2 threads = one line),
 * with R = x^(K-1-nat+24) mod g(x) of the natural position decided at that step; crc_sel picks the polynomial */
void fec_gen_pair_crc_table(const fec_geom_t *g, uint32_t poly, uint32_t *tab);
/* rate-matching rank table in device soft-buffer layout (3*SS+16 entries of 2 x rank, pad = 2 x (3K+12)) and the 4 rv
 * start ranks: transmitted index of soft position o for rv is (rank[o] - off[rv]) mod (3K+12). */
void fec_gen_rm_rank(const fec_geom_t *g, uint16_t *rank, uint32_t off[4]);
/* rank table re-based to the start rank `off` of one redundancy version: 2 x index of the received LLR, pad = 2 x (3K+12) */
void fec_rm_rank_rebase(const fec_geom_t *g, const uint16_t *rank, uint32_t off, uint16_t *out);
/* natural coded position t (0..3K+11) -> device soft-buffer offset, and -> reference soft-buffer offset */
/* transmit side: NULL-free circular-buffer position (counted from w[0], as rank) -> stream * stream_bits + index in d_stream,
 * index < K + 4 (36.212 5.1.4.1.1-2); n = 3K + 12 entries */
void fec_gen_rm_txsrc(const fec_geom_t *g, uint32_t stream_bits, uint16_t *src);
int  fec_soft_offset_dev(const fec_geom_t *g, int t);
int  fec_soft_offset_ref(const fec_geom_t *g, int t, int sb_layout);
/* R[m] = x^(m+24) mod g(x): CRC of a message is the XOR of R[nbits-1-i] over its set bits i */
void fec_gen_crc_powers(uint32_t poly, uint32_t *R, int n);

#ifdef __cplusplus
}
#endif
#endif
