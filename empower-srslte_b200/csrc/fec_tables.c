/*
 * fec_tables.c — host-side table generation for libsrslte_fec_cuda (see fec_tables.h).
 * Replaces, for the decode path, the static tables the reference builds in
 * srslte_rm_turbo_gentables (rm_turbo.c:267-305), srslte_tc_interl_LTE_gen_interl (tc_interl_lte.c:78-120)
 * and gen_crc_table (crc.c:38-47) — but in the device layout, and only for the K actually used.
 */
#include "fec_tables.h"
#include "lte_qpp_table.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

int fec_k_index(uint32_t K)
{
  /* the 188 sizes are 40..512 step 8, 528..1024 step 16, 1056..2048 step 32, 2112..6144 step 64 (cbsegm.c:35-52) */
  if (K > FEC_MAX_K) return -1;
  if (K <= 40) return 0;
  if (K <= 512) return (int)((K - 40 + 7) / 8);
  if (K <= 1024) return 59 + (int)((K - 512 + 15) / 16);
  if (K <= 2048) return 91 + (int)((K - 1024 + 31) / 32);
  return 123 + (int)((K - 2048 + 63) / 64);
}

int fec_k_size(uint32_t idx) { return idx < FEC_NOF_K ? (int)lte_qpp_table[idx][0] : -1; }

int fec_k_subblocks(uint32_t K)
{
  if ((K % 16) == 0 && K > 800) return 16;
  if ((K % 8) == 0 && K > 400) return 8;
  return 0;
}

int fec_segm(uint32_t tbs, uint32_t *o)
{
  memset(o, 0, 9 * sizeof(uint32_t));
  if (tbs == 0) return 0;
  uint32_t B = tbs + 24, C, Bp;
  if (B <= FEC_MAX_K) { C = 1; Bp = B; }
  else { C = (uint32_t)ceilf((float)B / (float)(FEC_MAX_K - 24)); Bp = B + 24 * C; }
  int i1 = fec_k_index((Bp - 1) / C + 1);
  if (i1 < 0) return -1;
  uint32_t K1 = lte_qpp_table[i1][0], K2 = 0, K2i = 0, C1 = 1, C2 = 0;
  if (C > 1) {
    K2i = (uint32_t)i1 - 1; K2 = lte_qpp_table[K2i][0];
    C2 = (C * K1 - Bp) / (K1 - K2); C1 = C - C2;
  }
  o[0] = C1 * K1 + C2 * K2 - Bp; o[1] = C; o[2] = K1; o[3] = K2; o[4] = (uint32_t)i1; o[5] = K2i;
  o[6] = C1; o[7] = C2; o[8] = tbs;
  return 0;
}

int fec_k_subblocks8(uint32_t K)
{   /* srslte_tdec_autoimp_get_subblocks_8bit, AVX2 build (turbodecoder.c:392-406) */
  if (!(K % 32) && K > 2048) return 32;
  if (!(K % 16) && K > 800) return 16;
  if (!(K % 8) && K > 400) return 8;
  return 0;
}

int fec_geom(uint32_t K, fec_geom_t *g) { return fec_geom_p(K, fec_k_subblocks(K), g); }

int fec_geom_p(uint32_t K, int P, fec_geom_t *g)
{
  int idx = fec_k_index(K);
  if (idx < 0 || lte_qpp_table[idx][0] != K) return -1;
  g->K = (int)K; g->idx = idx; g->P = P;
  g->NL = g->P ? g->P : 1;
  g->L = g->K / g->NL;
  g->Lp = (g->L + 7) & ~7;
  if (g->P) { g->LpE = g->L; while ((g->LpE & 3) != 2) g->LpE++; }
  else g->LpE = g->K;
  g->SS = g->NL * g->Lp;
  g->Lw = g->L; while ((g->Lw & 7) != 4) g->Lw++;
  g->f1 = lte_qpp_table[idx][1]; g->f2 = lte_qpp_table[idx][2];
  return 0;
}

void fec_gen_step_table(const fec_geom_t *g, uint32_t *tab)
{
  const int K = g->K, L = g->L;
  memset(tab, 0, sizeof(uint32_t) * 2 * (size_t)g->SS);
  for (int par = 0; par < 2; par++)
    for (int j = 0; j < K; j++) {
      uint64_t jj = (uint64_t)j;
      int nat = par ? (int)(((uint64_t)g->f1 * jj + (uint64_t)g->f2 * jj * jj) % (uint64_t)K) : j;
      int ea = (nat / L) * g->LpE + nat % L;
      tab[(size_t)par * g->SS + (j / L) * g->Lp + j % L] = ((uint32_t)nat << 16) | (uint32_t)ea;
    }
}

static int qpp_nat(const fec_geom_t *g, int par, int j)
{
  uint64_t jj = (uint64_t)j;
  return par ? (int)(((uint64_t)g->f1 * jj + (uint64_t)g->f2 * jj * jj) % (uint64_t)g->K) : j;
}

void fec_gen_pair_addr_table(const fec_geom_t *g, uint32_t *tab)
{
  const int L = g->L;
  memset(tab, 0, sizeof(uint32_t) * (size_t)(g->P / 2) * g->Lp);
  for (int j = 0; j < g->P / 2; j++)
    for (int k = 0; k < L; k++) {
      uint32_t ea[2];
      for (int h = 0; h < 2; h++) {
        const int nat = qpp_nat(g, 1, (2 * j + h) * L + k), d = nat / L, kk = nat % L;
        ea[h] = (uint32_t)(2 * ((d >> 1) * g->Lw + kk) + (d & 1));
      }
      tab[(size_t)j * g->Lp + k] = (ea[1] << 16) | ea[0];
    }
}

void fec_gen_pair_crc_table(const fec_geom_t *g, uint32_t poly, uint32_t *tab)
{
  const int L = g->L, K = g->K;
  const size_t rows = (size_t)(g->P / 2) * g->Lp;
  uint32_t *R = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)K);
  fec_gen_crc_powers(poly, R, K);
  memset(tab, 0, sizeof(uint32_t) * 2 * 2 * rows);
  for (int par = 0; par < 2; par++)
    for (int j = 0; j < g->P / 2; j++)
      for (int k = 0; k < L; k++)
        for (int h = 0; h < 2; h++)
          tab[2 * ((size_t)par * rows + (size_t)(((k >> 1) * (g->P / 2) + j) * 2 + (k & 1))) + h] = R[K - 1 - qpp_nat(g, par, (2 * j + h) * L + k)];
  free(R);
}

static const uint8_t rm_colperm[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                       1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};

int fec_soft_offset_dev(const fec_geom_t *g, int t)
{
  if (t >= 3 * g->K) return 3 * g->SS + (t - 3 * g->K);
  int s = t % 3, x = t / 3, d = x / g->L, k = x % g->L;
  if (g->P == 0) return s * g->SS + k;
  return s * g->SS + 2 * ((((k >> 2) * (g->P / 2) + (d >> 1)) << 2) + (k & 3)) + (d & 1);
}

int fec_soft_offset_ref(const fec_geom_t *g, int t, int sb_layout)
{
  if (!sb_layout || g->P == 0) return t;
  if (t >= 3 * g->K) return 3 * (g->K + 32) + (t - 3 * g->K);
  int s = t % 3, x = t / 3;
  return s * (g->K + 32) + (x % g->L) * g->P + x / g->L;
}

/* 36.212 5.1.4.1: three sub-block interleavers (32 columns, R rows, ND leading <NULL>s), bit collection
 * w = [v0 | v1,v2 interlaced], circular read from k0(rv) skipping <NULL>s. The rank of a coded bit is its
 * position among the non-NULL entries of w counted from w[0]. */
void fec_gen_rm_rank(const fec_geom_t *g, uint16_t *rank, uint32_t off[4])
{
  const int D = g->K + 4, R = (D - 1) / 32 + 1, KP = 32 * R, ND = KP - D, NCB = 3 * KP;
  int k0[4];
  for (int rv = 0; rv < 4; rv++) k0[rv] = R * (2 * (int)ceilf((float)NCB / (float)(8 * R)) * rv + 2);
  /* entries are BYTE offsets into the rate de-matcher's shared-memory tile (2 * position); device offsets no coded bit maps
   * to (lane padding) point at tile entry N = 3K + 12, which the kernel keeps at zero */
  for (int i = 0; i < 3 * g->SS + 16; i++) rank[i] = (uint16_t)(2 * (3 * g->K + 12));
  int cnt = 0;
  for (int w = 0; w < NCB; w++) {
    for (int rv = 0; rv < 4; rv++) if (w == k0[rv]) off[rv] = (uint32_t)cnt;
    int y, stream, m;
    if (w < KP) { stream = 0; y = (w % R) * 32 + rm_colperm[w / R]; }
    else if (((w - KP) & 1) == 0) { m = (w - KP) / 2; stream = 1; y = (m % R) * 32 + rm_colperm[m / R]; }
    else { m = (w - KP) / 2; stream = 2; y = (rm_colperm[m / R] + 32 * (m % R) + 1) % KP; }
    if (y < ND) continue;
    rank[fec_soft_offset_dev(g, 3 * (y - ND) + stream)] = (uint16_t)(2 * cnt++);
  }
}

/* The same table re-based to the start of one redundancy version: entry = 2 x (index i of the received LLR e[i] that lands on
 * this soft position, i in [0, 3K+12) for the first pass through the circular buffer), pad entries stay at 2 x (3K+12).
 * With it the rate de-matcher's tile is simply e[] in the order received (one bulk copy) followed by zeros. */
void fec_rm_rank_rebase(const fec_geom_t *g, const uint16_t *rank, uint32_t off, uint16_t *out)
{
  const uint32_t N = 3 * (uint32_t)g->K + 12;
  for (int i = 0; i < 3 * g->SS + 16; i++) {
    const uint32_t c = rank[i] / 2u;
    out[i] = (uint16_t)(c >= N ? 2 * N : 2 * ((c + N - off) % N));
  }
}

void fec_gen_rm_txsrc(const fec_geom_t *g, uint32_t stream_bits, uint16_t *src)
{
  const int D = g->K + 4, R = (D - 1) / 32 + 1, KP = 32 * R, ND = KP - D, NCB = 3 * KP;
  int cnt = 0;
  for (int w = 0; w < NCB; w++) {
    int y, stream, m;
    if (w < KP) { stream = 0; y = (w % R) * 32 + rm_colperm[w / R]; }
    else if (((w - KP) & 1) == 0) { m = (w - KP) / 2; stream = 1; y = (m % R) * 32 + rm_colperm[m / R]; }
    else { m = (w - KP) / 2; stream = 2; y = (rm_colperm[m / R] + 32 * (m % R) + 1) % KP; }
    if (y < ND) continue;
    src[cnt++] = (uint16_t)((uint32_t)stream * stream_bits + (uint32_t)(y - ND));
  }
}

void fec_gen_crc_powers(uint32_t poly, uint32_t *Rt, int n)
{
  uint32_t r = poly & 0xFFFFFFu;         /* x^24 mod g */
  for (int m = 0; m < n; m++) {
    Rt[m] = r;
    r <<= 1;
    if (r & 0x1000000u) r ^= poly;
    r &= 0xFFFFFFu;
  }
}
