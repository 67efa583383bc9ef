/*
 * fec_tx.cu — the transmit side of the transport-block chain: transport-block CRC, code-block
 * segmentation + CRC, turbo encoding and rate matching, i.e. the data part of srslte_dlsch_encode / srslte_ulsch_encode:
 *
 *   encode_tb_off                  sch.c:188-296        per-CB loop, E split (the ENCODER's: i <= C - gamma - 1)
 *   srslte_tcod_encode_lut         turbocoder.c:196-330 CRC-24A over the TB, CRC-24B per CB, two constituent encoders + tails
 *   srslte_rm_turbo_tx_lut         rm_turbo.c:332-376   sub-block interleavers, bit collection, circular read from k0(rv)
 *
 * One CTA per code block. The recursive encoder (3-bit state, GF(2)-linear) is parallelised over 32-bit words: every
 * thread runs its word from the zero state (zero-state response + end state), a short scan chains the end states
 * (state after a word = T32[state before] ^ zero-state end state), and the response of the true start state — one of 8
 * precomputed 32-bit "free responses" — is XOR-ed in. Decoder 2's input word is gathered through the QPP polynomial.
 * Rate matching is a gather: output bit m is bit src[(off(rv) + m) mod N] of the three d-streams held in shared memory,
 * src being the NULL-free circular buffer of 36.212 5.1.4.1.2 (the same enumeration the receive side's rank table inverts).
 * Output: packed bits, MSB first, code blocks back to back at bit granularity (first / last partial byte by atomicOr into
 * the zeroed buffer). HBM traffic: K/8 bytes in, E/8 out per code block — negligible next to the decode side.
 */
#include "fec_kernels.cuh"

#define TX_THREADS 256

/* state = r1 | r2 << 1 | r3 << 2; one step of g0 = 1 + D^2 + D^3 (feedback), g1 = 1 + D + D^3 (turbocoder.c:82-165) */
__device__ __forceinline__ uint32_t tx_step(uint32_t &st, uint32_t u)
{
  const uint32_t r1 = st & 1u, r2 = (st >> 1) & 1u, r3 = st >> 2;
  const uint32_t fb = u ^ r2 ^ r3;
  st = fb | (r1 << 1) | (r2 << 2);
  return fb ^ r1 ^ r3;
}

__device__ __forceinline__ uint32_t tx_bit(const uint32_t *w, uint32_t i) { return (w[i >> 5] >> (31u - (i & 31u))) & 1u; }

/* CRC-24A of every transport block -> 3 bytes, MSB first (crc.c:144-158 through the x^(24+m) mod g table) */
__global__ void __launch_bounds__(TX_THREADS) fec_tx_tbcrc_kernel(const FecTxTbDesc *descs, const uint32_t *r24a)
{
  const FecTxTbDesc d = descs[blockIdx.x];
  __shared__ uint32_t s_red[TX_THREADS / 32];
  const int tid = threadIdx.x;
  uint32_t acc = 0;
  /* the encode kernel ORs the bytes two code blocks share into the output: start from zero */
  for (uint32_t w = tid; w < d.e_words; w += TX_THREADS) d.e_out[w] = 0;
  for (uint32_t b = tid; b < d.tbs / 8; b += TX_THREADS) {
    uint32_t v = d.data[b];
    while (v) {
      const int hb = 31 - __clz(v);
      acc ^= r24a[d.tbs - 1 - (8 * b + (7 - hb))];
      v &= ~(1u << hb);
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) acc ^= __shfl_xor_sync(0xFFFFFFFFu, acc, s);
  if ((tid & 31) == 0) s_red[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    uint32_t c = 0;
    for (int i = 0; i < TX_THREADS / 32; i++) c ^= s_red[i];
    d.crc3[0] = (uint8_t)(c >> 16); d.crc3[1] = (uint8_t)(c >> 8); d.crc3[2] = (uint8_t)c;
  }
}

__global__ void __launch_bounds__(TX_THREADS) fec_tx_encode_kernel(const FecTxCbDesc *descs, const uint32_t *r24b)
{
  const FecTxCbDesc d = descs[blockIdx.x];
  __shared__ uint32_t sall[3 * FEC_TX_W];          /* d0 (systematic) | d1 (parity) | d2 (parity'), K + 4 bits each, MSB first */
  __shared__ uint32_t s_red[TX_THREADS / 32], s_free[8];
  __shared__ uint8_t s_fin[2][FEC_TX_W], s_ini[2][FEC_TX_W], s_t32[8], s_tail[12];
  uint32_t *su = sall, *s1 = sall + FEC_TX_W, *s2 = sall + 2 * FEC_TX_W;
  const int tid = threadIdx.x;
  const uint32_t K = d.K, nW = (K + 31) / 32;

  /* ---- the code block's bits: payload bytes, transport-block CRC behind the last block's payload ---- */
  for (uint32_t w = tid; w < FEC_TX_W; w += TX_THREADS) {
    uint32_t v = 0;
#pragma unroll
    for (uint32_t b = 0; b < 4; b++) {
      const uint32_t i = 4 * w + b;
      uint32_t x = 0;
      if (i < d.nsrc) x = d.src[i];
      else if (i < d.nsrc + d.ntb) x = d.tbcrc[i - d.nsrc];
      v = (v << 8) | x;
    }
    su[w] = v; s1[w] = 0; s2[w] = 0;
  }
  if (tid < 8) {                                     /* zero-input response of each start state over 32 steps */
    uint32_t st = tid, z = 0;
    for (int i = 0; i < 32; i++) z |= tx_step(st, 0u) << (31 - i);
    s_free[tid] = z; s_t32[tid] = (uint8_t)st;
  }
  __syncthreads();
  /* ---- code-block CRC-24B over the first K - 24 bits (turbocoder.c:255-265) ---- */
  if (d.cbcrc) {
    const uint32_t n = K - 24;
    uint32_t acc = 0;
    for (uint32_t w = tid; w < nW; w += TX_THREADS) {
      uint32_t v = su[w];
      while (v) {
        const int hb = 31 - __clz(v);
        const uint32_t pos = 32 * w + (31 - hb);
        if (pos < n) acc ^= r24b[n - 1 - pos];
        v &= ~(1u << hb);
      }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc ^= __shfl_xor_sync(0xFFFFFFFFu, acc, s);
    if ((tid & 31) == 0) s_red[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
      uint32_t c = 0;
      for (int i = 0; i < TX_THREADS / 32; i++) c ^= s_red[i];
      for (uint32_t i = 0; i < 3; i++) {
        const uint32_t by = n / 8 + i, x = (c >> (16 - 8 * i)) & 0xFFu;
        su[by >> 2] |= x << (8 * (3 - (by & 3)));
      }
    }
    __syncthreads();
  }
  /* ---- the two constituent encoders, word-parallel: zero-state response + end state of every word ---- */
  for (uint32_t job = tid; job < 2 * nW; job += TX_THREADS) {
    const uint32_t enc = job >= nW, w = enc ? job - nW : job;
    const uint32_t nb = min(32u, K - 32 * w);
    uint32_t in = su[w];
    if (enc) {
      /* input of the second encoder: c'_i = c_pi(i), pi(i) = (f1 i + f2 i^2) mod K, stepped by its first difference */
      const unsigned long long i0 = 32ull * w;
      uint32_t pi = (uint32_t)((d.f1 * i0 + (unsigned long long)d.f2 * i0 % K * i0) % K);
      uint32_t g = (uint32_t)((d.f1 + (unsigned long long)d.f2 * (2 * i0 + 1)) % K);
      const uint32_t g2 = (2 * d.f2) % K;
      in = 0;
      for (uint32_t i = 0; i < nb; i++) {
        in |= tx_bit(su, pi) << (31 - i);
        pi += g; if (pi >= K) pi -= K;
        g += g2; if (g >= K) g -= K;
      }
    }
    uint32_t st = 0, z = 0;
    for (uint32_t i = 0; i < nb; i++) z |= tx_step(st, (in >> (31 - i)) & 1u) << (31 - i);
    (enc ? s2 : s1)[w] = z;
    s_fin[enc][w] = (uint8_t)st;
  }
  __syncthreads();
  /* ---- chain the end states; the last word may be short, its end state is stepped out ---- */
  if (tid == 0 || tid == 32) {
    const uint32_t enc = tid >> 5;
    uint32_t st = 0;
    for (uint32_t w = 0; w < nW; w++) {
      s_ini[enc][w] = (uint8_t)st;
      const uint32_t nb = min(32u, K - 32 * w);
      if (nb == 32) st = s_t32[st];
      else for (uint32_t i = 0; i < nb; i++) tx_step(st, 0u);
      st ^= s_fin[enc][w];
    }
    /* trellis termination: three steps with the switch in the lower position (input = feedback), 36.212 5.1.3.2.2 */
    for (int k = 0; k < 3; k++) {
      const uint32_t r1 = st & 1u, r2 = (st >> 1) & 1u, r3 = st >> 2;
      s_tail[6 * enc + 2 * k] = (uint8_t)(r2 ^ r3);
      s_tail[6 * enc + 2 * k + 1] = (uint8_t)(r1 ^ r3);
      st = (r1 << 1) | (r2 << 2);
    }
  }
  __syncthreads();
  for (uint32_t job = tid; job < 2 * nW; job += TX_THREADS) {
    const uint32_t enc = job >= nW, w = enc ? job - nW : job;
    const uint32_t nb = min(32u, K - 32 * w), st = s_ini[enc][w];
    if (st) (enc ? s2 : s1)[w] ^= s_free[st] & (nb == 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> nb));
  }
  __syncthreads();
  if (tid == 0) {
    /* d_s[K + j] = tail[3 j + s] with tail = xK zK xK+1 zK+1 xK+2 zK+2 x'K z'K x'K+1 z'K+1 x'K+2 z'K+2 (36.212 5.1.3.2.2) */
    for (uint32_t j = 0; j < 4; j++)
      for (uint32_t s = 0; s < 3; s++)
        if (s_tail[3 * j + s]) sall[s * FEC_TX_W + ((K + j) >> 5)] |= 1u << (31u - ((K + j) & 31u));
  }
  __syncthreads();
  /* ---- rate matching: circular read of the NULL-free buffer from k0(rv), packed into the transport block's output ---- */
  const uint32_t first = d.wp >> 3, last = (d.wp + d.E - 1) >> 3, end = d.wp + d.E;
  for (uint32_t b = first + tid; b <= last; b += TX_THREADS) {
    uint32_t byte = 0;
    uint32_t o = 8 * b, p = 0;
    bool have = false;
#pragma unroll
    for (uint32_t jb = 0; jb < 8; jb++, o++) {
      if (o >= d.wp && o < end) {
        if (!have) { p = (d.off + (o - d.wp)) % d.N; have = true; }
        const uint32_t v = d.txsrc[p];
        byte |= tx_bit(sall, v) << (7 - jb);
        if (++p == d.N) p = 0;
      }
    }
    if ((b == first && (d.wp & 7)) || (b == last && (end & 7)))
      atomicOr(reinterpret_cast<unsigned int *>(d.e_out + (b & ~3u)), byte << (8 * (b & 3u)));
    else d.e_out[b] = (uint8_t)byte;
  }
}

cudaError_t fec_launch_tx(const FecTxTbDesc *tbd, uint32_t ntb, const FecTxCbDesc *cbd, uint32_t ncb,
                          const uint32_t *r24a, const uint32_t *r24b, cudaStream_t st)
{
  if (ntb) fec_tx_tbcrc_kernel<<<ntb, TX_THREADS, 0, st>>>(tbd, r24a);
  if (ncb) fec_tx_encode_kernel<<<ncb, TX_THREADS, 0, st>>>(cbd, r24b);
  return cudaGetLastError();
}
