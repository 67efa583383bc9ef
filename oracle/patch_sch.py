#!/usr/bin/env python
"""Level-2 integration patch of INTEGRATION.md, applied to a COPY of the reference's lib/src/phy/phch/sch.c (test
infrastructure: nothing of the reference is committed here, the edit is made by anchor, not by diff).

    python oracle/patch_sch.py /root/reference/lib/src/phy/phch/sch.c <out.c>      (oracle/Makefile compiles the copy and removes it)

Two edits, both behind -DSRSLTE_FEC_CUDA:
 1. decode_tb (sch.c:437-486) hands the whole transport block to libsrslte_fec_cuda (srslte_sch_decode_tb_cuda[_8bit]) as its
    first statement; decode_tb_cb (sch.c:307-422) becomes dead code. srslte_ulsch_decode / srslte_ulsch_uci_decode
    (sch.c:883-985) reach the GPU through it.
 2. srslte_dlsch_decode and srslte_dlsch_decode2 (sch.c:488-517) are compiled out: the library exports them itself.
"""
import re
import sys

src, dst = sys.argv[1], sys.argv[2]
lines = open(src).read().split("\n")


def find(pattern, start=0):
    for i in range(start, len(lines)):
        if re.match(pattern, lines[i]):
            return i
    raise SystemExit(f"patch_sch: anchor {pattern!r} not found in {src}")


def body_open(i):
    while "{" not in lines[i]:
        i += 1
    return i


def func_end(i):
    """line index of the closing brace (column 0) of the function that starts at or after line i"""
    return find(r"^}", body_open(i))


HOOK = """#ifdef SRSLTE_FEC_CUDA
  {
    extern int srslte_sch_decode_tb_cuda(srslte_softbuffer_rx_t *, uint32_t, uint32_t, uint32_t, uint32_t, int16_t *, uint8_t *, uint32_t, uint32_t *);
    extern int srslte_sch_decode_tb_cuda_8bit(srslte_softbuffer_rx_t *, uint32_t, uint32_t, uint32_t, uint32_t, int8_t *, uint8_t *, uint32_t, uint32_t *);
    if (!q || !data || !softbuffer || !e_bits || !cb_segm) return SRSLTE_ERROR_INVALID_INPUTS;
    if (q->llr_is_8bit)
      return srslte_sch_decode_tb_cuda_8bit(softbuffer, cb_segm->tbs, Qm, rv, nof_e_bits, (int8_t *)e_bits, data, q->max_iterations, &q->nof_iterations);
    return srslte_sch_decode_tb_cuda(softbuffer, cb_segm->tbs, Qm, rv, nof_e_bits, e_bits, data, q->max_iterations, &q->nof_iterations);
  }
#endif"""

# 2 first (it lies behind 1, so the line numbers of 1 stay valid): compile the two DL entry points out
a = find(r"^int srslte_dlsch_decode\(")
b = func_end(find(r"^int srslte_dlsch_decode2\("))
lines.insert(b + 1, "#endif /* SRSLTE_FEC_CUDA: srslte_dlsch_decode[2] come from libsrslte_fec_cuda */")
lines.insert(a, "#ifndef SRSLTE_FEC_CUDA")
# 1
o = body_open(find(r"^static int decode_tb\("))
lines.insert(o + 1, HOOK)
open(dst, "w").write("\n".join(lines))
print(f"patch_sch: wrote {dst}")
