"""Row b of SURVEY section 8: the reference's OWN test programs, compiled from /root/reference by `make -C oracle linktest`
and linked against libsrslte_fec_cuda.so in place of the decoder objects (INTEGRATION.md level 1: turbodecoder*.o, the
receive half of rm_turbo.o, the rx half of softbuffer.o; level 2: sch.c with decode_tb handing over and srslte_dlsch_decode[2]
taken from the library). They are built in the container (the reference is not on the GPU box) and travel with the snapshot
under oracle/_ref/link/. Command lines are the reference's ctest lines (lib/src/phy/fec/test/CMakeLists.txt:35-47,
lib/src/phy/phch/test/CMakeLists.txt:85-134,270-291)."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

LNK = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "link")


def run(prog, *args, timeout=280):
    exe = os.path.join(LNK, prog)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (needs /root/reference: make -C oracle linktest)")
    r = subprocess.run([exe, *args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    return r.returncode, r.stdout


def last_ber(out):
    """the BER of the last progress line (turbodecoder_test.c:287 overwrites it with \\r)"""
    m = re.findall(r"BER: ([0-9.e+-]+)", out)
    assert m, out[-400:]
    return m[-1]


@pytest.mark.parametrize("args", [("-n", "100", "-s", "1", "-l", "504", "-e", "1.0", "-t"),
                                  ("-n", "100", "-s", "1", "-l", "504", "-e", "2.0", "-t"),
                                  ("-n", "100", "-s", "1", "-l", "6144", "-e", "1.5", "-t"),
                                  ("-n", "1", "-s", "1", "-k", "-e", "0.5"),
                                  # manual decoder types (srslte_tdec_init_manual, turbodecoder.c:153-190): 2 SSE, 3 SSE-window, 4 AVX-window
                                  ("-n", "20", "-s", "1", "-l", "1024", "-e", "1.0", "-d", "2"),
                                  ("-n", "20", "-s", "1", "-l", "6144", "-e", "1.0", "-d", "3"),
                                  ("-n", "20", "-s", "1", "-l", "672", "-e", "1.5", "-d", "4")],
                         ids=lambda a: "".join(a))
def test_turbodecoder_test_prints_the_reference_ber(args):
    """same seed, same random data, same AWGN (srand/rand): the drop-in build must print the BER the unmodified reference prints"""
    rc, out = run("turbodecoder_test", *args)
    assert rc == 0, out[-800:]
    rc_ref, out_ref = run("turbodecoder_test_ref", *args)
    assert rc_ref == 0
    assert last_ber(out) == last_ber(out_ref), (last_ber(out), last_ber(out_ref))
    assert ("Errors" in out) == ("Errors" in out_ref)


@pytest.mark.parametrize("e", ["1920", "8192"])
def test_rm_turbo_test(e):
    """rm_turbo_test.c:140-193: for all 188 K x 4 rv the LUT receive path (here: the GPU) equals the float reference"""
    rc, out = run("rm_turbo_test", "-e", e)
    assert rc == 0, out[-800:]
    assert out.count("OK RX") == 188 * 4


@pytest.mark.parametrize("args", [("-m", "28", "-n", "100"), ("-n", "100"), ("-m", "20", "-n", "100", "-r", "2"), ("-m", "10", "-n", "50", "-r", "1"),
                                  ("-x", "multiplex", "-a", "2", "-t", "0", "-m", "28", "-M", "28", "-n", "100"),
                                  ("-x", "multiplex", "-a", "2", "-t", "0", "-m", "28", "-M", "28", "-n", "100", "-j"),
                                  ("-x", "diversity", "-a", "2", "-n", "50"), ("-m", "28", "-n", "100", "-b")],
                         ids=lambda a: "".join(a))
def test_pdsch_test(args):
    """pdsch_test.c:489-538: encode -> srslte_pdsch_decode (-> srslte_dlsch_decode2 from the library) -> byte compare;
    -j runs the PDSCH co-worker thread (two decoders at once, pdsch.c:837-864), -b the 8-bit LLR mode"""
    rc, out = run("pdsch_test", *args)
    assert rc == 0, out[-800:]


@pytest.mark.parametrize("args", [("-n", "100", "-L", "96", "-m", "20"), ("-n", "50", "-L", "25", "-m", "10", "-p", "uci_ack", "1"),
                                  ("-n", "25", "-L", "10", "-m", "16", "-p", "uci_cqi", "wideband")],
                         ids=lambda a: "".join(a))
def test_pusch_test(args):
    """pusch_test: srslte_pusch_encode -> srslte_pusch_decode -> srslte_ulsch_uci_decode -> decode_tb -> the library"""
    rc, out = run("pusch_test", *args)
    assert rc == 0, out[-800:]
