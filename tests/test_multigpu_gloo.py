"""world_size-2 check (gloo, CPU) of the multi-GPU host logic: transport blocks are partitioned by cell so
that every retransmission of a HARQ process meets its soft buffer on the same rank, nothing is exchanged on the data
path, and the whole-job figures are SUM(bits) / MAX(time).  The per-rank decode is done by the CPU oracle here (the
checker standing in for the GPU that is absent); the GPU run of the same logic is bench.py --gpus N."""
import ctypes as C
import importlib
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from _libs import P, oracle

shard = importlib.import_module("empower-srslte_b200.shard")
synth = importlib.import_module("empower-srslte_b200.synth")

TBS, QM, NBITS = 6200, 2, 9000          # two code blocks of K = 3136, rate ~0.69: rv 0 at sigma 0.9 fails, the rv 2 retransmission rescues it
CELLS, PIDS = 5, 2


def _workload():
    rng = np.random.default_rng(11)
    keys = [(cell, pid) for cell in range(CELLS) for pid in range(PIDS)]
    payload = rng.integers(0, 256, (len(keys), TBS // 8), dtype=np.uint8)
    tx = []
    for rv, sigma in ((0, 0.9), (2, 0.9)):
        e = synth.encode_tbs(payload, TBS, QM, rv, NBITS)
        tx.append((rv, synth.awgn_llr(e, sigma, np.random.default_rng(100 + rv))))
    return keys, payload, tx


def _decode(orc, idx, tx):
    """decode the transport blocks `idx` through both transmissions with one soft buffer each; returns per-TB
    (ret after tx0, ret after tx1, payload bytes)"""
    out = {}
    for i in idx:
        ot = C.c_void_p(orc.orc_tb_new())
        orc.orc_tb_reset(ot, TBS)
        rets = []
        for rv, llr in tx:
            data = np.zeros(TBS // 8 + 3 + 1024, np.uint8)
            crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
            rets.append(orc.orc_tb_decode(ot, TBS, QM, rv, NBITS, P(np.ascontiguousarray(llr[i])), 8, P(data), P(crc), P(its)))
            if rets[-1] == 0:
                break
        orc.orc_tb_free(ot)
        out[i] = (tuple(rets), bytes(data[: TBS // 8]))
    return out


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    keys, payload, tx = _workload()
    mine = shard.partition(keys, rank, world)
    res = _decode(oracle(), mine, tx)
    bits = sum(TBS for r in res.values() if r[0][-1] == 0)
    ms, sums = shard.reduce_job(dist, "cpu", 10.0 + rank, bits=bits, tbs=len(mine))
    gathered = [None] * world
    dist.all_gather_object(gathered, res)
    dist.barrier()
    if rank == 0:
        q.put((ms, sums, gathered, mine))
    dist.destroy_process_group()


def test_world_size_2_sharding():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ms, sums, gathered, mine0 = q.get(timeout=300)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    keys, payload, tx = _workload()
    # partition: disjoint, complete, cell-affine
    i0, i1 = set(gathered[0]), set(gathered[1])
    assert not (i0 & i1) and (i0 | i1) == set(range(len(keys)))
    assert all(shard.owner(keys[i][0], 2) == 0 for i in i0) and all(shard.owner(keys[i][0], 2) == 1 for i in i1)
    # same results as one process decoding everything (no exchange needed), HARQ combining exercised
    whole = _decode(oracle(), range(len(keys)), tx)
    merged = {**gathered[0], **gathered[1]}
    assert merged == whole
    assert any(len(r[0]) == 2 and r[0][0] == -1 and r[0][1] == 0 for r in whole.values()), "no retransmission was exercised"
    for i, r in whole.items():
        if r[0][-1] == 0:
            assert r[1] == bytes(payload[i])
    # whole-job reduction: SUM of bits and TBs, MAX of time
    assert ms == 11.0
    assert sums["tbs"] == len(keys) and sums["bits"] == sum(TBS for r in whole.values() if r[0][-1] == 0)


def test_single_process_reduce_is_identity():
    assert shard.reduce_job(None, "cpu", 3.5, a=2, b=7) == (3.5, {"a": 2.0, "b": 7.0})
    assert shard.partition([(0, 0), (1, 0), (2, 1), (3, 1)], 1, 2) == [1, 3]
