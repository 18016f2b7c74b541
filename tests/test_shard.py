"""Multi-GPU host logic on CPU: packets partitioned by rank (no data-path collective), results gathered in packet
order -- world_size 2 over gloo.  The decoder inside each rank is the oracle here (no GPU in this container);
on the GPU box the same partition feeds one w80211_ctx per rank (bench.py)."""
import os
import sys

import numpy as np
import pytest

import w80211rx

gp, g, shard = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.shard
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_is_a_balanced_cover():
    rng = np.random.default_rng(5)
    work = rng.integers(100, 33000, 1000)
    for world in (1, 2, 3, 8):
        parts = shard.partition(work, world)
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(1000))
        loads = [work[p].sum() for p in parts]
        assert max(loads) - min(loads) <= work.max()
    assert [len(p) for p in shard.partition([], 4)] == [0, 0, 0, 0]


def _batches():
    return [gp.gen_batch(g.F_L, 5, 1, 300, 3, snr_db=30, seed=41), gp.gen_batch(g.F_HT, 4, 1, 900, 3, snr_db=30, seed=42),
            gp.gen_batch(g.F_VHT, 6, 1, 150, 3, snr_db=30, seed=43), gp.gen_batch(g.F_HT, 7, 1, 2000, 2, snr_db=30, seed=44)]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import refgr
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pk = []
    for b in _batches():
        pk += gp.to_packets(b)
    parts = shard.partition([len(p["iq"][0]) for p in pk], world)
    mine = parts[rank]
    msgs = []
    for i in mine:  # one message list per packet (a packet may publish 0..n MPDUs)
        llr = refgr.run_demod(refgr.make_sig_stream([pk[i]]))
        msgs.append(refgr.run_decode(llr))
    out = shard.gather_in_packet_order(mine, msgs, dist)
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_over_gloo_give_the_single_process_result(refgr):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pk = []
    for b in _batches():
        pk += gp.to_packets(b)
    want = [refgr.run_decode(refgr.run_demod(refgr.make_sig_stream([p]))) for p in pk]
    assert got == want
    assert sum(len(m) for m in got) == len(pk)
