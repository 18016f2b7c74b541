"""Sharding of a batch of packets over the GPUs of one box.

Packets are independent units (own channel estimate, own trellis, own FCS), so a batch is partitioned BY PACKET,
one process per GPU (rank), greedy-balanced on the work per packet (trellis length ~ L-SIG duration), with no
collective on the data path.  The only communication is the gather of the decoded messages to rank 0 in packet
order, through torch.distributed (NCCL on GPUs, gloo in the CPU tests).
"""
import heapq

import numpy as np


def partition(work, world):
    """Longest-processing-time-first assignment of packets to `world` ranks.
    work: per-packet cost (e.g. number of samples or trellis steps).  -> list of `world` sorted index arrays."""
    work = np.asarray(work, np.float64)
    order = np.argsort(-work, kind="stable")
    heap = [(0.0, r) for r in range(world)]
    heapq.heapify(heap)
    parts = [[] for _ in range(world)]
    for i in order:
        load, r = heapq.heappop(heap)
        parts[r].append(int(i))
        heapq.heappush(heap, (load + float(work[i]), r))
    return [np.array(sorted(p), np.int64) for p in parts]


def gather_in_packet_order(local_indices, local_items, dist, dst=0):
    """Every rank passes the global indices of its packets and one item (any picklable object) per packet;
    rank `dst` gets the items of all ranks ordered by global packet index, the others get None."""
    world = dist.get_world_size()
    payload = (list(map(int, local_indices)), list(local_items))
    gathered = [None] * world if dist.get_rank() == dst else None
    dist.gather_object(payload, gathered, dst=dst)
    if dist.get_rank() != dst:
        return None
    pairs = []
    for idx, items in gathered:
        pairs += list(zip(idx, items))
    pairs.sort(key=lambda t: t[0])
    return [it for _, it in pairs]
