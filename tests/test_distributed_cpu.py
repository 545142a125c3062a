"""Host logic of the N>1 path on CPU: utterance sharding and the one exchange step (tensor
all-gather back into item order) with world_size 2 over gloo."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, D, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from speakerverification_b200.model import ModelHandling as MH
        # round-robin shard (fixed-length utterances)
        mine = MH.shard(n_items, rank, world)
        local = torch.stack([torch.full((D,), float(i)) for i in mine]) if mine else torch.zeros((0, D))
        full = MH.gather_sharded(local, n_items, rank, world)
        ok_rr = bool(torch.equal(full, torch.arange(n_items, dtype=torch.float32)[:, None].expand(n_items, D)))
        # length-balanced shard (ragged utterances) with an explicit assignment
        lengths = np.random.default_rng(5).integers(32000, 320000, size=n_items)
        parts = MH.shard_balanced(lengths, world)
        mine = parts[rank]
        local = torch.stack([torch.full((D,), float(i)) for i in mine]) if len(mine) else torch.zeros((0, D))
        full2 = MH.gather_assigned(local, parts, rank, world)
        ok_bal = bool(torch.equal(full2, torch.arange(n_items, dtype=torch.float32)[:, None].expand(n_items, D)))
        loads = [int(lengths[p].sum()) for p in parts]
        q.put((rank, ok_rr, ok_bal, loads))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [7, 64])
def test_shard_and_gather_world2(n_items):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, 5, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_rr, ok_bal, loads in res:
        assert ok_rr, "round-robin gather mismatch on rank %d" % rank
        assert ok_bal, "balanced gather mismatch on rank %d" % rank
        assert max(loads) - min(loads) <= 320000        # greedy LPT: imbalance below one utterance


def test_shard_balanced_properties():
    from speakerverification_b200.model import ModelHandling as MH
    lengths = np.random.default_rng(2).uniform(2.0, 20.0, size=1000) * 16000
    for world in (1, 2, 4, 8):
        parts = MH.shard_balanced(lengths, world)
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(1000))          # a partition: every utterance exactly once
        loads = np.array([lengths[p].sum() for p in parts])
        assert loads.max() / loads.mean() < 1.01
    assert [list(p) for p in MH.shard_balanced([5, 1, 1], 4)][3] == []  # more ranks than items: empty shard
