"""Host/collective logic of batch-mode F2F fine-tuning (frame2frame_b200/batch_f2f.py) on CPU:
sequence sharding, the gradient exchange and the ordered BatchNorm running-stat update, with a
world-size-2 (and 3) gloo process group.  No kernel is called here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from frame2frame_b200 import batch_f2f as bf

P = 4099          # stand-in for the 556 032-float gradient vector


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _seq_grad(s):
    g = torch.Generator().manual_seed(100 + s)
    return torch.randn(P, generator=g) * (10.0 ** (s % 3))


def _seq_stats(s):
    g = torch.Generator().manual_seed(500 + s)
    return torch.rand(bf.NBN, 2, bf.NCH, generator=g)


def _worker(rank, world, port, nseq, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        local = bf.shard_sequences(nseq, rank, world)
        rows = torch.stack([_seq_grad(s) for s in local]) if local else torch.zeros(0, P)
        det = bf.exchange_gradients(rows, nseq, deterministic=True)
        fast = bf.exchange_gradients(rows.sum(0) if local else torch.zeros(P), nseq, deterministic=False)
        stats = torch.stack([_seq_stats(s) for s in local]) if local else torch.zeros(0, bf.NBN, 2, bf.NCH)
        allstats = bf.gather_rows(stats, nseq)
        running = torch.ones(bf.NBN, 2, bf.NCH)
        nbt = torch.zeros(bf.NBN, dtype=torch.int64)
        bf.apply_bn_ema(running, nbt, allstats)
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), det=det.numpy(), fast=fast.numpy(), running=running.numpy(),
                 nbt=nbt.numpy())
    finally:
        dist.destroy_process_group()


def _single(nseq):
    rows = torch.stack([_seq_grad(s) for s in range(nseq)])
    det = bf.exchange_gradients(rows, nseq, deterministic=True)
    running = torch.ones(bf.NBN, 2, bf.NCH)
    nbt = torch.zeros(bf.NBN, dtype=torch.int64)
    bf.apply_bn_ema(running, nbt, torch.stack([_seq_stats(s) for s in range(nseq)]))
    return det, running, nbt


def test_shard_sequences_partition():
    for nseq in (1, 5, 8, 13):
        for world in (1, 2, 3, 4, 8):
            blocks = [bf.shard_sequences(nseq, r, world) for r in range(world)]
            assert sum(blocks, []) == list(range(nseq))                      # contiguous, ordered, complete
            assert max(map(len, blocks)) - min(map(len, blocks)) <= 1
    with pytest.raises(ValueError):
        bf.shard_sequences(4, 4, 4)


def test_single_process_exchange_is_ordered_sum():
    det, running, nbt = _single(8)
    ref = _seq_grad(0).clone()
    for s in range(1, 8):
        ref += _seq_grad(s)
    assert torch.equal(det, ref)
    r = torch.ones(bf.NBN, 2, bf.NCH)
    for s in range(8):                                                       # nn.BatchNorm2d's update, SURVEY A.4
        r = (1 - torch.tensor(0.1)) * r + torch.tensor(0.1) * _seq_stats(s)
    assert torch.allclose(running, r, rtol=0, atol=1e-6)
    assert int(nbt[0]) == 8


def test_batch_stats_from_sums():
    n = 1000.0
    z = torch.randn(bf.NBN, bf.NCH, int(n), dtype=torch.float64) * 3 + 1
    sums = torch.stack([z.sum(-1), (z * z).sum(-1)], 1)
    st = bf.batch_stats_from_sums(sums, n)
    assert torch.allclose(st[:, 0], z.mean(-1).float(), atol=1e-6)
    assert torch.allclose(st[:, 1], z.var(-1, unbiased=True).float(), rtol=1e-5)


@pytest.mark.parametrize("world,nseq", [(2, 8), (3, 8), (2, 1)])
def test_multi_rank_equals_single_process(tmp_path, world, nseq):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, nseq, str(tmp_path)), nprocs=world, join=True)
    det, running, nbt = _single(nseq)
    plain = torch.stack([_seq_grad(s) for s in range(nseq)]).sum(0)
    for r in range(world):
        d = np.load(os.path.join(str(tmp_path), "r%d.npz" % r))
        assert np.array_equal(d["det"], det.numpy())             # bit-identical to the single-process loop
        assert np.array_equal(d["running"], running.numpy())     # BN running stats too
        assert np.array_equal(d["nbt"], nbt.numpy())
        assert np.allclose(d["fast"], plain.numpy(), rtol=1e-5, atol=1e-4)   # all-reduce: same sum up to association
    a = np.load(os.path.join(str(tmp_path), "r0.npz"))["fast"]
    for r in range(1, world):
        assert np.array_equal(np.load(os.path.join(str(tmp_path), "r%d.npz" % r))["fast"], a)   # identical on every rank
