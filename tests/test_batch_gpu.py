"""Batch-mode F2F fine-tuning (BASELINE.json configs[3]) on the GPU:
  * one process, S sequences, against the FP32 CPU oracle loop "for each sequence: forward, loss,
    backward (gradients accumulate); then one Adam step" -- the definition of configs[3]'s
    reference, built from the same stock torch modules as the reference;
  * two processes sharing the work (sequences sharded, gradient exchange, ordered BatchNorm EMA)
    give BIT-IDENTICAL weights, running statistics and outputs to the one-process run in
    deterministic mode, and agree to float association in all-reduce mode.
The gloo two-rank runs work on CUDA tensors of one device so they need a single GPU; the NCCL two-rank
run (one process per GPU, the production configuration) needs two and is skipped on a one-GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

H, W, NSEQ, ITERS = 72, 104, 4, 3


def _inputs():
    from frame2frame_b200 import synth
    currs = [synth.noisy_frame(H, W, 1, seq=s) / np.float32(255.0) for s in range(NSEQ)]
    prevs = [synth.noisy_frame(H, W, 0, seq=s) / np.float32(255.0) for s in range(NSEQ)]
    flow = synth.backward_flow(H, W, 1)
    return currs, prevs, flow


def _net(dev):
    from frame2frame_b200 import synth
    from frame2frame_b200.dncnn import DnCNN
    net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[dev.index or 0]).to(dev)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()})
    return net


def _run(dev, group_world, rank, deterministic):
    from frame2frame_b200.batch_f2f import BatchF2F
    currs, prevs, flow = _inputs()
    eng = BatchF2F(_net(dev), NSEQ, deterministic=deterministic)
    t = lambda a: torch.from_numpy(a)[None, None]
    losses = []
    eng.finetune([t(currs[s]) for s in eng.local], [t(prevs[s]) for s in eng.local],
                 [torch.from_numpy(flow)[None] for _ in eng.local], ITERS, losses)
    den = eng.denoise(t(currs[0]))
    torch.cuda.synchronize()
    return (eng.net.flat_params.detach().cpu().numpy(), eng.net._running.cpu().numpy(), den.cpu().numpy(),
            np.array(losses))


def _worker(rank, world, port, deterministic, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p, r, d, l = _run(torch.device("cuda:0"), world, rank, deterministic)
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), p=p, r=r, d=d, l=l)
    finally:
        dist.destroy_process_group()


def _worker_nccl(rank, world, port, deterministic, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        p, r, d, l = _run(dev, world, rank, deterministic)
        np.savez(os.path.join(out_dir, "n%d.npz" % rank), p=p, r=r, d=d, l=l)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_batch_matches_oracle_accumulate_loop(cuda):
    from oracle import dncnn_oracle as do
    from oracle import f2f_oracle as fo
    from frame2frame_b200 import synth
    currs, prevs, flow = _inputs()
    ref = do.load_numpy_state(do.OracleDnCNN(residual=False), synth.pretrained_state_dict())
    opt = do.make_adam(ref)
    crit = torch.nn.L1Loss(reduction="sum")
    pre = []
    for s in range(NSEQ):
        warped, _, mask = fo.warp_and_mask(prevs[s], flow, f64=True)
        pre.append(tuple(torch.from_numpy(a)[None, None] for a in (currs[s], warped, mask)))
    ref.train()
    l_ref = []
    for _ in range(ITERS):
        opt.zero_grad()
        tot = 0.0
        for tc, tw, tm in pre:
            loss = crit(tm * (tc - ref(tc)), tm * tw)
            loss.backward()                       # gradients accumulate over the sequences
            tot += float(loss.detach())
        opt.step()
        l_ref.append(tot)
    p, r, d, l = _run(cuda, 1, 0, False)
    assert np.allclose(l, l_ref, rtol=2e-2), (l, l_ref)
    rm = torch.stack([torch.stack([m.running_mean, m.running_var]) for m in ref.dncnn if isinstance(m, torch.nn.BatchNorm2d)])
    assert np.allclose(r, rm.numpy(), rtol=5e-2, atol=5e-3)
    pf = do.flat_params(ref).numpy()
    assert np.abs(p - pf).max() <= 3 * 5e-5 * ITERS      # every Adam step moves a weight by at most ~lr


@pytest.mark.parametrize("deterministic", [True, False])
def test_two_ranks_equal_one(cuda, tmp_path, deterministic):
    p1, r1, d1, l1 = _run(cuda, 1, 0, deterministic)
    mp.spawn(_worker, args=(2, _free_port(), deterministic, str(tmp_path)), nprocs=2, join=True)
    for rank in range(2):
        z = np.load(os.path.join(str(tmp_path), "r%d.npz" % rank))
        if deterministic:
            assert np.array_equal(z["r"], r1)                      # ordered EMA: identical for any world size
            assert np.array_equal(z["p"], p1) and np.array_equal(z["d"], d1) and np.array_equal(z["l"], l1)
        else:
            assert np.allclose(z["l"], l1, rtol=1e-6)
            assert np.abs(z["p"] - p1).max() <= 1e-4
            assert np.allclose(z["r"], r1, rtol=1e-4, atol=1e-6)
    a, b = (np.load(os.path.join(str(tmp_path), "r%d.npz" % k)) for k in range(2))
    assert np.array_equal(a["p"], b["p"])                          # ranks hold identical weights afterwards


@pytest.mark.parametrize("deterministic", [True, False])
def test_two_ranks_nccl_equal_one(cuda, tmp_path, deterministic):
    """The production exchange: one process per GPU, gradients and BatchNorm statistics over NCCL
    (batch_f2f.exchange_gradients / gather_rows take the CUDA tensors as they are).  Deterministic mode must be
    bit-identical to the one-rank run; all-reduce mode agrees to float association."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (NCCL, one process per GPU)")
    p1, r1, d1, l1 = _run(cuda, 1, 0, deterministic)
    mp.spawn(_worker_nccl, args=(2, _free_port(), deterministic, str(tmp_path)), nprocs=2, join=True)
    a, b = (np.load(os.path.join(str(tmp_path), "n%d.npz" % k)) for k in range(2))
    assert np.array_equal(a["p"], b["p"]) and np.array_equal(a["r"], b["r"])
    for z in (a, b):
        if deterministic:
            assert np.array_equal(z["p"], p1) and np.array_equal(z["r"], r1) and np.array_equal(z["l"], l1)
        else:
            assert np.allclose(z["l"], l1, rtol=1e-6)
            assert np.abs(z["p"] - p1).max() <= 1e-4
            assert np.allclose(z["r"], r1, rtol=1e-4, atol=1e-6)
    assert np.array_equal(a["d"], d1) if deterministic else np.allclose(a["d"], d1, atol=1e-3)
