"""Hardware check of the data-parallel gradient exchange: two ranks on two GPUs over NCCL.  Each rank runs
one training step on its own shard through GradSync; the averaged gradients every rank ends up with must equal the mean of the
two ranks' single-rank gradients (computed in the same processes with the hook detached and exchanged with all_gather).
Skips below two GPUs (the driver's round-end box has one; run with `gpurun --gpus 2`).  Each shard has 70 utterances, so the
decoder bucket arrives twice per backward."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        from tacotron2_b200 import synth
        from tacotron2_b200.model import Tacotron2
        from tacotron2_b200.parallel import GradSync

        args, B, L, T = synth.HIFI, 70, 23, 12   # the same number of decoder chunks (= collectives) on every rank
        m = Tacotron2(**args.kwargs()).to(dev).train()
        m.load_state_dict(synth.synth_state_dict(args, seed=0))
        b = {k: v.to(dev) for k, v in synth.synth_batch(args, B, L, T, seed=100 + rank).items()}
        masks = synth.synth_masks(args, B, L, T, seed=200 + rank, train=True)

        def step():
            m.zero_grad(set_to_none=True)
            mels, post, gates, _ = m(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
                                     speaker_id=b["speaker_id"], dropout_masks=masks)
            m.loss(mels, post, gates, b["mel_spectrogram"], b["gate"]).backward()

        step()  # single-rank gradients (no hook)
        local = torch.cat([p.grad.reshape(-1) for p in m.parameters()]).clone()
        both = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(both, local)
        want = sum(both) / world
        rs0 = {k: v.clone() for k, v in m.state_dict().items() if "running_" in k}
        m.load_state_dict(synth.synth_state_dict(args, seed=0))   # undo the running-statistics update of the first step
        sync = GradSync(m)
        step()
        n = sync.finish()
        got = torch.cat([p.grad.reshape(-1) for p in m.parameters()])
        torch.cuda.synchronize()
        rel = float((got - want).norm() / want.norm())
        mx = float((got - want).abs().max() / want.abs().max())
        same_stats = all(torch.equal(rs0[k], v) for k, v in m.state_dict().items() if "running_" in k)
        q.put((rank, rel, mx, n == sum(p.numel() for p in m.parameters()) + sum(p.numel() for p in m._group_params("decoder").values()),
               same_stats))
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"{type(e).__name__}: {e}", None, False, False))


def test_nccl_gradsync_equals_mean_of_single_rank_gradients():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(60)
    for rank, rel, mx, counted, same_stats in res:
        assert isinstance(rel, float), (rank, rel)
        assert rel <= 1e-6 and mx <= 1e-6, (rank, rel, mx)   # the average of two fp32 vectors: one rounding
        assert counted, 'bucket element count'
        assert same_stats, "BatchNorm statistics must stay per-GPU (no SyncBN, SURVEY 8e)"
