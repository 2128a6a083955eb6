"""Host logic of the data-parallel gradient exchange (tacotron2_b200.parallel.GradSync) on CPU: world_size 2, gloo."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tacotron2_b200.parallel import GradSync

    class Dummy:  # stands in for Tacotron2: GradSync only needs the hook slot
        _grad_bucket_hook = None

    m = Dummy()
    gs = GradSync(m)
    assert m._grad_bucket_hook is not None and gs.world == world
    torch.manual_seed(rank)
    buckets = []
    for sizes in ((3, 5), (7,), (2, 2, 2)):  # three buckets in backward order
        params = [torch.nn.Parameter(torch.zeros(n)) for n in sizes]
        flat = torch.arange(sum(sizes), dtype=torch.float32) * (rank + 1)
        views, off = [], 0
        for p in params:
            views.append(flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        m._grad_bucket_hook(flat, params, views)
        buckets.append((params, sum(sizes), 1.0))
    # the first parameter group arrives a second time (a second decoder chunk of a batch above 64 utterances): contributions add
    params, total, _ = buckets[0]
    flat2 = torch.arange(total, dtype=torch.float32) * (rank + 1) * 2.0
    views2, off = [], 0
    for p in params:
        views2.append(flat2[off:off + p.numel()].view_as(p))
        off += p.numel()
    m._grad_bucket_hook(flat2, params, views2)
    buckets[0] = (params, total, 3.0)
    n = gs.finish()
    ok = n == 8 + 7 + 6 + 8
    for params, total, mult in buckets:
        want = torch.arange(total, dtype=torch.float32) * (mult * sum(range(1, world + 1)) / world)  # mean over ranks
        got = torch.cat([p.grad.reshape(-1) for p in params])
        ok = ok and torch.allclose(got, want)
    ok = ok and not gs.pending
    q.put((rank, bool(ok) if ok else f"n={n} grads={[p.grad.tolist() for ps, _ in buckets for p in ps]}"))
    dist.destroy_process_group()


def test_gradsync_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]
