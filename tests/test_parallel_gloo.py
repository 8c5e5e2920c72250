"""Host-side multi-rank logic on CPU: world_size 2, gloo.  The compute injected into the sharding
helpers is the CPU oracle (the product kernels need a GPU; the plumbing is what is under test)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import dmnerf_oracle as orc
        from dm_nerf_b200.parallel import shard_range, gather_rows, render_sharded, allreduce_grads
        torch.set_num_threads(1)
        # 1. sharded render of 101 rays (ragged split 51/50) == single-process render
        n, k = 101, 8
        g = torch.Generator().manual_seed(0)
        raw = torch.randn(n, 16, k + 5, generator=g)
        z = torch.sort(torch.rand(n, 16, generator=g) * 5 + 1, -1)[0]
        rd = torch.randn(n, 3, generator=g)

        def render_fn(a, cnt):
            rgb, _, depth, ins = orc.render_train(raw[a:a + cnt], z[a:a + cnt], rd[a:a + cnt])
            return {"rgb": rgb, "depth": depth, "labels": ins.argmax(-1).to(torch.uint8)}
        frame = render_sharded(render_fn, n)
        ok = True
        if rank == 0:
            full = render_fn(0, n)
            ok = all(torch.equal(frame[kk], full[kk]) for kk in full)
        else:
            ok = frame is None
        a, b = shard_range(n, rank, world)
        ok = ok and (b - a) == (51 if rank == 0 else 50)
        # 1b. the same frame gathered point to point into rank 0's preallocated buffers (no padding, no cat); rank 0
        # renders straight into its rows through `out`
        def render_into(a, cnt, out=None):
            res = render_fn(a, cnt)
            if out is None:
                return res
            for kk in out:
                out[kk].copy_(res[kk])
            return out
        bufs = None
        if rank == 0:
            bufs = {"rgb": torch.full((n, 3), -1.0), "depth": torch.full((n,), -1.0),
                    "labels": torch.full((n,), 255, dtype=torch.uint8)}
        frame2 = render_sharded(render_into, n, frame=bufs)
        if rank == 0:
            ok = ok and frame2 is bufs and all(torch.equal(frame2[kk], full[kk]) for kk in full)
        else:
            ok = ok and frame2 is None
        # 2. gradient all-reduce == gradient of the mean loss over the concatenated batch
        torch.manual_seed(1)
        lin = torch.nn.Linear(6, 3)
        x = torch.randn(8, 6, generator=torch.Generator().manual_seed(5))
        lo, hi = shard_range(8, rank, world)
        (lin(x[lo:hi]) ** 2).mean().backward()
        allreduce_grads(list(lin.parameters()))
        ref = torch.nn.Linear(6, 3)
        ref.load_state_dict(lin.state_dict())
        (ref(x) ** 2).mean().backward()
        ok = ok and all(torch.allclose(p.grad, r.grad, atol=1e-6) for p, r in zip(lin.parameters(), ref.parameters()))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_sharded_render_and_grad_allreduce_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    assert res == {0: True, 1: True}


def test_shard_range_covers_everything():
    sys.path.insert(0, ROOT)
    from dm_nerf_b200.parallel import shard_range
    for total in (0, 1, 7, 307200, 1254528):
        for world in (1, 2, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
