"""CPU, world_size 2 over gloo: the host-side multi-GPU logic (camera sharding, gradient bucket all-reduce)."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lodgs import parallel


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 1000, 1001):
        for world in (1, 2, 3, 8):
            cover = []
            for r in range(world):
                lo, hi = parallel.shard_range(n, r, world)
                cover += list(range(lo, hi))
                assert hi - lo in (n // world, n // world + 1)
            assert cover == list(range(n))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(100 + rank)
    grads = [torch.randn(50, 3, generator=g), None, torch.randn(50, 16, 3, generator=g), torch.randn(50, 1, generator=g)]
    want = None
    # rank 0 takes the flat-bucket path for every tensor, rank 1 would too; the second call forces the in-place path for the
    # 9600-byte tensor (threshold 1000 bytes) and keeps the bucket for the others — both must give the same sums
    again = [None if x is None else x.clone() for x in grads]
    grads0 = [x.clone() for x in grads if x is not None]
    nbytes = parallel.allreduce_gradients(grads)
    nbytes2 = parallel.allreduce_gradients(again, inplace_min_bytes=1000)
    assert nbytes2 == nbytes and all(a is None or torch.equal(a, b) for a, b in zip(again, grads))
    # per-tensor handles (the optimizer waits for each parameter's gradient on its own)
    third = [x.clone() for x in grads0]
    handles, nbytes3 = parallel.allreduce_gradients_async(third)
    for h in handles:
        h.wait()
    assert nbytes3 == nbytes and all(torch.equal(a, b) for a, b in zip(third, [g for g in grads if g is not None]))
    lo, hi = parallel.shard_range(11, rank, world)
    t = parallel.gather_timings(torch.arange(lo, hi, dtype=torch.float64))
    # numpy arrays travel by value; torch tensors would be passed as shared-memory handles that die with this process
    q.put((rank, [None if x is None else x.numpy().copy() for x in grads], nbytes, [x.tolist() for x in t]))
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_allreduce_two_ranks():
    world, port = 2, 29513
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    exp = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        exp.append([torch.randn(50, 3, generator=g), torch.randn(50, 16, 3, generator=g), torch.randn(50, 1, generator=g)])
    sums = [exp[0][i] + exp[1][i] for i in range(3)]
    for rank, grads, nbytes, timings in res:
        got = [torch.from_numpy(x) for x in grads if x is not None]
        assert nbytes == 4 * (150 + 2400 + 50)
        for a, b in zip(got, sums):
            assert torch.allclose(a, b, atol=0, rtol=0)
        assert timings == [[0.0, 1.0, 2.0, 3.0, 4.0, 5.0], [6.0, 7.0, 8.0, 9.0, 10.0]]


def _basis_deg1(d):
    """SH basis values of degree <= 1 for unit directions d [P,3] (computeColorFromSH, HR/cuda_rasterizer/forward.cu:20-40): [P,4]"""
    c0, c1 = 0.28209479177387814, 0.4886025119029199
    return torch.stack((torch.full_like(d[:, 0], c0), -c1 * d[:, 1], c1 * d[:, 2], -c1 * d[:, 0]), dim=1)


def _worker_factored(rank, world, port, q):
    from lodgs import dp
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = 40
    g = torch.Generator().manual_seed(7)
    means = torch.randn(P, 3, generator=g)                       # replicated scene
    g = torch.Generator().manual_seed(200 + rank)                # this rank's view
    campos, flat, drgb = torch.randn(3, generator=g), torch.randn(11 * P, generator=g), torch.randn(P, 3, generator=g)
    d = means - campos
    basis = _basis_deg1(d / d.norm(dim=1, keepdim=True))         # [P,4]
    dsh = basis[:, :, None] * drgb[:, None, :]                   # what the materialised backward would write: [P,4,3]
    flat2, drgb_all = flat.clone(), torch.empty(world, P, 3)
    campos_all = torch.empty(world, 3)
    dist.all_gather_into_tensor(campos_all, campos.reshape(1, 3))
    w1, w2, nb_f = dp.exchange_gradients(flat, drgb, drgb_all)   # factored: 11 + 3 floats per Gaussian
    w1.wait(); w2.wait()
    w3, w4, nb_m = dp.exchange_gradients(flat2, dsh=dsh)         # materialised: 11 + 12 floats per Gaussian at degree 1
    w3.wait(); w4.wait()
    # every replica rebuilds the summed SH gradient from the gathered colour gradients, views in rank order
    rebuilt = torch.zeros(P, 4, 3)
    for v in range(world):
        dv = means - campos_all[v]
        rebuilt += _basis_deg1(dv / dv.norm(dim=1, keepdim=True))[:, :, None] * drgb_all[v][:, None, :]
    q.put((rank, nb_f, nb_m, torch.equal(flat, flat2), float((rebuilt - dsh).abs().max()), rebuilt.numpy().copy(), drgb_all.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_factored_sh_gradient_exchange_two_ranks():
    """lodgs.dp.exchange_gradients: all-reduce of the flat geometry block + all-gather of dL_drgb carries the same information as the
    all-reduce of the materialised dL_dsh (rebuilt per replica, identical on every replica) in 56 instead of 92 B per Gaussian at degree 1
    (236 B at degree 3)."""
    world, port = 2, 29517
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_factored, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, nb_f, nb_m, flat_same, err, rebuilt, gathered in res:
        assert nb_f == 4 * 40 * (11 + 3) and nb_m == 4 * 40 * (11 + 12)
        assert flat_same and err < 1e-6
    assert (res[0][5] == res[1][5]).all() and (res[0][6] == res[1][6]).all()      # replicas bit-identical
