"""Multi-GPU plumbing. One process per GPU, torch.distributed for the only exchange the path has.

* Eval / viewer: frames are independent -> `shard_range` gives each rank a contiguous camera range; the hierarchy and the
  Gaussians are replicated; no collective on the data path (only `gather_timings` for reports).
* Multi-view training of one chunk: views are split across ranks, parameters replicated; after backward the dense
  parameter gradients are summed across ranks (`allreduce_gradients`: large tensors in place, small ones through one flat bucket), then every rank applies
  the same fused Adam step on identical sums (visibility = dL_dopacity != 0 after the reduce, train_single.py:171-172).
"""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous [lo, hi) of rank's share of n_items (first n_items % world ranks get one extra)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


INPLACE_MIN_BYTES = 1 << 20


def allreduce_gradients(grads, group=None, inplace_min_bytes=INPLACE_MIN_BYTES, coalesce=False):
    """Sum a list of gradient tensors across ranks (NCCL all-reduce over NVLink on a B200 box, gloo on CPU in the tests).
    Contiguous tensors of at least `inplace_min_bytes` are reduced where they lie, back to back on the collective's stream (no
    staging copy: packing 59 floats per Gaussian into a bucket and unpacking it again costs two extra passes over 236 MB at 1 M
    Gaussians); the small ones share one flat bucket so that each does not pay a collective's latency. Tensors are updated in
    place; returns the number of bytes reduced. `coalesce` issues the in-place calls inside one NCCL group (one launch for all
    tensors) through torch's coalescing manager."""
    grads = [g for g in grads if g is not None]
    if not grads or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0
    if coalesce:
        from torch.distributed.distributed_c10d import _coalescing_manager
        for g in grads:
            if not g.is_contiguous():
                raise ValueError("coalesced all-reduce needs contiguous gradient tensors")
        with _coalescing_manager(group=group, device=grads[0].device, async_ops=True) as cm:
            for g in grads:
                dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
        cm.wait()
        return sum(g.numel() * g.element_size() for g in grads)
    big = [g for g in grads if g.is_contiguous() and g.numel() * g.element_size() >= inplace_min_bytes]
    small = [g for g in grads if not (g.is_contiguous() and g.numel() * g.element_size() >= inplace_min_bytes)]
    works = [dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group, async_op=True) for g in big]
    nbytes = sum(g.numel() * g.element_size() for g in big)
    if small:
        flat = torch.cat([g.reshape(-1) for g in small])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        off = 0
        for g in small:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n
        nbytes += flat.numel() * flat.element_size()
    for w in works:
        w.wait()
    return nbytes


def allreduce_gradients_async(grads, group=None):
    """Start one in-place sum all-reduce per (contiguous) gradient tensor, in list order, and return the work handles without waiting:
    `handle.wait()` makes the CURRENT STREAM wait for that tensor only, so the optimizer can update the tensors that have arrived
    while the later ones are still on the wire (`SparseGaussianAdam.step_overlapped`). NCCL runs the calls in issue order: put
    the tensor the step needs first (dL_dopacity, the visibility mask) first and the largest next. -> (handles, bytes)"""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [None] * len(grads), 0
    for g in grads:
        if g is None or not g.is_contiguous():
            raise ValueError("allreduce_gradients_async needs contiguous gradient tensors")
    works = [dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group, async_op=True) for g in grads]
    return works, sum(g.numel() * g.element_size() for g in grads)


def gather_timings(values, group=None):
    """All ranks contribute a 1-D float64 tensor of per-frame timings; rank 0 gets the list of all."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [values]
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, values.tolist(), group=group)
    return [torch.tensor(o, dtype=torch.float64) for o in out]
