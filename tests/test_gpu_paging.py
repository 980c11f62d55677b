"""GPU: parameter / moment paging (csrc/page.cu, lodgs.paging.Pager) against the torch indexing it replaces, written as in
train_post.py:440-479 — bit-identical resident tensors and host storage, over several iterations, incl. empty and full masks."""
import pytest
import torch

from lodgs.paging import Pager

pytestmark = pytest.mark.gpu
SHAPES = [(3,), (1,), (1, 3), (15, 3), (3,), (4,)]     # xyz, opacity, f_dc, f_rest, scaling, rotation


def _reference_swap(tensors, storages, keep_mask, write_back_indices, load_indices):
    out = []
    for t, st in zip(tensors, storages):
        st[write_back_indices, :] = t[~keep_mask, :].detach().to("cpu")                 # train_post.py:440-445
        kept = t[keep_mask]                                                           # :448-453
        out.append(torch.cat((kept.detach(), st[load_indices].cuda())).contiguous())  # :457-462
    return out


@pytest.mark.parametrize("n_total,n_res,seed", [(5000, 1200, 0), (200000, 60000, 1), (100, 0, 2)])
def test_pager_matches_torch_indexing(n_total, n_res, seed):
    g = torch.Generator().manual_seed(seed)
    st_a = [torch.randn((n_total,) + s, generator=g) for s in SHAPES]
    st_b = [t.clone() for t in st_a]
    perm = torch.randperm(n_total, generator=g)
    render = perm[:n_res].clone()
    res_a = [st[render].cuda() + 1.0 for st in st_a]          # resident copies have been trained: they differ from storage
    res_b = [t.clone() for t in res_a]
    pager = Pager("cuda")
    for it in range(4):
        n = render.numel()
        keep = torch.rand(n, generator=g) < (0.0 if it == 2 else 0.7)
        if it == 3:
            keep[:] = True
        outside = perm[~torch.isin(perm, render)]
        load = outside[torch.randperm(outside.numel(), generator=g)[: min(outside.numel(), 37 + n_total // 20)]]
        wb = render[~keep]
        res_a = _reference_swap(res_a, st_a, keep.cuda(), wb, load)
        res_b = pager.swap(res_b, st_b, keep.cuda(), wb, load)
        torch.cuda.synchronize()
        for a, b in zip(res_a, res_b):
            assert a.shape == b.shape and torch.equal(a, b)
        for a, b in zip(st_a, st_b):
            assert torch.equal(a, b)
        render = torch.cat((render[keep], load))
        res_a = [t * 1.5 for t in res_a]
        res_b = [t * 1.5 for t in res_b]
