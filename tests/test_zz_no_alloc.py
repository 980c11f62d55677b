"""GPU: lodgs_page_plan_static, the allocation-free variant of lodgs_page_plan (scan scratch in a static __device__ array of the library;
include/lodgs.h). Called through the C ABI, twice in a row with different masks (the scratch keeps the first call's contents), and compared
bit for bit with the existing entry point; sizes beyond the static maximum are refused. Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

import lodgs
from lodgs import lib

pytestmark = pytest.mark.gpu


def _plan_existing(mask):
    n = mask.numel()
    ws = torch.empty(int(lib.lodgs_page_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
    dst = torch.full((max(n, 1),), 12345, dtype=torch.int32, device="cuda")
    counts = (C.c_int32 * 2)()
    rc = lib.lodgs_page_plan(mask.data_ptr(), n, dst.data_ptr(), ws.data_ptr(), ws.numel(), counts, None)
    assert rc == 0, lodgs.last_error()
    torch.cuda.synchronize()
    return dst[:n].clone(), (counts[0], counts[1])


def _plan_static(mask, dst, counts_dev):
    rc = lib.lodgs_page_plan_static(mask.data_ptr(), mask.numel(), dst.data_ptr(), counts_dev.data_ptr(), None)
    assert rc == 0, lodgs.last_error()
    torch.cuda.synchronize()


def test_page_plan_static_matches_page_plan_twice_in_a_row():
    g = torch.Generator().manual_seed(7)
    # a long mask first, then a shorter, differently filled one: stale block sums of call 1 lie behind the words call 2 uses
    masks = [(torch.rand(300_001, generator=g) < 0.7), (torch.rand(70_000, generator=g) < 0.2), torch.zeros(0, dtype=torch.bool)]
    masks = [m.to(torch.uint8).cuda() for m in masks]
    want = [_plan_existing(m) for m in masks]
    # everything the new entry point touches is allocated BEFORE it is called; the calls themselves must not allocate
    dsts = [torch.full((max(m.numel(), 1),), -777, dtype=torch.int32, device="cuda") for m in masks]
    counts_dev = [torch.full((2,), -1, dtype=torch.int32, device="cuda") for _ in masks]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for m, d, c in zip(masks, dsts, counts_dev):
        _plan_static(m, d, c)
    assert torch.cuda.memory_allocated() == before
    for m, d, c, (wd, wc) in zip(masks, dsts, counts_dev, want):
        n = m.numel()
        assert torch.equal(d[:n], wd)                      # integer work: bit-exact
        assert tuple(c.tolist()) == wc == (int(m.sum()), n - int(m.sum()))
    # and once more over the first mask, after the scratch has seen the other sizes
    d0 = torch.full_like(dsts[0], -777)
    _plan_static(masks[0], d0, counts_dev[0])
    assert torch.equal(d0, want[0][0]) and tuple(counts_dev[0].tolist()) == want[0][1]


def test_page_plan_static_refuses_sizes_beyond_its_scratch():
    limit = int(lib.lodgs_page_plan_static_max_rows())
    assert limit == 16 * 1024 * 1024
    one = torch.zeros(8, dtype=torch.uint8, device="cuda")
    counts_dev = torch.zeros(2, dtype=torch.int32, device="cuda")
    rc = lib.lodgs_page_plan_static(one.data_ptr(), limit + 1, one.data_ptr(), counts_dev.data_ptr(), None)   # refused before any launch
    assert rc == -2 and "maximum" in lodgs.last_error()
    rc = lib.lodgs_page_plan_static(one.data_ptr(), 8, one.data_ptr(), None, None)
    assert rc != 0
