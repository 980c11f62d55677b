"""CPU <-> GPU paging of Gaussian parameters and Adam moments between training iterations: the host side of csrc/page.cu, with the
semantics of train_post.py:440-479 (write-back of the rows that leave the render set, `tensor[keep_mask]`, `cat(kept, storage[load].cuda())`).
One mask scan per iteration, one device pass and one transfer each way per tensor, pinned staging. No CPU fallback for the device side."""
import ctypes as C

import torch

from . import check, current_stream, lib


class Pager:
    """swap() for one iteration; staging buffers are kept and grown between calls."""

    def __init__(self, device):
        self.dev = torch.device(device)
        self._ws = None
        self._stage = {}

    def _buf(self, key, numel, pinned):
        t = self._stage.get(key)
        if t is None or t.numel() < numel:
            t = torch.empty(max(numel, 1), dtype=torch.float32, pin_memory=True) if pinned else torch.empty(max(numel, 1), device=self.dev)
            self._stage[key] = t
        return t[:numel]

    def plan(self, keep_mask):
        """-> (dst i32[n] on the device, kept, evicted)"""
        m = keep_mask.to(device=self.dev).contiguous()
        if m.dtype not in (torch.bool, torch.uint8):
            m = m != 0
        n = m.numel()
        need = int(lib.lodgs_page_workspace_bytes(n))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
        dst = torch.empty(n, dtype=torch.int32, device=self.dev)
        counts = (C.c_int32 * 2)()
        check(lib.lodgs_page_plan(m.data_ptr(), n, dst.data_ptr(), self._ws.data_ptr(), self._ws.numel(), counts, current_stream()))
        return dst, counts[0], counts[1]

    def swap(self, tensors, storages, keep_mask, write_back_indices, load_indices):
        """tensors: resident CUDA tensors [n, ...]; storages: the matching host tensors [N_total, ...] (float32, contiguous; pinned or not).
        Rows with keep_mask False are written back to storage[write_back_indices] (write_back_indices = render_indices[~keep_mask], int64,
        host), rows storage[load_indices] are appended. -> list of new resident tensors [kept + len(load_indices), ...]."""
        dst, n_keep, n_evict = self.plan(keep_mask)
        wb = write_back_indices.to("cpu", torch.int64).contiguous()
        ld = load_indices.to("cpu", torch.int64).contiguous()
        if wb.numel() != n_evict:
            raise RuntimeError("swap: %d write-back indices for %d evicted rows" % (wb.numel(), n_evict))
        n, n_load = dst.numel(), ld.numel()
        out = []
        stream = current_stream()
        for k, (t, st) in enumerate(zip(tensors, storages)):
            t = t.detach().contiguous()
            width = t[0].numel() if n else st[0].numel()
            if st.dtype != torch.float32 or not st.is_contiguous() or st.device.type != "cpu":
                raise RuntimeError("swap: storages must be contiguous float32 host tensors")
            h_in, d_in = self._buf(("hin", k), n_load * width, True), self._buf(("din", k), n_load * width, False)
            if n_load:
                check(lib.lodgs_host_gather_rows(st.data_ptr(), st.size(0), width * 4, ld.data_ptr(), n_load, h_in.data_ptr()))
                d_in.copy_(h_in, non_blocking=True)                                    # one cudaMemcpyAsync
            d_ev = self._buf(("dev", k), n_evict * width, False)
            new = torch.empty((n_keep + n_load,) + tuple(t.shape[1:] if n else st.shape[1:]), device=self.dev)
            check(lib.lodgs_page_apply(t.data_ptr() if n else None, n, width, dst.data_ptr() if n else None, new.data_ptr() if new.numel() else None,
                                       d_ev.data_ptr() if n_evict else None, d_in.data_ptr() if n_load else None, n_load, n_keep, stream))
            if n_evict:
                h_ev = self._buf(("hev", k), n_evict * width, True)
                h_ev.copy_(d_ev, non_blocking=True)                                    # one cudaMemcpyAsync
                torch.cuda.current_stream().synchronize()
                check(lib.lodgs_host_scatter_rows(st.data_ptr(), st.size(0), width * 4, wb.data_ptr(), n_evict, h_ev.data_ptr()))
            out.append(new)
        return out
