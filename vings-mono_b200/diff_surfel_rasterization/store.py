"""Surfel add / prune churn on the GPU.

* ``knn_mean_dist2``  drop-in for ``distCUDA2`` (scripts/gaussian/gaussian_utils.py:12-16) without the scipy KD-tree
  and its GPU<->CPU round trip;
* ``SurfelStore``     every per-surfel array of the mapper (five raw parameters, their ten Adam moments, score
  bookkeeping) in capacity-doubling device buffers held in pairs: ``prune`` is ONE stable compaction of all of
  them from one buffer set into the other (prune_tensors_from_optimizer + update_records("prune"),
  scripts/gaussian/gaussian_base.py:96-112), ``append`` writes the new rows behind the live ones
  (cat_tensors_to_optimizer + update_records("densify"), :74-94) -- no per-tensor ``torch.cat`` / boolean-index
  reallocation, so P can change every step without touching the allocator.
"""
import ctypes

import torch

from . import _check, _load, _ptr, _stream, _vp


def knn_mean_dist2(points):
    """(KDTree(points).query(points, k=4)[0][:, 1:] ** 2).mean(1) on the device; points [N,3] float32 CUDA."""
    lib = _load()
    if not points.is_cuda or points.dtype != torch.float32 or points.ndim != 2 or points.shape[1] != 3:
        raise RuntimeError("knn_mean_dist2 needs a float32 CUDA tensor [N,3]")
    pts = points.detach().contiguous()
    out = torch.empty(pts.shape[0], dtype=torch.float32, device=pts.device)
    _check(lib.sb200_knn_mean_dist2(pts.shape[0], _ptr(pts), _ptr(out), _stream(pts.device)), "sb200_knn_mean_dist2")
    return out


class SurfelStore:
    """Per-surfel arrays with a common live row count ``P`` inside buffers of ``capacity`` rows.

    ``spec``: {name: (columns, dtype)}.  ``view(name)`` is the live ``[P, columns]`` window
    of the current buffer (a view: valid until the next ``prune`` / growing ``append``)."""

    def __init__(self, spec, device, capacity=1 << 16):
        self.spec = dict(spec)
        self.device = device
        self.capacity = int(capacity)
        self.P = 0
        self.cur = 0
        self._bufs = [self._alloc(self.capacity), self._alloc(self.capacity)]
        self._scratch = torch.empty(0, dtype=torch.uint8, device=device)
        self._kept = torch.zeros(1, dtype=torch.int32).pin_memory() if torch.cuda.is_available() else torch.zeros(1, dtype=torch.int32)

    def _alloc(self, cap):
        return {n: torch.zeros((cap, c), dtype=dt, device=self.device) for n, (c, dt) in self.spec.items()}

    def view(self, name):
        return self._bufs[self.cur][name][:self.P]

    def views(self):
        return {n: self.view(n) for n in self.spec}

    def reserve(self, rows):
        """Grow both buffer sets geometrically so that ``rows`` fit; live rows are kept."""
        if rows <= self.capacity:
            return
        cap = self.capacity
        while cap < rows:
            cap *= 2
        new = [self._alloc(cap), self._alloc(cap)]
        for n in self.spec:
            new[0][n][:self.P].copy_(self._bufs[self.cur][n][:self.P])
        self._bufs, self.cur, self.capacity = new, 0, cap

    def append(self, rows):
        """rows: {name: tensor [n, columns]} for some of the arrays; the others get zero rows (fresh Adam moments,
        scores: cat_tensors_to_optimizer's ``zeros_like`` and update_records("densify"))."""
        n = next(iter(rows.values())).shape[0]
        self.reserve(self.P + n)
        buf = self._bufs[self.cur]
        for name in self.spec:
            dst = buf[name][self.P:self.P + n]
            if name in rows:
                dst.copy_(rows[name].reshape(n, -1))
            else:
                dst.zero_()
        self.P += n
        return self.P

    def prune(self, delete_mask):
        """x = x[~delete_mask] for every array at once; returns the new P (one 4-byte read-back)."""
        lib = _load()
        P = self.P
        if delete_mask.numel() != P:
            raise RuntimeError("delete_mask must have P elements")
        if P == 0:
            return 0
        mask = delete_mask.contiguous()
        if mask.dtype != torch.bool:
            mask = mask != 0
        need = lib.sb200_prune_scratch_bytes(P)
        if self._scratch.numel() < need:
            self._scratch = torch.empty(int(need * 2), dtype=torch.uint8, device=self.device)
        names = list(self.spec)
        n = len(names)
        src, dst = self._bufs[self.cur], self._bufs[1 - self.cur]
        sp = (ctypes.c_void_p * n)(*[src[k].data_ptr() for k in names])
        dp = (ctypes.c_void_p * n)(*[dst[k].data_ptr() for k in names])
        rb = (ctypes.c_int * n)(*[src[k].element_size() * src[k].shape[1] for k in names])
        _check(lib.sb200_prune_rows(P, _ptr(mask), n, sp, dp, rb, _ptr(self._scratch), _vp(self._kept.data_ptr()),
                                    _stream(self.device)), "sb200_prune_rows")
        self.P = int(self._kept[0])
        self.cur = 1 - self.cur
        return self.P
