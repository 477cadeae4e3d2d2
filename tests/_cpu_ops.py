"""TEST DOUBLE for the multi-GPU host logic: the local steps of cuda_samples_b200.multi_gpu done on CPU
tensors with the oracle, so that the collective plumbing (shard carry, splitters, count exchange,
all-to-all layout) can be exercised with gloo and world_size 2 where there is no GPU.
Lives in tests/ on purpose: the product package has no CPU path."""
import numpy as np
import torch

import oracle


def _np(t, n=None):
    a = t.numpy()
    return a if n is None else a[:n]


class OracleOps:
    device_type = "cpu"

    def histogram_fused(self, data, nbytes):
        d = _np(data, nbytes)
        bins = np.concatenate([oracle.histogram256(d), oracle.histogram64(d)]).view(np.int32)
        return torch.from_numpy(bins.copy())

    def reduce(self, x, n):
        v = oracle.reduce(_np(x, n))
        return torch.tensor([v], dtype=x.dtype)

    def reduce_partial(self, x, n, slot):
        slot[0] = float(oracle.reduce(_np(x, n)))

    def scan_total(self, src, n):
        s = int(_np(src, n).view(np.uint32).astype(np.uint64).sum() % 2**32)
        return torch.from_numpy(np.array([s], np.uint32).view(np.int32).copy())

    def scan_with_carry(self, dst, src, n, carry):
        c = 0 if carry is None else int(carry.numpy().view(np.uint32)[0])
        out = oracle.scan_exclusive(_np(src, n).view(np.uint32), 1, n, c)
        dst[:n] = torch.from_numpy(out.view(np.int32).copy())

    def partition(self, keys, vals, n, splitters, buckets, direction):
        k, v = _np(keys, n).view(np.uint32), _np(vals, n).view(np.uint32)
        flip = np.uint32(0 if direction else 0xFFFFFFFF)
        spl = _np(splitters).view(np.uint32)[:buckets - 1]
        b = np.searchsorted(spl, k ^ flip, side="right")
        order = np.argsort(b, kind="stable")
        dk, dv = torch.empty_like(keys), torch.empty_like(vals)
        dk[:n] = torch.from_numpy(k[order].view(np.int32).copy())
        dv[:n] = torch.from_numpy(v[order].view(np.int32).copy())
        counts = torch.from_numpy(np.bincount(b, minlength=buckets).astype(np.int32))
        return dk, dv, counts

    def sort(self, keys, vals, n, direction):
        sk, sv = oracle.stable_sort_kv(_np(keys, n).view(np.uint32), _np(vals, n).view(np.uint32), direction)
        dk, dv = torch.empty_like(keys), torch.empty_like(vals)
        dk[:n] = torch.from_numpy(sk.view(np.int32).copy())
        dv[:n] = torch.from_numpy(sv.view(np.int32).copy())
        return dk, dv
