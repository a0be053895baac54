"""Worker for tests/test_ranges_cpu.py: every rank histograms only ITS index range (numpy), the histograms are
all-reduced over gloo, and every rank must arrive at the same key threshold as a sort of the whole vector."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from fricc_b200.ranges import index_range, locate_threshold  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    rng = np.random.default_rng(8)
    n = 30011
    x = rng.standard_normal(n) * 10.0 ** (-6.0 * rng.random(n) ** 2)
    j0, cnt = index_range(n, rank, world)
    mine = np.abs(x[j0:j0 + cnt]).view(np.uint64)

    def hist_fn(shift, prefix, prefix_bits):
        nbits = 3 if shift == 0 else 12
        sel = mine if prefix_bits == 0 else mine[(mine >> np.uint64(shift + nbits)) == np.uint64(prefix)]
        d = ((sel >> np.uint64(shift)) & np.uint64((1 << nbits) - 1)).astype(np.int64)
        h = torch.from_numpy(np.bincount(d, minlength=4096).astype(np.int64))
        dist.all_reduce(h)
        return h.numpy()

    keys = np.sort(np.abs(x).view(np.uint64))[::-1]
    ok = True
    for want in (1, 100, 65536 // 4, n):
        key_min, count = locate_threshold(hist_fn, want, small_bin=256)
        ok &= count == int((keys >= np.uint64(key_min)).sum()) and count >= want
        ok &= int((mine >= np.uint64(key_min)).sum()) <= count
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("THRESHOLD_OK" if int(flag) == 1 else "THRESHOLD_FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
