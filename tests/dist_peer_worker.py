"""Worker for tests/test_peer_ipc_gpu.py: the sharded step of bench.py (ViewEngine: plugin calls on L-slice views, stores
into the peers' copies through CUDA IPC mappings of a SymmArena) as SEPARATE PROCESSES, small shapes.

  * one GPU per process when the box has that many: the peer stores cross NVLink and the barrier is the device-side
    one (fricc_peer_barrier: epoch flags in the arena) -- kernels that wait on each other run on DIFFERENT GPUs;
  * fewer GPUs than processes: the processes share cuda:0, the IPC mapping and the stores are the same code, but the
    barrier is a host one (stream synchronize + gloo barrier): kernels of two processes time-slice on one GPU, so
    nothing on the device may spin on another process's flag.
Every rank then compares its OWN copy of every output with the unsharded contraction from the oracle port."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from conftest import TERMS  # noqa: E402
from fricc_b200 import _lib, py_rccsd  # noqa: E402
from fricc_b200 import dist as fd  # noqa: E402
from oracle import oracle  # noqa: E402


class HostBarrierEngine(fd.ViewEngine):
    def barrier(self):
        self.join()
        torch.cuda.synchronize()
        dist.barrier()


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    _lib.require_cuda()
    ngpu = torch.cuda.device_count()
    own_gpu = ngpu >= world
    dev_index = rank if own_gpu else 0
    torch.cuda.set_device(dev_index)
    device = torch.device("cuda", dev_index)
    _lib.check(_lib.lib().fricc_set_device(dev_index))
    fd.use_torch_stream()
    comm = fd.Comm(rank, world)
    P = oracle.port()
    rng = np.random.default_rng(21)                     # the same stream on every rank: identical inputs
    no, nv = 4, 12
    ext = {"o": no, "v": nv}
    n = no * no * nv * nv
    x = (rng.standard_normal(n) * 10.0 ** (-4.0 * rng.random(n) ** 2)).reshape(no, no, nv, nv)
    draws = rng.integers(0, 2**31 - 1, size=n).astype(np.int32)
    T = py_rccsd.SparseTensor4d(x, x.shape, n // 5, "fri", "systematic", False, draws=draws)
    flat, vals = T.export()
    host_w = {t: rng.random([ext[c] for c in TERMS[t][0]]) for t in TERMS}
    ok = True
    # (kernel form, peer stores on a side stream, copy engines: 0 = every slice through the store kernel (the default),
    #  1 = every slice that collapses to rows of >= 8 bytes through cudaMemcpy(2D)Async, 2 = the same on one private
    #  stream per peer)
    for tc_mode, side, copy_engine in ((0, False, 0), (2, True, 0), (2, True, 1), (2, True, 2)):
        py_rccsd.set_option("tc_mode", tc_mode)
        py_rccsd.set_option("push_copy_engine", copy_engine)
        py_rccsd.set_option("push_min_row_bytes", 8)
        copies0 = py_rccsd.get_stat("push_copies")
        arena = fd.SymmArena(8 * (8 * no * no * nv * nv + no ** 4) + 65536, comm, device)
        side_stream = fd.SideStream(device) if side else None
        eng = (fd.ViewEngine if own_gpu else HostBarrierEngine)(comm, device, arena if own_gpu else None, side_stream)
        eng.arena = arena
        outs = {}
        for t in sorted(TERMS):
            outs[t] = arena.tensor([ext[c] for c in TERMS[t][1]]).zero_()
        torch.cuda.synchronize()
        dist.barrier()
        for rep in range(3):                             # epochs advance; the outputs are re-zeroed behind a barrier
            for t in outs:
                outs[t].zero_()
            eng.barrier()
            for t in sorted(TERMS):
                Wd = torch.from_numpy(host_w[t]).to(device)
                if t == "2323":                          # the ladder: the rank holds only its slice of the operand
                    lo, hi = eng.my_range(nv)
                    part = Wd[lo:hi].contiguous()
                    eng.publish(eng.contract(t, part, T, outs[t], w_range=(lo, hi)))
                else:
                    eng.publish(eng.contract(t, Wd, T, outs[t]))
            eng.barrier()
            torch.cuda.synchronize()
            for t in sorted(TERMS):
                want = np.zeros([ext[c] for c in TERMS[t][1]])
                P.contract(t, host_w[t], flat, vals, x.shape, want)
                got = outs[t].cpu().numpy()
                good = np.abs(got - want).max() <= 1e-10 * max(1.0, np.abs(want).max())
                if not good:
                    print(f"rank {rank}: term {t} tc_mode {tc_mode} rep {rep}: max err {np.abs(got - want).max():.3e}", flush=True)
                ok &= bool(good)
            dist.barrier()                               # nobody re-zeroes while a peer is still reading its copy
        if world > 1 and (py_rccsd.get_stat("push_copies") > copies0) != (copy_engine >= 1):
            print(f"rank {rank}: copy-engine path {'not ' if copy_engine else ''}taken with push_copy_engine={copy_engine}", flush=True)
            ok = False
        del outs, eng
        torch.cuda.synchronize()
        dist.barrier()
        arena.close()
    py_rccsd.set_option("tc_mode", 1)
    py_rccsd.set_option("push_min_row_bytes", 4096)
    py_rccsd.set_option("push_copy_engine", 0)
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(("PEER_OK " if int(flag) == 1 else "PEER_FAIL ") + ("device-barrier" if own_gpu else "host-barrier"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
