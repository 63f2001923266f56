"""One rank of tests/test_gpu_dist.py::test_peer_gather_across_processes (not a test module itself).

Every rank sketches its own genomes on cuda:LOCAL_RANK, publishes the shard in its peer buffer and pulls the shards of all
ranks through CUDA-IPC mappings (kssd_b200_peer_init / _peer_connect / _sketch_publish / _sketch_gather_peers); then builds
the index, runs its block of triangle rows and the hit sets of all ranks are compared with the oracle on rank 0.
torch.distributed (gloo, CPU) only carries the IPC handles and the final hit lists; the per-step barrier is the library's own
(flags stored into the peers' memory by the publish kernel, awaited by the concat kernel)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import oracle  # noqa: E402
import util  # noqa: E402
from rabbitkssd_b200 import api, multigpu, synth  # noqa: E402


class View:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def dev_np(ptr, n, typestr, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return torch.as_tensor(View(ptr, n, typestr), device="cuda").cpu().numpy().view(dtype)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    k, s, l = 10, 6, 3
    n_local = 6
    fams = [g for f in range(world * 2) for g in synth.family(300 + f, 3, 90000 + 7000 * f, 0.03, n_contigs=2)]
    assert len(fams) == world * n_local
    ctx = api.Context(dev)
    ctx.set_params(k, s, l, util.shuffled_dim(k, s, l))
    # expected: all genomes sketched in one session on this rank's own context
    seq, _, off = synth.pack_batch(fams)
    want_h, want_off = ctx.sketch(seq, off)
    want_h, want_off = np.array(want_h, copy=True), np.array(want_off, copy=True)
    n_total = len(fams)

    handles = [None] * world
    dist.all_gather_object(handles, ctx.peer_init(n_local, 2 * len(want_h) // world + 64))
    ctx.peer_connect(b"".join(handles), world, rank)
    mine = fams[rank * n_local:(rank + 1) * n_local]
    seq_l, _, off_l = synth.pack_batch(mine)
    bounds = multigpu.triangle_bounds(n_total, world)
    for step in range(4):
        if step % 2 == rank % 2:
            time.sleep(0.05 * (1 + rank))  # skew the ranks: a reader must WAIT for the slower publishers (peer flags, no host barrier)
        ctx.sketch(seq_l, off_l)
        ctx.sketch_publish()
        ph, po, total, overflow = ctx.sketch_gather_peers()
        assert not overflow and total == len(want_h), (total, len(want_h))
        got_h = dev_np(ph, total, "<i4", np.uint32)
        got_o = dev_np(po, n_total + 1, "<i8", np.int64)
        assert np.array_equal(got_h, want_h), f"rank {rank} step {step}: gathered hashes differ"
        assert np.array_equal(got_o.astype(np.uint64), want_off.astype(np.uint64)), f"rank {rank} step {step}: offsets differ"
        ctx.index_build_device(ph, po, n_total)
        d_hits, n_hits = ctx.dist_rows_device(ph, po, n_total, True, 2 * k, 0.2, 0, rows=(bounds[rank], bounds[rank + 1]))
        hits = dev_np(d_hits, n_hits * 3, "<i4", np.uint32).reshape(-1, 3)
        every = [None] * world
        dist.all_gather_object(every, hits)
        if rank == 0:
            sizes = np.diff(want_off).astype(np.int64)
            allh, _, _ = api.refine_hits(np.concatenate(every), sizes, sizes, 2 * k, 0.2, True)
            hr, hc, hm = oracle.alldist(want_h, want_off, 2 * k, 0.2)
            o = np.lexsort((allh[:, 1], allh[:, 0]))
            assert len(hr) > 5 and np.array_equal(allh[o].astype(np.int64), np.stack([hr, hc, hm], axis=1).astype(np.int64))
    # a shard larger than the peer buffers is reported as overflow on every rank
    dist.barrier()
    ctx.close()
    print("PEER_OK", rank)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
