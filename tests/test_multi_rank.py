"""N > 1 host logic on CPU: two `gloo` ranks shard the reads with `_lib.shard_bounds`, each scores its own contiguous
range (here with the CPU oracle standing in for the per-GPU kernel — this is a test of the sharding / concatenation
logic, not of the kernels), rank 0 gathers and concatenates with `_lib.concat_results`, and the result must equal the
unsharded run.  Also covers the max-over-ranks reduction bench.py uses for its timing.  No data-path collective
exists in the product; torch.distributed is plumbing for the test / bench only."""
import os
import socket
import sys

import numpy as np
import pytest

import util

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_seg, out_dir):
    sys.path.insert(0, util.ROOT)
    sys.path.insert(0, os.path.join(util.ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        from sctagger_b200 import _lib
        from sctagger_b200.match_trie import MatchResult
        rng = np.random.default_rng(31)              # same data on every rank, like one input file
        bc, sb, off = util.random_case(rng, 200, n_seg, bl=16, dirty=0.1)
        a, b = _lib.shard_bounds(n_seg, world)[rank]
        part = MatchResult(*oracle.match(bc, sb, off[a:b + 1], 2, threads=16, algo="dp"))
        # bench.py's timing reduction: max over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == float(world)
        gathered = [None] * world
        dist.gather_object((part.min_ed, part.n_ties, part.tie_offsets, part.tie_ids), gathered if rank == 0 else None, dst=0)
        if rank == 0:
            whole = _lib.concat_results([MatchResult(*g) for g in gathered])
            full = MatchResult(*oracle.match(bc, sb, off, 2, threads=16, algo="dp"))
            assert whole.equals(full)
            assert int(whole.tie_offsets[-1]) == whole.tie_ids.size
            with open(os.path.join(out_dir, "ok"), "w") as f:
                f.write(str(int((whole.min_ed >= 0).sum())))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_seg", [(2, 501), (2, 1)])
def test_sharded_equals_unsharded_two_gloo_ranks(tmp_path, world, n_seg):
    from oracle import oracle
    oracle.build()      # once, before the ranks race to build it
    mp.spawn(_worker, args=(world, _free_port(), n_seg, str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / "ok").exists()


def test_shard_bounds_cover_and_are_contiguous():
    from sctagger_b200 import _lib
    for n in (0, 1, 7, 8, 1000, 10_000_019):
        for parts in (1, 2, 3, 4, 8):
            b = _lib.shard_bounds(n, parts)
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(parts - 1))
            sizes = [y - x for x, y in b]
            assert max(sizes) - min(sizes) <= 1
