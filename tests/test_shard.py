"""N > 1 host logic on CPU: world_size-2 gloo processes shard a corpus, each indexes + sizes its own files,
and the gathered table covers every file exactly once (no data-path collective exists to test)."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from avrecode_b200 import shard


def test_assign_is_a_balanced_disjoint_cover():
    w = [5, 1, 9, 3, 3, 7, 2, 8]
    for world in (1, 2, 3, 8, 11):
        parts = shard.assign(w, world)
        assert sorted(i for p in parts for i in p) == list(range(len(w)))
        loads = [sum(w[i] for i in p) for p in parts]
        assert max(loads) <= sum(w) / world + max(w)            # LPT bound
        assert parts == shard.assign(w, world)                    # deterministic
    assert shard.assign([], 4) == [[], [], [], []]
    with pytest.raises(ValueError):
        shard.assign(w, 0)


def _worker(rank, world, port, golden_dir, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import avrecode_b200 as A
        names = sorted(n for n in os.listdir(golden_dir) if n.endswith(".h264"))
        sizes = [os.path.getsize(os.path.join(golden_dir, n)) for n in names]
        mine = shard.my_items(sizes, rank, world)
        local = {}
        for i in mine:   # the host half of the path runs without a GPU: slice index of my files
            with open(os.path.join(golden_dir, names[i]), "rb") as f:
                ents = A.index_h264(f.read())
            local[i] = (len(ents), sum(e.desc.payload_size for e in ents))
        table = shard.gather_sizes(local)
        t = shard.max_over_ranks(float(rank + 1))
        s = shard.sum_over_ranks(float(len(mine)))
        q.put((rank, mine, table, t, s, len(names)))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding(built):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    golden = os.path.join(ROOT, "tests", "golden")
    procs = [ctx.Process(target=_worker, args=(r, 2, port, golden, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort()
    (r0, m0, t0, x0, s0, n), (r1, m1, t1, x1, s1, _) = got
    assert sorted(m0 + m1) == list(range(n)) and not set(m0) & set(m1)
    assert t0 == t1 and sorted(t0) == list(range(n))              # every rank sees the full metadata table
    assert all(cnt > 0 and nbytes > 0 for cnt, nbytes in t0.values())
    assert x0 == x1 == 2.0 and s0 == s1 == float(n)               # max / sum over ranks
