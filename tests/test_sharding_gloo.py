"""The N > 1 host logic on CPU: two ranks over gloo, each holding one shard of the offers.  Every rank derives its shard's
class rows (with the CPU oracle standing in for the per-shard kernel), the rows are all-gathered, laid out in global offer
order and committed; the result must equal the un-sharded oracle cycle."""
import os
import socket

import numpy as np
import pytest

from dcos_commons_b200 import sharded, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, n_offers, n_reqs, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    sys.path.insert(0, os.path.dirname(here))
    import torch
    import torch.distributed as dist
    from oracle_binding import Oracle
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cl = synth.make_cluster(n_offers, n_reqs, 8, 3, 4, 64, "c5")
    t = cl.tables
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    lo, hi = bounds[rank]
    stride = sharded.local_stride_words(shard_rows, hi - lo)
    # class rows of this shard: one fresh requirement per class is enough (fresh instances share their class row)
    first_req = [int(np.flatnonzero(t.class_id == c)[0]) for c in range(t.n_classes)]
    rows = np.zeros((t.n_classes, stride), np.uint32)
    if hi > lo:
        orc = Oracle()
        orc.tables = t
        orc.load_programs(t); orc.load_tasks(t); orc.load_offers(t, lo, hi); orc.load_requirements(t)
        for c, j in enumerate(first_req):
            r = orc.feasibility_rows(j, j + 1)[0]
            rows[c, : len(r)] = r[:stride]
        orc.close()
    gathered = [torch.zeros(rows.shape, dtype=torch.int32) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(rows.view(np.int32)))
    g = np.stack([x.numpy().view(np.uint32) for x in gathered])
    glob = sharded.global_rows(g, shard_rows, t.n_offers)
    dec = sharded.first_fit(glob, t.class_id, t.n_offers)
    np.save(os.path.join(out_dir, f"dec_{rank}.npy"), dec)
    dist.barrier()
    dist.destroy_process_group()


def test_shard_plan_properties():
    for n, w in [(1, 2), (1023, 2), (1024, 2), (1025, 2), (10_000, 4), (1_000_000, 8), (5, 8)]:
        shard_rows, bounds = sharded.shard_plan(n, w)
        assert shard_rows % 1024 == 0 and len(bounds) == w
        assert bounds[0][0] == 0 and bounds[-1][1] == n
        for r, (lo, hi) in enumerate(bounds):
            assert lo == min(n, r * shard_rows) and 0 <= hi - lo <= shard_rows
        assert sum(hi - lo for lo, hi in bounds) == n
    assert sharded.shard_plan(77, 1) == (0, [(0, 77)])


@pytest.mark.parametrize("n_offers", [3000, 1025])
def test_two_ranks_reach_the_unsharded_decisions(tmp_path, n_offers):
    import torch.multiprocessing as mp
    from oracle_binding import Oracle
    world, n_reqs = 2, 200
    port = _free_port()
    mp.spawn(_rank_main, args=(world, port, n_offers, n_reqs, str(tmp_path)), nprocs=world, join=True)
    cl = synth.make_cluster(n_offers, n_reqs, 8, 3, 4, 64, "c5")
    orc = Oracle().load(cl.tables)
    want, _ = orc.evaluate_cycle()
    for rank in range(world):
        got = np.load(tmp_path / f"dec_{rank}.npy")
        assert np.array_equal(got, want["offer_row"]), f"rank {rank}"
