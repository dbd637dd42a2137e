"""Sharded engines (world > 1) against the un-sharded engine and the oracle.  The ranks are emulated as several engine
handles on cuda:0, run one after the other — exactly the data flow of bench.py --gpus N, with the NCCL all-gather
replaced by a copy through the host."""
import numpy as np
import pytest

from dcos_commons_b200 import Engine, abi, sharded, synth
from fuzz_tables import random_tables
from oracle_binding import Oracle

pytestmark = pytest.mark.gpu


def _sharded_cycle(t, world):
    import torch
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    engines = []
    for rank, (lo, hi) in enumerate(bounds):
        e = Engine(device=0, rank=rank, world=world)
        e.load(t, lo, hi, shard_rows)
        engines.append(e)
    rows = [e.class_feasibility()[0] for e in engines]              # K0 + K1 per shard
    gathered = np.ascontiguousarray(np.stack(rows))                 # [world, rows, stride]: what the all-gather produces
    decisions = []
    for e in engines:
        ptr, nbytes = e.device_buffer(abi.BUF_CLASS_BITS_GLOBAL)
        assert nbytes == gathered.nbytes
        dst = torch.as_tensor(sharded._CudaView(ptr, nbytes // 4), device="cuda")
        dst.copy_(torch.from_numpy(gathered.view(np.int32).reshape(-1)))
        torch.cuda.synchronize()
        d, st = e.evaluate_cycle()
        decisions.append(d)
    for e in engines:
        e.close()
    return decisions, gathered, shard_rows


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_decisions_equal_unsharded(world):
    cl = synth.make_cluster(9000, 700, 8, 3, 4, 160, "c5")
    t = cl.tables
    want, _ = Oracle().load(t).evaluate_cycle()
    decisions, gathered, shard_rows = _sharded_cycle(t, world)
    # offer_row is global; dynamic ports / mount slots are only materialised by the rank that holds the winning row
    for rank, d in enumerate(decisions):
        assert np.array_equal(d["offer_row"], want["offer_row"]), f"rank {rank}"
    lo_hi = sharded.shard_plan(t.n_offers, world)[1]
    for rank, d in enumerate(decisions):
        lo, hi = lo_hi[rank]
        mine = (want["offer_row"] >= lo) & (want["offer_row"] < hi)
        assert d[mine].tobytes() == want[mine].tobytes(), f"rank {rank}: winners on its own shard must be fully materialised"
    # the gathered rows, laid out globally, are the un-sharded class rows
    single = Engine(device=0).load(t)
    f, _ = single.class_feasibility()
    glob = sharded.global_rows(gathered, shard_rows, t.n_offers)
    assert np.array_equal(glob, f[:, : glob.shape[1]])
    single.close()


@pytest.mark.parametrize("seed", range(3000, 3012))
def test_sharded_fuzz(seed):
    t = random_tables(seed, n_offers=2500 + seed % 300, n_reqs=50, n_tasks=40, with_reserved=False)
    want, _ = Oracle().load(t).evaluate_cycle()
    decisions, _, _ = _sharded_cycle(t, 2)
    for d in decisions:
        assert np.array_equal(d["offer_row"], want["offer_row"])
