"""Sharded engines (world > 1) against the un-sharded engine and the oracle.  The ranks are emulated as several engine
handles on cuda:0, run one after the other — exactly the data flow of bench.py --gpus N, with the NCCL all-gather
replaced by a copy through the host, or with the engines attached to each other (oe_peer_attach) so that the pass itself
publishes the rows into every shard's exchange region.  tests/test_gpu_peer_exchange.py runs the attached path with one
process per GPU when the box has two."""
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


# ---- peers attached: K1 stores its rows into every shard's exchange region, evaluate_cycle waits for the flags ------------
def _attached_engines(t, world, max_rows=None):
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    engines = []
    for rank, (lo, hi) in enumerate(bounds):
        e = Engine(device=0, rank=rank, world=world)
        e.load(t, lo, hi, shard_rows)
        engines.append(e)
    sharded.attach_local(engines, max_rows or engines[0].factored_rows())
    return engines, bounds, shard_rows


@pytest.mark.parametrize("world", [2, 3, 8])
def test_peer_exchange_decisions_equal_unsharded(world):
    cl = synth.make_cluster(9000, 700, 8, 3, 4, 160, "c5")
    t = cl.tables
    want, _ = Oracle().load(t).evaluate_cycle()
    engines, bounds, shard_rows = _attached_engines(t, world)
    for cycle in range(3):   # three passes: both buffer parities, and the first one again
        for e in engines:
            e.class_feasibility(fetch=False)          # the pass of every shard: publishes into all exchange regions
        for rank, e in enumerate(engines):
            d, st = e.evaluate_cycle()                # waits for every shard's flag of this pass, then commits
            assert np.array_equal(d["offer_row"], want["offer_row"]), f"cycle {cycle} rank {rank}"
            lo, hi = bounds[rank]
            mine = (want["offer_row"] >= lo) & (want["offer_row"] < hi)
            assert d[mine].tobytes() == want[mine].tobytes()
        # the merged table: every shard's winners (ports, MOUNT slots, fresh reservation ids) on every rank that asks for it
        for rank, e in enumerate(engines):
            assert e.gather_decisions().tobytes() == want.tobytes(), f"cycle {cycle} rank {rank}"
    for e in engines:
        e.close()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_replay_records_launches_on_every_rank(world):
    """Several cycles with the launches recorded in between (oe_apply_launches) on sharded, peer-attached engines: every rank
    must end up with the oracle's task table although a winning offer's placement columns live on one shard only."""
    import copy
    cl = synth.make_cluster(6000, 900, 8, 3, 4, 120, "c4")
    t = cl.tables
    per = 300
    orc = Oracle()
    engines = None
    for cyc in range(3):
        sub = copy.copy(t)
        sl = slice(cyc * per, (cyc + 1) * per)
        sub.class_id, sub.r_pod_type_id, sub.r_pod_index = t.class_id[sl].copy(), t.r_pod_type_id[sl].copy(), t.r_pod_index[sl].copy()
        sub.launch_begin = (t.launch_begin[sl.start: sl.stop + 1] - t.launch_begin[sl.start]).astype(np.int32)
        sub.launch_name_id = t.launch_name_id[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.launch_filter_bits = t.launch_filter_bits[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.normalise()
        if cyc == 0:
            orc.load(sub)
            engines, bounds, shard_rows = _attached_engines(sub, world, max_rows=64)
        else:
            orc.tables = sub
            orc.load_requirements(sub)
            for e in engines:
                e.tables = sub
                e.load_requirements(sub)
        want, _ = orc.evaluate_cycle()
        for e in engines:
            e.class_feasibility(fetch=False)
        for e in engines:
            assert np.array_equal(e.evaluate_cycle()[0]["offer_row"], want["offer_row"]), f"cycle {cyc}"
        orc.apply_launches()
        for e in engines:
            e.apply_launches()
        tc = orc.export_tasks()
        for rank, e in enumerate(engines):
            tg = e.export_tasks()
            for k in tc:
                assert np.array_equal(tc[k], tg[k]), f"cycle {cyc} rank {rank}: task column {k}"
    for e in engines:
        e.close()
    orc.close()


@pytest.mark.parametrize("seed", range(3100, 3108))
def test_peer_exchange_fuzz(seed):
    t = random_tables(seed, n_offers=2500 + seed % 300, n_reqs=50, n_tasks=40, with_reserved=False)
    want, _ = Oracle().load(t).evaluate_cycle()
    engines, _, _ = _attached_engines(t, 2 + seed % 3)
    for e in engines:
        e.class_feasibility(fetch=False)
    for e in engines:
        d, _ = e.evaluate_cycle()
        assert np.array_equal(d["offer_row"], want["offer_row"])
    for e in engines:
        e.close()


def test_peer_exchange_new_tables_between_passes():
    """Different offers and requirements from one pass to the next (as in the scheduler): the exchange regions are reused."""
    world = 4
    cl = synth.make_cluster(6000, 300, 8, 3, 4, 100, "c5")
    engines, bounds, shard_rows = _attached_engines(cl.tables, world, max_rows=64)
    for k in range(4):
        t = synth.make_cluster(6000, 300 + 17 * k, 8, 3, 4, 100 + 10 * k, "c5", seed=77 + k).tables
        want, _ = Oracle().load(t).evaluate_cycle()
        for e, (lo, hi) in zip(engines, bounds):
            e.load(t, lo, hi, shard_rows)
            e.class_feasibility(fetch=False)
        for e in engines:
            d, _ = e.evaluate_cycle()
            assert np.array_equal(d["offer_row"], want["offer_row"]), f"pass {k}"
    for e in engines:
        e.close()


def test_peer_exchange_single_call_cycle_from_one_thread_per_shard():
    """oe_evaluate_cycle alone as the whole sharded cycle (it runs the pass itself), every shard driven by its own host
    thread as separate processes would: the device-side wait really waits here."""
    import threading
    world = 2
    cl = synth.make_cluster(9000, 700, 8, 3, 4, 160, "c5")
    t = cl.tables
    want, _ = Oracle().load(t).evaluate_cycle()
    engines, bounds, shard_rows = _attached_engines(t, world)
    for e in engines:               # sizes every buffer first: a cudaFree inside the threaded cycle would wait for the whole
        e.class_feasibility(fetch=False)   # device, i.e. for the other shard's waiting kernel (one process, one device only)
    for e in engines:
        e.evaluate_cycle()
    got, errs = [None] * world, []

    def run(rank):
        try:
            e = engines[rank]
            lo, hi = bounds[rank]
            for _ in range(3):
                e.load_requirements(t)      # makes the class rows stale: the next cycle call runs the pass itself
                got[rank], _ = e.evaluate_cycle()
        except Exception as ex:   # noqa: BLE001
            errs.append((rank, ex))

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errs, errs
    for rank in range(world):
        assert np.array_equal(got[rank]["offer_row"], want["offer_row"]), f"rank {rank}"
    for e in engines:
        e.close()


def test_peer_exchange_rejects_more_rows_than_exported_and_reports_a_missing_peer(monkeypatch):
    cl = synth.make_cluster(3000, 100, 8, 3, 4, 40, "c5")
    t = cl.tables
    shard_rows, bounds = sharded.shard_plan(t.n_offers, 2)
    engines = [Engine(device=0, rank=r, world=2) for r in range(2)]
    for e, (lo, hi) in zip(engines, bounds):
        e.load(t, lo, hi, shard_rows)
    sharded.attach_local(engines, 2)                      # 8 classes do not fit 2 rows
    with pytest.raises(abi.EngineError) as ei:
        engines[0].class_feasibility(fetch=False)
    assert ei.value.code == -5 and "export" in str(ei.value)
    monkeypatch.setenv("OE_PEER_TIMEOUT_MS", "200")
    sharded.attach_local(engines, 64)
    engines[0].class_feasibility(fetch=False)             # shard 1 never runs its pass
    with pytest.raises(abi.EngineError) as ei:
        engines[0].evaluate_cycle()
    assert ei.value.code == -2 and "shard 1" in str(ei.value)
    # detached again, the engines take the caller-side gather as before
    import torch
    for e in engines:
        e.peer_detach()
    want, _ = Oracle().load(t).evaluate_cycle()
    rows = [e.class_feasibility()[0] for e in engines]
    gathered = np.ascontiguousarray(np.stack(rows))
    for e in engines:
        ptr, nbytes = e.device_buffer(abi.BUF_CLASS_BITS_GLOBAL)
        dst = torch.as_tensor(sharded._CudaView(ptr, nbytes // 4), device="cuda")
        dst.copy_(torch.from_numpy(gathered.view(np.int32).reshape(-1)))
        torch.cuda.synchronize()
        assert np.array_equal(e.evaluate_cycle()[0]["offer_row"], want["offer_row"])
    for e in engines:
        e.close()
