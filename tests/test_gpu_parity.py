"""GPU parity: the CUDA engine, driven through the C ABI, against the CPU oracle on identical encoded tables."""
import os

import numpy as np
import pytest

from dcos_commons_b200 import Engine, abi, synth, unpack_bits
from fuzz_tables import random_tables
from oracle_binding import Oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    e = Engine(device=0)
    yield e
    e.close()


def _compare_all(eng, t, two_cycles=True):
    orc = Oracle().load(t)
    eng.load(t)
    N = t.n_offers
    m_cpu, _ = orc.feasibility()
    m_gpu, st = eng.feasibility()
    assert m_gpu.shape == m_cpu.shape
    assert np.array_equal(unpack_bits(m_gpu, N), unpack_bits(m_cpu, N))
    assert not m_gpu[:, (N + 31) // 32:].any(), "padding words must be zero"
    d_cpu, s_cpu = orc.evaluate_cycle()
    d_gpu, s_gpu = eng.evaluate_cycle()
    assert d_gpu.tobytes() == d_cpu.tobytes()
    assert s_gpu["n_placed"] == s_cpu["n_placed"]
    if two_cycles and t.launch_begin is not None:
        orc.apply_launches()
        eng.apply_launches()
        tc, tg = orc.export_tasks(), eng.export_tasks()
        for k in tc:
            assert np.array_equal(tc[k], tg[k]), f"task column {k} differs after apply_launches"
        d_cpu2, _ = orc.evaluate_cycle()
        d_gpu2, _ = eng.evaluate_cycle()
        assert d_gpu2.tobytes() == d_cpu2.tobytes()
    orc.close()


@pytest.mark.parametrize("seed", range(120))
def test_fuzz_matrix_cycle_and_replay(engine, seed):
    _compare_all(engine, random_tables(seed))


@pytest.mark.parametrize("seed", range(1000, 1010))
def test_fuzz_wider(engine, seed):
    _compare_all(engine, random_tables(seed, n_offers=1500 + seed % 7, n_reqs=60, n_tasks=150, big_dicts=True))


@pytest.mark.parametrize("seed", range(2000, 2060))
def test_fuzz_new_pod_pipelines(engine, seed):
    """Pipelines of new pods only — the shape the engine folds (scalar chains -> thresholds, ports -> counts)."""
    _compare_all(engine, random_tables(seed, n_offers=300 + seed % 50, new_pods_only=True))


def test_scalar_chain_boundaries(engine):
    """Offers that sit exactly on, one ulp below and one ulp above what a chain of deductions needs."""
    from dcos_commons_b200 import tables as tb
    chains = [[0.1, 0.2], [0.1, 0.1, 0.1], [0.3], [1.0, 0.1, 0.2, 0.7], [0.1] * 7, [256.0, 0.1 + 0.2, 32.0]]
    prog = tb.ProgramBuilder()
    for ch in chains:
        prog.add_class([("exec", -1)] + [("scalar", 0, a) for a in ch] + [("scalar", 1, 1.0), ("nop",)])
    rows = []
    for ch in chains:
        exact = 0.0
        for a in reversed(ch):          # smallest start that leaves exactly zero, walking the chain backwards
            exact = exact + a
        for base in (exact, sum(ch), float(np.sum(np.array(ch)))):
            for k in range(-3, 4):
                v = base
                for _ in range(abs(k)):
                    v = np.nextafter(v, np.inf if k > 0 else -np.inf)
                rows.append(dict(scalars={(0, 0): float(v), (0, 1): 1.0}))
    rows.append(dict(scalars={(0, 0): float("nan"), (0, 1): 1.0}))
    rows.append(dict(scalars={(0, 0): float("inf"), (0, 1): 1.0}))
    rows.append(dict(scalars={(0, 1): 1.0}))                      # no cpus entry at all
    offers = tb.offers_from_rows(rows, S=1, K=4, R=1, D=0, A=1)
    M = len(chains) * 3
    class_id = np.arange(M) % len(chains)
    t = tb.make_tables(dict(n_agents=len(rows), n_hostnames=len(rows), n_pod_types=1), prog, tb.empty_tasks(), offers, class_id,
                       np.zeros(M, np.int32), np.arange(M))
    _compare_all(engine, t, two_cycles=False)


def test_unsorted_port_ranges_are_rejected(engine):
    t = random_tables(11, n_offers=20, n_reqs=3, n_tasks=0)
    if t.port_range.shape[1] < 2:
        t = random_tables(12, n_offers=20, n_reqs=3, n_tasks=0)
    t.port_range = t.port_range.copy()
    if t.port_range.shape[1] >= 2:
        t.port_range[0, 0, 3] = (2000) | (2100 << 16)
        t.port_range[0, 1, 3] = (2050) | (2200 << 16)      # overlaps the first
        with pytest.raises(abi.EngineError):
            engine.load(t)


@pytest.mark.parametrize("seed", range(40))
def test_failed_stage_masks(engine, seed):
    """Per-stage outcomes (every stage runs, failed stages do not deduct) against the oracle's literal pipeline."""
    t = random_tables(seed)
    orc = Oracle().load(t)
    engine.load(t)
    want, bits = orc.stage_failures()
    got = engine.stage_failures()
    assert np.array_equal(got, want)
    assert np.array_equal(got == 0, unpack_bits(bits, t.n_offers))
    if t.n_reqs > 2:
        assert np.array_equal(engine.stage_failures(1, t.n_reqs - 1), want[1:-1])
    orc.close()


@pytest.mark.parametrize("seed", range(30))
def test_unexpected_resource_scan(engine, seed):
    """DefaultScheduler.getUnexpectedResources: reservations whose id is not in the keep set, on offers the cycle left unused."""
    t = random_tables(seed, n_offers=150 + seed)
    rng = np.random.default_rng(seed)
    keep = rng.integers(0, 2 ** 32, (t.dict_sizes["n_res_ids"] + 31) // 32, dtype=np.uint64).astype(np.uint32)
    orc = Oracle().load(t)
    engine.load(t)
    for unused_only in (False, True):
        assert engine.unexpected_resources(keep, unused_only)[0].tolist() == orc.unexpected_resources(keep, unused_only)[0].tolist()
    orc.evaluate_cycle()
    engine.evaluate_cycle()
    for unused_only in (False, True):
        got, n_got = engine.unexpected_resources(keep, unused_only)
        want, n_want = orc.unexpected_resources(keep, unused_only)
        assert got.tolist() == want.tolist() and n_got == n_want == int(want.sum())
    orc.close()


def test_async_loads_give_the_same_answers_and_still_validate():
    eng = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
    for seed in (1, 2, 3):
        t = random_tables(seed, n_offers=400)
        orc = Oracle().load(t)
        eng.load(t)
        assert eng.evaluate_cycle()[0].tobytes() == orc.evaluate_cycle()[0].tobytes()
        eng.load(t)
        assert np.array_equal(eng.feasibility()[0], orc.feasibility()[0])
        orc.close()
    bad = random_tables(3, n_offers=50, n_reqs=5, n_tasks=10)
    bad.hostname_id = bad.hostname_id.copy()
    bad.hostname_id[7] = 10 ** 6
    eng.load(bad)                      # enqueued only ...
    with pytest.raises(abi.EngineError):
        eng.evaluate_cycle()           # ... reported by the evaluation, before any kernel gathers through the id
    with pytest.raises(abi.EngineError):
        eng.evaluate_cycle()           # and the handle stays unusable until tables are loaded again
    t = random_tables(4)
    eng.load(t)
    assert eng.evaluate_cycle()[0].tobytes() == Oracle().load(t).evaluate_cycle()[0].tobytes()
    eng.close()


@pytest.mark.parametrize("seed", range(48))
def test_async_loads_find_redeployed_pods_on_the_device(seed):
    """OE_CFG_ASYNC_LOADS: oe_load_requirements does nothing per requirement on the host; k_find_twins counts the requirements whose
    pod is deployed, and only when there are some does the evaluation call run the host-side search.  Fuzzed tables (some with
    re-deployed pods, some without), two cycles with the launches recorded in between (after which every placed pod is deployed)."""
    eng = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
    _compare_all(eng, random_tables(seed, n_offers=200 + 13 * seed))
    eng.close()


@pytest.mark.parametrize("seed", range(6))
def test_async_loads_with_more_match_records_than_the_buffer_holds(seed):
    """Many requirements that name the SAME deployed pod, which has several tasks: k_match_tasks finds tasks x requirements records,
    more than its buffer holds, and the evaluation call falls back to the host-side search over the caller's tables."""
    t = random_tables(seed, n_offers=300, n_reqs=120, n_tasks=40, with_reserved=False)
    t.normalise()
    ty, ix = int(t.t_pod_type_id[0]) if t.t_pod_type_id[0] >= 0 else 0, 5
    t.t_pod_type_id = t.t_pod_type_id.copy(); t.t_pod_index = t.t_pod_index.copy()
    t.t_pod_type_id[:30] = ty; t.t_pod_index[:30] = ix                  # one pod of 30 tasks ...
    t.r_pod_type_id = np.full_like(t.r_pod_type_id, ty); t.r_pod_index = np.full_like(t.r_pod_index, ix)   # ... named by all 120 requirements
    eng = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
    _compare_all(eng, t)
    eng.close()


@pytest.mark.parametrize("redeploy_every", [0, 1, 7])
def test_async_loads_keep_no_pointer_into_the_callers_tables(redeploy_every):
    """The caller's task and requirement columns must stay valid until the evaluation call after the loads returns, and no longer:
    they are overwritten right after it, and everything that reads the host-side copies afterwards (requirement rows, launches,
    the re-deployed pods of the next cycle) must still see what was loaded."""
    import copy
    t = synth.make_cluster(3000, 240, 8, 3, 4, 120, "c5", seed=77, redeploy_every=redeploy_every).tables
    t.normalise()
    ref = Engine(device=0).load(copy.deepcopy(t))
    eng = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS).load(t)
    want, _ = ref.evaluate_cycle()
    got, _ = eng.evaluate_cycle()
    assert got.tobytes() == want.tobytes()
    for name in ("class_id", "r_pod_type_id", "r_pod_index", "t_name_id", "t_pod_type_id", "t_pod_index", "launch_begin", "launch_name_id"):
        getattr(t, name)[...] = 0x5a5a5a
    rows_w, n_w = ref.requirement_rows()
    rows_g, n_g = eng.requirement_rows()
    assert n_w == n_g and np.array_equal(rows_w, rows_g)
    assert (n_w > 8) == (redeploy_every > 0)
    ref.apply_launches(); eng.apply_launches()
    tw, tg = ref.export_tasks(), eng.export_tasks()
    for k in tw:
        assert np.array_equal(tw[k], tg[k]), k
    want2, _ = ref.evaluate_cycle()
    got2, _ = eng.evaluate_cycle()
    assert got2.tobytes() == want2.tobytes()
    assert np.array_equal(ref.requirement_rows()[0], eng.requirement_rows()[0])
    eng.close(); ref.close()


@pytest.mark.parametrize("flags", [abi.OE_CFG_ASYNC_LOADS, abi.OE_CFG_ASYNC_LOADS | abi.OE_CFG_MATERIALIZE_MATRIX])
def test_async_offer_table_arrives_and_is_evaluated_in_row_chunks(flags, monkeypatch):
    """OE_CFG_ASYNC_LOADS: the offer table goes over the bus in row chunks and the pass evaluates (and, with the matrix
    requested, expands) every chunk as it lands.  Same answers as a load that waits; a bad id in a late chunk is reported."""
    monkeypatch.setenv("OE_OFFER_CHUNKS", "5")
    eng = Engine(device=0, flags=flags)
    ref = Engine(device=0)
    for seed, n in ((11, 40000), (12, 70001), (13, 8192 * 3)):
        cl = synth.make_cluster(n, 900, 8, 3, 4, 200, "c5", seed=seed)
        t = cl.tables
        ref.load(t)
        want, _ = ref.evaluate_cycle()
        eng.load(t)
        got, st = eng.evaluate_cycle()
        assert got.tobytes() == want.tobytes()
        if flags & abi.OE_CFG_MATERIALIZE_MATRIX:     # the matrix the chunked expansion left in device memory
            import torch
            ptr, nbytes = eng.device_buffer(abi.BUF_MATRIX)
            from dcos_commons_b200 import sharded
            m = torch.as_tensor(sharded._CudaView(ptr, nbytes // 4), device="cuda").cpu().numpy().view(np.uint32).reshape(t.n_reqs, -1)
            full, _ = ref.feasibility()
            assert np.array_equal(unpack_bits(m, t.n_offers), unpack_bits(full, t.n_offers))
        eng.load(t)
        f1, _ = eng.feasibility()
        f2, _ = ref.feasibility()
        assert np.array_equal(f1, f2)
        eng.load(t)
        c1, _ = eng.class_feasibility()
        c2, _ = ref.class_feasibility()
        assert np.array_equal(c1, c2)
    # fuzzed tables (general interpreter, reserved side table) through the chunked path
    for seed in (21, 22, 23):
        t = random_tables(seed, n_offers=20000)
        orc = Oracle().load(t)
        eng.load(t)
        assert eng.evaluate_cycle()[0].tobytes() == orc.evaluate_cycle()[0].tobytes()
        orc.close()
    cl = synth.make_cluster(40000, 100, 8, 3, 4, 50, "c5", seed=5)
    bad = cl.tables
    bad.zone_id = bad.zone_id.copy()
    bad.zone_id[39000] = 10 ** 6              # last chunk
    eng.load(bad)
    with pytest.raises(abi.EngineError) as ei:
        eng.evaluate_cycle()
    assert "13" in str(ei.value)
    with pytest.raises(abi.EngineError):
        eng.evaluate_cycle()                  # offers must be loaded again
    eng.load(bad)
    with pytest.raises(abi.EngineError):
        eng.stage_failures(0, 1)              # an entry that is not a pass waits for the whole table and reports, too
    eng.load(cl.tables if False else synth.make_cluster(40000, 100, 8, 3, 4, 50, "c5", seed=5).tables)
    assert eng.evaluate_cycle()[0].tobytes() == ref.load(synth.make_cluster(40000, 100, 8, 3, 4, 50, "c5", seed=5).tables).evaluate_cycle()[0].tobytes()
    eng.close(); ref.close()


def test_small_tables_fetched_from_pinned_host_memory_while_offers_arrive():
    """OE_CFG_ASYNC_LOADS with every column in pinned memory (what bench.py and a JNI adapter with registered direct buffers do): the
    task and requirement tables are read by k_fetch_host while the offer chunks are on the bus.  Odd sizes: 16-byte tails, columns
    whose slots are not 16-byte aligned among themselves."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    eng = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
    ref = Engine(device=0)
    for seed, (n, m, dep) in enumerate(((30001, 3001, 2501), (17003, 1999, 4099), (9000, 1500, 1237))):
        t = synth.make_cluster(n, m, 8, 3, 4, dep, "c5", seed=40 + seed).tables
        t.normalise()
        pt = bench.pin_tables(t)
        want, _ = ref.load(t).evaluate_cycle()
        for order in ("tasks-first", "offers-first"):
            eng.tables = pt
            eng.load_programs(pt)
            if order == "tasks-first":
                eng.load_tasks(pt)
            eng.load_offers(pt)
            if order != "tasks-first":
                eng.load_tasks(pt)
            eng.load_requirements(pt)
            got, _ = eng.evaluate_cycle()
            assert got.tobytes() == want.tobytes(), (seed, order)
    eng.close(); ref.close()


def _patched(t, rows, src_rows):
    """A copy of the tables whose offer rows `rows` hold what rows `src_rows` held."""
    import copy
    p = copy.copy(t)
    p._keep = []
    for name in abi.CycleTables.OFFER_COLUMNS:
        col = getattr(t, name).copy()
        col[..., rows] = getattr(t, name)[..., src_rows]
        setattr(p, name, col)
    return p


@pytest.mark.parametrize("seed", range(24))
def test_update_offers_equals_a_full_reload(engine, seed):
    """oe_update_offers (delta upload of the rows that changed between two cycles) against loading the patched table whole."""
    rng = np.random.default_rng(1000 + seed)
    t = random_tables(seed, n_offers=int(rng.integers(2, 400)), with_reserved=False)
    t.normalise()
    n = t.n_offers
    k = int(rng.integers(1, max(2, n // 3)))
    rows = rng.choice(n, size=k, replace=False).astype(np.int64)     # unordered on purpose
    src = rng.integers(0, n, size=k)                                 # what the re-offered agents look like now
    want = _patched(t, rows, src)
    orc = Oracle().load(want)
    engine.load(t)
    engine.evaluate_cycle()                                          # a cycle on the old table, then the delta
    engine.update_offers(rows, t, value_rows=src)
    engine.tables = want
    assert np.array_equal(engine.feasibility()[0], orc.feasibility()[0])
    assert engine.evaluate_cycle()[0].tobytes() == orc.evaluate_cycle()[0].tobytes()
    orc.close()


@pytest.mark.parametrize("flags", [0, abi.OE_CFG_ASYNC_LOADS], ids=["sync", "async"])
@pytest.mark.parametrize("rules,novel", [("c5", 0), ("c5", 300), ("c5-mount", 300), ("c5", 5000)])
def test_update_offers_keeps_the_range_list_dictionary(flags, rules, novel):
    """A table loaded in its packed form names each offer's `ports` ranges by a list id; oe_update_offers brings full range lists.
    The replaced rows get their ids back from the resident dictionary (k_reintern_tuples), lists nobody had are appended, and a
    dictionary that runs full (5000 new lists) falls back to per-offer port tests.  Same answers as loading the patched table."""
    import copy
    n = 7000
    a = synth.make_cluster(n, 500, 8, 3, 4, 250, rules, seed=5).tables
    b = synth.make_cluster(n, 500, 8, 3, 4, 250, rules, seed=6).tables
    a.normalise(); b.normalise()
    rng = np.random.default_rng(novel + 1)
    rows = np.sort(rng.choice(n, size=max(novel, 600), replace=False)).astype(np.int64)
    b.port_range = b.port_range.copy()
    for k, r in enumerate(rows[:novel]):           # range lists that no offer of the resident table carries
        b.port_range[0, 0, r] = (1 + k % 900) | ((1 + k % 900 + k // 900) << 16)
    want = copy.copy(a)
    want._keep = []
    for name in abi.CycleTables.OFFER_COLUMNS:
        col = getattr(a, name).copy()
        col[..., rows] = getattr(b, name)[..., rows]
        setattr(want, name, col)
    ref = Engine(device=0)
    eng = Engine(device=0, flags=flags)
    eng.tables = a
    eng.load_programs(a); eng.load_tasks(a); eng.load_offers_packed(abi.PackedOffers(a)); eng.load_requirements(a)
    assert eng.evaluate_cycle()[0].tobytes() == ref.load(a).evaluate_cycle()[0].tobytes()
    eng.update_offers(rows, b)
    eng.tables = want
    eng.load_requirements(a)
    ref.load(want)
    assert eng.evaluate_cycle()[0].tobytes() == ref.evaluate_cycle()[0].tobytes()
    assert np.array_equal(eng.class_feasibility()[0], ref.class_feasibility()[0])
    eng.update_offers(rows, a)                     # the old rows again: lists the dictionary has had from the start
    eng.tables = a
    eng.load_requirements(a)
    ref.load(a)
    assert eng.evaluate_cycle()[0].tobytes() == ref.evaluate_cycle()[0].tobytes()
    assert np.array_equal(eng.feasibility()[0], ref.feasibility()[0])
    eng.close(); ref.close()


def test_update_offers_rejects_what_it_cannot_patch(engine):
    t = random_tables(7, n_offers=60, with_reserved=False)
    engine.load(t)
    with pytest.raises(abi.EngineError):
        engine.update_offers(np.array([60], np.int64), t, value_rows=[0])          # not a row of the resident table
    bad = _patched(t, [3], [4])
    bad.hostname_id = bad.hostname_id.copy()
    bad.hostname_id[3] = 10 ** 6
    with pytest.raises(abi.EngineError):
        engine.update_offers(np.array([3], np.int64), bad)                          # ids are validated like those of a load
    with pytest.raises(abi.EngineError):
        engine.evaluate_cycle()                                                     # ... and the resident table is not trusted any more
    r = random_tables(8, n_offers=60, with_reserved=True)
    engine.load(r)
    if r.reserved_begin is not None and len(r.reserved_id):
        with pytest.raises(abi.EngineError):
            engine.update_offers(np.array([1], np.int64), r)                        # a reserved side table cannot be patched in place


def test_empty_inputs(engine):
    t = random_tables(5, n_offers=40, n_reqs=6, n_tasks=0)
    _compare_all(engine, t)
    t = random_tables(6, n_offers=1, n_reqs=1, n_tasks=3)
    _compare_all(engine, t)


def test_c4_shape_small(engine):
    cl = synth.make_cluster(2000, 300, 8, 3, 4, 0, "c4")
    _compare_all(engine, cl.tables)


def test_c5_shape_small_with_deployed_tasks(engine):
    cl = synth.make_cluster(4000, 320, 32, 8, 6, 640, "c5")
    _compare_all(engine, cl.tables)


def test_c4_multi_cycle_replay(engine):
    """Config 4's replay: 500 steps per cycle (scaled down), tallies updated between cycles."""
    cl = synth.make_cluster(1500, 400, 8, 3, 4, 0, "c4")
    t = cl.tables
    orc = Oracle()
    per = 100
    import copy
    for cyc in range(4):
        sub = copy.copy(t)
        sl = slice(cyc * per, (cyc + 1) * per)
        sub.class_id, sub.r_pod_type_id, sub.r_pod_index = t.class_id[sl].copy(), t.r_pod_type_id[sl].copy(), t.r_pod_index[sl].copy()
        sub.launch_begin = (t.launch_begin[sl.start: sl.stop + 1] - t.launch_begin[sl.start]).astype(np.int32)
        sub.launch_name_id = t.launch_name_id[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.launch_filter_bits = t.launch_filter_bits[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.normalise()
        if cyc == 0:
            orc.load(sub)
            engine.load(sub)
        else:
            for h in (orc, engine):
                h.tables = sub
                h.load_requirements(sub)
        d_cpu, _ = orc.evaluate_cycle()
        d_gpu, _ = engine.evaluate_cycle()
        assert d_gpu.tobytes() == d_cpu.tobytes(), f"cycle {cyc}"
        orc.apply_launches()
        engine.apply_launches()
    tc, tg = orc.export_tasks(), engine.export_tasks()
    assert len(tc["name_id"]) == len(tg["name_id"]) > 0
    for k in tc:
        assert np.array_equal(tc[k], tg[k])


def test_c4_full_size_multi_cycle_replay(engine):
    """BASELINE config 4's replay at full size: 10 000 offers, 500 steps per cycle, tallies updated between cycles
    (oe_apply_launches); every cycle's decisions against the oracle."""
    import copy
    cl = synth.config_c4()
    t = cl.tables
    orc = Oracle()
    per = 500
    for cyc in range(6):
        sub = copy.copy(t)
        sl = slice(cyc * per, (cyc + 1) * per)
        sub.class_id, sub.r_pod_type_id, sub.r_pod_index = t.class_id[sl].copy(), t.r_pod_type_id[sl].copy(), t.r_pod_index[sl].copy()
        sub.launch_begin = (t.launch_begin[sl.start: sl.stop + 1] - t.launch_begin[sl.start]).astype(np.int32)
        sub.launch_name_id = t.launch_name_id[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.launch_filter_bits = t.launch_filter_bits[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.normalise()
        if cyc == 0:
            orc.load(sub)
            engine.load(sub)
        else:
            for h in (orc, engine):
                h.tables = sub
                h.load_requirements(sub)
        d_cpu, _ = orc.evaluate_cycle()
        d_gpu, _ = engine.evaluate_cycle()
        assert d_gpu.tobytes() == d_cpu.tobytes(), f"cycle {cyc}"
        assert (d_gpu["offer_row"] >= 0).sum() > 0
        orc.apply_launches()
        engine.apply_launches()
    tc, tg = orc.export_tasks(), engine.export_tasks()
    assert len(tc["name_id"]) == len(tg["name_id"]) > 2000
    for k in tc:
        assert np.array_equal(tc[k], tg[k])
    orc.close()


def test_class_rows_expand_to_matrix(engine):
    cl = synth.make_cluster(5000, 500, 8, 3, 4, 80, "c5")
    engine.load(cl.tables)
    m, _ = engine.feasibility()
    f, _ = engine.class_feasibility()
    assert f.shape[0] == engine.factored_rows()
    # fresh instances (no deployed twin) copy their class row
    for j in range(0, cl.tables.n_reqs, 37):
        assert np.array_equal(m[j], f[cl.tables.class_id[j]])


def test_c4_full_size_sampled_rows_and_properties(engine):
    """Config 4 at full size: sampled matrix rows against the oracle, plus size-independent properties."""
    cl = synth.config_c4()
    t = cl.tables
    engine.load(t)
    m, _ = engine.feasibility()
    N = t.n_offers
    orc = Oracle().load(t)
    rng = np.random.default_rng(1)
    for j in rng.integers(0, t.n_reqs, 12):
        row = orc.feasibility_rows(int(j), int(j) + 1)
        assert np.array_equal(unpack_bits(m[j:j + 1], N), unpack_bits(row, N)), f"row {j}"
    dec, st = engine.evaluate_cycle()
    _check_commit_properties(t, m, dec)
    d_cpu, _ = orc.evaluate_cycle()
    assert dec.tobytes() == d_cpu.tobytes()


def _check_commit_properties(t, m, dec):
    """Every winner is feasible, no offer is taken twice, and the choice is first-fit given earlier winners."""
    N = t.n_offers
    bits = unpack_bits(m, N)
    used = np.zeros(N, bool)
    for j in range(t.n_reqs):
        row = bits[j] & ~used
        first = int(np.argmax(row)) if row.any() else -1
        assert dec["offer_row"][j] == first, f"requirement {j}: got {dec['offer_row'][j]}, first fit is {first}"
        if first >= 0:
            used[first] = True


def test_c5_tenth_scale_properties(engine):
    cl = synth.config_c5(0.1)
    t = cl.tables
    engine.load(t)
    m, st = engine.feasibility()
    assert st["n_pairs"] == t.n_offers * t.n_reqs
    dec, _ = engine.evaluate_cycle()
    # linear-time replay of the first-fit property using per-class rows
    f, _ = engine.class_feasibility()
    N = t.n_offers
    fb = unpack_bits(f, N)
    for j in range(0, t.n_reqs, 997):
        assert np.array_equal(unpack_bits(m[j:j + 1], N)[0], fb[t.class_id[j]])
    used = np.zeros(N, bool)
    cursors = np.zeros(fb.shape[0], np.int64)
    order = [np.flatnonzero(fb[c]) for c in range(fb.shape[0])]
    for j in range(t.n_reqs):
        c = t.class_id[j]
        cand = order[c]
        k = cursors[c]
        while k < len(cand) and used[cand[k]]:
            k += 1
        cursors[c] = k
        want = int(cand[k]) if k < len(cand) else -1
        assert dec["offer_row"][j] == want
        if want >= 0:
            used[want] = True
    # determinism: a second run gives the same bytes
    dec2, _ = engine.evaluate_cycle()
    assert dec.tobytes() == dec2.tobytes()


def _factored_row_of(t):
    """Bit row of every requirement, as the engine numbers them: its class, or — when the pod is already deployed (a task of
    the same type and index exists) and the class reads tallies — a row of its own behind the classes, in plan order."""
    twins = set(zip(t.t_pod_type_id.tolist(), t.t_pod_index.tolist()))
    is_var = np.fromiter(((a, b) in twins for a, b in zip(t.r_pod_type_id.tolist(), t.r_pod_index.tolist())), bool, t.n_reqs)
    row = t.class_id.astype(np.int64).copy()
    row[is_var] = t.n_classes + np.arange(int(is_var.sum()))
    return row, is_var


def _full_size_against_cached_oracle(engine, cl, variant_rows=8, shards=0, arriving=True):
    """A workload at its full size against the oracle's cached mode (pinned to the literal one by tests/test_oracle_modes.py):
    EVERY bit of every class row, sampled rows of re-deployed pods, and all decisions byte for byte (offer, dynamic ports,
    MOUNT slots, number of fresh reservation ids)."""
    from dcos_commons_b200 import sharded
    t = cl.tables
    N, M = t.n_offers, t.n_reqs
    orc = Oracle(cached=True, threads=os.cpu_count() or 1).load(t)
    engine.load(t)
    f, st = engine.class_feasibility()
    assert st["n_pairs"] == N * M
    _, is_var = _factored_row_of(t)
    row_of, n_rows = engine.requirement_rows()     # re-deployed requirements with identical rule verdicts share a bit row
    assert engine.factored_rows() == n_rows <= t.n_classes + int(is_var.sum()) and np.array_equal(row_of[~is_var], t.class_id[~is_var])
    assert f.shape[0] == n_rows
    words = (N + 31) // 32
    fresh = np.flatnonzero(~is_var)
    for c in range(t.n_classes):
        js = fresh[t.class_id[fresh] == c]
        if len(js) == 0:
            continue
        want = orc.feasibility_row(int(js[0]))
        assert np.array_equal(f[c, :words], want[:words]), f"class row {c}"
        assert not f[c, words:].any()
    rng = np.random.default_rng(9)
    var_js = np.flatnonzero(is_var)
    for j in (rng.choice(var_js, min(variant_rows, len(var_js)), replace=False) if len(var_js) else []):
        want = orc.feasibility_row(int(j))
        assert np.array_equal(f[row_of[j], :words], want[:words]), f"row of re-deployed requirement {j}"
    d_cpu, s_cpu = orc.evaluate_cycle()
    d_gpu, s_gpu = engine.evaluate_cycle()
    assert np.array_equal(d_gpu["offer_row"], d_cpu["offer_row"])
    assert d_gpu.tobytes() == d_cpu.tobytes()
    assert s_gpu["n_placed"] == s_cpu["n_placed"] > 0
    assert engine.evaluate_cycle()[0].tobytes() == d_cpu.tobytes()      # run to run
    orc.close()
    if arriving:   # the table evaluated while it arrives in chunks, committed in two passes — plain and packed wire form
        eng2 = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
        eng2.load(t)
        assert eng2.evaluate_cycle()[0].tobytes() == d_cpu.tobytes()
        _load_packed(eng2, t)
        assert np.array_equal(eng2.class_feasibility()[0], f)
        _load_packed(eng2, t)
        assert eng2.evaluate_cycle()[0].tobytes() == d_cpu.tobytes()
        eng2.close()
    if shards:     # shards on this device, class rows exchanged through the peer regions
        shard_rows, bounds = sharded.shard_plan(N, shards)
        engs = [Engine(device=0, rank=r, world=shards) for r in range(shards)]
        for e, (lo, hi) in zip(engs, bounds):
            e.load(t, lo, hi, shard_rows)
        sharded.attach_local(engs, engs[0].factored_rows())
        for e in engs:
            e.class_feasibility(fetch=False)
        for e in engs:
            assert np.array_equal(e.evaluate_cycle()[0]["offer_row"], d_cpu["offer_row"])
        for e in engs:
            e.close()
    return d_cpu


def test_c5_full_size_every_class_row_and_every_decision(engine):
    """BASELINE config 5 at full size (1 M offers x 100 k requirements, 100 k deployed tasks, 10^11 pairs): all 32 class rows
    x 1 M offers and all 100 k decisions against the oracle; the same from the chunked / two-pass cycle and from eight shards."""
    _full_size_against_cached_oracle(engine, synth.config_c5(1.0), shards=8)


@pytest.mark.parametrize("which", ["hostname", "mount", "one-region", "redeploy"])
def test_c5_variants_full_size(engine, which):
    """The C5-sized workloads that leave the headline's fast paths (bench.py --workload c5-<which>)."""
    _full_size_against_cached_oracle(engine, synth.config_c5_variant(which, 1.0), arriving=which != "redeploy")


@pytest.mark.parametrize("which", ["hostname", "mount", "one-region", "redeploy"])
def test_c5_variants_small_against_the_literal_oracle(engine, which):
    cl = synth.config_c5_variant(which, 0.004)
    _compare_all(engine, cl.tables)


def _load_packed(eng, t, packed=None, **kw):
    eng.tables = t
    eng.load_programs(t)
    eng.load_tasks(t)
    eng.load_offers_packed(packed or abi.PackedOffers(t), **kw)
    eng.load_requirements(t)
    return eng


@pytest.mark.parametrize("flags", [0, abi.OE_CFG_ASYNC_LOADS], ids=["sync", "async"])
def test_packed_offer_table_gives_what_the_plain_table_gives(engine, flags):
    """oe_load_offers_packed: narrow ids, fp32 scalars where exact, `ports` as ids of distinct range lists — widened on the
    device; matrix, failed-stage masks and decisions must be those of the plain load and of the oracle."""
    eng = Engine(device=0, flags=flags)
    for seed in list(range(40)) + list(range(2000, 2020)):
        t = random_tables(seed, new_pods_only=seed >= 2000)
        if seed % 3 == 0:          # scalars that do not survive fp32: that column must travel as fp64
            t.scalar = t.scalar + 1e-9
        orc = Oracle().load(t)
        packed = abi.PackedOffers(t)
        assert not (packed.scalar_is_f32 and seed % 3 == 0 and t.scalar.any())
        _load_packed(eng, t, packed)
        m_cpu, _ = orc.feasibility()
        m_gpu, _ = eng.feasibility()
        assert np.array_equal(unpack_bits(m_gpu, t.n_offers), unpack_bits(m_cpu, t.n_offers)), seed
        assert eng.evaluate_cycle()[0].tobytes() == orc.evaluate_cycle()[0].tobytes(), seed
        assert np.array_equal(eng.stage_failures(), orc.stage_failures()[0]), seed
        orc.close()
    # the C5 shape: few distinct range lists, so the port stages run once per list (k_port_tables) instead of once per offer
    for cl in (synth.make_cluster(6000, 500, 8, 3, 4, 160, "c5"), synth.config_c5_variant("mount", 0.004), synth.config_c4(0.3)):
        t = cl.tables
        packed = abi.PackedOffers(t)
        assert packed.port_tuple_bytes == 2 and len(packed.cols["port_tuples"]) < 200
        engine.load(t)
        want_f, _ = engine.class_feasibility()
        want_d, _ = engine.evaluate_cycle()
        _load_packed(eng, t, packed)
        got_f, _ = eng.class_feasibility()
        assert np.array_equal(got_f, want_f)
        assert eng.evaluate_cycle()[0].tobytes() == want_d.tobytes()
        # rows replaced in place afterwards: their range lists change, the per-list lookup must not be used any more
        rows = np.arange(0, t.n_offers, 7)
        t2 = synth.make_cluster(t.n_offers, 10, 8, 3, 4, 0, "c5", seed=99).tables if cl.name != "c5-mount-1M-offers-x-100k-instances" else None
        if t2 is not None and t2.port_range.shape == t.port_range.shape and t2.mount_size.shape == t.mount_size.shape:
            t2.port_range = t2.port_range.copy()
            t2.port_range[0, 0, rows] = (1025) | (1030 << 16)
            eng.update_offers(rows, t2)
            engine.update_offers(rows, t2)
            eng.load_requirements(t); engine.load_requirements(t)
            assert np.array_equal(eng.class_feasibility()[0], engine.class_feasibility()[0])
    # a list id outside the dictionary is reported, not followed
    t = synth.make_cluster(3000, 20, 4, 2, 2, 0, "c5").tables
    packed = abi.PackedOffers(t)
    packed.cols["port_tuple_id"] = packed.cols["port_tuple_id"].copy()
    packed.cols["port_tuple_id"][0, 17] = 60000
    with pytest.raises(abi.EngineError, match="range list"):
        _load_packed(eng, t, packed)
        eng.evaluate_cycle()
    eng.close()


def test_invalid_tables_are_rejected_not_faulted(engine):
    t = random_tables(3, n_offers=50, n_reqs=5, n_tasks=10)
    t.hostname_id = t.hostname_id.copy()
    t.hostname_id[7] = 10 ** 6
    with pytest.raises(abi.EngineError):
        engine.load(t)
    t2 = random_tables(3, n_offers=50, n_reqs=5, n_tasks=10)
    t2.class_id = t2.class_id.copy()
    t2.class_id[0] = 99
    with pytest.raises(abi.EngineError):
        engine.load(t2)
    # a reservation named by two stages of one pipeline, or an id outside the dictionary: the device pipeline does not track
    # consumed reservations the way MesosResourcePool.consumeReserved does, so the load refuses what it could not evaluate
    from dcos_commons_b200 import tables as tb
    for ops in ([("scalar_reserved", 0, 1.0, 2), ("scalar_reserved", 1, 1.0, 2)], [("mount_reserved", 1), ("free", 0, 1.0, 1, 0)],
                [("scalar_reserved", 0, 1.0, 8)], [("port_reserved", -1, 1000)]):
        prog = tb.ProgramBuilder()
        prog.add_class([("exec", -1)] + ops + [("nop",)])
        offers = tb.offers_from_rows([dict(scalars={(0, 0): 4.0, (0, 1): 4.0}, reserved=[(2, 0, 1.0)])], S=1, K=4, R=1, D=0, A=1)
        t3 = tb.make_tables(dict(n_res_ids=8), prog, tb.empty_tasks(), offers, [0], [0], [0])
        with pytest.raises(abi.EngineError, match="resource id"):
            engine.load(t3)
    # a shard that is not full although shards follow it
    t4 = random_tables(5, n_offers=3000, n_reqs=5, n_tasks=10, with_reserved=False)
    e2 = Engine(device=0, rank=0, world=2)
    with pytest.raises(abi.EngineError, match="must hold"):
        e2.load(t4, 0, 1000, 2048)
    e2.close()
    # the engine is still usable afterwards
    _compare_all(engine, random_tables(4))
