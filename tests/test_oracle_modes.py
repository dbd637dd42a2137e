"""The oracle's cached mode against its literal mode (oracle/oe_oracle.cpp): same matrix, same failed-stage masks, same
decisions — first cycle and, after the launches were recorded, second cycle — on every fuzz table and on config C4.
The literal mode follows the reference loop by loop (every rule re-walks every task for every pair); the cached mode
hoists those walks out of the loop over offers and is what the full-size C5 parity tests compare the engine with."""
import copy
import os

import numpy as np
import pytest

from dcos_commons_b200 import synth, unpack_bits
from fuzz_tables import random_tables
from oracle_binding import Oracle


def _same(t, two_cycles=True, masks=True):
    lit, fast = Oracle().load(t), Oracle(cached=True, threads=2).load(t)
    if masks:
        f_lit, b_lit = lit.stage_failures()
        f_fast, b_fast = fast.stage_failures()
        assert np.array_equal(f_lit, f_fast)
        assert np.array_equal(b_lit, b_fast)
    else:
        assert np.array_equal(lit.feasibility()[0], fast.feasibility()[0])
    d_lit, s_lit = lit.evaluate_cycle()
    d_fast, s_fast = fast.evaluate_cycle()
    assert d_lit.tobytes() == d_fast.tobytes()
    assert s_lit["n_placed"] == s_fast["n_placed"]
    if two_cycles and t.launch_begin is not None:
        lit.apply_launches()
        fast.apply_launches()
        a, b = lit.export_tasks(), fast.export_tasks()
        for k in a:
            assert np.array_equal(a[k], b[k])
        assert np.array_equal(lit.feasibility()[0], fast.feasibility()[0])
        assert lit.evaluate_cycle()[0].tobytes() == fast.evaluate_cycle()[0].tobytes()
    lit.close()
    fast.close()


@pytest.mark.parametrize("seed", range(120))
def test_fuzz(seed):
    _same(random_tables(seed))


@pytest.mark.parametrize("seed", range(1000, 1006))
def test_fuzz_wider(seed):
    _same(random_tables(seed, n_offers=700 + seed % 7, n_reqs=40, n_tasks=150, big_dicts=True), masks=False)


@pytest.mark.parametrize("seed", range(2000, 2030))
def test_fuzz_new_pod_pipelines(seed):
    _same(random_tables(seed, n_offers=300 + seed % 50, new_pods_only=True))


def test_many_redeployed_pods():
    """Requirements whose pod is already deployed (the rules must not count the pod's own tasks): a quarter of them."""
    cl = synth.make_cluster(1200, 240, 8, 3, 4, 480, "c5", redeploy_every=4)
    _same(cl.tables, masks=False)
    cl = synth.make_cluster(900, 200, 8, 3, 4, 400, "c4", redeploy_every=3)
    _same(cl.tables, masks=False)


def test_c4_full_size_two_replayed_cycles():
    """Config C4 at its stated size (10 000 offers), 2 x 500 steps with the tallies updated in between."""
    t = synth.config_c4().tables
    lit, fast = Oracle(), Oracle(cached=True, threads=os.cpu_count() or 1)
    per = 500
    for cyc in range(2):
        sub = copy.copy(t)
        sl = slice(cyc * per, (cyc + 1) * per)
        sub.class_id, sub.r_pod_type_id, sub.r_pod_index = t.class_id[sl].copy(), t.r_pod_type_id[sl].copy(), t.r_pod_index[sl].copy()
        sub.launch_begin = (t.launch_begin[sl.start: sl.stop + 1] - t.launch_begin[sl.start]).astype(np.int32)
        sub.launch_name_id = t.launch_name_id[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.launch_filter_bits = t.launch_filter_bits[t.launch_begin[sl.start]: t.launch_begin[sl.stop]].copy()
        sub.normalise()
        for h in (lit, fast):
            if cyc == 0:
                h.load(sub)
            else:
                h.tables = sub
                h.load_requirements(sub)
        assert lit.evaluate_cycle()[0].tobytes() == fast.evaluate_cycle()[0].tobytes(), f"cycle {cyc}"
        N = t.n_offers
        for j in (0, 123, 499):
            assert np.array_equal(lit.feasibility_row(j), fast.feasibility_row(j)), f"cycle {cyc}, row {j}"
        lit.apply_launches()
        fast.apply_launches()
    lit.close()
    fast.close()
