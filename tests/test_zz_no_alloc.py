"""oe_first_fit_static — the entry point for hosts that cannot let the library allocate device memory (several
engines sharing one GPU, where a cudaMalloc/cudaFree would stall everyone's cycle; or a process whose device memory
is owned by a pool or capped by a budget).  It must give the same placement column as oe_evaluate_cycle, call after
call, although every call re-uses one fixed scratch array that still holds the previous call's state.

This file sorts last on purpose: a `pytest -x` run reaches it only after every other test has run.
"""
import numpy as np
import pytest

from dcos_commons_b200 import Engine, abi, sharded, synth
from oracle_binding import Oracle


def test_sizes_beyond_the_fixed_scratch_are_refused_with_a_reason(engine_lib):
    # limits are checked before any device is touched, so this part also runs on a box without a GPU
    with pytest.raises(abi.EngineError) as e:
        abi.first_fit_static(0, 0, 0, (abi.OE_STATIC_MAX_OFFERS + 32) // 32, 1, abi.OE_STATIC_MAX_OFFERS + 1, 0, 1, 0)
    assert e.value.code == -5 and "OE_STATIC_MAX_OFFERS" in str(e.value)
    with pytest.raises(abi.EngineError) as e:
        abi.first_fit_static(0, 0, 0, 4, abi.OE_STATIC_MAX_ROWS + 1, 100, 0, 1, 0)
    assert e.value.code == -5 and "OE_STATIC_MAX_ROWS" in str(e.value)
    with pytest.raises(abi.EngineError) as e:
        abi.first_fit_static(0, 0, 0, 1, 1, 100, 0, 1, 0)   # one word cannot hold 100 offers
    assert e.value.code == -1


@pytest.mark.gpu
def test_same_placements_as_evaluate_cycle_on_consecutive_calls():
    import torch
    eng = Engine(device=0)
    # clusters of different sizes and rule mixes, so that stale scratch from one call would show up in the next;
    # none of their requirements has a deployed twin, hence requirement -> bit row is simply the class id
    clusters = [synth.make_cluster(4000, 320, 32, 8, 6, 640, "c5"),
                synth.make_cluster(9000, 1500, 8, 3, 4, 0, "c4"),
                synth.make_cluster(700, 900, 8, 3, 4, 0, "c4")]     # more requirements than offers: some get -1
    stream = torch.cuda.current_stream().cuda_stream
    for cl in clusters + clusters[:1]:
        t = cl.tables
        eng.load(t)
        want, _ = eng.evaluate_cycle()                     # existing entry point: leaves the class bit rows on the device
        assert eng.factored_rows() == t.n_classes
        # ... and independently of any CUDA code: the same greedy order replayed in numpy over the rows, and the CPU oracle's
        # own cycle on the same tables
        rows, _ = eng.class_feasibility()
        replay = sharded.first_fit(rows, t.class_id, t.n_offers)
        assert np.array_equal(want["offer_row"], replay), cl.name
        orc = Oracle(cached=True).load(t)
        assert np.array_equal(orc.evaluate_cycle()[0]["offer_row"], replay), cl.name
        orc.close()
        bits_ptr, _ = eng.device_buffer(abi.BUF_CLASS_BITS_LOCAL)
        req_row = torch.from_numpy(t.class_id.astype(np.int32)).cuda()
        for _ in range(2):                                 # twice in a row on the same inputs
            out = torch.full((t.n_reqs,), -7, dtype=torch.int32, device="cuda")
            abi.first_fit_static(0, stream, bits_ptr, eng.stride(), t.n_classes, t.n_offers, req_row.data_ptr(), t.n_reqs,
                                 out.data_ptr())
            torch.cuda.synchronize()
            assert np.array_equal(out.cpu().numpy(), want["offer_row"]), cl.name   # integer result: exact
            assert np.array_equal(out.cpu().numpy(), replay), cl.name
    eng.close()
