"""One engine, the C5 tables in pinned memory, a few end-to-end cycles with OE_CYCLE_TRACE=1: where a cycle's wall time goes.
   python tools/cycle_probe.py [matrix|decisions] [chunks|auto] [offers-first|tasks-first|small-first]
   OE_PROBE_WORKLOAD=c5|hostname|mount|redeploy|one-region picks the tables, OE_PROBE_WIRE=plain the unpacked offer table."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from dcos_commons_b200 import Engine, abi, synth  # noqa: E402


def main():
    import torch
    mode = sys.argv[1] if len(sys.argv) > 1 else "matrix"
    if len(sys.argv) > 2 and sys.argv[2] != "auto":
        os.environ["OE_OFFER_CHUNKS"] = sys.argv[2]
    order = sys.argv[3] if len(sys.argv) > 3 else "offers-first"
    which = os.environ.get("OE_PROBE_WORKLOAD", "c5")
    t = (synth.config_c5(1.0) if which == "c5" else synth.config_c5_variant(which)).tables
    t.normalise()
    pt = bench.pin_tables(t)
    ppk = abi.PackedOffers(t).pin() if os.environ.get("OE_PROBE_WIRE", "packed") == "packed" else None
    out = torch.empty(t.n_reqs * abi.DECISION_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    flags = abi.OE_CFG_ASYNC_LOADS | (abi.OE_CFG_MATERIALIZE_MATRIX if mode == "matrix" else 0)
    eng = Engine(device=0, flags=flags)
    for k in range(7):
        if k == 4:
            os.environ["OE_CYCLE_TRACE"] = "1"
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.tables = pt
        eng.load_programs(pt)
        if order in ("tasks-first", "small-first"):
            eng.load_tasks(pt)
        if order == "small-first":          # every small table in front of the offers on the bus
            eng.load_requirements(pt)
        t1 = time.perf_counter()
        eng.load_offers_packed(ppk) if ppk is not None else eng.load_offers(pt)
        t2 = time.perf_counter()
        if order == "offers-first":
            eng.load_tasks(pt)
        if order != "small-first":
            eng.load_requirements(pt)
        t3 = time.perf_counter()
        st = eng.evaluate_cycle_into(out.data_ptr())
        t4 = time.perf_counter()
        if k >= 4:
            print(f"{which} {mode} {order} chunks={os.environ.get('OE_OFFER_CHUNKS', 'auto')}: cycle {1e3 * (t4 - t0):.3f} ms = progs+tasks {1e3 * (t1 - t0):.3f} + offers {1e3 * (t2 - t1):.3f} + reqs {1e3 * (t3 - t2):.3f} "
                  f"+ evaluate {1e3 * (t4 - t3):.3f}; device " + " ".join(f"{k2[3:]}={v:.3f}" for k2, v in st.items() if k2.startswith("ms_")), flush=True)
    os.environ.pop("OE_CYCLE_TRACE", None)
    eng.close()


if __name__ == "__main__":
    main()
