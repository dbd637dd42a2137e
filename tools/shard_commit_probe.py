"""Eight shards of the C5 table as eight engine handles on ONE GPU (peers attached): what the commit path of a sharded cycle
costs on the device when nobody has to wait for anybody (OE_COMMIT_TRACE=1 adds the replay kernel's own counters)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcos_commons_b200 import Engine, abi, sharded, synth  # noqa: E402


def main():
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    t = synth.config_c5(1.0).tables
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    engs = [Engine(device=0, rank=r, world=world) for r in range(world)]
    for e, (lo, hi) in zip(engs, bounds):
        e.load(t, lo, hi, shard_rows)
    sharded.attach_local(engs, 64)
    for k in range(4):
        for e in engs:
            e.class_feasibility(fetch=False)
        for r, e in enumerate(engs):
            _, st = e.evaluate_cycle()
            if r in (0, world - 1):
                print(f"pass {k} rank {r}: " + " ".join(f"{a[3:]}={v:.3f}" for a, v in st.items() if a.startswith("ms_")), flush=True)
    single = Engine(device=0).load(t)
    for k in range(3):
        _, st = single.evaluate_cycle()
    print("unsharded:  " + " ".join(f"{a[3:]}={v:.3f}" for a, v in st.items() if a.startswith("ms_")))


if __name__ == "__main__":
    main()
