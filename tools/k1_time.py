"""K1 / K1x device times on the resident C5 tables (and on one shard of `world`): python tools/k1_time.py [world]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcos_commons_b200 import Engine, sharded, synth  # noqa: E402


def main():
    import torch
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    t = synth.config_c5(1.0).tables
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    lo, hi = bounds[0]
    eng = Engine(device=0, rank=0, world=world)
    eng.load(t, lo, hi, shard_rows)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    k0, k1, kx = [], [], []
    for k in range(14):
        flush.zero_(); torch.cuda.synchronize()
        _, st = eng.feasibility(fetch=False)
        if k >= 4:
            k0.append(st["ms_counters"]); k1.append(st["ms_feasibility"]); kx.append(st["ms_expand"])
    m = lambda v: 1e3 * sum(v) / len(v)
    print(f"world {world} offers {hi - lo}: K0 {m(k0):.1f} us  K1 {m(k1):.1f} us (min {1e3 * min(k1):.1f})  K1x {m(kx):.1f} us", flush=True)


if __name__ == "__main__":
    main()
