"""K1 time against the number of row tiles (OE_K1_TILES) on one shard of the C5 workload: python tools/k1_tiles_probe.py [world]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcos_commons_b200 import Engine, sharded, synth  # noqa: E402


def main():
    import torch
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    t = synth.config_c5(1.0).tables
    shard_rows, bounds = sharded.shard_plan(t.n_offers, world)
    lo, hi = bounds[0]
    eng = Engine(device=0, rank=0, world=world)
    eng.load(t, lo, hi, shard_rows)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for tiles in ("", "1", "2", "3", "4", "5", "6", "8", "11", "16", "32"):
        if tiles:
            os.environ["OE_K1_TILES"] = tiles
        ms = []
        for k in range(12):
            flush.zero_(); torch.cuda.synchronize()
            _, st = eng.class_feasibility(fetch=False)
            if k >= 2:
                ms.append(st["ms_feasibility"])
        print(f"world {world} offers {hi - lo} tiles {tiles or 'default'}: K1 {1e3 * sum(ms) / len(ms):.1f} us (min {1e3 * min(ms):.1f})", flush=True)


if __name__ == "__main__":
    main()
