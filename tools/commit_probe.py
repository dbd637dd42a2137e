"""Times one oe_evaluate_cycle per workload and prints the engine's own per-phase device times (debugging aid for the commit kernel)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from dcos_commons_b200 import abi, synth


def main():
    eng = abi.Engine(0)
    for name in sys.argv[1:] or ["c5", "c4"]:
        t = (synth.config_c5() if name == "c5" else synth.config_c4() if name == "c4" else synth.config_c5_variant(name[3:])).tables
        eng.load(t)
        out = []
        for _ in range(4):
            dec, st = eng.evaluate_cycle()
            out.append(round(st["ms_commit"], 4))
        print(json.dumps({"workload": name, "ms_commit": out, "placed": int(st["n_placed"])}))


if __name__ == "__main__":
    main()
