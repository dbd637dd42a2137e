"""A short program for `ncu --set full`: the C5 tables in their packed wire form, two feasibility passes (K0, K0c, K0p, lean K1,
K1x) and two placement cycles (commit path) on one GPU.  Nothing is timed here."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcos_commons_b200 import Engine, abi, synth  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c5"
    t = (synth.config_c5() if which == "c5" else synth.config_c5_variant(which[3:])).tables
    eng = Engine(device=0)
    eng.tables = t
    eng.load_programs(t); eng.load_tasks(t); eng.load_offers_packed(abi.PackedOffers(t)); eng.load_requirements(t)
    for _ in range(2):
        eng.feasibility(fetch=False)
    for _ in range(2):
        eng.evaluate_cycle()
    eng.close()


if __name__ == "__main__":
    main()
