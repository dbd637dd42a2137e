import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
from dcos_commons_b200 import Engine, unpack_bits
from fuzz_tables import random_tables
from oracle_binding import Oracle
for seed in range(6):
    t = random_tables(seed)
    orc = Oracle().load(t); eng = Engine(device=0).load(t)
    m_cpu, _ = orc.feasibility(); m_gpu, _ = eng.feasibility()
    a, b = unpack_bits(m_gpu, t.n_offers), unpack_bits(m_cpu, t.n_offers)
    bad = np.flatnonzero((a != b).any(axis=1))
    rows, n = eng.requirement_rows()
    print("seed", seed, "C", t.n_classes, "bad reqs", bad.tolist(), "classes", t.class_id[bad].tolist(), "rows", rows.tolist(), n, flush=True)
    f, _ = eng.class_feasibility()
    d_gpu, _ = eng.evaluate_cycle(); d_cpu, _ = orc.evaluate_cycle()
    print("   decisions equal", d_gpu.tobytes() == d_cpu.tobytes())
