import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
from dcos_commons_b200 import Engine, unpack_bits, tables as tb
from oracle_binding import Oracle
chains = [[0.1, 0.2], [0.1, 0.1, 0.1], [0.3], [1.0, 0.1, 0.2, 0.7], [0.1] * 7, [256.0, 0.1 + 0.2, 32.0]]
prog = tb.ProgramBuilder()
for ch in chains:
    prog.add_class([("exec", -1)] + [("scalar", 0, a) for a in ch] + [("scalar", 1, 1.0), ("nop",)])
rows = []
for ch in chains:
    exact = 0.0
    for a in reversed(ch):
        exact = exact + a
    for base in (exact, sum(ch), float(np.sum(np.array(ch)))):
        for k in range(-3, 4):
            v = base
            for _ in range(abs(k)):
                v = np.nextafter(v, np.inf if k > 0 else -np.inf)
            rows.append(dict(scalars={(0, 0): float(v), (0, 1): 1.0}))
rows.append(dict(scalars={(0, 0): float("nan"), (0, 1): 1.0}))
rows.append(dict(scalars={(0, 0): float("inf"), (0, 1): 1.0}))
rows.append(dict(scalars={(0, 1): 1.0}))
offers = tb.offers_from_rows(rows, S=1, K=4, R=1, D=0, A=1)
M = len(chains) * 3
class_id = np.arange(M) % len(chains)
t = tb.make_tables(dict(n_agents=len(rows), n_hostnames=len(rows), n_pod_types=1), prog, tb.empty_tasks(), offers, class_id, np.zeros(M, np.int32), np.arange(M))
orc = Oracle().load(t); eng = Engine(device=0).load(t)
f_cpu = unpack_bits(orc.feasibility()[0], t.n_offers)
f_gpu = unpack_bits(eng.class_feasibility()[0], t.n_offers)
for c in range(len(chains)):
    bad = np.flatnonzero(f_gpu[c] != f_cpu[c])
    print("class", c, "bad offers", bad.tolist(), [rows[i]["scalars"].get((0, 0)) for i in bad[:4]], "gpu", f_gpu[c][bad[:4]].tolist())
