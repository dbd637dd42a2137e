"""The ordered commit (S/scheduler/plan/PlanScheduler.java:56-67) on feasibility rows chosen by the test.

Every class's pipeline is one placement rule, `AttributeRule` over an attribute that is unique per offer, so the matcher's bit
mask over the attribute dictionary IS the class's feasibility row: the test decides which offers each row lists (dense, sparse,
nested, chained, disjoint) and in which order the requirements of the rows arrive.  Decisions are checked against the greedy
order replayed in numpy on those same rows — exact, at sizes where the candidate windows of the commit kernel move, clog,
get topped up and run out, which the small fuzzed tables hardly reach.
"""
import numpy as np
import pytest

from dcos_commons_b200 import Engine, abi
from dcos_commons_b200 import tables as tb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["resident", "arriving"])
def engine(request):
    """"arriving": OE_CFG_ASYNC_LOADS with the offer table cut into row chunks, so that the cycle commits in two passes — over
    the chunks that have landed, then the requirements still open over the last chunk — and must reach the same decisions."""
    import os
    if request.param == "arriving":
        old = os.environ.get("OE_OFFER_CHUNKS")
        os.environ["OE_OFFER_CHUNKS"] = "3"
        e = Engine(device=0, flags=abi.OE_CFG_ASYNC_LOADS)
    else:
        e = Engine(device=0)
    yield e
    e.close()
    if request.param == "arriving":
        if old is None:
            os.environ.pop("OE_OFFER_CHUNKS", None)
        else:
            os.environ["OE_OFFER_CHUNKS"] = old


def tables_for_rows(rows: np.ndarray, class_seq: np.ndarray) -> abi.CycleTables:
    """rows: [C, N] bool, class_seq: [M] class of every requirement in plan order."""
    C, N = rows.shape
    prog = tb.ProgramBuilder()
    for c in range(C):
        prog.add_class([("exec", -1), ("rule", ("attr_any", rows[c].copy())), ("nop",)])
    ids = np.arange(N, dtype=np.int32)
    offers = dict(scalar=np.zeros((1, 4, N)), scalar_present=np.zeros(N, np.uint32),
                  port_range=np.full((1, 1, N), abi.OE_PORT_EMPTY, np.uint32), mount_size=np.zeros((0, N)),
                  mount_profile=np.zeros((0, N), np.int32), alloc_role=np.full(N, -1, np.int32), agent_id=ids, hostname_id=ids,
                  zone_id=np.full(N, -1, np.int32), region_id=np.full(N, -1, np.int32), attr_id=ids[None, :].copy(),
                  rr_key=np.zeros((0, N), np.int32), exec_id=np.zeros((0, N), np.int32))
    M = len(class_seq)
    ds = dict(n_agents=N, n_hostnames=N, n_attrs=N, n_pod_types=C, n_task_names=0, n_filters=1)
    return tb.make_tables(ds, prog, tb.empty_tasks(TA=1, Q=0, FW=1), offers, class_seq.astype(np.int32), class_seq.astype(np.int32),
                          np.arange(M, dtype=np.int32))


def greedy(rows: np.ndarray, class_seq: np.ndarray) -> np.ndarray:
    """First fit in plan order: the lowest-index offer of the row that no earlier requirement took."""
    C, N = rows.shape
    lists = [np.flatnonzero(rows[c]) for c in range(C)]
    cursor = [0] * C
    used = np.zeros(N, bool)
    out = np.full(len(class_seq), -1, np.int32)
    for j, c in enumerate(class_seq):
        lst, k = lists[c], cursor[c]
        while k < len(lst) and used[lst[k]]:
            k += 1
        cursor[c] = k          # everything before it is taken for good
        if k < len(lst):
            out[j] = lst[k]
            used[lst[k]] = True
    return out


def check(engine, rows, class_seq):
    t = tables_for_rows(rows, class_seq)
    engine.load(t)
    dec, st = engine.evaluate_cycle()
    want = greedy(rows, class_seq)
    got = dec["offer_row"]
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, f"{bad.size} of {len(want)} decisions differ, first at requirement {bad[0]}: got {got[bad[0]]}, first fit is {want[bad[0]]}"
    assert st["n_placed"] == int((want >= 0).sum())
    return want


def test_one_row(engine):
    rng = np.random.default_rng(1)
    rows = (rng.random((1, 9000)) < 0.4)
    want = check(engine, rows, np.zeros(5000, np.int64))       # more requirements than offers the row lists
    assert (want < 0).sum() > 1000


def test_eight_nested_rows_clog_the_window(engine):
    """Rows of falling density over the same offers (the C4 shape): the dense rows strip the window of everything the
    sparse rows list, and the entries only the sparse rows list stay behind."""
    rng = np.random.default_rng(2)
    N = 30000
    u = rng.random(N)
    rows = np.stack([u < d for d in (0.9, 0.7, 0.5, 0.35, 0.2, 0.1, 0.05, 0.02)])
    seq = rng.integers(0, 8, 6000)
    check(engine, rows, seq)
    check(engine, rows, np.sort(seq))                 # all of row 0 first, ...
    check(engine, rows, np.sort(seq)[::-1].copy())    # ... and the sparsest first


def test_sparse_row_inside_a_dense_union(engine):
    """One row lists 1 offer in 200 of what the other lists, and asks as often: almost all of its picks lie beyond the window."""
    rng = np.random.default_rng(3)
    N = 40000
    dense = rng.random(N) < 0.8
    sparse = dense & (rng.random(N) < 0.005)
    rows = np.stack([dense, sparse])
    seq = rng.integers(0, 2, 3000)
    want = check(engine, rows, seq)
    assert (want[seq == 1] < 0).sum() > 0            # the sparse row runs out, the dense one does not
    assert (want[seq == 0] < 0).sum() == 0


def test_components_and_chains(engine):
    """32 rows: four disjoint groups of offers; inside a group the rows overlap as a chain (A-B, B-C, ...), not all pairwise."""
    rng = np.random.default_rng(4)
    N, C = 32768, 32
    rows = np.zeros((C, N), bool)
    for c in range(C):
        g, k = divmod(c, 8)                          # group g owns offers [g*8192, (g+1)*8192); row k a band of it
        lo = g * 8192 + k * 900
        band = np.zeros(N, bool)
        band[lo: lo + 1500] = True                   # neighbours share 600 offers
        rows[c] = band & (rng.random(N) < 0.7)
    seq = rng.integers(0, C, 40000)                  # more than the groups hold: the late requirements find nothing
    want = check(engine, rows, seq)
    assert (want >= 0).sum() > 8000 and (want < 0).sum() > 0


def test_more_rows_than_the_union_path_takes(engine):
    """40 rows: the commit falls back to per-row lists (k_commit); same answers."""
    rng = np.random.default_rng(5)
    N, C = 12000, 40
    rows = rng.random((C, N)) < rng.uniform(0.02, 0.5, (C, 1))
    check(engine, rows, rng.integers(0, C, 5000))


def test_tiny_lists_and_empty_rows(engine):
    rng = np.random.default_rng(6)
    N, C = 5000, 6
    rows = np.zeros((C, N), bool)
    rows[0, [7, 4000]] = True                        # two offers
    rows[1, 7] = True                                # one, shared with row 0
    rows[3, :70] = True                              # a little more than one window
    rows[4] = rng.random(N) < 0.3                    # rows 2 and 5 list nothing
    seq = rng.integers(0, C, 400)
    check(engine, rows, seq)
    check(engine, rows, np.array([1, 0, 0, 0, 1, 3] * 20))


@pytest.mark.parametrize("seed", range(6))
def test_random_shapes(engine, seed):
    rng = np.random.default_rng(100 + seed)
    C = int(rng.integers(2, 33))
    N = int(rng.integers(200, 20000))
    base = rng.random(N)
    rows = np.stack([(base < rng.uniform(0.05, 1.0)) & (rng.random(N) < rng.uniform(0.1, 1.0)) for _ in range(C)])
    check(engine, rows, rng.integers(0, C, int(rng.integers(1, 6000))))
