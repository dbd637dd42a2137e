"""Randomised small tables that touch every stage opcode and rule node of include/oe_abi.h."""
import numpy as np

from dcos_commons_b200 import abi, tables as tb

GRID = [0.0, 0.1, 0.2, 0.3, 0.5, 1.0, 1.5, 2.0, 4.0, 0.1 + 0.2, 32.0, 256.0, 1024.0]


def random_tables(seed: int, n_offers=None, n_reqs=None, n_tasks=None, with_reserved=True, big_dicts=False, new_pods_only=False) -> abi.CycleTables:
    rng = np.random.default_rng(seed)
    easy = rng.random() < 0.6  # short pipelines over generous offers, so that a good share of pairs is feasible
    if new_pods_only:
        # only the stages of a new pod's pipeline (OfferEvaluator.java:411-522): the shape the engine folds into one
        # threshold per pool entry and one port count; long scalar chains on few pool entries stress the folding
        with_reserved = False
    S, K = int(rng.integers(1, 3)), int(rng.integers(4, 6))
    R, D, A = int(rng.integers(1, 7)), int(rng.integers(0, 4)), int(rng.integers(1, 5))
    Q, E = int(rng.integers(0, 3)), int(rng.integers(0, 3))
    N = int(rng.integers(1, 200)) if n_offers is None else n_offers
    M = int(rng.integers(1, 40)) if n_reqs is None else n_reqs
    T = int(rng.integers(0, 60)) if n_tasks is None else n_tasks
    n_agents = max(1, N + 3)
    n_host = int(rng.integers(1, N + 2)) if not big_dicts else N + 50
    n_zones, n_regions = int(rng.integers(0, 5)), int(rng.integers(0, 3))
    if n_zones == 0 or n_regions == 0:
        n_zones = n_regions = 0
    n_attrs, n_rrv = int(rng.integers(1, 13)), int(rng.integers(1, 6))
    n_types, n_filters = int(rng.integers(1, 5)), int(rng.integers(1, 40 if rng.random() < 0.3 else 7))
    n_prof, n_exec, n_roles, n_res = 3, 3, 3, 8
    FW = (n_filters + 31) // 32
    TA = A + int(rng.integers(0, 2))

    def bits(n, p=0.5):
        return rng.random(max(n, 1)) < p

    # ---- rule trees
    def leaf():
        k = int(rng.integers(0, 18))
        f = int(rng.integers(0, n_filters))
        if k == 0: return ("true",)
        if k == 1: return ("false",)
        if k == 2: return ("agent", int(rng.integers(0, n_agents)))
        if k == 3: return ("attr_any", bits(n_attrs))
        if k == 4: return ("host_in", bits(n_host))
        if k == 5: return ("zone_in", bits(n_zones))
        if k == 6: return ("region_in", bits(n_regions))
        if k == 7: return ("local_region", int(rng.integers(0, max(n_regions, 1))), bool(rng.integers(0, 2)))
        if k == 8: return ("maxper_host", f, int(rng.integers(1, 4)))
        if k == 9: return ("maxper_zone", f, int(rng.integers(1, 4)))
        if k == 10: return ("maxper_region", f, int(rng.integers(1, 6)))
        if k == 11: return ("maxper_attr", f, int(rng.integers(1, 4)), bits(n_attrs))
        if k in (12, 13, 14):
            ks = [abi.KS_HOSTNAME, abi.KS_ZONE, abi.KS_REGION][k - 12]
            return ("rr", ks, 0, f, None if rng.random() < 0.4 else int(rng.integers(1, 5)))
        if k == 15:
            if Q == 0: return ("true",)
            return ("rr", abi.KS_RR_ATTR, int(rng.integers(0, Q)), f, None if rng.random() < 0.4 else int(rng.integers(1, 5)))
        if k == 16: return ("type_avoid", int(rng.integers(0, n_types)))
        return ("type_colocate", int(rng.integers(0, n_types)))

    def tree(depth):
        r = rng.random()
        if depth == 0 or r < 0.45:
            return leaf()
        if r < 0.65:
            return ("and", [tree(depth - 1) for _ in range(int(rng.integers(0, 4)))])
        if r < 0.85:
            return ("or", [tree(depth - 1) for _ in range(int(rng.integers(0, 4)))])
        return ("not", tree(depth - 1))

    # ---- classes
    prog = tb.ProgramBuilder()
    n_classes = int(rng.integers(1, 7))
    amt = lambda: float(GRID[int(rng.integers(0, len(GRID)))])
    for c in range(n_classes):
        ops, used_res = [], set()
        exec_id = int(rng.integers(-1, n_exec)) if rng.random() < 0.3 and not new_pods_only else -1
        ops.append(("exec", exec_id))
        if rng.random() < 0.7:
            ops.append(("rule", tree(1 if easy else 3)))
        for _ in range(int(rng.integers(1, 5 if easy else 12))):
            k = int(rng.integers(0, 12))
            slot = int(rng.integers(0, S))
            if new_pods_only:
                k = int(rng.integers(0, 6))
                if k == 5:
                    ops.append(("dynport", None, 0))
                    continue
                if k == 4:
                    slot = 0
            if k <= 3: ops.append(("scalar", int(rng.integers(0, K)), amt(), slot))
            elif k == 4: ops.append(("port", int(rng.integers(1000, 1060)), slot))
            elif k == 5:
                rs = None if rng.random() < 0.5 else [(int(b), int(b + rng.integers(0, 30))) for b in rng.integers(990, 1070, int(rng.integers(1, 3)))]
                ops.append(("dynport", rs, slot))
            elif k == 6 and D > 0:
                mask = int(rng.integers(0, 1 << n_prof)) if rng.random() < 0.5 else 0
                ops.append(("mount", amt(), mask))
            elif k in (7, 8, 9, 10) and with_reserved and rng.random() < (0.3 if easy else 1.0):
                free = [r for r in range(n_res) if r not in used_res]
                if not free:
                    continue
                res = int(free[int(rng.integers(0, len(free)))])
                used_res.add(res)
                if k == 7:
                    flags = abi.OPF_SKIP_IF_RUNNING_EXECUTOR if (E > 0 and rng.random() < 0.3) else 0
                    ops.append(("scalar_reserved", int(rng.integers(0, K)), amt(), res, slot, flags, int(rng.integers(0, n_exec))))
                elif k == 8: ops.append(("port_reserved", res, int(rng.integers(1000, 1060)), slot))
                elif k == 9: ops.append(("mount_reserved", res))
                else:
                    which = rng.random()
                    if which < 0.5: ops.append(("free", int(rng.integers(0, K)), amt(), res, slot))
                    elif which < 0.8: ops.append(("free", abi.OE_KIND_PORTS, 0.0, res, slot, int(rng.integers(1000, 1060))))
                    else: ops.append(("free", abi.OE_KIND_MOUNT, amt(), res, 0, int(rng.integers(0, n_prof + 1))))
            else:
                ops.append(("nop",))
        excl = None if rng.random() < 0.5 else [int(x) for x in rng.integers(1000, 1060, int(rng.integers(0, 5)))]
        prog.add_class(ops, pod_role=int(rng.integers(-1, n_roles)), port_excl=excl)

    # ---- offers
    rows = []
    for i in range(N):
        r = dict(agent=int(rng.integers(0, n_agents)), host=int(rng.integers(0, n_host)))
        r["scalars"] = {(s, k): amt() * (float(rng.integers(1, 8)) if easy else 1.0) for s in range(S) for k in range(K)
                        if rng.random() < (0.97 if easy else 0.85)}
        ports = {}
        for s in range(S):
            cur, lst = int(rng.integers(985, 1010)), []
            for _ in range(int(rng.integers(R if easy else 0, R + 1))):
                b = cur + int(rng.integers(0, 12))
                e = b + int(rng.integers(0, 25))
                lst.append((b, e))
                cur = e + 2 + int(rng.integers(0, 8))  # disjoint and non-adjacent, as IntervalSet would hold them
            ports[s] = lst
        r["ports"] = ports
        r["mounts"] = [(amt(), int(rng.integers(-1, n_prof))) for _ in range(int(rng.integers(0, D + 1)))]
        r["alloc_role"] = int(rng.integers(-1, n_roles)) if rng.random() < 0.4 else -1
        if n_zones and rng.random() < 0.8:
            r["zone"], r["region"] = int(rng.integers(0, n_zones)), int(rng.integers(0, n_regions))
        r["attrs"] = [int(rng.integers(0, n_attrs)) for _ in range(int(rng.integers(0, A + 1)))]
        r["rr"] = [int(rng.integers(-1, n_rrv)) for _ in range(Q)]
        r["execs"] = [int(rng.integers(0, n_exec)) for _ in range(int(rng.integers(0, E + 1)))]
        if with_reserved:
            ids = rng.permutation(n_res)[: int(rng.integers(0, 4))]
            r["reserved"] = [(int(x), int(rng.integers(0, K)), float(rng.integers(1000, 1060)) if rng.random() < 0.3 else amt()) for x in ids]
        rows.append(r)
    offers = tb.offers_from_rows(rows, S=S, K=K, R=R, D=D, A=A, Q=Q, E=E)
    if with_reserved and "reserved_begin" not in offers:
        pass

    # ---- requirements (small index range so that some deployed tasks are "equivalent")
    class_id = rng.integers(0, n_classes, M)
    r_type = rng.integers(0, n_types, M)
    r_index = rng.integers(0, 4, M)

    # ---- tasks
    trows = []
    for t in range(T):
        tr = dict(name=t, type=int(rng.integers(-1, n_types)) if rng.random() < 0.1 else int(rng.integers(0, n_types)),
                  agent=int(rng.integers(0, n_agents)))
        tr["index"] = abi.OE_NO_INDEX if rng.random() < 0.05 else int(rng.integers(0, 5))
        tr["host"] = int(rng.integers(-1, n_host)) if rng.random() < 0.1 else int(rng.integers(0, n_host))
        if n_zones and rng.random() < 0.8:
            tr["zone"], tr["region"] = int(rng.integers(0, n_zones)), int(rng.integers(0, n_regions))
        tr["attrs"] = [int(rng.integers(0, n_attrs)) for _ in range(int(rng.integers(0, TA + 1)))]
        tr["rr"] = [int(rng.integers(-1, n_rrv)) for _ in range(Q)]
        tr["filters"] = {f for f in range(n_filters) if rng.random() < 0.5}
        trows.append(tr)
    tasks = tb.tasks_from_rows(trows, TA, Q, FW) if T else tb.empty_tasks(TA, Q, FW)

    # ---- launch rows: 1-2 stored tasks per requirement; some names collide with deployed tasks (overwrite path)
    lb, ln, lf = [0], [], []
    next_name = T
    for j in range(M):
        for _ in range(int(rng.integers(1, 3))):
            if T and rng.random() < 0.2:
                ln.append(int(rng.integers(0, T)))
            else:
                ln.append(next_name)
                next_name += 1
            w = np.zeros(FW, np.uint32)
            for f in range(n_filters):
                if rng.random() < 0.5:
                    w[f >> 5] |= np.uint32(1 << (f & 31))
            lf.append(w)
        lb.append(len(ln))
    # a task name may be stored by one requirement only within a cycle (state-store key)
    seen, keep_ln = set(), []
    for x in ln:
        while x in seen:
            x = next_name
            next_name += 1
        seen.add(x)
        keep_ln.append(x)
    ds = dict(n_agents=n_agents, n_hostnames=n_host, n_zones=n_zones, n_regions=n_regions, n_attrs=n_attrs, n_rr_values=n_rrv,
              n_pod_types=n_types, n_task_names=next_name, n_filters=n_filters, n_exec_ids=n_exec, n_roles=n_roles, n_res_ids=n_res)
    return tb.make_tables(ds, prog, tasks, offers, class_id, r_type, r_index, launch_begin=np.array(lb, np.int32),
                          launch_name_id=np.array(keep_ln, np.int32), launch_filter_bits=np.array(lf, np.uint32).reshape(len(keep_ln), FW))
