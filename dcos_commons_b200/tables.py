"""Assembling encoded tables (include/oe_abi.h) from Python-level descriptions.

Used by the synthetic cluster generators and by tests.  A rule tree is a nested tuple:
  ("and", [children]) ("or", [children]) ("not", child) ("true",) ("false",)
  ("agent", agent_id) ("attr_any", bits) ("host_in", bits) ("zone_in", bits) ("region_in", bits)
  ("local_region", region_id, known)
  ("maxper_host", filter_id, max) ("maxper_zone", filter_id, max) ("maxper_region", filter_id, max)
  ("maxper_attr", filter_id, max, bits)
  ("rr", keyspace, rr_col, filter_id, distinct_or_None)
  ("type_avoid", pod_type_id) ("type_colocate", pod_type_id)
where `bits` is a boolean array over the relevant dictionary (matcher.matches(value) for each value).
A stage op is a tuple:
  ("nop",) ("exec", exec_id_or_-1) ("rule", tree) ("scalar", kind, amount[, slot]) ("port", port[, slot])
  ("dynport", [ (b,e), ... ] or None[, slot]) ("mount", amount, profile_mask)
  ("scalar_reserved", kind, amount, res_id[, slot, flags, exec_id]) ("port_reserved", res_id, port[, slot])
  ("mount_reserved", res_id) ("free", kind, amount, res_id, slot[, aux])   kind OE_KIND_PORTS: aux = port; OE_KIND_MOUNT: aux = profile + 1
"""
from __future__ import annotations

import numpy as np

from . import abi
from .abi import (COUNTER_SET_DTYPE, OE_PORT_EMPTY, RULE_NODE_DTYPE, STAGE_OP_DTYPE, CycleTables)


def pack_range(b, e):
    return (int(b) & 0xFFFF) | ((int(e) & 0xFFFF) << 16)


class ProgramBuilder:
    def __init__(self):
        self.ops, self.op_begin = [], [0]
        self.nodes = []
        self.pod_role = []
        self.excl, self.excl_begin = [], [0]
        self.rs, self.rs_begin = [], [0]
        self.csets, self.tsets = [], []
        self.mwords = []

    def matcher(self, bits) -> int:
        bits = np.asarray(bits, dtype=bool)
        off = len(self.mwords)
        n_words = max(1, (len(bits) + 31) // 32)
        padded = np.zeros(n_words * 32, dtype=np.uint8)
        padded[: len(bits)] = bits
        self.mwords.extend(np.packbits(padded, bitorder="little").view("<u4").tolist())
        return off

    def counter_set(self, keyspace, filter_id, rr_col=0) -> int:
        key = (int(keyspace), int(rr_col), int(filter_id))
        if key not in self.csets:
            self.csets.append(key)
        return self.csets.index(key)

    def type_set(self, pod_type) -> int:
        if pod_type not in self.tsets:
            self.tsets.append(int(pod_type))
        return self.tsets.index(pod_type)

    def range_set(self, ranges) -> int:
        self.rs.extend(pack_range(b, e) for b, e in ranges)
        self.rs_begin.append(len(self.rs))
        return len(self.rs_begin) - 2

    def _emit(self, tree):
        k = tree[0]
        if k in ("and", "or"):
            for ch in tree[1]:
                self._emit(ch)
            self.nodes.append((abi.R_AND if k == "and" else abi.R_OR, len(tree[1]), 0, 0))
        elif k == "not":
            self._emit(tree[1])
            self.nodes.append((abi.R_NOT, 0, 0, 0))
        elif k == "true":
            self.nodes.append((abi.R_TRUE, 0, 0, 0))
        elif k == "false":
            self.nodes.append((abi.R_FALSE, 0, 0, 0))
        elif k == "agent":
            self.nodes.append((abi.R_AGENT_EQ, tree[1], 0, 0))
        elif k in ("attr_any", "host_in", "zone_in", "region_in"):
            op = {"attr_any": abi.R_ATTR_ANY, "host_in": abi.R_HOST_IN, "zone_in": abi.R_ZONE_IN, "region_in": abi.R_REGION_IN}[k]
            self.nodes.append((op, self.matcher(tree[1]), 0, 0))
        elif k == "local_region":
            self.nodes.append((abi.R_LOCAL_REGION, tree[1], 1 if tree[2] else 0, 0))
        elif k in ("maxper_host", "maxper_zone", "maxper_region"):
            op, ks = {"maxper_host": (abi.R_MAXPER_HOST, abi.KS_HOSTNAME), "maxper_zone": (abi.R_MAXPER_ZONE, abi.KS_ZONE),
                      "maxper_region": (abi.R_MAXPER_REGION, abi.KS_REGION)}[k]
            self.nodes.append((op, self.counter_set(ks, tree[1]), tree[2], 0))
        elif k == "maxper_attr":
            self.nodes.append((abi.R_MAXPER_ATTR, self.counter_set(abi.KS_ATTR, tree[1]), tree[2], self.matcher(tree[3])))
        elif k == "rr":
            _, ks, col, f, distinct = tree
            self.nodes.append((abi.R_RR, self.counter_set(ks, f, col), -1 if distinct is None else distinct, 0))
        elif k == "type_avoid":
            self.nodes.append((abi.R_TYPE_AVOID, self.type_set(tree[1]), 0, 0))
        elif k == "type_colocate":
            self.nodes.append((abi.R_TYPE_COLOCATE, self.type_set(tree[1]), 0, 0))
        else:
            raise ValueError(f"unknown rule node {k}")

    def add_class(self, ops, pod_role=-1, port_excl=None) -> int:
        """Appends one class; `port_excl` defaults to the static ports named by the ops."""
        statics = []
        for op in ops:
            k = op[0]
            rec = dict(opcode=0, kind=0, slot=0, flags=0, arg=0, amount=0.0, mask=0, aux=0)
            if k == "nop":
                rec["opcode"] = abi.OP_NOP
            elif k == "exec":
                rec.update(opcode=abi.OP_EXEC_ID, arg=op[1])
            elif k == "rule":
                b = len(self.nodes)
                self._emit(op[1])
                rec.update(opcode=abi.OP_RULE, arg=b, aux=len(self.nodes) - b)
            elif k == "scalar":
                rec.update(opcode=abi.OP_SCALAR_NEW, kind=op[1], amount=op[2], slot=op[3] if len(op) > 3 else 0)
            elif k == "port":
                rec.update(opcode=abi.OP_PORT_STATIC, arg=op[1], slot=op[2] if len(op) > 2 else 0)
                statics.append(op[1])
            elif k == "dynport":
                rs = -1 if op[1] is None else self.range_set(op[1])
                rec.update(opcode=abi.OP_PORT_DYNAMIC, arg=rs, slot=op[2] if len(op) > 2 else 0)
            elif k == "mount":
                rec.update(opcode=abi.OP_VOL_MOUNT, amount=op[1], mask=op[2])
            elif k == "scalar_reserved":
                rec.update(opcode=abi.OP_SCALAR_RESERVED, kind=op[1], amount=op[2], arg=op[3], slot=op[4] if len(op) > 4 else 0,
                           flags=op[5] if len(op) > 5 else 0, aux=op[6] if len(op) > 6 else 0)
            elif k == "port_reserved":
                rec.update(opcode=abi.OP_PORT_RESERVED, arg=op[1], aux=op[2], slot=op[3] if len(op) > 3 else 0)
            elif k == "mount_reserved":
                rec.update(opcode=abi.OP_MOUNT_RESERVED, arg=op[1])
            elif k == "free":
                rec.update(opcode=abi.OP_FREE, kind=op[1], amount=op[2], arg=op[3], slot=op[4], aux=op[5] if len(op) > 5 else 0)
            else:
                raise ValueError(f"unknown stage op {k}")
            self.ops.append(tuple(rec[f] for f in STAGE_OP_DTYPE.names))
        self.op_begin.append(len(self.ops))
        self.pod_role.append(pod_role)
        excl = sorted(set(statics if port_excl is None else port_excl))
        self.excl.extend(excl)
        self.excl_begin.append(len(self.excl))
        return len(self.pod_role) - 1

    def arrays(self):
        cs = np.zeros(len(self.csets), dtype=COUNTER_SET_DTYPE)
        for i, (ks, col, f) in enumerate(self.csets):
            cs[i] = (ks, col, f, 0)
        return dict(
            op_begin=np.array(self.op_begin, np.int32),
            ops=np.array(self.ops, dtype=STAGE_OP_DTYPE) if self.ops else np.zeros(0, STAGE_OP_DTYPE),
            rule_nodes=np.array(self.nodes, dtype=RULE_NODE_DTYPE) if self.nodes else np.zeros(0, RULE_NODE_DTYPE),
            pod_role=np.array(self.pod_role, np.int32),
            port_excl_begin=np.array(self.excl_begin, np.int32),
            port_excl=np.array(self.excl, np.uint16),
            range_set_begin=np.array(self.rs_begin, np.int32),
            range_set=np.array(self.rs, np.uint32),
            counter_sets=cs,
            type_set_pod_type=np.array(self.tsets, np.int32),
            matcher_bits=np.array(self.mwords, np.uint32),
        )


def empty_tasks(TA=1, Q=0, FW=1):
    z = np.zeros(0, np.int32)
    return dict(t_name_id=z, t_pod_type_id=z, t_pod_index=z, t_agent_id=z, t_hostname_id=z, t_zone_id=z, t_region_id=z,
                t_attr_id=np.zeros((TA, 0), np.int32), t_rr_key=np.zeros((Q, 0), np.int32),
                t_filter_bits=np.zeros((FW, 0), np.uint32))


def tasks_from_rows(rows, TA, Q, FW):
    """rows: dicts with name, type, index, agent, host, zone, region, attrs (list), rr (list), filters (set of filter ids)."""
    T = len(rows)
    out = dict(t_name_id=np.array([r["name"] for r in rows], np.int32).reshape(T),
               t_pod_type_id=np.array([r["type"] for r in rows], np.int32).reshape(T),
               t_pod_index=np.array([r.get("index", abi.OE_NO_INDEX) for r in rows], np.int32).reshape(T),
               t_agent_id=np.array([r["agent"] for r in rows], np.int32).reshape(T),
               t_hostname_id=np.array([r.get("host", -1) for r in rows], np.int32).reshape(T),
               t_zone_id=np.array([r.get("zone", -1) for r in rows], np.int32).reshape(T),
               t_region_id=np.array([r.get("region", -1) for r in rows], np.int32).reshape(T))
    attr = np.full((TA, T), -1, np.int32)
    rr = np.full((Q, T), -1, np.int32)
    fb = np.zeros((FW, T), np.uint32)
    for t, r in enumerate(rows):
        for a, v in enumerate(r.get("attrs", [])):
            attr[a, t] = v
        for q, v in enumerate(r.get("rr", [])):
            rr[q, t] = v
        for f in r.get("filters", ()):
            fb[f >> 5, t] |= np.uint32(1 << (f & 31))
    out.update(t_attr_id=attr, t_rr_key=rr, t_filter_bits=fb)
    return out


def offers_from_rows(rows, S=1, K=4, R=4, D=0, A=1, Q=0, E=0):
    """rows: dicts with scalars {(slot,kind): value}, ports {slot: [(b,e),...]}, mounts [(size, profile)],
    alloc_role, agent, host, zone, region, attrs, rr, execs, reserved [(res_id, kind, value)]."""
    N = len(rows)
    scalar = np.zeros((S, K, N))
    present = np.zeros(N, np.uint32)
    port = np.full((S, R, N), OE_PORT_EMPTY, np.uint32)
    msize = np.full((D, N), -np.inf)
    mprof = np.full((D, N), -1, np.int32)
    alloc = np.full(N, -1, np.int32)
    agent = np.zeros(N, np.int32)
    host = np.zeros(N, np.int32)
    zone = np.full(N, -1, np.int32)
    region = np.full(N, -1, np.int32)
    attr = np.full((A, N), -1, np.int32)
    rr = np.full((Q, N), -1, np.int32)
    ex = np.full((E, N), -1, np.int32)
    rbegin, rid, rkind, rval = [0], [], [], []
    for i, r in enumerate(rows):
        for (s, k), v in r.get("scalars", {}).items():
            scalar[s, k, i] = v
            present[i] |= np.uint32(1 << (s * K + k))
        for s, ranges in r.get("ports", {}).items():
            for q, (b, e) in enumerate(ranges):
                port[s, q, i] = pack_range(b, e)
        for d, (size, prof) in enumerate(r.get("mounts", [])):
            msize[d, i], mprof[d, i] = size, prof
        alloc[i] = r.get("alloc_role", -1)
        agent[i] = r.get("agent", i)
        host[i] = r.get("host", i)
        zone[i] = r.get("zone", -1)
        region[i] = r.get("region", -1)
        for a, v in enumerate(r.get("attrs", [])):
            attr[a, i] = v
        for q, v in enumerate(r.get("rr", [])):
            rr[q, i] = v
        for e, v in enumerate(r.get("execs", [])):
            ex[e, i] = v
        for (res, kind, val) in r.get("reserved", []):
            rid.append(res); rkind.append(kind); rval.append(val)
        rbegin.append(len(rid))
    out = dict(scalar=scalar, scalar_present=present, port_range=port, mount_size=msize, mount_profile=mprof, alloc_role=alloc,
               agent_id=agent, hostname_id=host, zone_id=zone, region_id=region, attr_id=attr, rr_key=rr, exec_id=ex)
    if rid:
        out.update(reserved_begin=np.array(rbegin, np.int64), reserved_id=np.array(rid, np.int32),
                   reserved_kind=np.array(rkind, np.int32), reserved_value=np.array(rval, np.float64))
    return out


def make_tables(dict_sizes, prog: ProgramBuilder, tasks: dict, offers: dict, class_id, pod_type_id, pod_index, **extra) -> CycleTables:
    ds = dict(n_agents=1, n_hostnames=1, n_zones=0, n_regions=0, n_attrs=0, n_rr_values=0, n_pod_types=1, n_task_names=0,
              n_filters=1, n_exec_ids=0, n_roles=1, n_res_ids=0)
    ds.update(dict_sizes)
    t = CycleTables(dict_sizes=ds, **prog.arrays(), **tasks, **offers, class_id=np.asarray(class_id, np.int32),
                    r_pod_type_id=np.asarray(pod_type_id, np.int32), r_pod_index=np.asarray(pod_index, np.int32), **extra)
    return t.normalise()
