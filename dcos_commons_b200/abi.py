"""ctypes mirror of include/oe_abi.h plus thin Python wrappers.

Everything here is plumbing for tests, the benchmark and the multi-GPU driver: the product is the C-ABI library
``liboe_b200.so`` (dcos_commons_b200/csrc).  ``Engine`` refuses to work without it — there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ENGINE_LIB = os.path.join(HERE, "liboe_b200.so")

OE_ABI_VERSION = 1
OE_PORT_EMPTY = 0x0000FFFF
OE_NO_INDEX = -(2 ** 31)
OE_KIND_PORTS, OE_KIND_MOUNT = 100, 101
OE_CFG_MATERIALIZE_MATRIX = 1
OE_CFG_ASYNC_LOADS = 2
OE_MAX_DYN_PORTS, OE_MAX_MOUNT_OPS = 8, 4

# opcodes (oe_opcode)
(OP_NOP, OP_EXEC_ID, OP_RULE, OP_SCALAR_NEW, OP_PORT_STATIC, OP_PORT_DYNAMIC, OP_VOL_MOUNT, OP_SCALAR_RESERVED,
 OP_PORT_RESERVED, OP_MOUNT_RESERVED, OP_FREE) = range(11)
OPF_SKIP_IF_RUNNING_EXECUTOR = 1
# rule nodes (oe_rule_op)
(R_TRUE, R_FALSE, R_AND, R_OR, R_NOT, R_AGENT_EQ, R_ATTR_ANY, R_HOST_IN, R_ZONE_IN, R_REGION_IN, R_LOCAL_REGION,
 R_MAXPER_HOST, R_MAXPER_ZONE, R_MAXPER_REGION, R_MAXPER_ATTR, R_RR, R_TYPE_AVOID, R_TYPE_COLOCATE) = range(18)
KS_HOSTNAME, KS_ZONE, KS_REGION, KS_ATTR, KS_RR_ATTR = range(5)
BUF_CLASS_BITS_LOCAL, BUF_CLASS_BITS_GLOBAL, BUF_MATRIX, BUF_DECISIONS = range(4)

STAGE_OP_DTYPE = np.dtype([("opcode", "u1"), ("kind", "u1"), ("slot", "u1"), ("flags", "u1"), ("arg", "<i4"),
                           ("amount", "<f8"), ("mask", "<u8"), ("aux", "<u8")])
RULE_NODE_DTYPE = np.dtype([("op", "<i4"), ("a", "<i4"), ("b", "<i4"), ("c", "<i4")])
COUNTER_SET_DTYPE = np.dtype([("keyspace", "<i4"), ("rr_col", "<i4"), ("filter_id", "<i4"), ("reserved0", "<i4")])
DECISION_DTYPE = np.dtype([("offer_row", "<i4"), ("n_new_reservations", "<i4"), ("dyn_port", "<u2", (OE_MAX_DYN_PORTS,)),
                           ("mount_slot", "i1", (OE_MAX_MOUNT_OPS,)), ("reserved0", "<u4")])
assert STAGE_OP_DTYPE.itemsize == 32 and RULE_NODE_DTYPE.itemsize == 16 and DECISION_DTYPE.itemsize == 32

_p = C.c_void_p


class OeConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("flags", C.c_uint32), ("stream", _p)]


class OeDictSizes(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_agents", "n_hostnames", "n_zones", "n_regions", "n_attrs", "n_rr_values",
                                          "n_pod_types", "n_task_names", "n_filters", "n_exec_ids", "n_roles", "n_res_ids")]


class OeOfferTable(C.Structure):
    _fields_ = [("n_offers", C.c_int64), ("row_offset", C.c_int64), ("n_offers_global", C.c_int64), ("shard_rows", C.c_int64),
                ("n_role_slots", C.c_int32), ("n_scalar_kinds", C.c_int32), ("n_port_ranges", C.c_int32),
                ("n_mount_slots", C.c_int32), ("n_attr_slots", C.c_int32), ("n_rr_cols", C.c_int32),
                ("n_exec_slots", C.c_int32), ("reserved0", C.c_int32),
                ("scalar", _p), ("scalar_present", _p), ("port_range", _p), ("mount_size", _p), ("mount_profile", _p),
                ("alloc_role", _p), ("agent_id", _p), ("hostname_id", _p), ("zone_id", _p), ("region_id", _p),
                ("attr_id", _p), ("rr_key", _p), ("exec_id", _p),
                ("n_reserved", C.c_int64), ("reserved_begin", _p), ("reserved_id", _p), ("reserved_kind", _p),
                ("reserved_value", _p)]


class OeOfferPacking(C.Structure):
    _fields_ = [("scalar_is_f32", C.c_uint8), ("present_bytes", C.c_uint8), ("zone_bytes", C.c_uint8), ("region_bytes", C.c_uint8),
                ("attr_bytes", C.c_uint8), ("rr_bytes", C.c_uint8), ("mount_profile_bytes", C.c_uint8), ("exec_bytes", C.c_uint8),
                ("hostname_is_agent", C.c_uint8), ("alloc_role_is_const", C.c_uint8), ("port_tuple_bytes", C.c_uint8),
                ("reserved0", C.c_uint8 * 5), ("alloc_role_const", C.c_int32), ("n_port_tuples", C.c_int32),
                ("port_tuples", _p), ("port_tuple_id", _p)]


class OeTaskTable(C.Structure):
    _fields_ = [("n_tasks", C.c_int64), ("n_attr_slots", C.c_int32), ("n_rr_cols", C.c_int32),
                ("n_filter_words", C.c_int32), ("reserved0", C.c_int32),
                ("name_id", _p), ("pod_type_id", _p), ("pod_index", _p), ("agent_id", _p), ("hostname_id", _p),
                ("zone_id", _p), ("region_id", _p), ("attr_id", _p), ("rr_key", _p), ("filter_bits", _p)]


class OeProgramTable(C.Structure):
    _fields_ = [("n_classes", C.c_int32), ("n_ops", C.c_int32), ("n_rule_nodes", C.c_int32), ("n_counter_sets", C.c_int32),
                ("n_type_sets", C.c_int32), ("n_range_sets", C.c_int32), ("n_matcher_words", C.c_int64),
                ("op_begin", _p), ("ops", _p), ("rule_nodes", _p), ("pod_role", _p), ("port_excl_begin", _p),
                ("port_excl", _p), ("range_set_begin", _p), ("range_set", _p), ("counter_sets", _p),
                ("type_set_pod_type", _p), ("matcher_bits", _p), ("dict", OeDictSizes)]


class OeReqTable(C.Structure):
    _fields_ = [("n_reqs", C.c_int64), ("class_id", _p), ("pod_type_id", _p), ("pod_index", _p), ("self_begin", _p),
                ("self_rows", _p), ("launch_begin", _p), ("launch_name_id", _p), ("launch_filter_bits", _p)]


class OeStats(C.Structure):
    _fields_ = [("ms_h2d", C.c_float), ("ms_counters", C.c_float), ("ms_feasibility", C.c_float), ("ms_expand", C.c_float),
                ("ms_commit", C.c_float), ("ms_d2h", C.c_float), ("ms_total", C.c_float), ("n_placed", C.c_int32),
                ("n_pairs", C.c_int64), ("kernel_launches", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def _ptr(a: Optional[np.ndarray]):
    if a is None or a.size == 0:
        return None
    assert a.flags["C_CONTIGUOUS"], "table columns must be C-contiguous"
    return a.ctypes.data


def _col(a, dtype, shape=None):
    a = np.ascontiguousarray(a, dtype=dtype)
    if shape is not None:
        a = a.reshape(shape)
    return a


@dataclass
class CycleTables:
    """The four tables of one cycle as numpy arrays (slot-major 2-D columns: shape [slots, rows])."""
    # dictionaries
    dict_sizes: dict
    # programs
    op_begin: np.ndarray
    ops: np.ndarray            # STAGE_OP_DTYPE
    rule_nodes: np.ndarray     # RULE_NODE_DTYPE
    pod_role: np.ndarray
    port_excl_begin: np.ndarray
    port_excl: np.ndarray      # u2
    range_set_begin: np.ndarray
    range_set: np.ndarray      # u4 packed
    counter_sets: np.ndarray   # COUNTER_SET_DTYPE
    type_set_pod_type: np.ndarray
    matcher_bits: np.ndarray   # u4
    # tasks
    t_name_id: np.ndarray
    t_pod_type_id: np.ndarray
    t_pod_index: np.ndarray
    t_agent_id: np.ndarray
    t_hostname_id: np.ndarray
    t_zone_id: np.ndarray
    t_region_id: np.ndarray
    t_attr_id: np.ndarray      # [TA, T]
    t_rr_key: np.ndarray       # [Q, T]
    t_filter_bits: np.ndarray  # [FW, T]
    # offers
    scalar: np.ndarray         # [S, K, N] f8
    scalar_present: np.ndarray # [N] u4
    port_range: np.ndarray     # [S, R, N] u4
    mount_size: np.ndarray     # [D, N] f8
    mount_profile: np.ndarray  # [D, N] i4
    alloc_role: np.ndarray
    agent_id: np.ndarray
    hostname_id: np.ndarray
    zone_id: np.ndarray
    region_id: np.ndarray
    attr_id: np.ndarray        # [A, N]
    rr_key: np.ndarray         # [Q, N]
    exec_id: np.ndarray        # [E, N]
    # requirements
    class_id: np.ndarray
    r_pod_type_id: np.ndarray
    r_pod_index: np.ndarray
    reserved_begin: Optional[np.ndarray] = None  # [N+1] i8
    reserved_id: Optional[np.ndarray] = None
    reserved_kind: Optional[np.ndarray] = None
    reserved_value: Optional[np.ndarray] = None
    self_begin: Optional[np.ndarray] = None
    self_rows: Optional[np.ndarray] = None
    launch_begin: Optional[np.ndarray] = None
    launch_name_id: Optional[np.ndarray] = None
    launch_filter_bits: Optional[np.ndarray] = None  # [n_launch, FW]
    # set by cut_offers(): the offer columns hold only rows [offer_lo, offer_lo + n_offers) of a table of n_offers_global rows
    offer_lo: int = 0
    n_offers_global: Optional[int] = None
    _keep: list = field(default_factory=list, repr=False)

    def normalise(self):
        """Coerce dtypes/contiguity so raw pointers can be handed to C."""
        i4, u4, f8 = np.int32, np.uint32, np.float64
        for n in ("op_begin", "pod_role", "port_excl_begin", "range_set_begin", "type_set_pod_type", "t_name_id",
                  "t_pod_type_id", "t_pod_index", "t_agent_id", "t_hostname_id", "t_zone_id", "t_region_id", "t_attr_id",
                  "t_rr_key", "mount_profile", "alloc_role", "agent_id", "hostname_id", "zone_id", "region_id", "attr_id",
                  "rr_key", "exec_id", "class_id", "r_pod_type_id", "r_pod_index"):
            setattr(self, n, _col(getattr(self, n), i4))
        for n in ("matcher_bits", "range_set", "t_filter_bits", "scalar_present", "port_range"):
            setattr(self, n, _col(getattr(self, n), u4))
        self.port_excl = _col(self.port_excl, np.uint16)
        self.scalar = _col(self.scalar, f8)
        self.mount_size = _col(self.mount_size, f8)
        self.ops = _col(self.ops, STAGE_OP_DTYPE)
        self.rule_nodes = _col(self.rule_nodes, RULE_NODE_DTYPE)
        self.counter_sets = _col(self.counter_sets, COUNTER_SET_DTYPE)
        for n in ("self_begin", "self_rows", "launch_begin", "launch_name_id", "reserved_id", "reserved_kind"):
            if getattr(self, n) is not None:
                setattr(self, n, _col(getattr(self, n), i4))
        if self.launch_filter_bits is not None:
            self.launch_filter_bits = _col(self.launch_filter_bits, u4)
        if self.reserved_begin is not None:
            self.reserved_begin = _col(self.reserved_begin, np.int64)
            self.reserved_value = _col(self.reserved_value, f8)
        return self

    # ---- sizes
    @property
    def n_offers(self): return int(self.agent_id.shape[0])
    @property
    def n_reqs(self): return int(self.class_id.shape[0])
    @property
    def n_tasks(self): return int(self.t_name_id.shape[0])
    @property
    def n_classes(self): return int(self.op_begin.shape[0] - 1)

    OFFER_COLUMNS = ("scalar", "scalar_present", "port_range", "mount_size", "mount_profile", "alloc_role", "agent_id", "hostname_id",
                     "zone_id", "region_id", "attr_id", "rr_key", "exec_id")

    def cut_offers(self, lo: int, hi: int) -> "CycleTables":
        """A copy of the tables whose offer columns hold only rows [lo, hi): what one rank of a sharded run keeps in (pinned)
        host memory.  Programs, tasks and requirements are shared with the original."""
        import copy
        if self.reserved_begin is not None and len(self.reserved_id):
            raise NotImplementedError("sharding a table with a reserved side table")
        c = copy.copy(self)
        c._keep = []
        for name in self.OFFER_COLUMNS:
            setattr(c, name, np.ascontiguousarray(getattr(self, name)[..., lo:hi]))
        c.offer_lo, c.n_offers_global = lo, self.n_offers
        return c

    # ---- C views
    def c_programs(self) -> OeProgramTable:
        t = OeProgramTable()
        t.n_classes = self.n_classes
        t.n_ops = len(self.ops)
        t.n_rule_nodes = len(self.rule_nodes)
        t.n_counter_sets = len(self.counter_sets)
        t.n_type_sets = len(self.type_set_pod_type)
        t.n_range_sets = len(self.range_set_begin) - 1
        t.n_matcher_words = len(self.matcher_bits)
        t.op_begin, t.ops, t.rule_nodes = _ptr(self.op_begin), _ptr(self.ops), _ptr(self.rule_nodes)
        t.pod_role, t.port_excl_begin, t.port_excl = _ptr(self.pod_role), _ptr(self.port_excl_begin), _ptr(self.port_excl)
        t.range_set_begin, t.range_set = _ptr(self.range_set_begin), _ptr(self.range_set)
        t.counter_sets, t.type_set_pod_type, t.matcher_bits = _ptr(self.counter_sets), _ptr(self.type_set_pod_type), _ptr(self.matcher_bits)
        for k, v in self.dict_sizes.items():
            setattr(t.dict, k, int(v))
        return t

    def c_tasks(self) -> OeTaskTable:
        t = OeTaskTable()
        t.n_tasks = self.n_tasks
        t.n_attr_slots, t.n_rr_cols, t.n_filter_words = self.t_attr_id.shape[0], self.t_rr_key.shape[0], self.t_filter_bits.shape[0]
        t.name_id, t.pod_type_id, t.pod_index = _ptr(self.t_name_id), _ptr(self.t_pod_type_id), _ptr(self.t_pod_index)
        t.agent_id, t.hostname_id, t.zone_id, t.region_id = _ptr(self.t_agent_id), _ptr(self.t_hostname_id), _ptr(self.t_zone_id), _ptr(self.t_region_id)
        t.attr_id, t.rr_key, t.filter_bits = _ptr(self.t_attr_id), _ptr(self.t_rr_key), _ptr(self.t_filter_bits)
        return t

    def c_offers(self, lo: int = 0, hi: Optional[int] = None, shard_rows: int = 0, row_offset: Optional[int] = None) -> OeOfferTable:
        """Offer table for rows [lo, hi) (a shard when lo/hi are given)."""
        n_all = self.n_offers
        if self.n_offers_global is not None:  # pre-cut shard (cut_offers): every row present belongs to this rank
            lo, hi, whole = self.offer_lo, self.offer_lo + n_all, True
            n_all = self.n_offers_global
        else:
            hi = n_all if hi is None else hi
            whole = lo == 0 and hi == n_all

        def cut(a):
            if whole:
                return a
            b = np.ascontiguousarray(a[..., lo:hi])
            self._keep.append(b)
            return b
        t = OeOfferTable()
        # an empty trailing shard still sits at its nominal offset (rank * shard_rows)
        t.n_offers, t.row_offset, t.n_offers_global, t.shard_rows = hi - lo, (row_offset if row_offset is not None else lo), n_all, shard_rows
        S, K, _ = self.scalar.shape
        t.n_role_slots, t.n_scalar_kinds = S, K
        t.n_port_ranges = self.port_range.shape[1]
        t.n_mount_slots = self.mount_size.shape[0]
        t.n_attr_slots, t.n_rr_cols, t.n_exec_slots = self.attr_id.shape[0], self.rr_key.shape[0], self.exec_id.shape[0]
        t.scalar, t.scalar_present, t.port_range = _ptr(cut(self.scalar)), _ptr(cut(self.scalar_present)), _ptr(cut(self.port_range))
        t.mount_size, t.mount_profile = _ptr(cut(self.mount_size)), _ptr(cut(self.mount_profile))
        t.alloc_role, t.agent_id, t.hostname_id = _ptr(cut(self.alloc_role)), _ptr(cut(self.agent_id)), _ptr(cut(self.hostname_id))
        t.zone_id, t.region_id = _ptr(cut(self.zone_id)), _ptr(cut(self.region_id))
        t.attr_id, t.rr_key, t.exec_id = _ptr(cut(self.attr_id)), _ptr(cut(self.rr_key)), _ptr(cut(self.exec_id))
        if self.reserved_begin is not None and len(self.reserved_id):
            if not whole:
                raise NotImplementedError("sharding a table with a reserved side table")
            t.n_reserved = len(self.reserved_id)
            t.reserved_begin, t.reserved_id = _ptr(self.reserved_begin), _ptr(self.reserved_id)
            t.reserved_kind, t.reserved_value = _ptr(self.reserved_kind), _ptr(self.reserved_value)
        return t

    OFFER_COLUMNS = ("scalar", "scalar_present", "port_range", "mount_size", "mount_profile", "alloc_role", "agent_id", "hostname_id",
                     "zone_id", "region_id", "attr_id", "rr_key", "exec_id")

    def c_offer_rows(self, rows) -> OeOfferTable:
        """The given rows of the offer table as a small table of their own (what oe_update_offers takes)."""
        rows = np.asarray(rows, np.int64)
        t = OeOfferTable()
        t.n_offers, t.row_offset, t.n_offers_global, t.shard_rows = len(rows), 0, len(rows), 0
        S, K, _ = self.scalar.shape
        t.n_role_slots, t.n_scalar_kinds = S, K
        t.n_port_ranges = self.port_range.shape[1]
        t.n_mount_slots = self.mount_size.shape[0]
        t.n_attr_slots, t.n_rr_cols, t.n_exec_slots = self.attr_id.shape[0], self.rr_key.shape[0], self.exec_id.shape[0]
        for name in self.OFFER_COLUMNS:
            b = np.ascontiguousarray(getattr(self, name)[..., rows])
            self._keep.append(b)
            setattr(t, name, _ptr(b))
        return t

    def c_reqs(self) -> OeReqTable:
        t = OeReqTable()
        t.n_reqs = self.n_reqs
        t.class_id, t.pod_type_id, t.pod_index = _ptr(self.class_id), _ptr(self.r_pod_type_id), _ptr(self.r_pod_index)
        if self.self_begin is not None:
            t.self_begin = self.self_begin.ctypes.data
            t.self_rows = _ptr(self.self_rows)
        if self.launch_begin is not None:
            t.launch_begin = self.launch_begin.ctypes.data
            t.launch_name_id, t.launch_filter_bits = _ptr(self.launch_name_id), _ptr(self.launch_filter_bits)
        return t


def _narrow_ids(a: np.ndarray, dict_size: int):
    """Narrowest unsigned type whose all-ones value can stand for -1 and that holds every id of the dictionary."""
    for dt in (np.uint8, np.uint16):
        if dict_size < np.iinfo(dt).max:
            return np.where(a < 0, np.iinfo(dt).max, a).astype(dt), dt().itemsize
    return np.ascontiguousarray(a, np.int32), 4


class PackedOffers:
    """The compact wire form of (a shard of) an offer table — what an encoder that sees every value would emit instead of
    oe_offer_table's full-width columns (oe_offer_packing in include/oe_abi.h).  Built once per table; `pin()` moves the
    columns into page-locked memory."""

    def __init__(self, t: CycleTables, lo: int = 0, hi: Optional[int] = None):
        if t.n_offers_global is not None:    # a pre-cut shard: all of its rows
            lo_, hi_ = 0, t.n_offers
            self.row_lo, self.n_global = t.offer_lo, t.n_offers_global
        else:
            hi = t.n_offers if hi is None else hi
            lo_, hi_ = lo, hi
            self.row_lo, self.n_global = lo, t.n_offers
        cut = lambda a: np.ascontiguousarray(a[..., lo_:hi_])
        d = t.dict_sizes
        self.n = hi_ - lo_
        self.shape = (t.scalar.shape[0], t.scalar.shape[1], t.port_range.shape[1], t.mount_size.shape[0], t.attr_id.shape[0],
                      t.rr_key.shape[0], t.exec_id.shape[0])
        S, K, R = self.shape[:3]
        self.cols = {}
        scalar = cut(t.scalar)
        f32 = scalar.astype(np.float32)
        self.scalar_is_f32 = bool(np.array_equal(f32.astype(np.float64), scalar, equal_nan=True))
        self.cols["scalar"] = f32 if self.scalar_is_f32 else scalar
        pb = 1 if S * K <= 8 else 2 if S * K <= 16 else 4
        self.cols["scalar_present"], self.present_bytes = cut(t.scalar_present).astype({1: np.uint8, 2: np.uint16, 4: np.uint32}[pb]), pb
        self.cols["mount_size"] = cut(t.mount_size)
        self.cols["mount_profile"], self.mount_profile_bytes = _narrow_ids(cut(t.mount_profile), 64)
        self.cols["zone_id"], self.zone_bytes = _narrow_ids(cut(t.zone_id), d["n_zones"])
        self.cols["region_id"], self.region_bytes = _narrow_ids(cut(t.region_id), d["n_regions"])
        self.cols["attr_id"], self.attr_bytes = _narrow_ids(cut(t.attr_id), d["n_attrs"])
        self.cols["rr_key"], self.rr_bytes = _narrow_ids(cut(t.rr_key), d["n_rr_values"])
        self.cols["exec_id"], self.exec_bytes = _narrow_ids(cut(t.exec_id), d["n_exec_ids"])
        agent, host, alloc = cut(t.agent_id), cut(t.hostname_id), cut(t.alloc_role)
        self.cols["agent_id"] = agent
        self.hostname_is_agent = bool(np.array_equal(agent, host))
        if not self.hostname_is_agent:
            self.cols["hostname_id"] = host
        self.alloc_role_is_const = bool(self.n == 0 or (alloc == alloc[0]).all())
        self.alloc_role_const = int(alloc[0]) if self.n and self.alloc_role_is_const else -1
        if not self.alloc_role_is_const:
            self.cols["alloc_role"] = alloc
        # the distinct `ports` range lists and every (slot, offer)'s list
        ports = cut(t.port_range)                                      # [S, R, n]
        self.port_tuple_bytes = 0
        if R > 0 and self.n > 0:
            lists = np.ascontiguousarray(ports.transpose(0, 2, 1).reshape(S * self.n, R))
            uniq, inv = np.unique(lists, axis=0, return_inverse=True)
            if len(uniq) * 4 <= max(self.n, 4096):
                self.port_tuple_bytes = 2 if len(uniq) <= 65536 else 4
                self.cols["port_tuples"] = np.ascontiguousarray(uniq, np.uint32)
                self.cols["port_tuple_id"] = np.ascontiguousarray(inv.reshape(S, self.n), np.uint16 if self.port_tuple_bytes == 2 else np.uint32)
        if not self.port_tuple_bytes:
            self.cols["port_range"] = ports
        self.reserved = None
        if t.reserved_begin is not None and len(t.reserved_id):
            if lo_ != 0 or hi_ != t.n_offers:
                raise NotImplementedError("sharding a table with a reserved side table")
            self.reserved = (t.reserved_begin, t.reserved_id, t.reserved_kind, t.reserved_value)

    def pin(self):
        import torch
        self._pinned = {k: torch.from_numpy(v.copy()).pin_memory() for k, v in self.cols.items() if v.size}
        self.cols.update({k: v.numpy() for k, v in self._pinned.items()})
        return self

    @property
    def nbytes(self) -> int:
        return int(sum(v.nbytes for v in self.cols.values()))

    def c_structs(self, shard_rows: int = 0, row_offset: Optional[int] = None):
        c, t, pk = self.cols, OeOfferTable(), OeOfferPacking()
        t.n_offers, t.row_offset, t.n_offers_global, t.shard_rows = self.n, (self.row_lo if row_offset is None else row_offset), self.n_global, shard_rows
        (t.n_role_slots, t.n_scalar_kinds, t.n_port_ranges, t.n_mount_slots, t.n_attr_slots, t.n_rr_cols, t.n_exec_slots) = self.shape
        for name in ("scalar", "scalar_present", "port_range", "mount_size", "mount_profile", "alloc_role", "agent_id", "hostname_id", "zone_id",
                     "region_id", "attr_id", "rr_key", "exec_id"):
            setattr(t, name, _ptr(c.get(name)))
        if self.reserved is not None:
            rb, rid, rk, rv = self.reserved
            t.n_reserved, t.reserved_begin, t.reserved_id, t.reserved_kind, t.reserved_value = len(rid), _ptr(rb), _ptr(rid), _ptr(rk), _ptr(rv)
        pk.scalar_is_f32, pk.present_bytes = int(self.scalar_is_f32), self.present_bytes
        pk.zone_bytes, pk.region_bytes, pk.attr_bytes, pk.rr_bytes = self.zone_bytes, self.region_bytes, self.attr_bytes, self.rr_bytes
        pk.mount_profile_bytes, pk.exec_bytes = self.mount_profile_bytes, self.exec_bytes
        pk.hostname_is_agent, pk.alloc_role_is_const, pk.alloc_role_const = int(self.hostname_is_agent), int(self.alloc_role_is_const), self.alloc_role_const
        pk.port_tuple_bytes = self.port_tuple_bytes
        if self.port_tuple_bytes:
            pk.n_port_tuples, pk.port_tuples, pk.port_tuple_id = len(c["port_tuples"]), _ptr(c["port_tuples"]), _ptr(c["port_tuple_id"])
        return t, pk


class EngineError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"oe status {code}: {text}")
        self.code = code


class BoundLibrary:
    """Binds one shared library that exports the oe_abi.h entry points under `prefix`."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        self.lib = C.CDLL(path)
        self.prefix = prefix
        f = self.fn
        f("abi_version").restype = C.c_int32
        f("create").argtypes = [C.POINTER(OeConfig), C.POINTER(_p)]
        f("destroy").argtypes = [_p]
        f("destroy").restype = None
        f("last_error").argtypes = [_p]
        f("last_error").restype = C.c_char_p
        f("load_programs").argtypes = [_p, C.POINTER(OeProgramTable)]
        f("load_tasks").argtypes = [_p, C.POINTER(OeTaskTable)]
        f("load_offers").argtypes = [_p, C.POINTER(OeOfferTable)]
        f("load_requirements").argtypes = [_p, C.POINTER(OeReqTable)]
        f("matrix_stride_words").argtypes = [_p]
        f("matrix_stride_words").restype = C.c_int64
        f("feasibility").argtypes = [_p, _p, C.POINTER(OeStats)]
        f("evaluate_cycle").argtypes = [_p, _p, C.POINTER(OeStats)]
        f("apply_launches").argtypes = [_p]
        got = f("abi_version")()
        if got != OE_ABI_VERSION:
            raise RuntimeError(f"{path}: ABI version {got}, expected {OE_ABI_VERSION}")

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def has(self, name):
        return hasattr(self.lib, self.prefix + name)


class Handle:
    """One engine (or, in tests, one oracle) instance driven through the C ABI."""

    def __init__(self, bound: BoundLibrary, device=0, rank=0, world=1, flags=0, stream=0):
        self.b = bound
        self.h = _p()
        cfg = OeConfig(device, rank, world, flags, stream or None)
        rc = bound.fn("create")(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise EngineError(rc, (bound.fn("last_error")(None) or b"").decode())
        self.tables = None
        self.rank, self.world = rank, world

    def close(self):
        if self.h:
            self.b.fn("destroy")(self.h)
            self.h = _p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise EngineError(rc, (self.b.fn("last_error")(self.h) or b"").decode())

    # ---- loading
    def load_programs(self, t: CycleTables):
        self._cp = t.c_programs()
        self._check(self.b.fn("load_programs")(self.h, C.byref(self._cp)))

    def load_tasks(self, t: CycleTables):
        self._ct = t.c_tasks()
        self._check(self.b.fn("load_tasks")(self.h, C.byref(self._ct)))

    def load_offers(self, t: CycleTables, lo=0, hi=None, shard_rows=0):
        self._co = t.c_offers(lo, hi, shard_rows, self.rank * shard_rows if (shard_rows and self.world > 1) else None)
        self._n_local = int(self._co.n_offers)
        self._check(self.b.fn("load_offers")(self.h, C.byref(self._co)))

    def load_offers_packed(self, packed: PackedOffers, shard_rows=0):
        """oe_load_offers_packed: the same table in its compact wire form (widened on the device)."""
        self._co, self._cpk = packed.c_structs(shard_rows, self.rank * shard_rows if (shard_rows and self.world > 1) else None)
        self._n_local = int(self._co.n_offers)
        f = self.b.fn("load_offers_packed")
        f.argtypes = [_p, _p, _p]
        self._check(f(self.h, C.byref(self._co), C.byref(self._cpk)))

    def update_offers(self, rows, values: CycleTables, value_rows=None):
        """oe_update_offers: resident row rows[k] <- row value_rows[k] of `values` (default: the same row numbers)."""
        rows = np.ascontiguousarray(rows, np.int64)
        values.normalise()
        co = values.c_offer_rows(rows if value_rows is None else value_rows)
        f = self.b.fn("update_offers")
        f.argtypes = [_p, _p, C.c_int64, _p]
        self._check(f(self.h, rows.ctypes.data if len(rows) else None, len(rows), C.byref(co)))
        self._upd = (rows, co)

    def load_requirements(self, t: CycleTables):
        self._cr = t.c_reqs()
        self._check(self.b.fn("load_requirements")(self.h, C.byref(self._cr)))

    def load(self, t: CycleTables, lo=0, hi=None, shard_rows=0):
        t.normalise()
        self.tables = t
        self.load_programs(t)
        self.load_tasks(t)
        self.load_offers(t, lo, hi, shard_rows)
        self.load_requirements(t)
        return self

    # ---- evaluation
    def stride(self) -> int:
        return int(self.b.fn("matrix_stride_words")(self.h))

    def feasibility(self, fetch=True):
        """Full [n_reqs, stride] uint32 bit matrix (or None when fetch is False) and the stats."""
        st = OeStats()
        out = np.zeros((self.tables.n_reqs, self.stride()), dtype=np.uint32) if fetch else None
        self._check(self.b.fn("feasibility")(self.h, out.ctypes.data if fetch else None, C.byref(st)))
        return out, st.as_dict()

    def class_feasibility(self, fetch=True, rows=None):
        st = OeStats()
        f = self.b.fn("class_feasibility")
        f.argtypes = [_p, _p, C.POINTER(OeStats)]
        out = None
        if fetch:
            out = np.zeros((rows if rows is not None else self.factored_rows(), self.stride()), dtype=np.uint32)
        self._check(f(self.h, out.ctypes.data if fetch else None, C.byref(st)))
        if fetch and rows is None:
            out = out[: self.factored_rows()]   # the pass may merge the rows of re-deployed requirements (oe_requirement_rows)
        return out, st.as_dict()

    def requirement_rows(self):
        """(bit row of every requirement, number of bit rows) as of the last feasibility pass (oe_requirement_rows)."""
        f = self.b.fn("requirement_rows")
        f.argtypes = [_p, _p, C.POINTER(C.c_int32)]
        rows, n = np.zeros(self.tables.n_reqs, np.int32), C.c_int32()
        self._check(f(self.h, rows.ctypes.data if rows.size else None, C.byref(n)))
        return rows, int(n.value)

    def factored_rows(self) -> int:
        ptr, nbytes = self.device_buffer(BUF_CLASS_BITS_LOCAL)
        return nbytes // (4 * self.stride())

    def evaluate_cycle(self):
        st = OeStats()
        out = np.zeros(self.tables.n_reqs, dtype=DECISION_DTYPE)
        self._check(self.b.fn("evaluate_cycle")(self.h, out.ctypes.data, C.byref(st)))
        return out, st.as_dict()

    def gather_decisions(self):
        """oe_gather_decisions: the merged decision table of a sharded cycle (peers attached)."""
        out = np.zeros(self.tables.n_reqs, dtype=DECISION_DTYPE)
        f = self.b.fn("gather_decisions")
        f.argtypes = [_p, _p]
        self._check(f(self.h, out.ctypes.data if out.size else None))
        return out

    def evaluate_cycle_into(self, out_ptr):
        """Decisions written to caller-owned host memory (e.g. a pinned buffer); None on a peer-attached shard that leaves the
        table to oe_gather_decisions on another rank."""
        st = OeStats()
        f = self.b.fn("evaluate_cycle")
        f.argtypes = [_p, _p, C.POINTER(OeStats)]
        self._check(f(self.h, out_ptr, C.byref(st)))
        return st.as_dict()

    def gather_decisions_into(self, out_ptr):
        f = self.b.fn("gather_decisions")
        f.argtypes = [_p, _p]
        self._check(f(self.h, out_ptr))

    def stage_failures(self, j0=0, j1=None):
        """[j1 - j0, n_local_offers] uint64 failed-stage masks (oe_stage_failures)."""
        j1 = self.tables.n_reqs if j1 is None else j1
        f = self.b.fn("stage_failures")
        f.argtypes = [_p, C.c_int64, C.c_int64, _p]
        out = np.zeros((j1 - j0, self._n_local), np.uint64)
        self._check(f(self.h, j0, j1, out.ctypes.data if out.size else None))
        return out

    def unexpected_resources(self, keep_bits: np.ndarray, only_unused_offers=True):
        """Per reserved side-table entry: 1 iff its resource id is not in the keep set (oe_unexpected_resources)."""
        f = self.b.fn("unexpected_resources")
        f.argtypes = [_p, _p, C.c_int32, _p, C.POINTER(C.c_int64)]
        n = 0 if self.tables.reserved_id is None else len(self.tables.reserved_id)
        out = np.zeros(n, np.uint8)
        cnt = C.c_int64()
        keep = np.ascontiguousarray(keep_bits, np.uint32)
        self._check(f(self.h, keep.ctypes.data if keep.size else None, int(only_unused_offers), out.ctypes.data if n else None, C.byref(cnt)))
        return out, int(cnt.value)

    def apply_launches(self):
        self._check(self.b.fn("apply_launches")(self.h))

    def device_buffer(self, which: int):
        f = self.b.fn("device_buffer")
        f.argtypes = [_p, C.c_int32, C.POINTER(_p), C.POINTER(C.c_int64)]
        ptr, n = _p(), C.c_int64()
        self._check(f(self.h, which, C.byref(ptr), C.byref(n)))
        return int(ptr.value or 0), int(n.value)

    def peer_export(self, max_rows: int):
        """oe_peer_export: (64-byte IPC handle, device pointer) of this shard's exchange region."""
        f = self.b.fn("peer_export")
        f.argtypes = [_p, C.c_int32, _p, C.POINTER(_p)]
        handle = C.create_string_buffer(PEER_HANDLE_BYTES)
        ptr = _p()
        self._check(f(self.h, max_rows, handle, C.byref(ptr)))
        return handle.raw, int(ptr.value or 0)

    def peer_attach(self, handles=None, same_process_ptrs=None):
        """oe_peer_attach: handles = the shards' IPC handles in rank order (bytes, world x 64) and / or, for shards whose
        engine lives in this process, their exchange regions' device pointers (0 = use the handle)."""
        f = self.b.fn("peer_attach")
        f.argtypes = [_p, _p, _p]
        hb = C.create_string_buffer(b"".join(handles), self.world * PEER_HANDLE_BYTES) if handles is not None else None
        pa = (_p * self.world)(*[(q or None) for q in same_process_ptrs]) if same_process_ptrs is not None else None
        self._check(f(self.h, hb, pa))

    def peer_detach(self):
        f = self.b.fn("peer_detach")
        f.argtypes = [_p]
        self._check(f(self.h))

    def cuda_stream(self) -> int:
        f = self.b.fn("stream")
        f.argtypes, f.restype = [_p], _p
        return int(f(self.h) or 0)

    def export_tasks(self):
        """Current task table as a dict of numpy columns (after apply_launches)."""
        f = self.b.fn("export_tasks")
        t = self.tables
        TA, Q, FW = t.t_attr_id.shape[0], t.t_rr_key.shape[0], t.t_filter_bits.shape[0]
        if self.b.prefix == "oe_":
            f.argtypes = [_p, C.POINTER(C.c_int64)] + [_p] * 10
            n = C.c_int64()
            self._check(f(self.h, C.byref(n), *([None] * 10)))
            n = n.value
        else:
            cnt = self.b.fn("task_count")
            cnt.argtypes, cnt.restype = [_p], C.c_int64
            n = cnt(self.h)
            f.argtypes = [_p] + [_p] * 10
        cols = {k: np.zeros(n, np.int32) for k in ("name_id", "pod_type_id", "pod_index", "agent_id", "hostname_id", "zone_id", "region_id")}
        cols["attr_id"] = np.zeros((TA, n), np.int32)
        cols["rr_key"] = np.zeros((Q, n), np.int32)
        cols["filter_bits"] = np.zeros((FW, n), np.uint32)
        ptrs = [cols[k].ctypes.data if cols[k].size else None for k in
                ("name_id", "pod_type_id", "pod_index", "agent_id", "hostname_id", "zone_id", "region_id", "attr_id", "rr_key", "filter_bits")]
        if self.b.prefix == "oe_":
            n2 = C.c_int64()
            self._check(f(self.h, C.byref(n2), *ptrs))
        else:
            self._check(f(self.h, *ptrs))
        return cols


_engine_lib: Optional[BoundLibrary] = None


def engine_library() -> BoundLibrary:
    """The CUDA engine library.  Raises if it has not been built — the product has no other compute path."""
    global _engine_lib
    if _engine_lib is None:
        _engine_lib = BoundLibrary(ENGINE_LIB, "oe_")
    return _engine_lib


class Engine(Handle):
    """B200 offer-evaluation engine (liboe_b200.so) on one CUDA device."""

    def __init__(self, device=0, rank=0, world=1, flags=0, stream=0):
        super().__init__(engine_library(), device, rank, world, flags, stream)


def unpack_bits(matrix: np.ndarray, n_offers: int) -> np.ndarray:
    """[rows, stride] uint32 words → [rows, n_offers] bool."""
    b = np.unpackbits(matrix.view(np.uint8), axis=1, bitorder="little")
    return b[:, :n_offers].astype(bool)


OE_STATIC_MAX_OFFERS, OE_STATIC_MAX_ROWS = 4 * 1024 * 1024, 8192
PEER_HANDLE_BYTES, PEER_MAX_WORLD = 64, 16


def first_fit_static(device: int, stream: int, d_row_bits: int, stride_words: int, n_rows: int, n_offers: int, d_req_row: int,
                     n_reqs: int, d_offer_row: int) -> None:
    """oe_first_fit_static (include/oe_abi.h): ordered first-fit over bit rows that already sit in device memory, for hosts
    that cannot let the library allocate.  All three pointers are raw DEVICE addresses owned by the caller; the kernel is
    only enqueued on `stream` — synchronise before reading the result.  Raises EngineError with the library's text."""
    f = engine_library().fn("first_fit_static")
    f.argtypes = [C.c_int32, _p, _p, C.c_int64, C.c_int32, C.c_int64, _p, C.c_int64, _p, C.c_char_p, C.c_int32]
    f.restype = C.c_int32
    err = C.create_string_buffer(512)
    rc = f(device, stream or None, d_row_bits or None, stride_words, n_rows, n_offers, d_req_row or None, n_reqs, d_offer_row or None,
           err, len(err))
    if rc != 0:
        raise EngineError(rc, err.value.decode())
