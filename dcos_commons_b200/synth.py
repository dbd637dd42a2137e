"""Synthetic clusters of the shapes BASELINE.json names, generated directly as encoded tables.

All draws come from numpy.random.Generator(PCG64(seed)) with seed 20240607 unless overridden.  The generator plays the
role of the host encoder for data that never existed as protobufs: it emits what the encoder would emit for

  * one offer per agent, scalar pools cpus/mem/disk/gpus in the unreserved role "*", a six-range `ports` resource in
    the shape of a stock Mesos agent with 3 % of the agents missing one of the statically requested ports,
    attributes rack:r{0..199} and tier:{gold,silver,bronze}, a fault domain (region, zone);
  * pod classes whose stage pipeline is that of a one-task pod (S/offer/evaluate/OfferEvaluator.java:411-522):
    executor id, placement rule, executor overhead cpus 0.1 / disk 256 / mem 32
    (S/scheduler/SchedulerConfig.java:160-170,358-367), two static ports, one dynamic port, cpus, mem, one ROOT volume,
    launch;
  * placement rules as P/MarathonConstraintParser.java:61-87 would build them, task filter "<podType>-.*".
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import abi, tables as tb

SEED = 20240607
KIND_CPUS, KIND_MEM, KIND_DISK, KIND_GPUS = 0, 1, 2, 3
N_RACKS, N_TIERS = 200, 3
BASE_PORT_RANGES = [(1025, 2180), (2182, 3887), (3889, 5049), (5052, 8079), (8082, 8180), (8182, 32000)]
MOUNT_SIZES = (10240.0, 20480.0, 51200.0)
EXECUTOR_OVERHEAD = ((KIND_CPUS, 0.1), (KIND_DISK, 256.0), (KIND_MEM, 32.0))  # TreeMap order: cpus, disk, mem


@dataclass
class Cluster:
    tables: abi.CycleTables
    n_regions: int
    n_zones: int          # total zones
    class_cpus: np.ndarray
    class_mem: np.ndarray
    class_disk: np.ndarray
    class_ports: list
    name: str


def _class_rule(kind: str, c: int, n_regions: int, zones_per_region: int, maxper_zone: int):
    """Rule tree of pod class c."""
    f = c  # task filter id == pod type id ("<type>-.*")
    rack_bits = np.zeros(N_RACKS + N_TIERS, bool)
    rack_bits[100:200] = True  # "rack:r1[0-9][0-9]" over the attribute dictionary
    like = ("attr_any", rack_bits)
    group_by = ("rr", abi.KS_ZONE, 0, f, 12)
    max_per = ("maxper_host", f, 2)
    if kind == "c4":
        pick = c % 4
        if pick == 0:
            return like
        if pick == 1:
            return group_by
        if pick == 2:
            return max_per
        return ("and", [like, group_by, max_per])
    if kind == "c5-hostname":
        # no zone / region leaf: MaxPerHostname reads a tally per hostname (a gather over as many keys as there are agents)
        # and the AttributeRule walks the offer's attribute slots (P/MaxPerHostnameRule.java:112-124, P/AttributeRule.java:30-60)
        return ("and", [max_per, like])
    if kind in ("c5", "c5-mount"):
        region_bits = np.zeros(n_regions, bool)
        region_bits[c % n_regions] = True
        return ("and", [("region_in", region_bits), ("maxper_zone", f, maxper_zone), ("rr", abi.KS_ZONE, 0, f, zones_per_region)])
    raise ValueError(kind)


def make_cluster(n_offers: int, n_reqs: int, n_classes: int, n_regions: int, zones_per_region: int, n_deployed: int = 0,
                 rules: str = "c4", seed: int = SEED, name: str = "synthetic", redeploy_every: int = 0) -> Cluster:
    """rules: "c4" | "c5" | "c5-hostname" (MaxPerHostname + AttributeRule, no zone / region leaf) | "c5-mount" (c5's rules,
    plus one MOUNT volume with a profile list and a range-constrained dynamic port per pod; the agents carry two MOUNT disks).
    redeploy_every = k > 0: every k-th requirement re-deploys a pod that is already running (same type and index as a
    deployed task, S/offer/TaskUtils.java:136-145), so its rules must not count the pod's own tasks."""
    rng = np.random.Generator(np.random.PCG64(seed))
    N, M, Cn = n_offers, n_reqs, n_classes
    n_zones = n_regions * zones_per_region

    # ---- offers: one per agent
    cpus = rng.integers(2, 65, N).astype(np.float64)
    mem = (rng.integers(4, 513, N) * 1024).astype(np.float64)        # MiB
    disk = (rng.integers(50, 2001, N) * 1024).astype(np.float64)     # MiB
    gpus = np.where(rng.integers(0, 4, N) == 3, 8.0, 0.0)
    scalar = np.stack([cpus, mem, disk, gpus])[None, :, :]            # [S=1, K=4, N]
    present = np.where(gpus > 0, 0b1111, 0b0111).astype(np.uint32)    # Mesos omits zero-valued resources
    zone = rng.integers(0, n_zones, N).astype(np.int32)
    region = (zone // zones_per_region).astype(np.int32)
    rack = rng.integers(0, N_RACKS, N).astype(np.int32)
    tier = (N_RACKS + rng.integers(0, N_TIERS, N)).astype(np.int32)
    ids = np.arange(N, dtype=np.int32)

    # ---- classes
    cpu_choices, mem_choices = (0.5, 1.0, 2.0, 4.0), (1024.0, 4096.0, 16384.0)
    class_cpus = np.array([cpu_choices[c % 4] for c in range(Cn)])
    class_mem = np.array([mem_choices[(c // 4) % 3] for c in range(Cn)])
    class_disk = np.array([1024.0 * (1 + c % 5) for c in range(Cn)])
    class_ports = [(7000 + c, 9042 + c) for c in range(Cn)]
    per_class = max(1, M // Cn)
    maxper_zone = max(2, (2 * per_class + zones_per_region - 1) // zones_per_region)
    prog = tb.ProgramBuilder()
    for c in range(Cn):
        ops = [("exec", -1), ("rule", _class_rule(rules, c, n_regions, zones_per_region, maxper_zone))]
        ops += [("scalar", k, v) for k, v in EXECUTOR_OVERHEAD]
        dyn_ranges = [(1200 + 10 * c, 1299 + 10 * c), (3000, 3049)] if rules == "c5-mount" else None
        ops += [("port", class_ports[c][0]), ("port", class_ports[c][1]), ("dynport", dyn_ranges)]
        ops += [("scalar", KIND_CPUS, float(class_cpus[c])), ("scalar", KIND_MEM, float(class_mem[c]))]
        ops += [("scalar", KIND_DISK, float(class_disk[c]))]
        if rules == "c5-mount":   # VolumeSpec MOUNT with a profile list (S/specification/VolumeSpec.java:13-38): profiles {c % 3, 2}
            ops += [("mount", float(MOUNT_SIZES[c % 3]), (1 << (c % 3)) | 4)]
        ops += [("nop",)]
        prog.add_class(ops)

    # ---- ports: stock ranges; 3 % of the agents have one statically requested port already taken
    R = 8
    port = np.full((1, R, N), abi.OE_PORT_EMPTY, np.uint32)
    for q, (b, e) in enumerate(BASE_PORT_RANGES):
        port[0, q, :] = tb.pack_range(b, e)
    holes = np.nonzero(rng.random(N) < 0.03)[0]
    all_static = np.array([p for pair in class_ports for p in pair])
    hole_port = all_static[rng.integers(0, len(all_static), len(holes))]
    for i, p in zip(holes, hole_port):
        ranges = []
        for (b, e) in BASE_PORT_RANGES:
            if b <= p <= e:
                if b <= p - 1:
                    ranges.append((b, p - 1))
                if p + 1 <= e:
                    ranges.append((p + 1, e))
            else:
                ranges.append((b, e))
        port[0, :, i] = abi.OE_PORT_EMPTY
        for q, (b, e) in enumerate(ranges):
            port[0, q, i] = tb.pack_range(b, e)

    if rules == "c5-mount":   # two MOUNT disks per agent: sizes around what the classes ask for, a profile on three of four
        msize = np.array(MOUNT_SIZES + (8192.0,))[rng.integers(0, 4, (2, N))]
        mprof = rng.integers(-1, 3, (2, N)).astype(np.int32)
    else:
        msize, mprof = np.zeros((0, N)), np.zeros((0, N), np.int32)
    offers = dict(scalar=scalar, scalar_present=present, port_range=port, mount_size=msize, mount_profile=mprof,
                  alloc_role=np.full(N, -1, np.int32), agent_id=ids, hostname_id=ids, zone_id=zone, region_id=region,
                  attr_id=np.stack([rack, tier]), rr_key=np.zeros((0, N), np.int32), exec_id=np.zeros((0, N), np.int32))

    # ---- requirements in plan order: a seeded interleaving of the classes; indices continue after the deployed pods
    class_id = rng.permutation(np.arange(M) % Cn).astype(np.int32)
    deployed_per_class = n_deployed // Cn
    pod_index = np.zeros(M, np.int32)
    nxt = np.full(Cn, deployed_per_class, np.int64)
    redo = np.zeros(Cn, np.int64)      # deployed pods of the class already chosen for a re-deployment
    redeployed = np.zeros(M, bool)
    for j in range(M):
        c = class_id[j]
        if redeploy_every > 0 and j % redeploy_every == 0 and redo[c] < deployed_per_class:
            pod_index[j] = redo[c]
            redo[c] += 1
            redeployed[j] = True
        else:
            pod_index[j] = nxt[c]
            nxt[c] += 1

    # ---- deployed tasks: `deployed_per_class` pods of every class on random agents (one task per pod)
    T = deployed_per_class * Cn
    FW = (Cn + 31) // 32
    if T > 0:
        t_type = np.repeat(np.arange(Cn, dtype=np.int32), deployed_per_class)
        t_index = np.tile(np.arange(deployed_per_class, dtype=np.int32), Cn)
        if rules in ("c5", "c5-mount"):
            # pods deployed earlier obeyed the same RegionRule: class c lives in region c % n_regions
            by_region = [np.flatnonzero(region == r) for r in range(n_regions)]
            t_agent = np.empty(T, np.int32)
            for c in range(Cn):
                pool = by_region[c % n_regions]
                sel = slice(c * deployed_per_class, (c + 1) * deployed_per_class)
                t_agent[sel] = pool[rng.integers(0, len(pool), deployed_per_class)] if len(pool) else rng.integers(0, N, deployed_per_class)
        else:
            t_agent = rng.integers(0, N, T).astype(np.int32)
        fb = np.zeros((FW, T), np.uint32)
        fb[t_type >> 5, np.arange(T)] = (np.uint32(1) << (t_type & 31).astype(np.uint32))
        tasks = dict(t_name_id=np.arange(T, dtype=np.int32), t_pod_type_id=t_type, t_pod_index=t_index, t_agent_id=t_agent,
                     t_hostname_id=t_agent, t_zone_id=zone[t_agent], t_region_id=region[t_agent],
                     t_attr_id=np.stack([rack[t_agent], tier[t_agent]]), t_rr_key=np.zeros((0, T), np.int32), t_filter_bits=fb)
    else:
        tasks = tb.empty_tasks(TA=2, Q=0, FW=FW)

    # launch rows for multi-cycle replay: one stored task per pod, named <type>-<index>-server
    launch_begin = np.arange(M + 1, dtype=np.int32)
    launch_name = (T + np.arange(M)).astype(np.int32)
    # a re-deployed pod overwrites its stored task (same name): row class * deployed_per_class + index of the task table
    launch_name[redeployed] = (class_id[redeployed].astype(np.int64) * deployed_per_class + pod_index[redeployed]).astype(np.int32)
    lf = np.zeros((M, FW), np.uint32)
    lf[np.arange(M), class_id >> 5] = (np.uint32(1) << (class_id & 31).astype(np.uint32))

    ds = dict(n_agents=N, n_hostnames=N, n_zones=n_zones, n_regions=n_regions, n_attrs=N_RACKS + N_TIERS, n_rr_values=0,
              n_pod_types=Cn, n_task_names=T + M, n_filters=Cn, n_exec_ids=0, n_roles=1, n_res_ids=0)
    t = tb.make_tables(ds, prog, tasks, offers, class_id, class_id, pod_index, launch_begin=launch_begin, launch_name_id=launch_name,
                       launch_filter_bits=lf)
    return Cluster(t, n_regions, n_zones, class_cpus, class_mem, class_disk, class_ports, name)


def config_c4(scale: float = 1.0, seed: int = SEED) -> Cluster:
    """BASELINE config 4: 10k agents x 5k pod instances, 8 classes, 3 regions x 4 zones, T = 0 at cycle start."""
    return make_cluster(int(10_000 * scale), int(5_000 * scale), 8, 3, 4, 0, "c4", seed, "c4-10k-agents-x-5k-instances")


def config_c5_variant(which: str, scale: float = 1.0, seed: int = SEED) -> Cluster:
    """C5-sized workloads that leave the headline configuration's fast paths (same 1 M offers x 100 k instances, 32 classes,
    100 k deployed pods): "hostname" = MaxPerHostname(2) + AttributeRule LIKE instead of the zone / region rules; "mount" = one
    MOUNT volume with profiles and a range-constrained dynamic port per pod; "redeploy" = 5 % of the requirements re-deploy a
    running pod (5 000 bit rows of their own); "one-region" = all classes compete for the same offers (one commit component)."""
    N, M, T = int(1_000_000 * scale), int(100_000 * scale), int(100_000 * scale)
    if which == "hostname":
        return make_cluster(N, M, 32, 8, 6, T, "c5-hostname", seed, "c5-hostname-1M-offers-x-100k-instances")
    if which == "mount":
        return make_cluster(N, M, 32, 8, 6, T, "c5-mount", seed, "c5-mount-1M-offers-x-100k-instances")
    if which == "redeploy":
        return make_cluster(N, M, 32, 8, 6, T, "c5", seed, "c5-redeploy-1M-offers-x-100k-instances", redeploy_every=20)
    if which == "one-region":
        return make_cluster(N, M, 32, 1, 6, T, "c5", seed, "c5-one-region-1M-offers-x-100k-instances")
    raise ValueError(which)


def config_c5(scale: float = 1.0, seed: int = SEED) -> Cluster:
    """BASELINE config 5: 1M offers x 100k pod instances, 32 classes, 8 regions x 6 zones.  100k pods (scaled) of the
    same services are already deployed so that the MaxPerZone / RoundRobinByZone tallies are not trivially empty."""
    return make_cluster(int(1_000_000 * scale), int(100_000 * scale), 32, 8, 6, int(100_000 * scale), "c5", seed,
                        "c5-1M-offers-x-100k-instances")
