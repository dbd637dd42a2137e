"""Offer sharding across the GPUs of one box: contiguous row blocks, so that global first-fit order is
(shard, local row); programs, tasks and requirements are replicated; the one exchange step gathers the class-factored
bit rows of every shard, after which every rank replays the same commit.

This module is host-side plumbing shared by bench.py and the tests: the shard plan, the layout of the gathered rows
and the NCCL all-gather over the engine's own device buffers.
"""
from __future__ import annotations

import numpy as np

from . import abi


def shard_plan(n_offers: int, world: int):
    """(shard_rows, [(lo, hi)] per rank).  shard_rows is a multiple of 1024 (32 words x 32 offers) so that every shard's
    bit rows start on a word boundary of the global row and have the same padded stride on every rank."""
    if world <= 1:
        return 0, [(0, n_offers)]
    shard_rows = ((n_offers + world - 1) // world + 1023) // 1024 * 1024
    return shard_rows, [(min(n_offers, r * shard_rows), min(n_offers, (r + 1) * shard_rows)) for r in range(world)]


def local_stride_words(shard_rows: int, n_local: int) -> int:
    """Row stride of one shard's bit rows: ceil(rows / 32) rounded up to 128 words (oe_matrix_stride_words)."""
    rows = shard_rows if shard_rows > 0 else n_local
    return max(128, ((rows + 31) // 32 + 127) // 128 * 128)


def global_rows(gathered: np.ndarray, shard_rows: int, n_offers: int) -> np.ndarray:
    """[world, rows, stride] gathered shard rows -> [rows, ceil(n_offers / 32)] rows in global offer order."""
    world, rows, stride = gathered.shape
    shard_words = shard_rows // 32
    total_words = (n_offers + 31) // 32
    out = np.zeros((rows, total_words), np.uint32)
    for r in range(world):
        w0 = r * shard_words
        w1 = min(total_words, w0 + shard_words)
        if w1 > w0:
            out[:, w0:w1] = gathered[r, :, : w1 - w0]
    return out


def first_fit(rows: np.ndarray, req_row: np.ndarray, n_offers: int) -> np.ndarray:
    """Reference commit over bit rows (S/scheduler/plan/PlanScheduler.java:56-67): requirements in order, each takes the
    lowest-index offer of its row that no earlier requirement took.  O(M x candidates); for tests."""
    bits = np.unpackbits(rows.view(np.uint8), axis=1, bitorder="little")[:, :n_offers].astype(bool)
    cand = [np.flatnonzero(bits[r]) for r in range(bits.shape[0])]
    cursor = np.zeros(bits.shape[0], np.int64)
    used = np.zeros(n_offers, bool)
    out = np.full(len(req_row), -1, np.int32)
    for j, r in enumerate(req_row):
        c, k = cand[r], cursor[r]
        while k < len(c) and used[c[k]]:
            k += 1
        cursor[r] = k
        if k < len(c):
            out[j] = c[k]
            used[c[k]] = True
    return out


def attach_peers(engine: "abi.Engine", dist, max_rows: int) -> bool:
    """One process per GPU: every rank exports its exchange region (oe_peer_export), the 64-byte IPC handles go round
    through the process group, and every rank maps all of them (oe_peer_attach).  From then on the engine's feasibility
    pass stores its class rows into every shard's region itself and oe_evaluate_cycle is the whole sharded cycle: no
    all-gather call.  Collective: all ranks call it, after loading their offer shard.  Returns False — on EVERY rank, with
    every rank detached again — when any rank could not export or map (no peer access between the GPUs): the caller then
    keeps gather_class_bits between oe_class_feasibility and oe_evaluate_cycle."""
    handle, err = None, None
    try:
        handle, _ = engine.peer_export(max_rows)
    except abi.EngineError as ex:
        err = ex
    handles = [None] * engine.world
    dist.all_gather_object(handles, handle)
    ok = err is None and all(h is not None for h in handles)
    if ok:
        try:
            engine.peer_attach(handles=handles)
        except abi.EngineError as ex:
            err, ok = ex, False
    verdicts = [None] * engine.world
    dist.all_gather_object(verdicts, ok)      # also: nobody publishes before everybody has mapped everybody
    if all(verdicts):
        return True
    if ok:
        engine.peer_detach()
    elif err is not None:
        import sys
        print(f"rank {engine.rank}: peer exchange unavailable: {err}", file=sys.stderr)
    return False


def attach_local(engines, max_rows: int) -> None:
    """The same wiring for shards whose engines all live in THIS process (tests: several handles on one device, or one
    handle per device of the box): the exchange regions are handed over as plain device pointers."""
    ptrs = [e.peer_export(max_rows)[1] for e in engines]
    for e in engines:
        e.peer_attach(same_process_ptrs=[0 if r == e.rank else ptrs[r] for r in range(len(engines))])


class _CudaView:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch.distributed can gather engine memory."""

    def __init__(self, ptr, n_words):
        self.__cuda_array_interface__ = {"shape": (n_words,), "typestr": "<i4", "data": (ptr, False), "version": 2}


def gather_class_bits(engine: "abi.Engine", dist) -> None:
    """All-gather of this rank's class-factored rows (OE_BUF_CLASS_BITS_LOCAL) into every rank's
    OE_BUF_CLASS_BITS_GLOBAL over NCCL.  The engine's stream was synchronised by the call that produced the rows."""
    import torch
    lp, lb = engine.device_buffer(abi.BUF_CLASS_BITS_LOCAL)
    gp, gb = engine.device_buffer(abi.BUF_CLASS_BITS_GLOBAL)
    local = torch.as_tensor(_CudaView(lp, lb // 4), device="cuda")
    glob = torch.as_tensor(_CudaView(gp, gb // 4), device="cuda")
    dist.all_gather_into_tensor(glob, local)
    torch.cuda.current_stream().synchronize()
