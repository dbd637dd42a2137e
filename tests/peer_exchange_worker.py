"""One rank of tests/test_gpu_peer_exchange.py (launched by torch.distributed.run, one process per GPU): the sharded cycle
with peers attached — the feasibility pass stores its class rows into every rank's exchange region over NVLink peer
memory and oe_evaluate_cycle alone is the cycle — against the CPU oracle and against the NCCL all-gather path."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from dcos_commons_b200 import Engine, abi, sharded, synth
    from oracle_binding import Oracle

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n_offers = int(os.environ.get("OE_TEST_OFFERS", "60000"))
    eng = Engine(device=local, rank=rank, world=world, flags=abi.OE_CFG_MATERIALIZE_MATRIX)
    ref = Engine(device=local, rank=rank, world=world)           # the same cycle through the NCCL all-gather
    shard_rows, bounds = sharded.shard_plan(n_offers, world)
    lo, hi = bounds[rank]
    attached = False
    for k in range(5):
        t = synth.make_cluster(n_offers, 2000 + 100 * k, 16, 4, 4, 500 + 50 * k, "c5", seed=4242 + k).tables
        want = None
        if rank == 0 and k < 2:                                  # the oracle walks every task for every pair: two passes of it are enough
            sub = synth.make_cluster(n_offers, 2000 + 100 * k, 16, 4, 4, 500 + 50 * k, "c5", seed=4242 + k).tables
            want, _ = Oracle().load(sub).evaluate_cycle()
        eng.load(t, lo, hi, shard_rows)
        if not attached:
            assert sharded.attach_peers(eng, dist, max_rows=256), "these GPUs should reach each other over NVLink"
            attached = True
        got, st = eng.evaluate_cycle()                           # pass + exchange + commit, one call
        ref.load(t, lo, hi, shard_rows)
        ref.class_feasibility(fetch=False)
        sharded.gather_class_bits(ref, dist)
        exp, _ = ref.evaluate_cycle()
        assert np.array_equal(got["offer_row"], exp["offer_row"]), f"rank {rank} pass {k}: peer exchange and NCCL gather disagree"
        mine = (exp["offer_row"] >= lo) & (exp["offer_row"] < hi)
        assert got[mine].tobytes() == exp[mine].tobytes()
        if want is not None:
            assert np.array_equal(got["offer_row"], want["offer_row"]), f"pass {k}: engine and oracle disagree"
        # the merged table (every shard's winners, all fields): asked for on ranks 0 and world - 1 only, the others skip the copy
        if rank in (0, world - 1):
            merged = eng.gather_decisions()
            assert np.array_equal(merged["offer_row"], exp["offer_row"])
            assert (merged["n_new_reservations"][merged["offer_row"] >= 0] > 0).all(), "a winner of another shard was not filled in"
            if want is not None:
                assert merged.tobytes() == want.tobytes(), f"pass {k}: merged decisions differ from the oracle's"
        # every rank reached the same decisions
        mineT = torch.from_numpy(got["offer_row"].astype(np.int32)).cuda()
        allT = [torch.empty_like(mineT) for _ in range(world)]
        dist.all_gather(allT, mineT)
        for r in range(world):
            assert torch.equal(allT[r], mineT), f"ranks {rank} and {r} differ in pass {k}"
    dist.barrier()
    print(f"peer exchange ok: rank {rank} of {world}, {n_offers} offers, last cycle {st['ms_total']:.3f} ms, {st['n_placed']} placed", flush=True)
    eng.close(); ref.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
