"""Host-side multi-GPU logic (b2g16.dist) with world_size 2 over gloo on CPU: point-range sharding of
a multiexp with all-gather + sum of the partials, and partition -> rank scheduling with proof gather
(SURVEY 8e).  The per-rank compute is stood in by the CPU oracle here -- on the GPU box the same
functions are driven by the CUDA library (tests/test_gpu_parity.py::test_sharded_* and bench.py)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from b2g16 import codec, dist as bd
    from oracle import coracle
    coracle.set_threads(2)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        rng = np.random.Generator(np.random.PCG64(31337))          # same inputs on every rank
        n = 1001
        bases = coracle.g1_gen_mul(codec.random_fr(n, rng))
        scal = codec.random_fr(n, rng)
        first, last = bd.shard_range(n, rank, world)
        part = coracle.msm_g1(bases[first:last], scal[first:last])

        def sum_fn(pts):
            acc = pts[0]
            for p in pts[1:]:
                acc = coracle.g1_add(acc, p)
            return acc
        full = bd.msm_sharded(part, sum_fn)
        ok_msm = np.array_equal(full, coracle.msm_g1(bases, scal))
        # partitions: 5 partitions over 2 ranks, "proof" of partition k is a recognisable 192-byte string
        local = {k: bytes([k + 1]) * 192 for k in bd.my_partitions(5, rank, world)}
        multi = bd.gather_proofs(local, 5)
        ok_proofs = multi == b"".join(bytes([k + 1]) * 192 for k in range(5))
        # scheduled multi-proof: 5 partitions on 2 ranks = two full waves + one proof spread over both ranks.
        # The device calls are stubbed (a "pointer" is the partition index); what is under test is the sub-group
        # creation, who proves what, and the order of the gathered 192-byte records.

        class StubCtx:
            def prove_device(self, pk, ptr, r, s):
                return bytes([ptr + 1]) * 192

            def prove_partial(self, pk, ptr, part, parts):
                out = np.zeros(576, dtype=np.uint8)
                out[0], out[1], out[2] = ptr, part, parts
                return out

            def prove_assemble(self, pk, partials, r, s):
                p = np.asarray(partials).reshape(-1, 576)
                assert p.shape[0] == p[0, 2] and sorted(p[:, 1].tolist()) == list(range(p.shape[0])) and len(set(p[:, 0].tolist())) == 1
                return bytes([int(p[0, 0]) + 1]) * 192
        sched = bd.PartitionScheduler(5)
        multi2 = sched.prove_all(StubCtx(), None, lambda k: k, [None] * 5, [None] * 5)
        ok_sched = multi2 == b"".join(bytes([k + 1]) * 192 for k in range(5)) and [ranks for _, ranks in sched.plan] == [[0], [1], [0], [1], [0, 1]]
        q.put((rank, ok_msm, ok_proofs and ok_sched, (first, last)))
    finally:
        dist.destroy_process_group()


def test_shard_ranges_cover_everything():
    sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
    from b2g16 import dist as bd
    for n in (0, 1, 7, 8, 1001, 1 << 26):
        for world in (1, 2, 4, 8):
            r = [bd.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    assert [bd.partition_owner(k, 8) for k in range(10)] == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]   # C5: 10 partitions on 8 GPUs
    assert bd.my_partitions(10, 1, 8) == [1, 9]


def test_partition_schedule():
    sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
    from b2g16 import dist as bd
    # C5: 10 partitions on 8 GPUs -> one full wave, then 2 partitions on 4 GPUs each
    plan = bd.partition_schedule(10, 8)
    assert plan[:8] == [(k, [k]) for k in range(8)]
    assert plan[8:] == [(8, [0, 1, 2, 3]), (9, [4, 5, 6, 7])]
    assert bd.partition_schedule(1, 8) == [(0, list(range(8)))]                 # C4: one partition, all GPUs
    assert bd.partition_schedule(3, 8) == [(0, [0, 1]), (1, [2, 3]), (2, [4, 5])]
    assert bd.partition_schedule(5, 2) == [(0, [0]), (1, [1]), (2, [0]), (3, [1]), (4, [0, 1])]
    assert bd.partition_schedule(4, 1) == [(k, [0]) for k in range(4)]
    for p in range(0, 20):
        for w in (1, 2, 4, 8):
            plan = bd.partition_schedule(p, w)
            assert [k for k, _ in plan] == list(range(p))
            assert all(0 <= r < w for _, ranks in plan for r in ranks)


def test_world_size_2_gloo(oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), "sharded multiexp differs from the full one"
    assert all(r[2] for r in res), "gathered multi-proof is not the in-order concatenation"
