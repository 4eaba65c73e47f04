"""b2g16.dist.ntt_sharded with world_size 2 over gloo on CPU: the exchange sequence and
index arithmetic of the sharded evaluation-domain transform, with the per-rank device steps stood in by the
CPU oracle (oracle.coracle.ntt for the local transform, big-int arithmetic for the cross step).  The
result must equal the oracle's transform of the whole vector.  On the GPU box the same function runs with
b2g16.dist.DeviceNttOps (tests/multi_gpu_check.py, bench.py --gpus N)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleNttOps:
    def __init__(self, coracle, ref):
        self.o, self.ref = coracle, ref

    def ntt_local(self, t, log_m, inverse):
        import torch
        t.copy_(torch.from_numpy(self.o.ntt(t.numpy(), inverse=inverse, coset=False)))

    def ntt_cross(self, m, log_n, log_g, chunk, inverse, coset):
        import torch
        R, G = self.ref.R, 1 << log_g
        cols = m.shape[1]
        wn, wg = self.ref.fr_root_of_unity(log_n), self.ref.fr_root_of_unity(log_g)
        if inverse:
            wn, wg = pow(wn, R - 2, R), pow(wg, R - 2, R)
        scale = 1 if not inverse else ((1 << (log_n - log_g)) if coset else pow(G, R - 2, R))
        a = m.numpy()
        out = np.zeros_like(a)
        for jj in range(cols):
            j = chunk * cols + jj
            v = [int.from_bytes(a[r, jj].tobytes(), "little") for r in range(G)]
            if inverse:
                v = [v[k] * pow(wn, j * k, R) % R for k in range(G)]
            y = [sum(v[r] * pow(wg, r * k, R) for r in range(G)) % R for k in range(G)]
            if not inverse:
                y = [y[k] * pow(wn, j * k, R) % R for k in range(G)]
            for k in range(G):
                out[k, jj] = np.frombuffer((y[k] * scale % R).to_bytes(32, "little"), dtype=np.uint8)
        return torch.from_numpy(out)

    def fr_coset_scale(self, t, log_n, offset, inverse):
        import torch
        R = self.ref.R
        g = self.ref.FR_GENERATOR if not inverse else pow(self.ref.FR_GENERATOR, R - 2, R)
        pre = 1 if not inverse else pow(1 << log_n, R - 2, R)
        a = t.numpy()
        for i in range(a.shape[0]):
            v = int.from_bytes(a[i].tobytes(), "little") * pre * pow(g, offset + i, R) % R
            a[i] = np.frombuffer(v.to_bytes(32, "little"), dtype=np.uint8)

    def fr_transpose(self, t, rows, cols):
        return t.view(rows, cols, 32).transpose(0, 1).contiguous().view(rows * cols, 32)


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    from b2g16 import codec, dist as bd
    from oracle import coracle
    from oracle.pyref import bls12_381 as ref
    coracle.set_threads(1)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        ops = OracleNttOps(coracle, ref)
        results = []
        for log_n in (2 * (world.bit_length() - 1), 6, 9):
            n, M = 1 << log_n, (1 << log_n) // world
            x = codec.random_fr(n, np.random.Generator(np.random.PCG64(100 + log_n)))    # same on every rank
            mine = torch.from_numpy(x[rank * M:(rank + 1) * M].copy())
            for inverse in (False, True):
                for coset in (False, True):
                    want = coracle.ntt(x, inverse=inverse, coset=coset)
                    got = bd.ntt_sharded(ops, mine.clone(), log_n, inverse=inverse, coset=coset, order="natural").numpy()
                    results.append(np.array_equal(got, want[rank * M:(rank + 1) * M]))
            # cyclic spectrum order: rank r holds X[r + G k2]; forward then inverse is the identity
            spec = bd.ntt_sharded(ops, mine.clone(), log_n, order="cyclic")
            full = coracle.ntt(x)
            results.append(np.array_equal(spec.numpy(), full[rank::world]))
            back = bd.ntt_sharded(ops, spec, log_n, inverse=True, order="cyclic")
            results.append(np.array_equal(back.numpy(), mine.numpy()))
        q.put((rank, results))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_ntt_sharded_gloo(oracle, world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() + world) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == list(range(world))
    for rank, results in res:
        assert all(results), "rank %d: sharded transform differs from the oracle: %s" % (rank, results)
