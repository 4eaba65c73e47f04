"""Multi-GPU parity check, one process per GPU over NCCL:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/multi_gpu_check.py
Every rank compares the sharded results with the same computation done whole on its own GPU:
  * b2g16.dist.ntt_sharded (forward / inverse x coset x natural / cyclic order) against b2g16_ntt_device
    of the full vector, bit for bit;
  * b2g16.dist.msm_resident_sharded against the full multiexp;
  * b2g16.dist.prove_sharded: one proof with the five multiexps range-sharded over the ranks equals b2g16_prove's;
  * b2g16.dist.prove_partitions: 2 N + 1 partitions under one key, the gathered 192 (2N+1)-byte multi-proof
    equals the proofs made one by one and passes b2g16_verify_batch.
tests/test_gpu_verify.py-style pytest wrapper: tests/test_gpu_multi.py (skips with fewer than 2 GPUs)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np
import torch
import torch.distributed as dist

from b2g16 import codec, dist as bd, lib, synth


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    ctx = lib.Context(local, ts.cuda_stream)
    ops = bd.DeviceNttOps(ctx)
    log_g = world.bit_length() - 1
    checks = 0
    # ---- transform
    for log_n in sorted({2 * log_g, 7, 12, 16, 21}):
        if log_n < 2 * log_g:
            continue
        n, M = 1 << log_n, (1 << log_n) // world
        x = torch.from_numpy(codec.random_fr(n, np.random.Generator(np.random.PCG64(40 + log_n)))).to(dev)   # same on every rank
        mine = x[rank * M:(rank + 1) * M].contiguous()
        for inverse in (False, True):
            for coset in (False, True):
                full = x.clone()
                ctx.ntt_device(full.data_ptr(), log_n, inverse=inverse, coset=coset)
                got = bd.ntt_sharded(ops, mine.clone(), log_n, inverse=inverse, coset=coset, order="natural")
                assert torch.equal(got, full[rank * M:(rank + 1) * M]), ("ntt natural", log_n, inverse, coset, rank)
                checks += 1
        full = x.clone()
        ctx.ntt_device(full.data_ptr(), log_n)
        spec = bd.ntt_sharded(ops, mine.clone(), log_n, order="cyclic")
        assert torch.equal(spec, full[rank::world].contiguous()), ("ntt cyclic", log_n, rank)
        assert torch.equal(bd.ntt_sharded(ops, spec, log_n, inverse=True, order="cyclic"), mine), ("ntt cyclic inverse", log_n, rank)
        checks += 2
    # ---- the same transform with the cross step fused with its exchanges over NVLink peer memory
    for log_n in sorted({2 * log_g, 12, 21}):
        n, M = 1 << log_n, (1 << log_n) // world
        x = torch.from_numpy(codec.random_fr(n, np.random.Generator(np.random.PCG64(90 + log_n)))).to(dev)
        pn = bd.PeerNtt(ctx, log_n)
        for coset in (False, True):
            full = x.clone()
            ctx.ntt_device(full.data_ptr(), log_n, inverse=False, coset=coset)
            pn.x.copy_(x[rank * M:(rank + 1) * M])
            spec = pn.forward(coset=coset)
            assert torch.equal(spec, full[rank::world].contiguous()), ("p2p ntt forward", log_n, coset, rank)
            back = pn.inverse(coset=coset)                                # y still holds the spectrum
            assert torch.equal(back, x[rank * M:(rank + 1) * M]), ("p2p ntt inverse", log_n, coset, rank)
            checks += 2
        pn.close()
    # ---- multiexp by point range
    n = 50000 + 7
    rng = np.random.Generator(np.random.PCG64(5))
    ks = torch.from_numpy(codec.random_fr(n, rng)).to(dev)
    sc = torch.from_numpy(codec.random_fr(n, rng)).to(dev)
    bases = ctx.bases_generate(1, ks.data_ptr(), n)
    want = ctx.msm_resident(bases, sc.data_ptr(), n)
    first, last = bd.shard_range(n, rank, world)
    mine_b = ctx.bases_upload(1, bases.download(first, last - first))
    got = bd.msm_resident_sharded(ctx, mine_b, sc[first:last].contiguous().data_ptr(), last - first)
    assert np.array_equal(got, want), ("msm sharded", rank)
    checks += 1
    # ---- partitions
    s = synth.SynthR1cs(10, 3, 2027, n_free=8)

    class Cs:
        log_m, n_in, n_aux, n_cons, mats = s.log_m, s.n_in, s.n_aux, s.n_cons, (s.A, s.B, s.C)
    rng = np.random.Generator(np.random.PCG64(6))
    toxic = codec.random_fr(5, rng)
    pk, vk = ctx.pk_generate(Cs, toxic, want_vk=True)
    n_part = 2 * world + 1
    zs = []
    for k in range(n_part):
        dz = torch.zeros((1 + s.n_in + s.n_aux, 32), dtype=torch.uint8, device=dev)
        head = s.z_head.copy()
        head[1:] = codec.random_fr(head.shape[0] - 1, np.random.Generator(np.random.PCG64(700 + k)))
        dz[: head.shape[0]] = torch.from_numpy(head).to(dev)
        ctx.synth_witness(pk, dz.data_ptr(), s.n_free)
        zs.append(dz.cpu().numpy())
    rs, ss = codec.random_fr(n_part, rng), codec.random_fr(n_part, rng)
    multi = bd.prove_partitions(ctx, pk, zs, rs, ss)
    assert len(multi) == 192 * n_part
    assert multi == b"".join(ctx.prove(pk, zs[k], rs[k], ss[k]) for k in range(n_part)), ("multi-proof", rank)
    # one proof spread over all ranks: same 192 bytes as the single-GPU prover
    dz0 = torch.from_numpy(zs[0]).to(dev)
    assert bd.prove_sharded(ctx, pk, dz0.data_ptr(), rs[0], ss[0]) == multi[:192], ("prove_sharded", rank)
    checks += 1
    # scheduled multi-proof: full waves one partition per GPU, the remainder spread over sub-groups
    sched = bd.PartitionScheduler(n_part, ctx=ctx, log_m=s.log_m)          # remainder proofs shard the witness map too
    dzs = {}

    def d_for(k):
        dzs[k] = torch.from_numpy(zs[k]).to(dev)
        return dzs[k].data_ptr()
    assert sched.prove_all(ctx, pk, d_for, rs, ss) == multi, ("scheduled multi-proof", rank)
    if world >= 4:
        n3 = world + world // 2                                       # remainder world/2 -> groups of 2
        sched3 = bd.PartitionScheduler(n3)
        assert sched3.prove_all(ctx, pk, d_for, rs, ss) == multi[:192 * n3], ("scheduled multi-proof, groups of 2", rank)
    sched.close()
    checks += 1
    # ... and with the witness map sharded too (transforms on the fused peer-memory path)
    pn = bd.PeerNtt(ctx, s.log_m)
    assert bd.prove_sharded_full(ctx, pk, pn, dz0.data_ptr(), rs[0], ss[0]) == multi[:192], ("prove_sharded_full", rank)
    assert bd.prove_sharded_full(ctx, pk, pn, d_for(1), rs[1], ss[1]) == multi[192:384], ("prove_sharded_full, reuse", rank)
    pn.close()
    checks += 1
    pvk = ctx.vk_prepare(vk)
    assert ctx.verify_batch(pvk, np.stack([z[1:1 + s.n_in] for z in zs]), multi)
    checks += 2
    pvk.free(); pk.free(); ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    print("rank %d/%d: %d multi-GPU checks passed" % (rank, world, checks), flush=True)


if __name__ == "__main__":
    main()
