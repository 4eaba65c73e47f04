"""Multi-GPU host logic (one process per GPU, torch.distributed for the plumbing).

Two sharding units, both from SURVEY 8(e):
  * proof partitions (compound_proof::partition_count,
    libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:85-87):
    independent, partition k -> rank k mod G, no data-path collective; the 192-byte proofs
    (libs/filecoin/include/nil/filecoin/proofs/constants.hpp:93) are gathered and concatenated in
    partition order (libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:304-308);
  * one multiexp by contiguous point range: every rank reduces its slice to one affine point, the
    G partials are all-gathered (96/192 bytes each) and summed on every rank.
The compute itself is always the CUDA library (b2g16.lib); the functions here take the local
result / callables so the collective logic is testable with the gloo backend on CPU.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """Contiguous [first, last) of n units owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(n, world)
    first = rank * base + min(rank, rem)
    return first, first + base + (1 if rank < rem else 0)


def partition_owner(k, world):
    return k % world


def my_partitions(n_partitions, rank, world):
    return [k for k in range(n_partitions) if partition_owner(k, world) == rank]


def _device(group=None):
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")


def all_gather_bytes(local, group=None):
    """local: uint8 numpy array (same shape on every rank) -> array [world, ...]."""
    dev = _device(group)
    t = torch.from_numpy(np.ascontiguousarray(local, dtype=np.uint8)).to(dev)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size(group))]
    dist.all_gather(out, t, group=group)
    return np.stack([o.cpu().numpy() for o in out])


def msm_sharded(local_partial, sum_fn, group=None):
    """local_partial: this rank's affine MSM result over its point range (uint8[96] or [192]);
    sum_fn(points[world, 96|192]) -> their group sum (b2g16 Context.g1_sum / g2_sum).
    Returns the full multiexp result, identical on every rank."""
    return sum_fn(all_gather_bytes(local_partial, group))


def msm_resident_sharded(ctx, bases, d_scalars_ptr, n_local, group=None):
    """Each rank holds its slice of the bases resident and its slice of the scalars on the device."""
    part = ctx.msm_resident(bases, d_scalars_ptr, n_local)
    return msm_sharded(part, ctx.g1_sum if bases.group == 1 else ctx.g2_sum, group)


def gather_proofs(local, n_partitions, group=None):
    """local: {partition index: 192 proof bytes} for the partitions this rank proved.
    Returns the multi-proof byte string (192 * n_partitions) on every rank."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = (n_partitions + world - 1) // world
    buf = np.zeros((per, 192), dtype=np.uint8)
    mine = my_partitions(n_partitions, rank, world)
    assert sorted(local) == mine, "rank %d proved %s, owns %s" % (rank, sorted(local), mine)
    for slot, k in enumerate(mine):
        buf[slot] = np.frombuffer(local[k], dtype=np.uint8)
    allb = all_gather_bytes(buf, group)                  # [world, per, 192]
    out = bytearray()
    for k in range(n_partitions):
        out += allb[partition_owner(k, world), k // world].tobytes()
    return bytes(out)


def prove_partitions(ctx, pk, assignments, rs, ss, group=None):
    """assignments[k], rs[k], ss[k] for every partition k (host arrays, same on every rank or at
    least present on the owner).  Each rank proves the partitions it owns; all ranks return the
    concatenated multi-proof."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    local = {k: ctx.prove(pk, assignments[k], rs[k], ss[k]) for k in my_partitions(len(assignments), rank, world)}
    return gather_proofs(local, len(assignments), group)
