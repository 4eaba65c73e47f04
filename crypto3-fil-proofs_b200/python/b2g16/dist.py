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


def prove_sharded(ctx, pk, d_assignment_ptr, r, s, group=None):
    """One proof over all ranks of the group (latency mode): every rank holds the key and the
    witness on its device, reduces its slice of the five queries (b2g16_prove_partial), the 576-byte partial
    sets are all-gathered and every rank assembles the same 192 bytes (b2g16_prove_assemble)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    part = ctx.prove_partial(pk, d_assignment_ptr, rank, world)
    return ctx.prove_assemble(pk, all_gather_bytes(part, group), r, s)


def partition_schedule(n_partitions, world):
    """Waves of a multi-partition proof on `world` GPUs (C5 = 10 partitions on 8 GPUs).
    Full waves run one partition per GPU; the remaining R < world partitions are each spread over a group of
    world // R' GPUs (R' = R rounded up to a power of two) instead of leaving world - R GPUs idle.
    Returns [(partition, [ranks])] in execution order; entries of one wave are consecutive."""
    out, k = [], 0
    while n_partitions - k >= world:
        out += [(k + r, [r]) for r in range(world)]
        k += world
    rem = n_partitions - k
    if rem:
        groups = 1
        while groups < rem:
            groups *= 2
        size = max(1, world // groups)
        out += [(k + g, list(range(g * size, (g + 1) * size))) for g in range(rem)]
    return out


class PartitionScheduler:
    """Creates the sub-groups of partition_schedule once (collective: every rank must construct it) and proves
    all partitions; d_assignment_for(k) returns the device pointer of partition k's witness on this rank (it is
    called only for partitions this rank takes part in)."""

    def __init__(self, n_partitions, group=None, ctx=None, log_m=None):
        """With ctx and log_m given, every sub-group also gets a PeerNtt so its proofs shard the witness map too
        (prove_sharded_full); otherwise they repeat it on every rank (prove_sharded)."""
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n = n_partitions
        self.plan = partition_schedule(n_partitions, self.world)
        self.groups = {}
        for k, ranks in self.plan:
            if len(ranks) > 1 and tuple(ranks) not in self.groups:
                self.groups[tuple(ranks)] = dist.new_group(ranks=ranks)        # every rank calls new_group for every group
        self.parent = group
        self.peers = {}
        if ctx is not None and log_m is not None:
            for ranks, g in self.groups.items():
                if self.rank in ranks and len(ranks) in (2, 4, 8) and log_m >= 2 * (len(ranks).bit_length() - 1):
                    self.peers[ranks] = PeerNtt(ctx, log_m, group=g)

    def close(self):
        for p in self.peers.values():
            p.close()
        self.peers = {}

    def prove_all(self, ctx, pk, d_assignment_for, rs, ss):
        local = {}
        for k, ranks in self.plan:
            if self.rank not in ranks:
                continue
            if len(ranks) == 1:
                local[k] = ctx.prove_device(pk, d_assignment_for(k), rs[k], ss[k])
            elif tuple(ranks) in self.peers:
                local[k] = prove_sharded_full(ctx, pk, self.peers[tuple(ranks)], d_assignment_for(k), rs[k], ss[k], group=self.groups[tuple(ranks)])
            else:
                local[k] = prove_sharded(ctx, pk, d_assignment_for(k), rs[k], ss[k], group=self.groups[tuple(ranks)])
        # owner of partition k for the gather = first rank of its group
        per = max(sum(1 for k, ranks in self.plan if ranks[0] == r) for r in range(self.world))
        buf = np.zeros((max(per, 1), 192), dtype=np.uint8)
        slot_of, counters = {}, [0] * self.world
        for k, ranks in self.plan:
            slot_of[k] = (ranks[0], counters[ranks[0]])
            counters[ranks[0]] += 1
        for k, (owner, slot) in slot_of.items():
            if owner == self.rank:
                buf[slot] = np.frombuffer(local[k], dtype=np.uint8)
        allb = all_gather_bytes(buf, self.parent)
        return b"".join(allb[slot_of[k][0], slot_of[k][1]].tobytes() for k in range(self.n))


# ------------------------------------------------------------------ sharded Fr transform
# One evaluation-domain transform of size N = 2^log_n over G ranks (replaces a
# single crypto3-math evaluation_domain<Fr>::fft / inverse_fft call when the vector is spread over
# several GPUs).  Rank r holds x[r M, (r+1) M), M = N / G.  With n = r M + j and k = k1 + G k2:
#     X[k1 + G k2] = sum_j w_M^(j k2) [ w_N^(j k1) sum_r w_G^(r k1) x[r M + j] ]
# forward:  all-to-all (columns -> ranks) | cross step | all-to-all (k1 -> ranks) | local size-M
#           transform: rank k1 now holds X[k1 + G k2] ("cyclic" order) | for "natural" order one more
#           all-to-all + a local G x M/G transpose.
# inverse:  the same steps backwards with inverse roots.
# `ops` supplies the per-rank device steps (b2g16.lib.Context on a GPU; the CPU tests pass an
# oracle-backed stand-in), tensors are uint8 [M, 32] canonical scalars on the backend's device.
def _a2a(t, world, group):
    """t: uint8 [world, rows, 32]; slice i goes to rank i; returns what every rank sent here."""
    out = torch.empty_like(t)
    dist.all_to_all_single(out.view(-1), t.contiguous().view(-1), group=group)
    return out


def ntt_sharded(ops, x_local, log_n, inverse=False, coset=False, order="natural", group=None):
    """x_local: this rank's block (uint8 tensor [M, 32]).  Forward: natural block in -> spectrum out;
    inverse: spectrum in -> natural block out.  order: layout of the spectrum side -- "natural"
    (rank r holds X[r M, (r+1) M)) or "cyclic" (rank r holds X[r + G k2], saving one exchange)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    log_g = world.bit_length() - 1
    assert (1 << log_g) == world and 1 <= log_g <= 3, "ntt_sharded: 2, 4 or 8 ranks"
    assert log_n >= 2 * log_g and order in ("natural", "cyclic")
    M = (1 << log_n) >> log_g
    cols = M >> log_g
    assert tuple(x_local.shape) == (M, 32) and x_local.dtype == torch.uint8
    if not inverse:
        x = x_local.clone() if coset else x_local
        if coset:
            ops.fr_coset_scale(x, log_n, rank * M, inverse=False)                     # x[n] *= g^n
        m1 = _a2a(x.view(world, cols, 32), world, group)                               # [r][j'] for my column chunk
        m2 = ops.ntt_cross(m1, log_n, log_g, rank, inverse=False, coset=coset)         # [k1][j']
        y = _a2a(m2, world, group).view(M, 32)                                         # rank k1: all j, natural
        ops.ntt_local(y, log_n - log_g, inverse=False)                                 # Y[k2] = X[k1 + G k2]
        if order == "cyclic":
            return y
        z = _a2a(y.view(world, cols, 32), world, group)                                # [k1][u] = X[rank M + u G + k1]
        return ops.fr_transpose(z.view(world * cols, 32), world, cols)                 # -> [u G + k1]
    if order == "natural":
        z = ops.fr_transpose(x_local, cols, world)                                     # [k1][u]
        y = _a2a(z.view(world, cols, 32), world, group).view(M, 32)                    # rank k1: Y[k2] natural
    else:
        y = x_local.clone()
    ops.ntt_local(y, log_n - log_g, inverse=True)                                      # u_{k1}[j], scaled by 1/M
    m1 = _a2a(y.view(world, cols, 32), world, group)                                   # [k1][j'] for my column chunk
    m2 = ops.ntt_cross(m1, log_n, log_g, rank, inverse=True, coset=coset)              # [r][j']
    x = _a2a(m2, world, group).view(M, 32)                                             # rank r: x[r M + j]
    if coset:
        ops.fr_coset_scale(x, log_n, rank * M, inverse=True)                           # x[n] *= n^-1 g^-n (cross left 1/n out)
    return x


class DeviceNttOps:
    """The per-rank steps of ntt_sharded on the GPU: every call is a kernel of libb2g16 on ctx's stream."""

    def __init__(self, ctx):
        self.ctx = ctx

    def ntt_local(self, t, log_m, inverse):
        self.ctx.ntt_device(t.data_ptr(), log_m, inverse=inverse, coset=False)

    def ntt_cross(self, m, log_n, log_g, chunk, inverse, coset):
        out = torch.empty_like(m)
        self.ctx.ntt_cross_device(m.data_ptr(), out.data_ptr(), log_n, log_g, chunk, inverse, coset)
        return out

    def fr_coset_scale(self, t, log_n, offset, inverse):
        self.ctx.fr_coset_scale_device(t.data_ptr(), t.shape[0], log_n, offset, inverse)

    def fr_transpose(self, t, rows, cols):
        out = torch.empty_like(t)
        self.ctx.fr_transpose_device(t.data_ptr(), out.data_ptr(), rows, cols)
        return out


class _RawDeviceBytes:
    """torch.as_tensor adaptor for a raw device pointer owned by the library"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerNtt:
    """The sharded transform with the cross step and both of its exchanges fused into one kernel over NVLink
    peer memory (b2g16_ntt_cross_p2p_device): every rank owns two IPC-exported buffers, `x` (its natural
    block, M scalars) and `y` (its cyclic spectrum block); the kernel loads its column chunk from all x's
    and stores row k1 into rank k1's y (inverse: the other way round).  What is left of the collective is
    two barriers.  NCCL is used only for the handle exchange at construction and the barriers.
        forward : x (natural block) -> y = X[rank + G k2]      inverse : y -> x
    x and y are uint8 [M, 32] torch views of the library's buffers; fill x (or y) in place."""

    def __init__(self, ctx, log_n, group=None):
        self.ctx, self.log_n, self.group = ctx, log_n, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.log_g = self.world.bit_length() - 1
        assert (1 << self.log_g) == self.world and 1 <= self.log_g <= 3 and log_n >= 2 * self.log_g
        self.M = (1 << log_n) >> self.log_g
        nbytes = self.M * 32
        self._own, handles = [], []
        for _ in range(2):
            ptr, h = ctx.peer_alloc(nbytes)
            self._own.append(ptr)
            handles.append(h)
        allh = [None] * self.world
        dist.all_gather_object(allh, handles, group=group)
        self._opened = []
        self.x_ptrs, self.y_ptrs = [], []
        for r in range(self.world):
            for which, lst in ((0, self.x_ptrs), (1, self.y_ptrs)):
                if r == self.rank:
                    lst.append(self._own[which])
                else:
                    p = ctx.peer_open(allh[r][which])
                    self._opened.append(p)
                    lst.append(p)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.x = torch.as_tensor(_RawDeviceBytes(self._own[0], nbytes), device=dev).view(self.M, 32)
        self.y = torch.as_tensor(_RawDeviceBytes(self._own[1], nbytes), device=dev).view(self.M, 32)

    def _barrier(self):
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self.group)

    def forward(self, coset=False):
        if coset:
            self.ctx.fr_coset_scale_device(self.x.data_ptr(), self.M, self.log_n, self.rank * self.M, False)
        self._barrier()                                                  # every x written, every y free
        self.ctx.ntt_cross_p2p_device(self.x_ptrs, self.y_ptrs, self.log_n, self.log_g, self.rank, False, coset)
        self._barrier()                                                  # every store into my y has landed
        self.ctx.ntt_device(self.y.data_ptr(), self.log_n - self.log_g, inverse=False, coset=False)
        return self.y

    def inverse(self, coset=False):
        self.ctx.ntt_device(self.y.data_ptr(), self.log_n - self.log_g, inverse=True, coset=False)
        self._barrier()
        self.ctx.ntt_cross_p2p_device(self.y_ptrs, self.x_ptrs, self.log_n, self.log_g, self.rank, True, coset)
        self._barrier()
        if coset:
            self.ctx.fr_coset_scale_device(self.x.data_ptr(), self.M, self.log_n, self.rank * self.M, True)
        return self.x

    def close(self):
        self._barrier()
        for p in self._opened:
            self.ctx.peer_close(p)
        self._opened = []
        self._barrier()
        for p in self._own:
            self.ctx.peer_free(p)
        self._own = []


def prove_sharded_full(ctx, pk, peer, d_assignment_ptr, r, s, group=None):
    """One proof over all ranks with the witness map sharded as well (prove_sharded repeats it on every rank).
    peer: a PeerNtt for the key's domain (log_n = log_m), built once and reused.  Rank k1 evaluates the rows
    k1 + G k2 of a, b, c (the cyclic order the sharded inverse transform consumes), the seven transforms run on the
    fused peer-memory path, the quotient by Z is pointwise, and the rank's block of H coefficients
    [rank M, (rank + 1) M) is exactly its share of the H query."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    M, log_m = peer.M, peer.log_n
    dev = peer.x.device
    a, b, c = (torch.empty((M, 32), dtype=torch.uint8, device=dev) for _ in range(3))
    ctx.abc_evals_strided_device(pk, d_assignment_ptr, rank, world, M, a.data_ptr(), b.data_ptr(), c.data_ptr())
    for v in (a, b, c):
        peer.y.copy_(v)
        peer.inverse()                                   # evaluations on the domain -> coefficients (natural block)
        peer.forward(coset=True)                         # -> evaluations on the coset g D (cyclic)
        v.copy_(peer.y)
    ctx.h_pointwise_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), M, log_m, peer.y.data_ptr())
    peer.inverse(coset=True)                             # -> H coefficients [rank M, (rank + 1) M)
    part = ctx.prove_partial_h(pk, d_assignment_ptr, rank, world, peer.x.data_ptr(), rank * M, M)
    return ctx.prove_assemble(pk, all_gather_bytes(part, group), r, s)
