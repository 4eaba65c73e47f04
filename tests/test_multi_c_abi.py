"""Several devices behind the C ABI (b2g16_init_multi / b2g16_prove_batch_multi / b2g16_prove_multi, csrc/multi.cu):
the partition loop of compound_proof::prove
(libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:89-95,127-137) spread over the box's
GPUs from ONE process.  The schedule is checked on however many devices the lease has; with one device the same
device is listed several times (each entry gets its own context, stream and host thread), which exercises the full
waves, the latency-mode remainder wave and the assembly exactly as separate devices would."""
import numpy as np
import pytest

from b2g16 import lib
import util


def test_init_multi_without_a_device_is_an_error():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(lib.B2g16Error) as e:
        lib.MultiContext([0, 1])
    assert e.value.code == -4            # B2G16_E_CUDA: no CPU fallback
    with pytest.raises(lib.B2g16Error) as e:
        lib.MultiContext([])
    assert e.value.code == -1


@pytest.mark.parametrize("g,parts", [(8, 10), (8, 8), (8, 3), (4, 6), (3, 4), (2, 5), (1, 3), (8, 1), (5, 0)])
def test_schedule_covers_every_device_once_per_wave(g, parts):
    """b2g16_multi_schedule (host-only): in the w full waves device i takes partitions [i w, (i + 1) w); the last partial wave splits ALL g
    devices into one contiguous group per remaining partition, sizes differing by at most one -- the 8 x 1 + 2 x 4 plan
    for the 10 partitions of a 32 GiB sector on 8 GPUs (constants.hpp:70-73)."""
    import ctypes as C
    plan = (C.c_uint32 * (3 * max(parts, 1)))()
    assert lib.load().b2g16_multi_schedule(g, C.c_size_t(parts), plan) == 0
    rows = [tuple(plan[3 * p:3 * p + 3]) for p in range(parts)]
    waves = parts // g
    for p in range(waves * g):
        assert rows[p] == (p % waves, p // waves, 1)                     # device i: the contiguous run [i w, (i + 1) w)
    rem = rows[waves * g:]
    if rem:
        assert all(w == waves for w, _, _ in rem)
        cover = []
        for _, first, size in rem:
            cover += list(range(first, first + size))
        assert cover == list(range(g))                                   # every device used exactly once, in order
        sizes = [size for _, _, size in rem]
        assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    if (g, parts) == (8, 10):
        assert rows[8] == (1, 0, 4) and rows[9] == (1, 4, 4)
    assert lib.load().b2g16_multi_schedule(0, C.c_size_t(1), plan) == -1


@pytest.mark.gpu
@pytest.mark.parametrize("slots,n_part", [(2, 5), (3, 4), (4, 2), (1, 2)])
def test_prove_batch_multi_equals_single_device(oracle, slots, n_part):
    import torch
    n_dev = torch.cuda.device_count()
    devices = [i % n_dev for i in range(slots)]
    log_m, n_in, n_free, seed = 12, 5, 8, 99
    toxic, _, _ = util.toxic_from_seed(seed)
    rng = np.random.Generator(np.random.PCG64(seed))
    from b2g16 import codec
    rs, ss = codec.random_fr(n_part, rng), codec.random_fr(n_part, rng)

    # single-device reference proofs, one witness per partition (different inputs / free variables)
    ctx = lib.Context(0)
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
    nv = pk.shape()["n_vars"]
    zs = np.zeros((n_part, nv, 32), dtype=np.uint8)
    want = []
    for p in range(n_part):
        dz = torch.zeros((nv, 32), dtype=torch.uint8, device="cuda:0")
        head = codec.random_fr(1 + n_in + n_free, rng); head[0] = 0; head[0, 0] = 1
        dz[: head.shape[0]] = torch.from_numpy(head).cuda()
        ctx.synth_witness(pk, dz.data_ptr(), n_free)
        zs[p] = dz.cpu().numpy()
        want.append(ctx.prove(pk, zs[p], rs[p], ss[p]))
    pvk = ctx.vk_prepare(vk)
    assert all(ctx.verify(pvk, zs[p][1:1 + n_in], want[p]) for p in range(n_part))
    pvk.free(); pk.free(); ctx.close()

    m = lib.MultiContext(devices)
    k = m.key_generate_synth(log_m, n_in, n_free, seed, toxic)
    got = m.prove_batch(k, zs, rs, ss)
    assert got == want                                       # partition order, bit-exact
    assert m.prove(k, zs[0], rs[0], ss[0]) == want[0]        # one proof over all contexts (latency mode)
    bad = zs[:1].copy(); bad[0, 3, :] = 0xFF                 # non-canonical scalar: error with the device named, no proof
    with pytest.raises(lib.B2g16Error) as e:
        m.prove_batch(k, bad, rs[:1], ss[:1])
    assert e.value.code == -2 and "device" in str(e.value)
    m.key_free(k)
    m.close()


@pytest.mark.gpu
def test_priority_proof_runs_between_partitions_of_a_batch(oracle):
    """compound_proof.hpp:134 `priority`: a PoSt proof submitted while a sector's partitions are being proved is admitted
    ahead of the batch's remaining partitions (one admission per partition), and both results are still bit-exact."""
    import threading
    import time
    import torch
    from b2g16 import codec
    log_m, n_in, n_free, seed, n_part = 16, 3, 8, 5, 6
    toxic, r, sr = util.toxic_from_seed(seed)
    ctx = lib.Context(0)
    pk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic)
    nv = pk.shape()["n_vars"]
    dz = torch.zeros((nv, 32), dtype=torch.uint8, device="cuda:0")
    head = codec.random_fr(1 + n_in + n_free, np.random.Generator(np.random.PCG64(seed))); head[0] = 0; head[0, 0] = 1
    dz[: head.shape[0]] = torch.from_numpy(head).cuda()
    ctx.synth_witness(pk, dz.data_ptr(), n_free)
    z = dz.cpu().numpy()
    want = ctx.prove(pk, z, r, sr)
    zs = np.stack([z] * n_part); rs = np.stack([r] * n_part); ss = np.stack([sr] * n_part)
    t = {}

    def batch():
        t["batch0"] = time.perf_counter()
        t["batch"] = ctx.prove_batch(pk, zs, rs, ss)
        t["batch1"] = time.perf_counter()

    th = threading.Thread(target=batch)
    th.start()
    while "batch0" not in t:
        time.sleep(0.0005)
    time.sleep(0.01)                                   # the batch is inside its first partitions
    got = ctx.prove(pk, z, r, sr, priority=True)
    t["hi1"] = time.perf_counter()
    th.join()
    assert got == want and t["batch"] == want * n_part
    assert t["hi1"] < t["batch1"], "the priority proof waited for the whole batch"
    pk.free(); ctx.close()
