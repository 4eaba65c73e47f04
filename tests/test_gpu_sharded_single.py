"""The kernels behind the multi-GPU paths, exercised on ONE GPU so the single-GPU tier covers them: the ranks of
a sharded transform are emulated by separate buffers on the same device (the all-to-all exchanges become tensor
indexing, the peer pointers of the fused cross step point at local buffers), and a proof is assembled from
partial multiexp sets computed one after the other.  Expected values: the library's own single-call results,
which tests/test_gpu_parity.py pins to the oracle.  The real N-rank runs are tests/test_gpu_multi.py."""
import numpy as np
import pytest
import torch

import util
from b2g16 import codec, lib

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def oracle():
    from oracle import coracle
    coracle.build()
    return coracle


def _full(ctx, x, log_n, inverse, coset):
    f = x.clone()
    ctx.ntt_device(f.data_ptr(), log_n, inverse=inverse, coset=coset)
    ctx.synchronize()
    return f


@pytest.mark.parametrize("log_g", [1, 2, 3])
@pytest.mark.parametrize("log_n", [6, 13, 18])
def test_cross_step_emulated_ranks(ctx, oracle, log_g, log_n):
    G, n = 1 << log_g, 1 << log_n
    M, cols = n >> log_g, n >> (2 * log_g)
    x = torch.from_numpy(codec.random_fr(n, np.random.Generator(np.random.PCG64(log_n * 8 + log_g)))).cuda()
    if log_n <= 13:                                                   # anchor the expected value to the oracle once
        assert np.array_equal(_full(ctx, x, log_n, False, False).cpu().numpy(), oracle.ntt(x.cpu().numpy()))
    blocks = [x[r * M:(r + 1) * M].contiguous() for r in range(G)]
    for coset in (False, True):
        want = _full(ctx, x, log_n, False, coset)
        xs = [b.clone() for b in blocks]
        if coset:
            for r in range(G):
                ctx.fr_coset_scale_device(xs[r].data_ptr(), M, log_n, r * M, False)
        ys = []
        m2 = []
        for c in range(G):                                            # "rank" c: its column chunk of every block
            m1 = torch.stack([xs[r].view(G, cols, 32)[c] for r in range(G)]).contiguous()
            out = torch.empty_like(m1)
            ctx.ntt_cross_device(m1.data_ptr(), out.data_ptr(), log_n, log_g, c, False, coset)
            m2.append(out)
        for k1 in range(G):
            y = torch.stack([m2[c][k1] for c in range(G)]).contiguous().view(M, 32)
            ctx.ntt_device(y.data_ptr(), log_n - log_g)
            ctx.synchronize()
            assert torch.equal(y, want[k1::G].contiguous()), (coset, k1)
            ys.append(y)
        # fused form: the same kernel with per-rank base pointers (here all local), forward then inverse
        yp = [torch.empty((M, 32), dtype=torch.uint8, device="cuda") for _ in range(G)]
        for c in range(G):
            ctx.ntt_cross_p2p_device([t.data_ptr() for t in xs], [t.data_ptr() for t in yp], log_n, log_g, c, False, coset)
        for k1 in range(G):
            ctx.ntt_device(yp[k1].data_ptr(), log_n - log_g)
        ctx.synchronize()
        assert all(torch.equal(yp[k], ys[k]) for k in range(G))
        for k1 in range(G):
            ctx.ntt_device(yp[k1].data_ptr(), log_n - log_g, inverse=True)
        xb = [torch.empty((M, 32), dtype=torch.uint8, device="cuda") for _ in range(G)]
        for c in range(G):
            ctx.ntt_cross_p2p_device([t.data_ptr() for t in yp], [t.data_ptr() for t in xb], log_n, log_g, c, True, coset)
        for r in range(G):
            if coset:
                ctx.fr_coset_scale_device(xb[r].data_ptr(), M, log_n, r * M, True)
        ctx.synchronize()
        assert all(torch.equal(xb[r], blocks[r]) for r in range(G)), ("inverse", coset)
        # natural block order of the spectrum: chunk rho of every rank's cyclic block, transposed
        for rho in range(G):
            z = torch.stack([ys[k1].view(G, cols, 32)[rho] for k1 in range(G)]).contiguous().view(G * cols, 32)
            nat = torch.empty_like(z)
            ctx.fr_transpose_device(z.data_ptr(), nat.data_ptr(), G, cols)
            ctx.synchronize()
            assert torch.equal(nat, want[rho * M:(rho + 1) * M]), ("natural", rho)


def test_cross_step_rejects_bad_arguments(ctx):
    a = torch.zeros((64, 32), dtype=torch.uint8, device="cuda")
    b = torch.empty_like(a)
    for args in ((6, 4, 0), (3, 2, 0), (6, 2, 4)):                     # 16 ranks; 2^3 < 4^2; chunk >= G
        with pytest.raises(lib.B2g16Error):
            ctx.ntt_cross_device(a.data_ptr(), b.data_ptr(), args[0], args[1], args[2])
    bad = torch.full((64, 32), 0xFF, dtype=torch.uint8, device="cuda")   # >= r: not canonical
    with pytest.raises(lib.B2g16Error):
        ctx.ntt_cross_device(bad.data_ptr(), b.data_ptr(), 6, 1, 0)


@pytest.mark.parametrize("parts", [1, 3, 8])
def test_proof_from_partial_sets(ctx, oracle, parts):
    s, cs, z = util.synth_instance(oracle, 11, 5, seed=91)
    toxic, r, sr = util.toxic_from_seed(91)
    pk = ctx.pk_generate(cs, toxic)
    want = ctx.prove(pk, z, r, sr)
    okey = oracle.groth16_setup(cs, toxic)
    assert want == oracle.groth16_prove(okey, z, r, sr, want_trace=False)[0]
    dz = torch.from_numpy(z).cuda()
    partials = np.stack([ctx.prove_partial(pk, dz.data_ptr(), k, parts) for k in range(parts)])
    assert ctx.prove_assemble(pk, partials, r, sr) == want
    assert ctx.prove_assemble(pk, partials[::-1].copy(), r, sr) == want          # any order
    with pytest.raises(lib.B2g16Error):
        ctx.prove_partial(pk, dz.data_ptr(), parts, parts)                       # part out of range
    pk.free()


def test_witness_map_pieces_single_gpu(ctx, oracle):
    """The device steps of b2g16.dist.prove_sharded_full with one 'rank': strided constraint evaluation, the
    transform chain through the public transform entry point, the pointwise quotient, and the partial prove that
    takes the H coefficients from outside -- every intermediate against the oracle's trace."""
    s, cs, z = util.synth_instance(oracle, 10, 3, seed=17)
    toxic, r, sr = util.toxic_from_seed(17)
    pk = ctx.pk_generate(cs, toxic)
    okey = oracle.groth16_setup(cs, toxic)
    want, tr = oracle.groth16_prove(okey, z, r, sr, want_trace=True)
    m, log_m = 1 << s.log_m, s.log_m
    dz = torch.from_numpy(z).cuda()
    a, b, c = (torch.empty((m, 32), dtype=torch.uint8, device="cuda") for _ in range(3))
    ctx.abc_evals_strided_device(pk, dz.data_ptr(), 0, 1, m, a.data_ptr(), b.data_ptr(), c.data_ptr())
    ctx.synchronize()
    for got, name in ((a, "a"), (b, "b"), (c, "c")):
        assert np.array_equal(got.cpu().numpy(), tr[name]), name
    # two interleaved halves (what two ranks would compute) give the same values
    half = [tuple(torch.empty((m // 2, 32), dtype=torch.uint8, device="cuda") for _ in range(3)) for _ in range(2)]
    for k1 in range(2):
        ctx.abc_evals_strided_device(pk, dz.data_ptr(), k1, 2, m // 2, *(t.data_ptr() for t in half[k1]))
    ctx.synchronize()
    assert torch.equal(half[0][0], a[0::2].contiguous()) and torch.equal(half[1][2], c[1::2].contiguous())
    for v in (a, b, c):
        ctx.ntt_device(v.data_ptr(), log_m, inverse=True)
        ctx.ntt_device(v.data_ptr(), log_m, inverse=False, coset=True)
    h = torch.empty((m, 32), dtype=torch.uint8, device="cuda")
    ctx.h_pointwise_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), m, log_m, h.data_ptr())
    ctx.ntt_device(h.data_ptr(), log_m, inverse=True, coset=True)
    ctx.synchronize()
    assert np.array_equal(h.cpu().numpy(), tr["h"])
    part = ctx.prove_partial_h(pk, dz.data_ptr(), 0, 1, h.data_ptr(), 0, m)
    assert ctx.prove_assemble(pk, part, r, sr) == want
    # two H blocks with the other four queries split in two as well
    parts = np.stack([ctx.prove_partial_h(pk, dz.data_ptr(), k, 2, h[k * (m // 2):].data_ptr(), k * (m // 2), m // 2) for k in range(2)])
    assert ctx.prove_assemble(pk, parts, r, sr) == want
    bad = h.clone(); bad[5] = 0xFF
    with pytest.raises(lib.B2g16Error):
        ctx.prove_partial_h(pk, dz.data_ptr(), 0, 1, bad.data_ptr(), 0, m)
    pk.free()
