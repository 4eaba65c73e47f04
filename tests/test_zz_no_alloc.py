"""b2g16_fr_dot_device_noalloc (csrc/no_alloc.cu): the entry point that allocates nothing -- its scratch is a file-scope
device array -- against the existing b2g16_fr_dot_device, through the C ABI.  No counterpart in the reference.
Integer arithmetic: the tolerance is the one every parity test here uses, bit-exact.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from b2g16 import codec, lib

pytestmark = pytest.mark.gpu
R = codec.R_MOD


def _rand_fr(torch, n, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    x = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device="cuda", generator=g)
    x[:, 31] &= 0x3F                      # < 2^254 < r : canonical
    return x


def test_no_alloc_entry_point_matches_the_allocating_one_twice_in_a_row():
    import torch
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)                                  # torch work and the library share one stream
    ctx = lib.Context(0, ts.cuda_stream)
    try:
        out = torch.zeros(32, dtype=torch.uint8, device="cuda")
        status = torch.full((1,), -99, dtype=torch.int32, device="cuda")
        # two calls in a row with different sizes and data: the second finds the first call's partial sums and flag
        # in the shared scratch (sizes: several blocks + ragged tail; then fewer blocks than before)
        for n, seed in ((3 * 4096 + 77, 11), (4096 + 5, 12), (1, 13), (0, 14)):
            a, b = _rand_fr(torch, max(n, 1), seed)[:n], _rand_fr(torch, max(n, 1), seed + 100)[:n]
            a, b = a.contiguous(), b.contiguous()
            want = ctx.fr_dot_device(a.data_ptr(), b.data_ptr(), n)         # existing entry point (allocates, synchronises)
            out.fill_(0xEE); status.fill_(-99)
            ctx.fr_dot_device_noalloc(a.data_ptr(), b.data_ptr(), n, out.data_ptr(), status.data_ptr())
            torch.cuda.synchronize()
            got = bytes(out.cpu().numpy())
            assert got == bytes(want), (n, got.hex(), bytes(want).hex())
            assert int(status.item()) == 0
            if 0 < n <= 5000:                                               # and against big-integer arithmetic
                ai, bi = codec.fr_to_ints(a.cpu().numpy()), codec.fr_to_ints(b.cpu().numpy())
                assert int.from_bytes(got, "little") == sum(x * y for x, y in zip(ai, bi)) % R
        # a non-canonical operand sets the status word; the next call clears the flag again
        n = 2 * 4096
        a, b = _rand_fr(torch, n, 21), _rand_fr(torch, n, 22)
        bad = a.clone(); bad[4097] = 0xFF
        ctx.fr_dot_device_noalloc(bad.data_ptr(), b.data_ptr(), n, out.data_ptr(), status.data_ptr())
        torch.cuda.synchronize()
        assert int(status.item()) == -2                                     # B2G16_E_NOT_CANONICAL
        ctx.fr_dot_device_noalloc(a.data_ptr(), b.data_ptr(), n, out.data_ptr(), status.data_ptr())
        torch.cuda.synchronize()
        assert int(status.item()) == 0 and bytes(out.cpu().numpy()) == bytes(ctx.fr_dot_device(a.data_ptr(), b.data_ptr(), n))
        # beyond the stated maximum: refused before anything is launched (the pointers are never dereferenced)
        with pytest.raises(lib.B2g16Error) as e:
            ctx.fr_dot_device_noalloc(a.data_ptr(), b.data_ptr(), (1 << 24) + 1, out.data_ptr(), status.data_ptr())
        assert e.value.code == -1 and "B2G16_NOALLOC_MAX_N" in str(e.value)
    finally:
        ctx.close()
        torch.cuda.set_stream(torch.cuda.default_stream())
