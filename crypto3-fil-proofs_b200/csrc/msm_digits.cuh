// Signed-digit window decomposition of canonical Fr scalars for the Pippenger MSM
// (host + device; the host build is exercised by tests/host_emu).
#pragma once
#include "fp.cuh"

namespace b2 {

B2_HD int msm_windows(int c) { return (256 + c - 1) / c; }
B2_HD uint32_t msm_bit(const uint32_t* s, int i) { return (i < 0 || i >= 256) ? 0u : (s[i >> 5] >> (i & 31)) & 1u; }
// digit of window w: u_w + bit(wc-1) - 2^c bit(wc+c-1); sum_w d_w 2^(cw) = s for s < 2^255
B2_HD int msm_signed_digit(const uint32_t* s, int w, int c) {
    const int lo = w * c;
    const int i = lo >> 5, sh = lo & 31;
    uint32_t u = 0;
    if (i < 8) {
        u = s[i] >> sh;
        if (sh && i + 1 < 8) u |= s[i + 1] << (32 - sh);
    }
    u &= (1u << c) - 1;
    return (int)u + (int)msm_bit(s, lo - 1) - (int)(msm_bit(s, lo + c - 1) << c);
}

// Bit-decomposition tail of the bucket reduction (msm_reduce_bits_kernel): the j-th index, in increasing order, whose
// bit b is set -- the high part of j moves up by one position and bit b is forced on.
B2_HD uint32_t msm_bits_index(uint32_t j, int b) {
    const uint32_t low = (1u << b) - 1u;
    return ((j >> b) << (b + 1)) | (1u << b) | (j & low);
}

// chunk partition of the 2^(c-1) buckets of a window for the running-sum reduction
B2_HD uint32_t msm_reduce_chunks(int c) {
    uint32_t B = 1u << (c - 1), T = B / 8;
    if (T > 2048) T = 2048;
    if (T < 1) T = 1;
    return T;
}

}  // namespace b2
