// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11), the counter-based generator
// behind the reparameterisation noise (loss.cu; replaces torch.randn_like at base_vae.py:85) and the dropout keep
// masks (feed.cu; nn.Dropout at encoder_decoder.py:164).  Host + device so the shipped arithmetic can be checked
// against the published known-answer vectors without a GPU (tests/philox_emu.cu, tests/test_philox_cpu.py).
#pragma once
#include <stdint.h>

namespace mvae {

struct PhiloxWords {
    uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ uint32_t philox_mulhi(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

// one block: 128-bit counter {c0, c1, c2, c3}, 64-bit key {k0, k1}, ten rounds, key bumped by the Weyl constants
__host__ __device__ __forceinline__ PhiloxWords philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                              uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = philox_mulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = philox_mulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    PhiloxWords o = {c0, c1, c2, c3};
    return o;
}

// key = 64-bit seed; counter = {ctr_lo (element / block index), ctr_hi (stream offset)}
struct Philox {
    uint32_t k0, k1;
    __host__ __device__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
    __host__ __device__ __forceinline__ PhiloxWords operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
        return philox4x32_10((uint32_t)ctr_lo, (uint32_t)(ctr_lo >> 32), (uint32_t)ctr_hi, (uint32_t)(ctr_hi >> 32), k0, k1);
    }
};

// uniform in (0, 1) from the top 24 bits (never 0: safe under the logarithm of Box-Muller)
__host__ __device__ __forceinline__ float philox_u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }
// uniform in [0, 1) on the 2^-24 grid: what the dropout keep test compares with p
__host__ __device__ __forceinline__ float philox_u01_grid(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// the two words element `e` of a noise stream draws from block e >> 1 of (seed, offset): (x, y) for even e, (z, w) for odd
__host__ __device__ __forceinline__ void philox_normal_words(uint64_t seed, uint64_t offset, uint64_t e, uint32_t& a,
                                                             uint32_t& b) {
    const PhiloxWords r = Philox(seed)(e >> 1, offset);
    a = (e & 1) ? r.z : r.x;
    b = (e & 1) ? r.w : r.y;
}

}  // namespace mvae
