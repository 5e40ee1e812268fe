// TEST INFRASTRUCTURE (not shipped): host-side calls into medvae_disentangled_multimodal_b200/csrc/philox.cuh — the
// very header the reparameterisation (loss.cu) and dropout (feed.cu) kernels compile — so the shipped Philox4x32-10
// arithmetic is checked against the published known-answer vectors and the numpy restatement (oracle/philox_oracle.py)
// without a GPU.  Built by tests/test_philox_cpu.py with nvcc (host code only).
#include "../medvae_disentangled_multimodal_b200/csrc/philox.cuh"

using namespace mvae;

extern "C" {

void emu_philox_raw(const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
    const PhiloxWords r = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

void emu_philox_block(uint64_t seed, uint64_t ctr_lo, uint64_t ctr_hi, uint32_t* out) {
    const PhiloxWords r = Philox(seed)(ctr_lo, ctr_hi);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

// dropout keep mask exactly as feed.cu:dropout_kernel derives it: word j of block e / 4 of stream (seed, offset)
void emu_keep_mask(long long n, float p, uint64_t seed, uint64_t offset, uint8_t* mask) {
    const Philox ph(seed);
    for (long long q = 0; q < (n + 3) / 4; ++q) {
        const PhiloxWords r = ph((uint64_t)q, offset);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
        for (int j = 0; j < 4 && q * 4 + j < n; ++j) mask[q * 4 + j] = philox_u01_grid(w[j]) >= p ? 1 : 0;
    }
}

// the two uniforms element e of a noise stream feeds to Box-Muller (loss.cu:philox_normal), as fp32
void emu_normal_uniforms(long long n, uint64_t seed, uint64_t offset, float* u1, float* u2) {
    for (long long e = 0; e < n; ++e) {
        uint32_t a, b;
        philox_normal_words(seed, offset, (uint64_t)e, a, b);
        u1[e] = philox_u01(a);
        u2[e] = philox_u01(b);
    }
}

}  // extern "C"
