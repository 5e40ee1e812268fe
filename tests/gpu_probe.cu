// TEST INFRASTRUCTURE: a few-second device check of the reparameterisation noise addressing through the C ABI, no
// Python: the noise of a batch drawn in one call equals the noise of its two halves drawn with elem_base.
//   nvcc -o tests/_build/gpu_probe tests/gpu_probe.cu -Iinclude -Lmedvae_disentangled_multimodal_b200 -lmedvae_b200
#include <cstdio>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "medvae_b200.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); return 2; } } while (0)

int main() {
    const int B = 8, zc = 5, R = 3, Z = zc * R * R, split = 3;
    if (!mvae_device_ok()) { printf("no sm_100 device\n"); return 3; }
    float *mom, *eps, *mean, *lv, *z;
    CK(cudaMalloc(&mom, sizeof(float) * B * R * R * 2 * zc));
    CK(cudaMemset(mom, 0, sizeof(float) * B * R * R * 2 * zc));
    CK(cudaMalloc(&eps, sizeof(float) * 2 * B * Z));
    CK(cudaMalloc(&mean, sizeof(float) * B * Z));
    CK(cudaMalloc(&lv, sizeof(float) * B * Z));
    CK(cudaMalloc(&z, sizeof(float) * B * Z));
    const uint64_t seed = 99, offset = 7;
    int rc = mvae_reparam_kl_fwd(mom, NULL, eps, mean, lv, NULL, z, NULL, NULL, B, zc, R, 0.f, seed, offset, 0, NULL, MVAE_F32, NULL);
    rc |= mvae_reparam_kl_fwd(mom, NULL, eps + B * Z, mean, lv, NULL, z, NULL, NULL, split, zc, R, 0.f, seed, offset, 0, NULL, MVAE_F32, NULL);
    rc |= mvae_reparam_kl_fwd(mom, NULL, eps + B * Z + split * Z, mean, lv, NULL, z, NULL, NULL, B - split, zc, R, 0.f, seed, offset,
                              (uint64_t)split * Z, NULL, MVAE_F32, NULL);
    if (rc) { printf("call failed: %s\n", mvae_last_error()); return 4; }
    CK(cudaDeviceSynchronize());
    std::vector<float> h(2 * B * Z);
    CK(cudaMemcpy(h.data(), eps, sizeof(float) * 2 * B * Z, cudaMemcpyDeviceToHost));
    const bool same = memcmp(h.data(), h.data() + B * Z, sizeof(float) * B * Z) == 0;
    const bool distinct = memcmp(h.data(), h.data() + split * Z, sizeof(float) * split * Z) != 0;
    double s = 0, s2 = 0;
    for (int i = 0; i < B * Z; ++i) { s += h[i]; s2 += (double)h[i] * h[i]; }
    printf("gpu_probe: split-independent %d, halves distinct %d, mean %.3f, var %.3f, eps[0..3] %.6f %.6f %.6f %.6f\n", same, distinct,
           s / (B * Z), s2 / (B * Z), h[0], h[1], h[2], h[3]);
    return (same && distinct) ? 0 : 1;
}
