/* TEST INFRASTRUCTURE: a plain C99 host of the C ABI — what a maintainer binding libmedvae_b200.so from C (or through
 * cgo / JNI / any FFI that consumes C headers) compiles.  Built and run by tests/test_abi.py; no GPU needed: it only
 * calls the host-side helpers and checks that a compute entry point refuses to run without a device. */
#include <stdio.h>
#include <string.h>
#include "medvae_b200.h"

int main(void) {
    mvae_conv_desc d = {2, 28, 28, 8, 8, 3, MVAE_CONV_DOWN, MVAE_F32, MVAE_ENGINE_AUTO};
    int32_t oh = 0, ow = 0;
    if (mvae_abi_version() != MVAE_ABI_VERSION) return 1;
    if (mvae_conv2d_out_hw(&d, &oh, &ow) != 0 || oh != 14 || ow != 14) return 2;   /* pad (0,1,0,1) + 3x3 s2 */
    if (mvae_conv2d_num_slabs(&d) != 9) return 3;
    d.mode = MVAE_CONV_UP;
    if (mvae_conv2d_out_hw(&d, &oh, &ow) != 0 || oh != 56 || mvae_conv2d_num_slabs(&d) != 16) return 4;
    d.ksize = 5;
    if (mvae_conv2d_out_hw(&d, &oh, &ow) == 0 || strstr(mvae_last_error(), "bad conv descriptor") == NULL) return 5;
    if (!mvae_device_ok()) {
        /* no device: a compute call must fail with a message, never fall back */
        float x[4] = {0, 0, 0, 0}, y[4];
        if (mvae_nhwc_to_nchw_f32(x, y, 1, 2, 2, 1, 0, 1, MVAE_F32, NULL) == 0) return 6;
        if (strlen(mvae_last_error()) == 0) return 7;
    }
    printf("c host ok: abi %d, device_ok %d\n", mvae_abi_version(), mvae_device_ok());
    return 0;
}
