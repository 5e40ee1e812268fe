// TEST INFRASTRUCTURE (not shipped, not linked into libmedvae_b200.so): a host-only interpreter of the "tap
// programs" that medvae_disentangled_multimodal_b200/csrc/conv_lower.h lowers the reference's three conv
// geometries to (encoder_decoder.py:128-146 nn.Conv2d s1, :184-188 Downsample, :205-209 Upsample) and their
// input / weight gradients.  tests/test_lowering_cpu.py runs it against torch's own conv + autograd, so the
// lowering both tcgen05 and SIMT launchers consume is checked without a GPU.  Compiled by the test with nvcc
// (host code only; the header pulls in CUDA types).
#include <vector>
#include "../medvae_disentangled_multimodal_b200/csrc/conv_lower.h"

using namespace mvae;

// a view element, or 0 outside the sub-grid (what TMA out-of-bounds fill / the SIMT bounds check deliver)
static inline double view_load(const float* t, const View& v, int n, int y, int x, int c) {
    if (y < 0 || y >= v.Hs || x < 0 || x >= v.Ws) return 0.0;
    return (double)t[v.off + (long long)n * v.sn + (long long)y * v.sy + (long long)x * v.sx + c];
}

// out[oview](n,y,x,:) = sum_t in[view_t](n, y+dy_t, x+dx_t, :) * W[slab_t]^T, W packed [slab][N][K]
static void run_gather(const GatherProblem& p, const float* in, const float* w, float* out) {
    for (int n = 0; n < p.B; ++n)
        for (int y = 0; y < p.OH; ++y)
            for (int x = 0; x < p.OW; ++x) {
                if (y >= p.out_view.Hs || x >= p.out_view.Ws) continue;
                float* o = out + p.out_view.off + (long long)n * p.out_view.sn + (long long)y * p.out_view.sy +
                           (long long)x * p.out_view.sx;
                for (int co = 0; co < p.N; ++co) {
                    double acc = 0.0;
                    for (int t = 0; t < p.taps.ntaps; ++t) {
                        const Tap& tp = p.taps.t[t];
                        const float* ws = w + ((long long)tp.slab * p.N + co) * p.K;
                        for (int k = 0; k < p.K; ++k)
                            acc += view_load(in, p.in_views[tp.view], n, y + tp.dy, x + tp.dx, k) * (double)ws[k];
                    }
                    o[co] = (float)acc;
                }
            }
}

extern "C" {

// problems / taps the lowering produced (for assertions on the shape of the program itself)
int emu_program_shape(const mvae_conv_desc* d, int which, int* nproblems, int* ntaps_total) {
    ConvGeom g;
    if (conv_geom(*d, g)) return 1;
    GatherProblem pr[4];
    int np = 0, nt = 0;
    if (which == 0) np = lower_fprop(g, pr);
    else if (which == 1) np = lower_dgrad(g, pr);
    else { ReduceProblem rp; lower_wgrad(g, rp); *nproblems = 1; *ntaps_total = rp.taps.ntaps; return 0; }
    for (int i = 0; i < np; ++i) nt += pr[i].taps.ntaps;
    *nproblems = np; *ntaps_total = nt;
    return 0;
}

// y (NHWC, P x Q x Cout) = conv(x) with packed weights [slab][Cout][Cin]
int emu_fprop(const mvae_conv_desc* d, const float* x, const float* w_fprop, float* y) {
    ConvGeom g;
    if (conv_geom(*d, g)) return 1;
    GatherProblem pr[4];
    const int np = lower_fprop(g, pr);
    for (int i = 0; i < np; ++i) run_gather(pr[i], x, w_fprop, y);
    return 0;
}

// dx (NHWC, H x W x Cin) from dy with packed weights [slab][Cin][Cout]
int emu_dgrad(const mvae_conv_desc* d, const float* dy, const float* w_dgrad, float* dx) {
    ConvGeom g;
    if (conv_geom(*d, g)) return 1;
    GatherProblem pr[4];
    const int np = lower_dgrad(g, pr);
    for (int i = 0; i < np; ++i) run_gather(pr[i], dy, w_dgrad, dx);
    return 0;
}

// dW per slab [slab][Cout][Cin] = sum_{n,y,x} dy[oview](n,y,x,co) * in[view](n, y+dy, x+dx, ci)
int emu_wgrad(const mvae_conv_desc* d, const float* x, const float* dy, float* dw_slabs) {
    ConvGeom g;
    if (conv_geom(*d, g)) return 1;
    ReduceProblem p;
    lower_wgrad(g, p);
    for (int t = 0; t < p.taps.ntaps; ++t) {
        const Tap& tp = p.taps.t[t];
        const View& dv = p.dy_views[tp.oview];
        std::vector<double> acc((size_t)p.Cout * p.Cin, 0.0);
        for (int n = 0; n < p.B; ++n)
            for (int y = 0; y < p.GH; ++y)
                for (int xx = 0; xx < p.GW; ++xx)
                    for (int co = 0; co < p.Cout; ++co) {
                        const double gdy = view_load(dy, dv, n, y, xx, co);
                        if (gdy == 0.0) continue;
                        for (int ci = 0; ci < p.Cin; ++ci)
                            acc[(size_t)co * p.Cin + ci] +=
                                gdy * view_load(x, p.in_views[tp.view], n, y + tp.dy, xx + tp.dx, ci);
                    }
        float* o = dw_slabs + (long long)tp.slab * p.Cout * p.Cin;
        for (size_t i = 0; i < acc.size(); ++i) o[i] = (float)acc[i];
    }
    return 0;
}

}  // extern "C"
