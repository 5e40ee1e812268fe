// Host-side lowering of the reference's three conv geometries (encoder_decoder.py:128-146
// nn.Conv2d s1, :184-188 Downsample, :205-209 Upsample) and their input/weight gradients to
// "tap programs" over strided NHWC views.  Shared by the SIMT and the tcgen05 launchers.
#pragma once
#include "../../include/medvae_b200.h"
#include "common.cuh"

namespace mvae {

struct ConvGeom {
    int B, H, W, Cin, Cout, R, mode;
    int P, Q;   // output spatial size
    int pad;    // (R-1)/2 for S1
};

inline int conv_geom(const mvae_conv_desc& d, ConvGeom& g) {
    g.B = d.batch; g.H = d.in_h; g.W = d.in_w; g.Cin = d.in_c; g.Cout = d.out_c; g.R = d.ksize;
    g.mode = d.mode; g.pad = (d.ksize - 1) / 2;
    if (d.batch <= 0 || d.in_h <= 0 || d.in_w <= 0 || d.in_c <= 0 || d.out_c <= 0) return 1;
    if (d.ksize != 1 && d.ksize != 3) return 1;
    if (d.mode == MVAE_CONV_S1) { g.P = g.H; g.Q = g.W; }
    else if (d.mode == MVAE_CONV_DOWN) {
        if (d.ksize != 3) return 1;
        g.P = (g.H + 1 - 3) / 2 + 1; g.Q = (g.W + 1 - 3) / 2 + 1;
    } else if (d.mode == MVAE_CONV_UP) {
        if (d.ksize != 3) return 1;
        g.P = 2 * g.H; g.Q = 2 * g.W;
    } else return 1;
    return 0;
}

inline int conv_num_slabs(const ConvGeom& g) { return g.mode == MVAE_CONV_UP ? 16 : g.R * g.R; }

inline View full_view(int H, int W, int C) {
    View v; v.off = 0; v.Hs = H; v.Ws = W;
    v.sn = (long long)H * W * C; v.sy = (long long)W * C; v.sx = C; return v;
}
// sub-grid {(2i+a, 2j+b)} of an H x W x C tensor
inline View parity_view(int H, int W, int C, int a, int b) {
    View v; v.off = ((long long)a * W + b) * C; v.Hs = (H - a + 1) / 2; v.Ws = (W - b + 1) / 2;
    v.sn = (long long)H * W * C; v.sy = 2LL * W * C; v.sx = 2LL * C; return v;
}
inline int up_slab(int a, int b, int ui, int vi) { return ((a * 2 + b) * 2 + ui) * 2 + vi; }

// ---- y = conv(x): 1 problem (S1, DOWN) or 4 parity problems (UP)
inline int lower_fprop(const ConvGeom& g, GatherProblem* out) {
    if (g.mode == MVAE_CONV_S1) {
        GatherProblem& p = out[0];
        p.B = g.B; p.OH = g.P; p.OW = g.Q; p.K = g.Cin; p.N = g.Cout; p.nviews = 1;
        p.in_views[0] = full_view(g.H, g.W, g.Cin);
        p.out_view = full_view(g.P, g.Q, g.Cout);
        p.taps.ntaps = g.R * g.R;
        for (int r = 0; r < g.R; ++r) for (int s = 0; s < g.R; ++s) {
            Tap& t = p.taps.t[r * g.R + s];
            t.view = 0; t.dy = r - g.pad; t.dx = s - g.pad; t.slab = r * g.R + s; t.oview = 0;
        }
        return 1;
    }
    if (g.mode == MVAE_CONV_DOWN) {
        // y(p,q) = sum W[r,s] x(2p+r, 2q+s), rows/cols == H/W read as zero (the (0,1,0,1) pad)
        GatherProblem& p = out[0];
        p.B = g.B; p.OH = g.P; p.OW = g.Q; p.K = g.Cin; p.N = g.Cout; p.nviews = 4;
        for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
            p.in_views[a * 2 + b] = parity_view(g.H, g.W, g.Cin, a, b);
        p.out_view = full_view(g.P, g.Q, g.Cout);
        p.taps.ntaps = 9;
        for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) {
            Tap& t = p.taps.t[r * 3 + s];
            t.view = (r & 1) * 2 + (s & 1); t.dy = r >> 1; t.dx = s >> 1; t.slab = r * 3 + s; t.oview = 0;
        }
        return 1;
    }
    // UP: out(2i+a, 2j+b) = sum_{ui,vi} Weff[a,b,ui,vi] x(i+u, j+v), u = ui-1+a, v = vi-1+b
    for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b) {
        GatherProblem& p = out[a * 2 + b];
        p.B = g.B; p.OH = g.H; p.OW = g.W; p.K = g.Cin; p.N = g.Cout; p.nviews = 1;
        p.in_views[0] = full_view(g.H, g.W, g.Cin);
        p.out_view = parity_view(g.P, g.Q, g.Cout, a, b);
        p.taps.ntaps = 4;
        for (int ui = 0; ui < 2; ++ui) for (int vi = 0; vi < 2; ++vi) {
            Tap& t = p.taps.t[ui * 2 + vi];
            t.view = 0; t.dy = ui - 1 + a; t.dx = vi - 1 + b; t.slab = up_slab(a, b, ui, vi); t.oview = 0;
        }
    }
    return 4;
}

// ---- dx = conv^T(dy): the gather runs over dy with K = Cout, N = Cin
inline int lower_dgrad(const ConvGeom& g, GatherProblem* out) {
    if (g.mode == MVAE_CONV_S1) {
        GatherProblem& p = out[0];
        p.B = g.B; p.OH = g.H; p.OW = g.W; p.K = g.Cout; p.N = g.Cin; p.nviews = 1;
        p.in_views[0] = full_view(g.P, g.Q, g.Cout);
        p.out_view = full_view(g.H, g.W, g.Cin);
        p.taps.ntaps = g.R * g.R;
        for (int r = 0; r < g.R; ++r) for (int s = 0; s < g.R; ++s) {
            Tap& t = p.taps.t[r * g.R + s];
            t.view = 0; t.dy = g.pad - r; t.dx = g.pad - s; t.slab = r * g.R + s; t.oview = 0;
        }
        return 1;
    }
    if (g.mode == MVAE_CONV_DOWN) {
        // dx(2i+a, 2j+b) = sum_{r%2==a, s%2==b} dy(i - r/2, j - s/2) W[r,s]^T
        for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b) {
            GatherProblem& p = out[a * 2 + b];
            View ov = parity_view(g.H, g.W, g.Cin, a, b);
            p.B = g.B; p.OH = ov.Hs; p.OW = ov.Ws; p.K = g.Cout; p.N = g.Cin; p.nviews = 1;
            p.in_views[0] = full_view(g.P, g.Q, g.Cout);
            p.out_view = ov;
            int n = 0;
            for (int r = a; r < 3; r += 2) for (int s = b; s < 3; s += 2) {
                Tap& t = p.taps.t[n++];
                t.view = 0; t.dy = -(r >> 1); t.dx = -(s >> 1); t.slab = r * 3 + s; t.oview = 0;
            }
            p.taps.ntaps = n;
        }
        return 4;
    }
    // UP: dx(i,j) = sum_{a,b,ui,vi} dy(2(i-u)+a, 2(j-v)+b) Weff[a,b,ui,vi]^T
    GatherProblem& p = out[0];
    p.B = g.B; p.OH = g.H; p.OW = g.W; p.K = g.Cout; p.N = g.Cin; p.nviews = 4;
    for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
        p.in_views[a * 2 + b] = parity_view(g.P, g.Q, g.Cout, a, b);
    p.out_view = full_view(g.H, g.W, g.Cin);
    int n = 0;
    for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
        for (int ui = 0; ui < 2; ++ui) for (int vi = 0; vi < 2; ++vi) {
            Tap& t = p.taps.t[n++];
            t.view = a * 2 + b; t.dy = -(ui - 1 + a); t.dx = -(vi - 1 + b);
            t.slab = up_slab(a, b, ui, vi); t.oview = 0;
        }
    p.taps.ntaps = n;
    return 1;
}

// ---- dW: one tap per packed slab
inline void lower_wgrad(const ConvGeom& g, ReduceProblem& p) {
    p.B = g.B; p.Cout = g.Cout; p.Cin = g.Cin;
    if (g.mode == MVAE_CONV_S1) {
        p.GH = g.P; p.GW = g.Q; p.n_inviews = 1; p.n_dyviews = 1;
        p.in_views[0] = full_view(g.H, g.W, g.Cin);
        p.dy_views[0] = full_view(g.P, g.Q, g.Cout);
        p.taps.ntaps = g.R * g.R;
        for (int r = 0; r < g.R; ++r) for (int s = 0; s < g.R; ++s) {
            Tap& t = p.taps.t[r * g.R + s];
            t.view = 0; t.oview = 0; t.dy = r - g.pad; t.dx = s - g.pad; t.slab = r * g.R + s;
        }
    } else if (g.mode == MVAE_CONV_DOWN) {
        p.GH = g.P; p.GW = g.Q; p.n_inviews = 4; p.n_dyviews = 1;
        for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
            p.in_views[a * 2 + b] = parity_view(g.H, g.W, g.Cin, a, b);
        p.dy_views[0] = full_view(g.P, g.Q, g.Cout);
        p.taps.ntaps = 9;
        for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) {
            Tap& t = p.taps.t[r * 3 + s];
            t.view = (r & 1) * 2 + (s & 1); t.oview = 0; t.dy = r >> 1; t.dx = s >> 1; t.slab = r * 3 + s;
        }
    } else {
        p.GH = g.H; p.GW = g.W; p.n_inviews = 1; p.n_dyviews = 4;
        p.in_views[0] = full_view(g.H, g.W, g.Cin);
        for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
            p.dy_views[a * 2 + b] = parity_view(g.P, g.Q, g.Cout, a, b);
        int n = 0;
        for (int a = 0; a < 2; ++a) for (int b = 0; b < 2; ++b)
            for (int ui = 0; ui < 2; ++ui) for (int vi = 0; vi < 2; ++vi) {
                Tap& t = p.taps.t[n++];
                t.view = 0; t.oview = a * 2 + b; t.dy = ui - 1 + a; t.dx = vi - 1 + b;
                t.slab = up_slab(a, b, ui, vi);
            }
        p.taps.ntaps = n;
    }
}

// Optional GroupNorm fusion in the epilogue of the tcgen05 gather kernel (conv_umma.cu):
//   mode 1 (fprop):  sums[n][g] += (sum y, sum y^2) over the stored outputs: the statistics pass of
//                    the GroupNorm that consumes y (encoder_decoder.py:28-33) costs no HBM pass.
//   mode 2 (dgrad):  the accumulator is dL/da with a = [silu](GN(x)); the epilogue turns it into
//                    du = dL/du (u = xhat*gamma + beta) using x and (mean, rstd), stores du instead, and
//                    accumulates sums[n][g] += (sum du*gamma, sum du*gamma*xhat): the reduction pass
//                    of the GroupNorm backward.  x is addressed like the output.
struct GnHook {
    int mode;
    int G, cpg_log2, silu;
    float* sums;          // [B][G][2], accumulated with fp32 atomics (caller zeroes)
    const void* x;        // mode 2
    const float* stat;    // mode 2: [B][G][2] (mean, rstd)
    const float* gamma;   // mode 2
    const float* beta;    // mode 2
};

// launchers (conv_simt.cu / conv_umma.cu)
int simt_gather(const GatherProblem& p, const void* in, const void* w, const float* bias,
                const void* residual, void* out, int dtype, cudaStream_t st);
int simt_reduce(const ReduceProblem& p, const void* in, const void* dy, float* partial,
                int dtype, cudaStream_t st);
bool umma_gather_ok(const GatherProblem& p);
int umma_gather(const GatherProblem& p, const void* in, const void* w, int nslabs, const float* bias,
                const void* residual, void* out, cudaStream_t st, const GnHook* hook = nullptr);
bool umma_reduce_ok(const ReduceProblem& p);
int umma_reduce_splits(const ReduceProblem& p);
int umma_reduce(const ReduceProblem& p, const void* in, const void* dy, float* partial, int splits,
                cudaStream_t st);

}  // namespace mvae
