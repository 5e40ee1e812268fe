// extern "C" surface of libmedvae_b200.so (see include/medvae_b200.h).  Argument validation,
// engine selection, error reporting.  No CPU fallback anywhere: without a device every compute
// entry fails with an error string.
#include <cstdarg>
#include <cstring>

#include "conv_lower.h"

namespace mvae {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static unsigned long long g_launches = 0;
void count_launch() { __atomic_fetch_add(&g_launches, 1ull, __ATOMIC_RELAXED); }
static int g_knobs[32] = {0};
int debug_knob(int key) { return (key >= 0 && key < 32) ? g_knobs[key] : 0; }

// implemented in the other translation units
int pack_weights(const ConvGeom& g, const float* w, void* wf, void* wd, int Ctot, int co_off, int dtype, cudaStream_t st);
int pack_weights_batched(const mvae_pack_job* jobs_dev, int n_jobs, int total_tiles, int dtype, cudaStream_t st);
int wgrad_finalize(const ConvGeom& g, const float* partial, float* dw, int splits, float beta, cudaStream_t st);
int colsum_parts(long long M);
int colsum(const void* dy, float* db, float* scratch, long long M, int C, int dtype, float beta, cudaStream_t st);
int nhwc_to_nchw(const void* src, float* dst, int B, int HW, int Cs, int c_off, int Cd, int dtype, cudaStream_t st);
int nchw_to_nhwc(const float* src, void* dst, int B, int HW, int C, int dtype, cudaStream_t st);
int input_pack_fwd(const float* x, const float* cond, const float* w, const float* b, void* out, int B, int C, int H,
                   int W, int cond_dim, int out_c, int dtype, cudaStream_t st);
int input_pack_bwd(const void* d_in, const float* cond, const float* w, const float* b, float* dw, float* db, int B,
                   int C, int H, int W, int cond_dim, int out_c, int dtype, cudaStream_t st);
int dis_input_fwd(const float* x, const long long* idx, const float* pw, const float* pb, const int* is_gray, void* out,
                  int B, int xc, int HW, int n_mod, int out_c, int dtype, cudaStream_t st);
int dis_input_bwd(const void* d_in, const float* x, const long long* idx, const int* is_gray, float* dpw, float* dpb,
                  int B, int xc, int HW, int n_mod, int out_c, int dtype, cudaStream_t st);
int dis_head_fwd(const void* r, int r_stride, const long long* idx, const float* w1, const float* b1, const float* w2,
                 const float* b2, const float* pw, const float* pb, const int* is_gray, float* hid, float* y, float* out,
                 int B, int H, int W, int n_mod, int out_c, int dtype, cudaStream_t st);
int dis_head_bwd(const float* d_out, const void* r, int r_stride, const float* hid, const float* y, const long long* idx,
                 const float* w1, const float* w2, const float* pw, const int* is_gray, void* d_r, float* gflat,
                 float* scratch, int B, int H, int W, int n_mod, int out_c, int dtype, cudaStream_t st);
long long dis_aux_floats(int B, int m, int n_mod);
int dis_aux_fwd(const float* z, const long long* idx, float* aux, int B, long long D, int s0, int m, int n_mod, float tau,
                cudaStream_t st);
int dis_aux_bwd(const float* z, const long long* idx, const float* aux, const float* g_sep, const float* g_con, float* dz,
                int B, long long D, int s0, int m, int n_mod, float tau, cudaStream_t st);
int dis_total(const float* t0, const float* t1, const float* t2, const float* t3, float w0, float w1, float w2, float w3,
              float* out, float* mask, cudaStream_t st);
int gn_splits(int B, int HW, int C);
long long gn_scratch_floats(int B, int HW, int C, int G, int backward);
int gn_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, float* scratch, int B, int HW,
           int C, int G, float eps, int silu, int dtype, cudaStream_t st);
int gn_bwd(const void* dy, const void* x, const float* gamma, const float* beta, const float* stats, const void* dx_add,
           void* dx, float* dgamma, float* dbeta, float* scratch, int B, int HW, int C, int G, int silu, int dtype,
           cudaStream_t st);
bool gn_pre_ok(int C, int G);
int gn_bwd_pre_grid(int B, int HW, int C);
int gn_apply_pre(const void* x, const float* gamma, const float* beta, void* y, const float* sums, float* stats, int B,
                 int HW, int C, int G, float eps, int silu, int dtype, cudaStream_t st);
int gn_bwd_pre(const void* du, const void* x, const float* gamma, const float* stats, const float* s12,
               const void* dx_add, void* dx, float* out3, float* scratch, int B, int HW, int C, int G, int dtype,
               cudaStream_t st);
int attn_fwd(const void* qkv, void* out, float* probs, int B, int Tn, int C, int dtype, cudaStream_t st);
int attn_bwd(const void* qkv, const float* probs, const void* dout, void* dqkv, float* ds, int B, int Tn, int C,
             int dtype, cudaStream_t st);
int reparam_fwd(const void* mom, const float* eps, float* eps_out, float* mean, float* logvar, float* stdv,
                float* z_nchw, void* z_nhwc, double* kl_sum, int B, int zc, int R, float clampv, uint64_t seed, uint64_t offset,
                uint64_t elem_base, const double* dyn, int dtype, cudaStream_t st);
int step_state_advance(double* dyn, double rng_inc, cudaStream_t st);
int reparam_bwd(const float* d_mean, const float* d_logvar, const float* d_z, const float* logvar, const float* eps,
                const void* mom, void* d_mom, int B, int zc, int R, float clampv, int dtype, cudaStream_t st);
int recon_loss(const float* r, const float* x, double* loss_sum, float* d_unit, long long n, int kind, cudaStream_t st);
int kl_fwd(const float* mean, const float* logvar, double* kl_sum, long long n, cudaStream_t st);
int kl_bwd(const float* mean, const float* logvar, const float* coef, float* d_mean, float* d_logvar, long long n,
           cudaStream_t st);
int scale_dev(const float* in, const float* s, float* out, long long n, cudaStream_t st);
int grad_sqnorm(const float* const* grads, const long long* sizes, int nt, double* out, cudaStream_t st);
int adamw_step(float* const* params, const float* const* grads, float* const* m1, float* const* m2,
               const long long* sizes, int nt, const double* sqnorm, float max_norm, float lr, float b1, float b2,
               float eps, float wd, int decoupled, int step, cudaStream_t st);

int grad_sqnorm_chunked(const float* const* grads, const int* ch_tensor, const long long* ch_off, const int* ch_len,
                        int n_chunks, double* tensor_sq, int* tensor_bad, cudaStream_t st);
int adamw_step_chunked(float* const* params, const float* const* grads, float* const* m1, float* const* m2,
                       const int* ch_tensor, const long long* ch_off, const int* ch_len, int n_chunks,
                       const double* tensor_sq, const int* tensor_bad, const int* tensor_skip, int n_tensors,
                       double* sqnorm_out, float max_norm, float lr, float b1, float b2, float eps, float wd,
                       int decoupled, int step, const double* dyn, cudaStream_t st);

bool attn_fused_ok(int Tn, int C);
int attn_fused_fwd(const bf16* qkv, bf16* out, float* lse2, int B, int Tn, int C, cudaStream_t st);
int attn_fused_bwd(const bf16* qkv, const bf16* out, const float* lse2, const bf16* dout, bf16* dqkv, float* dscratch,
                   int B, int Tn, int C, cudaStream_t st);
int linattn_fwd(const void* qkv, void* out, float* ks, float* ctx, int B, int Tn, int C, int dtype, cudaStream_t st);
int linattn_bwd(const void* qkv, const float* ks, const float* ctx, const void* dout, void* dqkv, float* dctx, float* dks,
                int B, int Tn, int C, int dtype, cudaStream_t st);
int dropout_apply(const void* x, void* y, long long n, float p, uint64_t seed, uint64_t offset, uint64_t elem_base,
                  const double* dyn,
                  const uint8_t* mask_in, uint8_t* mask_out, int dtype, cudaStream_t st);
int film_fwd(const float* f, const float* cond, const float* ws, const float* bs, const float* wt, const float* bt, float* out,
             float* st, int B, int C, int HW, int K, cudaStream_t s);
int film_bwd(const float* dy, const float* f, const float* cond, const float* ws, const float* wt, const float* st, float* df,
             float* dst, float* dws, float* dbs, float* dwt, float* dbt, float* dcond, int B, int C, int HW, int K,
             cudaStream_t s);
int batch_feed_u8(const uint8_t* src, const int* src_c, const int* tgt_c, const long long* idx, float* x, float* onehot, int B,
                  int H, int W, int SC, int OC, int n_mod, cudaStream_t st);
long long recon_metrics_ws_doubles(int planes, int H, int W);
long long stat_metrics_ws_doubles();
int recon_metrics(const float* x, const float* r, int planes, int H, int W, float data_range, double* ws, float* out,
                  cudaStream_t st);
int kl_metrics(const float* mean, const float* logvar, int B, int C, int S, double* ws, float* out, cudaStream_t st);
int latent_metrics(const float* z, int B, long long D, float thr, double* ws, float* out, cudaStream_t st);

static bool dtype_ok(int d) { return d == MVAE_F32 || d == MVAE_BF16; }

static bool use_umma_gather(const mvae_conv_desc& d, const GatherProblem* ps, int np) {
    if (d.dtype != MVAE_BF16 || d.engine == MVAE_ENGINE_SIMT) return false;
    for (int i = 0; i < np; ++i)
        if (!umma_gather_ok(ps[i])) return false;
    return true;
}

}  // namespace mvae

using namespace mvae;

#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

const char* mvae_last_error(void) { return g_err; }
int mvae_abi_version(void) { return MVAE_ABI_VERSION; }

int mvae_device_ok(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        return 0;
    }
    int dev = 0, major = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    return major == 10 ? 1 : 0;
}

uint64_t mvae_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int mvae_set_debug(int key, int value) {
    MVAE_CHECK(key >= 0 && key < 32, "debug key out of range");
    g_knobs[key] = value;
    return 0;
}

int mvae_conv2d_out_hw(const mvae_conv_desc* d, int32_t* oh, int32_t* ow) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0, "bad conv descriptor");
    *oh = g.P; *ow = g.Q;
    return 0;
}
int mvae_conv2d_num_slabs(const mvae_conv_desc* d) {
    ConvGeom g;
    if (!d || conv_geom(*d, g)) return -1;
    return conv_num_slabs(g);
}
int mvae_conv2d_uses_umma(const mvae_conv_desc* d) {
    ConvGeom g;
    if (!d || conv_geom(*d, g)) return 0;
    GatherProblem ps[4];
    int np = lower_fprop(g, ps);
    return use_umma_gather(*d, ps, np) ? 1 : 0;
}

int mvae_conv2d_pack_weights(const mvae_conv_desc* d, const float* w, void* wf, void* wd, int32_t out_c_total,
                             int32_t out_c_offset, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(w && (wf || wd), "pack_weights: null pointer");
    MVAE_CHECK(out_c_offset >= 0 && out_c_offset + g.Cout <= out_c_total, "pack_weights: bad channel window");
    return pack_weights(g, w, wf, wd, out_c_total, out_c_offset, d->dtype, ST(stream));
}

int mvae_conv2d_pack_weights_batched(const mvae_pack_job* jobs_dev, int32_t n_jobs, int32_t total_tiles, int32_t dtype,
                                     void* stream) {
    MVAE_CHECK(jobs_dev && n_jobs > 0 && total_tiles > 0 && dtype_ok(dtype), "pack_weights_batched: bad argument");
    return pack_weights_batched(jobs_dev, n_jobs, total_tiles, dtype, ST(stream));
}

int mvae_conv2d_fprop(const mvae_conv_desc* d, const void* x, const void* wf, const float* bias, const void* residual,
                      void* y, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(x && wf && y, "conv2d_fprop: null pointer");
    GatherProblem ps[4];
    int np = lower_fprop(g, ps);
    bool um = use_umma_gather(*d, ps, np);
    MVAE_CHECK(um || d->engine != MVAE_ENGINE_UMMA, "conv2d_fprop: tcgen05 engine requested for an unsupported shape");
    for (int i = 0; i < np; ++i) {
        int rc = um ? umma_gather(ps[i], x, wf, conv_num_slabs(g), bias, residual, y, ST(stream))
                    : simt_gather(ps[i], x, wf, bias, residual, y, d->dtype, ST(stream));
        if (rc) return rc;
    }
    return 0;
}

int mvae_conv2d_dgrad(const mvae_conv_desc* d, const void* dy, const void* wd, void* dx, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(dy && wd && dx, "conv2d_dgrad: null pointer");
    GatherProblem ps[4];
    int np = lower_dgrad(g, ps);
    bool um = use_umma_gather(*d, ps, np);
    MVAE_CHECK(um || d->engine != MVAE_ENGINE_UMMA, "conv2d_dgrad: tcgen05 engine requested for an unsupported shape");
    for (int i = 0; i < np; ++i) {
        int rc = um ? umma_gather(ps[i], dy, wd, conv_num_slabs(g), nullptr, nullptr, dx, ST(stream))
                    : simt_gather(ps[i], dy, wd, nullptr, nullptr, dx, d->dtype, ST(stream));
        if (rc) return rc;
    }
    return 0;
}

// GroupNorm over `c` channels in `groups` groups can ride in the tcgen05 gather epilogue
static int gn_cpg_log2(int c, int groups) {
    if (groups <= 0 || c % groups != 0) return -1;
    int cpg = c / groups, l = 0;
    while ((1 << l) < cpg) ++l;
    return ((1 << l) == cpg && l >= 2) ? l : -1;
}

int mvae_conv2d_fuses_gn(const mvae_conv_desc* d, int32_t groups, int32_t direction) {
    ConvGeom g;
    if (!d || conv_geom(*d, g)) return 0;
    GatherProblem ps[4];
    const int np = direction ? lower_dgrad(g, ps) : lower_fprop(g, ps);
    if (!use_umma_gather(*d, ps, np)) return 0;
    const int c = direction ? g.Cin : g.Cout;
    if (direction) {
        // tuning knob 10: 1 = never fuse the GroupNorm backward, 2 / 3 = only behind a reduction longer than 1152 / 2304
        const int knob = debug_knob(10);
        const long long kred = (long long)g.Cout * d->ksize * d->ksize;
        if (knob == 1 || (knob == 2 && kred <= 1152) || (knob == 3 && kred <= 2304)) return 0;
    }
    return (gn_cpg_log2(c, groups) >= 0 && gn_pre_ok(c, groups)) ? 1 : 0;
}

int mvae_conv2d_fprop_gn(const mvae_conv_desc* d, const void* x, const void* wf, const float* bias, const void* residual,
                         void* y, float* gn_sums, int32_t groups, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(x && wf && y && gn_sums, "conv2d_fprop_gn: null pointer");
    MVAE_CHECK(mvae_conv2d_fuses_gn(d, groups, 0), "conv2d_fprop_gn: this conv / group count cannot fuse GroupNorm statistics");
    GatherProblem ps[4];
    int np = lower_fprop(g, ps);
    MVAE_CUDA(cudaMemsetAsync(gn_sums, 0, sizeof(float) * 2 * (size_t)g.B * groups, ST(stream)));
    GnHook h;
    memset(&h, 0, sizeof(h));
    h.mode = 1; h.G = groups; h.cpg_log2 = gn_cpg_log2(g.Cout, groups); h.sums = gn_sums;
    for (int i = 0; i < np; ++i)
        if (umma_gather(ps[i], x, wf, conv_num_slabs(g), bias, residual, y, ST(stream), &h)) return 1;
    return 0;
}

int mvae_conv2d_dgrad_gn(const mvae_conv_desc* d, const void* dy, const void* wd, void* du, const void* gn_x,
                         const float* gn_stats, const float* gamma, const float* beta, int32_t groups, int32_t apply_silu,
                         float* gn_s12, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(dy && wd && du && gn_x && gn_stats && gamma && beta && gn_s12, "conv2d_dgrad_gn: null pointer");
    MVAE_CHECK(mvae_conv2d_fuses_gn(d, groups, 1), "conv2d_dgrad_gn: this conv / group count cannot fuse the GroupNorm backward");
    GatherProblem ps[4];
    int np = lower_dgrad(g, ps);
    MVAE_CUDA(cudaMemsetAsync(gn_s12, 0, sizeof(float) * 2 * (size_t)g.B * groups, ST(stream)));
    GnHook h;
    memset(&h, 0, sizeof(h));
    h.mode = 2; h.G = groups; h.cpg_log2 = gn_cpg_log2(g.Cin, groups); h.silu = apply_silu; h.sums = gn_s12;
    h.x = gn_x; h.stat = gn_stats; h.gamma = gamma; h.beta = beta;
    for (int i = 0; i < np; ++i)
        if (umma_gather(ps[i], dy, wd, conv_num_slabs(g), nullptr, nullptr, du, ST(stream), &h)) return 1;
    return 0;
}

int mvae_groupnorm_fused_ok(int32_t c, int32_t groups) { return gn_pre_ok(c, groups) ? 1 : 0; }
int64_t mvae_groupnorm_apply_bwd_scratch_floats(int32_t batch, int32_t hw, int32_t c) {
    return (int64_t)gn_bwd_pre_grid(batch, hw, c) * 3 * c;
}
int mvae_groupnorm_apply_fwd(const void* x, const float* gamma, const float* beta, void* y, const float* sums,
                             float* stats, int32_t batch, int32_t hw, int32_t c, int32_t groups, float eps,
                             int32_t apply_silu, int32_t dtype, void* stream) {
    MVAE_CHECK(x && gamma && beta && y && sums && stats && dtype_ok(dtype), "groupnorm_apply_fwd: bad argument");
    return gn_apply_pre(x, gamma, beta, y, sums, stats, batch, hw, c, groups, eps, apply_silu, dtype, ST(stream));
}
int mvae_groupnorm_apply_bwd(const void* du, const void* x, const float* gamma, const float* stats, const float* s12,
                             const void* dx_add, void* dx, float* out3, float* scratch, int32_t batch, int32_t hw,
                             int32_t c, int32_t groups, int32_t dtype, void* stream) {
    MVAE_CHECK(du && x && gamma && stats && s12 && dx && out3 && scratch && dtype_ok(dtype),
               "groupnorm_apply_bwd: bad argument");
    return gn_bwd_pre(du, x, gamma, stats, s12, dx_add, dx, out3, scratch, batch, hw, c, groups, dtype, ST(stream));
}

static bool use_umma_reduce(const mvae_conv_desc& d, const ReduceProblem& p) {
    return d.dtype == MVAE_BF16 && d.engine != MVAE_ENGINE_SIMT && umma_reduce_ok(p);
}

size_t mvae_conv2d_wgrad_workspace_bytes(const mvae_conv_desc* d) {
    ConvGeom g;
    if (!d || conv_geom(*d, g)) return 0;
    ReduceProblem p;
    lower_wgrad(g, p);
    int splits = use_umma_reduce(*d, p) ? umma_reduce_splits(p) : 1;
    size_t partial = sizeof(float) * (size_t)splits * conv_num_slabs(g) * g.Cout * g.Cin;
    size_t cs = sizeof(float) * (size_t)colsum_parts((long long)g.B * g.P * g.Q) * g.Cout;
    return partial + cs + 256;
}

static int wgrad_impl(const mvae_conv_desc* d, const void* x, const void* dy, float* dw, float* dbias, void* workspace,
                      size_t workspace_bytes, float beta, void* stream) {
    ConvGeom g;
    MVAE_CHECK(d && conv_geom(*d, g) == 0 && dtype_ok(d->dtype), "bad conv descriptor");
    MVAE_CHECK(x && dy && dw && workspace, "conv2d_wgrad: null pointer");
    MVAE_CHECK(workspace_bytes >= mvae_conv2d_wgrad_workspace_bytes(d), "conv2d_wgrad: workspace too small");
    ReduceProblem p;
    lower_wgrad(g, p);
    bool um = use_umma_reduce(*d, p);
    MVAE_CHECK(um || d->engine != MVAE_ENGINE_UMMA, "conv2d_wgrad: tcgen05 engine requested for an unsupported shape");
    int splits = um ? umma_reduce_splits(p) : 1;
    float* partial = reinterpret_cast<float*>(workspace);
    size_t partial_elems = (size_t)splits * conv_num_slabs(g) * g.Cout * g.Cin;
    int rc = um ? umma_reduce(p, x, dy, partial, splits, ST(stream)) : simt_reduce(p, x, dy, partial, d->dtype, ST(stream));
    if (rc) return rc;
    if (wgrad_finalize(g, partial, dw, splits, beta, ST(stream))) return 1;
    if (dbias) {
        float* cs = partial + ((partial_elems + 63) / 64) * 64;
        return colsum(dy, dbias, cs, (long long)g.B * g.P * g.Q, g.Cout, d->dtype, beta, ST(stream));
    }
    return 0;
}

int mvae_conv2d_wgrad(const mvae_conv_desc* d, const void* x, const void* dy, float* dw, float* dbias, void* workspace,
                      size_t workspace_bytes, void* stream) {
    return wgrad_impl(d, x, dy, dw, dbias, workspace, workspace_bytes, 0.f, stream);
}
int mvae_conv2d_wgrad_acc(const mvae_conv_desc* d, const void* x, const void* dy, float* dw, float* dbias, void* workspace,
                          size_t workspace_bytes, float beta, void* stream) {
    return wgrad_impl(d, x, dy, dw, dbias, workspace, workspace_bytes, beta, stream);
}

int mvae_groupnorm_splits(int32_t batch, int32_t hw, int32_t c) { return gn_splits(batch, hw, c); }
int64_t mvae_groupnorm_scratch_floats(int32_t batch, int32_t hw, int32_t c, int32_t groups, int32_t backward) {
    return gn_scratch_floats(batch, hw, c, groups, backward);
}

int mvae_groupnorm_silu_fwd(const void* x, const float* gamma, const float* beta, void* y, float* stats, float* scratch,
                            int32_t batch, int32_t hw, int32_t c, int32_t groups, float eps, int32_t apply_silu,
                            int32_t dtype, void* stream) {
    MVAE_CHECK(x && gamma && beta && y && stats && scratch && dtype_ok(dtype), "groupnorm_silu_fwd: bad argument");
    return gn_fwd(x, gamma, beta, y, stats, scratch, batch, hw, c, groups, eps, apply_silu, dtype, ST(stream));
}
int mvae_groupnorm_silu_bwd(const void* dy, const void* x, const float* gamma, const float* beta, const float* stats,
                            const void* dx_add, void* dx, float* dgamma, float* dbeta, float* scratch, int32_t batch,
                            int32_t hw, int32_t c, int32_t groups, int32_t apply_silu, int32_t dtype, void* stream) {
    MVAE_CHECK(dy && x && gamma && beta && stats && dx && dgamma && dbeta && scratch && dtype_ok(dtype),
               "groupnorm_silu_bwd: bad argument");
    return gn_bwd(dy, x, gamma, beta, stats, dx_add, dx, dgamma, dbeta, scratch, batch, hw, c, groups, apply_silu, dtype,
                  ST(stream));
}

int mvae_attn_core_fwd(const void* qkv, void* out, float* probs, int32_t batch, int32_t tokens, int32_t c, int32_t dtype,
                       void* stream) {
    MVAE_CHECK(qkv && out && probs && dtype_ok(dtype), "attn_core_fwd: bad argument");
    return attn_fwd(qkv, out, probs, batch, tokens, c, dtype, ST(stream));
}
int mvae_attn_core_bwd(const void* qkv, const float* probs, const void* dout, void* dqkv, float* dscratch, int32_t batch,
                       int32_t tokens, int32_t c, int32_t dtype, void* stream) {
    MVAE_CHECK(qkv && probs && dout && dqkv && dscratch && dtype_ok(dtype), "attn_core_bwd: bad argument");
    return attn_bwd(qkv, probs, dout, dqkv, dscratch, batch, tokens, c, dtype, ST(stream));
}

int mvae_input_pack_fwd(const float* x, const float* cond, const float* w, const float* b, void* out, int32_t batch,
                        int32_t c, int32_t h, int32_t w_, int32_t cond_dim, int32_t out_c, int32_t dtype, void* stream) {
    MVAE_CHECK(x && out && dtype_ok(dtype), "input_pack_fwd: bad argument");
    MVAE_CHECK((w == nullptr) || (cond && b && out_c >= 2 * c), "input_pack_fwd: conditioning needs cond, w, b");
    MVAE_CHECK(out_c >= c, "input_pack_fwd: out_c < c");
    return input_pack_fwd(x, cond, w, b, out, batch, c, h, w_, cond_dim, out_c, dtype, ST(stream));
}
int mvae_input_pack_bwd(const void* d_in, const float* cond, const float* w, const float* b, float* dw, float* db,
                        int32_t batch, int32_t c, int32_t h, int32_t w_, int32_t cond_dim, int32_t out_c, int32_t dtype,
                        void* stream) {
    MVAE_CHECK(d_in && cond && w && b && dw && db && dtype_ok(dtype), "input_pack_bwd: bad argument");
    return input_pack_bwd(d_in, cond, w, b, dw, db, batch, c, h, w_, cond_dim, out_c, dtype, ST(stream));
}

int mvae_nhwc_to_nchw_f32(const void* src, float* dst, int32_t batch, int32_t h, int32_t w, int32_t c_src, int32_t c_off,
                          int32_t c_dst, int32_t dtype, void* stream) {
    MVAE_CHECK(src && dst && dtype_ok(dtype) && c_off >= 0 && c_off + c_dst <= c_src, "nhwc_to_nchw: bad argument");
    return nhwc_to_nchw(src, dst, batch, h * w, c_src, c_off, c_dst, dtype, ST(stream));
}
int mvae_nchw_f32_to_nhwc(const float* src, void* dst, int32_t batch, int32_t h, int32_t w, int32_t c, int32_t dtype,
                          void* stream) {
    MVAE_CHECK(src && dst && dtype_ok(dtype), "nchw_to_nhwc: bad argument");
    return nchw_to_nhwc(src, dst, batch, h * w, c, dtype, ST(stream));
}

int mvae_reparam_kl_fwd(const void* moments, const float* eps, float* eps_out, float* mean, float* logvar,
                        float* std_out, float* z_nchw, void* z_nhwc, double* kl_sum, int32_t batch, int32_t zc, int32_t r, float clamp, uint64_t seed,
                        uint64_t offset, uint64_t elem_base, const double* step_state, int32_t dtype, void* stream) {
    MVAE_CHECK(moments && mean && logvar && z_nchw && dtype_ok(dtype), "reparam_kl_fwd: bad argument");
    return reparam_fwd(moments, eps, eps_out, mean, logvar, std_out, z_nchw, z_nhwc, kl_sum, batch, zc, r, clamp, seed, offset,
                       elem_base, step_state, dtype, ST(stream));
}
int mvae_reparam_kl_bwd(const float* d_mean, const float* d_logvar, const float* d_z, const float* mean_unused,
                        const float* logvar, const float* eps, const void* moments, void* d_moments, int32_t batch,
                        int32_t zc, int32_t r, float clamp, int32_t dtype, void* stream) {
    (void)mean_unused;
    MVAE_CHECK(logvar && eps && moments && d_moments && dtype_ok(dtype), "reparam_kl_bwd: bad argument");
    return reparam_bwd(d_mean, d_logvar, d_z, logvar, eps, moments, d_moments, batch, zc, r, clamp, dtype, ST(stream));
}

int mvae_recon_loss_fwd_bwd(const float* recon, const float* x, double* loss_sum, float* d_unit, int64_t n, int32_t kind,
                            void* stream) {
    MVAE_CHECK(recon && x && loss_sum && n > 0 && kind >= 0 && kind <= 2, "recon_loss: bad argument");
    return recon_loss(recon, x, loss_sum, d_unit, n, kind, ST(stream));
}
int mvae_kl_fwd(const float* mean, const float* logvar, double* kl_sum, int64_t n, void* stream) {
    MVAE_CHECK(mean && logvar && kl_sum && n > 0, "kl_fwd: bad argument");
    return kl_fwd(mean, logvar, kl_sum, n, ST(stream));
}
int mvae_kl_bwd(const float* mean, const float* logvar, const float* coef_dev, float* d_mean, float* d_logvar, int64_t n,
                void* stream) {
    MVAE_CHECK(mean && logvar && coef_dev && d_mean && d_logvar && n > 0, "kl_bwd: bad argument");
    return kl_bwd(mean, logvar, coef_dev, d_mean, d_logvar, n, ST(stream));
}
int mvae_scale_by_device_scalar(const float* in, const float* scale_dev_, float* out, int64_t n, void* stream) {
    MVAE_CHECK(in && scale_dev_ && out && n > 0, "scale: bad argument");
    return scale_dev(in, scale_dev_, out, n, ST(stream));
}

int mvae_grad_sqnorm(const float* const* grads, const int64_t* sizes, int32_t n_tensors, double* sqnorm_dev, void* stream) {
    MVAE_CHECK(grads && sizes && sqnorm_dev && n_tensors > 0, "grad_sqnorm: bad argument");
    return grad_sqnorm(grads, reinterpret_cast<const long long*>(sizes), n_tensors, sqnorm_dev, ST(stream));
}
int mvae_adamw_clip_step(float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                         const int64_t* sizes, int32_t n_tensors, const double* sqnorm_dev, float max_norm, float lr,
                         float beta1, float beta2, float eps, float weight_decay, int32_t decoupled, int32_t step,
                         void* stream) {
    MVAE_CHECK(params && grads && exp_avg && exp_avg_sq && sizes && n_tensors > 0 && step >= 1, "adamw: bad argument");
    return adamw_step(params, grads, exp_avg, exp_avg_sq, reinterpret_cast<const long long*>(sizes), n_tensors, sqnorm_dev,
                      max_norm, lr, beta1, beta2, eps, weight_decay, decoupled, step, ST(stream));
}

int mvae_grad_sqnorm_chunked(const float* const* grads, const int32_t* chunk_tensor, const int64_t* chunk_offset,
                             const int32_t* chunk_len, int32_t n_chunks, double* tensor_sq, int32_t* tensor_bad,
                             void* stream) {
    MVAE_CHECK(grads && chunk_tensor && chunk_offset && chunk_len && tensor_sq && tensor_bad && n_chunks > 0,
               "grad_sqnorm_chunked: bad argument");
    return grad_sqnorm_chunked(grads, chunk_tensor, reinterpret_cast<const long long*>(chunk_offset), chunk_len, n_chunks,
                               tensor_sq, tensor_bad, ST(stream));
}
int mvae_adamw_clip_step_chunked(float* const* params, const float* const* grads, float* const* exp_avg,
                                 float* const* exp_avg_sq, const int32_t* chunk_tensor, const int64_t* chunk_offset,
                                 const int32_t* chunk_len, int32_t n_chunks, const double* tensor_sq,
                                 const int32_t* tensor_bad, const int32_t* tensor_skip, int32_t n_tensors,
                                 double* sqnorm_out, float max_norm, float lr, float beta1, float beta2, float eps,
                                 float weight_decay, int32_t decoupled, int32_t step, const double* step_state,
                                 void* stream) {
    MVAE_CHECK(params && grads && exp_avg && exp_avg_sq && chunk_tensor && chunk_offset && chunk_len && n_chunks > 0 &&
                   tensor_sq && tensor_bad && n_tensors > 0 && (step >= 1 || step_state), "adamw_chunked: bad argument");
    return adamw_step_chunked(params, grads, exp_avg, exp_avg_sq, chunk_tensor,
                              reinterpret_cast<const long long*>(chunk_offset), chunk_len, n_chunks, tensor_sq, tensor_bad,
                              tensor_skip, n_tensors, sqnorm_out, max_norm, lr, beta1, beta2, eps, weight_decay, decoupled,
                              step, step_state, ST(stream));
}

int mvae_dis_input_fwd(const float* x, const int64_t* idx, const float* pw, const float* pb, const int32_t* is_gray,
                       void* out, int32_t batch, int32_t xc, int32_t h, int32_t w, int32_t n_mod, int32_t out_c,
                       int32_t dtype, void* stream) {
    MVAE_CHECK(x && idx && pw && pb && is_gray && out && dtype_ok(dtype) && out_c >= 3, "dis_input_fwd: bad argument");
    return dis_input_fwd(x, reinterpret_cast<const long long*>(idx), pw, pb, is_gray, out, batch, xc, h * w, n_mod, out_c,
                         dtype, ST(stream));
}
int mvae_dis_input_bwd(const void* d_in, const float* x, const int64_t* idx, const int32_t* is_gray, float* dpw, float* dpb,
                       int32_t batch, int32_t xc, int32_t h, int32_t w, int32_t n_mod, int32_t out_c, int32_t dtype,
                       void* stream) {
    MVAE_CHECK(d_in && x && idx && is_gray && dpw && dpb && dtype_ok(dtype), "dis_input_bwd: bad argument");
    return dis_input_bwd(d_in, x, reinterpret_cast<const long long*>(idx), is_gray, dpw, dpb, batch, xc, h * w, n_mod,
                         out_c, dtype, ST(stream));
}

int mvae_dis_head_fwd(const void* r_nhwc, int32_t r_channels, const int64_t* idx, const float* w1, const float* b1,
                      const float* w2, const float* b2, const float* proj_w, const float* proj_b, const int32_t* is_gray,
                      float* hid, float* y, float* out_nchw, int32_t batch, int32_t h, int32_t w, int32_t n_mod,
                      int32_t out_c, int32_t dtype, void* stream) {
    MVAE_CHECK(r_nhwc && idx && w1 && b1 && w2 && b2 && proj_w && proj_b && is_gray && hid && y && out_nchw &&
                   dtype_ok(dtype) && r_channels >= 3 && out_c >= 1 && n_mod >= 1,
               "dis_head_fwd: bad argument");
    return dis_head_fwd(r_nhwc, r_channels, reinterpret_cast<const long long*>(idx), w1, b1, w2, b2, proj_w, proj_b,
                        is_gray, hid, y, out_nchw, batch, h, w, n_mod, out_c, dtype, ST(stream));
}
int mvae_dis_head_bwd(const float* d_out_nchw, const void* r_nhwc, int32_t r_channels, const float* hid, const float* y,
                      const int64_t* idx, const float* w1, const float* w2, const float* proj_w, const int32_t* is_gray,
                      void* d_r_nhwc, float* grads, float* scratch, int32_t batch, int32_t h, int32_t w, int32_t n_mod,
                      int32_t out_c, int32_t dtype, void* stream) {
    MVAE_CHECK(d_out_nchw && r_nhwc && hid && y && idx && w1 && w2 && proj_w && is_gray && d_r_nhwc && grads && scratch &&
                   dtype_ok(dtype) && r_channels >= 3,
               "dis_head_bwd: bad argument");
    return dis_head_bwd(d_out_nchw, r_nhwc, r_channels, hid, y, reinterpret_cast<const long long*>(idx), w1, w2, proj_w,
                        is_gray, d_r_nhwc, grads, scratch, batch, h, w, n_mod, out_c, dtype, ST(stream));
}
int64_t mvae_dis_aux_floats(int32_t batch, int32_t modality_dim, int32_t n_mod) {
    return dis_aux_floats(batch, modality_dim, n_mod);
}
int mvae_dis_aux_loss_fwd(const float* z_flat, const int64_t* idx, float* aux, int32_t batch, int64_t flat_dim,
                          int32_t shared_dim, int32_t modality_dim, int32_t n_mod, float temperature, void* stream) {
    MVAE_CHECK(z_flat && idx && aux && temperature > 0.f, "dis_aux_loss_fwd: bad argument");
    return dis_aux_fwd(z_flat, reinterpret_cast<const long long*>(idx), aux, batch, flat_dim, shared_dim, modality_dim,
                       n_mod, temperature, ST(stream));
}
int mvae_dis_aux_loss_bwd(const float* z_flat, const int64_t* idx, const float* aux, const float* g_sep_dev,
                          const float* g_con_dev, float* dz_flat, int32_t batch, int64_t flat_dim, int32_t shared_dim,
                          int32_t modality_dim, int32_t n_mod, float temperature, void* stream) {
    MVAE_CHECK(z_flat && idx && aux && dz_flat && temperature > 0.f, "dis_aux_loss_bwd: bad argument");
    return dis_aux_bwd(z_flat, reinterpret_cast<const long long*>(idx), aux, g_sep_dev, g_con_dev, dz_flat, batch, flat_dim,
                       shared_dim, modality_dim, n_mod, temperature, ST(stream));
}
int mvae_dis_loss_total(const float* recon, const float* kl, const float* sep, const float* con, float w_recon, float w_kl,
                        float w_sep, float w_con, float* out5, float* mask4, void* stream) {
    MVAE_CHECK(recon && kl && sep && con && out5 && mask4, "dis_loss_total: bad argument");
    return dis_total(recon, kl, sep, con, w_recon, w_kl, w_sep, w_con, out5, mask4, ST(stream));
}

int mvae_attn_fused_ok(int32_t tokens, int32_t c, int32_t dtype) {
    return (dtype == MVAE_BF16 && attn_fused_ok(tokens, c) && debug_knob(11) == 0) ? 1 : 0;
}
int mvae_attn_fused_fwd(const void* qkv, void* out, float* lse2, int32_t batch, int32_t tokens, int32_t c, int32_t dtype,
                        void* stream) {
    MVAE_CHECK(qkv && out && lse2 && batch > 0, "attn_fused_fwd: bad argument");
    MVAE_CHECK(dtype == MVAE_BF16, "attn_fused_fwd: bf16 only (the fp32 verification mode uses mvae_attn_core_fwd)");
    return attn_fused_fwd((const bf16*)qkv, (bf16*)out, lse2, batch, tokens, c, ST(stream));
}
int mvae_attn_fused_bwd(const void* qkv, const void* out, const float* lse2, const void* dout, void* dqkv, float* dscratch,
                        int32_t batch, int32_t tokens, int32_t c, int32_t dtype, void* stream) {
    MVAE_CHECK(qkv && out && lse2 && dout && dqkv && dscratch && batch > 0, "attn_fused_bwd: bad argument");
    MVAE_CHECK(dtype == MVAE_BF16, "attn_fused_bwd: bf16 only (the fp32 verification mode uses mvae_attn_core_bwd)");
    return attn_fused_bwd((const bf16*)qkv, (const bf16*)out, lse2, (const bf16*)dout, (bf16*)dqkv, dscratch, batch, tokens,
                          c, ST(stream));
}
int mvae_linattn_core_fwd(const void* qkv, void* out, float* ks, float* ctx, int32_t batch, int32_t tokens, int32_t c,
                          int32_t dtype, void* stream) {
    MVAE_CHECK(qkv && out && ks && ctx && dtype_ok(dtype) && batch > 0 && tokens > 0 && c > 0, "linattn_core_fwd: bad argument");
    return linattn_fwd(qkv, out, ks, ctx, batch, tokens, c, dtype, ST(stream));
}
int mvae_linattn_core_bwd(const void* qkv, const float* ks, const float* ctx, const void* dout, void* dqkv, float* dctx,
                          float* dks, int32_t batch, int32_t tokens, int32_t c, int32_t dtype, void* stream) {
    MVAE_CHECK(qkv && ks && ctx && dout && dqkv && dctx && dks && dtype_ok(dtype), "linattn_core_bwd: bad argument");
    return linattn_bwd(qkv, ks, ctx, dout, dqkv, dctx, dks, batch, tokens, c, dtype, ST(stream));
}
int mvae_dropout_apply(const void* x, void* y, int64_t n, float p, uint64_t seed, uint64_t offset, uint64_t elem_base,
                       const double* step_state, const uint8_t* mask_in, uint8_t* mask_out, int32_t dtype, void* stream) {
    MVAE_CHECK(x && y && n > 0 && dtype_ok(dtype), "dropout_apply: bad argument");
    MVAE_CHECK(p >= 0.f && p < 1.f, "dropout_apply: p must be in [0, 1), got %f", (double)p);
    return dropout_apply(x, y, n, p, seed, offset, elem_base, step_state, mask_in, mask_out, dtype, ST(stream));
}
int mvae_step_state_advance(double* step_state, double rng_increment, void* stream) {
    MVAE_CHECK(step_state && rng_increment >= 0.0, "step_state_advance: bad argument");
    return step_state_advance(step_state, rng_increment, ST(stream));
}
int mvae_film_fwd(const float* features, const float* cond, const float* w_scale, const float* b_scale, const float* w_shift,
                  const float* b_shift, float* out, float* st, int32_t batch, int32_t c, int32_t hw, int32_t k, void* stream) {
    MVAE_CHECK(features && cond && w_scale && b_scale && w_shift && b_shift && out && st && batch > 0 && c > 0 && hw > 0 && k > 0,
               "film_fwd: bad argument");
    return film_fwd(features, cond, w_scale, b_scale, w_shift, b_shift, out, st, batch, c, hw, k, ST(stream));
}
int mvae_film_bwd(const float* d_out, const float* features, const float* cond, const float* w_scale, const float* w_shift,
                  const float* st, float* d_features, float* dst, float* d_w_scale, float* d_b_scale, float* d_w_shift,
                  float* d_b_shift, float* d_cond, int32_t batch, int32_t c, int32_t hw, int32_t k, void* stream) {
    MVAE_CHECK(d_out && features && cond && w_scale && w_shift && st && d_features && dst && d_w_scale && d_b_scale &&
                   d_w_shift && d_b_shift, "film_bwd: bad argument");
    return film_bwd(d_out, features, cond, w_scale, w_shift, st, d_features, dst, d_w_scale, d_b_scale, d_w_shift, d_b_shift,
                    d_cond, batch, c, hw, k, ST(stream));
}
int mvae_batch_feed_u8(const uint8_t* src, const int32_t* src_c, const int32_t* tgt_c, const int64_t* idx, float* x,
                       float* onehot, int32_t batch, int32_t h, int32_t w, int32_t src_stride_c, int32_t out_c, int32_t n_mod,
                       void* stream) {
    MVAE_CHECK(src && src_c && tgt_c && x && batch > 0 && h > 0 && w > 0, "batch_feed_u8: bad argument");
    MVAE_CHECK(src_stride_c == 1 || src_stride_c == 3, "batch_feed_u8: staged batches are 1 or 3 channels wide, got %d", src_stride_c);
    MVAE_CHECK(out_c == 1 || out_c == 3, "batch_feed_u8: out_c must be 1 or 3, got %d", out_c);
    MVAE_CHECK(onehot == nullptr || (idx && n_mod > 0), "batch_feed_u8: one-hot output needs idx and n_mod");
    return batch_feed_u8(src, src_c, tgt_c, reinterpret_cast<const long long*>(idx), x, onehot, batch, h, w, src_stride_c,
                         out_c, n_mod, ST(stream));
}
int64_t mvae_recon_metrics_ws_doubles(int32_t planes, int32_t h, int32_t w) { return recon_metrics_ws_doubles(planes, h, w); }
int64_t mvae_stat_metrics_ws_doubles(void) { return stat_metrics_ws_doubles(); }
int mvae_recon_metrics(const float* original, const float* recon, int32_t planes, int32_t h, int32_t w, float data_range,
                       double* ws, float* out4, void* stream) {
    MVAE_CHECK(original && recon && ws && out4 && planes > 0 && h > 0 && w > 0 && data_range > 0.f, "recon_metrics: bad argument");
    return recon_metrics(original, recon, planes, h, w, data_range, ws, out4, ST(stream));
}
int mvae_kl_metrics(const float* mean, const float* logvar, int32_t batch, int32_t c, int32_t s, double* ws, float* out4,
                    void* stream) {
    MVAE_CHECK(mean && logvar && ws && out4 && batch > 0 && c > 0 && s > 0, "kl_metrics: bad argument");
    return kl_metrics(mean, logvar, batch, c, s, ws, out4, ST(stream));
}
int mvae_latent_metrics(const float* z, int32_t batch, int64_t d, float thr, double* ws, float* out3, void* stream) {
    MVAE_CHECK(z && ws && out3 && batch > 0 && d > 0, "latent_metrics: bad argument");
    return latent_metrics(z, batch, d, thr, ws, out3, ST(stream));
}

}  // extern "C"
