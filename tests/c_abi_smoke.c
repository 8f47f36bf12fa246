/* Plain-C caller of libdincae_b200: proves that include/dincae_b200.h is a usable C ABI (no Python, no torch, no C++).
 * Built by __graft_entry__.build() with `gcc tests/c_abi_smoke.c -ldl`, run on the GPU box by tests/test_gpu_c_abi.py.
 *
 *   c_abi_smoke <path/to/libdincae_b200.so>
 *
 * Walks the sequence a DINCAE host performs (src/model.jl:479-651): context -> model -> parameters in -> one forward ->
 * three training steps on host buffers -> parameters out; then the points path: interp_adjoint on the reference's own
 * golden vectors (test/test_interp_adjoint.jl:20-47) and its adjoint identity (:49-65).  Exit code 0 = all checks passed. */
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../include/dincae_b200.h"

#define LOAD(name) \
    __typeof__(&name) p_##name = (__typeof__(&name))dlsym(h, #name); \
    if (!p_##name) { fprintf(stderr, "missing symbol %s\n", #name); return 2; }

#define CHECK(call) \
    do { int32_t st_ = (call); if (st_ != DINCAE_OK) { fprintf(stderr, "%s -> %d: %s\n", #call, st_, p_dincae_last_error()); return 3; } } while (0)

#define EXPECT(cond) \
    do { if (!(cond)) { fprintf(stderr, "check failed: %s (line %d)\n", #cond, __LINE__); return 4; } } while (0)

static float frand(unsigned* s) { *s = *s * 1664525u + 1013904223u; return (float)((*s >> 8) & 0xFFFF) / 65536.0f - 0.5f; }

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s libdincae_b200.so\n", argv[0]); return 1; }
    void* h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 1; }
    LOAD(dincae_last_error) LOAD(dincae_ctx_create) LOAD(dincae_ctx_destroy) LOAD(dincae_sync) LOAD(dincae_version)
    LOAD(dincae_model_create) LOAD(dincae_model_destroy) LOAD(dincae_model_num_params) LOAD(dincae_model_init_glorot)
    LOAD(dincae_model_get_params) LOAD(dincae_model_set_params) LOAD(dincae_forward) LOAD(dincae_loss) LOAD(dincae_train_step)
    LOAD(dincae_interp) LOAD(dincae_interp_adjoint) LOAD(dincae_launch_count)

    EXPECT(p_dincae_version() >= 100);
    dincae_ctx* ctx = NULL;
    CHECK(p_dincae_ctx_create(0, &ctx));

    /* ---- model: genmodel((W,H), ninput, noutput; enc_nfilter_internal = [32,64], loss_weights_refine = (0.3,0.7)) */
    enum { W = 24, H = 20, CIN = 10, B = 3 };
    dincae_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.W = W; cfg.H = H; cfg.ninput = CIN; cfg.noutput = 1; cfg.nlevels = 2; cfg.nfilter[0] = 32; cfg.nfilter[1] = 64;
    cfg.skip_mask = 0x2u;                         /* skipconnections = 2:(L+1) minus the identity level */
    cfg.upsampling = DINCAE_UP_NEAREST; cfg.nchains = 2; cfg.loss_weights[0] = 0.3f; cfg.loss_weights[1] = 0.7f;
    cfg.gamma = 10.0f; cfg.l2_beta = 1e-3f; cfg.clip_grad = 5.0f;
    cfg.adam_beta1 = 0.9f; cfg.adam_beta2 = 0.999f; cfg.adam_eps = 1e-8f;
    cfg.precision = DINCAE_PREC_TF32; cfg.max_batch = B;
    dincae_model* m = NULL;
    CHECK(p_dincae_model_create(ctx, &cfg, &m));
    int64_t np = 0;
    CHECK(p_dincae_model_num_params(m, &np));
    /* sum(length, Flux.trainables): per chain 9*(cin*32 + 32*64 + 64*32 + 32*2) + 32+64+32+2 */
    EXPECT(np == (9 * (10 * 32 + 32 * 64 + 64 * 32 + 32 * 2) + 130) + (9 * (12 * 32 + 32 * 64 + 64 * 32 + 32 * 2) + 130));
    CHECK(p_dincae_model_init_glorot(m, 7));
    float* w0 = (float*)malloc(sizeof(float) * np);
    float* w1 = (float*)malloc(sizeof(float) * np);
    CHECK(p_dincae_model_get_params(m, w0, np));
    CHECK(p_dincae_model_set_params(m, w0, np));          /* Flux-order round trip */
    CHECK(p_dincae_model_get_params(m, w1, np));
    EXPECT(memcmp(w0, w1, sizeof(float) * np) == 0);

    /* ---- one mini-batch, Julia layout (W,H,C,B) column-major */
    const size_t nin = (size_t)W * H * CIN * B, nt = (size_t)W * H * 2 * B;
    float* xin = (float*)malloc(sizeof(float) * nin);
    float* xt = (float*)malloc(sizeof(float) * nt);
    float* out = (float*)malloc(sizeof(float) * nt);
    unsigned seed = 12345u;
    for (size_t i = 0; i < nin; i++) xin[i] = 2.0f * frand(&seed);
    for (int b = 0; b < B; b++)
        for (int p = 0; p < W * H; p++) {
            float inv = frand(&seed) > -0.1f ? 1.0f + frand(&seed) : 0.0f;     /* 1/sigma^2, 0 = cloud */
            xt[((size_t)b * 2 + 1) * W * H + p] = inv;
            xt[((size_t)b * 2 + 0) * W * H + p] = inv * frand(&seed);
        }
    CHECK(p_dincae_forward(m, xin, B, out, DINCAE_HOST));
    for (size_t i = 0; i < nt; i++) EXPECT(isfinite(out[i]));
    for (int b = 0; b < B; b++)
        for (int p = 0; p < W * H; p++) EXPECT(out[((size_t)b * 2 + 1) * W * H + p] > 0.0f);    /* sigma^2 > 0 */
    double l0 = 0;
    CHECK(p_dincae_loss(m, xin, xt, B, DINCAE_HOST, &l0));
    float l[3];
    for (int k = 0; k < 3; k++) CHECK(p_dincae_train_step(m, xin, xt, B, 1e-3f, DINCAE_HOST, &l[k]));
    EXPECT(fabs((double)l[0] - l0) <= 1e-4 * fabs(l0));     /* the step reports the loss before its update */
    EXPECT(l[2] < l[0]);                                    /* three Adam steps on one batch reduce its loss */
    CHECK(p_dincae_model_get_params(m, w1, np));
    int changed = 0;
    for (int64_t i = 0; i < np; i++) changed += w0[i] != w1[i];
    EXPECT(changed > np / 2);
    EXPECT(p_dincae_launch_count(ctx) > 50);                /* kernels of THIS library ran */

    /* ---- points path: test/test_interp_adjoint.jl:20-47 */
    {
        const float pos[4] = {1.5f, 2.0f, 5.0f, 2.3f};      /* (x,y) interleaved, 1-based */
        const float vals[2] = {3.0f, 4.0f};
        float A[6 * 7];
        CHECK(p_dincae_interp_adjoint(ctx, pos, vals, 2, 6, 7, A, DINCAE_HOST));
        float ref[6 * 7];
        memset(ref, 0, sizeof ref);
        ref[0 + 6 * 1] = 1.5f; ref[1 + 6 * 1] = 1.5f; ref[4 + 6 * 1] = 0.7f * 4; ref[4 + 6 * 2] = 0.3f * 4;
        for (int i = 0; i < 42; i++) EXPECT(fabsf(A[i] - ref[i]) <= 1e-6f);
        /* adjoint identity <interp(F), v> == <F, interp_adj(v)> (test/test_interp_adjoint.jl:49-65) */
        float F[42], g[2];
        for (int i = 0; i < 42; i++) F[i] = frand(&seed);
        CHECK(p_dincae_interp(ctx, pos, 2, F, 6, 7, g, DINCAE_HOST));
        double lhs = (double)g[0] * vals[0] + (double)g[1] * vals[1], rhs = 0;
        for (int i = 0; i < 42; i++) rhs += (double)F[i] * A[i];
        EXPECT(fabs(lhs - rhs) <= 1e-5);
    }
    CHECK(p_dincae_sync(ctx));
    CHECK(p_dincae_model_destroy(m));
    CHECK(p_dincae_ctx_destroy(ctx));
    free(w0); free(w1); free(xin); free(xt); free(out);
    printf("c_abi_smoke: OK (%lld parameters, losses %.6f %.6f %.6f)\n", (long long)np, l[0], l[1], l[2]);
    return 0;
}
