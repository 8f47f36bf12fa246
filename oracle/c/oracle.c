/* Plain-C restatement of the DINCAE hot path -- TEST INFRASTRUCTURE ONLY
 * (see oracle/__init__.py for the pinning status).  Arrays use the reference's
 * own memory order: Julia column-major WHCN, 0-based here.
 * Second, independent restatement beside oracle/model.py (torch ops): the two
 * are checked against each other in tests/test_oracle_golden.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define IDX4(i, j, c, b, W, H, C) ((size_t)(i) + (size_t)(W) * ((size_t)(j) + (size_t)(H) * ((size_t)(c) + (size_t)(C) * (size_t)(b))))

/* src/points.jl:6-26: bilinear gather; pos 1-based interleaved (x,y). vec must be zeroed by caller. */
void orc_interpnd(const float* pos, int64_t P, const float* A, int W, int H, float* vec) {
    (void)H;
    for (int64_t p = 0; p < P; p++) {
        float px = pos[2 * p], py = pos[2 * p + 1];
        int ix = (int)floorf(px), iy = (int)floorf(py);
        float cx = px - (float)ix, cy = py - (float)iy;
        for (int oj = 0; oj < 2; oj++)
            for (int oi = 0; oi < 2; oi++) {
                float cc = (oi ? cx : 1.0f - cx) * (oj ? cy : 1.0f - cy);
                vec[p] += A[(ix - 1 + oi) + (size_t)W * (iy - 1 + oj)] * cc;
            }
    }
}

/* src/points.jl:52-71: adjoint scatter-add on a zeroed grid (sequential order). */
void orc_interp_adjn(const float* pos, const float* values, int64_t P, int W, int H, float* A2) {
    memset(A2, 0, sizeof(float) * (size_t)W * H);
    for (int64_t p = 0; p < P; p++) {
        float px = pos[2 * p], py = pos[2 * p + 1];
        int ix = (int)floorf(px), iy = (int)floorf(py);
        float cx = px - (float)ix, cy = py - (float)iy;
        for (int oj = 0; oj < 2; oj++)
            for (int oi = 0; oi < 2; oi++) {
                float cc = (oi ? cx : 1.0f - cx) * (oj ? cy : 1.0f - cy);
                A2[(ix - 1 + oi) + (size_t)W * (iy - 1 + oj)] += values[p] * cc;
            }
    }
}

/* Flux Conv((3,3), cin=>cout, pad=1) = NNlib true convolution (kernel flipped):
 * out[x,y,co] = b[co] + sum_{a,b,ci} Wt[a,b,ci,co] * in[x+1-a, y+1-b, ci]; act: 0 identity, 1 relu.
 * Accumulates in double (the "exact" answer). */
void orc_conv3x3(const float* in, int W, int H, int Cin, int B, const float* Wt, const float* bias,
                 int Cout, int act, float* out) {
    for (int b = 0; b < B; b++)
        for (int co = 0; co < Cout; co++)
            for (int y = 0; y < H; y++)
                for (int x = 0; x < W; x++) {
                    double s = bias[co];
                    for (int ci = 0; ci < Cin; ci++)
                        for (int kb = 0; kb < 3; kb++) {
                            int yy = y + 1 - kb;
                            if (yy < 0 || yy >= H) continue;
                            for (int ka = 0; ka < 3; ka++) {
                                int xx = x + 1 - ka;
                                if (xx < 0 || xx >= W) continue;
                                s += (double)Wt[ka + 3 * (kb + 3 * (ci + (size_t)Cin * co))] *
                                     (double)in[IDX4(xx, yy, ci, b, W, H, Cin)];
                            }
                        }
                    if (act == 1 && s < 0) s = 0;
                    out[IDX4(x, y, co, b, W, H, Cout)] = (float)s;
                }
}

/* MeanPool((2,2), pad=odd): window 2, stride 2, symmetric pad, divide by 4 always. */
void orc_meanpool2(const float* in, int W, int H, int C, int B, float* out) {
    int ow = W % 2, oh = H % 2, W2 = W / 2 + ow, H2 = H / 2 + oh;
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++)
            for (int y = 0; y < H2; y++)
                for (int x = 0; x < W2; x++) {
                    double s = 0;
                    for (int dy = 0; dy < 2; dy++)
                        for (int dx = 0; dx < 2; dx++) {
                            int xx = 2 * x + dx - ow, yy = 2 * y + dy - oh;
                            if (xx >= 0 && xx < W && yy >= 0 && yy < H) s += in[IDX4(xx, yy, c, b, W, H, C)];
                        }
                    out[IDX4(x, y, c, b, W2, H2, C)] = (float)(s / 4.0);
                }
}

/* Upsample(scale=2) nearest (method 0) or bilinear align_corners=true (method 1), then
 * croppadding: keep the first Wo x Ho (src/model.jl:232-235). */
void orc_upsample2_crop(const float* in, int W, int H, int C, int B, int method, int Wo, int Ho, float* out) {
    int W2 = 2 * W, H2 = 2 * H;
    for (int b = 0; b < B; b++)
        for (int c = 0; c < C; c++)
            for (int y = 0; y < Ho; y++)
                for (int x = 0; x < Wo; x++) {
                    double v;
                    if (method == 0) {
                        v = in[IDX4(x / 2, y / 2, c, b, W, H, C)];
                    } else {
                        double sx = (W2 > 1) ? (double)x * (W - 1) / (double)(W2 - 1) : 0.0;
                        double sy = (H2 > 1) ? (double)y * (H - 1) / (double)(H2 - 1) : 0.0;
                        int x0 = (int)floor(sx), y0 = (int)floor(sy);
                        int x1 = x0 + 1 < W ? x0 + 1 : W - 1, y1 = y0 + 1 < H ? y0 + 1 : H - 1;
                        double fx = sx - x0, fy = sy - y0;
                        v = (1 - fy) * ((1 - fx) * in[IDX4(x0, y0, c, b, W, H, C)] + fx * in[IDX4(x1, y0, c, b, W, H, C)]) +
                            fy * ((1 - fx) * in[IDX4(x0, y1, c, b, W, H, C)] + fx * in[IDX4(x1, y1, c, b, W, H, C)]);
                    }
                    out[IDX4(x, y, c, b, Wo, Ho, C)] = (float)v;
                }
}

/* transform_m sigma2 (src/model.jl:16-41) on (W,H,2*nout,B) in place layout -> out. */
void orc_head(const float* x, int64_t npix, int nout, int B, float gamma, float* out) {
    for (int b = 0; b < B; b++)
        for (int v = 0; v < nout; v++) {
            const float* a = x + npix * (size_t)(2 * v + 2 * nout * b);
            const float* l = a + npix;
            float* m = out + npix * (size_t)(2 * v + 2 * nout * b);
            float* s = m + npix;
            for (int64_t p = 0; p < npix; p++) {
                float inv = expf(fminf(l[p], gamma));
                float s2 = 1.0f / fmaxf(inv, 1e-3f);
                m[p] = a[p] * s2;
                s[p] = s2;
            }
        }
}

/* costfun (src/model.jl:54-148), gridded, no Laplacian; double accumulation. */
double orc_costfun(const float* xrec, const float* xtrue, int64_t npix, int nout, int B, int truth_uncertain) {
    double total = 0;
    for (int v = 0; v < nout; v++) {
        double slog = 0, sd = 0, n = 0;
        for (int b = 0; b < B; b++) {
            size_t o = npix * (size_t)(2 * v + 2 * nout * b);
            for (int64_t p = 0; p < npix; p++) {
                float m = xrec[o + p], s2 = xrec[o + npix + p];
                float t1 = xtrue[o + p], t2 = xtrue[o + npix + p];
                if (t2 == 0) continue;
                float s2t = 1.0f / fmaxf(t2, 1e-3f);
                float mt = t1 * s2t;
                double d = (double)m - mt;
                n += 1;
                if (truth_uncertain) { slog += log((double)s2 / s2t); sd += (d * d + s2t) / s2; }
                else { slog += log((double)s2); sd += d * d / s2; }
            }
        }
        total += (slog + sd) / n;
    }
    return total;
}

/* ClipGrad + Adam, src/flux.jl:10-14 (Optimisers.jl rule, float32 state). bt1/bt2 = beta^t AFTER this step. */
void orc_adam(float* w, const float* g, float* m, float* v, int64_t n, float lr, float clip,
              float b1, float b2, float eps, float bt1, float bt2) {
    for (int64_t i = 0; i < n; i++) {
        float gi = g[i];
        if (clip > 0) gi = fminf(fmaxf(gi, -clip), clip);
        m[i] = b1 * m[i] + (1 - b1) * gi;
        v[i] = b2 * v[i] + (1 - b2) * gi * gi;
        float mh = m[i] / (1 - bt1), vh = v[i] / (1 - bt2);
        w[i] -= lr * mh / (sqrtf(vh) + eps);
    }
}

/* getxy! without perturbation (src/data.jl:391-466, :504-514): x is (W,H,V,N,2). ind 0-based. */
void orc_getxy(const float* x, int W, int H, int V, int N, int T, const float* lon_s, const float* lat_s,
               float tcos, float tsin, int ind, const uint8_t* isoutput, float* xin, float* xtrue) {
    size_t np_ = (size_t)W * H;
    for (int j = 0; j < H; j++)
        for (int i = 0; i < W; i++) {
            xin[i + (size_t)W * j] = lon_s[i];
            xin[np_ + i + (size_t)W * j] = lat_s[j];
            xin[2 * np_ + i + (size_t)W * j] = tcos;
            xin[3 * np_ + i + (size_t)W * j] = tsin;
        }
    int half = (T - 1) / 2, off = 4;
    for (int v = 0; v < V; v++)
        for (int k = 0; k < T; k++) {
            int n = ind - half + k;
            if (n < 0) n = 0;
            if (n > N - 1) n = N - 1;
            for (int e = 0; e < 2; e++)
                memcpy(xin + (size_t)(off + e) * np_, x + np_ * ((size_t)v + (size_t)V * ((size_t)n + (size_t)N * e)),
                       sizeof(float) * np_);
            off += 2;
        }
    int o = 0;
    for (int v = 0; v < V; v++)
        if (isoutput[v]) {
            for (int e = 0; e < 2; e++)
                memcpy(xtrue + (size_t)(o + e) * np_, x + np_ * ((size_t)v + (size_t)V * ((size_t)ind + (size_t)N * e)),
                       sizeof(float) * np_);
            o += 2;
        }
}
