/*
 * pva_oracle.c -- plain-C restatement of the fused arithmetic of the reference
 * transform chain.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): it is the
 * checker in tests/ and smoke(), never a product path and never a fallback.
 *
 * What it restates, per output element:
 *
 *   run.py:82    UniformTemporalSubsample  -> frame gather through idx1[]
 *   run.py:83    Div255                    -> u / 255.0f          (IEEE fp32 division)
 *   run.py:84    Normalize(mean, std)      -> (x - mean) / std    (IEEE fp32 division)
 *   run.py:89/95 (Random)ShortSideScale    -> ATen upsample_bilinear2d, align_corners=False:
 *                                             src = max(0, fma(scale, dst + 0.5, -0.5)),
 *                                             x-pass then y-pass, each fma(w0, v0, w1 * v1)
 *   run.py:90/96 RandomCrop / CenterCrop   -> window (top, left, crop)
 *   run.py:91    RandomHorizontalFlip      -> column mirror inside the window
 *   run.py:59-63 PackPathway               -> slow[j] = fast[idx2[j]]
 *
 * The temporal index tables (torch.linspace(...).long()) are inputs: they are not
 * closed-form in integer arithmetic and come from torch on the host.
 *
 * Build: see oracle/Makefile (gcc -O2 -mfma -ffp-contract=off).  Contraction is
 * switched off so that the only fused operations are the explicit fmaf() calls.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

/* lut[c*256 + u] = ((u / 255) - mean_c) / std_c, all in fp32 with true divisions. */
void pva_oracle_build_lut(const float* mean, const float* std, int C, float* lut) {
    for (int c = 0; c < C; ++c) {
        for (int u = 0; u < 256; ++u) {
            volatile float q = (float)u / 255.0f;
            volatile float d = q - mean[c];
            lut[c * 256 + u] = d / std[c];
        }
    }
}

/* ATen area_pixel_compute_source_index for align_corners=False, cubic=False. */
static inline float src_coord(float scale, int dst) {
    float s = fmaf(scale, (float)dst + 0.5f, -0.5f);
    return s < 0.0f ? 0.0f : s;
}

static inline void taps(float scale, int dst, int in_size, int* i0, int* i1, float* l0, float* l1) {
    float s = src_coord(scale, dst);
    int a = (int)floorf(s);
    if (a > in_size - 1) a = in_size - 1;
    int b = a + 1;
    if (b > in_size - 1) b = in_size - 1;
    float w1 = s - (float)a;
    /* ATen clamps lambda to [0, 1] */
    if (w1 < 0.0f) w1 = 0.0f;
    if (w1 > 1.0f) w1 = 1.0f;
    *i0 = a;
    *i1 = b;
    *l1 = w1;
    *l0 = 1.0f - w1;
}

/*
 * One clip.  src: uint8 [C][T][H][W].  fast: fp32 [C][n_fast][crop][crop] (contiguous,
 * the reference's own layout).  slow: fp32 [C][n_slow][crop][crop] or NULL.
 * (oh, ow) is the resized frame size, (top, left) the crop origin inside it.
 * Returns 0, or -1 on an impossible window.
 */
int pva_oracle_clip(const uint8_t* src, int C, int T, int H, int W,
                    const int32_t* idx1, int n_fast, const int32_t* idx2, int n_slow,
                    int oh, int ow, int top, int left, int flip, int crop,
                    const float* lut, float* fast, float* slow) {
    if (top < 0 || left < 0 || top + crop > oh || left + crop > ow) return -1;
    const float sy = (float)H / (float)oh;
    const float sx = (float)W / (float)ow;
    for (int c = 0; c < C; ++c) {
        const float* l = lut + c * 256;
        for (int k = 0; k < n_fast; ++k) {
            int t = idx1[k];
            if (t < 0 || t >= T) return -1;
            const uint8_t* plane = src + ((size_t)c * T + t) * H * W;
            float* dst = fast + ((size_t)c * n_fast + k) * crop * crop;
            for (int y = 0; y < crop; ++y) {
                int y0, y1;
                float ly0, ly1;
                taps(sy, top + y, H, &y0, &y1, &ly0, &ly1);
                for (int x = 0; x < crop; ++x) {
                    int col = left + (flip ? crop - 1 - x : x);
                    int x0, x1;
                    float lx0, lx1;
                    taps(sx, col, W, &x0, &x1, &lx0, &lx1);
                    float v00 = l[plane[(size_t)y0 * W + x0]];
                    float v01 = l[plane[(size_t)y0 * W + x1]];
                    float v10 = l[plane[(size_t)y1 * W + x0]];
                    float v11 = l[plane[(size_t)y1 * W + x1]];
                    float tp = fmaf(lx0, v00, lx1 * v01);
                    float bt = fmaf(lx0, v10, lx1 * v11);
                    dst[(size_t)y * crop + x] = fmaf(ly0, tp, ly1 * bt);
                }
            }
        }
        if (slow) {
            for (int j = 0; j < n_slow; ++j) {
                int k = idx2[j];
                if (k < 0 || k >= n_fast) return -1;
                const float* s = fast + ((size_t)c * n_fast + k) * crop * crop;
                float* d = slow + ((size_t)c * n_slow + j) * crop * crop;
                for (size_t i = 0; i < (size_t)crop * crop; ++i) d[i] = s[i];
            }
        }
    }
    return 0;
}

/* Source footprint of the crop window: distinct source rows / cols it touches.
 * Used for the algorithmic-bytes figure. */
void pva_oracle_footprint(int H, int W, int oh, int ow, int top, int left, int crop, int* fh, int* fw) {
    const float sy = (float)H / (float)oh;
    const float sx = (float)W / (float)ow;
    int a0, a1, b0, b1;
    float u, v;
    taps(sy, top, H, &a0, &a1, &u, &v);
    taps(sy, top + crop - 1, H, &b0, &b1, &u, &v);
    *fh = b1 - a0 + 1;
    taps(sx, left, W, &a0, &a1, &u, &v);
    taps(sx, left + crop - 1, W, &b0, &b1, &u, &v);
    *fw = b1 - a0 + 1;
}
