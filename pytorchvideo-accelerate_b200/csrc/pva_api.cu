// pva_api.cu -- the C ABI of include/pva_preproc.h: validation, descriptor packing, dispatch.
//
// Host-side only.  One call = one or more asynchronous kernel launches on the caller's
// stream; nothing is allocated, copied or kept.  The geometry that the reference computes in
// Python doubles (resized size, crop origin; run.py:89-96) arrives precomputed in the
// PvaClipDesc; what is derived here is only what ATen derives in fp32 (the two scale
// factors) and the normalisation constants.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/pva_preproc.h"
#include "pva_kernels.cuh"

namespace {

thread_local char g_err[512] = "";
thread_local PvaLaunchInfo g_info = {};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// ATen: area_pixel_compute_source_index + guard_index_and_lambda (host mirror, for footprints).
void host_taps(float scale, int dst, int in_size, int* i0, int* i1) {
    float s = fmaf(scale, (float)dst + 0.5f, -0.5f);
    if (s < 0.0f) s = 0.0f;
    int a = (int)floorf(s);
    if (a > in_size - 1) a = in_size - 1;
    int b = a + 1;
    if (b > in_size - 1) b = in_size - 1;
    *i0 = a;
    *i1 = b;
}

}  // namespace

extern "C" {

int32_t pva_version(void) { return (PVA_VERSION_MAJOR << 16) | PVA_VERSION_MINOR; }

const char* pva_last_error_string(void) { return g_err; }

int pva_last_launch_info(PvaLaunchInfo* info) {
    if (!info) return fail(PVA_ERR_INVALID_ARGUMENT, "pva_last_launch_info: info is NULL");
    *info = g_info;
    return PVA_OK;
}

int pva_build_norm_lut(const float* mean, const float* std, int32_t channels, float* lut) {
    if (!mean || !std || !lut) return fail(PVA_ERR_INVALID_ARGUMENT, "pva_build_norm_lut: NULL argument");
    if (channels < 1 || channels > PVA_MAX_CHANNELS)
        return fail(PVA_ERR_INVALID_ARGUMENT, "pva_build_norm_lut: channels=%d outside 1..%d", channels,
                    PVA_MAX_CHANNELS);
    for (int c = 0; c < channels; ++c) {
        if (std[c] == 0.0f)
            return fail(PVA_ERR_INVALID_ARGUMENT, "std evaluated to zero after conversion to float32 (channel %d)", c);
        for (int u = 0; u < 256; ++u) {
            // Div255: x / 255.0 ; Normalize: (x - mean) / std -- true IEEE divisions, in this order.
            volatile float q = (float)u / 255.0f;
            volatile float d = q - mean[c];
            lut[c * 256 + u] = d / std[c];
        }
    }
    return PVA_OK;
}

int pva_source_footprint(int32_t h, int32_t w, int32_t oh, int32_t ow, int32_t top, int32_t left, int32_t crop,
                         int32_t* fh, int32_t* fw) {
    if (!fh || !fw || h < 1 || w < 1 || oh < 1 || ow < 1 || crop < 1 || top < 0 || left < 0 || top + crop > oh ||
        left + crop > ow)
        return fail(PVA_ERR_INVALID_ARGUMENT, "pva_source_footprint: bad window");
    const float sy = (float)h / (float)oh, sx = (float)w / (float)ow;
    int a0, a1, b0, b1;
    host_taps(sy, top, h, &a0, &a1);
    host_taps(sy, top + crop - 1, h, &b0, &b1);
    *fh = b1 - a0 + 1;
    host_taps(sx, left, w, &a0, &a1);
    host_taps(sx, left + crop - 1, w, &b0, &b1);
    *fw = b1 - a0 + 1;
    return PVA_OK;
}

int pva_upload_selected_frames(const void* host_src, void* dev_dst, int32_t channels, int32_t t_in, int32_t h,
                               int32_t w, const int32_t* frame_idx, int32_t n_frames, void* stream_v) {
    if (!host_src || !dev_dst || !frame_idx)
        return fail(PVA_ERR_INVALID_ARGUMENT, "pva_upload_selected_frames: NULL argument");
    if (channels < 1 || t_in < 1 || h < 1 || w < 1 || n_frames < 1)
        return fail(PVA_ERR_INVALID_ARGUMENT, "pva_upload_selected_frames: empty shape");
    for (int j = 0; j < n_frames; ++j)
        if (frame_idx[j] < 0 || frame_idx[j] >= t_in)
            return fail(PVA_ERR_INVALID_ARGUMENT, "frame_idx[%d]=%d outside [0, %d)", j, frame_idx[j], t_in);
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    const size_t plane = (size_t)h * w;
    const unsigned char* src = static_cast<const unsigned char*>(host_src);
    unsigned char* dst = static_cast<unsigned char*>(dev_dst);
    for (int c = 0; c < channels; ++c) {
        int j = 0;
        while (j < n_frames) {
            // longest run j, j+1, ... whose source frames are equally spaced (forward)
            int len = 1;
            if (j + 1 < n_frames && frame_idx[j + 1] > frame_idx[j]) {
                const int step = frame_idx[j + 1] - frame_idx[j];
                len = 2;
                while (j + len < n_frames && frame_idx[j + len] - frame_idx[j + len - 1] == step) ++len;
            }
            const size_t spitch = len > 1 ? (size_t)(frame_idx[j + 1] - frame_idx[j]) * plane : plane;
            cudaError_t ce = cudaMemcpy2DAsync(dst + ((size_t)c * n_frames + j) * plane, plane,
                                               src + ((size_t)c * t_in + frame_idx[j]) * plane, spitch, plane,
                                               (size_t)len, cudaMemcpyHostToDevice, stream);
            if (ce != cudaSuccess) {
                (void)cudaGetLastError();
                return fail(PVA_ERR_CUDA, "cudaMemcpy2DAsync failed: %s", cudaGetErrorString(ce));
            }
            j += len;
        }
    }
    return PVA_OK;
}

int pva_preprocess(const PvaClipDesc* clips, int32_t n_clips, const PvaPlan* plan, void* fast_out, void* slow_out,
                   void* stream_v) {
    using namespace pva;
    g_info = PvaLaunchInfo{};
    if (!plan) return fail(PVA_ERR_INVALID_ARGUMENT, "plan is NULL");
    if (n_clips < 0) return fail(PVA_ERR_INVALID_ARGUMENT, "n_clips=%d is negative", n_clips);
    if (plan->channels < 1 || plan->channels > PVA_MAX_CHANNELS)
        return fail(PVA_ERR_UNSUPPORTED, "channels=%d outside 1..%d", plan->channels, PVA_MAX_CHANNELS);
    if (plan->n_fast < 1 || plan->n_fast > kMaxFrames)
        return fail(PVA_ERR_UNSUPPORTED, "n_fast=%d outside 1..%d", plan->n_fast, kMaxFrames);
    if (plan->n_slow < 0 || plan->n_slow > kMaxFrames)
        return fail(PVA_ERR_UNSUPPORTED, "n_slow=%d outside 0..%d", plan->n_slow, kMaxFrames);
    if (plan->crop < 1) return fail(PVA_ERR_INVALID_ARGUMENT, "crop=%d must be positive", plan->crop);
    if (plan->out_dtype < PVA_F32 || plan->out_dtype > PVA_BF16)
        return fail(PVA_ERR_INVALID_ARGUMENT, "out_dtype=%d is not a PvaDType", plan->out_dtype);
    if (plan->out_layout != PVA_LAYOUT_CTHW && plan->out_layout != PVA_LAYOUT_THWC)
        return fail(PVA_ERR_INVALID_ARGUMENT, "out_layout=%d is not a PvaLayout", plan->out_layout);
    if (plan->norm_mode != PVA_NORM_EXACT && plan->norm_mode != PVA_NORM_AFFINE)
        return fail(PVA_ERR_INVALID_ARGUMENT, "norm_mode=%d is not a PvaNormMode", plan->norm_mode);
    if (plan->kernel < PVA_KERNEL_AUTO || plan->kernel > PVA_KERNEL_TILED)
        return fail(PVA_ERR_INVALID_ARGUMENT, "kernel=%d is not a PvaKernel", plan->kernel);
    if (plan->src_type != PVA_SRC_U8 && plan->src_type != PVA_SRC_F32)
        return fail(PVA_ERR_INVALID_ARGUMENT, "src_type=%d is not a PvaSrcType", plan->src_type);
    if (plan->max_ctas < 0) return fail(PVA_ERR_INVALID_ARGUMENT, "max_ctas=%d is negative", plan->max_ctas);
    if (plan->flags & ~(PVA_FLAG_CLIPS_READY | PVA_FLAG_OUTPUTS_DISJOINT))
        return fail(PVA_ERR_INVALID_ARGUMENT, "flags=0x%x has bits that are not PVA_FLAG_*", plan->flags);
    if ((plan->n_slow == 0) != (slow_out == nullptr))
        return fail(PVA_ERR_INVALID_ARGUMENT, "slow_out must be NULL exactly when n_slow == 0 (n_slow=%d)",
                    plan->n_slow);
    if (plan->n_slow > 0 && !plan->slow_idx) return fail(PVA_ERR_INVALID_ARGUMENT, "slow_idx is NULL with n_slow > 0");
    for (int j = 0; j < plan->n_slow; ++j)
        if (plan->slow_idx[j] < 0 || plan->slow_idx[j] >= plan->n_fast)
            return fail(PVA_ERR_INVALID_ARGUMENT, "slow_idx[%d]=%d outside [0, %d)", j, plan->slow_idx[j],
                        plan->n_fast);
    if (n_clips == 0) return PVA_OK;  // an empty batch is a no-op
    if (!clips) return fail(PVA_ERR_INVALID_ARGUMENT, "clips is NULL");
    if (!fast_out) return fail(PVA_ERR_INVALID_ARGUMENT, "fast_out is NULL");

    const int C = plan->channels, crop = plan->crop;
    for (int i = 0; i < n_clips; ++i) {
        const PvaClipDesc& d = clips[i];
        if (!d.src) return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: src is NULL", i);
        if (!d.frame_idx) return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: frame_idx is NULL", i);
        if (d.t_in < 1 || d.h < 1 || d.w < 1)
            return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: empty clip %dx%dx%d", i, d.t_in, d.h, d.w);
        if (d.t_in > 65535) return fail(PVA_ERR_UNSUPPORTED, "clip %d: t_in=%d exceeds 65535", i, d.t_in);
        if (d.oh < 1 || d.ow < 1)
            return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: resized size %dx%d must be positive", i, d.oh, d.ow);
        if (d.oh < crop || d.ow < crop)
            return fail(PVA_ERR_CROP_TOO_LARGE, "Required crop size (%d, %d) is larger than input image size (%d, %d)",
                        crop, crop, d.oh, d.ow);
        if (d.top < 0 || d.left < 0 || d.top + crop > d.oh || d.left + crop > d.ow)
            return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: crop window top=%d left=%d size=%d outside %dx%d", i,
                        d.top, d.left, crop, d.oh, d.ow);
        for (int k = 0; k < plan->n_fast; ++k)
            if (d.frame_idx[k] < 0 || d.frame_idx[k] >= d.t_in)
                return fail(PVA_ERR_INVALID_ARGUMENT, "clip %d: frame_idx[%d]=%d outside [0, %d)", i, k,
                            d.frame_idx[k], d.t_in);
    }

    KParams kp;
    memset(&kp, 0, sizeof(kp));
    kp.C = C;
    kp.n_fast = plan->n_fast;
    kp.n_slow = plan->n_slow;
    kp.crop = crop;
    kp.src_f32 = plan->src_type == PVA_SRC_F32 ? 1 : 0;
    kp.clips_ready = (plan->flags & PVA_FLAG_CLIPS_READY) ? 1 : 0;
    kp.outputs_disjoint = (plan->flags & PVA_FLAG_OUTPUTS_DISJOINT) ? 1 : 0;
    kp.max_ctas = plan->max_ctas;
    kp.fast_clip_elems = (long long)C * plan->n_fast * crop * crop;
    kp.slow_clip_elems = (long long)C * plan->n_slow * crop * crop;
    for (int j = 0; j < plan->n_slow; ++j) kp.slow_idx[j] = (uint16_t)plan->slow_idx[j];
    for (int k = 0; k < kMaxFrames; ++k) kp.slow_of[k] = kp.slow_next[k] = kNoSlot;
    for (int j = plan->n_slow - 1; j >= 0; --j) {  // back to front: the chains come out in ascending slot order
        const int k = plan->slow_idx[j];
        kp.slow_next[j] = kp.slow_of[k];
        kp.slow_of[k] = (uint16_t)j;
    }
    {
        // the 256-entry table (768 IEEE divisions) is only read in EXACT mode on uint8 clips
        if (plan->norm_mode == PVA_NORM_EXACT && plan->src_type == PVA_SRC_U8) {
            int rc = pva_build_norm_lut(plan->mean, plan->std, C, kp.lut);
            if (rc != PVA_OK) return rc;
        }
        for (int c = 0; c < C; ++c) {
            if (plan->std[c] == 0.0f)
                return fail(PVA_ERR_INVALID_ARGUMENT, "std evaluated to zero after conversion to float32 (channel %d)", c);
            const double m = (double)plan->mean[c], s = (double)plan->std[c];
            kp.mean[c] = plan->mean[c];
            kp.std[c] = plan->std[c];
            kp.na[c] = (float)(1.0 / (255.0 * s));
            kp.nb[c] = (float)(-m / s);
        }
    }

    const size_t esize = plan->out_dtype == PVA_F32 ? 4 : 2;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    int done = 0;
    while (done < n_clips) {
        // Pack as many clips as fit one parameter block (clip slots, distinct index tables).
        int rows = 0, n = 0;
        while (done + n < n_clips && n < kMaxClipsPerLaunch) {
            const PvaClipDesc& d = clips[done + n];
            uint16_t row[kMaxFrames];
            for (int k = 0; k < plan->n_fast; ++k) row[k] = (uint16_t)d.frame_idx[k];
            int r = 0;
            for (; r < rows; ++r)
                if (memcmp(kp.frame_idx[r], row, sizeof(uint16_t) * plan->n_fast) == 0) break;
            if (r == rows) {
                if (rows == kMaxIdxRows) break;  // table full: flush, continue in the next launch
                memcpy(kp.frame_idx[rows++], row, sizeof(uint16_t) * plan->n_fast);
            }
            ClipGeom& g = kp.clips[n];
            g.src = static_cast<const uint8_t*>(d.src);
            g.t_in = d.t_in;
            g.h = d.h;
            g.w = d.w;
            g.oh = d.oh;
            g.ow = d.ow;
            g.top = d.top;
            g.left = d.left;
            g.sy = (float)d.h / (float)d.oh;
            g.sx = (float)d.w / (float)d.ow;
            g.idx_row = (uint16_t)r;
            g.flip = d.flip ? 1 : 0;
            ++n;
        }
        kp.n_clips = n;
        kp.fast_out = static_cast<char*>(fast_out) + (size_t)done * kp.fast_clip_elems * esize;
        kp.slow_out = slow_out ? static_cast<char*>(slow_out) + (size_t)done * kp.slow_clip_elems * esize : nullptr;

        bool use_tiled = false;
        cudaError_t ce = cudaSuccess;
        if (plan->kernel != PVA_KERNEL_GENERIC) {
            TiledConfig cfg = {};
            const char* why = nullptr;
            ce = launch_tiled(kp, plan->out_dtype, plan->out_layout, plan->norm_mode, stream, &cfg, &why);
            use_tiled = why == nullptr;  // planned and launched (or the launch itself failed)
            if (!use_tiled) {
                if (plan->kernel == PVA_KERNEL_TILED)
                    return fail(PVA_ERR_UNSUPPORTED, "tiled kernel cannot take this launch: %s", why);
                ce = cudaSuccess;
            } else {
                g_info.kernel = PVA_KERNEL_TILED;
                g_info.grid = cfg.grid;
                g_info.block = cfg.block;
                g_info.smem_bytes = cfg.smem_bytes;
                g_info.rows_per_tile = cfg.rows_per_tile;
                g_info.tiles += cfg.tiles;
            }
        }
        if (!use_tiled) {
            int grid = 0, block = 0;
            ce = launch_generic(kp, plan->out_dtype, plan->out_layout, plan->norm_mode, stream, &grid, &block);
            g_info.kernel = PVA_KERNEL_GENERIC;
            g_info.grid = grid;
            g_info.block = block;
            g_info.smem_bytes = 0;
            g_info.rows_per_tile = 0;
            g_info.tiles += grid;
        }
        if (ce != cudaSuccess) {
            (void)cudaGetLastError();  // do not leave the error for an unrelated later call to trip over
            return fail(PVA_ERR_CUDA, "%s kernel launch failed: %s", use_tiled ? "tiled" : "generic",
                        cudaGetErrorString(ce));
        }
        g_info.launches += 1;
        done += n;
    }
    return PVA_OK;
}

}  // extern "C"
