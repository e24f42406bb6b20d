// pva_kernels.cuh -- shared between the host dispatch (pva_api.cu) and the kernels.
//
// Everything a launch needs travels BY VALUE as one __grid_constant__ kernel argument
// (CUDA 12 allows 32 KB of parameters on sm_70+): clip descriptors, the temporal index
// tables, the normalisation constants.  Hence no device-side tables to allocate, upload
// or keep alive, and the library stays stateless (SURVEY.md section 8b "threading").
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pva {

constexpr int kMaxClipsPerLaunch = 128;  // more clips -> the host issues several launches
constexpr int kMaxFrames = 128;          // n_fast limit
constexpr int kMaxIdxRows = 8;           // distinct temporal index tables per launch
constexpr int kMaxChannels = 4;
constexpr uint16_t kNoSlot = 0xffff;

// Geometry of one clip, precomputed on the host (pva_api.cu) from a PvaClipDesc.
struct ClipGeom {
    const uint8_t* src;  // device, uint8 [C][t_in][h][w]
    int t_in, h, w;
    int oh, ow;          // resized frame (reference: short_side_scale, run.py:89/95)
    int top, left;       // crop origin in the resized frame (run.py:90/96)
    float sy, sx;        // ATen area_pixel_compute_scale: float(in) / float(out)
    uint16_t idx_row;    // row of KParams::frame_idx this clip uses
    uint16_t flip;       // run.py:91
};

struct KParams {
    void* fast_out;
    void* slow_out;  // nullptr for a single pathway
    int n_clips, C, n_fast, n_slow, crop;
    int src_f32;        // source clips are float32 (0..255 valued) instead of uint8: generic kernel only
    int clips_ready;    // PVA_FLAG_CLIPS_READY: reading the clips need not wait for the previous kernel
    int outputs_disjoint;  // PVA_FLAG_OUTPUTS_DISJOINT: the previous kernel does not touch this launch's outputs
    int max_ctas;       // PvaPlan.max_ctas: grid cap of the tiled kernel (0 = one CTA per SM)
    long long fast_clip_elems;  // C * n_fast * crop * crop
    long long slow_clip_elems;
    float na[kMaxChannels];  // AFFINE: out = fma(p, na, nb), p = bilinear over raw bytes
    float nb[kMaxChannels];
    float mean[kMaxChannels];  // for float sources, where no table applies
    float std[kMaxChannels];
    float lut[kMaxChannels * 256];  // EXACT: ((u / 255) - mean) / std, built on the host
    uint16_t frame_idx[kMaxIdxRows][kMaxFrames];  // UniformTemporalSubsample tables (run.py:82)
    uint16_t slow_idx[kMaxFrames];                // PackPathway table (run.py:59-63)
    // the same table inverted: first slow slot fed by fast frame k (kNoSlot if none), and the next slot fed
    // by the same frame (a caller's table may repeat a frame; PackPathway's never does)
    uint16_t slow_of[kMaxFrames];
    uint16_t slow_next[kMaxFrames];
    ClipGeom clips[kMaxClipsPerLaunch];
};

enum : int { kLayoutCTHW = 0, kLayoutTHWC = 1 };
enum : int { kNormExact = 0, kNormAffine = 1 };

// ATen's area_pixel_compute_source_index (align_corners=False) + guarded index/lambda:
// src = max(0, fma(scale, dst + 0.5, -0.5)); i0 = min(floor(src), in-1); i1 = min(i0+1, in-1);
// l1 = clamp(src - i0, 0, 1); l0 = 1 - l1.   
__device__ __forceinline__ void bilinear_taps(float scale, int dst, int in_size, int& i0, int& i1, float& l0,
                                              float& l1) {
    float s = fmaf(scale, (float)dst + 0.5f, -0.5f);
    s = fmaxf(s, 0.0f);
    int a = (int)s;  // s >= 0: truncation == floor
    a = min(a, in_size - 1);
    i0 = a;
    i1 = min(a + 1, in_size - 1);
    float w1 = s - (float)a;
    w1 = fminf(fmaxf(w1, 0.0f), 1.0f);
    l1 = w1;
    l0 = 1.0f - w1;
}

template <typename T>
__device__ __forceinline__ T to_out(float v);
template <>
__device__ __forceinline__ float to_out<float>(float v) {
    return v;
}
template <>
__device__ __forceinline__ __half to_out<__half>(float v) {
    return __float2half_rn(v);
}
template <>
__device__ __forceinline__ __nv_bfloat16 to_out<__nv_bfloat16>(float v) {
    return __float2bfloat16_rn(v);
}

// Host-side launchers (one per translation unit).
cudaError_t launch_generic(const KParams& p, int out_dtype, int layout, int norm, cudaStream_t stream, int* grid,
                           int* block);

struct TiledConfig {
    int rows_per_tile;
    int smem_bytes;
    int grid;
    int block;
    int tiles;
};
// Plans (once) and launches.  When the tiled kernel cannot take this launch nothing is enqueued,
// *why names the reason and cudaErrorNotSupported comes back; otherwise *why is nullptr and the
// result is the launch's.
cudaError_t launch_tiled(const KParams& p, int out_dtype, int layout, int norm, cudaStream_t stream, TiledConfig* cfg,
                         const char** why);

}  // namespace pva
