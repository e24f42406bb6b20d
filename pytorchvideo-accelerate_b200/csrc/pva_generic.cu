// pva_generic.cu -- the shape-agnostic kernel: one thread per output pixel, all channels.
//
// It implements the whole chain of reference run.py:81-100 for any geometry the ABI accepts
// (odd widths, down- or up-scaling, any crop, 1..4 channels, both layouts, three dtypes).
// It is the kernel the fp32 bit-exactness tests run (EXACT mode reproduces ATen's arithmetic,
// SURVEY.md App. A.5) and what PVA_KERNEL_AUTO falls back to when the tiled bulk-copy kernel
// (pva_tiled.cu) does not support a shape.  Taps are read straight from global memory
// through the read-only path; neighbouring threads share sectors, so L1/L2 absorb the reuse.
#include "pva_kernels.cuh"

namespace pva {

template <typename OutT, int kLayout, int kNorm, typename SrcT>
__global__ void __launch_bounds__(256) pva_generic_kernel(const __grid_constant__ KParams p) {
    __shared__ float s_lut[kMaxChannels * 256];
    if (kNorm == kNormExact && sizeof(SrcT) == 1) {
        for (int i = threadIdx.x; i < p.C * 256; i += blockDim.x) s_lut[i] = p.lut[i];
        __syncthreads();
    }
    const int clip_i = blockIdx.z;
    const int k = blockIdx.y;  // fast frame
    const ClipGeom& g = p.clips[clip_i];
    const int crop = p.crop;
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= crop * crop) return;
    const int y = pix / crop;
    const int x = pix - y * crop;

    int y0, y1, x0, x1;
    float ly0, ly1, lx0, lx1;
    bilinear_taps(g.sy, g.top + y, g.h, y0, y1, ly0, ly1);
    const int col = g.left + ((g.flip & 1) ? crop - 1 - x : x);
    bilinear_taps(g.sx, col, g.w, x0, x1, lx0, lx1);

    const int t_src = p.frame_idx[g.idx_row][k];
    const size_t plane_sz = (size_t)g.h * g.w;
    OutT* fast = reinterpret_cast<OutT*>(p.fast_out) + (size_t)clip_i * p.fast_clip_elems;
    OutT* slow = p.slow_out ? reinterpret_cast<OutT*>(p.slow_out) + (size_t)clip_i * p.slow_clip_elems : nullptr;

    for (int c = 0; c < p.C; ++c) {
        const SrcT* plane = reinterpret_cast<const SrcT*>(g.src) + ((size_t)c * g.t_in + t_src) * plane_sz;
        const SrcT u00 = __ldg(plane + (size_t)y0 * g.w + x0);
        const SrcT u01 = __ldg(plane + (size_t)y0 * g.w + x1);
        const SrcT u10 = __ldg(plane + (size_t)y1 * g.w + x0);
        const SrcT u11 = __ldg(plane + (size_t)y1 * g.w + x1);
        float r;
        if (kNorm == kNormExact) {
            float v00, v01, v10, v11;
            if (sizeof(SrcT) == 1) {
                const float* l = s_lut + c * 256;
                v00 = l[(int)u00], v01 = l[(int)u01], v10 = l[(int)u10], v11 = l[(int)u11];
            } else {
                // float frames: Div255 and Normalize as the reference evaluates them, per tap
                const float m = p.mean[c], sd = p.std[c];
                v00 = __fdiv_rn(__fdiv_rn((float)u00, 255.0f) - m, sd);
                v01 = __fdiv_rn(__fdiv_rn((float)u01, 255.0f) - m, sd);
                v10 = __fdiv_rn(__fdiv_rn((float)u10, 255.0f) - m, sd);
                v11 = __fdiv_rn(__fdiv_rn((float)u11, 255.0f) - m, sd);
            }
            const float tp = fmaf(lx0, v00, lx1 * v01);
            const float bt = fmaf(lx0, v10, lx1 * v11);
            r = fmaf(ly0, tp, ly1 * bt);
        } else {
            const float tp = fmaf(lx0, (float)u00, lx1 * (float)u01);
            const float bt = fmaf(lx0, (float)u10, lx1 * (float)u11);
            const float pr = fmaf(ly0, tp, ly1 * bt);
            r = fmaf(pr, p.na[c], p.nb[c]);
        }
        const OutT o = to_out<OutT>(r);
        size_t off;
        if (kLayout == kLayoutCTHW) {
            off = (((size_t)c * p.n_fast + k) * crop + y) * crop + x;
        } else {
            off = ((((size_t)k * crop + y) * crop) + x) * p.C + c;
        }
        fast[off] = o;
        if (slow) {
            for (int j = 0; j < p.n_slow; ++j) {
                if (p.slow_idx[j] == k) {
                    size_t soff;
                    if (kLayout == kLayoutCTHW) {
                        soff = (((size_t)c * p.n_slow + j) * crop + y) * crop + x;
                    } else {
                        soff = ((((size_t)j * crop + y) * crop) + x) * p.C + c;
                    }
                    slow[soff] = o;
                }
            }
        }
    }
}

template <typename OutT>
static cudaError_t launch_generic_t(const KParams& p, int layout, int norm, cudaStream_t stream, dim3 grid,
                                    dim3 block) {
#define PVA_GENERIC_CASE(L, N)                                                          \
    if (layout == L && norm == N) {                                                     \
        if (p.src_f32)                                                                  \
            pva_generic_kernel<OutT, L, N, float><<<grid, block, 0, stream>>>(p);       \
        else                                                                            \
            pva_generic_kernel<OutT, L, N, uint8_t><<<grid, block, 0, stream>>>(p);     \
        return cudaGetLastError();                                                      \
    }
    PVA_GENERIC_CASE(kLayoutCTHW, kNormExact)
    PVA_GENERIC_CASE(kLayoutCTHW, kNormAffine)
    PVA_GENERIC_CASE(kLayoutTHWC, kNormExact)
    PVA_GENERIC_CASE(kLayoutTHWC, kNormAffine)
#undef PVA_GENERIC_CASE
    return cudaErrorInvalidValue;
}

cudaError_t launch_generic(const KParams& p, int out_dtype, int layout, int norm, cudaStream_t stream, int* grid_out,
                           int* block_out) {
    const dim3 block(256);
    const dim3 grid((unsigned)((p.crop * p.crop + 255) / 256), (unsigned)p.n_fast, (unsigned)p.n_clips);
    if (grid_out) *grid_out = (int)(grid.x * grid.y * grid.z);
    if (block_out) *block_out = 256;
    switch (out_dtype) {
        case 0: return launch_generic_t<float>(p, layout, norm, stream, grid, block);
        case 1: return launch_generic_t<__half>(p, layout, norm, stream, grid, block);
        case 2: return launch_generic_t<__nv_bfloat16>(p, layout, norm, stream, grid, block);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace pva
