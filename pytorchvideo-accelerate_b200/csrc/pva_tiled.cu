// pva_tiled.cu -- the production kernel: persistent, warp-specialised, bulk-copy staged.
//
// Work item ("tile") = (clip, fast frame k, band of R output rows), all C=3 channels.
// Per tile:
//   producer warp  : cp.async.bulk (TMA, 1-D) of the source row segments the band touches,
//                    three planes, into a shared-memory stage; completion on an mbarrier.
//                    Temporal subsampling (run.py:82) is folded into the source address.
//   consumer warps : each thread owns 4 adjacent output columns x 3 channels.  The x-pass is
//                    computed once per SOURCE row (two aligned 32-bit shared loads per plane,
//                    taps picked with PRMT) and kept in registers; consecutive output rows
//                    reuse it for the y-pass (bilinear, run.py:89/95).  Crop and flip
//                    (run.py:90-91) are folded into the per-thread column tables, Div255 +
//                    Normalize (run.py:83-84) into the epilogue (AFFINE) or a table (EXACT).
//   store          : outputs are packed to fp16/bf16/fp32 into a shared staging strip in final
//                    memory order and written with cp.async.bulk shared->global: once to the
//                    fast pathway and again to the slow pathway when the frame belongs to it
//                    (PackPathway, run.py:59-63) -- the source is read once.
//
// The source row pitch (340 B) is not a multiple of 16, so a 2-D tensor map is not legal; each
// row segment is copied from its 16-byte-aligned superset and the consumers add the per-row
// shift.  The library only ever reads 16-byte granules that contain at least one byte of the clip.
#include <type_traits>

#include "pva_kernels.cuh"

namespace pva {

namespace {

constexpr int kConsumerThreads = 256;
constexpr int kBlockThreads = kConsumerThreads + 32;
constexpr int kStages = 3;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kPx = 4;  // output columns per thread

struct TileGeom {
    // fixed per launch, computed on the host
    int R;          // output rows per tile
    int rps;        // rows per strip
    int tps;        // threads per strip (multiple of 32)
    int n_strips;
    int bands;      // ceil(crop / R)
    int rows_cap;   // source rows per plane a stage can hold
    int pitch;      // bytes per staged source row (multiple of 16)
    int stage_bytes;
    int strip_bytes;   // staging bytes of one strip buffer
    int bulk_store;    // 1: cp.async.bulk stores are legal (alignment), 0: plain stores
    int total_tiles;
    int off_ytab, off_in, off_out, off_lut;
};

struct TiledArgs {
    KParams p;
    TileGeom g;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_bar(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// Source row / column range of a tile, from the same tap formula the consumers use.
__device__ __forceinline__ void src_range(float scale, int first_dst, int last_dst, int in_size, int& a, int& n) {
    int i0, i1;
    float l0, l1;
    bilinear_taps(scale, first_dst, in_size, i0, i1, l0, l1);
    a = i0;
    bilinear_taps(scale, last_dst, in_size, i0, i1, l0, l1);
    n = i1 - a + 1;
}

template <typename OutT>
struct Pack;
template <>
struct Pack<__half> {
    static __device__ __forceinline__ uint32_t two(float a, float b) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};
template <>
struct Pack<__nv_bfloat16> {
    static __device__ __forceinline__ uint32_t two(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};

// Write the thread's 4 px x 3 ch results of one output row into the staging strip.
// THWC: 12 consecutive elements at element offset row*crop*3 + 12*g.
// CTHW: 4 consecutive elements per channel at c*rows*crop + row*crop + 4*g.
template <typename OutT, int kLayout>
__device__ __forceinline__ void stage_row(unsigned char* strip, int row, int rows, int crop, int g, bool full,
                                          const float (&o)[kPx][3]) {
    if (kLayout == kLayoutTHWC) {
        OutT* base = reinterpret_cast<OutT*>(strip) + (size_t)row * crop * 3 + 12 * g;
        if (full) {
            if (sizeof(OutT) == 4) {
                float4* d = reinterpret_cast<float4*>(base);
                d[0] = make_float4(o[0][0], o[0][1], o[0][2], o[1][0]);
                d[1] = make_float4(o[1][1], o[1][2], o[2][0], o[2][1]);
                d[2] = make_float4(o[2][2], o[3][0], o[3][1], o[3][2]);
            } else {
                using P = Pack<typename std::conditional<sizeof(OutT) == 2, OutT, __half>::type>;
                uint2* d = reinterpret_cast<uint2*>(base);
                d[0] = make_uint2(P::two(o[0][0], o[0][1]), P::two(o[0][2], o[1][0]));
                d[1] = make_uint2(P::two(o[1][1], o[1][2]), P::two(o[2][0], o[2][1]));
                d[2] = make_uint2(P::two(o[2][2], o[3][0]), P::two(o[3][1], o[3][2]));
            }
        } else {
            const int npx = crop - kPx * g;
#pragma unroll
            for (int i = 0; i < kPx; ++i)
                if (i < npx)
#pragma unroll
                    for (int c = 0; c < 3; ++c) base[i * 3 + c] = to_out<OutT>(o[i][c]);
        }
    } else {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            OutT* base = reinterpret_cast<OutT*>(strip) + ((size_t)c * rows + row) * crop + kPx * g;
            if (full && (crop & 3) == 0) {
                if (sizeof(OutT) == 4) {
                    *reinterpret_cast<float4*>(base) = make_float4(o[0][c], o[1][c], o[2][c], o[3][c]);
                } else {
                    using P = Pack<typename std::conditional<sizeof(OutT) == 2, OutT, __half>::type>;
                    *reinterpret_cast<uint2*>(base) = make_uint2(P::two(o[0][c], o[1][c]), P::two(o[2][c], o[3][c]));
                }
            } else {
                const int npx = crop - kPx * g;
#pragma unroll
                for (int i = 0; i < kPx; ++i)
                    if (i < npx) base[i] = to_out<OutT>(o[i][c]);
            }
        }
    }
}

template <typename OutT, int kLayout, int kNorm>
__global__ void __launch_bounds__(kBlockThreads, 1) pva_tiled_kernel(const __grid_constant__ TiledArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const KParams& p = a.p;
    const TileGeom& tg = a.g;
    __shared__ __align__(8) unsigned long long s_full[kStages];
    __shared__ __align__(8) unsigned long long s_empty[kStages];

    const int tid = threadIdx.x;
    const int crop = p.crop;
    // contiguous tile range of this CTA
    const int t_begin = (int)(((long long)tg.total_tiles * blockIdx.x) / gridDim.x);
    const int t_end = (int)(((long long)tg.total_tiles * (blockIdx.x + 1)) / gridDim.x);
    const int tiles_per_clip = p.n_fast * tg.bands;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(smem_u32(&s_full[s]), 1);
            mbar_init(smem_u32(&s_empty[s]), kConsumerThreads / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (kNorm == kNormExact) {
        float* lut = reinterpret_cast<float*>(smem + tg.off_lut);
        for (int i = tid; i < 3 * 256; i += blockDim.x) lut[i] = p.lut[i];
    }
    __syncthreads();

    const uint32_t in_base = smem_u32(smem + tg.off_in);

    if (tid >= kConsumerThreads) {
        // ------------------------------ producer warp ------------------------------
        const int lane = tid - kConsumerThreads;
        int cur_clip = -1, ca = 0, ncols = 0;
        for (int t = t_begin, it = 0; t < t_end; ++t, ++it) {
            const int clip_i = t / tiles_per_clip;
            const int rem = t - clip_i * tiles_per_clip;
            const int k = rem / tg.bands;
            const int band = rem - k * tg.bands;
            const ClipGeom& g = p.clips[clip_i];
            if (clip_i != cur_clip) {
                cur_clip = clip_i;
                src_range(g.sx, g.left, g.left + crop - 1, g.w, ca, ncols);
            }
            const int yb = band * tg.R;
            const int ye = min(yb + tg.R, crop);
            int ra, nrows;
            src_range(g.sy, g.top + yb, g.top + ye - 1, g.h, ra, nrows);
            const int s = it % kStages;
            const uint32_t ph = (uint32_t)(it / kStages) & 1u;
            mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
            const int t_src = p.frame_idx[g.idx_row][k];
            const size_t plane_sz = (size_t)g.h * g.w;
            const int n_copies = 3 * nrows;
            // bytes first (expect_tx must cover every copy of the phase), then the copies
            uint32_t my_bytes = 0;
            for (int j = lane; j < n_copies; j += 32) {
                const int c = j / nrows, r = j - c * nrows;
                const uintptr_t ga = (uintptr_t)(g.src + ((size_t)c * g.t_in + t_src) * plane_sz + (size_t)(ra + r) * g.w + ca);
                const uintptr_t al = ga & ~(uintptr_t)15;
                my_bytes += (uint32_t)(((ga + ncols - al) + 15) & ~(uintptr_t)15);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) my_bytes += __shfl_xor_sync(0xffffffffu, my_bytes, o);
            const uint32_t bar = smem_u32(&s_full[s]);
            if (lane == 0) mbar_arrive_expect_tx(bar, my_bytes);
            __syncwarp();
            const uint32_t st_base = in_base + (uint32_t)s * tg.stage_bytes;
            for (int j = lane; j < n_copies; j += 32) {
                const int c = j / nrows, r = j - c * nrows;
                const uintptr_t ga = (uintptr_t)(g.src + ((size_t)c * g.t_in + t_src) * plane_sz + (size_t)(ra + r) * g.w + ca);
                const uintptr_t al = ga & ~(uintptr_t)15;
                const uint32_t bytes = (uint32_t)(((ga + ncols - al) + 15) & ~(uintptr_t)15);
                bulk_load(st_base + (uint32_t)((c * tg.rows_cap + r) * tg.pitch), reinterpret_cast<const void*>(al), bytes, bar);
            }
        }
        return;
    }

    // -------------------------------- consumers --------------------------------
    const int strip = tid / tg.tps;
    const int g_col = tid - strip * tg.tps;  // column group of this thread
    const int gx = (crop + kPx - 1) / kPx;
    const bool active_col = g_col < gx && strip < tg.n_strips;
    const bool full_group = (kPx * g_col + kPx) <= crop && (crop & 3) == 0;  // vector staging stores need crop % 4 == 0
    const int strip_threads = tg.tps;
    int4* ytab = reinterpret_cast<int4*>(smem + tg.off_ytab);
    unsigned char* out_base = smem + tg.off_out + (size_t)strip * 2 * tg.strip_bytes;
    const float* lut = reinterpret_cast<const float*>(smem + tg.off_lut);
    const size_t esz = sizeof(OutT);

    // per-clip, per-thread column state
    int cur_clip = -1;
    int ca = 0;
    bool fastx = false;
    int woff = 0;             // fast path: byte offset (from column ca) of the thread's aligned 8-byte window
    uint32_t sel0 = 0, sel1 = 0;
    int off0[kPx], off1[kPx];  // byte path: offsets from column ca
    float lx0[kPx], lx1[kPx];
#pragma unroll
    for (int i = 0; i < kPx; ++i) off0[i] = off1[i] = 0, lx0[i] = lx1[i] = 0.f;

    int strip_it = 0;  // staging double-buffer index of this strip

    for (int t = t_begin, it = 0; t < t_end; ++t, ++it) {
        const int clip_i = t / tiles_per_clip;
        const int rem = t - clip_i * tiles_per_clip;
        const int k = rem / tg.bands;
        const int band = rem - k * tg.bands;
        const ClipGeom& g = p.clips[clip_i];

        if (clip_i != cur_clip) {
            // every strip must have left the previous clip before the row table is rebuilt
            named_bar(15, kConsumerThreads);
            cur_clip = clip_i;
            for (int y = tid; y < crop; y += kConsumerThreads) {
                int y0, y1;
                float l0, l1;
                bilinear_taps(g.sy, g.top + y, g.h, y0, y1, l0, l1);
                ytab[y] = make_int4(y0, y1, __float_as_int(l0), __float_as_int(l1));
            }
            int dummy;
            src_range(g.sx, g.left, g.left + crop - 1, g.w, ca, dummy);
            fastx = (g.flip >> 8) != 0;  // host: sx <= 1, w % 4 == 0, src % 4 == 0
            int xmin = 0x7fffffff;
            int x0v[kPx], x1v[kPx];
#pragma unroll
            for (int i = 0; i < kPx; ++i) {
                const int x = min(kPx * g_col + i, crop - 1);
                const int col = g.left + ((g.flip & 1) ? crop - 1 - x : x);
                bilinear_taps(g.sx, col, g.w, x0v[i], x1v[i], lx0[i], lx1[i]);
                xmin = min(xmin, x0v[i]);
            }
            const int wc = xmin & ~3;  // aligned window start column
            woff = wc - ca;
            sel0 = sel1 = 0;
#pragma unroll
            for (int i = 0; i < kPx; ++i) {
                off0[i] = x0v[i] - ca;
                off1[i] = x1v[i] - ca;
                sel0 |= (uint32_t)((x0v[i] - wc) & 7) << (4 * i);
                sel1 |= (uint32_t)((x1v[i] - wc) & 7) << (4 * i);
            }
            named_bar(15, kConsumerThreads);
        }

        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        const int yb = band * tg.R;
        const int ye = min(yb + tg.R, crop);
        const int y_begin = yb + strip * tg.rps;
        const int y_end = min(y_begin + tg.rps, ye);
        const int rows = max(y_end - y_begin, 0);
        unsigned char* stage_out = out_base + (size_t)(strip_it & 1) * tg.strip_bytes;

        mbar_wait(smem_u32(&s_full[s]), ph);

        if (active_col && rows > 0) {
            const int ra = ytab[yb].x;  // first source row staged for this tile
            const int t_src = p.frame_idx[g.idx_row][k];
            const uint32_t st_base = in_base + (uint32_t)s * tg.stage_bytes;
            // low 4 address bits of (plane c, row 0, column ca); a row adds r * w
            uint32_t k_c[3];
#pragma unroll
            for (int c = 0; c < 3; ++c)
                k_c[c] = (uint32_t)((uintptr_t)(g.src + ((size_t)c * g.t_in + t_src) * (size_t)g.h * g.w + ca));

            // x-pass of one source row into h[i][c]
            auto hrow = [&](int r, float (&h)[kPx][3]) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t shift = (k_c[c] + (uint32_t)r * (uint32_t)g.w) & 15u;
                    const uint32_t row_addr = st_base + (uint32_t)((c * tg.rows_cap + (r - ra)) * tg.pitch) + shift;
                    uint32_t b0[kPx], b1[kPx];
                    if (fastx) {
                        const uint32_t w0 = lds32(row_addr + woff);
                        const uint32_t w1 = lds32(row_addr + woff + 4);
                        const uint32_t t0 = __byte_perm(w0, w1, sel0);
                        const uint32_t t1 = __byte_perm(w0, w1, sel1);
#pragma unroll
                        for (int i = 0; i < kPx; ++i) {
                            b0[i] = (t0 >> (8 * i)) & 0xffu;
                            b1[i] = (t1 >> (8 * i)) & 0xffu;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < kPx; ++i) {
                            b0[i] = lds8(row_addr + off0[i]);
                            b1[i] = lds8(row_addr + off1[i]);
                        }
                    }
#pragma unroll
                    for (int i = 0; i < kPx; ++i) {
                        float v0, v1;
                        if (kNorm == kNormExact) {
                            v0 = lut[c * 256 + b0[i]];
                            v1 = lut[c * 256 + b1[i]];
                        } else {
                            v0 = (float)b0[i];
                            v1 = (float)b1[i];
                        }
                        h[i][c] = fmaf(lx0[i], v0, lx1[i] * v1);
                    }
                }
            };
            // y-pass + normalisation epilogue + staging of one output row
            auto emit = [&](int y, const int4& e, const float (&top)[kPx][3], const float (&bot)[kPx][3]) {
                const float ly0 = __int_as_float(e.z), ly1 = __int_as_float(e.w);
                float o[kPx][3];
#pragma unroll
                for (int i = 0; i < kPx; ++i)
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const float v = fmaf(ly0, top[i][c], ly1 * bot[i][c]);
                        o[i][c] = (kNorm == kNormExact) ? v : fmaf(v, p.na[c], p.nb[c]);
                    }
                stage_row<OutT, kLayout>(stage_out, y - y_begin, rows, crop, g_col, full_group, o);
            };

            float hA[kPx][3], hB[kPx][3];
            int y = y_begin;
            int4 e = ytab[y];
            int r = e.x;
            hrow(r, hA);
            // Source-row-driven walk; X holds source row r.  Returns true when the strip is done,
            // false after moving on to row r + 1 (held in Y): the caller swaps the roles.
            auto step = [&](float (&X)[kPx][3], float (&Y)[kPx][3]) -> bool {
                while (true) {
                    while (y < y_end && e.y == r) {  // both taps on row r (clamped at the bottom edge)
                        emit(y, e, X, X);
                        if (++y < y_end) e = ytab[y];
                    }
                    if (y >= y_end) return true;
                    if (e.x != r) {  // down-scaling skipped a source row
                        r = e.x;
                        hrow(r, X);
                        continue;
                    }
                    hrow(r + 1, Y);
                    while (y < y_end && e.x == r) {  // taps on rows r, r + 1
                        emit(y, e, X, Y);
                        if (++y < y_end) e = ytab[y];
                    }
                    r = r + 1;
                    return y >= y_end;
                }
            };
            while (true) {
                if (step(hA, hB)) break;
                if (step(hB, hA)) break;
            }
        }

        // this warp no longer reads the input stage
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(smem_u32(&s_empty[s]));

        if (strip < tg.n_strips) {
            // ---- strip store: staging -> global (fast pathway, and slow when frame k belongs to it)
            OutT* fast = reinterpret_cast<OutT*>(p.fast_out) + (size_t)clip_i * p.fast_clip_elems;
            OutT* slow = p.slow_out ? reinterpret_cast<OutT*>(p.slow_out) + (size_t)clip_i * p.slow_clip_elems : nullptr;
            if (tg.bulk_store) {
                fence_async_smem();  // generic-proxy writes of the strip -> visible to the async proxy
                if (g_col == 0) bulk_wait_read_all();  // the other staging buffer's previous store has drained
                named_bar(1 + strip, strip_threads);
                if (g_col == 0 && rows > 0) {
                    const uint32_t src = smem_u32(stage_out);
                    if (kLayout == kLayoutTHWC) {
                        const uint32_t bytes = (uint32_t)((size_t)rows * crop * 3 * esz);
                        bulk_store(fast + ((size_t)k * crop + y_begin) * crop * 3, src, bytes);
                        if (slow)
                            for (int j = 0; j < p.n_slow; ++j)
                                if (p.slow_idx[j] == k)
                                    bulk_store(slow + ((size_t)j * crop + y_begin) * crop * 3, src, bytes);
                    } else {
                        const uint32_t bytes = (uint32_t)((size_t)rows * crop * esz);
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            bulk_store(fast + (((size_t)c * p.n_fast + k) * crop + y_begin) * crop, src + c * bytes, bytes);
                            if (slow)
                                for (int j = 0; j < p.n_slow; ++j)
                                    if (p.slow_idx[j] == k)
                                        bulk_store(slow + (((size_t)c * p.n_slow + j) * crop + y_begin) * crop,
                                                   src + c * bytes, bytes);
                        }
                    }
                    bulk_commit();
                }
                ++strip_it;
            } else {
                // alignment does not allow bulk stores: cooperative element copy by the strip
                named_bar(1 + strip, strip_threads);
                const OutT* st = reinterpret_cast<const OutT*>(stage_out);
                const int per_c = rows * crop;
                const int n_el = per_c * 3;
                for (int idx = g_col; idx < n_el; idx += strip_threads) {
                    size_t off_f, off_s_base;
                    if (kLayout == kLayoutTHWC) {
                        off_f = ((size_t)k * crop + y_begin) * crop * 3 + idx;
                        off_s_base = (size_t)y_begin * crop * 3 + idx;
                    } else {
                        const int c = idx / per_c, w = idx - c * per_c;
                        off_f = (((size_t)c * p.n_fast + k) * crop + y_begin) * crop + w;
                        off_s_base = ((size_t)c * p.n_slow * crop + y_begin) * crop + w;
                    }
                    const OutT v = st[idx];
                    fast[off_f] = v;
                    if (slow)
                        for (int j = 0; j < p.n_slow; ++j)
                            if (p.slow_idx[j] == k) {
                                const size_t fr = (kLayout == kLayoutTHWC) ? (size_t)j * crop * crop * 3
                                                                           : (size_t)j * crop * crop;
                                slow[off_s_base + fr] = v;
                            }
                }
                named_bar(1 + strip, strip_threads);  // staging may be overwritten by the next tile
            }
        }
    }
    if (tg.bulk_store && g_col == 0 && strip < tg.n_strips) bulk_wait_all();
}

int round_up(int v, int m) { return (v + m - 1) / m * m; }

void host_taps(float scale, int dst, int in_size, int* i0, int* i1) {
    float s = fmaf(scale, (float)dst + 0.5f, -0.5f);
    if (s < 0.0f) s = 0.0f;
    int a = (int)s;
    if (a > in_size - 1) a = in_size - 1;
    int b = a + 1;
    if (b > in_size - 1) b = in_size - 1;
    *i0 = a;
    *i1 = b;
}

// Choose rows-per-strip and the shared-memory carve-up; false if nothing fits.
bool plan_tiles(const KParams& p, int out_dtype, int layout, TileGeom* tg, const char** why) {
    (void)layout;
    if (p.C != 3) {
        *why = "tiled kernel handles 3-channel clips";
        return false;
    }
    const int crop = p.crop;
    const int gx = (crop + kPx - 1) / kPx;
    const int tps = round_up(gx, 32);
    if (tps > kConsumerThreads) {
        *why = "crop wider than 1024";
        return false;
    }
    const int n_strips = kConsumerThreads / tps;
    const int esz = out_dtype == 0 ? 4 : 2;
    const int row_bytes = crop * 3 * esz;

    int max_cols = 0;
    for (int i = 0; i < p.n_clips; ++i) {
        const ClipGeom& g = p.clips[i];
        int a0, a1, b0, b1;
        host_taps(g.sx, g.left, g.w, &a0, &a1);
        host_taps(g.sx, g.left + crop - 1, g.w, &b0, &b1);
        max_cols = max_cols > (b1 - a0 + 1) ? max_cols : (b1 - a0 + 1);
    }
    const int pitch = round_up(max_cols + 15 + 8, 16);  // + row shift + the 8-byte tap window overhang
    const int ytab_bytes = round_up(crop * 16, 128);
    const int lut_bytes = 3 * 256 * 4;

    for (int rps = 8; rps >= 1; rps >>= 1) {
        const int R = rps * n_strips;
        const int bands = (crop + R - 1) / R;
        int rows_cap = 0;
        for (int i = 0; i < p.n_clips; ++i) {
            const ClipGeom& g = p.clips[i];
            for (int b = 0; b < bands; ++b) {
                const int yb = b * R, ye = (yb + R < crop ? yb + R : crop);
                int a0, a1, b0, b1;
                host_taps(g.sy, g.top + yb, g.h, &a0, &a1);
                host_taps(g.sy, g.top + ye - 1, g.h, &b0, &b1);
                const int n = b1 - a0 + 1;
                rows_cap = rows_cap > n ? rows_cap : n;
            }
        }
        const int stage_bytes = round_up(3 * rows_cap * pitch, 128);
        const int strip_bytes = round_up(rps * row_bytes, 128);
        int off = 0;
        tg->off_ytab = off;
        off += ytab_bytes;
        tg->off_lut = off;
        off += lut_bytes;
        off = round_up(off, 128);
        tg->off_in = off;
        off += kStages * stage_bytes;
        tg->off_out = off;
        off += n_strips * 2 * strip_bytes;
        if (off + 1024 > kMaxSmem) continue;
        tg->R = R;
        tg->rps = rps;
        tg->tps = tps;
        tg->n_strips = n_strips;
        tg->bands = bands;
        tg->rows_cap = rows_cap;
        tg->pitch = pitch;
        tg->stage_bytes = stage_bytes;
        tg->strip_bytes = strip_bytes;
        // bulk stores need 16-byte aligned global addresses and sizes for every strip
        const bool thwc = layout == kLayoutTHWC;
        // every strip offset and size is a multiple of `unit` (bytes of one output row, per channel for CTHW)
        const long long unit = thwc ? (long long)row_bytes : (long long)crop * esz;
        bool ok = (unit % 16) == 0;
        ok = ok && ((uintptr_t)p.fast_out % 16 == 0) && (!p.slow_out || (uintptr_t)p.slow_out % 16 == 0);
        tg->bulk_store = ok ? 1 : 0;
        tg->total_tiles = p.n_clips * p.n_fast * bands;
        return true;
    }
    *why = "tile does not fit shared memory (source rows too wide for this scale)";
    return false;
}

template <typename OutT>
cudaError_t launch_t(const TiledArgs& a, int layout, int norm, int grid, int smem, cudaStream_t stream) {
#define PVA_TILED_CASE(L, N)                                                                              \
    if (layout == L && norm == N) {                                                                       \
        auto kern = pva_tiled_kernel<OutT, L, N>;                                                         \
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);     \
        if (e != cudaSuccess) return e;                                                                   \
        kern<<<grid, kBlockThreads, smem, stream>>>(a);                                                   \
        return cudaGetLastError();                                                                        \
    }
    PVA_TILED_CASE(kLayoutCTHW, kNormExact)
    PVA_TILED_CASE(kLayoutCTHW, kNormAffine)
    PVA_TILED_CASE(kLayoutTHWC, kNormExact)
    PVA_TILED_CASE(kLayoutTHWC, kNormAffine)
#undef PVA_TILED_CASE
    return cudaErrorInvalidValue;
}

}  // namespace

bool tiled_supported(const KParams& p, int out_dtype, int layout, const char** why) {
    TileGeom tg;
    return plan_tiles(p, out_dtype, layout, &tg, why);
}

cudaError_t launch_tiled(KParams& p, int out_dtype, int layout, int norm, cudaStream_t stream, TiledConfig* cfg) {
    static thread_local TiledArgs a;  // 13 KB: keep it off the stack
    const char* why = "";
    if (!plan_tiles(p, out_dtype, layout, &a.g, &why)) return cudaErrorInvalidConfiguration;
    // per-clip fast x-path flag (bit 8 of flip): up-scaling or identity in x, 4-byte aligned rows
    for (int i = 0; i < p.n_clips; ++i) {
        ClipGeom& g = p.clips[i];
        const bool fastx = (g.w <= g.ow) && (g.w % 4 == 0) && ((uintptr_t)g.src % 4 == 0);
        g.flip = (uint16_t)((g.flip & 1) | (fastx ? 0x100 : 0));
    }
    a.p = p;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int grid = sms < a.g.total_tiles ? sms : a.g.total_tiles;
    const int smem = a.g.off_out + a.g.n_strips * 2 * a.g.strip_bytes;
    cfg->rows_per_tile = a.g.R;
    cfg->smem_bytes = smem;
    cfg->grid = grid;
    cfg->block = kBlockThreads;
    cfg->tiles = a.g.total_tiles;
    switch (out_dtype) {
        case 0: return launch_t<float>(a, layout, norm, grid, smem, stream);
        case 1: return launch_t<__half>(a, layout, norm, grid, smem, stream);
        case 2: return launch_t<__nv_bfloat16>(a, layout, norm, grid, smem, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace pva
