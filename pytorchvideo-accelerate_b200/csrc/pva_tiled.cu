// pva_tiled.cu -- the production kernel: persistent, warp-specialised, bulk-copy staged.
//
// Work item ("tile") = (clip, fast frame k, band of R output rows), all C=3 channels.
// One CTA per SM (on half of the SMs when launches stream, see launch_tiled), each with a contiguous share
// of the tiles.  Per tile:
//   producer warp  : one cp.async.bulk (TMA, 1-D) per channel plane over the whole band of source
//                    rows the tile touches (rows are contiguous, so a band is one linear range),
//                    into a stage of a 3-deep shared-memory ring; completion on an mbarrier.
//                    Temporal subsampling (run.py:82) is folded into the source address.
//                    Measured alternatives: per-row bulk copies (~90 cycles per 256-byte row on
//                    the SM's TMA unit: 57 us for the loads alone) and LDGSTS producers (slower).
//   consumer warps : 16 warps in two groups of 8 that alternate over the tiles.  A group's four
//                    strips take 8 output rows each; each thread owns 4 adjacent output columns x
//                    3 channels.  The x-pass is computed once per SOURCE row (two aligned 32-bit
//                    shared loads per plane, taps picked with PRMT, bytes turned into floats as
//                    fp16 subnormals: a zero-extended byte IS the half b * 2^-24) and kept in
//                    registers; consecutive output rows reuse it for the y-pass (bilinear,
//                    run.py:89/95).  Crop and flip (run.py:90-91) are folded into the per-thread
//                    column tables, Div255 + Normalize (run.py:83-84) into the epilogue (AFFINE)
//                    or a table (EXACT).
//   store          : outputs are packed to fp16/bf16/fp32 into a shared staging strip in final
//                    memory order and written with cp.async.bulk shared->global: once to the
//                    fast pathway and again to the slow pathway when the frame belongs to it
//                    (PackPathway, run.py:59-63) -- the source is read once.
//
// The source row pitch (340 B) is not a multiple of 16, so a 2-D tensor map is not legal.  The
// staged band keeps the source pitch and starts at the 16-byte granule of its first needed byte,
// which makes the staged address of (row, column) LINEAR in the row index: no per-row shift.
// The library only ever reads 16-byte granules that contain at least one byte of the clip.
// (Two column-cropped load paths were built and measured slower on the headline shape -- 128-bit LDGSTS
// rows, and four tensor maps per clip over the rows of one residue class mod 4; profiles/r1_ablation.md,
// profiles/r2_ablation.md.  They are no longer in the source.)
//
// Launches are chained by programmatic dependent launch; how much of a launch may overlap its predecessor
// is the caller's call (PVA_FLAG_CLIPS_READY, PVA_FLAG_OUTPUTS_DISJOINT; TileGeom::pdl).
//
// Layout: T x S x S x C (channels-last, the default) or the reference's C x T x S x S.  The two differ
// only in how a strip is staged and stored: one contiguous range of whole output rows, or three
// ranges, one per channel plane (template parameter kPlanar).
#include <stdlib.h>

#include <atomic>
#include <type_traits>

#include "pva_kernels.cuh"

namespace pva {

namespace {

// Experiment switches.  The production library reads NOTHING from the environment: every knob is a
// compile-time constant and the measurement-only branches of the kernel (stores / loads / arithmetic
// switched off, alternative load paths) are compiled out.  `make experiments` builds a second
// library, scripts/libpva_preproc_exp.so, with -DPVA_EXPERIMENTS: there the same knobs are the
// PVA_TILED_* environment variables the A/B scripts under scripts/ use.
#ifdef PVA_EXPERIMENTS
constexpr bool kExperiments = true;
int knob_env(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}
#define PVA_KNOB(name, dflt) knob_env("PVA_TILED_" name, dflt)
#else
constexpr bool kExperiments = false;
#define PVA_KNOB(name, dflt) (dflt)
#endif

constexpr int kConsumerThreads = 512;
// 16 consumer warps + one producer WARPGROUP (only its first warp works): setmaxnreg trades
// registers between the roles in warpgroup units, which lets the consumers have 112 registers
// although 20 warps x 96 is all the register file holds at launch.  setmaxnreg.inc draws only on
// what setmaxnreg.dec released inside the CTA: 4 warps x 32 x (96 - 24) = 9216 registers released,
// 16 warps x 32 x (112 - 96) = 8192 claimed (were it more, the consumers would wait forever).
constexpr int kBlockThreads = kConsumerThreads + 128;
constexpr int kProducerRegs = 24, kConsumerRegs = 112;
static_assert(4 * (96 - kProducerRegs) >= 16 * (kConsumerRegs - 96), "consumers would wait forever for registers");
constexpr int kMaxStages = 3;
constexpr int kMaxStrips = 8;  // named barriers 1..8, one per strip
constexpr int kMaxSmem = 227 * 1024;
constexpr int kPx = 4;  // output columns per thread

struct TileGeom {
    // fixed per launch, computed on the host
    int R;          // output rows per tile
    int rps;        // rows per strip
    int tps;        // threads per strip (multiple of 32)
    int n_strips;   // strips per consumer group
    int groups;     // consumer groups (1 or 2); group g takes this CTA's tiles t_begin + g, + groups, ...
    int bands;      // ceil(crop / R)
    int rows_cap;   // source rows per plane a stage can hold
    int plane_stride;  // bytes between the staged planes of a tile
    int stage_bytes;
    int strip_bytes;   // staging bytes of one strip buffer
    int bulk_store;    // 1: cp.async.bulk stores are legal (alignment), 0: plain stores
    int total_tiles;
    int off_ytab, off_in, off_out, off_lut;
    int pdl;         // programmatic dependent launch: 0 off, 1 wait before the first store, 2 wait at the start,
                     // 3 wait only before the CTA exits (PVA_FLAG_OUTPUTS_DISJOINT)
    int tile0_first; // the producer lets a CTA's first tile land before it requests the next ones
    int src_esz;     // bytes per source sample: 1 (uint8 clips) or 4 (float32 clips, kSrcF32)
    int l2hint;      // bit 0: loads evict_first, bit 1: stores evict_first
    int stages;      // input ring depth (2 or 3)
    int out_bufs;    // staging buffers per strip (2 = double-buffered, 1 = wait before rewriting)
    int na_uniform;  // AFFINE: all channels share one scale, so it is folded into the row table
    int debug;       // PVA_TILED_DEBUG bits (experiments only): 1 no stores, 2 no loads, 4 no compute,
                     // 16 no proxy fence (wrong results)
    // experiments only: per-CTA time stamps (8 x uint64 %globaltimer ns per CTA) of this launch, or null
    unsigned long long* trace;
};

constexpr int kMaxGrid = 160;

struct TiledArgs {
    KParams p;
    TileGeom g;
    // first tile of every CTA; tile_begin[grid] == total_tiles.  Equal tile counts by default: the
    // kernel is bound by its writes, which are the same for every tile.  (A compute-cost-weighted
    // split, PVA_TILED_BALANCE=1, measured 2 us slower.)
    int tile_begin[kMaxGrid + 1];
};

static_assert(sizeof(TiledArgs) <= 32764, "kernel parameter block limit");

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
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_load_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void bulk_store_hint(void* dst, uint32_t src, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst), "r"(src),
                 "r"(bytes), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_but_one() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_bar(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

constexpr int kTraceWords = 128;  // uint64 words per CTA and launch in the experiments' trace buffer
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Experiments build: stamp slot `slot` of this CTA's trace record (first writer wins when once == true).
#define PVA_TRACE(slot)                                                             \
    do {                                                                            \
        if (kExperiments && tg.trace) tg.trace[(size_t)blockIdx.x * kTraceWords + (slot)] = global_ns(); \
    } while (0)
// Per-tile stamps of each consumer group's first thread (strip 0's issuing thread): phase 0 tile loop top,
// 1 staging slot drained, 2 tile landed, 3 strip computed, 4 strip handed to the store engine; the
// group's tiles 4..15 (steady state), group g at word 8 + g * 60.
#define PVA_TRACE_TILE(n, phase)                                                                              \
    do {                                                                                                      \
        if (kExperiments && tg.trace && gtid == 0 && group < 2 && (n) >= 4 && (n) < 16)                       \
            tg.trace[(size_t)blockIdx.x * kTraceWords + 8 + group * 60 + ((n) - 4) * 5 + (phase)] = global_ns(); \
    } while (0)

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

// Write the thread's 4 px x 3 ch results of one output row into the staging strip, in final
// T x S x S x C order: 12 consecutive elements at element offset row * crop * 3 + 12 * g.
template <typename OutT>
__device__ __forceinline__ void stage_row(unsigned char* strip, int row, int crop, int g, bool full,
                                          const float (&o)[kPx][3]) {
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
}

// The same results for the reference's C x T x S x S order: the strip is staged as three channel
// planes of rps x crop elements, each of which is one contiguous range of the output.
template <typename OutT>
__device__ __forceinline__ void stage_row_planar(unsigned char* strip, int row, int rps, int crop, int g, bool full,
                                                 const float (&o)[kPx][3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        OutT* base = reinterpret_cast<OutT*>(strip) + ((size_t)c * rps + row) * crop + kPx * g;
        if (full) {
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

// Two bytes of t (compile-time positions) as a half2 of fp16 SUBNORMALS: the 16-bit pattern
// 0x00bb is exactly b * 2^-24, so zero-extension is the whole integer->float conversion and
// cvt.f32.f16 finishes it.  The 2^24 is folded into the row weights.
template <int lo, int hi>
__device__ __forceinline__ __half2 bytes_as_half2(uint32_t t) {
    const uint32_t v = __byte_perm(t, 0u, 0x4040u | lo | (hi << 8));
    return *reinterpret_cast<const __half2*>(&v);
}

// kNormT: 0 = EXACT, 1 = AFFINE with one scale for all channels (folded into the row table),
// 2 = AFFINE with per-channel scales.  kPlanar: write C x T x S x S (the reference's order) instead of
// T x S x S x C.
// kSrcF32: the clips hold float32 samples (what pytorchvideo's decoder hands over when a frame is not a
// whole number of bytes any more): the bands are staged as floats (4x the bytes, so smaller tiles) and
// the x-pass reads its taps as floats; EXACT evaluates Div255 and Normalize per tap with the reference's
// two IEEE divisions, AFFINE interpolates the raw values.
template <typename OutT, int kNormT, bool kFastX, bool kPlanar, bool kSrcF32>
__global__ void __launch_bounds__(kBlockThreads, 1) pva_tiled_kernel(const __grid_constant__ TiledArgs a) {
    static_assert(!(kSrcF32 && kFastX), "the aligned-window x-pass is for byte samples");
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int kNorm = kNormT == 0 ? kNormExact : kNormAffine;
    constexpr bool kUniformNa = kNormT == 1;
    const KParams& p = a.p;
    const TileGeom& tg = a.g;
    // measurement-only switches (see kExperiments): compile-time constants in the production library
    const int dbg = kExperiments ? tg.debug : 0;
    const int l2hint = kExperiments ? tg.l2hint : 3;
    // one "tile has landed" barrier per (stage, consumer group): with an odd ring depth a stage
    // alternates between the groups, and a group polling a shared barrier by parity could run two
    // phases ahead of it (bulk copies complete out of order) and take the other group's phase for its own
    __shared__ __align__(8) unsigned long long s_full[kMaxStages * 2];
    __shared__ __align__(8) unsigned long long s_empty[kMaxStages];

    const int tid = threadIdx.x;
    if (tid == 0) PVA_TRACE(0);  // kernel entry
    const int crop = p.crop;
    const int n_stages = tg.stages;
    // contiguous tile range of this CTA (tiles are ordered clip, frame, band)
    const int t_begin = a.tile_begin[blockIdx.x];
    const int t_end = a.tile_begin[blockIdx.x + 1];
    const int tiles_per_clip = p.n_fast * tg.bands;

    if (tid == 0) {
        for (int s = 0; s < kMaxStages * 2; ++s) {
            // the producer's arrive.expect_tx; the copies complete the tx count
            mbar_init(smem_u32(&s_full[s]), 1);
            if (s < kMaxStages) mbar_init(smem_u32(&s_empty[s]), kConsumerThreads / 32 / tg.groups);  // the warps of the group that consumed the tile
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (kNorm == kNormExact) {
        float* lut = reinterpret_cast<float*>(smem + tg.off_lut);
        for (int i = tid; i < 3 * 256; i += blockDim.x) lut[i] = p.lut[i];
    }
    __syncthreads();
    if (tg.pdl) {
        // Programmatic dependent launch.  This CTA may run while the previous kernel of the stream
        // is still finishing (it starts on an SM that kernel has already left).  Reading the clips
        // and computing is safe then -- the only kernel that signals its dependents early is this
        // one, and it never writes clip bytes -- but nothing may be WRITTEN before the previous
        // kernel has completed (the caller may be reusing the output buffers): the threads that
        // issue stores wait right before their first store, everybody else never has to.
        //
        // pdl == 3 (the caller vouches that the previous kernel neither reads nor writes this launch's
        // outputs): nothing is waited for before the stores either, so the launch overlaps the previous
        // one's tail completely.  One thread per CTA still executes the wait before the CTA exits, which
        // keeps completion transitive: this grid cannot complete before its predecessor has.
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        if (!tg.bulk_store || tg.pdl == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
    }

    const uint32_t in_base = smem_u32(smem + tg.off_in);
    if (tid == 0) PVA_TRACE(1);  // set-up done

    if (tid >= kConsumerThreads) {
        // ------------------------------ producer warp ------------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kProducerRegs));
        if (tid >= kConsumerThreads + 32) return;  // the other three warps only existed to give registers away
        const int lane = tid - kConsumerThreads;
        const uint64_t pol = l2_evict_first_policy();
        int cur_clip = -1, ca = 0, ncols = 0, W = 0, H = 0, T_in = 0, top = 0, idx_row = 0;
        float sy = 0.f;
        const uint8_t* src = nullptr;
        int s = 0;
        uint32_t ph = 0;
        for (int t = t_begin; t < t_end; ++t) {
            const int grp = (t - t_begin) % tg.groups;  // the consumer group this tile is for
            const int clip_i = t / tiles_per_clip;
            const int rem = t - clip_i * tiles_per_clip;
            const int k = rem / tg.bands;
            const int band = rem - k * tg.bands;
            if (clip_i != cur_clip) {
                cur_clip = clip_i;
                const ClipGeom& g = p.clips[clip_i];
                src = g.src, W = g.w, H = g.h, T_in = g.t_in, sy = g.sy, top = g.top, idx_row = g.idx_row;
                src_range(g.sx, g.left, g.left + crop - 1, g.w, ca, ncols);
            }
            const int yb = band * tg.R;
            const int ye = min(yb + tg.R, crop);
            int ra, nrows;
            src_range(sy, top + yb, top + ye - 1, H, ra, nrows);
            mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
            const int t_src = p.frame_idx[idx_row][k];
            const uint32_t st_base = in_base + (uint32_t)s * tg.stage_bytes;
            const uint32_t bar = smem_u32(&s_full[2 * s + grp]);
            {
                // One bulk copy per plane: the whole band of source rows, first needed byte (ra, ca)
                // to last needed byte, rounded out to 16-byte granules.  Rows are contiguous in
                // memory, so this is a single linear range; staged byte (r, col) sits at
                //   plane base + (g0 & 15) + ((r - ra) * W + (col - ca)) * bytes-per-sample.
                const uint32_t esz_src = (uint32_t)tg.src_esz;  // 1 or 4 bytes per sample
                const uint8_t* g0 = src + ((((size_t)min(lane, 2) * T_in + t_src) * H + ra) * W + ca) * esz_src;
                const uint32_t sh = (uint32_t)((uintptr_t)g0 & 15);
                const uint32_t bytes = (sh + ((uint32_t)(nrows - 1) * (uint32_t)W + (uint32_t)ncols) * esz_src + 15u) & ~15u;
                uint32_t total = lane < 3 ? bytes : 0u;
                total += __shfl_xor_sync(0xffffffffu, total, 1);
                total += __shfl_xor_sync(0xffffffffu, total, 2);
                if (dbg & 2) total = 0;
                if (lane == 0) mbar_arrive_expect_tx(bar, total);
                __syncwarp();
                if (lane < 3 && !(dbg & 2)) {
                    if (l2hint & 1) bulk_load_hint(st_base + (uint32_t)lane * tg.plane_stride, g0 - sh, bytes, bar, pol);
                    else bulk_load(st_base + (uint32_t)lane * tg.plane_stride, g0 - sh, bytes, bar);
                }
            }
            // At the start of a launch every SM asks for its whole ring at once (148 x 3 tiles = 13 MB)
            // and a CTA's first tile used to land only after ~3 us, behind everybody's second and third.
            // Let tile 0 land before asking for more: it arrives in about half the time and the first
            // consumer group starts that much earlier; steady state is unaffected.
            if (tg.tile0_first && t == t_begin) mbar_wait(bar, 0u);
            if (++s == n_stages) s = 0, ph ^= 1u;
        }
        if (lane == 0) PVA_TRACE(7);  // all loads issued
        return;
    }

    // -------------------------------- consumers --------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kConsumerRegs));
    // Two consumer groups alternate over the CTA's tiles: each gives its strips twice the rows of
    // a tile (more reuse of the x-pass, half the per-tile set-up per warp), and while one group
    // waits -- for its tile to land or for its single staging slot to drain -- the other computes.
    const int group_threads = kConsumerThreads / tg.groups;
    const int group = tid / group_threads;
    const int gtid = tid - group * group_threads;
    const int strip = gtid / tg.tps;          // strip index inside the group
    const int g_col = gtid - strip * tg.tps;  // column group of this thread
    const int gx = (crop + kPx - 1) / kPx;
    const bool in_strip = strip < tg.n_strips;
    const int bar_strip = 1 + group * tg.n_strips + strip;  // named barrier of this strip (1..8)
    const bool active_col = g_col < gx && in_strip;
    const bool full_group = (kPx * g_col + kPx) <= crop && (crop & 3) == 0;  // vector staging stores need crop % 4 == 0
    const int strip_threads = tg.tps;
    // Threads that hand a strip to the store engine: thread 0 of the strip; in the C x T x S x S order the
    // strip's first three threads, one channel plane each, so that the three copies are issued side by
    // side instead of one after the other (36.6 -> 35.5 us per headline launch; six threads, with the slow
    // pathway's copies on their own, measured the same).
    const bool issuer = kPlanar ? g_col < 3 : (g_col == 0 || (g_col == 32 && p.slow_out != nullptr && tg.tps > 32));
    int4* ytab = reinterpret_cast<int4*>(smem + tg.off_ytab) + (size_t)group * crop;  // one row table per group
    unsigned char* out_base = smem + tg.off_out + (size_t)(group * tg.n_strips + strip) * tg.out_bufs * tg.strip_bytes;
    const float* lut = reinterpret_cast<const float*>(smem + tg.off_lut);
    const size_t esz = sizeof(OutT);

    // per-clip state, cached in registers
    int cur_clip = -1;
    int ca = 0, W = 0, H = 0, T_in = 0, idx_row = 0;
    const uint8_t* src = nullptr;
    int woff = 0;               // kFastX: byte offset (from column ca) of the thread's aligned 8-byte window
    uint32_t sel0 = 0, sel1 = 0;  // kFastX: PRMT selectors picking the 4 left / right taps out of the window
    int off0[kPx], off1[kPx];   // !kFastX: byte offsets of the taps from column ca
    float lx0[kPx], lx1[kPx];
#pragma unroll
    for (int i = 0; i < kPx; ++i) off0[i] = off1[i] = 0, lx0[i] = lx1[i] = 0.f;

    int strip_it = 0;  // staging buffer index of this strip
    bool may_store = !(tg.pdl == 1 && tg.bulk_store);  // false until griddepcontrol.wait has been executed

    // Where this group's tiles are: (clip, frame, band), input stage and barrier phase are carried from
    // tile to tile instead of being derived from the tile number -- four integer divisions by run-time
    // values sat on the critical path between one tile's store and the next tile's first instruction.
    // A (stage, group) barrier completes once every `period` tiles of the CTA.
    const int period = (n_stages % tg.groups == 0) ? n_stages : n_stages * tg.groups;
    int clip_i, k, band, s, pc;
    uint32_t ph;
    {
        const int t0 = t_begin + group, it0 = group;
        clip_i = t0 / tiles_per_clip;
        const int rem = t0 - clip_i * tiles_per_clip;
        k = rem / tg.bands;
        band = rem - k * tg.bands;
        s = it0 % n_stages;
        pc = it0 % period;
        ph = (uint32_t)(it0 / period) & 1u;
    }
    for (int t = t_begin + group; t < t_end; t += tg.groups,
             band += tg.groups, s += tg.groups, pc += tg.groups) {
        // carry (the step, tg.groups, is at most 2; bands, stages and period are at least 1)
        while (band >= tg.bands) {
            band -= tg.bands;
            if (++k == p.n_fast) k = 0, ++clip_i;
        }
        while (s >= n_stages) s -= n_stages;
        while (pc >= period) pc -= period, ph ^= 1u;

        if (clip_i != cur_clip) {
            // A strip rebuilds only ITS OWN rows of the row table (the rows it computes in every band),
            // so a clip change needs no barrier beyond the strip's own: its threads all passed the
            // strip barrier of the previous tile's store, i.e. none of them still reads the old rows.
            // (A group-wide barrier here cost the CTAs that cross a clip boundary ~3 us: it lines up
            // strips that otherwise drift apart by up to a tile, which is what keeps the stores flowing.)
            cur_clip = clip_i;
            const ClipGeom& g = p.clips[clip_i];
            src = g.src, W = g.w, H = g.h, T_in = g.t_in, idx_row = g.idx_row;
            // AFFINE: the row weights absorb the 2^24 of the subnormal trick and, when all
            // channels share it, the normalisation scale 1 / (255 std)
            const float wscale = (kNorm == kNormAffine) ? (kSrcF32 ? 1.0f : 16777216.0f) * (kUniformNa ? p.na[0] : 1.0f) : 1.0f;
            for (int i = g_col; i < tg.bands * tg.rps && in_strip; i += strip_threads) {
                const int y = (i / tg.rps) * tg.R + strip * tg.rps + (i % tg.rps);
                if (y >= crop) continue;
                int y0, y1;
                float l0, l1;
                bilinear_taps(g.sy, g.top + y, g.h, y0, y1, l0, l1);
                ytab[y] = make_int4(y0, y1, __float_as_int(l0 * wscale), __float_as_int(l1 * wscale));
            }
            int dummy;
            src_range(g.sx, g.left, g.left + crop - 1, g.w, ca, dummy);
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
            if (in_strip) named_bar(bar_strip, strip_threads);  // the strip's rows of the table are complete
        }

        const int yb = band * tg.R;
        const int ye = min(yb + tg.R, crop);
        const int y_begin = yb + strip * tg.rps;
        const int y_end = min(y_begin + tg.rps, ye);
        const int rows = max(y_end - y_begin, 0);
        unsigned char* stage_out = out_base + (size_t)strip_it * tg.strip_bytes;

        // Staged address of (plane c, source row r, window start) = rowbase[c] + r * W.  Worked out BEFORE
        // the waits below so that its table look-ups and multiplications are not on the path between
        // "tile has landed" and the first shared-memory load.  Only the low four bits of the global
        // address matter (the band starts at its 16-byte granule), so 32-bit arithmetic will do.
        uint32_t rowbase[3];
        {
            int ra;  // first source row staged for this tile (row yb of the table belongs to strip 0)
            {
                const ClipGeom& g = p.clips[clip_i];
                int rb;
                float l0, l1;
                bilinear_taps(g.sy, g.top + yb, g.h, ra, rb, l0, l1);
            }
            const uint32_t t_src = p.frame_idx[idx_row][k];
            constexpr uint32_t kEsz = kSrcF32 ? 4u : 1u;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const uint32_t lin = ((((uint32_t)c * (uint32_t)T_in + t_src) * (uint32_t)H + (uint32_t)ra) * (uint32_t)W + (uint32_t)ca) * kEsz;
                const uint32_t gc = (uint32_t)(uintptr_t)src + lin;
                rowbase[c] = in_base + (uint32_t)s * tg.stage_bytes + (uint32_t)c * tg.plane_stride + (gc & 15u) +
                             (uint32_t)(kFastX ? woff : 0) - (uint32_t)ra * (uint32_t)W * kEsz;
            }
        }

        const int tile_no = (t - t_begin) / tg.groups;  // this group's n-th tile
        PVA_TRACE_TILE(tile_no, 0);
        if (tg.out_bufs == 1 && tg.bulk_store && in_strip) {
            // single staging buffer: the previous tile's store must have drained it
            if (issuer) bulk_wait_read_all();
            named_bar(bar_strip, strip_threads);
        }

        PVA_TRACE_TILE(tile_no, 1);
        // every thread polls: measured faster than one poller per warp + __syncwarp (43.4 vs 46.0 us)
        mbar_wait(smem_u32(&s_full[2 * s + group]), ph);
        if (tid == 0 && t == t_begin) PVA_TRACE(2);  // first tile landed
        PVA_TRACE_TILE(tile_no, 2);

        if (active_col && rows > 0 && !(dbg & 4)) {
            // x-pass of one source row into h[i][c].
            //   EXACT : h = fma(lx0, lut[b0], lx1 * lut[b1])            (ATen's arithmetic)
            //   AFFINE: h = fma(lx1, b1 - b0, b0) * 2^-24               (same value, one rounding)
            auto hrow = [&](int r, float (&h)[kPx][3]) {
                const uint32_t ro = (uint32_t)r * (uint32_t)W * (kSrcF32 ? 4u : 1u);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t addr = rowbase[c] + ro;
                    if constexpr (kSrcF32) {
                        const float m = p.mean[c], sd = p.std[c];
#pragma unroll
                        for (int i = 0; i < kPx; ++i) {
                            const float f0 = lds_f32(addr + 4u * (uint32_t)off0[i]);
                            const float f1 = lds_f32(addr + 4u * (uint32_t)off1[i]);
                            if (kNorm == kNormExact) {
                                // Div255 then Normalize per tap, true divisions, as the reference ops do (run.py:83-84)
                                const float v0 = __fdiv_rn(__fdiv_rn(f0, 255.0f) - m, sd);
                                const float v1 = __fdiv_rn(__fdiv_rn(f1, 255.0f) - m, sd);
                                h[i][c] = fmaf(lx0[i], v0, lx1[i] * v1);
                            } else {
                                h[i][c] = fmaf(lx1[i], f1 - f0, f0);
                            }
                        }
                    } else if (kFastX) {
                        const uint32_t w0 = lds32(addr);
                        const uint32_t w1 = lds32(addr + 4);
                        const uint32_t t0 = __byte_perm(w0, w1, sel0);
                        const uint32_t t1 = __byte_perm(w0, w1, sel1);
                        if (kNorm == kNormExact) {
#pragma unroll
                            for (int i = 0; i < kPx; ++i) {
                                const float v0 = lut[c * 256 + ((t0 >> (8 * i)) & 0xffu)];
                                const float v1 = lut[c * 256 + ((t1 >> (8 * i)) & 0xffu)];
                                h[i][c] = fmaf(lx0[i], v0, lx1[i] * v1);
                            }
                        } else {
                            const __half2 a01 = bytes_as_half2<0, 1>(t0), a23 = bytes_as_half2<2, 3>(t0);
                            const __half2 b01 = bytes_as_half2<0, 1>(t1), b23 = bytes_as_half2<2, 3>(t1);
                            const float2 v01 = __half22float2(a01), v23 = __half22float2(a23);
                            const float2 d01 = __half22float2(__hsub2(b01, a01));  // exact: multiples of 2^-24
                            const float2 d23 = __half22float2(__hsub2(b23, a23));
                            h[0][c] = fmaf(lx1[0], d01.x, v01.x);
                            h[1][c] = fmaf(lx1[1], d01.y, v01.y);
                            h[2][c] = fmaf(lx1[2], d23.x, v23.x);
                            h[3][c] = fmaf(lx1[3], d23.y, v23.y);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < kPx; ++i) {
                            const uint32_t b0 = lds8(addr + off0[i]);
                            const uint32_t b1 = lds8(addr + off1[i]);
                            if (kNorm == kNormExact) {
                                h[i][c] = fmaf(lx0[i], lut[c * 256 + b0], lx1[i] * lut[c * 256 + b1]);
                            } else {
                                const float v0 = (float)b0 * 5.9604644775390625e-08f;  // 2^-24, as on the fast path
                                const float dd = (float)((int)b1 - (int)b0) * 5.9604644775390625e-08f;
                                h[i][c] = fmaf(lx1[i], dd, v0);
                            }
                        }
                    }
                }
            };
            // y-pass + normalisation epilogue + staging of one output row
            auto emit = [&](int y, const int4& e, const float (&tp)[kPx][3], const float (&bt)[kPx][3]) {
                const float ly0 = __int_as_float(e.z), ly1 = __int_as_float(e.w);
                float o[kPx][3];
                if (kNorm == kNormExact) {
#pragma unroll
                    for (int i = 0; i < kPx; ++i)
#pragma unroll
                        for (int c = 0; c < 3; ++c) o[i][c] = fmaf(ly0, tp[i][c], ly1 * bt[i][c]);
                } else {
                    // out = (ly0 top + ly1 bot) * na + nb, na folded into the row weights
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        float w0 = ly0, w1 = ly1;
                        if (!kUniformNa) w0 *= p.na[c], w1 *= p.na[c];
                        const float b = p.nb[c];
#pragma unroll
                        for (int i = 0; i < kPx; ++i) o[i][c] = fmaf(w0, tp[i][c], fmaf(w1, bt[i][c], b));
                    }
                }
                if (kPlanar) stage_row_planar<OutT>(stage_out, y - y_begin, tg.rps, crop, g_col, full_group, o);
                else stage_row<OutT>(stage_out, y - y_begin, crop, g_col, full_group, o);
            };
            // One output row.  T holds source row rT (the row's upper tap), B holds rB.  When the
            // upper tap moves on to the row B already holds, the buffers swap roles (returns true)
            // instead of being copied.
            auto row_step = [&](int y, const int4& e, float (&T)[kPx][3], float (&B)[kPx][3], int& rT, int& rB) -> bool {
                if (e.x == rB && e.x != rT) {
                    if (e.y != rT) {
                        hrow(e.y, T);
                        rT = e.y;
                    }
                    emit(y, e, B, T);
                    return true;
                }
                if (e.x != rT) {
                    hrow(e.x, T);
                    rT = e.x;
                }
                if (e.y != rB) {
                    hrow(e.y, B);
                    rB = e.y;
                }
                emit(y, e, T, B);
                return false;
            };

            float hA[kPx][3], hB[kPx][3];
            int rA = -1, rB = -1;
            bool top_is_a = true;
#pragma unroll 1
            for (int y = y_begin; y < y_end; ++y) {
                const int4 e = ytab[y];
                if (top_is_a) {
                    if (row_step(y, e, hA, hB, rA, rB)) top_is_a = false;
                } else {
                    if (row_step(y, e, hB, hA, rB, rA)) top_is_a = true;
                }
            }
        }

        PVA_TRACE_TILE(tile_no, 3);
        // this warp no longer reads the input stage
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(smem_u32(&s_empty[s]));

        if (in_strip) {
            // ---- strip store: staging -> global (fast pathway, and slow when frame k belongs to it)
            OutT* fast = reinterpret_cast<OutT*>(p.fast_out) + (size_t)clip_i * p.fast_clip_elems;
            OutT* slow = p.slow_out ? reinterpret_cast<OutT*>(p.slow_out) + (size_t)clip_i * p.slow_clip_elems : nullptr;
            if (tg.bulk_store) {
                if (!(dbg & 16)) fence_async_smem();  // generic-proxy writes of the strip -> visible to the async proxy
                // double-buffered: the other buffer's store (issued a tile ago) must have drained
                // before anyone runs ahead into the next tile and overwrites it
                if (issuer) {
                    if (tg.out_bufs == 2) bulk_wait_read_all();
                    if (tg.out_bufs == 3) bulk_wait_read_but_one();  // the store before last has drained
                }
                named_bar(bar_strip, strip_threads);
                if (issuer && rows > 0 && !(dbg & 1)) {
                    if (tid == 0 && t == t_begin) PVA_TRACE(3);  // first strip computed, about to store
                    if (!may_store) {
                        asm volatile("griddepcontrol.wait;" ::: "memory");
                        may_store = true;
                    }
                    if (tid == 0 && t == t_begin) PVA_TRACE(4);  // previous kernel complete (PDL wait returned)
                    const uint32_t ssrc = smem_u32(stage_out);
                    if (kPlanar) {
                        // one contiguous range per channel plane: rows x crop elements
                        const uint32_t bytes = (uint32_t)((size_t)rows * crop * esz);
                        const uint32_t sstep = (uint32_t)((size_t)tg.rps * crop * esz);
                        const uint64_t pol = l2_evict_first_policy();
                        const int c = g_col;  // this thread's plane
                        {
                            OutT* d = fast + (((size_t)c * p.n_fast + k) * crop + y_begin) * crop;
                            if (l2hint & 2) bulk_store_hint(d, ssrc + c * sstep, bytes, pol);
                            else bulk_store(d, ssrc + c * sstep, bytes);
                        }
                        if (slow)
                            for (int j = p.slow_of[k]; j != kNoSlot; j = p.slow_next[j]) {
                                OutT* d = slow + (((size_t)c * p.n_slow + j) * crop + y_begin) * crop;
                                if (l2hint & 2) bulk_store_hint(d, ssrc + c * sstep, bytes, pol);
                                else bulk_store(d, ssrc + c * sstep, bytes);
                            }
                    } else {
                    const uint32_t bytes = (uint32_t)((size_t)rows * crop * 3 * esz);
                    // thread 0 of the strip sends the fast pathway's copy, the first thread of the strip's second
                    // warp the slow pathway's (a strip of one warp: thread 0 sends both)
                    const bool do_fast = g_col == 0, do_slow = slow && (g_col == 32 || tg.tps <= 32);
                    if (l2hint & 2) {
                        const uint64_t pol = l2_evict_first_policy();
                        if (do_fast) bulk_store_hint(fast + ((size_t)k * crop + y_begin) * crop * 3, ssrc, bytes, pol);
                        if (do_slow)
                            for (int j = p.slow_of[k]; j != kNoSlot; j = p.slow_next[j])
                                bulk_store_hint(slow + ((size_t)j * crop + y_begin) * crop * 3, ssrc, bytes, pol);
                    } else {
                        if (do_fast) bulk_store(fast + ((size_t)k * crop + y_begin) * crop * 3, ssrc, bytes);
                        if (do_slow)
                            for (int j = p.slow_of[k]; j != kNoSlot; j = p.slow_next[j])
                                bulk_store(slow + ((size_t)j * crop + y_begin) * crop * 3, ssrc, bytes);
                    }
                    }
                    bulk_commit();
                }
                PVA_TRACE_TILE(tile_no, 4);
                if (++strip_it >= tg.out_bufs) strip_it = 0;
            } else {
                // alignment does not allow bulk stores: cooperative element copy by the strip
                named_bar(bar_strip, strip_threads);
                const OutT* st = reinterpret_cast<const OutT*>(stage_out);
                const int n_el = rows * crop * 3;
                if (kPlanar) {
                    const int plane_el = rows * crop;
                    for (int idx = g_col; idx < n_el; idx += strip_threads) {
                        const int c = idx / plane_el, rem = idx - c * plane_el;
                        const OutT v = st[(size_t)c * tg.rps * crop + rem];
                        fast[(((size_t)c * p.n_fast + k) * crop + y_begin) * crop + rem] = v;
                        if (slow)
                            for (int j = p.slow_of[k]; j != kNoSlot; j = p.slow_next[j])
                                slow[(((size_t)c * p.n_slow + j) * crop + y_begin) * crop + rem] = v;
                    }
                } else {
                for (int idx = g_col; idx < n_el; idx += strip_threads) {
                    const OutT v = st[idx];
                    fast[((size_t)k * crop + y_begin) * crop * 3 + idx] = v;
                    if (slow)
                        for (int j = p.slow_of[k]; j != kNoSlot; j = p.slow_next[j])
                            slow[((size_t)j * crop + y_begin) * crop * 3 + idx] = v;
                }
                }
                named_bar(bar_strip, strip_threads);  // staging may be overwritten by the next tile
            }
        }
    }
    // the staging buffers must outlive the stores' reads; the writes themselves complete with the kernel
    if (tid == 0) PVA_TRACE(5);  // last tile of group 0 handed to the store engine
    if (tg.pdl == 3 && tid == 0) asm volatile("griddepcontrol.wait;" ::: "memory");  // completion stays transitive
    if (tg.bulk_store && issuer && in_strip) bulk_wait_read_all();
    if (tid == 0) PVA_TRACE(6);  // staging drained: CTA about to exit
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

// Choose rows-per-strip, ring depth and the shared-memory carve-up; false if nothing fits.
bool plan_tiles(const KParams& p, int out_dtype, int layout, TileGeom* tg, const char** why) {
    if (layout != kLayoutTHWC && layout != kLayoutCTHW) {
        *why = "unknown output layout";
        return false;
    }
    if (p.C != 3) {
        *why = "tiled kernel handles 3-channel clips";
        return false;
    }
    const int esz_src = p.src_f32 ? 4 : 1;
    for (int i = 0; i < p.n_clips; ++i)
        if ((uintptr_t)p.clips[i].src % esz_src) {
            *why = "float32 clip not 4-byte aligned";
            return false;
        }
    const int crop = p.crop;
    const int gx = (crop + kPx - 1) / kPx;
    const int tps = round_up(gx, 32);
    if (tps > kConsumerThreads) {
        *why = "crop wider than 2048";
        return false;
    }
    int strips_total = kConsumerThreads / tps;  // over all consumer groups
    if (strips_total > kMaxStrips) strips_total = kMaxStrips;
    const int esz = out_dtype == 0 ? 4 : 2;
    const int row_bytes = crop * 3 * esz;

    const int lut_bytes = 3 * 256 * 4;
    // a bulk store covers whole output rows (T x S x S x C) or whole rows of one channel plane (C x T x S x S)
    const int store_unit = layout == kLayoutCTHW ? crop * esz : row_bytes;
    const bool bulk_ok = (store_unit % 16) == 0 && ((uintptr_t)p.fast_out % 16 == 0) &&
                         (!p.slow_out || (uintptr_t)p.slow_out % 16 == 0);

    // candidate configurations in order of preference:
    // (consumer groups, rows per strip, input stages, staging buffers per strip)
    // {2, 4, 2, 1} is for float32 clips only (their bands are 4x the bytes: 16-row tiles, one stage per group)
    static const int kCand[][4] = {{2, 8, 3, 1}, {2, 4, 2, 1}, {1, 4, 2, 3}, {1, 4, 3, 2}, {1, 4, 2, 2}, {1, 4, 2, 1}, {1, 2, 3, 2},
                                   {1, 2, 2, 2}, {1, 2, 2, 1}, {1, 1, 2, 2}, {1, 1, 2, 1}};
    const int force_rps = PVA_KNOB("RPS", 0), force_bufs = PVA_KNOB("OUTBUFS", 0),
              force_stages = PVA_KNOB("STAGES", 0), force_groups = PVA_KNOB("GROUPS", 0);
    for (const auto& cand : kCand) {
        const int groups = cand[0], rps = cand[1], stages = cand[2], out_bufs = bulk_ok ? cand[3] : 1;
        if (groups == 2 && rps == 4 && !p.src_f32) continue;
        if (force_groups && groups != force_groups) continue;
        if (force_rps && rps != force_rps) continue;
        if (force_bufs && cand[3] != force_bufs) continue;
        if (force_stages && stages != force_stages) continue;
        if (groups == 2 && (strips_total < 2 || strips_total % 2 || !bulk_ok)) continue;
        const int n_strips = strips_total / groups;  // per group
        const int ytab_bytes = round_up(groups * crop * 16, 128);
        const int R = rps * n_strips;
        const int bands = (crop + R - 1) / R;
        int rows_cap = 0, band_bytes = 0;
        for (int i = 0; i < p.n_clips; ++i) {
            const ClipGeom& g = p.clips[i];
            int c0, c1, d0, d1;
            host_taps(g.sx, g.left, g.w, &c0, &c1);
            host_taps(g.sx, g.left + crop - 1, g.w, &d0, &d1);
            const int ncols = d1 - c0 + 1;
            for (int b = 0; b < bands; ++b) {
                const int yb = b * R, ye = (yb + R < crop ? yb + R : crop);
                int a0, a1, b0, b1;
                host_taps(g.sy, g.top + yb, g.h, &a0, &a1);
                host_taps(g.sy, g.top + ye - 1, g.h, &b0, &b1);
                const int n = b1 - a0 + 1;
                rows_cap = rows_cap > n ? rows_cap : n;
                const int bytes = ((n - 1) * g.w + ncols) * esz_src;
                band_bytes = band_bytes > bytes ? band_bytes : bytes;
            }
        }
        // + 15 leading alignment bytes, rounded to a granule, + the 8-byte tap window overhang
        const int plane_stride = round_up(band_bytes + 15 + 16 + 8, 16);
        const int stage_bytes = round_up(3 * plane_stride, 128);
        const int strip_bytes = round_up(rps * row_bytes, 128);
        int off = 0;
        tg->off_ytab = off;
        off += ytab_bytes;
        tg->off_lut = off;
        off += lut_bytes;
        off = round_up(off, 128);
        tg->off_in = off;
        off += stages * stage_bytes;
        tg->off_out = off;
        off += groups * n_strips * out_bufs * strip_bytes;
        if (off + 1024 > kMaxSmem) continue;
        tg->groups = groups;
        tg->R = R;
        tg->rps = rps;
        tg->tps = tps;
        tg->n_strips = n_strips;
        tg->bands = bands;
        tg->rows_cap = rows_cap;
        tg->plane_stride = plane_stride;
        tg->stage_bytes = stage_bytes;
        tg->strip_bytes = strip_bytes;
        tg->stages = stages;
        tg->out_bufs = out_bufs;
        // bulk stores need 16-byte aligned global addresses and sizes: every strip offset and
        // size is a multiple of the bytes of one output row
        tg->bulk_store = bulk_ok ? 1 : 0;
        tg->total_tiles = p.n_clips * p.n_fast * bands;
        tg->na_uniform = (p.na[0] == p.na[1] && p.na[1] == p.na[2]) ? 1 : 0;
        tg->src_esz = esz_src;
        tg->debug = PVA_KNOB("DEBUG", 0);
        tg->trace = nullptr;
        return true;
    }
    *why = "tile does not fit shared memory (source rows too wide for this scale)";
    return false;
}

// Per-device facts, looked up once per process (relaxed atomics: racing threads store the same value).
constexpr int kMaxDevices = 64;
int sm_count(int dev) {
    static std::atomic<int> cached[kMaxDevices];
    if (dev >= 0 && dev < kMaxDevices) {
        const int v = cached[dev].load(std::memory_order_relaxed);
        if (v > 0) return v;
    }
    int sms = 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
        (void)cudaGetLastError();
        sms = 148;
    }
    if (dev >= 0 && dev < kMaxDevices) cached[dev].store(sms, std::memory_order_relaxed);
    return sms;
}

template <typename OutT, int N, bool F, bool kPlanar, bool kSrcF32 = false>
cudaError_t launch_one(const TiledArgs& a, int dev, int grid, int smem, cudaStream_t stream) {
    auto kern = pva_tiled_kernel<OutT, N, F, kPlanar, kSrcF32>;
    // The opt-in shared-memory ceiling only ever needs raising (per device, per variant).  This is
    // the one piece of process-wide state the library keeps: a high-water mark per device, updated
    // with relaxed atomics (two threads racing here both make the same idempotent call).
    static std::atomic<int> configured[kMaxDevices];
    const bool tracked = dev >= 0 && dev < kMaxDevices;
    if (!tracked || configured[dev].load(std::memory_order_relaxed) < smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (tracked) {
            int seen = configured[dev].load(std::memory_order_relaxed);
            while (seen < smem && !configured[dev].compare_exchange_weak(seen, smem, std::memory_order_relaxed)) {
            }
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kBlockThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = a.g.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, a);
}

template <typename OutT, bool kPlanar>
cudaError_t launch_t(const TiledArgs& a, int norm, bool fastx, int dev, int grid, int smem, cudaStream_t stream) {
    if (a.p.src_f32) {
        if (norm == 0) return launch_one<OutT, 0, false, kPlanar, true>(a, dev, grid, smem, stream);
        if (norm == 1) return launch_one<OutT, 1, false, kPlanar, true>(a, dev, grid, smem, stream);
        if (norm == 2) return launch_one<OutT, 2, false, kPlanar, true>(a, dev, grid, smem, stream);
        return cudaErrorInvalidValue;
    }
#define PVA_TILED_CASE(N, F) \
    if (norm == N && fastx == F) return launch_one<OutT, N, F, kPlanar>(a, dev, grid, smem, stream);
    PVA_TILED_CASE(0, false)
    PVA_TILED_CASE(0, true)
    PVA_TILED_CASE(1, false)
    PVA_TILED_CASE(1, true)
    PVA_TILED_CASE(2, false)
    PVA_TILED_CASE(2, true)
#undef PVA_TILED_CASE
    return cudaErrorInvalidValue;
}

}  // namespace

cudaError_t launch_tiled(const KParams& p, int out_dtype, int layout, int norm, cudaStream_t stream, TiledConfig* cfg,
                         const char** why) {
    // The argument block (~14 KB) lives on this call's stack: nothing survives the call.
    TiledArgs a;
    *why = nullptr;
    const char* reason = "";
    // one plan per launch: it both decides whether the tiled kernel can take the shapes and carves
    // up shared memory
    if (!plan_tiles(p, out_dtype, layout, &a.g, &reason)) {
        *why = reason;
        return cudaErrorNotSupported;
    }
    // the fast x-path needs, for every clip of the launch: up-scaling or identity in x (so that
    // the taps of 4 adjacent outputs fit an aligned 8-byte window) and 4-byte aligned rows
    bool fastx = !p.src_f32;
    for (int i = 0; i < p.n_clips; ++i) {
        const ClipGeom& g = p.clips[i];
        fastx = fastx && (g.w <= g.ow) && (g.w % 4 == 0) && ((uintptr_t)g.src % 4 == 0);
    }
    a.p = p;
    // 2: wait for the previous kernel before touching memory at all; 1 (PVA_FLAG_CLIPS_READY): only
    // before the first store; 3 (+ PVA_FLAG_OUTPUTS_DISJOINT): only before the CTA exits.  Only the
    // caller's explicit flags ever relax the ordering.
    a.g.pdl = p.clips_ready ? (p.outputs_disjoint ? 3 : 1) : 2;
#ifdef PVA_EXPERIMENTS
    a.g.pdl = knob_env("PVA_TILED_PDL", a.g.pdl);  // 0 = no programmatic dependent launch at all
#endif
    a.g.tile0_first = PVA_KNOB("TILE0_FIRST", 1);
    a.g.l2hint = PVA_KNOB("L2HINT", 3);  // clip bytes are read once, outputs are consumed later: stream both
#ifdef PVA_EXPERIMENTS
    {
        // PVA_TILED_TRACE=<hex device address of a uint64 buffer, kMaxGrid * 128 entries per launch>,
        // PVA_TILED_TRACE_SLOTS=<launch records the buffer holds>: consecutive launches fill
        // consecutive records (scripts/trace_launch.py)
        static std::atomic<unsigned> launch_no{0};
        const char* tp = getenv("PVA_TILED_TRACE");
        if (tp) {
            const unsigned slots = (unsigned)knob_env("PVA_TILED_TRACE_SLOTS", 1);
            const unsigned n = launch_no.fetch_add(1);
            a.g.trace = reinterpret_cast<unsigned long long*>(strtoull(tp, nullptr, 16)) +
                        (size_t)(n % slots) * kMaxGrid * kTraceWords;
        }
    }
#endif
    int dev = 0;
    cudaGetDevice(&dev);
    const int sms = sm_count(dev);
    int grid = sms < a.g.total_tiles ? sms : a.g.total_tiles;
    if (grid > kMaxGrid) grid = kMaxGrid;
    // Streaming regime (both flags: clips ready, outputs in a ring): a launch takes HALF the SMs and the
    // next one starts beside it at once, so two launches are always in flight, half a launch apart.
    // A CTA then works through twice as many tiles, which halves the share of its life spent waiting
    // for its first tile and draining its last stores, and one launch's fill always overlaps the
    // other's steady state.  Measured on B200, headline batch: 31.2 vs 32.6 us per launch in a long
    // run, 33.4 vs 34.7 us over 20 launches -- at the price of a longer isolated launch (69 vs 42 us),
    // which a caller that does not stream never pays because it does not set the flags.
    if (a.g.pdl == 3 && p.max_ctas == 0 && grid == sms && PVA_KNOB("HALF_GRID", 1)) grid = (sms + 1) / 2;
    if (p.max_ctas > 0 && grid > p.max_ctas) grid = p.max_ctas;  // PvaPlan.max_ctas
    {
        const int cap = PVA_KNOB("MAXGRID", 0);
        if (cap > 0 && grid > cap) grid = cap;
    }
    {
        // Equal tile counts per CTA: the kernel is bound by its writes, which are the same for every
        // tile.  (Experiments: PVA_TILED_BALANCE=1 weights tiles by a compute-cost model instead --
        // emit + control per output row, x-pass per source row of a strip; measured 2 us slower.)
        const int tiles_per_clip = p.n_fast * a.g.bands;
        const bool balance = PVA_KNOB("BALANCE", 0) != 0;
        double total = 0.0;
        double w[kMaxClipsPerLaunch];
        for (int i = 0; i < p.n_clips; ++i) {
            const double src_rows_per_strip = a.g.rps * (double)p.clips[i].sy + 1.5;
            w[i] = balance ? 62.0 * a.g.rps + 70.0 * src_rows_per_strip : 1.0;
            total += w[i] * tiles_per_clip;
        }
        int clip = 0, t = 0;
        double acc = 0.0;
        a.tile_begin[0] = 0;
        for (int b = 1; b < grid; ++b) {
            const double target = total * b / grid;
            while (t < a.g.total_tiles && acc + 0.5 * w[clip] <= target) {
                acc += w[clip];
                if (++t % tiles_per_clip == 0) ++clip;
            }
            a.tile_begin[b] = t;
        }
        a.tile_begin[grid] = a.g.total_tiles;
    }
    const int smem = a.g.off_out + a.g.groups * a.g.n_strips * a.g.out_bufs * a.g.strip_bytes;
    const int kind = norm == kNormExact ? 0 : (a.g.na_uniform ? 1 : 2);
    cfg->rows_per_tile = a.g.R;
    cfg->smem_bytes = smem;
    cfg->grid = grid;
    cfg->block = kBlockThreads;
    cfg->tiles = a.g.total_tiles;
    const bool planar = layout == kLayoutCTHW;
    switch (out_dtype) {
        case 0: return planar ? launch_t<float, true>(a, kind, fastx, dev, grid, smem, stream)
                              : launch_t<float, false>(a, kind, fastx, dev, grid, smem, stream);
        case 1: return planar ? launch_t<__half, true>(a, kind, fastx, dev, grid, smem, stream)
                              : launch_t<__half, false>(a, kind, fastx, dev, grid, smem, stream);
        case 2: return planar ? launch_t<__nv_bfloat16, true>(a, kind, fastx, dev, grid, smem, stream)
                              : launch_t<__nv_bfloat16, false>(a, kind, fastx, dev, grid, smem, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace pva
