/*
 * pva_preproc.h -- C ABI of libpva_preproc.so: the B200 (sm_100a) replacement for the
 * per-clip video preprocessing path of nateraw/pytorchvideo-accelerate.
 *
 * The reference has no native interface: the path is a Python callable built by
 * make_transform (reference run.py:68-102) and run per clip on CPU DataLoader workers.
 * This header is the boundary a binding of that callable talks to.  Each entry point
 * names the reference code it replaces.  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *   - plain C types only; no torch / CUDA runtime types in the signatures
 *     (a stream is passed as void*, i.e. a cudaStream_t value; NULL = legacy default stream)
 *   - the library allocates nothing and frees nothing on the device; the caller owns
 *     the clips, the outputs and keeps them alive until the stream has run the launch
 *   - every function returns PVA_OK (0) or a negative PvaStatus and never throws or exits;
 *     pva_last_error_string() gives the text for the calling thread's last failure
 *   - launches are asynchronous on the caller's stream.  No device-side state and no
 *     allocations: descriptors and small tables travel as kernel arguments.  What the
 *     library does keep, on the host: per calling thread, the text of the last error and the
 *     PvaLaunchInfo of the last call; per process, two small per-device caches filled on first
 *     use (SM count; the largest dynamic shared-memory size already opted into per kernel
 *     variant), updated with atomics, so calls from several threads are safe.
 *   - the library reads no environment variables (the PVA_TILED_* experiment switches exist
 *     only in the separate -DPVA_EXPERIMENTS build that scripts/ uses)
 */
#ifndef PVA_PREPROC_H_
#define PVA_PREPROC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PVA_VERSION_MAJOR 0
#define PVA_VERSION_MINOR 1

typedef enum PvaStatus {
    PVA_OK = 0,
    PVA_ERR_INVALID_ARGUMENT = -1, /* NULL pointer, non-positive size, bad enum                       */
    PVA_ERR_CROP_TOO_LARGE = -2,   /* crop window does not fit the resized frame: the reference raises
                                      ValueError (torchvision transforms.py:647-648); no zero padding  */
    PVA_ERR_UNSUPPORTED = -3,      /* shape outside the limits listed below                           */
    PVA_ERR_CUDA = -4              /* a CUDA runtime call failed; text holds cudaGetErrorString       */
} PvaStatus;

typedef enum PvaDType { PVA_F32 = 0, PVA_F16 = 1, PVA_BF16 = 2 } PvaDType;

/* Output memory order of one clip.  The reference produces NCDHW-without-N, i.e.
 * C x T x S x S contiguous (run.py:64-65).  NDHWC is T x S x S x C: the
 * channels_last_3d order the SlowFast stem convolutions consume. */
typedef enum PvaLayout { PVA_LAYOUT_CTHW = 0, PVA_LAYOUT_THWC = 1 } PvaLayout;

/* How Div255 + Normalize (run.py:83-84) are evaluated.
 *   EXACT : per-tap table lut[u] = ((u / 255) - mean) / std with two IEEE fp32 divisions, then
 *           ATen's bilinear arithmetic: fp32 results are bit-identical to the reference ops.
 *   AFFINE: bilinear on the raw byte values, then one fused multiply-add per output with
 *           a = 1 / (255 std), b = -mean / std.  Same function, different rounding: differs from
 *           EXACT by at most a few fp32 ulps (< 4e-6 absolute), far inside the 1e-3 budget. */
typedef enum PvaNormMode { PVA_NORM_EXACT = 0, PVA_NORM_AFFINE = 1 } PvaNormMode;

/* Kernel selection.  AUTO picks the tiled bulk-copy kernel when the shapes allow it and the
 * generic one otherwise; the other values force one (used by the parity tests). */
typedef enum PvaKernel { PVA_KERNEL_AUTO = 0, PVA_KERNEL_GENERIC = 1, PVA_KERNEL_TILED = 2 } PvaKernel;

/* Element type of the source clip.  U8 is what the dataloader hands over; F32 holds the
 * same 0..255 values as floats (what pytorchvideo's decoder produces). */
typedef enum PvaSrcType { PVA_SRC_U8 = 0, PVA_SRC_F32 = 1 } PvaSrcType;

#define PVA_MAX_CHANNELS 4

/* PvaPlan.flags.
 * PVA_FLAG_CLIPS_READY: the caller asserts that the clip bytes were complete before the previous
 * KERNEL of this stream was launched -- they came from a copy, or from a kernel further back --
 * and that no clip aliases the outputs of the previous pva_preprocess call.  Consecutive calls
 * on one stream then overlap: the next launch reads clips and computes while the previous one is
 * still draining, and waits for it only before its first store (output buffers may be reused
 * freely; that ordering is always kept).  Without the flag a launch waits for the previous
 * kernel before it touches memory at all. */
#define PVA_FLAG_CLIPS_READY 1
/* PVA_FLAG_OUTPUTS_DISJOINT (only honoured together with PVA_FLAG_CLIPS_READY): the caller asserts
 * that the previous kernel of this stream neither reads nor writes any byte of this call's fast_out /
 * slow_out -- e.g. consecutive batches go to different buffers of a ring, as a prefetching data
 * pipeline has them anyway.  The launch then does not wait for the previous kernel at all before it
 * stores: it fills the SMs the previous launch's slowest blocks have not reached yet, and the two
 * overlap completely; unless max_ctas says otherwise it also takes only half of the SMs, so that two
 * consecutive launches run side by side, half a launch apart (better throughput for a stream of
 * launches, a longer single launch).  Stream order as seen from outside is unchanged (a block of this launch does
 * not exit before the previous kernel has completed, so whatever follows on the stream still sees
 * both finished).  Do NOT set it when the same output buffer is handed to consecutive calls, or when
 * outputs come fresh from a caching allocator that may recycle the previous call's block. */
#define PVA_FLAG_OUTPUTS_DISJOINT 2
/* Limits: 1..4 channels; n_fast, n_slow <= 128; t_in <= 65535; any h, w, crop.  More than 128 clips
 * or 8 distinct frame_idx tables per call are handled with several launches.  The tiled kernel takes
 * 3-channel uint8 or float32 clips, either layout and crops up to 2048; everything else runs on the
 * generic kernel (PVA_KERNEL_AUTO decides; PVA_KERNEL_TILED returns PVA_ERR_UNSUPPORTED instead). */

/*
 * One clip and its per-clip draws.  Replaces the arguments the reference passes implicitly
 * through the op chain of run.py:81-97:
 *   src, t_in, h, w  : the 'video' entry of the sample dict, C x T x H x W, contiguous, in DEVICE
 *                      memory; uint8, or float32 when plan->src_type == PVA_SRC_F32 (any float values;
 *                      EXACT evaluates them with the reference's two divisions per tap; both kernels
 *                      take them, the tiled one stages the bands as floats).
 *                      src may also point into PAGE-LOCKED HOST memory that is mapped into the
 *                      device's address space (cudaHostAlloc / cudaHostRegister, a pinned torch
 *                      tensor): the kernel then reads the frames and rows it needs straight over
 *                      PCIe and no host-to-device copy of the clip is needed at all -- for the usual
 *                      64 -> 32 frame case less than half of the clip ever crosses the bus.  A few
 *                      CTAs saturate the link: see PvaPlan.max_ctas.
 *   frame_idx        : HOST array of plan->n_fast indices into [0, t_in): the result of
 *                      UniformTemporalSubsample's linspace(0, T-1, n).long() (run.py:82).
 *                      Not closed-form in integers, hence a table.  Read during the call only.
 *   oh, ow           : frame size after (Random)ShortSideScale (run.py:89,95), computed by the
 *                      host with the reference's Python-double formula
 *   top, left        : crop origin inside the resized frame (RandomCrop / CenterCrop, run.py:90,96)
 *   flip             : nonzero = mirror columns inside the crop (RandomHorizontalFlip, run.py:91)
 */
typedef struct PvaClipDesc {
    const void* src;
    const int32_t* frame_idx;
    int32_t t_in, h, w;
    int32_t oh, ow;
    int32_t top, left;
    int32_t flip;
} PvaClipDesc;

/*
 * What is common to all clips of a call.  Replaces the keyword arguments of
 * make_transform (run.py:68-79) that shape the output:
 *   channels         : C (1..PVA_MAX_CHANNELS); len(mean) in the reference
 *   n_fast           : num_frames
 *   n_slow, slow_idx : 0 / NULL for a single pathway (is_slowfast=False); otherwise
 *                      n_fast // alpha and the HOST table linspace(0, n_fast-1, n_slow).long()
 *                      of PackPathway (run.py:59-63): slow[j] = fast[slow_idx[j]]
 *   crop             : crop_size; outputs are crop x crop
 *   mean, std        : Normalize (run.py:84)
 *   max_ctas         : 0 = one CTA per SM.  A positive value caps the tiled kernel's grid; meant for
 *                      launches that read their clips from mapped host memory, which are bound by
 *                      PCIe (4 CTAs already saturate it) and should leave the other SMs to whatever
 *                      else runs on the device.  No counterpart in the reference.
 */
typedef struct PvaPlan {
    int32_t channels;
    int32_t n_fast;
    int32_t n_slow;
    int32_t crop;
    const int32_t* slow_idx;
    float mean[PVA_MAX_CHANNELS];
    float std[PVA_MAX_CHANNELS];
    int32_t out_dtype;  /* PvaDType   */
    int32_t out_layout; /* PvaLayout  */
    int32_t norm_mode;  /* PvaNormMode */
    int32_t kernel;     /* PvaKernel  */
    int32_t src_type;   /* PvaSrcType */
    int32_t flags;      /* PVA_FLAG_* */
    int32_t max_ctas;   /* 0 = all SMs */
    int32_t reserved;   /* must be 0 */
} PvaPlan;

/* What the last pva_preprocess call of this thread launched (for tests and bench.py). */
typedef struct PvaLaunchInfo {
    int32_t kernel;       /* PvaKernel actually used                      */
    int32_t launches;     /* kernel launches issued by the call           */
    int32_t grid, block;  /* of the last launch                           */
    int32_t smem_bytes;   /* dynamic shared memory of the last launch     */
    int32_t rows_per_tile;
    int32_t tiles;        /* work items over all launches                 */
    int32_t reserved;
} PvaLaunchInfo;

/*
 * Run the whole transform chain for n_clips clips on the current device:
 *   UniformTemporalSubsample -> Div255 -> Normalize -> (Random)ShortSideScale ->
 *   RandomCrop/CenterCrop -> RandomHorizontalFlip -> PackPathway
 * i.e. everything ApplyTransformToKey('video', Compose([...])) does at run.py:81-102,
 * followed by the stacking default_collate does at run.py:170-183.
 *
 *   fast_out : DEVICE, n_clips clips of C x n_fast x crop x crop elements (layout/dtype per plan),
 *              clip i at element offset i * C * n_fast * crop * crop
 *   slow_out : DEVICE, same with n_slow frames; must be NULL iff plan->n_slow == 0
 *   stream   : cudaStream_t value the launches are enqueued on
 */
int pva_preprocess(const PvaClipDesc* clips, int32_t n_clips, const PvaPlan* plan,
                   void* fast_out, void* slow_out, void* stream);

/* Fill `lut` (HOST, channels * 256 floats) with ((u / 255) - mean[c]) / std[c] exactly as
 * Div255 + Normalize evaluate it (run.py:83-84): two IEEE fp32 divisions. */
int pva_build_norm_lut(const float* mean, const float* std, int32_t channels, float* lut);

/* Source footprint of a crop window: number of distinct source rows (*fh) and columns (*fw)
 * the bilinear taps of the window touch.  Used for the algorithmic-bytes figure. */
int pva_source_footprint(int32_t h, int32_t w, int32_t oh, int32_t ow, int32_t top, int32_t left,
                         int32_t crop, int32_t* fh, int32_t* fw);

/*
 * Host -> device copy of only the frames UniformTemporalSubsample (run.py:82) would select,
 * for one clip: the frame gather is folded into the DMA addressing, so half of a 64-frame clip
 * never crosses PCIe.  Replaces the device placement of the batch (run.py:196-198) for callers
 * that keep decoded clips in pinned host memory.
 *
 *   host_src  : HOST (pinned, for the copy to be asynchronous), uint8 [channels][t_in][h][w]
 *   dev_dst   : DEVICE, uint8 [channels][n_frames][h][w]; preprocess it with t_in = n_frames and
 *               frame_idx = 0, 1, ..., n_frames - 1
 *   frame_idx : HOST, n_frames indices into [0, t_in)
 * Issues one cudaMemcpy2DAsync per run of equally spaced indices and channel on `stream`;
 * allocates nothing.
 */
int pva_upload_selected_frames(const void* host_src, void* dev_dst, int32_t channels, int32_t t_in, int32_t h,
                               int32_t w, const int32_t* frame_idx, int32_t n_frames, void* stream);

int pva_last_launch_info(PvaLaunchInfo* info);

/* (major << 16) | minor */
int32_t pva_version(void);

/* Text of the calling thread's most recent failure ("" if none).  Never NULL. */
const char* pva_last_error_string(void);

#ifdef __cplusplus
}
#endif
#endif /* PVA_PREPROC_H_ */
