"""Reference-facing interface of the GPU preprocessing path.

``make_transform`` keeps the signature, defaults and dict-in/dict-out behaviour of the
reference factory (``run.py:68-102``); what it returns runs the whole op chain as ONE fused
CUDA launch instead of seven CPU passes.  Documented deviations from the reference output
(SURVEY.md section 8b): dtype fp16 by default (not fp32), device cuda, and -- unless
``channels_last=False`` -- the tensors are ``C x T x S x S`` *views* over ``T x S x S x C``
memory (``channels_last_3d`` once batched), which is what the SlowFast stem consumes.

The batched hand-over replaces the DataLoader/collate boundary (``run.py:170-183``):

* ``SampleParams``       worker side: makes the random draws with the reference's RNG calls,
                         leaves the clip as raw uint8 (optionally frame-subsampled)
* ``collate_raw_clips``  ``collate_fn``: stacks raw clips + an int32 draw table
* ``GpuPreprocess``      training-process side: one launch per batch on the rank's device,
                         applied where the reference reads ``batch["video"]`` (``run.py:251-253``)

There is no CPU implementation here; without the CUDA library every call raises.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Union

import torch

from . import _lib
from .geometry import ClipParams, pack_indices, sample_train_params, temporal_indices, val_params

_DTYPES = {torch.float32: _lib.PVA_F32, torch.float16: _lib.PVA_F16, torch.bfloat16: _lib.PVA_BF16}
_NORM = {"exact": _lib.PVA_NORM_EXACT, "affine": _lib.PVA_NORM_AFFINE}
_KERNEL = {"auto": _lib.PVA_KERNEL_AUTO, "generic": _lib.PVA_KERNEL_GENERIC, "tiled": _lib.PVA_KERNEL_TILED}

PARAMS_KEY = "pva_params"


class PackPathway(torch.nn.Module):
    """Same contract as the reference module (``run.py:38-65``): ``frames -> [slow, fast]``.

    Kept for callers that pack an already-preprocessed tensor; inside ``make_transform``
    the split is folded into the fused kernel's store addressing instead.
    """

    def __init__(self, alpha: int):
        super().__init__()
        self.alpha = alpha

    def forward(self, frames: torch.Tensor) -> List[torch.Tensor]:
        idx = torch.tensor(pack_indices(frames.shape[1], self.alpha), dtype=torch.long, device=frames.device)
        return [torch.index_select(frames, 1, idx), frames]


class _IdxTable:
    """Host int32 array kept alive for the duration of a call."""

    def __init__(self, values: Sequence[int]):
        self.n = len(values)
        self.array = (ctypes.c_int32 * max(self.n, 1))(*values)

    @property
    def ptr(self):
        return ctypes.cast(self.array, ctypes.POINTER(ctypes.c_int32))


class GpuClipTransform:
    """Callable returned by :func:`make_transform`.

    ``transform(sample_dict)`` mirrors ``Compose([ApplyTransformToKey('video', ...),
    RemoveKey('audio')])`` (``run.py:102``): it mutates and returns the same dict, drops
    ``'audio'``, and replaces ``'video'`` (uint8 ``C x T x H x W``) by ``[slow, fast]`` -- or by a
    single tensor when ``is_slowfast=False``.
    """

    def __init__(
        self,
        min_short_side_scale: int = 256,
        max_short_side_scale: int = 320,
        mean: Sequence[float] = (0.45, 0.45, 0.45),
        std: Sequence[float] = (0.225, 0.225, 0.225),
        crop_size: int = 256,
        num_frames: int = 32,
        slowfast_alpha: int = 4,
        is_slowfast: bool = True,
        horizontal_flip_p: float = 0.5,
        training: bool = False,
        *,
        device: Union[None, str, torch.device] = None,
        out_dtype: torch.dtype = torch.float16,
        channels_last: bool = True,
        norm_mode: str = "affine",
        kernel: str = "auto",
    ):
        if len(mean) != len(std):
            raise ValueError("mean and std must have the same length")
        if not 1 <= len(mean) <= _lib.PVA_MAX_CHANNELS:
            raise NotImplementedError(f"{len(mean)} channels: the kernels handle 1..{_lib.PVA_MAX_CHANNELS}")
        if out_dtype not in _DTYPES:
            raise TypeError(f"out_dtype must be one of {list(_DTYPES)}")
        if norm_mode not in _NORM or kernel not in _KERNEL:
            raise ValueError(f"norm_mode in {list(_NORM)}, kernel in {list(_KERNEL)}")
        if num_frames < 1 or crop_size < 1:
            raise ValueError("num_frames and crop_size must be positive")
        if is_slowfast and num_frames // slowfast_alpha < 1:
            raise ValueError("num_frames // slowfast_alpha must be at least 1")
        self.min_short_side_scale = int(min_short_side_scale)
        self.max_short_side_scale = int(max_short_side_scale)
        self.mean = [float(m) for m in mean]
        self.std = [float(s) for s in std]
        self.crop_size = int(crop_size)
        self.num_frames = int(num_frames)
        self.slowfast_alpha = int(slowfast_alpha)
        self.is_slowfast = bool(is_slowfast)
        self.horizontal_flip_p = float(horizontal_flip_p)
        self.training = bool(training)
        self.device = torch.device(device) if device is not None else None
        self.out_dtype = out_dtype
        self.channels_last = bool(channels_last)
        self.norm_mode = norm_mode
        self.kernel = kernel
        self.n_slow = self.num_frames // self.slowfast_alpha if self.is_slowfast else 0
        self._slow_tab = _IdxTable(pack_indices(self.num_frames, self.slowfast_alpha) if self.is_slowfast else ())
        self._idx_cache: Dict[int, _IdxTable] = {}
        # the reference converts mean/std to fp32 tensors (torchvision normalize): round once, here
        self._mean32 = [float(torch.tensor(m, dtype=torch.float32)) for m in self.mean]
        self._std32 = [float(torch.tensor(v, dtype=torch.float32)) for v in self.std]
        self._lib = _lib.load_library()  # loud failure here, not at first use

    # ---- host-side draws -------------------------------------------------------------
    def sample_params(self, h: int, w: int) -> ClipParams:
        """Draws for one clip of ``h x w`` frames (global torch RNG when training)."""
        if self.training:
            return sample_train_params(h, w, self.min_short_side_scale, self.max_short_side_scale, self.crop_size,
                                       self.horizontal_flip_p)
        return val_params(h, w, self.min_short_side_scale, self.crop_size)

    # ---- launch ----------------------------------------------------------------------
    def _upload_table(self, t_in: int) -> _IdxTable:
        """Frame table for pva_upload_selected_frames (same indices the kernel would gather)."""
        return self._frame_table(t_in)

    def _frame_table(self, t_in: int) -> _IdxTable:
        tab = self._idx_cache.get(t_in)
        if tab is None:
            tab = self._idx_cache[t_in] = _IdxTable(temporal_indices(t_in, self.num_frames))
        return tab

    def _plan(self, src_dtype: torch.dtype = torch.uint8, clips_ready: bool = False, max_ctas: int = 0,
              outputs_disjoint: bool = False) -> _lib.PvaPlan:
        plan = _lib.PvaPlan()
        plan.flags = (_lib.PVA_FLAG_CLIPS_READY if clips_ready else 0) | (
            _lib.PVA_FLAG_OUTPUTS_DISJOINT if outputs_disjoint else 0)
        plan.max_ctas = int(max_ctas)
        plan.channels = len(self.mean)
        plan.n_fast = self.num_frames
        plan.n_slow = self.n_slow
        plan.crop = self.crop_size
        plan.slow_idx = self._slow_tab.ptr if self.n_slow else None
        for c, (m, s) in enumerate(zip(self._mean32, self._std32)):
            plan.mean[c] = m
            plan.std[c] = s
        plan.out_dtype = _DTYPES[self.out_dtype]
        plan.out_layout = _lib.PVA_LAYOUT_THWC if self.channels_last else _lib.PVA_LAYOUT_CTHW
        plan.norm_mode = _NORM[self.norm_mode]
        plan.kernel = _KERNEL[self.kernel]
        plan.src_type = _lib.PVA_SRC_F32 if src_dtype == torch.float32 else _lib.PVA_SRC_U8
        return plan

    def _check_clip(self, video: torch.Tensor) -> None:
        if not isinstance(video, torch.Tensor):
            raise TypeError(f"video should be a torch.Tensor, got {type(video)}")
        if video.dim() != 4:
            raise ValueError(f"video should be C x T x H x W, got shape {tuple(video.shape)}")
        if video.dtype not in (torch.uint8, torch.float32):
            # uint8 is the raw-frame hand-over; float32 0..255 is what pytorchvideo's decoder produces
            raise TypeError(f"video should be uint8 or float32 frames in 0..255, got {video.dtype}")
        if video.shape[0] != len(self.mean):
            raise ValueError(f"video has {video.shape[0]} channels, mean/std have {len(self.mean)}")
        if video.numel() == 0:
            raise ValueError(f"empty clip of shape {tuple(video.shape)}")

    def _alloc(self, n: int, frames: int, device: torch.device) -> torch.Tensor:
        c, s = len(self.mean), self.crop_size
        if self.channels_last:
            return torch.empty((n, frames, s, s, c), dtype=self.out_dtype, device=device)
        return torch.empty((n, c, frames, s, s), dtype=self.out_dtype, device=device)

    def _logical(self, t: torch.Tensor) -> torch.Tensor:
        """``N x C x T x S x S`` view whatever the memory order."""
        return t.permute(0, 4, 1, 2, 3) if self.channels_last else t

    def prepare_batch(self, clips: Union[torch.Tensor, Sequence[torch.Tensor]], params: Sequence[ClipParams],
                      out: Optional[Sequence[torch.Tensor]] = None, clips_ready: bool = False,
                      max_ctas: int = 0, device: Union[None, str, torch.device] = None,
                      outputs_disjoint: bool = False) -> "PreparedBatch":
        """Validate a batch and build its descriptor table once; ``.launch()`` then costs a
        single C call.  ``clips``: a ``B x C x T x H x W`` uint8 cuda tensor or a sequence of
        ``C x T x H x W`` uint8 cuda tensors (ragged shapes allowed).  ``out`` may supply raw
        output buffers from :meth:`alloc_outputs`.

        PINNED host clips (``tensor.pin_memory()``) are accepted as well and read IN PLACE: the
        kernel fetches the frames and rows it needs over PCIe, no copy of the clip is made
        (``include/pva_preproc.h``, PvaClipDesc.src).  Such a launch is PCIe-bound; ``max_ctas``
        (e.g. 8) keeps it from occupying every SM while it waits for the bus.  The caller keeps
        the host clips unchanged until the launch has run.

        ``clips_ready=True`` sets ``PVA_FLAG_CLIPS_READY`` (``include/pva_preproc.h``): say so when
        the clip bytes were complete before the previous kernel on the launch stream started (they
        came from a copy, or have been resident for a while) -- consecutive launches then overlap.
        Leave it off when a kernel enqueued right before this call produces the clips.

        ``outputs_disjoint=True`` adds ``PVA_FLAG_OUTPUTS_DISJOINT``: the caller supplies ``out`` and
        vouches that the previous kernel on the launch stream neither reads nor writes those buffers
        (consecutive batches go to different slots of a ring).  The launch then overlaps the previous
        one's tail completely.  Never combine it with outputs allocated per call.
        """
        if outputs_disjoint and out is None:
            raise ValueError("outputs_disjoint needs caller-owned output buffers (out=...)")
        clip_list = list(clips.unbind(0)) if isinstance(clips, torch.Tensor) and clips.dim() == 5 else list(clips)
        n = len(clip_list)
        if len(params) != n:
            raise ValueError(f"{n} clips but {len(params)} parameter rows")
        if n == 0:
            raise ValueError("empty batch")
        src_device = clip_list[0].device
        if src_device.type == "cuda":
            device = src_device
        else:
            if not all(v.is_pinned() for v in clip_list):
                raise ValueError("host clips must be page-locked (pin_memory()) to be read in place; "
                                 "use __call__ / GpuPreprocess to have pageable clips copied")
            device = torch.device(device) if device is not None else self._device_for(clip_list[0])
        descs = (_lib.PvaClipDesc * n)()
        keep = []
        for i, (v, p) in enumerate(zip(clip_list, params)):
            self._check_clip(v)
            if v.device != src_device:
                raise ValueError("all clips of a batch must live on one device")
            if not v.is_contiguous():
                if not v.is_cuda:
                    # .contiguous() would hand the kernel a fresh PAGEABLE copy, which it cannot read
                    raise ValueError("host clips that are read in place must be contiguous (got a strided view); "
                                     "make them contiguous before pinning")
                v = v.contiguous()
            if v.dtype != clip_list[0].dtype:
                raise TypeError("all clips of a batch must share one dtype")
            keep.append(v)
            _c, t_in, h, w = v.shape
            if p.oh < self.crop_size or p.ow < self.crop_size:
                raise ValueError(
                    f"Required crop size {(self.crop_size, self.crop_size)} is larger than input image size {(p.oh, p.ow)}")
            d = descs[i]
            d.src = v.data_ptr()
            d.frame_idx = self._frame_table(t_in).ptr
            d.t_in, d.h, d.w = t_in, h, w
            d.oh, d.ow, d.top, d.left, d.flip = p.oh, p.ow, p.top, p.left, int(p.flip)
        if out is None:
            out = self.alloc_outputs(n, device)
        return PreparedBatch(self, descs, n, keep, list(out), device, clip_list[0].dtype, clips_ready, max_ctas,
                             outputs_disjoint)

    def run_batch(self, clips: Union[torch.Tensor, Sequence[torch.Tensor]], params: Sequence[ClipParams],
                  out: Optional[Sequence[torch.Tensor]] = None,
                  stream: Optional[torch.cuda.Stream] = None) -> List[torch.Tensor]:
        """Preprocess device-resident uint8 clips with explicit draws in one fused launch.
        Returns ``[slow, fast]`` (``[fast]`` for a single pathway), each ``B x C x T' x S x S``."""
        return self.prepare_batch(clips, params, out).launch(stream)

    def alloc_outputs(self, n: int, device: torch.device) -> List[torch.Tensor]:
        """Raw output buffers ``[slow_raw, fast_raw]`` (``[fast_raw]`` single pathway) for ``n`` clips."""
        bufs = [self._alloc(n, self.num_frames, device)]
        if self.n_slow:
            bufs.insert(0, self._alloc(n, self.n_slow, device))
        return bufs

    def _device_for(self, video: torch.Tensor) -> torch.device:
        if video.is_cuda:
            return video.device
        if self.device is not None:
            return self.device
        if not torch.cuda.is_available():
            raise _lib.PvaError("no CUDA device: this path has no CPU implementation")
        return torch.device("cuda", torch.cuda.current_device())

    def explicit(self, video: torch.Tensor, params: ClipParams) -> Union[torch.Tensor, List[torch.Tensor]]:
        """The transform with the draws given (the parity-test form): one clip in,
        ``[slow, fast]`` (or a tensor) of shape ``C x T' x S x S`` out."""
        self._check_clip(video)
        dev = self._device_for(video)
        v = video if video.is_cuda else video.to(dev, non_blocking=True)
        res = [t[0] for t in self.run_batch([v], [params])]
        return res if self.is_slowfast else res[0]

    def __call__(self, sample: dict) -> dict:
        video = sample["video"]
        self._check_clip(video)
        params = self.sample_params(int(video.shape[2]), int(video.shape[3]))
        sample["video"] = self.explicit(video, params)
        if "audio" in sample:
            del sample["audio"]
        return sample


class PreparedBatch:
    """A validated batch: descriptor table, plan and output buffers, ready to (re)launch."""

    def __init__(self, transform: GpuClipTransform, descs, n: int, clips: List[torch.Tensor],
                 out: List[torch.Tensor], device: torch.device, src_dtype: torch.dtype = torch.uint8,
                 clips_ready: bool = False, max_ctas: int = 0, outputs_disjoint: bool = False):
        self.t = transform
        self.descs = descs
        self.n = n
        self.clips = clips
        self.raw_out = out
        self.device = device
        self.plan = transform._plan(src_dtype, clips_ready, max_ctas, outputs_disjoint)
        self._plan_ref = ctypes.byref(self.plan)
        self.slow_raw, self.fast_raw = (out[0], out[1]) if transform.n_slow else (None, out[-1])
        self._fast_ptr = self.fast_raw.data_ptr()
        self._slow_ptr = self.slow_raw.data_ptr() if self.slow_raw is not None else None
        self._fn = transform._lib.pva_preprocess

    @property
    def outputs(self) -> List[torch.Tensor]:
        res = [self.t._logical(self.fast_raw)]
        if self.slow_raw is not None:
            res.insert(0, self.t._logical(self.slow_raw))
        return res

    def launch_raw(self, cuda_stream: int) -> None:
        """Enqueue on a raw ``cudaStream_t`` value; the caller guarantees the current device
        is this batch's device and keeps the clips alive."""
        rc = self._fn(self.descs, self.n, self._plan_ref, self._fast_ptr, self._slow_ptr, cuda_stream)
        if rc != 0:
            _lib.raise_for_status(rc)

    def launch(self, stream: Optional[torch.cuda.Stream] = None) -> List[torch.Tensor]:
        with torch.cuda.device(self.device):
            s = stream if stream is not None else torch.cuda.current_stream(self.device)
            self.launch_raw(s.cuda_stream)
        for v in self.clips:  # inputs must outlive the asynchronous launch on this stream
            if v.is_cuda:
                v.record_stream(s)  # (pinned host clips read in place: the caller keeps them, see prepare_batch)
        return self.outputs


def make_transform(
    min_short_side_scale=256,
    max_short_side_scale=320,
    mean=[0.45, 0.45, 0.45],
    std=[0.225, 0.225, 0.225],
    crop_size=256,
    num_frames=32,
    slowfast_alpha=4,
    is_slowfast=True,
    horizontal_flip_p=0.5,
    training=False,
    **gpu_options,
) -> GpuClipTransform:
    """Drop-in for the reference factory (``run.py:68-102``): same keyword arguments, same
    defaults, same dict-in/dict-out callable.  ``gpu_options`` (``device``, ``out_dtype``,
    ``channels_last``, ``norm_mode``, ``kernel``) are the only additions."""
    return GpuClipTransform(min_short_side_scale, max_short_side_scale, mean, std, crop_size, num_frames,
                            slowfast_alpha, is_slowfast, horizontal_flip_p, training, **gpu_options)


# --------------------------------------------------------------------------------------
# Batched hand-over: DataLoader workers draw, the training process preprocesses on device.
# --------------------------------------------------------------------------------------
class SampleParams:
    """Worker-side stand-in for the transform (``run.py:156,165``).

    Makes exactly the RNG calls the reference's train transform makes (so a seeded worker
    reproduces its draws), stores them as an int32 row under ``'pva_params'``, removes
    ``'audio'`` and leaves ``'video'`` as raw uint8.  With ``subsample_on_worker=True`` the
    uint8 frame gather of ``UniformTemporalSubsample`` is done here, halving the bytes that
    cross PCIe for the usual 64 -> 32 frame case; the kernel then reads every frame it gets.

    pytorchvideo's decoder hands over float32 frames holding the decoded bytes (0..255, whole
    numbers).  ``narrow_float_frames=True`` (default) turns such a clip back into uint8 here
    when, and only when, that is lossless -- the result of the transform is unchanged, the clip
    crosses PCIe at a quarter of the size and takes the uint8 kernel.  A float clip with
    fractional or out-of-range values is shipped as float32.
    """

    def __init__(self, transform: GpuClipTransform, subsample_on_worker: bool = False,
                 narrow_float_frames: bool = True):
        self.t = transform
        self.subsample_on_worker = subsample_on_worker
        self.narrow_float_frames = narrow_float_frames

    def __call__(self, sample: dict) -> dict:
        video = sample["video"]
        self.t._check_clip(video)
        p = self.t.sample_params(int(video.shape[2]), int(video.shape[3]))
        if self.subsample_on_worker:
            idx = torch.tensor(temporal_indices(int(video.shape[1]), self.t.num_frames), dtype=torch.long)
            video = sample["video"] = torch.index_select(video, 1, idx)
        if self.narrow_float_frames and video.dtype == torch.float32:
            as_bytes = video.to(torch.uint8)
            if torch.equal(as_bytes.to(torch.float32), video):  # whole numbers in 0..255: nothing is lost
                sample["video"] = as_bytes
        sample[PARAMS_KEY] = torch.tensor(p.as_row(), dtype=torch.int32)
        if "audio" in sample:
            del sample["audio"]
        return sample


def collate_raw_clips(samples: Sequence[dict]) -> dict:
    """``collate_fn`` for raw clips (replaces ``default_collate`` at ``run.py:170-183``).

    Same-shape clips are stacked into one ``B x C x T x H x W`` uint8 tensor; ragged clips
    stay a list.  The draw rows become an int32 ``B x 6`` table; every other key goes
    through ``default_collate`` unchanged.
    """
    from torch.utils.data import default_collate

    videos = [s["video"] for s in samples]
    rows = torch.stack([s[PARAMS_KEY] for s in samples]).to(torch.int32)
    rest = [{k: v for k, v in s.items() if k not in ("video", PARAMS_KEY)} for s in samples]
    batch = default_collate(rest) if rest and rest[0] else {}
    same = all(v.shape == videos[0].shape for v in videos)
    batch["video"] = torch.stack(videos) if same else videos
    batch[PARAMS_KEY] = rows
    return batch


class GpuPreprocess:
    """Device side of the hand-over: ``batch -> batch`` with ``batch['video'] = [slow, fast]``
    (``B x C x T' x S x S`` each), applied right where the reference reads ``batch["video"]``
    (``run.py:251-253``).  Host clips are copied to the rank's device first (pinned memory
    makes that copy asynchronous).  Each rank preprocesses only its own batches: the path
    shards by clip and needs no collective.

    ``read_in_place=True``: a pinned uint8 batch (``DataLoader(pin_memory=True)``) is not copied
    at all -- the kernel reads the selected frames' rows from host memory over PCIe on ``max_ctas``
    SMs, which leaves the rest of the device to the model.  Pageable or float batches take the
    copy path as before.
    """

    def __init__(self, transform: GpuClipTransform, device: Union[None, str, torch.device] = None,
                 upload_selected_frames: bool = False, read_in_place: bool = False, max_ctas: int = 4):
        self.t = transform
        self.device = torch.device(device) if device is not None else transform.device
        # copy only the frames UniformTemporalSubsample selects (strided DMA, include/pva_preproc.h):
        # half the PCIe bytes for 64 -> 32 frames; needs a stacked, contiguous host batch
        self.upload_selected_frames = bool(upload_selected_frames)
        # read_in_place: no copy at all when the batch is pinned (DataLoader(pin_memory=True)) -- the
        # kernel fetches the frames and rows it needs over PCIe, on max_ctas SMs
        self.read_in_place = bool(read_in_place)
        self.max_ctas = int(max_ctas)
        self._in_flight: List[tuple] = []  # (event, host clips) of launches that may still be reading

    def _upload_selected(self, video: torch.Tensor, dev: torch.device) -> torch.Tensor:
        b, c, t_in, h, w = video.shape
        n = self.t.num_frames
        tab = self.t._upload_table(t_in)
        out = torch.empty((b, c, n, h, w), dtype=torch.uint8, device=dev)
        src, dst = video.data_ptr(), out.data_ptr()
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            for i in range(b):
                rc = self.t._lib.pva_upload_selected_frames(src + i * c * t_in * h * w, dst + i * c * n * h * w,
                                                            c, t_in, h, w, tab.ptr, n, stream)
                _lib.raise_for_status(rc)
        return out

    def __call__(self, batch: dict) -> dict:
        dev = self.device
        if dev is None:
            if not torch.cuda.is_available():
                raise _lib.PvaError("no CUDA device: this path has no CPU implementation")
            dev = torch.device("cuda", torch.cuda.current_device())
        video = batch["video"]
        rows = batch[PARAMS_KEY]
        params = [ClipParams.from_row(r) for r in rows.tolist()]
        # clips that this call copies from the host are complete before any kernel it launches (a copy
        # is not a kernel: the launch is ordered after it in full); device tensors may have just been
        # produced by a kernel of the caller, so they get the conservative ordering
        from_host = (not video.is_cuda) if isinstance(video, torch.Tensor) else all(not v.is_cuda for v in video)
        host_list = [video] if isinstance(video, torch.Tensor) else list(video)
        if (self.read_in_place and from_host and all(v.is_pinned() and v.dtype == torch.uint8 for v in host_list)):
            self._in_flight = [(e, v) for e, v in self._in_flight if not e.query()]
            out = self.t.prepare_batch(video, params, clips_ready=True, max_ctas=self.max_ctas, device=dev).launch()
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(dev))
            self._in_flight.append((done, video))  # the pinned clips stay ours until the kernel has read them
            batch["video"] = out if self.t.is_slowfast else out[0]
            del batch[PARAMS_KEY]
            return batch
        raw_dma = (self.upload_selected_frames and isinstance(video, torch.Tensor) and not video.is_cuda
                   and video.dtype == torch.uint8 and video.is_contiguous())
        if raw_dma:
            clips = self._upload_selected(video, dev)
        elif isinstance(video, torch.Tensor):
            clips = video.to(dev, non_blocking=True)
        else:
            clips = [v.to(dev, non_blocking=True) for v in video]
        if raw_dma:
            # The strided copies were queued through the C ABI, behind torch's back: its host allocator
            # does not know that the pinned batch is still being read and would recycle the block as
            # soon as the batch dict drops it.  Keep the batch until the copies have run.
            self._in_flight = [(e, v) for e, v in self._in_flight if not e.query()]
            copied = torch.cuda.Event()
            copied.record(torch.cuda.current_stream(dev))
            self._in_flight.append((copied, video))
        out = self.t.prepare_batch(clips, params, clips_ready=from_host).launch()
        batch["video"] = out if self.t.is_slowfast else out[0]
        del batch[PARAMS_KEY]
        return batch
