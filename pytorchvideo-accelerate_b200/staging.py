"""Overlapped host -> device -> host staging around the fused kernel.

The reference moves fp32 batches to the device after preprocessing them on the CPU
(``run.py:196-198``).  Here the raw uint8 batch crosses PCIe instead and the copy is taken
off the critical path: ``StagedPreprocess`` keeps ``depth`` device-side slots and three
streams, so that while the kernel works on batch ``i`` the clips of batch ``i+1`` are already
in flight and -- when the caller asks for host copies of the result -- batch ``i-1`` is on its
way back.  PCIe is full duplex, so the two copy directions overlap as well.

This is plumbing (torch streams, events and ``copy_``); the arithmetic is the same single
``pva_preprocess`` launch as everywhere else.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

import ctypes

from . import _lib
from .geometry import ClipParams, temporal_indices
from .transforms import GpuClipTransform


def _flat_outputs(t: GpuClipTransform, batch: int, device, pin: bool):
    """[slow_raw, fast_raw] (or [fast_raw]) as views of ONE contiguous byte buffer, so that a result
    crosses PCIe as a single copy instead of one per pathway.  Returns (flat, views)."""
    c, s = len(t.mean), t.crop_size
    esz = torch.empty((), dtype=t.out_dtype).element_size()
    frames = ([t.n_slow] if t.n_slow else []) + [t.num_frames]
    sizes = [batch * f * s * s * c * esz for f in frames]
    offs, total = [], 0
    for b in sizes:
        offs.append(total)
        total += -(-b // 256) * 256  # every view starts 256-byte aligned (bulk stores need 16)
    flat = torch.empty(total, dtype=torch.uint8, pin_memory=True) if pin else torch.empty(total, dtype=torch.uint8, device=device)
    views = []
    for f, o, b in zip(frames, offs, sizes):
        shape = (batch, f, s, s, c) if t.channels_last else (batch, c, f, s, s)
        views.append(flat[o:o + b].view(t.out_dtype).view(shape))
    return flat, views


class _Slot:
    def __init__(self, t: GpuClipTransform, batch: int, clip_shape: Sequence[int], device: torch.device,
                 host_out: bool, selected_frames: bool, in_place: bool):
        c, t_in, h, w = clip_shape
        frames = t.num_frames if selected_frames else t_in
        # no device copy of the clips when the kernel reads the pinned host batch in place
        self.dev_in = None if in_place else torch.empty((batch, c, frames, h, w), dtype=torch.uint8, device=device)
        self.flat_out, self.raw_out = _flat_outputs(t, batch, device, pin=False)
        self.host_flat, self.host_out = _flat_outputs(t, batch, device, pin=True) if host_out else (None, None)
        self.in_ready = torch.cuda.Event()
        self.kernel_done = torch.cuda.Event()
        self.out_ready = torch.cuda.Event()
        self.released = None  # event recorded by release(): the consumer is done with raw_out
        self.host_clips = None  # the pinned batch a queued copy or kernel may still be reading
        self.consumer_stream = None  # stream that was current when result() handed out raw_out
        self.used = False


class StagedPreprocess:
    """``submit(host_clips, params)`` enqueues H2D + kernel (+ D2H) for one batch and returns a
    ticket; ``result(ticket)`` waits for that batch only.  ``host_clips`` must be a pinned
    ``B x C x T x H x W`` uint8 tensor of the shape given at construction.

    With ``upload_selected_frames=True`` only the frames ``UniformTemporalSubsample`` selects
    (``run.py:82``) are copied to the device -- ``pva_upload_selected_frames`` folds the gather
    into strided DMA copies -- which halves the PCIe traffic of the usual 64 -> 32 frame case; the
    kernel then reads every frame of the compact device clip.

    With ``read_in_place=True`` there is no H2D copy at all: the kernel is pointed at the pinned host
    batch and fetches the selected frames' rows over PCIe itself (``include/pva_preproc.h``,
    PvaClipDesc.src), on ``max_ctas`` SMs.  Fewer bytes cross the bus than with either copy (only the
    rows the crop windows touch), and no copy engine is involved.

    Lifetime rules.  In every upload mode the host batch is read ASYNCHRONOUSLY (by a copy or by the
    kernel itself): the caller must leave a submitted host batch unchanged until its ticket has been
    collected with ``result()``; the pipeline keeps a reference to it until then, so the tensor is
    not freed (and its pinned block not recycled by torch's host allocator) under a queued read.
    With ``host_out=False`` ``result()`` hands out views of the slot's device buffers; the slot is
    rewritten ``depth`` submits later.  ``submit`` therefore makes the kernel stream wait for the
    stream that was current when the ticket was collected -- or, if the consumer works on other
    streams, for the event of an explicit ``release(ticket)`` -- before it overwrites the slot."""

    def __init__(self, transform: GpuClipTransform, batch: int, clip_shape: Sequence[int],
                 device: Optional[torch.device] = None, depth: int = 3, host_out: bool = False,
                 upload_selected_frames: bool = False, read_in_place: bool = False, max_ctas: int = 4):
        if depth < 2:
            raise ValueError("depth must be at least 2 to overlap anything")
        self.t = transform
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.batch = batch
        self.clip_shape = tuple(clip_shape)
        self.host_out = host_out
        self.in_place = bool(read_in_place)
        self.max_ctas = int(max_ctas)
        self.selected = bool(upload_selected_frames) and not self.in_place
        self.slots = [_Slot(transform, batch, clip_shape, self.device, host_out, self.selected, self.in_place)
                      for _ in range(depth)]
        idx = temporal_indices(self.clip_shape[1], transform.num_frames)
        self._idx = (ctypes.c_int32 * len(idx))(*idx)
        self._lib = _lib.load_library()
        self.s_in = torch.cuda.Stream(self.device)
        self.s_run = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        self._n = 0
        self._reading: list = []  # in place: (kernel-done event, pinned batch) of kernels that may still be reading
        # bytes that cross PCIe per batch: the copy's size, or (in place) what the last submitted batch's
        # kernel reads -- full-width source rows of the selected frames that its crop windows touch
        self.h2d_bytes = 0 if self.in_place else self.slots[0].dev_in.numel()
        self.d2h_bytes = self.slots[0].flat_out.numel() if host_out else 0  # one copy of [slow | fast]

    def submit(self, host_clips: torch.Tensor, params: Sequence[ClipParams]) -> int:
        if tuple(host_clips.shape) != (self.batch,) + self.clip_shape or host_clips.dtype != torch.uint8:
            raise ValueError(f"expected uint8 clips of shape {(self.batch,) + self.clip_shape}")
        slot = self.slots[self._n % len(self.slots)]
        if self.in_place:
            return self._submit_in_place(slot, host_clips, params)
        slot.host_clips = host_clips  # stays referenced while the copy queued below may still read it
        with torch.cuda.stream(self.s_in):
            if slot.used:
                self.s_in.wait_event(slot.kernel_done)  # the kernel that read this slot has finished
            if self.selected:
                if not host_clips.is_contiguous():
                    raise ValueError("upload_selected_frames needs a contiguous host batch")
                c, t_in, h, w = self.clip_shape
                n = self.t.num_frames
                src, dst = host_clips.data_ptr(), slot.dev_in.data_ptr()
                with torch.cuda.device(self.device):
                    for b in range(self.batch):
                        rc = self._lib.pva_upload_selected_frames(src + b * c * t_in * h * w, dst + b * c * n * h * w,
                                                                  c, t_in, h, w, self._idx, n, self.s_in.cuda_stream)
                        _lib.raise_for_status(rc)
            else:
                slot.dev_in.copy_(host_clips, non_blocking=True)
            slot.in_ready.record(self.s_in)
        with torch.cuda.stream(self.s_run):
            self.s_run.wait_event(slot.in_ready)
            self._wait_for_consumer(slot)
            # the clips arrived through a copy on another stream (event-ordered): nothing the previous
            # kernel on s_run did can have produced them
            # consecutive kernels on s_run write different slots (depth >= 2): outputs are disjoint
            prepared = self.t.prepare_batch(slot.dev_in, params, out=slot.raw_out, clips_ready=True,
                                            outputs_disjoint=True)
            prepared.launch_raw(self.s_run.cuda_stream)
            slot.kernel_done.record(self.s_run)
        if self.host_out:
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(slot.kernel_done)
                slot.host_flat.copy_(slot.flat_out, non_blocking=True)  # [slow | fast] in one copy
                slot.out_ready.record(self.s_out)
        slot.used = True
        ticket = self._n
        self._n += 1
        return ticket

    def _submit_in_place(self, slot: _Slot, host_clips: torch.Tensor, params: Sequence[ClipParams]) -> int:
        if not host_clips.is_pinned():
            raise ValueError("read_in_place needs a pinned host batch (pin_memory())")
        c, _t_in, h, w = self.clip_shape
        rows = 0
        for p in params:
            fh, _fw = _lib.source_footprint(h, w, p.oh, p.ow, p.top, p.left, self.t.crop_size)
            rows += fh
        self.h2d_bytes = c * self.t.num_frames * rows * w
        slot.host_clips = host_clips
        with torch.cuda.stream(self.s_run):
            self._wait_for_consumer(slot)
            prepared = self.t.prepare_batch(host_clips, params, out=slot.raw_out, clips_ready=True,
                                            max_ctas=self.max_ctas, device=self.device, outputs_disjoint=True)
            prepared.launch_raw(self.s_run.cuda_stream)
            slot.kernel_done = torch.cuda.Event()
            slot.kernel_done.record(self.s_run)
            # the pinned batch stays referenced until the kernel that reads it has finished
            self._reading = [(e, v) for e, v in self._reading if not e.query()]
            self._reading.append((slot.kernel_done, host_clips))
        if self.host_out:
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(slot.kernel_done)
                slot.host_flat.copy_(slot.flat_out, non_blocking=True)  # [slow | fast] in one copy
                slot.out_ready.record(self.s_out)
        slot.used = True
        ticket = self._n
        self._n += 1
        return ticket

    def _wait_for_consumer(self, slot: _Slot) -> None:
        """Called on s_run before a slot's outputs are overwritten."""
        if not slot.used:
            return
        if self.host_out:
            self.s_run.wait_event(slot.out_ready)  # the previous result has left the slot
        elif slot.released is not None:
            self.s_run.wait_event(slot.released)  # whoever took result() has finished reading raw_out
            slot.released = None
        elif slot.consumer_stream is not None:
            self.s_run.wait_stream(slot.consumer_stream)  # everything the consumer has enqueued there so far
            slot.consumer_stream = None

    def result(self, ticket: int) -> List[torch.Tensor]:
        """Block until batch ``ticket`` is complete.  Returns the pinned host tensors (``host_out=True``,
        raw ``N x T x S x S x C`` order) or the device tensors as ``[slow, fast]`` logical views.

        Device views stay valid until the slot is reused (``depth`` submits later).  Work the caller
        enqueues on the stream that is current NOW is waited for automatically before that reuse; a
        consumer on another stream calls ``release(ticket)`` when it has enqueued its last read."""
        if ticket < self._n - len(self.slots) or ticket >= self._n:
            raise ValueError("ticket is no longer (or not yet) resident in a slot")
        slot = self.slots[ticket % len(self.slots)]
        if self.host_out:
            slot.out_ready.synchronize()
            slot.host_clips = None
            return slot.host_out
        slot.kernel_done.synchronize()
        slot.host_clips = None
        slot.consumer_stream = torch.cuda.current_stream(self.device)
        return [self.t._logical(r) for r in slot.raw_out]

    def release(self, ticket: int, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Declare that everything enqueued so far on ``stream`` (default: the current stream) is all
        that will ever read the device tensors of ``ticket``."""
        slot = self.slots[ticket % len(self.slots)]
        ev = torch.cuda.Event()
        ev.record(stream if stream is not None else torch.cuda.current_stream(self.device))
        slot.released = ev
        slot.consumer_stream = None

    def drain(self) -> None:
        for s in (self.s_in, self.s_run, self.s_out):
            s.synchronize()
