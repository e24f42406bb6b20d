"""Overlapped staging (host clips -> device -> host) returns the same bytes as a plain call."""
import pytest
import torch

import pva_b200

from _util import synth_clip

pytestmark = pytest.mark.gpu


def test_upload_selected_frames_is_the_temporal_gather(cuda_device):
    """The strided DMA upload lands exactly index_select(clip, 1, linspace(0, T-1, n).long())."""
    import ctypes

    from pva_b200 import _lib

    lib = _lib.load_library()
    for t_in, n in [(64, 32), (80, 16), (16, 16), (5, 8), (33, 7)]:
        clip = synth_clip((3, t_in, 40, 52), t_in * 100 + n).pin_memory()
        idx = pva_b200.temporal_indices(t_in, n)
        dst = torch.zeros((3, n, 40, 52), dtype=torch.uint8, device=cuda_device)
        arr = (ctypes.c_int32 * n)(*idx)
        rc = lib.pva_upload_selected_frames(clip.data_ptr(), dst.data_ptr(), 3, t_in, 40, 52, arr, n,
                                            torch.cuda.current_stream().cuda_stream)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert torch.equal(dst.cpu(), clip.index_select(1, torch.tensor(idx)))
    arr = (ctypes.c_int32 * 2)(0, 9)
    assert lib.pva_upload_selected_frames(clip.data_ptr(), dst.data_ptr(), 3, 5, 40, 52, arr, 2, None) == -1


@pytest.mark.parametrize("upload", ["all", "selected", "inplace"])
@pytest.mark.parametrize("host_out", [False, True])
def test_staged_pipeline_matches_direct_calls(host_out, upload, cuda_device):
    kw = dict(min_short_side_scale=128, max_short_side_scale=160, crop_size=128, num_frames=8, training=True)
    t = pva_b200.make_transform(**kw)
    shape = (3, 16, 128, 172)
    staged = pva_b200.StagedPreprocess(t, 4, shape, cuda_device, depth=2, host_out=host_out,
                                       upload_selected_frames=(upload == "selected"),
                                       read_in_place=(upload == "inplace"), max_ctas=5)
    torch.manual_seed(3)
    batches = []
    for b in range(5):  # more batches than slots: slots are reused
        clips = torch.stack([synth_clip(shape, 100 * b + i) for i in range(4)]).pin_memory()
        params = [t.sample_params(128, 172) for _ in range(4)]
        batches.append((clips, params))
    tickets = []
    for i, (clips, params) in enumerate(batches):
        tickets.append(staged.submit(clips, params))
        if i >= 1:
            check(staged, t, tickets[i - 1], batches[i - 1], host_out)
    check(staged, t, tickets[-1], batches[-1], host_out)
    with pytest.raises(ValueError):
        staged.result(tickets[0])  # long gone from its slot
    staged.drain()


def check(staged, t, ticket, batch, host_out):
    clips, params = batch
    got = staged.result(ticket)
    want = t.run_batch(clips.cuda(), params)
    torch.cuda.synchronize()
    for g, w in zip(got, want):
        if host_out:
            g = g.permute(0, 4, 1, 2, 3)  # raw N x T x S x S x C -> logical
        assert torch.equal(g.cpu(), w.cpu())


def test_gpu_preprocess_with_selected_frame_upload(cuda_device):
    kw = dict(min_short_side_scale=128, max_short_side_scale=160, crop_size=128, num_frames=8, training=True)
    t = pva_b200.make_transform(**kw)
    sp = pva_b200.SampleParams(t)
    torch.manual_seed(4)
    samples = [sp({"video": synth_clip((3, 16, 128, 170), 60 + i), "label": i}) for i in range(3)]
    batch = pva_b200.collate_raw_clips(samples)
    batch["video"] = batch["video"].pin_memory()
    a = pva_b200.GpuPreprocess(t, cuda_device, upload_selected_frames=True)(dict(batch))["video"]
    b = pva_b200.GpuPreprocess(t, cuda_device)(dict(batch))["video"]
    torch.cuda.synchronize()
    assert all(torch.equal(x, y) for x, y in zip(a, b))


def test_pinned_host_clips_are_read_in_place(cuda_device):
    """A pinned host clip is a legal PvaClipDesc.src: same bits as from a device copy, on both kernels,
    with a capped grid, ragged batch; pageable host memory is refused before any launch."""
    from pva_b200 import _lib

    kw = dict(num_frames=8, slowfast_alpha=4, crop_size=96, min_short_side_scale=96, max_short_side_scale=128,
              training=True)
    torch.manual_seed(21)
    host = [synth_clip((3, 12 + i, 100 + 4 * i, 132), 300 + i).pin_memory() for i in range(3)]
    for kernel in ("tiled", "generic"):
        t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
        params = [t.sample_params(v.shape[2], v.shape[3]) for v in host]
        want = [o.clone() for o in t.run_batch([v.cuda() for v in host], params)]
        got = t.prepare_batch(host, params, max_ctas=3, device=cuda_device).launch()
        torch.cuda.synchronize()
        assert all(torch.equal(g, w) for g, w in zip(got, want))
        if kernel == "tiled":
            assert _lib.last_launch_info()["grid"] == 3
    with pytest.raises(ValueError, match="page-locked"):
        t.prepare_batch([synth_clip((3, 12, 100, 132), 1)], params[:1])
    # GpuPreprocess: a pinned batch is not copied, a pageable one is -- same result
    t = pva_b200.make_transform(**kw)
    sp = pva_b200.SampleParams(t)
    samples = [sp({"video": synth_clip((3, 16, 128, 172), 70 + i), "label": i}) for i in range(3)]
    batch = pva_b200.collate_raw_clips(samples)
    pre = pva_b200.GpuPreprocess(t, cuda_device, read_in_place=True, max_ctas=4)
    a = pre(dict(batch, video=batch["video"].pin_memory()))["video"]
    assert _lib.last_launch_info()["grid"] == 4
    b = pre(dict(batch))["video"]  # pageable: copy path
    c = pva_b200.GpuPreprocess(t, cuda_device)(dict(batch))["video"]
    torch.cuda.synchronize()
    assert all(torch.equal(x, y) and torch.equal(x, z) for x, y, z in zip(a, b, c))


def test_back_to_back_launches_into_the_same_outputs_keep_stream_order(cuda_device):
    """PVA_FLAG_CLIPS_READY lets a launch start while the previous one drains, but it must not
    store before that one has completed: the last launch into a reused output buffer wins."""
    kw = dict(num_frames=32, slowfast_alpha=4, training=True)
    t = pva_b200.make_transform(**kw)
    torch.manual_seed(12)
    sets = []
    for s in range(3):
        clips = torch.stack([synth_clip((3, 64, 256, 340), 900 + 10 * s + i) for i in range(4)]).cuda()
        params = [t.sample_params(256, 340) for _ in range(4)]
        sets.append((clips, params))
    want = [[o.clone() for o in t.run_batch(c, p)] for c, p in sets]
    shared = t.alloc_outputs(4, cuda_device)
    prepared = [t.prepare_batch(c, p, out=shared, clips_ready=True) for c, p in sets]
    torch.cuda.synchronize()
    for rounds in range(20):
        last = None
        for i in range(7):  # no synchronisation in between: 7 launches race through one output buffer
            last = (rounds + i) % 3
            prepared[last].launch()
        torch.cuda.synchronize()
        got = prepared[last].outputs
        assert all(torch.equal(g, w) for g, w in zip(got, want[last])), (rounds, last)


def test_disjoint_outputs_overlap_fully_and_stay_in_stream_order(cuda_device):
    """PVA_FLAG_OUTPUTS_DISJOINT: launches into a ring of distinct output buffers do not wait for one
    another before storing.  Results must be those of isolated launches, and whatever follows on the
    stream (here: a device-to-host copy of the FIRST launch's output, enqueued after a chain of seven
    launches) must see every launch complete -- a big launch followed by small ones is the case where
    a later, shorter kernel could otherwise finish first."""
    kw = dict(num_frames=32, slowfast_alpha=4, training=True)
    t = pva_b200.make_transform(**kw)
    torch.manual_seed(13)
    big = torch.stack([synth_clip((3, 64, 256, 340), 300 + i) for i in range(8)]).cuda()
    big_p = [t.sample_params(256, 340) for _ in range(8)]
    small = [torch.stack([synth_clip((3, 64, 256, 340), 400 + 10 * j + i) for i in range(1)]).cuda() for j in range(3)]
    small_p = [[t.sample_params(256, 340)] for _ in range(3)]
    want_big = [o.clone() for o in t.run_batch(big, big_p)]
    want_small = [[o.clone() for o in t.run_batch(c, p)] for c, p in zip(small, small_p)]
    ring_big = [t.alloc_outputs(8, cuda_device) for _ in range(2)]
    ring_small = [t.alloc_outputs(1, cuda_device) for _ in range(3)]
    with pytest.raises(ValueError, match="caller-owned"):
        t.prepare_batch(big, big_p, clips_ready=True, outputs_disjoint=True)
    pb = [t.prepare_batch(big, big_p, out=o, clips_ready=True, outputs_disjoint=True) for o in ring_big]
    ps = [t.prepare_batch(c, p, out=o, clips_ready=True, outputs_disjoint=True)
          for c, p, o in zip(small, small_p, ring_small)]
    host = [torch.empty_like(o, device="cpu").pin_memory() for o in ring_big[0]]
    torch.cuda.synchronize()
    for rounds in range(10):
        for o in ring_big[rounds % 2] + [x for r in ring_small for x in r]:
            o.zero_()
        torch.cuda.synchronize()
        first = pb[rounds % 2]
        first.launch()
        for j in range(6):
            ps[j % 3].launch()
        for h, o in zip(host, first.raw_out):
            h.copy_(o, non_blocking=True)  # stream-ordered after all seven launches
        torch.cuda.synchronize()
        for h, w in zip(host, want_big):
            assert torch.equal(h.permute(0, 4, 1, 2, 3), w.cpu()), rounds
        for j in range(3):
            assert all(torch.equal(g, w) for g, w in zip(ps[j].outputs, want_small[j])), (rounds, j)


def test_device_results_are_not_overwritten_under_a_running_consumer(cuda_device):
    """host_out=False hands out views of the slot's device buffers.  A consumer that is still reading
    them on the current stream when the slot comes round again must be waited for: here the consumer
    is a long chain of kernels enqueued right after result(), and the slot is resubmitted at once."""
    kw = dict(min_short_side_scale=128, max_short_side_scale=160, crop_size=128, num_frames=8, training=True)
    t = pva_b200.make_transform(**kw)
    shape = (3, 16, 128, 172)
    staged = pva_b200.StagedPreprocess(t, 4, shape, cuda_device, depth=2, host_out=False)
    torch.manual_seed(8)
    batches = []
    for b in range(4):
        clips = torch.stack([synth_clip(shape, 500 * b + i) for i in range(4)]).pin_memory()
        batches.append((clips, [t.sample_params(128, 172) for _ in range(4)]))
    want = [[o.clone() for o in t.run_batch(c.cuda(), p)] for c, p in batches]
    torch.cuda.synchronize()
    spin = torch.empty(64 << 20, dtype=torch.float32, device=cuda_device)
    t0 = staged.submit(*batches[0])
    slow, fast = staged.result(t0)
    for _ in range(30):  # ~ms of queued work on the current stream ahead of the consumer's read
        spin.add_(1.0)
    late_copy = [slow.clone(), fast.clone()]  # the consumer's read, queued behind the spinning
    staged.submit(*batches[1])
    staged.submit(*batches[2])  # depth 2: this one reuses ticket 0's slot
    torch.cuda.synchronize()
    assert torch.equal(late_copy[0], want[0][0]) and torch.equal(late_copy[1], want[0][1])
    # an explicit release on a side stream is honoured the same way
    side = torch.cuda.Stream(cuda_device)
    t3 = staged.submit(*batches[3])  # reuses ticket 1's slot
    slow3, fast3 = staged.result(t3)
    with torch.cuda.stream(side):
        for _ in range(30):
            spin.add_(1.0)
        got = fast3.clone()
        staged.release(t3, side)
    staged.submit(*batches[0])
    staged.submit(*batches[1])  # reuses ticket 3's slot: must wait for `side`
    torch.cuda.synchronize()
    assert torch.equal(got, want[3][1])
    staged.drain()


def test_strided_pinned_clip_is_rejected_not_copied(cuda_device):
    """A non-contiguous pinned clip cannot be read in place, and making it contiguous would silently
    produce a pageable copy the kernel cannot reach: ValueError instead."""
    t = pva_b200.make_transform(min_short_side_scale=64, crop_size=64, num_frames=4, slowfast_alpha=2)
    base = synth_clip((3, 8, 64, 160), 1).pin_memory()
    view = base[:, :, :, :80]
    assert view.is_pinned() and not view.is_contiguous()
    with pytest.raises(ValueError, match="contiguous"):
        t.prepare_batch([view], [t.sample_params(64, 80)], device=cuda_device)
