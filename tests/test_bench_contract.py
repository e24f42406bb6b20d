"""bench.py's line contract, checked on the CPU (no timing here): both arms describe the workload with
the same `config` object, the rotation over buffer sets exceeds L2, and the clock sampler's child process
protocol works without a GPU (it reports no samples, not an exception)."""
import json

import bench


def test_both_arms_print_the_same_config():
    for name, wl in bench.WORKLOADS.items():
        for world in (1, 2, 8):
            a = bench.workload_config(name, wl, world)
            b = bench.workload_config(name, dict(wl), world)
            assert a == b and json.dumps(a)  # serialisable, deterministic
            assert a["global_batch"] == wl["batch"] * world and a["workload"] == name
            assert "model" not in a and "L2" in a["l2"]


def test_rotation_working_set_exceeds_l2():
    for wl in bench.WORKLOADS.values():
        n_sets, in_bytes, out_bytes = bench.rotation_sets(wl)
        assert n_sets >= 4 and n_sets * (in_bytes + out_bytes) > 2 * bench.L2_BYTES
    # headline: 8 clips of 3x64x256x340 in, fp16 slow (8 frames) + fast (32 frames) at 256x256 out
    _, i, o = bench.rotation_sets(bench.WORKLOADS["slowfast_train_b8"])
    assert i == 8 * 3 * 64 * 256 * 340 and o == 8 * 3 * 40 * 256 * 256 * 2


def test_workloads_are_the_baseline_configs():
    w = bench.WORKLOADS
    assert w["slowfast_train_b8"]["kwargs"] == dict(num_frames=32, slowfast_alpha=4, training=True)
    assert w["slowfast_val_b8"]["kwargs"]["training"] is False
    assert bench.workload_geometry(w["single_pathway_b64"]) == (224, 16, 0)
    assert bench.workload_geometry(w["slowfast_train_b8"]) == (256, 32, 8)


def test_clock_sampler_degrades_without_a_gpu():
    s = bench.ClockSampler(0, period=0.01)
    s.mark_start()
    s.mark_end()
    out = s.finish()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons", "samples"}
    off = bench.ClockSampler(0, enabled=False)
    off.mark_start(); off.mark_end()
    assert off.finish()["samples"] == 0


def test_staged_outputs_are_views_of_one_buffer():
    """StagedPreprocess returns [slow, fast] as views of ONE contiguous buffer (one D2H copy per batch):
    256-byte aligned offsets, no overlap, both memory orders (host side only: no GPU needed)."""
    import torch

    import pva_b200
    from pva_b200.staging import _flat_outputs

    for channels_last in (True, False):
        t = pva_b200.make_transform(training=True, crop_size=100, num_frames=8, slowfast_alpha=4,
                                    channels_last=channels_last)
        flat, (slow, fast) = _flat_outputs(t, 3, "cpu", pin=False)
        want_slow = (3, 2, 100, 100, 3) if channels_last else (3, 3, 2, 100, 100)
        want_fast = (3, 8, 100, 100, 3) if channels_last else (3, 3, 8, 100, 100)
        assert tuple(slow.shape) == want_slow and tuple(fast.shape) == want_fast
        assert slow.dtype == fast.dtype == torch.float16 and slow.is_contiguous() and fast.is_contiguous()
        o_slow, o_fast = slow.data_ptr() - flat.data_ptr(), fast.data_ptr() - flat.data_ptr()
        assert o_slow == 0 and o_fast % 256 == 0 and o_fast >= slow.numel() * 2
        assert o_fast + fast.numel() * 2 <= flat.numel()
    single = pva_b200.make_transform(is_slowfast=False, num_frames=4, crop_size=64)
    flat, views = _flat_outputs(single, 2, "cpu", pin=False)
    assert len(views) == 1 and views[0].data_ptr() == flat.data_ptr()
