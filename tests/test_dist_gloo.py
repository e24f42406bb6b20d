"""World-size-2 run of the host-side multi-GPU logic on CPU (gloo): the path shards by clip,
every rank draws and prepares only its own batches, and nothing but the benchmark's timing
reduction crosses ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pva_b200


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, num_batches, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert pva_b200.rank_and_world() == (rank, world, rank)
        mine = pva_b200.batches_for_rank(num_batches, rank, world)
        t = pva_b200.make_transform(training=True)
        torch.manual_seed(pva_b200.rank_seed(42, rank))
        sp = pva_b200.SampleParams(t)
        rows = []
        for b in mine:
            g = torch.Generator().manual_seed(1000 + b)
            samples = [sp({"video": torch.randint(0, 256, (3, 8, 64, 85), dtype=torch.uint8, generator=g), "label": b})
                       for _ in range(2)]
            batch = pva_b200.collate_raw_clips(samples)
            assert batch["video"].dtype == torch.uint8 and batch["video"].shape[0] == 2
            rows.append(batch["pva_params"])
        # what bench.py does across ranks: barrier + MAX of the per-rank time
        elapsed = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.barrier()
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
        gathered = [None] * world
        dist.all_gather_object(gathered, (mine, torch.cat(rows).tolist()))
        if rank == 0:
            out.put((float(elapsed.item()), gathered))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_clips_without_data_collectives():
    world, num_batches = 2, 7
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, num_batches, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    elapsed, gathered = out.get()
    assert elapsed == 11.0  # max over ranks
    all_batches = sorted(b for mine, _ in gathered for b in mine)
    assert all_batches == list(range(6))  # every batch of the full rounds exactly once, remainder dropped
    assert [len(m) for m, _ in gathered] == [3, 3]
    # different ranks draw from different RNG streams
    assert gathered[0][1] != gathered[1][1]


def test_shard_helper_edges():
    assert pva_b200.batches_for_rank(8, 1, 4) == [1, 5]
    assert pva_b200.batches_for_rank(3, 0, 4) == []
    with pytest.raises(ValueError):
        pva_b200.batches_for_rank(8, 4, 4)
