"""N > 1 logic on CPU: world_size 2 over gloo.

What the multi-GPU path relies on: (1) the counter-based read generator is seamless, so
rank r generating reads [r*n, (r+1)*n) yields exactly the reads a single rank would generate for that range;
(2) the per-rank counter blocks are plain integer sums, so ONE sum-allreduce of the raw (not yet finalized)
block followed by the finalize step equals the single-rank block; (3) the host's round-robin deal of batches
(b mod N) covers every batch exactly once.  The oracle stands in for the device here (this file runs without
a GPU); tests/test_gpu_cli.py repeats (2) on hardware when two GPUs are visible."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORLD = 2
N_PER_RANK = 6000
READ_LEN = 100
FILTER = dict(min_read_length=50, min_read_quality=20, max_N=5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    import oracle_binding as ob

    q = g.load_package().hpgq
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = q.synth_cfg(READ_LEN)
        par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS, filter=FILTER)
        lay = ob.layout(q.Layout, 128, kmers=1)
        # rank-local shard of the read index space, two batches dealt like the host does (b mod N)
        seq, qual, idx = q.synth_host(cfg, rank * N_PER_RANK, N_PER_RANK)
        ctr, mask, _ = ob.process_batch(par, lay, seq, qual, idx, finalize=False)
        t = torch.from_numpy(ctr.view(np.int64).copy())
        dist.all_reduce(t)  # the one collective of the path
        merged = t.numpy().view(np.uint64).copy()
        ob.lib().oracle_finalize(C.byref(lay), merged.ctypes.data_as(C.c_void_p))
        # masks are gathered in input order (rank order = read order)
        masks = [torch.zeros(N_PER_RANK, dtype=torch.uint8) for _ in range(world)]
        dist.all_gather(masks, torch.from_numpy(mask.copy()))
        if rank == 0:
            np.save(os.path.join(out_dir, "merged.npy"), merged)
            np.save(os.path.join(out_dir, "mask.npy"), torch.cat(masks).numpy())
    finally:
        dist.destroy_process_group()


def test_two_rank_shards_sum_to_single_rank_block(q, ob, tmp_path):
    import torch.multiprocessing as mp

    port = _free_port()
    mp.spawn(_rank_main, args=(WORLD, port, str(tmp_path)), nprocs=WORLD, join=True)
    merged = np.load(tmp_path / "merged.npy")
    gmask = np.load(tmp_path / "mask.npy")

    cfg = q.synth_cfg(READ_LEN)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS, filter=FILTER)
    lay = ob.layout(q.Layout, 128, kmers=1)
    seq, qual, idx = q.synth_host(cfg, 0, WORLD * N_PER_RANK)
    exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx, finalize=True)
    assert np.array_equal(gmask, emask)
    skip = {q.S["BATCHES"], q.S["SLOW_READS"], q.S["CG_FIXED_READS"], q.S["CG_MOVED_WORDS"]}
    bad = [int(i) for i in np.nonzero(merged != exp)[0] if int(i) not in skip]
    assert not bad, f"merged block differs at words {bad[:10]}"
    assert int(merged[q.S["NUM_PASSED"]] + merged[q.S["NUM_FAILED"]]) == WORLD * N_PER_RANK
    assert int(merged[q.S["NUM_READS"]]) == int(emask.sum())


def test_sharded_generation_is_seamless(q):
    cfg = q.synth_cfg(READ_LEN)
    whole = q.synth_host(cfg, 0, 2000)
    a = q.synth_host(cfg, 0, 1000)
    b = q.synth_host(cfg, 1000, 1000)
    assert np.array_equal(whole[0], np.concatenate([a[0], b[0]]))
    assert np.array_equal(whole[1], np.concatenate([a[1], b[1]]))
    assert np.array_equal(whole[2][1001:], b[2][1:] + a[2][-1])


@pytest.mark.parametrize("n_batches,n_gpus", [(1, 2), (7, 2), (16, 8), (5, 8)])
def test_round_robin_deal_covers_every_batch_once(n_batches, n_gpus):
    # host/main.c deals batch b to GPU b mod N and slot (b / N) mod 2; outputs are written in batch order
    owner = [b % n_gpus for b in range(n_batches)]
    per_gpu = [[b for b in range(n_batches) if owner[b] == g] for g in range(n_gpus)]
    assert sorted(sum(per_gpu, [])) == list(range(n_batches))
    for lst in per_gpu:
        slots = [(b // n_gpus) % 2 for b in lst]
        assert all(slots[i] != slots[i + 1] for i in range(len(slots) - 1))
