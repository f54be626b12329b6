"""world_size-2 gloo run of the multi-GPU host logic (no GPU): voice shards, global-voice Philox keys, gathers, timing."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, T, hop, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ddsp_pytorch_b200.sharding import gather_voices, max_over_ranks, voice_shard
        from oracle import ddsp_oracle as O
        from oracle.philox import noise_uniform
        lo, hi = voice_shard(B, world, rank)
        noise = torch.from_numpy(noise_uniform(hi - lo, T, hop, seed=11, offset=3, voice_offset=lo))
        f0, c, a, H = O.synthetic_controls(hi - lo, T, 12, 9, 16000, hop, seed=5, voice_offset=lo)
        y = O.oscillator_bank(f0, c, a, 12, 16000, hop) + O.filtered_noise(H, hop, noise)
        full = gather_voices(y, B)
        slow = max_over_ranks(1.0 + rank)
        if rank == 0:
            ret["y"] = full.numpy()
            ret["slow"] = slow
            ret["ranges"] = [voice_shard(B, world, r) for r in range(world)]
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_reproduces_the_unsharded_result():
    B, T, hop = 5, 3, 64                                    # odd voice count: shards of 3 and 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), B, T, hop, ret), nprocs=2, join=True)
    from oracle import ddsp_oracle as O
    from oracle.philox import noise_uniform
    noise = torch.from_numpy(noise_uniform(B, T, hop, seed=11, offset=3, voice_offset=0))
    f0, c, a, H = O.synthetic_controls(B, T, 12, 9, 16000, hop, seed=5)
    want = O.oscillator_bank(f0, c, a, 12, 16000, hop) + O.filtered_noise(H, hop, noise)
    assert ret["ranges"] == [(0, 3), (3, 5)]
    assert ret["slow"] == 2.0
    np.testing.assert_array_equal(ret["y"], want.numpy())


def test_voice_shard_covers_everything():
    from ddsp_pytorch_b200.sharding import voice_shard
    for B in (1, 7, 8, 1024):
        for world in (1, 2, 3, 8):
            ranges = [voice_shard(B, world, r) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == B
            assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in ranges) - min(h - l for l, h in ranges) <= 1


def test_package_input_generator_draws_the_oracle_voices():
    """bench.py generates its inputs with ddsp_pytorch_b200.synthetic (the product arm never imports oracle/); the values
    are the oracle generator's, for any sharding."""
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    from oracle import ddsp_oracle as O
    want = O.synthetic_controls(3, 7, 5, 4, 16000, 64, seed=9, voice_offset=2)
    got = synthetic_controls(3, 7, 5, 4, 16000, 64, seed=9, voice_offset=2)
    assert all(torch.equal(a, b) for a, b in zip(got, want))


def test_work_balanced_shards():
    """balanced_voice_shards: contiguous cover, nobody empty, bottleneck no worse than the equal-count split and within one
    voice of the mean; voice_work counts what the kernels evaluate (harmonics at or below sample_rate//2 in either frame of
    a hop segment) -- checked against the brute-force mask of harmonic_oscillator.py:31."""
    from ddsp_pytorch_b200.sharding import balanced_voice_shards, voice_shard, voice_work
    from ddsp_pytorch_b200.synthetic import synthetic_f0
    B, T, K, sr, hop = 96, 20, 64, 48000, 512
    f0 = synthetic_f0(B, T, sr, hop, seed=1238)
    w = voice_work(f0, K, sr, hop, per_sample_cost=0.0)
    k = torch.arange(1, K + 1, dtype=torch.float32).view(1, 1, -1)
    kact = ((k * f0) <= float(sr // 2)).sum(-1).double()
    left, right = torch.cat([kact[:, :1], kact], 1), torch.cat([kact, kact[:, -1:]], 1)
    n = torch.full((T + 1,), float(hop), dtype=torch.float64)
    n[0] = n[-1] = hop / 2
    assert torch.equal(w, (torch.maximum(left, right) * n).sum(1))
    for world in (1, 2, 3, 8):
        r = balanced_voice_shards(w.tolist(), world)
        assert len(r) == world and r[0][0] == 0 and r[-1][1] == B
        assert all(r[i][1] == r[i + 1][0] for i in range(world - 1)) and all(hi > lo for lo, hi in r)
        loads = [float(w[lo:hi].sum()) for lo, hi in r]
        equal = [float(w[slice(*voice_shard(B, world, i))].sum()) for i in range(world)]
        assert max(loads) <= max(equal) + 1e-9
        assert max(loads) <= sum(loads) / world + float(w.max())
    assert balanced_voice_shards([1.0, 1.0], 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]          # fewer voices than ranks: empty tails
    assert balanced_voice_shards([5.0, 1.0, 1.0, 1.0, 1.0, 1.0], 2) == [(0, 1), (1, 6)]
