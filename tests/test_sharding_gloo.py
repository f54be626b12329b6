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
