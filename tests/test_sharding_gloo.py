"""Host side of the multi-GPU path on CPU: world-size-2 `gloo` process group.  Each rank takes its shard of a
small W5 batch (by GLOBAL voice index), renders its voices with the oracle (the checker stands in for the
device here), sums them into a local stereo mix and the mixes are reduced onto rank 0 -- the same sharding,
parameter generation and collective call (`dusp_b200.mixdown`) the GPU bench uses with NCCL."""
import os
import socket

import numpy as np
import pytest

from dusp_b200 import mixdown

N_VOICES, SECONDS = 6, 0.02


def test_shard_covers_every_voice_once():
    for n in (0, 1, 7, 100000):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                v0, k = mixdown.shard(n, r, world)
                got += list(range(v0, v0 + k))
            assert got == list(range(n))
            sizes = [mixdown.shard(n, r, world)[1] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        mixdown.shard(10, 2, 2)


def _local_mix(indices):
    from dusp_b200 import workloads as W
    from dusp_b200.render import circuit_blob, n_samples_for
    from oracle import oracle
    n = n_samples_for(SECONDS)
    mix = np.zeros((2, n), dtype=np.float64)
    for i in indices:
        chans = oracle.render(circuit_blob(W.w5_voice_circuit(i))[0], SECONDS)
        for c in range(2):
            mix[c] += chans[c % len(chans)]  # mono voices feed both channels (Sum's modulo broadcast, Sum.js:33-44)
    return mix.astype(np.float32)


def _worker(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    v0, n = mixdown.shard(N_VOICES, rank, world)
    t = torch.from_numpy(_local_mix(range(v0, v0 + n)))
    work = mixdown.reduce_mix(t, dst=0, async_op=True)
    work.wait()
    # transport selection of the renderer: the one-sided NVLink path is for CUDA ranks only; CPU ranks reduce through
    # torch.distributed, and an explicit request is honoured
    assert mixdown.MixdownRenderer(None, 256, "cpu").collective == "nccl"
    assert mixdown.MixdownRenderer(None, 256, "cpu", collective="nccl").collective == "nccl"
    if rank == 0:
        np.save(out_path, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_mixdown_matches_single_process(oracle, tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "mix.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    want = _local_mix(range(N_VOICES))
    assert got.shape == want.shape
    assert np.max(np.abs(got.astype(np.float64) - want)) <= 1e-5 * max(1.0, float(np.max(np.abs(want))))


def test_reduce_without_process_group_is_a_noop():
    import torch
    t = torch.ones(2, 4)
    assert mixdown.reduce_mix(t) is None and float(t.sum()) == 8.0


def test_renderer_without_process_group_has_no_collective():
    r = mixdown.MixdownRenderer(None, 256, "cpu")
    assert r.collective == "none" and r.world == 1 and r.rank == 0 and r.acc is None and r.ticket is None
    r.close()
