"""Host-side logic of the two multi-GPU modes, on CPU with world_size-2 gloo process groups:
coordinate sharding (shard ranges + the one allreduce of [grads | loss]) and the clip-parallel partition."""
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


def test_shard_ranges_cover_exactly_and_align():
    from inr_for_audio_b200.engine import shard_range
    for T in (1, 127, 128, 129, 88200, 264600, 1323000, 2646000):
        for world in (1, 2, 3, 4, 8):
            covered, prev_end = 0, 0
            for r in range(world):
                i0, n, per = shard_range(T, r, world)
                assert per % 128 == 0
                assert i0 == min(prev_end, T) or n == 0
                assert n == 0 or i0 % 128 == 0
                covered += n
                prev_end = i0 + n
            assert covered == T


def test_clip_partition():
    from inr_for_audio_b200 import ClipBatchFitter
    for K in (1, 7, 64, 1024):
        for world in (1, 2, 3, 8):
            spans = [ClipBatchFitter.clip_range(K, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == K
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert ClipBatchFitter.clip_range(1024, 7, 8) == (896, 1024)


def _worker(rank, world, port, T, P, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from inr_for_audio_b200.engine import allreduce_step_state, shard_range
    i0, n, _ = shard_range(T, rank, world)
    # a stand-in for the backward kernels: per-sample "gradient" rows and squared errors, summed over this shard
    rng = np.random.default_rng(0)
    rows = rng.standard_normal((T, P))
    err = rng.standard_normal(T)
    grads = torch.from_numpy(rows[i0:i0 + n].sum(axis=0))
    loss_acc = torch.zeros(4, dtype=torch.float64)
    loss_acc[0] = float((err[i0:i0 + n] ** 2).sum())
    loss_acc[1] = 0.125                      # whole-output (spectral) loss: evaluated identically on every rank
    loss_acc[2] = 7.0                        # scratch slots are rank-local and must not be reduced
    allreduce_step_state(grads, loss_acc, dist.group.WORLD, world, True)
    # the same exchange as ONE collective over [grads | loss hi, lo] in fp32 (what the device path issues)
    from inr_for_audio_b200.engine import allreduce_fused
    gbuf = torch.zeros(P + 4, dtype=torch.float32)
    gbuf[:P] = torch.from_numpy(rows[i0:i0 + n].sum(axis=0)).float()
    acc2 = torch.zeros(4, dtype=torch.float64)
    acc2[0] = float((err[i0:i0 + n] ** 2).sum())
    acc2[1] = 0.125
    acc2[2] = 7.0
    allreduce_fused(gbuf, P, acc2, dist.group.WORLD, world, True)
    out_q.put((rank, grads.numpy(), loss_acc.numpy(), gbuf.numpy(), acc2.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_allreduce_equals_single_process():
    T, P, world = 1000, 17, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, T, P, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(0)
    rows = rng.standard_normal((T, P))
    err = rng.standard_normal(T)
    for rank, g, acc, gb, acc2 in res:
        np.testing.assert_allclose(g, rows.sum(axis=0), rtol=1e-12, atol=1e-12)
        assert abs(acc[0] - (err ** 2).sum()) < 1e-9
        assert acc[1] == 0.125 and acc[2] == 7.0
        np.testing.assert_allclose(gb[:P], rows.sum(axis=0), rtol=1e-5, atol=1e-5)
        assert np.all(gb[P:] == 0.0)                                      # the tail is handed back clean
        assert abs(acc2[0] - (err ** 2).sum()) < 2e-7 * (err ** 2).sum()  # fp32-grade: the loss the scheduler sees is a float32
        assert abs(acc2[1] - 0.125) < 1e-12 and acc2[2] == 7.0
    assert np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])   # replicas stay in lock-step
    assert np.array_equal(res[0][3], res[1][3]) and np.array_equal(res[0][4], res[1][4])
