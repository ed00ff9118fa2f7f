"""Multi-GPU host logic: frames shard into contiguous ranges, no data-path collective.
CPU: b200v_shard_range properties; a world_size-2 gloo run in which every rank takes its range, produces a
per-frame result and the ranks' slices are gathered into one output (what bench.py's ranks and the host
"concatenation" do). GPU: b200v_sharded_decode (one instance + host thread per shard) equals a single-instance
decode and the oracle."""
import os
import socket

import numpy as np
import pytest

import ka9q_viterbi_comparison_b200 as kv


def test_shard_ranges_partition_the_batch():
    for n_frames in (0, 1, 7, 8, 1000, 1 << 20, (1 << 20) + 3):
        for n_shards in (1, 2, 3, 4, 8):
            covered, sizes = 0, []
            for g in range(n_shards):
                f0, cnt = kv.shard_range(n_frames, n_shards, g)
                assert f0 == covered and cnt >= 0
                covered += cnt
                sizes.append(cnt)
            assert covered == n_frames
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        kv.shard_range(10, 2, 2)
    with pytest.raises(ValueError):
        kv.shard_range(10, 0, 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, n_frames, ret):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        f0, cnt = kv.shard_range(n_frames, world, rank)
        # per-frame "result": a function of the GLOBAL frame index only (stands in for the decoded bytes)
        mine = torch.arange(f0, f0 + cnt, dtype=torch.int64) * 2654435761 % 251
        sizes = [kv.shard_range(n_frames, world, g)[1] for g in range(world)]
        pad = max(sizes)
        buf = torch.zeros(pad, dtype=torch.int64)
        buf[:cnt] = mine
        gathered = [torch.zeros(pad, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(gathered, buf)
        whole = torch.cat([g[:n] for g, n in zip(gathered, sizes)])
        # the timing reduction bench.py uses: max over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            ret["whole"] = whole.numpy().copy()
            ret["tmax"] = float(t.item())
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_shards_and_gathers():
    import torch.multiprocessing as mp
    n_frames, world = 1001, 2
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_rank_main, args=(world, _free_port(), n_frames, ret), nprocs=world, join=True)
        expect = np.arange(n_frames, dtype=np.int64) * 2654435761 % 251
        assert np.array_equal(ret["whole"], expect)
        assert ret["tmax"] == float(world)


@pytest.mark.gpu
@pytest.mark.parametrize("name,codec", [("ka9q27", kv.KA9Q27), ("spiral29", kv.SPIRAL29)])
def test_sharded_decode_matches_single_instance(name, codec):
    import _oracle as o
    c = o.codec_by_name(name)
    nbytes, n_frames = 16, 301
    nbits = nbytes * 8
    T = nbits + c.K - 1
    syms = np.stack([o.make_frame(c, nbytes, 0.7, seed=5, frame=f)[1] for f in range(n_frames)])
    ndev = kv.load_library().b200v_device_count()
    devices = [g % ndev for g in range(3)]             # three shards, also when only one GPU is present
    with kv.ShardedDecoder(codec, T, n_frames, devices) as sd:
        assert sd.shards() == [kv.shard_range(n_frames, 3, g) for g in range(3)]
        out = sd.decode(syms, nbits)
        # numpy arrays are pageable: the decode works and says what it cost
        assert b"pageable" in sd.lib.b200v_sharded_last_error(sd.h)
    with kv.BatchDecoder(codec, T, n_frames) as dec:
        assert np.array_equal(out, dec.decode(syms, nbits))
    for f in (0, 100, 101, 200, 300):
        q = o.Oracle(c, T)
        qo, _ = q.decode(syms[f], nbits)
        assert np.array_equal(out[f], qo)
        q.close()


def test_weighted_shard_ranges_partition_the_batch():
    """b200v_shard_range_weighted: contiguous, every frame exactly once, proportional to the weights; equal weights agree
    with b200v_shard_range up to one frame; bad weights are rejected."""
    for n_frames, weights in ((1000, [20.4] * 4 + [35.2] * 4), (7, [1, 1, 1]), (1 << 20, [1, 3]), (5, [0, 1, 0, 2]), (3, [5.0])):
        pos, sizes = 0, []
        for g in range(len(weights)):
            f0, cnt = kv.shard_range_weighted(n_frames, weights, g)
            assert f0 == pos and cnt >= 0
            pos += cnt
            sizes.append(cnt)
        assert pos == n_frames
        tot = float(sum(weights))
        for g, w in enumerate(weights):
            assert abs(sizes[g] - n_frames * w / tot) <= 1.0, (n_frames, weights, sizes)
    for g in range(3):
        a, b = kv.shard_range_weighted(301, [2, 2, 2], g), kv.shard_range(301, 3, g)
        assert abs(a[0] - b[0]) <= 1 and abs(a[1] - b[1]) <= 1
    for bad in ([1, -1], [0, 0], [float("nan"), 1]):
        with pytest.raises(ValueError):
            kv.shard_range_weighted(10, bad, 0)


@pytest.mark.gpu
def test_weighted_sharded_decode_matches_single_instance():
    import _oracle as o
    c = o.codec_by_name("ka9q27")
    nbytes, n_frames = 16, 257
    nbits = nbytes * 8
    T = nbits + c.K - 1
    syms = np.stack([o.make_frame(c, nbytes, 0.7, seed=6, frame=f)[1] for f in range(n_frames)])
    ndev = kv.load_library().b200v_device_count()
    devices = [g % ndev for g in range(3)]
    weights = [20.4, 35.2, 0.0]                              # a shard may be empty
    with kv.ShardedDecoder(kv.KA9Q27, T, n_frames, devices, weights=weights) as sd:
        assert sd.shards() == [kv.shard_range_weighted(n_frames, weights, g) for g in range(3)]
        assert sd.shards()[2][1] == 0
        out = sd.decode(syms, nbits)
    with kv.BatchDecoder(kv.KA9Q27, T, n_frames) as dec:
        assert np.array_equal(out, dec.decode(syms, nbits))
