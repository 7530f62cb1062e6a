"""CPU suite, part 4: the host-side multi-GPU logic (frame blocks, query-block sharded all-vs-all with a chunked
all-gather) on world_size-2 and -3 gloo process groups.  The scoring kernel is replaced by a numpy popcount scorer
defined HERE (test code), so only the partitioning / gather / assembly logic of the product is exercised."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from orb_slam_cuda_b200 import sharding as S


def test_block_range_partitions_exactly():
    for n in (0, 1, 7, 8, 4096, 4099):
        for world in (1, 2, 3, 4, 8):
            covered = []
            for r in range(world):
                b, e = S.block_range(n, r, world)
                assert 0 <= b <= e <= n
                covered += list(range(b, e))
                for u in range(b, e):
                    assert S.owner_of(u, n, world) == r
            assert covered == list(range(n))
            sizes = [S.block_range(n, r, world)[1] - S.block_range(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    assert list(S.frames_for_rank(10, 1, 4)) == [3, 4, 5]


def test_chunk_plan():
    assert S.chunk_plan(10, 4) == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert S.chunk_plan(2, 4) == [(0, 1), (1, 2)]
    assert S.chunk_plan(0, 4) == []
    for k in range(1, 40):
        for c in range(1, 9):
            plan = S.chunk_plan(k, c)
            assert plan[0][0] == 0 and plan[-1][1] == k and all(a[1] == b[0] for a, b in zip(plan, plan[1:]))


_POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


def _score_numpy(q, db, out, th=50, ratio=0.6):
    """accepted best matches of every query keyframe in every database keyframe (the semantics of
    orbx_match_allpairs_device), in numpy."""
    qn, dn = q.numpy(), db.numpy()
    for i in range(qn.shape[0]):
        for j in range(dn.shape[0]):
            d = _POP[qn[i][:, None, :] ^ dn[j][None, :, :]].sum(-1)  # [n, n]
            s = np.sort(d, axis=1)
            b1, b2 = s[:, 0], (s[:, 1] if d.shape[1] > 1 else np.full(d.shape[0], 256))
            out[i, j] = int(((b1 <= th) & (b1.astype(np.float32) < np.float32(ratio) * b2.astype(np.float32))).sum())


def _make_db(k_total, n, seed=0):
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (k_total, n, 32), dtype=np.uint8)
    for k in range(1, k_total):  # every keyframe shares a few noisy descriptors with its predecessor
        m = n // 3
        noise = rng.integers(0, 256, (m, 32), dtype=np.uint8) & rng.integers(0, 256, (m, 32), dtype=np.uint8) & \
            rng.integers(0, 256, (m, 32), dtype=np.uint8)
        db[k, :m] = db[k - 1, n - m:] ^ noise
    return db


def _worker(rank, world, port, k_total, n, chunks, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        db = _make_db(k_total, n)
        b, e = S.block_range(k_total, rank, world)
        local = torch.from_numpy(db[b:e].copy())
        scores, moved = S.allpairs_sharded(local, k_total, _score_numpy, chunks=chunks)
        ret[rank] = (b, e, scores.numpy().copy(), moved)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,k_total,chunks", [(2, 6, 2), (2, 5, 4), (3, 7, 1)])
def test_allpairs_sharded_gloo(world, k_total, chunks):
    n = 24
    db = _make_db(k_total, n)
    want = np.zeros((k_total, k_total), np.int32)
    _score_numpy(torch.from_numpy(db), torch.from_numpy(db), want)
    assert want.max() > 0
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(world, port, k_total, n, chunks, ret), nprocs=world, join=True)
    got = np.zeros_like(want)
    for r in range(world):
        b, e, sc, moved = ret[r]
        assert sc.shape == (e - b, k_total)
        got[b:e] = sc
        assert moved >= k_total * n * 32
    assert np.array_equal(got, want)


def test_allpairs_single_process():
    db = _make_db(3, 16, seed=4)
    want = np.zeros((3, 3), np.int32)
    _score_numpy(torch.from_numpy(db), torch.from_numpy(db), want)
    got, moved = S.allpairs_sharded(torch.from_numpy(db), 3, _score_numpy)
    assert moved == 0 and np.array_equal(got.numpy(), want)
