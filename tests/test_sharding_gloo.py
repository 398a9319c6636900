"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: unit round-robin, database slices,
key packing and the all-gather + merge of the sharded 2-NN.  The local top-2 is supplied by the
oracle here (tests may use it); on the GPU box the same driver runs liborbx kernels + NCCL."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, nq, ndb, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import orbx_pkg
    pkg = orbx_pkg.load()
    from oracle import oracle as O
    sh = pkg.sharding
    q, db, _ = pkg.synth.descriptor_sets(nq, ndb, seed=12)
    db[ndb - 3] = db[2]; q[1] = db[2]                      # duplicate rows in different shards: tie -> lowest index
    start, cnt = sh.db_slice(ndb, rank, world)

    def local_top2(qq, dbl, base):
        d1, i1, d2, i2 = O.knn2(qq, dbl, nthreads=2)
        g1 = np.where(i1 >= 0, i1 + base, -1); g2 = np.where(i2 >= 0, i2 + base, -1)
        return np.stack([sh.pack_keys(d1, g1), sh.pack_keys(d2, g2)], axis=1)

    def all_gather(keys):
        t = torch.from_numpy(keys.view(np.int64).copy())
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        return torch.stack(parts).numpy().view(np.uint64)

    best, second = sh.knn2_sharded(q, db[start:start + cnt], start, local_top2, all_gather, sh.merge_keys)
    d1, i1 = sh.unpack_keys(best); d2, i2 = sh.unpack_keys(second)
    o = O.knn2(q, db, nthreads=2)
    ok = all(np.array_equal(a, b) for a, b in zip((d1, i1, d2, i2), o))
    ok = ok and (i1[1], i2[1]) == (min(2, ndb - 3), max(2, ndb - 3))
    frames = sh.units_for_rank(11, rank, world)
    t = torch.zeros(11, dtype=torch.int32); t[torch.from_numpy(frames)] = 1
    dist.all_reduce(t)
    ok = ok and bool((t == 1).all())                        # every unit exactly once
    ret[rank] = ok
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nq,ndb", [(200, 1001), (64, 3)])
def test_knn2_sharded_world2(nq, ndb):
    world = 2
    port = 29500 + (os.getpid() % 400) + ndb % 7
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, nq, ndb, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_slices_and_units():
    sys.path.insert(0, ROOT)
    import orbx_pkg
    sh = orbx_pkg.load().sharding
    for ndb in (0, 1, 7, 16_000_000):
        for world in (1, 2, 3, 8):
            sl = [sh.db_slice(ndb, r, world) for r in range(world)]
            assert sl[0][0] == 0 and sum(c for _, c in sl) == ndb
            assert all(sl[r][0] + sl[r][1] == sl[r + 1][0] for r in range(world - 1))
            assert max(c for _, c in sl) - min(c for _, c in sl) <= 1
    got = np.sort(np.concatenate([sh.units_for_rank(37, r, 8) for r in range(8)]))
    assert np.array_equal(got, np.arange(37))
    k = sh.pack_keys([3, 257], [5, -1])
    assert sh.unpack_keys(k)[0].tolist() == [3, 257] and sh.unpack_keys(k)[1].tolist() == [5, -1]
