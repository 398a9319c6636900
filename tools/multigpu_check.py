#!/usr/bin/env python3
"""Run under torchrun on N GPUs: database-sharded 2-NN with the NCCL all-gather of packed top-2
keys must equal the single-GPU search of the whole database (rank 0 checks and prints)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg  # noqa: E402

pkg = orbx_pkg.load()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nq, ndb = 4096, 200003
q, db, _ = pkg.synth.descriptor_sets(nq, ndb, seed=5)
m = pkg.orbx.ORBmatcher(device=local)
lo, cnt = pkg.sharding.db_slice(ndb, rank, world)
hi = lo + cnt
dev = torch.device("cuda", local)
q_dev = torch.from_numpy(q).to(dev)
db_dev = torch.from_numpy(np.ascontiguousarray(db[lo:hi])).to(dev)
got = pkg.sharding.knn2_sharded_cuda(m, q_dev, db_dev, lo, 0.8, dist)
torch.cuda.synchronize()
if rank == 0:
    ref = m.knn2(q, db, 0.8)
    ok = all(np.array_equal(g.cpu().numpy().astype(np.int64), np.asarray(r).astype(np.int64)) for g, r in zip(got, ref))
    print("knn2 sharded over", world, "GPUs == single GPU:", ok, flush=True)
    assert ok
dist.barrier()
dist.destroy_process_group()
