"""Multi-GPU host logic (one process per GPU).

* Extraction, stereo and the per-frame searches shard over INDEPENDENT units (frames, stereo
  pairs, KeyFrame pairs): unit i -> rank i mod world, no data-path collective.
* Only the database-sharded brute-force 2-NN has an exchange step: every rank holds a
  contiguous slice of the database and all queries, computes its local top-2 as packed
  (dist << 32 | global index) keys, all-gathers the 16 B/query key pairs, and merges.

This module is backend-agnostic on purpose: the CUDA path passes the liborbx kernels and
torch.distributed/NCCL collectives; the CPU tests (gloo, world_size 2) pass stand-ins, so the
index arithmetic and merge order are covered without a GPU."""
import numpy as np

NONE_KEY = np.uint64(0xFFFFFFFFFFFFFFFF)


def units_for_rank(n_units, rank, world):
    """Round-robin assignment of independent units (frames / pairs): unit i -> rank i % world.
    L/R frames of a stereo pair form ONE unit, so a pair never straddles GPUs."""
    return np.arange(rank, n_units, world)


def db_slice(ndb, rank, world):
    """Contiguous database slice [start, start+count) of `rank`; sizes differ by at most one row."""
    base, rem = divmod(ndb, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def pack_keys(d, idx):
    """(distance, global index) -> packed u64 keys; idx < 0 -> NONE_KEY."""
    d = np.asarray(d, np.uint64); i = np.asarray(idx, np.int64)
    k = (d << np.uint64(32)) | i.astype(np.uint64)
    k[i < 0] = NONE_KEY
    return k


def merge_keys(parts):
    """parts: (nparts, nq, 2) u64 -> best, second (nq,) u64.  Key order == (distance, index), so
    ties go to the lowest global index regardless of which rank found them."""
    parts = np.asarray(parts, np.uint64)
    flat = np.sort(parts.transpose(1, 0, 2).reshape(parts.shape[1], -1), axis=1)
    return flat[:, 0], flat[:, 1]


def unpack_keys(k):
    k = np.asarray(k, np.uint64)
    none = k == NONE_KEY
    d = (k >> np.uint64(32)).astype(np.int64); i = (k & np.uint64(0xFFFFFFFF)).astype(np.int64)
    d[none] = 257; i[none] = -1
    return d.astype(np.int32), i.astype(np.int32)


def knn2_sharded(queries, db_local, index_base, local_top2, all_gather, merge):
    """Generic driver.  local_top2(queries, db_local, index_base) -> (nq, 2) keys;
    all_gather(keys) -> (world, nq, 2); merge(parts) -> result."""
    return merge(all_gather(local_top2(queries, db_local, index_base)))


# ------------------------------------------------------------------ CUDA / NCCL path
def knn2_sharded_cuda(matcher, q_dev, db_dev, index_base, ratio, dist=None):
    """q_dev: (nq, 32) u8 CUDA tensor (replicated); db_dev: (ndb_local, 32) u8 CUDA tensor (this
    rank's slice).  Returns d1, i1, d2, i2, pass CUDA tensors (identical on every rank)."""
    import torch
    nq = q_dev.shape[0]
    dev = q_dev.device
    stream = torch.cuda.ExternalStream(matcher.stream(), device=dev)
    keys = torch.empty((nq, 2), dtype=torch.int64, device=dev)
    stream.wait_stream(torch.cuda.current_stream(dev))
    matcher.knn2_device(q_dev.data_ptr(), nq, db_dev.data_ptr(), db_dev.shape[0], int(index_base), keys.data_ptr())
    world = dist.get_world_size() if dist is not None else 1
    if world > 1:
        parts = torch.empty((world, nq, 2), dtype=torch.int64, device=dev)
        torch.cuda.current_stream(dev).wait_stream(stream)
        dist.all_gather_into_tensor(parts, keys)              # NCCL over NVLink: 16 B per query per rank
        stream.wait_stream(torch.cuda.current_stream(dev))
    else:
        parts = keys
    out = [torch.empty(nq, dtype=torch.int32, device=dev) for _ in range(4)]
    ps = torch.empty(nq, dtype=torch.uint8, device=dev)
    matcher.knn2_merge_device(parts.data_ptr(), world, nq, ratio, out[0].data_ptr(), out[1].data_ptr(),
                              out[2].data_ptr(), out[3].data_ptr(), ps.data_ptr())
    torch.cuda.current_stream(dev).wait_stream(stream)
    return out[0], out[1], out[2], out[3], ps
