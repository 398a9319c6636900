"""orbx_knn2_device_noalloc (include/orbx.h): the allocation-free variant of the brute-force 2-NN
search (partial keys in a static device array inside liborbx.so) must return exactly what
orbx_knn2_device returns -- integer keys, so the tolerance is bit-exact as in test_gpu_matcher.py --
on two calls in a row (the array keeps the first call's contents).  The entry point has no
counterpart in the reference; it exists for callers that cannot call cudaMalloc.
Named test_zz_* so that pytest collects it last: a failure here must not hide other tests under -x."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _download(ox, buf, n):
    out = np.empty(n, np.uint64)
    st = ox.lib().orbx_memcpy_d2h(out.ctypes.data_as(C.c_void_p), buf.ptr, out.nbytes)
    assert st == 0, ox.lib().orbx_last_error()
    return out


def test_knn2_noalloc_matches_knn2_device_twice(pkg, synth):
    ox = pkg.orbx
    m = ox.ORBmatcher()
    cap = ox.lib().orbx_knn2_noalloc_scratch_bytes()
    assert cap == 8 << 20
    # two different problems back to back; both cut the database into several grid.y slices, so
    # both really go through the static scratch array (slices * nq * 16 bytes, asserted below)
    cases = [(3000, 300_000, 5, 0), (1500, 120_001, 6, 1 << 20)]
    bufs = []
    try:
        for nq, ndb, seed, base in cases:
            q, db, _ = synth.descriptor_sets(nq, ndb, seed=seed)
            db[7] = db[3]; q[0] = db[3]                      # exact duplicates: the tie goes to the lowest index
            dq = ox.DeviceBuffer(q.nbytes); dq.upload(q)
            ddb = ox.DeviceBuffer(db.nbytes); ddb.upload(db)
            k_ref = ox.DeviceBuffer(nq * 16); k_new = ox.DeviceBuffer(nq * 16)
            bufs += [dq, ddb, k_ref, k_new]
            k_new.upload(np.full(2 * nq, 0xDEADBEEF, np.uint64))     # stale output must be overwritten
            m.knn2_device(dq.ptr, nq, ddb.ptr, ndb, base, k_ref.ptr)
            assert m.last_launches() == 2                    # search + merge of the slices: scratch in use
            m.knn2_device_noalloc(dq.ptr, nq, ddb.ptr, ndb, base, k_new.ptr)
            assert m.last_launches() == 2
            m.synchronize()
            ref, new = _download(ox, k_ref, 2 * nq), _download(ox, k_new, 2 * nq)
            assert np.array_equal(ref, new)
            # and the keys mean what they should: (dist << 32 | global index), best of the duplicate pair first
            assert ref[0] == base + 3 and ref[1] == base + 7
            d1 = (ref[0::2] >> np.uint64(32)).astype(np.int64); d2 = (ref[1::2] >> np.uint64(32)).astype(np.int64)
            assert np.all(d1 <= d2) and d1.max() <= 256
            k = nq // 2
            d = np.unpackbits(q[k][None] ^ db, axis=1).sum(1)
            o = np.lexsort((np.arange(ndb), d))[:2]
            assert int(ref[2 * k]) == (int(d[o[0]]) << 32 | base + int(o[0])) and int(ref[2 * k + 1]) == (int(d[o[1]]) << 32 | base + int(o[1]))
    finally:
        for b in bufs:
            b.free()
        m.close()


def test_knn2_noalloc_rejects_oversize(pkg):
    """Beyond the static array the call fails with a clear error before touching the GPU buffers."""
    ox = pkg.orbx
    m = ox.ORBmatcher()
    try:
        nq, ndb = 1 << 20, (1 << 24)                         # 4 slices x 1 M queries x 16 B = 64 MiB > 8 MiB
        with pytest.raises(ox.OrbxError, match="static array"):
            m.knn2_device_noalloc(0x1000, nq, 0x1000, ndb, 0, 0x1000)    # never dereferenced: rejected on the host
    finally:
        m.close()
