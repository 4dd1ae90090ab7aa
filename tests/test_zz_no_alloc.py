"""hs_bucket_histogram_noalloc — the entry point for hosts that cannot take memory from cudaMalloc
(their GPU belongs to another allocator's pool, or they may not cause the device-wide stall an
allocation brings). It keeps its histogram counters in a fixed array inside the library, shared by all
calls, so the interesting failure is a second call seeing the first call's counts. This file is
collected last on purpose: if it fails under `-x`, every other GPU test has already run.

Checked against the ordinary entry point hs_hash_bucket (bit-exact bucket ids) on two calls in a row,
the second with other data and fewer buckets."""
import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu


def _device_keys(torch, abi, k64, k32):
    return abi.DeviceColumnSet([
        {"type": abi.HS_INT32, "length": k32.numel(), "data": k32.data_ptr()},
        {"type": abi.HS_INT64, "length": k64.numel(), "data": k64.data_ptr()}])


def test_noalloc_matches_hash_bucket_twice_in_a_row(hs):
    import torch
    from hyperspace_1_b200 import abi
    dev = torch.device("cuda", 0)
    ctx = hs.Context(0)
    try:
        for call, (n, nb, seed) in enumerate([(300_000, 8192, 1), (123_457, 200, 2)]):
            g = torch.Generator(device=dev)
            g.manual_seed(seed)
            k64 = torch.randint(-2**62, 2**62, (n,), generator=g, device=dev, dtype=torch.int64)
            k32 = torch.randint(-2**31, 2**31 - 1, (n,), generator=g, device=dev, dtype=torch.int64).to(torch.int32)
            ids = torch.full((n,), -1, dtype=torch.int32, device=dev)
            counts = np.full(nb, 0xFFFFFFFF, dtype=np.uint32)          # host memory
            torch.cuda.synchronize()
            ctx.bucket_histogram_noalloc(_device_keys(torch, abi, k64, k32), nb, ids.data_ptr(), counts.ctypes.data)
            # the existing entry point, same keys (host copies): ids must agree bit for bit
            want = ctx.hash_bucket([pa.array(k32.cpu().numpy()), pa.array(k64.cpu().numpy())], nb)
            got = ids.cpu().numpy()
            assert np.array_equal(got, want), f"call {call}: bucket ids differ from hs_hash_bucket"
            assert np.array_equal(counts.astype(np.int64), np.bincount(want, minlength=nb)), \
                f"call {call}: counts are wrong (stale counters from the previous call?)"
            assert int(counts.sum()) == n
        # device-resident counts work too
        dcounts = torch.zeros(200, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        ctx.bucket_histogram_noalloc(_device_keys(torch, abi, k64, k32), 200, ids.data_ptr(), dcounts.data_ptr())
        assert np.array_equal(dcounts.cpu().numpy().astype(np.int64), np.bincount(want, minlength=200))
    finally:
        ctx.close()


def test_noalloc_rejects_what_it_cannot_do(hs):
    import torch
    from hyperspace_1_b200 import abi
    dev = torch.device("cuda", 0)
    ctx = hs.Context(0)
    try:
        k = torch.arange(100, dtype=torch.int64, device=dev)
        ids = torch.empty(100, dtype=torch.int32, device=dev)
        counts = np.zeros(abi_max := 8192, dtype=np.uint32)
        keys = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": 100, "data": k.data_ptr()}])
        torch.cuda.synchronize()
        with pytest.raises(hs.HyperspaceException, match="beyond"):
            ctx.bucket_histogram_noalloc(keys, abi_max + 1, ids.data_ptr(), np.zeros(abi_max + 1, np.uint32).ctypes.data)
        host_keys = abi.ColumnSet([pa.array(np.arange(100, dtype=np.int64))])   # host buffers: would need staging
        with pytest.raises(hs.HyperspaceException, match="device memory"):
            ctx.bucket_histogram_noalloc(host_keys, 16, ids.data_ptr(), counts.ctypes.data)
    finally:
        ctx.close()
