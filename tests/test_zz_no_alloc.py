"""b200pt_trace_batch_device_static — the closest-hit batch call that allocates nothing (its scratch is a file-scope
static device array, include/b200pt.h) — against the existing entry point b200pt_trace_batch_device, through the C ABI.

Collected last on purpose (the name sorts after every other test file).
Bar: the records are bit-identical (integer / index work and the same float operations in the same kernels)."""
import numpy as np
import pytest

import oracle_binding as ob
from scenes import camera_rays, random_rays

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from ray_tracer_b200 import build
    build.build()
    import ray_tracer_b200 as rt
    return rt


def _records(t, n):
    from ray_tracer_b200.backend import HIT_DTYPE
    return t[: n * 48].cpu().numpy().view(HIT_DTYPE)


def _same(a, b):
    return a.tobytes() == b.tobytes()


def test_static_scratch_entry_matches_the_allocating_one(gpu, cornell_blobs):
    import torch
    from ray_tracer_b200.backend import BackendError
    blobs = cornell_blobs
    nobj = blobs["objects"].size // 80
    pc = ob.default_push_constants(256, 256, 10, nobj)
    cap = gpu.ComputePass.STATIC_MAX_RAYS
    first = np.concatenate([camera_rays(pc, 128, 128), random_rays(cap - 128 * 128, seed=11)])      # exactly the maximum
    second = random_rays(20011, seed=12)                                                               # smaller: stale contents stay behind it
    assert len(first) == cap
    cp = gpu.ComputePass(64, 64)
    cp.copy_buffers(blobs)
    dev = torch.device("cuda", 0)
    d_first = torch.from_numpy(first).to(dev)
    d_second = torch.from_numpy(second).to(dev)
    want_first = torch.zeros(cap * 48, dtype=torch.uint8, device=dev)
    want_second = torch.zeros(cap * 48, dtype=torch.uint8, device=dev)
    got = torch.zeros(cap * 48, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    # the existing entry point (scratch grown with cudaMalloc)
    cp.trace_batch_device(pc, d_first.data_ptr(), len(first), want_first.data_ptr())
    cp.trace_batch_device(pc, d_second.data_ptr(), len(second), want_second.data_ptr())
    cp.sync()
    w1, w2 = _records(want_first, len(first)), _records(want_second, len(second))
    assert (w1["kind"] != 0).mean() > 0.5 and (w2["kind"] != 0).mean() > 0.5          # the comparison below is not about misses
    # the oracle agrees with the existing entry point on a slice (so both are checked against the checker, not only each other)
    o, _ = ob.OracleScene(blobs).trace_batch(pc, second[:4000])
    assert np.array_equal(o["kind"], w2["kind"][:4000]) and np.array_equal(o["dst"].view(np.uint32), w2["dst"][:4000].view(np.uint32))

    free_before, _ = torch.cuda.mem_get_info(dev)
    # two calls in a row on the shared static scratch: the full size, then a smaller batch
    cp.trace_batch_device_static(pc, d_first.data_ptr(), len(first), got.data_ptr())
    cp.sync()
    g1 = _records(got, len(first)).copy()
    got.fill_(0xAB)
    torch.cuda.synchronize()
    cp.trace_batch_device_static(pc, d_second.data_ptr(), len(second), got.data_ptr())
    cp.sync()
    g2 = _records(got, len(second)).copy()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert _same(g1, w1), "first call: %d of %d records differ" % (int((g1 != w1).sum()), len(w1))
    assert _same(g2, w2), "second call: %d of %d records differ" % (int((g2 != w2).sum()), len(w2))
    assert bytes(got[len(second) * 48: len(second) * 48 + 48].cpu().numpy()) == b"\xab" * 48      # nothing written past n records
    assert free_after >= free_before, "device memory in use grew by %d bytes across the two calls" % (free_before - free_after)
    # again the first batch: the array holds the second call's contents now
    cp.trace_batch_device_static(pc, d_first.data_ptr(), len(first), got.data_ptr())
    cp.sync()
    assert _same(_records(got, len(first)), w1)

    # beyond the maximum: rejected with a message, nothing launched
    with pytest.raises(BackendError, match="B200PT_STATIC_MAX_RAYS"):
        cp.trace_batch_device_static(pc, d_first.data_ptr(), cap + 1, got.data_ptr())
    # n = 0 is a no-op
    cp.trace_batch_device_static(pc, 0, 0, 0)
    cp.sync()
