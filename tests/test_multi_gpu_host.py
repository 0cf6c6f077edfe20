"""Multi-GPU host logic on CPU: world_size 2 under gloo. The CUDA renderer is replaced by a stand-in with the
ComputePass.render_frames signature that renders with the oracle, so what is tested is the partition + collective
plumbing of ray_tracer_b200/multi_gpu.py (frame split f = r mod N + reduce; row bands + gather), which bench.py
runs under NCCL on the GPUs."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

from ray_tracer_b200 import multi_gpu  # noqa: E402

W, H, FRAMES = 40, 31, 5


class OracleRenderer:
    """ComputePass-shaped stand-in: accumulates oracle frames into a torch tensor (H x W x 4)."""

    def __init__(self, blobs, accum):
        import oracle_binding as ob
        self.ob, self.sc, self.accum = ob, ob.OracleScene(blobs), accum

    def render_frames(self, pc, first_frame, n_frames, frame_stride=1, rows=None):
        r0, r1 = rows if rows is not None else (0, H)
        for k in range(n_frames):
            img, _ = self.sc.render_frames(pc, W, H, first_frame + k * frame_stride, 1, threads=1)
            self.accum[r0:r1] += torch.from_numpy(img[r0:r1])


def _worker(rank, world, port, blobs, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle_binding as ob
        pc = ob.default_push_constants(W, H, 10, blobs["objects"].size // 80)
        accum = torch.zeros(H, W, 4)
        total = torch.zeros(H, W, 4) if rank == 0 else None
        r = OracleRenderer(blobs, accum)
        # two batches of the sample split (uneven: 5 frames over 2 ranks, then 1 frame: rank 1 idle)
        multi_gpu.render_frames_split(r, pc, 0, FRAMES, accum, total)
        multi_gpu.render_frames_split(r, pc, FRAMES, 1, accum, total)
        assert float(accum.abs().sum()) == 0.0              # step buffers are handed back zeroed
        # one frame split in row bands
        multi_gpu.render_frame_bands(r, pc, 7, accum)
        if rank == 0:
            out["split"] = total.numpy().copy()
            out["bands"] = accum.numpy().copy()
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_frame_share_and_row_bands_partition():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 5, 8, 1024):
            seen = []
            for r in range(world):
                first, count, stride = multi_gpu.frame_share(10, n, r, world)
                seen += [first + k * stride for k in range(count)]
            assert sorted(seen) == list(range(10, 10 + n))
        for h in (1, 7, 1080, 2160):
            bands = [multi_gpu.row_band(h, r, world) for r in range(world)]
            assert bands[0][0] == 0 and bands[-1][1] == h
            assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in bands]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_world_size_2_gloo(cornell_blobs):
    import oracle_binding as ob
    ob.lib()                                              # build once before forking
    pc = ob.default_push_constants(W, H, 10, cornell_blobs["objects"].size // 80)
    sc = ob.OracleScene(cornell_blobs)
    want, _ = sc.render_frames(pc, W, H, 0, FRAMES + 1, threads=1)
    want_band, _ = sc.render_frames(pc, W, H, 7, 1, threads=1)
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(2, _free_port(), cornell_blobs, out), nprocs=2, join=True)
        got, bands = out["split"], out["bands"]
    assert np.array_equal(got[..., 3], np.full((H, W), FRAMES + 1, np.float32))
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)   # equal up to fp32 summation order
    assert np.array_equal(bands, want_band)               # bands are exact
