"""GPU tests of the library's multi-GPU entries (include/b200pt.h "multi-GPU", csrc/comm.cu) and of the interleaved-row
split they rest on. The exactness claims: a pixel's seed depends only on (x, y, W, frame) (shaders/raytrace.comp:562-564),
so ANY subset of rows renders the same pixel values; frames have independent seeds, so a sample split changes only the
order of the float sums.

One GPU is enough for the row-stride tests and for a one-rank communicator (NCCL bound, initialised and called);
the two-rank tests need two devices and skip otherwise (`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`)."""
import threading

import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from ray_tracer_b200 import build
    build.build()
    import ray_tracer_b200 as rt
    return rt


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _pass(gpu, blobs, W, H, device=0):
    cp = gpu.ComputePass(W, H, device=device)
    cp.copy_buffers(blobs)
    return cp


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_interleaved_rows_tile_the_frame_exactly(gpu, bunny_blobs, nranks):
    W, H = 160, 101                                   # H not a multiple of any nranks
    pc = ob.default_push_constants(W, H, 10, bunny_blobs["objects"].size // 80)
    cp = _pass(gpu, bunny_blobs, W, H)
    cp.render_frames(pc, 5, 2)
    whole = cp.read_accum().copy()
    cp.clear()
    for r in range(nranks):                           # virtual ranks, one after the other, into the same accumulator
        cp.render_frames(pc, 5, 2, rows=(r, H), row_stride=nranks)
    split = cp.read_accum()
    assert np.array_equal(bits(split), bits(whole))
    # a band of interleaved rows: rows 10, 13, 16 ... below 50 and nothing else
    cp.clear()
    cp.render_frames(pc, 5, 2, rows=(10, 50), row_stride=3)
    part = cp.read_accum()
    touched = np.zeros(H, bool); touched[10:50:3] = True
    assert np.array_equal(bits(part[touched]), bits(whole[touched]))
    assert not part[~touched].any()
    cp.close()


def test_one_rank_communicator_goes_through_nccl(gpu, cornell_blobs):
    from ray_tracer_b200 import backend
    W = H = 96
    pc = ob.default_push_constants(W, H, 10, 9)
    cp = _pass(gpu, cornell_blobs, W, H)
    cp.render_frames(pc, 0, 3)
    want = cp.read_accum().copy()
    cp.clear()
    assert cp.comm_info() == (1, 0)
    cp.comm_init(1, 0, backend.comm_unique_id())
    assert cp.comm_info() == (1, 0)
    with pytest.raises(backend.BackendError):
        cp.comm_init(1, 0, backend.comm_unique_id())      # one communicator per context
    cp.render_frames_split(pc, 0, 3)
    cp.reduce_accum(0)                                    # ncclReduce of one rank onto itself
    assert np.array_equal(bits(cp.read_accum()), bits(want))
    cp.clear()
    cp.render_frame_interleaved(pc, 1, 0)
    one = cp.read_accum().copy()
    cp.clear()
    cp.render_frames(pc, 1, 1)
    assert np.array_equal(bits(one), bits(cp.read_accum()))
    cp.comm_destroy()
    assert cp.comm_info() == (1, 0)
    cp.close()


def _two_ranks(gpu, blobs, W, H, body):
    """Runs body(rank, cp) on two threads, one ComputePass per device, joined by a communicator."""
    from ray_tracer_b200 import backend
    uid = backend.comm_unique_id()
    out, errs = [None, None], []

    def work(r):
        try:
            cp = _pass(gpu, blobs, W, H, device=r)
            cp.comm_init(2, r, uid)
            out[r] = body(r, cp)
            cp.sync()
            cp.comm_destroy()
            cp.close()
        except Exception as e:      # noqa: BLE001
            errs.append((r, repr(e)))

    th = [threading.Thread(target=work, args=(r,)) for r in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    return out


def test_two_gpus_sample_split_and_interleaved_frame(gpu, bunny_blobs):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    W, H = 192, 108
    pc = ob.default_push_constants(W, H, 10, bunny_blobs["objects"].size // 80)
    single = _pass(gpu, bunny_blobs, W, H)
    single.render_frames(pc, 0, 7)
    want = single.read_accum().copy()
    single.clear()
    single.render_frames(pc, 3, 1)
    frame3 = single.read_accum().copy()
    single.close()

    def split(r, cp):
        cp.render_frames_split(pc, 0, 7)              # rank 0: frames 0 2 4 6, rank 1: 1 3 5
        cp.reduce_accum(0)
        return cp.read_accum().copy()
    got = _two_ranks(gpu, bunny_blobs, W, H, split)
    assert np.array_equal(got[0][..., 3], want[..., 3])                        # 7 frames counted in every pixel
    assert np.allclose(got[0][..., :3], want[..., :3], rtol=2e-6, atol=1e-7)   # same samples, another summation order
    assert not got[1].any()                                                    # zero_others

    def inter(r, cp):
        cp.render_frame_interleaved(pc, 3, 0)
        return cp.read_accum().copy()
    got = _two_ranks(gpu, bunny_blobs, W, H, inter)
    assert np.array_equal(bits(got[0]), bits(frame3))                          # the root holds the whole frame, bit for bit
    assert np.array_equal(bits(got[1][1::2]), bits(frame3[1::2])) and not got[1][0::2].any()
