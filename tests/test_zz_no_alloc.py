"""GPU test of me_search_static_scratch, the entry point for callers that cannot allocate.

Who that is: code running inside another framework's memory pool (the pool owns the GPU, a stray
cudaMalloc either fails or fragments it), a latency-critical loop where cudaMalloc's device-wide
synchronisation would stall unrelated streams, or a worker that only ever receives buffers from its
parent.  Such a caller hands over device frames and a device output array it already owns; the
library's working memory comes from a fixed array inside the loaded module.

The file is named so that pytest collects it last: a failure here cannot hide the other tests of an
``-x`` run.  The check is the library against itself: the same pairs through the ordinary
context-based entry point (me_search_device + me_fetch_pairs) must give byte-identical records
(tolerance 0, as in tests/test_gpu_parity.py), twice in a row with different inputs so that stale
contents of the shared array from the first call would show up in the second.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _static_search(me, frames_t, pairs, cfg, out_t, stream):
    L = me._lib.lib()
    p = np.ascontiguousarray(np.asarray(pairs, np.int32).reshape(-1, 2))
    prm = cfg.as_struct()
    rc = L.me_search_static_scratch(C.c_void_p(frames_t.data_ptr()), frames_t.stride(1), frames_t.shape[0],
                                    frames_t.shape[1], frames_t.shape[2], C.byref(prm),
                                    p.ctypes.data_as(C.POINTER(C.c_int32)), p.shape[0],
                                    C.c_void_p(out_t.data_ptr()), C.c_void_p(stream))
    return rc


def test_static_scratch_entry_point_equals_context_entry_point(me):
    import torch
    from motion_estimation_b200 import synth

    H, W = 96, 256
    cfg = me.SearchConfig(block=16, range=16)
    stream = torch.cuda.current_stream().cuda_stream
    with me.MotionEstimator(H, W, cfg, max_frames=2, max_pairs=4) as est:
        for call, (seed, pairs) in enumerate([(0, [(0, 1), (1, 2), (2, 3)]), (7, [(3, 0), (2, 1), (0, 2), (1, 1)])]):
            frames, _ = synth.translating_frames(4, H, W, 14, seed_texture=seed, seed_motion=seed + 1)
            t = torch.from_numpy(frames).cuda().contiguous()
            out = torch.zeros(len(pairs) * 8, dtype=torch.int32, device="cuda")     # 32-byte records
            assert _static_search(me, t, pairs, cfg, out, stream) == 0, me._lib.lib().me_last_error()
            torch.cuda.synchronize()
            got = out.cpu().numpy().view(me.PAIR_DTYPE)
            est.search_device(t.data_ptr(), t.stride(1), t.shape[0], pairs, stream=stream)
            want = est.fetch_pairs(len(pairs))
            assert got.tobytes() == want.tobytes(), f"call {call}: {got} != {want}"
            assert want["flag"].all() and want["nvalid_x"].min() > 10


def test_static_scratch_against_the_oracle_and_without_allocating(me):
    """The entry point checked directly against the CPU oracle (not only against the context path), and its promise checked
    too: free device memory as the driver reports it (cudaMemGetInfo) is the same before and after a call that follows the
    first one -- the first call may still load the kernel module, later ones must not touch the allocator at all."""
    import torch
    from motion_estimation_b200 import synth
    from oracle import sad as osad
    H, W = 128, 320
    cfg = me.SearchConfig(block=16, range=24, range_y=6)
    stream = torch.cuda.current_stream().cuda_stream
    frames, _ = synth.translating_frames(3, H, W, 10, seed_texture=5, seed_motion=6)
    t = torch.from_numpy(frames).cuda().contiguous()
    pairs = [(0, 1), (1, 2), (2, 0)]
    out = torch.zeros(len(pairs) * 8, dtype=torch.int32, device="cuda")
    assert _static_search(me, t, pairs, cfg, out, stream) == 0, me._lib.lib().me_last_error()       # first call: one-time set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(2):
        assert _static_search(me, t, pairs, cfg, out, stream) == 0, me._lib.lib().me_last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, f"device memory changed by {free0 - free1} bytes across calls that must not allocate"
    got = out.cpu().numpy().view(me.PAIR_DTYPE)
    for i, (a, b) in enumerate(pairs):
        o = osad.search(frames[a], frames[b], 16, 24, range_y=6)
        assert (got[i]["pdx"], got[i]["pdy"], got[i]["flag"], got[i]["nvalid_x"], got[i]["nvalid_y"], got[i]["n_textured"],
                got[i]["mean_dx"], got[i]["mean_dy"]) == (o["pdx"], o["pdy"], o["flag"], o["nvalid_x"], o["nvalid_y"],
                                                          o["n_textured"], o["mean_dx"], o["mean_dy"]), i


def test_static_scratch_with_candidate_slabs(me):
    """A range large enough to be cut into several CTAs per tile uses two more pieces of the shared array."""
    import torch
    from motion_estimation_b200 import synth
    H, W = 96, 208
    cfg = me.SearchConfig(block=16, range=100)
    stream = torch.cuda.current_stream().cuda_stream
    frames, _ = synth.translating_frames(3, H, W, 24)
    t = torch.from_numpy(frames).cuda().contiguous()
    pairs = [(0, 1), (1, 2)]
    with me.MotionEstimator(H, W, cfg, max_frames=2, max_pairs=2) as est:
        assert est.geometry["slabs_x"] * est.geometry["slabs_y"] > 1
        for _ in range(2):
            out = torch.zeros(len(pairs) * 8, dtype=torch.int32, device="cuda")
            assert _static_search(me, t, pairs, cfg, out, stream) == 0, me._lib.lib().me_last_error()
            torch.cuda.synchronize()
            est.search_device(t.data_ptr(), t.stride(1), t.shape[0], pairs, stream=stream)
            assert out.cpu().numpy().view(me.PAIR_DTYPE).tobytes() == est.fetch_pairs(len(pairs)).tobytes()


def test_static_scratch_with_multi_row_tiles(me):
    """A small range on a large frame is planned with tiles of several block rows; same records as the context path."""
    import torch
    from motion_estimation_b200 import synth
    H, W = 720, 1280
    cfg = me.SearchConfig(block=16, range=8)
    stream = torch.cuda.current_stream().cuda_stream
    frames, _ = synth.translating_frames(3, H, W, 7)
    t = torch.from_numpy(frames).cuda().contiguous()
    pairs = [(0, 1), (1, 2), (2, 0), (1, 0)]
    with me.MotionEstimator(H, W, cfg, max_frames=3, max_pairs=4) as est:
        assert est.geometry["tile_blocks_y"] > 1
        for _ in range(2):
            out = torch.zeros(len(pairs) * 8, dtype=torch.int32, device="cuda")
            assert _static_search(me, t, pairs, cfg, out, stream) == 0, me._lib.lib().me_last_error()
            torch.cuda.synchronize()
            est.search_device(t.data_ptr(), t.stride(1), t.shape[0], pairs, stream=stream)
            assert out.cpu().numpy().view(me.PAIR_DTYPE).tobytes() == est.fetch_pairs(len(pairs)).tobytes()


def test_static_scratch_rejects_oversized_requests(me):
    import torch
    L = me._lib.lib()
    cfg = me.SearchConfig(block=8, range=8)
    H, W = 1080, 1920                                     # 32 400 blocks x 17 B per pair: 8 pairs do not fit 4 MiB
    t = torch.zeros((2, H, W), dtype=torch.uint8, device="cuda")
    out = torch.zeros(8 * 8, dtype=torch.int32, device="cuda")
    rc = _static_search(me, t, [(0, 1)] * 8, cfg, out, torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = L.me_last_error().decode()
    assert "ME_STATIC_SCRATCH_BYTES" in msg and "4194304" in msg
    assert _static_search(me, t, [(0, 3)], cfg, out, 0) == -1 and "outside" in L.me_last_error().decode()
