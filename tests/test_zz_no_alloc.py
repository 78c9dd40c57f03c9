"""pt_scene_render_device_static (include/pt_render.h): the wavefront render out of the library's static device
workspace — the entry point for callers that may not allocate.  It has no counterpart in the reference; its check is
the existing entry point (pt_scene_render_device_w, itself pinned on the oracle and the goldens in
test_gpu_parity.py): same arguments, results equal bit for bit, on calls in a row that share the array with different
layouts.  Collected last on purpose."""
import numpy as np
import pytest

import ptb200
import scenes

pytestmark = pytest.mark.gpu


def assert_bits_equal(a, b, what=""):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype, what
    ne = a.view(np.uint8) != b.view(np.uint8)
    assert not ne.any(), f"{what}: {int(ne.reshape(a.size, -1).any(1).sum())} of {a.size} values differ"


def frame_buffers(cfg):
    import torch
    H, W = cfg.height, cfg.width
    # poisoned: the entry points must write every value they return
    return (torch.full((H, W, 3), float("nan"), device="cuda"), torch.full((H, W, 3), float("nan"), device="cuda"),
            torch.full((H, W), -7, dtype=torch.int32, device="cuda"), torch.full((H, W), float("nan"), device="cuda"))


def render(ds, cfg, static):
    import torch
    sm, s2, cnt, wt = frame_buffers(cfg)
    call = ds.render_device_static if static else ds.render_device
    call(cfg, sm.data_ptr(), s2.data_ptr(), cnt.data_ptr(), 0, wt.data_ptr())
    torch.cuda.synchronize()
    return sm.cpu().numpy(), s2.cpu().numpy(), cnt.cpu().numpy(), wt.cpu().numpy()


CASES = {
    # one chunk: 3072 pixels x 16 samples per step = 49152 path slots
    "one_chunk": dict(width=64, height=48, spp=16, max_depth=8, error=-1.0, seed=5),
    # 12800 pixels x 32 per step > 65536 slots: pixel chunks of 2048, 37 spp = a full and a partial step
    "chunked": dict(width=128, height=100, spp=37, max_depth=6, error=-1.0, seed=6),
    # Russian roulette (alive lists, the weight plane) and a sample partition
    "roulette": dict(width=96, height=64, spp=24, max_depth=8, error=-1.0, seed=7, rr_depth=3, sample_begin=1, sample_step=2),
    "watertight": dict(width=80, height=60, spp=8, max_depth=5, error=-1.0, seed=8, intersector=ptb200.PT_ISECT_WATERTIGHT),
    "brute": dict(width=32, height=32, spp=4, max_depth=4, error=-1.0, seed=9, accel=ptb200.PT_ACCEL_BRUTE),
}


def test_static_workspace_render_equals_the_pooled_one(lib):
    import torch
    sc = scenes.torus_room(24, 12)
    handles, want = {}, {}
    for name, kw in CASES.items():  # the existing entry point first (it allocates its workspace from the pool)
        cfg = ptb200.config(**kw)
        handles[name] = (lib.scene_create(sc, cfg), cfg)
        want[name] = render(handles[name][0], cfg, static=False)
        assert want[name][2].max() >= 1  # samples_count counts the samples that reached an emitter
    order = ["one_chunk", "chunked", "one_chunk", "roulette", "watertight", "brute", "chunked", "roulette"]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for i, name in enumerate(order):  # calls in a row: each finds the previous call's contents in the shared array
        ds, cfg = handles[name]
        got = render(ds, cfg, static=True)
        for g, w, plane in zip(got, want[name], ("sum", "sum2", "count", "weight")):
            assert_bits_equal(g, w, f"call {i} ({name}) {plane}")
    # nothing was allocated on the device by the static calls (the test's own frame buffers come from torch's cache,
    # filled by the first loop)
    assert free0 - torch.cuda.mem_get_info()[0] <= (1 << 20)
    for ds, _ in handles.values():
        ds.close()


def test_static_workspace_accepts_null_outputs_and_a_stream(lib):
    import torch
    sc = scenes.torus_room(16, 8)
    cfg = ptb200.config(width=48, height=40, spp=12, max_depth=6, error=-1.0, seed=3)
    ds = lib.scene_create(sc, cfg)
    want = render(ds, cfg, static=False)
    st = torch.cuda.Stream()
    sm = torch.full((40, 48, 3), float("nan"), device="cuda")
    cnt = torch.full((40, 48), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):
        ds.render_device_static(cfg, sm.data_ptr(), 0, cnt.data_ptr(), st.cuda_stream)
        st.synchronize()
        assert_bits_equal(sm.cpu().numpy(), want[0], "sum")
        assert_bits_equal(cnt.cpu().numpy(), want[2], "count")
        sm.fill_(float("nan"))
        torch.cuda.synchronize()
    ds.close()


def test_static_workspace_rejects_what_it_cannot_hold(lib):
    sc = scenes.torus_room(16, 8)
    ok = ptb200.config(width=128, height=128, spp=4, max_depth=4, error=-1.0)
    ds = lib.scene_create(sc, ok)
    ds.render_device_static(ok)  # exactly the maximum: 16384 pixels
    for kw, word in ((dict(width=129, height=128, spp=4, error=-1.0), "exceeds the maximum"),
                     (dict(width=1920, height=1080, spp=4, error=-1.0), "exceeds the maximum"),
                     (dict(width=64, height=64, spp=20, error=0.001), "flat-spp"),
                     (dict(width=64, height=64, spp=4, error=-1.0, mode=1), "megakernel"),
                     (dict(width=64, height=64, spp=8, error=-1.0, resume=1, sample_begin=4), "resume")):
        with pytest.raises(ptb200.PtError) as e:
            ds.render_device_static(ptb200.config(**kw))
        assert word in str(e.value), str(e.value)
    import torch
    torch.cuda.synchronize()
    ds.close()
