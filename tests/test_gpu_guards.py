"""GPU: out-of-bounds WRITE detection without compute-sanitizer (closed on this pool).  Every output / workspace tensor the
ctypes wrappers allocate is placed between two 4 KB guard zones filled with 0xA5; after each call the guard zones must be
intact.  Shapes are chosen off the kernels' vector widths and tile sizes (ragged segments, odd counts, tails)."""
import numpy as np
import pytest
import torch

from cvpce_b200 import synth

pytestmark = pytest.mark.gpu
GUARD = 4096


class GuardedTorch:
    """Stands in for the `torch` name inside cvpce_b200.ops: CUDA `empty` / `zeros` come with guard zones, all else is torch."""

    def __init__(self):
        self.allocs = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _alloc(self, size, dtype, device, zero):
        if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
            size = tuple(size[0])
        dtype = dtype or torch.float32
        if device is None or torch.device(device).type != "cuda":
            return (torch.zeros if zero else torch.empty)(size, dtype=dtype, device=device)
        nbytes = int(np.prod(size, dtype=np.int64)) * torch.empty((), dtype=dtype).element_size()
        buf = torch.full((GUARD + nbytes + GUARD,), 0xA5, dtype=torch.uint8, device=device)
        view = buf[GUARD:GUARD + nbytes].view(dtype).view(tuple(int(s) for s in size))
        if zero:
            view.zero_()
        self.allocs.append((buf, nbytes))
        return view

    def empty(self, *size, dtype=None, device=None, **kw):
        return self._alloc(size, dtype, device, False)

    def zeros(self, *size, dtype=None, device=None, **kw):
        return self._alloc(size, dtype, device, True)

    def check(self, what):
        torch.cuda.synchronize()
        assert self.allocs, what
        for buf, nbytes in self.allocs:
            assert bool((buf[:GUARD] == 0xA5).all()), "%s: write BEFORE a %d-byte output" % (what, nbytes)
            assert bool((buf[GUARD + nbytes:] == 0xA5).all()), "%s: write PAST a %d-byte output" % (what, nbytes)
        n = len(self.allocs)
        self.allocs = []
        return n


@pytest.fixture()
def guarded(cuda, monkeypatch):
    from cvpce_b200 import _lib, ops
    assert _lib.load().cvpce_device_check() == 0, _lib.last_error()
    g = GuardedTorch()
    monkeypatch.setattr(ops, "torch", g)
    return ops, g


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_detection_kernels_stay_inside_their_outputs(guarded):
    ops, g = guarded
    # NMS: ragged CSR batch (empty segment, 1 box, 63 / 64 / 65 boxes = panel edges, a large one)
    sizes = [0, 1, 63, 64, 65, 1000, 7]
    boxes = np.concatenate([synth.random_boxes(10 + i, n, extent=200.0)[0] for i, n in enumerate(sizes)] + [np.zeros((0, 4), np.float32)])
    scores = np.concatenate([synth.random_boxes(10 + i, n, extent=200.0)[1] for i, n in enumerate(sizes)] + [np.zeros(0, np.float32)])
    seg = np.cumsum([0] + sizes).astype(np.int32)
    scale = np.tile(np.array([[1.5, 0.75]], np.float32), (len(sizes), 1))
    ops.nms_batched(dev(boxes), dev(scores), dev(seg), max(sizes), 0.5, None, 0, dev(scale), True, True)
    assert g.check("nms_batched") == 4
    ops.nms_batched(dev(boxes), dev(scores), dev(seg), max(sizes), 0.3, 0.2, 5, None, True, True)
    g.check("nms_batched max_keep")
    stride = 1000
    sb = np.zeros((3 * stride, 4), np.float32)
    ss = np.zeros(3 * stride, np.float32)
    for i, n in enumerate((1000, 0, 333)):
        sb[i * stride:i * stride + n], ss[i * stride:i * stride + n] = synth.random_boxes(30 + i, n, extent=150.0)
    ops.nms_segments(dev(sb), dev(ss), stride, dev(np.array([1000, 0, 333], np.int32)), 0.5, None, 200, None, True, True)
    g.check("nms_segments")
    # candidate selection on a small multi-level head
    for name in list(synth.HEAD_CASES)[:3]:
        logits, regression, anchors, image_sizes, C, topk = synth.head_case(name)
        ops.select_candidates([dev(x) for x in logits], [dev(x) for x in regression], [dev(a) for a in anchors], image_sizes, 0.05, topk)
        g.check("select_candidates " + name)
    # box IoU / matching: row pitches with and without the vector store, several bit-words of columns
    for na, nb in ((1, 1), (3, 5), (17, 8), (33, 1023), (40, 1025)):
        a, b = synth.random_boxes(1, na, extent=300.0)[0], synth.random_boxes(2, nb, extent=300.0)[0]
        ops.box_iou(dev(a), dev(b))
        g.check("box_iou %dx%d" % (na, nb))
        ops.iou_match(dev(a), dev(b), 0.3)
        g.check("iou_match %dx%d" % (na, nb))
    a, b = synth.random_boxes(3, 70, extent=300.0)[0], synth.random_boxes(4, 2100, extent=300.0)[0]
    ops.iou_match(dev(a), dev(b), 0.2, row_offsets=dev(np.array([0, 9, 9, 70], np.int32)), col_offsets=dev(np.array([0, 1031, 2099, 2100], np.int32)),
                  max_cols=1068)
    g.check("iou_match batched")


def test_image_kernels_stay_inside_their_outputs(guarded):
    ops, g = guarded
    img = dev(synth.shelf_image(5, 300, 410))
    table = torch.tensor([[img.data_ptr(), 300, 410, img.stride(0), img.stride(1)]], dtype=torch.int64).cuda()
    boxes = dev(np.array([[3, 4, 200, 280], [0, 0, 410, 300], [100, 100, 101, 250], [50.5, 60.5, 90.2, 75.9], [7, 7, 7, 7]], np.float32))
    idx = torch.zeros(5, dtype=torch.int32, device="cuda")
    for S in (256, 224, 64, 37, 30):                       # band kernel (S % 4 == 0) and direct kernel
        for dtype, cl in ((torch.float32, False), (torch.float32, True), (torch.bfloat16, True), (torch.float16, True)):
            ops.crop_resize(table, 1, boxes, idx, S, True, 0.5, None, dtype, cl, return_valid=True)
            g.check("crop_resize S=%d %s cl=%s" % (S, dtype, cl))
        ops.crop_resize(table, 1, boxes, idx, S, False, 0.5, (2.0, -1.0, 2.0, -1.0, 2.0, -1.0))
        g.check("crop_resize plain bilinear S=%d" % S)
    gen = torch.Generator(device="cuda").manual_seed(1)
    for shape in ((3, 64, 10, 14), (1, 8, 7, 9), (2, 512, 32, 32)):   # odd H / W: floor mode drops the last row / column
        for dtype in (torch.float32, torch.bfloat16, torch.float16):
            x = torch.randn(shape, device="cuda", generator=gen).to(dtype).contiguous(memory_format=torch.channels_last)
            if ops.maxpool2x2_supported(x):
                ops.maxpool2x2(x)
                g.check("maxpool2x2 %s %s" % (shape, dtype))
    for B, C, hw, cl in ((3, 512, 16, False), (3, 512, 16, True), (1, 24, 5, False), (5, 1024, 8, True)):
        f4 = torch.rand((B, C, 2 * hw, 2 * hw), device="cuda", generator=gen)
        f5 = torch.rand((B, C, hw, hw), device="cuda", generator=gen)
        if cl:
            f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
        ops.region_pool_embed([f4, f5])
        g.check("region_pool_embed MAC B=%d C=%d cl=%s" % (B, C, cl))
        ops.region_pool_embed([f4, f5], regions=[ops.rmac_regions(2 * hw, 2 * hw, 3), ops.rmac_regions(hw, hw, 3)], per_region_l2=True)
        g.check("region_pool_embed R-MAC B=%d C=%d cl=%s" % (B, C, cl))
        from cvpce_b200 import _lib
        with _lib.tuning(rmac_kernel=2):  # the persistent TMA-staged kernel (automatic only from B >= SMs/2)
            ops.region_pool_embed([f4, f5], regions=[ops.rmac_regions(2 * hw, 2 * hw, 3), ops.rmac_regions(hw, hw, 3)], per_region_l2=True)
        g.check("region_pool_embed R-MAC staged kernel B=%d C=%d cl=%s" % (B, C, cl))


def test_search_kernels_stay_inside_their_outputs(guarded):
    ops, g = guarded
    gen = torch.Generator(device="cuda").manual_seed(2)
    for Q, G, D, k in ((1, 1, 64, 1), (5, 77, 64, 3), (129, 300, 128, 5), (257, 1000, 1024, 5), (130, 40, 512, 32), (700, 5000, 512, 20)):
        q = torch.randn((Q, D), device="cuda", generator=gen)
        gal = torch.randn((G, D), device="cuda", generator=gen)
        qb, qi = ops.rows_prepare(q)
        gb, gi = ops.rows_prepare(gal)
        g.check("rows_prepare %d/%d x %d" % (Q, G, D))
        kk = min(k, 32)
        ke = min(k, 24)                                       # the exact re-score keeps k + 8 <= 32 candidates per row
        ops.sim_topk(qb, gb, ke, 7, q, gal, qi, gi)
        g.check("sim_topk exact Q=%d G=%d D=%d k=%d" % (Q, G, D, ke))
        if kk > 24:
            with pytest.raises(RuntimeError, match="k \\+ 8"):
                ops.sim_topk(qb, gb, kk, 7, q, gal, qi, gi)
        ops.sim_topk(qb, gb, kk, 0)
        g.check("sim_topk bf16 Q=%d G=%d D=%d k=%d" % (Q, G, D, kk))
        ops.sim_topk_f32(q, gal, min(k, G), 3)
        g.check("sim_topk_f32 Q=%d G=%d D=%d" % (Q, G, D))
    s = torch.randn((3, 41, 5), device="cuda", generator=gen).sort(dim=2, descending=True).values
    i = torch.randint(0, 10_000, (3, 41, 5), device="cuda", generator=gen)
    ops.topk_merge(s, i)
    g.check("topk_merge")
    x = torch.randn((7, 100), device="cuda", generator=gen)
    ops.l2norm_rows_(x)


def test_streaming_mac_stays_inside_an_exactly_sized_workspace(guarded):
    """cvpce_region_pool_embed_ws with a workspace of EXACTLY cvpce_region_pool_workspace_bytes between guard zones (the Python
    wrapper's cached workspace is rounded up, which would hide an overrun of the per-segment rows): batch sizes from one image
    (fewer chunks than CTAs) to images straddling many CTA ranges, other grid sizes, ragged pixel counts."""
    import ctypes
    from cvpce_b200 import _lib
    ops, g = guarded
    lib = _lib.load()
    gen = torch.Generator(device="cuda").manual_seed(4)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for B, shapes in ((1, [(512, 16, 16), (512, 16, 16)]), (2, [(1024, 3, 3)]), (37, [(512, 9, 7), (256, 5, 3), (1024, 2, 2)]),
                      (300, [(512, 16, 16), (512, 8, 8)]), (1000, [(64, 5, 5)])):
        feats = [torch.randn((B, c, h, w), device="cuda", generator=gen).contiguous(memory_format=torch.channels_last) for c, h, w in shapes]
        maps = (_lib.FeatureMap * len(feats))()
        for i, f in enumerate(feats):
            maps[i] = _lib.FeatureMap(f.data_ptr(), f.shape[1], f.shape[2], f.shape[3], f.stride(0), f.stride(1), f.stride(2), f.stride(3))
        want = torch.cat([f.amax(dim=(2, 3)) for f in feats], dim=1)
        for per_sm in (0, 1, 5):
            with _lib.tuning(mac_ctas_per_sm=per_sm):
                need = int(lib.cvpce_region_pool_workspace_bytes(maps, len(feats), 0, B))
                assert need > 0
                ws = g.zeros((need,), dtype=torch.uint8, device="cuda")
                out = g.empty((B, want.shape[1]), dtype=torch.float32, device="cuda")
                for _ in range(2):                            # the counters are left at zero: a second launch on the same workspace
                    rc = lib.cvpce_region_pool_embed_ws(maps, len(feats), 0, B, None, None, 0, 0, 1e-8, ctypes.c_void_p(out.data_ptr()),
                                                        out.stride(0), ctypes.c_void_p(ws.data_ptr()), need, stream)
                    assert rc == 0, _lib.last_error()
                    torch.cuda.synchronize()
                    assert torch.equal(out, want), (B, shapes, per_sm)
                assert int(ws[:4 * B].view(torch.int32).abs().sum()) == 0
                g.check("streaming MAC B=%d %s ctas/sm=%d" % (B, shapes, per_sm))
                rc = lib.cvpce_region_pool_embed_ws(maps, len(feats), 0, B, None, None, 0, 0, 1e-8, ctypes.c_void_p(out.data_ptr()),
                                                    out.stride(0), ctypes.c_void_p(ws.data_ptr()), need - 1, stream)
                assert rc == 1 and "workspace" in _lib.last_error()      # one byte short: refused, nothing launched


def test_tcgen05_search_stays_inside_the_smallest_workspace_it_accepts(guarded):
    """cvpce_sim_topk_workspace_bytes is a generous upper bound (any tuning override must fit), which would hide a candidate list
    written past the slots the launch plan really uses.  The entry point refuses a workspace smaller than its plan needs, so the
    smallest accepted size is found by bisection and the search then runs in exactly that much memory between guard zones:
    linear item ranges (several slots per row, a varying number per query block), split grids, both list forms."""
    import ctypes
    from cvpce_b200 import _lib
    ops, g = guarded
    gen = torch.Generator(device="cuda").manual_seed(6)
    for Q, G, D, k, knobs in ((1300, 5000, 1024, 5, {}), (700, 3000, 256, 3, {"sim_plan": 2}), (2000, 3000, 512, 5, {}), (257, 40000, 512, 5, {}),
                              (900, 9000, 128, 5, {"sim_split": 5}), (1500, 6000, 256, 20, {})):
        q = torch.randn((Q, D), device="cuda", generator=gen)
        gal = torch.randn((G, D), device="cuda", generator=gen)
        qb, qi = ops.rows_prepare(q)
        gb, gi = ops.rows_prepare(gal)
        g.check("rows_prepare")
        lib = _lib.load()
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        P = lambda t: ctypes.c_void_p(t.data_ptr())

        def call(ws_tensor, nbytes, out_s, out_i):
            return lib.cvpce_sim_topk(P(qb), P(gb), Q, G, D, k, 3, P(q), P(gal), P(qi), P(gi), P(out_s), P(out_i), P(ws_tensor), nbytes, stream)

        with _lib.tuning(**knobs):
            hi = ops.sim_topk_workspace_bytes(Q, G, D, k)
            big = torch.empty(hi, dtype=torch.uint8, device="cuda")
            ws_s, ws_i = torch.empty((Q, k), device="cuda"), torch.empty((Q, k), dtype=torch.int64, device="cuda")
            assert call(big, hi, ws_s, ws_i) == 0, _lib.last_error()
            lo = 0                                             # invariant: lo bytes are refused, hi bytes are accepted
            while hi - lo > 1:
                mid = (lo + hi) // 2
                rc = call(big, mid, ws_s, ws_i)
                assert rc in (0, 5), _lib.last_error()         # CVPCE_ERR_WORKSPACE: refused before any launch
                if rc == 0:
                    hi = mid
                else:
                    lo = mid
            torch.cuda.synchronize()
            ws = g.empty((hi,), dtype=torch.uint8, device="cuda")
            got_s, got_i = g.empty((Q, k), dtype=torch.float32, device="cuda"), g.empty((Q, k), dtype=torch.int64, device="cuda")
            assert call(ws, hi, got_s, got_i) == 0, _lib.last_error()
            torch.cuda.synchronize()
            assert torch.equal(got_s, ws_s) and torch.equal(got_i, ws_i)
            g.check("sim_topk in its smallest workspace (%d bytes) Q=%d G=%d D=%d k=%d %s" % (hi, Q, G, D, k, knobs))


def test_guard_zones_do_detect_an_overrun(guarded):
    _, g = guarded
    v = g.empty((5, 3), dtype=torch.float32, device="cuda")
    buf, nbytes = g.allocs[-1]
    assert nbytes == 60 and v.shape == (5, 3) and v.data_ptr() == buf.data_ptr() + GUARD
    buf[GUARD + nbytes] = 0                      # one byte past the end
    with pytest.raises(AssertionError, match="PAST"):
        g.check("self-test")
