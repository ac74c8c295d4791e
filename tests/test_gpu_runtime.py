"""GPU: run-time behaviour of the library calls -- bit-identical results from run to run (the kernels use atomics only where
the final order does not depend on them), correct results on a non-default stream, and concurrent calls on two streams."""
import numpy as np
import pytest
import torch

from cvpce_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(cuda):
    from cvpce_b200 import _lib, ops as _ops
    assert _lib.load().cvpce_device_check() == 0, _lib.last_error()
    return _ops


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _workload(ops):
    """One call of every kernel family on fixed inputs -> list of result tensors."""
    out = []
    b, s = synth.random_boxes(5, 3000, extent=400.0, tie_every=9)
    seg = dev(np.array([0, 1000, 1000, 3000], np.int32))
    keep, counts, kb, ks = ops.nms_batched(dev(b), dev(s), seg, 2000, 0.5, want_boxes=True, want_scores=True)
    c = counts.cpu().tolist()
    out += [counts] + [keep[o:o + n] for o, n in zip((0, 1000, 1000), c)] + [kb[o:o + n] for o, n in zip((0, 1000, 1000), c)]
    name = list(synth.HEAD_CASES)[0]
    logits, regression, anchors, sizes, C, topk = synth.head_case(name)
    bx, sc, lb, segc, lvl, stride = ops.select_candidates([dev(x) for x in logits], [dev(x) for x in regression], [dev(a) for a in anchors],
                                                          sizes, 0.05, topk)
    n = segc.cpu().tolist()
    out += [segc, lvl] + [bx[i * stride:i * stride + m] for i, m in enumerate(n)] + [sc[i * stride:i * stride + m] for i, m in enumerate(n)]
    t, p = synth.detection_eval_case(8, 150, 400)
    out += list(ops.iou_match(dev(p), dev(t), 0.5)) + [ops.box_iou(dev(p), dev(t))]
    img = dev(synth.shelf_image(6, 400, 520))
    boxes = dev(np.array([[10, 20, 300, 390], [0, 0, 520, 400], [100.5, 50.2, 180.9, 300.1]], np.float32))
    out.append(ops.crop_resize(ops.make_image_table([img]), 1, boxes, torch.zeros(3, dtype=torch.int32, device="cuda"), 256))
    gen = torch.Generator(device="cuda").manual_seed(4)
    f4 = torch.rand((9, 512, 16, 16), device="cuda", generator=gen).contiguous(memory_format=torch.channels_last)
    f5 = torch.rand((9, 512, 16, 16), device="cuda", generator=gen).contiguous(memory_format=torch.channels_last)
    emb = ops.region_pool_embed([f4, f5])
    out += [emb, ops.maxpool2x2(f4)]
    gal = torch.nn.functional.normalize(torch.rand((3000, 1024), device="cuda", generator=gen), dim=1)
    q = torch.cat([emb, gal[:300] + 0.01 * torch.rand((300, 1024), device="cuda", generator=gen)])
    qb, qi = ops.rows_prepare(q)
    gb, gi = ops.rows_prepare(gal)
    out += list(ops.sim_topk(qb, gb, 5, 0, q, gal, qi, gi)) + list(ops.sim_topk(qb, gb, 5, 0)) + [qb, qi]
    return out


def _same(a, b):
    assert len(a) == len(b)
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.shape == y.shape and x.dtype == y.dtype, i
        xv, yv = (x.view(torch.int16), y.view(torch.int16)) if x.dtype in (torch.bfloat16, torch.float16) else (x, y)
        if x.dtype in (torch.float32,):
            xv, yv = x.view(torch.int32), y.view(torch.int32)      # bit pattern, so that NaN == NaN
        assert torch.equal(xv, yv), "result %d differs" % i


def test_results_are_bit_identical_from_run_to_run_and_stream_to_stream(ops):
    ref = _workload(ops)
    torch.cuda.synchronize()
    for _ in range(2):
        _same(_workload(ops), ref)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        got = _workload(ops)
    side.synchronize()
    _same(got, ref)


def test_concurrent_calls_on_two_streams(ops):
    gen = torch.Generator(device="cuda").manual_seed(7)
    data = []
    for i in range(2):
        gal = torch.nn.functional.normalize(torch.randn((20000, 512), device="cuda", generator=gen), dim=1)
        q = torch.nn.functional.normalize(torch.randn((4096, 512), device="cuda", generator=gen), dim=1)
        b, s = synth.random_boxes(40 + i, 4000, extent=500.0)
        data.append((q, gal, dev(b), dev(s), dev(np.array([0, 4000], np.int32))))

    def run(d):
        q, gal, b, s, seg = d
        qb, qi = ops.rows_prepare(q)
        gb, gi = ops.rows_prepare(gal)
        sim, idx = ops.sim_topk(qb, gb, 5, 0, q, gal, qi, gi)
        keep, counts, _, _ = ops.nms_batched(b, s, seg, 4000, 0.5)
        return sim, idx, keep, counts

    want = [run(d) for d in data]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    got = [None, None]
    for rep in range(3):
        for i in (0, 1):
            with torch.cuda.stream(streams[i]):
                got[i] = run(data[i])
        for st in streams:
            st.synchronize()
        for i in (0, 1):
            n = int(want[i][3][0])
            assert int(got[i][3][0]) == n and torch.equal(got[i][2][:n], want[i][2][:n])
            assert torch.equal(got[i][0], want[i][0]) and torch.equal(got[i][1], want[i][1])
