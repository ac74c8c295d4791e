"""Allocation-free MAC pooling (cvpce_region_pool_embed_static, include/cvpce_b200.h): no counterpart in the reference, added
for callers that cannot call cudaMalloc.  Called through the C ABI (ctypes, raw pointers) and compared with the existing entry
point cvpce_region_pool_embed on the same inputs: maxima bit-equal, L2-normalised descriptors within the 2e-6 relative the MAC
parity tests use.  Two calls in a row on the shared static scratch (different inputs, different batch sizes), the size limit,
and the layouts the entry point rejects.  The file name makes pytest collect it last."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _maps(feats):
    from cvpce_b200 import _lib
    maps = (_lib.FeatureMap * len(feats))()
    for i, f in enumerate(feats):
        maps[i] = _lib.FeatureMap(f.data_ptr(), f.shape[1], f.shape[2], f.shape[3], f.stride(0), f.stride(1), f.stride(2), f.stride(3))
    return maps


def _call(lib, name, feats, final_l2, out, static):
    maps = _maps(feats)
    B = feats[0].shape[0]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    optr = ctypes.c_void_p(out.data_ptr())
    if static:
        return lib.cvpce_region_pool_embed_static(maps, len(feats), 0, B, 1 if final_l2 else 0, 1e-8, optr, out.stride(0), stream)
    return lib.cvpce_region_pool_embed(maps, len(feats), 0, B, None, None, 0, 1 if final_l2 else 0, 1e-8, optr, out.stride(0), stream)


def test_static_scratch_mac_equals_existing_entry_point_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from cvpce_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(11)
    cases = [(37, [(512, 9, 7), (256, 5, 3)]), (256, [(512, 16, 16), (512, 16, 16)]), (5, [(1024, 4, 4)])]
    for final_l2 in (False, True):
        for B, shapes in cases:                               # consecutive calls share the static array: B and shapes change
            feats = [torch.randn((B, c, h, w), generator=gen, device=dev).contiguous(memory_format=torch.channels_last) for c, h, w in shapes]
            D = sum(c for c, _, _ in shapes)
            want = torch.full((B, D), 7.0, device=dev)
            assert _call(lib, "embed", feats, final_l2, want, static=False) == 0, _lib.last_error()
            for rep in range(2):                              # two calls in a row on the same inputs
                got = torch.full((B, D), -3.0, device=dev)
                assert _call(lib, "static", feats, final_l2, got, static=True) == 0, _lib.last_error()
                torch.cuda.synchronize()
                if final_l2:
                    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=2e-6, atol=1e-7)
                else:
                    assert torch.equal(got, want), (B, shapes, rep)
    # back-to-back calls WITHOUT a synchronisation in between (stream-ordered: the second reuses the scratch of the first)
    a = [torch.randn((64, 512, 8, 8), generator=gen, device=dev).contiguous(memory_format=torch.channels_last)]
    b = [torch.randn((48, 512, 8, 8), generator=gen, device=dev).contiguous(memory_format=torch.channels_last)]
    oa, ob = torch.empty((64, 512), device=dev), torch.empty((48, 512), device=dev)
    assert _call(lib, "static", a, False, oa, static=True) == 0 and _call(lib, "static", b, False, ob, static=True) == 0
    torch.cuda.synchronize()
    assert torch.equal(oa, a[0].amax(dim=(2, 3))) and torch.equal(ob, b[0].amax(dim=(2, 3)))


def test_static_scratch_mac_rejects_what_it_cannot_serve():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from cvpce_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    # beyond the static array: the needed scratch is computed from shapes alone, so a small real buffer suffices to ask
    f = torch.zeros((4, 512, 16, 16), device=dev).contiguous(memory_format=torch.channels_last)
    maps = _maps([f])
    out = torch.empty((4, 512), device=dev)
    B_big = 20000
    need = lib.cvpce_region_pool_workspace_bytes(maps, 1, 0, B_big)
    assert need > 8 * 1024 * 1024
    rc = lib.cvpce_region_pool_embed_static(maps, 1, 0, B_big, 1, 1e-8, ctypes.c_void_p(out.data_ptr()), 512, None)
    assert rc == 2 and "static array" in _lib.last_error()                        # CVPCE_ERR_UNSUPPORTED, limit in the message
    nchw = torch.zeros((4, 512, 16, 16), device=dev)                              # not channels-last
    rc = lib.cvpce_region_pool_embed_static(_maps([nchw]), 1, 0, 4, 1, 1e-8, ctypes.c_void_p(out.data_ptr()), 512, None)
    assert rc == 2 and "channels-last" in _lib.last_error()
    assert lib.cvpce_region_pool_embed_static(maps, 1, 0, 0, 1, 1e-8, None, 512, None) == 0      # empty batch: nothing to do
