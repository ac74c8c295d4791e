"""CPU: the C-ABI library builds for sm_100a, loads, exports every symbol include/cvpce_b200.h declares, and refuses to
run without a B200 (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT
from cvpce_b200 import _lib, build, ops


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cvpce_b200.h")).read()
    return sorted(set(re.findall(r"CVPCE_API\s+[\w\s\*]+?\b(cvpce_\w+)\s*\(", text)))


def test_header_symbols_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), "libcvpce_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "no ctypes signature for %s" % n
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_struct_layout(lib):
    assert lib.cvpce_version() == 100
    assert ctypes.sizeof(_lib.FeatureMap) == 56  # void* + 3*int32 (+pad) + 4*int64


def test_binary_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _lib.SO_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_fails_loudly_without_gpu(lib):
    assert lib.cvpce_device_check() == 4  # CVPCE_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.cvpce_last_error()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.nms(torch.zeros(3, 4), torch.zeros(3), 0.5)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.sim_topk_f32(torch.zeros(2, 8), torch.zeros(4, 8), 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.region_pool_embed([torch.zeros(1, 2, 3, 3)])


def test_argument_validation_on_host(lib):
    # validation happens before any device work, so it can be exercised without a GPU
    assert lib.cvpce_nms_batched(None, None, None, -1, 0, 0.5, 0.0, 0, None, None, None, None, None, None) == 1
    assert lib.cvpce_nms_batched(None, None, None, 0, 0, 0.5, 0.0, 0, None, None, None, None, None, None) == 0  # empty batch ok
    assert lib.cvpce_crop_resize(None, 0, None, None, 0, 256, 1, 0.5, None, 0, 0, None, None, None) == 0        # zero crops ok
    assert lib.cvpce_crop_resize(None, 0, None, None, 4, 0, 1, 0.5, None, 0, 0, None, None, None) == 1
    assert lib.cvpce_sim_topk_f32(None, None, 0, 10, 8, 1, 0, 1, 1e-8, None, None, None) == 0                  # zero queries ok
    assert lib.cvpce_sim_topk_f32(None, None, 4, 10, 8, 100, 0, 1, 1e-8, None, None, None) == 2                # k too large
    assert lib.cvpce_topk_merge(None, None, 0, 4, 1, None, None, None) == 1


def test_workspace_planners_on_host(lib):
    """The workspace-size entry points are pure host arithmetic (the launch plans): exercised here without a GPU."""
    from cvpce_b200 import _lib
    # tcgen05 search: grows with Q, covers the linear-range plan's slots at the C3 shape and the split plan of the 1M-gallery shard
    c3 = lib.cvpce_sim_topk_workspace_bytes(6400, 8192, 1024, 5)
    c5 = lib.cvpce_sim_topk_workspace_bytes(100000, 125000, 512, 5)
    assert c3 >= 6400 * 16 * 8 and c5 >= 100000 * 16 * 8 and c5 > c3
    assert lib.cvpce_sim_topk_workspace_bytes(0, 10, 64, 1) == 0 and lib.cvpce_sim_topk_workspace_bytes(10, 10, 64, 33) == 0
    # MAC streaming kernel: only dense channels-last maps have a streaming plan; the size grows with the batch
    def fmap(C, H, W, nhwc=True, ptr=0x10000):
        return _lib.FeatureMap(ptr, C, H, W, C * H * W, 1 if nhwc else H * W, W * C if nhwc else W, C if nhwc else 1)
    maps = (_lib.FeatureMap * 2)(fmap(512, 16, 16), fmap(512, 16, 16))
    w256 = lib.cvpce_region_pool_workspace_bytes(maps, 2, 0, 256)
    w512 = lib.cvpce_region_pool_workspace_bytes(maps, 2, 0, 512)
    assert 65536 < w256 < w512 <= 8 * 1024 * 1024            # both fit the static scratch of cvpce_region_pool_embed_static
    assert lib.cvpce_region_pool_workspace_bytes((_lib.FeatureMap * 1)(fmap(512, 16, 16, nhwc=False)), 1, 0, 256) == 0
    assert lib.cvpce_region_pool_workspace_bytes((_lib.FeatureMap * 1)(fmap(512, 16, 16, ptr=0x10004)), 1, 0, 256) == 0   # misaligned
    assert lib.cvpce_region_pool_workspace_bytes((_lib.FeatureMap * 1)(fmap(6, 3, 3)), 1, 0, 4) == 0                     # C not a vector multiple
    assert lib.cvpce_region_pool_workspace_bytes(maps, 2, 0, 0) == 0
    # the allocation-free entry point validates on the host before touching the device
    assert lib.cvpce_region_pool_embed_static(None, 1, 0, 4, 1, 1e-8, None, 0, None) == 1
    assert lib.cvpce_region_pool_embed_static(maps, 2, 0, 0, 1, 1e-8, None, 0, None) == 0
    assert lib.cvpce_region_pool_embed_static((_lib.FeatureMap * 1)(fmap(512, 16, 16, nhwc=False)), 1, 0, 4, 1, 1e-8, None, 0, None) == 2
    assert lib.cvpce_region_pool_embed_static(maps, 2, 0, 100000, 1, 1e-8, None, 0, None) == 2 and "static array" in _lib.last_error()


def test_threshold_rounding_helper():
    import numpy as np
    assert ops.threshold_as_cpu_op(0.5) == 0.5
    f = ops.threshold_as_cpu_op(1 / 3)
    assert f <= 1 / 3 and float(np.nextafter(np.float32(f), np.float32(1))) > 1 / 3
    f = ops.threshold_as_cpu_op(0.7)
    assert f <= 0.7 and float(np.nextafter(np.float32(f), np.float32(1))) > 0.7


def test_rmac_regions_host_helper_matches_oracle(lib):
    import oracle
    for H, W, L in ((16, 16, 3), (14, 14, 3), (32, 32, 2), (14, 20, 3), (7, 7, 1)):
        assert ops.rmac_regions(H, W, L) == [tuple(r) for r in oracle.rmac_regions(H, W, L).tolist()]


def test_affine_fold_matches_reference_composition():
    from cvpce_b200 import classification
    x = torch.rand(2, 3, 5, 5)
    a = torch.tensor(classification.tanh_normalize_affine()).view(3, 2)
    got = x * a[:, 0].view(1, 3, 1, 1) + a[:, 1].view(1, 3, 1, 1)
    want = ((x * 2 - 1) - torch.tensor(classification.NORM_MEAN).view(1, 3, 1, 1)) / torch.tensor(classification.NORM_STD).view(1, 3, 1, 1)
    assert torch.allclose(got, want, atol=1e-6)


def test_header_is_plain_c_and_a_c_program_can_bind(lib, tmp_path):
    """The boundary is a C ABI: the header must compile as C99 (no C++-isms, no torch types) and a C host must be able to
    dlopen the library and call through it -- here the calls that need no GPU (version, argument validation, last error)."""
    import os
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "host.c"
    src.write_text(r'''
#include <dlfcn.h>
#include <stdio.h>
#include <string.h>
#include "cvpce_b200.h"
typedef int (*version_fn)(void);
typedef const char* (*err_fn)(void);
typedef int (*iou_fn)(const float*, int, const float*, int, float*, void*);
typedef int (*match_fn)(const float*, const int64_t*, const int32_t*, const float*, const int64_t*, const int32_t*, int, int, float,
                        uint8_t*, uint8_t*, void*);
int main(int argc, char** argv) {
  void* h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
  if (!h) { fprintf(stderr, "%s\n", dlerror()); return 2; }
  version_fn version = (version_fn)dlsym(h, "cvpce_version");
  err_fn last_error = (err_fn)dlsym(h, "cvpce_last_error");
  iou_fn box_iou = (iou_fn)dlsym(h, "cvpce_box_iou");
  match_fn iou_match = (match_fn)dlsym(h, "cvpce_iou_match");
  if (!version || !last_error || !box_iou || !iou_match) return 3;
  if (version() != CVPCE_B200_VERSION) return 4;
  if (box_iou(NULL, 0, NULL, 0, NULL, NULL) != CVPCE_OK) return 5;                 /* empty input: valid, nothing launched */
  if (box_iou(NULL, 3, NULL, 2, NULL, NULL) != CVPCE_ERR_INVALID_ARG) return 6;    /* null pointers: refused on the host */
  if (strstr(last_error(), "box_iou") == NULL) return 7;
  if (iou_match(NULL, NULL, NULL, NULL, NULL, NULL, 1, CVPCE_MATCH_MAX_COLS + 1, 0.5f, NULL, NULL, NULL) == CVPCE_OK) return 8;
  printf("ok %d\n", version());
  return 0;
}
''')
    exe = tmp_path / "host"
    hdr = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c",
                          os.path.join(root, "include", "cvpce_b200.h")], capture_output=True, text=True)
    assert hdr.returncode == 0, hdr.stderr
    # (no -pedantic for the host: ISO C has no object-pointer -> function-pointer conversion, which dlsym needs)
    cc = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(root, "include"), str(src), "-o", str(exe), "-ldl"],
                        capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr
    run = subprocess.run([str(exe), _lib.SO_PATH], capture_output=True, text=True)
    assert run.returncode == 0 and run.stdout.strip() == "ok 100", (run.returncode, run.stdout, run.stderr)
