"""gsr_ssim_forward_noalloc: the allocation-free variant of gsr_ssim_forward (scratch in a static device array inside
the library; no counterpart in the reference).  Called through the
C-ABI and compared with the existing entry point on two calls in a row; the file sorts last on purpose."""
import ctypes as C

import pytest
import torch

from gaussian_splatting_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_TILES = 65536          # GSR_SSIM_NOALLOC_MAX_TILES in include/gsrast.h


def _params(img1, img2, crop=0):
    B, CH, H, W = img1.shape
    return _lib.GsrSsimParams(img1.data_ptr(), img2.data_ptr(), B, CH, H, W, 0.01 ** 2, 0.03 ** 2, crop)


def _existing(lib, img1, img2, crop, stream):
    B, CH, H, W = img1.shape
    maps = [torch.empty_like(img1) for _ in range(4)]
    mean = torch.empty(1, device=img1.device)
    partials = torch.empty(max(1, lib.gsr_ssim_partials(B, CH, H, W)), device=img1.device)
    p = _params(img1, img2, crop)
    _lib.check(lib.gsr_ssim_forward(C.byref(p), *[m.data_ptr() for m in maps], mean.data_ptr(), partials.data_ptr(), stream))
    return maps, mean


def _noalloc(lib, img1, img2, crop, stream, maps, mean):
    p = _params(img1, img2, crop)
    return lib.gsr_ssim_forward_noalloc(C.byref(p), *[m.data_ptr() for m in maps], mean.data_ptr(), stream)


@pytest.mark.parametrize("crop", [0, 5])
def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(crop):
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(11)
    # two different sizes back to back: the second call finds the first call's partial sums in the shared array
    shapes = [(1, 3, 270, 481), (2, 3, 97, 64)]
    # outputs are allocated up front: the entry point itself must not need any allocation
    cases = []
    for shp in shapes:
        a = torch.rand(*shp, device=dev, generator=g)
        b = (a + 0.1 * torch.randn(*shp, device=dev, generator=g)).clamp(0, 1)
        cases.append((a, b, [torch.full_like(a, float("nan")) for _ in range(4)], torch.full((1,), float("nan"), device=dev)))
    for rep in range(2):
        for a, b, maps, mean in cases:
            want_maps, want_mean = _existing(lib, a, b, crop, stream)
            before = torch.cuda.memory_allocated(dev)
            _lib.check(_noalloc(lib, a, b, crop, stream, maps, mean))
            assert torch.cuda.memory_allocated(dev) == before
            torch.cuda.synchronize()
            # same kernels, same reduction order: the results are identical (tolerance of the SSIM tests: 1e-6)
            assert abs(float(mean) - float(want_mean)) <= 1e-6, (rep, a.shape, float(mean), float(want_mean))
            for got, want in zip(maps, want_maps):
                assert torch.allclose(got, want, rtol=0, atol=1e-6)
            mean.fill_(float("nan"))


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    # 3 x 4800 x 4800 -> 3 * 150 * 150 = 67 500 tiles > 65 536; the call must fail before touching the images
    img = torch.zeros(1, 3, 4800, 4800, device=dev)
    mean = torch.full((1,), 7.0, device=dev)
    p = _params(img, img)
    rc = lib.gsr_ssim_forward_noalloc(C.byref(p), None, None, None, None, mean.data_ptr(), stream)
    assert rc == _lib.GSR_ERR_CAPACITY
    assert b"GSR_SSIM_NOALLOC_MAX_TILES" in lib.gsr_last_error()
    torch.cuda.synchronize()
    assert float(mean) == 7.0
    # mean_out is what needs the scratch: without it the call is refused
    rc = lib.gsr_ssim_forward_noalloc(C.byref(_params(img[:, :, :64, :64].contiguous(), img[:, :, :64, :64].contiguous())),
                                      None, None, None, None, None, stream)
    assert rc == _lib.GSR_ERR_INVALID_ARG
