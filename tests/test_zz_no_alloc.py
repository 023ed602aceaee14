"""capdec_gemm_static: the allocation-free variant of capdec_debug_gemm (scratch in a file-scope device array).
Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it creates its own engine handles and assumes nothing about the static array's contents."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-4  # the tolerance test_gemm uses for the same computation


def _engine(precision, dev):
    from image_caption_b200 import CaptionDecodeEngine
    from oracle import torch_port
    dec = torch_port.build_decoder(1, num_layers=1, max_len=24).to(dev)
    return CaptionDecodeEngine(dec, precision=precision, max_images=1, max_mem=1, max_len=4)


def _static_gemm(eng, a, w, b, relu):
    from image_caption_b200 import _lib
    c = torch.empty((a.shape[0], w.shape[0]), dtype=torch.float32, device=a.device)  # caller-owned output
    rc = eng._lib.capdec_gemm_static(eng._h, a.data_ptr(), w.data_ptr(), b.data_ptr(), c.data_ptr(), a.shape[0], w.shape[0],
                                     a.shape[1], int(relu), torch.cuda.current_stream(a.device).cuda_stream)
    _lib.check(eng._h, rc)
    return c


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_gemm_static_matches_debug_gemm_twice(precision):
    dev = torch.device("cuda:0")
    eng = _engine(precision, dev)
    g = torch.Generator().manual_seed(7)
    # two shapes in a row on the shared scratch: the second with smaller, different operands (stale contents of the first
    # call lie beyond and under them) and the other activation
    for m, n, k, relu in [(300, 1536, 512, False), (37, 512, 2048, True)]:
        a = torch.randn(m, k, generator=g).to(dev)
        w = (torch.randn(n, k, generator=g) * 0.05).to(dev)
        b = torch.randn(n, generator=g).to(dev)
        want = eng.debug_gemm(a, w, b, relu=relu)  # the existing entry point (cudaMalloc'd scratch)
        got = _static_gemm(eng, a, w, b, relu)     # (the first use of a kernel variant may load its code lazily)
        out = torch.empty_like(got)
        torch.cuda.synchronize(dev)
        free0 = torch.cuda.mem_get_info(dev)[0]
        rc = eng._lib.capdec_gemm_static(eng._h, a.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(), m, n, k, int(relu),
                                         torch.cuda.current_stream(dev).cuda_stream)
        torch.cuda.synchronize(dev)
        free1 = torch.cuda.mem_get_info(dev)[0]
        assert rc == 0 and free1 == free0, "capdec_gemm_static changed the device's free memory"
        assert torch.equal(out, got)  # same inputs, same scratch: same bits
        assert got.shape == want.shape
        assert (got - want).abs().max().item() <= TOL
        ref = (a.bfloat16().double() if precision == "bf16" else a.double()) @ \
              (w.bfloat16().double() if precision == "bf16" else w.double()).T + b.double()
        ref = ref.clamp_min(0) if relu else ref
        assert (got.double() - ref).abs().max().item() <= TOL
    eng.close()


def test_gemm_static_rejects_sizes_beyond_the_scratch():
    from image_caption_b200 import _lib
    dev = torch.device("cuda:0")
    eng = _engine("bf16", dev)
    m, n, k = 2048, 2048, 512  # (m + n) * k = 2 Mi elements: exactly the limit
    a = torch.zeros(m, k, device=dev)
    w = torch.zeros(n, k, device=dev)
    b = torch.zeros(n, device=dev)
    assert float(_static_gemm(eng, a, w, b, False).abs().max()) == 0.0
    a2 = torch.zeros(m + 1, k, device=dev)  # one row more
    with pytest.raises(_lib.CapdecError, match="static scratch") as ei:
        _static_gemm(eng, a2, w, b, False)
    assert ei.value.code == -4  # CAPDEC_ERR_CAPACITY
    eng.close()
