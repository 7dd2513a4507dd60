"""GPU tests of the deformation network's MLP half: the tcgen05 linear layer against float64 torch."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fdls():
    import importlib
    return importlib.import_module("4dlangsplat_b200")


@pytest.mark.parametrize("rows,N,K,relu", [(128, 128, 128, 0), (300, 128, 48, 1), (1000, 8, 128, 0), (77, 64, 16, 1),
                                           (4096, 128, 128, 1)])
def test_umma_linear_matches_float64(fdls, rows, N, K, relu):
    """Y = act(X W^T + b) on the tensor cores with bf16 x 3 operand splits reproduces the fp32 result: every element
    within 2e-6 of the row's scale of the float64 product (fp32 rounding of the 128-term sums is ~1e-6 itself)."""
    lib = fdls._lib.load()
    g = torch.Generator().manual_seed(rows * 7 + N + K)
    X = (torch.randn(rows, K, generator=g) * 2.0).cuda()
    W = (torch.randn(N, K, generator=g) * 0.3).cuda()
    b = torch.randn(N, generator=g).cuda()
    Y = torch.full((rows, N), float("nan"), device="cuda")
    rc = lib.fdls_umma_linear(rows, N, K, X.data_ptr(), W.data_ptr(), b.data_ptr(), Y.data_ptr(), relu,
                              C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    ref = X.double() @ W.double().t() + b.double()
    scale = (X.double().abs() @ W.double().abs().t() + b.double().abs())  # magnitude of the summed terms
    if relu:
        ref = ref.clamp_min(0)
    err = (Y.double() - ref).abs()
    assert torch.isfinite(Y).all()
    assert bool((err <= 2e-6 * scale + 1e-30).all()), "max err/scale = %.3g" % float((err / scale).max())
