"""Known-answer test of the tcgen05 building blocks (csrc/nsb_tc.cuh): shared-memory matrix
descriptors, the instruction descriptor, TMEM-resident A operands and the 3xTF32 split."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
@pytest.mark.parametrize("K,N", [(32, 32), (96, 64), (8, 16), (64, 32)])
def test_tf32_gemm(mode, K, N):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200._lib import check, lib
    g = torch.Generator().manual_seed(K * 100 + N + mode)
    A = torch.randn(128, K, generator=g).cuda()
    W = torch.randn(N, K, generator=g).cuda()
    D = torch.zeros(128, N, device="cuda")
    check(lib().nsb_tc_selftest(C.c_void_p(A.data_ptr()), C.c_void_p(W.data_ptr()), C.c_void_p(D.data_ptr()), K, N,
                                mode, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "tc_selftest")
    torch.cuda.synchronize()
    exact = A.double() @ W.double().t()
    if mode & 2:       # 3xTF32: fp32-class accuracy
        err = (D.double() - exact).abs().max().item() / exact.abs().max().item()
        assert err < 2e-6, err
    else:              # single pass: equals the product of the tf32-truncated operands
        ref = _tf32(A).double() @ _tf32(W).double().t()
        err = (D.double() - ref).abs().max().item() / ref.abs().max().item()
        assert err < 1e-5, err
