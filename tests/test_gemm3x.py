"""The chained 3xTF32 tcgen05 GEMM (csrc/nsb_gemm.cu) against fp64 torch matmuls: both operand
addressings, ragged M / N / K, every epilogue, split-K with atomics and the column sums."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _run(A_mk, B_nk, a_rc, b_rc, epi=0, bias=None, mask=None, n_split=1, colsum=False, C0=None, packed=False):
    """A_mk [M,K], B_nk [N,K] logical operands; stored transposed when *_rc.  packed: B goes through the
    pre-split operand images + bulk copies (the route of the iMAP* chains' weight matrices)."""
    from nice_slam_annotation_b200._lib import check, lib
    M, K = A_mk.shape
    N = B_nk.shape[0]
    A = (A_mk.t().contiguous() if a_rc else A_mk.contiguous()).cuda()
    B = (B_nk.t().contiguous() if b_rc else B_nk.contiguous()).cuda()
    lda = M if a_rc else K
    ldb = N if b_rc else K
    Cm = torch.zeros(M, N, device="cuda") if C0 is None else C0.clone().cuda()
    cs = torch.zeros(M, device="cuda") if colsum else None
    bias_d = None if bias is None else bias.cuda()
    mask_d = None if mask is None else mask.contiguous().cuda()
    scratch = torch.empty(lib().nsb_gemm3x_pack_floats(N, K), device="cuda") if packed else None
    check(lib().nsb_gemm3x(_p(A), lda, a_rc, _p(B), ldb, b_rc, _p(Cm), N, M, N, K, epi, _p(bias_d), _p(mask_d),
                           N, _p(cs), n_split, _p(scratch), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "gemm3x")
    torch.cuda.synchronize()
    return Cm.cpu(), (None if cs is None else cs.cpu())


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("a_rc", [0, 1])
@pytest.mark.parametrize("b_rc", [0, 1])
@pytest.mark.parametrize("M,N,K", [(128, 256, 256), (300, 256, 93), (1000, 96, 256), (77, 16, 64), (513, 4, 256),
                                   (4, 256, 1000), (256, 93, 777), (64, 64, 4), (40000, 256, 256), (260, 160, 600)])
def test_gemm_store(a_rc, b_rc, M, N, K, packed):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + a_rc * 2 + b_rc)
    A, B = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g)
    out, _ = _run(A, B, a_rc, b_rc, packed=packed)
    ref = A.double() @ B.double().t()
    assert _rel(out, ref) < 1e-5


def test_gemm_epilogues():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    g = torch.Generator().manual_seed(3)
    M, N, K = 700, 256, 256
    A, B = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g)
    bias, mask = torch.randn(N, generator=g), torch.randn(M, N, generator=g)
    ref = A.double() @ B.double().t()
    out, _ = _run(A, B, 0, 0, epi=1, bias=bias)
    assert _rel(out, torch.relu(ref + bias.double())) < 1e-5
    out, _ = _run(A, B, 0, 0, epi=2, bias=bias)
    assert _rel(out, ref + bias.double()) < 1e-5
    out, _ = _run(A, B, 0, 1, epi=3, mask=mask)
    assert _rel(out, ref * (mask > 0)) < 1e-5
    assert bool((out[mask <= 0] == 0).all())
    out, _ = _run(A, B, 0, 0, epi=1, bias=bias, packed=True)
    assert _rel(out, torch.relu(ref + bias.double())) < 1e-5
    out, _ = _run(A, B, 0, 1, epi=3, mask=mask, packed=True)
    assert _rel(out, ref * (mask > 0)) < 1e-5


@pytest.mark.parametrize("M,N", [(256, 93), (256, 256), (4, 256), (200, 96)])
@pytest.mark.parametrize("n_split", [1, 5, 37])
def test_gemm_split_k_atomic_colsum(n_split, M, N):
    """The weight-gradient shape: M, N = features, K = points, both operands row-contiguous (MN-major images)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    g = torch.Generator().manual_seed(n_split + M + N)
    K = 5000
    A, B = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g)
    C0 = torch.randn(M, N, generator=g)
    out, cs = _run(A, B, 1, 1, epi=4, n_split=n_split, colsum=True, C0=C0)
    ref = C0.double() + A.double() @ B.double().t()
    assert _rel(out, ref) < 2e-5
    assert _rel(cs, A.double().sum(1)) < 2e-5
