"""GPU: the tcgen05 TF32 GEMM building block and the native DeformNetwork path
against plain PyTorch fp32 (tolerances stated per test: TF32 keeps 10 mantissa
bits of each operand, accumulation is fp32)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def tf32_trunc(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)


def run_gemm(native_lib, A, B, bias=None, relu=False, Z=None, D=None):
    M, K = A.shape
    N = B.shape[0]
    if D is None:
        D = torch.full((M, ((N + 3) // 4) * 4), float("nan"), device=A.device)[:, :N]
    rc = native_lib.trase_b200_gemm_tf32(A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), D.data_ptr(),
                                         D.stride(0), M, N, K, bias.data_ptr() if bias is not None else None,
                                         int(relu), Z.data_ptr() if Z is not None else None,
                                         Z.stride(0) if Z is not None else 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, native_lib.trase_b200_last_error()
    torch.cuda.synchronize()
    return D


@pytest.mark.parametrize("M,N,K", [(128, 256, 32), (128, 256, 256), (1000, 256, 256), (4099, 256, 84),
                                   (4099, 256, 340), (777, 16, 256), (777, 10, 256), (50000, 256, 16),
                                   (300000, 256, 256)])
def test_gemm_plain(native_lib, M, N, K):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    ldk = ((K + 3) // 4) * 4
    A = torch.randn(M, ldk, device=DEV, generator=g)[:, :K]
    B = torch.randn(N, ldk, device=DEV, generator=g)[:, :K]
    D = run_gemm(native_lib, A, B)
    ref_t = tf32_trunc(A.contiguous()).double() @ tf32_trunc(B.contiguous()).double().T
    ref = A.double() @ B.double().T
    scale = ref.abs().max().item()
    assert torch.isfinite(D).all()
    err_t = (D.double() - ref_t).abs().max().item() / scale
    err = (D.double() - ref).abs().max().item() / scale
    # vs TF32-truncated operands only fp32 accumulation order differs; vs exact fp32 the TF32 rounding shows
    assert err_t <= 2e-5 or err <= 2e-3, (err_t, err)
    assert err <= 5e-3, err


def test_gemm_bias_relu_mask_and_strided_output(native_lib):
    g = torch.Generator(device=DEV).manual_seed(5)
    M, N, K = 3001, 256, 256
    A = torch.randn(M, K, device=DEV, generator=g)
    B = torch.randn(N, K, device=DEV, generator=g) * 0.1
    bias = torch.randn(N, device=DEV, generator=g)
    Z = torch.randn(M, N, device=DEV, generator=g)
    ref = A @ B.T + bias
    D = run_gemm(native_lib, A, B, bias=bias, relu=True)
    assert (D - torch.relu(ref)).abs().max() <= 2e-2
    assert (D >= 0).all()
    D = run_gemm(native_lib, A, B, Z=Z)
    want = (A @ B.T) * (Z > 0)
    assert (D - want).abs().max() <= 2e-2
    assert (D[Z <= 0] == 0).all()
    # output written into columns [84:340] of a wider buffer (the skip-connection layout)
    wide = torch.zeros(M, 340, device=DEV)
    run_gemm(native_lib, A, B, bias=bias, relu=True, D=wide[:, 84:340])
    assert (wide[:, 84:] - torch.relu(ref)).abs().max() <= 2e-2
    assert not wide[:, :84].any()
