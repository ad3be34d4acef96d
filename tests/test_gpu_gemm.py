"""tcgen05/TMEM GEMM kernel against a plain PyTorch fp32 reference of the same (bf16-rounded) operands.
Tolerance: the result is bf16, so |d| <= 2^-8 * |ref|max (one bf16 ulp at the top of the range) + f32
accumulation-order noise."""
import pytest

pytestmark = pytest.mark.gpu

SHAPES = [
    (128, 64, 64, 0), (64, 384, 384, 0), (1, 1152, 384, 0), (300, 1152, 384, 0), (3000, 384, 1536, 0),
    (1500, 1536, 384, 2), (64, 51864, 384, 0), (257, 384, 240, 0), (3000, 768, 384, 0), (4096, 512, 512, 0),
    (1000, 2304, 768, 0), (777, 3072, 768, 2), (2049, 192, 3072, 0),
]


@pytest.mark.parametrize("M,N,K,flags", SHAPES)
def test_gemm_bf16_vs_torch(M, N, K, flags):
    import torch
    from whisper_tflite_b200 import capi
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    b = torch.randn(N, device="cuda", generator=g)
    c = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    capi.debug_gemm_bf16(a, w, b, c, flags)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().T + b
    if flags & 2:
        ref = torch.nn.functional.gelu(ref)
    assert torch.isfinite(c.float()).all()
    err = (c.float() - ref).abs().max().item()
    assert err <= ref.abs().max().item() * 2 ** -8 + 1e-3
