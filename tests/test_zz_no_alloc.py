"""wt_gemm_argmax_noalloc — the C-ABI entry point that allocates nothing on the device (its scratch is a static
__device__ array; include/whisper_b200.h) — against the existing allocating path, wt_debug_gemm_ex with the fused
(max, last index) epilogue, on the same operands. Indices are integers: the comparison is exact. Two calls in a row with
different operands and shapes prove that what the shared scratch keeps from one call does not leak into the next. The file
is named to be collected last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference_ids(a, w, bias, bn):
    """The existing entry point: per-(row, column tile) packed winners, merged here (one integer max per row)."""
    import torch
    from whisper_tflite_b200 import capi
    M, K = a.shape
    N = w.shape[0]
    n_part = (N + bn - 1) // bn
    part = torch.zeros(M, n_part, device="cuda", dtype=torch.int64)
    capi.debug_gemm_ex(A=a, a_dims=[K, M, 1], a_strides=[K * 2, K * 2 * M], rows_per_batch=M, n_batches=1, W=w, bias=bias,
                       N=N, K_total=K, taps=1, c_batch_stride=0, c_rows=M, c_row_stride=N, argmax_part=part,
                       flags=capi.EPI_BIAS | capi.EPI_ARGMAX)
    torch.cuda.synchronize()
    p = part.cpu().numpy().view(np.uint64)
    return (p.max(axis=1) & np.uint64(0xffffffff)).astype(np.int64)


def _operands(M, N, K, seed):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    return a, w, bias


def test_noalloc_gemm_argmax_matches_allocating_path_twice_in_a_row():
    import torch
    from whisper_tflite_b200 import capi
    # the decoder's tied logits at tiny (64 clips x 51 864 of the vocabulary, bn = 128), then a smaller, differently
    # tiled problem (bn = 64) on the same scratch
    cases = [(64, 51864, 384, 128, 1), (100, 4096, 256, 64, 2)]
    for M, N, K, bn, seed in cases:
        a, w, bias = _operands(M, N, K, seed)
        want = _reference_ids(a, w, bias, bn)
        ids = torch.full((M,), -7, device="cuda", dtype=torch.int32)
        before = torch.cuda.memory_allocated()
        capi.gemm_argmax_noalloc(a, w, bias, ids)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = ids.cpu().numpy().astype(np.int64)
        np.testing.assert_array_equal(got, want)
        # and against plain fp32 arithmetic on the same bf16 operands wherever the top-2 margin is not a rounding question
        logits = a.float() @ w.float().T + bias
        top2 = logits.topk(2, dim=1).values
        clear = ((top2[:, 0] - top2[:, 1]) > 1e-3).cpu().numpy()
        assert clear.sum() >= M // 2
        np.testing.assert_array_equal(got[clear], logits.argmax(dim=1).cpu().numpy()[clear])
    # the first problem again: the scratch now holds the second call's partials and scheduler state
    M, N, K, bn, seed = cases[0]
    a, w, bias = _operands(M, N, K, seed)
    ids = torch.full((M,), -7, device="cuda", dtype=torch.int32)
    capi.gemm_argmax_noalloc(a, w, None, ids)
    torch.cuda.synchronize()
    want = _reference_ids(a, w, torch.zeros(N, device="cuda"), bn)
    np.testing.assert_array_equal(ids.cpu().numpy().astype(np.int64), want)


def test_noalloc_rejects_sizes_beyond_its_static_scratch():
    import torch
    from whisper_tflite_b200 import capi
    M, N, K = 4096, 51864, 64          # 4096 rows x 406 column tiles of 128 > 131 072 partial winners
    a = torch.zeros(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.zeros(N, K, device="cuda", dtype=torch.bfloat16)
    ids = torch.zeros(M, device="cuda", dtype=torch.int32)
    with pytest.raises(capi.WhisperB200Error, match="exceed the static scratch"):
        capi.gemm_argmax_noalloc(a, w, None, ids)
