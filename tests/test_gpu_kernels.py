"""Kernel-level parity of the production tcgen05 kernels against plain PyTorch fp32 references of the same
(bf16-rounded) operands, through the C ABI's debug entries (wt_debug_attention_bf16, wt_debug_gemm_ex):

  * attention3_sm100_kernel for S in {92, 128, 129, 1500} (masked last key block; 1500 = 11*128 + 92), and a row whose
    score jumps by more than 2^64 against everything seen before (the redo branch of the streamed reference maximum);
  * gemm_bf16_sm100_kernel's addressing / epilogues the plain GEMM test does not reach: the residual add IN PLACE,
    the 3-tap convolution addressing (against torch.conv1d, strides 1 and 2), the cross-K/V scatter into the decoder's
    XOR-swizzled chunk stream, the fused (max, last index) with constructed ties.
Tolerances: outputs are bf16: 2^-8 of the reference's max (one bf16 ulp at the top of the range) + accumulation noise;
attention additionally carries P rounded to bf16 (2^-9 relative per weight, averaged by the row sum).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _attn_ref(qkv, H):
    import torch
    B, S, D3 = qkv.shape
    d = D3 // 3
    q, k, v = qkv.float().split(d, dim=2)
    sh = lambda t: t.view(B, S, H, 64).transpose(1, 2)   # [B][H][S][64]
    s = (sh(q) @ sh(k).transpose(2, 3)) * 0.125
    o = torch.softmax(s, dim=-1) @ sh(v)
    return o.transpose(1, 2).reshape(B, S, d)


@pytest.mark.parametrize("S", [92, 128, 129, 1500])
def test_attention_vs_torch(S):
    import torch
    from whisper_tflite_b200 import capi
    H, B = 2, 2
    g = torch.Generator(device="cuda").manual_seed(S)
    qkv = torch.randn(B, S, 3 * H * 64, device="cuda", generator=g)
    qkv[..., :H * 64] *= 2.0      # spread the scores: the softmax is not flat
    qkv = qkv.bfloat16()
    out = torch.full((B, S, H * 64), float("nan"), device="cuda", dtype=torch.bfloat16)
    capi.debug_attention_bf16(qkv, out, H)
    torch.cuda.synchronize()
    ref = _attn_ref(qkv, H)
    assert torch.isfinite(out.float()).all()
    err = (out.float() - ref).abs().max().item()
    assert err <= 1.5e-2 * ref.abs().max().item(), (S, err, ref.abs().max().item())


def test_attention_score_jump_redo_branch():
    """Keys of the first blocks score ~0 for query row 5; key 700 (block 5) scores 8*8*64*0.125 = 512 natural-log units
    = 739 in the exp2 domain: more than 64 above the running reference, so the block is redone with its exact maximum
    (attn3_sm100.cuh). The softmax of that row is one-hot on key 700; every other row is ordinary."""
    import torch
    from whisper_tflite_b200 import capi
    H, B, S = 1, 1, 1500
    g = torch.Generator(device="cuda").manual_seed(5)
    qkv = torch.randn(B, S, 3 * 64, device="cuda", generator=g) * 0.5
    qkv[0, 5, 0:64] = 8.0          # q row 5
    qkv[0, :, 64:128] *= 0.05      # small keys ...
    qkv[0, 700, 64:128] = 8.0      # ... except key 700
    qkv = qkv.bfloat16()
    out = torch.full((B, S, 64), float("nan"), device="cuda", dtype=torch.bfloat16)
    capi.debug_attention_bf16(qkv, out, H)
    torch.cuda.synchronize()
    ref = _attn_ref(qkv, H)
    assert torch.isfinite(out.float()).all()
    np.testing.assert_allclose(out[0, 5].float().cpu().numpy(), qkv[0, 700, 128:].float().cpu().numpy(), rtol=0, atol=2e-2)
    err = (out.float() - ref).abs().max().item()
    assert err <= 1.5e-2 * ref.abs().max().item(), err


def _tol(ref):
    return ref.abs().max().item() * 2 ** -8 + 1e-3


@pytest.mark.parametrize("M,N,K", [(300, 384, 384), (3000, 384, 1536), (1500 * 3, 512, 512)])
def test_gemm_residual_add_in_place(M, N, K):
    """x += A . W^T + b with C == addend (the encoder's out-proj / fc2 residual updates: engine.cu encode())."""
    import torch
    from whisper_tflite_b200 import capi
    g = torch.Generator(device="cuda").manual_seed(M + N)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    b = torch.randn(N, device="cuda", generator=g)
    x = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    ref = x.float() + a.float() @ w.float().T + b
    capi.debug_gemm_ex(A=a, a_dims=[K, M, 1], a_strides=[K * 2, K * 2 * M], rows_per_batch=M, n_batches=1, W=w, bias=b,
                       N=N, K_total=K, taps=1, C=x, c_batch_stride=0, c_rows=M, c_row_stride=N, addend=x,
                       add_batch_stride=0, add_rows=M, add_row_stride=N, flags=capi.EPI_BIAS | capi.EPI_ADD)
    torch.cuda.synchronize()
    assert (x.float() - ref).abs().max().item() <= _tol(ref)


@pytest.mark.parametrize("add", [0, 1])
@pytest.mark.parametrize("T,N,K", [(300, 384, 384), (1500, 384, 1536), (92, 1152, 384)])
def test_gemm_tma_store_clips_at_clip_end(T, N, K, add):
    """Per-clip batches into a padded C [B][T + 8][N] (and, with add, a residual read from a differently padded buffer):
    the epilogue's bulk tensor stores must clip at row T of every clip — the guard rows keep their sentinel — and the
    ragged last 128-row tile of a clip must not leak into the next clip."""
    import torch
    from whisper_tflite_b200 import capi
    B = 3
    g = torch.Generator(device="cuda").manual_seed(T + N + add)
    a = (torch.randn(B, T, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    b = torch.randn(N, device="cuda", generator=g)
    out = torch.full((B, T + 8, N), 7.0, device="cuda", dtype=torch.bfloat16)
    res = torch.randn(B, T + 16, N, device="cuda", generator=g).bfloat16()
    ref = a.float() @ w.float().T + b
    kw = dict(A=a, a_dims=[K, T, B], a_strides=[K * 2, T * K * 2], rows_per_batch=T, n_batches=B, W=w, bias=b, N=N,
              K_total=K, taps=1, C=out, c_batch_stride=(T + 8) * N, c_rows=T, c_row_stride=N, flags=capi.EPI_BIAS)
    if add:
        ref = ref + res[:, :T].float()
        kw.update(addend=res, add_batch_stride=(T + 16) * N, add_rows=T, add_row_stride=N, flags=capi.EPI_BIAS | capi.EPI_ADD)
    capi.debug_gemm_ex(**kw)
    torch.cuda.synchronize()
    assert (out[:, T:].float() == 7.0).all()
    assert (out[:, :T].float() - ref).abs().max().item() <= _tol(ref)


@pytest.mark.parametrize("stride,cin,T", [(1, 80, 3000), (2, 384, 3000), (1, 80, 200), (2, 128, 256)])
def test_gemm_three_tap_conv_vs_torch_conv1d(stride, cin, T):
    """conv1 (stride 1, 80 -> d) and conv2 (stride 2, d -> d) as 3-tap GEMMs over the zero-padded time-major activation
    [B][T+2][cin] (engine.cu encode()): stride 1 reads rows t, t+1, t+2; stride 2 views the buffer as PAIRS of rows
    and reads (pair p, col 0), (pair p, col cin), (pair p+1, col 0). GELU and, for conv2, the positional rows."""
    import torch
    from whisper_tflite_b200 import capi
    B, d = 3, 384
    g = torch.Generator(device="cuda").manual_seed(T + cin)
    xin = torch.zeros(B, T + 2, cin, device="cuda")
    xin[:, 1:T + 1] = torch.randn(B, T, cin, device="cuda", generator=g)
    xin = xin.bfloat16()
    wt_ = (torch.randn(d, cin, 3, device="cuda", generator=g) * (1.0 / (3 * cin) ** 0.5)).bfloat16()   # torch conv layout
    w_tap = wt_.permute(0, 2, 1).contiguous().view(d, 3 * cin)                                       # [n][tap][c_in]
    bias = torch.randn(d, device="cuda", generator=g)
    T_out = T // stride
    pos = torch.randn(T_out, d, device="cuda", generator=g).bfloat16()
    conv = torch.nn.functional.conv1d(xin[:, 1:T + 1].float().transpose(1, 2), wt_.float(), bias, stride=stride, padding=1)
    ref = torch.nn.functional.gelu(conv).transpose(1, 2)                                              # [B][T_out][d]
    out = torch.full((B, T_out, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    if stride == 1:
        capi.debug_gemm_ex(A=xin, a_dims=[cin, T + 2, B], a_strides=[cin * 2, (T + 2) * cin * 2], rows_per_batch=T,
                           n_batches=B, W=w_tap, bias=bias, N=d, K_total=3 * cin, taps=3, tap_k=cin, tap_c0=[0, 0, 0],
                           tap_row=[0, 1, 2], C=out, c_batch_stride=T_out * d, c_rows=T_out, c_row_stride=d,
                           flags=capi.EPI_BIAS | capi.EPI_GELU)
    else:
        ref = ref + pos.float()
        capi.debug_gemm_ex(A=xin, a_dims=[2 * cin, (T + 2) // 2, B], a_strides=[2 * cin * 2, (T + 2) * cin * 2],
                           rows_per_batch=T_out, n_batches=B, W=w_tap, bias=bias, N=d, K_total=3 * cin, taps=3, tap_k=cin,
                           tap_c0=[0, cin, 0], tap_row=[0, 0, 1], C=out, c_batch_stride=T_out * d, c_rows=T_out,
                           c_row_stride=d, addend=pos, add_batch_stride=0, add_rows=T_out, add_row_stride=d,
                           flags=capi.EPI_BIAS | capi.EPI_GELU | capi.EPI_ADD)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    assert (out.float() - ref).abs().max().item() <= _tol(ref)


@pytest.mark.parametrize("chunked", [1, 0])
def test_gemm_split_kv_scatter(chunked):
    """Cross-attention K/V projection with the scatter epilogue: [clip*S + s][k | v] -> per-(clip, head) streams.
    chunked: [clip][H][12 chunks][K rows 128x64 | V rows 128x64], the 16-byte segments of a key's row XOR-swizzled by
    (key & 7) — un-swizzled here and compared; plain: K and V as [clip][H][S][64]."""
    import torch
    from whisper_tflite_b200 import capi
    B, S, H, K = 2, 1500, 2, 128
    d = H * 64
    g = torch.Generator(device="cuda").manual_seed(3 + chunked)
    a = (torch.randn(B, S, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(2 * d, K, device="cuda", generator=g) * 0.1).bfloat16()
    bias = torch.randn(2 * d, device="cuda", generator=g)
    ref = (a.float() @ w.float().T + bias).view(B, S, 2, H, 64)      # [clip][s][k|v][head][e]
    cpp = (S + 127) // 128
    common = dict(A=a, a_dims=[K, S, B], a_strides=[K * 2, K * 2 * S], rows_per_batch=S, n_batches=B, W=w, bias=bias,
                  N=2 * d, K_total=K, taps=1, c_batch_stride=0, c_rows=B * S, c_row_stride=2 * d, kv_S=S, kv_H=H,
                  flags=capi.EPI_BIAS | capi.EPI_SPLIT_KV)
    if chunked:
        kv = torch.zeros(B, H, cpp, 2, 128, 64, device="cuda", dtype=torch.bfloat16)
        capi.debug_gemm_ex(kv_out=kv, kv_chunked=1, **common)
        torch.cuda.synchronize()
        seg = kv.view(B, H, cpp, 2, 128, 8, 8)                       # [..][key][segment][8 elements]
        key = torch.arange(128, device="cuda")
        idx = (torch.arange(8, device="cuda")[None, :] ^ (key[:, None] & 7))       # stored segment of logical segment j
        unsw = torch.gather(seg, 5, idx[None, None, None, None, :, :, None].expand(B, H, cpp, 2, 128, 8, 8))
        got = unsw.reshape(B, H, cpp, 2, 128, 64)
        for which in (0, 1):
            full = got[:, :, :, which].reshape(B, H, cpp * 128, 64)[:, :, :S].float()      # [clip][head][s][e]
            want = ref[:, :, which].permute(0, 2, 1, 3)
            assert (full - want).abs().max().item() <= _tol(want), which
        # keys past the clip's end (1500..1535 of the last chunk) are never written
        assert got[:, :, cpp - 1, :, S - (cpp - 1) * 128:].abs().max().item() == 0.0
    else:
        kx = torch.full((B, H, S, 64), float("nan"), device="cuda", dtype=torch.bfloat16)
        vx = torch.full((B, H, S, 64), float("nan"), device="cuda", dtype=torch.bfloat16)
        capi.debug_gemm_ex(kv_out=kx, kv_out_v=vx, kv_chunked=0, **common)
        torch.cuda.synchronize()
        for which, got in ((0, kx), (1, vx)):
            want = ref[:, :, which].permute(0, 2, 1, 3)
            assert (got.float() - want).abs().max().item() <= _tol(want), which


@pytest.mark.parametrize("M,bn", [(64, 128), (5, 128), (200, 64)])
def test_gemm_fused_argmax_with_ties(M, bn):
    """EPI_ARGMAX: per (row, N tile) the packed (value, index) maximum with whisper.cpp:346-361's rule. Rows of W are
    duplicated so that exact ties exist inside a tile and across tiles; the merged winner per row must be the LAST
    maximal index of the f32 logits (computed here from the same bf16 operands; ties are exact because identical weight
    rows give bit-identical accumulations)."""
    import torch
    from whisper_tflite_b200 import capi
    N, K = 51865 // 8 * 8, 128
    g = torch.Generator(device="cuda").manual_seed(M)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    logits = a.float() @ w.float().T
    win = logits.argmax(dim=1)
    # duplicate every row's winning weight row at two later indices: same tile (+1..) and a far tile
    dup = {}
    for r in range(min(M, 8)):
        t0 = int(win[r])
        for j in ((t0 + 3) % N, (t0 + 17 * bn + 5) % N):
            w[j] = w[t0]
            dup.setdefault(r, [t0]).append(j)
    n_tiles = (N + bn - 1) // bn
    part = torch.zeros(M, n_tiles, device="cuda", dtype=torch.int64)
    capi.debug_gemm_ex(A=a, a_dims=[K, M, 1], a_strides=[K * 2, K * 2 * M], rows_per_batch=M, n_batches=1, W=w, N=N,
                       K_total=K, taps=1, c_batch_stride=0, c_rows=M, c_row_stride=N, argmax_part=part,
                       flags=capi.EPI_ARGMAX, force_bn=bn)
    torch.cuda.synchronize()
    p = part.cpu().numpy().view(np.uint64)
    best = p.max(axis=1)
    got_idx = (best & np.uint64(0xffffffff)).astype(np.int64)
    ref_logits = (a.float() @ w.float().T).cpu().numpy()
    for r in range(M):
        row = ref_logits[r]
        m = row.max()
        cands = np.flatnonzero(row >= m - 1e-3 * abs(m))       # f32 accumulation-order noise between torch and the kernel
        assert got_idx[r] in cands, (r, got_idx[r], cands[:5])
        if r in dup:
            assert got_idx[r] == max(dup[r]), (r, got_idx[r], dup[r])
