"""Encoder attention kernel alone on the bench shape (64 clips x 6 heads x 1500 keys, head_dim 64), CUDA events,
inputs of 2 x 221 MB rotating (> L2).   usage: python tools/attn_bench.py [B] [H] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from whisper_tflite_b200 import capi  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
H = int(sys.argv[2]) if len(sys.argv) > 2 else 6
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
S = 1500
g = torch.Generator(device="cuda").manual_seed(0)
qkvs = [torch.randn(B, S, 3 * H * 64, device="cuda", generator=g).bfloat16() for _ in range(2)]
out = torch.empty(B, S, H * 64, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    for i in range(3):
        capi.debug_attention_bf16(qkvs[i % 2], out, H, stream=st.cuda_stream)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.synchronize()
    a.record(st)
    for i in range(iters):
        capi.debug_attention_bf16(qkvs[i % 2], out, H, stream=st.cuda_stream)
    b.record(st)
    st.synchronize()
ms = a.elapsed_time(b) / iters
flop = 4.0 * B * H * S * S * 64
ref = torch.softmax((qkvs[(iters - 1) % 2][:2, :, :64].float() @ qkvs[(iters - 1) % 2][:2, :, H * 64:H * 64 + 64].float().transpose(1, 2)) * 0.125, -1) \
    @ qkvs[(iters - 1) % 2][:2, :, 2 * H * 64:2 * H * 64 + 64].float()
err = (out[:2, :, :64].float() - ref).abs().max().item() / ref.abs().max().item()
print(f"attention3_sm100_kernel B={B} H={H}: {ms:.4f} ms per launch, {flop / ms / 1e9:.1f} TFLOP/s, "
      f"rel-to-max err (head 0, 2 clips) {err:.2e}")
