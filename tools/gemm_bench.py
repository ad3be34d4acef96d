"""The encoder's GEMM shapes one by one through the tcgen05 kernel (wt_debug_gemm_ex), CUDA events, operands rotating
over > L2 worth of buffers.  usage: python tools/gemm_bench.py [size] [B] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from whisper_tflite_b200 import capi, synth  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "tiny"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = synth.DIMS[size].d_model
M = B * 1500
g = torch.Generator(device="cuda").manual_seed(0)
st = torch.cuda.Stream()
shapes = [("qkv", 3 * d, d, capi.EPI_BIAS), ("out+res", d, d, capi.EPI_BIAS | capi.EPI_ADD),
          ("fc1+gelu", 4 * d, d, capi.EPI_BIAS | capi.EPI_GELU), ("fc2+res", d, 4 * d, capi.EPI_BIAS | capi.EPI_ADD),
          ("plain d->d", d, d, 0), ("plain d->4d", 4 * d, d, 0)]
tot = 0.0
for name, N, K, flags in shapes:
    a = [(torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16() for _ in range(2)]
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    c = [torch.zeros(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
    def run(i):
        kw = dict(A=a[i % 2], a_dims=[K, M, 1], a_strides=[K * 2, K * 2 * M], rows_per_batch=M, n_batches=1, W=w, bias=bias,
                  N=N, K_total=K, taps=1, C=c[i % 2], c_batch_stride=0, c_rows=M, c_row_stride=N, flags=flags)
        if flags & capi.EPI_ADD:
            kw.update(addend=c[i % 2], add_batch_stride=0, add_rows=M, add_row_stride=N)
        capi.debug_gemm_ex(stream=st.cuda_stream, **kw)
    with torch.cuda.stream(st):
        for i in range(3):
            run(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize()
        e0.record(st)
        for i in range(iters):
            run(i)
        e1.record(st)
        st.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flop = 2.0 * M * N * K
    byt = M * K * 2 + M * N * 2 * (2 if flags & capi.EPI_ADD else 1)
    if name in ("qkv", "out+res", "fc1+gelu", "fc2+res"):
        tot += ms
    print(f"{name:12s} M={M} N={N:5d} K={K:5d}: {ms * 1e3:8.1f} us  {flop / ms / 1e9:7.1f} TFLOP/s  {byt / ms / 1e6:7.1f} GB/s algorithmic")
print(f"one encoder layer's four GEMMs: {tot * 1e3:.1f} us")
