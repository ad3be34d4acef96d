"""Where the C stores of the sixteen-epilogue-warp GEMM cost: the same [96 000 x 384] x [1536 x 384]^T product with C
(a) as usual (295 MB to HBM) and (b) with every 128-row tile written over the same L2-resident rows (750 batches of 128
rows whose C batch stride is 16 elements). usage: python tools/probes/gemm_store_probe.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from whisper_tflite_b200 import capi  # noqa: E402

M, N, K = 96000, 1536, 384
g = torch.Generator(device="cuda").manual_seed(0)
a = [(torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16() for _ in range(2)]
w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
c_full = [torch.zeros(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
c_small = torch.zeros(128 * N + 750 * 16 + 64, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.Stream()


def run(i, small):
    if small:
        capi.debug_gemm_ex(stream=st.cuda_stream, A=a[i % 2], a_dims=[K, 128, 750], a_strides=[K * 2, K * 2 * 128], rows_per_batch=128,
                           n_batches=750, W=w, N=N, K_total=K, taps=1, C=c_small, c_batch_stride=16, c_rows=128, c_row_stride=N, flags=0)
    else:
        capi.debug_gemm_ex(stream=st.cuda_stream, A=a[i % 2], a_dims=[K, 128, 750], a_strides=[K * 2, K * 2 * 128], rows_per_batch=128,
                           n_batches=750, W=w, N=N, K_total=K, taps=1, C=c_full[i % 2], c_batch_stride=128 * N, c_rows=128,
                           c_row_stride=N, flags=0)


for small in (False, True):
    with torch.cuda.stream(st):
        for i in range(3):
            run(i, small)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize()
        e0.record(st)
        for i in range(20):
            run(i, small)
        e1.record(st)
        st.synchronize()
    print(f"C {'over the same L2-resident rows' if small else 'to HBM (295 MB)'}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
