"""Mainloop sweep of the tcgen05 GEMM on clean shapes: block_n x CTA-pair mode, CUDA-event timed (tuning aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def run(name, mk, flops):
    b = Builder(dev); mk(b)
    s = torch.cuda.current_stream().cuda_stream
    b.prog.run(s); torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.prog.run(s); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"{name:44s} {best*1000:9.1f} us {flops/best/1e9:8.1f} TFLOP/s", flush=True)
shapes = [(8192, 8192, 8192), (16384, 2560, 2560), (65536, 320, 2880), (65536, 1280, 1280)]
for M, N, K in shapes:
    x, w = bf(M, K), bf(N, K)
    for bn in (128, 160, 256):
        if N % bn: continue
        for pair in (1, 2):
            run(f"lin {M}x{N}x{K} bn={bn} pair={pair}", lambda b: b.linear(x, w, block_n=bn, cta_pair=pair), 2.0 * M * N * K)
    a = x.float() if False else None
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    y = x @ w.t(); torch.cuda.synchronize(); best = 1e9
    for _ in range(5):
        flush.zero_(); t0.record(); y = x @ w.t(); t1.record(); torch.cuda.synchronize(); best = min(best, t0.elapsed_time(t1))
    print(f"cuBLAS {M}x{N}x{K}: {best*1000:9.1f} us {2.0*M*N*K/best/1e9:8.1f} TFLOP/s", flush=True)
for NB, H, C, Co in [(16, 64, 320, 320), (16, 32, 640, 640), (16, 64, 320, 256)]:
    x, w = bf(NB, H, H, C), bf(Co, 9 * C)
    for bn in (128, 160, 256):
        if Co % bn: continue
        for pair in (1, 2):
            run(f"conv {NB}x{H}x{H} {C}->{Co} bn={bn} pair={pair}", lambda b: b.conv(x, w, block_n=bn, cta_pair=pair), 2.0 * NB * H * H * Co * 9 * C)
