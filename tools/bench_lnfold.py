"""In-isolation A/B of the LayerNorm-folded consumer GEMMs against the two-kernel form (LayerNorm kernel + plain GEMM).
CUDA events around one program run, L2 flushed between repetitions."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder
from stablediffusion_b200.packing import fold_layernorm, geglu_block_n, geglu_interleave

dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(b, reps=7):
    s = torch.cuda.current_stream().cuda_stream
    b.prog.run(s); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.prog.run(s); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e3

for M, C, N, geglu in [(65536, 320, 960, False), (65536, 320, 2560, True), (65536, 320, 320, False), (16384, 640, 1920, False),
                       (16384, 640, 5120, True), (4096, 1280, 3840, False), (4096, 1280, 10240, True)]:
    x = bf(M, C)
    w = torch.randn(N, C, device=dev) / math.sqrt(C)
    bias = torch.randn(N, device=dev)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    wf, cs, bfold = fold_layernorm(w, bias, gamma, beta)
    wp, bp = w.to(torch.bfloat16), bias
    bn = 0
    if geglu:
        bn = geglu_block_n(N)
        wf, cs, bfold = [geglu_interleave(t, bn).contiguous() for t in (wf, cs, bfold)]
        wp, bp = geglu_interleave(wp, bn).contiguous(), geglu_interleave(bp, bn).contiguous()
    slots = 2 * ((C + 127) // 128)
    res = {}
    b = Builder(dev); b.linear(x, wp, bias=bp, geglu=geglu, block_n=bn); res["plain gemm (bias)"] = timeit(b)
    b = Builder(dev); b.linear(x, wp, geglu=geglu, block_n=bn); res["plain gemm (no bias)"] = timeit(b)
    b = Builder(dev); st = b.zeros([slots, M, 2], torch.float32); b.linear(x, wf, bias=bfold, geglu=geglu, block_n=bn, ln_fold=(st, cs, 1e-5)); res["folded gemm"] = timeit(b)
    b = Builder(dev); b.layernorm(x, gamma, beta, 1e-5); res["layernorm kernel"] = timeit(b)
    print(f"{M}x{N}x{C} geglu={int(geglu)}: " + "  ".join(f"{k} {v:.1f} us" for k, v in res.items()), flush=True)
