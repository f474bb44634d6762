"""Exhaustive tiling sweep (block_n x CTA pair x split-K) of the short projections of a UNet step against the planner's own
choice and cuBLAS (tuning aid).  Each case: bias + residual epilogue like the real call sites, L2 flushed between runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def run(mk):
    b = Builder(dev); mk(b)
    s = torch.cuda.current_stream().cuda_stream
    b.prog.run(s); torch.cuda.synchronize()
    best = 1e9
    for _ in range(7):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.prog.run(s); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1000
shapes = [(65536, 320, 320), (16384, 640, 640), (4096, 1280, 1280), (1024, 1280, 1280), (65536, 960, 320), (65536, 320, 1280),
          (16384, 640, 2560), (4096, 1280, 5120)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for M, N, K in shapes:
    x, w, bias, res = bf(M, K), bf(N, K), torch.randn(N, device=dev), bf(M, N)
    auto = run(lambda b: b.linear(x, w, bias=bias, residual=res))
    res_t = []
    for bn in (32, 64, 128, 160, 256):
        if N % bn and bn > N: continue
        for pair in (1, 2):
            for sp in (1, 2, 3, 4):
                try:
                    t = run(lambda b: b.linear(x, w, bias=bias, residual=res, block_n=bn, cta_pair=pair, splits=sp))
                except Exception as e:
                    continue
                res_t.append((t, bn, pair, sp))
    res_t.sort()
    y = torch.addmm(res.float(), x.float(), w.float().t()) if False else None
    best = 1e9
    for _ in range(7):
        flush.zero_(); t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record(); yy = torch.nn.functional.linear(x, w); t1.record(); torch.cuda.synchronize(); best = min(best, t0.elapsed_time(t1) * 1000)
    top = ", ".join(f"{t:.1f}us(bn{bn},p{p},s{s})" for t, bn, p, s in res_t[:4])
    print(f"{M}x{N}x{K}: auto {auto:.1f} us | best: {top} | cuBLAS (no epilogue) {best:.1f} us", flush=True)
    if os.environ.get("SWEEP_ALL"):
        print("   all:", " ".join(f"{bn}/{p}/{s_}:{t:.1f}" for t, bn, p, s_ in sorted(res_t, key=lambda r: (r[1], r[2], r[3]))), flush=True)
