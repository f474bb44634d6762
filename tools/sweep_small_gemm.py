"""Tiling sweep for the small-M GEMMs of the SD1.5 b8 step (block_n x CTA pair x split-K), warm caches, in-graph:
K copies of the op in one CUDA-graph launch (tools/bench_chain.py explains why).  Compares with the chooser's pick and
with cuBLAS (torch.matmul) on the same shape.  Tuning aid."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder

ap = argparse.ArgumentParser()
ap.add_argument("--copies", type=int, default=8); ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--out", default="gpurun_out/sweep_small_gemm.txt"); ap.add_argument("--set", default="small", choices=["small", "n640"])
a = ap.parse_args()
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)

def time_prog(build):
    b = Builder(dev)
    try:
        for _ in range(a.copies): build(b)
    except (ValueError, RuntimeError) as e:
        return None
    st = torch.cuda.Stream(device=dev); s = st.cuda_stream
    with torch.cuda.stream(st):
        b.prog.run(s); torch.cuda.synchronize()
        b.prog.graph_capture(s)
        for _ in range(2): b.prog.graph_launch(s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(a.reps): b.prog.graph_launch(s)
        e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (a.reps * a.copies)

out = []
def emit(line):
    print(line, flush=True); out.append(line)

if a.set == "n640":  # the 32x32 level (N = 640 = 2.5 tiles of 256): is 4 x 160 better than 3 x 256?
    lin_shapes = [(16384, 640, 640, True), (16384, 640, 2560, True), (16384, 1920, 640, False)]
    conv_shapes_sel = [(16, 32, 640, 640), (16, 32, 1280, 640), (16, 32, 960, 640), (16, 64, 320, 320), (16, 64, 640, 320)]
else:
    conv_shapes_sel = None
lin_shapes = lin_shapes if a.set == "n640" else [(4096, 1280, 1280, True), (16384, 640, 640, True), (4096, 1280, 5120, True), (4096, 3840, 1280, False), (1024, 1280, 1280, True),
              (16384, 1920, 640, False), (16384, 640, 2560, True)]
for M, N, K, res in lin_shapes:
    xs = [bf(M, K) for _ in range(a.copies)]; w = bf(N, K); r = bf(M, N) if res else None
    it = iter(range(10 ** 9))
    def mk(bn=0, pair=0, sp=0):
        def build(b):
            b.linear(xs[next(it) % a.copies], w, bias=torch.zeros(N, device=dev), residual=r, block_n=bn, cta_pair=pair, splits=sp)
        return build
    flops = 2.0 * M * N * K
    t = time_prog(mk())
    emit(f"lin {M}x{N}x{K} res={int(res)} chooser: {t:7.1f} us {flops / t / 1e6:7.0f} TFLOP/s")
    for bn in (128, 160, 256):
        if N % bn and bn not in (160, 256): continue
        for pair in (1, 2):
            for sp in ((1,) if a.set == "n640" else (1, 2, 3, 4)):
                t = time_prog(mk(bn, pair, sp))
                if t is not None:
                    emit(f"    bn={bn} pair={pair} splits={sp}: {t:7.1f} us {flops / t / 1e6:7.0f} TFLOP/s")
    y = None
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3): y = xs[0] @ w.t()
    e0.record()
    for i in range(a.reps * a.copies): y = xs[i % a.copies] @ w.t()
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e3 / (a.reps * a.copies)
    emit(f"    cuBLAS (no epilogue, eager launches): {t:7.1f} us {flops / t / 1e6:7.0f} TFLOP/s")

conv_shapes = conv_shapes_sel or [(16, 8, 1280, 1280), (16, 16, 1280, 1280), (16, 16, 640, 1280), (16, 8, 2560, 1280)]
for NB, H, C, Co in conv_shapes:
    xs = [bf(NB, H, H, C) for _ in range(a.copies)]; w = bf(Co, 9 * C)
    it = iter(range(10 ** 9))
    def mkc(bn=0, pair=0, sp=0):
        def build(b):
            b.conv(xs[next(it) % a.copies], w, bias=torch.zeros(Co, device=dev), block_n=bn, cta_pair=pair, splits=sp)
        return build
    flops = 2.0 * NB * H * H * Co * 9 * C
    t = time_prog(mkc())
    emit(f"conv {NB}x{H}x{H} {C}->{Co} chooser: {t:7.1f} us {flops / t / 1e6:7.0f} TFLOP/s")
    for bn in (128, 160, 256):
        for pair in (1, 2):
            for sp in ((1,) if a.set == "n640" else (1, 2, 3, 4, 6, 8)):
                t = time_prog(mkc(bn, pair, sp))
                if t is not None:
                    emit(f"    bn={bn} pair={pair} splits={sp}: {t:7.1f} us {flops / t / 1e6:7.0f} TFLOP/s")
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
open(a.out, "w").write("\n".join(out) + "\n")
