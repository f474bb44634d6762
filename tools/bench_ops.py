"""Times representative single ops of the SD1.5 b8 step in isolation (CUDA events, L2 flushed between reps).
Used for kernel tuning and as the short command wrapped by `ncu --set full`."""
import argparse, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder
from stablediffusion_b200.packing import geglu_interleave, geglu_block_n

ap = argparse.ArgumentParser()
ap.add_argument("--only", default=""); ap.add_argument("--reps", type=int, default=5); ap.add_argument("--out", default="gpurun_out/bench_ops.txt"); ap.add_argument("--pair", type=int, default=0)
a = ap.parse_args()
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
cases = []
def lin(name, M, N, K, res=False, geglu=False):
    def mk(b):
        x, w = bf(M, K), bf(N, K)
        bias = torch.randn(N, device=dev)
        if geglu:
            bn = geglu_block_n(N); w = geglu_interleave(w, bn).contiguous(); bias = geglu_interleave(bias, bn).contiguous()
            b.linear(x, w, bias=bias, geglu=True, block_n=bn, cta_pair=a.pair)
        else:
            b.linear(x, w, bias=bias, residual=bf(M, N) if res else None, cta_pair=a.pair)
    cases.append((name, mk, 2.0 * M * N * K, (M * K + N * K + M * (N // 2 if geglu else N) * (2 if res else 1)) * 2))
def conv(name, NB, H, C, Co, res=False):
    def mk(b):
        b.conv(bf(NB, H, H, C), bf(Co, 9 * C), bias=torch.randn(Co, device=dev), residual=bf(NB, H, H, Co) if res else None, cta_pair=a.pair)
    cases.append((name, mk, 2.0 * NB * H * H * Co * 9 * C, (NB * H * H * (C + Co * (2 if res else 1)) + Co * 9 * C) * 2))
def att(name, B, h, Nq, Nk, d):
    def mk(b):
        b.attention(bf(B, Nq, h * d), bf(B, Nk, h * d), bf(B, Nk, h * d), h, d, d ** -0.5)
    cases.append((name, mk, 4.0 * B * h * Nq * Nk * d, (2 * B * Nq + 2 * B * Nk) * h * d * 2))
def gn(name, NB, HW, C):
    def mk(b):
        b.groupnorm(bf(NB, HW, C), torch.ones(C, device=dev), torch.zeros(C, device=dev), 1e-5, True)
    cases.append((name, mk, 0, 4 * NB * HW * C))
def ln(name, rows, C):
    def mk(b):
        b.layernorm(bf(rows, C), torch.ones(C, device=dev), torch.zeros(C, device=dev), 1e-5)
    cases.append((name, mk, 0, 4 * rows * C))

lin("lin_65536x320x320_res", 65536, 320, 320, res=True)
lin("lin_65536x960x320", 65536, 960, 320)
lin("geglu_65536x2560x320", 65536, 2560, 320, geglu=True)
lin("lin_65536x320x1280_res", 65536, 320, 1280, res=True)
lin("lin_16384x640x640_res", 16384, 640, 640, res=True)
lin("geglu_16384x5120x640", 16384, 5120, 640, geglu=True)
lin("lin_4096x1280x1280_res", 4096, 1280, 1280, res=True)
conv("conv_16x64x64_320_320", 16, 64, 320, 320)
conv("conv_16x64x64_320_320_res", 16, 64, 320, 320, res=True)
conv("conv_16x32x32_640_640", 16, 32, 640, 640)
conv("conv_16x16x16_1280_1280", 16, 16, 1280, 1280)
conv("conv_16x8x8_1280_1280", 16, 8, 1280, 1280)
att("attn_self_4096_d40", 16, 8, 4096, 4096, 40)
att("attn_self_1024_d80", 16, 8, 1024, 1024, 80)
att("attn_cross_4096_d40", 16, 8, 4096, 77, 40)
att("attn_cross_1024_d80", 16, 8, 1024, 77, 80)
att("attn_self_256_d160", 16, 8, 256, 256, 160)
att("attn_self_64_d160", 16, 8, 64, 64, 160)
att("attn_self_4096_d40_b2", 2, 8, 4096, 4096, 40)
att("attn_self_1024_d80_b2", 2, 8, 1024, 1024, 80)
att("attn_xl_4096_d64", 4, 10, 4096, 4096, 64)
att("attn_xl_1024_d64", 4, 20, 1024, 1024, 64)
gn("gn_16x4096x320", 16, 4096, 320)
gn("gn_16x1024x640", 16, 1024, 640)
gn("gn_16x256x1280", 16, 256, 1280)
gn("gn_16x64x1280", 16, 64, 1280)
gn("gn_16x64x2560", 16, 64, 2560)
gn("gn_16x256x2560", 16, 256, 2560)
gn("gn_16x1024x1280", 16, 1024, 1280)
gn("gn_2x4096x320", 2, 4096, 320)
gn("gn_2x1024x640", 2, 1024, 640)
gn("gn_2x256x1280", 2, 256, 1280)
gn("gn_2x64x1280", 2, 64, 1280)
ln("ln_65536x320", 65536, 320)

flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = []
for name, mk, flops, bytes_ in cases:
    if a.only and not any(o in name for o in a.only.split(',')):
        continue
    b = Builder(dev)
    mk(b)
    s = torch.cuda.current_stream().cuda_stream
    b.prog.run(s); torch.cuda.synchronize()
    best = 1e9
    for _ in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.prog.run(s); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    line = f"{name:32s} {best*1000:9.1f} us  " + (f"{flops/best/1e9:8.1f} TFLOP/s  " if flops else " " * 18) + f"{bytes_/best/1e6:8.1f} GB/s (min traffic)"
    out.append(line); print(line)
open(a.out, "w").write("\n".join(out) + "\n")
