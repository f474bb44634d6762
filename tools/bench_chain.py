"""In-graph time of small ops: K copies of one op on K different tensors as ONE CUDA-graph launch, time / K.
Per-op CUDA events cannot resolve ops of a few microseconds (the event pair itself has a floor of several us);
inside a graph with programmatic dependent launch the kernels run back to back as they do in the sampler step."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200.lowering import Builder

ap = argparse.ArgumentParser()
ap.add_argument("--copies", type=int, default=24); ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--out", default="gpurun_out/bench_chain.txt"); ap.add_argument("--only", default="")
a = ap.parse_args()
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)

cases = []
def gn(NB, HW, C, C2=0):
    def mk(b):
        b.groupnorm(bf(NB, HW, C), torch.ones(C + C2, device=dev), torch.zeros(C + C2, device=dev), 1e-5, True,
                    x2=bf(NB, HW, C2) if C2 else None)
    cases.append((f"gn_{NB}x{HW}x{C}+{C2}", mk, 4 * NB * HW * (C + C2)))
def ln(rows, C):
    def mk(b):
        b.layernorm(bf(rows, C), torch.ones(C, device=dev), torch.zeros(C, device=dev), 1e-5)
    cases.append((f"ln_{rows}x{C}", mk, 4 * rows * C))
def att(B, h, Nq, Nk, d):
    def mk(b):
        b.attention(bf(B, Nq, h * d), bf(B, Nk, h * d), bf(B, Nk, h * d), h, d, d ** -0.5)
    cases.append((f"att_{B}x{h}x{Nq}x{Nk}d{d}", mk, (2 * B * Nq + 2 * B * Nk) * h * d * 2))

for NB in (16, 2):
    gn(NB, 4096, 320); gn(NB, 1024, 640); gn(NB, 1024, 320); gn(NB, 256, 1280); gn(NB, 64, 1280); gn(NB, 64, 1280, 1280); gn(NB, 256, 1280, 1280)
    gn(NB, 1024, 640, 640); gn(NB, 1024, 1280)
    ln(NB * 4096, 320); ln(NB * 1024, 640); ln(NB * 256, 1280)
    att(NB, 8, 4096, 77, 40); att(NB, 8, 1024, 77, 80); att(NB, 8, 256, 77, 160); att(NB, 8, 64, 77, 160)

out = []
for name, mk, bytes_ in cases:
    if a.only and not any(o in name for o in a.only.split(',')):
        continue
    copies = max(2, min(a.copies, int((6 << 30) // max(bytes_, 1))))
    b = Builder(dev)
    for _ in range(copies):
        mk(b)
    st = torch.cuda.Stream(device=dev)
    s = st.cuda_stream
    with torch.cuda.stream(st):
        b.prog.run(s); torch.cuda.synchronize()
        b.prog.graph_capture(s)
        for _ in range(3): b.prog.graph_launch(s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(a.reps): b.prog.graph_launch(s)
        e1.record(st)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (a.reps * copies)
    line = f"{name:28s} {us:8.2f} us/op  {bytes_ / us / 1e3:8.1f} GB/s  ({copies} copies, {b.prog.num_launches // copies} launch/op)"
    print(line); out.append(line)
    del b
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
open(a.out, "w").write("\n".join(out) + "\n")
