"""Prints CTA (0,0,0)'s clock-stamp trace of one attention launch (library built with -DB200SD_ATT_TRACE; tuning aid)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200 import _native as N
from stablediffusion_b200.lowering import Builder
B, h, Nq, Nk, d = [int(x) for x in sys.argv[1:6]]
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
b = Builder(dev)
b.attention(bf(B, Nq, h * d), bf(B, Nk, h * d), bf(B, Nk, h * d), h, d, d ** -0.5)
s = torch.cuda.current_stream().cuda_stream
b.prog.run(s); torch.cuda.synchronize()
b.prog.run(s); torch.cuda.synchronize()
buf = (C.c_ulonglong * 6144)()
N.lib().b200sd_debug_att_trace(buf)
names = {0: "tma", 1: "mma", 2: "sm0", 3: "sm1", 4: "sm2", 5: "sm3"}
ev = []
for role in range(6):
    for i in range(1024):
        v = buf[role * 1024 + i]
        if v: ev.append((v & 0xFFFFFFFFFFFF, role, v >> 48))
ev.sort(); t0 = ev[0][0]
lo, hi = int(os.environ.get("TRACE_LO", "0")), int(os.environ.get("TRACE_HI", "100000"))
for t, role, idn in ev[lo:hi]:
    print(f"{t - t0:8d} {'         ' * role}{names[role]}:{idn}")
