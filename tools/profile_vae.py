"""Per-op device times of the VAE decode program (CUDA events around every op; profiling aid, not a bench)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from stablediffusion_b200 import _native as N
from stablediffusion_b200.models.stable_diffusion import SDModelWrapper
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8); ap.add_argument("--res", type=int, default=512)
ap.add_argument("--out", default="gpurun_out/profile_vae.txt")
a = ap.parse_args()
dev = "cuda:0"
m = SDModelWrapper.from_random("sd15", device=dev)
z = torch.randn(a.batch, 4, a.res // 8, a.res // 8, device=dev)
img = m.vae.decode(z / 0.18215, return_dict=False)[0]
torch.cuda.synchronize()
st = [v for k, v in m.vae._programs.items() if k[0] == "dec"][0]
b = st["b"]
s = torch.cuda.current_stream().cuda_stream
best = None
for _ in range(3):
    ms = b.prog.run_timed(s)
    best = ms if best is None else [min(x, y) for x, y in zip(best, ms)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); b.prog.run(s); e1.record(); torch.cuda.synchronize()
agg = {}
for i, (t, k) in enumerate(zip(best, b.prog.kinds)):
    md = b.meta.get(i, {})
    name = md.get("op", N.OP_NAMES[k])
    key = name + (":" + "x".join(str(md.get(q)) for q in ("M", "N", "K")) if "M" in md else "") + (f":{md.get('Nq')}x{md.get('Nkv')}d{md.get('d')}" if name == "attention" else "") + (f":{md.get('HW', md.get('rows'))}x{md.get('C')}" if "bytes" in md else "")
    e = agg.setdefault(key, [0, 0.0, 0.0, 0.0]); e[0] += 1; e[1] += t; e[2] += md.get("flops", 0); e[3] += md.get("bytes", 0)
tot = sum(best)
out = [f"# VAE decode sd15 res {a.res} batch {a.batch}; sum of per-op times {tot:.3f} ms over {len(best)} ops; one plain run {e0.elapsed_time(e1):.3f} ms"]
for key, (n, t, fl, by) in sorted(agg.items(), key=lambda x: -x[1][1]):
    extra = f"{fl / (t * 1e-3) / 1e12:8.1f} TFLOP/s" if fl else (f"{by / (t * 1e-3) / 1e9:8.1f} GB/s" if by else "")
    out.append(f"{t:8.3f} ms {100 * t / tot:5.1f}% n={n:3d} {t / n * 1000:8.1f} us/op {extra}  {key}")
open(a.out, "w").write("\n".join(out) + "\n")
print("\n".join(out[:40]))
