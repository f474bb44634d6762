"""`b200sd_groupnorm_static_scratch` — the allocation-free GroupNorm(+SiLU) entry point (include/b200sd.h; no counterpart
in the reference: added for callers that cannot call cudaMalloc).

Called through the C ABI (ctypes) on raw device pointers and compared with the existing entry point for the same op
(B200SD_OP_GROUPNORM through an op program with a caller workspace), on two calls in a row with different shapes and data —
the second call finds the first call's partial sums in the shared static scratch and must not be disturbed by them.
Tolerance: the element-wise bound tests/test_kernels_gpu.py::test_groupnorm uses (2^-7 of the value + 2^-7 of the tensor's
scale, rel-L2 < 1e-2), against the existing entry point AND against torch's fp32 group_norm.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _check(out, ref, name, rtol=2 ** -7, atol_scale=2 ** -7):
    out, ref = out.float(), ref.float()
    assert torch.isfinite(out).all(), f"{name}: non-finite output"
    scale = ref.abs().max().item() + 1e-6
    err = (out - ref).abs()
    bad = (err > atol_scale * scale + rtol * ref.abs()).sum().item()
    rel_l2 = (err.pow(2).sum().sqrt() / (ref.pow(2).sum().sqrt() + 1e-12)).item()
    assert bad == 0 and rel_l2 < 1e-2, f"{name}: {bad} elements out of tolerance, max err {err.max().item():.4g}, rel-L2 {rel_l2:.3g}"
    return rel_l2


def _case(dev, NB, HW, C1, C2, silu, eps, seed):
    g = torch.Generator().manual_seed(seed)
    x1 = (torch.randn(NB, HW, C1, generator=g) * 2.0 + 0.5).to(torch.bfloat16).to(dev)
    x2 = torch.randn(NB, HW, C2, generator=g).to(torch.bfloat16).to(dev) if C2 else None
    gamma = (torch.randn(C1 + C2, generator=g) * 0.2 + 1.0).to(dev)
    beta = (torch.randn(C1 + C2, generator=g) * 0.2).to(dev)
    return x1, x2, gamma, beta, silu, eps


def _existing_entry_point(dev, x1, x2, gamma, beta, silu, eps):
    from stablediffusion_b200 import lowering as L

    b = L.Builder(dev)
    out = b.groupnorm(x1, gamma, beta, eps, silu, x2=x2)
    b.prog.run(torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize()
    return out.clone()


def _static_scratch_op(x1, x2, gamma, beta, silu, eps, out):
    from stablediffusion_b200 import _native as N

    NB, HW, C1 = x1.shape
    p = N.GroupNorm()
    p.x1, p.ld1, p.C1 = x1.data_ptr(), x1.stride(1), C1
    if x2 is not None:
        p.x2, p.ld2, p.C2 = x2.data_ptr(), x2.stride(1), x2.shape[2]
    p.NB, p.HW, p.groups, p.silu, p.eps = NB, HW, 32, int(silu), float(eps)
    p.gamma, p.beta, p.out = gamma.data_ptr(), beta.data_ptr(), out.data_ptr()
    p.partials, p.partials_bytes = None, 0      # the scratch is the library's static device array
    return p


def _torch_ref(x1, x2, gamma, beta, silu, eps):
    x = x1.float() if x2 is None else torch.cat([x1.float(), x2.float()], dim=2)
    ref = F.group_norm(x.transpose(1, 2), 32, gamma, beta, eps).transpose(1, 2)
    return F.silu(ref) if silu else ref


def test_groupnorm_static_scratch_matches_the_program_op_on_two_calls_in_a_row(cuda_device):
    from stablediffusion_b200 import _native as N

    dev = cuda_device
    lib = N.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    # two different shapes back to back: a level-0 resnet norm with a skip concat (many chunks), then a small 8x8-level one
    # (the program op takes its single-launch cluster kernel there; the static-scratch variant always uses the scratch)
    cases = [_case(dev, 2, 4096, 640, 320, True, 1e-5, seed=1), _case(dev, 3, 64, 1280, 0, False, 1e-6, seed=2)]
    refs = [_existing_entry_point(dev, *c) for c in cases]
    outs = [torch.full_like(r, float("nan")) for r in refs]
    ops = [_static_scratch_op(*c, out=o) for c, o in zip(cases, outs)]
    torch.cuda.synchronize()
    # first call (sets the entry point up on this device: symbol address, kernel attribute)
    N.check(lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(ops[0]), C.c_void_p(stream)))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    # ... and the two calls in a row, on one stream, no synchronisation in between
    outs[0].fill_(float("nan"))
    N.check(lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(ops[0]), C.c_void_p(stream)))
    N.check(lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(ops[1]), C.c_void_p(stream)))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "the entry point changed the device's free memory: it allocated"
    for i, (c, o, r) in enumerate(zip(cases, outs, refs)):
        a = _check(o, r, f"call {i}: static scratch vs program op")
        b = _check(o, _torch_ref(*c), f"call {i}: static scratch vs torch fp32")
        print(f"NO-ALLOC GROUPNORM call {i} {tuple(o.shape)}: rel-L2 vs program op {a:.3g} (bit-identical: {torch.equal(o, r)}), vs torch {b:.3g}")
    # a third call with the FIRST shape again: stale partials of the small call in the scratch change nothing
    first = outs[0].clone()
    outs[0].fill_(float("nan"))
    N.check(lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(ops[0]), C.c_void_p(stream)))
    torch.cuda.synchronize()
    assert torch.equal(outs[0], first)


def test_groupnorm_static_scratch_rejects_what_it_cannot_serve(cuda_device):
    from stablediffusion_b200 import _native as N

    dev = cuda_device
    lib = N.lib()
    x1, x2, gamma, beta, silu, eps = _case(dev, 2, 64, 64, 0, True, 1e-5, seed=3)
    out = torch.empty_like(x1)
    op = _static_scratch_op(x1, x2, gamma, beta, silu, eps, out)
    # beyond the static maximum: NB * chunks * 32 groups * 8 bytes > 1 MiB (pure host arithmetic: nothing is launched)
    op.NB, op.HW = 8192, 64
    assert lib.b200sd_groupnorm_chunks(8192, 64, 64) * 8192 * 32 * 8 > N.GN_STATIC_SCRATCH_BYTES
    rc = lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(op), None)
    assert rc == N.ERR_INVALID and "static scratch holds" in N.last_error()
    op.NB = 2
    ws = torch.empty(1 << 16, dtype=torch.uint8, device=dev)
    op.partials = ws.data_ptr()                 # a caller workspace is a misuse of this entry point
    rc = lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(op), None)
    assert rc == N.ERR_INVALID and "must be NULL" in N.last_error()
    op.partials = None
    with pytest.raises(ValueError):
        N.check(lib.b200sd_groupnorm_static_scratch(dev.index, None, None))
    N.check(lib.b200sd_groupnorm_static_scratch(dev.index, C.byref(op), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    torch.cuda.synchronize()
    _check(out, _torch_ref(x1, x2, gamma, beta, silu, eps), "after the rejected calls")
