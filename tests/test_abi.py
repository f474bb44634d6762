"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol that
include/b200sd.h declares, its struct layouts match the ctypes mirror, host-only planning helpers work, and
compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from stablediffusion_b200 import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200sd.h")


def _declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"B200SD_API\s+[\w\s\*]+?\b(b200sd_\w+)\s*\(", src)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = N.lib()
    syms = _declared_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200sd.h but not exported"
    assert set(syms) == set(N.ENTRY_POINTS), "ctypes binding and header disagree on the entry points"
    out = subprocess.run(["nm", "-D", "--defined-only", N.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (b200sd_\w+)", out))
    assert exported == set(syms), f"exported-but-undeclared or missing: {exported ^ set(syms)}"


def test_library_has_no_libcuda_link_dependency():
    out = subprocess.run(["ldd", N.LIB_PATH], capture_output=True, text=True).stdout
    assert "libcuda.so" not in out  # the driver entry point is fetched at run time


def test_struct_layouts_match_header():
    lib = N.lib()
    assert lib.b200sd_abi_version() == N.ABI_VERSION == 3
    assert lib.b200sd_sizeof_op() == C.sizeof(N.Op)
    for kind, st in N.STRUCT_OF_KIND.items():
        assert lib.b200sd_sizeof_struct(kind) == C.sizeof(st), st.__name__


def test_enum_values_match_header():
    src = open(HEADER).read()
    for name, val in re.findall(r"(B200SD_OP_\w+)\s*=\s*(\d+)", src):
        assert getattr(N, name.replace("B200SD_", "")) == int(val)
    for name, val in re.findall(r"(B200SD_ST_\w+)\s*=\s*(\d+)", src):
        assert getattr(N, name.replace("B200SD_", "")) == int(val)
    for name, shift in re.findall(r"(B200SD_GEMM_\w+)\s*=\s*1\s*<<\s*(\d+)", src):
        assert getattr(N, name.replace("B200SD_", "")) == 1 << int(shift)
    assert int(re.search(r"#define B200SD_STEP_ROW (\d+)", src).group(1)) == N.STEP_ROW


def test_sass_contains_blackwell_native_instructions():
    out = subprocess.run(["cuobjdump", "-sass", N.LIB_PATH], capture_output=True, text=True).stdout
    if not out:
        pytest.skip("cuobjdump unavailable")
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM", "STTM"):
        assert mnemonic in out, f"{mnemonic} missing: the kernels are not tcgen05/TMA code"
    assert "HMMA." not in out.replace("UTCHMMA", ""), "legacy mma.sync path found"


def test_planning_helpers_run_without_gpu():
    lib = N.lib()
    g = N.Gemm()
    g.nseg, g.M, g.N = 1, 128, 1280
    g.seg[0].channels, g.seg[0].taps = 11520, 1
    ws = lib.b200sd_gemm_workspace_bytes(C.byref(g))
    assert ws > 0 and ws % (128 * 128 * 4) == 0      # the 8x8 UNet level is split along K
    g.M = 16384
    assert lib.b200sd_gemm_workspace_bytes(C.byref(g)) == 0
    g.seg[0].channels = 100                           # not a multiple of 8
    assert lib.b200sd_gemm_workspace_bytes(C.byref(g)) == N.ERR_INVALID
    assert "multiple of 8" in N.last_error()
    assert 1 <= lib.b200sd_groupnorm_chunks(2, 4096, 320) <= 4096 // 16


def test_batch_invariant_mode_pins_reduction_order_choices():
    """In the batch-invariant mode the split-K factor and the GroupNorm chunking follow the per-image shape only."""
    lib = N.lib()

    def splits(M, rps, K=11520, Nn=1280):
        g = N.Gemm()
        g.nseg, g.M, g.N, g.rows_per_sample = 1, M, Nn, rps
        g.seg[0].channels, g.seg[0].taps = K, 1
        ws = lib.b200sd_gemm_workspace_bytes(C.byref(g))
        assert ws >= 0
        return ws

    def conv_ws(NB, H=8, W=8, Cin=1280, Nn=1280):
        g = N.Gemm()
        g.spatial, g.nseg, g.N, g.NB, g.H, g.W, g.M = 1, 1, Nn, NB, H, W, NB * H * W
        g.seg[0].channels, g.seg[0].taps = Cin, 9
        ws = lib.b200sd_gemm_workspace_bytes(C.byref(g))
        assert ws >= 0
        return ws

    prev = N.set_batch_invariant(False)
    try:
        assert splits(128, 64) > 0 and splits(16384, 64) == 0            # default: the factor follows M
        N.set_batch_invariant(True)
        assert N.batch_invariant()
        per_row = {splits(M, 64) // ((M + 127) // 128) for M in (128, 256, 1024, 4096)}
        assert len(per_row) == 1, per_row                                 # same split factor at every batch (workspace ~ M tiles)
        per_img = {conv_ws(NB) // ((NB * 64 + 127) // 128) for NB in (2, 4, 16, 32)}
        assert len(per_img) == 1, per_img
        ch = {lib.b200sd_groupnorm_chunks(NB, 4096, 320) for NB in (1, 2, 8, 16, 64)}
        assert len(ch) == 1, ch
    finally:
        N.set_batch_invariant(prev)
    assert len({lib.b200sd_groupnorm_chunks(NB, 4096, 320) for NB in (1, 64)}) == 2


def test_compute_fails_loudly_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = N.lib().b200sd_program_create(C.byref(h), 0)
    assert rc == N.ERR_NO_DEVICE and "no CPU fallback" in N.last_error()
    from stablediffusion_b200.lowering import Builder

    with pytest.raises(RuntimeError):
        Builder(torch.device("cpu"))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "stablediffusion_b200")
    allowed = {os.path.join(pkg, "smoke.py")}  # __graft_entry__.smoke(): the oracle as checker only
    bad = []
    for dp, _, files in os.walk(pkg):
        for f in files:
            p = os.path.join(dp, f)
            if f.endswith((".py", ".cu", ".cuh", ".h")) and p not in allowed:
                if re.search(r"^\s*(from|import)\s+oracle\b", open(p).read(), re.M):
                    bad.append(p)
    assert not bad, f"product files import the oracle: {bad}"
