"""ctypes binding of libb200sd.so (declared in include/b200sd.h).

The structures below mirror the C header field for field; `tests/test_abi.py` checks every size against
`b200sd_sizeof_struct` so a drift between the two is caught on CPU.  There is deliberately no fallback:
if the shared library is missing, `lib()` raises, and every compute entry point fails without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200SD_LIB_DIR: tuning aid — load a variant build (csrc/build.sh with B200SD_LIB_OUT / B200SD_EXTRA_NVCC_FLAGS) for same-box A/Bs
LIB_PATH = os.path.join(os.environ.get("B200SD_LIB_DIR") or os.path.join(_HERE, "_lib"), "libb200sd.so")
BUILD_SCRIPT = os.path.join(_HERE, "csrc", "build.sh")

MAX_SEG = 3
MAX_TAPS = 9
STEP_ROW = 16
IPC_HANDLE_BYTES = 64
GN_STATIC_SCRATCH_BYTES = 1 << 20  # include/b200sd.h B200SD_GN_STATIC_SCRATCH_BYTES

# enum b200sd_op_kind
OP_GEMM, OP_GROUPNORM, OP_LAYERNORM, OP_ATTENTION, OP_TIMESTEP_EMBED, OP_PREP_LATENTS = 1, 2, 3, 4, 5, 6
OP_UPSAMPLE2X, OP_PHASE_SPLIT, OP_SOFTMAX_ROWS, OP_CFG_STEP, OP_AXPBY, OP_IMAGE_OUT, OP_STEP_ADVANCE = 7, 8, 9, 10, 11, 12, 13
OP_IMAGE_U8, OP_EMBED, OP_GAUSS_SAMPLE, OP_PAIR_EXCHANGE = 14, 15, 16, 17
ABI_VERSION = 3  # include/b200sd.h B200SD_ABI_VERSION
OP_NAMES = {
    1: "gemm", 2: "groupnorm", 3: "layernorm", 4: "attention", 5: "timestep_embed", 6: "prep_latents",
    7: "upsample2x", 8: "phase_split", 9: "softmax_rows", 10: "cfg_step", 11: "axpby", 12: "image_out",
    13: "step_advance", 14: "image_u8", 15: "embed", 16: "gauss_sample", 17: "pair_exchange",
}
# enum b200sd_dtype
BF16, F32, F16 = 0, 1, 2
# gemm flags
GEMM_BIAS, GEMM_BIAS_ROW, GEMM_ROWVEC, GEMM_RESIDUAL, GEMM_SILU, GEMM_GEGLU, GEMM_OUT_F32 = 1, 2, 4, 8, 16, 32, 64
GEMM_QUICK_GELU, GEMM_GELU = 128, 256
# step-table columns
ST_TIMESTEP, ST_IN_SCALE, ST_K1, ST_K2, ST_CX, ST_CE, ST_C0, ST_C1, ST_GUIDANCE = range(9)
ST_CN, ST_H1, ST_H2, ST_H3, ST_FLAGS, ST_BA, ST_BS = 9, 10, 11, 12, 13, 14, 15
STF_PUSH_EPS, STF_SAVE_X, STF_USE_SAVED_X, STF_UNIPC, STF_UNIPC_CORRECT = 1, 2, 4, 8, 16

ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_UNSUPPORTED = -1, -2, -3, -4


class GemmSeg(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("w", C.c_void_p), ("lda", C.c_int64), ("channels", C.c_int32), ("taps", C.c_int32),
        ("Hs", C.c_int32), ("Ws", C.c_int32), ("phases", C.c_int32),
        ("tap_phase", C.c_int8 * MAX_TAPS), ("tap_dh", C.c_int8 * MAX_TAPS), ("tap_dw", C.c_int8 * MAX_TAPS),
        ("pad_", C.c_int8 * 1),
    ]


class Gemm(C.Structure):
    _fields_ = [
        ("spatial", C.c_int32), ("nseg", C.c_int32), ("seg", GemmSeg * MAX_SEG), ("ldw", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("NB", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("out", C.c_void_p), ("ldc", C.c_int64),
        ("OH", C.c_int32), ("OW", C.c_int32), ("sy", C.c_int32), ("sx", C.c_int32), ("oy", C.c_int32), ("ox", C.c_int32),
        ("bias", C.c_void_p), ("rowvec", C.c_void_p), ("ld_rowvec", C.c_int64),
        ("rows_per_group", C.c_int32), ("flags", C.c_int32), ("residual", C.c_void_p), ("ldr", C.c_int64),
        ("alpha", C.c_float), ("splits", C.c_int32), ("workspace", C.c_void_p), ("workspace_bytes", C.c_int64),
        ("block_n", C.c_int32), ("cta_pair", C.c_int32), ("rows_per_sample", C.c_int32), ("pad2_", C.c_int32),
    ]


class GroupNorm(C.Structure):
    _fields_ = [
        ("x1", C.c_void_p), ("x2", C.c_void_p), ("ld1", C.c_int64), ("ld2", C.c_int64),
        ("C1", C.c_int32), ("C2", C.c_int32), ("NB", C.c_int32), ("HW", C.c_int32), ("groups", C.c_int32),
        ("silu", C.c_int32), ("eps", C.c_float), ("pad_", C.c_int32),
        ("gamma", C.c_void_p), ("beta", C.c_void_p), ("out", C.c_void_p), ("partials", C.c_void_p),
        ("partials_bytes", C.c_int64),
    ]


class LayerNorm(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("out", C.c_void_p), ("ldx", C.c_int64), ("ldo", C.c_int64), ("rows", C.c_int32),
        ("C", C.c_int32), ("eps", C.c_float), ("pad_", C.c_int32), ("gamma", C.c_void_p), ("beta", C.c_void_p),
    ]


class Attention(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("out", C.c_void_p),
        ("ldq", C.c_int64), ("ldk", C.c_int64), ("ldv", C.c_int64), ("ldo", C.c_int64),
        ("B", C.c_int32), ("heads", C.c_int32), ("Nq", C.c_int32), ("Nkv", C.c_int32), ("d", C.c_int32),
        ("scale", C.c_float), ("causal", C.c_int32), ("pad_", C.c_int32),
    ]


class TimestepEmbed(C.Structure):
    _fields_ = [
        ("t", C.c_void_p), ("step_table", C.c_void_p), ("step_counter", C.c_void_p), ("t_stride", C.c_int32),
        ("rows", C.c_int32), ("dim", C.c_int32), ("flip_sin_to_cos", C.c_int32), ("freq_shift", C.c_float),
        ("max_period", C.c_float), ("out", C.c_void_p),
    ]


class PrepLatents(C.Structure):
    _fields_ = [
        ("latents", C.c_void_p), ("out", C.c_void_p), ("step_table", C.c_void_p), ("step_counter", C.c_void_p),
        ("fixed_scale", C.c_float), ("in_dtype", C.c_int32), ("B", C.c_int32), ("C", C.c_int32), ("hw", C.c_int32),
        ("Cpad", C.c_int32), ("rep", C.c_int32),
    ]


class Upsample2x(C.Structure):
    _fields_ = [("x", C.c_void_p), ("out", C.c_void_p), ("NB", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32)]


class PhaseSplit(C.Structure):
    _fields_ = [("x", C.c_void_p), ("out", C.c_void_p), ("NB", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32)]


class SoftmaxRows(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("out", C.c_void_p), ("ldx", C.c_int64), ("ldo", C.c_int64), ("rows", C.c_int32),
        ("cols", C.c_int32), ("scale", C.c_float), ("pad_", C.c_int32),
    ]


class CfgStep(C.Structure):
    _fields_ = [
        ("eps", C.c_void_p), ("ld_eps", C.c_int64), ("x", C.c_void_p), ("x_out", C.c_void_p), ("x0_prev", C.c_void_p),
        ("next_in", C.c_void_p), ("step_table", C.c_void_p), ("step_counter", C.c_void_p),
        ("guidance_scale", C.c_float), ("cfg", C.c_int32), ("x_dtype", C.c_int32), ("B", C.c_int32), ("C", C.c_int32),
        ("hw", C.c_int32), ("Cpad", C.c_int32), ("rep", C.c_int32), ("nsteps", C.c_int32), ("gs_from_table", C.c_int32),
        ("noise", C.c_void_p), ("eps_hist", C.c_void_p), ("x_saved", C.c_void_p),
        ("blend_mask", C.c_void_p), ("blend_image", C.c_void_p), ("blend_noise", C.c_void_p),
    ]


class Axpby(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("y", C.c_void_p), ("out", C.c_void_p), ("a", C.c_float), ("b", C.c_float),
        ("x_dtype", C.c_int32), ("y_dtype", C.c_int32), ("out_dtype", C.c_int32), ("pad_", C.c_int32), ("n", C.c_int64),
        ("coef", C.c_void_p),
    ]


class ImageOut(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("out", C.c_void_p), ("ldx", C.c_int64), ("B", C.c_int32), ("C", C.c_int32),
        ("HW", C.c_int32), ("out_dtype", C.c_int32),
    ]


class StepAdvance(C.Structure):
    _fields_ = [("step_counter", C.c_void_p), ("ack_remote", C.c_void_p), ("state", C.c_void_p)]


class PairExchange(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst_remote", C.c_void_p), ("bytes", C.c_int64), ("sync_local", C.c_void_p),
                ("sync_remote", C.c_void_p), ("state", C.c_void_p), ("timeout_ms", C.c_int32), ("pad_", C.c_int32)]


class ImageU8(C.Structure):
    _fields_ = [("x", C.c_void_p), ("out", C.c_void_p), ("B", C.c_int32), ("C", C.c_int32), ("HW", C.c_int32), ("x_dtype", C.c_int32)]


class Embed(C.Structure):
    _fields_ = [("ids", C.c_void_p), ("tok", C.c_void_p), ("pos", C.c_void_p), ("out", C.c_void_p),
                ("rows", C.c_int32), ("L", C.c_int32), ("D", C.c_int32), ("vocab", C.c_int32)]


class GaussSample(C.Structure):
    _fields_ = [("moments", C.c_void_p), ("noise", C.c_void_p), ("out", C.c_void_p), ("scale", C.c_float),
                ("m_dtype", C.c_int32), ("n_dtype", C.c_int32), ("out_dtype", C.c_int32),
                ("B", C.c_int32), ("C", C.c_int32), ("hw", C.c_int32), ("pad_", C.c_int32)]


class _OpUnion(C.Union):
    _fields_ = [
        ("gemm", Gemm), ("groupnorm", GroupNorm), ("layernorm", LayerNorm), ("attention", Attention),
        ("timestep_embed", TimestepEmbed), ("prep_latents", PrepLatents), ("upsample2x", Upsample2x),
        ("phase_split", PhaseSplit), ("softmax_rows", SoftmaxRows), ("cfg_step", CfgStep), ("axpby", Axpby),
        ("image_out", ImageOut), ("step_advance", StepAdvance), ("image_u8", ImageU8), ("embed", Embed),
        ("gauss_sample", GaussSample), ("pair_exchange", PairExchange),
    ]


class Op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("tag", C.c_int32), ("u", _OpUnion)]


STRUCT_OF_KIND = {
    OP_GEMM: Gemm, OP_GROUPNORM: GroupNorm, OP_LAYERNORM: LayerNorm, OP_ATTENTION: Attention,
    OP_TIMESTEP_EMBED: TimestepEmbed, OP_PREP_LATENTS: PrepLatents, OP_UPSAMPLE2X: Upsample2x,
    OP_PHASE_SPLIT: PhaseSplit, OP_SOFTMAX_ROWS: SoftmaxRows, OP_CFG_STEP: CfgStep, OP_AXPBY: Axpby,
    OP_IMAGE_OUT: ImageOut, OP_STEP_ADVANCE: StepAdvance, OP_IMAGE_U8: ImageU8, OP_EMBED: Embed, OP_GAUSS_SAMPLE: GaussSample, OP_PAIR_EXCHANGE: PairExchange, 100: GemmSeg,
}
FIELD_OF_KIND = {
    OP_GEMM: "gemm", OP_GROUPNORM: "groupnorm", OP_LAYERNORM: "layernorm", OP_ATTENTION: "attention",
    OP_TIMESTEP_EMBED: "timestep_embed", OP_PREP_LATENTS: "prep_latents", OP_UPSAMPLE2X: "upsample2x",
    OP_PHASE_SPLIT: "phase_split", OP_SOFTMAX_ROWS: "softmax_rows", OP_CFG_STEP: "cfg_step", OP_AXPBY: "axpby",
    OP_IMAGE_OUT: "image_out", OP_STEP_ADVANCE: "step_advance", OP_IMAGE_U8: "image_u8", OP_EMBED: "embed", OP_GAUSS_SAMPLE: "gauss_sample", OP_PAIR_EXCHANGE: "pair_exchange",
}

ENTRY_POINTS = {
    "b200sd_abi_version": (C.c_int, []),
    "b200sd_last_error": (C.c_char_p, []),
    "b200sd_sizeof_op": (C.c_size_t, []),
    "b200sd_sizeof_struct": (C.c_size_t, [C.c_int32]),
    "b200sd_device_check": (C.c_int, [C.c_int]),
    "b200sd_set_batch_invariant": (C.c_int, [C.c_int]),
    "b200sd_get_batch_invariant": (C.c_int, []),
    "b200sd_peer_alloc": (C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "b200sd_peer_open": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    "b200sd_peer_close": (C.c_int, [C.c_void_p]),
    "b200sd_peer_free": (C.c_int, [C.c_void_p]),
    "b200sd_groupnorm_static_scratch": (C.c_int, [C.c_int, C.POINTER(GroupNorm), C.c_void_p]),
    "b200sd_groupnorm_chunks": (C.c_int, [C.c_int32, C.c_int32, C.c_int32]),
    "b200sd_gemm_workspace_bytes": (C.c_int64, [C.POINTER(Gemm)]),
    "b200sd_program_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "b200sd_program_add": (C.c_int, [C.c_void_p, C.POINTER(Op)]),
    "b200sd_program_num_ops": (C.c_int, [C.c_void_p]),
    "b200sd_program_num_launches": (C.c_int, [C.c_void_p]),
    "b200sd_program_run": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200sd_program_run_range": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "b200sd_program_run_timed": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.c_int]),
    "b200sd_program_graph_capture": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200sd_program_graph_launch": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200sd_program_destroy": (C.c_int, [C.c_void_p]),
}

_lib = None
_lock = threading.Lock()


def build(force: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into _lib/libb200sd.so (nvcc cross-compiles without a GPU)."""
    if force or not os.path.exists(LIB_PATH) or _stale():
        subprocess.run(["bash", BUILD_SCRIPT], check=True)
    return LIB_PATH


def _stale() -> bool:
    try:
        t = os.path.getmtime(LIB_PATH)
        src_dir = os.path.join(_HERE, "csrc")
        srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir)] + [os.path.join(_HERE, "..", "include", "b200sd.h")]
        return any(os.path.getmtime(s) > t for s in srcs if os.path.exists(s))
    except OSError:
        return True


def lib() -> C.CDLL:
    """The loaded shared library.  Raises (never falls back) when it is missing."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(stablediffusion_b200 has no CPU / eager fallback)")
            l = C.CDLL(LIB_PATH)
            for name, (res, args) in ENTRY_POINTS.items():
                fn = getattr(l, name)
                fn.restype = res
                fn.argtypes = args
            if l.b200sd_abi_version() != ABI_VERSION:
                raise RuntimeError("libb200sd ABI version mismatch")
            if l.b200sd_sizeof_op() != C.sizeof(Op):
                raise RuntimeError(f"b200sd_op layout mismatch: C {l.b200sd_sizeof_op()} vs ctypes {C.sizeof(Op)}")
            _lib = l
    return _lib


def last_error() -> str:
    return lib().b200sd_last_error().decode("utf-8", "replace")


def check(rc: int) -> int:
    """Map a status code to the reference's exception convention (ValueError for bad arguments —
    pipelines/sd_unified_pipeline.py:303,364,391,527; RuntimeError otherwise)."""
    if rc >= 0:
        return rc
    msg = last_error()
    if rc == ERR_INVALID:
        raise ValueError(msg)
    raise RuntimeError(f"libb200sd error {rc}: {msg}")


def set_batch_invariant(on: bool) -> bool:
    """Process-wide switch (include/b200sd.h): reduction-order choices from the per-image shape only, so an image's result
    does not depend on the batch it is computed in.  Affects ops planned AFTER the call; returns the previous setting."""
    return bool(lib().b200sd_set_batch_invariant(1 if on else 0))


def batch_invariant() -> bool:
    return bool(lib().b200sd_get_batch_invariant())


def make_op(kind: int, payload: C.Structure, tag: int = 0) -> Op:
    op = Op()
    op.kind = kind
    op.tag = tag
    setattr(op.u, FIELD_OF_KIND[kind], payload)
    return op


class Program:
    """An op program: add ops once (validated and planned in C), then run / graph-replay on a stream."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        check(lib().b200sd_program_create(C.byref(self._h), int(device)))
        self.device = device
        self.kinds: list[int] = []
        self.tags: list[int] = []
        self._keep: list = []  # tensors that own the memory the ops point at

    def add(self, kind: int, payload: C.Structure, tag: int = 0, keep=()) -> int:
        op = make_op(kind, payload, tag)
        idx = check(lib().b200sd_program_add(self._h, C.byref(op)))
        self.kinds.append(kind)
        self.tags.append(tag)
        self._keep.extend(keep)
        return idx

    def keep(self, *tensors) -> None:
        self._keep.extend(tensors)

    @property
    def num_ops(self) -> int:
        return lib().b200sd_program_num_ops(self._h)

    @property
    def num_launches(self) -> int:
        return lib().b200sd_program_num_launches(self._h)

    def run(self, stream: int) -> None:
        check(lib().b200sd_program_run(self._h, C.c_void_p(stream)))

    def run_range(self, first: int, last: int, stream: int) -> None:
        check(lib().b200sd_program_run_range(self._h, first, last, C.c_void_p(stream)))

    def run_timed(self, stream: int) -> list[float]:
        n = self.num_ops
        buf = (C.c_float * n)()
        check(lib().b200sd_program_run_timed(self._h, C.c_void_p(stream), buf, n))
        return list(buf)

    def graph_capture(self, stream: int) -> None:
        check(lib().b200sd_program_graph_capture(self._h, C.c_void_p(stream)))

    def graph_launch(self, stream: int) -> None:
        check(lib().b200sd_program_graph_launch(self._h, C.c_void_p(stream)))

    def close(self) -> None:
        if self._h:
            lib().b200sd_program_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
