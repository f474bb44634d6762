"""The fused denoising loop: one CUDA-graph launch per sampler step.

What the reference executes per step in Python + a few hundred library launches
(pipelines/sd_unified_pipeline.py:465-489: cat([latents]*2), scale_model_input, UNet, chunk, CFG combine,
scheduler.step) is here ONE captured op program:

    [UNet ops on the already scaled / duplicated NHWC input]
    -> CFG_STEP  (eps = u + s(c-u); scheduler update from the device step table; ALSO writes the next step's
                  scaled, duplicated UNet input, so a1/a2 of SURVEY §8a never exist as separate work)
    -> STEP_ADVANCE (++device step counter)

Latents stay on the device in fp32 for the whole trajectory (the reference rounds them to fp16 every step —
SURVEY App. B.1; keeping an fp32 master is strictly more accurate and is the documented deviation).

Optional cond/uncond GPU-pair split (BASELINE config 5): each rank of a pair runs the UNet on ONE half
of the CFG batch, the eps halves are exchanged once per step, and both ranks then run the same CFG_STEP so latents
stay replicated.  Two exchanges:
  * "peer" (default when the two GPUs can map each other's memory): the exchange is INSIDE the step graph — conv_out
    stores this rank's half into its slot of a peer window, OP_PAIR_EXCHANGE copies the slot into the peer's window over
    NVLink and waits for the peer's half (csrc/peer.cu); still one graph launch per step, no NCCL kernel, no host step
  * "nccl": UNet graph | one all_gather_into_tensor | CFG_STEP graph — the tested fallback
"""
from __future__ import annotations

from typing import Dict, Optional

import ctypes as C

import torch

from . import _native as N
from . import lowering as L
from .unet import B200UNet, UNetLowering, cfg_is_sdxl

MAX_STEPS = 1000


class _RawCuda:
    """Device memory owned by the library (a peer window) seen as a torch tensor (__cuda_array_interface__)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


class PeerWindow:
    """A zero-filled device buffer the pair's other process maps and stores into (include/b200sd.h peer windows).
    Layout: [payload ... | 256 B: int32 sync words {flag, ack} written by the peer]."""
    SYNC_BYTES = 256

    def __init__(self, device: torch.device, payload_bytes: int, group):
        import torch.distributed as dist

        self.device, self.group = device, group
        self.payload_bytes = (payload_bytes + 255) // 256 * 256
        nbytes = self.payload_bytes + self.SYNC_BYTES
        ptr, handle = C.c_void_p(), C.create_string_buffer(N.IPC_HANDLE_BYTES)
        N.check(N.lib().b200sd_peer_alloc(device.index, nbytes, C.byref(ptr), handle))
        self.ptr = ptr.value
        self.remote = None
        mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(device)
        both = [torch.empty_like(mine) for _ in range(2)]
        dist.all_gather(both, mine, group=group)       # (also a barrier: both windows are zeroed before anyone maps them)
        peer = both[1 - dist.get_rank(group)].cpu().numpy().tobytes()
        rptr = C.c_void_p()
        rc = N.lib().b200sd_peer_open(device.index, peer, C.byref(rptr))
        ok = torch.tensor([1 if rc == 0 else 0], device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)   # both ranks take the same path
        if rc == 0:
            self.remote = rptr.value
        if int(ok.item()) == 0:
            err = N.last_error() if rc != 0 else "the peer could not map this rank's window"
            self.close()
            raise RuntimeError(f"peer window unavailable: {err}")
        self.bytes = torch.as_tensor(_RawCuda(self.ptr, nbytes), device=device)

    @property
    def sync_local(self) -> int:
        return self.ptr + self.payload_bytes

    @property
    def sync_remote(self) -> int:
        return self.remote + self.payload_bytes

    def close(self):
        if getattr(self, "remote", None):
            N.lib().b200sd_peer_close(C.c_void_p(self.remote))
            self.remote = None
        if getattr(self, "ptr", None):
            N.lib().b200sd_peer_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DenoiseEngine:
    def __init__(self, unet: B200UNet, B: int, h: int, w: int, do_cfg: bool, ctx_dim: int, uses_history: bool,
                 pair_group=None, pair_rank: int = 0, use_graph: bool = True, eps_history: bool = False,
                 noise_steps: int = 0, blend: bool = False, pair_exchange: str = "auto"):
        if unet.packed is None:
            raise RuntimeError("UNet has no weights on a CUDA device")
        self.unet, self.B, self.h, self.w, self.do_cfg = unet, B, h, w, do_cfg
        self.device = unet.device
        self.pair_group, self.pair_rank = pair_group, pair_rank
        self.split = pair_group is not None
        if self.split and not do_cfg:
            raise ValueError("the cond/uncond pair split needs classifier-free guidance")
        cfg = unet.config
        C = cfg["in_channels"]
        rep_total = 2 if do_cfg else 1
        self.NB = B if self.split else rep_total * B      # UNet batch on THIS rank
        NB = self.NB
        pk = unet.packed
        self.packed = pk  # the programs below point into these tensors: the engine owns a reference
        self.use_graph = use_graph

        # ---- static device state ----------------------------------------------------------------------
        pre = L.Builder(self.device, workspace_bytes=16)
        self.latents = pre.empty([B, C, h, w], torch.float32)
        self.ctx = pre.empty([NB, unet.ctx_len, ctx_dim])
        self.table = pre.zeros([MAX_STEPS, N.STEP_ROW], torch.float32)
        self.counter = pre.zeros([1], torch.int32)
        self.unet_in_full = pre.empty([rep_total * B, h * w, pk.cin_pad])  # both CFG halves (identical rows)
        self.hist = pre.zeros([B, C, h * w], torch.float32) if uses_history else None
        # PNDM / PLMS: three earlier eps + the remembered sample; ancestral / SDE samplers: one standard-normal draw per
        # step, generated for the whole trajectory before the loop (the captured graph indexes it by the step counter)
        self.eps_hist = pre.zeros([3, B, C, h * w], torch.float32) if eps_history else None
        self.x_saved = pre.zeros([B, C, h * w], torch.float32) if eps_history else None
        self.noise = pre.zeros([noise_steps, B, C, h * w], torch.float32) if noise_steps > 0 else None
        # inpainting (pipeline :492-506): mask (1 = repaint), latents of the original image, the start noise
        self.blend = (pre.zeros([B, h * w], torch.float32), pre.zeros([B, C, h * w], torch.float32),
                      pre.zeros([B, C, h * w], torch.float32)) if blend else None
        pre.prep_latents(self.latents, self.unet_in_full, rep=rep_total, step_table=self.table, step_counter=self.counter)
        self.prep = pre

        # ---- the step program -------------------------------------------------------------------------
        b = L.Builder(self.device)
        self.b = b
        unet_in = self.unet_in_full[:NB]  # with the pair split each rank feeds one (identical) half
        kw = {}
        if cfg_is_sdxl(cfg):
            self.text_embeds = b.empty([NB, cfg["projection_class_embeddings_input_dim"] - 6 * cfg["addition_time_embed_dim"]])
            self.time_ids = b.empty([NB, 6], torch.float32)
            kw = dict(text_embeds=self.text_embeds, time_ids=self.time_ids)
        low = UNetLowering(b, pk, NB, h, w, unet.ctx_len)
        # the prompt's cross-attention K/V do not depend on the timestep: projected ONCE per trajectory in the prep program
        # (the reference's attn2.to_k / to_v re-project the same prompt every step, pipeline :475-482)
        kw["ctx_kv"] = low.project_context(pre, self.ctx)
        self.window = None
        self.exchange_kind = None
        eps_shape = [rep_total * B, h * w, cfg["out_channels"]]
        if self.split:
            if pair_exchange not in ("auto", "peer", "nccl"):
                raise ValueError("pair_exchange must be 'auto', 'peer' or 'nccl'")
            self.exchange_kind = "nccl"
            if pair_exchange != "nccl":
                try:
                    self.window = PeerWindow(self.device, 2 * B * h * w * cfg["out_channels"] * 2, pair_group)
                    self.exchange_kind = "peer"
                except RuntimeError:
                    if pair_exchange == "peer":
                        raise
        if self.window is not None:
            n_el = eps_shape[0] * eps_shape[1] * eps_shape[2]
            self.eps_full = self.window.bytes[:2 * n_el].view(torch.bfloat16).view(eps_shape)
            b.prog.keep(self.window)
            self.xstate = b.zeros([4], torch.int32)      # {seq, err, arrival counter, -}
            mine = self.eps_full[pair_rank * B:(pair_rank + 1) * B]
            eps_local = low.emit(unet_in, self.ctx, step_table=self.table, step_counter=self.counter, eps_out=mine, **kw)
            off = mine.data_ptr() - self.window.ptr
            b.pair_exchange(mine, self.window.remote + off, self.window.sync_local, self.window.sync_remote, self.xstate)
        else:
            self.eps_full = b.empty(eps_shape)
            eps_local = low.emit(unet_in, self.ctx, step_table=self.table, step_counter=self.counter, **kw)
        self.eps_local = eps_local
        self.unet_ops = b.prog.num_ops
        if not self.split:
            self.eps_for_step = eps_local
            self._emit_tail(b, eps_local)
            self.tail = None
        elif self.window is not None:
            # ONE graph per step: UNet -> PAIR_EXCHANGE (peer stores + flag) -> CFG_STEP -> STEP_ADVANCE (+ ack)
            self._emit_tail(b, self.eps_full, ack_remote=self.window.sync_remote + 4, state=self.xstate)
            self.tail = None
        else:
            # UNet graph | NCCL all-gather of eps halves | CFG_STEP graph
            tail = L.Builder(self.device, workspace_bytes=16)
            self._emit_tail(tail, self.eps_full)
            self.tail = tail
        self.unet_flops, self.gemm_flops, self.attn_flops = b.flops, b.gemm_flops, b.attn_flops
        self._captured = False
        self._stream = torch.cuda.Stream(self.device)
        # the step loop of every run() is bracketed by two events on the engine's stream (bench.py reads the last one)
        self._ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))

    def _emit_tail(self, b: L.Builder, eps, ack_remote: int = 0, state=None):
        b.cfg_step(eps, self.latents, self.latents, self.table, self.counter, 1.0, self.do_cfg, MAX_STEPS,
                   x0_prev=self.hist, next_in=self.unet_in_full, rep=2 if self.do_cfg else 1, gs_from_table=True,
                   noise=self.noise, eps_hist=self.eps_hist, x_saved=self.x_saved, blend=self.blend)
        b.step_advance(self.counter, ack_remote=ack_remote, state=state)

    @property
    def launches_per_step(self) -> int:
        return self.b.prog.num_launches + (self.tail.prog.num_launches if self.tail else 0)

    # ---- one trajectory ---------------------------------------------------------------------------------
    def load(self, latents, prompt_embeds, step_rows: torch.Tensor, guidance_scale: float, text_embeds=None, time_ids=None,
             generator=None, inpaint=None):
        """Stage one request: latents (any float dtype, NCHW), context embeddings, the scheduler's table rows.
        `generator`: device RNG the per-step noise of a stochastic sampler is drawn from (None: the global one, which
        is what the reference's `scheduler.step(noise_pred, t, latents)` uses, pipeline :489)."""
        n = step_rows.shape[0]
        if n > MAX_STEPS:
            raise ValueError(f"at most {MAX_STEPS} sampler steps")
        if self.noise is not None:
            if n > self.noise.shape[0]:
                raise ValueError(f"engine built for {self.noise.shape[0]} noisy steps, {n} requested")
            self.noise[:n].copy_(torch.randn((n,) + tuple(self.noise.shape[1:]), generator=generator, device=self.device,
                                             dtype=torch.float32))
        elif bool((step_rows[:, N.ST_CN] != 0).any()):
            raise ValueError("this sampler needs per-step noise: build the engine with noise_steps > 0")
        if self.eps_hist is not None:
            self.eps_hist.zero_()
        if (self.blend is None) != (inpaint is None):
            raise ValueError("inpaint inputs and an engine built with blend=True go together")
        if inpaint is not None:  # (mask [B,1,h,w] in {0,1}, image latents [B,C,h,w], start noise [B,C,h,w])
            for dst, src in zip(self.blend, inpaint):
                dst.copy_(src.reshape(dst.shape))
        rows = step_rows.clone()
        rows[:, N.ST_GUIDANCE] = float(guidance_scale)
        rows[0, N.ST_C1] = 0.0  # a multistep solver has no history on the first executed step (img2img starts mid-table)
        self.table[:n].copy_(rows, non_blocking=True)
        self.counter.zero_()
        if self.hist is not None:
            self.hist.zero_()
        self.latents.copy_(latents)
        if self.split:  # rank 0 of a pair holds the uncond half, rank 1 the cond half ([negative; positive] order, :438)
            half = prompt_embeds.shape[0] // 2
            prompt_embeds = prompt_embeds[self.pair_rank * half:(self.pair_rank + 1) * half]
            if text_embeds is not None:
                text_embeds = text_embeds[self.pair_rank * half:(self.pair_rank + 1) * half]
                time_ids = time_ids[self.pair_rank * half:(self.pair_rank + 1) * half]
        self.ctx.copy_(prompt_embeds)
        if text_embeds is not None:
            self.text_embeds.copy_(text_embeds)
            self.time_ids.copy_(time_ids)
        self.n_steps = n

    def run(self, n_steps: Optional[int] = None):
        """Launches prep + n step graphs on the engine's stream (ordered after the caller's current stream)."""
        n = self.n_steps if n_steps is None else n_steps
        s = self._stream
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            self.prep.prog.run(s.cuda_stream)
            if self.use_graph and not self._captured:
                self.b.prog.graph_capture(s.cuda_stream)
                if self.tail is not None:
                    self.tail.prog.graph_capture(s.cuda_stream)
                self._captured = True
            self._ev[0].record(s)
            for _ in range(n):
                self.step_once(s)
            self._ev[1].record(s)
        torch.cuda.current_stream(self.device).wait_stream(s)
        if self.window is not None:  # a spin in the exchange kernel timed out (peer gone / fell out of step): fail loudly
            err = int(self.xstate[1].item())
            if err:
                raise RuntimeError(f"cond/uncond pair exchange timed out (code {err}: 1 = ack, 2 = peer's eps half never arrived)")
        return self.latents

    def last_loop_ms(self) -> float:
        """Device time of the step loop of the most recent run() (synchronises on its end)."""
        self._ev[1].synchronize()
        return self._ev[0].elapsed_time(self._ev[1])

    def step_once(self, s):
        if self.use_graph:
            self.b.prog.graph_launch(s.cuda_stream)
        else:
            self.b.prog.run(s.cuda_stream)
        if self.tail is not None:
            import torch.distributed as dist

            B = self.B
            half = self.eps_full.view(2, B, *self.eps_full.shape[1:])
            # one exchange per step: message = B*4*h*w*2 bytes
            dist.all_gather_into_tensor(self.eps_full, self.eps_local.contiguous(), group=self.pair_group)
            del half
            if self.use_graph:
                self.tail.prog.graph_launch(s.cuda_stream)
            else:
                self.tail.prog.run(s.cuda_stream)

    def run_unet_timed(self):
        """Per-op device times of one step program (profiling aid; leaves latents/counter advanced by one step)."""
        s = self._stream
        with torch.cuda.stream(s):
            return self.b.prog.run_timed(s.cuda_stream)
