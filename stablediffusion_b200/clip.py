"""CLIP text encoders on the B200 kernels.

Drop-in for what the reference's `encode_prompt` touches on `model.text_encoder` / `model.text_encoder_2`
(pipelines/sd_unified_pipeline.py:587-609, :655-665; classes loaded at models/stable_diffusion.py:124-145):

    out = text_encoder(input_ids, output_hidden_states=True)
    out[0]                      last_hidden_state (CLIPTextModel) / text_embeds (CLIPTextModelWithProjection)
    out.hidden_states[-2]       penultimate hidden state (SDXL, and clip-skip)
    out[-1][-(clip_skip + 1)]   the hidden-state tuple again
    text_encoder.text_model.final_layer_norm(h)
    text_encoder.dtype, text_encoder.config

One op program per batch size: embed -> L x [LN, fused QKV GEMM, causal tcgen05 attention, out-proj GEMM (+residual),
LN, fc1 GEMM (+quick_gelu / gelu), fc2 GEMM (+residual)] -> final LN.  The pooled row is picked by index arithmetic on
the token ids (argmax for the legacy eos_token_id == 2 configs, first EOS otherwise — transformers' rule) and, for the
projection variant, goes through one more GEMM.  Weights live as bf16 on the device (q/k/v concatenated), biases and
LayerNorm parameters as fp32.  There is no CPU path.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import lowering as L


def random_clip_state_dict(cfg: Dict, with_projection: bool = False, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Random weights of the exact architecture with transformers' key names (synthetic benchmarks / tests: no
    checkpoints are reachable offline).  N(0, 1/fan_in) matrices so that activations stay O(1) through the stack."""
    g = torch.Generator().manual_seed(seed)
    D, I, V, P = cfg["hidden_size"], cfg["intermediate_size"], cfg["vocab_size"], cfg["max_position_embeddings"]
    rn = lambda *s, std=1.0: torch.randn(*s, generator=g) * std
    sd = {
        "text_model.embeddings.token_embedding.weight": rn(V, D, std=0.5),
        "text_model.embeddings.position_embedding.weight": rn(P, D, std=0.5),
        "text_model.final_layer_norm.weight": 1 + rn(D, std=0.1),
        "text_model.final_layer_norm.bias": rn(D, std=0.1),
    }
    for i in range(cfg["num_hidden_layers"]):
        p = f"text_model.encoder.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            sd[p + f"self_attn.{n}.weight"] = rn(D, D, std=D ** -0.5)
            sd[p + f"self_attn.{n}.bias"] = rn(D, std=0.1)
        sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"] = rn(I, D, std=D ** -0.5), rn(I, std=0.1)
        sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"] = rn(D, I, std=I ** -0.5), rn(D, std=0.1)
        for n in ("layer_norm1", "layer_norm2"):
            sd[p + n + ".weight"], sd[p + n + ".bias"] = 1 + rn(D, std=0.1), rn(D, std=0.1)
    if with_projection:
        sd["text_projection.weight"] = rn(cfg["projection_dim"], D, std=D ** -0.5)
    return sd


class CLIPOutput:
    """Indexable like transformers' ModelOutput (None fields are skipped), with the same attribute names."""

    def __init__(self, **fields):
        self._order = [k for k, v in fields.items() if v is not None]
        for k, v in fields.items():
            setattr(self, k, v)

    def __getitem__(self, i):
        if isinstance(i, str):
            return getattr(self, i)
        return getattr(self, self._order[i])

    def __len__(self):
        return len(self._order)


class _Config(dict):
    __getattr__ = dict.get


class _TextModel:
    """`text_encoder.text_model` as far as the reference uses it: `.final_layer_norm(hidden)` (:608)."""

    def __init__(self, owner: "B200CLIPTextModel"):
        self._owner = owner

    def final_layer_norm(self, hidden: torch.Tensor) -> torch.Tensor:
        return self._owner._final_ln(hidden)


class B200CLIPTextModel:
    def __init__(self, config: Dict, weights: Dict[str, torch.Tensor], with_projection: bool, device):
        self.config = _Config(config)
        self.with_projection = with_projection
        self.device = torch.device(device)
        self.dtype = torch.bfloat16
        self.W = weights
        self.text_model = _TextModel(self)
        self._programs: Dict[int, dict] = {}
        self._ln_programs: Dict[tuple, dict] = {}
        d = config["hidden_size"] // config["num_attention_heads"]
        if config["hidden_size"] % 8 or d % 8 or config["intermediate_size"] % 8:
            raise ValueError("CLIP dims must be multiples of 8")
        if config["hidden_act"] not in ("quick_gelu", "gelu"):
            raise ValueError(f"unsupported hidden_act {config['hidden_act']}")

    # ---- construction -------------------------------------------------------------------------------
    @staticmethod
    def _pack(config: Dict, sd: Dict[str, torch.Tensor], with_projection: bool, dev) -> Dict[str, torch.Tensor]:
        bf = lambda t: t.detach().to(device=dev, dtype=torch.bfloat16).contiguous()
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        W = {
            "tok": bf(sd["text_model.embeddings.token_embedding.weight"]),
            "pos": bf(sd["text_model.embeddings.position_embedding.weight"]),
            "lnf.w": f32(sd["text_model.final_layer_norm.weight"]), "lnf.b": f32(sd["text_model.final_layer_norm.bias"]),
        }
        for i in range(config["num_hidden_layers"]):
            p = f"text_model.encoder.layers.{i}."
            W[f"{i}.qkv.w"] = bf(torch.cat([sd[p + f"self_attn.{n}_proj.weight"] for n in "qkv"], 0))
            W[f"{i}.qkv.b"] = f32(torch.cat([sd[p + f"self_attn.{n}_proj.bias"] for n in "qkv"], 0))
            W[f"{i}.out.w"], W[f"{i}.out.b"] = bf(sd[p + "self_attn.out_proj.weight"]), f32(sd[p + "self_attn.out_proj.bias"])
            W[f"{i}.fc1.w"], W[f"{i}.fc1.b"] = bf(sd[p + "mlp.fc1.weight"]), f32(sd[p + "mlp.fc1.bias"])
            W[f"{i}.fc2.w"], W[f"{i}.fc2.b"] = bf(sd[p + "mlp.fc2.weight"]), f32(sd[p + "mlp.fc2.bias"])
            for k, n in (("ln1", "layer_norm1"), ("ln2", "layer_norm2")):
                W[f"{i}.{k}.w"], W[f"{i}.{k}.b"] = f32(sd[p + n + ".weight"]), f32(sd[p + n + ".bias"])
        if with_projection:
            W["proj.w"] = bf(sd["text_projection.weight"])
        return W

    @classmethod
    def from_state_dict(cls, config: Dict, sd: Dict[str, torch.Tensor], with_projection: bool = False, device="cuda"):
        dev = torch.device(device)
        return cls(config, cls._pack(config, sd, with_projection, dev), with_projection, dev)

    @classmethod
    def from_transformers(cls, module, device="cuda"):
        """module: transformers.CLIPTextModel / CLIPTextModelWithProjection (what `load_hf_checkpoint` reads).  The
        torch module is kept as the host master copy of the weights: text-encoder LoRA is merged into it
        (models/stable_diffusion.py:275-309 semantics) and `refresh_weights()` re-packs the device copy in place."""
        c = module.config
        cfg = {k: getattr(c, k) for k in ("vocab_size", "hidden_size", "intermediate_size", "num_hidden_layers",
                                          "num_attention_heads", "max_position_embeddings", "hidden_act",
                                          "layer_norm_eps", "eos_token_id", "projection_dim")}
        sd = module.state_dict()
        self = cls.from_state_dict(cfg, sd, with_projection="text_projection.weight" in sd, device=device)
        self.torch_module = module
        return self

    def named_modules(self):  # LoRA bookkeeping addresses the torch master copy by module path
        return self.torch_module.named_modules()

    def refresh_weights(self):
        """Re-pack the device weights from the torch master copy, in place (cached programs keep their pointers)."""
        fresh = self._pack(self.config, self.torch_module.state_dict(), self.with_projection, self.device)
        for k, v in fresh.items():
            self.W[k].copy_(v)

    def to(self, *a, **k):  # the wrapper calls .to(device) on every sub-model
        return self

    # ---- programs -----------------------------------------------------------------------------------
    def _program(self, B: int, Lseq: int) -> dict:
        st = self._programs.get((B, Lseq))
        if st is not None:
            return st
        c, W = self.config, self.W
        D, H = c["hidden_size"], c["num_attention_heads"]
        d = D // H
        eps = c.get("layer_norm_eps", 1e-5)
        b = L.Builder(self.device)
        st = {"ids": torch.zeros(B * Lseq, dtype=torch.int32, device=self.device)}
        b.prog.keep(st["ids"])
        x = b.embed(st["ids"], W["tok"], W["pos"][:Lseq])
        hidden = [x]
        for i in range(c["num_hidden_layers"]):
            h = b.layernorm(x, W[f"{i}.ln1.w"], W[f"{i}.ln1.b"], eps)
            qkv = b.linear(h, W[f"{i}.qkv.w"], bias=W[f"{i}.qkv.b"]).view(B, Lseq, 3 * D)
            a = b.attention(qkv[:, :, :D], qkv[:, :, D:2 * D], qkv[:, :, 2 * D:], H, d, d ** -0.5, causal=True)
            x = b.linear(a.view(B * Lseq, D), W[f"{i}.out.w"], bias=W[f"{i}.out.b"], residual=x)
            h = b.layernorm(x, W[f"{i}.ln2.w"], W[f"{i}.ln2.b"], eps)
            h = b.linear(h, W[f"{i}.fc1.w"], bias=W[f"{i}.fc1.b"], act=c["hidden_act"])
            x = b.linear(h, W[f"{i}.fc2.w"], bias=W[f"{i}.fc2.b"], residual=x)
            hidden.append(x)
        st["last"] = b.layernorm(x, W["lnf.w"], W["lnf.b"], eps)
        st["hidden"] = hidden
        st["prog"] = b.prog
        if self.with_projection:
            pb = L.Builder(self.device)
            st["pooled_in"] = pb.empty([B, D])
            st["text_embeds"] = pb.linear(st["pooled_in"], W["proj.w"])
            st["proj_prog"] = pb.prog
        self._programs[(B, Lseq)] = st
        return st

    def _final_ln(self, hidden: torch.Tensor) -> torch.Tensor:
        if not hidden.is_cuda:
            raise RuntimeError("final_layer_norm: the hidden state must be on the sm_100 device (no CPU fallback)")
        key = tuple(hidden.shape)
        st = self._ln_programs.get(key)
        if st is None:
            b = L.Builder(self.device)
            st = {"x": b.empty([hidden.numel() // hidden.shape[-1], hidden.shape[-1]])}
            st["y"] = b.layernorm(st["x"], self.W["lnf.w"], self.W["lnf.b"], self.config.get("layer_norm_eps", 1e-5))
            st["prog"] = b.prog
            self._ln_programs[key] = st
        st["x"].copy_(hidden.reshape(st["x"].shape))
        st["prog"].run(torch.cuda.current_stream(self.device).cuda_stream)
        return st["y"].view(hidden.shape).clone()

    # ---- forward ------------------------------------------------------------------------------------
    def __call__(self, input_ids: torch.Tensor, attention_mask=None, output_hidden_states: Optional[bool] = None, **kw):
        if attention_mask is not None:
            raise ValueError("attention_mask is not supported (the reference never passes one: :587-589)")
        if input_ids.dim() != 2:
            raise ValueError("input_ids must be [batch, tokens]")
        if not input_ids.is_cuda:
            raise RuntimeError("the B200 text encoder needs input_ids on the sm_100 device (no CPU fallback)")
        B, Lseq = input_ids.shape
        if Lseq > self.config["max_position_embeddings"]:
            raise ValueError(f"sequence of {Lseq} tokens exceeds max_position_embeddings")
        st = self._program(B, Lseq)
        stream = torch.cuda.current_stream(self.device)
        st["ids"].copy_(input_ids.reshape(-1))
        st["prog"].run(stream.cuda_stream)
        D = self.config["hidden_size"]
        last = st["last"].view(B, Lseq, D).clone()
        hidden = tuple(h.view(B, Lseq, D).clone() for h in st["hidden"]) if output_hidden_states else None
        # pooled row: index arithmetic on the ids (transformers: argmax for the legacy eos id 2, else the first EOS)
        ids = input_ids.to(torch.int64)
        if self.config.get("eos_token_id", 2) == 2:
            eos = ids.argmax(dim=-1)
        else:
            eos = (ids == self.config["eos_token_id"]).int().argmax(dim=-1)
        pooled = last[torch.arange(B, device=last.device), eos]
        if self.with_projection:
            st["pooled_in"].copy_(pooled)
            st["proj_prog"].run(stream.cuda_stream)
            return CLIPOutput(text_embeds=st["text_embeds"].clone(), last_hidden_state=last, hidden_states=hidden)
        return CLIPOutput(last_hidden_state=last, pooler_output=pooled, hidden_states=hidden)
