"""CPU oracle of the CLIP text encoders — TEST INFRASTRUCTURE (see oracle/__init__.py for who may import it).

The reference calls `text_encoder(text_input_ids, output_hidden_states=True)` (pipelines/sd_unified_pipeline.py:587-589,
:655-658) on `transformers.CLIPTextModel` / `CLIPTextModelWithProjection` (models/stable_diffusion.py:124-145) and
reads `[0]`, `.hidden_states[-2]`, `[-1][-(clip_skip + 1)]` and `text_model.final_layer_norm` (:595-609).  The
arithmetic lives in the third-party package `transformers` (requirements.txt pins 4.x; this image carries 5.5 — the
text tower is unchanged between them).  This file restates it in plain fp32 torch with the same state-dict keys.

PINNED: tests/golden/clip_*.pt hold inputs, weights and outputs produced by the installed `transformers` classes
(tests/golden/make_clip_golden.py); tests/test_clip.py checks this restatement against them to fp32 round-off.
"""
import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F


def quick_gelu(x):  # transformers.activations.QuickGELUActivation
    return x * torch.sigmoid(1.702 * x)


ACT = {"quick_gelu": quick_gelu, "gelu": lambda x: F.gelu(x)}


class CLIPTextOracle:
    """config keys: vocab_size, hidden_size, intermediate_size, num_hidden_layers, num_attention_heads,
    max_position_embeddings, hidden_act, layer_norm_eps, eos_token_id, projection_dim (with_projection only)."""

    def __init__(self, config: Dict, state_dict: Dict[str, torch.Tensor], with_projection: bool = False):
        self.config = dict(config)
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items()}
        self.with_projection = with_projection

    def _ln(self, x, prefix):
        return F.layer_norm(x, (x.shape[-1],), self.sd[prefix + ".weight"], self.sd[prefix + ".bias"],
                            self.config.get("layer_norm_eps", 1e-5))

    def final_layer_norm(self, x):
        return self._ln(x.float(), "text_model.final_layer_norm")

    def _lin(self, x, prefix):
        return F.linear(x, self.sd[prefix + ".weight"], self.sd.get(prefix + ".bias"))

    def _layer(self, x, i):
        c = self.config
        B, L, D = x.shape
        H = c["num_attention_heads"]
        d = D // H
        p = f"text_model.encoder.layers.{i}"
        h = self._ln(x, p + ".layer_norm1")
        q = self._lin(h, p + ".self_attn.q_proj") * d ** -0.5          # CLIPAttention scales q
        k = self._lin(h, p + ".self_attn.k_proj")
        v = self._lin(h, p + ".self_attn.v_proj")
        q, k, v = (t.view(B, L, H, d).transpose(1, 2) for t in (q, k, v))
        s = q @ k.transpose(-1, -2)
        causal = torch.full((L, L), float("-inf")).triu(1)              # _create_4d_causal_attention_mask
        a = torch.softmax(s + causal, dim=-1) @ v
        a = a.transpose(1, 2).reshape(B, L, D)
        x = x + self._lin(a, p + ".self_attn.out_proj")
        h = self._ln(x, p + ".layer_norm2")
        h = ACT[c["hidden_act"]](self._lin(h, p + ".mlp.fc1"))
        return x + self._lin(h, p + ".mlp.fc2")

    def __call__(self, input_ids: torch.Tensor, num_layers: Optional[int] = None):
        c = self.config
        ids = input_ids.cpu().long()
        B, L = ids.shape
        x = self.sd["text_model.embeddings.token_embedding.weight"][ids] + \
            self.sd["text_model.embeddings.position_embedding.weight"][:L][None]
        hidden = [x]
        for i in range(c["num_hidden_layers"] if num_layers is None else num_layers):
            x = self._layer(x, i)
            hidden.append(x)
        last = self.final_layer_norm(x)
        if c.get("eos_token_id", 2) == 2:   # legacy configs: the EOS token has the highest id
            eos = ids.argmax(dim=-1)
        else:                               # first occurrence of eos_token_id
            eos = (ids == c["eos_token_id"]).int().argmax(dim=-1)
        pooled = last[torch.arange(B), eos]
        out = {"last_hidden_state": last, "pooler_output": pooled, "hidden_states": tuple(hidden)}
        if self.with_projection:
            out["text_embeds"] = F.linear(pooled, self.sd["text_projection.weight"])
        return out
