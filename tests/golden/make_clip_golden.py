"""Generates tests/golden/clip_l_tiny.pt and clip_g_tiny.pt from the INSTALLED transformers CLIP text towers.

Run in the build container (`python tests/golden/make_clip_golden.py`); the GPU box only reads the fixtures.
Two miniature towers cover both variants the reference loads (models/stable_diffusion.py:124-145):
  clip_l_tiny  CLIPTextModel,               quick_gelu, eos_token_id = 2 (legacy: pooled row = argmax of the ids)
  clip_g_tiny  CLIPTextModelWithProjection, gelu,       eos_token_id = 99 (pooled row = first EOS), projection
"""
import os

import torch
import transformers
from transformers import CLIPTextConfig, CLIPTextModel, CLIPTextModelWithProjection

HERE = os.path.dirname(os.path.abspath(__file__))


def make(name, cls, seed, **kw):
    torch.manual_seed(seed)
    cfg = CLIPTextConfig(vocab_size=100, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                         num_attention_heads=2, max_position_embeddings=77, layer_norm_eps=1e-5,
                         bos_token_id=0, pad_token_id=1, **kw)
    model = cls(cfg).eval().float()
    with torch.no_grad():  # the default init is tiny (std 0.02): widen it so that every term matters
        for n, p in model.named_parameters():
            if p.dim() == 2 and "embedding" not in n:
                p.normal_(0, p.shape[1] ** -0.5)
            elif "bias" in n:
                p.normal_(0, 0.1)
            elif "embedding" in n:
                p.normal_(0, 0.5)
            elif "layer_norm" in n and n.endswith("weight"):
                p.normal_(1.0, 0.1)
    g = torch.Generator().manual_seed(seed + 1)
    ids = torch.randint(3, 98, (3, 77), generator=g)
    ids[:, 0] = 0
    eos = cfg.eos_token_id
    for b, n in enumerate((5, 40, 76)):   # prompt lengths: EOS at n, padding (= EOS for the second tokenizer) after
        ids[b, n] = eos if eos != 2 else 99
        ids[b, n + 1:] = eos if eos != 2 else 1
    # (legacy pooling takes the arg-max id: 99 is the largest id present and marks the end of the prompt)
    with torch.no_grad():
        out = model(input_ids=ids, output_hidden_states=True)
    blob = {
        "transformers_version": transformers.__version__,
        "config": {k: getattr(cfg, k) for k in ("vocab_size", "hidden_size", "intermediate_size", "num_hidden_layers",
                                                "num_attention_heads", "max_position_embeddings", "hidden_act",
                                                "layer_norm_eps", "eos_token_id", "projection_dim")},
        "with_projection": cls is CLIPTextModelWithProjection,
        "state_dict": {k: v.clone() for k, v in model.state_dict().items() if "position_ids" not in k},
        "input_ids": ids,
        "last_hidden_state": out.last_hidden_state,
        "hidden_states": torch.stack(out.hidden_states),
    }
    if cls is CLIPTextModelWithProjection:
        blob["text_embeds"] = out.text_embeds
    else:
        blob["pooler_output"] = out.pooler_output
    torch.save(blob, os.path.join(HERE, name + ".pt"))
    print(name, {k: (tuple(v.shape) if torch.is_tensor(v) else v) for k, v in blob.items() if k != "state_dict"})


if __name__ == "__main__":
    make("clip_l_tiny", CLIPTextModel, 11, hidden_act="quick_gelu", eos_token_id=2, projection_dim=64)
    make("clip_g_tiny", CLIPTextModelWithProjection, 12, hidden_act="gelu", eos_token_id=99, projection_dim=96)
