"""Deterministic retriever weights under the reference's ``state_dict`` key names.

Key list and shapes follow ``Retriever.__init__`` (src/models/components/retriever.py:130-182);
column order of ``state_net.0.weight`` is ``[iv | sc | err | dist]`` (retriever.py:470-481).
Values come from numpy's PCG64 stream (platform independent), so the golden-vector generator
(which loads them into the *reference* module) and the tests on the GPU box rebuild identical
tensors without shipping 38 MB of weights.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch


def state_dict_shapes(emb_dim: int = 1024, hidden_dim: int = 1024, num_rounds: int = 2, num_rev: int = 2):
    s = 2 * 2 * (1 + num_rounds + num_rev)
    d, h = emb_dim, hidden_dim
    return {
        "entity_proj.network.0.weight": (d, d),
        "entity_proj.network.0.bias": (d,),
        "relation_proj.network.0.weight": (d, d),
        "relation_proj.network.0.bias": (d,),
        "query_proj.network.0.weight": (d, d),
        "query_proj.network.0.bias": (d,),
        "non_text_entity_emb.weight": (1, d),
        "q_gate.0.weight": (d, d),
        "q_gate.0.bias": (d,),
        "q_bias.0.weight": (d, d),
        "q_bias.0.bias": (d,),
        "struct_proj.0.weight": (d, s),
        "struct_proj.0.bias": (d,),
        "struct_proj.1.weight": (d,),
        "struct_proj.1.bias": (d,),
        "struct_gate_net.0.weight": (1, d),
        "struct_gate_net.0.bias": (1,),
        "state_net.0.weight": (h, 3 * d + 1),
        "state_net.0.bias": (h,),
        "state_net.1.weight": (h,),
        "state_net.1.bias": (h,),
        "state_net.4.weight": (h, h),
        "state_net.4.bias": (h,),
        "score_head.weight": (1, h),
        "score_head.bias": (1,),
    }


def make_state_dict(
    seed: int = 0,
    emb_dim: int = 1024,
    hidden_dim: int = 1024,
    num_rounds: int = 2,
    num_rev: int = 2,
    trained_like: bool = False,
) -> Dict[str, torch.Tensor]:
    """nn.Linear-style U(-1/sqrt(fan_in), 1/sqrt(fan_in)) weights; LayerNorm gamma ~ 1 +/- 0.1,
    beta ~ +/- 0.05 so the affine part is exercised. ``trained_like`` rescales the score head so
    logits have std of a few units (random init gives a 0.03-0.5 logit span)."""
    rng = np.random.default_rng([seed, emb_dim, hidden_dim, num_rounds, num_rev])
    sd: Dict[str, torch.Tensor] = {
        "parity_meta": torch.tensor([1, 2, num_rounds, num_rev], dtype=torch.long)
    }
    for name, shape in state_dict_shapes(emb_dim, hidden_dim, num_rounds, num_rev).items():
        u = rng.random(shape, dtype=np.float64) * 2.0 - 1.0
        if name in ("struct_proj.1.weight", "state_net.1.weight"):
            v = 1.0 + 0.1 * u
        elif name in ("struct_proj.1.bias", "state_net.1.bias"):
            v = 0.05 * u
        elif name == "non_text_entity_emb.weight":
            v = 1.7 * u
        else:
            fan_in = shape[-1] if len(shape) == 2 else emb_dim
            v = u / np.sqrt(fan_in)
        sd[name] = torch.from_numpy(v.astype(np.float32))
    if trained_like:
        sd["score_head.weight"] = sd["score_head.weight"] * 40.0
    return sd
