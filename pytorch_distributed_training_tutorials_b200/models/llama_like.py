"""A LLaMA-shaped decoder stack with random weights, for the auto-placement demo.

Same module tree as the ``LlamaForCausalLM`` the reference prints (03.model_parallel.ipynb:67-92):
``model.embed_tokens``, ``model.layers[i].{self_attn.{q,k,v,o}_proj, mlp.{gate,up,down}_proj,
input_layernorm, post_attention_layernorm}``, ``model.norm``, ``lm_head``; 9 parameter tensors per
layer, 291 for 32 layers.  Pretrained weights and int8 storage are out of scope (no network, no
bitsandbytes); sizes are configurable so tests stay small.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.eps = eps

    def forward(self, x):
        v = x.float().pow(2).mean(-1, keepdim=True)
        return (x.float() * torch.rsqrt(v + self.eps)).to(x.dtype) * self.weight


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int) -> None:
        super().__init__()
        self.heads = heads
        self.q_proj = nn.Linear(dim, dim, bias=False)
        self.k_proj = nn.Linear(dim, dim, bias=False)
        self.v_proj = nn.Linear(dim, dim, bias=False)
        self.o_proj = nn.Linear(dim, dim, bias=False)

    def forward(self, x):
        B, T, D = x.shape
        h = self.heads
        q = self.q_proj(x).view(B, T, h, D // h).transpose(1, 2)
        k = self.k_proj(x).view(B, T, h, D // h).transpose(1, 2)
        v = self.v_proj(x).view(B, T, h, D // h).transpose(1, 2)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.o_proj(y.transpose(1, 2).reshape(B, T, D))


class MLP(nn.Module):
    def __init__(self, dim: int, hidden: int) -> None:
        super().__init__()
        self.gate_proj = nn.Linear(dim, hidden, bias=False)
        self.up_proj = nn.Linear(dim, hidden, bias=False)
        self.down_proj = nn.Linear(hidden, dim, bias=False)

    def forward(self, x):
        return self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x))


class DecoderLayer(nn.Module):
    def __init__(self, dim: int, heads: int, hidden: int) -> None:
        super().__init__()
        self.self_attn = Attention(dim, heads)
        self.mlp = MLP(dim, hidden)
        self.input_layernorm = RMSNorm(dim)
        self.post_attention_layernorm = RMSNorm(dim)

    def forward(self, x):
        x = x + self.self_attn(self.input_layernorm(x))
        return x + self.mlp(self.post_attention_layernorm(x))


class LlamaModel(nn.Module):
    def __init__(self, vocab: int, dim: int, layers: int, heads: int, hidden: int) -> None:
        super().__init__()
        self.embed_tokens = nn.Embedding(vocab, dim)
        self.layers = nn.ModuleList([DecoderLayer(dim, heads, hidden) for _ in range(layers)])
        self.norm = RMSNorm(dim)

    def forward(self, ids):
        x = self.embed_tokens(ids)
        for layer in self.layers:
            x = layer(x)
        return self.norm(x)


class LlamaLikeForCausalLM(nn.Module):
    def __init__(self, vocab: int = 32000, dim: int = 4096, layers: int = 32, heads: int = 32,
                 hidden: int = 11008) -> None:
        super().__init__()
        self.model = LlamaModel(vocab, dim, layers, heads, hidden)
        self.lm_head = nn.Linear(dim, vocab, bias=False)

    def placement_blocks(self):
        """Blocks in execution order for ``auto_place``: embedding, each decoder layer, norm, head."""
        blocks = [("model.embed_tokens", self.model.embed_tokens)]
        blocks += [(f"model.layers.{i}", l) for i, l in enumerate(self.model.layers)]
        blocks += [("model.norm", self.model.norm), ("lm_head", self.lm_head)]
        return blocks

    def forward(self, ids):
        return self.lm_head(self.model(ids))
