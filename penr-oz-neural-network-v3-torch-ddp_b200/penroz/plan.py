"""Recognise the GPT-2 module tree that mappers.py builds (mappers.py:121-176, and the /model/ DSL of main.py:57-83)
and expose its parameters as a table, without renaming or replacing a single nn.Parameter: state-dict keys
(`layers.2.0.1.weight` ...) are the on-disk format and stay as the reference wrote them.

Tree (top-level `layers` list):
    [0]            Summation(nn.Embedding, PositionEmbedding)
    [1]            nn.Dropout                                     (optional)
    [2 .. 2+L)     ResidualConnection(
                       Sequential(LayerNorm, Linear(C,3C), CausalSelfAttention, Linear(C,C), [Dropout]),
                       Sequential(LayerNorm, Linear(C,F), GELU, Linear(F,C), [Dropout]))
    [2+L]          LayerNorm
    [3+L]          Linear(C, V)                                   (bias optional; untied from wte — SURVEY F1)
    [4+L]          SoftmaxOnLast                                  (optional)
"""
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.nn as nn

import neural_net_layers as nnl


@dataclass
class BlockParams:
    ln1: nn.LayerNorm
    qkv: nn.Linear
    attn: "nnl.CausalSelfAttention"
    proj: nn.Linear
    ln2: nn.LayerNorm
    fc: nn.Linear
    gelu_tanh: bool
    proj2: nn.Linear


@dataclass
class Gpt2Plan:
    wte: nn.Embedding
    wpe: "nnl.PositionEmbedding"
    blocks: list
    lnf: nn.LayerNorm
    head: nn.Linear
    n_head: int
    n_embd: int
    vocab: int
    block_size: int
    has_softmax_last: bool
    dropouts: list = field(default_factory=list)
    n_top_level: int = 0

    @property
    def n_layer(self) -> int:
        return len(self.blocks)

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head

    def active_dropout(self, training: bool) -> bool:
        return training and any(d.p > 0.0 for d in self.dropouts)


def _is_plain(module, cls) -> bool:
    return type(module) is cls


def _ln_ok(m, C) -> bool:
    return (_is_plain(m, nn.LayerNorm) and tuple(m.normalized_shape) == (C,) and m.elementwise_affine
            and m.weight is not None)


def _match_branch(seq, C):
    """Sequential(LayerNorm, Linear, <mid>, Linear, [Dropout]) -> (ln, lin_a, mid, lin_b, dropout|None) or None."""
    if not _is_plain(seq, nn.Sequential):
        return None
    mods = list(seq)
    drop = None
    if mods and _is_plain(mods[-1], nn.Dropout):
        drop = mods.pop()
    if len(mods) != 4:
        return None
    ln, a, mid, b = mods
    if not (_ln_ok(ln, C) and _is_plain(a, nn.Linear) and _is_plain(b, nn.Linear)):
        return None
    if a.in_features != C or b.out_features != C:
        return None
    return ln, a, mid, b, drop


def match_gpt2(layers) -> Optional[Gpt2Plan]:
    """Return a Gpt2Plan when `layers` (the model's top-level ModuleList) is the GPT-2 tree, else None."""
    mods = list(layers)
    if len(mods) < 4:
        return None
    first = mods[0]
    if not (_is_plain(first, nnl.Summation) and len(first) == 2):
        return None
    wte, wpe = first[0], first[1]
    if not (_is_plain(wte, nn.Embedding) and _is_plain(wpe, nnl.PositionEmbedding)):
        return None
    if wte.padding_idx is not None or wte.max_norm is not None or wte.embedding_dim != wpe.embedding_dim:
        return None
    C = wte.embedding_dim
    dropouts = []
    i = 1
    if _is_plain(mods[i], nn.Dropout):
        dropouts.append(mods[i])
        i += 1
    blocks = []
    n_head = None
    while i < len(mods) and _is_plain(mods[i], nnl.ResidualConnection):
        res = mods[i]
        if len(res) != 2:
            return None
        a = _match_branch(res[0], C)
        m = _match_branch(res[1], C)
        if a is None or m is None:
            return None
        ln1, qkv, attn, proj, d1 = a
        ln2, fc, act, proj2, d2 = m
        if not _is_plain(attn, nnl.CausalSelfAttention) or not _is_plain(act, nn.GELU):
            return None
        if attn.rope_theta is not None or attn.num_kv_heads != attn.num_heads:
            return None
        if qkv.out_features != 3 * C or proj.in_features != C or fc.out_features != proj2.in_features:
            return None
        if C % attn.num_heads != 0 or (n_head is not None and attn.num_heads != n_head):
            return None
        n_head = attn.num_heads
        if act.approximate not in ("none", "tanh"):
            return None
        blocks.append(BlockParams(ln1, qkv, attn, proj, ln2, fc, act.approximate == "tanh", proj2))
        dropouts += [d for d in (d1, d2) if d is not None]
        if attn.dropout and attn.dropout > 0.0:
            dropouts.append(nn.Dropout(attn.dropout))
        i += 1
    if not blocks or i + 2 > len(mods):
        return None
    lnf, head = mods[i], mods[i + 1]
    if not (_ln_ok(lnf, C) and _is_plain(head, nn.Linear) and head.in_features == C):
        return None
    rest = mods[i + 2:]
    has_sm = False
    if rest:
        if len(rest) != 1 or not _is_plain(rest[0], nnl.SoftmaxOnLast):
            return None
        has_sm = True
    return Gpt2Plan(wte=wte, wpe=wpe, blocks=blocks, lnf=lnf, head=head, n_head=n_head, n_embd=C,
                    vocab=head.out_features, block_size=wpe.num_embeddings, has_softmax_last=has_sm,
                    dropouts=dropouts, n_top_level=len(mods))
