"""Stand-in for the reference's `mappers.Mapper` when that module is not importable.

`mappers.py` is NOT one of the four modules this package replaces: inside the reference service the real one is
used (neural_net_model imports it first). On a box without the reference checkout (the GPU box, bench.py, tests)
this module builds the same nn.Module tree from the same JSON layer DSL, drawing from torch's global RNG in the
same order (construction first, then the init overrides, mappers.py:72-99), so `torch.manual_seed(s)` gives
bit-identical initial weights — tests/test_host_logic.py pins that against the golden init state dict.

Only the DSL -> modules and DSL -> optimizer directions are provided (what NeuralNetworkModel.__init__ and
deserialize need). HuggingFace config/state-dict mapping stays in the reference's mappers.py.
"""
from typing import Any, Iterable

import torch
import torch.nn as nn
import torch.optim as optim

import neural_net_layers as nnl

_LAYER_KINDS = {
    "embedding": nn.Embedding, "linear": nn.Linear, "flatten": nn.Flatten, "batchnorm1d": nn.BatchNorm1d,
    "relu": nn.ReLU, "gelu": nn.GELU, "sigmoid": nn.Sigmoid, "softmax": nn.Softmax, "tanh": nn.Tanh,
    "dropout": nn.Dropout, "sequential": nn.Sequential, "layernorm": nn.LayerNorm,
    "attention": nnl.CausalSelfAttention, "summation": nnl.Summation, "residual": nnl.ResidualConnection,
    "position": nnl.PositionEmbedding, "softmaxlast": nnl.SoftmaxOnLast, "rmsnorm": nnl.RMSNorm,
    "gatedmlp": nnl.GatedMLP, "scaledembedding": nnl.ScaledEmbedding, "transformerblock": nnl.TransformerBlock,
}
_WEIGHT_INITS = {"xavier_uniform": nn.init.xavier_uniform_, "kaiming_uniform": nn.init.kaiming_uniform_,
                 "normal": nn.init.normal_}
_BIAS_INITS = {"zeros": nn.init.zeros_}
_OPTIMIZERS = {"adam": optim.Adam, "adamw": optim.AdamW, "sgd": optim.SGD}


def _first_known(spec: dict, table: dict):
    for key, val in spec.items():
        if key in table:
            return table[key], val
    return None, None


def build_layer(spec: dict) -> nn.Module:
    ctor, args = _first_known(spec, _LAYER_KINDS)
    if ctor is None:
        raise ValueError(f"Unsupported layer: {spec}")
    if isinstance(args, dict):
        # nested layer specs are built (and draw their RNG) before the parent is constructed
        for name, val in list(args.items()):
            if isinstance(val, dict):
                args[name] = build_layer(val)
        module = ctor(**args)
    elif isinstance(args, list):
        module = ctor(*[build_layer(a) if isinstance(a, dict) else a for a in args])
    else:
        module = ctor(args)
    w_init, w_args = _first_known(spec, _WEIGHT_INITS)
    if w_init is not None:
        for sub in _post_order(module):
            if hasattr(sub, "weight"):
                w_init(sub.weight, **w_args)
    b_init, b_args = _first_known(spec, _BIAS_INITS)
    if b_init is not None:
        for sub in _post_order(module):
            if hasattr(sub, "bias"):
                b_init(sub.bias, **b_args)
    confidence = spec.get("confidence")
    if confidence is not None:
        with torch.no_grad():
            for sub in _post_order(module):
                sub.weight *= confidence
    return module


def _post_order(module: nn.Module):
    """Children before the module itself: the visiting order of nn.Module.apply."""
    for child in module.children():
        yield from _post_order(child)
    yield module


class Mapper:
    def __init__(self, layers: list[dict], optimizer: dict):
        self.layers = layers
        self.optimizer = optimizer

    def to_layers(self) -> list[nn.Module]:
        return [build_layer(spec) for spec in self.layers]

    def to_optimizer(self, params: Iterable[Any]):
        ctor, args = _first_known(self.optimizer, _OPTIMIZERS)
        if ctor is None:
            raise ValueError(f"Unsupported optimizer: {self.optimizer}")
        if "betas" in args:
            args["betas"] = tuple(args["betas"])
        return ctor(params, **args)


def gpt2_layers(n_layer: int, n_head: int, n_embd: int, block_size: int, vocab: int, *, tanh_gelu=False,
                ln_bias=True, dropout=0.0) -> list[dict]:
    """The /model/ request body documented at main.py:57-83 for an (untied-head) GPT-2 of the given size."""
    import math
    proj_std = 0.02 / math.sqrt(2 * n_layer)
    ln = {"normalized_shape": n_embd} if ln_bias else {"normalized_shape": n_embd, "bias": False}
    gelu = {"gelu": {"approximate": "tanh"}} if tanh_gelu else {"gelu": {}}

    def lin(i, o, std):
        return {"linear": {"in_features": i, "out_features": o}, "normal": {"mean": 0.0, "std": std}, "zeros": {}}

    spec = [{"summation": [
        {"embedding": {"num_embeddings": vocab, "embedding_dim": n_embd}, "normal": {"mean": 0.0, "std": 0.02}},
        {"position": {"num_embeddings": block_size, "embedding_dim": n_embd}, "normal": {"mean": 0.0, "std": 0.02}}]},
        {"dropout": {"p": dropout}}]
    for _ in range(n_layer):
        spec.append({"residual": [
            {"sequential": [{"layernorm": dict(ln)}, lin(n_embd, 3 * n_embd, 0.02),
                            {"attention": {"num_heads": n_head, "dropout": dropout}},
                            lin(n_embd, n_embd, proj_std), {"dropout": {"p": dropout}}]},
            {"sequential": [{"layernorm": dict(ln)}, lin(n_embd, 4 * n_embd, 0.02), dict(gelu),
                            lin(4 * n_embd, n_embd, proj_std), {"dropout": {"p": dropout}}]}]})
    spec += [{"layernorm": dict(ln)},
             {"linear": {"in_features": n_embd, "out_features": vocab, "bias": False}},
             {"softmaxlast": {"dim": -1}}]
    return spec


DEFAULT_ADAMW = {"adamw": {"lr": 6e-4, "betas": [0.9, 0.95], "eps": 1e-8}}
