"""Shared test helpers: build models through the drop-in modules inside a scratch working directory."""
import copy
import os
import tempfile

import numpy as np
import torch


def scratch_cwd():
    d = tempfile.mkdtemp(prefix="penroz_test_")
    os.chdir(d)
    os.makedirs("shm", exist_ok=True)
    import neural_net_model as nnm
    nnm.NeuralNetworkModel.SHM_PATH = os.path.join(d, "shm")
    return d


def build_gpt(model_id, L, H, C, T, V, seed=0, **kw):
    import neural_net_model as nnm
    from penroz import dsl
    torch.manual_seed(seed)
    return nnm.NeuralNetworkModel(model_id, dsl.Mapper(dsl.gpt2_layers(L, H, C, T, V, **kw),
                                                       copy.deepcopy(dsl.DEFAULT_ADAMW)))


def load_state(model, g, prefix):
    sd = {k: torch.from_numpy(g[prefix + k]) for k in model.state_dict().keys()}
    model.load_state_dict(sd)
