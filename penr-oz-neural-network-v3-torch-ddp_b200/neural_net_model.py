"""Drop-in for the reference's neural_net_model.py: same class, methods, persistence format and status/progress/stats
schemas (reference: neural_net_model.py:28-779), with the GPT-2 training step and KV-cache decode step routed
through the fused sm_100a engine (penroz/engine.py) whenever the parameters live on a CUDA device.

Routing
  * GPT-2 tree (penroz/plan.py) + CUDA parameters  -> penroz engine (hand-written kernels behind the C ABI). This is
    the product path; it raises if the CUDA library is missing or the step cannot be fused — it never falls back.
  * anything else (other layer DSLs such as MLPs or the Gemma family; parameters on the CPU) -> the module
    composition the reference itself performs with torch modules built by mappers.py. Out of scope for kernels
    (SURVEY.md §8), kept so the service's other model families and its CPU test-suite behave as before.

Deliberate deviations from the reference, all behaviour-preserving fixes:
  * stats histograms cast bf16 tensors to fp32 first (the reference crashes on bf16 after the last epoch);
  * /generate/, /output/, /evaluate/ move a GPT-2 model to the GPU when one is visible (the reference leaves it on
    the CPU); sampling still draws from torch's CPU generator so seeded generations keep their stream.
Reference semantics that change numbers are kept: untied LM head, weight decay on every parameter, no clipping,
micro-step == full (batch_size, block_size) batch, accumulation sync as in the reference (PENROZ_ACCUM_SYNC=correct
opts into syncing on the last micro-step instead).
"""
import logging
import multiprocessing
import os
import platform
import random
import shutil
import tempfile
import threading
import time
from collections import OrderedDict
from contextlib import nullcontext
from datetime import datetime as dt
from typing import Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn
from torch import Tensor
from torch.optim import Optimizer

import ddp
from kv_cache import KVCache, TurboQuantKVCache, create_kv_cache
from neural_net_layers import CausalSelfAttention, PositionEmbedding, SoftmaxOnLast, TransformerBlock

try:  # inside the reference service these are the reference's own (unreplaced) modules
    from loaders import Loader
except Exception:  # no reference checkout / no `datasets`: same window arithmetic, restated
    from penroz.shards import Loader
try:
    from mappers import Mapper
except Exception:
    from penroz.dsl import Mapper
try:
    from transformers import AutoConfig, AutoModelForCausalLM
except Exception:  # pragma: no cover
    AutoConfig = AutoModelForCausalLM = None

from penroz.plan import match_gpt2

log = logging.getLogger(__name__)
MODELS_FOLDER = "models"

# ---------------------------------------------------------------------------------------------------- residency
# The service deserializes the model on EVERY /generate/, /output/, /evaluate/, /progress/ and /stats/ request
# (main.py:405,415,424,478,489): torch.load of weights + AdamW moments (2.6 GB for 124M) and, for inference, a move to
# the GPU -- seconds, against tens of milliseconds for the generation itself. SURVEY 8f row n2: keep the last few
# models resident in the process. An entry is valid while the cached file is byte-for-byte the one it was loaded from
# (same mtime_ns and size), so a training process rewriting the file invalidates it; the on-disk format is unchanged.
# Those requests only read the model, but generation attaches its KV cache to the layers, so calls that run the model
# take the model's run lock. The lock lives on the model object (a caller may still hold the object after the LRU has
# dropped its entry) and is a plain non-reentrant Lock: the service drives generate_tokens_stream through Starlette's
# StreamingResponse, one threadpool dispatch per next() (main.py:427-431), so the thread that finishes or closes the
# stream need not be the one that started it -- an RLock could not be released there, and would let a second request
# scheduled onto the owner's thread walk straight in. PENROZ_MODEL_CACHE=0|1 (default: on when a GPU is visible),
# PENROZ_MODEL_CACHE_SIZE (2).
_RESIDENT: "OrderedDict[str, tuple]" = OrderedDict()      # cached-file path -> (mtime_ns, size, model)
_RESIDENT_GUARD = threading.Lock()


def _residency_on() -> bool:
    flag = os.environ.get("PENROZ_MODEL_CACHE")
    if flag is not None:
        return flag not in ("0", "false", "False", "")
    return torch.cuda.is_available()


def _file_stamp(path: str):
    st = os.stat(path)
    return st.st_mtime_ns, st.st_size


class _Exclusive:
    """Context manager: the model's run lock when `model` was handed out by the residency cache (and so may be shared
    between request threads), else nothing. Acquire and release may happen on different threads."""

    def __init__(self, model):
        self.lock = model.__dict__.get("_run_lock")

    def __enter__(self):
        if self.lock is not None:
            self.lock.acquire()
        return self

    def __exit__(self, *exc):
        if self.lock is not None:
            self.lock.release()
        return False


def _now_status(code: str, message: str) -> dict:
    return {"code": code, "dt": dt.now().isoformat(), "message": message}


class NeuralNetworkModel(nn.Module):

    @staticmethod
    def _detect_shm_path() -> str:
        """RAM-backed folder for the model cache: /dev/shm (Linux), /Volumes/RAMDisk (macOS), else the temp dir."""
        candidates = {"Linux": "/dev/shm", "Darwin": "/Volumes/RAMDisk"}
        path = candidates.get(platform.system())
        if path and os.path.isdir(path) and os.access(path, os.W_OK):
            return path
        return tempfile.gettempdir()

    SHM_PATH = _detect_shm_path()

    def __init__(self, model_id: str, mapper: Mapper):
        super().__init__()
        self.model_id = model_id
        self.mapper = mapper
        self.layers = nn.ModuleList(self.mapper.to_layers())
        self._is_softmax_last = isinstance(self.layers[-1], nn.Softmax)
        self.optimizer: Optimizer = self.mapper.to_optimizer(self.parameters())
        self.progress = []
        self.avg_cost = None
        self.avg_cost_history = []
        self.stats = None
        self.status = _now_status("Created", "Model created but not yet trained.")
        self._plan = match_gpt2(self.layers)
        self._fused = None

    # ------------------------------------------------------------------ small properties
    @property
    def _weights(self) -> list[Tensor]:
        """One entry per parameter: the tensor itself when it is a 2-D weight, else None."""
        return [p if p.ndim == 2 else None for p in self.parameters()]

    @property
    def num_params(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def to(self, device: str = None, dtype: torch.dtype = None):
        if device is not None:
            if device == 'cuda' and ddp.is_ddp():
                device = f"cuda:{ddp.ddp_local_rank()}"
                torch.cuda.set_device(device)
            super().to(device)
            self._fused = None
        if dtype is not None:
            super().to(dtype=dtype)
            self._fused = None

    # ------------------------------------------------------------------ fused engine access
    def _engine(self):
        """The fused engine for this model, or None when the tree is not GPT-2 or the parameters are not on CUDA."""
        if self._plan is None:
            return None
        first = self._plan.wte.weight
        if first.device.type != "cuda":
            return None
        if self._fused is None or self._fused.device != first.device:
            from penroz.engine import Gpt2Engine
            with torch.inference_mode(False):  # engine-owned tensors must outlive an inference_mode() caller
                self._fused = Gpt2Engine(self._plan, list(self.parameters()), self.optimizer)
        return self._fused

    def _maybe_accelerate(self):
        """F5b: serve /generate/ /output/ /evaluate/ from the GPU when there is one (the reference stays on CPU)."""
        if self._plan is None or os.environ.get("PENROZ_INFERENCE_DEVICE", "cuda") == "cpu":
            return
        if self._plan.wte.weight.device.type == "cpu" and torch.cuda.is_available():
            self.to("cuda")

    # ------------------------------------------------------------------ persistence (format == reference)
    @classmethod
    def get_model_path(cls, model_id):
        return os.path.join(MODELS_FOLDER, f"model_{model_id}.pth")

    def serialize(self):
        os.makedirs(MODELS_FOLDER, exist_ok=True)
        os.makedirs(os.path.join(self.SHM_PATH, MODELS_FOLDER), exist_ok=True)
        disk_path = self.get_model_path(self.model_id)
        shm_path = os.path.join(self.SHM_PATH, disk_path)
        payload = {
            "layers": self.mapper.layers,
            "state": self.state_dict(),
            "optim": self.mapper.optimizer,
            "optim_state": self.optimizer.state_dict(),
            "progress": self.progress,
            "average_cost": self.avg_cost,
            "average_cost_history": self.avg_cost_history,
            "stats": self.stats,
            "status": self.status,
        }
        chatty = ddp.master_proc()
        if chatty:
            log.info(f"Caching model to {shm_path}...")
        torch.save(payload, shm_path, pickle_protocol=5)
        with _RESIDENT_GUARD:
            entry = _RESIDENT.get(shm_path)
            if entry is not None:
                if entry[2] is self:        # the file now holds exactly this object's state: keep it resident
                    _RESIDENT[shm_path] = _file_stamp(shm_path) + (self,)
                else:
                    del _RESIDENT[shm_path]
        if chatty:
            log.info(f"Model cached successfully: {shm_path}")
            log.info(f"Offload flushing model cache {shm_path} to {disk_path}...")
        multiprocessing.Process(target=shutil.copyfile, args=(shm_path, disk_path)).start()

    @classmethod
    def deserialize(cls, model_id: str):
        disk_path = cls.get_model_path(model_id)
        shm_path = os.path.join(cls.SHM_PATH, disk_path)
        chatty = ddp.master_proc()
        try:
            if not os.path.exists(shm_path):
                if chatty:
                    log.info(f"Cache miss: copying from {disk_path}")
                    os.makedirs(os.path.join(cls.SHM_PATH, MODELS_FOLDER), exist_ok=True)
                    shutil.copyfile(disk_path, shm_path)
                if ddp.is_ddp() and dist.is_available() and dist.is_initialized():
                    dist.barrier()
            resident = _residency_on()
            if resident:
                stamp = _file_stamp(shm_path)
                with _RESIDENT_GUARD:
                    entry = _RESIDENT.get(shm_path)
                    if entry is not None and entry[:2] == stamp:
                        _RESIDENT.move_to_end(shm_path)
                        if chatty:
                            log.info(f"Model {model_id} is resident (unchanged {shm_path})")
                        return entry[2]
            if chatty:
                log.info(f"Retrieving model from {shm_path}...")
            data = torch.load(shm_path, weights_only=False)
        except FileNotFoundError as e:
            log.error(f"File not found error occurred: {str(e)}")
            raise KeyError(f"Model {model_id} not created yet.")
        model = cls(model_id, Mapper(data["layers"], data["optim"]))
        # a checkpoint saved in reduced precision (HF import keeps bf16) must be restored in that precision:
        # load_state_dict's copy_ would otherwise upcast into the fresh fp32 parameters
        stored = next((v.dtype for v in data["state"].values() if torch.is_tensor(v) and v.is_floating_point()), None)
        if stored is not None and stored != torch.float32:
            model.to(dtype=stored)
            if chatty:
                log.info(f"Restored model {model_id} dtype to {stored}")
        model.load_state_dict(data["state"])
        model.optimizer.load_state_dict(data["optim_state"])
        model.progress = data["progress"]
        model.avg_cost = data["average_cost"]
        model.avg_cost_history = data["average_cost_history"]
        model.stats = data["stats"]
        model.status = data["status"]
        if chatty:
            log.info(f"Model {model_id} is fully loaded")
        if resident and _file_stamp(shm_path) == stamp:       # not rewritten while it was being read
            model.__dict__["_resident_key"] = shm_path
            model.__dict__["_run_lock"] = threading.Lock()
            with _RESIDENT_GUARD:
                _RESIDENT[shm_path] = stamp + (model,)
                _RESIDENT.move_to_end(shm_path)
                keep = max(1, int(os.environ.get("PENROZ_MODEL_CACHE_SIZE", "2")))
                while len(_RESIDENT) > keep:
                    _RESIDENT.popitem(last=False)
        return model

    @classmethod
    def from_huggingface(cls, model_id: str, hf_repo_id: str, revision: Optional[str] = None,
                         device: str = "cpu") -> "NeuralNetworkModel":
        """Import a HuggingFace checkpoint (GPT-2 / Gemma families) into the internal format, keep it in bf16,
        serialise it and return it (reference: neural_net_model.py:176-237)."""
        log.info(f"Fetching HuggingFace config for {hf_repo_id} (revision={revision})")
        hf_config = AutoConfig.from_pretrained(hf_repo_id, revision=revision)
        log.info(f"Downloading HuggingFace model weights for {hf_repo_id}")
        hf_model = AutoModelForCausalLM.from_pretrained(hf_repo_id, revision=revision, dtype=torch.bfloat16,
                                                        low_cpu_mem_usage=True)
        hf_sd = hf_model.state_dict()
        del hf_model
        n_layer = Mapper.detect_hf_n_layer(hf_sd)
        if n_layer == 0:
            text_cfg = getattr(hf_config, "text_config", hf_config)
            n_layer = getattr(text_cfg, "n_layer", None) or getattr(text_cfg, "num_hidden_layers", None)
        log.info(f"Detected {n_layer} transformer layers from HuggingFace state dict")
        layer_specs = Mapper.from_hf_config(hf_config, n_layer_override=n_layer)
        mapper = Mapper(layer_specs, {"adamw": {"lr": 6e-4, "betas": [0.9, 0.95], "eps": 1e-8}})
        model = cls(model_id, mapper)
        model.to(dtype=torch.bfloat16)
        model.to(device)
        mapped = Mapper.map_hf_state_dict_to_custom(hf_sd, n_layer, hf_config)
        del hf_sd
        model.load_state_dict(mapped, strict=True)
        del mapped
        log.info(f"Loaded HuggingFace weights into model {model_id}")
        model.status = _now_status("Imported", f"Model imported from HuggingFace: {hf_repo_id}")
        model.serialize()
        return model

    @classmethod
    def delete(cls, model_id: str):
        disk_path = cls.get_model_path(model_id)
        with _RESIDENT_GUARD:
            _RESIDENT.pop(os.path.join(cls.SHM_PATH, disk_path), None)
        try:
            os.remove(os.path.join(cls.SHM_PATH, disk_path))
            if os.path.exists(disk_path):
                os.remove(disk_path)
        except FileNotFoundError as e:
            log.warning(f"Failed to delete: {str(e)}")

    # ------------------------------------------------------------------ forward
    def forward(self, input_tensor: Tensor, target: Tensor = None, skip_softmax=False) -> Tuple[list[Tensor], Tensor]:
        """Returns (output of every top-level layer, cost). cost is an empty tensor without a target; cross entropy
        on the logits when the model ends in a Softmax layer, else MSE (reference: neural_net_model.py:250-271)."""
        engine = self._engine()
        # A KV cache attached to the attention layers (the int8 TurboQuantKVCache, or a caller's own cache object) lives
        # in the layers' forward: only the module composition honours it. The fused decode step never comes through
        # here -- _generate_next_token drives the engine directly with its paged cache.
        cache_attached = any(layer._kv_cache is not None for layer in self._find_attention_layers())
        if (engine is not None and input_tensor.is_cuda and not input_tensor.is_floating_point() and not cache_attached
                and not self.__dict__.get("_torch_training")):
            return self._forward_fused(engine, input_tensor, target, skip_softmax)
        run = self.layers[:-1] if (skip_softmax and self._is_softmax_last) else self.layers
        outputs = []
        current = before_last = input_tensor
        for layer in run:
            before_last = current
            current = layer(current)
            outputs.append(current)
        if target is None:
            return outputs, torch.empty(0)
        if not self._is_softmax_last:
            return outputs, nn.functional.mse_loss(current, target)
        logits = current if skip_softmax else before_last
        if logits.ndim > 2 and target.ndim > 1:
            logits, target = logits.view(-1, logits.size(-1)), target.view(-1)
        return outputs, nn.functional.cross_entropy(logits, target)

    def _forward_fused(self, engine, input_tensor, target, skip_softmax):
        if target is not None and not (target.ndim == input_tensor.ndim):
            raise ValueError("fused forward expects a target of the same shape as the token input")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            if target is None or ddp.ddp_world_size() > 1:
                raise NotImplementedError(
                    "forward() under autograd on the fused CUDA path needs a target (the cost is what backward() "
                    "starts from) and a single process; train through train_model() for DDP, or call forward() "
                    "inside torch.no_grad()/inference_mode()")
            return self._forward_fused_autograd(engine, input_tensor, target, skip_softmax)
        logits, loss, acts = engine.forward_infer(input_tensor, target, pos_offset=self._plan.wpe.position_offset,
                                                   collect=True)
        if self._is_softmax_last and not skip_softmax:
            acts.append(self.layers[-1](logits))
        cost = loss if loss is not None else torch.empty(0)
        return acts, cost

    def _forward_fused_autograd(self, engine, input_tensor, target, skip_softmax):
        """`acts, cost = model(x, y); cost.backward(); optimizer.step()` on CUDA (the reference's own step,
        neural_net_model.py:629-653,677, written by hand by a caller): forward + hand-written backward run now,
        cost.backward() folds the pending gradients into .grad (engine.forward_with_pending_grads). Activations come
        back detached; under torch.autocast(bfloat16) the bf16 path runs, else the exact fp32 one."""
        use_amp = torch.is_autocast_enabled("cuda") if hasattr(torch, "is_autocast_enabled") else False
        loss, kept = engine.forward_with_pending_grads(input_tensor, target, use_amp)
        acts = [a for a, _ in _FusedStepper._stat_pairs(kept)]
        if self._is_softmax_last and not skip_softmax:
            acts.append(self.layers[-1](acts[-1]))
        anchor = next(p for p in self.parameters() if p.requires_grad)
        return acts, _PendingGradCost.apply(anchor, engine, loss)

    @torch.no_grad()
    def compute_output(self, input_data: list, target: list | int | None = None) -> Tuple[list, float]:
        """Activated output of the last layer and, with a target, the cost — no training (/output/)."""
        with _Exclusive(self):
            return self._compute_output(input_data, target)

    def _compute_output(self, input_data, target):
        if self.training or any(m.training for m in self.modules()):    # eval() walks ~170 modules: skip when already there
            self.eval()
        self.layers.training = False
        self._maybe_accelerate()
        first_param = next(self.parameters())
        device, model_dtype = first_param.device, first_param.dtype
        log.info(f"Computing model output using device {device}")
        x = torch.tensor(input_data, device=device)
        if x.is_floating_point():
            x = x.to(model_dtype)  # e.g. bf16 checkpoints
        y = torch.tensor(target, device=device) if target is not None else None
        activations, cost = self(x, y)
        if self._engine() is not None:
            self._engine().raise_on_index_errors()
        return activations[-1].tolist(), (cost.item() if cost.numel() > 0 else None)

    @torch.no_grad()
    def evaluate_model(self, dataset_id: str, target_dataset_id: str | None, shard: int,
                       epochs: int, batch_size: int, block_size: int, step_size: int) -> float:
        """Average cost over `epochs` loader windows (each re-used for every step), no training (/evaluate/)."""
        with _Exclusive(self):
            return self._evaluate_model(dataset_id, target_dataset_id, shard, epochs, batch_size, block_size, step_size)

    def _evaluate_model(self, dataset_id, target_dataset_id, shard, epochs, batch_size, block_size, step_size):
        if self.training or any(m.training for m in self.modules()):    # eval() walks ~170 modules: skip when already there
            self.eval()
        self.layers.training = False
        self._maybe_accelerate()
        window = batch_size * block_size
        world = ddp.ddp_world_size()
        num_steps = max(1, window // (step_size * block_size * world))
        first, stride = window * ddp.ddp_rank(), window * world
        if ddp.master_proc():
            log.info(f"Evaluation batch size: {batch_size}, buffer size: {window}, step size: {step_size}")
            log.info(f"Evaluation calc to be done in {num_steps} step(s) repeated {epochs} epoch(s)")
        loader = Loader(dataset_id, shard, first, window, stride)
        target_loader = Loader(target_dataset_id, shard, first, window, stride) if target_dataset_id is not None else None
        device = next(self.parameters()).device
        if ddp.master_proc():
            log.info(f"Evaluating model using device {device}")
        total = torch.tensor(0.0).to(device)
        for _ in range(epochs):
            if target_loader is None:
                xs, ys = loader.next_batch()
            else:
                xs = loader.next_batch(target_offset=0)
                ys = target_loader.next_batch(target_offset=0)
            for _ in range(num_steps):
                x = torch.tensor(xs, dtype=torch.long).view(batch_size, block_size).to(device)
                y = torch.tensor(ys, dtype=torch.long).view(batch_size, block_size).to(device)
                _, step_cost = self(x, y, skip_softmax=True)
                total += step_cost / (epochs * num_steps)
        if ddp.use_ddp(device.type):
            ddp.ddp_all_reduce(total)
        avg_cost = total.item()
        if ddp.master_proc():
            log.info(f"Model {self.model_id}: For {epochs} evaluation(s)  Avg Cost: {avg_cost:.4f}")
        return avg_cost

    # ------------------------------------------------------------------ generation
    def _find_attention_layers(self) -> list[CausalSelfAttention]:
        found = self.__dict__.get("_attention_layers")       # the module tree is fixed once the model is built
        if found is None:
            found = self.__dict__["_attention_layers"] = [m for m in self.modules() if isinstance(m, CausalSelfAttention)]
        return found

    def _find_position_embeddings(self) -> list[PositionEmbedding]:
        found = self.__dict__.get("_position_layers")
        if found is None:
            found = self.__dict__["_position_layers"] = [m for m in self.modules() if isinstance(m, PositionEmbedding)]
        return found

    def _attach_kv_cache(self) -> tuple[KVCache | None, list[PositionEmbedding]]:
        """Create one KV cache for all attention layers and attach it; returns (cache, position embeddings)."""
        attn = self._find_attention_layers()
        positions = self._find_position_embeddings()
        if not attn:
            return None, positions
        cache = create_kv_cache(len(attn))
        for i, layer in enumerate(attn):
            layer.set_kv_cache(cache, i)
        return cache, positions

    def _detach_kv_cache(self, pos_embeddings: list[PositionEmbedding] | None = None):
        for layer in self._find_attention_layers():
            layer.set_kv_cache(None, 0)
        for pe in (pos_embeddings or self._find_position_embeddings()):
            pe.position_offset = 0

    def _prepare_generation(self, input_context: list, max_new_tokens: int, temperature: float, top_k: int | None):
        """eval mode, context tensor on the parameter device, and the softmax used for sampling."""
        if self.training or any(m.training for m in self.modules()):    # eval() walks ~170 modules: skip when already there
            self.eval()
        self.layers.training = False
        self._maybe_accelerate()
        first_param = next(self.parameters())
        device = first_param.device
        context = torch.tensor(input_context, dtype=torch.long, device=device)
        top_k_msg = "" if top_k is None else f" top {top_k}"
        log.info(f"Generating at most {max_new_tokens}{top_k_msg} tokens with {temperature} temperature"
                 f" using device {device}")
        softmax_layer = self.layers[-1] if self._is_softmax_last else SoftmaxOnLast
        return context, softmax_layer

    @torch.inference_mode()
    def _generate_next_token(self, context: Tensor, block_size: int, temperature: float, top_k: int | None,
                             softmax_layer, kv_cache: KVCache | None = None,
                             pos_embeddings: list[PositionEmbedding] | None = None) -> Tensor:
        """One decode step: choose prefill / incremental / overflow re-prefill, run the model, pick the next token
        (reference: neural_net_model.py:360-406)."""
        positions = pos_embeddings or []
        cached = kv_cache.seq_len() if kv_cache is not None else 0
        offset = 0
        if cached > 0 and cached < block_size:
            model_input = context[:, -1:]          # incremental: only the newest token
            offset = cached
        else:
            if cached >= block_size:               # window full: drop the cache, re-prefill the cropped context
                kv_cache.clear()
            model_input = context[:, -block_size:]
        if cached > 0:
            for pe in positions:
                pe.position_offset = offset
        engine = self._engine()
        # the exact cache, or the int8 one (TURBO_QUANT_KV_CACHE=1): both have a paged store behind the fused step
        fused = (engine is not None and model_input.is_cuda and kv_cache is not None
                 and type(kv_cache) in (KVCache, TurboQuantKVCache))
        if fused:
            B = model_input.shape[0]
            pl = self._plan
            kv_cache.paged(B, pl.n_head, pl.head_dim, pl.wte.weight.dtype, model_input.device,
                           capacity_hint=min(block_size, pl.block_size))
            greedy_ids = None
            if model_input.shape[1] == 1:      # incremental step: one native call enqueues the whole token step
                logits, greedy_ids = engine.decode_step(model_input, offset, kv_cache)
            elif offset == 0 and os.environ.get("PENROZ_PREFILL", "native") == "native":
                logits, greedy_ids = engine.prefill(model_input, kv_cache)      # the whole prompt pass, one native call
            else:
                logits, _, _ = engine.forward_infer(model_input.contiguous(), None, pos_offset=offset, kv=kv_cache,
                                                    last_only=True)
        else:
            activations, _ = self(model_input, skip_softmax=True)
            logits = activations[-1]
        if temperature == 0.0:
            if fused:
                if greedy_ids is not None:
                    return greedy_ids
                from penroz import ops
                return ops.argmax_rows(logits.view(logits.shape[0], -1)).view(-1, 1)
            return logits[:, -1, :].argmax(dim=-1, keepdim=True)
        if top_k is not None:
            picked = logits.topk(top_k, dim=-1) if top_k < logits.size(-1) else logits.sort(dim=-1, descending=True)
            probs = softmax_layer.forward(picked.values / temperature)
            choice = self._draw(probs)
            return picked.indices[:, -1, :].gather(dim=1, index=choice)
        return self._draw(softmax_layer.forward(logits / temperature))

    @staticmethod
    def _draw(probs: Tensor) -> Tensor:
        """torch.multinomial on the CPU generator (the reference always decodes on the CPU, so a seeded request keeps
        its random stream), result back on the probabilities' device."""
        if probs.is_cuda:
            return torch.multinomial(probs.float().cpu(), num_samples=1).to(probs.device)
        return torch.multinomial(probs.float(), num_samples=1)

    def _token_stream(self, input_context, block_size, max_new_tokens, temperature, top_k, stop_token, sink,
                      emit=True):
        """Shared decode loop: yields each new token id (row 0) and leaves the full context in sink[0]. A model shared
        through the residency cache is held exclusively until the stream ends (generation attaches its KV cache to
        the layers)."""
        with _Exclusive(self):
            yield from self._token_stream_locked(input_context, block_size, max_new_tokens, temperature, top_k,
                                                 stop_token, sink, emit)

    def _token_stream_locked(self, input_context, block_size, max_new_tokens, temperature, top_k, stop_token, sink,
                             emit=True):
        context, softmax_layer = self._prepare_generation(input_context, max_new_tokens, temperature, top_k)
        cache, positions = self._attach_kv_cache()
        try:
            for _ in range(max_new_tokens):
                nxt = self._generate_next_token(context, block_size, temperature, top_k, softmax_layer, cache,
                                                positions)
                context = torch.cat((context, nxt), dim=1)
                sink[0] = context
                if not emit:
                    continue        # nobody looks at this token before the end: no device->host read, the host runs ahead
                token = nxt[0].item()
                yield token
                if stop_token is not None and token == stop_token:
                    break
        finally:
            sink[0] = context
            if cache is not None:
                cache.log_metrics()
            self._detach_kv_cache(positions)

    @torch.inference_mode()
    def generate_tokens(self, input_context: list, block_size: int, max_new_tokens: int,
                        temperature=1.0, top_k: int | None = None, stop_token: int | None = None) -> list:
        sink = [None]
        # greedy decoding without a stop token needs no token on the host until the end (sampling draws on the host)
        emit = not (temperature == 0.0 and stop_token is None)
        for _ in self._token_stream(input_context, block_size, max_new_tokens, temperature, top_k, stop_token, sink,
                                    emit=emit):
            pass
        if sink[0] is None:  # max_new_tokens == 0
            sink[0] = torch.tensor(input_context, dtype=torch.long)
        return sink[0][0].tolist()

    @torch.inference_mode()
    def generate_tokens_stream(self, input_context: list, block_size: int, max_new_tokens: int,
                               temperature=1.0, top_k: int | None = None, stop_token: int | None = None):
        """Yields token ids one at a time as they are produced."""
        log.info("Streaming token generation started")
        yield from self._token_stream(input_context, block_size, max_new_tokens, temperature, top_k, stop_token, [None])
        log.info("Streaming token generation completed")

    # ------------------------------------------------------------------ training
    @classmethod
    def train_model_on_device(cls, model_id: str, device: str, dataset_id: str, shard: int,
                              epochs: int, batch_size: int, block_size: int, step_size: int):
        """Worker entry point (one process per GPU under torch-elastic): join the process group, load the model,
        move it and its optimizer state to the device, train, leave the group."""
        if ddp.is_ddp():
            ddp.reconfig_logging()
            backend = 'nccl' if device == 'cuda' else 'gloo'
            log.info(f"DDP local rank {ddp.ddp_local_rank()} - training model {model_id} on device {device} "
                     f"with backend {backend}")
            dist.init_process_group(backend=backend)
        model = cls.deserialize(model_id)
        model.to(device)
        if ddp.master_proc():
            log.info(f"Moved model {model_id} to device {device}")
        where = next(model.parameters()).device
        for slot in model.optimizer.state.values():
            for key, val in slot.items():
                if isinstance(val, torch.Tensor):
                    slot[key] = val.to(where)
        model.train_model(dataset_id, shard, epochs, batch_size, block_size, step_size)
        if ddp.is_ddp():
            log.info(f"Process {ddp.ddp_local_rank()} - cleaning up")
            dist.destroy_process_group()

    def train_model(self, dataset_id: str, shard: int, epochs: int, batch_size: int, block_size: int, step_size: int):
        """Train on the dataset shards: one optimizer step per epoch, `num_steps` micro-batches of the full
        (batch_size, block_size) window each (reference: neural_net_model.py:552-722)."""
        device = next(self.parameters()).device
        master = ddp.master_proc()
        if master:
            log.info(f"Training model using device {device}")
        amp_ctx, amp_scaler, amp_dtype = nullcontext(), None, None
        if device.type == 'cuda':
            if torch.cuda.is_bf16_supported():
                amp_dtype = torch.bfloat16
            else:
                amp_dtype = torch.float16
                amp_scaler = torch.amp.GradScaler('cuda')
            amp_ctx = torch.amp.autocast('cuda', dtype=amp_dtype)
            if master:
                log.info(f"AMP enabled with dtype {amp_dtype}")

        world = ddp.ddp_world_size()
        window = batch_size * block_size
        num_steps = max(1, window // (step_size * block_size * world))
        first, stride = window * ddp.ddp_rank(), window * world
        log.info(f"Training starts from idx {first} of ds {dataset_id} at shard {shard} offset every {stride}")
        if master:
            log.info(f"Training batch size: {batch_size}, buffer size: {window}, step size: {step_size}")
            log.info(f"Training calc to be done in {num_steps} step(s) repeated {epochs} epoch(s)")
        loader = Loader(dataset_id, begin_shard=shard, begin_idx=first, buffer_size=window, idx_offset=stride)

        self.progress = []
        self.stats = None
        self.status = _now_status("Training", "Model is currently being trained.")
        last_saved = None
        if master:
            self.serialize()
            last_saved = time.time()

        engine = self._engine() if device.type == 'cuda' else None
        if engine is not None and (engine.plan.active_dropout(True) or engine.param_mode() != "f32"):
            # An HF-imported GPT-2 (bf16 parameters, dropout 0.1: neural_net_model.py:219-222, mappers.py:140-142) is
            # outside what the fused step implements (fp32 master weights, no dropout: torch's Philox stream would have
            # to be reproduced for parity). The reference trains such a model with plain torch; so does the drop-in --
            # /import/ then /train/ keeps working -- and it says that this run is not the fused path.
            if master:
                log.warning(f"Model {self.model_id}: parameters are {engine.param_mode()} / dropout is active -- training "
                            f"runs through torch autograd (the reference's own step), not the fused sm_100a step")
            engine = None
        if engine is not None:
            stepper = _FusedStepper(self, engine, device, use_amp=(os.environ.get("PENROZ_TRAIN_PRECISION", "amp") != "f32"))
        else:
            stepper = _TorchStepper(self, device, amp_ctx, amp_scaler)
        self.__dict__["_torch_training"] = isinstance(stepper, _TorchStepper)
        stepper.begin()
        if hasattr(stepper, "warn_if_ranks_will_drift"):
            stepper.warn_if_ranks_will_drift(num_steps)
        self.layers.training = True

        pairs = []
        for epoch in range(epochs):
            t_begin = time.time()
            slow = last_saved is not None and (t_begin - last_saved >= 10)
            want_stats = master and (epoch + 1 == epochs or slow)
            records = master and epoch % max(1, epochs // 100) == 0      # at most ~100 progress points
            stepper.measure_update = records
            stepper.more_epochs = epoch + 1 < epochs
            # the fused AdamW pass measures its own update; only the torch stepper needs the reference's weight clones
            before = ([None if w is None else w.clone().detach() for w in self._weights]
                      if master and not stepper.measures_update else [])
            try:
                cost, pairs = stepper.run_epoch(loader, num_steps, batch_size, block_size, want_stats)
            except Exception as exc:
                self.__dict__["_torch_training"] = False
                if master:
                    log.error(f"Model {self.model_id}: Training Epoch {epoch + 1} failed: {str(exc)}")
                    self.status = _now_status("Error", f"Training epoch {epoch + 1} failed: {str(exc)}")
                    self.serialize()
                raise exc
            if ddp.use_ddp(device.type):
                ddp.ddp_all_reduce(cost)
            stepper.apply_update(pairs)
            if device.type == 'cuda':  # the epoch time below must cover the device work
                torch.cuda.synchronize()
            elif device.type == 'mps':
                torch.mps.synchronize()
            stepper.after_sync()
            if master:
                secs = time.time() - t_begin
                speed = window / secs
                cost_value = cost.item()
                if records:
                    ratios = stepper.update_ratios() if stepper.measures_update else self._weight_update_ratios(before)
                    self.progress.append({"dt": dt.now().isoformat(), "epoch": epoch + 1, "durationInSecs": secs,
                                          "speedPerSec": speed, "cost": cost_value, "weight_upd_ratio": ratios})
                log.info(f"Model {self.model_id}: Training Epoch {epoch + 1}, Cost: {cost_value:.4f}, "
                         f"Duration: {secs:.2f} secs, Speed: {speed:.2f} tokens/sec")
            if master and slow:  # pragma: no cover
                self._record_training_overall_progress(pairs)
                self.serialize()
                last_saved = time.time()
        stepper.end()
        self.__dict__["_torch_training"] = False
        if master:
            self.status = _now_status("Trained", f"Model trained for {epochs} epochs.")
            log.info(f"Model {self.model_id}: Done training for {epochs} epochs.")
            self._record_training_overall_progress(pairs)
            self.serialize()

    # ------------------------------------------------------------------ diagnostics (rank 0)
    @torch.no_grad()
    def _weight_update_ratios(self, before: list) -> list:
        """std(w_new - w_old) / (std(w_new) + 1e-8) per 2-D weight, None elsewhere — the reference's arithmetic
        (neural_net_model.py:696-699) with ONE device->host read for the whole list instead of one sync per weight."""
        now = self._weights
        picked = [i for i, (pw, w) in enumerate(zip(before, now)) if w is not None and pw is not None]
        out = [None] * len(now)
        if picked:
            vals = torch.stack([(now[i] - before[i]).data.std() / (now[i].data.std() + 1e-8) for i in picked])
            for i, v in zip(picked, vals.tolist()):
                out[i] = v
        return out


    @staticmethod
    def _density_histogram(t: Tensor):
        """torch.histogram(density=True) as ([left bin edges], [densities]); bf16 is widened first."""
        h = torch.histogram(t.detach().float().cpu(), density=True)
        return h.bin_edges[:-1].tolist(), h.hist.tolist()

    @torch.no_grad()
    def _record_training_overall_progress(self, activations):
        """Fold this run's costs into the running average and rebuild `stats` (per-layer activation / gradient
        summaries, per-weight gradient summaries). `activations` is a list of tensors (their .grad is used) or of
        (activation, gradient) pairs."""
        costs = [entry["cost"] for entry in self.progress]
        mean_cost = sum(costs) / len(self.progress)
        self.avg_cost = ((self.avg_cost or mean_cost) + mean_cost) / 2.0
        self.avg_cost_history.append(self.avg_cost)
        if len(self.avg_cost_history) > 100:
            self.avg_cost_history.pop(random.randint(1, 98))

        pairs = [a if isinstance(a, tuple) else (a, a.grad) for a in activations]

        def saturation(kind: str, a: Tensor) -> float:
            if kind == "embedding":
                hit = torch.norm(a, dim=-1) > 5.0
            elif kind == "batchnorm1d":
                hit = a.abs() > 3.0
            elif kind in ("tanh", "sigmoid"):
                hit = a.abs() > 0.97
            elif kind == "relu":
                hit = a <= 0
            elif kind == "softmax":
                hit = a.max(dim=-1).values > 0.97
            else:
                hit = a.abs() > 5.0
            return hit.float().mean().item()

        # |a| > threshold is the saturation test of every kind except these three (their own reductions below)
        abs_threshold = {"batchnorm1d": 3.0, "tanh": 0.97, "sigmoid": 0.97}

        def summary(t: Tensor, kind: str | None = None):
            """(mean, std, saturated | None, histogram x, histogram y). CUDA tensors: one pass pair on the GPU and one
            small host read (penroz_tensor_stats); host tensors: the reference's torch calls."""
            if t.is_cuda and t.dtype in (torch.float32, torch.bfloat16) and t.numel() > 0:
                from penroz import ops
                own = kind is not None and kind not in ("embedding", "relu", "softmax")
                mean, std, _, _, sat, lo, hi, counts = ops.tensor_stats(t, 100, abs_threshold.get(kind, 5.0) if own else 5.0)
                edges = torch.linspace(lo, hi, 101, dtype=torch.float32)
                width = (edges[1:] - edges[:-1]).double()
                dens = (torch.tensor(counts, dtype=torch.float64) / (float(t.numel()) * width)).float()
                if kind is not None and not own:
                    sat = saturation(kind, t.float())
                return mean, std, sat, edges[:-1].tolist(), dens.tolist()
            x, y = self._density_histogram(t)
            tf = t.float()
            return tf.mean().item(), tf.std().item(), (saturation(kind, tf) if kind is not None else None), x, y

        layer_stats = []
        for layer, (a, g) in zip(self.layers, pairs):
            kind = layer.__class__.__name__.lower()
            mean, std, sat, ax, ay = summary(a, kind)
            entry = {"algo": kind,
                     "activation": {"mean": mean, "std": std, "saturated": sat, "histogram": {"x": ax, "y": ay}},
                     "gradient": None}
            if g is not None:
                mean, std, _, gx, gy = summary(g)
                entry["gradient"] = {"mean": mean, "std": std, "histogram": {"x": gx, "y": gy}}
            layer_stats.append(entry)
        weight_stats = []
        for w in self._weights:
            if w is None:
                weight_stats.append(None)
                continue
            gmean, gstd, _, gx, gy = summary(w.grad)
            if w.is_cuda and w.dtype in (torch.float32, torch.bfloat16):
                from penroz import ops
                wmean, wstd = ops.tensor_stats(w.data, 1, 5.0)[:2]
            else:
                wmean, wstd = w.mean().item(), w.std().item()
            weight_stats.append({"shape": str(tuple(w.shape)),
                                 "data": {"mean": wmean, "std": wstd},
                                 "gradient": {"mean": gmean, "std": gstd, "histogram": {"x": gx, "y": gy}}})
        self.stats = {"layers": layer_stats, "weights": weight_stats}
        log.info(f"Model {self.model_id} - Cost: {mean_cost:.4f} Overall Cost: {self.avg_cost:.4f}")


class _PendingGradCost(torch.autograd.Function):
    """The cost of a fused forward as an autograd node: its backward adds grad_output x (the gradients the engine
    already computed) to the parameters' .grad buffer."""

    @staticmethod
    def forward(ctx, anchor, engine, loss):
        ctx.engine = engine
        return loss.detach().clone()

    @staticmethod
    def backward(ctx, grad_out):
        ctx.engine.backward_pending(grad_out)
        return None, None, None


class _TorchStepper:
    """The reference's own training step (torch autograd, optional AMP/GradScaler, stock DDP) for trees and devices
    the fused engine does not cover."""
    measures_update = False
    measure_update = False

    def __init__(self, model: NeuralNetworkModel, device, amp_ctx, amp_scaler):
        self.model, self.device, self.amp_ctx, self.scaler = model, device, amp_ctx, amp_scaler
        self.wrapped = model

    def begin(self):
        if ddp.use_ddp(self.device.type):
            self.wrapped = nn.parallel.DistributedDataParallel(self.model)
        self.wrapped.train()

    def end(self):
        pass

    def after_sync(self):
        pass

    def run_epoch(self, loader, num_steps, batch_size, block_size, want_stats):
        model = self.model
        model.optimizer.zero_grad()
        cost = torch.tensor(0.0).to(self.device)
        kept = []
        for step in range(num_steps):
            x, y = (torch.tensor(arr, dtype=torch.long).view(batch_size, block_size).to(self.device)
                    for arr in loader.next_batch())
            with self.amp_ctx:
                outs, step_cost = self.wrapped(x, y, skip_softmax=True)
                share = step_cost / num_steps
            cost += share.detach()
            if ddp.master_proc():
                kept.extend(outs)
                if want_stats:
                    for a in outs:
                        a.retain_grad()
            if ddp.use_ddp(self.device.type):
                # the reference sets this after forward (so it takes effect one micro-step late)
                self.wrapped.require_backward_grad_sync = (step == num_steps - 1)
            if self.scaler is not None:
                self.scaler.scale(share).backward()
            else:
                share.backward()
        return cost, kept

    def apply_update(self, kept):
        if self.scaler is None:
            self.model.optimizer.step()
            return
        self.scaler.step(self.model.optimizer)
        if ddp.master_proc():  # un-scale retained activation gradients before the scale changes
            inv = 1.0 / self.scaler.get_scale()
            for a in kept:
                if a.grad is not None:
                    a.grad.mul_(inv)
        self.scaler.update()


class _FusedStepper:
    """Training step on the penroz engine: hand-written forward/backward kernels, bucketed NCCL gradient exchange
    overlapped with backward, one fused AdamW pass."""

    measures_update = True       # weight_upd_ratio comes out of the AdamW pass (penroz_adamw_flat_stats): no weight clones

    def __init__(self, model: NeuralNetworkModel, engine, device, use_amp: bool):
        self.model, self.engine, self.device, self.use_amp = model, engine, device, use_amp
        self.measure_update = False
        self._keep_grads = True
        self.sync_mode = os.environ.get("PENROZ_ACCUM_SYNC", "reference")
        self._pinned = None      # two pinned int32 [2,B,T] staging buffers, reused round-robin
        self._copied = None      # CUDA events: the H2D copy out of each buffer has finished
        self._slot = 0
        self._ahead = None       # a window staged ahead of time by prefetch()
        self.more_epochs = False

    def _stage(self, arrays, batch_size, block_size):
        """Loader window (two int32 numpy arrays) -> one pinned int32 [2,B,T] staging buffer; returns its slot. Host
        work only: it can run while the GPU is still busy with the previous step."""
        shape = (2, batch_size, block_size)
        if self._pinned is None or self._pinned[0].shape != shape:
            self._pinned = [torch.empty(shape, dtype=torch.int32).pin_memory() for _ in range(2)]
            self._copied = [None, None]
            self._dev32 = torch.empty(shape, dtype=torch.int32, device=self.device)
            self._dev64 = torch.empty(shape, dtype=torch.long, device=self.device)
        slot = self._slot
        self._slot ^= 1
        if self._copied[slot] is not None:
            self._copied[slot].synchronize()      # the previous copy out of this buffer is done: safe to overwrite
        host = self._pinned[slot]
        for i, arr in enumerate(arrays):
            host[i].view(-1).copy_(torch.from_numpy(arr))          # int32 -> int32 (no widening on the host)
        return slot

    def _upload(self, arrays, batch_size, block_size, slot=None):
        """int32 window through pinned memory, asynchronous H2D on the compute stream, widened to the int64 ids the
        kernels index with ON the device (penroz_widen_tokens). The reference widens on the host and does a
        synchronous pageable copy per tensor (neural_net_model.py:630-631)."""
        from penroz import ops
        if slot is None:
            slot = self._stage(arrays, batch_size, block_size)
        self._dev32.copy_(self._pinned[slot], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._copied[slot] = ev
        ops.widen_tokens(self._dev32, self._dev64)
        return self._dev64[0], self._dev64[1]

    def prefetch(self, loader, batch_size, block_size):
        """Read and stage the NEXT epoch's first window now -- after this epoch's kernels are enqueued, before the host
        blocks on them -- so the GPU does not wait for the loader and the pinned copy after the epoch's synchronize."""
        try:
            self._ahead = ("ok", self._stage(loader.next_batch(), batch_size, block_size))
        except Exception as exc:      # surfaces where the window would have been read
            self._ahead = ("error", exc)

    def begin(self):
        from penroz._lib import PenrozError
        opt = self.model.optimizer
        if type(opt) is not torch.optim.AdamW:
            raise PenrozError(f"the fused training step implements AdamW (the reference's GPT-2 optimizer), "
                              f"found {type(opt).__name__}")
        eng = self.engine
        eng.optimizer = opt
        eng.setup_training(self.use_amp)
        self.model.train()
        if ddp.use_ddp(self.device.type) and dist.is_available() and dist.is_initialized():
            from penroz.grad_sync import GradSync, gpt2_segments
            eng.grad_sync = GradSync(eng.flat.g, gpt2_segments(eng.flat, eng.plan))
            eng.grad_sync.broadcast_(eng.flat.p)      # DDP's constructor broadcast from rank 0
            if eng.flat.shadow is not None:
                eng.flat.refresh_shadow()
        else:
            eng.grad_sync = None

    def warn_if_ranks_will_drift(self, num_steps: int):
        """With accumulation under DDP the reference exchanges gradients in micro-step 0 only (it arms DDP's flag after
        the forward that reads it): every rank then adds its own later micro-steps and the replicas' weights
        drift apart, silently. The default reproduces that for parity; say so once, and name the switch."""
        if (self.engine.grad_sync is not None and num_steps > 1 and self.sync_mode == "reference"
                and ddp.master_proc()):
            log.warning(f"DDP with {num_steps} accumulation micro-steps: like the reference, only micro-step 0's gradients "
                        f"are all-reduced and the ranks' weights will diverge; set PENROZ_ACCUM_SYNC=correct to "
                        f"exchange the accumulated gradient after the last micro-step instead")

    def end(self):
        self.engine.release_workspace()

    def after_sync(self):
        # out-of-range token / target ids are flagged on the device; look once the epoch has synchronised anyway
        self.engine.raise_on_index_errors()

    def run_epoch(self, loader, num_steps, batch_size, block_size, want_stats):
        eng = self.engine
        eng.zero_grad()
        self._keep_grads = bool(want_stats)
        cost = torch.empty((), device=self.device, dtype=torch.float32)   # written by the first micro-step's loss reduction
        pairs = []
        for step in range(num_steps):
            ahead, self._ahead = self._ahead, None
            if ahead is not None and ahead[0] == "error":
                raise ahead[1]
            if ahead is not None:
                x, y = self._upload(None, batch_size, block_size, slot=ahead[1])
            else:
                x, y = self._upload(loader.next_batch(), batch_size, block_size)
            keep = want_stats and step == 0
            eng.keep_for_stats = {} if keep else None
            # which micro-step carries the gradient exchange: the reference arms DDP one step late, so with
            # accumulation it syncs in micro-step 0 only; "correct" syncs after the last one
            sync_here = eng.grad_sync is not None and (
                (step == 0) if (self.sync_mode == "reference") else (step == num_steps - 1))
            if sync_here and self.sync_mode == "reference" and num_steps > 1:
                pass  # later micro-steps add un-synced local gradients on top, exactly as the reference does
            eng.train_micro_step(x, y, 1.0 / num_steps, sync_grads=sync_here, keep_stats=keep, cost_out=cost,
                                 cost_first=(step == 0))
            if sync_here and step != num_steps - 1:
                eng.grad_sync.wait_all()   # later micro-steps accumulate into the reduced buffer
            if keep:
                pairs = self._stat_pairs(eng.keep_for_stats)
                eng.keep_for_stats = None
        self._next = (loader, batch_size, block_size)
        return cost, pairs

    @staticmethod
    def _stat_pairs(kept: dict):
        xs, gx = kept["x"], kept["x_grad"]
        pairs = [(xs[0], gx[0]), (xs[0], gx[0])]          # summation output and the (p = 0) dropout's output
        pairs += [(xs[l], gx[l]) for l in range(1, len(xs))]
        pairs.append((kept["lnf"], kept["lnf_grad"]))
        pairs.append((kept["logits"], kept["logits_grad"]))
        return pairs

    def apply_update(self, pairs):
        # on a statistics epoch the per-weight gradient summaries are read after the step: keep the gradients then
        self.engine.optimizer_step(track_update=self.measure_update, clear_grads=not self._keep_grads)
        if self.more_epochs:
            self.prefetch(*self._next)

    def update_ratios(self):
        return self.engine.update_ratios()
