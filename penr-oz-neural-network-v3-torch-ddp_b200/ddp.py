"""Drop-in for the reference's ddp.py: rank helpers read from the environment at call time, the single-node elastic
launcher, and the scalar all-reduce (reference: ddp.py:13-114). Data parallelism — one process per GPU, NCCL over
NVLink — is the only parallelism of this path; the bucketed gradient exchange itself lives in penroz/grad_sync.py.

Module-scope names (cuda, mps, cpu_count, elastic_launch, LaunchConfig, all_reduce, get_backend, ReduceOp, Path)
are part of the surface: callers and the reference's tests patch them here.
"""
import logging
import os
import sys
from multiprocessing import cpu_count
from pathlib import Path
from typing import Callable

from torch import Tensor, cuda, mps
from torch.distributed import ReduceOp, all_reduce, get_backend
from torch.distributed.launcher.api import LaunchConfig, elastic_launch

log = logging.getLogger(__name__)


def _env_int(name: str, default: int) -> int:
    return int(os.environ.get(name, default))


def is_ddp() -> bool:
    return _env_int("RANK", -1) != -1


def ddp_rank() -> int:
    return _env_int("RANK", 0)


def ddp_local_rank() -> int:
    return _env_int("LOCAL_RANK", 0)


def ddp_world_size() -> int:
    return _env_int("WORLD_SIZE", 1)


def master_proc() -> bool:
    return ddp_rank() == 0


def running_on_linux() -> bool:
    return sys.platform.startswith("linux")


def detect_active_ip_family() -> str:
    """'ipv6' when the loopback answers on ::1, else 'ipv4'."""
    import socket
    if not socket.has_ipv6:
        return "ipv4"
    try:
        with socket.socket(socket.AF_INET6, socket.SOCK_DGRAM) as probe:
            probe.settimeout(0.5)
            probe.connect(("::1", 1))
        return "ipv6"
    except Exception:
        return "ipv4"


def _worker_count(device: str) -> int:
    if device == "cuda":
        return cuda.device_count()
    if device == "mps":
        return mps.device_count()
    return max(1, cpu_count() // 2)


def launch_single_node_ddp(run_id: str, device: str, worker_op: Callable[..., None], *args):
    """Spawn one worker per device on this node through torch-elastic (c10d rendezvous, no restarts)."""
    if device == "mps":
        os.environ["PYTORCH_ENABLE_MPS_FALLBACK"] = "1"
        log.warning("MPS device detected: enabled PYTORCH_ENABLE_MPS_FALLBACK=1 "
                    "so unsupported ops (e.g. c10d::allgather_) fall back to CPU.")
    nproc = _worker_count(device)
    log.info(f"Launching single node DDP run {run_id} with {nproc} processes on device {device}")
    settings = {"min_nodes": 1, "max_nodes": 1, "nproc_per_node": nproc, "rdzv_backend": "c10d",
                "max_restarts": 0, "monitor_interval": 5, "run_id": run_id}
    if not running_on_linux():
        # loopback rendezvous: avoids hostname / IPv6 lookups outside Linux
        v6 = detect_active_ip_family() == "ipv6"
        loopback = "::1" if v6 else "127.0.0.1"
        os.environ["MASTER_ADDR"] = loopback
        os.environ["GLOO_USE_IPV6"] = "1" if v6 else "0"
        settings["local_addr"] = loopback
        settings["rdzv_endpoint"] = f"[{loopback}]:0" if v6 else f"{loopback}:0"
        if v6 and sys.platform == "darwin":
            os.environ["GLOO_SOCKET_IFNAME"] = "lo0"
    elastic_launch(LaunchConfig(**settings), entrypoint=worker_op)(*args)


def use_ddp(device: str) -> bool:
    if device == "mps" and is_ddp() and ddp_world_size() == 1:
        return False
    return is_ddp()


def ddp_all_reduce(tensor: Tensor):
    """In-place mean over ranks (NCCL has AVG; other backends sum then divide)."""
    if get_backend() == "nccl":
        all_reduce(tensor, op=ReduceOp.AVG)
        return
    all_reduce(tensor, op=ReduceOp.SUM)
    tensor.div_(ddp_world_size())


def reconfig_logging():
    """(Re)apply log_config.json in a worker; outside Linux also tee each rank to logs/ddp_rankNN.log."""
    import json
    import logging.config
    with open("log_config.json", "r") as f:
        cfg = json.load(f)
    if is_ddp() and not running_on_linux():
        logging.getLogger("torch.distributed.elastic.multiprocessing.redirects").setLevel(logging.ERROR)
        folder = Path("logs")
        folder.mkdir(parents=True, exist_ok=True)
        cfg["handlers"]["ddp_file"] = {
            "level": "INFO", "class": "logging.handlers.RotatingFileHandler", "formatter": "default",
            "filename": str(folder / f"ddp_rank{ddp_rank():02d}.log"), "maxBytes": 10_485_760, "backupCount": 3}
        if "ddp_file" not in cfg["root"]["handlers"]:
            cfg["root"]["handlers"].append("ddp_file")
    logging.config.dictConfig(cfg)
