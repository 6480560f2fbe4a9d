"""Rank body for tests/test_ddp_gpu.py, launched by torch.distributed.run with one process per GPU:
    ddp_worker.py <work_dir> <model_id> <epochs> <step_size> <out_dir>
Joins the NCCL group, loads the serialised model, trains through the drop-in's train_model (the fused fp32 path) and
writes this rank's epoch costs and final parameters."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "penr-oz-neural-network-v3-torch-ddp_b200"))


def main():
    work, model_id, epochs, step_size, out_dir = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    os.environ["PENROZ_TRAIN_PRECISION"] = "f32"
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    os.chdir(work)
    import torch
    import torch.distributed as dist
    import ddp
    import neural_net_model as nnm
    from penroz import _lib
    nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, "shm")
    _lib.require_device()
    torch.cuda.set_device(ddp.ddp_local_rank())
    dist.init_process_group("nccl", device_id=torch.device("cuda", ddp.ddp_local_rank()))
    model = nnm.NeuralNetworkModel.deserialize(model_id)
    model.to("cuda")
    n0 = _lib.lib().launch_count()
    model.train_model("synth65", 0, epochs, 16, 64, step_size)
    assert _lib.lib().launch_count() - n0 > epochs * 40, "the CUDA kernels did not run"
    out = {"final/" + k: v.detach().float().cpu().numpy() for k, v in model.state_dict().items()}
    out["cost"] = np.array([p["cost"] for p in model.progress], dtype=np.float64)
    np.savez(os.path.join(out_dir, f"r{ddp.ddp_rank()}.npz"), **out)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
