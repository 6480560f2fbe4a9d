"""Golden vectors for the data-parallel step: the UNMODIFIED reference under 2-rank gloo DDP on CPU (fp32).

    python tests/golden/make_golden_ddp.py          # build container only; writes tests/golden/golden_ddp.npz

Each rank builds the tiny GPT from the same seed, joins a gloo process group and calls the reference's
NeuralNetworkModel.train_model (neural_net_model.py:552-722), which wraps itself in DistributedDataParallel (:609),
reads its own loader windows (:581-594) and all-reduces the epoch cost (ddp.py:80-85).
Two cases: num_steps = 1 (plain DDP) and num_steps = 2, where the reference assigns require_backward_grad_sync after
forward (:647-648) so gradients are exchanged in micro-step 0 only and the ranks' weights drift apart.
Recorded: rank-0 per-epoch cost, and every rank's final embeddings, first attention block, last MLP projection,
final LayerNorm and head (KEEP below).
"""
import os
import sys
import tempfile

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


KEEP = ("final/layers.0.", "final/layers.2.0.", "final/layers.3.1.3.", "final/layers.4.", "final/layers.5.")


def worker(rank, world, port, work, out_dir, step_size, epochs, tag):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    os.chdir(work)
    sys.path.insert(0, REF)
    import copy
    import torch
    import torch.distributed as dist
    torch.set_num_threads(2)
    from make_golden import OPTIM, gpt_dsl
    import neural_net_model as ref_nnm
    from mappers import Mapper
    ref_nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, f"shm{rank}")
    os.makedirs(ref_nnm.NeuralNetworkModel.SHM_PATH, exist_ok=True)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = ref_nnm.NeuralNetworkModel(f"ddp_{tag}_{rank}", Mapper(copy.deepcopy(gpt_dsl(2, 4, 128, 64, 65)), copy.deepcopy(OPTIM)))
    model.train_model("synth65", 0, epochs, 16, 64, step_size)
    out = {f"final/{k}": v.detach().numpy().copy() for k, v in model.state_dict().items()}
    if rank == 0:
        out["cost"] = np.array([p["cost"] for p in model.progress], dtype=np.float64)
    np.savez(os.path.join(out_dir, f"{tag}_r{rank}.npz"), **out)
    dist.destroy_process_group()


def main():
    import socket
    import torch.multiprocessing as mp
    from make_golden import synth_shard
    work = tempfile.mkdtemp(prefix="penroz_golden_ddp_")
    os.makedirs(os.path.join(work, "data"))
    np.save(os.path.join(work, "data", "synth65_000000"), synth_shard(65, 0, 2 ** 17))
    merged = {}
    for tag, step_size, epochs in (("ns1", 8, 10), ("ns2", 4, 6)):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        mp.spawn(worker, args=(2, port, work, work, step_size, epochs, tag), nprocs=2, join=True)
        for r in range(2):
            d = np.load(os.path.join(work, f"{tag}_r{r}.npz"))
            for k in d.files:
                if k == "cost" or k.startswith(KEEP):       # a subset keeps the fixture small
                    merged[f"{tag}/r{r}/{k}"] = d[k]
    np.savez_compressed(os.path.join(HERE, "golden_ddp.npz"), **merged)
    print("wrote golden_ddp.npz", len(merged), "arrays;",
          "ns1 cost", merged["ns1/r0/cost"][:3], "ns2 cost", merged["ns2/r0/cost"][:3])
    a, b = merged["ns1/r0/final/layers.5.weight"], merged["ns1/r1/final/layers.5.weight"]
    c, d = merged["ns2/r0/final/layers.5.weight"], merged["ns2/r1/final/layers.5.weight"]
    print("ranks identical after training: num_steps=1", bool(np.array_equal(a, b)), " num_steps=2", bool(np.array_equal(c, d)))


if __name__ == "__main__":
    main()
