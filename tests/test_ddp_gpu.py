"""Two-GPU parity of the data-parallel step: the drop-in under torch.distributed.run (one process per GPU, NCCL)
against the reference's own 2-rank DDP run (tests/golden/golden_ddp.npz, made by make_golden_ddp.py) -- plain DDP
and the accumulation case where only micro-step 0 exchanges gradients. Needs two GPUs; skipped otherwise
(`gpurun --gpus 2 -- python -m pytest tests/test_ddp_gpu.py -m gpu`)."""
import os
import socket
import subprocess
import sys
import tempfile

import numpy as np
import pytest
import torch

from helpers import build_gpt, scratch_cwd

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("tag,step_size,epochs", [("ns1", 8, 10), ("ns2", 4, 6)])
def test_two_gpu_training_matches_reference_ddp(tag, step_size, epochs):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from penroz import shards
    work = scratch_cwd()
    shards.write_synthetic_shards("synth65", 65, 1, 2 ** 17)
    m = build_gpt(f"ddp_{tag}", 2, 4, 128, 64, 65)
    m.serialize()
    out_dir = tempfile.mkdtemp(prefix="penroz_ddp_out_")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(HERE, "ddp_worker.py"), work, f"ddp_{tag}", str(epochs), str(step_size), out_dir]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    d = np.load(os.path.join(HERE, "golden", "golden_ddp.npz"))
    got = [np.load(os.path.join(out_dir, f"r{k}.npz")) for k in range(2)]
    np.testing.assert_allclose(got[0]["cost"], d[f"{tag}/r0/cost"], rtol=1e-4)
    for k in range(2):
        pre = f"{tag}/r{k}/"
        for name in [n for n in d.files if n.startswith(pre + "final/")]:
            ref = d[name]
            assert np.abs(got[k][name[len(pre):]] - ref).max() < 1e-4 * max(1.0, np.abs(ref).max()), name
    same = np.array_equal(got[0]["final/layers.5.weight"], got[1]["final/layers.5.weight"])
    assert same == (tag == "ns1")       # with accumulation the reference's ranks drift apart; so must ours
