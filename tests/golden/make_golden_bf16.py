"""Fold the outputs of oracle/ref_cuda.py (the UNMODIFIED reference run on a B200 box through its own torch-CUDA path)
into the small committed fixture tests/golden/golden_bf16_124m.npz.

    # on the GPU box (gpurun), with oracle/_ref staged by oracle/stage_ref.py:
    python oracle/ref_cuda.py train  --config 124m --epochs 60 --out gpurun_out/ref_train_n1.json
    python oracle/ref_cuda.py decode --out gpurun_out/ref_decode.json
    # back in the build container:
    python tests/golden/make_golden_bf16.py gpurun_out

What is pinned: the 60 per-epoch losses of train_model_on_device('cuda') on GPT-2 124M (seed-0 weights, synthetic shard,
32 x 1024 tokens per step, bf16 autocast); optionally the 8-rank DDP losses; greedy ids, top-2 logit gaps and seeded
top-k ids of generate_tokens for the bf16 and the fp32 124M model.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main(src):
    out = {}
    meta = {}
    for tag in ("n1", "n8"):
        path = os.path.join(src, f"ref_train_{tag}.json")
        if not os.path.exists(path):
            continue
        d = json.load(open(path))
        out[f"train_{tag}/losses"] = np.array(d["losses"], dtype=np.float64)
        for k in ("batch_size", "block_size", "num_params", "world", "num_steps"):
            out[f"train_{tag}/{k}"] = np.array(d[k])
        meta[f"train_{tag}"] = {k: d[k] for k in ("what", "config", "torch", "gpu", "tokens_per_sec", "guards")}
    dec = os.path.join(src, "ref_decode.npz")
    if os.path.exists(dec):
        z = np.load(dec)
        for k in z.files:
            out["decode/" + k] = z[k]
        meta["decode"] = json.load(open(os.path.join(src, "ref_decode.json")))
    np.savez_compressed(os.path.join(HERE, "golden_bf16_124m.npz"), **out)
    with open(os.path.join(HERE, "golden_bf16_124m_meta.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("wrote golden_bf16_124m.npz with", sorted(out))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out")
