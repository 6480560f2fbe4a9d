"""Stage the UNMODIFIED reference next to the oracle so it can run where /root/reference does not exist.

TEST INFRASTRUCTURE, not product code: nothing under penr-oz-neural-network-v3-torch-ddp_b200/ imports this.

    python oracle/stage_ref.py [--src /root/reference] [--dst oracle/_ref]

The reference (derinworks/penr-oz-neural-network-v3-torch-ddp) is pure Python: "building" it is a byte-for-byte copy of
its source files. The copy lands in oracle/_ref/, which is git-ignored (reference sources never enter this repository's
history) but is NOT gpurun-ignored, so it travels to the GPU box with the snapshot like a built .so does. It is what
  * oracle/ref_cuda.py runs on the B200 box (the reference's own torch-CUDA training step and decode: the bf16 oracle and
    the performance bar, SURVEY 8c tier 2 / BASELINE.md section 3 "B-GPU"), and
  * bench.py --impl reference times on the host cores (the reference's own CPU train_model).
A manifest with the sha256 of every staged file is written beside it so a run can state exactly which reference it used.
"""
import argparse
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_SRC = "/root/reference"
DEFAULT_DST = os.path.join(HERE, "_ref")
# the service modules the hot path lives in plus the modules they import at module scope; tests, docs, the dashboard
# and the launch scripts are not needed to run the path
WANTED = ("neural_net_model.py", "neural_net_layers.py", "ddp.py", "kv_cache.py", "mappers.py", "loaders.py",
          "gpt_tokenizers.py", "main.py", "log_config.json", "requirements.txt", "LICENSE")


def _sha(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for block in iter(lambda: f.read(1 << 20), b""):
            h.update(block)
    return h.hexdigest()


def stage(src: str = DEFAULT_SRC, dst: str = DEFAULT_DST, verbose: bool = True) -> dict:
    """Copy the reference's files unmodified; returns the manifest {file: sha256}. Raises if `src` is absent."""
    if not os.path.isdir(src):
        raise FileNotFoundError(f"reference checkout not found at {src}")
    os.makedirs(dst, exist_ok=True)
    manifest = {}
    for name in WANTED:
        s = os.path.join(src, name)
        if not os.path.exists(s):
            continue
        d = os.path.join(dst, name)
        shutil.copyfile(s, d)
        manifest[name] = _sha(d)
        if manifest[name] != _sha(s):
            raise RuntimeError(f"staged copy of {name} differs from the source")
    with open(os.path.join(dst, "MANIFEST.json"), "w") as f:
        json.dump({"source": src, "files": manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print(f"[oracle] staged {len(manifest)} reference files into {dst}")
    return manifest


def staged_path(dst: str = DEFAULT_DST):
    """The staged reference's directory if it is present and complete, else None."""
    need = ("neural_net_model.py", "neural_net_layers.py", "ddp.py", "kv_cache.py", "mappers.py", "loaders.py")
    return dst if all(os.path.exists(os.path.join(dst, n)) for n in need) else None


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default=DEFAULT_SRC)
    ap.add_argument("--dst", default=DEFAULT_DST)
    a = ap.parse_args()
    stage(a.src, a.dst)
    sys.exit(0)
