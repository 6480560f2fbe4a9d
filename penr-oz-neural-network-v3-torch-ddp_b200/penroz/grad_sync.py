"""Bucketed gradient all-reduce overlapped with the hand-written backward.

Replaces what nn.parallel.DistributedDataParallel(self) does in the reference (neural_net_model.py:609, all
defaults): parameters are broadcast from rank 0 once, and during backward the fp32 gradients are averaged across
ranks bucket by bucket. Here gradients already live in ONE flat buffer (engine.FlatState), so a bucket is a
contiguous slice: no copy-in/copy-out. Buckets follow the order in which our backward finishes them — LM head +
final LayerNorm first, then each transformer block from last to first, embeddings last — and each bucket's
all-reduce is issued on a side stream the moment the engine calls ready(name), so NCCL (NVLink 5 / NVSwitch) runs
under the remaining backward kernels. Small adjacent buckets are merged up to `bucket_bytes`.

Backend: NCCL for CUDA tensors (ReduceOp.AVG), Gloo for the CPU test tier (SUM then divide), through
torch.distributed — the same c10d process group the reference initialises (neural_net_model.py:532-535).
"""
import torch
import torch.distributed as dist


class GradSync:
    def __init__(self, flat_grad: torch.Tensor, segments: list, bucket_bytes: int = 64 << 20, group=None):
        """segments: [(name, start, end)] element ranges of flat_grad in the order backward completes them
        (names are what the engine passes to ready()). Adjacent segments are merged into buckets of at most
        bucket_bytes (a segment larger than that is its own bucket)."""
        self.g = flat_grad
        self.group = group
        self.world = dist.get_world_size(group)
        self.backend = dist.get_backend(group)
        self.buckets = []          # [(lo, hi, last_segment_name)]
        self._trigger = {}         # segment name -> bucket index launched when that segment is ready
        cur_lo = cur_hi = None
        cur_names = []
        limit = max(1, bucket_bytes // flat_grad.element_size())

        def flush():
            nonlocal cur_lo, cur_hi, cur_names
            if cur_names:
                self._trigger[cur_names[-1]] = len(self.buckets)
                self.buckets.append((cur_lo, cur_hi, cur_names[-1]))
            cur_lo = cur_hi = None
            cur_names = []

        for name, lo, hi in segments:
            contiguous = cur_names and (hi == cur_lo or lo == cur_hi)
            if cur_names and (not contiguous or (max(hi, cur_hi) - min(lo, cur_lo)) > limit):
                flush()
            if not cur_names:
                cur_lo, cur_hi = lo, hi
            else:
                cur_lo, cur_hi = min(lo, cur_lo), max(hi, cur_hi)
            cur_names.append(name)
        flush()
        self._pending = []
        spans = sorted((lo, hi) for lo, hi, _ in self.buckets)
        # True when the buckets tile the whole flat buffer, so a per-bucket optimizer pass touches every element once
        self.covers_all = bool(spans) and spans[0][0] == 0 and spans[-1][1] == flat_grad.numel() and all(
            a[1] == b[0] for a, b in zip(spans, spans[1:]))
        self._cuda = flat_grad.is_cuda
        self._stream = torch.cuda.Stream(device=flat_grad.device) if self._cuda else None
        self.launched = []         # bucket indices in launch order (inspected by tests)

    def broadcast_(self, tensor: torch.Tensor, src: int = 0):
        """DDP's constructor-time parameter broadcast from rank 0."""
        dist.broadcast(tensor, src=src, group=self.group)

    def ready(self, segment_name: str):
        b = self._trigger.get(segment_name)
        if b is None:
            return
        lo, hi, _ = self.buckets[b]
        view = self.g[lo:hi]
        self.launched.append(b)
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            with torch.cuda.stream(self._stream):
                self._stream.wait_event(ev)
                work = self._reduce(view)
            self._pending.append((work, view, lo, hi))
        else:
            self._pending.append((self._reduce(view), view, lo, hi))

    def _reduce(self, view):
        if self.backend == "nccl":
            return dist.all_reduce(view, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        return dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def wait_all(self):
        """Make the compute stream (or the host, for Gloo) wait for every launched bucket."""
        for _ in self.drain():
            pass

    def all_in_flight(self) -> bool:
        """Every bucket has been launched and none has been waited for yet."""
        return len(self._pending) == len(self.buckets) and len(set(self.launched)) == len(self.buckets)

    def drain(self):
        """Yield (lo, hi) for each launched bucket in launch order, right after making the compute stream (or the
        host, for Gloo) wait for THAT bucket only: the caller can update a reduced bucket's parameters while the
        later buckets are still on the wire."""
        pending, self._pending = self._pending, []
        self.launched.clear()
        for work, view, lo, hi in pending:
            work.wait()
            if self.backend != "nccl":
                view.div_(self.world)
            yield lo, hi


def gpt2_segments(flat, plan) -> list:
    """Element ranges of the flat gradient buffer per backward stage, in completion order."""
    def span(params):
        los = [flat.offsets[flat._index[id(p)]] for p in params]
        his = [lo + _pad(p.numel()) for lo, p in zip(los, params)]
        return min(los), max(his)

    def _pad(n):
        return (n + 63) // 64 * 64

    segs = []
    head_params = [p for p in (plan.lnf.weight, plan.lnf.bias, plan.head.weight, plan.head.bias) if p is not None]
    segs.append(("head",) + span(head_params))
    for l in range(plan.n_layer - 1, -1, -1):
        blk = plan.blocks[l]
        ps = [p for m in (blk.ln1, blk.qkv, blk.proj, blk.ln2, blk.fc, blk.proj2) for p in (m.weight, m.bias)
              if p is not None]
        segs.append((f"block{l}",) + span(ps))
    segs.append(("embed",) + span([plan.wte.weight, plan.wpe.weight]))
    return segs
