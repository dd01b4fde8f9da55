"""One process per GPU: the little plumbing bench.py needs from torch.distributed.

Windows are independent units, so the data path has no collective; ranks only agree on (a) who owns which windows
(``pipeline.shard_bounds``), (b) a barrier around the timed region and (c) max / sum of a few scalars.  Backend
``nccl`` on GPUs, ``gloo`` in the CPU tests (tests/test_sharding_gloo.py)."""
import os


class Ranks:
    def __init__(self, backend=None, device=None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        self.device = device
        if self.world > 1:
            import torch
            import torch.distributed as dist

            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            kwargs = {}
            if backend == "nccl":
                self.device = torch.device("cuda", self.local_rank) if device is None else device
                kwargs["device_id"] = self.device
            else:
                self.device = torch.device("cpu")
            dist.init_process_group(backend, **kwargs)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def _reduce(self, x, op):
        if self.dist is None:
            return float(x)
        import torch

        t = torch.tensor([float(x)], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x):
        return self._reduce(x, None if self.dist is None else self.dist.ReduceOp.MAX)

    def sum(self, x):
        return self._reduce(x, None if self.dist is None else self.dist.ReduceOp.SUM)

    def gather_ints(self, values):
        """All ranks' lists of ints (equal lengths), concatenated in rank order."""
        if self.dist is None:
            return list(values)
        import torch

        t = torch.tensor(list(values), dtype=torch.int64, device=self.device)
        out = [torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [int(v) for part in out for v in part.tolist()]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()
            self.dist = None
