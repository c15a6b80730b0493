#!/usr/bin/env python
"""Where does a short step go on N GPUs?  Runs one bench block (dcgan | resnet50) exactly as bench.py's run_block does — one
batch per RUN_STEP workload through PyTorchTrialController — and reports, from rank 0: device-timed ms/step, a cProfile of the
host side of the timed steps (top cumulative entries), and a kernel table of three traced steps (torch profiler; never a
timing source). Launch like bench.py: python -m torch.distributed.run --nproc-per-node N ... scripts/gpu_block_probe.py dcgan"""
import cProfile
import io
import json
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main() -> None:
    name = sys.argv[1] if len(sys.argv) > 1 else "dcgan"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(device)
    os.environ.setdefault("DET_REQUIRE_P2P", "1")
    class A:
        warmup, steps_ = 3, steps
    A.steps = steps
    prof = cProfile.Profile() if rank == 0 else None
    from determined_b200.pytorch import _controller as C

    orig = C.PyTorchTrialController._train_for_step
    seen = {"n": 0}

    def timed(self, *a, **k):
        seen["n"] += 1
        if prof is not None and seen["n"] > A.warmup + 6:
            return prof.runcall(orig, self, *a, **k)
        return orig(self, *a, **k)

    C.PyTorchTrialController._train_for_step = timed
    out = bench.run_block(name, A, device, rank, world)
    if rank == 0:
        print(json.dumps({"block": name, "world": world, "omp_threads": torch.get_num_threads(), **{k: out[k] for k in ("value", "ms_per_step", "gpu_launches")}}), flush=True)
        buf = io.StringIO()
        pstats.Stats(prof, stream=buf).sort_stats("cumulative").print_stats(40)
        print(buf.getvalue()[:9000], flush=True)
    if world > 1:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
