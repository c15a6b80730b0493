"""Multi-GPU check of the peer-memory collectives (run under torch.distributed.run, one rank per GPU):
fused all-reduce (scale + sumsq + found_inf) vs NCCL, broadcast, metric reduce, and bandwidth timing."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402
from determined_b200.parallel.comm import make_communicator  # noqa: E402


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import time

    t0 = time.time()
    comm = make_communicator(dev)
    t_comm = time.time() - t0
    world = comm.size
    if rank == 0:
        print(f"communicator ready in {t_comm:.2f} s (backend {dist.get_backend()})", flush=True)
    ok = True
    if rank == 0:
        print("comm:", comm.describe(), flush=True)
    results = {}
    ALGOS = {"auto": 0, "nvls": 1, "lsu": 2, "bulk": 3}
    for algo_name, n in [(a, n) for a in ALGOS for n in (4096, 1 << 20, 25 * (1 << 20) + 64)]:
        algo = ALGOS[algo_name]
        torch.manual_seed(100 + rank)
        g = comm.alloc(n, torch.float32, "grad", dev)
        g.copy_(torch.randn(n, device=dev))
        ref = g.clone()
        dist.all_reduce(ref)
        ref *= 0.125
        sumsq = torch.zeros(1, device=dev)
        finf = torch.zeros(1, dtype=torch.int32, device=dev)
        applied = comm.allreduce_grad_bucket(g, 0.125, sumsq_dev=sumsq, found_inf_dev=finf, algo=algo)
        torch.cuda.synchronize()
        err = ((g - ref).abs().max() / ref.abs().max()).item()
        ss_ref = ref.double().pow(2).sum().item()
        ss_err = abs(sumsq.item() - ss_ref) / ss_ref
        good = applied and err < 1e-5 and ss_err < 1e-3 and finf.item() == 0
        ok &= good
        if rank == 0:
            print(f"{'PASS' if good else 'FAIL'} allreduce[{algo_name}] n={n} applied={applied} err={err:.2e} sumsq_err={ss_err:.2e}", flush=True)
        # non-finite propagation
        g.copy_(torch.randn(n, device=dev))
        if rank == world - 1:
            g[n // 2] = float("nan")
        finf.zero_()
        comm.allreduce_grad_bucket(g, 1.0, found_inf_dev=finf, algo=algo)
        torch.cuda.synchronize()
        flags = comm.allgather_object(int(finf.item()))
        # only the owner of the slice containing n//2 scans it, but the flag is exchanged inside the kernel: every rank
        # must have raised it, or ranks would disagree about skipping the step
        good = all(f == 1 for f in flags)
        ok &= good
        if rank == 0:
            print(f"{'PASS' if good else 'FAIL'} allreduce[{algo_name}] found_inf flags={flags}", flush=True)
    # bf16 on the wire (optimizations.gradient_compression): multimem bf16x2 path, then the peer-load path
    for label, env in (("nvls", None), ("lsu", "1")):
        if env:
            os.environ["DET_DISABLE_NVLS"] = env
        for n in (4096, 25 * (1 << 20) + 64):
            torch.manual_seed(300 + rank)
            g = comm.alloc(n, torch.float32, "grad", dev)
            g.copy_(torch.randn(n, device=dev))
            ref = g.to(torch.bfloat16).float()          # what each rank puts on the wire
            dist.all_reduce(ref)
            ref *= 0.125
            sumsq = torch.zeros(1, device=dev)
            finf = torch.zeros(1, dtype=torch.int32, device=dev)
            comm.compress_gradients = True
            applied = comm.allreduce_grad_bucket(g, 0.125, sumsq_dev=sumsq, found_inf_dev=finf)
            comm.compress_gradients = False
            torch.cuda.synchronize()
            # two bf16 roundings (the switch's output, the scaled result): relative error of a few 2^-8 per element
            err = ((g - ref).abs().max() / ref.abs().max()).item()
            ss_ref = ref.double().pow(2).sum().item()
            ss_err = abs(sumsq.item() - ss_ref) / ss_ref
            same = comm.allgather_object(float(g.double().sum().item()))
            good = applied and err < 2e-2 and ss_err < 1e-2 and finf.item() == 0 and all(x == same[0] for x in same)
            ok &= good
            if rank == 0:
                print(f"{'PASS' if good else 'FAIL'} allreduce[bf16wire/{label}] n={n} err={err:.2e} sumsq_err={ss_err:.2e} identical_on_all_ranks={all(x == same[0] for x in same)}", flush=True)
            g.copy_(torch.randn(n, device=dev))
            if rank == world - 1:
                g[n // 2] = float("inf")
            finf.zero_()
            comm.compress_gradients = True
            comm.allreduce_grad_bucket(g, 1.0, found_inf_dev=finf)
            comm.compress_gradients = False
            torch.cuda.synchronize()
            flags = comm.allgather_object(int(finf.item()))
            good = all(f == 1 for f in flags)
            ok &= good
            if rank == 0:
                print(f"{'PASS' if good else 'FAIL'} allreduce[bf16wire/{label}] found_inf flags={flags}", flush=True)
        os.environ.pop("DET_DISABLE_NVLS", None)
    # pipelined form: reduce-scatter + statistics, then the all-gather slice by slice in rank order (side stream), with the
    # consumer waiting per slice — here the "optimizer" is a scale-by-2 of each slice as it lands
    os.environ["DET_AR_PIPELINE"] = "1"
    for label, env in (("nvls", None), ("lsu", "1")):
        if env:
            os.environ["DET_DISABLE_NVLS"] = env
        for rep in range(3):
            n = 25 * (1 << 20) + 64
            torch.manual_seed(500 + rank + 10 * rep)
            g = comm.alloc(n, torch.float32, "grad", dev)
            g.copy_(torch.randn(n, device=dev))
            ref = g.clone()
            dist.all_reduce(ref)
            ref *= 0.125
            bounds = comm.pipeline_plan(g)
            assert bounds is not None, "bucket does not qualify for the pipelined form"
            C = len(bounds)
            sumsq = torch.zeros(1, device=dev)
            finf = torch.zeros(1, dtype=torch.int32, device=dev)
            main = torch.cuda.current_stream(dev)
            side = torch.cuda.Stream(dev, priority=-1)
            comm.reduce_scatter_stats(g, C, 0.125, sumsq_dev=sumsq, found_inf_dev=finf)
            side.wait_stream(main)
            sent = []
            with torch.cuda.stream(side):
                for c in range(C):
                    comm.allgather_send(g, C, c)
                    ev = torch.cuda.Event()
                    ev.record(side)
                    sent.append(ev)
            for c, (lo, hi) in enumerate(bounds):
                main.wait_event(sent[c])
                comm.wait_chunk(c)
                g[lo:hi].mul_(2.0)
            torch.cuda.synchronize()
            err = ((g - 2.0 * ref).abs().max() / ref.abs().max()).item()
            ss_ref = ref.double().pow(2).sum().item()
            ss_err = abs(sumsq.item() - ss_ref) / ss_ref
            good = err < 1e-5 and ss_err < 1e-3 and finf.item() == 0
            ok &= good
            if rank == 0:
                print(f"{'PASS' if good else 'FAIL'} pipelined[{label}] rep={rep} err={err:.2e} sumsq_err={ss_err:.2e}", flush=True)
        os.environ.pop("DET_DISABLE_NVLS", None)
    # broadcast
    p = comm.alloc(1 << 20, torch.float32, "param", dev)
    p.fill_(float(rank + 1))
    comm.broadcast_tensors([p], root=0)
    torch.cuda.synchronize()
    good = bool((p == 1.0).all().item())
    ok &= good
    if rank == 0:
        print(f"{'PASS' if good else 'FAIL'} broadcast", flush=True)
    # metric reduce
    v = torch.tensor([float(rank), float(rank) * 2 + 1, -float(rank)], device=dev)
    avg = comm.reduce_metrics(v, F.RED_AVG)
    mx = comm.reduce_metrics(v, F.RED_MAX)
    wavg = comm.reduce_metrics(v, F.RED_AVG, weights=[float(r + 1) for r in range(world)])
    torch.cuda.synchronize()
    exp_avg = sum(range(world)) / world
    wsum = sum(r + 1 for r in range(world))
    exp_w = sum(r * (r + 1) for r in range(world)) / wsum
    good = abs(avg[0].item() - exp_avg) < 1e-5 and abs(mx[0].item() - (world - 1)) < 1e-6 and abs(wavg[0].item() - exp_w) < 1e-5
    ok &= good
    if rank == 0:
        print(f"{'PASS' if good else 'FAIL'} metric reduce avg={avg.tolist()} max={mx.tolist()} wavg={wavg.tolist()}", flush=True)

    # timing: BERT-base sized bucket (110M fp32 = 440 MB) fused vs NCCL
    n = 110 * (1 << 20)
    g = comm.alloc(n, torch.float32, "grad", dev)
    g.normal_()
    nccl_buf = torch.randn(n, device=dev)

    def timeit(fn, iters=10):
        for _ in range(3):
            fn()
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    sumsq = torch.zeros(1, device=dev)
    finf = torch.zeros(1, dtype=torch.int32, device=dev)
    t_fused = timeit(lambda: comm.allreduce_grad_bucket(g, 1.0 / world, sumsq_dev=sumsq, found_inf_dev=finf))

    def nccl_path():
        dist.all_reduce(nccl_buf)
        nccl_buf.mul_(1.0 / world)
        torch.isfinite(nccl_buf).all()
        nccl_buf.pow(2).sum()

    t_nccl = timeit(nccl_path)
    t_nccl_only = timeit(lambda: dist.all_reduce(nccl_buf))
    bytes_moved = 2 * (world - 1) / world * n * 4
    by_algo = {}
    for algo_name, algo in ALGOS.items():
        if algo == 0:
            continue
        for mb in ((0,) if algo != 3 else (24, 48, 64, 96, 128, 148)):
            t = timeit(lambda: comm.allreduce_grad_bucket(g, 1.0 / world, sumsq_dev=sumsq, found_inf_dev=finf, algo=algo, max_blocks=mb), iters=8)
            by_algo[f"{algo_name}_b{mb}"] = {"ms": round(t, 3), "busbw_gbs": round(bytes_moved / t / 1e6, 1)}
    side = torch.cuda.Stream(dev, priority=-1)

    def pipelined():
        main = torch.cuda.current_stream(dev)
        C = comm.PIPE_CHUNKS
        comm.reduce_scatter_stats(g, C, 1.0 / world, sumsq_dev=sumsq, found_inf_dev=finf)
        side.wait_stream(main)
        with torch.cuda.stream(side):
            for c in range(C):
                comm.allgather_send(g, C, c)
        main.wait_stream(side)
        for c in range(C):
            comm.wait_chunk(c)

    t = timeit(pipelined, iters=8)
    by_algo["pipelined_rs_then_ag_in_chunks"] = {"ms": round(t, 3), "busbw_gbs": round(bytes_moved / t / 1e6, 1)}
    t = timeit(lambda: comm.reduce_scatter_stats(g, comm.PIPE_CHUNKS, 1.0 / world, sumsq_dev=sumsq, found_inf_dev=finf), iters=8)
    by_algo["reduce_scatter_stats_only"] = {"ms": round(t, 3)}
    comm.compress_gradients = True
    t = timeit(lambda: comm.allreduce_grad_bucket(g, 1.0 / world, sumsq_dev=sumsq, found_inf_dev=finf), iters=8)
    comm.compress_gradients = False
    by_algo["bf16wire_nvls"] = {"ms": round(t, 3), "busbw_gbs_fp32_equiv": round(bytes_moved / t / 1e6, 1)}
    if rank == 0:
        print("allreduce by algorithm:", json.dumps(by_algo), flush=True)
    # bandwidth against the number of CTAs: what a collective squeezed onto a few reserved SMs can deliver
    by_blocks = {}
    for mb in (2, 4, 8, 16, 32, 64, 0):
        t = timeit(lambda: comm.allreduce_grad_bucket(g, 1.0 / world, max_blocks=mb), iters=5)
        by_blocks[str(mb or "all")] = {"ms": round(t, 3), "busbw_gbs": round(bytes_moved / t / 1e6, 1)}
    if rank == 0:
        print("allreduce vs CTAs:", json.dumps(by_blocks), flush=True)
    results = {"world": world, "bucket_mb": n * 4 / 1e6, "fused_ms": round(t_fused, 3), "nccl_allreduce_only_ms": round(t_nccl_only, 3),
               "nccl_plus_scale_finite_norm_ms": round(t_nccl, 3), "fused_busbw_gbs": round(bytes_moved / t_fused / 1e6, 1),
               "nccl_busbw_gbs": round(bytes_moved / t_nccl_only / 1e6, 1), "by_algo": by_algo, "comm": comm.describe()}
    if rank == 0:
        print(json.dumps(results), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/comm_timing_w{world}.json", "w") as f:
            json.dump(results, f, indent=1)
        print("SUMMARY", "PASS" if ok else "FAIL", flush=True)
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
