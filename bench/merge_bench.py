"""K-AVG merge bandwidth sweep: the fused merge kernel (P2P and NVLS variants)
against NCCL all_reduce + divide over the same buffers, for model sizes from LeNet to VGG-16.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29531 bench/merge_bench.py [--check]

Every number is device time (CUDA events on the launching stream), max over ranks, after warm-up.
Rooflines (bytes that must cross each GPU's NVLink ports per direction, over the measured
770 GB/s per direction, /opt/skills/guides/B200_PROFILING.md):
  P2P   a rank reads its 1/P slice from the P-1 peers and writes the average back to them, and it
        serves the same traffic to its peers: 2*(P-1)/P*B out and the same in   -> t >= 2(P-1)/P*B/BW
  NVLS  the switch reduces and replicates: every byte leaves a GPU once ((P-1)/P*B pulled by
        multimem.ld_reduce + B/P multicast store) and enters it once            -> t >= B/BW
(NCCL's ring/tree all-reduce moves the P2P volume.)
"""
import argparse, ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

SIZES = {"lenet": 44_426, "resnet34": 21_797_672, "vgg16": 138_357_544, "1M": 1 << 20, "8M": 8 << 20}
NVLINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--sizes", default="lenet,1M,resnet34,vgg16")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from kubeml_b200 import _native
    from kubeml_b200.parallel.arena import ArenaLayout, WeightArena
    from kubeml_b200.parallel.device_sync import SpmdDeviceSyncGroup
    lib = _native.lib()
    out = []
    for name in args.sizes.split(","):
        n = (SIZES[name] + 255) // 256 * 256
        mod = torch.nn.Module()
        mod.w = torch.nn.Parameter(torch.empty(n))
        group = SpmdDeviceSyncGroup(rank, world, job_tag=len(out))
        rs = group.rs
        mod.to(dev)
        torch.manual_seed(1000 + rank)
        with torch.no_grad():
            mod.w.copy_(torch.randn(n, device=dev))
        rs._ensure_arena(mod, dev, copy_in=True)
        L = rs.layout
        expected = None
        if args.check:
            ws = [torch.empty(n, device=dev) for _ in range(world)]
            dist.all_gather(ws, rs.arena.weights[:n].clone())
            acc = ws[0].clone()
            for w in ws[1:]:
                acc += w
            expected = acc / world
        gen = [0]

        def launch(use_mc):
            gen[0] += 1
            p = _native.MergeParams()
            for r, a in rs.peers.weights.items():
                p.replica[r] = a; p.ireplica[r] = rs.peers.counters[r]; p.flags[r] = rs.peers.flags[r]
            p.n_f32, p.n_i64, p.n_sgd, p.my_rank = L.n_f32, L.n_i64, 0, rank
            p.participants = p.write_mask = (1 << world) - 1
            p.generation = gen[0] if world > 1 else 0
            p.mcast = rs.peers.mcast if (use_mc and rs.peers.mcast) else None
            p.status = rs.arena.ptr(L.off_status)
            _native.check(lib.kml_kavg_merge(ctypes.byref(p), _native.stream_ptr()), "merge")

        def timed(fn, iters):
            for _ in range(3):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        row = {"model": name, "params": n, "mbytes": round(n * 4 / 1e6, 2), "ranks": world,
               "nvls_available": bool(rs.peers.mcast)}
        if args.check:
            launch(False); torch.cuda.synchronize(); dist.barrier()
            err = float((rs.arena.weights[:n] - expected).abs().max())
            row["p2p_max_abs_err"] = err
            if rs.peers.mcast:
                with torch.no_grad():
                    torch.manual_seed(2000 + rank); rs.arena.weights[:n].copy_(torch.randn(n, device=dev))
                torch.cuda.synchronize(); dist.barrier()
                ws = [torch.empty(n, device=dev) for _ in range(world)]
                dist.all_gather(ws, rs.arena.weights[:n].clone())
                exp2 = sum(ws) / world
                launch(True); torch.cuda.synchronize(); dist.barrier()
                row["nvls_max_abs_err"] = float((rs.arena.weights[:n] - exp2).abs().max())
        row["p2p_ms"] = round(timed(lambda: launch(False), args.iters), 4)
        if rs.peers.mcast:
            row["nvls_ms"] = round(timed(lambda: launch(True), args.iters), 4)
        buf = rs.arena.weights[:n]
        def nccl():
            dist.all_reduce(buf); buf.div_(world)
        row["nccl_allreduce_div_ms"] = round(timed(nccl, args.iters), 4)
        B = n * 4
        p2p_roof = (2 * B * (world - 1) / world) / (NVLINK_GBS * 1e9) * 1e3 if world > 1 else 0.0
        nvls_roof = B / (NVLINK_GBS * 1e9) * 1e3 if world > 1 else 0.0
        row["p2p_roofline_ms"] = round(p2p_roof, 4)
        row["p2p_roofline_fraction"] = round(p2p_roof / row["p2p_ms"], 3)
        if "nvls_ms" in row:
            row["nvls_roofline_ms"] = round(nvls_roof, 4)
            row["nvls_roofline_fraction"] = round(nvls_roof / row["nvls_ms"], 3)
        best = min(row["p2p_ms"], row.get("nvls_ms", 1e9))
        row["speedup_vs_nccl"] = round(row["nccl_allreduce_div_ms"] / best, 2)
        out.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        dist.barrier()
        group.close()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
