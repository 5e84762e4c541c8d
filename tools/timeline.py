"""Kernel timeline of ONE train step as the CUDA graph replays it (CUPTI activity records through
torch.profiler: start, duration and stream of every kernel / memset, in the step's own time base).
usage: python tools/timeline.py [--config N] [--out file.md]     (under torchrun: one file per rank)"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    from torch.profiler import ProfilerActivity, profile
    import induction_b200  # noqa: F401
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.trainer import EpisodeTrainer
    from induction_b200.util.data_loader import JSONFileDataLoader
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=1)
    ap.add_argument("--scaling", default="weak")
    ap.add_argument("--stream", default="global")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = bench.WORKLOADS[args.config]
    B = w["B"] // world if args.scaling == "strong" else w["B"]
    cfg = EpisodeConfig(N=w["N"], K=w["K"], Q=bench.Q, B=B, L=w["L"], E=bench.E, u=w["u"], S=bench.S, keep_prob=0.9, mode="bf16")
    tables = bench.make_tables_for(w)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=0)
    smp = JSONFileDataLoader.from_tables(tables)._device()
    smp.seed(0 if args.stream == "global" else rank)
    tr = EpisodeTrainer(pipe, smp, rank=rank, world=world, stream_mode=args.stream)
    tr.prime()
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")
    for _ in range(5):
        tr.step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(3):
            with torch.cuda.stream(tr.stream):
                flush.fill_(float(i))
            tr.step()
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # the last step = everything after the last L2-flush fill
    fills = [i for i, e in enumerate(evs) if "FillFunctor" in e.name]
    evs = evs[fills[-1] + 1:]
    t0 = evs[0].time_range.start
    end = max(e.time_range.end for e in evs)
    streams = {}
    lines = []
    busy = 0.0
    for e in evs:
        sid = getattr(e, "device_resource_id", None)
        if sid is None:
            sid = getattr(e, "stream", -1)
        lane = streams.setdefault(sid, len(streams))
        name = e.name.replace("indnet::", "").replace("void ", "").split("(")[0][:56]
        lines.append("| %8.1f | %7.1f | %d | %s |" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, lane, name))
        busy += e.time_range.end - e.time_range.start
    out = ["# One train step as the CUDA graph replays it - config %d (%s), rank %d of %d" % (args.config, w["name"] if "name" in w else "", rank, world), "",
           "`python tools/timeline.py --config %d` (torch.profiler / CUPTI kernel records; L2 flushed before the step; "
           "times in us from the step's first kernel; lane = stream the graph node ran on).  Step span %.1f us, "
           "sum of kernel durations %.1f us over %d launches." % (args.config, end - t0, busy, len(evs)), "",
           "| start us | dur us | lane | kernel |", "|---|---|---|---|"] + lines
    text = "\n".join(out) + "\n"
    path = args.out or os.path.join(ROOT, "gpurun_out", "timeline_c%d_r%d.md" % (args.config, rank))
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write(text)
    if rank == 0:
        print(text)
    tr.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
