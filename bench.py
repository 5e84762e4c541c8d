#!/usr/bin/env python
"""bench.py - train-step episodes/sec of the Induction-Network episode pipeline on B200.

  python bench.py --gpus N --steps K --warmup W                      # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K --warmup W     # CPU arm (oracle port; TF 1.x absent)

A "step" is one pass of the hot path over one batch of synthetic episodes: device sampler + gather ->
forward -> loss -> backward -> gradient all-reduce (N > 1) -> TF Adam, replayed as one CUDA graph by
``induction_b200.trainer.EpisodeTrainer`` - the same object ``Base.train`` / ``train.py`` drive.

Workloads are numbered as in SURVEY.md section 8 / BASELINE.md (config i = BASELINE.json configs[i-1]).
The default is config 1, the configuration the metric is quoted on (5-way 5-shot).  ``--scaling weak``
(default): every rank runs the config's batch; ``--scaling strong``: the config's batch is the GLOBAL batch,
sharded over the ranks.  Either way all ranks walk the same sampler streams and gather their own episodes, so
the global batch is bit-identical for any N.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# SURVEY section 8 numbering.  Q = 5 everywhere (train.py:33); E = 50; S = 100.
WORKLOADS = {
    1: dict(name="config 1: 5-way 5-shot, 4 episodes, L=40, u=128", N=5, K=5, B=4, L=40, u=128),
    2: dict(name="config 2: 5-way 1-shot, 64 episodes, L=40, u=128", N=5, K=1, B=64, L=40, u=128),
    3: dict(name="config 3: 10-way 5-shot, 128 episodes, L=40, u=128", N=10, K=5, B=128, L=40, u=128),
    4: dict(name="config 4: 10-way 10-shot, 1024 episodes, L=40, u=128", N=10, K=10, B=1024, L=40, u=128),
    5: dict(name="config 5: encoder+induction inference, 20-way 10-shot, L=128, u=256, 512 episodes per GPU "
                 "(4096 across 8)", N=20, K=10, B=512, L=128, u=256, inference=True, chunk=32),
}
E, S, Q, VOCAB, RELS, PER_REL = 50, 100, 5, 400000, 64, 700
MIN_TIMED_S = 0.5          # the --steps block is repeated until this much GPU time has been timed


def fwd_flops_per_episode(N, K, L, u, relation=True):
    """SURVEY.md section 8 FLOP formula (multiply-add = 2), forward, one episode."""
    D, H, n_s, n_q = E + 10, 2 * u, N * (K + Q), N * Q
    f = (2 * n_s * L * D * 4 * u * 2 + 2 * n_s * L * u * 4 * u * 2 + 2 * n_s * L * H * H + 4 * n_s * L * H
         + 2 * N * K * H * H + 3 * 4 * N * K * H)
    if relation:
        f += 2 * N * H * H * S + 2 * N * n_q * H * S + 2 * n_q * N * N * S
    return f


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML every few ms while ``armed`` (the timed
    region).  In-process NVML: spawning nvidia-smi per sample is too slow for sub-second regions."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False
        self.ready = self.armed = False
        self.max_mhz = None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = int(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                nv.nvmlDeviceGetCurrentClocksThrottleReasons
        except Exception:
            return
        self.ready = True
        while not self.stop_flag:
            if self.armed:
                try:
                    self.samples.append((int(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), int(get_reasons(h))))
                except Exception:
                    pass
            time.sleep(0.004)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unsampled"]}
        mhz = sorted(s[0] for s in self.samples)
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                "sw_power_cap": 0x4, "hw_power_brake_slowdown": 0x80}
        reasons = [n for n, b in bits.items() if any(s[1] & b for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.samples), "source": "NVML, 4 ms period, timed region only"}


def make_tables_for(w, seed=0):
    from induction_b200.util.synthetic import make_tables
    return make_tables(RELS, PER_REL, w["L"], VOCAB, E, seed=seed)


def host_episodes(w, tables, B, pool, seed=1):
    """``pool`` host batches of B episodes drawn with a private RandomState (same distribution as the
    sampler's): the pinned-host inputs of the e2e leg and the CPU arm's inputs."""
    import numpy as np
    rs = np.random.RandomState(seed)
    n_s, n_q = w["N"] * (w["K"] + Q), w["N"] * Q
    batches = []
    for _ in range(pool):
        rows = np.zeros((B, n_s), dtype=np.int64)
        label = np.zeros((B, n_q), dtype=np.int32)
        for b in range(B):
            classes = rs.choice(RELS, w["N"], replace=False)
            sup, qry, lab = [], [], []
            for i, c in enumerate(classes):
                idx = c * PER_REL + rs.choice(PER_REL, w["K"] + Q, replace=False)
                sup += list(idx[: w["K"]]); qry += list(idx[w["K"]:]); lab += [i] * Q
            perm = rs.permutation(n_q)
            rows[b] = np.array(sup + [qry[p] for p in perm])
            label[b] = np.array([lab[p] for p in perm])
        batches.append(({k: np.ascontiguousarray(tables[k][rows]) for k in ("word", "pos1", "pos2", "mask")}, label))
    return batches


# ------------------------------------------------------------------------------------------------
def cpu_episodes_per_sec(w, tables, steps, warmup, episodes_per_step, budget_s=None):
    """The CPU arm: what the reference's loop body does per episode (model/base.py:136-152) - the host sampler
    ``next_one_tf`` (restated in oracle/sampler_oracle.py, pinned to the reference's own outputs) + forward +
    autograd backward + TF-form Adam (oracle port, torch-CPU fp32) - on a bounded sample of the workload, all
    host threads.  Returns (episodes/s, cores, steps timed, sampler share of the time)."""
    import random
    import numpy as np
    import torch
    from oracle import induction_oracle as orc
    from oracle import sampler_oracle as so
    from induction_b200.engine import init_tf_params, tf_param_shapes
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    params = init_tf_params(tf_param_shapes(w["N"], w["L"], E, w["u"], S), seed=0)
    m = {k: torch.zeros_like(v) for k, v in params.items()}
    v = {k: torch.zeros_like(x) for k, x in params.items()}
    wt = torch.from_numpy(tables["word_vec_mat"])
    scopes = [(int(a), int(b)) for a, b in tables["rel2scope"].values()]
    Bc = episodes_per_step
    random.seed(0); np.random.seed(0)
    t_smp = [0.0]

    def one(step):
        t0 = time.perf_counter()
        eps = [so.next_one_tf_stdlib(tables, scopes, w["N"], w["K"], Q) for _ in range(Bc)]
        t_smp[0] += time.perf_counter() - t0
        ti = {k: torch.from_numpy(np.stack([e[0][k] for e in eps])) for k in ("word", "pos1", "pos2", "mask")}
        lab = torch.from_numpy(np.stack([e[1] for e in eps])).long()
        _, g = orc.loss_and_grads(params, wt, ti["word"], ti["pos1"], ti["pos2"], ti["mask"], lab, w["N"], w["K"], Q)
        orc.tf_adam_step(params, g, m, v, global_step=step)

    for s in range(warmup):
        one(s)
    t_smp[0] = 0.0
    t0 = time.perf_counter()
    done = 0
    for s in range(steps):
        one(warmup + s)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return Bc * done / dt, cores, done, t_smp[0] / dt


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if w.get("inference"):
        print(json.dumps({"impl": "reference", "unavailable": "config 5 is an inference sweep; the CPU arm times the train step (configs 1-4)"}))
        return
    tables = make_tables_for(w)
    Bc = min(w["B"], 4)
    warm = min(args.warmup, 2)
    eps, cores, done, smp_share = cpu_episodes_per_sec(w, tables, args.steps, warm, Bc, budget_s=150.0)
    sample = ("%d episodes/step (of B=%d) x %d steps: host next_one_tf + fwd + autograd bwd + TF Adam, torch-CPU fp32 "
              "oracle port (sampler = %.1f %% of the time)" % (Bc, w["B"], done, 100 * smp_share))
    print(json.dumps({
        "impl": "reference", "metric": "train-step episodes/sec", "value": eps, "unit": "episodes/s",
        "n_gpus": args.gpus, "steps": done, "warmup": warm, "ms_per_step": 1000.0 * Bc / eps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "note": "reference TF 1.x graph unavailable on this box; CPU oracle port timed"},
        "cpu_baseline": {"value": eps, "unit": "episodes/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": eps, "unit": "episodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------------------------
def timed_blocks(torch, stream, flush, run_step, steps, world, dist, min_s=MIN_TIMED_S, max_blocks=400):
    """Times blocks of exactly ``steps`` steps (CUDA events on the launching stream, an L2-flushing write
    between steps outside each step's event pair) until ``min_s`` of GPU time is covered.  Returns the per-step
    times in ms (this rank) and the number of blocks."""
    times = []
    blocks = 0
    while True:
        evs = []
        for i in range(steps):
            with torch.cuda.stream(stream):
                flush.fill_(float(i))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            run_step()
            e1.record(stream)
            evs.append((e0, e1))
        stream.synchronize()
        times += [a.elapsed_time(b) for a, b in evs]
        blocks += 1
        # every rank must run the same number of blocks (collectives inside the step): decide together
        t = torch.tensor([sum(times)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if float(t[0]) / 1000.0 >= min_s or blocks >= max_blocks:
            break
    return times, blocks


def summarize(times_ms):
    s = sorted(times_ms)
    return sum(s) / len(s), s[len(s) // 2]


def run_b200(args, w):
    import numpy as np
    import torch
    import torch.distributed as dist
    import induction_b200  # noqa: F401
    from induction_b200._lib import LIB
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.trainer import EpisodeTrainer
    from induction_b200.util.data_loader import JSONFileDataLoader

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL's own logging (NCCL_DEBUG / NCCL_DEBUG_FILE) is left exactly as the caller exported it
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    if w.get("inference"):
        return run_inference(args, w, world, rank, local, peaks)

    B = w["B"]
    if args.scaling == "strong":
        if B % world:
            raise SystemExit("--scaling strong: the global batch %d is not divisible by %d ranks" % (B, world))
        B //= world
    cfg = EpisodeConfig(N=w["N"], K=w["K"], Q=Q, B=B, L=w["L"], E=E, u=w["u"], S=S, keep_prob=args.keep_prob,
                        mode=args.mode)
    tables = make_tables_for(w)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=0)
    loader = JSONFileDataLoader.from_tables(tables)
    smp = loader._device()
    # --stream global: random.seed(0); np.random.seed(0) on EVERY rank = one global episode stream;
    # --stream per-rank: rank r seeds its own streams with r (the serial walk then covers only the rank's episodes)
    smp.seed(0 if args.stream == "global" else rank)
    trainer = EpisodeTrainer(pipe, smp, rank=rank, world=world, cuda_graph=not args.no_graph, stream_mode=args.stream)
    trainer.prime()
    stream = trainer.stream
    POOL = 4
    batches = host_episodes(w, tables, cfg.B, POOL, seed=1 + rank)
    dev_pool = [({k: torch.from_numpy(v).cuda() for k, v in inp.items()}, torch.from_numpy(lab).cuda()) for inp, lab in batches]
    # host episodes: pinned, in the device input buffer's layout (EpisodePipeline.pack_inputs) = one H2D copy per step
    host_pool = [(pipe.pack_inputs(inp, lab), None) for inp, lab in batches]
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")   # 256 MB > 126 MB L2

    # kernels launched by one step (counted on an eager pass of the same body)
    with torch.cuda.stream(stream):
        l0 = LIB.launch_count()
        trainer.step_body()
        launches_per_step = LIB.launch_count() - l0
    stream.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        trainer.step()
    stream.synchronize()
    t_wait = time.time()
    while not sampler.ready and time.time() - t_wait < 3.0:
        time.sleep(0.01)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.armed = True
    times, blocks = timed_blocks(torch, stream, flush, trainer.step, args.steps, world, dist)
    torch.cuda.synchronize()
    time.sleep(0.03)                      # let the last in-flight sample of the region land
    sampler.armed = False
    sampler.stop_flag = True
    if world > 1:
        dist.barrier()
    loss_after, acc_after = trainer.read()
    trainer.check_sampler()

    # ---- same steps without the sampler in the loop (episodes already resident in HBM)
    pool_i = [0]

    def resident_step():
        with torch.cuda.stream(stream):
            pipe.set_inputs(*dev_pool[pool_i[0] % POOL])
        pool_i[0] += 1
        trainer.step_on()

    res_times, _ = timed_blocks(torch, stream, flush, resident_step, args.steps, world, dist, min_s=0.2)

    # ---- keep_prob = 1.0 (dropout off) for comparison with round 1, resident inputs, own graph
    keep1 = None
    if args.keep_prob < 1.0 and world == 1:
        pipe.cfg.keep_prob = 1.0
        g1 = None
        with torch.cuda.stream(stream):
            trainer.train_body()
            stream.synchronize()
            if not args.no_graph:
                g1 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g1, stream=stream):
                    trainer.train_body()

        def keep1_step():
            with torch.cuda.stream(stream):
                g1.replay() if g1 is not None else trainer.train_body()

        k1_times, _ = timed_blocks(torch, stream, flush, keep1_step, args.steps, world, dist, min_s=0.2)
        pipe.cfg.keep_prob = args.keep_prob
        g1 = None
        keep1 = {"keep_prob": 1.0, "ms_per_step": summarize(k1_times)[0],
                 "value": cfg.B / (summarize(k1_times)[0] / 1000.0), "unit": "episodes/s",
                 "note": "same step with dropout off, episodes resident (the round-1 setting)"}

    # ---- roofline leg: per-call CUDA-event timing of the same step, eager launches
    prof_steps = min(args.steps, 5)
    with torch.cuda.stream(stream):
        pipe.profile = []
        smp_ev = []
        for i in range(prof_steps):
            flush.fill_(1.0)
            trainer._gather()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            trainer._walk()                      # serial here: timed on its own
            e1.record()
            smp_ev.append((e0, e1))
            trainer.train_body()
        prof = pipe.profile_summary()
        pipe.profile = None
        sampler_walk_ms = sum(a.elapsed_time(b) for a, b in smp_ev) / len(smp_ev)
    trainer.check_sampler()

    # ---- encoder leg: sentences/sec of gather -> embedding -> BiLSTM -> attention pool, forward
    enc_ev = []
    with torch.cuda.stream(stream):
        for i in range(max(10, min(args.steps, 50))):
            flush.fill_(2.0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            trainer._gather()
            pipe.encode(train=False)
            e1.record()
            enc_ev.append((e0, e1))
    stream.synchronize()
    enc_ms = summarize([a.elapsed_time(b) for a, b in enc_ev[2:]])[0]

    # ---- e2e leg: the user-facing call with HOST buffers (pinned): H2D of the step's episodes + step + D2H
    e2e_steps = max(10, min(args.steps, 50))
    h2d = host_pool[0][0].numel() * 4
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        trainer.step_on(*host_pool[i % POOL])         # H2D from pinned memory, then the step graph
        res = trainer.read()                          # D2H {loss, accuracy} (synchronises)
    torch.cuda.synchronize()
    e2e_sync_ms = (time.perf_counter() - t0) * 1000.0
    # the same loop with every step's {loss, accuracy} still read by the host, one step behind (pinned slot +
    # event): the host never drains the GPU, so its launch work hides under the previous step
    if world > 1:
        dist.barrier()
    e2e_losses = []
    t0 = time.perf_counter()
    pend = None
    for i in range(e2e_steps):
        trainer.step_on(*host_pool[i % POOL])
        nxt = trainer.read_async()
        if pend is not None:
            e2e_losses.append(pend.get())
        pend = nxt
    e2e_losses.append(pend.get())
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1000.0
    assert len(e2e_losses) == e2e_steps and all(l == l for l, _ in e2e_losses)
    # ... and the loop Base.train runs: device sampler in the step, {loss, accuracy} read back every step
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        trainer.step()
        res = trainer.read()
    torch.cuda.synchronize()
    e2e_smp_ms = (time.perf_counter() - t0) * 1000.0
    trainer.check_sampler()

    mean_ms, med_ms = summarize(times)
    res_mean, _ = summarize(res_times)
    t = torch.tensor([mean_ms, e2e_ms, res_mean, e2e_smp_ms, med_ms, e2e_sync_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    mean_ms, e2e_ms, res_mean, e2e_smp_ms, med_ms, e2e_sync_ms = (float(x) for x in t)
    value = world * cfg.B / (mean_ms / 1000.0)
    e2e_value = world * cfg.B * e2e_steps / (e2e_ms / 1000.0)

    # ---- config 1 only: the same 5-way 5-shot step at the per-GPU batch that fills the chip
    filled = None
    if args.config == 1 and world == 1 and not args.no_filled:
        cfgF = EpisodeConfig(N=w["N"], K=w["K"], Q=Q, B=64, L=w["L"], E=E, u=w["u"], S=S, keep_prob=args.keep_prob,
                             mode=args.mode)
        storeF = ParamStore(tf_param_shapes(cfgF.N, cfgF.L, cfgF.E, cfgF.u, cfgF.S))
        storeF.load_state_dict(init_tf_params(storeF.tf_shapes, seed=0))
        pipeF = EpisodePipeline(cfgF, storeF, torch.from_numpy(tables["word_vec_mat"]), seed=0)
        trF = EpisodeTrainer(pipeF, smp, cuda_graph=not args.no_graph)
        smp.seed(0)
        trF.prime()
        for _ in range(3):
            trF.step()
        f_times, _ = timed_blocks(torch, trF.stream, flush, trF.step, args.steps, 1, dist, min_s=0.3)
        f_mean, f_med = summarize(f_times)
        filled = {"workload": "5-way 5-shot, 64 episodes per GPU (the batch that fills 148 SMs), device sampler in the step",
                  "value": 64 / (f_mean / 1000.0), "unit": "episodes/s", "ms_per_step": f_mean, "ms_per_step_median": f_med,
                  "step_tflops": 64 / (f_mean / 1000.0) * 3 * fwd_flops_per_episode(cfg.N, cfg.K, cfg.L, cfg.u) / 1e12}
        trF.close()

    if rank == 0:
        roof, share = roofline(args, cfg, store, prof, prof_steps, peaks, sampler.summary())
        if os.environ.get("INDNET_SHARES_OUT"):        # full per-call table of the eager profile leg (tuning aid)
            with open(os.environ["INDNET_SHARES_OUT"], "w") as f:
                json.dump({k: {"calls": v[0], "ms_total": v[1], "ms_per_call": v[1] / v[0], "ms_per_step": v[1] / prof_steps}
                           for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])}, f, indent=1)
        step_tflops = value * 3 * fwd_flops_per_episode(cfg.N, cfg.K, cfg.L, cfg.u) / 1e12
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            Bc = min(cfg.B, 4)
            eps, cores, done, smp_share = cpu_episodes_per_sec(w, tables, steps=50, warmup=1, episodes_per_step=Bc,
                                                               budget_s=args.cpu_seconds)
            cpu = {"value": eps, "unit": "episodes/s", "cores": cores, "kind": "port",
                   "sample": "%d episodes/step x %d steps of the same workload (host next_one_tf + fwd + autograd bwd + "
                             "TF Adam; sampler %.1f %% of the time), torch-CPU fp32 oracle; reference TF 1.x unavailable "
                             "on this box" % (Bc, done, 100 * smp_share)}
        line = json.dumps({
            "metric": "train-step episodes/sec", "value": value, "unit": "episodes/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": mean_ms, "ms_per_step_median": med_ms,
            "timed_blocks": blocks, "timed_steps": blocks * args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32" if args.mode == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "survey_config": args.config, "episodes_per_gpu": cfg.B,
                       "global_batch": cfg.B * world, "Q": Q, "E": E, "S": S, "routing_iters": 3,
                       "keep_prob": args.keep_prob, "mode": args.mode, "vocab": VOCAB, "relations": RELS,
                       "instances_per_relation": PER_REL,
                       "parallelism": ("episode-sharded dp%d (one global sampler stream, rank r owns episodes [r*B,(r+1)*B))" % world)
                       if args.stream == "global" else ("episode-sharded dp%d (one sampler stream per rank)" % world),
                       "sampler_stream": args.stream,
                       "cuda_graph": trainer.graph_step is not None,
                       "l2": "256 MB buffer written between timed iterations (outside each step's event pair)",
                       "timing": "blocks of --steps graph replays repeated until >= %.1f s of GPU time; value = B / mean step time" % MIN_TIMED_S},
            "e2e": {"value": e2e_value, "unit": "episodes/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                    "steps": e2e_steps,
                    "timer": "host perf_counter around: pinned-host episodes H2D + step graph + D2H of {loss, acc}, every "
                             "step; the host reads step t's result while step t+1 runs (EpisodeTrainer.read_async)",
                    "synchronous": {"value": world * cfg.B * e2e_steps / (e2e_sync_ms / 1000.0), "unit": "episodes/s",
                                    "note": "same loop, the host drains the GPU after every step (trainer.read())"},
                    "with_device_sampler": {"value": world * cfg.B * e2e_steps / (e2e_smp_ms / 1000.0), "unit": "episodes/s",
                                            "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 8,
                                            "note": "the loop Base.train runs: bit-exact device sampler inside the step graph "
                                                    "(no host episodes to copy), {loss, acc} read back every step"}},
            "gpu_launches": int(launches_per_step * args.steps * blocks),
            "launches_per_step": int(launches_per_step),
            "clocks": sampler.summary(),
            "roofline": roof,
            "step_share": share,
            "step_tflops": step_tflops,
            "train_only": {"value": world * cfg.B / (res_mean / 1000.0), "unit": "episodes/s", "ms_per_step": res_mean,
                           "note": "fwd+bwd+all-reduce+Adam on episodes already resident in HBM (sampler not in the loop)"},
            "keep_prob_1": keep1,
            "filled_chip": filled,
            "encoder": {"value": world * cfg.X / (enc_ms / 1000.0), "unit": "sentences/s", "ms": enc_ms,
                        "sentences_per_gpu": cfg.X,
                        "what": "gather -> embedding -> BiLSTM -> self-attention pool, forward only, eval mode (SURVEY 8d)"},
            "sampler": {"where": "device, bit-exact MT19937 x2 (csrc/sampler.cu); the walk for step t+1 runs on a side "
                                 "stream inside the step graph", "walk_ms_serial": sampler_walk_ms,
                        "walked_episodes": trainer.global_B},
            "cpu_baseline": cpu,
            "loss_after": loss_after, "accuracy_after": acc_after,
        })
    teardown(trainer, world, line if rank == 0 else None)


def roofline(args, cfg, store, prof, prof_steps, peaks, clocks):
    """Roofline of the step's dominant call from its eager CUDA-event time (DESIGN.md sections 4-5)."""
    dom = max(prof.items(), key=lambda kv: kv[1][1])
    tot = sum(x[1] for x in prof.values())
    share = {k: round(v[1] / tot, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])[:8]}
    dom_ms = dom[1][1] / dom[1][0]
    calls_per_step = dom[1][0] / prof_steps
    T, H2 = cfg.T, cfg.H * 2
    flops_tab = {   # algorithmic FLOPs per launch of the calls that can dominate
        "indnet_lstm_fwd": 2 * T * cfg.u * 4 * cfg.u * 2, "indnet_lstm_bwd": 2 * T * cfg.u * 4 * cfg.u * 2,
        "indnet_lstm_fwd_tc": 2 * T * (64 + cfg.u) * 4 * cfg.u * 2, "indnet_lstm_bwd_tc": 2 * T * cfg.u * 4 * cfg.u * 2,
        "tc_gemm:attn_dense": 2 * T * cfg.H * cfg.H, "tc_gemm:attn_dW": 2 * T * cfg.H * cfg.H,
        "tc_gemm:attn_dh": 2 * T * cfg.H * cfg.H, "tc_gemm:lstm_dW": 2 * T * (cfg.D + cfg.u) * 4 * cfg.u,
        "tc_gemm:ntn_stage1": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S, "tc_gemm:ntn_dM": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
        "tc_gemm:ntn_dcv": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
        "sgemm:attn_dense": 2 * T * cfg.H * cfg.H, "sgemm:attn_dW": 2 * T * cfg.H * cfg.H, "sgemm:attn_dh": 2 * T * cfg.H * cfg.H,
        "sgemm:inproj": 2 * T * cfg.D * 4 * cfg.u, "sgemm:lstm_dWx": 2 * T * cfg.D * 4 * cfg.u,
        "sgemm:lstm_dWh": 2 * T * cfg.u * 4 * cfg.u, "sgemm:ntn_stage1": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
        "sgemm:ntn_dM": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S, "sgemm:ntn_dcv": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
    }
    bytes_tab = {   # algorithmic HBM bytes per launch: streams only, resident weights ignored
        "indnet_lstm_fwd_tc": T * (2 * 64 * 2 + 2 * cfg.u * 14), "indnet_lstm_bwd_tc": T * (2 * cfg.u * 20 + 2 * 64),
        "tc_gemm:lstm_dW": T * (4 * cfg.u + 64 + cfg.u) * 2,
        "indnet_attn_pool_bwd_bf16": T * H2 * 3, "indnet_attn_pool_fwd_bf16": T * H2 * 2,
        "tc_gemm:attn_dense": T * H2 * 2, "tc_gemm:attn_dh": T * H2 * 2, "tc_gemm:attn_dW": T * H2 * 2,
        "indnet_attn_pool_bwd": cfg.X * (cfg.L * cfg.H * 4 * 4 + cfg.H * 8), "indnet_attn_pool_fwd": cfg.X * 42304,
        "indnet_embed_bwd": T * 48, "indnet_adam_step": store.numel * 30,
    }
    peak_tf = peaks.get("bf16_tflops_sustained")
    peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    if peak_tf is None:
        peak_tf, peak_src = 1400.0, "fallback (B200_PROFILING.md sustained)"
    pk_bw = peaks.get("hbm_gbs", 6650.0)
    t_flop = flops_tab[dom[0]] / (peak_tf * 1e12) if dom[0] in flops_tab else 0.0
    t_byte = bytes_tab[dom[0]] / (pk_bw * 1e9) if dom[0] in bytes_tab else 0.0
    base = {"kernel": dom[0], "avg_launch_ms": dom_ms, "launches_per_step": calls_per_step, "traffic": None}
    if t_flop == 0.0 and t_byte == 0.0:
        roof = dict(base, bound="hbm", achieved=None, peak=pk_bw, unit="GB/s", frac=None,
                    note="no algorithmic byte / flop count tabulated for this call")
    elif t_flop >= t_byte:
        ach = flops_tab[dom[0]] / (dom_ms * 1e-3) / 1e12
        roof = dict(base, bound="tensor", achieved=ach, peak=peak_tf, unit="TFLOP/s", frac=ach / peak_tf, peak_source=peak_src)
    else:
        ach = bytes_tab[dom[0]] / (dom_ms * 1e-3) / 1e9
        roof = dict(base, bound="hbm", achieved=ach, peak=pk_bw, unit="GB/s", frac=ach / pk_bw,
                    peak_source="measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)",
                    algorithmic_bytes_per_launch=bytes_tab[dom[0]])
    if dom[0] in ("indnet_lstm_fwd_tc", "indnet_lstm_bwd_tc"):
        mhz = clocks.get("sm_mhz") or 1965
        ctas = 2 * ((cfg.X + 31) // 32)
        roof["latency_cycles_per_lstm_step"] = dom_ms * 1e-3 * mhz * 1e6 / cfg.L
        roof["note"] = ("persistent recurrence, %d CTAs on 148 SMs: the launch lasts L = %d dependent steps whatever the "
                        "batch, so below a full chip its duration is a latency (%.0f SM cycles per step), not a byte "
                        "count; byte floor %.1f us, flop floor %.1f us" %
                        (ctas, cfg.L, roof["latency_cycles_per_lstm_step"], t_byte * 1e6, t_flop * 1e6))
    try:        # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture, same workload
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tr = json.load(f)
        key = {"indnet_lstm_fwd_tc": "tc_lstm_fwd_kernel", "indnet_lstm_bwd_tc": "tc_lstm_bwd_kernel"}.get(dom[0], dom[0])
        if args.config != 1:
            key = "%s@config %d" % (key, args.config)
        if key in tr and tr[key].get("workload") == "config %d" % args.config and args.mode == "bf16":
            roof["traffic"] = tr[key]["dram_bytes_per_launch"]
            roof["traffic_source"] = tr[key].get("source", "profiles/ (dram__bytes_read.sum + dram__bytes_write.sum)")
    except Exception:
        pass
    return roof, share


def run_inference(args, w, world, rank, local, peaks):
    """config 5: encoder + induction forward over 512 episodes per GPU (replicas, no collective), in chunks."""
    import torch
    import torch.distributed as dist
    from induction_b200._lib import LIB
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes
    from induction_b200.util.data_loader import JSONFileDataLoader
    chunk = w["chunk"]
    cfg = EpisodeConfig(N=w["N"], K=w["K"], Q=Q, B=chunk, L=w["L"], E=E, u=w["u"], S=S, keep_prob=1.0, mode=args.mode,
                        inference=True)
    tables = make_tables_for(w)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=0)
    loader = JSONFileDataLoader.from_tables(tables)
    smp = loader._device()
    smp.seed(rank)               # replicas over independent episode shards
    inputs = dict(word=pipe.word, pos1=pipe.pos1, pos2=pipe.pos2, mask=pipe.mask, label=pipe.label)
    nchunks = w["B"] // chunk
    stream = torch.cuda.Stream(priority=-1)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")

    def body():
        smp.enqueue(cfg.B, cfg.N, cfg.K, cfg.Q, out=inputs)
        pipe.encode(train=False)
        pipe.induction()

    with torch.cuda.stream(stream):
        body()
        stream.synchronize()
        l0 = LIB.launch_count()
        body()
        launches = LIB.launch_count() - l0
        stream.synchronize()
        graph = None
        if not args.no_graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                body()

    def run_step():
        with torch.cuda.stream(stream):
            for _ in range(nchunks):
                graph.replay() if graph is not None else body()

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        run_step()
    stream.synchronize()
    t_wait = time.time()
    while not sampler.ready and time.time() - t_wait < 3.0:
        time.sleep(0.01)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.armed = True
    times, blocks = timed_blocks(torch, stream, flush, run_step, args.steps, world, dist)
    sampler.armed = False
    sampler.stop_flag = True
    smp.check()
    # per-call shares (eager) for the roofline of the dominant call
    with torch.cuda.stream(stream):
        pipe.profile = []
        for _ in range(2):
            pipe.encode(train=False)
            pipe.induction()
        prof = pipe.profile_summary()
        pipe.profile = None
    # e2e: host episodes (pinned) H2D + encode + induction + D2H of the class vectors, per chunk
    batches = host_episodes(w, tables, cfg.B, 2, seed=1 + rank)
    # host episodes: pinned, in the device input buffer's layout (EpisodePipeline.pack_inputs) = one H2D copy per step
    host_pool = [(pipe.pack_inputs(inp, lab), None) for inp, lab in batches]
    out_host = torch.empty(cfg.B * cfg.N, cfg.H, dtype=torch.float32).pin_memory()
    h2d = host_pool[0][0].numel() * 4 * nchunks
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_steps = 3
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for i in range(e2e_steps * nchunks):
            pipe.set_inputs(*host_pool[i % 2])
            pipe.encode(train=False)
            pipe.induction()
            out_host.copy_(pipe.cvec, non_blocking=True)
        stream.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1000.0
    mean_ms, med_ms = summarize(times)
    t = torch.tensor([mean_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    mean_ms, e2e_ms = float(t[0]), float(t[1])
    sent_per_step = world * w["B"] * cfg.n_s
    if rank == 0:
        value = sent_per_step / (mean_ms / 1000.0)
        flops = fwd_flops_per_episode(cfg.N, cfg.K, cfg.L, cfg.u, relation=False) * w["B"] * world
        peak_tf = peaks.get("bf16_tflops_sustained", 1400.0)
        dom = max(prof.items(), key=lambda kv: kv[1][1])
        tot = sum(x[1] for x in prof.values())
        ach = flops / world / (mean_ms * 1e-3) / 1e12
        line = json.dumps({
            "metric": "encoder sentences/sec", "value": value, "unit": "sentences/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": mean_ms, "ms_per_step_median": med_ms, "timed_blocks": blocks,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.mode == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "survey_config": 5, "episodes_per_gpu": w["B"], "chunk_episodes": chunk,
                       "sentences_per_gpu": w["B"] * cfg.n_s, "mode": args.mode, "parallelism": "replicas (no collective), dp%d" % world,
                       "cuda_graph": graph is not None,
                       "l2": "256 MB buffer written between timed iterations; the working set of one chunk exceeds L2"},
            "episodes_per_sec": world * w["B"] / (mean_ms / 1000.0),
            "e2e": {"value": world * w["B"] * cfg.n_s * e2e_steps / (e2e_ms / 1000.0), "unit": "sentences/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": cfg.B * cfg.N * cfg.H * 4 * nchunks,
                    "timer": "host perf_counter around pinned H2D of each chunk's episodes + encode + induction + D2H of class vectors"},
            "gpu_launches": int(launches * nchunks * args.steps * blocks), "launches_per_step": int(launches * nchunks),
            "clocks": sampler.summary(),
            "roofline": {"bound": "tensor", "kernel": "whole step (encoder + induction forward)", "achieved": ach,
                         "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf, "traffic": None,
                         "dominant_call": dom[0], "dominant_share": dom[1][1] / tot},
            "step_share": {k: round(v[1] / tot, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])[:6]},
            "cpu_baseline": None,
        })
    teardown(None, world, line if rank == 0 else None)


def teardown(trainer, world, line):
    """Ordered: drain, drop the graphs, destroy the process group - and only then print the JSON line, so that it is
    the LAST line on stdout even when NCCL_DEBUG=INFO makes NCCL log its shutdown there.  A watchdog turns a
    communicator that will not shut down into a plain exit (the line is printed first) instead of a hung job."""
    import torch

    def emit():
        if line is not None:
            print(line)
            sys.stdout.flush()

    if trainer is not None:
        trainer.close()
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()

        def bail():
            emit()
            os._exit(0)

        t = threading.Timer(30.0, bail)
        t.daemon = True
        t.start()
        dist.destroy_process_group()
        t.cancel()
    emit()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=sorted(WORKLOADS),
                    help="SURVEY section 8 config number (1 = 5-way 5-shot, the metric's configuration)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--stream", default="global", choices=["global", "per-rank"],
                    help="global: one episode stream for the job, every rank walks all of it (bit-identical global batch for "
                         "any N); per-rank: independent streams, the walk covers only the rank's own episodes")
    ap.add_argument("--mode", default="bf16", choices=["fp32", "bf16"],
                    help="bf16 = tcgen05 tensor-core mode (default); fp32 = FFMA parity mode")
    ap.add_argument("--keep-prob", type=float, default=0.9, help="LSTM output dropout as the reference trains (base.py:26,149)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-filled", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    w = WORKLOADS[args.config]
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_b200(args, w)


if __name__ == "__main__":
    main()
