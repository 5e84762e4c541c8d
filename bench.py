#!/usr/bin/env python
"""bench.py - train-step episodes/sec of the Induction-Network episode pipeline on B200.

  python bench.py --gpus N --steps K --warmup W          # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K --warmup W   # CPU arm (oracle port; TF 1.x absent)

A "step" is one pass of the hot path over one batch of B synthetic episodes per GPU: (device
sampler + gather when available) -> forward -> loss -> backward -> gradient all-reduce (N > 1) -> TF
Adam.  Weak scaling: every rank runs B episodes.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs[i] -> episode shape (Q = 5 everywhere, train.py:33; E = 50; S = 100)
WORKLOADS = [
    dict(name="cfg0: 5-way 5-shot, 4 episodes, L=40, u=128", N=5, K=5, B=4, L=40, u=128),
    dict(name="cfg1: 5-way 1-shot, 64 episodes, L=40, u=128", N=5, K=1, B=64, L=40, u=128),
    dict(name="cfg2: 10-way 5-shot, 128 episodes, L=40, u=128", N=10, K=5, B=128, L=40, u=128),
    dict(name="cfg3: 10-way 10-shot, 1024 episodes, L=40, u=128", N=10, K=10, B=1024, L=40, u=128),
]
E, S, Q, VOCAB, RELS, PER_REL = 50, 100, 5, 400000, 64, 700


def fwd_flops_per_episode(N, K, L, u):
    """SURVEY.md section 8 FLOP formula (multiply-add = 2), forward, one episode."""
    D, H, n_s, n_q = E + 10, 2 * u, N * (K + Q), N * Q
    return (2 * n_s * L * D * 4 * u * 2 + 2 * n_s * L * u * 4 * u * 2 + 2 * n_s * L * H * H + 4 * n_s * L * H
            + 2 * N * K * H * H + 3 * 4 * N * K * H + 2 * N * H * H * S + 2 * N * n_q * H * S + 2 * n_q * N * N * S)


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


def make_inputs(w, pool, seed=0):
    """Synthetic FewRel-shaped tables and ``pool`` pre-sampled episode batches (host)."""
    import numpy as np
    from induction_b200.util.synthetic import make_tables
    tables = make_tables(RELS, PER_REL, w["L"], VOCAB, E, seed=seed)
    rs = np.random.RandomState(seed + 1)
    n_s, n_q = w["N"] * (w["K"] + Q), w["N"] * Q
    batches = []
    for _ in range(pool):
        rows = np.zeros((w["B"], n_s), dtype=np.int64)
        label = np.zeros((w["B"], n_q), dtype=np.int32)
        for b in range(w["B"]):
            classes = rs.choice(RELS, w["N"], replace=False)
            sup, qry, lab = [], [], []
            for i, c in enumerate(classes):
                idx = c * PER_REL + rs.choice(PER_REL, w["K"] + Q, replace=False)
                sup += list(idx[: w["K"]]); qry += list(idx[w["K"]:]); lab += [i] * Q
            perm = rs.permutation(n_q)
            rows[b] = np.array(sup + [qry[p] for p in perm])
            label[b] = np.array([lab[p] for p in perm])
        batches.append(({k: np.ascontiguousarray(tables[k][rows]) for k in ("word", "pos1", "pos2", "mask")}, label))
    return tables, batches


# ------------------------------------------------------------------------------------------------
def cpu_episodes_per_sec(w, tables, batches, steps, warmup, episodes_per_step, budget_s=None):
    """The CPU arm: oracle port (torch-CPU fp32 autograd + TF-form Adam) on a bounded sample of the
    same workload, all host threads.  Returns (episodes/s, cores, steps actually timed)."""
    import torch
    from oracle import induction_oracle as orc
    from induction_b200.engine import init_tf_params, tf_param_shapes
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    params = init_tf_params(tf_param_shapes(w["N"], w["L"], E, w["u"], S), seed=0)
    m = {k: torch.zeros_like(v) for k, v in params.items()}
    v = {k: torch.zeros_like(x) for k, x in params.items()}
    wt = torch.from_numpy(tables["word_vec_mat"])
    Bc = episodes_per_step

    def one(step):
        inp, lab = batches[step % len(batches)]
        ti = {k: torch.from_numpy(x[:Bc]) for k, x in inp.items()}
        _, g = orc.loss_and_grads(params, wt, ti["word"], ti["pos1"], ti["pos2"], ti["mask"],
                                  torch.from_numpy(lab[:Bc]).long(), w["N"], w["K"], Q)
        orc.tf_adam_step(params, g, m, v, global_step=step)

    for s in range(warmup):
        one(s)
    t0 = time.perf_counter()
    done = 0
    for s in range(steps):
        one(warmup + s)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return Bc * done / dt, cores, done


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    tables, batches = make_inputs(w, pool=2)
    Bc = min(w["B"], 4)
    eps, cores, done = cpu_episodes_per_sec(w, tables, batches, args.steps, min(args.warmup, 2), Bc)
    sample = "%d episodes/step (of B=%d) x %d steps, torch-CPU fp32 oracle port, fwd+autograd bwd+TF Adam" % (Bc, w["B"], done)
    print(json.dumps({
        "impl": "reference", "metric": "train-step episodes/sec", "value": eps, "unit": "episodes/s",
        "n_gpus": args.gpus, "steps": done, "warmup": min(args.warmup, 2), "ms_per_step": 1000.0 * Bc / eps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "note": "reference TF 1.x graph unavailable on this box; CPU oracle port timed"},
        "cpu_baseline": {"value": eps, "unit": "episodes/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": eps, "unit": "episodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------------------------
def run_b200(args, w):
    import numpy as np
    import torch
    import torch.distributed as dist
    import induction_b200  # noqa: F401
    from induction_b200._lib import LIB
    from induction_b200.engine import EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its version banner on stdout at NCCL_DEBUG=VERSION/INFO; stdout carries the JSON line only
        os.environ["NCCL_DEBUG"] = os.environ.get("INDNET_NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass

    cfg = EpisodeConfig(N=w["N"], K=w["K"], Q=Q, B=w["B"], L=w["L"], E=E, u=w["u"], S=S,
                        keep_prob=args.keep_prob, mode=args.mode)
    POOL = 4
    tables, batches = make_inputs(w, pool=POOL, seed=rank)
    store = ParamStore(tf_param_shapes(cfg.N, cfg.L, cfg.E, cfg.u, cfg.S))
    store.load_state_dict(init_tf_params(store.tf_shapes, seed=0))
    pipe = EpisodePipeline(cfg, store, torch.from_numpy(tables["word_vec_mat"]), seed=rank)
    # device episode sampler: per-rank streams seeded random.seed(rank) / np.random.seed(rank)
    from induction_b200.util.data_loader import JSONFileDataLoader
    loader = JSONFileDataLoader.from_tables(tables)
    smp = loader._device()
    smp.seed(rank)
    pipe_inputs = dict(word=pipe.word, pos1=pipe.pos1, pos2=pipe.pos2, mask=pipe.mask, label=pipe.label)
    # the sampler of step t+1 runs at the lower priority: its table kernel fills the GPU and would otherwise take
    # SMs from the training kernels of step t, which are the critical path
    side = torch.cuda.Stream(priority=0)
    # resident input pool in HBM + pinned host copies for the e2e leg
    dev_pool, host_pool = [], []
    for inp, lab in batches:
        dev_pool.append(({k: torch.from_numpy(v).cuda() for k, v in inp.items()}, torch.from_numpy(lab).cuda()))
        host_pool.append(({k: torch.from_numpy(v).pin_memory() for k, v in inp.items()}, torch.from_numpy(lab).pin_memory()))
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")   # 256 MB > 126 MB L2

    # Gradient all-reduce in two pieces: M (first slot of the flat buffer, 95 % of the bytes) is final right
    # after the relation-module backward, so its all-reduce is issued then on a side stream and hides under
    # the attention / BPTT backward; the remaining slots go in one trailing all-reduce.
    comm = torch.cuda.Stream(priority=-1)
    m_end = store.offsets["w"]          # slots are laid out M, w, b, ... (engine.ParamStore)

    def on_grad_ready(name):
        if world > 1 and name == "M":
            comm.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(comm):
                dist.all_reduce(store.grads[:m_end])

    pipe.grad_ready_hook = on_grad_ready

    def grad_sync():
        if world > 1:
            dist.all_reduce(store.grads[m_end:])
            torch.cuda.current_stream().wait_stream(comm)

    def train_body():
        pipe.forward(train=True)
        pipe.backward()
        grad_sync()
        pipe.adam(grad_scale=1.0 / world)

    def step_body():
        # picks drawn during the previous step -> this step's inputs; then the serial stream walk for
        # the NEXT step runs on a side stream underneath this step's forward/backward/update
        smp.enqueue_gather(cfg.B, cfg.N, cfg.K, cfg.Q, out=pipe_inputs)
        cur = torch.cuda.current_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            smp.enqueue_walk(cfg.B, cfg.N, cfg.K, cfg.Q)
        train_body()
        cur.wait_stream(side)

    stream = torch.cuda.Stream(priority=-1)
    graph = None
    with torch.cuda.stream(stream):
        pipe.set_inputs(*dev_pool[0])
        smp.enqueue_walk(cfg.B, cfg.N, cfg.K, cfg.Q)      # prologue: picks for the first step
        for _ in range(2):          # allocates backward buffers, sets func attributes, warms NCCL
            step_body()
        stream.synchronize()
        l0 = LIB.launch_count()
        step_body()
        launches_per_step = LIB.launch_count() - l0
        stream.synchronize()
        graph_train = None
        if not args.no_graph:
            try:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=stream):
                    step_body()
                graph_train = torch.cuda.CUDAGraph()       # host-fed variant for the e2e leg
                with torch.cuda.graph(graph_train, stream=stream):
                    train_body()
            except Exception as e:   # e.g. NCCL capture refused: fall back to eager launches
                if rank == 0:
                    print("cuda graph capture failed, eager launches: %r" % (e,), file=sys.stderr)
                graph = graph_train = None
                torch.cuda.synchronize()

        def run_step(i):
            if graph is not None:
                graph.replay()
            else:
                step_body()

        sampler = ClockSampler(local)
        sampler.start()
        for i in range(args.warmup):
            run_step(i)
        stream.synchronize()
        t_wait = time.time()
        while not sampler.ready and time.time() - t_wait < 3.0:
            time.sleep(0.01)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        sampler.armed = True
        evs = []
        for i in range(args.steps):
            flush.fill_(float(i))                   # L2 flush between timed iterations (outside the event pair)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            run_step(args.warmup + i)
            e1.record(stream)
            evs.append((e0, e1))
        stream.synchronize()
        torch.cuda.synchronize()
        time.sleep(0.03)                      # let the last in-flight sample of the region land
        sampler.armed = False
        sampler.stop_flag = True
        if world > 1:
            dist.barrier()
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        loss_after = pipe.loss()

        # ---- same K steps without the sampler in the loop (episodes already resident in HBM)
        evs2 = []
        for i in range(args.steps):
            flush.fill_(float(i))
            pipe.set_inputs(*dev_pool[i % POOL])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            if graph_train is not None:
                graph_train.replay()
            else:
                train_body()
            e1.record(stream)
            evs2.append((e0, e1))
        stream.synchronize()
        resident_ms = sum(a.elapsed_time(b) for a, b in evs2)

        # ---- roofline leg: per-call CUDA-event timing of the same step, eager launches
        pipe.profile = []
        smp_ev = []
        for i in range(min(args.steps, 5)):
            flush.fill_(1.0)
            smp.enqueue_gather(cfg.B, cfg.N, cfg.K, cfg.Q, out=pipe_inputs)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            smp.enqueue_walk(cfg.B, cfg.N, cfg.K, cfg.Q)      # serial here: timed on its own
            e1.record()
            smp_ev.append((e0, e1))
            train_body()
        prof = pipe.profile_summary()
        pipe.profile = None
        prof_steps = min(args.steps, 5)
        sampler_walk_ms = sum(a.elapsed_time(b) for a, b in smp_ev) / len(smp_ev)
        smp.check()

        # ---- e2e leg: the user-facing call with HOST buffers (pinned), H2D + D2H inside the timed region
        e2e_steps = max(3, min(args.steps, 20))
        h2d = sum(v.numel() * 4 for v in host_pool[0][0].values()) + host_pool[0][1].numel() * 4
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            pipe.set_inputs(*host_pool[i % POOL])         # H2D from pinned memory
            if graph_train is not None:
                graph_train.replay()
            else:
                train_body()
            res = pipe.out2.cpu()                        # D2H {loss, accuracy} (synchronises)
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1000.0

    t = torch.tensor([total_ms, e2e_ms, resident_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, resident_ms = float(t[0]), float(t[1]), float(t[2])
    ms_per_step = total_ms / args.steps
    value = world * cfg.B * args.steps / (total_ms / 1000.0)
    e2e_value = world * cfg.B * e2e_steps / (e2e_ms / 1000.0)

    if rank == 0:
        # dominant call of the step and its roofline
        dom = max(prof.items(), key=lambda kv: kv[1][1])
        if os.environ.get("INDNET_SHARES_OUT"):        # full per-call table of the eager profile leg (tuning aid)
            with open(os.environ["INDNET_SHARES_OUT"], "w") as f:
                json.dump({k: {"calls": v[0], "ms_total": v[1], "ms_per_call": v[1] / v[0], "ms_per_step": v[1] / prof_steps}
                           for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])}, f, indent=1)
        share = {k: round(v[1] / sum(x[1] for x in prof.values()), 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])[:8]}
        dom_ms = dom[1][1] / dom[1][0]
        calls_per_step = dom[1][0] / prof_steps
        T = cfg.T
        flops_tab = {   # algorithmic FLOPs per launch of the calls that can dominate
            "indnet_lstm_fwd": 2 * T * cfg.u * 4 * cfg.u * 2,
            "indnet_lstm_bwd": 2 * T * cfg.u * 4 * cfg.u * 2,
            # tcgen05 recurrence: input projection (K = 64 padded) + recurrent part, both directions
            "indnet_lstm_fwd_tc": 2 * T * (64 + cfg.u) * 4 * cfg.u * 2,
            "indnet_lstm_bwd_tc": 2 * T * cfg.u * 4 * cfg.u * 2,
            "tc_gemm:attn_dense": 2 * T * cfg.H * cfg.H, "tc_gemm:attn_dW": 2 * T * cfg.H * cfg.H,
            "tc_gemm:attn_dh": 2 * T * cfg.H * cfg.H, "tc_gemm:lstm_dWx": 2 * T * cfg.D * 4 * cfg.u,
            "tc_gemm:lstm_dWh": 2 * T * cfg.u * 4 * cfg.u, "tc_gemm:lstm_dx": 2 * T * 10 * 4 * cfg.u,
            "tc_gemm:ntn_stage1": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
            "tc_gemm:ntn_dM": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
            "tc_gemm:ntn_dcv": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
            "sgemm:attn_dense": 2 * T * cfg.H * cfg.H, "sgemm:attn_dW": 2 * T * cfg.H * cfg.H,
            "sgemm:attn_dh": 2 * T * cfg.H * cfg.H,
            "sgemm:inproj": 2 * T * cfg.D * 4 * cfg.u, "sgemm:lstm_dWx": 2 * T * cfg.D * 4 * cfg.u,
            "sgemm:lstm_dWh": 2 * T * cfg.u * 4 * cfg.u, "sgemm:lstm_dx": 2 * T * 10 * 4 * cfg.u,
            "sgemm:ntn_stage1": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S, "sgemm:ntn_dM": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
            "sgemm:ntn_dcv": 2 * cfg.B * cfg.N * cfg.H * cfg.H * cfg.S,
        }
        peak_tf = peaks.get("bf16_tflops_sustained")
        peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        if peak_tf is None:
            peak_tf, peak_src = 1400.0, "fallback (B200_PROFILING.md sustained)"
        pk_bw = peaks.get("hbm_gbs", 6650.0)
        H2 = cfg.H * 2
        # algorithmic HBM bytes per launch: streams only, resident weights ignored
        bytes_tab = {
            # fwd: x in (64 bf16 per token, once per direction); out per (dir, token, unit): gates 8 + c 2 + h_prev 2 + h 2
            "indnet_lstm_fwd_tc": T * (2 * 64 * 2 + 2 * cfg.u * 14),
            # bwd: in gates 8 + c_prev 2 + dh 2, out dz 8, per (dir, token, unit)
            "indnet_lstm_bwd_tc": T * (2 * cfg.u * 20 + 2 * 64),     # + the 16-float dx rows per direction
            "tc_gemm:lstm_dx": T * (4 * cfg.u * 2 + 10 * 4), "tc_gemm:lstm_dWx": T * (4 * cfg.u + 64) * 2,
            "tc_gemm:lstm_dWh": T * (4 * cfg.u + cfg.u) * 2,
            "indnet_attn_pool_bwd_bf16": T * H2 * 3, "indnet_attn_pool_fwd_bf16": T * H2 * 2,
            "tc_gemm:attn_dense": T * H2 * 2, "tc_gemm:attn_dh": T * H2 * 2, "tc_gemm:attn_dW": T * H2 * 2,
            "indnet_attn_pool_bwd": cfg.X * (cfg.L * cfg.H * 4 * 4 + cfg.H * 8),   # h, proj in; dh, dpre out
            "indnet_attn_pool_fwd": cfg.X * 42304,
            "indnet_embed_bwd": T * 48, "indnet_adam_step": store.numel * 30,
            "sgemm:ntn_dqe": cfg.B * (cfg.n_q * cfg.N * cfg.S + cfg.N * cfg.S * cfg.H + cfg.n_q * cfg.H) * 4,
            "sgemm:ntn_stage2": cfg.B * (cfg.n_q * cfg.N * cfg.S + cfg.N * cfg.S * cfg.H + cfg.n_q * cfg.H) * 4,
        }
        # the binding roofline of the dominant call is the larger of its two lower-bound times
        t_flop = flops_tab[dom[0]] / (peak_tf * 1e12) if dom[0] in flops_tab else 0.0
        t_byte = bytes_tab[dom[0]] / (pk_bw * 1e9) if dom[0] in bytes_tab else 0.0
        if t_flop == 0.0 and t_byte == 0.0:
            roof = {"bound": "hbm", "kernel": dom[0], "achieved": None, "peak": pk_bw, "unit": "GB/s", "frac": None,
                    "traffic": None, "avg_launch_ms": dom_ms, "launches_per_step": calls_per_step,
                    "note": "no algorithmic byte / flop count tabulated for this call"}
        elif t_flop >= t_byte:
            ach = flops_tab[dom[0]] / (dom_ms * 1e-3) / 1e12
            roof = {"bound": "tensor", "kernel": dom[0], "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                    "frac": ach / peak_tf, "traffic": None, "peak_source": peak_src,
                    "avg_launch_ms": dom_ms, "launches_per_step": calls_per_step,
                    "note": "FP32 parity-mode kernels run on FFMA, reported against the bf16 tensor peak"
                    if args.mode == "fp32" else "bf16 tcgen05 mode"}
        else:
            ach = bytes_tab[dom[0]] / (dom_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": dom[0], "achieved": ach, "peak": pk_bw, "unit": "GB/s",
                    "frac": ach / pk_bw, "traffic": None,
                    "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)",
                    "algorithmic_bytes_per_launch": bytes_tab[dom[0]],
                    "avg_launch_ms": dom_ms, "launches_per_step": calls_per_step,
                    "note": "byte time %.1f us > flop time %.1f us at the measured peaks: HBM is the binding roofline"
                            % (t_byte * 1e6, t_flop * 1e6)}
        # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (profiles/), same workload
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tr = json.load(f)
            key = {"indnet_lstm_fwd_tc": "tc_lstm_fwd_kernel", "indnet_lstm_bwd_tc": "tc_lstm_bwd_kernel"}.get(dom[0])
            if key in tr and tr[key]["workload"] == "cfg%d" % args.config and args.mode == "bf16":
                roof["traffic"] = tr[key]["dram_bytes_per_launch"]
                roof["traffic_source"] = "profiles/r1_tc_lstm_ncu.md (dram__bytes_read.sum + dram__bytes_write.sum)"
        except Exception:
            pass
        roof["step_share"] = share
        # whole-step achieved TFLOP/s (train = 3 x forward)
        step_tflops = value * 3 * fwd_flops_per_episode(cfg.N, cfg.K, cfg.L, cfg.u) / 1e12
        # ---- CPU baseline (oracle port) on a bounded sample, rank 0, N = 1 only
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            Bc = min(cfg.B, 4)
            eps, cores, done = cpu_episodes_per_sec(w, tables, batches, steps=50, warmup=1,
                                                    episodes_per_step=Bc, budget_s=args.cpu_seconds)
            cpu = {"value": eps, "unit": "episodes/s", "cores": cores, "kind": "port",
                   "sample": "%d episodes/step x %d steps of the same workload (fwd + autograd bwd + TF Adam), "
                             "torch-CPU fp32 oracle; reference TF 1.x unavailable on this box" % (Bc, done)}
        print(json.dumps({
            "metric": "train-step episodes/sec", "value": value, "unit": "episodes/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.mode == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "episodes_per_gpu": cfg.B, "global_batch": cfg.B * world,
                       "Q": Q, "E": E, "S": S, "routing_iters": 3, "keep_prob": args.keep_prob, "mode": args.mode,
                       "vocab": VOCAB, "relations": RELS, "instances_per_relation": PER_REL,
                       "parallelism": "episode-sharded dp%d" % world,
                       "cuda_graph": graph is not None,
                       "l2": "256 MB buffer written between timed iterations (outside each step's event pair); "
                             "step working set itself is >1 GB"},
            "e2e": {"value": e2e_value, "unit": "episodes/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                    "steps": e2e_steps, "timer": "host perf_counter around pinned H2D + step + D2H of {loss, acc}"},
            "gpu_launches": int(launches_per_step * args.steps),
            "launches_per_step": int(launches_per_step),
            "clocks": sampler.summary(),
            "roofline": roof,
            "step_tflops": step_tflops,
            "train_only": {"value": world * cfg.B * args.steps / (resident_ms / 1000.0), "unit": "episodes/s",
                           "ms_per_step": resident_ms / args.steps,
                           "note": "fwd+bwd+all-reduce+Adam on episodes already resident in HBM (sampler not in the loop)"},
            "sampler": {"where": "device, bit-exact MT19937 x2 (csrc/sampler.cu), walk for step t+1 overlapped on a "
                                 "side stream inside the step graph", "walk_ms_serial": sampler_walk_ms},
            "cpu_baseline": cpu,
            "loss_after": loss_after,
        }))
    if world > 1:
        # tearing down a communicator that captured CUDA graphs still reference can hang at exit:
        # drop the graphs, drain the device, line the ranks up, then leave without NCCL destructors
        graph = graph_train = None
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=1, help="index into BASELINE.json configs (0-3)")
    ap.add_argument("--mode", default="bf16", choices=["fp32", "bf16"],
                    help="bf16 = tcgen05 tensor-core mode (default); fp32 = FFMA parity mode")
    ap.add_argument("--keep-prob", type=float, default=1.0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
