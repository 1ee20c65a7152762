#!/usr/bin/env python
"""Headline benchmark: RNN-T joint + loss fwd+bwd, lattice cells/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

N>1 is launched by the driver as `python -m torch.distributed.run --nproc-per-node N ... bench.py
--gpus N ...` (one rank per GPU, NCCL). Rank 0 prints ONE JSON line.

Workloads
  aishell_sweep_B32_T500_U100 (default; BASELINE.json configs[2], the shape the metric is quoted on) and the
  other joint shapes: a "step" is one pass of the hot path (model.py:68,70 + train.py:45) over one batch of
  synthetic input: forward joint+loss, backward to enc_state, dec_state and the four joint parameters; at N>1
  the global batch of N*B utterances is sharded by utterance (rnnt_b200 shard_batch; each rank owns B: weak
  scaling) and the step ends with ONE NCCL all-reduce of the joint parameter gradients with the loss sum folded
  in.
  aishell_train_step_B32_T500_U50 (BASELINE.json configs[3]): a "step" is one full training step of the AISHELL
  model (stock 4x320 biLSTM encoder + 1x512 LSTM prediction network + fused joint/loss, DDP gradient all-reduce,
  grad-norm clip, SGD momentum), per-GPU batch B, train.py:24-52,175-177.

Timed regions: `value` = K steps with device-resident inputs, nothing but the step inside the region (no
per-kernel events); `e2e` = K steps from pinned HOST buffers (H2D of every step's inputs and D2H of its loss inside
the region); the per-kernel breakdown behind `roofline` / `path` comes from a separate pass of K steps with CUDA
events around every own kernel (on the launching stream).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[2]: the shape the metric is quoted on (AISHELL joint, V=4232)
    "aishell_sweep_B32_T500_U100": dict(B=32, T=500, U=100, E=320, D=320, H=512, V=4232),
    "aishell_cpu_B8_T200_U30": dict(B=8, T=200, U=30, E=320, D=320, H=512, V=4232),       # configs[0]
    "timit_B32_T300_U60": dict(B=32, T=300, U=60, E=128, D=128, H=128, V=49),            # configs[1]
    "aishell_long_B16_T1500_U200": dict(B=16, T=1500, U=200, E=320, D=320, H=512, V=4232),  # configs[4]
    # configs[3]: full training step; T/U = data.max_input_length / max_target_length (aishell.yaml:11-12)
    "aishell_train_step_B32_T500_U50": dict(B=32, T=500, U=50, E=320, D=320, H=512, V=4232, train_step=True),
}
DEFAULT_WORKLOAD = "aishell_sweep_B32_T500_U100"
METRIC = "rnnt_joint_loss_fwd_bwd_lattice_cells_per_s"
UNIT = "cells/s"
AISHELL_MODEL = dict(enc=dict(type="lstm", hidden_size=320, output_size=320, n_layers=4, bidirectional=True),
                     dec=dict(type="lstm", hidden_size=512, output_size=320, n_layers=1),
                     joint=dict(input_size=640, inner_size=512), vocab_size=4232, share_weight=False,
                     feature_dim=160, dropout=0.3)   # config/aishell.yaml:21-39


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p["bf16_tflops_sustained"], "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def make_config(args, wl, world):
    """The `config` object of the JSON line: identical for both arms (--impl ours / reference)."""
    B, T, U = wl["B"], wl["T"], wl["U"]
    cfg = {"workload": args.workload, **{k: v for k, v in wl.items() if k != "train_step"}, "per_gpu_batch": B,
           "global_batch": B * world, "lengths": "ragged" if args.ragged else "full",
           "cells_per_gpu": B * T * (U + 1), "parallelism": f"dp{world} by utterance"}
    if wl.get("train_step"):
        cfg["model"] = "AISHELL transducer: 4x320 biLSTM encoder, 1x512 LSTM decoder, joint 640->512->4232"
        cfg["l2"] = "working set per step (activations of 32 x 500 frames, h stash 0.84 GB) exceeds the 126 MB L2"
    else:
        gb = B * T * (U + 1) * 2e-9
        cfg["l2"] = (f"working set per step (h stash {gb * wl['H']:.2f} GB, exponentials / logit gradients "
                     f"{gb * wl['V']:.1f} GB) exceeds the 126 MB L2")
    return cfg


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []      # (receive time, csv line)
        self.windows = []    # [t0, t1] of the timed regions; only samples received inside them count

    def window_begin(self):
        self.windows.append([time.monotonic(), None])

    def window_end(self):
        if self.windows and self.windows[-1][1] is None:
            self.windows[-1][1] = time.monotonic()

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.monotonic(), line.strip()))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, pw, mx, reasons = [], [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for tr, ln in self.lines:
            # a sample is the state ~one polling period before it is received
            if not any(t0 <= tr and (t1 is None or tr <= t1 + 0.05) for t0, t1 in self.windows):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
                pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        pw.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_median": pw[len(pw) // 2] if pw else None}


# --------------------------------------------------------------------------------------- reference arm
def run_reference(args, wl, rank, world):
    """--impl reference: the reference's own CPU implementation of the path, timed on the box's host cores.
    The reference's loss (warprnnt_pytorch CPU) cannot be installed offline, so this is the oracle port
    (oracle/cpu_path.py: torch-CPU joint exactly as model.py:18-37 + the C float64 loss with OpenMP; for the
    training-step workload also stock torch LSTMs on the CPU): kind "port". Rank 0 alone runs. Each step is a
    bounded sample of the workload (whole utterances of the configured shape, or -- when one utterance would
    not let the K+W steps end within a few minutes -- its first frames), chosen from one untimed probe pass;
    cells/s is a rate, the sample is stated in `cpu_baseline.sample`."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)   # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every core
    from oracle import c_oracle, cpu_path
    c_oracle.build()
    budget_s = float(os.environ.get("BENCH_REF_BUDGET_S", "200"))
    make = cpu_path.make_train_step_sample if wl.get("train_step") else cpu_path.make_joint_sample
    probe = make(wl, wl["T"], cores)
    t0 = time.perf_counter()
    probe()
    t_probe = time.perf_counter() - t0
    nstep = args.steps + args.warmup
    T_s = wl["T"]
    if t_probe * nstep > budget_s:
        T_s = max(8, int(wl["T"] * budget_s / (t_probe * nstep)))
    fn = probe if T_s == wl["T"] else make(wl, T_s, cores)
    cells = T_s * (wl["U"] + 1)
    for _ in range(args.warmup):
        fn()
    secs = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        fn()
        secs.append(time.perf_counter() - t0)
    secs.sort()
    med = secs[len(secs) // 2]
    value = cells / med
    sample = (f"1 of {wl['B']} utterances per step (B=1, first {T_s} of T={wl['T']} frames, U={wl['U']}, "
              f"full label length); median of {args.steps} timed steps after {args.warmup} warm-up; "
              f"{cores} threads (torch + OpenMP)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * med, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 joint / f64 lattice", "data": "synthetic",
        "config": make_config(args, wl, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "ms_per_step_min_max": [1e3 * secs[0], 1e3 * secs[-1]]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- our arm
class Ctx:
    pass


def setup(args):
    import torch
    import torch.distributed as dist
    c = Ctx()
    c.rank = int(os.environ.get("RANK", "0"))
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    c.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(c.local_rank)
    c.dev = torch.device("cuda", c.local_rank)
    # Anything a library prints on stdout (NCCL prints its version there on first use) would
    # break the one-JSON-line contract: route fd 1 to stderr until the result line is printed.
    sys.stdout.flush()
    c.saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if c.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=c.dev)
    return c


def sync_all(c):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if c.world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def timed(c, fn, steps):
    """K steps bracketed by barrier + synchronize on both sides, CUDA events on the current stream.
    Returns (max-over-ranks ms per step, [ms per step of every rank], last result)."""
    import torch
    import torch.distributed as dist
    sync_all(c)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    out = None
    for _ in range(steps):
        out = fn()
    e1.record()
    sync_all(c)
    ms = e0.elapsed_time(e1) / steps
    per_rank = [ms]
    if c.world > 1:
        t = torch.tensor([ms], device=c.dev)
        allr = [torch.zeros_like(t) for _ in range(c.world)]
        dist.all_gather(allr, t)
        per_rank = [float(x) for x in allr]
    return max(per_rank), per_rank, out


def emit(c, line):
    import torch.distributed as dist
    sys.stdout.flush()
    os.dup2(c.saved_stdout, 1)
    if c.rank == 0:
        print(json.dumps(line), flush=True)
    if c.world > 1:
        dist.destroy_process_group()


def roofline_objects(R, wl, cells, ks, kern_ms, ms_step, steps, workload):
    """`roofline` (own tensor-core kernel with the largest share of the step), `roofline_joint_fwd`, `path`."""
    H, V = wl["H"], wl["V"]
    peaks = load_peaks()
    peak, burst = peaks["bf16_tflops_sustained"], peaks["bf16_tflops"]   # kernels are timed inside a long step
    flops = 2.0 * H * V * cells                                          # algorithmic: 2*H*V per lattice cell
    traffic_tab, traffic_src = {}, None
    for name in ("r02_ncu_traffic.json", "r01_ncu_traffic.json"):
        try:   # dram bytes per launch from a committed ncu capture of the same workload (not measured in this run)
            with open(os.path.join(ROOT, "profiles", name)) as f:
                traffic_tab = json.load(f).get(workload, {})
            traffic_src = "profiles/" + name
            if traffic_tab:
                break
        except Exception:
            continue
    from rnnt_transducer_b200 import _lib
    bn = _lib.lib().rnnt_b200_vocab_tile(V)                              # UMMA N of the vocabulary tiles
    own = {"joint_fwd": (f"joint_kernel<{bn},0,0> (tcgen05 vocab GEMM + online LSE)", "joint_kernel_fwd_dram_bytes", 1),
           "joint_fwd_store": (f"joint_kernel<{bn},2,0> (tcgen05 vocab GEMM + LSE; stores bf16 exponentials for the backward)",
                               "joint_kernel_fwd_store_dram_bytes", 1),
           "joint_bwd_ds": (f"joint_kernel<{bn},1,0> (tcgen05 vocab GEMM recompute + dS epilogue)",
                            "joint_kernel_bwd_dram_bytes_per_chunk", 1),
           "dh_gemm": ("dh_gemm_kernel (tcgen05 dS.W2, tanh-derivative epilogue)", "dh_gemm_dram_bytes_per_chunk", 1),
           "dw2_gemm": ("dw2_wide_kernel / dw2_gemm_kernel (tcgen05 dS^T.[h|1], MN-major operands)",
                        "dw2_gemm_dram_bytes_per_chunk", 1),
           # the two consumers of a dS chunk launched side by side on half of the SMs each, timed as one region:
           # 2 x 2*H*V flop per cell
           "dh_dw2_pair": ("dh_gemm_kernel || dw2_wide_kernel (side by side on two streams, half of the CTA pairs each)",
                           "dh_dw2_pair_dram_bytes_per_chunk", 2)}

    def roof(k):
        ms = kern_ms.get(k, 0.0)
        n = max(len(ks.get(k, [])) // steps, 1)                          # launches per step (dS chunks)
        ach = own[k][2] * flops / (ms * 1e-3) / 1e12 if ms > 0 else 0.0  # = flop per launch / ms per launch
        return {"bound": "tensor", "kernel": own[k][0], "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                "frac": ach / peak, "frac_of_burst_peak": ach / burst,
                "peak_source": peaks["source"] + " bf16_tflops_sustained (burst: bf16_tflops)",
                "traffic": traffic_tab.get(own[k][1]), "traffic_source": f"{traffic_src} (committed ncu capture)",
                "kernel_ms": ms / n, "launches_per_step": n, "share_of_step": ms / ms_step}

    stored = kern_ms.get("joint_fwd_store", 0.0) > 0    # the backward ran from the forward's stored exponentials
    if stored:
        own["dh_dw2_pair"] = (own["dh_dw2_pair"][0], "dh_dw2_pair_stored_dram_bytes_per_chunk", 2)
    # the kernel (or side-by-side pair) with the largest time per 2*H*V flop/cell of work it does
    dominant = max(own, key=lambda k: kern_ms.get(k, 0.0) / own[k][2])
    total_alg = 6.0 * H * V * cells
    path = {"backward": ("from the forward's stored exponentials: no second vocabulary GEMM (DESIGN.md 3.7)" if stored
                         else "logit gradient recomputed in a second GEMM pass"),
            "algorithmic_tflops": total_alg / (ms_step * 1e-3) / 1e12,
            "frac_of_sustained_peak": total_alg / (ms_step * 1e-3) / 1e12 / peak,
            "kernel_ms_one_step": {k: round(v, 4) for k, v in kern_ms.items()},
            # every vocabulary-sized GEMM of the step against the same peak (2*H*V flop per cell each;
            # joint_bwd_ds is the recompute: hardware work, not counted in the algorithmic 6*H*V)
            "gemm_frac_of_sustained_peak": {k: round(own.get(k, (0, 0, 1))[2] * flops / (kern_ms[k] * 1e-3) / 1e12 / peak, 4)
                                            for k in ("joint_fwd", "joint_fwd_store", "joint_bwd_ds", "dh_gemm", "lib_gemm_dh",
                                                      "dw2_gemm", "dh_dw2_pair") if kern_ms.get(k, 0.0) > 0}}
    return roof(dominant), roof("joint_fwd_store" if kern_ms.get("joint_fwd_store", 0.0) > 0 else "joint_fwd"), path


def run_joint(args, wl, c):
    import torch
    import torch.distributed as dist
    import rnnt_transducer_b200 as R
    dev, rank, world = c.dev, c.rank, c.world
    B, T, U, E, D, H, V = (wl[k] for k in "BTUEDHV")
    joint = R.JointNet(E + D, H, V).to(dev)
    if world > 1:  # identical parameters on every rank
        for p_ in joint.parameters():
            dist.broadcast(p_.data, 0)
    # The GLOBAL batch (N*B utterances) is drawn identically on every rank; shard_batch deals the utterances to
    # the ranks in snake order of their cell counts (what a sharded loader does once per batch, outside the step).
    gen = torch.Generator().manual_seed(2019)            # the reference's seed (aishell.yaml:46)
    GB = B * world
    if args.ragged:
        from tests.helpers import synth_lengths
        gtl, gul = synth_lengths(GB, T, U, "ragged", gen)
    else:
        gtl = torch.full((GB,), T, dtype=torch.int32)
        gul = torch.full((GB,), U, dtype=torch.int32)
    class S:    # this rank's share of the global batch (rebuilt once if the ranks' speeds differ, see below)
        pass

    def deal(shares):
        mine = torch.tensor(shares[rank])
        gen_r = torch.Generator().manual_seed(2019 + 1000 * (rank + 1))
        S.n = len(mine)
        S.tl, S.ul = gtl[mine].contiguous(), gul[mine].contiguous()
        Tm, Um = int(S.tl.max()), int(S.ul.max())            # a rank pads to its own longest utterance
        S.h_enc = torch.randn(S.n, Tm, E, generator=gen_r).pin_memory()
        S.h_dec = torch.randn(S.n, Um + 1, D, generator=gen_r).pin_memory()
        y = torch.randint(1, V, (S.n, Um), generator=gen_r, dtype=torch.int32)
        for b in range(S.n):
            y[b, int(S.ul[b]):] = 0
        S.h_y, S.h_tl, S.h_ul = y.pin_memory(), S.tl.pin_memory(), S.ul.pin_memory()
        S.cells = int((S.tl.long() * (S.ul.long() + 1)).sum())
        S.cells_btu = int((S.tl.long() * S.ul.long()).sum())
        S.d_enc, S.d_dec = S.h_enc.to(dev), S.h_dec.to(dev)
        S.d_y, S.d_tl, S.d_ul = S.h_y.to(dev), S.h_tl.to(dev), S.h_ul.to(dev)

    deal(R.shard_batch(gtl.tolist(), gul.tolist(), world))
    assert S.n == B
    params = list(joint.parameters())
    # one flat gradient buffer: the parameters' .grad are views of it, the loss sum rides in the last element,
    # so the step ends in ONE all-reduce with no concatenation and no copy-back
    n_flat = sum(p_.numel() for p_ in params)
    flat = R.bind_flat_grads(params, extra=1)
    fused_none = R.JointRNNTLoss(joint, reduction="none")
    ar_events = []

    def step(enc, dec, yy, tll, ull, time_ar=False, allreduce=True):
        """fwd + bwd (+ NCCL all-reduce of joint grads and the loss sum at N>1)."""
        enc.requires_grad_(True)
        dec.requires_grad_(True)
        flat.zero_()
        loss_b = fused_none(enc, dec, yy, tll, ull, host_lens=(S.tl, S.ul))
        loss_sum = loss_b.sum()
        (loss_sum / GB).backward()
        flat[-1] = loss_sum.detach()
        if world > 1 and allreduce:
            if time_ar:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.all_reduce(flat)
                e1.record()
                ar_events.append((e0, e1))
            else:
                dist.all_reduce(flat)
        return flat[-1] / GB

    # ---------------- device-resident (value) ----------------
    def step_resident(**kw):
        return step(S.d_enc.detach(), S.d_dec.detach(), S.d_y, S.d_tl, S.d_ul, **kw)

    # Speed-weighted sharding (N > 1, --balance): the GPUs of a box run at different clocks under their power caps
    # and a synchronous step waits for the slowest, so the global batch is dealt in proportion to each rank's measured
    # speed. Calibration = untimed steps of the real, synchronous step: a rank's compute time is the step minus its
    # wait in the all-reduce; half of the measured difference is corrected (a rank that is given less work idles,
    # regains power headroom and speeds up: shard_batch_weighted). The global batch and the loss / gradient
    # normalisation do not change; utterances per rank are reported in `path`.
    balance = None
    CAL_WARM, CAL_STEPS = 4, 8
    if world > 1 and args.balance:
        try:
            for _ in range(CAL_WARM):
                step_resident()
            torch.cuda.synchronize()
            dist.barrier()
            ar_events.clear()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(CAL_STEPS):
                step_resident(time_ar=True)
            e1.record()
            torch.cuda.synchronize()
            wait = sum(a.elapsed_time(b) for a, b in ar_events) / CAL_STEPS
            ar_events.clear()
            t = torch.tensor([S.cells / (e0.elapsed_time(e1) / CAL_STEPS - wait)], device=dev, dtype=torch.float64)
            rates = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(rates, t)
            rates = [float(x) for x in rates]
            shares = R.shard_batch_weighted(gtl.tolist(), gul.tolist(), rates, min_spread=0.03, damping=0.5)
            balance = {"cells_per_ms_per_rank": [round(x, 1) for x in rates],
                       "utterances_per_rank": [len(s_) for s_ in shares], "damping": 0.5,
                       "calibration": f"{CAL_WARM}+{CAL_STEPS} untimed synchronous steps (step - all-reduce wait)"}
            if len(shares[rank]) != S.n or sorted(shares[rank]) != sorted(R.shard_batch(gtl.tolist(), gul.tolist(), world)[rank]):
                deal(shares)
        except Exception as exc:      # never let the balancing take the measurement down: keep the equal split
            balance = {"error": repr(exc)[:200]}
            deal(R.shard_batch(gtl.tolist(), gul.tolist(), world))

    cells, cells_btu = S.cells, S.cells_btu
    if world > 1:
        tot = torch.tensor([cells, cells_btu], device=dev, dtype=torch.int64)
        dist.all_reduce(tot)
        cells_all, cells_btu_all = int(tot[0]), int(tot[1])
    else:
        cells_all, cells_btu_all = cells, cells_btu
    sampler = ClockSampler(c.local_rank) if rank == 0 else None
    if sampler:
        sampler.start()      # started ahead of the warm-up so that it is polling when the timed region begins
    for _ in range(args.warmup):
        step_resident()
    if sampler:
        sampler.window_begin()
    with R.KernelTimer(events=False) as kt_count:    # counts this package's launches, records no event
        ms_step, ms_ranks, loss = timed(c, step_resident, args.steps)
    if sampler:
        sampler.window_end()
    launches = kt_count.launches // args.steps

    # ---------------- end to end through the public API with HOST buffers (e2e) ----------------
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
    # Host inputs are staged the way a training loop's loader does it: the pinned batch of step
    # i+1 is copied on a side stream while step i computes (every step still copies its own inputs
    # host->device inside the timed region and reads its loss back).
    copy_stream = torch.cuda.Stream()
    main_stream = torch.cuda.current_stream()
    hosts = (S.h_enc, S.h_dec, S.h_y, S.h_tl, S.h_ul)
    # two device-side input slots, allocated once (a fresh allocation per step on the copy stream can end in a
    # cudaMalloc + synchronisation in the middle of the timed region): step i computes from slot i % 2 while the
    # batch of step i + 1 lands in the other one
    slots = [[torch.empty(t.shape, dtype=t.dtype, device=dev) for t in hosts] for _ in range(2)]
    slot_free = [None, None]                 # recorded on the main stream behind the step that read the slot
    e2e = {"i": 0, "staged": None}

    def stage_inputs(k):
        with torch.cuda.stream(copy_stream):
            if slot_free[k] is not None:
                copy_stream.wait_event(slot_free[k])
            for d_, h_ in zip(slots[k], hosts):
                d_.copy_(h_, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return ev

    def step_e2e():
        k = e2e["i"] & 1
        ev = e2e["staged"] if e2e["staged"] is not None else stage_inputs(k)
        main_stream.wait_event(ev)
        e2e["staged"] = stage_inputs(k ^ 1)  # next step's batch, overlapped with this step
        enc_, dec_, y_, tl_, ul_ = slots[k]
        l = step(enc_.detach(), dec_.detach(), y_, tl_, ul_)
        slot_free[k] = torch.cuda.Event()
        slot_free[k].record(main_stream)
        loss_host.copy_(l.reshape(1), non_blocking=True)
        e2e["i"] += 1
        return l

    step_e2e()
    step_e2e()                               # both slots have been through a step
    torch.cuda.synchronize()
    e2e["staged"] = None                     # the timed region starts with no copy in flight or done ahead of it
    if sampler:
        sampler.window_begin()               # second timed region (same work per step)
    ms_e2e, _, _ = timed(c, step_e2e, args.steps)
    if sampler:
        sampler.window_end()
    clocks = sampler.stop() if sampler else None
    h2d = sum(t.numel() * t.element_size() for t in (S.h_enc, S.h_dec, S.h_y, S.h_tl, S.h_ul))

    # ---------------- per-kernel breakdown: a separate pass with CUDA events around every own kernel ----------------
    with R.KernelTimer() as kt:
        ms_break, _, _ = timed(c, lambda: step_resident(time_ar=True), args.steps)
    ks = kt.summary()
    kern_ms = {k: sum(v) / args.steps for k, v in ks.items()}      # ms per step, per C-ABI call site
    roofline, roofline_fwd, path = roofline_objects(R, wl, cells, ks, kern_ms, ms_break, args.steps, args.workload)
    path["breakdown_pass_ms_per_step"] = ms_break
    path["step_ms_per_rank"] = [round(x, 4) for x in ms_ranks]
    if world > 1:
        # time between a rank's last backward kernel and the end of the all-reduce: mostly the wait for the slowest
        # rank (the GPUs sit at different clocks under their power caps); the rank that waits least is the limiter
        ar = sum(a.elapsed_time(b) for a, b in ar_events) / max(len(ar_events), 1)
        t = torch.tensor([ar], device=dev)
        allr = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allr, t)
        path["allreduce_ms"] = round(ar, 4)
        path["allreduce_ms_per_rank"] = [round(float(x), 4) for x in allr]
        path["compute_ms_per_rank (breakdown pass: step - all-reduce wait)"] = [round(ms_break - float(x), 4) for x in allr]
        path["allreduce_bytes"] = 4 * (n_flat + 1)

    cfg = make_config(args, wl, world)
    if balance is not None:
        path["speed_weighted_sharding"] = balance
    line = {
        "metric": METRIC, "value": cells_all / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16 GEMM operands / f32 accumulate / f64 lattice",
        "data": "synthetic", "config": cfg,
        "value_btu": cells_btu_all / (ms_step * 1e-3),
        "e2e": {"value": cells_all / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e},
        "gpu_launches": kt_count.launches,
        "gpu_launches_per_step": launches,
        "roofline": roofline, "roofline_joint_fwd": roofline_fwd, "path": path, "clocks": clocks, "loss": float(loss),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(wl)
    emit(c, line)


def cpu_baseline(wl):
    """The oracle port on the box's host cores: a bounded sample (about 10-30 s) of the same workload."""
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    from oracle import c_oracle, cpu_path
    c_oracle.build()
    make = cpu_path.make_train_step_sample if wl.get("train_step") else cpu_path.make_joint_sample
    fn = make(wl, wl["T"], cores)
    secs = []
    for i in range(4):                      # one warm-up + three timed passes
        t0 = time.perf_counter()
        fn()
        if i:
            secs.append(time.perf_counter() - t0)
    secs.sort()
    cells = wl["T"] * (wl["U"] + 1)
    what = ("stock torch-CPU LSTMs + " if wl.get("train_step") else "")
    return {"value": cells / secs[1], "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"1 of {wl['B']} utterances (B=1,T={wl['T']},U={wl['U']}), median of 3 passes after 1 warm-up; "
                      f"{what}torch-CPU joint as model.py:18-37 + oracle/rnnt_oracle.c (OpenMP), {cores} threads"}


def run_train_step(args, wl, c):
    """BASELINE.json configs[3]: train.py:24-52 with the model of config/aishell.yaml, one process per GPU, DDP."""
    import torch
    import torch.distributed as dist
    import rnnt_transducer_b200 as R
    dev, rank, world = c.dev, c.rank, c.world
    B, T, U, V = wl["B"], wl["T"], wl["U"], wl["V"]
    F = AISHELL_MODEL["feature_dim"]
    torch.manual_seed(2019)
    model = R.Transducer(AISHELL_MODEL).to(dev).train()
    ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[c.local_rank], gradient_as_bucket_view=True) \
        if world > 1 else model
    opt = R.Optimizer(model.parameters(), dict(type="sgd", lr=1e-4, momentum=0.9)).optimizer   # aishell.yaml:59-66
    g = torch.Generator().manual_seed(2019 + rank)
    xl = torch.full((B,), T, dtype=torch.long)       # lengths stay on the host (train.py:30-31 reads them back)
    yl = torch.full((B,), U, dtype=torch.long)
    h_x = torch.randn(B, T, F, generator=g).pin_memory()
    h_y = torch.randint(1, V, (B, U), generator=g).pin_memory()
    d_x, d_y = h_x.to(dev), h_y.to(dev)
    cells = B * T * (U + 1)
    nparam = sum(p_.numel() for p_ in model.parameters())

    def step(x, y, sync_grads=True):
        opt.zero_grad(set_to_none=True)
        if world > 1 and not sync_grads:
            with ddp.no_sync():
                loss = ddp(x, xl, y, yl)
                loss.backward()
        else:
            loss = ddp(x, xl, y, yl)
            loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 200.0, foreach=True)   # train.py:49-50, max_grad_norm 200
        opt.step()
        return loss.detach()

    sampler = ClockSampler(c.local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    for _ in range(args.warmup):
        step(d_x, d_y)
    if sampler:
        sampler.window_begin()
    with R.KernelTimer(events=False) as kt_count:
        ms_step, ms_ranks, loss = timed(c, lambda: step(d_x, d_y), args.steps)
    if sampler:
        sampler.window_end()
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()

    def step_e2e():
        l = step(h_x.to(dev, non_blocking=True), h_y.to(dev, non_blocking=True))
        loss_host.copy_(l.reshape(1), non_blocking=True)
        return l

    step_e2e()
    if sampler:
        sampler.window_begin()
    ms_e2e, _, _ = timed(c, step_e2e, args.steps)
    if sampler:
        sampler.window_end()
    clocks = sampler.stop() if sampler else None
    # what limits scaling: the same step without the gradient all-reduce (DDP no_sync), and the own-kernel breakdown
    ms_nosync = None
    if world > 1:
        for _ in range(2):
            step(d_x, d_y, sync_grads=False)
        ms_nosync, _, _ = timed(c, lambda: step(d_x, d_y, sync_grads=False), args.steps)
    with R.KernelTimer() as kt:
        ms_break, _, _ = timed(c, lambda: step(d_x, d_y), args.steps)
    ks = kt.summary()
    kern_ms = {k: sum(v) / args.steps for k, v in ks.items()}
    roofline, roofline_fwd, path = roofline_objects(R, wl, cells, ks, kern_ms, ms_break, args.steps, args.workload)
    own_ms = sum(kern_ms.values())
    path.update({"breakdown_pass_ms_per_step": ms_break, "step_ms_per_rank": [round(x, 4) for x in ms_ranks],
                 "own_kernels_ms_per_step": round(own_ms, 4),
                 "rest_ms_per_step (stock cuDNN LSTMs, first-layer GEMMs, optimizer, host)": round(ms_break - own_ms, 4),
                 "gradient_allreduce_bytes": 4 * nparam,
                 "host_syncs_per_step": "0 in the step (lengths are host tensors: no .cpu()/.item() read-back; "
                                        "the reference makes 3: train.py:30,31,47)"})
    if ms_nosync is not None:
        path["step_ms_without_gradient_allreduce"] = ms_nosync
        path["allreduce_exposed_ms"] = round(ms_step - ms_nosync, 4)
    cfg = make_config(args, wl, world)
    cfg["cells_per_gpu"] = cells
    line = {
        "metric": METRIC, "value": cells * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 LSTMs (stock) / bf16 joint GEMM operands / f32 accumulate / f64 lattice",
        "data": "synthetic", "config": cfg, "value_btu": B * T * U * world / (ms_step * 1e-3),
        "utterances_per_s": B * world / (ms_step * 1e-3),
        "e2e": {"value": cells * world / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": h_x.numel() * 4 + h_y.numel() * 8, "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e},
        "gpu_launches": kt_count.launches, "gpu_launches_per_step": kt_count.launches // args.steps,
        "roofline": roofline, "roofline_joint_fwd": roofline_fwd, "path": path, "clocks": clocks, "loss": float(loss),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(wl)
    emit(c, line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ragged", action="store_true", help="ragged lengths instead of full (parity-style inputs)")
    ap.add_argument("--balance", action="store_true",
                    help="N > 1: speed-weighted sharding of the global batch (off by default: measured to change "
                         "nothing, profiles/r02_notes.md)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args, wl, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    c = setup(args)
    if wl.get("train_step"):
        run_train_step(args, wl, c)
    else:
        run_joint(args, wl, c)


if __name__ == "__main__":
    main()
