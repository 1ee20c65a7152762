#!/usr/bin/env python
"""Headline benchmark: RNN-T joint + loss fwd+bwd, lattice cells/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

N>1 is launched by the driver as `python -m torch.distributed.run --nproc-per-node N ... bench.py
--gpus N ...` (one rank per GPU, NCCL). A "step" is one pass of the hot path (model.py:68,70 +
train.py:45) over one batch of synthetic input: forward joint+loss, backward to enc_state,
dec_state and the four joint parameters; at N>1 the batch is sharded by utterance (each rank
owns B utterances: weak scaling) and the step ends with the NCCL all-reduce of the joint
parameter gradients and the loss sum. Rank 0 prints ONE JSON line.
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
}
DEFAULT_WORKLOAD = "aishell_sweep_B32_T500_U100"
METRIC = "rnnt_joint_loss_fwd_bwd_lattice_cells_per_s"
UNIT = "cells/s"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p["bf16_tflops_sustained"], "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


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
        sm, mx, reasons = [], None, set()
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
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args, wl, rank, world):
    """--impl reference: the reference's own CPU implementation of the path, timed on the box's
    host cores. The reference's loss (warprnnt_pytorch CPU) cannot be installed offline, so this
    is the oracle port (oracle/cpu_path.py: torch-CPU joint exactly as model.py:18-37 + the C
    float64 loss with OpenMP). Rank 0 alone runs; each step is a bounded sample of the workload."""
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host core (stated in `cores`)
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    from oracle import c_oracle, cpu_path
    c_oracle.build()
    n_utts = 1
    r = cpu_path.time_cpu_path(wl["E"], wl["D"], wl["H"], wl["V"], wl["T"], wl["U"], n_utts=n_utts,
                               steps=args.steps, warmup=min(args.warmup, 1))
    ms = 1e3 * sum(r["seconds"]) / len(r["seconds"])
    sample = f"{n_utts} of {wl['B']} utterances per step (B=1,T={wl['T']},U={wl['U']}), full lengths"
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 joint / f64 lattice", "data": "synthetic",
        "config": {"workload": args.workload, **wl, "note": "CPU; sample per step: " + sample},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ragged", action="store_true", help="ragged lengths instead of full (parity-style inputs)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return

    import torch
    import torch.distributed as dist
    import rnnt_transducer_b200 as R

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # Anything a library prints on stdout (NCCL prints its version there on first use) would
    # break the one-JSON-line contract: route fd 1 to stderr until the result line is printed.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    args.warmup = max(args.warmup, 3)

    B, T, U, E, D, H, V = (wl[k] for k in "BTUEDHV")
    gen = torch.Generator().manual_seed(2019 + rank)     # the reference's seed (aishell.yaml:46)
    joint = R.JointNet(E + D, H, V).to(dev)
    if world > 1:  # identical parameters on every rank
        for p_ in joint.parameters():
            dist.broadcast(p_.data, 0)
    # host (pinned) inputs of one step
    h_enc = torch.randn(B, T, E, generator=gen).pin_memory()
    h_dec = torch.randn(B, U + 1, D, generator=gen).pin_memory()
    if args.ragged:
        from tests.helpers import synth_lengths
        tl, ul = synth_lengths(B, T, U, "ragged", gen)
    else:
        tl = torch.full((B,), T, dtype=torch.int32)
        ul = torch.full((B,), U, dtype=torch.int32)
    y = torch.randint(1, V, (B, U), generator=gen, dtype=torch.int32)
    for b in range(B):
        y[b, int(ul[b]):] = 0
    h_y, h_tl, h_ul = y.pin_memory(), tl.pin_memory(), ul.pin_memory()
    cells = int((tl.long() * (ul.long() + 1)).sum())
    cells_btu = int((tl.long() * ul.long()).sum())
    params = list(joint.parameters())
    flat_n = sum(p_.numel() for p_ in params) + 1

    def step(enc, dec, yy, tll, ull):
        """fwd + bwd (+ NCCL all-reduce of joint grads and the loss sum at N>1)."""
        enc.requires_grad_(True)
        dec.requires_grad_(True)
        for p_ in params:
            p_.grad = None
        loss_b = fused_none(enc, dec, yy, tll, ull, host_lens=(tl, ul))
        loss_sum = loss_b.sum()
        (loss_sum / (B * world)).backward()
        if world > 1:
            flat = torch.cat([p_.grad.reshape(-1) for p_ in params] + [loss_sum.detach().reshape(1)])
            dist.all_reduce(flat)
            off = 0
            for p_ in params:
                p_.grad.copy_(flat[off:off + p_.numel()].view_as(p_.grad))
                off += p_.numel()
            return flat[-1] / (B * world)
        return loss_sum.detach() / B

    fused_none = R.JointRNNTLoss(joint, reduction="none")
    d_enc, d_dec = h_enc.to(dev), h_dec.to(dev)
    d_y, d_tl, d_ul = h_y.to(dev), h_tl.to(dev), h_ul.to(dev)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        sync_all()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms / steps, out

    # ---------------- device-resident (value) ----------------
    def step_resident():
        return step(d_enc.detach(), d_dec.detach(), d_y, d_tl, d_ul)

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()      # started ahead of the warm-up so that it is polling when the timed region begins
    for _ in range(args.warmup):
        step_resident()
    if sampler:
        sampler.window_begin()
    # per-kernel CUDA events (on the launching stream, no synchronisation) ride along in the timed
    # region: the roofline numbers below are averages over exactly these K steps
    with R.KernelTimer() as kt:
        ms_step, loss = timed(step_resident, args.steps)
    if sampler:
        sampler.window_end()
    ks = kt.summary()
    launches = kt.launches // args.steps

    # ---------------- end to end through the public API with HOST buffers (e2e) ----------------
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()

    # Host inputs are staged the way a training loop's loader does it: the pinned batch of step
    # i+1 is copied on a side stream while step i computes (every step still copies its own inputs
    # host->device inside the timed region and reads its loss back).
    copy_stream = torch.cuda.Stream()
    main_stream = torch.cuda.current_stream()
    pending = []

    def stage_inputs():
        with torch.cuda.stream(copy_stream):
            ts = [t.to(dev, non_blocking=True) for t in (h_enc, h_dec, h_y, h_tl, h_ul)]
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        pending.append((ts, ev))

    def step_e2e():
        if not pending:
            stage_inputs()
        ts, ev = pending.pop(0)
        main_stream.wait_event(ev)
        for t in ts:
            t.record_stream(main_stream)
        stage_inputs()                       # next step's batch, overlapped with this step
        l = step(*ts)
        loss_host.copy_(l.reshape(1), non_blocking=True)
        return l

    step_e2e()
    torch.cuda.synchronize()
    pending.clear()                          # the timed region starts with no copy in flight
    if sampler:
        sampler.window_begin()               # second timed region (same work per step)
    ms_e2e, _ = timed(step_e2e, args.steps)
    if sampler:
        sampler.window_end()
    clocks = sampler.stop() if sampler else None
    h2d = sum(t.numel() * t.element_size() for t in (h_enc, h_dec, h_y, h_tl, h_ul))

    # ---------------- roofline of the dominant kernel ----------------
    # The step's own tensor-core kernels all do 2*H*V algorithmic flop per lattice cell (the backward dS
    # kernel's is the recompute); the one with the largest share of the measured step is reported, the
    # forward joint GEMM (the kernel north_star's 60 % target names) beside it.
    kern_ms = {k: sum(v) / args.steps for k, v in ks.items()}      # ms per step, per C-ABI call site
    peaks = load_peaks()
    peak = peaks["bf16_tflops_sustained"]                            # kernels are timed inside a long step
    flops_fwd = 2.0 * H * V * cells                                  # algorithmic: 2*H*V per lattice cell
    try:   # dram bytes per launch from the committed ncu capture of the same workload
        with open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")) as f:
            traffic_tab = json.load(f).get(args.workload, {})
    except Exception:
        traffic_tab = {}
    from rnnt_transducer_b200 import _lib
    bn = _lib.lib().rnnt_b200_vocab_tile(V)                          # UMMA N of the vocabulary tiles (64 / 128 / 256)
    own = {"joint_fwd": (f"joint_kernel<{bn},0,0> (tcgen05 vocab GEMM + online LSE)", "joint_kernel_fwd_dram_bytes"),
           "joint_bwd_ds": (f"joint_kernel<{bn},1,0> (tcgen05 vocab GEMM recompute + dS epilogue)",
                            "joint_kernel_bwd_dram_bytes_per_chunk"),
           "dw2_gemm": ("dw2_gemm_kernel (tcgen05 dS^T.[h|1], MN-major operands)", "dw2_gemm_dram_bytes_per_chunk")}

    def roof(k):
        ms = kern_ms.get(k, 0.0)
        n = max(len(ks.get(k, [])) // args.steps, 1)                 # launches per step (dS chunks)
        ach = flops_fwd / (ms * 1e-3) / 1e12 if ms > 0 else 0.0      # = flop per launch / ms per launch
        return {"bound": "tensor", "kernel": own[k][0], "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                "frac": ach / peak, "peak_source": peaks["source"] + " bf16_tflops_sustained",
                "traffic": traffic_tab.get(own[k][1]), "kernel_ms": ms / n, "launches_per_step": n,
                "share_of_step": ms / ms_step}

    dominant = max(own, key=lambda k: kern_ms.get(k, 0.0))
    roofline = roof(dominant)
    roofline_fwd = roof("joint_fwd")
    total_alg = 6.0 * H * V * cells
    path = {"algorithmic_tflops": total_alg / (ms_step * 1e-3) / 1e12,
            "frac_of_sustained_peak": total_alg / (ms_step * 1e-3) / 1e12 / peak,
            "kernel_ms_one_step": {k: round(v, 4) for k, v in kern_ms.items()},
            # every vocabulary-sized GEMM of the step against the same peak (2*H*V flop per cell each;
            # joint_bwd_ds is the recompute: hardware work, not counted in the algorithmic 6*H*V)
            "gemm_frac_of_sustained_peak": {k: round(flops_fwd / (kern_ms[k] * 1e-3) / 1e12 / peak, 4)
                                            for k in ("joint_fwd", "joint_bwd_ds", "lib_gemm_dh", "dw2_gemm")
                                            if kern_ms.get(k, 0.0) > 0}}

    line = {
        "metric": METRIC, "value": cells * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16 GEMM operands / f32 accumulate / f64 lattice",
        "data": "synthetic",
        "config": {"workload": args.workload, **wl, "per_gpu_batch": B, "global_batch": B * world,
                   "lengths": "ragged" if args.ragged else "full", "cells_per_gpu": cells,
                   "l2": "working set per step (h stash 1.7 GB, dS staging up to 4 GB) exceeds the 126 MB L2",
                   "parallelism": f"dp{world} by utterance"},
        "value_btu": cells_btu * world / (ms_step * 1e-3),
        "e2e": {"value": cells * world / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e},
        "gpu_launches": kt.launches,
        "gpu_launches_per_step": launches,
        "roofline": roofline, "roofline_joint_fwd": roofline_fwd, "path": path, "clocks": clocks, "loss": float(loss),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import c_oracle, cpu_path
        c_oracle.build()
        r = cpu_path.time_cpu_path(E, D, H, V, T, U, n_utts=1, steps=2, warmup=0)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                "sample": f"1 of {B} utterances (B=1,T={T},U={U}), median of 2 passes; "
                                          "torch-CPU joint as model.py:18-37 + oracle/rnnt_oracle.c (OpenMP)"}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
