"""Every fused-path parity measurement of tests/test_fused_gpu.py in one run, with the TF32 first-layer backward on
and off, written to gpurun_out/parity_report.json (committed as profiles/r02_parity.json). The stated bf16-path
tolerances in tests/test_fused_gpu.py are set from the worst case found here."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import gpu_checks as G

CASES = {
    "golden_timit_like": lambda: G.fused_golden("timit_like"),
    "golden_aishell_like": lambda: G.fused_golden("aishell_like"),
    "golden_edge_u0_t1": lambda: G.fused_golden("edge_u0_t1"),
    "synth_aishell_dims": lambda: G.fused_synth(2, 20, 8, 320, 320, 512, 4232, mode="ragged"),
    "synth_timit_dims_edge": lambda: G.fused_synth(4, 30, 12, 128, 128, 128, 49, mode="edge"),
    "synth_single_cell": lambda: G.fused_synth(1, 1, 0, 64, 64, 64, 300, mode="full"),
    "synth_u130": lambda: G.fused_synth(2, 17, 130, 64, 64, 256, 1000, mode="ragged"),
    "synth_v270": lambda: G.fused_synth(2, 9, 4, 64, 64, 64, 270, mode="ragged"),
    "synth_v460": lambda: G.fused_synth(2, 9, 4, 64, 64, 128, 460, mode="full"),
    "cfg0_B8_T200_U30": lambda: G.fused_vs_large_oracle(8, 200, 30, 320, 320, 512, 4232, mode="full"),
    "cfg1_timit_B32_T300_U60": lambda: G.fused_vs_large_oracle(32, 300, 60, 128, 128, 128, 49, mode="ragged"),
    "cfg2_4utt_T500_U100": lambda: G.fused_vs_large_oracle(4, 500, 100, 320, 320, 512, 4232,
                                                         lens=([500, 431, 1, 377], [100, 0, 73, 58])),
}
out = {}
for tf32 in ("1", "0"):
    os.environ["RNNT_B200_TF32_BWD"] = tf32
    for name, fn in CASES.items():
        if tf32 == "0" and not name.startswith(("cfg0", "cfg2", "synth_aishell")):
            continue
        m = fn()
        out[f"{name}|tf32_bwd={tf32}"] = m
        keys = [k for k in m if k.startswith("loss_rel") or "rel_l2" in k]
        print(name, "tf32_bwd", tf32, {k: float(f"{m[k]:.3g}") for k in keys}, flush=True)
# the H = 512 cases once more through the recomputing backward (the default for them stores the exponentials)
os.environ["RNNT_B200_TF32_BWD"] = "1"
os.environ["RNNT_B200_BWD"] = "recompute"
for name in ("golden_aishell_like", "synth_aishell_dims", "cfg0_B8_T200_U30", "cfg2_4utt_T500_U100"):
    m = CASES[name]()
    out[f"{name}|bwd=recompute"] = m
    keys = [k for k in m if k.startswith("loss_rel") or "rel_l2" in k]
    print(name, "bwd=recompute", {k: float(f"{m[k]:.3g}") for k in keys}, flush=True)
del os.environ["RNNT_B200_BWD"]
worst = {}
for name, m in out.items():
    if name.endswith("=1"):
        for k, v in m.items():
            if k.startswith("loss_rel") or "rel_l2" in k:
                worst[k] = max(worst.get(k, 0.0), v)
print("WORST (tf32_bwd=1):", {k: float(f"{v:.3g}") for k, v in sorted(worst.items())})
out["worst_tf32_bwd=1"] = worst
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "parity_report.json"), "w"), indent=1)
