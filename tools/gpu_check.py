"""Runs the GPU parity measurements one per subprocess (a trapped kernel poisons its CUDA
context, so isolation keeps the rest of the list alive) and writes gpurun_out/gpu_check.json.

    python tools/gpu_check.py [name-substring ...]
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHECKS = {
    "gemm_bn256_k64": "G.selftest_gemm(128, 256, 64)",
    "gemm_bn256": "G.selftest_gemm(256, 512, 512)",
    "gemm_bn64": "G.selftest_gemm(128, 192, 512)",
    "dropin_tiny": "G.dropin_golden('tiny')",
    "dropin_timit_like": "G.dropin_golden('timit_like')",
    "dropin_edge": "G.dropin_golden('edge_u0_t1')",
    "dropin_v4232": "G.dropin_synth(2, 40, 10, 4232)",
    "dropin_cfg1_init": "G.dropin_synth(2, 200, 30, 4232)",
    "dropin_cfg1_trained": "G.dropin_synth(2, 200, 30, 4232, trained_like=True)",
    "fused_fwd_timit_like": "G.fused_golden('timit_like', grads=False)",
    "fused_fwd_aishell_like": "G.fused_golden('aishell_like', grads=False)",
    "fused_fwd_aishell_dims": "G.fused_synth(2, 20, 8, 320, 320, 512, 4232, grads=False)",
    "fused_timit_like": "G.fused_golden('timit_like')",
    "fused_aishell_like": "G.fused_golden('aishell_like')",
    "fused_edge": "G.fused_golden('edge_u0_t1')",
    "fused_aishell_dims": "G.fused_synth(2, 20, 8, 320, 320, 512, 4232)",
    "fused_timit_dims": "G.fused_synth(4, 30, 12, 128, 128, 128, 49, mode='edge')",
    "fused_vs_dropin": "G.fused_matches_dropin(4, 120, 40, 320, 320, 512, 4232)",
}


def main():
    want = sys.argv[1:]
    results = {}
    for name, expr in CHECKS.items():
        if want and not any(w in name for w in want):
            continue
        code = ("import sys, json; sys.path.insert(0, %r)\n"
                "from tests import gpu_checks as G\n"
                "print('RESULT ' + json.dumps(%s))\n" % (ROOT, expr))
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
            line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            if p.returncode == 0 and line:
                results[name] = json.loads(line[-1][7:])
            else:
                results[name] = {"error": (p.stderr or p.stdout)[-1500:], "returncode": p.returncode}
        except subprocess.TimeoutExpired:
            results[name] = {"error": "timeout"}
        results[name]["seconds"] = round(time.time() - t0, 1)
        print(name, json.dumps(results[name])[:1200], flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
