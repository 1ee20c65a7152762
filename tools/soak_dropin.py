"""Soak run of boundary 1 (RNNTLoss on materialised logits) against the float64 C oracle on randomly
drawn shapes, both input regimes of SURVEY.md 8(d). Usage: python tools/soak_dropin.py [N=300] [seed=0]."""
import os, random, sys, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import gpu_checks as G

N = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for i in range(N):
    B = rng.randint(1, 6)
    T = rng.choice([1, 2, 3, 9, 31, 32, 33, 64, 100, 257])
    U = rng.choice([0, 1, 2, 7, 30, 31, 32, 63, 64, 100, 129])
    V = rng.choice([2, 3, 4, 5, 7, 8, 9, 31, 33, 49, 127, 128, 129, 130, 515, 1027, 4232])
    while B * T * (U + 1) * V > 6_000_000:
        if T > 9: T = 9
        elif U > 7: U = 7
        else: V = 49
    mode = rng.choice(["ragged", "edge", "full"])
    trained = rng.random() < 0.4
    case = dict(B=B, T=T, U=U, V=V, mode=mode, trained_like=trained, seed=5000 + i)
    try:
        m = G.dropin_synth(B, T, U, V, mode=mode, seed=5000 + i, trained_like=trained)
        assert m["loss_rel"] < 1e-5 and m["grad_max_abs"] < 1e-4 and m["grad_outside_window_max"] == 0.0, m
    except AssertionError as e:
        bad += 1
        print("FAIL", case, str(e)[:600], flush=True)
    except Exception:
        bad += 1
        print("ERROR", case, traceback.format_exc()[-600:], flush=True)
        break
print(f"soak_dropin: {N} cases, {bad} failed", flush=True)
sys.exit(1 if bad else 0)
