"""Eager against CUDA-graphed fused step (JointRNNTLoss / GraphedJointRNNTLoss) at the BASELINE shapes."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import gpu_checks as G
for name, dims in [("configs[0] B8 T200 U30 aishell", (8, 200, 30, 320, 320, 512, 4232)),
                   ("configs[1] B32 T300 U60 timit", (32, 300, 60, 128, 128, 128, 49)),
                   ("configs[2] B32 T500 U100 aishell", (32, 500, 100, 320, 320, 512, 4232))]:
    m = G.graphed_matches_eager(*dims)
    print(name, json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in m.items()}), flush=True)
