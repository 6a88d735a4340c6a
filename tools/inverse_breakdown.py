"""One SPD inverse of N=16, d=4096 (for a per-kernel time list under ncu) -- diagnostic."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gadmm_b200.ops import linalg as LA

dev = torch.device("cuda", 0)
N, d = 16, 4096
A = torch.randn((N, d, d), dtype=torch.float64, device=dev)
G = A @ A.transpose(1, 2) / d
del A
LA.shifted_inverse(G, d, 1.0)
torch.cuda.synchronize()
print("ok")
