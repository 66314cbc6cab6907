"""cProfile of the host side of sequential small solves (what holds the GIL in bench.py --mode batch_small)."""
import cProfile, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from learningsparsepreconditioner4gpu_b200 import systems

samples = [systems.poisson_sample(g, seed=i, device="cuda") for i, g in enumerate([16, 20, 25, 32] * 4)]
ws = systems.default_workspace(samples[0])
for s in samples:
    ws.solve(s, rtol=1e-6)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    for s in samples:
        ws.solve(s, rtol=1e-6)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
