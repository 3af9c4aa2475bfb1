"""Single-GPU emulation of one rank of a sharded run: objective before / after a half step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nmf_b200 import engine as E
E.allreduce_max_ = lambda *a, **k: None
import bench

m, n, k, world, rank = [int(x) for x in sys.argv[1:6]]
eng = E.Engine(m, n, k, rank=rank, world=world)
bench.synth_fill_engine(eng, seed=0)
bench.synth_init(eng, seed=1)
torch.cuda.synchronize()
for it in range(3):
    eng.objective_eu(slot=0)
    torch.cuda.synchronize()
    print("objective", it, eng.read_scalar(0), flush=True)
    eng.mur_eu_half_step(0, 0.0)
    torch.cuda.synchronize()
    eng.mur_eu_half_step(1, 0.0)
    torch.cuda.synchronize()
print("OK", m, n, k, world, rank, flush=True)
