"""A few eager refinement iterations (M=16) for an ncu launch list: python profiles/refine_step.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.argv = [sys.argv[0], "--iters", "3"]
exec(open(os.path.join(ROOT, "profiles", "bench_refine.py")).read().split("res = {}")[0])
import torch
it16, cam, P = build(16, 3)
for _ in range(8):
    it16()
torch.cuda.synchronize()
print("ok")
