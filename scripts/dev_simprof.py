import os, sys, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hddm_b200 import simulators as S
f = lambda: S.simulator([1.0, 1.5, 0.5, 0.3], model="ddm", n_samples=1000)
for _ in range(5): f()
pr = cProfile.Profile(); pr.enable()
for _ in range(100): f()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
