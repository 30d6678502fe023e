import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hddm_b200 import simulators as S
defaults = {"v": 0.8, "a": 1.3, "z": 0.5, "t": 0.3, "theta": 0.3, "alpha": 1.8, "beta": 2.0, "sz": 0.1, "sv": 0.5, "st": 0.1,
            "g": 0.3, "b": 0.2, "shape": 2.0, "scale": 1.5, "c": 1.0, "tinit": 0.5, "dinit": 0.5, "tslope": 1.0, "dslope": 1.0,
            "tfixedp": 1.0, "tcoh": 0.5, "dcoh": 0.5, "angle": 0.3, "vh": 0.8, "vl1": 0.5, "vl2": -0.5, "zh": 0.5, "zl1": 0.5,
            "zl2": 0.5, "d": 0.5, "alphagamma": 2.0, "scalegamma": 0.6}
bad = 0
for name, cfg in S.model_config.items():
    th = []
    for p in cfg["params"]:
        if p in defaults: th.append(defaults[p])
        elif p.startswith("v"): th.append(0.5 + 0.3 * int(p[1:]))
        elif p.startswith("z"): th.append(0.3)
        else: raise KeyError(p)
    if name == "levy": th[cfg["params"].index("alpha")] = 1.7
    out = S.simulator(th, model=name, n_samples=2000, random_state=1)
    rts, ch, meta = out
    okc = set(np.unique(ch)).issubset(set(meta["possible_choices"]) | {0})
    fin = np.isfinite(rts).all() and (rts > 0).all()
    if not (okc and fin and rts.shape == (2000, 1)): bad += 1
    print(f"{name:34s} mean rt {rts.mean():6.3f} choices {dict(zip(*np.unique(ch, return_counts=True)))} {'OK' if okc and fin else 'BAD'}")
print("models", len(S.model_config), "bad", bad)
