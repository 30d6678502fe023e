"""Drop-in for `hddm.simulators.simulator` (/root/reference/hddm/simulators/basic_simulator.py:6-36), which forwards
to the compiled simulators of un-vendored `ssm-simulators`; here the walks run on the GPU (hddm_b200/csrc/sim.cu).

Behaviour follows the in-tree ancestor of that dispatch (/root/reference/graveyard/basic_simulator_old.py:168-1073):
theta coercion to float32 and 2-D (:208-224), `s = 0 if no_noise else 1` (:227-230), theta column order per model,
output squeeze rules (:1041-1045), `metadata["model"]` (:1047), RT x choice binning (:70-119, :1049-1053).
Metadata carries every name in the model's parameter list plus the keys the reference's simulators return
(/root/reference/src/cddm_data_simulation.pyx:364-375, :540-553).

There is no CPU path: without the built library / a GPU every call raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

# name -> (model id, theta columns, reference simulator name, boundary function name, multiplicative, choices)
_M = {
    "ddm_hddm_base": (0, ["v", "a", "z", "t"], "ddm", "constant", True, [0, 1]),
    "full_ddm_hddm_base": (1, ["v", "a", "z", "t", "sz", "sv", "st"], "full_ddm_hddm_base", "constant", True, [0, 1]),
    "ddm": (2, ["v", "a", "z", "t"], "ddm_flexbound", "constant", True, [-1, 1]),
    "angle": (3, ["v", "a", "z", "t", "theta"], "ddm_flexbound", "angle", False, [-1, 1]),
    "weibull": (4, ["v", "a", "z", "t", "alpha", "beta"], "ddm_flexbound", "weibull_cdf", True, [-1, 1]),
    "full_ddm": (5, ["v", "a", "z", "t", "sz", "sv", "st"], "full_ddm", "constant", True, [-1, 1]),
    "ddm_sdv": (6, ["v", "a", "z", "t", "sv"], "ddm_sdv", "constant", True, [-1, 1]),
    "ornstein": (7, ["v", "a", "z", "g", "t"], "ornstein_uhlenbeck", "constant", True, [-1, 1]),
    "levy": (8, ["v", "a", "z", "alpha", "t"], "levy_flexbound", "constant", True, [-1, 1]),
    "gamma_drift": (9, ["v", "a", "z", "t", "shape", "scale", "c"], "ddm_flex", "constant", True, [-1, 1]),
    "gamma_drift_angle": (10, ["v", "a", "z", "t", "theta", "shape", "scale", "c"], "ddm_flex", "angle", False, [-1, 1]),
    "ds_conflict_drift": (11, ["a", "z", "t", "tinit", "dinit", "tslope", "dslope", "tfixedp", "tcoh", "dcoh"],
                          "ddm_flex", "constant", True, [-1, 1]),
    "ds_conflict_drift_angle": (12, ["a", "z", "t", "tinit", "dinit", "tslope", "dslope", "tfixedp", "tcoh", "dcoh", "angle"],
                                "ddm_flex", "angle", False, [-1, 1]),
    "ddm_cov": (13, ["v", "a", "z", "t"], "ddm", "constant", True, [-1, 1]),
}
# multi-accumulator models: name -> (lca, n_particles, angle, theta columns) -- column orders of the reference's dispatch
# (/root/reference/graveyard/basic_simulator_old.py:476-760) and hddm/model_config.py:939-1200
def _race_entries():
    out = {}
    for npar in (3, 4):
        vs = [f"v{i}" for i in range(npar)]
        zs = [f"z{i}" for i in range(npar)]
        out[f"race_{npar}"] = (0, npar, 0, vs + ["a"] + zs + ["t"])
        out[f"race_no_bias_{npar}"] = (0, npar, 0, vs + ["a", "z", "t"])
        out[f"race_no_bias_angle_{npar}"] = (0, npar, 1, vs + ["a", "z", "t", "theta"])
        out[f"lca_{npar}"] = (1, npar, 0, vs + ["a"] + zs + ["g", "b", "t"])
        out[f"lca_no_bias_{npar}"] = (1, npar, 0, vs + ["a", "z", "g", "b", "t"])
        out[f"lca_no_bias_angle_{npar}"] = (1, npar, 1, vs + ["a", "z", "g", "b", "t", "theta"])
    return out


_RACE = _race_entries()

# two-stage 4-choice model ddm_flexbound_par2: name -> (bound kind, theta columns); column orders of
# /root/reference/graveyard/basic_simulator_old.py:795-873 (no_bias variants fix every starting point at 0.5)
_PAR2 = {
    "ddm_par2": (0, ["vh", "vl1", "vl2", "a", "zh", "zl1", "zl2", "t"]),
    "ddm_par2_no_bias": (0, ["vh", "vl1", "vl2", "a", "t"]),
    "ddm_par2_angle_no_bias": (1, ["vh", "vl1", "vl2", "a", "t", "theta"]),
    "ddm_par2_weibull_no_bias": (2, ["vh", "vl1", "vl2", "a", "t", "alpha", "beta"]),
}

# siblings of par2 (kind 1 seq2, 2 mic2_adj, 3 tradeoff): name -> (kind, bound kind, theta columns); column orders of
# /root/reference/graveyard/basic_simulator_old.py:715-793, 875-1040.  Bound kind 3 is the conflict-gamma bound.
def _hier_entries():
    out = {}
    for kind, stem, extra in ((1, "ddm_seq2", []), (2, "ddm_mic2_adj", ["d"]), (3, "tradeoff", ["d"])):
        base = ["vh", "vl1", "vl2", "a"]
        if kind != 3:
            out[stem] = (kind, 0, base + ["zh", "zl1", "zl2"] + extra + ["t"])
        out[stem + "_no_bias"] = (kind, 0, base + extra + ["t"])
        out[stem + "_angle_no_bias"] = (kind, 1, base + extra + ["t", "theta"])
        out[stem + "_weibull_no_bias"] = (kind, 2, base + extra + ["t", "alpha", "beta"])
    out["tradeoff_conflict_gamma_no_bias"] = (3, 3, ["vh", "vl1", "vl2", "d", "t", "a", "theta", "scale", "alphagamma",
                                                     "scalegamma"])
    return out


_HIER = _hier_entries()
_BOUND_NAMES = ("constant", "angle", "weibull_cdf", "conflict_gamma_bound")

_ALIAS = {"ddm_legacy": "ddm_hddm_base", "full_ddm_legacy": "full_ddm_hddm_base", "test": "ddm",
          "weibull_cdf": "weibull"}

# minimal stand-in for the part of hddm.model_config the simulator boundary needs (params and choices per model)
model_config = {name: {"params": list(cols), "choices": list(ch), "boundary_name": bnd, "boundary_multiplicative": mult}
                for name, (_, cols, _, bnd, mult, ch) in _M.items()}
model_config.update({name: {"params": list(cols), "choices": list(range(npar)), "boundary_name": "angle" if ang else "constant",
                            "boundary_multiplicative": not ang}
                     for name, (_, npar, ang, cols) in _RACE.items()})


model_config.update({name: {"params": list(cols), "choices": [0, 1, 2, 3],
                            "boundary_name": ("constant", "angle", "weibull_cdf")[kind], "boundary_multiplicative": kind != 1}
                     for name, (kind, cols) in _PAR2.items()})


model_config.update({name: {"params": list(cols), "choices": [0, 1, 2, 3], "boundary_name": _BOUND_NAMES[bk],
                            "boundary_multiplicative": bk in (0, 2)}
                     for name, (kind, bk, cols) in _HIER.items()})


def _resolve(model):
    name = _ALIAS.get(model, model)
    if name in _RACE or name in _PAR2 or name in _HIER:
        return name
    if name not in _M:
        raise ValueError(f"unknown simulator model '{model}'")
    return name


def _coerce_theta(theta, name):
    """basic_simulator_old.py:208-224."""
    cols = (_RACE[name][3] if name in _RACE else _PAR2[name][1] if name in _PAR2 else _HIER[name][2] if name in _HIER
            else _M[name][1])
    if isinstance(theta, list):
        theta = np.asarray(theta).astype(np.float32)
    elif isinstance(theta, np.ndarray):
        theta = theta.astype(np.float32)
    elif hasattr(theta, "columns") and hasattr(theta, "values"):  # pandas DataFrame
        theta = theta[cols].values.astype(np.float32)
    elif hasattr(theta, "numpy"):  # torch tensor
        theta = theta.detach().cpu().numpy().astype(np.float32)
    else:
        raise TypeError(f"theta of type {type(theta).__name__} is not supported")
    if theta.ndim < 2:
        theta = np.expand_dims(theta, axis=0)
    if theta.ndim != 2 or theta.shape[1] != len(cols):
        raise ValueError(f"model '{name}' takes theta with {len(cols)} columns {cols}, got shape {theta.shape}")
    return np.ascontiguousarray(theta)


def _boundary_array(name, theta_row, delta_t, max_t):
    """The `boundary` metadata entry: the last parameter set's bound on t_s = arange(0, max_t + dt, dt), float32
    (cddm_data_simulation.pyx:491-509; the reference overwrites one array per trial, so the last one survives)."""
    t_s = np.arange(0, max_t + delta_t, delta_t).astype(np.float32)
    cols = _M[name][1]
    a = theta_row[cols.index("a")]
    bnd = _M[name][3]
    if bnd == "angle":
        th = theta_row[cols.index("theta") if "theta" in cols else cols.index("angle")]
        return np.add(a, np.multiply(t_s, (-np.sin(th) / np.cos(th)))).astype(np.float32)
    if bnd == "weibull_cdf":
        alpha, beta = theta_row[4], theta_row[5]
        return np.multiply(a, np.exp(-np.power(np.divide(t_s, beta), alpha))).astype(np.float32)
    return np.multiply(a, np.ones_like(t_s)).astype(np.float32)


def _seed(random_state):
    if random_state is None:
        return int(np.random.randint(0, 2 ** 62))  # np.random.seed(...) upstream makes the run reproducible
    return int(random_state) & (2 ** 64 - 1)


def _run(name, theta, n_samples, delta_t, max_t, s, seed, want_raw, nbins, sample_offset=0, want_traj=False):
    L = _lib.load()
    mid = _M[name][0]
    n_sets, n_par = theta.shape
    rts = choices = hist = traj = None
    steps = np.zeros(1, dtype=np.uint64)
    if want_raw:
        rts = np.empty((n_samples, n_sets, 1), dtype=np.float32)
        choices = np.empty((n_samples, n_sets, 1), dtype=np.int32)
    if nbins:
        hist = np.zeros((n_sets, nbins, 2), dtype=np.uint32)
    traj_len = 0
    if want_traj:
        traj_len = int(max_t / delta_t) + 1
        traj = np.full((traj_len, 1), -999, dtype=np.float32)

    def p(a):
        return a.ctypes.data_as(C.c_void_p) if a is not None else None

    _lib.check(L.wfpt_sim_run_host(mid, p(theta), n_par, n_sets, int(n_samples), int(sample_offset), float(delta_t),
                                   float(max_t), float(s), seed, p(rts), p(choices), p(hist), int(nbins or 0),
                                   float(max_t), p(steps), p(traj), traj_len))
    return rts, choices, hist, traj, int(steps[0])


def _race_canonical(name, theta):
    """Per-model theta columns -> the kernel's layout v[np], a, z[np], g, b, t, angle (a shared z is replicated)."""
    lca, npar, ang, cols = _RACE[name]
    col = {c: theta[:, i] for i, c in enumerate(cols)}
    zero = np.zeros(theta.shape[0], dtype=np.float32)
    out = [col[f"v{i}"] for i in range(npar)] + [col["a"]]
    out += [col.get(f"z{i}", col.get("z")) for i in range(npar)]
    out += [col.get("g", zero), col.get("b", zero), col["t"], col.get("theta", zero)]
    return np.ascontiguousarray(np.stack(out, axis=1), dtype=np.float32)


def _run_race(name, theta, n_samples, delta_t, max_t, s, seed, want_raw, nbins, sample_offset=0):
    L = _lib.load()
    lca, npar, ang, cols = _RACE[name]
    canon = _race_canonical(name, theta)
    n_sets = canon.shape[0]
    rts = choices = hist = None
    steps = np.zeros(1, dtype=np.uint64)
    if want_raw:
        rts = np.empty((n_samples, n_sets, 1), dtype=np.float32)
        choices = np.empty((n_samples, n_sets, 1), dtype=np.int32)
    if nbins:
        hist = np.zeros((n_sets, nbins, npar), dtype=np.uint32)

    def p(a):
        return a.ctypes.data_as(C.c_void_p) if a is not None else None

    _lib.check(L.wfpt_sim_race_run_host(lca, npar, ang, p(canon), n_sets, int(n_samples), int(sample_offset), float(delta_t),
                                        float(max_t), float(s), seed, p(rts), p(choices), p(hist), int(nbins or 0),
                                        float(max_t), p(steps)))
    return rts, choices, hist, int(steps[0])


def _simulator_race(name, model, theta, n_samples, delta_t, max_t, s, seed, bin_dim, binned):
    lca, npar, ang, cols = _RACE[name]
    n_trials = theta.shape[0]
    rts, choices, hist, steps = _run_race(name, theta, n_samples, delta_t, max_t, s, seed, True, int(bin_dim) if binned else 0)
    meta = {c: theta[:, i].copy() for i, c in enumerate(cols)}
    if "z" in meta:  # the reference also reports per-accumulator starting points (cddm_data_simulation.pyx:1350-1358)
        for i in range(npar):
            meta.setdefault(f"z{i}", meta["z"])
    t_s = np.arange(0, max_t + delta_t, delta_t).astype(np.float32)
    a_last = theta[-1, cols.index("a")]
    bnd = np.add(a_last, np.multiply(t_s, -np.sin(theta[-1, cols.index("theta")]) / np.cos(theta[-1, cols.index("theta")]))) \
        if ang else np.multiply(a_last, np.ones_like(t_s))
    meta.update({"s": s, "delta_t": delta_t, "max_t": max_t, "n_samples": n_samples,
                 "simulator": "lca" if lca else "race_model", "boundary_fun_type": "angle" if ang else "constant",
                 "possible_choices": list(range(npar)), "trajectory": None, "boundary": bnd.astype(np.float32),
                 "model": model, "n_euler_steps": steps, "seed": seed})
    if binned:
        return (hist[0].astype(np.float64) / n_samples, meta)
    if n_trials == 1:
        rts, choices = np.squeeze(rts, axis=1), np.squeeze(choices, axis=1)
    elif n_samples == 1:
        rts, choices = np.squeeze(rts, axis=0), np.squeeze(choices, axis=0)
    return (rts, choices, meta)


def _two_stage(name):
    """(kind, bound kind, theta columns) of a two-stage 4-choice model; kind 0 = par2."""
    if name in _PAR2:
        return (0,) + tuple(_PAR2[name])
    return _HIER[name]


def _par2_canonical(name, theta):
    """Per-model columns -> v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1 (starting points 0.5 where the model has none);
    the par2 siblings append d, b2, b3 (13 columns)."""
    kind, bk, cols = _two_stage(name)
    col = {c: theta[:, i] for i, c in enumerate(cols)}
    half = np.full(theta.shape[0], 0.5, dtype=np.float32)
    zero = np.zeros(theta.shape[0], dtype=np.float32)
    b0 = col.get("theta", col.get("alpha", zero))
    b1 = col.get("scale", col.get("beta", zero))
    out = [col["vh"], col["vl1"], col["vl2"], col["a"], col.get("zh", half), col.get("zl1", half), col.get("zl2", half),
           col["t"], b0, b1]
    if kind:
        out += [col.get("d", zero), col.get("alphagamma", zero), col.get("scalegamma", zero)]
    return np.ascontiguousarray(np.stack(out, axis=1), dtype=np.float32)


def _run_par2(name, theta, n_samples, delta_t, max_t, s, seed, want_raw, nbins, sample_offset=0):
    L = _lib.load()
    kind, bk, _ = _two_stage(name)
    canon = _par2_canonical(name, theta)
    n_sets = canon.shape[0]
    rts = choices = hist = None
    steps = np.zeros(1, dtype=np.uint64)
    if want_raw:
        rts = np.empty((n_samples, n_sets, 1), dtype=np.float32)
        choices = np.empty((n_samples, n_sets, 1), dtype=np.int32)
    if nbins:
        hist = np.zeros((n_sets, nbins, 4), dtype=np.uint32)

    def p(a):
        return a.ctypes.data_as(C.c_void_p) if a is not None else None

    tail = (p(canon), n_sets, int(n_samples), int(sample_offset), float(delta_t), float(max_t), float(s), seed, p(rts),
            p(choices), p(hist), int(nbins or 0), float(max_t), p(steps))
    _lib.check(L.wfpt_sim_par2_run_host(bk, *tail) if kind == 0 else L.wfpt_sim_hier_run_host(kind, bk, *tail))
    return rts, choices, hist, int(steps[0])


def _simulator_par2(name, model, theta, n_samples, delta_t, max_t, s, seed, bin_dim, binned):
    kind, bk, cols = _two_stage(name)
    n_trials = theta.shape[0]
    rts, choices, hist, steps = _run_par2(name, theta, n_samples, delta_t, max_t, s, seed, True, int(bin_dim) if binned else 0)
    canon = _par2_canonical(name, theta)
    meta = {c: theta[:, i].copy() for i, c in enumerate(cols)}
    for key, j in (("zh", 4), ("zl1", 5), ("zl2", 6)):   # the reference's metadata always carries them (:1771-1777)
        meta.setdefault(key, canon[:, j].copy())
    meta.update({"s": s, "delta_t": delta_t, "max_t": max_t, "n_samples": n_samples,
                 "simulator": "ddm_flexbound_mic2_adj" if kind == 3 else "ddm_flexbound",   # sic (:2131)
                 "boundary_fun_type": _BOUND_NAMES[bk], "possible_choices": [0, 1, 2, 3],
                 "trajectory": "This simulator does not yet allow for trajectory simulation", "model": model,
                 "n_euler_steps": steps, "seed": seed})
    if binned:
        return (hist[0].astype(np.float64) / n_samples, meta)
    if n_trials == 1:
        rts, choices = np.squeeze(rts, axis=1), np.squeeze(choices, axis=1)
    elif n_samples == 1:
        rts, choices = np.squeeze(rts, axis=0), np.squeeze(choices, axis=0)
    return (rts, choices, meta)


def simulator(theta, model="angle", n_samples=1000, delta_t=0.001, max_t=20, no_noise=False, bin_dim=None,
              bin_pointwise=False, random_state=None, **ignored):
    """(rts, choices, metadata), or (histogram, metadata) when bin_dim > 0 -- see the module docstring.

    rts float32 / choices int32 with shape (n_samples, n_trials, 1); n_trials == 1 -> (n_samples, 1);
    n_trials > 1 and n_samples == 1 -> (n_trials, 1).  `random_state` (an integer) fixes the Philox seed; when it is
    None the seed is drawn from NumPy's global RNG."""
    name = _resolve(model)
    theta = _coerce_theta(theta, name)
    n_trials = theta.shape[0]
    n_samples = int(n_samples)
    s = 0.0 if no_noise else 1.0
    seed = _seed(random_state)
    binned = bin_dim is not None and bin_dim != 0
    if binned and bin_dim < 0:
        return "invalid bin_dim"
    if binned and bin_pointwise and n_trials > 1 and n_samples > 1:
        return "currently n_trials > 1 and n_samples > 1 will not work together with bin_pointwise"
    if binned and bin_pointwise:
        # basic_simulator_old.py:1054-1068: simulate as usual, then replace every RT by its bin number
        out = simulator(theta, model=model, n_samples=n_samples, delta_t=delta_t, max_t=max_t, no_noise=no_noise,
                        random_state=seed)
        b = bin_simulator_output_pointwise(out, nbins=int(bin_dim))
        return (np.expand_dims(b[:, 0], axis=1), np.expand_dims(b[:, 1], axis=1), out[2])
    if binned and n_trials > 1:
        return "currently binned outputs not implemented for multi-trial simulators"

    if name in _RACE:
        return _simulator_race(name, model, theta, n_samples, delta_t, max_t, s, seed, bin_dim, binned)
    if name in _PAR2 or name in _HIER:
        return _simulator_par2(name, model, theta, n_samples, delta_t, max_t, s, seed, bin_dim, binned)
    rts, choices, hist, traj, steps = _run(name, theta, n_samples, delta_t, max_t, s, seed, True,
                                           int(bin_dim) if binned else 0, want_traj=True)
    _, cols, sim_name, bnd, mult, ch = _M[name]
    meta = {c: theta[:, i].copy() for i, c in enumerate(cols)}
    meta.update({"s": s, "delta_t": delta_t, "max_t": max_t, "n_samples": n_samples, "simulator": sim_name,
                 "boundary_fun_type": bnd, "possible_choices": list(ch), "trajectory": traj, "model": model,
                 "n_euler_steps": steps, "seed": seed})
    if sim_name not in ("ddm", "full_ddm_hddm_base"):
        meta["boundary"] = _boundary_array(name, theta[-1], delta_t, max_t)
    if binned:
        return (hist[0].astype(np.float64) / n_samples, meta)
    if n_trials == 1:
        rts, choices = np.squeeze(rts, axis=1), np.squeeze(choices, axis=1)
    elif n_samples == 1:
        rts, choices = np.squeeze(rts, axis=0), np.squeeze(choices, axis=0)
    return (rts, choices, meta)


def simulate_histograms(theta, model="angle", n_samples=100000, nbins=512, delta_t=0.001, max_t=20, no_noise=False,
                        random_state=None, sample_offset=0):
    """RT x choice count histograms for many parameter sets at once: uint32 [n_sets, nbins, 2] and the total number of
    Euler steps.  The per-set rule is bin_simulator_output's (basic_simulator_old.py:70-119) with freq_cnt=True; the
    reference only offers it for one parameter set per call.  No per-particle output is materialised.
    `sample_offset` shifts the particle indices that key the random streams (ranks sharding particles pass theirs and
    merge with hddm_b200.distributed.merge_histograms)."""
    name = _resolve(model)
    theta = _coerce_theta(theta, name)
    if name in _RACE or name in _PAR2 or name in _HIER:
        run = _run_race if name in _RACE else _run_par2
        _, _, hist, steps = run(name, theta, int(n_samples), delta_t, max_t, 0.0 if no_noise else 1.0,
                                _seed(random_state), False, int(nbins), sample_offset=sample_offset)
        return hist, steps
    _, _, hist, _, steps = _run(name, theta, int(n_samples), delta_t, max_t, 0.0 if no_noise else 1.0,
                                _seed(random_state), False, int(nbins), sample_offset=sample_offset)
    return hist, steps


def simulate_histograms_device(theta, model="angle", n_samples=100000, nbins=512, delta_t=0.001, max_t=20, no_noise=False,
                               random_state=0, sample_offset=0, set_offset=0, set_stride=1, hist=None, steps=None):
    """simulate_histograms with everything left on the device (the batching boundary of the simulators): `theta` is a
    float32 CUDA tensor [n_sets, n_par]; the launch goes to torch's current stream and nothing synchronises.  Returns
    (hist int32[n_sets, nbins, 2] -- the uint32 counts, ADDED to `hist` when one is passed -- and steps int64[1]).
    `set_offset` / `set_stride` name the global index of the local sets (rank r of W sharding sets round-robin passes
    r, W), `sample_offset` the global index of the first particle: both only key the random streams, so a shard's counts
    equal the same sets' / particles' counts in an unsharded run bit for bit, and an NCCL all-gather / all-reduce issued
    on the same stream right behind this call merges them without a host round trip
    (hddm_b200.distributed.simulate_histograms_sharded)."""
    import torch

    name = _resolve(model)
    if name in _RACE or name in _PAR2 or name in _HIER:
        raise NotImplementedError("device-resident histograms exist for the single-accumulator models (wfpt_sim_run)")
    if not (isinstance(theta, torch.Tensor) and theta.is_cuda and theta.dtype == torch.float32 and theta.dim() == 2):
        raise TypeError("theta must be a float32 CUDA tensor [n_sets, n_par]")
    theta = theta.contiguous()
    n_sets, n_par = theta.shape
    if hist is None:
        hist = torch.zeros((n_sets, int(nbins), 2), dtype=torch.int32, device=theta.device)
    if steps is None:
        steps = torch.zeros(1, dtype=torch.int64, device=theta.device)
    L = _lib.load()
    with torch.cuda.device(theta.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(L.wfpt_sim_run_sharded(_M[name][0], theta.data_ptr(), n_par, n_sets, int(set_offset), int(set_stride),
                                          int(n_samples), int(sample_offset), float(delta_t), float(max_t),
                                          0.0 if no_noise else 1.0, _seed(random_state), None, None, hist.data_ptr(), int(nbins),
                                          float(max_t), steps.data_ptr(), None, 0, st))
    return hist, steps


def bin_simulator_output(out, bin_dt=0.04, nbins=0, max_t=-1, freq_cnt=False):
    """Host restatement of basic_simulator_old.py:70-119 for already-materialised (rts, choices, metadata)."""
    if max_t == -1:
        max_t = out[2]["max_t"]
    if nbins == 0:
        nbins = int(max_t / bin_dt)
    bins = np.zeros(nbins + 1)
    bins[:nbins] = np.linspace(0, max_t, nbins)
    bins[nbins] = np.inf
    counts = np.zeros((nbins, len(out[2]["possible_choices"])))
    for cnt, choice in enumerate(out[2]["possible_choices"]):
        counts[:, cnt] = np.histogram(out[0][out[1] == choice], bins=bins)[0]
    if not freq_cnt:
        counts = counts / out[2]["n_samples"]
    return counts


def bin_simulator_output_pointwise(out, bin_dt=0.04, nbins=0):
    """Host restatement of basic_simulator_old.py:27-67: int32 [n, 2] = (bin number of each RT, choice with -1 -> 0).
    An RT lying exactly on a bin edge keeps its value, as in the reference's strict comparisons."""
    max_t = out[2]["max_t"]
    if nbins == 0:
        nbins = int(max_t / bin_dt)
    bins = np.zeros(nbins + 1)
    bins[:nbins] = np.linspace(0, max_t, nbins)
    bins[nbins] = np.inf
    rt = np.array(out[0], dtype=np.float64, copy=True)
    j = np.searchsorted(bins, rt, side="left")            # bins[j-1] < rt <= bins[j]
    inside = (j >= 1) & (j <= nbins) & (rt < bins[np.minimum(j, nbins)])
    binned = np.where(inside, j - 1, rt)
    ch = np.array(out[1], copy=True)
    ch[ch == -1] = 0
    return np.concatenate([binned.astype(out[0].dtype), ch.astype(out[0].dtype)], axis=-1).astype(np.int32)


def hddm_preprocess(simulator_data=None, subj_id="none", keep_negative_responses=False, add_model_parameters=False,
                    keep_subj_idx=True):
    """`(rts, choices, metadata)` -> the data frame HDDM fits (`hddm/simulators/hddm_dataset_generators.py:12-53`):
    `rt` float64, `response` int with -1 mapped to 0 unless asked otherwise, optional `subj_idx` and one column per
    model parameter (scalar for a single parameter set, per-row otherwise)."""
    import pandas as pd

    rts, choices, meta = simulator_data
    df = pd.DataFrame({"rt": np.asarray(rts, dtype=np.double).reshape(-1),
                       "response": np.asarray(choices).astype(int).reshape(-1)})
    if not keep_negative_responses:
        df.loc[df["response"] == -1, "response"] = 0
    if keep_subj_idx:
        df["subj_idx"] = subj_id
    if add_model_parameters:
        for name in model_config[meta["model"]]["params"]:
            val = np.atleast_1d(meta[name])
            df[name] = val if val.shape[0] > 1 else val[0]
    return df


def _add_outliers(sim_out=None, p_outlier=None, max_rt_outlier=10.0):
    """Replace the last Binomial(n, p_outlier) samples by outliers: RT uniform on [0, max_rt_outlier], choice uniform
    over `possible_choices` (`hddm_dataset_generators.py:56-101`).  Draws come from NumPy's global generator in the
    reference's order (binomial, uniform, choice), so a seeded caller gets the reference's outliers; in place."""
    if p_outlier == 0:
        return sim_out
    k = np.random.binomial(n=sim_out[0].shape[0], p=p_outlier)
    if k > 0:
        rt_out = np.random.uniform(low=0.0, high=max_rt_outlier, size=k)
        ch_out = np.random.choice(sim_out[2]["possible_choices"], size=k)
        sim_out[0][-k:, 0] = rt_out
        sim_out[1][-k:, 0] = ch_out
    return sim_out


def _smoke():
    rts, ch, meta = simulator([1.0, 1.0, 0.5, 0.3, 0.2], model="angle", n_samples=20000, random_state=1)
    assert rts.shape == (20000, 1) and ch.shape == (20000, 1) and meta["possible_choices"] == [-1, 1]
    p_up = float(np.mean(ch == 1))
    assert 0.80 < p_up < 0.90 and 0.8 < float(rts.mean()) < 1.1, (p_up, float(rts.mean()))
