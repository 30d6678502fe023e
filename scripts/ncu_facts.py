"""Turns ncu captures (scripts/profile_all.sh) into profiles/traffic.json entries: per kernel the executed instruction
counts per unit of work, pipe utilisation, registers, local-memory traffic and DRAM bytes of ONE launch.  bench.py combines
them with launch times it measures itself (roofline blocks).
    python scripts/ncu_facts.py <out.json> <key> <rep.ncu-rep> <units_per_launch> [<key> <rep> <units> ...]"""
import csv
import json
import os
import subprocess
import sys

out_path = sys.argv[1]
facts = json.load(open(out_path)) if os.path.exists(out_path) else {}
args = sys.argv[2:]
for key, rep, units in zip(args[0::3], args[1::3], args[2::3]):
    units = float(units)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    d = dict(zip(rows[0], rows[2]))
    u = dict(zip(rows[0], rows[1]))

    def num(name, default=None):
        try:
            return float(d[name].replace(",", ""))
        except (KeyError, ValueError):
            return default

    def nbytes(name):
        v = num(name)
        if v is None:
            return None
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u.get(name, "byte"), 1)

    fp64_w = num("smsp__inst_executed_pipe_fp64.sum")
    inst_w = num("smsp__inst_executed.sum")
    thr = num("smsp__thread_inst_executed.sum")
    lld, lst = num("smsp__sass_inst_executed_op_local_ld.sum", 0.0), num("smsp__sass_inst_executed_op_local_st.sum", 0.0)
    rd, wr = nbytes("dram__bytes_read.sum"), nbytes("dram__bytes_write.sum")
    e = {
        "kernel": d.get("Kernel Name"), "units_per_launch": units, "evals_per_launch": units,
        "capture_time_ms": num("gpu__time_duration.sum", 0.0) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}.get(u.get("gpu__time_duration.sum"), 1e-6),
        "registers": num("launch__registers_per_thread"),
        "fp64_thread_inst_per_eval": fp64_w * 32 / units if fp64_w else None,
        "thread_inst_per_eval": inst_w * 32 / units if inst_w else None,
        "active_thread_inst_per_eval": thr / units if thr else None,
        "thread_inst_per_step": inst_w * 32 / units if inst_w else None,
        "fp64_pipe_active": num("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 0.0) / 100,
        "issue_active": num("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0) / 100,
        "lanes_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "local_ldst_per_eval": (lld + lst) * 32 / units,
        "dram_bytes_per_launch": (rd + wr) if rd is not None and wr is not None else None,
        "dram_read_bytes": rd, "dram_write_bytes": wr,
        "source": os.path.basename(rep),
    }
    facts[key] = e
    print(key, json.dumps(e))
json.dump(facts, open(out_path, "w"), indent=1)
