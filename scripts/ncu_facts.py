#!/usr/bin/env python
"""Counter facts of a workload's dominant kernel, from an `ncu --set full` capture, into profiles/r02_ncu_facts.json
(read by bench.py: roofline.traffic and roofline.ncu come from there — measured, never typed in) and the raw counter
export into profiles/r02_ncu_raw_<workload>.csv.

  python scripts/ncu_facts.py <workload> <report.ncu-rep> <chains> <transitions per launch> [--commit <hash>]

Runs here (no GPU needed): `ncu -i <rep> --page raw --csv`.
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6,
        "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main():
    wl, rep, chains, trans = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    commit = sys.argv[sys.argv.index("--commit") + 1] if "--commit" in sys.argv else \
        subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    lines = [l for l in raw.splitlines() if l.startswith('"')]
    rows = list(csv.reader(io.StringIO("\n".join(lines))))
    hdr, units, vals = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(hdr)}

    def get(name, scale_units=False):
        i = col.get(name)
        if i is None:
            return None
        try:
            v = float(vals[i].replace(",", ""))
        except ValueError:
            return None
        if scale_units:
            v *= UNIT.get(units[i], 1.0)
        return v
    dur_us = get("gpu__time_duration.sum", True)
    rd, wr = get("dram__bytes_read.sum", True), get("dram__bytes_write.sum", True)
    facts = dict(commit=commit, kernel=vals[col["Kernel Name"]] if "Kernel Name" in col else None, chains=chains, transitions=trans,
                 duration_us=dur_us, dram_bytes_read=rd, dram_bytes_write=wr,
                 dram_bytes_per_chain_transition=(rd + wr) / (chains * max(trans, 1)) if rd is not None and wr is not None else None,
                 dram_gbs=(rd + wr) / (dur_us * 1e-6) / 1e9 if rd is not None and dur_us else None,
                 issue_slots_busy_pct=get("smsp__issue_active.avg.pct_of_peak_sustained_active") or get("sm__inst_issued.avg.pct_of_peak_sustained_active"),
                 eligible_warps_per_scheduler=get("smsp__warps_eligible.avg.per_cycle_active"),
                 warps_active_pct=get("sm__warps_active.avg.pct_of_peak_sustained_active"),
                 fp64_pipe_pct=get("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") or get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                 tensor_pipe_pct=get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                 registers=get("launch__registers_per_thread"), grid=get("launch__grid_size"), block=get("launch__block_size"),
                 waves_per_sm=get("launch__waves_per_multiprocessor"), achieved_occupancy_pct=get("sm__warps_active.avg.pct_of_peak_sustained_active"),
                 l1_hit_pct=get("l1tex__t_sector_hit_rate.pct"), l2_hit_pct=get("lts__t_sector_hit_rate.pct"),
                 ipc=get("sm__inst_executed.avg.per_cycle_active"),
                 smem_bank_conflict_wavefronts=get("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"))
    # the stall reason with the largest share (warp state sampling)
    stalls = {h: get(h) for h in hdr if h.startswith("smsp__average_warp_latency_issue_stalled_") or h.startswith("smsp__average_warps_issue_stalled_")}
    stalls = {k: v for k, v in stalls.items() if v}
    if stalls:
        k = max(stalls, key=stalls.get)
        facts["top_stall"] = k.split("stalled_")[1].split(".")[0] if "stalled_" in k else k
    out = os.path.join(ROOT, "profiles", "r02_ncu_facts.json")
    allf = json.load(open(out)) if os.path.exists(out) else {}
    allf[wl] = {k: v for k, v in facts.items() if v is not None}
    json.dump(allf, open(out, "w"), indent=1, sort_keys=True)
    with open(os.path.join(ROOT, "profiles", f"r02_ncu_raw_{wl}.csv"), "w") as f:       # the raw export itself: metric, unit, value
        f.write("metric,unit,value\n")
        for h, u, v in zip(hdr, units, vals):
            f.write(f'"{h}","{u}","{v}"\n')
    print(wl, json.dumps(allf[wl])[:600])


if __name__ == "__main__":
    main()
