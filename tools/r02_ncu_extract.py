"""python tools/r02_ncu_extract.py: gpurun_out/ncu/*.ncu-rep -> profiles/r02_ncu_<workload>.txt (key metrics) and
profiles/r02_ncu_metrics.json (what bench.py's roofline block reads)."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "gpurun_out", "ncu")
M = {"dur_ns": "gpu__time_duration.sum", "dram_r": "dram__bytes_read.sum", "dram_w": "dram__bytes_write.sum",
     "inst": "smsp__inst_executed.sum", "tinst": "smsp__thread_inst_executed.sum",
     "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "tpi": "smsp__thread_inst_executed_per_inst_executed.ratio",
     "l1": "l1tex__t_sector_hit_rate.pct", "l2": "lts__t_sector_hit_rate.pct", "elig": "smsp__warps_eligible.avg.per_cycle_active",
     "wact": "sm__warps_active.avg.pct_of_peak_sustained_active", "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "lat": "smsp__average_warp_latency_per_inst_issued.ratio", "regs": "launch__registers_per_thread", "lts": "lts__t_bytes.sum",
     "fma": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "alu": "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
     "smact": "smsp__cycles_active.avg", "elapsed": "sm__cycles_elapsed.max"}
UNIT = {"us": 1e3, "ms": 1e6, "ns": 1.0, "s": 1e9, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}


def num(v, unit):
    x = float(v.replace(",", ""))
    return x * UNIT.get(unit, 1.0)


out_json = {}
path_json = os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")
if os.path.exists(path_json):
    out_json = json.load(open(path_json))
for f in sorted(os.listdir(SRC)):
    if not f.endswith(".ncu-rep"):
        continue
    wl = f[:-8]
    rep = os.path.join(SRC, f)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, r = rows[0], rows[1], rows[2]
    g = {k: num(r[hdr.index(m)], units[hdr.index(m)]) for k, m in M.items() if m in hdr}
    if "tinst" not in g:   # not in the full set: instructions x average active threads per instruction
        g["tinst"] = g["inst"] * g["tpi"]
    kernel = r[hdr.index("Kernel Name")]
    summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], stdout=subprocess.PIPE, text=True).stdout
    txt = os.path.join("profiles", f"r02_ncu_{wl}.txt")
    with open(os.path.join(ROOT, txt), "w") as fh:
        fh.write(f"# ncu --set full --clock-control none, one launch of the dominant kernel of bench.py workload {wl} (tools/r02_ncu.sh,\n"
                 f"# VORTEX_B200_SCHEDULE=0: row-major tile order; same instruction and byte counts as the ordered schedule)\n" + summ)
    bound = "issue"
    if g["issue"] < 55 and g["dram_pct"] < 40:
        bound = "latency (neither issue slots nor DRAM are near their peak: %.0f %% issue-active, %.1f eligible warps per cycle)" % (g["issue"], g["elig"])
    out_json[wl] = {"file": txt, "kernel": kernel[:60], "bound": bound, "gpus": 1, "scale_with_gpus": "incoherent" in wl,
                    "thread_inst_executed": g["tinst"], "inst_executed": g["inst"], "dram_bytes": g["dram_r"] + g["dram_w"],
                    "duration_us_under_ncu": g["dur_ns"] / 1e3, "issue_active_pct": g["issue"], "threads_per_inst": g["tpi"],
                    "issue_x_simt": g["issue"] / 100 * g["tpi"] / 32, "l1_hit_pct": g["l1"], "l2_hit_pct": g["l2"],
                    "eligible_warps_per_cycle": g["elig"], "warps_active_pct": g["wact"], "dram_throughput_pct": g["dram_pct"],
                    "warp_latency_cycles_per_inst": g["lat"], "registers": g["regs"], "l2_bytes": g.get("lts"),
                    "pipe_fma_pct": g.get("fma"), "pipe_alu_pct": g.get("alu"),
                    "sm_active_frac": (g["smact"] / g["elapsed"]) if g.get("elapsed") else None}
    print(wl, json.dumps(out_json[wl])[:400])
json.dump(out_json, open(path_json, "w"), indent=1)
for f in os.listdir(SRC):
    if f.endswith(".csv"):
        open(os.path.join(ROOT, "profiles", f), "w").write(open(os.path.join(SRC, f)).read())
