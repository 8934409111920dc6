"""Summaries of the two ncu passes of /opt/skills/guides/B200_PROFILING.md (run here, no GPU needed):
    python profiles/summarize_ncu.py launches gpurun_out/r01f_launches.csv "<command line>"  > profiles/..._summary.txt
    python profiles/summarize_ncu.py full gpurun_out/r01f_conv_tc_full.ncu-rep               > profiles/..._ncu_full.txt
    python profiles/summarize_ncu.py dram gpurun_out/r02f_conv_tc_dram.csv "<what>" profiles/conv_tc_traffic.json > ..._dram_summary.txt
`full` also writes <rep stem>_traffic.json: per-launch average DRAM bytes of the captured kernel (bench.py's
roofline.traffic reads profiles/conv_tc_traffic.json)."""
import csv
import io
import json
import subprocess
import sys
from collections import OrderedDict


def launches(path, cmd):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0}.get(unit, 1e-6)
            rows.append((r["Kernel Name"], ms))
    tot = sum(ms for _, ms in rows)
    agg = OrderedDict()
    for k, ms in rows:
        k = k.split("(")[0][:72]
        n, t = agg.get(k, (0, 0.0))
        agg[k] = (n + 1, t + ms)
    print("# ncu --metrics gpu__time_duration.sum --clock-control none: %s" % cmd)
    print("# %d launches, %.1f ms total (cold-cache, serialised: compare shares)" % (len(rows), tot))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        print("%-74s %6d %10.3f ms %5.1f%%" % (k, n, t, 100 * t / tot))


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def to_bytes(v, unit):
    u = unit.lower()
    return v * (1e9 if u.startswith("g") else 1e6 if u.startswith("m") else 1e3 if u.startswith("k") else 1.0)


def full(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rd = csv.reader(io.StringIO(out))
    header = next(rd)
    units = next(rd)
    idx = {h: i for i, h in enumerate(header)}
    rows = list(rd)
    name_i = idx["Kernel Name"]
    per = OrderedDict()
    tot_r = tot_w = tot_t = 0.0
    for r in rows:
        k = r[name_i].split("(")[0]
        d = per.setdefault(k, {"n": 0})
        d["n"] += 1
        for m in WANT:
            if m in idx and r[idx[m]] not in ("", "n/a"):
                v = float(r[idx[m]].replace(",", ""))
                if m.startswith("dram__bytes") or m.startswith("lts__t_bytes"):
                    v = to_bytes(v, units[idx[m]])
                elif m == "gpu__time_duration.sum":
                    u = units[idx[m]].lower()
                    v = v * (1e-3 if u.startswith("n") else 1.0 if u.startswith("u") else 1e3)   # -> us
                d[m] = d.get(m, 0.0) + v
        tot_r += to_bytes(float(r[idx["dram__bytes_read.sum"]].replace(",", "")), units[idx["dram__bytes_read.sum"]])
        tot_w += to_bytes(float(r[idx["dram__bytes_write.sum"]].replace(",", "")), units[idx["dram__bytes_write.sum"]])
    print("# ncu --set full --clock-control none, %d launches in %s (sums over launches unless marked avg)" % (len(rows), rep))
    for k, d in per.items():
        n = d["n"]
        print("----\n%s   launches = %d" % (k, n))
        for m in WANT:
            if m in d:
                avg = ("pct" in m) or m.startswith("launch__")
                print("  %-62s %s%.4g" % (m, "avg " if avg else "sum ", d[m] / n if avg else d[m]))
    traffic = {"launches": len(rows), "dram_bytes_read": tot_r, "dram_bytes_write": tot_w,
               "dram_bytes_per_launch": (tot_r + tot_w) / max(len(rows), 1), "source": rep}
    print("----\nall captured launches: DRAM read %.1f MB + write %.1f MB = %.2f MB per launch" %
          (tot_r / 1e6, tot_w / 1e6, traffic["dram_bytes_per_launch"] / 1e6))
    with open(rep.rsplit(".", 1)[0] + "_traffic.json", "w") as f:
        json.dump(traffic, f)


def dram(path, cmd, out_json):
    """per-variant sums of the dram / L2 / tensor-pipe pass over every conv_tc launch of one forward (csv log of
    `ncu --metrics dram__bytes_read.sum,... -k regex:conv_tc_kernel`); writes the traffic json bench.py reads"""
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    per_id = OrderedDict()
    for r in csv.DictReader(io.StringIO("".join(lines))):
        d = per_id.setdefault(r["ID"], {"k": r["Kernel Name"].split("(")[0]})
        v = float(r["Metric Value"].replace(",", ""))
        m, u = r["Metric Name"], r.get("Metric Unit", "")
        if m.startswith("dram__bytes") or m.startswith("lts__t_bytes"):
            v = to_bytes(v, u[:1] if u.lower() not in ("byte", "bytes") else "")
        elif m == "gpu__time_duration.sum":
            ul = u.lower()
            v = v * (1e-3 if ul.startswith("n") else 1.0 if ul.startswith("u") else 1e3)
        d[m] = v
    agg = OrderedDict()
    for d in per_id.values():
        a = agg.setdefault(d["k"], {"n": 0, "t": 0.0, "r": 0.0, "w": 0.0, "l2": 0.0, "tp": 0.0})
        t = d.get("gpu__time_duration.sum", 0.0)
        a["n"] += 1; a["t"] += t; a["r"] += d.get("dram__bytes_read.sum", 0.0); a["w"] += d.get("dram__bytes_write.sum", 0.0)
        a["l2"] += d.get("lts__t_bytes.sum", 0.0)
        a["tp"] += t * d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0)
    n = sum(a["n"] for a in agg.values()); t = sum(a["t"] for a in agg.values())
    r = sum(a["r"] for a in agg.values()); w = sum(a["w"] for a in agg.values())
    print("# %s" % cmd)
    print("# template arguments: <BN, EPI, HALO, DS, PAIR, GNF>")
    print("# %d launches, %.1f us total, DRAM read %.1f MB + write %.1f MB = %.2f MB per launch, L2 traffic %.1f MB, "
          "tensor pipe active %.1f%% (time-weighted over all launches)" %
          (n, t, r / 1e6, w / 1e6, (r + w) / max(n, 1) / 1e6, sum(a["l2"] for a in agg.values()) / 1e6,
           sum(a["tp"] for a in agg.values()) / max(t, 1e-9)))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
        print("%-46s n=%3d %9.1f us  dram r %8.1f MB w %8.1f MB  L2 %9.1f MB  tensor pipe active %.1f%% (time-weighted)" %
              (k, a["n"], a["t"], a["r"] / 1e6, a["w"] / 1e6, a["l2"] / 1e6, a["tp"] / max(a["t"], 1e-9)))
    if out_json:
        with open(out_json, "w") as f:
            json.dump({"launches": n, "dram_bytes_read": r, "dram_bytes_write": w, "dram_bytes_per_launch": (r + w) / max(n, 1),
                       "source": cmd}, f)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    elif sys.argv[1] == "dram":
        dram(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "", sys.argv[4] if len(sys.argv) > 4 else None)
    else:
        full(sys.argv[2])
