#!/usr/bin/env python3
"""gpurun_out/ of tools/gpu_profile_job_r02.sh -> the summaries committed under profiles/ (tag given on the command line).
    python tools/make_profiles.py r02a
Writes: <tag>_bench*.json (the bench lines as printed), <tag>_phase_*.txt, <tag>_launches_summary.txt,
<tag>_lm_kernel_ncu.txt / <tag>_lm_kernel_q8_ncu.txt (key metrics of the ncu --set full capture of one frame launch),
<tag>_vgemm_umma_ncu.txt, <tag>_sass_summary.txt, and refreshes profiles/lm_kernel_traffic.json (bench.py's roofline.traffic)."""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r02a"


def copy(src, dst):
    p = os.path.join(OUT, src)
    if os.path.exists(p) and os.path.getsize(p) > 0:
        shutil.copy(p, os.path.join(PROF, dst))
        return True
    return False


def raw_rows(path):
    """ncu --page raw --csv -> list of {metric: (unit, value)} per launch (+ 'Kernel Name')."""
    rows = [r for r in csv.reader(l for l in open(path, errors="replace") if l.startswith('"'))]
    if len(rows) < 3:
        return []
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {}
        for h, u, v in zip(hdr, units, r):
            d[h.split(".TriageCompute.")[-1] if ".TriageCompute." in h else h] = (u, v)
        out.append(d)
    return out


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return float("nan")


def to_bytes(unit, v):
    s = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return num(v) * s.get(unit, 1)


LM_KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fmaheavy.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
           "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio"]


def lm_summary(raw_csv, dst, title, traffic_key, traffic):
    p = os.path.join(OUT, raw_csv)
    if not os.path.exists(p):
        return
    rows = raw_rows(p)
    if not rows:
        return
    d = rows[0]
    with open(os.path.join(PROF, dst), "w") as f:
        f.write(f"# {title}\n# kernel: {d.get('Kernel Name', ('', '?'))[1]}\n")
        for k in LM_KEYS:
            if k in d:
                f.write(f"{k:90s} {d[k][1]:>18s} {d[k][0]}\n")
    if "dram__bytes_read.sum" in d:
        traffic[traffic_key] = {"dram_bytes_read": int(to_bytes(*d["dram__bytes_read.sum"])),
                                "dram_bytes_write": int(to_bytes(*d["dram__bytes_write.sum"]))}


def main():
    os.makedirs(PROF, exist_ok=True)
    for src, dst in [("bench_r02.json", f"{tag}_bench.json"), ("bench_r02_q8.json", f"{tag}_bench_q8_0.json"),
                     ("bench_r02_config3.json", f"{tag}_bench_config3_q8_0_sampled.json"),
                     ("bench_r02_long.json", f"{tag}_bench_4096.json"), ("phase_r02_batch1.txt", f"{tag}_phase_batch1.txt"),
                     ("phase_r02_q8.txt", f"{tag}_phase_q8.txt"), ("phase_r02_batch8.txt", f"{tag}_phase_batch8.txt"),
                     ("phase_r02_batch16.txt", f"{tag}_phase_batch16.txt"),
                     ("pytest_gpu_r02.log", f"{tag}_pytest_gpu.log"), ("smoke_r02.log", f"{tag}_smoke.log")]:
        copy(src, dst)
    lc = os.path.join(OUT, "launches_r02.csv")
    if os.path.exists(lc):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "summarize_launches.py"), lc,
                            "ncu --metrics gpu__time_duration.sum --clock-control none -c 1500: python bench.py --steps 1 --warmup 1 --frames 24 "
                            "--no-cpu-baseline (cold-cache, serialised per-launch times: shares, not absolutes)"],
                           capture_output=True, text=True)
        open(os.path.join(PROF, f"{tag}_launches_summary.txt"), "w").write(r.stdout)
    traffic = {"source": f"profiles/{tag}_lm_kernel_ncu.txt / {tag}_lm_kernel_q8_ncu.txt (ncu --set full --clock-control none, one frame launch of a 24-frame run)"}
    lm_summary("lm_kernel_r02_raw.csv", f"{tag}_lm_kernel_ncu.txt", "ncu --set full --clock-control none -k regex:lm_kernel -s 40 -c 1 (F16 model, one frame launch)", "f16", traffic)
    lm_summary("lm_kernel_q8_r02_raw.csv", f"{tag}_lm_kernel_q8_ncu.txt", "ncu --set full --clock-control none -k regex:lm_kernel -s 40 -c 1 (Q8_0 model, one frame launch)", "q8_0", traffic)
    if "f16" in traffic:
        json.dump(traffic, open(os.path.join(PROF, "lm_kernel_traffic.json"), "w"), indent=1)
    copy("lm_kernel_r02_details.txt", f"{tag}_lm_kernel_ncu_details.txt")
    vg = os.path.join(OUT, "vgemm_umma_r02_raw.csv")
    if os.path.exists(vg):
        keys = ["gpu__time_duration.sum", "launch__grid_size", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
                "sm__inst_executed_pipe_tensor.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
                "sm__warps_active.avg.pct_of_peak_sustained_active"]
        with open(os.path.join(PROF, f"{tag}_vgemm_umma_ncu.txt"), "w") as f:
            f.write("# ncu --set full --clock-control none -k regex:vgemm_umma -s 130 -c 12: python tools/dev/dev_voc_only.py (320-frame decode, decoder-block launches)\n")
            f.write(" | ".join(["kernel"] + keys) + "\n")
            for d in raw_rows(vg):
                f.write(" | ".join([d.get("Kernel Name", ("", "?"))[1][:32]] + [f"{d[k][1]} {d[k][0]}" if k in d else "-" for k in keys]) + "\n")
    sass_summary(os.path.join(PROF, f"{tag}_sass_summary.txt"))


def sass_summary(dst):
    lib = os.path.join(ROOT, "qwen3-tts.cpp_b200", "libq3t_b200.so")
    if not os.path.exists(lib):
        return
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    mn = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTCCP", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "LDGSTS", "HMMA", "IMMA",
          "FHFMA", "IDP.4A", "REDUX", "DFMA", "DADD"]
    per = collections.OrderedDict()
    cur = None
    tot = collections.Counter()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for k in mn:
            if re.search(r"\b" + re.escape(k) + r"\b", line) or (k + ".") in line:
                per[cur][k] += 1
                tot[k] += 1
    names = {}
    try:
        dem = subprocess.run(["c++filt"] + list(per.keys()), capture_output=True, text=True).stdout.splitlines()
        names = dict(zip(per.keys(), dem))
    except Exception:
        pass
    with open(dst, "w") as f:
        f.write("# cuobjdump -sass qwen3-tts.cpp_b200/libq3t_b200.so | count of each mnemonic   (built for sm_100a)\n"
                "# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld (TMEM), UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA bulk copy, no tensor map),\n"
                "# UTMALDG = cp.async.bulk.tensor (TMA tensor-map load), SYNCS = mbarrier ops, LDGSTS = cp.async, FHFMA = fma.rn.f32.f16, IDP.4A = dp4a\n\n"
                "whole library:\n")
        for k in mn:
            f.write(f"  {k:16s} {tot[k]}\n")
        f.write("\nper kernel (non-zero counts):\n")
        for fn, c in per.items():
            if sum(c.values()) == 0:
                continue
            f.write(f"  {names.get(fn, fn)[:110]}\n      " + ", ".join(f"{k} {c[k]}" for k in mn if c[k]) + "\n")


if __name__ == "__main__":
    main()
