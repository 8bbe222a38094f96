"""profiles/ncu_traffic.json from an `ncu --set full` report: DRAM bytes per launch (dram__bytes_read.sum +
dram__bytes_write.sum), duration and pipe utilisation of one kernel at one shape. bench.py reads `roofline.traffic` from
this file (never a literal), so it must be regenerated from a capture of the SAME kernel version it is committed with.

    python scripts/ncu_traffic.py <report.ncu-rep> <kernel key> <shape key> [kernel-name regex]
    e.g. python scripts/ncu_traffic.py gpurun_out/attn_v51.ncu-rep attn_prefill_sm100 b16_t2048 attn_prefill_sm100
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_traffic.json")
WANT = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "sm__cycles_elapsed.avg.per_second", "sm__issue_active.avg.pct_of_peak_sustained_active"]


def to_bytes(v, unit):
    return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    rep, kernel, shape = sys.argv[1], sys.argv[2], sys.argv[3]
    pat = re.compile(sys.argv[4] if len(sys.argv) > 4 else kernel)
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    iname = hdr.index("Kernel Name")
    launches = [r for r in rows[2:] if pat.search(r[iname])]
    assert launches, "no launch matches"
    r = launches[-1]
    rec = {"report": os.path.basename(rep), "launch": r[iname][:120], "launches_in_report": len(launches)}
    for m in WANT:
        if m in hdr:
            i = hdr.index(m)
            v = r[i].replace(",", "")
            rec[m] = (to_bytes(v, units[i]) if "bytes" in m else float(v)) if v else None
            if "bytes" not in m:
                rec[m + ".unit"] = units[i]
    rec["dram_bytes_per_launch"] = rec["dram__bytes_read.sum"] + rec["dram__bytes_write.sum"]
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    data.setdefault(kernel, {})[shape] = rec
    json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)
    print(json.dumps(rec, indent=1))


if __name__ == "__main__":
    main()
