"""Summarise an `ncu --set full` report of tools/profile_case.py for profiles/ (read here, on the CPU box).

usage: python tools/ncu_summary.py gpurun_out/r1k_full_raw.csv|x.ncu-rep profiles/r1k_ncu_full_summary.txt [profiles/roofline_traffic.json]

One line per kernel launch with the metrics DESIGN.md quotes; with a third argument also rewrites the per-direction
DRAM traffic / warp-instruction table that bench.py's `roofline.traffic` and `roofline.bounds.issue` read."""
import csv, io, json, subprocess, sys

KEEP = [
    ("gpu__time_duration.sum", "gpu__time_duration"),
    ("dram__bytes_read.sum", "dram__bytes_read"),
    ("dram__bytes_write.sum", "dram__bytes_write"),
    ("smsp__inst_executed.sum", "smsp__inst_executed"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "sm__warps_active"),
    ("smsp__issue_active.avg.pct", "smsp__issue_active"),
    ("launch__registers_per_thread", "launch__registers_per_thread"),
    ("launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_dynamic"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stalled_short_scoreboard_per_issue"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stalled_long_scoreboard_per_issue"),
    ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stalled_branch_resolving_per_issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stalled_wait_per_issue"),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "stalled_no_instruction_per_issue"),
    ("smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio", "stalled_sleeping_per_issue"),
    ("smsp__average_warp_latency_per_inst_issued.ratio", "warp_latency_per_inst_issued"),
    ("l1tex__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate"),
    ("lts__t_sector_hit_rate.pct", "lts__t_sector_hit_rate"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput"),
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "inst": 1.0}


def rows(rep):
    if rep.endswith(".csv"):   # already exported on the GPU box (`ncu -i x.ncu-rep --page raw --csv`): the report itself is > 64 MiB
        out = open(rep).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    out = out[out.index('"ID"'):]
    r = list(csv.reader(io.StringIO(out)))
    head, units, body = r[0], r[1], r[2:]
    return head, units, body


def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return None


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    head, units, body = rows(rep)
    col = {h: i for i, h in enumerate(head)}
    lines, per_kernel, inst = [], {}, {}
    for b in body:
        name = b[col["Kernel Name"]]
        parts = ["k=%s" % name[:60], "Grid Size=%s" % b[col["Grid Size"]], "Block Size=%s" % b[col["Block Size"]]]
        for m, short in KEEP:
            if m in col:
                parts.append("%s=%s %s" % (short, b[col[m]], units[col[m]]))
        lines.append(" | ".join(parts))
        if "avr_" not in name:
            continue
        key = ("wide " if "wide" in name else "job ") + ("decompress" if "Decompress" in name else "compress")
        rd, wr = (num(b[col[m]]) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        ok = rd is not None and wr is not None and rd == rd and wr == wr
        per_kernel[key] = (rd * SCALE.get(units[col["dram__bytes_read.sum"]], 1.0) + wr * SCALE.get(units[col["dram__bytes_write.sum"]], 1.0)) if ok else None
        i = num(b[col["smsp__inst_executed.sum"]])
        inst[key] = i if (i is not None and i == i) else None
    with open(dst, "w") as f:
        f.write("ncu --set full --clock-control none --import-source on, %s; one line per launch, serialised by ncu\n" % rep)
        f.write("\n".join(lines) + "\n")
    if len(sys.argv) > 3:
        def both(tab, d):
            v = [tab.get("wide " + d), tab.get("job " + d)]
            v = [x for x in v if x is not None] if ("wide " + d) not in tab else v
            return None if (not v or any(x is None for x in v)) else float(sum(v))
        js = {"compress": both(per_kernel, "compress"), "decompress": both(per_kernel, "decompress"),
              "warp_inst_per_launch": {"compress": both(inst, "compress"), "decompress": both(inst, "decompress")},
              "per_kernel": per_kernel, "per_kernel_warp_inst": inst,
              "source": "%s via tools/ncu_summary.py: dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum of the LAST "
                        "launch of each kernel, summed over the two kernels of a direction (they run concurrently in the bench)" % dst}
        with open(sys.argv[3], "w") as f:
            json.dump(js, f, indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
