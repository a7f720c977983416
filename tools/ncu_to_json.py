"""Record the ncu figures bench.py quotes, from a capture of THIS tree:
    python tools/ncu_to_json.py KEY REPORT.ncu-rep [--kernel REGEX] [--spp N] [--dram REPORT_OR_CSV --dram-spp N]
KEY = workload name (C2, C3, C5 ...; C2_pool for the pool pipeline).  Merges into profiles/ncu_current.json; bench.py reads that
file and nothing else (no hard-coded captures).  --dram: a second capture of the bench-sized launch with the two DRAM byte
counters (`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`), as .ncu-rep."""
import argparse, csv, io, json, os, re, subprocess, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
OUT = os.path.join(ROOT, "profiles", "ncu_current.json")
WANT = {
    "gpu__time_duration.sum": "time",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_warp_instruction",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "pipe_fma_cycles_active_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "inst_executed_pipe_alu_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "inst_executed_pipe_lsu_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "lts__t_sectors.sum": "l2_sectors",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wavefronts",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio": "stall_barrier_per_issue",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio": "stall_no_instruction_per_issue",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "stall_long_scoreboard_per_issue",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio": "stall_short_scoreboard_per_issue",
    "launch__registers_per_thread": "registers_per_thread",
}
UNIT_SCALE = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "msecond": 1.0, "usecond": 1e-3, "second": 1e3, "nsecond": 1e-6,
              "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}


def raw_rows(rep, kernel):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    name_col = hdr.index("Kernel Name")
    for r in rows[2:]:
        if re.search(kernel, r[name_col]):
            return hdr, units, r
    raise SystemExit("no kernel matching %r in %s" % (kernel, rep))


def num(v, unit):
    x = float(v.replace(",", ""))
    return x * UNIT_SCALE.get(unit, 1.0)


ap = argparse.ArgumentParser()
ap.add_argument("key"); ap.add_argument("report"); ap.add_argument("--kernel", default="k_render|k_wf_extend"); ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--dram"); ap.add_argument("--dram-spp", type=int, default=0); ap.add_argument("--profile-file", default="")
a = ap.parse_args()
hdr, units, row = raw_rows(a.report, a.kernel)
entry = {"source": a.profile_file or os.path.relpath(a.report, ROOT), "kernel": row[hdr.index("Kernel Name")], "spp_of_the_captured_launch": a.spp}
for m, short in WANT.items():
    if m in hdr:
        i = hdr.index(m)
        entry[short + ("_ms" if short == "time" else "")] = num(row[i], units[i])
if a.dram:
    h2, u2, r2 = raw_rows(a.dram, a.kernel)
    entry["dram_bytes"] = {"spp": a.dram_spp, "read": num(r2[h2.index("dram__bytes_read.sum")], u2[h2.index("dram__bytes_read.sum")]),
                           "write": num(r2[h2.index("dram__bytes_write.sum")], u2[h2.index("dram__bytes_write.sum")]),
                           "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum of the %d-spp launch (%s)" % (a.dram_spp, os.path.basename(a.dram))}
cur = {}
if os.path.exists(OUT):
    cur = json.load(open(OUT))
cur[a.key] = entry
json.dump(cur, open(OUT, "w"), indent=1, sort_keys=True)
print(json.dumps(entry, indent=1))
