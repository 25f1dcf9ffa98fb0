"""Digest of ncu CSV pages: key metrics from a --page raw export and the per-opcode stall table from --page source."""
import collections
import csv
import sys

KEYS = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "launch__grid_size", "launch__block_size"]


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[0]
    units = rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        for k in KEYS:
            if k in d:
                print("{} [{}] = {}".format(k, u.get(k, ""), d[k]))
        for k in hdr:
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and float(d[k] or 0) > 0.1:
                print("{} = {}".format(k, d[k]))


def source(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot, execd, st = collections.Counter(), collections.Counter(), collections.Counter()
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        op = r[ix["Source"]].split()
        if not op:
            continue
        o = op[1] if op[0].startswith("@") else op[0]
        o = o.split(".")[0]
        tot[o] += int(r[ix["# Samples"]] or 0)
        execd[o] += int(r[ix["Instructions Executed"]] or 0)
        for c in stall_cols:
            st[(o, c)] += int(r[ix[c]] or 0)
    S, E = sum(tot.values()), sum(execd.values())
    print("samples {} warp-instructions {}".format(S, E))
    for o, s in tot.most_common(14):
        top = sorted(((st[(o, c)], c) for c in stall_cols), reverse=True)[:3]
        print("{:8s} samples {:5.1f}%  executed {:5.1f}% ({})  top stalls: {}".format(
            o, 100.0 * s / S, 100.0 * execd[o] / E, execd[o], ", ".join("{} {}".format(c, n) for n, c in top)))


if __name__ == "__main__":
    (raw if sys.argv[1] == "raw" else source)(sys.argv[2])
