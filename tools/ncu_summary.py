"""Summarise an ncu report (raw + source pages exported as CSV) into a short text file for profiles/.
usage: python tools/ncu_summary.py <raw.csv> <source.csv> <out.txt>"""
import collections
import csv
import re
import sys

RAW_KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
            "smsp__warps_eligible.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
            "smsp__cycles_active.avg", "sm__cycles_elapsed.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__sass_inst_executed_op_local_ld.sum",
            "smsp__sass_inst_executed_op_local_st.sum"]


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


def main(raw, src, out):
    o = []
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    o.append("kernels captured: " + "; ".join(sorted(set(r[ki] for r in data))))
    for k in RAW_KEYS:
        if k in hdr:
            i = hdr.index(k)
            o.append(f"{k} [{units[i]}]: " + ", ".join(r[i] for r in data))
    rows = list(csv.reader(open(src)))
    blocks = []
    cur = None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "data": []}; blocks.append(cur); continue
        if cur is None:
            continue
        if cur["hdr"] is None:
            cur["hdr"] = r; continue
        if len(r) == len(cur["hdr"]):
            cur["data"].append(r)
    b = blocks[0]
    ix = {h: i for i, h in enumerate(b["hdr"])}
    ops = collections.Counter(); tot = 0
    for r in b["data"]:
        s = re.sub(r"^@!?U?P\d+\s+", "", r[ix["Source"]].strip())
        op = s.split()[0].split(".")[0] if s else "?"
        n = I(r[ix["Instructions Executed"]]); ops[op] += n; tot += n
    o.append(f"\nSASS of {b['name']}: {len(b['data'])} instructions, {tot} warp-instructions executed")
    o.append("opcode mix (share of executed warp-instructions): " +
             ", ".join(f"{k} {v / tot:.3f}" for k, v in ops.most_common(14)))
    st = [h for h in b["hdr"] if h.startswith("stall_") and "Not Issued" not in h]
    t = {h: sum(I(r[ix[h]]) for r in b["data"]) for h in st}
    S = sum(t.values()) or 1
    o.append("warp stall samples: " + ", ".join(f"{k[6:]} {v / S:.3f}" for k, v in sorted(t.items(), key=lambda kv: -kv[1])[:8]))
    segs = []
    for n, r in enumerate(b["data"]):
        c = I(r[ix["Instructions Executed"]])
        if segs and segs[-1][2] == c:
            segs[-1][1] = n
        else:
            segs.append([n, n, c])
    o.append("hot straight-line regions (SASS index range, length, executions, share):")
    for a, e, c in sorted(segs, key=lambda s: -(s[1] - s[0] + 1) * s[2])[:8]:
        cn = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", b["data"][i][ix["Source"]].strip()).split()[0].split(".")[0]
                                 for i in range(a, e + 1))
        o.append(f"  [{a}-{e}] len {e - a + 1} x {c} = {(e - a + 1) * c / tot:.3f}  top: " +
                 ", ".join(f"{k} {v}" for k, v in cn.most_common(5)))
    open(out, "w").write("\n".join(o) + "\n")
    print("\n".join(o))


if __name__ == "__main__":
    main(*sys.argv[1:4])
