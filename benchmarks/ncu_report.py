"""Condense an .ncu-rep into the numbers that matter for a roofline discussion (+ top stall lines)."""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "sm__cycles_elapsed.max"]


def main(path, title=""):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = dict(zip(hdr, zip(units, vals)))
    print(f"# {title or path}")
    print(f"kernel: {d.get('Kernel Name', ('', '?'))[1]}")
    for k in KEYS:
        if k in d:
            print(f"{k:72s} {d[k][1]:>18s} {d[k][0]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    if len(srows) > 2:
        h = srows[1]
        ix = {n: i for i, n in enumerate(h)}
        stall = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
        tot = {c: 0 for c in stall}
        lines = []
        for r in srows[2:]:
            try:
                n = int(r[ix["# Samples"]])
            except (ValueError, KeyError, IndexError):
                continue
            dd = {c: int(r[ix[c]] or 0) for c in stall}
            for c in stall:
                tot[c] += dd[c]
            lines.append((n, r[ix["Source"]].strip(), dd))
        S = sum(tot.values()) or 1
        print("stall reasons (all warps): " + ", ".join(f"{c[6:]} {100 * v / S:.0f}%" for c, v in
                                                       sorted(tot.items(), key=lambda kv: -kv[1])[:6]))
        print("hottest instructions:")
        for n, s, dd in sorted(lines, key=lambda x: -x[0])[:8]:
            top = max(dd.items(), key=lambda kv: kv[1])
            print(f"  {n:6d}  {s[:78]:78s} {top[0][6:]}")
    if len(rows) > 3:
        print("all kernels in this report (us, DRAM % of peak, tensor pipe % active, L1/LSU % of peak):")
        ix = {n: i for i, n in enumerate(hdr)}
        cols = ["gpu__time_duration.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
        for r in rows[2:]:
            if len(r) < len(hdr):
                continue
            name = r[ix["Kernel Name"]].replace("void <unnamed>::", "").split("(")[0]
            nums = []
            for c in cols:
                try:
                    nums.append(f"{float(r[ix[c]].replace(',', '')):8.1f}")
                except (KeyError, ValueError):
                    nums.append("       -")
            print(f"  {name:44s} " + " ".join(nums))
    print()


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
