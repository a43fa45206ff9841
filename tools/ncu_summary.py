"""Summarise an .ncu-rep (read here, no GPU): key raw metrics + top stall lines -> text for profiles/."""
import csv, io, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "lts__t_bytes.sum ",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread ",
        "launch__grid_size", "launch__block_size", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.avg ", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled"]

def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout

def main(rep):
    raw = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units = raw[0], raw[1]
    for vals in raw[2:]:
        print("== kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
        for h, u, v in zip(hdr, units, vals):
            if any((h + " ").startswith(k) or h.startswith(k) for k in KEYS):
                print(f"  {h} [{u}] = {v}")
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv"]))))
    hi = next(i for i, r in enumerate(src) if r and r[0] == "Address")
    h = src[hi]; data = src[hi + 1:]
    ix = {k: i for i, k in enumerate(h)}
    def f(r, k):
        try: return float(r[ix[k]])
        except Exception: return 0.0
    tot = sum(f(r, "# Samples") for r in data) or 1.0
    byop = {}
    for r in data:
        t = r[1].split()
        op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")).split(".")[0]
        byop[op] = byop.get(op, 0) + f(r, "# Samples")
    print("== warp-stall samples by opcode (all warps):")
    for k, v in sorted(byop.items(), key=lambda kv: -kv[1])[:12]:
        print(f"  {k:10s} {100 * v / tot:5.1f}%")
    print("== top lines:")
    st = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:14]:
        top = sorted(((k, f(r, k)) for k in st), key=lambda kv: -kv[1])[:2]
        print(f"  {100 * f(r, '# Samples') / tot:5.1f}%  {r[1][:72]:72s} {top[0][0]}={int(top[0][1])} exec={r[ix['Instructions Executed']]}")

if __name__ == "__main__":
    main(sys.argv[1])
