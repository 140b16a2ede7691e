"""Summarise an ncu report (run here, no GPU needed):  python profiles/summarize_ncu.py gpurun_out/x.ncu-rep > profiles/x.md
Per profiled launch: duration, tensor-pipe %, DRAM bytes read+written, L2 %, registers, SM clock; then the top stall
sites of the first launch of every kernel (source page, needs -lineinfo)."""
import csv
import io
import subprocess
import sys

WANT = [("gpu__time_duration.sum", "time"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("sm__cycles_elapsed.avg.per_second", "sm_clk")]


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True, check=True).stdout


def main(rep):
    rows = list(csv.reader(io.StringIO(ncu("-i", rep, "--page", "raw", "--csv"))))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu summary of `{rep}`\n")
    print("| # | kernel | " + " | ".join(n for _, n in WANT) + " |")
    print("|---|---|" + "---|" * len(WANT))
    for n, r in enumerate(rows[2:]):
        cells = []
        for key, _ in WANT:
            if key in ix:
                v = r[ix[key]]
                try:
                    v = f"{float(v):.4g}"
                except ValueError:
                    pass
                cells.append(f"{v} {units[ix[key]]}".strip())
            else:
                cells.append("-")
        print(f"| {n} | `{r[ix['Kernel Name']][:60]}` | " + " | ".join(cells) + " |")
    src = list(csv.reader(io.StringIO(ncu("-i", rep, "--page", "source", "--csv"))))
    sections, cur, h = [], None, None
    for r in src:
        if len(r) >= 2 and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            sections.append(cur)
        elif r and r[0] == "Address":
            h = r
            cur["hdr"] = r
        elif cur is not None and h and len(r) == len(h):
            cur["rows"].append(r)
    seen = set()
    for s in sections:
        if s["name"] in seen or not s["rows"]:
            continue
        seen.add(s["name"])
        six = {k: i for i, k in enumerate(s["hdr"])}
        stalls = [k for k in s["hdr"] if k.startswith("stall_") and "Not Issued" not in k]
        total = sum(int(r[six["# Samples"]] or 0) for r in s["rows"])
        print(f"\n## stall sites: `{s['name'][:90]}` ({total} samples)\n")
        agg = {k: sum(int(r[six[k]] or 0) for r in s["rows"]) for k in stalls}
        print("totals: " + ", ".join(f"{k[6:]} {v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:6]) + "\n")
        print("| samples | SASS | top reasons |\n|---|---|---|")
        for r in sorted(s["rows"], key=lambda r: -int(r[six["# Samples"]] or 0))[:12]:
            st = {k[6:]: int(r[six[k]] or 0) for k in stalls if int(r[six[k]] or 0) > 0}
            top = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2])
            print(f"| {r[six['# Samples']]} | `{r[six['Source']].strip()[:70]}` | {top} |")


if __name__ == "__main__":
    main(sys.argv[1])
