#!/usr/bin/env python
"""Joins an ncu SASS source page with nvdisasm line info: where a kernel's issue slots and stall samples go.

  ncu -i prof.ncu-rep --page source --csv > src.csv
  cuobjdump -xelf all lib.so ; nvdisasm -g -c kernels.sm_100a.cubin > kernels.sass
  python tools/ncu_lines.py src.csv kernels.sass <mangled-kernel-substring> [top_n] [launch index in the report, 1-based]

Runs on the CPU box (no GPU needed)."""
import collections
import csv
import re
import sys


def load_lines(sass_path, kernel_sub):
    """address -> (file, line) for the first function whose section name contains kernel_sub."""
    out, cur, active = {}, ("?", 0), False
    for ln in open(sass_path, errors="replace"):
        if ln.startswith(".text."):
            active = kernel_sub in ln
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur, m.group(2))
    return out


def main():
    src_csv, sass, kernel_sub = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    want = int(sys.argv[5]) if len(sys.argv) > 5 else 1      # which captured launch of the report (1-based)
    lines = load_lines(sass, kernel_sub)
    rows = list(csv.reader(open(src_csv)))
    # the page holds one table per kernel instance: header row starts with "Address"
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])   # samples, inst, thread_inst, local ops
    by_file = collections.defaultdict(lambda: [0, 0, 0])
    stalls = collections.defaultdict(lambda: collections.Counter())
    hdr = None
    tables = 0
    first_addr = None
    for r in rows:
        if r and r[0] == "Address":
            hdr = {h: i for i, h in enumerate(r)}
            tables += 1
            continue
        if hdr is None or tables != want or not r or not re.match(r"^(0x)?[0-9a-f]+$", r[0]):
            continue
        addr = int(r[0], 16)
        if first_addr is None:
            first_addr = addr            # ncu prints device addresses: offsets are relative to the kernel's first instruction
        key, _ = lines.get(addr - first_addr, (("?", 0), ""))
        f = lambda name: float(r[hdr[name]] or 0) if name in hdr else 0.0
        smp, inst, tinst = f("# Samples"), f("Instructions Executed"), f("Thread Instructions Executed")
        a = agg[key]
        a[0] += smp; a[1] += inst; a[2] += tinst
        if r[hdr.get("Address Space", 0)] == "Local":
            a[3] += inst
        b = by_file[key[0]]
        b[0] += smp; b[1] += inst; b[2] += tinst
        for h, i in hdr.items():
            if h.startswith("stall_") and "Not Issued" not in h and r[i]:
                stalls[key][h] += float(r[i])
    tot_s = sum(a[0] for a in agg.values()) or 1
    tot_i = sum(a[1] for a in agg.values()) or 1
    print(f"total samples {tot_s:.0f}, warp instructions {tot_i:.3e}")
    print("\nby file: samples%  inst%  threads/inst")
    for k, b in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:28s} {100 * b[0] / tot_s:6.2f} {100 * b[1] / tot_i:6.2f} {b[2] / max(b[1], 1):6.2f}")
    print(f"\ntop {top} lines: samples%  inst%  threads/inst  local-ops%  top stalls")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        st = ", ".join(f"{h[6:]} {100 * v / max(a[0], 1):.0f}%" for h, v in stalls[k].most_common(3))
        print(f"  {k[0]:22s}:{k[1]:<5d} {100 * a[0] / tot_s:6.2f} {100 * a[1] / tot_i:6.2f} {a[2] / max(a[1], 1):6.2f} {100 * a[3] / max(a[1], 1):6.1f}  {st}")


if __name__ == "__main__":
    main()
