#!/usr/bin/env python3
"""Per-source-line table of an ncu capture: joins `ncu -i X.ncu-rep --page source --csv` (one row per SASS
instruction: stall samples, instructions executed) with the line table of the same kernel from
`nvdisasm -gi <cubin>` (needs -lineinfo at compile time).  The two listings hold the kernel's instructions in
the same order, so they are joined by position.

  cuobjdump -xelf all libdfchem_b200.so; nvdisasm -gi dfchem_capi.sm_100a.cubin > dis.txt
  ncu -i prof.ncu-rep --page source --csv > src.csv
  python scripts/ncu_source_hotspots.py src.csv dis.txt <mangled kernel name> [top]
"""
import collections
import csv
import re
import sys


def line_table(dis_path, kernel):
    rows, cur, block, inside = [], ("?", 0), [], False
    for ln in open(dis_path, errors="replace"):
        if ln.startswith(".text."):
            inside = ln.strip() == f".text.{kernel}:"
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            block.append((m.group(1), int(m.group(2))))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            if block:
                cur, block = block[0], []      # innermost frame comes first
            rows.append(cur)
    return rows


def main():
    src_csv, dis, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lines = line_table(dis, kernel)
    rd = csv.reader(open(src_csv))
    next(rd)
    head = next(rd)
    col = {h: i for i, h in enumerate(head)}
    stall_cols = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    data = [r for r in rd if len(r) == len(head)]
    assert len(data) == len(lines), (len(data), len(lines))
    by_line = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    by_file = collections.defaultdict(lambda: [0, 0])
    ops = collections.Counter()
    tot_s = tot_i = 0
    for r, (f, l) in zip(data, lines):
        s, n = int(r[col["# Samples"]]), int(r[col["Instructions Executed"]])
        f = f.split("/")[-1]
        e = by_line[(f, l)]
        e[0] += s
        e[1] += n
        for h in stall_cols:
            v = int(r[col[h]])
            if v:
                e[2][h[6:]] += v
        by_file[f][0] += s
        by_file[f][1] += n
        ops[re.sub(r"^@!?U?P\d+\s+", "", r[col["Source"]].strip()).split()[0].split(".")[0]] += n
        tot_s += s
        tot_i += n
    print(f"kernel {kernel}: {len(data)} SASS instructions, {tot_i} warp instructions executed, {tot_s} stall samples")
    print("\nby file: samples %  instructions %")
    for f, (s, n) in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print(f"  {f:28s} {100 * s / tot_s:6.2f} {100 * n / tot_i:6.2f}")
    print(f"\ntop {top} source lines by stall samples: file:line  samples%  instr%  top stall reasons")
    for (f, l), (s, n, st) in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
        why = ", ".join(f"{k} {100 * v / max(1, sum(st.values())):.0f}%" for k, v in st.most_common(3))
        print(f"  {f}:{l:<5d} {100 * s / tot_s:6.2f} {100 * n / tot_i:6.2f}   {why}")
    print("\nexecuted opcode mix (% of warp instructions):")
    print("  " + ", ".join(f"{k} {100 * v / tot_i:.1f}" for k, v in ops.most_common(14)))


if __name__ == "__main__":
    main()
