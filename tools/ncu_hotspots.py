"""Per-source-line stall samples of one kernel from an ncu report.

  ncu -i rep.ncu-rep --page source --csv > sass.csv            (per-instruction samples, SASS order)
  python tools/ncu_hotspots.py sass.csv <cubin> <mangled-kernel-substring> [top]

The csv has no line numbers; `nvdisasm -g -c` of the cubin built with -lineinfo does ("//## File ..., line N" markers).
Instruction k of the csv = instruction k of the function in the disassembly (the SASS text is compared)."""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def main():
    sass_csv, cubin, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(sass_csv)))
    hdr = rows[1]
    i_s, i_all, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
    inst = [(r[i_s].strip(), int(r[i_all] or 0), int(r[i_ex] or 0)) for r in rows[2:] if len(r) > i_ex]
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    lines, cur, inside = [], None, False
    for ln in dis:
        if ln.startswith(".text.") or re.match(r"\s*\.section\s+\.text\.", ln):
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)), "inlined" in m.group(3))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            lines.append((cur, m.group(2).strip()))
    if len(lines) != len(inst):
        print(f"warning: {len(inst)} instructions in the report, {len(lines)} in the cubin", file=sys.stderr)
    per = defaultdict(lambda: [0, 0])
    total = sum(s for _, s, _ in inst)
    for k, (txt, s, ex) in enumerate(inst):
        if k >= len(lines):
            break
        key = lines[k][0][:2] if lines[k][0] else ("?", 0)
        per[key][0] += s
        per[key][1] += ex
    src_cache = {}
    print(f"total samples {total}")
    for key, (s, ex) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        try:
            if key[0] not in src_cache:
                import glob
                cands = glob.glob(f"duckdb_arrow_b200/src/**/{key[0]}", recursive=True)
                src_cache[key[0]] = open(cands[0]).read().splitlines() if cands else []
            text = src_cache[key[0]][key[1] - 1].strip()[:100]
        except Exception:  # noqa: BLE001
            pass
        print(f"{s:7d} {100 * s / max(total, 1):5.1f}%  {key[0]}:{key[1]:<5d} warp-inst {ex:>10d}  {text}")


if __name__ == "__main__":
    main()
