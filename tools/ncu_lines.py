"""Attribute ncu's per-instruction stall samples to source lines.

    python tools/ncu_lines.py <report.ncu-rep> <kernel regex> [top N]

Uses `ncu --page source --csv` (SASS with sample counts) and `nvdisasm -g` of the cubin inside
mercury_b200/libmercury_b200.so (line info), matched by instruction offset.
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(kernel_rx):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "mercury_b200", "libmercury_b200.so")], cwd=tmp,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    out = {}
    for f in os.listdir(tmp):
        if not f.endswith(".cubin") or "-" in f:
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, inl = None, None, None
        for ln in txt.splitlines():
            m = re.match(r"\.text\.(\S+):", ln)
            if m:
                cur = m.group(1) if re.search(kernel_rx, m.group(1)) else None
                continue
            if cur is None:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
            if m:
                line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.setdefault(cur, {})[int(m.group(1), 16)] = (line, m.group(2).strip())
    return out


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}", "-c", "1"],
                            capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(csvtxt)))
    h = next(i for i, r in enumerate(rows) if "Address" in r[:2])
    hdr = rows[h]
    ai, si, ii = hdr.index("Address"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
    table = sass_lines(rx)
    fn = max(table.values(), key=len)
    addrs = [int(r[ai], 16) for r in rows[h + 1:] if r and r[ai].startswith("0x")]
    base = min(addrs)
    per_line, total = {}, 0
    for r in rows[h + 1:]:
        if not r or not r[ai].startswith("0x"):
            continue
        off = int(r[ai], 16) - base
        n = int(r[si] or 0)
        total += n
        key = fn.get(off, (None, "?"))[0]
        d = per_line.setdefault(key, [0, 0])
        d[0] += n
        d[1] += int(r[ii] or 0)
    print(f"total samples {total}")
    src_cache = {}
    for key, (n, ex) in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        if key:
            p = os.path.join(ROOT, "mercury_b200", "csrc", key[0])
            if p not in src_cache and os.path.exists(p):
                src_cache[p] = open(p).read().splitlines()
            if p in src_cache and key[1] <= len(src_cache[p]):
                text = src_cache[p][key[1] - 1].strip()[:100]
        print(f"{n:6d} {100.0 * n / max(total, 1):5.1f}%  inst {ex:8d}  {key}  {text}")


if __name__ == "__main__":
    main()
