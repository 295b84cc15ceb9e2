#!/usr/bin/env python
"""Join an ncu SASS source page with nvdisasm -g line info and aggregate per source line.

  python tools/ncu_lines.py <report.ncu-rep> <lib.so> <kernel mangled-name substring> [top]

Needs -lineinfo at compile time.  Prints the share of executed warp instructions and of
stall samples per file:line (innermost inlined location), sorted by instructions."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(so, kernel_sub):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, check=True, capture_output=True)
    out = []
    for f in sorted(os.listdir(d)):
        txt = subprocess.run(["nvdisasm", "-g", os.path.join(d, f)], capture_output=True, text=True).stdout
        cur, inside, fn = None, False, None
        for ln in txt.splitlines():
            if ln.startswith(".text."):
                inside = kernel_sub in ln
                fn = None
                continue
            mf = re.match(r"\$[^$]+\$(\S+):\s*$", ln)
            if mf and inside:
                fn = mf.group(1)
                continue
            if ln.lstrip().startswith(".section"):
                inside = False
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), cur, m.group(2).strip(), fn))
        if out:
            break
    return out


def main():
    rep, so, ksub = sys.argv[1:4]
    so = os.path.abspath(so)
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci = {n: i for i, n in enumerate(hdr)}
    body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
    sl = sass_lines(so, ksub)
    if len(sl) != len(body):
        print(f"warning: {len(sl)} SASS instructions in the cubin vs {len(body)} in the report", file=sys.stderr)
    inst = collections.Counter()
    samp = collections.Counter()
    ops = collections.defaultdict(collections.Counter)
    n = min(len(sl), len(body))
    for k in range(n):
        loc = sl[k][1]
        e = int(body[k][ci["Instructions Executed"]] or 0)
        s = int(body[k][ci["# Samples"]] or 0)
        inst[loc] += e
        samp[loc] += s
        ops[loc][sl[k][2].split()[0] if not sl[k][2].startswith("@") else sl[k][2].split()[1]] += e
    ti, ts = sum(inst.values()), sum(samp.values())
    print(f"total warp instructions {ti}, samples {ts}")
    byfile = collections.Counter()
    for (loc, e) in inst.items():
        byfile[loc[0] if loc else None] += e
    print("by file:", {k: f"{100*v/ti:.1f}%" for k, v in byfile.items()})
    print(f"{'file:line':38s} {'inst%':>7s} {'samp%':>7s}  top opcodes")
    for loc, e in inst.most_common(top):
        o = ", ".join(f"{a}:{100*b/max(e,1):.0f}%" for a, b in ops[loc].most_common(4))
        print(f"{(loc[0]+':'+str(loc[1])) if loc else '?':38s} {100*e/ti:7.2f} {100*samp[loc]/max(ts,1):7.2f}  {o}")
    # by code region: kernel body vs called device functions (CALL targets), body split by the
    # enclosing source line of the kernel's main file (carried forward over inlined helpers)
    reg_i, reg_s = collections.Counter(), collections.Counter()
    _dm = {}
    main_file = os.environ.get("NCU_LINES_MAIN", "lmpc_warp_solve.cuh")
    bucket = int(os.environ.get("NCU_LINES_BUCKET", "50"))
    cur = 0
    for k in range(n):
        addr, loc = sl[k][0], sl[k][1]
        e = int(body[k][ci["Instructions Executed"]] or 0)
        sm = int(body[k][ci["# Samples"]] or 0)
        if sl[k][3] is None:
            if loc and loc[0] == main_file:
                cur = (loc[1] // bucket) * bucket
            key = f"body {main_file}:{cur}-{cur + bucket - 1}"
        else:
            key = "func " + subprocess.run(["c++filt", sl[k][3]], capture_output=True, text=True).stdout.strip()[:60] if sl[k][3] not in _dm else _dm[sl[k][3]]
            _dm[sl[k][3]] = key
        reg_i[key] += e
        reg_s[key] += sm
    print("by region:")
    for key, e in reg_i.most_common(30):
        print(f"  {key:44s} inst {100*e/ti:6.2f}%  samples {100*reg_s[key]/max(ts,1):6.2f}%")
    # opcode totals
    tot = collections.Counter()
    for loc in ops:
        tot.update(ops[loc])
    print("opcode totals:", ", ".join(f"{a}:{100*b/ti:.1f}%" for a, b in tot.most_common(25)))


if __name__ == "__main__":
    main()
