#!/usr/bin/env python
"""Sizes and opcode mix of a kernel's body and of the device functions it calls.
  python tools/sass_funcs.py <lib.so> <kernel substring>"""
import collections, re, subprocess, sys
so, ksub = sys.argv[1:3]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
rows, inside = [], False
for ln in txt.splitlines():
    if "Function :" in ln:
        inside = ksub in ln
        continue
    if not inside:
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", ln)
    if m:
        rows.append((int(m.group(1), 16), m.group(2).strip()))
pat = re.compile(r"CALL\.REL\.NOINC (0x[0-9a-f]+)")
calls = collections.Counter(int(m.group(1), 16) for a, op in rows for m in [pat.search(op)] if m)
starts = [0] + sorted(calls) + [rows[-1][0] + 16]
for s, e in zip(starts, starts[1:]):
    ops = collections.Counter((op.split()[1] if op.startswith("@") else op.split()[0]) for a, op in rows if s <= a < e)
    print(f"{s:#8x} {(e - s) / 1024:6.1f} KB calls={calls.get(s, 0):3d} ", dict(ops.most_common(9)))
