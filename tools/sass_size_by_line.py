#!/usr/bin/env python
"""Code bytes of one kernel (with its inlined/called device functions) per source file:line bucket.
  python tools/sass_size_by_line.py <lib.so> <kernel substring> [bucket]"""
import collections, sys
sys.path.insert(0, __import__("os").path.dirname(__file__))
from ncu_lines import sass_lines
so, ksub = sys.argv[1:3]
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 20
rows = sass_lines(so, ksub)
by = collections.Counter()
for addr, loc, op, fn in rows:
    f, l = loc if loc else ("?", 0)
    by[(f, (l // bucket) * bucket)] += 16
tot = sum(by.values())
print(f"total {tot/1024:.1f} KB, {len(rows)} instructions")
for (f, l), b in sorted(by.items(), key=lambda kv: -kv[1])[:40]:
    print(f"{f}:{l}-{l+bucket-1}  {b/1024:.1f} KB")
