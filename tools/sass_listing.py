"""Full SASS listing of libb200ps.so (instructions only, demangled kernel names) -> profiles/libb200ps.sass.gz."""
import gzip
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "b200ps", "_lib", "libb200ps.so")],
                      capture_output=True, text=True).stdout.splitlines()
out = []
for line in sass:
    if "Function :" in line:
        out.append(subprocess.run(["c++filt"], input=line.strip(), capture_output=True, text=True).stdout.strip())
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?;)\s*/\*", line)
    if m:
        out.append(f"  /*{m.group(1)}*/ {m.group(2)}")
with gzip.open(os.path.join(ROOT, "profiles", "libb200ps.sass.gz"), "wt", compresslevel=9) as f:
    f.write("\n".join(out) + "\n")
print(len(out), "lines")
