"""Code bytes of one kernel by source line: nvdisasm -g -c output -> 16 B per SASS instruction,
attributed to the `//## File "...", line N` marker above it (innermost inlined location).
usage: python tools/sass_lines.py <object.o> <kernel-name-substring> [top]"""
import collections
import signal
import re
import subprocess
import sys
import tempfile
from pathlib import Path

signal.signal(signal.SIGPIPE, signal.SIG_DFL)   # quiet when piped into head
obj, want = Path(sys.argv[1]).resolve(), sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", str(obj)], cwd=td, check=True, capture_output=True)
    cubin = next(Path(td).glob("*.cubin"))
    text = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], check=True, capture_output=True, text=True).stdout
by_line = collections.Counter()
by_file = collections.Counter()
cur = None
inside = False
total = 0
for ln in text.splitlines():
    if ln.startswith(".text."):
        inside = want in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (Path(m.group(1)).name, int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", ln):
        by_line[cur] += 16
        by_file[cur[0] if cur else None] += 16
        total += 16
print("total", total, "bytes")
for f, b in by_file.most_common():
    print("%7d  %s" % (b, f))
print()
for (k, b) in by_line.most_common(top):
    print("%6d  %s:%s" % (b, k[0] if k else None, k[1] if k else None))
