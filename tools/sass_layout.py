"""Where a kernel's code comes from, in 2 KB steps of its address range (nvdisasm -g -c line markers),
and where its out-of-line device functions start.  usage: python tools/sass_layout.py <object.o> <kernel-substring>"""
import collections
import signal
import re
import subprocess
import sys
import tempfile
from pathlib import Path

signal.signal(signal.SIGPIPE, signal.SIG_DFL)   # quiet when piped into head
obj, want = Path(sys.argv[1]).resolve(), sys.argv[2]
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", str(obj)], cwd=td, check=True, capture_output=True)
    cubin = next(Path(td).glob("*.cubin"))
    text = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], check=True, capture_output=True, text=True).stdout
inside, cur, last = False, None, 0
buckets = collections.defaultdict(collections.Counter)
for ln in text.splitlines():
    if ln.startswith(".text."):
        inside = want in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = "%s:%s" % (Path(m.group(1)).name, m.group(2))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(\S+)", ln)
    if m:
        last = int(m.group(1), 16)
        buckets[last // 2048][cur.split(":")[0] if cur else None] += 16
    elif re.match(r"^\$\S+\$\S+:", ln):
        print("out-of-line function after 0x%x: %s" % (last, ln.strip()[-60:]))
for k in sorted(buckets):
    print("%6d  %s" % (k * 2048, dict(buckets[k].most_common(3))))
