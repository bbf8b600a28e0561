"""One C3-style sheet batch through the engine (for ncu captures of the sheet kernels).
usage: python tools/sheets_probe.py [n_sheets]"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import corpus  # noqa: E402


def main():
    from replisense_rfq_b200 import Engine, build
    build.build()
    eng = Engine(0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    pairs = [corpus.sheet_xml_doc(s, 50, 1000) for s in range(n)]
    docs, kind, sst_of = [], [], []
    for sx, ss in pairs:
        docs += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(docs) - 2]
    for _ in range(3):
        r = eng.sheets_batch(docs, kind, sst_of, 1000)
    print("sheets", n, "in MB", sum(len(d) for d in docs) / 1e6, "out bytes", int(r.total), "run ms", eng.last_run_ms())
    eng.set_profile(True)
    r = eng.sheets_batch(docs, kind, sst_of, 1000)
    for name, ms in eng.profile():
        print("   %-28s %8.3f ms" % (name, ms))


if __name__ == "__main__":
    main()
