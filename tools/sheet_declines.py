"""Run on the GPU box: how many sheets of the test corpora are declined, and why (RFQ_SHEET_WHY=1 makes
the library say it on stderr)."""
import os
import sys
from pathlib import Path

os.environ["RFQ_SHEET_WHY"] = "1"
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import corpus  # noqa: E402


def run(eng, name, cases):
    docs, kind, sst_of, ids = [], [], [], []
    for c in cases:
        sx, ss = c[0], c[1]
        dates, tds, d1904 = (c[2], c[3], c[4]) if len(c) > 2 else ((), (), False)
        docs += [ss, sx]; kind += [1, 2 if d1904 else 0]; sst_of += [-1, len(docs) - 2]
        ids += [(), [i | (0x80000000 if i in tds else 0) for i in dates]]
    print("==", name, len(cases), "sheets", flush=True)
    sys.stderr.write("== %s\n" % name); sys.stderr.flush()
    r = eng.sheets_batch(docs, kind, sst_of, 1000, ids)
    n = sum(1 for i in range(len(cases)) if int(r.status[2 * i + 1]) == 2)
    print("   declined", n, "= %.1f %%" % (100.0 * n / len(cases)), flush=True)


def main():
    from replisense_rfq_b200 import Engine, build
    build.build()
    eng = Engine(0)
    run(eng, "fuzz", [corpus.sheet_xml_fuzz(s) for s in range(400)])
    run(eng, "dates", [corpus.sheet_xml_dates(s) for s in range(500)])
    pairs = []
    for seed in range(300):
        sx, ss = corpus.sheet_xml_fuzz(seed) if seed % 2 else corpus.sheet_xml_doc(seed, 5, 120)
        pairs.append((corpus.strip_refs(sx, rows=seed % 3 != 1, cells=seed % 3 != 2), ss))
    run(eng, "no_refs", pairs)


if __name__ == "__main__":
    main()
