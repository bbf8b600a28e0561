"""Small inputs through every device path, for compute-sanitizer (memcheck) runs on the GPU box."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import __graft_entry__ as g  # noqa: E402
from tests import corpus  # noqa: E402


def main():
    g.smoke()
    from replisense_rfq_b200 import Engine, FileParser
    eng = Engine(0)
    fp = FileParser(engine=eng)
    docs = [corpus.json_doc(s) for s in range(300)] + [corpus.json_doc_broken(s) for s in range(300)]
    docs.append(("[" + ",".join(corpus.json_doc(s).decode() for s in range(40)) + "]").encode())
    out = fp.parse_many(docs, ["json"] * len(docs))
    print("json", sum(isinstance(o, str) for o in out), "texts")
    pairs = []
    for seed in range(60):
        sx, ss = corpus.sheet_xml_fuzz(seed) if seed % 3 == 0 else corpus.sheet_xml_doc(seed, 5, 60)
        pairs.append((corpus.strip_refs_partial(sx, seed), ss))
    d2, kind, sst_of = [], [], []
    for sx, ss in pairs:
        d2 += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(d2) - 2]
    r = eng.sheets_batch(d2, kind, sst_of, 1000)
    print("partial refs", int((r.status == 0).sum()), "ok")
    cases = [corpus.sheet_xml_dates(s, p_td=0.6) for s in range(60)]
    d3, kind, sst_of, ids = [], [], [], []
    for sx, ss, dates, tds, d1904 in cases:
        d3 += [ss, sx]; kind += [1, 2 if d1904 else 0]; sst_of += [-1, len(d3) - 2]
        ids += [(), [i | (0x80000000 if i in tds else 0) for i in dates]]
    r = eng.sheets_batch(d3, kind, sst_of, 1000, ids)
    print("timedelta sheets", int((r.status == 0).sum()), "ok")
    csvs = [corpus.fuzz_csv(s) for s in range(300)] + [corpus.rfq_csv(3000, seed=2)]
    out = eng.parse_csv_many(csvs, 1000, on_unsupported="return")
    print("csv", sum(isinstance(o, str) for o in out), "texts")


if __name__ == "__main__":
    main()
