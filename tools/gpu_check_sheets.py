"""Run on the GPU box: sheet parity sweep with a mismatch report under gpurun_out/."""
import json
import sys
import warnings
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import file_parser_oracle as O  # noqa: E402
from oracle import xlsx_reader as X  # noqa: E402
from tests import corpus  # noqa: E402


def main():
    from replisense_rfq_b200 import Engine, build
    build.build()
    eng = Engine(0)
    lo, hi = int(sys.argv[1]), int(sys.argv[2])
    pairs = [corpus.sheet_xml_fuzz(s) for s in range(lo, hi)]
    docs, kind, sst_of = [], [], []
    for sx, ss in pairs:
        docs += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(docs) - 2]
    r = eng.sheets_batch(docs, kind, sst_of, 1000)
    rep = []
    warnings.simplefilter("ignore")
    for i, (sx, ss) in enumerate(pairs):
        st = int(r.status[2 * i + 1])
        if st == 2:
            continue
        got = r.doc_bytes(2 * i + 1).decode("utf-8", "replace") if st == 0 else None
        df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=1000), 1000)
        want = None if df.empty else O.frame_to_text(df)
        if got != want:
            rep.append({"seed": lo + i, "status": st, "want": want, "got": got,
                        "xml": sx[sx.find(b"<sheetData>"):].decode("utf-8", "replace")[:1500]})
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "sheet_mismatch.json").write_text(json.dumps(rep[:30], indent=1, ensure_ascii=False))
    print("mismatches", len(rep), "of", len(pairs))
    for m in rep[:4]:
        print(json.dumps(m, ensure_ascii=False)[:2500])


if __name__ == "__main__":
    main()
