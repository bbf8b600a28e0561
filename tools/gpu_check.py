"""Run on the GPU box: parity sweep with a detailed mismatch report under gpurun_out/."""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import file_parser_oracle as O  # noqa: E402
from tests import corpus  # noqa: E402


def main():
    from replisense_rfq_b200 import Engine, build
    build.build()
    eng = Engine(0)
    lo, hi = int(sys.argv[1]), int(sys.argv[2])
    docs = [corpus.fuzz_csv(s) for s in range(lo, hi)]
    t0 = time.time()
    r = eng.csv_batch(docs, 1000)
    rep = []
    for i, d in enumerate(docs):
        try:
            want = O.parse_csv_bytes(d, 1000)
        except O.OracleParsingError:
            want = None
        st = int(r.status[i])
        got = r.doc_bytes(i).decode("utf-8", "replace") if st == 0 else ("UNSUP" if st == 2 else None)
        if got != want:
            rep.append({"seed": lo + i, "data": repr(d[:500]), "want": want, "got": got, "status": st})
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "csv_mismatch.json").write_text(json.dumps(rep[:40], indent=1, ensure_ascii=False))
    print("mismatches", len(rep), "of", len(docs), "in %.1fs" % (time.time() - t0))
    for m in rep[:5]:
        print(json.dumps(m, ensure_ascii=False)[:1500])


if __name__ == "__main__":
    main()
