"""Quick look at the row pipeline on a GPU box: small documents against the oracle, with diffs."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from oracle import file_parser_oracle as O
from replisense_rfq_b200 import Engine, build
from tests import corpus


def oracle(d, n=1000):
    try:
        return O.parse_csv_bytes(d, n)
    except O.OracleParsingError:
        return None


def main():
    build.build()
    eng = Engine(0)
    docs = [b"a,b\n1,2\n3,4\n", corpus.rfq_csv(5, seed=1), corpus.rfq_csv(300, seed=3), corpus.rfq_csv(2500, seed=4)]
    docs += [corpus.fuzz_csv(s) for s in range(200)]
    t0 = time.time()
    got = eng.parse_csv_many(docs, 1000, on_unsupported="return")
    print("ran", len(docs), "docs in", round(time.time() - t0, 3), "s", flush=True)
    bad = 0
    for i, (g, d) in enumerate(zip(got, docs)):
        w = oracle(d)
        if isinstance(g, Exception):
            if not corpus.bare_cr_blank_first(d):
                bad += 1
                print("DECLINED", i, repr(d[:120]))
            continue
        if g != w:
            bad += 1
            if bad <= 6:
                print("DIFF", i, repr(d[:200]), "\nWANT", repr(w)[:400], "\nGOT ", repr(g)[:400], flush=True)
    print("bad", bad, "of", len(docs))
    for n, cap in [(70000, 10 ** 6), (200000, 10 ** 6)]:
        d = corpus.rfq_csv(n, seed=n)
        t0 = time.time()
        g = eng.parse_csv_many([d], cap)[0]
        t1 = time.time()
        print(n, "rows:", "OK" if g == oracle(d, cap) else "MISMATCH", "gpu call %.1f ms" % ((t1 - t0) * 1e3), flush=True)
    return bad


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
