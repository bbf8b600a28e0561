"""SHA-256 of the oracle's full output for config C2 (1 M rows x 8 columns), seeds 0-7: what every
rank of `bench.py --gpus N` must reproduce byte for byte before its timing counts (rank r parses
corpus.rfq_csv(1_000_000, seed=r)).  About 25 s of pandas per seed on one core.

    python tests/golden/make_c2_hashes.py          # writes tests/golden/c2_sha256.json
"""
from __future__ import annotations

import hashlib
import json
import sys
from concurrent.futures import ProcessPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))

ROWS = 1_000_000


def one(seed: int):
    from oracle import file_parser_oracle as O
    from tests import corpus
    doc = corpus.rfq_csv(ROWS, seed=seed)
    text = O.parse_csv_bytes(doc, ROWS).encode("utf-8")
    return seed, {"input_sha256": hashlib.sha256(doc).hexdigest(), "input_bytes": len(doc),
                  "output_sha256": hashlib.sha256(text).hexdigest(), "output_bytes": len(text)}


def main():
    import pandas as pd
    with ProcessPoolExecutor(4) as ex:
        res = dict(ex.map(one, range(8)))
    out = {"rows": ROWS, "pandas": pd.__version__,
           "how": "oracle.file_parser_oracle.parse_csv_bytes(corpus.rfq_csv(1_000_000, seed), 1_000_000), utf-8 bytes",
           "seeds": {str(k): res[k] for k in sorted(res)}}
    (HERE / "c2_sha256.json").write_text(json.dumps(out, indent=1))
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
