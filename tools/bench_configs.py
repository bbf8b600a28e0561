"""Secondary measurements (not the bench.py headline): configs C3/C4/C5-style corpora of
BASELINE.json at reduced document counts (the generators are pure Python), device-resident and
end-to-end, beside the oracle on one host core for a sample.  Writes gpurun_out/configs.json."""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from oracle import file_parser_oracle as O  # noqa: E402
from oracle import xlsx_reader as X  # noqa: E402
from tests import corpus  # noqa: E402


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def main():
    from replisense_rfq_b200 import Engine, build
    from replisense_rfq_b200.engine import PinnedArena
    build.build()
    eng = Engine(0)
    out = {}
    n_txt = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
    n_sheet = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    n_csv = int(sys.argv[3]) if len(sys.argv) > 3 else 4000

    # ---- C4: text bodies
    docs = [corpus.text_body(s) for s in range(n_txt)]
    nbytes = sum(map(len, docs))
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    e2e = timed(lambda: eng.text_parse_arena(arena, off, ln))
    eng.stage(arena, off, ln)
    dev = timed(lambda: eng.text_run())
    eng.fetch()
    t0 = time.perf_counter()
    for d in docs[:20000]:
        O.parse_txt_bytes(d)
    cpu = (time.perf_counter() - t0) / sum(map(len, docs[:20000]))
    out["C4_text"] = {"docs": n_txt, "MB": nbytes / 1e6, "device_MBps": nbytes / 1e6 / dev,
                      "e2e_MBps": nbytes / 1e6 / e2e, "docs_per_s_e2e": n_txt / e2e,
                      "cpu_1core_MBps": 1e-6 / cpu}
    arena.free()

    # ---- C5-style small CSV documents
    docs = [corpus.rfq_csv(10 + (s * 37) % 1990, seed=s) for s in range(n_csv)]
    nbytes = sum(map(len, docs))
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    mr = np.full(len(docs), 1000, dtype=np.int64)
    e2e = timed(lambda: eng.csv_parse_arena(arena, off, ln, mr), 3)
    eng.stage(arena, off, ln)
    dev = timed(lambda: eng.csv_run(len(docs), mr), 3)
    eng.fetch()
    t0 = time.perf_counter()
    for d in docs[:40]:
        O.parse_csv_bytes(d, 1000)
    cpu = (time.perf_counter() - t0) / sum(map(len, docs[:40]))
    out["C5_small_csv"] = {"docs": n_csv, "MB": nbytes / 1e6, "device_MBps": nbytes / 1e6 / dev,
                           "e2e_MBps": nbytes / 1e6 / e2e, "docs_per_s_e2e": n_csv / e2e,
                           "cpu_1core_MBps": 1e-6 / cpu}
    arena.free()

    # ---- C3: sheets (+ their sharedStrings)
    pairs = [corpus.sheet_xml_doc(s, 50, 1000) for s in range(n_sheet)]
    docs, kind, sst_of = [], [], []
    for sx, ss in pairs:
        docs += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(docs) - 2]
    nbytes = sum(map(len, docs))
    e2e = timed(lambda: eng.sheets_batch(docs, kind, sst_of, 1000), 3)
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    e2e_arena = timed(lambda: eng.sheets_parse_arena(arena, off, ln, kind, sst_of, 1000), 3)
    eng.set_profile(True)
    t0 = time.perf_counter()
    eng.sheets_batch(docs, kind, sst_of, 1000)
    t_call = time.perf_counter() - t0
    sheet_dev_ms = eng.last_run_ms()            # first kernel .. last kernel of rfq_sheets_run (host syncs inside)
    sheet_prof = {k: round(v, 4) for k, v in eng.profile()}
    eng.set_profile(False)
    t0 = time.perf_counter()
    for sx, ss in pairs[:20]:
        df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=1000), 1000)
        O.frame_to_text(df)
    cpu = (time.perf_counter() - t0) / sum(len(a) + len(b) for a, b in pairs[:20])
    out["C3_sheets"] = {"sheets": n_sheet, "MB": nbytes / 1e6, "e2e_MBps_incl_python_packing": nbytes / 1e6 / e2e,
                        "sheets_per_s_e2e": n_sheet / e2e,
                        "e2e_MBps": nbytes / 1e6 / e2e_arena, "sheets_per_s_e2e_arena": n_sheet / e2e_arena, "cpu_1core_MBps_restated_reader": 1e-6 / cpu,
                        "run_ms_first_to_last_kernel": sheet_dev_ms, "call_ms": t_call * 1e3,
                        "device_MBps": nbytes / 1e3 / sheet_dev_ms, "stage_ms": sheet_prof}
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "configs.json").write_text(json.dumps(out, indent=1))
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
