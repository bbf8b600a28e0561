"""BASELINE.json configs C3 / C4 / C5 at their stated sizes, as extra keys of bench.py's JSON line (the
headline `value` stays C2).  Every rank builds the same seeded corpus, takes its shard (documents are
independent units: LPT partition by input bytes, no data-path collective — replisense_rfq_b200/shard.py),
parses it on its own GPU through the arena entry points of the C ABI, and the per-rank summaries are
gathered on the host.  Total work is fixed as the number of GPUs grows (strong scaling); the time of a
config is the maximum over ranks.

The generators are pure Python, so a corpus is a few hundred to a few thousand DISTINCT seeded documents
replicated to the stated count (stated in every entry); inputs are far larger than L2 either way.

  C3  10 000 sheets (50-1000 rows x 4-12 columns) + their sharedStrings, host-inflated XML
  C4  1 000 000 raw e-mail / PDF-text bodies
  C5  100 000 mixed documents: 40 % CSV (10-2000 rows), 30 % sheet XML, 30 % text, one rfq_mixed_parse
"""
from __future__ import annotations

import sys
import time
import warnings
import zlib
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _max_over_ranks(torch, dist, world, t):
    x = torch.tensor([t], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return float(x.item())


def _sum_over_ranks(torch, dist, world, v):
    x = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(x, op=dist.ReduceOp.SUM)
    return float(x.item())


def _shard(sizes, rank, world):
    from replisense_rfq_b200 import shard
    if world == 1:
        return list(range(len(sizes)))
    return shard.partition(sizes, world)[rank]


def _timed(torch, dist, world, fn, reps):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return _max_over_ranks(torch, dist, world, (time.perf_counter() - t0) / reps)


def _entry(name, n_docs, in_b, out_b, t_dev, t_e2e, hbm, unit_name, extra):
    e = {"docs": n_docs, "input_MB": in_b / 1e6, "output_MB": out_b / 1e6,
         "device": {"ms": t_dev * 1e3, "MB_per_s": in_b / 1e6 / t_dev, unit_name + "_per_s": n_docs / t_dev},
         "e2e": {"ms": t_e2e * 1e3, "MB_per_s": in_b / 1e6 / t_e2e, unit_name + "_per_s": n_docs / t_e2e,
                 "h2d_bytes": int(in_b), "d2h_bytes": int(out_b)},
         "roofline": {"bound": "hbm", "achieved": (in_b + out_b) / t_dev / 1e9, "peak": hbm, "unit": "GB/s",
                      "frac": (in_b + out_b) / t_dev / 1e9 / hbm,
                      "note": "whole pipeline: (input + output bytes) / device time"}}
    e.update(extra)
    return name, e


def run_c4(eng, torch, dist, rank, world, hbm, n_total=1_000_000, n_distinct=50_000, reps=3):
    from oracle import file_parser_oracle as O
    from replisense_rfq_b200.engine import PinnedArena
    from tests import corpus
    distinct = [corpus.text_body(s) for s in range(n_distinct)]
    sizes = [len(distinct[g % n_distinct]) for g in range(n_total)]
    mine = _shard(sizes, rank, world)
    docs = [distinct[g % n_distinct] for g in mine]
    in_b = sum(len(d) for d in docs)
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    r = eng.text_parse_arena(arena, off, ln)
    out_b = int(r.total)
    # parity: a sample of this rank's documents against the oracle, and replicas print alike
    step = max(1, len(docs) // 300)
    for k in range(0, len(docs), step):
        assert r.doc_bytes(k).decode("utf-8") == O.parse_txt_bytes(docs[k]), ("C4 parity", mine[k])
    t_e2e = _timed(torch, dist, world, lambda: eng.text_parse_arena(arena, off, ln), reps)
    eng.stage(arena, off, ln)

    def dev():
        eng.text_run()
    t_dev = _timed(torch, dist, world, dev, reps)
    eng.fetch()
    t0 = time.perf_counter()
    nb = 0
    for d in distinct[:20000]:
        O.parse_txt_bytes(d)
        nb += len(d)
    cpu = nb / 1e6 / (time.perf_counter() - t0)
    arena.free()
    tot_in, tot_out = _sum_over_ranks(torch, dist, world, in_b), _sum_over_ranks(torch, dist, world, out_b)
    return _entry("C4_text_bodies", n_total, tot_in, tot_out, t_dev, t_e2e, hbm * world, "bodies",
                  {"workload": f"{n_total} e-mail / PDF-text bodies ({n_distinct} distinct seeded bodies replicated), "
                               "text path (UTF-8 errors='ignore', universal newlines, strip)",
                   "parity": "every ~1/300th document of each shard == oracle text",
                   "cpu_sample": {"MB_per_s": cpu, "cores": 1, "kind": "port", "sample": "20000 bodies, oracle parse_txt_bytes"}})


def run_c3(eng, torch, dist, rank, world, hbm, n_total=10_000, n_distinct=400, reps=2):
    from oracle import file_parser_oracle as O
    from oracle import xlsx_reader as X
    from replisense_rfq_b200.engine import PinnedArena
    from tests import corpus
    distinct = [corpus.sheet_xml_doc(s, 50, 1000) for s in range(n_distinct)]
    sizes = [len(distinct[g % n_distinct][0]) + len(distinct[g % n_distinct][1]) for g in range(n_total)]
    mine = _shard(sizes, rank, world)
    docs, kind, sst_of = [], [], []
    for g in mine:
        sx, ss = distinct[g % n_distinct]
        docs += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(docs) - 2]
    in_b = sum(len(d) for d in docs)
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    r = eng.sheets_parse_arena(arena, off, ln, kind, sst_of, 1000)
    out_b = int(r.total)
    declined = int((np.asarray(r.status)[1::2] == 2).sum())
    step = max(1, len(mine) // 12)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for k in range(0, len(mine), step):
            sx, ss = distinct[mine[k] % n_distinct]
            if int(r.status[2 * k + 1]) != 0:
                continue
            df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=1000), 1000)
            assert r.doc_bytes(2 * k + 1).decode("utf-8") == O.frame_to_text(df), ("C3 parity", mine[k])
    t_e2e = _timed(torch, dist, world, lambda: eng.sheets_parse_arena(arena, off, ln, kind, sst_of, 1000), reps)
    t_dev = _max_over_ranks(torch, dist, world, eng.last_run_ms() / 1e3)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        t0 = time.perf_counter()
        nb = 0
        for sx, ss in distinct[:12]:
            O.frame_to_text(X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=1000), 1000))
            nb += len(sx) + len(ss)
        cpu = nb / 1e6 / (time.perf_counter() - t0)
    arena.free()
    # ---- the same sheets as real .xlsx containers (zip + DEFLATE): host zipfile inflate + sheet path, against
    # raw members uploaded and inflated on the device; rank 0, a bounded number of workbooks
    xlsx = None
    if rank == 0:
        from replisense_rfq_b200 import xlsx_host
        n_x = min(1500, n_total)
        cont = [corpus.make_xlsx([sx], ss) for sx, ss in distinct[:min(n_distinct, 150)]]
        files = [cont[i % len(cont)] for i in range(n_x)]
        xlsx_host.parse_workbooks(eng, [xlsx_host.scan_workbook(f) for f in files[:50]], 1000)   # warm-up
        t0 = time.perf_counter()
        host_txt = xlsx_host.parse_workbooks(eng, [xlsx_host.open_workbook(f) for f in files], 1000)
        t_host = time.perf_counter() - t0
        t0 = time.perf_counter()
        wbs = [xlsx_host.scan_workbook(f) for f in files]
        t_scan = time.perf_counter() - t0
        dev_txt = xlsx_host.parse_workbooks(eng, wbs, 1000)
        t_dev_x = time.perf_counter() - t0
        assert [str(a) for a in host_txt] == [str(b) for b in dev_txt], "device inflate: texts differ from the host-inflated path"
        members = [m for wb in wbs for m in ([wb.sst] + wb.sheets)]
        eng.inflate_many(members[:64])
        comp = sum(len(m.raw) for m in members)
        plain = sum(m.size for m in members)
        import numpy as _np
        a2 = PinnedArena(eng, sum((len(m.raw) + 15) & ~15 for m in members) + 4096)
        o2, l2, _ = a2.pack([m.raw for m in members])
        eng.stage(a2, o2, l2)
        meth = _np.fromiter((m.method for m in members), dtype=_np.uint8, count=len(members))
        size = _np.fromiter((m.size for m in members), dtype=_np.uint64, count=len(members))
        crc = _np.fromiter((m.crc for m in members), dtype=_np.uint32, count=len(members))
        for _ in range(2):
            eng._check(eng.lib.rfq_inflate_run(eng.ctx, meth.ctypes.data, size.ctypes.data, crc.ctypes.data), "rfq_inflate_run")
            eng._check(eng.lib.rfq_fetch_tables(eng.ctx), "rfq_fetch_tables")
        t_inf = eng.last_run_ms() / 1e3
        a2.free()
        xlsx = {"workbooks": n_x, "container_MB": sum(len(f) for f in files) / 1e6, "inflated_MB": plain / 1e6,
                "host_zipfile_path": {"s": t_host, "workbooks_per_s": n_x / t_host,
                                      "what": "zipfile + zlib inflate of every part on the host, then the sheet path"},
                "device_inflate_path": {"s": t_dev_x, "workbooks_per_s": n_x / t_dev_x, "host_directory_walk_s": t_scan,
                                        "what": "zip directory walk on the host, raw members uploaded, rfq_inflate_run + sheet path"},
                "speedup_end_to_end": t_host / t_dev_x,
                "inflate_kernel": {"ms": t_inf * 1e3, "members": len(members), "compressed_MB": comp / 1e6,
                                   "inflated_GB_per_s": plain / 1e9 / t_inf},
                "parity": "texts of both paths identical for every workbook"}
    tot_in, tot_out = _sum_over_ranks(torch, dist, world, in_b), _sum_over_ranks(torch, dist, world, out_b)
    name, e = _entry("C3_sheets", n_total, tot_in, tot_out, t_dev, t_e2e, hbm * world, "sheets",
                  {"workload": f"{n_total} worksheets of 50-1000 rows x 4-12 columns + their sharedStrings, host-inflated "
                               f"XML ({n_distinct} distinct seeded sheets replicated)",
                   "declined_sheets": int(_sum_over_ranks(torch, dist, world, declined)),
                   "device_time": "first to last kernel of rfq_sheets_run (its host synchronisations included)",
                   "parity": "~12 sheets of each shard == restated-reader oracle text",
                   "cpu_sample": {"MB_per_s": cpu, "cores": 1, "kind": "port", "sample": "12 sheets, restated openpyxl reader + pandas"}})
    if xlsx is not None:
        e["from_xlsx_containers"] = xlsx
    return name, e


def run_c5(eng, torch, dist, rank, world, hbm, n_total=100_000, n_distinct=1000, reps=2):
    """40 % CSV / 30 % sheet XML / 30 % text through rfq_mixed_parse, sharded by LPT over input bytes."""
    import random
    from oracle import file_parser_oracle as O
    from replisense_rfq_b200 import _lib
    from replisense_rfq_b200.engine import PinnedArena
    from tests import corpus
    r0 = random.Random(5)
    distinct = []   # (kind, payload)
    for j in range(n_distinct):
        x = r0.random()
        if x < 0.4:
            distinct.append(("csv", corpus.rfq_csv(r0.choice([10, 25, 60, 150, 400, 1000, 2000]), seed=j)))
        elif x < 0.7:
            distinct.append(("sheet", corpus.sheet_xml_doc(j)))
        else:
            distinct.append(("txt", corpus.text_body(j)))

    def size(j):
        k, p = distinct[j]
        return len(p[0]) + len(p[1]) if k == "sheet" else len(p)
    dsz = [size(j) for j in range(n_distinct)]
    sizes = [dsz[g % n_distinct] for g in range(n_total)]
    mine = _shard(sizes, rank, world)
    docs, kind, sst_of, owner = [], [], [], []
    for g in mine:
        k, p = distinct[g % n_distinct]
        if k == "sheet":
            docs += [p[1], p[0]]; kind += [_lib.KIND_SST, _lib.KIND_SHEET]; sst_of += [-1, len(docs) - 2]; owner += [-1, g]
        else:
            docs.append(p); kind.append(_lib.KIND_CSV if k == "csv" else _lib.KIND_TEXT); sst_of.append(-1); owner.append(g)
    in_b = sum(len(d) for d in docs)
    arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
    off, ln, _ = arena.pack(docs)
    r = eng.mixed_parse_arena(arena, off, ln, kind, 1000, sst_of)
    out_b = int(r.total)
    st = np.asarray(r.status)
    # parity on a sample: CSV and text against the oracle (sheets are covered by C3 and the GPU tests)
    step = max(1, len(docs) // 60)
    for k in range(0, len(docs), step):
        if kind[k] == _lib.KIND_CSV:
            assert int(st[k]) == 0 and r.doc_bytes(k).decode("utf-8") == O.parse_csv_bytes(docs[k], 1000), ("C5 parity", k)
        elif kind[k] == _lib.KIND_TEXT:
            assert r.doc_bytes(k).decode("utf-8") == O.parse_txt_bytes(docs[k]), ("C5 parity", k)
    # host gather: every rank's (document ids, bytes, checksum of lengths); rank 0 checks the cover
    summary = (mine, int(out_b), zlib.crc32(np.asarray(r.lengths).tobytes()))
    t_e2e = _timed(torch, dist, world, lambda: eng.mixed_parse_arena(arena, off, ln, kind, 1000, sst_of), reps)
    t_dev = _max_over_ranks(torch, dist, world, eng.last_run_ms() / 1e3)
    gathered = [summary]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, summary)
    if rank == 0:
        seen = np.zeros(n_total, dtype=np.uint8)
        for ids, _, _ in gathered:
            seen[np.asarray(ids, dtype=np.int64)] += 1
        assert (seen == 1).all(), "C5 gather: the shards do not cover the batch exactly once"
    arena.free()
    tot_in, tot_out = _sum_over_ranks(torch, dist, world, in_b), _sum_over_ranks(torch, dist, world, out_b)
    n_csv = sum(1 for g in range(n_total) if distinct[g % n_distinct][0] == "csv")
    n_sheet = sum(1 for g in range(n_total) if distinct[g % n_distinct][0] == "sheet")
    return _entry("C5_mixed", n_total, tot_in, tot_out, t_dev, t_e2e, hbm * world, "docs",
                  {"workload": f"{n_total} mixed documents ({n_csv} CSV of 10-2000 rows, {n_sheet} sheet XML + sharedStrings, "
                               f"{n_total - n_csv - n_sheet} text; {n_distinct} distinct seeded documents replicated), "
                               "one rfq_mixed_parse per rank: one arena, one upload, three pipelines",
                   "sharding": f"LPT partition over input bytes into {world} shard(s), host gather of the per-rank results, no collective",
                   "declined": int(_sum_over_ranks(torch, dist, world, int((st == 2).sum()))),
                   "device_time": "first to last kernel of rfq_mixed_run (host synchronisations of the three sub-runs included)",
                   "parity": "every ~1/60th CSV / text document of each shard == oracle text; shards cover the batch exactly once"})


def run_all(eng, torch, dist, rank, world, hbm, scale=1.0):
    out = {}
    for fn, n in ((run_c4, 1_000_000), (run_c3, 10_000), (run_c5, 100_000)):
        name, e = fn(eng, torch, dist, rank, world, hbm, n_total=max(100, int(n * scale)))
        out[name] = e
    return out


if __name__ == "__main__":   # stand-alone: python tools/bench_configs.py [scale]
    import json
    import torch
    from replisense_rfq_b200 import Engine, build
    build.build()
    print(json.dumps(run_all(Engine(0), torch, None, 0, 1, 6545.9, float(sys.argv[1]) if len(sys.argv) > 1 else 1.0), indent=1))
