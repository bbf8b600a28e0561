"""Where the time of FileParser's .xlsx batch path goes (tools/bench_configs.py C3 from containers)."""
import cProfile, pstats, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from replisense_rfq_b200 import Engine, xlsx_host
from tests import corpus

eng = Engine(0)
distinct = [corpus.sheet_xml_doc(s, 50, 1000) for s in range(150)]
cont = [corpus.make_xlsx([sx], ss) for sx, ss in distinct]
files = [cont[i % len(cont)] for i in range(1500)]
xlsx_host.parse_workbooks(eng, [xlsx_host.scan_workbook(f) for f in files[:50]], 1000)
for _ in range(2):
    t0 = time.perf_counter()
    books = [xlsx_host.scan_workbook(f) for f in files]
    t1 = time.perf_counter()
    out = xlsx_host.parse_workbooks(eng, books, 1000)
    t2 = time.perf_counter()
    print("scan %.1f ms, parse_workbooks %.1f ms (device run %.1f ms)" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, eng.last_run_ms()))
pr = cProfile.Profile(); pr.enable()
out = xlsx_host.parse_workbooks(eng, books, 1000)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
