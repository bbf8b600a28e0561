"""bench.py — BASELINE.json's metric on config C2 (synthetic RFQ CSV, 1M rows x 8 cols, 1 B200 per
rank): parsed input MB/s (+ cells/s) through the CSV hot path, beside the reference's CPU path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R]

One step = one pass of the whole CSV path (tokenise, type, dropna/fillna, to_string) over one
1M-row document per rank.  `value` = device-resident throughput (input already in HBM, CUDA
events around every step, L2 flushed between steps); `e2e` = the same through the C-ABI call
rfq_csv_parse with HOST (pinned) buffers: H2D + kernels + D2H inside the timed region.
N > 1: one process per GPU (torchrun), independent documents per rank, no data-path collective
(weak scaling); time = max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "csv_parsed_input_MB_per_s"
UNIT = "MB/s"


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons DURING the timed regions (NVML, ~2 ms period; nvidia-smi as a
    fallback)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.sm, self.mx, self.reasons, self.stop_flag = index, [], 0, set(), False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nvml = None

    def run(self):
        while not self.stop_flag:
            if self.nvml is not None:
                try:
                    self.sm.append(self.nvml.nvmlDeviceGetClockInfo(self.h, self.nvml.NVML_CLOCK_SM))
                    r = self.nvml.nvmlDeviceGetCurrentClocksEventReasons(self.h) \
                        if hasattr(self.nvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                        else self.nvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in self.REASONS.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
                time.sleep(0.002)
            else:
                try:
                    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm",
                                          "--format=csv,noheader,nounits", "-i", str(self.index)],
                                         capture_output=True, text=True, timeout=5).stdout.split(",")
                    self.sm.append(float(out[0])); self.mx = max(self.mx, float(out[1]))
                except Exception:
                    pass
                time.sleep(0.05)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def make_doc(rows: int, seed: int) -> bytes:
    from tests import corpus
    return corpus.rfq_csv(rows, seed=seed)


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path on all host cores, a bounded sample of the same
    workload per step: FileParser._parse_csv_sync of the reference's unmodified app/file_parser.py when
    oracle/_ref holds it (oracle/build_ref.py puts it there), else the oracle port of the same lines."""
    if rank != 0:
        return
    import multiprocessing as mp
    from oracle import ref_runner
    kind = "reference" if ref_runner.available() else "port"
    cores = os.cpu_count() or 1
    sample_rows = max(1000, min(args.rows, args.ref_sample_rows))
    doc = make_doc(sample_rows, 0)
    with mp.get_context("fork").Pool(cores) as pool:
        for _ in range(args.warmup):
            pool.map(_ref_one, [(doc, sample_rows, kind)] * cores)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_ref_one, [(doc, sample_rows, kind)] * cores)
        dt = time.perf_counter() - t0
    mb = len(doc) * cores * args.steps / 1e6
    val = mb / dt
    what = ("the reference's own FileParser._parse_csv_sync (oracle/_ref/app/file_parser.py, unmodified)"
            if kind == "reference" else "oracle port of FileParser._parse_csv_sync over pandas")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic", "config": config_dict(args, args.rows),
        "cells_per_s": sample_rows * 8 * cores * args.steps / dt,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{cores} processes x one {sample_rows}-row slice of the C2 CSV per step ({what})"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def _ref_one(a):
    doc, rows, kind = a
    if kind == "reference":
        from oracle import ref_runner
        return len(ref_runner.parse_csv_bytes(doc, rows))
    from oracle import file_parser_oracle as O
    return len(O.parse_csv_bytes(doc, rows))


def config_dict(args, rows):
    return {"workload": f"C2: synthetic RFQ CSV, {rows} rows x 8 cols (part_number, description[quoted, "
                        "embedded , and \"\"], quantity, target_price, currency, delivery_date, supplier, "
                        "notes[80% empty]), one document per GPU, FileParser(max_file_size_mb=4096, "
                        f"max_excel_rows={rows})",
            "rows": rows, "cols": 8, "l2": "flushed between timed steps (256 MiB device write)",
            "reference_arm": f"--impl reference runs {args.ref_sample_rows}-row slices of the same CSV, one per host "
                             "core per step (pandas' to_string is super-linear in rows: the slice favours the CPU)",
            "parallelism": f"documents sharded across {args.gpus} GPU(s), no collective"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--ref-sample-rows", type=int, default=20_000)
    ap.add_argument("--cpu-sample-rows", type=int, default=100_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config-scale", type=float, default=1.0,
                    help="document counts of the C3 / C4 / C5 entries relative to BASELINE.json (0 = skip them)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from replisense_rfq_b200 import Engine, build
    from replisense_rfq_b200.engine import PinnedArena

    # stdout must carry exactly ONE JSON line: anything libraries print while we work (NCCL prints
    # its version banner on stdout) goes to stderr until the result is ready
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)

    if rank == 0 or world == 1:
        build.build()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    build.build()
    eng = Engine(local_rank)
    # the kernels and the timing events must share ONE stream: a non-default torch stream (the
    # legacy default stream has handle 0, which rfq_set_stream reads as "use the ctx's own")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)

    doc = make_doc(args.rows, seed=rank)
    nbytes = len(doc)
    arena = PinnedArena(eng, nbytes + 4096)
    offsets, lengths, _ = arena.pack([doc])
    mr = np.array([args.rows], dtype=np.int64)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- parity gate, EVERY rank, before any timing counts: the SHA-256 of this rank's full output
    # against the hash the oracle produced for its document (tests/golden/c2_sha256.json, generator
    # tests/golden/make_c2_hashes.py); other sizes: the whole text against the oracle on a capped size
    import hashlib
    from oracle import file_parser_oracle as O
    gold = json.loads((ROOT / "tests" / "golden" / "c2_sha256.json").read_text())
    if args.rows == gold["rows"] and str(rank) in gold["seeds"]:
        g = gold["seeds"][str(rank)]
        assert hashlib.sha256(doc).hexdigest() == g["input_sha256"], "bench input differs from the hashed corpus"
        r0 = eng.csv_parse_arena(arena, offsets, lengths, mr)
        assert int(r0.status[0]) == 0, "parity gate: document not parsed"
        assert hashlib.sha256(r0.doc_bytes(0)).hexdigest() == g["output_sha256"], f"parity gate: rank {rank} output differs from the oracle's"
        parity = f"sha256 of the full {g['output_bytes']}-byte output == oracle's (seed {rank})"
    else:
        nsmall = min(args.rows, 20000)
        small = make_doc(nsmall, seed=rank)
        assert eng.parse_csv_many([small], nsmall)[0] == O.parse_csv_bytes(small, nsmall), "parity gate"
        parity = f"text of a {nsmall}-row document == oracle's (no committed hash for {args.rows} rows / seed {rank})"

    # ---- device-resident: stage once, time K runs with CUDA events, L2 flushed between steps
    eng.stage(arena, offsets, lengths)
    for _ in range(args.warmup):
        eng.csv_run(1, mr)
    res = eng.fetch()
    assert int(res.status[0]) == 0
    out_bytes = int(res.lengths[0])
    n_rows_out = int(res.rows[0])
    eng.set_profile(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    l0 = eng.launch_count()
    step_ms = []
    prof_acc = {}
    for _ in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.csv_run(1, mr)
        e1.record()
        torch.cuda.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        eng.lib.rfq_fetch(eng.ctx)  # finalises the per-kernel profile of this run (outside the timing)
        for name, ms in eng.profile():
            prof_acc[name] = prof_acc.get(name, 0.0) + ms
    launches = eng.launch_count() - l0
    barrier()
    eng.set_profile(False)
    t_dev = sum(step_ms) / 1e3
    tt = torch.tensor([t_dev], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_dev = float(tt.item())

    # ---- end to end through the C-ABI call with host (pinned) buffers
    for _ in range(2):
        eng.csv_parse_arena(arena, offsets, lengths, mr)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = eng.csv_parse_arena(arena, offsets, lengths, mr)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    te = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    t_e2e = float(te.item())
    # ---- the same call with two in flight (EnginePool: two ctx / streams on this GPU): the H2D copy
    # of one step overlaps the kernels and the D2H copy of the other (PCIe is full duplex).  Every
    # step still uploads its input from pinned host memory and downloads its whole result.
    from replisense_rfq_b200.engine import EnginePool
    # measured on B200 (C2, every engine warmed): depth 2 -> 3.98, 3 -> 3.54, 4 -> 3.25, 6 -> 3.09 ms per step;
    # the PCIe floor for 86.8 MB up + 135.1 MB down at once is 2.71 ms (tools/pcie_probe.py)
    depth = int(os.environ.get("RFQ_POOL_DEPTH", "4"))
    pool = EnginePool(local_rank, depth=depth)

    def one_step(e, _):
        r = e.csv_parse_arena(arena, offsets, lengths, mr)
        return int(r.lengths[0])
    for e in pool.engines:   # every engine twice: its buffers are sized on the first call
        for _ in range(2):
            one_step(e, 0)
    pool.map(one_step, range(2 * depth))
    barrier()
    t0 = time.perf_counter()
    outs = pool.map(one_step, range(args.steps))
    torch.cuda.synchronize()
    t_pipe = time.perf_counter() - t0
    assert all(o == out_bytes for o in outs)
    tp = torch.tensor([t_pipe], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tp, op=dist.ReduceOp.MAX)
    t_pipe = float(tp.item())
    # ---- e2e_prompt: what /upload/ consumes (app/main.py:243-273): the prompt and text_length; raw_text is
    # parsed and wrapped on the device and never crosses PCIe (D2H drops from the whole text to the prompt)
    from replisense_rfq_b200 import prompt as Pm
    p_head = (Pm.EXTRACTION_PROMPT_TEMPLATE + '\n"""\n').encode()
    p_tail, p_mark = Pm.PROMPT_TAIL.encode(), Pm.TRUNCATION_MARK.encode()

    def prompt_step(e, _):
        r = e.csv_prompt_arena(arena, offsets, lengths, mr, p_head, p_tail, p_mark, Pm.MAX_PROMPT_CHARS)
        return int(r.lengths[0]), int(r.raw_chars[0])
    for e in pool.engines:
        for _ in range(2):
            prompt_step(e, 0)
    barrier()
    t0 = time.perf_counter()
    pouts = pool.map(prompt_step, range(args.steps))
    torch.cuda.synchronize()
    t_prompt = time.perf_counter() - t0
    tq = torch.tensor([t_prompt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tq, op=dist.ReduceOp.MAX)
    t_prompt = float(tq.item())
    prompt_bytes, raw_chars = pouts[0]
    pool.close()
    # ---- what the host side of this box can move: every rank copies a step's input up and a step's output
    # down at the same time (two streams, pinned memory), all ranks at once; max over ranks.  e2e cannot be
    # faster than `both_ms`, e2e_prompt not faster than `h2d_ms`.
    def _copy_probe(h2d, d2h, reps=8):
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if h2d:
                with torch.cuda.stream(ps1):
                    p_di.copy_(p_hi, non_blocking=True)
            if d2h:
                with torch.cuda.stream(ps2):
                    p_ho.copy_(p_do, non_blocking=True)
        torch.cuda.synchronize()
        tt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())
    p_hi = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    p_ho = torch.empty(out_bytes, dtype=torch.uint8).pin_memory()
    p_di = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    p_do = torch.empty(out_bytes, dtype=torch.uint8, device="cuda")
    ps1, ps2 = torch.cuda.Stream(), torch.cuda.Stream()
    _copy_probe(True, True, 2)
    pc_h2d, pc_d2h, pc_both = _copy_probe(True, False), _copy_probe(False, True), _copy_probe(True, True)
    del p_hi, p_ho, p_di, p_do
    sampler.stop_flag = True
    sampler.join()
    # and once more all the way to a Python str (what FileParser._parse_csv_sync returns)
    t0 = time.perf_counter()
    res_ = eng.csv_parse_arena(arena, offsets, lengths, mr)
    t1 = time.perf_counter()
    txt = res_.doc_str(0)
    t_str = time.perf_counter() - t0
    t_str_only = time.perf_counter() - t1
    assert txt.isascii() == (len(txt) == out_bytes)
    del txt
    # e2e_python: file bytes -> str through the reference-facing method FileParser._parse_csv_sync
    # (file read, pack into pinned memory, H2D, kernels, D2H, str)
    from replisense_rfq_b200 import FileParser
    import tempfile
    tdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    fd, fpath = tempfile.mkstemp(suffix=".csv", dir=tdir)
    with os.fdopen(fd, "wb") as f:
        f.write(doc)
    fp = FileParser(max_file_size_mb=4096, max_excel_rows=args.rows, engine=eng)
    fp._parse_csv_sync(Path(fpath))
    t_py = []
    for _ in range(3):
        t0 = time.perf_counter()
        out_txt = fp._parse_csv_sync(Path(fpath))
        t_py.append(time.perf_counter() - t0)
        n_txt = len(out_txt)
        del out_txt
    os.unlink(fpath)
    t_py = min(t_py)

    # ---- BASELINE configs C3 / C4 / C5 at their stated sizes (extra keys; strong scaling over the ranks)
    configs = None
    if args.config_scale > 0:
        arena.free()
        del flush
        torch.cuda.empty_cache()
        from tools import bench_configs
        hbm_peak, _ = peaks()
        configs = bench_configs.run_all(eng, torch, dist if world > 1 else None, rank, world, hbm_peak, args.config_scale)
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    total_mb = nbytes * world * args.steps / 1e6
    value = total_mb / t_dev
    hbm, hbm_src = peaks()
    # dominant kernel (live CUDA-event timing on the launching stream, averaged per launch)
    kern = {k: v / args.steps for k, v in prof_acc.items() if k.endswith("_kernel")}
    dom = max(kern, key=kern.get)
    alg = {"csv2_emit_kernel": nbytes + out_bytes, "csv2_type_kernel": nbytes, "csv2_scan_kernel": nbytes,
           "csv_emit_kernel": nbytes + out_bytes, "csv_measure_kernel": nbytes,
           "csv_classify_kernel": nbytes, "csv_tilefn_kernel": nbytes, "csv_count_kernel": nbytes,
           "csv_index_kernel": nbytes}.get(dom, nbytes)
    achieved = alg / (kern[dom] / 1e3) / 1e9
    traffic = None
    tf = ROOT / "profiles" / "traffic_r02.json"
    if tf.exists() and args.rows == 1_000_000:
        traffic = json.loads(tf.read_text()).get(dom)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_dict(args, args.rows),
        "cells_per_s": args.rows * 8 * world * args.steps / t_dev,
        "input_bytes_per_doc": nbytes, "output_bytes_per_doc": out_bytes, "rows_out": n_rows_out,
        "pipeline_algorithmic_GBps": (nbytes + out_bytes) * world * args.steps / t_dev / 1e9,
        "pipeline_frac_of_hbm": (nbytes + out_bytes) * args.steps / t_dev / 1e9 / hbm,
        # headline e2e: rfq_csv_parse with host buffers, two calls in flight (EnginePool depth 2);
        # "serial_*" = one call at a time (its latency)
        "e2e": {"value": total_mb / t_pipe, "unit": UNIT, "h2d_bytes_per_step": nbytes,
                "d2h_bytes_per_step": out_bytes + 64, "ms_per_step": t_pipe / args.steps * 1e3,
                "in_flight": depth, "serial_value": total_mb / t_e2e,
                "serial_ms_per_step": t_e2e / args.steps * 1e3, "to_python_str_ms": t_str * 1e3, "str_build_only_ms": t_str_only * 1e3,
                "frac_of_copy_ceiling": pc_both / (t_pipe / args.steps)},
        # the copies alone on this box, all ranks at once (max over ranks): what e2e / e2e_prompt are bounded by
        "copy_ceiling": {"h2d_ms": pc_h2d * 1e3, "d2h_ms": pc_d2h * 1e3, "both_ms": pc_both * 1e3,
                         "h2d_GBps_per_gpu": nbytes / pc_h2d / 1e9, "d2h_GBps_per_gpu": out_bytes / pc_d2h / 1e9,
                         "both_GBps_all_gpus": (nbytes + out_bytes) * world / pc_both / 1e9,
                         "what": "pinned H2D of one step's input and D2H of one step's output on two streams, "
                                 "every rank at the same time"},
        # the same document through the Python method the reference's callers use (file -> str), this rank
        "e2e_python": {"value": nbytes / 1e6 / t_py, "unit": UNIT, "ms_per_call": t_py * 1e3, "chars": n_txt,
                       "call": "FileParser(max_file_size_mb=4096, max_excel_rows=rows)._parse_csv_sync(path)"},
        # parse + prompt assembly fused on the device (rfq_csv_prompt_parse): H2D of the file, D2H of the prompt only
        "e2e_prompt": {"value": total_mb / t_prompt, "unit": UNIT, "ms_per_step": t_prompt / args.steps * 1e3,
                       "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": prompt_bytes + 64, "in_flight": depth,
                       "text_length": raw_chars, "call": "rfq_csv_prompt_parse (C ABI), host buffers",
                       "frac_of_copy_ceiling": pc_h2d / (t_prompt / args.steps)},
        "parity_gate": parity,
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm, "unit": "GB/s",
                     "frac": achieved / hbm, "traffic": traffic, "peak_source": hbm_src,
                     "algorithmic_bytes_per_launch": alg, "kernel_ms": kern[dom]},
        "kernel_ms_per_step": {k: round(v / args.steps, 4) for k, v in sorted(prof_acc.items(), key=lambda x: -x[1])},
        "clocks": sampler.summary(),
    }
    if configs is not None:
        line["configs"] = configs
    if not args.no_cpu_baseline:
        rows = min(args.cpu_sample_rows, args.rows)
        sample = make_doc(rows, seed=0)
        from oracle import ref_runner
        use_ref = ref_runner.available()          # oracle/_ref: the reference's own file, travelled with the snapshot
        t0 = time.perf_counter()
        if use_ref:
            ref_runner.parse_csv_bytes(sample, rows)
        else:
            O.parse_csv_bytes(sample, rows)
        dt = time.perf_counter() - t0
        what = ("the reference's own FileParser._parse_csv_sync (oracle/_ref/app/file_parser.py, unmodified)" if use_ref
                else "the oracle port of FileParser._parse_csv_sync (pandas read_csv + to_string)")
        line["cpu_baseline"] = {"value": len(sample) / 1e6 / dt, "unit": UNIT, "cores": 1,
                                "kind": "reference" if use_ref else "port",
                                "sample": f"first {rows} rows of the C2 CSV through {what}, "
                                          f"{dt:.2f} s on one host core of {os.cpu_count()}",
                                "cells_per_s": rows * 8 / dt}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    print(json.dumps(line), flush=True)
    if world > 1:
        os.dup2(2, 1)
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
