#!/usr/bin/env python
"""bench.py — the contract benchmark of the B200 hash join (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One STEP = one whole join of the workload: build side pushed and built, probe side probed in chunks,
(build_idx, probe_idx) pairs materialised.  N = 1 runs config C2 (100 M-row int64 unique build x 1 B-row
uniform probe, 1:1 matches); N > 1 (launched by torchrun, one rank per GPU) runs the same per-GPU volume on
every rank over an N-times larger key space: both sides are hash-partitioned on the device and exchanged
over NVLink, then every GPU joins its own partition ("weak" scaling, C4's shape).

`value`      whole-job (build + probe) tuples/s, inputs resident in HBM, CUDA events on the join's stream,
             max over ranks.
`e2e`        same metric through the public host-buffer API (the call the Rust operator makes): every step
             copies the key buffers from pinned host memory and reads the index pairs back to pinned host memory.
`roofline`   the dominant kernel (probe), algorithmic bytes / its CUDA-event launch time / measured HBM peak.
`cpu_baseline`  oracle/cpu_join_mt.c (C restatement of the reference's version10 algorithm, all host threads)
             on a bounded sample of the same workload (rank 0, N = 1 only).
`--impl reference`  times that CPU implementation alone (the Rust crate cannot be built here: no cargo).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "build+probe tuples/sec"
UNIT = "tuples/s"
HBM_FALLBACK_GBS = 6650.0          # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
NVLINK_GBS = 770.0                 # measured peer copy per direction (B200_PROFILING.md)

# algorithmic bytes per tuple (SURVEY.md §8d / BASELINE.md §3)
B_BUILD = 24                       # 8 key read + 16 slot write
B_PROBE = 36                       # 8 key read + 16 slot read + 12 pair write
B_PROBE_KERNEL_RADIX = 24          # the probe KERNEL of the radix layout: 12 (key, row id) read + 12 pair write, table in L2
B_PROBE_DENSE = 16                 # dense probe kernel: 8 key read + 8 build-index write (row ids come from the partition pass)
B_PARTITION = 28                   # histogram 8 read + scatter 8 read + 12 write


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--build-rows", type=int, default=100_000_000)
    ap.add_argument("--probe-rows", type=int, default=1_000_000_000)
    ap.add_argument("--chunk-rows", type=int, default=1 << 30, help="probe rows per b200j_probe_batch call")
    ap.add_argument("--layout", default="auto")
    ap.add_argument("--radix-bits", type=int, default=0)
    ap.add_argument("--load-factor", type=float, default=0.0)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-probe-sample", type=int, default=250_000_000)
    ap.add_argument("--stock-scale", type=int, default=10, help="the Acero (stock HashJoinExec stand-in) baseline joins 1/this of the workload")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--overlap-build", type=int, default=0, help="N>1: 1 = exchange the probe side while the local table is built")
    ap.add_argument("--e2e-wide", action="store_true", help="e2e returns (u64, u32) pairs instead of B200J_OUT_BUILD_U32's (u32, u32)")
    ap.add_argument("--e2e-pairs", action="store_true", help="e2e returns index pairs instead of the probe-ordered build-index array (B200J_OUT_PROBE_ORDER)")
    ap.add_argument("--no-gather", action="store_true")
    ap.add_argument("--host-chunk-rows", type=int, default=1 << 27, help="N>1 e2e: probe rows per pipelined host chunk (DistributedJoin.probe_host_chunks)")
    ap.add_argument("--exchange-chunk-rows", type=int, default=0, help="N>1: > 0 = probe rows per exchange / local-join pipeline stage (DistributedJoin.probe_chunks)")
    ap.add_argument("--shuffle", default="pull", choices=["p2p", "nccl", "ce", "pull"], help="N>1: pull = single-pass local scatter, the receivers' partition pass pulls the regions over NVLink (default); p2p = fused peer-store scatter; nccl = NCCL all-to-all; ce = copy-engine transfers")
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------------
# the reference arm / cpu_baseline: oracle/cpu_join_mt.c on the host cores
# --------------------------------------------------------------------------------------------------
def cpu_join_sample(n_build: int, n_probe_full: int, probe_sample: int, reps: int = 1, warmup: int = 0):
    """Times the CPU join: full-size build, `probe_sample` probe rows; the whole-job figure scales the probe
    time to n_probe_full rows (the probe loop is a stream of independent 8192-row batches)."""
    import numpy as np
    import oracle
    oracle.build()
    bk = oracle.gen_permutation_i64(n_build, 3)
    pk = oracle.gen_uniform_i64(min(probe_sample, n_probe_full), n_build, 4)
    # all host threads this process may run on -- NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to its workers,
    # which would quietly turn the N > 1 reference arm into a single-threaded run
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or oracle.MtJoin.max_threads()
    out = []
    for it in range(warmup + reps):
        j = oracle.MtJoin(bk, threads=cores)
        tb = j.build_seconds
        tp, matched, _ = j.probe(pk)
        assert matched == pk.shape[0], (matched, pk.shape[0])
        del j
        if it >= warmup:
            out.append((tb, tp))
    tb = sum(x[0] for x in out) / len(out)
    tp = sum(x[1] for x in out) / len(out)
    build_tps = n_build / tb
    probe_tps = pk.shape[0] / tp
    whole = (n_build + n_probe_full) / (n_build / build_tps + n_probe_full / probe_tps)
    return {"value": whole, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": (f"full {n_build}-row build + the full {n_probe_full}-row probe side" if pk.shape[0] == n_probe_full else
                       f"PROJECTED from a {100.0 * pk.shape[0] / n_probe_full:.0f} % probe sample: full {n_build}-row build timed, "
                       f"{pk.shape[0]} of the {n_probe_full} probe rows timed and the probe time scaled to the full side")
                      + f"; oracle/cpu_join_mt.c, OpenMP x {cores} threads",
            "build_tuples_per_s": build_tps, "probe_tuples_per_s": probe_tps,
            "build_s": tb, "probe_sample_s": tp}, tb + tp


def acero_join_sample(n_build: int, n_probe_full: int, scale: int = 10):
    """Stand-in for DataFusion's stock HashJoinExec (the `None` join replacement, src/parse_sql.rs:37-54), which cannot be built
    here: pyarrow Acero's multi-threaded hash join on the same key distributions at 1/scale of the workload (inner join, one
    int64 payload per side, output table materialised)."""
    import numpy as np
    import pyarrow as pa
    import oracle
    nb, npr = max(1, n_build // scale), max(1, n_probe_full // scale)
    bk = oracle.gen_permutation_i64(nb, 3)
    pk = oracle.gen_uniform_i64(npr, nb, 4)
    bt = pa.table({"k": bk, "b_row": np.arange(nb, dtype=np.int64)})
    pt = pa.table({"k": pk, "p_row": np.arange(npr, dtype=np.int64)})
    t0 = time.perf_counter()
    out = pt.join(bt, keys="k", join_type="inner", use_threads=True)
    dt = time.perf_counter() - t0
    assert out.num_rows == npr, (out.num_rows, npr)
    return {"value": (nb + npr) / dt, "unit": UNIT, "cores": pa.cpu_count(), "kind": "stand-in",
            "engine": f"pyarrow {pa.__version__} Acero Table.join (stock HashJoinExec stand-in: DataFusion is not buildable here, no cargo)",
            "sample": f"1/{scale} scale of the workload: {nb}-row unique int64 build x {npr}-row uniform probe, inner join, output materialised",
            "seconds": dt}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    res, _ = cpu_join_sample(args.build_rows, args.probe_rows, args.cpu_probe_sample, reps=args.steps, warmup=args.warmup)
    per_step = (time.time() - t0) / max(1, args.steps + args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": res,
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference's own Rust crate is not buildable here (no cargo, nightly, aarch64-only version10): "
                "this is the C restatement of its version10 algorithm on all host threads",
    }
    print(json.dumps(line))


def workload_config(args, n_gpus):
    return {"workload": f"C2: {args.build_rows}-row int64 unique-key build x {args.probe_rows}-row uniform probe, 1:1 matches, "
                        f"{'one int64 build payload column, ' if n_gpus == 1 else ''}inner join, (build_idx u64, probe_idx u32) pairs out" + (f"; per GPU, {n_gpus} GPUs, key space x{n_gpus}, "
                        f"hash-partition + NVLink exchange of both sides (mode {args.shuffle})" if n_gpus > 1 else ""),
            "build_rows_per_gpu": args.build_rows, "probe_rows_per_gpu": args.probe_rows,
            "probe_chunk_rows": args.chunk_rows if (n_gpus == 1 or args.exchange_chunk_rows <= 0) else min(args.chunk_rows, args.exchange_chunk_rows),
            "l2": "inputs (8 GB probe keys per step) are far larger than the 126 MB L2; no flush needed"}


# --------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import datafusion_parallelism_b200 as d

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if d.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the join has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    n_build, n_probe, chunk = args.build_rows, args.probe_rows, args.chunk_rows
    hbm_gbs, peak_src = measured_peaks()
    stream = torch.cuda.Stream(device=dev)

    # ---- synthetic inputs, generated on the device (SURVEY §8d C2 / C4 shapes)
    g = torch.Generator(device=dev)
    g.manual_seed(3 + 1000 * rank)
    key_space = n_build * world
    if world == 1:
        bk = torch.randperm(n_build, device=dev, generator=g, dtype=torch.int64)
    else:
        # rank r owns the build keys congruent to r modulo world (unique across ranks), in random order
        bk = torch.randperm(n_build, device=dev, generator=g, dtype=torch.int64) * world + rank
    pk = torch.randint(0, key_space, (n_probe,), device=dev, generator=g, dtype=torch.int64)
    torch.cuda.synchronize()

    if world > 1:
        from datafusion_parallelism_b200.exchange import DistributedJoin
        join = DistributedJoin("inner", key_width=8, stream=stream, layout=args.layout, radix_bits=args.radix_bits,
                               load_factor=args.load_factor, expected_build_rows=int(n_build * 1.1), mode=args.shuffle, timing=True)
    else:
        # the build side carries one int64 payload column (here: the row number, so the gather can be checked)
        join = d.HashJoin("inner", key_widths=(8,), build_payload_types=(8,), layout=args.layout, radix_bits=args.radix_bits,
                          load_factor=args.load_factor, expected_build_rows=n_build, stream=stream.cuda_stream, timing=True)
    bpay = torch.arange(n_build, device=dev, dtype=torch.int64) if world == 1 else None

    def step_device():
        """one whole join, inputs resident in HBM; returns matched rows"""
        join.reset()
        if world == 1:
            join.build_push(0, [bk], [bpay])
        else:
            join.build_push(0, [bk])
        if world > 1 and args.overlap_build and args.exchange_chunk_rows <= 0 and chunk >= n_probe:
            # the probe side crosses NVLink while the local table is being built
            token = join.probe_exchange_begin([pk])
            join.build_finish(0)
            r = join.probe_batch_received(0, token, want=d.OUT_PAIRS)
            tot = r.n
            r.release()
            return tot
        join.build_finish(0)
        tot = 0
        if world > 1 and args.exchange_chunk_rows > 0:
            # optional: sub-chunks, the NVLink exchange of chunk k+1 overlapping the local join of chunk k. OFF by default:
            # measured on 2 x B200 it is slower (35.4 -> 50.7 ms/step at 128 M-row chunks) because the exchange kernels'
            # streams evict the radix partition's table from L2 under the probe kernel (ms_probe 7.0 -> 19 ms)
            sub = min(chunk, args.exchange_chunk_rows)
            for r in join.probe_chunks(0, [pk[lo:lo + sub] for lo in range(0, n_probe, sub)], want=d.OUT_PAIRS):
                tot += r.n
                r.release()
            return tot
        for lo in range(0, n_probe, chunk):
            r = join.probe_batch(0, [pk[lo:lo + chunk]], want=d.OUT_PAIRS)
            tot += r.n
            r.release()
        return tot

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    expect = None
    for _ in range(args.warmup):
        tot = step_device()
        expect = tot
    if world == 1 and expect is not None:
        assert expect == n_probe, f"1:1 workload produced {expect} rows for {n_probe} probe rows"
    verified = None
    if not args.no_verify:
        verified = (verify_single(args, d, torch, dev, join, bk, pk) if world == 1 else
                    verify_distributed(args, d, torch, dist, dev, join, bk, pk, rank, world))
        torch.cuda.empty_cache()
        step_device()                   # back to the steady state the warm-up left (block cache, L2)

    join.reset_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        step_device()
    ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    st = join.stats()
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = (n_build + n_probe) * world / (ms_step * 1e-3)

    # ---- phases (device time from the library's own CUDA events, per step)
    K = args.steps
    ph = {k: st[k] / K for k in ("ms_build_copy", "ms_build_partition", "ms_build_insert", "ms_probe_partition", "ms_probe")}
    for k in ("ms_exchange_build", "ms_exchange_probe"):
        if k in st:
            ph[k] = st[k] / K
    if world > 1:
        ph["ms_exchange_kernels"] = st["ms_exchange_partition"] / K       # the partition / scatter kernels inside the two exchange phases
    build_ms = ph["ms_build_copy"] + ph["ms_build_partition"] + ph["ms_build_insert"]
    probe_ms = ph["ms_probe_partition"] + ph["ms_probe"]
    radix = st["layout"] == 2
    probe_launch_ms = st["ms_probe"] / max(1, st["probe_launches"])
    rows_per_launch = n_probe * K / max(1, st["probe_launches"]) if world == 1 else st["probe_rows"] / max(1, st["probe_launches"])
    dense = st.get("dense_probe_launches", 0) == st["probe_launches"] and st["probe_launches"] > 0
    # the table term of the ALGORITHMIC bytes is counted at load factor 0.5 (32 B per distinct build key) whatever the table's
    # actual size: spreading the table thinner must not raise the roofline fraction
    table_alg = min(st["table_bytes"], 32 * max(1, st["distinct_keys"]) + 4096)
    if dense:        # 8 B key in + 8 B build index out per row, plus ONE pass over the table per launch (it is L2-resident per partition)
        seg = st.get("seg_probe_launches", 0) == st["probe_launches"]
        narrow = seg and st.get("narrow_probe_launches", 0) == st["probe_launches"]
        # ALGORITHMIC bytes: 8 B key in + 8 B build index out per row. (When every key fits 32 bits the partitioned stream carries
        # them as 4-byte values and the kernel moves 12 B per row; the algorithmic figure does not shrink with that.)
        bpt = B_PROBE_DENSE
        kname = ("probe_dense_seg_kernel<4,false,false,false>" if narrow else "probe_dense_seg_kernel<8,false,false,false>") if seg else "probe_dense_kernel<8,false,true,false>"
        table_b = table_alg
    elif radix:
        bpt, kname, table_b = B_PROBE_KERNEL_RADIX, "probe_fast_kernel<8,PM_PAIRS,false,true>", table_alg
    else:
        bpt, kname, table_b = B_PROBE, "probe_fast_kernel<8,PM_PAIRS,false,false>", 0
    bytes_per_launch = rows_per_launch * bpt + table_b
    achieved = bytes_per_launch / (probe_launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": achieved, "peak": hbm_gbs, "unit": "GB/s", "frac": achieved / hbm_gbs, "traffic": None,
                "peak_source": peak_src, "bytes_per_tuple": bpt, "stream_key_bytes": (4 if (dense and st.get("narrow_probe_launches", 0) == st["probe_launches"]) else 8),
                "table_bytes_per_launch": table_b,
                "tuples_per_launch": rows_per_launch, "ms_per_launch": probe_launch_ms}
    # DRAM bytes of that kernel per launch from the committed ncu --set full capture of THIS build (scripts/make_traffic_json.py);
    # only quoted when the capture is of the kernel that ran and of the same number of rows
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            with open(tr) as f:
                tj = json.load(f)
            if tj.get("kernel") == kname and abs(tj.get("rows_per_launch", 0) - rows_per_launch) <= 0.01 * rows_per_launch:
                roofline["traffic"] = tj.get("probe_kernel_dram_bytes_per_launch")
                roofline["traffic_source"] = tj.get("source")
        except Exception:
            pass
    local_build = st["build_rows"]
    phases = {
        "ms_per_step": ph,
        "build_tuples_per_s": local_build / (build_ms * 1e-3) if build_ms > 0 else None,
        "probe_tuples_per_s": (st["probe_rows"] / K) / (probe_ms * 1e-3) if probe_ms > 0 else None,
        "build_roofline_frac_24B": local_build * B_BUILD / (build_ms * 1e-3) / 1e9 / hbm_gbs if build_ms > 0 else None,
        "probe_roofline_frac_36B": (st["probe_rows"] / K) * B_PROBE / (probe_ms * 1e-3) / 1e9 / hbm_gbs if probe_ms > 0 else None,
        "layout": {1: "direct", 2: "radix"}.get(st["layout"], st["layout"]), "radix_bits": st["radix_bits"], "table_bytes": st["table_bytes"],
    }
    gpu_launches = int(st["kernel_launches"])

    # ---- output gather (a6, take_multiple_record_batch) as its own phase: build key + build payload at build_idx, probe key at
    #      probe_idx, device to device, for the whole C2 result ("reported as its own phase")
    if world == 1 and not args.no_gather:
        phases["gather"] = run_gather_phase(args, d, torch, dev, join, bk, bpay, pk, hbm_gbs)

    # ---- e2e: host buffers through the public API
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, d, torch, dist, dev, rank, world, bk, pk, join, barrier)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        del bk, pk
        torch.cuda.empty_cache()
        cpu, _ = cpu_join_sample(n_build, n_probe, args.cpu_probe_sample, reps=1, warmup=0)
        try:
            cpu["stock"] = acero_join_sample(n_build, n_probe, args.stock_scale)
        except Exception as exc:          # a host without the RAM for the 1/10-scale tables: say so, keep the line
            cpu["stock"] = {"unavailable": repr(exc)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches,
            "roofline": roofline, "cpu_baseline": cpu, "phases": phases,
            "verified": bool(verified and verified.get("ok")), "verification": verified,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def verify_single(args, d, torch, dev, join, bk, pk):
    """Full-size check of one whole join, outside the timed region: every (build_idx, probe_idx) pair joins equal keys, every
    probe row appears exactly once (1:1 workload), the pair count is the probe row count."""
    n_build, n_probe, chunk = args.build_rows, args.probe_rows, args.chunk_rows
    join.reset()
    join.build_push(0, [bk], [torch.arange(n_build, device=dev, dtype=torch.int64)])
    join.build_finish(0)
    total, S = 0, 1 << 27
    for lo in range(0, n_probe, chunk):
        m = min(chunk, n_probe - lo)
        r = join.probe_batch(0, [pk[lo:lo + m]], want=d.OUT_PAIRS)
        assert r.n == m, f"1:1 workload: {r.n} pairs for {m} probe rows"
        b_, p_ = r.device_arrays()
        tb, tp = torch.as_tensor(b_, device=dev), torch.as_tensor(p_, device=dev)
        psum = 0
        for s0 in range(0, r.n, S):
            bb = tb[s0:s0 + S]
            pp = tp[s0:s0 + S].long() & 0xFFFFFFFF
            assert int(bb.min()) >= 0 and int(bb.max()) < n_build and int(pp.max()) < m, "index out of range"
            assert torch.equal(bk[bb], pk[lo:lo + m][pp]), "a pair joins two different keys"
            psum += int(pp.sum())
            del bb, pp
        assert psum == m * (m - 1) // 2, "probe rows are not covered exactly once"
        total += r.n
        r.release()
    assert total == n_probe
    return {"ok": True, "pairs_checked": total, "checks": ["count == probe rows", "build_key[build_idx] == probe_key[probe_idx] on every pair",
                                                           "sum(probe_idx) == n(n-1)/2"]}


def verify_distributed(args, d, torch, dist, dev, join, bk, pk, rank, world):
    """N > 1, outside the timed region: (1) the pair counts of all ranks add up to world x probe rows; (2) the exchange lost and
    invented nothing: count and wrapping sum of the keys sent == of the keys received, over all ranks; (3) every received key
    hashes to this rank (b200j_hash_keys on a sample); (4) on a 2^24-row sample of this rank's result, the received build key
    at build_idx equals the received probe key at probe_idx."""
    from datafusion_parallelism_b200.exchange import _RawCudaArray
    n_build, n_probe = args.build_rows, args.probe_rows
    join.reset()
    join.build_push(0, [bk])
    join.build_finish(0)
    r = join.probe_batch(0, [pk], want=d.OUT_PAIRS)
    n_recv_b = int(join.build.plan.recv_rows[rank])
    n_recv_p = int(join.probe.plan.recv_rows[rank])
    rb = join.received(join.build, "keys")          # local copies of what this rank received (one piece per source in mode "pull")
    rp = join.received(join.probe, "keys")
    assert rb.shape[0] == n_recv_b and rp.shape[0] == n_recv_p
    t = torch.tensor([r.n, n_recv_b, n_recv_p, int(bk.sum()), int(rb.sum()), int(pk.sum()), int(rp.sum())], device=dev, dtype=torch.int64)
    dist.all_reduce(t)          # int64 sums wrap: a checksum of checksums
    tot, nb_, np_, sb, srb, sp, srp = [int(x) for x in t.tolist()]
    assert tot == n_probe * world, f"{tot} pairs over all ranks for {n_probe * world} probe rows"
    assert nb_ == n_build * world and np_ == n_probe * world, "rows lost or duplicated by the exchange"
    assert sb == srb and sp == srp, "key checksum differs between what was sent and what was received"
    S = min(1 << 22, n_recv_p)
    if S:
        h = torch.empty(S, dtype=torch.int64, device=dev)
        join.join.hash_keys_device([rp[:S]], h.data_ptr())
        torch.cuda.synchronize()
        dest = (((h >> 32) & 0xFFFFF) * world) >> 20
        assert bool((dest == rank).all()), "a received key does not hash to this rank"
    S = min(1 << 24, r.n)
    if S:
        b_, p_ = r.device_arrays()
        tb = torch.as_tensor(b_, device=dev)[:S]
        tp = torch.as_tensor(p_, device=dev)[:S].long() & 0xFFFFFFFF
        assert torch.equal(rb[tb], rp[tp]), "a pair joins two different keys"
    r.release()
    return {"ok": True, "pairs_all_ranks": tot, "checks": ["sum of pair counts == world x probe rows", "exchange conserves row count and key checksum",
                                                          "received keys hash to the receiving rank (2^22 sample)",
                                                          "build_key[build_idx] == probe_key[probe_idx] (2^24 sample per rank)"]}


def run_gather_phase(args, d, torch, dev, join, bk, bpay, pk, hbm_gbs):
    """b200j_gather, device to device, of the whole C2 result: once on plain pairs and once on pairs probed with
    B200J_OUT_GATHER_HINT (the probe leaves its partitioned keys and the slot of every pair with the result). Every output row is
    checked. The hinted probe's own extra time and the one-off slot-ordered payload copy are reported next to the gather's."""
    from datafusion_parallelism_b200.exchange import _RawCudaArray
    n_build, n_probe = args.build_rows, args.probe_rows
    m = min(n_probe, args.chunk_rows)
    out = {}
    for mode, want in (("plain", d.OUT_PAIRS), ("hinted", d.OUT_PAIRS | d.OUT_GATHER_HINT)):
        join.reset()
        join.build_push(0, [bk], [bpay])
        join.build_finish(0)
        probe_ms = []
        for rep in range(2):                                    # the second probe is the warm one
            s0 = join.stats()
            r = join.probe_batch(0, [pk[:m]], want=want)
            s1 = join.stats()
            probe_ms.append(s1["ms_probe"] - s0["ms_probe"] + s1["ms_probe_partition"] - s0["ms_probe_partition"])
            if rep == 0:
                r.release()
        times, ok = [], None
        for rep in range(3):
            st0 = join.stats()["ms_gather"]
            g = join.gather(r, [pk[:m]], want=0)                     # device in, device out
            times.append(join.stats()["ms_gather"] - st0)
            if rep == 0:
                cols = [g.column_buffers(i) for i in range(g.n_cols)]
                b_, p_ = r.device_arrays()
                ok = g.n_rows == r.n and all(c["null_count"] == 0 for c in cols[:3])
                S = 1 << 27
                for lo in range(0, g.n_rows, S):                    # every output row: keys equal, payload = the build row, probe key = the probe row's
                    k = min(S, g.n_rows - lo)
                    gbk, gbp, gpk = (torch.as_tensor(_RawCudaArray(c["values"] + lo * 8, k, "<i8"), device=dev) for c in cols[:3])
                    tb = torch.as_tensor(b_, device=dev)[lo:lo + k]
                    tp = torch.as_tensor(p_, device=dev)[lo:lo + k].long() & 0xFFFFFFFF
                    ok = ok and bool(torch.equal(gbk, gpk)) and bool(torch.equal(gbp, tb)) and bool(torch.equal(gpk, pk[:m][tp]))
                    del gbk, gbp, gpk, tb, tp
            g.release()
        st = join.stats()
        n_out = r.n
        r.release()
        out[mode] = {"ms": min(times), "ms_all": times, "probe_phase_ms": probe_ms[-1], "verified": ok,
                     "gather_hint_launches": st.get("gather_hint_launches", 0), "gather_seg_launches": st.get("gather_seg_launches", 0),
                     "ms_slot_order_once_per_build": st.get("ms_slot_order", 0.0)}
    hinted = bool(out["hinted"]["gather_hint_launches"])
    ms = out["hinted"]["ms"] if hinted else out["plain"]["ms"]
    # ALGORITHMIC bytes per output row. Plain pairs ("per output row sum(column widths) read + write", plus the pair):
    # 12 B index pair + three 8-byte values read + three written = 60 B. Hinted: what that path has to move -- the row's key in the
    # partitioned stream (4 B while the keys fit 32 bits, as here), its slot (4 B), one 8-byte payload value, three values written
    # = 40 B; the fraction on SURVEY's 60 B is given beside it and exceeds what the kernels really move.
    frac = lambda b, t: n_out * b / (t * 1e-3) / 1e9 / hbm_gbs
    bytes_per_row = 40 if hinted else 60
    return {"ms": ms, "mode": "B200J_OUT_GATHER_HINT" if hinted else "plain pairs", "rows": n_out,
            "columns": "build key i64 + build payload i64 (at build_idx) + probe key i64 (at probe_idx)",
            "bytes_per_row": bytes_per_row, "rows_per_s": n_out / (ms * 1e-3), "gather_roofline_frac": frac(bytes_per_row, ms),
            "frac_on_survey_60B": frac(60, ms), "plain_gather_roofline_frac": frac(60, out["plain"]["ms"]),
            "hint_extra_probe_ms": out["hinted"]["probe_phase_ms"] - out["plain"]["probe_phase_ms"],
            "plain": out["plain"], "hinted": out["hinted"], "verified": bool(out["plain"]["verified"] and out["hinted"]["verified"])}


def verify_host_result(np, r, build_keys, probe_keys, probe_order, sample=1 << 22):
    """Host-side check of one e2e result (N = 1): the row count, and key equality build_key[build_idx] == probe_key[probe_idx] on a
    random sample of the rows (every row of the 1:1 workload has a match)."""
    import ctypes as C
    n = r.n
    if n != probe_keys.shape[0] or r.n_build_null != 0:
        return False
    rng = np.random.default_rng(1)
    idx = rng.integers(0, n, min(sample, n))
    bt = C.c_uint32 if r.build_idx_width == 4 else C.c_uint64
    b = np.ctypeslib.as_array(C.cast(r.build_ptr, C.POINTER(bt)), shape=(n,))
    if probe_order:
        if not r.probe_order or r.probe_ptr:
            return False
        p = idx
    else:
        p = np.ctypeslib.as_array(C.cast(r.probe_ptr, C.POINTER(C.c_uint32)), shape=(n,))[idx].astype(np.int64)
    return bool(np.array_equal(build_keys[b[idx].astype(np.int64)], probe_keys[p]))


def run_e2e(args, d, torch, dist, dev, rank, world, bk, pk, join, barrier):
    """Same join through host buffers: pinned Arrow key buffers in, pinned index pairs out, every step."""
    import numpy as np
    n_build, n_probe, chunk = args.build_rows, args.probe_rows, args.chunk_rows
    # bound pinned memory by what the host has (keys 8 B + pairs 12 B per probe row)
    try:
        with open("/proc/meminfo") as f:
            avail = [int(l.split()[1]) * 1024 for l in f if l.startswith("MemAvailable")][0]
    except Exception:
        avail = 64 << 30
    hb = d.PinnedArray(n_build, np.int64)
    hp = d.PinnedArray(n_probe, np.int64)
    hpay = None
    if world == 1:
        hpay = d.PinnedArray(n_build, np.int64)              # the build side's payload column crosses PCIe as well
        hpay.array[:] = np.arange(n_build, dtype=np.int64)
    torch.from_numpy(hb.array).copy_(bk)
    torch.from_numpy(hp.array).copy_(pk)
    torch.cuda.synchronize()
    steps = max(1, min(args.e2e_steps, args.steps))
    checksum = 0

    host_chunk = min(chunk, args.host_chunk_rows)
    # What comes back over PCIe (N = 1). Default: B200J_OUT_PROBE_ORDER | B200J_OUT_BUILD_U32 -- the build keys are unique, so the result
    # is ONE u32 array in the order of the probe batch (build_idx[i] = the match of probe row i): 4 bytes per probe row.
    # --e2e-pairs: (u32 build_idx, u32 probe_idx) pairs, 8 bytes per pair; --e2e-wide: (u64, u32) pairs, 12 bytes.
    probe_order = world == 1 and not (args.e2e_wide or args.e2e_pairs)
    out_flags = 0 if args.e2e_wide else d.OUT_BUILD_U32
    if probe_order:
        out_flags |= d.OUT_PROBE_ORDER
    pair_bytes = 12 if args.e2e_wide else (4 if probe_order else 8)
    fmt = {12: "(u64 build_idx, u32 probe_idx) pairs", 8: "(u32 build_idx, u32 probe_idx) pairs [B200J_OUT_BUILD_U32]",
           4: "u32 build_idx per probe row, in probe order [B200J_OUT_PROBE_ORDER | B200J_OUT_BUILD_U32]"}[pair_bytes]
    check = {"ok": None}

    def step_host():
        if world > 1:
            with torch.cuda.stream(join.stream):
                return step_host_inner()
        return step_host_inner()

    def step_host_inner():
        nonlocal checksum
        join.reset()
        tot = 0
        if world > 1:
            # N > 1: DistributedJoin's host-batch calls; H2D of chunk k+1 and D2H of chunk k-1's pairs overlap chunk k's exchange + join
            join.build_push_host(0, hb.array)
            join.build_finish(0)
            for r in join.probe_host_chunks(0, hp.array, chunk_rows=host_chunk, want=d.OUT_PAIRS | out_flags):
                tot += r.n
            return tot
        join.build_push(0, [hb.array], [hpay.array])
        join.build_finish(0)
        for lo in range(0, n_probe, chunk):
            r = join.probe_batch(0, [hp.array[lo:lo + chunk]], want=d.OUT_PAIRS | d.OUT_HOST | out_flags)
            tot += r.n
            if check["ok"] is None and lo == 0:
                check["ok"] = verify_host_result(np, r, hb.array, hp.array[lo:lo + chunk], probe_order)     # first warm-up step only: outside the timed region
            r.release()
        return tot

    for _ in range(2):                # warm-up: pinned result buffers, pool growth, first-touch of the pinned pages
        step_host()
    barrier()
    t0 = time.perf_counter()
    ms_steps = []
    for _ in range(steps):
        ts = time.perf_counter()
        tot = step_host()
        ms_steps.append(round((time.perf_counter() - ts) * 1e3, 1))
    barrier()
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    hb.free()
    hp.free()
    if hpay is not None:
        hpay.free()
    return {"value": (n_build + n_probe) * world / (dt / steps), "unit": UNIT, "h2d_bytes_per_step": 8 * (n_build + n_probe) + (8 * n_build if world == 1 else 0),
            "d2h_bytes_per_step": pair_bytes * tot, "pair_format": fmt, "verified": check["ok"], "steps": steps, "ms_per_step": dt / steps * 1e3, "ms_steps_rank0": ms_steps,
            "api": ("HashJoin.build_push/probe_batch(host numpy over b200j_host_alloc pinned buffers, OUT_PAIRS|OUT_HOST" + ("|OUT_BUILD_U32" if pair_bytes <= 8 else "") + ("|OUT_PROBE_ORDER" if probe_order else "") + ") -> b200j_* C ABI" if world == 1 else
                    f"DistributedJoin.build_push_host/probe_host_chunks(pinned numpy in, pinned pairs out" + ("" if pair_bytes == 12 else ", OUT_BUILD_U32") + f", {host_chunk}-row chunks) -> b200j_* C ABI")}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
