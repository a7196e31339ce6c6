#!/usr/bin/env python
"""bench.py -- samples/sec of the decision-forest batch-predict hot path (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # our arm
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU predict

Headline workload (BASELINE.json configs[2], the one north_star names for 1/2/4/8 GPUs): Higgs-shaped
synthetic forest (28 features, 1000 trees, depth 12) on 10M rows IN TOTAL, rank g owning rows
[g*R/N, (g+1)*R/N) -- strong scaling; N=1 holds all 10M rows (1.12 GB).  A "step" = one
tahoe_predict_to over this rank's rows whose last-chunk epilogue stores the final scores straight
into rank 0's gathered buffer (CUDA-IPC mapped; no copy and no collective on the path).
The forest is replicated once (one broadcast, untimed).

At N=1 the same JSON line carries `configs` (C1..C4, one device-resident measurement each, with the
SURVEY 8(d) roofline fractions against the L2 / HBM / shared-memory peaks measured in this run) and
`sweep` (model pick vs exhaustive enumeration at a few batch sizes).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tahoe_b200 import synth  # noqa: E402

WORKLOAD = "C3"
METRIC = "samples/sec"
L2_FLUSH_BYTES = 256 << 20


def _hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic(cfg_id: str):
    """dram bytes per launch of the config's dominant kernel from the committed ncu capture, or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("dram_bytes_per_launch", {}).get(cfg_id)
    return None


def _ncu(cfg_id: str):
    """Counters of the config's dominant kernel from the committed `ncu --set full` capture (profiles/r02_ncu_full_*):
    DRAM GB/s, LSU data pipe, shared-memory wavefronts and bank-conflict wavefronts, threads per warp instruction."""
    name = f"r02_ncu_full_{cfg_id}_summary.json"
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        return None
    with open(path) as f:
        d = json.load(f)[0]

    def val(key):
        v = d.get(key)
        if not v or v[0] in ("", None):
            return None
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}.get(v[1], 1.0)
        return float(v[0]) * scale

    sec = val("gpu__time_duration.sum")
    dram = (val("dram__bytes_read.sum") or 0.0) + (val("dram__bytes_write.sum") or 0.0)
    wf, conf = val("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"), val("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
    return {"source": f"profiles/{name}", "kernel": d.get("kernel"), "kernel_ms_under_ncu": None if not sec else sec * 1e3,
            "dram_gbs": None if not sec else dram / sec / 1e9,
            "dram_pct_of_peak": val("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
            "lsu_data_pipe_pct": val("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
            "smem_wavefronts": wf, "smem_bank_conflict_wavefronts": conf,
            "smem_bank_conflict_share": None if not wf else (conf or 0.0) / wf,
            "threads_per_warp_instruction": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "stall_samples_pct": d.get("stall_samples_pct")}


def width_rule(cols: int) -> int:
    """reference node-width rule, Struct.h:1831-1852: 1 if max_fid < 32, 2 if < 8192, else 4"""
    return 1 if cols - 1 < 32 else (2 if cols - 1 < 8192 else 4)


def algorithmic_bytes(cfg, rows: int, mean_depth: float, forest_nodes: int):
    """B_sample = R(4C+4); B_node = R*T*(dbar*(4+b)+4); B_forest = T*N_real*(4+b)."""
    C, T = cfg["cols"], cfg["trees"]
    b = width_rule(C)
    b_sample = rows * (4.0 * C + 4.0)
    b_node = float(rows) * T * (mean_depth * (4 + b) + 4)
    b_forest = float(forest_nodes) * (4 + b)
    return b_sample, b_node, b_forest


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.stop_flag = index, threading.Event()
        self.sm, self.reasons, self.sm_max, self.ok = [], set(), None, False

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.sm_max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            self.ok = True
            while not self.stop_flag.is_set():
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
                time.sleep(0.02)
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def result(self):
        self.stop_flag.set()
        self.join(timeout=2)
        if not self.ok or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": ["nvml_unavailable"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except (ValueError, IndexError):
            pass
    return local_rank


def host_threads() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _cpu_runner(cfg, nodes, threads):
    """The reference's own CPU predict (oracle/_ref, OpenMP over rows) -- or the oracle port when the
    shim is absent.  The only place besides tests/ and smoke() that executes anything under oracle/."""
    from oracle.oracle import Oracle, RefShim
    if RefShim.available():
        shim = RefShim()
        return (lambda d: shim.predict(nodes, cfg["trees"], cfg["depth"], d, -999.0, nthreads=threads)), "reference"
    orc = Oracle()
    return (lambda d: orc.predict_dense(nodes, cfg["trees"], cfg["depth"], d, -999.0)), "port"


def cpu_reference_rate(cfg, nodes, target_s: float, threads: int):
    """Time the CPU predict on a bounded sample of the workload: (samples/s, kind, rows, seconds)."""
    run, kind = _cpu_runner(cfg, nodes, threads)
    probe = synth.gen_data(synth.DATA_SEED[WORKLOAD], 64 * threads, cfg["cols"])
    run(probe[:64])
    t0 = time.perf_counter(); run(probe); dt = time.perf_counter() - t0
    rows = int(max(64 * threads, min(2_000_000, probe.shape[0] / dt * target_s)))
    data = synth.gen_data(synth.DATA_SEED[WORKLOAD], rows, cfg["cols"])
    t0 = time.perf_counter(); run(data); dt = time.perf_counter() - t0
    return rows / dt, kind, rows, dt


def workload_name(cfg, rows_total):
    return (f"{cfg['name']} synthetic forest ({cfg['cols']} features, {cfg['trees']} trees, depth "
            f"{cfg['depth']}) on {rows_total} rows in total")


def run_reference(args, rank: int, world: int):
    """--impl reference: the reference's own CPU predict on the host cores (rank 0 only)."""
    if rank != 0:
        return
    cfg = synth.CONFIGS[WORKLOAD]
    threads = host_threads()
    nodes = synth.gen_dense_forest(synth.FOREST_SEED[WORKLOAD], cfg["trees"], cfg["depth"], cfg["cols"])
    run, kind = _cpu_runner(cfg, nodes, threads)
    # size one step to ~1.5 s so warmup+steps end within a few minutes
    probe = synth.gen_data(synth.DATA_SEED[WORKLOAD], 64 * threads, cfg["cols"])
    run(probe[:64])
    t0 = time.perf_counter(); run(probe); dt = time.perf_counter() - t0
    budget = min(1.5, 150.0 / max(1, args.steps + args.warmup))
    rows = int(max(64 * threads, min(2_000_000, probe.shape[0] / dt * budget)))
    data = synth.gen_data(synth.DATA_SEED[WORKLOAD], rows, cfg["cols"])
    for _ in range(args.warmup):
        run(data)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(data)
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    sample = f"{rows} rows of the {cfg['rows']}-row workload per step, {threads} OpenMP threads over rows"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(cfg, cfg["rows"]), "sampled_rows_per_step": rows},
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


_RESULT_FD = None


def _claim_stdout():
    """stdout carries ONE JSON line.  Libraries write there too (NCCL prints its version banner on
    fd 1), so fd 1 is pointed at stderr for the run and the line goes out on the saved descriptor."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


# ---------------------------------------------------------------------------------------------
def build_forest(cfg_id: str, nodes=None):
    """-> (Forest, real node count, levels).  Dense configs from the AoS nodes, config 4 from compact
    nodes; levels = (tree_offsets, level of every compact node, largest |leaf value|) for the sparse config, else None."""
    from tahoe_b200 import api
    cfg = synth.CONFIGS[cfg_id]
    if cfg["sparse"]:
        sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg_id], cfg["trees"], cfg["depth"], cfg["cols"])
        leaf_abs_max = float(np.abs(sp["val"][sp["bits"] < 0]).max()) if "bits" in sp.dtype.names else 1.0
        return (api.Forest.from_sparse(sp, offs, cfg["cols"], weights=w, missing=-999.0), int(offs[-1]),
                (offs, lv, leaf_abs_max))
    if nodes is None:
        nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg_id], cfg["trees"], cfg["depth"], cfg["cols"])
    f = api.Forest.from_dense(nodes, cfg["trees"], cfg["depth"], cfg["cols"], missing=-999.0)
    return f, cfg["trees"] * synth.tree_num_nodes(cfg["depth"]), None


def measured_mean_depth(forest, data, levels, sample=4096):
    """dbar = mean internal nodes visited per (row, tree).  For the unbalanced forest it is
    MEASURED: the level of the leaf every (row, tree) of a row sample reaches, from the leaf ids the
    DIRECT kernel reports (held bit-exact to the oracle's by tests/test_gpu_full_size.py)."""
    import torch
    from tahoe_b200 import api
    offs, lv = levels[0], levels[1]
    n = min(sample, int(data.shape[0]))
    T = offs.shape[0] - 1
    ids = torch.zeros((n, T), dtype=torch.int32, device="cuda")
    forest.predict(data[:n].contiguous(), strategy=api.DIRECT, leaf_ids=ids)
    torch.cuda.synchronize()
    node = np.asarray(offs[:-1], dtype=np.int64)[None, :] + ids.cpu().numpy()
    return float(np.asarray(lv)[node].mean())


KERNEL_OF = {  # dominant kernel per strategy (names as they appear in the ncu launch lists under profiles/)
    "direct": "k_direct", "shared_data_per_thread": "k_direct<smem_rows>", "shared_data": "k_shared_data",
    "shared_forest": "k_smem_forest", "split_shared_forest": "k_smem_forest",
}


def measure_config(cfg_id, forest, n_nodes, data, peaks, hbm_gbs, steps, strategy=None, flush=None, levels=None):
    """One device-resident measurement of one config: step time over `steps` predicts (CUDA events on
    the launch stream) and the dominant kernel's own time (events recorded around it inside the
    library).  Returns the sub-record with the SURVEY 8(d) fractions."""
    import torch
    from tahoe_b200 import api
    cfg = synth.CONFIGS[cfg_id]
    rows = int(data.shape[0])
    info = forest.info()
    s = forest.select(rows)[0] if strategy is None else strategy
    preds = torch.empty(rows, dtype=torch.float32, device="cuda")
    for _ in range(3):
        forest.predict(data, preds, strategy=s)
    torch.cuda.synchronize()
    forest.set_profiling(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    total = 0.0
    for _ in range(steps):
        if flush is not None:
            flush.zero_()                      # inputs smaller than L2 (config 1 is not; the others are larger anyway)
        e0.record(); forest.predict(data, preds, strategy=s); e1.record()
        torch.cuda.synchronize()
        total += e0.elapsed_time(e1)
    q_ms, t_ms, n = forest.profile_read()
    forest.set_profiling(False)
    step_ms = total / steps
    kernel_ms = t_ms / n if n else step_ms
    name = api.STRATEGY_NAMES[s]
    kernel = KERNEL_OF[name]
    if info.is_sparse and name in ("shared_forest", "split_shared_forest"):
        kernel = "k_smem_forest_deep" if info.sparse_top_active else "k_sparse_smem"
    mean_depth = measured_mean_depth(forest, data, levels) if levels is not None else float(cfg["depth"])
    b_sample, b_node, b_forest = algorithmic_bytes(cfg, rows, mean_depth, n_nodes)
    abytes = b_sample + b_node
    sec = kernel_ms * 1e-3
    rec = {
        "workload": workload_name(cfg, rows), "rows": rows, "strategy": name, "kernel": kernel,
        "kernel_ms": kernel_ms, "prepare_ms": (q_ms / n if n else 0.0), "step_ms": step_ms,
        "samples_per_s": rows / (step_ms * 1e-3),
        "algorithmic_bytes_per_launch": abytes, "achieved_gbs": abytes / sec / 1e9,
        "forest_fits_l2": bool(b_forest <= peaks["l2_bytes"]),
        "frac_l2": abytes / sec / 1e9 / peaks["l2_read_gbs"],
        "frac_hbm_only": (b_sample + b_forest) / sec / 1e9 / hbm_gbs,
        "mean_depth": mean_depth, "traffic": _traffic(cfg_id), "ncu": _ncu(cfg_id),
    }
    if name in ("shared_forest", "split_shared_forest") and not info.is_sparse:
        # the node and rank words of this kernel come out of SHARED MEMORY: 8 B per visit + 4 B per leaf
        smem_bytes = float(rows) * cfg["trees"] * (cfg["depth"] * 8 + 4)
        rec["frac_smem"] = smem_bytes / sec / 1e9 / peaks["smem_read_gbs"]
        rec["plan"] = {"warps": info.smem_warps, "staged": info.smem_staged, "group": info.smem_group,
                       "trees_per_chunk": info.trees_per_chunk, "chunks": info.num_chunks, "cluster": info.smem_cluster}
    # the sparse shared-memory path finishes its deep walks in an order that depends on the launch shape: its host
    # slabs and its whole-batch launch may differ in the last bits; bound = the tests' 1e-6 x the L1 mass of a row's
    # summands (<= trees x the largest |leaf|).  Every other path must match bit for bit.
    tol = 1e-6 * cfg["trees"] * levels[2] if (info.is_sparse and levels is not None and len(levels) > 2) else None
    rec["e2e"] = host_leg(forest, data, preds, s, rows / (step_ms * 1e-3), tol_abs=tol)
    return rec, s, preds


def host_leg(forest, data, preds, strategy, device_rate, reps=3, tol_abs=None):
    """The same predict through host buffers (tahoe_predict_host: pinned rows in, scores out, copies inside the timed
    region) and the same slab copies with the kernels left out: what the end-to-end number is bounded by."""
    import torch
    rows = int(data.shape[0])
    h_data = torch.empty(tuple(data.shape), dtype=torch.float32, pin_memory=True)
    h_data.copy_(data)
    h_preds = torch.empty(rows, dtype=torch.float32, pin_memory=True)
    out = {}
    for key, copy_only in (("copies_only_samples_per_s", True), ("value", False)):
        forest.predict_host(h_data, h_preds, strategy=strategy, copy_only=copy_only)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            forest.predict_host(h_data, h_preds, strategy=strategy, copy_only=copy_only)
        torch.cuda.synchronize()
        out[key] = rows * reps / (time.perf_counter() - t0)
    out["unit"] = "samples/s"
    out["h2d_bytes_per_step"] = int(data.numel()) * 4
    out["d2h_bytes_per_step"] = rows * 4
    out["bit_identical_to_device_path"] = bool(torch.equal(h_preds.cuda(), preds))
    out["max_abs_diff_vs_device_path"] = float((h_preds.cuda().double() - preds.double()).abs().max().item())
    out["tolerance_abs"] = tol_abs           # None: bit-identity is required (every path that sums in tree order)
    out["matches_device_path"] = out["bit_identical_to_device_path"] or (
        tol_abs is not None and out["max_abs_diff_vs_device_path"] <= tol_abs)
    out["frac_of_bound"] = out["value"] / min(device_rate, out["copies_only_samples_per_s"])
    return out


def time_strategy(forest, d, p, s, reps=3):
    import torch
    forest.predict(d, p, strategy=s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = float("inf")
    for _ in range(reps):
        e0.record(); forest.predict(d, p, strategy=s); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def sweep_record(cfg_id, forest, data, sizes):
    """BASELINE.json config 5 in small: every applicable strategy timed at a few batch sizes, the
    model's pick against the measured best (profiles/ holds the full sweeps)."""
    import torch
    from tahoe_b200 import api
    pts = []
    for R in sizes:
        if R > data.shape[0]:
            continue
        d = data[:R]
        p = torch.empty(R, dtype=torch.float32, device="cuda")
        times = {api.STRATEGY_NAMES[s]: time_strategy(forest, d, p, s) for s in api.ALL_STRATEGIES
                 if forest.applicable(s, R)}
        pick = api.STRATEGY_NAMES[forest.select(R)[0]]
        best = min(times, key=times.get)
        pts.append({"rows": R, "ms": times, "pick": pick, "best": best, "slowdown": times[pick] / times[best]})
    return {"config": cfg_id, "points": pts,
            "hit_rate": sum(p["pick"] == p["best"] for p in pts) / max(1, len(pts)),
            "worst_slowdown": max([p["slowdown"] for p in pts], default=1.0)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--strategy", type=int, default=-1, help="force a strategy id (default: model picks)")
    ap.add_argument("--rows", type=int, default=0, help="override TOTAL rows (debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs only)")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1/C2/C4 sub-records and the sweep")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    _claim_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from tahoe_b200 import api, shard
    api.lib()  # fail loudly before any work if the CUDA library is missing
    assert torch.cuda.is_available(), "bench.py (our arm) needs a GPU; there is no CPU path"
    torch.cuda.set_device(local_rank)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    cfg = dict(synth.CONFIGS[WORKLOAD])
    total_rows = args.rows or cfg["rows"]
    lo, hi = shard.row_range(rank, world, total_rows)
    rows = hi - lo
    n_nodes = cfg["trees"] * synth.tree_num_nodes(cfg["depth"])
    # ---- forest: built once on rank 0, broadcast once (the only collective; not timed)
    nodes = None
    if rank == 0:
        nodes = synth.gen_dense_forest(synth.FOREST_SEED[WORKLOAD], cfg["trees"], cfg["depth"], cfg["cols"])
    nodes = shard.broadcast_forest(nodes, n_nodes, src=0)
    forest = api.Forest.from_dense(nodes, cfg["trees"], cfg["depth"], cfg["cols"], missing=-999.0)
    info = forest.info()
    strategy = args.strategy if args.strategy >= 0 else forest.select(rows)[0]

    # ---- this rank's shard, generated in HBM (inputs resident before the timed region); the final
    # scores of every rank land in rank 0's buffer (peer-mapped), partial sums stay local
    data = api.gen_data_device(synth.DATA_SEED[WORKLOAD], rows, cfg["cols"], row0=lo)
    gather = shard.PeerGather(total_rows, dst=0)
    torch.cuda.synchronize()

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        gather.predict(forest, data, strategy=strategy)
    launches_per_step = forest.launches()
    forest.set_profiling(True)
    sampler = ClockSampler(physical_gpu_index(local_rank))
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        gather.predict(forest, data, strategy=strategy)
    e1.record()
    barrier()
    clocks = sampler.result()
    ms_total = e0.elapsed_time(e1)
    q_ms, t_ms, n_prof = forest.profile_read()
    forest.set_profiling(False)
    if distributed:
        t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = total_rows / (ms_step * 1e-3)
    # untimed check of the fused gather: every rank predicts its rows once more into LOCAL memory; the sum of those
    # (all-reduced) must be what rank 0 finds in the gathered buffer, and rank 0's own rows must match bit for bit
    local_check = forest.predict(data, strategy=strategy)
    local_sum = torch.tensor([float(local_check.double().sum().item())], dtype=torch.float64, device="cuda")
    if distributed:
        dist.all_reduce(local_sum, op=dist.ReduceOp.SUM)
    gathered_checksum = float(gather.full.double().sum().item()) if rank == 0 else None
    gathered_ok = None
    if rank == 0:
        gathered_ok = bool(torch.equal(gather.full[lo:hi], local_check)) and \
            abs(gathered_checksum - float(local_sum.item())) <= 1e-9 * max(1.0, abs(gathered_checksum))

    # ---- end to end: pinned host rows in, host scores out, every step (this rank's shard)
    e2e = None
    if not args.no_e2e:
        h_data = torch.empty((rows, cfg["cols"]), dtype=torch.float32, pin_memory=True)
        h_data.copy_(data)
        h_preds = torch.empty(rows, dtype=torch.float32, pin_memory=True)
        e2e_steps = max(3, min(args.steps, 5))

        def timed_host(copy_only):
            for _ in range(2):
                forest.predict_host(h_data, h_preds, strategy=strategy, copy_only=copy_only)
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                forest.predict_host(h_data, h_preds, strategy=strategy, copy_only=copy_only)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if distributed:
                tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            return dt

        copy_s = timed_host(True)       # the same slabs, copies only: the bound of this path
        e2e_s = timed_host(False)
        same = bool(torch.equal(h_preds.cuda(), local_check))
        if distributed:
            ok_t = torch.tensor([1 if same else 0], dtype=torch.int32, device="cuda")
            dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
            same = bool(ok_t.item())
        h2d = rows * cfg["cols"] * 4
        e2e = {"value": total_rows * e2e_steps / e2e_s, "unit": "samples/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": rows * 4,
               "steps": e2e_steps, "matches_device_path": same,
               "copies_only_samples_per_s": total_rows * e2e_steps / copy_s,
               "h2d_only_gbs": world * (h2d + rows * 4) * e2e_steps / copy_s / 1e9,
               "frac_of_bound": (total_rows * e2e_steps / e2e_s) / min(value, total_rows * e2e_steps / copy_s)}
        del h_data, h_preds

    if rank == 0:
        hbm_gbs, hbm_src = _hbm_peak()
        peaks = api.measure_peaks()         # L2 / shared-memory read bandwidth, measured now, on this GPU
        b_sample, b_node, b_forest = algorithmic_bytes(cfg, rows, float(cfg["depth"]), n_nodes)
        abytes = b_sample + b_node
        trav_ms = t_ms / n_prof if n_prof else ms_step
        sec = trav_ms * 1e-3
        achieved = abytes / sec / 1e9
        smem_bytes = float(rows) * cfg["trees"] * (cfg["depth"] * 8 + 4)
        fits_l2 = b_forest <= peaks["l2_bytes"]
        peak = peaks["l2_read_gbs"] if fits_l2 else hbm_gbs
        line = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": workload_name(cfg, total_rows),
                "strategy": api.STRATEGY_NAMES[strategy], "rows_total": total_rows, "rows_per_gpu": rows,
                "parallelism": f"rows sharded x{world}, forest replicated",
                "gathered": True,
                "gather": "final scores stored by the kernel epilogue into rank 0's buffer (CUDA IPC peer mapping)",
                "l2_policy": f"inputs_exceed_l2 ({rows * cfg['cols'] * 4 / 1e6:.0f} MB of rows per GPU per step "
                             f"vs {peaks['l2_bytes'] / 1e6:.0f} MB L2)",
                "forest_chunks": info.num_chunks, "trees_per_chunk": info.trees_per_chunk,
                "plan": {"warps": info.smem_warps, "staged": info.smem_staged, "group": info.smem_group,
                         "cluster": info.smem_cluster},
                "node_bytes": 4, "rank_bits": info.rank_bits, "gathered_checksum": gathered_checksum,
                "gathered_matches_local_predicts": gathered_ok},
            "roofline": {"bound": "l2" if fits_l2 else "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         # ncu capture at the full 10M rows, scaled to this rank's rows (the tile re-streaming is linear in rows)
                         "traffic": (None if _traffic(WORKLOAD) is None else int(_traffic(WORKLOAD) * rows / cfg["rows"])),
                         "kernel": "k_smem_forest_cluster" if info.smem_cluster > 1 else "k_smem_forest",
                         "kernel_ms": trav_ms, "quantize_ms": (q_ms / n_prof if n_prof else None),
                         "algorithmic_bytes_per_launch": abytes,
                         "peak_source": "L2 streaming-read bandwidth measured in this run (tahoe_measure_peaks); SURVEY 8(d): "
                                        "the packed forest fits L2" if fits_l2 else hbm_src,
                         "frac_hbm_only": (b_sample + b_forest) / sec / 1e9 / hbm_gbs,
                         "hbm_peak": hbm_gbs, "hbm_peak_source": hbm_src,
                         "note": "SURVEY 8(d) accounting: sample bytes + forest-node bytes touched, rank 0's shard; the node "
                                 "words are served from shared memory -- see smem for the bound the kernel runs against",
                         "smem": {"bound": "shared-memory bandwidth", "achieved": smem_bytes / sec / 1e9,
                                  "peak": peaks["smem_read_gbs"], "unit": "GB/s",
                                  "frac": smem_bytes / sec / 1e9 / peaks["smem_read_gbs"],
                                  "peak_source": "conflict-free LDS.32 stream measured in this run",
                                  "ideal_bytes_per_launch": smem_bytes}},
            "peaks": peaks,
            "e2e": e2e,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
        }
        if world == 1 and not args.no_configs:
            flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
            cfgs = {}
            rec, _, _ = measure_config("C3", forest, n_nodes, data, peaks, hbm_gbs, steps=3, strategy=strategy)
            cfgs["C3"] = rec
            sweeps = [sweep_record("C3", forest, data, [1, 1000, 100_000, 10_000_000])]
            del data
            forest.close()
            torch.cuda.empty_cache()
            for cid, steps in (("C1", 20), ("C2", 10), ("C4", 10)):
                c = synth.CONFIGS[cid]
                f, nn, levels = build_forest(cid)
                d = api.gen_data_device(synth.DATA_SEED[cid], c["rows"], c["cols"])
                rec, _, _ = measure_config(cid, f, nn, d, peaks, hbm_gbs, steps=steps, flush=flush, levels=levels)
                cfgs[cid] = rec
                if cid == "C2":
                    sweeps.append(sweep_record("C2", f, d, [1, 1000, 100_000, 5_000_000]))
                del d
                f.close()
                torch.cuda.empty_cache()
            line["configs"] = {k: cfgs[k] for k in sorted(cfgs)}
            line["sweep"] = sweeps
        if world == 1 and not args.no_cpu_baseline:
            threads = host_threads()
            v, kind, srows, sdt = cpu_reference_rate(cfg, nodes, 12.0, threads)
            line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": threads, "kind": kind,
                                    "sample": f"{srows} rows of the workload in {sdt:.1f} s, OpenMP over rows"}
        _emit(line)
    if distributed:
        dist.barrier()
    if rank != 0:
        gather.close()          # mappings first, then (after the barrier) the owner's buffer
    if distributed:
        dist.barrier()
    gather.close()
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
