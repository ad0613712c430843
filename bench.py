#!/usr/bin/env python
"""bench.py — queries/sec of exact top-k over a 10M x 1024 fp32 corpus (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl b200|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

A "step" is one pass of the hot path (index.search) over one batch of synthetic queries.
  value     queries/s with queries AND corpus resident in HBM (CUDA events over exactly K steps)
  e2e       the same through the C ABI with HOST buffers: pinned query rows go H2D, (D, I) come back D2H,
            every step, inside the timed region
  roofline  the scan kernel's algorithmic bytes (rows x row pitch, read once) / its mean launch time,
            measured live with CUDA events on the launching stream, against MEASURED_PEAKS.json
  cpu_baseline  the oracle's C restatement of faiss IndexFlat (OpenMP over queries only, like faiss below 20
            queries) timed on this box's host cores on a bounded row sample, scaled linearly in rows
N > 1: the fixed corpus is row-sharded over the ranks (strong scaling), rank 0's query batch reaches the other ranks
through NVLink peer memory, every rank scans its shard and stores its (score, id) lists into every rank's receive block
from inside its merge kernel, every rank merges (nv_genai_b200/sharded.py; NCCL broadcast / all-gather where there is
no peer access).
`--impl reference` times the CPU restatement alone (the reference's own implementation, faiss-cpu, is not
installable here: BASELINE.md section 2).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

UNIT = "queries/s"
# Seeds of the synthetic corpus / query streams.  The oracle keeps its own copy
# (oracle/flat.py CORPUS_SEED / QUERY_SEED); tests/test_abi.py checks that the two agree.
CORPUS_SEED = 0x5EED0001
QUERY_SEED = 0x5EED0002

# BASELINE.json configs, by name.  cfg4 is the 100M x 1024 bf16 corpus sharded over N >= 2 GPUs; at N = 1 it runs its
# weak-scaling reference point, the 12.5M-row shard one GPU holds at N = 8 (definition ii).
CONFIGS = {
    "cfg1": dict(rows=100_000, batch=1, k=4, dtype="float32", metric="l2", normalize=False,
                 label="BASELINE.json configs[0]: IndexFlatL2, the 5_mins_rag_no_gpu shape"),
    "cfg2": dict(rows=10_000_000, batch=1, k=10, dtype="float32", metric="ip", normalize=True,
                 label="BASELINE.json configs[1], batch 1"),
    "cfg2b16": dict(rows=10_000_000, batch=16, k=10, dtype="float32", metric="ip", normalize=True,
                    label="BASELINE.json configs[1], batch 16"),
    "cfg3": dict(rows=10_000_000, batch=4096, k=100, dtype="float32", metric="ip", normalize=True,
                 label="BASELINE.json configs[2], fp32 rows"),
    "cfg3bf16": dict(rows=10_000_000, batch=4096, k=100, dtype="bfloat16", metric="ip", normalize=True,
                     label="BASELINE.json configs[2], bf16 rows"),
    "cfg4": dict(rows=100_000_000, batch=1024, k=100, dtype="bfloat16", metric="ip", normalize=True,
                 label="BASELINE.json configs[3]"),
    "cfg5": dict(rows=10_000_000, batch=256, k=20, dtype="float32", metric="ip", normalize=True,
                 label="BASELINE.json configs[4]: the search of max_marginal_relevance_search (fetch_k = 20, batch 256); "
                       "ingest and the MMR kernel are under extra.cfg5_ingest_and_mmr"),
}


def metric_name(a):
    rows = f"{a.rows // 1_000_000}M" if a.rows % 1_000_000 == 0 else str(a.rows)
    kind = "IndexFlatIP" if a.metric == "ip" else "IndexFlatL2"
    return f"queries/sec, exact top-k ({kind}) over {rows} x {a.dim} {'fp32' if a.dtype == 'float32' else 'bf16'}"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS),
                    help="a BASELINE.json configuration by name (default cfg2 = configs[1], batch 1: the one the metric is "
                         "quoted on); --rows/--batch/--k/--dtype/--metric override single fields")
    ap.add_argument("--rows", type=int, default=None)
    ap.add_argument("--dim", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--dtype", default=None, choices=["float32", "bfloat16"])
    ap.add_argument("--metric", default=None, choices=["ip", "l2"])
    ap.add_argument("--cpu-sample-rows", type=int, default=1_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the planted-row / oracle-query check of the results")
    ap.add_argument("--extra", action="store_true",
                    help="also time the other BASELINE configs (batch 16, configs[2] fp32 and bf16, configs[4], configs[0]) under 'extra'")
    a = ap.parse_args()
    preset = dict(CONFIGS[a.config])
    a.label = preset.pop("label")
    a.overridden = [kk for kk in ("rows", "batch", "k", "dtype", "metric") if getattr(a, kk) is not None]
    for kk, v in preset.items():
        if getattr(a, kk, None) is None:
            setattr(a, kk, v)
    if a.config == "cfg4" and a.gpus == 1 and "rows" not in a.overridden:
        a.rows = 12_500_000
        a.label += " - weak-scaling reference point: the 12.5M-row shard one GPU holds at 8 GPUs"
    return a


def workload_name(a):
    kind = "IndexFlatIP" if a.metric == "ip" else "IndexFlatL2"
    how = "normalised " if a.normalize else ""
    s = (f"{kind} over {a.rows}x{a.dim} {a.dtype} {how}synthetic embeddings, query batch {a.batch}, k={a.k} "
         f"(--config {a.config}: {a.label}")
    if a.overridden:
        s += "; overridden on the command line: " + ", ".join(a.overridden)
    return s + ")"


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_id=None):
        self.samples = []  # (wall time, sm, sm_max, power, flags dict)
        self.proc = None
        cmd = ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"]
        if gpu_id is not None:
            cmd += ["-i", str(gpu_id)]
        try:
            self.proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm, smmax = float(parts[1]), float(parts[2])
            except ValueError:
                continue
            flags = {"hw_slowdown": parts[5], "hw_thermal_slowdown": parts[6], "sw_thermal_slowdown": parts[7],
                     "sw_power_cap": parts[8]}
            self.samples.append((time.time(), sm, smmax, parts[3], flags))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        inside = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples[-3:]
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted({name for s in inside for name, v in s[4].items() if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(s[1] for s in inside), "sm_max_mhz": max(s[2] for s in inside),
                "reasons": reasons, "samples": len(inside)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_tensor_peak(fp32_rows):
    """Dense tensor peak for a kernel timed inside a long step: the measured sustained bf16 cuBLAS figure; the
    tf32 pipe runs at half the bf16 rate (no measured tf32 number exists, so it is derived and labelled)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak, src = 1400.0, "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained bf16)"
    if os.path.exists(p):
        try:
            peak, src = float(json.load(open(p))["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        except Exception:
            pass
    if fp32_rows:
        return peak / 2.0, src + " / 2: tf32 issues at half the bf16 rate"
    return peak, src


def scan_tc_max_queries(k):
    """Queries one tensor-core scan pass takes at this k (b2_scan_tc.cu: scan_tc_max_queries)."""
    kpad = 32
    while kpad < k:
        kpad *= 2
    return 0 if kpad > 128 else min(32, 2048 // kpad)


def scan_kernel_name(batch, k, rows, dim, dtype):
    """Which scan kernel b2_index_search picks for an inner-product batch (b2_scan_tc.cu: scan_tc_supported)."""
    bf16 = dtype == "bfloat16"
    min_nq = 3 if bf16 else 5
    if rows * dim * (2 if bf16 else 4) >= float(os.environ.get("B2_SCAN_TC_MIN_BYTES", "8e9")) and (bf16 or k <= 64):
        min_nq = 1
    if "B2_SCAN_TC_MIN_NQ" in os.environ:
        min_nq = int(os.environ["B2_SCAN_TC_MIN_NQ"])
    if "B2_NO_TENSOR" in os.environ or "B2_NO_SCAN_TC" in os.environ or rows < 4096:
        return "scan_topk_kernel"
    return "scan_tc_kernel" if min_nq <= batch <= scan_tc_max_queries(k) else "scan_topk_kernel"


def uses_tensor_path(batch, k, rows):
    """True when b2_index_search sends this batch to the GEMM kernel (tensor-bound), not to a scan (HBM-bound)."""
    return batch >= 20 and k <= 128 and rows >= 4096 and batch > scan_tc_max_queries(k)


def ncu_traffic(kernel):
    """(dram bytes per launch, where the figure comes from) for the scan kernel of the default config.  bench.py cannot
    run ncu on itself: the figure is the one of the committed `ncu --set full` capture of this same command, and is
    reported only with its source beside it (None when no capture of this kernel exists)."""
    p = os.path.join(ROOT, "profiles", "scan_traffic.json")
    try:
        j = json.load(open(p))
        if not str(j.get("kernel", "")).startswith(kernel):
            return None, None
        return j.get("dram_bytes_per_launch"), "profiles/scan_traffic.json <- " + str(j.get("source"))
    except Exception:
        return None, None


# ------------------------------------------------------------------------------------------ CPU legs
def host_mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / (1 << 20)
    except Exception:
        pass
    return 0.0


def _oracle_metric(a, flat):
    return flat.METRIC_IP if a.metric == "ip" else flat.METRIC_L2


def _oracle_queries(a, flat, n, row0=0):
    q = flat.synth(row0, n, a.dim, QUERY_SEED)
    if a.normalize:
        flat.normalize_L2(q)
    return q


def _oracle_rows(a, flat, row0, n):
    x = flat.synth(row0, n, a.dim, CORPUS_SEED)
    if a.normalize:
        flat.normalize_L2(x)
    if a.dtype == "bfloat16":
        flat.round_bf16_inplace(x)  # what a bf16 row store holds
    return x


def _oracle_search(a, flat, x, q):
    """faiss's dispatch: the sequential routine (OpenMP over queries only) below 20 queries, the BLAS routine above."""
    if q.shape[0] < flat.BLAS_THRESHOLD:
        D, I, used = flat.search_seq(x, q, a.k, _oracle_metric(a, flat))
        return D, I, int(used)
    D, I = flat.search_blas(x, q, a.k, _oracle_metric(a, flat))
    return D, I, int(flat.lib().b2o_max_threads())


def cpu_full_pass(a, block=1_000_000):
    """The oracle over ALL rows of the workload for ONE batch of queries, streamed in blocks (the rows are generated
    block by block, untimed; only the scans are timed).  Returns (cpu_baseline dict, D, I): the timed CPU baseline
    and, from the same pass, the oracle's answer for the batch, which the caller checks the GPU results against."""
    from oracle import flat

    q = _oracle_queries(a, flat, a.batch)
    st = flat.StreamingSearch(q, a.k, _oracle_metric(a, flat))
    t_scan = 0.0
    for r0 in range(0, a.rows, block):
        x = _oracle_rows(a, flat, r0, min(block, a.rows - r0))
        t0 = time.perf_counter()
        st.feed(x, r0)
        t_scan += time.perf_counter() - t0
    D, I = st.result()
    cores = min(a.batch, int(flat.lib().b2o_max_threads()))
    cpu = {"value": a.batch / t_scan, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"all {a.rows} rows x {a.dim}, batch {a.batch}, k={a.k}, one pass in {t_scan:.2f} s of scan time "
                     f"(rows generated block by block outside the timing); C restatement of faiss IndexFlat "
                     f"(oracle/flat_oracle.c), OpenMP over queries only as in faiss below 20 queries; host has "
                     f"{os.cpu_count()} logical cores"}
    return cpu, D, I


def cpu_sample_timing(a, budget_s):
    """The oracle on a bounded row sample, scaled linearly in rows (a flat scan is linear in rows): the CPU baseline of
    the configurations whose full pass would take minutes (large batches: faiss's BLAS path on all cores)."""
    from oracle import flat

    q = _oracle_queries(a, flat, a.batch)
    rows = min(a.cpu_sample_rows, a.rows)
    if a.batch >= flat.BLAS_THRESHOLD:
        rows = min(rows, max(20_000, int(2.0e10 * budget_s / (a.batch * a.dim))))  # ~20 GFMA/s per pass as a first guess
    x = _oracle_rows(a, flat, 0, rows)
    _oracle_search(a, flat, x[: min(rows, 20000)], q)  # warm the library
    times, used = [], 1
    t_begin = time.time()
    while len(times) < 5 and (time.time() - t_begin < budget_s or not times):
        t0 = time.perf_counter()
        _, _, used = _oracle_search(a, flat, x, q)
        times.append(time.perf_counter() - t0)
    t_pass = statistics.median(times)
    scale = a.rows / rows
    path = ("sequential routine, OpenMP over queries only" if a.batch < flat.BLAS_THRESHOLD
            else "BLAS routine (numpy sgemm blocks + top-k), all cores")
    out = {"value": a.batch / (t_pass * scale), "unit": UNIT, "cores": min(int(used), max(a.batch, 1)) if a.batch < flat.BLAS_THRESHOLD else int(used),
           "kind": "port",
           "sample": f"{rows} of {a.rows} rows x {a.dim}, batch {a.batch}, k={a.k}, {len(times)} passes, median pass "
                     f"{t_pass * 1e3:.1f} ms, scaled x{scale:g} (a flat scan is linear in rows); faiss's {path}; host has "
                     f"{os.cpu_count()} logical cores"}
    if a.batch < flat.BLAS_THRESHOLD:
        # not faiss behaviour: the same scan with the ROWS split over all cores, as a fairer upper bound for small batches
        t0 = time.perf_counter()
        _, _, used_rp = flat.search_seq(x, q, a.k, _oracle_metric(a, flat), rowpar=True)
        t_rp = time.perf_counter() - t0
        out["all_cores_row_parallel"] = {"value": a.batch / (t_rp * scale), "unit": UNIT, "cores": int(used_rp),
                                         "note": "not what faiss does below 20 queries; the same sample, rows split over the cores"}
    return out


def run_reference(a):
    """--impl reference: the reference's CPU path for this workload.  faiss-cpu itself cannot be installed
    (BASELINE.md section 2), so this is the oracle's restatement of it, timed alone on the host cores with faiss's own
    threading.  Every step scans ALL rows when the corpus fits host memory and (warmup + steps) passes fit in about
    two minutes; otherwise each step scans a bounded sample and the step time is scaled linearly in rows (stated)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import flat

    flat.build()
    total_passes = a.steps + a.warmup
    # cost of one pass per row, from a small probe
    probe_rows = min(50_000, a.rows)
    qp = _oracle_queries(a, flat, a.batch)
    xp = _oracle_rows(a, flat, 0, probe_rows)
    _oracle_search(a, flat, xp[:5000], qp)
    t0 = time.perf_counter()
    _oracle_search(a, flat, xp, qp)
    per_row = (time.perf_counter() - t0) / probe_rows
    need_gb = a.rows * a.dim * 4 / (1 << 30)
    budget = 120.0
    full = per_row * a.rows * total_passes <= budget and host_mem_available_gb() >= need_gb * 1.1 + 4
    rows = a.rows if full else int(min(a.rows, max(20_000, budget / max(total_passes, 1) / per_row)))
    del xp
    x = np_rows_blockwise(a, flat, rows)
    qs = _oracle_queries(a, flat, a.batch * total_passes)
    used = 1
    for i in range(a.warmup):
        _oracle_search(a, flat, x, qs[i * a.batch:(i + 1) * a.batch])
    t0 = time.perf_counter()
    for i in range(a.warmup, total_passes):
        _, _, used = _oracle_search(a, flat, x, qs[i * a.batch:(i + 1) * a.batch])
    dt = time.perf_counter() - t0
    scale = a.rows / rows
    t_step = dt / a.steps * scale
    value = a.batch / t_step
    path = ("sequential routine, OpenMP over queries only" if a.batch < flat.BLAS_THRESHOLD
            else "BLAS routine (numpy sgemm blocks + top-k), all cores")
    cores = min(int(used), a.batch) if a.batch < flat.BLAS_THRESHOLD else int(used)
    sample = ((f"each step scans all {a.rows} rows" if full else
               f"each step scans {rows} of {a.rows} rows and the step time is scaled x{scale:g} (linear in rows)") +
              f", batch {a.batch}, k={a.k}; C restatement of faiss IndexFlat (oracle/flat_oracle.c), faiss's {path}; "
              f"host has {os.cpu_count()} logical cores")
    line = {"impl": "reference", "metric": metric_name(a), "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "weak" if (a.config == "cfg4" and a.gpus == 1) else "strong",
            "vs_baseline": None, "dtype": "f32" if a.dtype == "float32" else "bf16 rows, f32 accumulate", "data": "synthetic",
            "config": {"workload": workload_name(a), "rows": a.rows, "dim": a.dim, "batch": a.batch, "k": a.k,
                       "rows_scanned_per_step": rows},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def np_rows_blockwise(a, flat, rows, block=1_000_000):
    """The first `rows` rows of the workload's corpus as one host array, generated block by block."""
    import numpy as np

    x = np.empty((rows, a.dim), dtype=np.float32)
    for r0 in range(0, rows, block):
        m = min(block, rows - r0)
        x[r0:r0 + m] = _oracle_rows(a, flat, r0, m)
    return x


# ------------------------------------------------------------------------------------------ GPU arm
def build_shard(a, device, lo, hi, dtype=None):
    """Rows [lo, hi) of the synthetic corpus, generated on the device and ingested (with the fused normalise for the
    inner-product configs)."""
    import torch

    from nv_genai_b200 import IndexFlatIP, IndexFlatL2, synth_fill

    idx = (IndexFlatIP if a.metric == "ip" else IndexFlatL2)(a.dim, dtype=dtype or a.dtype, device=device)
    chunk = min(1_000_000, max(hi - lo, 1))
    buf = torch.empty((chunk, a.dim), dtype=torch.float32, device=f"cuda:{device}")
    for r0 in range(lo, hi, chunk):
        m = min(chunk, hi - r0)
        synth_fill(buf[:m], r0, CORPUS_SEED)
        idx.add(buf[:m], normalize=a.normalize)
    torch.cuda.synchronize(device)
    del buf
    torch.cuda.empty_cache()
    return idx


def cfg5(a, device, idx, peak, peak_src):
    """BASELINE configs[4]: add_embeddings ingest + normalise (kernel-only from device rows, and end to end from
    pinned host rows), then max_marginal_relevance_search fetch_k=20, k=4 over a batch of 256 queries."""
    import torch

    from nv_genai_b200 import IndexFlatIP, synth_fill

    rows = min(1_000_000, a.rows)
    out = {}
    for dtype in ("float32", "bfloat16"):
        tmp = IndexFlatIP(a.dim, dtype=dtype, device=device)
        buf = torch.empty((rows, a.dim), dtype=torch.float32, device=f"cuda:{device}")
        synth_fill(buf, 0, CORPUS_SEED)
        reps = 5
        for _ in range(reps):  # grow the row store to its final size first: mapping memory is host work, not the kernel
            tmp.add(buf, normalize=True)
        tmp.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            tmp.add(buf, normalize=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        bytes_alg = rows * a.dim * (4 + (4 if dtype == "float32" else 2))
        out[f"ingest_kernel_{dtype}"] = {"rows_per_s": rows / (ms * 1e-3), "ms_per_1M_rows": ms * 1e6 / rows,
                                         "roofline": {"bound": "hbm", "achieved": bytes_alg / (ms * 1e-3) / 1e9,
                                                      "peak": peak, "unit": "GB/s",
                                                      "frac": bytes_alg / (ms * 1e-3) / 1e9 / peak,
                                                      "kernel": "ingest_kernel", "algorithmic_bytes_per_launch": bytes_alg,
                                                      "peak_source": peak_src}}
        del tmp
        if dtype == "float32":
            host = torch.empty((rows, a.dim), dtype=torch.float32).pin_memory()
            host.copy_(buf)
            tmp = IndexFlatIP(a.dim, dtype=dtype, device=device)
            tmp.add(host[:1000].numpy(), normalize=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            tmp.add(host.numpy(), normalize=True)  # H2D staged in 64 MB chunks inside b2_index_add, synchronous
            dt = time.perf_counter() - t0
            out["ingest_e2e_pinned_host_float32"] = {"rows_per_s": rows / dt, "h2d_gbs": rows * a.dim * 4 / dt / 1e9,
                                                     "note": "PCIe-bound: host rows -> device staging -> fused normalise"}
            # what a langchain caller passes: pageable numpy rows (np.array(embeddings, dtype=np.float32)); they go
            # through the library's pinned ring (b2_index_add)
            pageable = host.numpy().copy()
            tmp.reset()
            tmp.add(pageable[:1000], normalize=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            tmp.add(pageable, normalize=True)
            dt = time.perf_counter() - t0
            out["ingest_e2e_pageable_host_float32"] = {"rows_per_s": rows / dt, "h2d_gbs": rows * a.dim * 4 / dt / 1e9,
                                                       "note": "pageable numpy rows -> pinned ring (up to 8 host threads) -> DMA -> fused normalise"}
            del tmp, host, pageable
        del buf
        torch.cuda.empty_cache()
    nq = 256
    q = make_queries(a, device, 1, nq)[0]
    idx.mmr(q, k=4, fetch_k=20, lambda_mult=0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        idx.mmr(q, k=4, fetch_k=20, lambda_mult=0.5)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out["mmr_batch256_fetch20_k4"] = {"queries_per_s": nq / (ms * 1e-3), "ms_per_batch": ms,
                                      "note": "search (tensor-core path, fetch_k=20) + row gather + MMR kernel, device resident"}
    return out


def cfg1(a, device):
    """BASELINE configs[0]: IndexFlatL2 over 100k x 1024 un-normalised rows, one query, k = 4 (the 5_mins_rag_no_gpu
    shape): the launch-latency regime, reported as latency."""
    import torch

    from nv_genai_b200 import IndexFlatL2, synth_fill

    n = 100_000
    idx = IndexFlatL2(a.dim, device=device)
    buf = torch.empty((n, a.dim), dtype=torch.float32, device=f"cuda:{device}")
    synth_fill(buf, 0, CORPUS_SEED)
    idx.add(buf)
    q = torch.empty((256, a.dim), dtype=torch.float32, device=f"cuda:{device}")
    synth_fill(q, 0, QUERY_SEED)
    qh = q.cpu().numpy()
    for i in range(10):
        idx.search(q[i:i + 1], 4)
        idx.search(qh[i:i + 1], 4)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 200
    e0.record()
    for i in range(reps):
        idx.search(q[i:i + 1], 4)
    e1.record()
    torch.cuda.synchronize()
    dev_us = e0.elapsed_time(e1) * 1e3 / reps
    lat = []
    for i in range(reps):
        t0 = time.perf_counter()
        idx.search(qh[i:i + 1], 4)  # host buffers in and out, synchronous
        lat.append((time.perf_counter() - t0) * 1e6)
    return {"rows": n, "device_resident_us_per_query": dev_us, "e2e_host_median_us": statistics.median(lat),
            "e2e_host_p99_us": sorted(lat)[int(0.99 * len(lat)) - 1],
            "hbm_floor_us": n * a.dim * 4 / (measured_peak()[0] * 1e9) * 1e6,
            "note": "409.6 MB per query: 62.6 us at the measured copy peak; scan + merge launches and the host call on top"}


def make_queries(a, device, count, batch):
    import torch

    from nv_genai_b200 import normalize_L2, synth_fill

    q = torch.empty((count * batch, a.dim), dtype=torch.float32, device=f"cuda:{device}")
    synth_fill(q, 0, QUERY_SEED)
    if a.normalize:
        normalize_L2(q)
    return q.view(count, batch, a.dim)


def compare_topk(Dg, Ig, Dr, Ir, rtol=1e-5, atol=1e-6):
    """The parity rule of tests/parity.py in short: scores within rtol position by position; ids equal except inside
    runs of scores that tie within the tolerance.  Returns "ok" or what differs."""
    import numpy as np

    Dg, Ig, Dr, Ir = (np.asarray(v) for v in (Dg, Ig, Dr, Ir))
    if Dg.shape != Dr.shape or Ig.shape != Ir.shape:
        return f"shape {Dg.shape} vs {Dr.shape}"
    tol = atol + rtol * np.abs(Dr.astype(np.float64))
    if not np.all(np.abs(Dg.astype(np.float64) - Dr.astype(np.float64)) <= tol):
        return "scores differ beyond 1e-5 relative"
    for r in range(Ig.shape[0]):
        for j in np.nonzero(Ig[r] != Ir[r])[0]:
            pos = np.nonzero(Ir[r] == Ig[r][j])[0]
            ref = Dr[r][pos[0]] if len(pos) else Dr[r][-1]  # an id from outside must tie the k-th score
            if abs(float(ref) - float(Dr[r][j])) > atol + 10 * rtol * abs(float(Dr[r][j])):
                return f"query {r} position {j}: id {int(Ig[r][j])} where the oracle has {int(Ir[r][j])} and the scores do not tie"
    return "ok"


def planted_ids(rows, world):
    """Eight global ids whose rows are searched for themselves: the ends, the middle and the shard boundaries."""
    from nv_genai_b200.sharded import shard_bounds

    ids = [0, 1, rows // 2 - 1, rows // 2, rows - 1]
    for r in range(1, world):
        lo, _ = shard_bounds(rows, world, r)
        ids += [lo - 1, lo]
    ids += [rows // 3, (2 * rows) // 3, rows // 5]
    out = []
    for i in ids:
        if 0 <= i < rows and i not in out:
            out.append(i)
    return out[:8]


def run_b200(a):
    import torch

    from nv_genai_b200 import _native
    from nv_genai_b200.sharded import ShardedFlatIndex, shard_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: nv_genai_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            # rank 0 prints ONE JSON line on stdout: NCCL's own lines (its version banner) go to the debug file
            os.environ["NCCL_DEBUG"] = "WARN"
            os.environ.setdefault("NCCL_DEBUG_FILE", os.devnull)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}: launch N>1 with torchrun"

    lo, hi = shard_bounds(a.rows, world, rank)
    idx = build_shard(a, local, lo, hi)
    total = a.warmup + a.steps
    sh = None
    if world > 1:
        sh = ShardedFlatIndex(a.dim, idx.metric_type, local_index=idx)
        sh._refresh_bases()
        assert sh.ntotal == a.rows

    pitch = ((a.dim * (2 if a.dtype == "bfloat16" else 4) + 511) // 512) * 512
    peak, peak_src = measured_peak()
    total = a.warmup + a.steps
    props = torch.cuda.get_device_properties(local)
    gpu_id = f"GPU-{props.uuid}" if getattr(props, "uuid", None) else local
    sampler = ClockSampler(gpu_id) if rank == 0 else None

    def one_config(batch, index=None, k=None, dtype=None, steps=None):
        """Returns dict with device-resident and end-to-end timings for this query batch size."""
        index = idx if index is None else index
        k = a.k if k is None else k
        dtype = a.dtype if dtype is None else dtype
        steps = a.steps if steps is None else steps
        warm = min(a.warmup, 3) if steps < a.steps else a.warmup
        tot = warm + steps
        shx = None
        if world > 1:
            shx = sh if index is idx else ShardedFlatIndex(a.dim, index.metric_type, local_index=index)
            if shx is not sh:
                shx._refresh_bases()
        q_dev = make_queries(a, local, tot, batch)
        q_host = torch.empty((tot, batch, a.dim), dtype=torch.float32).pin_memory()
        q_host.copy_(q_dev.cpu())
        q_np = q_host.numpy()
        stage = torch.empty((batch, a.dim), dtype=torch.float32, device=f"cuda:{local}")
        packed = torch.empty(batch * k * 12, dtype=torch.uint8, device=f"cuda:{local}")
        out_host = torch.empty(batch * k * 12, dtype=torch.uint8).pin_memory()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tensor = uses_tensor_path(batch, k, hi - lo)
        kern = _native.PROF_GEMM if tensor else _native.PROF_SCAN

        def step_resident(i):
            if shx is None:
                return index.search(q_dev[i], k)
            # rank 0 owns the batch; the others receive it by broadcast inside sh.search
            buf = q_dev[i] if rank == 0 else stage
            return shx.search(buf, k, replicated=False, src=0)

        def step_e2e(i):
            if shx is None:
                return index.search(q_np[i], k)  # host buffers in, host buffers out, through b2_index_search
            if rank == 0:
                stage.copy_(q_host[i], non_blocking=True)  # H2D from pinned memory
            D, I = shx.search(stage, k, replicated=False, src=0, out=packed)
            if rank == 0:
                out_host.copy_(packed)  # ONE D2H of the merged result, packed [I | D]; synchronises
            return D, I

        def timed(step_fn, with_prof):
            for i in range(warm):
                step_fn(i)
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
                torch.cuda.synchronize()
            if with_prof:
                _native.prof_read(reset=True, kernel=kern)
                _native.prof_enable(True)
            _native.launch_count_reset()
            t0 = time.time()
            e0.record()
            for i in range(warm, tot):
                step_fn(i)
            e1.record()
            torch.cuda.synchronize()
            t1 = time.time()
            launches = _native.launch_count()
            k_ms, k_n = (0.0, 0)
            if with_prof:
                _native.prof_enable(False)
                k_ms, k_n = _native.prof_read(reset=True, kernel=kern)
            if dist is not None:
                dist.barrier()
                torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local}")
            ln = torch.tensor([launches], dtype=torch.int64, device=f"cuda:{local}")
            if dist is not None:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                dist.all_reduce(ln, op=dist.ReduceOp.SUM)
            return float(ms.item()), int(ln.item()), k_ms, k_n, t0, t1

        ms, launches, k_ms, k_n, t0, t1 = timed(step_resident, True)
        ms_e, _, _, _, _, _ = timed(step_e2e, False)
        clocks = sampler.summary(t0, t1) if sampler is not None else None
        pitch_c = ((a.dim * (2 if dtype == "bfloat16" else 4) + 511) // 512) * 512
        k_avg_ms = k_ms / max(k_n, 1)
        k_step_ms = k_ms / steps  # all launches of the profiled kernel within one step
        if tensor:
            fp32_rows = dtype != "bfloat16"
            flops = 2.0 * batch * (hi - lo) * a.dim  # algorithmic: one product per (query, row, dim)
            # default: filter and refine = ONE approximate product per score on the tensor cores (the candidates near
            # the cut are re-scored in fp32 by the merge kernel); the all-exact arithmetic (B2_GEMM_REFINE=0 or an
            # explicit number of query parts) issues 2 (bf16 rows) / 3 (fp32 rows) products per score
            refine = os.environ.get("B2_GEMM_REFINE", "1") != "0" and "B2_GEMM_QUERY_PARTS" not in os.environ
            # fp32 rows under refine are filtered through their bf16 shadow: the approximate pass is a bf16 product
            shadow = fp32_rows and refine and "B2_NO_SHADOW" not in os.environ and "B2_NO_VMM" not in os.environ
            tpeak, tsrc = measured_tensor_peak(fp32_rows and not shadow)
            products = 1 if refine else (3 if fp32_rows else 2)
            achieved = flops / (k_step_ms * 1e-3) / 1e12 if k_n else None
            roof = {"bound": "tensor", "achieved": achieved, "peak": tpeak, "unit": "TFLOP/s",
                    "frac": (achieved / tpeak) if achieved else None, "traffic": None,
                    "kernel": "gemm_topk_kernel (%s)" % (("approximate pass of filter-and-refine" +
                                                          (", bf16 shadow of the fp32 rows" if shadow else "")) if refine else "all-exact"),
                    "algorithmic_flops_per_step": flops,
                    "launches_per_step": k_n / steps,
                    "launch_note": "a search launches the kernel twice when k >= 32: a threshold pre-pass over every 32nd corpus tile, "
                                   "then the main pass over the other tiles; achieved = algorithmic flop of the step / summed launch time",
                    "mma_products_per_score": products,
                    "issued_mma_tflops": (achieved * products * (-(-batch // 128) * 128) / batch) if achieved else None,
                    "kernel_ms_per_step": k_step_ms, "launches_timed": k_n, "peak_source": tsrc,
                    "kernel_share_of_step": (k_ms / ms) if k_n else None}
        else:
            shard_bytes = (hi - lo) * pitch_c
            achieved = shard_bytes / (k_avg_ms * 1e-3) / 1e9 if k_n else None
            per_step = 1 if batch <= max(16, scan_tc_max_queries(k)) else -(-batch // 16)  # scan passes per step
            roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": (achieved / peak) if achieved else None,
                    "traffic": None,
                    "kernel": scan_kernel_name(batch, k, hi - lo, a.dim, dtype),
                    "algorithmic_bytes_per_launch": shard_bytes,
                    "avg_launch_ms": k_avg_ms, "launches_timed": k_n, "peak_source": peak_src,
                    "scan_share_of_step": (k_avg_ms * k_n / ms) if k_n else None, "scan_passes_per_step": per_step}
            if batch == 1 and dtype == "float32" and world == 1 and a.rows == 10_000_000 and a.metric == "ip":
                roof["traffic"], roof["traffic_source"] = ncu_traffic(roof["kernel"])
        return {
            "batch": batch, "k": k, "dtype": dtype, "steps": steps, "ms_per_step": ms / steps,
            "value": batch * steps / (ms * 1e-3),
            "e2e_value": batch * steps / (ms_e * 1e-3), "e2e_ms_per_step": ms_e / steps,
            "launches": launches, "clocks": clocks, "roofline": roof,
            "h2d": batch * a.dim * 4, "d2h": batch * k * 12,
        }

    main = one_config(a.batch)
    extra = None
    if a.extra:
        extra = {}
        keep = ("batch", "k", "dtype", "steps", "value", "ms_per_step", "e2e_value", "roofline")
        if a.batch != 16 and a.k == 10:
            e = one_config(16)  # the other half of BASELINE configs[1]
            extra["batch16"] = {kk: e[kk] for kk in keep}
        if a.dtype == "float32" and a.batch != 4096 and a.metric == "ip":
            # BASELINE configs[2]: batch 4096, k = 100, fp32 rows and bf16 rows, tensor-core path
            e = one_config(4096, k=100, steps=max(3, min(a.steps, 5)))
            extra["cfg3_fp32_batch4096_k100"] = {kk: e[kk] for kk in keep}
            idx16 = build_shard(a, local, lo, hi, dtype="bfloat16")
            e = one_config(4096, index=idx16, k=100, dtype="bfloat16", steps=max(3, min(a.steps, 5)))
            extra["cfg3_bf16_batch4096_k100"] = {kk: e[kk] for kk in keep}
            e = one_config(1, index=idx16, dtype="bfloat16", steps=max(3, min(a.steps, 20)))
            extra["bf16_batch1_k10"] = {kk: e[kk] for kk in keep}
            del idx16
            torch.cuda.empty_cache()
        if world == 1 and a.metric == "ip":
            extra["cfg5_ingest_and_mmr"] = cfg5(a, local, idx, peak, peak_src)
            extra["cfg1_100k_l2_k4_latency"] = cfg1(a, local)
    elif a.config == "cfg5" and world == 1:
        extra = {"cfg5_ingest_and_mmr": cfg5(a, local, idx, peak, peak_src)}
    if sampler is not None:
        sampler.stop()

    # ---- correctness of what was timed (outside the timed region; every N): planted rows must find themselves, and
    # the first query batch must equal the oracle's answer over all rows.  The oracle pass is also the CPU baseline.
    check, cpu = None, None
    if not a.no_check:
        ids = planted_ids(a.rows, world)
        pq = torch.empty((len(ids), a.dim), dtype=torch.float32, device=f"cuda:{local}")
        from nv_genai_b200 import normalize_L2, synth_fill

        for j, gid in enumerate(ids):
            synth_fill(pq[j:j + 1], gid, CORPUS_SEED)
        if a.normalize:
            normalize_L2(pq)
        Dp, Ip = (sh.search(pq, a.k, replicated=True) if sh is not None else idx.search(pq, a.k))
        q0 = make_queries(a, local, 1, a.batch)[0]
        nchk = a.batch if a.batch < 20 else 4  # large batches: the first four queries of the same batch
        D0, I0 = (sh.search(q0, a.k, replicated=True) if sh is not None else idx.search(q0, a.k))
        torch.cuda.synchronize()
        if rank == 0:
            Ip_h, Dp_h = Ip.cpu().numpy(), Dp.cpu().numpy()
            best = 1.0 if a.metric == "ip" else 0.0
            tol = 1e-2 if a.dtype == "bfloat16" else 1e-4
            ok = bool((Ip_h[:, 0] == ids).all()) and (not a.normalize or bool((abs(Dp_h[:, 0] - best) <= tol).all()))
            check = {"planted": "ok" if ok else f"FAILED: ids {Ip_h[:, 0].tolist()} for planted rows {ids}",
                     "planted_ids": ids}
    if rank == 0 and not a.no_cpu_baseline:
        from oracle import flat

        flat.build()
        full_cost_s = a.rows / 1e6 * (1.8 + 0.35 * min(a.batch, 20) / max(1, min(min(a.batch, 20), os.cpu_count() or 1)))
        if a.no_check or full_cost_s > 60.0:
            if check is not None:
                check["oracle_query"] = f"skipped: a full oracle pass over {a.rows} rows would take ~{full_cost_s:.0f} s"
        else:
            sub = argparse.Namespace(**vars(a))
            sub.batch = nchk
            cpu_full, Dr, Ir = cpu_full_pass(sub)
            check["oracle_query"] = compare_topk(D0[:nchk].cpu().numpy(), I0[:nchk].cpu().numpy(), Dr, Ir)
            check["oracle_queries_checked"] = nchk
            if a.batch < 20 and world == 1:
                cpu = cpu_full
        if world == 1 and cpu is None:
            cpu = cpu_sample_timing(a, budget_s=12.0)
        elif world == 1 and a.batch < 20:
            rp = cpu_sample_timing(a, budget_s=2.0)
            if "all_cores_row_parallel" in rp:
                cpu["all_cores_row_parallel"] = rp["all_cores_row_parallel"]
                cpu["all_cores_row_parallel"]["sample"] = rp["sample"]

    if rank == 0:
        weak = a.config == "cfg4" and world == 1
        line = {
            "metric": metric_name(a), "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True,
            "scaling": "weak" if weak else "strong",
            "vs_baseline": None, "dtype": "f32" if a.dtype == "float32" else "bf16 rows, f32 accumulate",
            "data": "synthetic",
            "config": {"workload": workload_name(a), "name": a.config, "rows": a.rows, "dim": a.dim, "batch": a.batch, "k": a.k,
                       "rows_per_gpu": hi - lo, "partitioning": f"row-sharded x{world}" if world > 1 else "single GPU",
                       "l2_note": "inputs larger than L2: each step streams the whole row store "
                                  f"({(hi - lo) * pitch / 1e9:.2f} GB per GPU) against a 126 MB L2"},
            "e2e": {"value": main["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": main["h2d"],
                    "d2h_bytes_per_step": main["d2h"], "ms_per_step": main["e2e_ms_per_step"]},
            "gpu_launches": main["launches"], "clocks": main["clocks"], "roofline": main["roofline"],
        }
        if check is not None:
            line["check"] = check
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if a.config == "cfg4":
            line["cfg4_efficiency"] = cfg4_efficiency(world, main["value"], main["ms_per_step"])
        if extra is not None:
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def cfg4_efficiency(world, qps, ms):
    """(i) strong scaling against the smallest feasible point, QPS(G) / (QPS(2) G / 2); (ii) weak scaling,
    t1(12.5M rows on one GPU) / t8(100M rows on eight).  The other points come from the lines of earlier runs kept under
    profiles/r02_bench_cfg4_n<N>.json (this run's own line is what the caller saves there)."""
    out = {"definition_i": None, "definition_ii": None}
    def load(n):
        try:
            return json.load(open(os.path.join(ROOT, "profiles", f"r02_bench_cfg4_n{n}.json")))
        except Exception:
            return None
    two, one = load(2), load(1)
    if world >= 2:
        q2 = qps if world == 2 else (two or {}).get("value")
        if q2:
            out["definition_i"] = {"eff": qps / (q2 * world / 2), "qps_2gpu": q2, "formula": "QPS(G) / (QPS(2) * G / 2)"}
    if world == 8 and one:
        out["definition_ii"] = {"eff": one["ms_per_step"] / ms, "t1_ms_12.5M_rows": one["ms_per_step"], "t8_ms": ms,
                                "formula": "t1(12.5M rows, 1 GPU) / t8(100M rows, 8 GPUs)"}
    return out


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
