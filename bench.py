#!/usr/bin/env python
"""Benchmark of the ORFanage hot path on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (CUDA path through the C-ABI)
  python bench.py --impl reference [--gpus N] [--steps K] ...     the reference's own CPU code (oracle/_ref)

A "step" is one pass of the whole hot path (prep, candidate rows, codon scan, rescue + score, --mode
reduction) over one batch: the GENCODE-scale synthetic configuration (BASELINE.json configs[1]:
~250k query transcripts vs ~90k template CDSs on a 3 Gb random genome, --mode LONGEST_MATCH,BEST),
generated on the box by the seeded generator tools/orfx_synth.  At N > 1 every rank evaluates its own
configs[1]-sized locus shard (weak scaling; loci are independent, no collective on the data path).

One JSON line on stdout (rank 0).  `value` = query transcripts/s with the batch resident in HBM (device
time, CUDA events on the library's stream, max over ranks); `e2e` = the same through orfx_submit /
orfx_wait with the batch in pinned HOST memory, uploads and result downloads inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C2 = dict(seed=20261104, loci=20000, chroms=24, chrom_len=125000000, queries=13, templates=8, template_files=1)
# BASELINE.json configs[3]: long-read novel isoforms, ~2 M query transcripts (100 per locus), ONE workload sharded by locus
# over the GPUs of the box (strong scaling; the product's own shard_bundles deals the loci)
C4 = dict(seed=20261107, loci=20000, chroms=24, chrom_len=125000000, queries=100, templates=8, template_files=1)
MODE = "LONGEST_MATCH,BEST"
LAYOUT = "islands2d"         # islands of query-exon bases as 2-bit codes + exception list (ORFX_SEQ_NT2), chains through the
                             # segment dictionary (orfx_batch::seg_ref): the smallest upload
ALG_BYTES_PER_NT = 1.25      # codon scan: 1 B read per spliced nucleotide + 2 mask bits written


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def workload_dir(cfg, tag):
    base = os.environ.get("ORFX_BENCH_DIR", "/tmp/orfx_bench")
    return os.path.join(base, "%s_s%d_l%d_c%dx%d_q%d_t%d" % (tag, cfg["seed"], cfg["loci"], cfg["chroms"], cfg["chrom_len"],
                                                           cfg["queries"], cfg["templates"]))


def generate(cfg, tag):
    """Seeded synthetic workload (tools/orfx_synth); reused when already on disk."""
    d = workload_dir(cfg, tag)
    done = os.path.join(d, ".done")
    if os.path.exists(done):
        return d
    os.makedirs(d, exist_ok=True)
    t = time.time()
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", str(cfg["seed"]), "--loci", str(cfg["loci"]),
                    "--chroms", str(cfg["chroms"]), "--chrom-len", str(cfg["chrom_len"]), "--queries", str(cfg["queries"]),
                    "--templates", str(cfg["templates"]), "--template-files", str(cfg["template_files"])], check=True,
                   stderr=subprocess.DEVNULL)
    open(done, "w").write("ok\n")
    log("generated %s in %.1f s" % (d, time.time() - t))
    return d


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:   # NVML missing: report that, never invent clocks
            self.nv, self.err = None, str(e)

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown:
                 "hw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def result(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no NVML samples"}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ---- the reference's CPU implementation of the path (oracle/_ref) -----------------------------------
SAMPLE = dict(C2, loci=3336, chroms=4)     # same generator law and locus spacing as configs[1], 1/6 of the loci


def evaluated_queries(gtf_path):
    """Query transcripts the hot path evaluated, counted from an output GTF without any of this repo's code: the
    transcript rows that carry an orfanage_status attribute are exactly the queries of bundles that hold a template
    (orfanage.cpp:383-432 copies the others through without one)."""
    n = 0
    with open(gtf_path, "rb") as f:
        for line in f:
            if b"\ttranscript\t" in line and b"orfanage_status" in line:
                n += 1
    return n


def source_digest():
    """SHA-256 over the device code -- every .cu file that defines a kernel, the .cuh headers and the public headers: ties a
    profile (profiles/r02_traffic.json) to the kernels it measured.  orfx_api.cu (host code only: copies, launches, waits) is
    left out, so a host-side change does not orphan a kernel profile."""
    import hashlib
    h = hashlib.sha256()
    files = []
    for d in ("orfanage_b200/csrc", "include"):
        for f in sorted(os.listdir(os.path.join(ROOT, d))):
            if f.endswith((".cu", ".cuh", ".h")):
                path = os.path.join(ROOT, d, f)
                if f.endswith(".cu") and b"__global__" not in open(path, "rb").read():
                    continue
                files.append(path)
    for f in files:
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def file_sha256(path):
    import hashlib
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 22), b""):
            h.update(chunk)
    return h.hexdigest()


def cli_leg(d):
    """The thing a user runs: whole-process wall time of orfanage_b200/bin/orfanage on the configs[1] files (its own stage
    timers beside it), next to the reference binary's whole-process time on the same files with all host threads; the GTF
    is checked against the reference's (SHA-256 of the reference's --threads 1 output, tests/golden/fullsize_hashes.json:
    above one thread the reference's output order is nondeterministic)."""
    import orfanage_b200 as ob
    out = {"workload": "GENCODE-scale synthetic (BASELINE.json configs[1]) from the files on disk", "mode": MODE}
    args = ["--reference", "genome.fa", "--query", "query.gtf", "--mode", MODE]
    walls, stderr = [], ""
    for _ in range(2):   # the second run has the page cache the reference run below also gets
        t0 = time.time()
        p = subprocess.run([ob.cli_path()] + args + ["--output", "cli.gtf", "--timing", "template.gtf"], cwd=d, capture_output=True, text=True)
        walls.append(time.time() - t0)
        if p.returncode != 0:
            return {"error": p.stderr[-300:]}
        stderr = p.stderr
    nq = evaluated_queries(os.path.join(d, "cli.gtf"))
    out.update({"wall_s": min(walls), "wall_s_runs": walls, "evaluated_query_transcripts": nq, "transcripts_per_s": nq / min(walls),
                "stage_timers": [l[len("[timing] "):] for l in stderr.splitlines() if l.startswith("[timing]")]})
    try:
        want = json.load(open(os.path.join(ROOT, "tests", "golden", "fullsize_hashes.json")))["c2"]
        same_gen = all(want["generator"].get(k) == C2.get(k) for k in ("seed", "loci", "chroms", "chrom_len", "queries", "templates", "template_files"))
        out["gtf_identical_to_reference"] = bool(same_gen and file_sha256(os.path.join(d, "cli.gtf")) == want["gtf_sha256"])
    except Exception as e:
        out["gtf_identical_to_reference"] = None
        out["note"] = "no digest to compare with: %s" % e
    try:
        threads = os.cpu_count() or 1
        ev, whole = time_reference_once(d, threads)
        out["reference"] = {"binary": "oracle/_ref/orfanage_ref (the reference's own sources)", "threads": threads, "wall_s": whole, "main_loop_s": ev,
                            "evaluated_query_transcripts": evaluated_queries(os.path.join(d, "ref_out.gtf"))}
        out["speedup_whole_process"] = whole / min(walls)
    except Exception as e:
        out["reference"] = {"unavailable": str(e)}
    return out


def time_reference_once(d, threads):
    """One run of the reference binary on the workload in d.  Returns seconds spent in its main evaluation
    loop (between its own 'starting main evaluation' and 'Done' stderr lines, orfanage.cpp:369,819) and
    whole-process seconds."""
    ref = os.path.join(ROOT, "oracle", "_ref", "orfanage_ref")
    if not os.path.exists(ref):
        raise RuntimeError("oracle/_ref/orfanage_ref is not built (make oracle, in the container that has /root/reference)")
    out = os.path.join(d, "ref_out.gtf")
    t0 = time.time()
    p = subprocess.Popen([ref, "--reference", "genome.fa", "--query", "query.gtf", "--output", out, "--mode", MODE, "--threads",
                          str(threads), "template.gtf"], cwd=d, stderr=subprocess.PIPE, stdout=subprocess.DEVNULL, text=True, bufsize=1)
    t_start = t_done = None
    for line in p.stderr:
        if "starting main evaluation" in line:
            t_start = time.time()
        elif line.startswith("Done"):
            t_done = time.time()
    rc = p.wait()
    t1 = time.time()
    if rc != 0 or t_start is None or t_done is None:
        raise RuntimeError("reference run failed (rc %d)" % rc)
    return t_done - t_start, t1 - t0


def reference_arm(args):
    """bench.py --impl reference: the reference's own sources (compiled against the gclib stand-in) on the
    host cores, all threads, on a bounded sample of the workload.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # A step = one run of the reference binary (it reloads its inputs every time: ~14 s per run at configs[1]).  Runs of a few
    # steps take configs[1] itself; the driver's 20 + 5 steps take the bounded 1/6 sample so that the arm ends within minutes.
    full = args.steps + args.warmup <= 6
    d = generate(C2 if full else SAMPLE, "c2" if full else "sample")
    for _ in range(args.warmup):
        time_reference_once(d, threads)
    evals, wholes = [], []
    for _ in range(args.steps):
        e, w = time_reference_once(d, threads)
        evals.append(e)
        wholes.append(w)
    nq = evaluated_queries(os.path.join(d, "ref_out.gtf"))   # counted from the reference's own output
    ms = 1e3 * sum(evals) / len(evals)
    value = nq / (ms / 1e3)
    sample = ("configs[1] itself: %d loci / %d evaluated query transcripts" % (C2["loci"], nq)) if full else \
        "%d loci / %d query transcripts of the configs[1] generator (1/6 of the loci, same locus spacing)" % (SAMPLE["loci"], nq)
    line = {"impl": "reference", "metric": "query transcripts/sec", "value": value, "unit": "transcripts/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": "GENCODE-scale synthetic (BASELINE.json configs[1])" + ("" if full else " — bounded sample"), "mode": MODE, "sample": sample,
                       "timed": "reference main evaluation loop (orfanage.cpp:369-819), mean of %d runs" % args.steps,
                       "whole_process_s": sum(wholes) / len(wholes)},
            "cpu_baseline": {"value": value, "unit": "transcripts/s", "cores": threads, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "transcripts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def bind_to_gpu_numa_node(index):
    """Pins this process (and the pages it first touches: the pinned batch arena, the result buffers) to the NUMA node the
    GPU hangs off, so that eight ranks pushing their batches at once do not cross the socket interconnect.  Returns what was
    done for the JSON line; silently does nothing where the topology cannot be read."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:      # NVML prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read().strip())
        if node < 0:
            return {"gpu": index, "node": None}
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"gpu": index, "pci": bus, "node": node, "cpus": len(cpus)}
    except Exception as e:
        return {"gpu": index, "node": None, "note": str(e)[:80]}


# ---- our arm -------------------------------------------------------------------------------------------
def cpu_baseline_leg():
    """Reference binary on the bounded sample, all host threads, one warm-up + two timed runs."""
    threads = os.cpu_count() or 1
    d = generate(SAMPLE, "sample")
    time_reference_once(d, threads)
    ev = [time_reference_once(d, threads)[0] for _ in range(2)]
    e1 = time_reference_once(d, 1)[0]
    nq = evaluated_queries(os.path.join(d, "ref_out.gtf"))
    sample = "%d loci / %d query transcripts of the configs[1] generator; main evaluation loop only" % (SAMPLE["loci"], nq)
    return {"value": nq / (sum(ev) / len(ev)), "unit": "transcripts/s", "cores": threads, "kind": "reference", "sample": sample,
            "single_thread_value": nq / e1}


def ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import orfanage_b200 as ob

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 and not args.no_numa else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    c4 = args.config == "c4"
    cfg = dict(C4 if c4 else C2)
    if args.loci:
        cfg["loci"] = args.loci
        cfg["chroms"] = max(1, min(24, args.loci // 834 + 1))
    # every rank evaluates a configs[1]-sized locus shard of its own (distinct seed => distinct loci)
    # (when the scratch disk cannot hold one genome per rank, all ranks read rank 0's)
    import shutil
    base = os.environ.get("ORFX_BENCH_DIR", "/tmp/orfx_bench")
    os.makedirs(base, exist_ok=True)
    need = world * (cfg["chroms"] * cfg["chrom_len"] * 1.2 + 1e9)
    shared = world > 1 and (c4 or shutil.disk_usage(base).free < need)
    cfg["seed"] = cfg["seed"] + (0 if shared or c4 else rank) + args.seed_offset
    tag = "c4" if c4 else "c2"
    if shared:
        if rank == 0:
            generate(cfg, tag)
        barrier()
    d = generate(cfg, tag)
    t = time.time()
    w = ob.Workload(os.path.join(d, "query.gtf"), [os.path.join(d, "template.gtf")], os.path.join(d, "genome.fa"))
    # configs[1]: every rank its own workload (weak scaling); configs[3]: every rank loads the one workload and packs the
    # locus shard shard_bundles deals it (strong scaling)
    batch = w.pack(rank, world, pinned=True, layout=LAYOUT) if c4 else w.pack(0, 1, pinned=True, layout=LAYOUT)
    nq, nt_sum = batch.n_queries, (batch.query_nt() if c4 else w.count("query_nt"))
    h2d_bytes, seq_bytes = batch.h2d_bytes, int(batch.arrays["seq"].nbytes)
    log("rank %d: loaded + packed in %.1f s: %d queries, %d templates, %d bundles, %.1f MB windows, %.1f M spliced nt"
        % (rank, time.time() - t, nq, w.count("templates"), w.count("active_bundles"), batch.arrays["seq"].nbytes / 1e6, nt_sum / 1e6))
    if args.verbose_shard:   # what makes one rank's shard slower than another's: its largest loci
        import numpy as np
        a_ = batch.arrays
        per_tx = np.diff(a_["tx_seg_off"]).astype(np.int64)
        per_b = np.add.reduceat(per_tx, a_["bundle_tx_off"][:-1].astype(np.int64))
        ntx_b = np.diff(a_["bundle_tx_off"])
        top = np.argsort(-per_b)[:5]
        log("rank %d: largest bundles by segments: %s (transcripts %s); bundles with > 8192 segments: %d; transcripts with > 32 segments: %d"
            % (rank, per_b[top].tolist(), ntx_b[top].tolist(), int((per_b > 8192).sum()), int((per_tx > 32).sum())))
    params = ob.Params.make(mode=MODE, want_rows=False)
    cx = ob.Context(params, device=local)

    # ---- leg 1: batch resident in HBM -------------------------------------------------------------
    cx.upload(batch)
    for _ in range(args.warmup):
        cx.run_resident()
    stage_ms = {}
    launches = 0
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        launches += cx.run_resident()
        dev_ms += cx.last_run_ms()
        for k, v in cx.kernel_times().items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.result()
    res = cx.fetch()
    found = int((res.best["t_tx"] >= 0).sum())
    # ---- leg 1b: the same resident batch with every kernel alone on the one stream (orfx_debug_set "serialize"): per-kernel
    # launch times without anything running beside them -- the codon-scan roofline's kernel time (in the product path the scan
    # of the ordinary loci shares the GPU with the heavy bundles' rescue kernel)
    stage_ms_alone, alone_ms = {}, 0.0
    cx.debug_set("serialize", 1)
    for _ in range(2):
        cx.run_resident()
    for _ in range(args.steps):
        cx.run_resident()
        alone_ms += cx.last_run_ms()
        for k, v in cx.kernel_times().items():
            stage_ms_alone[k] = stage_ms_alone.get(k, 0.0) + v
    cx.debug_set("serialize", 0)
    assert cx.fetch().best.tobytes() == res.best.tobytes(), "serialized and overlapped runs give the same answers"
    # order-independent digest of the chosen ORFs over workload-wide transcript indices: the same whatever the sharding
    gl = w.tx_global(len(batch.arrays["tx_kind"]), rank if c4 else 0).astype(np.uint64)
    b_ = res.best
    tq = gl[b_["q_tx"]]
    tt_ = np.where(b_["t_tx"] >= 0, gl[np.maximum(b_["t_tx"], 0)], np.uint64(0xFFFFFFFF))
    mix = (tq * np.uint64(0x9E3779B97F4A7C15)) ^ (tt_ * np.uint64(0xC2B2AE3D27D4EB4F)) ^ (b_["cds_start"].astype(np.uint64) * np.uint64(0x165667B19E3779F9)) \
        ^ (b_["cds_end"].astype(np.uint64) * np.uint64(0x27D4EB2F165667C5)) ^ (b_["ilpi"].view(np.uint32).astype(np.uint64) << np.uint64(17)) \
        ^ (b_["duplicity"].astype(np.uint64) << np.uint64(49))
    mix ^= mix >> np.uint64(29)
    dig_lo, dig_hi = int((mix & np.uint64(0xFFFFFFFF)).sum()), int((mix >> np.uint64(32)).sum())

    # ---- leg 2: end to end through orfx_submit / orfx_wait, batch in pinned host memory -----------
    for _ in range(max(1, args.warmup // 2)):
        cx.run(batch)
    barrier()
    t0 = time.perf_counter()
    e2e_dev = [0.0, 0.0, 0.0]
    for _ in range(args.steps):
        cx.submit(batch)
        r = cx.wait(copy=False)
        launches += cx.last_run_launches()
        e2e_dev = [e2e_dev[0] + r.ms_h2d, e2e_dev[1] + r.ms_kernels, e2e_dev[2] + r.ms_d2h]
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    assert cx.last_upload_bytes() == h2d_bytes, "the library copied %d bytes per step, the batch arrays say %d" % (cx.last_upload_bytes(), h2d_bytes)
    d2h_bytes = int(nq * ob._capi.BEST_DTYPE.itemsize + 32)

    # ---- leg 2b: the path with the alignment stage enabled (params.align_enabled, --pi): every scored row's ORF protein against
    # its template's protein.  The batch carries raw bytes and the template CDS bases (layout islands_t).
    align_path = None
    if rank == 0 and not c4 and not args.no_align_path:
        tb = w.pack(0, 1, pinned=True, layout="islands_t")
        ap_ = ob.Params.make(mode=MODE, want_rows=True, align=True)
        cxa = ob.Context(ap_, device=local)
        cxa.upload(tb)
        for _ in range(2):
            cxa.run_resident()
        ms_a, st_a, n_a = 0.0, {}, 3
        for _ in range(n_a):
            cxa.run_resident()
            ms_a += cxa.last_run_ms()
            for k, v in cxa.kernel_times().items():
                st_a[k] = st_a.get(k, 0.0) + v
        ra = cxa.fetch()
        scored = ra.rows["qlen"] > 0
        tlen_codons = ra.rows["tlen"][scored].astype(np.int64) // 3
        cells_a = float(((ra.rows["qlen"][scored].astype(np.int64) // 3) * tlen_codons).sum())
        align_path = {"workload": "configs[1] with params.align_enabled (the CLI's --pi): %d scored rows aligned per step" % int(scored.sum()),
                      "ms_per_step": ms_a / n_a, "translate_ms": st_a.get("translate", 0.0) / n_a, "align_ms": st_a.get("align", 0.0) / n_a,
                      "cells_per_step": cells_a, "gcups_in_path": cells_a / (st_a.get("align", 1e-9) / n_a * 1e-3) / 1e9,
                      "sequence_upload_bytes": int(tb.arrays["seq"].nbytes), "checksum": int(np.bitwise_xor.reduce(ra.rows["aln_score"].astype(np.int64) * 31 + ra.rows["aln_max"]))}
        cxa.close()
        batch = w.pack(0, 1, pinned=True, layout=LAYOUT)   # packing shard 0 again released the arrays `batch` viewed

    # ---- leg 3: the same calls as a stream of batches: two contexts, the next batch is submitted while the previous one is
    # still being evaluated (orfx_submit returns once everything is enqueued; a context is waited for before it is reused).
    # Every step's inputs cross PCIe and every step's results are read back inside the timed region; what the double buffering
    # hides is the latency of one behind the other, as any caller that has more than one batch would.
    e2e3_ms = None
    if not args.no_pipelined:
        n_ring = max(2, args.stream_contexts)
        ring = [cx] + [ob.Context(params, device=local) for _ in range(n_ring - 1)]
        # more waiting threads on the box than cores: let them sleep instead of spin (include/orfx.h: blocking_sync)
        blocking = args.blocking_sync == 1 or (args.blocking_sync < 0 and world * n_ring > (os.cpu_count() or 1))
        for c_ in ring:
            c_.debug_set("blocking_sync", 1 if blocking else 0)
        for c_ in ring:
            c_.submit(batch)
        for c_ in ring:
            c_.wait(copy=False)
        # one host thread per context (as the CLI drives its shards): orfx_submit blocks on the batch's validation flag and on
        # the four totals that size the row buffers, and a second caller thread keeps the copy engine fed meanwhile
        import threading
        share = [args.steps // n_ring + (1 if k < args.steps % n_ring else 0) for k in range(n_ring)]
        errs = []

        def drive(c_, n):
            try:
                for _ in range(n):
                    c_.submit(batch)
                    c_.wait(copy=False)
            except Exception as e:   # noqa: BLE001
                errs.append(e)
        barrier()
        t0 = time.perf_counter()
        th = [threading.Thread(target=drive, args=(ring[k], share[k])) for k in range(n_ring)]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
        barrier()
        e2e3_ms = (time.perf_counter() - t0) * 1e3
        assert not errs, errs
        cx.debug_set("blocking_sync", 0)
        for c_ in ring[1:]:
            c_.close()

    # ---- leg 4 (last: packing the shards releases the whole-batch arrays `batch` views): the same call sequence on several contexts, each with its share of the loci (shard_bundles): the upload of one
    # share overlaps the kernels of another (orfx_submit returns once everything is enqueued).  Same work, same bytes, same API.
    e2e2_ms = None
    n_ctx = max(2, args.e2e_contexts)
    if not args.no_pipelined and not c4:
        parts = [w.pack(k, n_ctx, pinned=True, layout=LAYOUT) for k in range(n_ctx)]
        assert sum(h.n_queries for h in parts) == nq
        extra = [ob.Context(params, device=local) for _ in range(n_ctx - 1)]
        pair = list(zip([cx] + extra, parts))
        for _ in range(max(1, args.warmup // 2)):
            for c_, b_ in pair:
                c_.submit(b_)
            found2 = sum(int((c_.wait().best["t_tx"] >= 0).sum()) for c_, _ in pair)
        assert found2 == found, "the locus shards find the same ORFs as the whole batch"
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            for c_, b_ in pair:
                c_.submit(b_)
            for c_, _ in pair:
                c_.wait(copy=False)
        barrier()
        e2e2_ms = (time.perf_counter() - t0) * 1e3
        for c_ in extra:
            c_.close()

    # max over ranks of the timed regions, sum of the units
    tt = torch.tensor([dev_ms, wall_ms, e2e_ms, e2e2_ms if e2e2_ms is not None else 0.0, e2e3_ms if e2e3_ms is not None else 0.0], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([nq, launches, found, nt_sum, dig_lo, dig_hi], dtype=torch.int64, device="cuda")
    per_rank = None
    if world > 1:
        mine = tt.clone()
        gathered = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(gathered, mine)   # per-rank device / e2e times beside the max: shards differ (distinct loci per rank)
        per_rank = {"device_ms_per_step": [round(float(g[0]) / args.steps, 4) for g in gathered],
                    "e2e_ms_per_step": [round(float(g[2]) / args.steps, 4) for g in gathered],
                    "e2e_pipelined_ms_per_step": [round(float(g[3]) / args.steps, 4) for g in gathered],
                    "e2e_streamed_ms_per_step": [round(float(g[4]) / args.steps, 4) for g in gathered]}
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt)
    dev_ms, wall_ms, e2e_ms, e2e2_ms, e2e3_ms = [float(x) for x in tt.tolist()]
    nq_all, launches_all, found_all, nt_all, dig_lo, dig_hi = [int(x) for x in cnt.tolist()]
    digest = "%08x%08x" % (dig_hi & 0xFFFFFFFF, dig_lo & 0xFFFFFFFF)
    align = None
    if c4:
        pass
    elif args.align_pairs > 0 and world > 1:
        # C5 pairs are independent: dealt to the ranks by DP cells (ob.shard_pairs), no exchange step; device time = max over ranks
        import numpy as np
        t = time.time()
        aa, off, pq, pt = c5_pairs(args.align_pairs * world, 20261104)   # weak scaling: the 1-GPU sample per GPU
        mine = ob.shard_pairs(off, pq, pt, world)[rank]
        cx.align(aa, off, pq[mine[:64]], pt[mine[:64]])
        best_ms, cells, out = None, 0.0, None
        for _ in range(3):
            barrier()
            out, cells, ms = cx.align(aa, off, pq[mine], pt[mine])
            best_ms = ms if best_ms is None else min(best_ms, ms)
        ta = torch.tensor([best_ms], dtype=torch.float64, device="cuda")
        tc = torch.tensor([cells, float(int(np.bitwise_xor.reduce(out["score"].astype(np.int64) * 31 + out["max"])) & 0xFFFFFF)], dtype=torch.float64, device="cuda")
        gathered = [torch.zeros_like(ta) for _ in range(world)]
        dist.all_gather(gathered, ta)
        dist.all_reduce(ta, op=dist.ReduceOp.MAX)
        dist.all_reduce(tc)
        if rank == 0:
            align = {"workload": "C5-law protein pairs (BASELINE.json configs[4]), %d pairs over %d GPUs, dealt by DP cells (shard_pairs)"
                                 % (args.align_pairs * world, world),
                     "pairs": int(args.align_pairs * world), "n_gpus": world, "cells": float(tc[0]), "kernel_ms": float(ta[0]),
                     "gcups": float(tc[0]) / (float(ta[0]) * 1e-3) / 1e9, "kernel_ms_per_rank": [round(float(g[0]), 3) for g in gathered],
                     "scaling": "weak", "timing": "CUDA events around the kernel on each rank, best of 3, max over ranks"}
            log("alignment leg on %d GPUs: %.1f s, %.1f GCUPS" % (world, time.time() - t, align["gcups"]))
    elif rank == 0 and args.align_pairs > 0:
        t = time.time()
        align = alignment_leg(cx, args.align_pairs)
        log("alignment leg: %.1f s, %.1f GCUPS" % (time.time() - t, align["gcups"]))
    return dict(numa=numa, align_path=align_path, stream_contexts=max(2, args.stream_contexts), stream_blocking=bool(args.blocking_sync == 1 or (args.blocking_sync < 0 and world * max(2, args.stream_contexts) > (os.cpu_count() or 1))), e2e3_ms=e2e3_ms, stage_ms_alone=stage_ms_alone, alone_ms=alone_ms, c4=c4, digest=digest, nt_all=nt_all, e2e2_ms=e2e2_ms, per_rank=per_rank, align=align, rank=rank, world=world, cfg=cfg, nq=nq, nq_all=nq_all, nt_sum=nt_sum, dev_ms=dev_ms, wall_ms=wall_ms, e2e_ms=e2e_ms,
                e2e_dev=e2e_dev, stage_ms=stage_ms, launches=launches_all, found=found_all, clocks=clocks, h2d=h2d_bytes,
                d2h=d2h_bytes, windows=seq_bytes, n_templates=w.count("templates"),
                n_bundles=w.count("active_bundles"))


# ---- protein alignment leg (BASELINE.json configs[4]; metric's "alignment GCUPS vs host-CPU ref") ------------
AA_FREQ = [8.25, 5.53, 4.06, 5.45, 1.37, 3.93, 6.75, 7.07, 2.27, 5.96, 9.66, 5.84, 2.42, 3.86, 4.70, 6.56, 5.34, 1.08, 2.92, 6.87]


def c5_pairs(n_pairs, seed):
    """Seeded C5-law pairs: 99 % of lengths log-normal (median 400 aa, sigma_ln 0.9) clipped to
    [50, 35000], 1 % log-uniform in [5000, 35000], at least n/10000 titin-class pairs (30-35 k aa); residues
    i.i.d. with proteome background frequencies; the partner is a mutated copy (5-40 % substitutions, up to two
    deletions of 1-11 residues).  Returns aa (0..19), off, pair_q, pair_t."""
    import numpy as np
    rng = np.random.default_rng(seed)
    L = np.clip(np.exp(rng.normal(np.log(400.0), 0.9, n_pairs)), 50, 35000).astype(np.int64)
    tail = rng.random(n_pairs) < 0.01
    L[tail] = np.exp(rng.uniform(np.log(5000.0), np.log(35000.0), int(tail.sum()))).astype(np.int64)
    n_titin = max(1, n_pairs // 10000)
    L[rng.choice(n_pairs, n_titin, replace=False)] = rng.integers(30000, 35001, n_titin)
    freq = np.array(AA_FREQ) / sum(AA_FREQ)
    cdf = np.cumsum(freq)
    total = int(L.sum())
    q = np.minimum(np.searchsorted(cdf, rng.random(total, dtype=np.float32)), 19).astype(np.uint8)
    qoff = np.concatenate([[0], np.cumsum(L)])
    rate = np.repeat(rng.uniform(0.05, 0.4, n_pairs).astype(np.float32), L)
    sub = rng.random(total, dtype=np.float32) < rate
    t = q.copy()
    t[sub] = np.minimum(np.searchsorted(cdf, rng.random(int(sub.sum()), dtype=np.float32)), 19).astype(np.uint8)
    idx = np.arange(total, dtype=np.int64) - np.repeat(qoff[:-1], L)
    keep = np.ones(total, dtype=bool)
    for _ in range(2):
        p0 = np.repeat((rng.random(n_pairs) * L).astype(np.int64), L)
        k = np.repeat(rng.integers(0, 12, n_pairs), L)
        keep &= ~((idx >= p0) & (idx < p0 + k) & (idx > 0))
    tl = np.add.reduceat(keep.astype(np.int64), qoff[:-1])
    t = t[keep]
    aa = np.concatenate([q, t])
    lens = np.empty(2 * n_pairs, dtype=np.int64)
    lens[:n_pairs] = L
    lens[n_pairs:] = tl
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    pq = np.arange(n_pairs, dtype=np.int32)
    return aa, off, pq, pq + np.int32(n_pairs)


def alignment_leg(cx, n_pairs, cpu_seconds=12.0):
    """GCUPS of orfx_align_batch on a C5-law sample, against the measured integer-ALU issue rate; the scalar
    restatement of the same recurrence (oracle, OpenMP over pairs) on a subsample as the CPU baseline."""
    import ctypes as C
    import numpy as np
    aa, off, pq, pt = c5_pairs(n_pairs, 20261104)
    ql = (off[pq + 1] - off[pq]).astype(np.float64)
    tl = (off[pt + 1] - off[pt]).astype(np.float64)
    cx.align(aa, off, pq[:64], pt[:64])   # warm-up
    best = None
    for _ in range(3):
        out, cells, ms = cx.align(aa, off, pq, pt)
        if best is None or ms < best[2]:
            best = (out, cells, ms)
    out, cells, ms = best
    peak = cx.int_alu_peak()
    gcups = cells / (ms * 1e-3) / 1e9
    # per power-of-two bucket of max(qlen, tlen)
    buckets = {}
    mx = np.maximum(ql, tl)
    for lo in (32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768):
        sel = np.nonzero((mx >= lo) & (mx < 2 * lo))[0]
        if len(sel) == 0:
            continue
        _, c, m = cx.align(aa, off, pq[sel], pt[sel])
        buckets["%d-%d" % (lo, 2 * lo - 1)] = {"pairs": int(len(sel)), "gcups": c / (m * 1e-3) / 1e9 if m > 0 else None}
    line = {"workload": "C5-law protein pairs (BASELINE.json configs[4]), %d pairs, lengths 50-35000 aa" % n_pairs,
            "pairs": int(n_pairs), "cells": cells, "kernel_ms": ms, "gcups": gcups,
            "int_alu_peak_ops_per_s": peak, "ops_per_cell": 10,
            # Roofline at the instruction level: the recurrence costs 4.5 ALU-pipe instructions per cell in this formulation
            # (VIMNMX3 + VIADD + 2 VIADDMNMX + half a VIMNMX3; the substitution add runs on the FMA pipe as IDP.4A), against
            # the ALU-pipe issue rate measured on this GPU (VIADDMNMX chains, all SMs) -- ncu's sm__pipe_alu_cycles_active
            # is the same reading.  The "canonical" 10-operations-per-cell count of SURVEY.md 8d is kept beside it: fused
            # add-max / max3 instructions do two of those operations each, so that ratio can exceed 1.
            "roofline": {"bound": "int-alu", "achieved": 4.5 * gcups, "peak": peak["viaddmnmx"] / 2e9, "unit": "Ginst/s",
                         "frac": 4.5 * gcups * 1e9 / (peak["viaddmnmx"] / 2) if peak["viaddmnmx"] else None, "inst_per_cell": 4.5,
                         "peak_source": "measured on this GPU (orfx_int_alu_peak: VIADDMNMX issue rate, all SMs)"},
            "canonical_ops": {"ops_per_cell": 10, "achieved_gops": 10.0 * gcups, "iadd3_peak_gops": peak["iadd3"] / 1e9,
                              "ratio": 10.0 * gcups * 1e9 / peak["iadd3"] if peak["iadd3"] else None},
            "gcups_by_length": buckets,
            "checksum": int(np.bitwise_xor.reduce(out["score"].astype(np.int64) * 31 + out["max"]))}
    # CPU baseline: same recurrence, scalar int32, all cores, bounded subsample
    if cpu_seconds <= 0:
        return line
    try:
        lib = C.CDLL(os.path.join(ROOT, "oracle", "liborfx_oracle.so"))
        lib.orfx_oracle_align_batch.restype = C.c_int
        mat = np.ascontiguousarray(np.load(os.path.join(ROOT, "tests", "golden", "mat50.npy")), dtype=np.int8)
        order = np.arange(n_pairs)
        budget_cells = cpu_seconds * 0.25e9 * (os.cpu_count() or 1)   # ~0.25 GCUPS per core, scalar
        csum = np.cumsum(ql * tl)
        k = max(8, int(np.searchsorted(csum, budget_cells)))
        k = min(k, n_pairs)
        sel = order[:k]
        res = [np.zeros(k, dtype=np.int32) for _ in range(4)]
        sq, st = np.ascontiguousarray(pq[sel]), np.ascontiguousarray(pt[sel])
        t0 = time.perf_counter()
        used = lib.orfx_oracle_align_batch(C.c_void_p(aa.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(sq.ctypes.data),
                                           C.c_void_p(st.ctypes.data), C.c_int64(k), C.c_void_p(mat.ctypes.data), 24, 12, 2, 0,
                                           *[C.c_void_p(r.ctypes.data) for r in res])
        dt = time.perf_counter() - t0
        same = all(np.array_equal(res[i], out[n][:k]) for i, n in enumerate(("max", "score", "mqe", "mte")))
        line["cpu_baseline"] = {"value": float(csum[k - 1]) / dt / 1e9, "unit": "GCUPS", "cores": int(used), "kind": "port",
                                "sample": "first %d pairs of the same set (%.3g cells); scalar int32 restatement, OpenMP over pairs"
                                          % (k, float(csum[k - 1])), "identical_to_gpu": bool(same)}
    except Exception as e:
        line["cpu_baseline"] = {"value": None, "unit": "GCUPS", "sample": "unavailable: %s" % e}
    return line


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def profiled_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture of this workload (profiles/r02_traffic.json,
    written by tools/ncu_summary.py from a capture of THIS build: the file records the digest of the CUDA sources it was
    taken from and is refused when that differs from the sources next to this script)."""
    try:
        doc = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if doc.get("source_digest") != source_digest():
            return None, "profiles/r02_traffic.json was captured from other sources (digest %s, these are %s): refused" % (doc.get("source_digest"), source_digest())
        t = doc["kernels"][kernel]
        return t["dram_bytes_per_launch"], doc.get("source")
    except Exception as e:
        return None, "no usable profiles/r02_traffic.json (%s)" % e


def emit(line):
    """The one JSON line goes to the real stdout; everything else that libraries print to fd 1 (NCCL's
    version banner under torchrun, for one) was sent to stderr by quiet_stdout()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


_REAL_STDOUT = None


def quiet_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--loci", type=int, default=0, help="override the number of loci (development only; 0 = configs[1])")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--verbose-shard", action="store_true", help="log the largest loci of this rank's shard")
    ap.add_argument("--seed-offset", type=int, default=0, help="generate the shard rank k of a multi-GPU run would get (k = offset)")
    ap.add_argument("--no-pipelined", action="store_true", help="skip the multi-context end-to-end leg")
    ap.add_argument("--blocking-sync", type=int, default=-1, help="streamed leg: 1 = waiting threads sleep (blocking_sync), 0 = spin, "
                    "-1 = sleep when ranks x contexts exceed the host's cores")
    ap.add_argument("--stream-contexts", type=int, default=3, help="contexts (one caller thread each) of the streamed end-to-end leg")
    ap.add_argument("--e2e-contexts", type=int, default=2, help="contexts (locus shards) of the pipelined end-to-end leg")
    ap.add_argument("--align-pairs", type=int, default=-1, help="protein pairs of the alignment leg per GPU (0 = skip; default: BASELINE.json configs[4]'s "
                    "1 000 000 pairs on one GPU, 125 000 per GPU under torchrun)")
    ap.add_argument("--config", default="c2", choices=["c2", "c4"], help="c2 = BASELINE.json configs[1] (the driver's metric; weak scaling); "
                    "c4 = configs[3], one 2 M-query workload sharded by locus over the GPUs (strong scaling)")
    ap.add_argument("--no-align-path", action="store_true", help="skip the leg that runs the path with the alignment stage enabled")
    ap.add_argument("--no-numa", action="store_true", help="multi-GPU runs: do not bind the rank to its GPU's NUMA node")
    ap.add_argument("--no-cli", action="store_true", help="skip the whole-process CLI leg")
    ap.add_argument("--only-align", action="store_true", help="development: run the alignment leg alone and print its object")
    args = ap.parse_args()
    quiet_stdout()
    if args.align_pairs < 0:
        args.align_pairs = 1000000 if int(os.environ.get("WORLD_SIZE", "1")) == 1 else 125000
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        reference_arm(args)
        return
    if args.only_align:
        import orfanage_b200 as ob
        cx = ob.Context(ob.Params.make(mode=MODE, want_rows=False), device=int(os.environ.get("LOCAL_RANK", "0")))
        emit(alignment_leg(cx, args.align_pairs, cpu_seconds=0.0 if args.no_cpu_baseline else 12.0))
        return
    m = ours(args)
    if m["rank"] != 0:
        return
    K = args.steps
    ms_per_step = m["dev_ms"] / K
    value = m["nq_all"] / (ms_per_step / 1e3)
    e2e_value = m["nq_all"] / (m["e2e_ms"] / K / 1e3)
    e2e_api = "orfx_submit + orfx_wait on a pinned host batch (include/orfx.h)"
    e2e_single = {"value": e2e_value, "ms_per_step": m["e2e_ms"] / K}
    e2e_step_ms = m["e2e_ms"] / K
    if m.get("e2e2_ms"):   # several contexts, a share of the loci each: same calls, same bytes; uploads overlap kernels
        v2 = m["nq_all"] / (m["e2e2_ms"] / K / 1e3)
        nc = max(2, args.e2e_contexts)
        if v2 > e2e_value:
            e2e_value, e2e_step_ms = v2, m["e2e2_ms"] / K
            e2e_api = ("orfx_submit + orfx_wait on pinned host batches (include/orfx.h), %d contexts per GPU with a locus shard each, "
                       "submitted back to back: the upload of one shard overlaps the kernels of another" % nc)
        e2e_single["pipelined_contexts"] = {"contexts": nc, "value": v2, "ms_per_step": m["e2e2_ms"] / K}
    if m.get("e2e3_ms"):   # a stream of batches over two contexts
        v3 = m["nq_all"] / (m["e2e3_ms"] / K / 1e3)
        if v3 > e2e_value:
            e2e_value, e2e_step_ms = v3, m["e2e3_ms"] / K
            e2e_api = ("orfx_submit + orfx_wait on pinned host batches (include/orfx.h) as a stream of batches over %d contexts, one caller thread each: batch k + 1 is "
                       "uploaded while batch k is evaluated; every batch is uploaded and its results read back inside the timed region") % m.get("stream_contexts", 2)
        e2e_single["streamed"] = {"contexts": m.get("stream_contexts", 2), "host_wait": "sleep (blocking_sync)" if m.get("stream_blocking") else "spin", "value": v3, "ms_per_step": m["e2e3_ms"] / K}
    peak, peak_src = measured_peak()
    c4 = m["c4"]
    nt_per_launch = m["nt_sum"]          # spliced query nucleotides one rank's codon_scan launch covers
    scan_in_step_ms = m["stage_ms"].get("codon_scan", 0.0) / K
    scan_ms = m["stage_ms_alone"].get("codon_scan", 0.0) / K      # one launch over all transcripts, nothing beside it
    planes_ms = m["stage_ms_alone"].get("base_planes", 0.0) / K
    alg_bytes = ALG_BYTES_PER_NT * nt_per_launch
    achieved = alg_bytes / (scan_ms / 1e3) / 1e9 if scan_ms > 0 else 0.0
    stage_achieved = alg_bytes / ((scan_ms + planes_ms) / 1e3) / 1e9 if scan_ms + planes_ms > 0 else 0.0
    traffic, traffic_src = profiled_traffic("codon_scan_kernel")
    if c4:   # the capture is one configs[1] launch: it says nothing about a launch eight times that size
        traffic, traffic_src = None, "not captured for this workload (profiles/r02_traffic.json is a configs[1] launch)"
    dominant = max(m["stage_ms"], key=lambda k: m["stage_ms"][k]) if m["stage_ms"] else None
    line = {
        "metric": "query transcripts/sec", "value": value, "unit": "transcripts/s", "n_gpus": m["world"], "steps": K,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if c4 else "weak", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic",
        "config": {"workload": "long-read novel-isoform scale synthetic (BASELINE.json configs[3]): one workload, loci dealt to the GPUs by shard_bundles" if c4
                   else "GENCODE-scale synthetic (BASELINE.json configs[1])", "mode": MODE,
                   "query_transcripts_per_gpu": m["nq"], "query_transcripts_total": m["nq_all"], "templates_after_dedup": m["n_templates"], "bundles": m["n_bundles"],
                   "best_digest": m["digest"],
                   "genome": "%d x %d bp random, planted ORFs, seed %d%s" % (m["cfg"]["chroms"], m["cfg"]["chrom_len"], m["cfg"]["seed"], "" if c4 else "+rank"),
                   "spliced_query_nt_per_gpu": m["nt_sum"], "sequence_upload_bytes_per_gpu": m["windows"], "sequence_format": "islands, 2-bit base codes + exception list (ORFX_SEQ_NT2)", "chain_format": "segment dictionary (orfx_batch::seg_ref): 16-bit reference per segment + the distinct records of each locus",
                   "parallelism": "locus shards, one per GPU, no collective", "orfs_found": m["found"],
                   "l2": "inputs larger than L2 (windows + chains + masks > 126 MB); no explicit flush",
                   "timing": "CUDA events on the library stream around the whole path, summed over steps, max over ranks",
                   "wall_ms_per_step": m["wall_ms"] / K},
        "e2e": {"value": e2e_value, "unit": "transcripts/s", "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                "ms_per_step": e2e_step_ms, "device_ms_per_step": {"h2d": m["e2e_dev"][0] / K, "kernels": m["e2e_dev"][1] / K,
                                                                   "d2h": m["e2e_dev"][2] / K},
                "single_context": e2e_single, "api": e2e_api},
        "gpu_launches": m["launches"],
        "clocks": m["clocks"],
        # `achieved` / `frac` follow the contract: ALGORITHMIC bytes (1.25 B per spliced nucleotide) over the kernel's
        # launch time against the measured HBM copy bandwidth.  That is a streaming-equivalent rate, not DRAM utilisation: the
        # kernel reads bit planes, so its real DRAM traffic (`traffic`, ncu) is a fraction of the algorithmic bytes (`dram_frac`)
        # and what limits it is instruction issue on the ALU pipe (`limiter`).  `stage` is the same figure for the whole
        # bases -> masks stage (plane build + scan).
        "roofline": {"bound": "hbm", "kernel": "codon_scan_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak if peak else None, "algorithmic_frac": achieved / peak if peak else None,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "dram_frac": (traffic / (scan_ms / 1e3) / 1e9 / peak) if (traffic and scan_ms > 0 and peak) else None,
                     "limiter": "instruction issue (ALU pipe), not DRAM: see profiles/r02_ncu_full.md",
                     "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": scan_ms,
                     "kernel_ms_source": "CUDA events around the codon_scan_kernel launch, mean of %d launches on the resident batch with every kernel alone on "
                                         "one stream (orfx_debug_set serialize); `in_step` is the same stage inside the timed steps" % K,
                     "in_step": {"ms": scan_in_step_ms, "achieved": alg_bytes / (scan_in_step_ms / 1e3) / 1e9 if scan_in_step_ms > 0 else None,
                                 "frac": (alg_bytes / (scan_in_step_ms / 1e3) / 1e9 / peak) if (scan_in_step_ms > 0 and peak) else None,
                                 "note": "two launches (heavy bundles' transcripts first), the heavy bundles' rescue kernel running beside the second"},
                     "share_of_step": scan_in_step_ms / ms_per_step if ms_per_step else None,
                     "stage": {"kernels": "base_planes_nt2_kernel + plane_exceptions_kernel + codon_scan_kernel", "ms": scan_ms + planes_ms, "achieved": stage_achieved,
                               "frac": stage_achieved / peak if peak else None},
                     "dominant_stage": dominant,
                     "dominant": {"stage": dominant, "ms": m["stage_ms"][dominant] / K if dominant else None, "bound": "latency (dependent chain walks and mask scans; "
                                  "low DRAM and SM throughput, see profiles/r02_ncu_full.md)", "share_of_step": (m["stage_ms"][dominant] / K) / ms_per_step if dominant else None}},
        "stages_ms": {k: v / K for k, v in m["stage_ms"].items()},
        "stages_ms_alone": {k: v / K for k, v in m["stage_ms_alone"].items()},
        "ms_per_step_serialized": m["alone_ms"] / K,
        "stages_note": "heavy_bundles (CTA-per-bundle kernels of the titin-class loci) runs on a side stream beside fill_rows / rescue_score: the stages overlap and do not add up to the step",
    }
    if m["align"] is not None:
        line["alignment"] = m["align"]
    if m.get("align_path"):
        line["alignment_in_path"] = m["align_path"]
    if m.get("numa"):
        line["config"]["numa_binding_rank0"] = m["numa"]
    if m.get("per_rank"):
        line["per_rank"] = m["per_rank"]
        dm = m["per_rank"]["device_ms_per_step"]
        line["per_rank"]["imbalance_max_over_mean"] = max(dm) / (sum(dm) / len(dm)) if dm and sum(dm) > 0 else None
    if m["world"] == 1 and not c4 and not args.no_cli:
        try:
            line["cli"] = cli_leg(generate(dict(C2), "c2"))
        except Exception as e:
            line["cli"] = {"error": str(e)}
    if m["world"] == 1 and not args.no_cpu_baseline and not c4:
        try:
            line["cpu_baseline"] = cpu_baseline_leg()
        except Exception as e:   # the checker being absent must not hide the GPU number
            line["cpu_baseline"] = {"value": None, "unit": "transcripts/s", "cores": os.cpu_count(), "kind": "reference",
                                    "sample": "unavailable: %s" % e}
    emit(line)


if __name__ == "__main__":
    main()
