#!/usr/bin/env python3
"""bench.py -- throughput of the select_STR_reads hot path on B200 (reads/s, Gbases/s).

Workload at N GPUs (weak scaling): BASELINE.json configs[1] on every GPU -- 10 M synthetic
100 bp reads with planted 2-6mer STRs (5 M pairs: mate-1 reads then mate-2 reads), default
flags (-n 2 -f 31, 2..6mers) -- held as 1 M-read raw-FASTQ batches.  One step = one pass of
the whole GPU pipeline (line index -> scan -> ordered compaction -> format) over all of a
GPU's batches.

  value  kernel pipeline with the raw FASTQ resident in HBM (CUDA events on the launch
         stream; the 2.2 GB a step reads is far larger than the 126 MB L2, no flush needed)
  e2e    the same job through the C-ABI call bstr_scan_batch from pinned HOST buffers:
         H2D copy + kernels + D2H copy of the annotated FASTQ inside the timed region
  --impl reference   the unmodified reference (oracle/_ref, CPython + Cython + re) with
         -p <all cores> on a bounded prefix of the same workload, on the host CPU

One JSON line on stdout (rank 0).  torch is used only for the multi-rank barrier / max.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG, READ_LEN = 2, 100
READS_PER_GPU = 10_000_000
BATCH_READS = 1_000_000
FLAGS = dict(n=2, f=31, ks=None, rmhm=False, altalgo=True, illumina=False)
WORKLOAD = ("C2: synthetic 10M paired 100bp reads with planted 2-6mer STRs, "
            "default flags (-n 2 -f 31, 2-6mers), per GPU")
# the other BASELINE.json configs, for sweeps (--workload); the bench line of record is C2
WORKLOADS = {
    "C2": (2, 100, dict(FLAGS), WORKLOAD),
    "C3": (3, 150, dict(FLAGS, n=3, rmhm=True), "C3: synthetic 150bp reads, -n 3 -r, per GPU"),
    "C4": (4, 250, dict(FLAGS, ks=[2, 3], f=41), "C4: synthetic 250bp reads, --2mer --3mer -f 41, per GPU"),
    "C5": (5, 150, dict(FLAGS), "C5: synthetic 150bp reads, default flags, per GPU"),
}
METRIC = "reads/s scanned (select_STR_reads hot path: parse+scan+compact+format)"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML; nvidia-smi as fallback)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                     nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.05)
        except Exception as e:      # no NVML: record that, never fail the bench
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def bind_to_gpu_node(index):
    """Run this rank (and allocate its pinned batch buffers) on the CPUs NVML reports as local to
    its GPU, so that H2D/D2H traffic does not cross the socket interconnect."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        words = nv.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def dist_setup(ngpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    bind_to_gpu_node(local)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        return world, rank, local, dist
    return 1, 0, 0, None


def barrier(dist):
    if dist is not None:
        import torch
        dist.barrier()
        torch.cuda.synchronize()


def max_over_ranks(dist, x):
    from baitstr_b200 import shard
    return shard.reduce_scalar(dist, x, "max")


def sum_over_ranks(dist, x):
    from baitstr_b200 import shard
    return shard.reduce_scalar(dist, x, "sum")


def make_batches(rank, reads_per_gpu, batch_reads):
    """This rank's share of the workload as pinned raw-FASTQ batches (mate 1 then mate 2)."""
    from baitstr_b200 import capi, shard, synth
    rb = synth.record_bytes(CFG, READ_LEN)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    batches = []
    for mate in (1, 2):
        first, left = shard.read_range(world * (reads_per_gpu // 2), rank, world)
        while left > 0:
            cnt = min(batch_reads, left)
            buf = capi.PinnedBuffer(rb * cnt)
            synth.generate_into(buf.array, CFG, READ_LEN, first, cnt, mate)
            batches.append(buf)
            first += cnt
            left -= cnt
    return batches, rb


def run_ours(args):
    world, rank, local, dist = dist_setup(args.gpus)
    from baitstr_b200 import capi
    ctx = capi.Context(devices=[local], **FLAGS)
    batches, rb = make_batches(rank, args.reads, args.batch_reads)
    nreads = sum(b.nbytes for b in batches) // rb
    in_bytes = sum(b.nbytes for b in batches)

    # ---- kernel-only: inputs resident in HBM, CUDA events inside the library ------------
    sampler = ClockSampler(local)
    barrier(dist)
    sampler.start()
    res = ctx.bench_resident(batches, args.steps, args.warmup, flush_l2=False)
    clocks = sampler.finish()
    barrier(dist)
    ms_step = max_over_ranks(dist, res["ms_total"] / args.steps)
    total_reads = sum_over_ranks(dist, float(res["n_records"]))
    value = total_reads / (ms_step * 1e-3)

    # ---- end to end through the C ABI from pinned host buffers ---------------------------
    # bstr_submit_batch / bstr_collect_batch: every lane (3 per GPU) holds one sub-batch in
    # flight, so H2D, kernels and D2H of neighbouring sub-batches overlap; outputs land in
    # page-locked host memory in input order.
    nl = ctx.lanes()
    sub = max(1, args.e2e_batch_reads) * rb
    jobs = [(b, off, min(sub, b.nbytes - off)) for b in batches for off in range(0, b.nbytes, sub)]

    def e2e_pass():
        out_b, inflight = 0, []
        for i, (b, off, nb) in enumerate(jobs):
            if len(inflight) == nl:
                out_b += ctx.collect(inflight.pop(0), copy=False)[1]["out_bytes"]
            ctx.submit(i % nl, b, final=True, nbytes=nb, offset=off)
            inflight.append(i % nl)
        for lane in inflight:
            out_b += ctx.collect(lane, copy=False)[1]["out_bytes"]
        return out_b
    for _ in range(min(args.warmup, 2)):
        e2e_pass()
    barrier(dist)
    e_steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e_steps):
        out_bytes = e2e_pass()
    barrier(dist)
    e2e_s = max_over_ranks(dist, (time.perf_counter() - t0) / e_steps)
    e2e_value = total_reads / e2e_s

    if rank != 0:
        ctx.close()
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel --------------------------------------------------
    peak, peak_src = peaks()
    kern = res["kernels_ms"]
    dom = max(kern, key=kern.get)
    sel_frac = res["n_selected"] / max(1, res["n_records"])
    a_scan = READ_LEN + 16                                   # bytes per read
    a_pipe = (in_bytes + res["out_bytes"]) / nreads          # R_in + sel * R_out
    per_read = a_pipe if dom.startswith("k_fused") else a_scan
    achieved = per_read * nreads / (kern[dom] * 1e-3) / 1e9
    # DRAM traffic of the dominant kernel from the committed ncu --set full capture (per launch =
    # one 1 M-read batch of this workload), scaled to the batch size used here
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")) as fh:
            t = json.load(fh)
        if args.workload == "C2":
            traffic = round((t["dram_bytes_read"] + t["dram_bytes_write"]) * args.batch_reads / t["reads_per_launch"])
    except Exception:
        pass
    launches = len(batches)                                  # k_fused_tile launches per step
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "algorithmic_bytes_per_read": round(per_read, 1),
                "algorithmic_bytes_per_launch": round(per_read * nreads / launches),
                "kernel_us_per_launch": round(kern[dom] * 1e3 / launches, 1),
                "peak_source": peak_src,
                "kernel_ms_per_step": {k: round(v, 4) for k, v in kern.items()},
                "pipeline_bytes_per_read": round(a_pipe, 1),
                "pipeline_frac_of_hbm": round(a_pipe * nreads / (res["ms_total"] / args.steps * 1e-3) / 1e9 / peak, 4),
                "scan_only_bytes_per_read": a_scan,
                "scan_only_frac_of_hbm": round(a_scan * nreads / (kern[dom] * 1e-3) / 1e9 / peak, 4),
                "note": "achieved = (raw record bytes read + annotated bytes written) per launch / kernel time "
                        "(A_pipe of SURVEY.md 8d; the fused kernel parses, scans, compacts and formats in one pass); "
                        "scan_only_* counts L + 16 bytes per read (A_scan)"}

    # ---- CPU baseline (oracle port, 1 core) on a bounded sample + parity check -------------
    cpu = None
    if world == 1 and not args.no_cpu:
        from oracle import oracle
        sample = min(args.cpu_sample, args.batch_reads, nreads)
        raw = batches[0].array[:sample * rb].tobytes()
        t0 = time.perf_counter()
        want = oracle.select_buffer(raw, **FLAGS)
        dt = time.perf_counter() - t0
        got, _ = ctx.scan_batch(raw)
        cpu = {"value": round(sample / dt, 1), "unit": "reads/s", "cores": 1, "kind": "port",
               "sample": "first %d reads of the workload, oracle/str_oracle.c single thread" % sample,
               "parity_vs_gpu": "byte-identical" if got == want else "MISMATCH"}
        if got != want:
            print("bench.py: GPU output differs from the oracle on the sample", file=sys.stderr)
            sys.exit(1)

    # ---- file ingest (host side of bstr_run_files) on a bounded sample: plain and .gz ---------
    ingest = None
    if world == 1 and not args.no_ingest:
        ingest = measure_ingest(args.ingest_reads, local)

    line = {
        "metric": METRIC, "value": round(value, 1), "unit": "reads/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "reads_per_gpu": nreads, "read_len": READ_LEN, "batch_reads": args.batch_reads,
                   "raw_fastq_bytes_per_gpu": in_bytes, "selected_fraction": round(sel_frac, 4),
                   "l2_policy": "inputs larger than L2 (%.2f GB read per step), no flush" % (in_bytes / 1e9),
                   "parallelism": "reads sharded across GPUs, no collective"},
        "gbases_per_s": round(value * READ_LEN / 1e9, 3),
        "e2e": {"value": round(e2e_value, 1), "unit": "reads/s", "h2d_bytes_per_step": in_bytes,
                "d2h_bytes_per_step": int(out_bytes), "gbases_per_s": round(e2e_value * READ_LEN / 1e9, 3),
                "path": "bstr_submit_batch/bstr_collect_batch, %d lanes, %d sub-batches of %d reads per step "
                        "(pinned host in -> pinned host out)" % (nl, len(jobs), args.e2e_batch_reads)},
        "gpu_launches": res["launches_per_step"] * args.steps,
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "ingest": ingest,
        "out_checksum_fnv1a": "%016x" % res["out_checksum"],
    }
    print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def measure_ingest(nreads, device):
    """FILES -> bstr_run_files -> /dev/null for a prefix of the workload, as plain FASTQ and as
    gzip -1 (the C4/C5 input format): what the reader / parallel inflater / stitcher sustain when
    the GPU is not the limit.  Not part of `value` or `e2e`."""
    import shutil
    from baitstr_b200 import capi, synth
    tmp = tempfile.mkdtemp(prefix="bstr_ingest_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    cores = len(os.sched_getaffinity(0))
    out = {"reads": nreads, "host_threads": cores, "sink": "/dev/null", "source": "tmpfs, two mate files"}
    try:
        plain = []
        for mate in (1, 2):
            p = os.path.join(tmp, "m%d.fq" % mate)
            synth.generate(CFG, READ_LEN, 0, nreads // 2, mate).tofile(p)
            plain.append(p)
        procs = [subprocess.Popen("gzip -1 -c %s > %s.gz" % (p, p), shell=True) for p in plain]
        if any(pr.wait() for pr in procs):
            raise RuntimeError("gzip failed")
        ctx = capi.Context(devices=[device], host_threads=cores, **FLAGS)
        with open(os.devnull, "wb") as dn:
            for kind, files in (("plain", plain), ("gz", [p + ".gz" for p in plain])):
                best = None
                for _ in range(3):                                   # first pass allocates the page-locked buffers
                    t0 = time.perf_counter()
                    ctx.run_files(files, out_fd=dn.fileno(), err_fd=dn.fileno())
                    dt = time.perf_counter() - t0
                    best = dt if best is None else min(best, dt)
                out[kind] = {"reads_per_s": round(nreads / best, 1), "file_GB_per_s": round(sum(os.path.getsize(f) for f in files) / best / 1e9, 3),
                             "ms": round(best * 1e3, 1)}
        ctx.close()
    except Exception as e:                                           # never lose the bench line over the side measurement
        out["error"] = str(e)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def run_reference(args):
    """The unmodified reference select stage (oracle/_ref) on the host cores, -p <all cores>,
    on a bounded prefix of the same workload.  Rank 0 only; other ranks exit at once."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from baitstr_b200 import synth
    cores = os.cpu_count() or 1
    runner = os.path.join(ROOT, "oracle", "ref_select.py")
    rb = synth.record_bytes(CFG, READ_LEN)
    tmpdir = tempfile.mkdtemp(prefix="bstr_ref_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)

    def one_pass(nreads):
        half = nreads // 2
        paths = []
        for mate in (1, 2):
            p = os.path.join(tmpdir, "syn2_%d.fq" % mate)
            synth.generate(CFG, READ_LEN, 0, half, mate).tofile(p)
            paths.append(p)
        t0 = time.perf_counter()
        with open(os.devnull, "wb") as dn:
            subprocess.run([sys.executable, runner, "-p", str(cores)] + paths, stdout=dn,
                           stderr=dn, check=True)
        return time.perf_counter() - t0, 2 * half

    try:
        dt, n = one_pass(40000)                       # calibration (untimed)
        rate = n / dt
        budget = 200.0 / max(1, args.steps + args.warmup)
        sample = int(max(40000, min(2_000_000, rate * min(budget, 15.0))))
        for _ in range(args.warmup):
            one_pass(sample)
        total, reads = 0.0, 0
        for _ in range(args.steps):
            dt, n = one_pass(sample)
            total += dt
            reads += n
    finally:
        for f in os.listdir(tmpdir):
            os.unlink(os.path.join(tmpdir, f))
        os.rmdir(tmpdir)
    value = reads / total
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 1), "unit": "reads/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(total / args.steps * 1e3, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "read_len": READ_LEN,
                   "note": "each step = the first %d reads of the workload (prefix), files in tmpfs, "
                           "stdout to /dev/null; interpreter CPython %d.%d + Cython shim build "
                           "(Python 2 unavailable)" % (sample, sys.version_info[0], sys.version_info[1])},
        "gbases_per_s": round(value * READ_LEN / 1e9, 6),
        "cpu_baseline": {"value": round(value, 1), "unit": "reads/s", "cores": cores, "kind": "reference",
                         "sample": "first %d reads per step, select_STR_reads -p %d" % (sample, cores)},
        "e2e": {"value": round(value, 1), "unit": "reads/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=READS_PER_GPU, help="reads per GPU")
    ap.add_argument("--batch-reads", type=int, default=BATCH_READS)
    ap.add_argument("--cpu-sample", type=int, default=1_000_000)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ingest", action="store_true")
    ap.add_argument("--ingest-reads", type=int, default=8_000_000)
    ap.add_argument("--e2e-batch-reads", type=int, default=1_000_000)
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    args = ap.parse_args()
    global CFG, READ_LEN, FLAGS, WORKLOAD
    CFG, READ_LEN, FLAGS, WORKLOAD = WORKLOADS[args.workload]
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
