#!/usr/bin/env python3
"""bench.py -- throughput of the select_STR_reads hot path on B200 (reads/s, Gbases/s).

Workload of record at N GPUs (weak scaling): BASELINE.json configs[1] (C2) on every GPU -- 10 M
synthetic 100 bp reads with planted 2-6mer STRs (5 M pairs: mate-1 reads then mate-2 reads),
default flags (-n 2 -f 31, 2..6mers) -- held as 1 M-read raw-FASTQ batches.  One step = one pass
of the whole GPU pipeline (line index -> scan -> ordered compaction -> format) over all of a
GPU's batches.

  value        kernel pipeline with the raw FASTQ resident in HBM (CUDA events on the launch
               stream; the 2.2 GB a step reads is far larger than the 126 MB L2, no flush needed)
  e2e          the same job through the C ABI (bstr_submit_batch / bstr_collect_batch) from pinned
               HOST buffers: H2D copy + kernels + D2H copy of the annotated FASTQ in the timed region
  pcie_probe   the e2e leg's copies alone (bstr_pcie_probe: same buffers, same lanes, no kernels):
               the ceiling of e2e on this box, per N
  c3           the same three numbers for BASELINE.json configs[2] (C3: 150 bp, -n 3 -r; the config
               named for 1/2/4/8 GPUs), 12.5 M reads per GPU = the 100 M reads split over 8 GPUs
  other_configs (N = 1) C3, C4, C5: value, roofline fraction, e2e
  multi_gpu    (N > 1, rank 0) the in-process multi-GPU path of the product: ONE context with
               n_devices = N, bstr_run_files over N shard files, byte-compared with the 1-device
               output and with the oracle on a prefix; the bench fails on a mismatch
  parity_vs_reference (N = 1) GPU output cmp'd with the unmodified reference's output (oracle/_ref,
               -p <cores>) on a prefix of every config's workload, with that config's flags
  --impl reference   the unmodified reference (oracle/_ref, CPython + Cython + re) with
               -p <all cores> on a bounded prefix of the same workload, on the host CPU

One JSON line on stdout (rank 0).  torch is used only for the multi-rank barrier / max.
"""
import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

READS_PER_GPU = 10_000_000
BATCH_READS = 1_000_000
DEFAULT_FLAGS = dict(n=2, f=31, ks=None, rmhm=False, altalgo=True, illumina=False)
# name -> (synthetic config number, read length, flags, workload text, CLI flags of the reference)
WORKLOADS = {
    "C2": (2, 100, dict(DEFAULT_FLAGS),
           "C2: synthetic 10M paired 100bp reads with planted 2-6mer STRs, default flags (-n 2 -f 31, 2-6mers), per GPU", []),
    "C3": (3, 150, dict(DEFAULT_FLAGS, n=3, rmhm=True), "C3: synthetic 150bp reads, -n 3 -r, per GPU", ["-n", "3", "-r"]),
    "C4": (4, 250, dict(DEFAULT_FLAGS, ks=[2, 3], f=41), "C4: synthetic 250bp reads, --2mer --3mer -f 41, per GPU",
           ["--2mer", "--3mer", "-f", "41"]),
    "C5": (5, 150, dict(DEFAULT_FLAGS), "C5: synthetic 150bp reads, default flags, per GPU", []),
}
METRIC = "reads/s scanned (select_STR_reads hot path: parse+scan+compact+format)"
TRAFFIC_FILE = os.path.join("profiles", "r02_ncu_traffic.json")


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


class Dist:
    """torch.distributed plumbing: NCCL for the device barrier / scalar reductions of the timed
    legs, a gloo group for host-side waits (ranks that idle while rank 0 drives every GPU must
    not spin in an NCCL kernel on the GPUs rank 0 is using)."""

    def __init__(self):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = self.gloo = None
        bind_to_gpu_node(self.local)
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
            self.gloo = dist.new_group(backend="gloo")

    def barrier(self):
        if self.dist is not None:
            import torch
            self.dist.barrier()
            torch.cuda.synchronize()

    def host_barrier(self):
        if self.dist is not None:
            self.dist.barrier(group=self.gloo)

    def reduce(self, x, op):
        from baitstr_b200 import shard
        return shard.reduce_scalar(self.dist, x, op)

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def make_batches(cfg, read_len, rank, world, reads_per_gpu, batch_reads):
    """This rank's share of the workload as pinned raw-FASTQ batches (mate 1 then mate 2)."""
    from baitstr_b200 import capi, shard, synth
    rb = synth.record_bytes(cfg, read_len)
    batches, todo = [], []
    for mate in (1, 2):
        first, left = shard.read_range(world * (reads_per_gpu // 2), rank, world)
        while left > 0:
            cnt = min(batch_reads, left)
            buf = capi.PinnedBuffer(rb * cnt)
            todo.append((buf, first, cnt, mate))
            batches.append(buf)
            first += cnt
            left -= cnt
    # the generator is a pure function of the read index: fill the batches on several threads
    threads = [threading.Thread(target=lambda t=t: synth.generate_into(t[0].array, cfg, read_len, t[1], t[2], t[3]))
               for t in todo]
    width = max(1, min(8, len(os.sched_getaffinity(0))))
    for i in range(0, len(threads), width):
        for t in threads[i:i + width]:
            t.start()
        for t in threads[i:i + width]:
            t.join()
    return batches, rb


def measure_config(D, name, reads_per_gpu, batch_reads, steps, warmup, e2e_batch_reads, probe=True,
                   sample_clocks=False):
    """value / e2e / pcie_probe / roofline of one workload on this rank's GPU (all ranks call it)."""
    from baitstr_b200 import capi
    cfg, read_len, flags, workload, _ = WORKLOADS[name]
    ctx = capi.Context(devices=[D.local], **flags)
    batches, rb = make_batches(cfg, read_len, D.rank, D.world, reads_per_gpu, batch_reads)
    in_bytes = sum(b.nbytes for b in batches)
    nreads = in_bytes // rb

    # ---- kernel-only: inputs resident in HBM, CUDA events inside the library ------------
    sampler = ClockSampler(D.local) if sample_clocks else None
    D.barrier()
    if sampler:
        sampler.start()
    res = ctx.bench_resident(batches, steps, warmup, flush_l2=False)
    clocks = sampler.finish() if sampler else None
    D.barrier()
    ms_step = D.reduce(res["ms_total"] / steps, "max")
    total_reads = D.reduce(float(res["n_records"]), "sum")
    value = total_reads / (ms_step * 1e-3)

    # ---- end to end through the C ABI from pinned host buffers ---------------------------
    # bstr_submit_batch / bstr_collect_batch: every lane (3 per GPU) holds one sub-batch in
    # flight, so H2D, kernels and D2H of neighbouring sub-batches overlap; outputs land in
    # page-locked host memory in input order.
    nl = ctx.lanes()
    sub = max(1, e2e_batch_reads) * rb
    jobs = [(b, off, min(sub, b.nbytes - off)) for b in batches for off in range(0, b.nbytes, sub)]
    job_out = [0] * len(jobs)

    def e2e_pass():
        out_b, inflight = 0, []
        for i, (b, off, nb) in enumerate(jobs):
            if len(inflight) == nl:
                j = inflight.pop(0)
                job_out[j] = ctx.collect(j % nl, copy=False)[1]["out_bytes"]
                out_b += job_out[j]
            ctx.submit(i % nl, b, final=True, nbytes=nb, offset=off)
            inflight.append(i)
        for j in inflight:
            job_out[j] = ctx.collect(j % nl, copy=False)[1]["out_bytes"]
            out_b += job_out[j]
        return out_b
    for _ in range(max(1, min(warmup, 2))):
        e2e_pass()
    D.barrier()
    e_steps = max(1, min(steps, 5))
    t0 = time.perf_counter()
    for _ in range(e_steps):
        out_bytes = e2e_pass()
    D.barrier()
    e2e_s = D.reduce((time.perf_counter() - t0) / e_steps, "max")
    e2e_value = total_reads / e2e_s
    e2e = {"value": round(e2e_value, 1), "unit": "reads/s", "h2d_bytes_per_step": in_bytes,
           "d2h_bytes_per_step": int(out_bytes), "gbases_per_s": round(e2e_value * read_len / 1e9, 3),
           "h2d_GBps_all_gpus": round(D.reduce(float(in_bytes), "sum") / e2e_s / 1e9, 2),
           "d2h_GBps_all_gpus": round(D.reduce(float(out_bytes), "sum") / e2e_s / 1e9, 2),
           "path": "bstr_submit_batch/bstr_collect_batch, %d lanes, %d sub-batches of %d reads per step "
                   "(pinned host in -> pinned host out)" % (nl, len(jobs), e2e_batch_reads)}

    # ---- the copies of the e2e leg alone: same buffers, same lanes, same order, no kernels ----
    pcie = None
    if probe:
        pcie = {"what": "bstr_pcie_probe: the e2e leg's cudaMemcpyAsync calls without the kernels, all ranks at once; "
                        "GB/s summed over the GPUs", "reads_per_s_ceiling": None}
        sums = {"h2d": D.reduce(float(in_bytes), "sum"), "d2h": D.reduce(float(out_bytes), "sum")}
        for key, mode in (("h2d_only", 1), ("d2h_only", 2), ("both", 3)):
            ctx.pcie_probe(jobs, job_out, mode, 1)
            D.barrier()
            ms = D.reduce(ctx.pcie_probe(jobs, job_out, mode, 2) / 2, "max")
            D.barrier()
            if mode & 1:
                pcie[key + "_h2d_GBps"] = round(sums["h2d"] / ms / 1e6, 2)
            if mode & 2:
                pcie[key + "_d2h_GBps"] = round(sums["d2h"] / ms / 1e6, 2)
            if mode == 3:
                pcie["reads_per_s_ceiling"] = round(total_reads / (ms * 1e-3), 1)
                pcie["e2e_over_ceiling"] = round(e2e_value / (total_reads / (ms * 1e-3)), 3)

    # ---- roofline of the dominant kernel --------------------------------------------------
    peak, peak_src = peaks()
    kern = res["kernels_ms"]
    dom = max(kern, key=kern.get)
    sel_frac = res["n_selected"] / max(1, res["n_records"])
    a_scan = read_len + 16                                   # bytes per read
    a_pipe = (in_bytes + res["out_bytes"]) / nreads          # R_in + sel * R_out
    per_read = a_pipe if dom.startswith("k_fused") else a_scan
    achieved = per_read * nreads / (kern[dom] * 1e-3) / 1e9
    launches = len(batches)                                  # k_fused_tile launches per step
    traffic, traffic_src = None, None
    try:                                                     # DRAM bytes need ncu: a committed capture, labelled
        with open(os.path.join(ROOT, TRAFFIC_FILE)) as fh:
            t = json.load(fh)[name]
        traffic = round((t["dram_bytes_read"] + t["dram_bytes_write"]) * batch_reads / t["reads_per_launch"])
        traffic_src = "%s: %s, build %s (ncu --set full, not measured in this run)" % (TRAFFIC_FILE, t["source"], t["build"])
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "traffic_source": traffic_src,
                "algorithmic_bytes_per_read": round(per_read, 1),
                "algorithmic_bytes_per_launch": round(per_read * nreads / launches),
                "kernel_us_per_launch": round(kern[dom] * 1e3 / launches, 1),
                "peak_source": peak_src,
                "kernel_ms_per_step": {k: round(v, 4) for k, v in kern.items()},
                "pipeline_bytes_per_read": round(a_pipe, 1),
                "pipeline_frac_of_hbm": round(a_pipe * nreads / (res["ms_total"] / steps * 1e-3) / 1e9 / peak, 4),
                "scan_only_bytes_per_read": a_scan,
                "scan_only_frac_of_hbm": round(a_scan * nreads / (kern[dom] * 1e-3) / 1e9 / peak, 4),
                "note": "achieved = (raw record bytes read + annotated bytes written) per launch / kernel time "
                        "(A_pipe of SURVEY.md 8d; the fused kernel parses, scans, compacts and formats in one pass); "
                        "scan_only_* counts L + 16 bytes per read (A_scan).  kernel_ms_per_step / kernel_us_per_launch "
                        "come from a separate pass with the launches serialised on one stream and events between them; "
                        "ms_per_step (value) times the pass that alternates batches over two streams, so the two "
                        "differ by the overlap of consecutive launches"}
    out = {"name": name, "workload": workload, "ctx": ctx, "batches": batches, "rb": rb, "read_len": read_len,
           "flags": flags, "nreads": nreads, "in_bytes": in_bytes, "value": value, "ms_step": ms_step,
           "sel_frac": sel_frac, "e2e": e2e, "pcie_probe": pcie, "roofline": roofline, "clocks": clocks,
           "launches_per_step": res["launches_per_step"], "checksum": res["out_checksum"], "batch_reads": batch_reads}
    return out


def summary_block(m):
    """The short form of a measured config (c3 / other_configs)."""
    return {"workload": m["workload"], "reads_per_gpu": m["nreads"], "value": round(m["value"], 1), "unit": "reads/s",
            "ms_per_step": round(m["ms_step"], 4), "gbases_per_s": round(m["value"] * m["read_len"] / 1e9, 3),
            "selected_fraction": round(m["sel_frac"], 4), "roofline_frac": m["roofline"]["frac"],
            "roofline_achieved_GBps": m["roofline"]["achieved"],
            "kernel_us_per_launch": m["roofline"]["kernel_us_per_launch"],
            "algorithmic_bytes_per_read": m["roofline"]["algorithmic_bytes_per_read"],
            "e2e": {k: m["e2e"][k] for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step")},
            "pcie_probe": m["pcie_probe"], "out_checksum_fnv1a": "%016x" % m["checksum"]}


def release(m):
    m["ctx"].close()
    for b in m["batches"]:
        b.close()
    m["batches"] = []


def reference_output(paths, cli_flags, cores):
    """stdout of the unmodified reference (oracle/_ref through oracle/ref_select.py) on `paths`."""
    runner = os.path.join(ROOT, "oracle", "ref_select.py")
    with open(os.devnull, "wb") as dn:
        return subprocess.run([sys.executable, runner, "-p", str(cores)] + cli_flags + paths, stdout=subprocess.PIPE,
                              stderr=dn, check=True).stdout


def parity_vs_reference(local, nreads):
    """cmp of the GPU output with the reference's on the first `nreads` reads of every config, with
    that config's flags; the bench fails on a mismatch."""
    from baitstr_b200 import capi, synth
    cores = len(os.sched_getaffinity(0))
    tmp = tempfile.mkdtemp(prefix="bstr_par_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    out = {}
    try:
        for name, (cfg, read_len, flags, _, cli) in WORKLOADS.items():
            raw = synth.generate(cfg, read_len, 0, nreads, 1)
            p = os.path.join(tmp, "%s.fq" % name)
            raw.tofile(p)
            want = reference_output([p], cli, cores)
            with capi.Context(devices=[local], **flags) as ctx:
                got, info = ctx.scan_batch(raw.tobytes())
            ok = got == want
            out[name] = {"reads": nreads, "flags": " ".join(cli) or "(default)", "selected": info["n_selected"],
                         "out_bytes": len(got), "identical": ok, "md5": hashlib.md5(got).hexdigest()}
            if not ok:
                print("bench.py: GPU output differs from the reference on %s" % name, file=sys.stderr)
                sys.exit(1)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def file_md5(path):
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def multi_gpu_leg(ndev, reads_per_shard):
    """Rank 0 only: the product's in-process multi-GPU path.  ONE context with n_devices = ndev
    drives every GPU (dynamic dealing of batches + ordered writer, csrc/stream_host.cu) over ndev
    shard files; outputs are compared byte for byte with the 1-device run and, on a prefix, with
    the oracle.  Replaces Pool + ordered print of /root/reference/src/select_STR_reads.pyx:356-367."""
    from baitstr_b200 import capi, synth
    from oracle import oracle
    cfg, read_len, flags, _, _ = WORKLOADS["C2"]
    cores = len(os.sched_getaffinity(0))
    tmp = tempfile.mkdtemp(prefix="bstr_mg_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    res = {"devices": ndev, "shards": ndev, "reads_per_shard": reads_per_shard, "host_threads": cores,
           "source": "tmpfs shard files", "sink": "tmpfs file"}
    try:
        plain = [os.path.join(tmp, "shard%d.fq" % i) for i in range(ndev)]

        def gen(i):
            synth.generate(cfg, read_len, i * reads_per_shard, reads_per_shard, 1).tofile(plain[i])
        th = [threading.Thread(target=gen, args=(i,)) for i in range(ndev)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        procs = [subprocess.Popen("gzip -1 -c %s > %s.gz" % (p, p), shell=True) for p in plain]
        if any(pr.wait() for pr in procs):
            raise RuntimeError("gzip failed")
        gz = [p + ".gz" for p in plain]
        total = ndev * reads_per_shard

        def run(devices, files, batch_bytes, tag, repeat):
            """-> (output file, best seconds with /dev/null as the sink, seconds of the run into the file, stats)"""
            outp = os.path.join(tmp, "out_%s" % tag)
            best, stats = None, None
            with capi.Context(devices=devices, host_threads=cores, batch_bytes=batch_bytes, **flags) as ctx:
                with open(os.devnull, "wb") as dn:
                    for _ in range(repeat):                   # the first pass allocates the page-locked buffers
                        t0 = time.perf_counter()
                        ctx.run_files(files, out_fd=dn.fileno(), err_fd=dn.fileno())
                        dt = time.perf_counter() - t0
                        best = dt if best is None else min(best, dt)
                    ctx.reset_stats()
                    with open(outp, "wb") as fo:              # the run whose bytes are compared
                        t0 = time.perf_counter()
                        ctx.run_files(files, out_fd=fo.fileno(), err_fd=dn.fileno())
                        dt_file = time.perf_counter() - t0
                    stats = ctx.stats()
            return outp, best, dt_file, stats
        ref_out, ref_dt, ref_dt_file, _ = run([0], plain, 0, "1dev", 2)
        ref_md5 = file_md5(ref_out)
        res["sink_note"] = ("reads_per_s: sink /dev/null (as the 1-GPU `ingest` leg); reads_per_s_to_file: the run whose output is "
                            "compared, written by the one ordered writer into a tmpfs file (write(2) of one thread: ~4 GB/s)")
        res["one_device"] = {"reads_per_s": round(total / ref_dt, 1), "reads_per_s_to_file": round(total / ref_dt_file, 1),
                             "out_md5": ref_md5, "out_bytes": os.path.getsize(ref_out)}
        # oracle on a prefix of shard 0
        npre = min(200_000, reads_per_shard)
        rb = synth.record_bytes(cfg, read_len)
        with open(plain[0], "rb") as fh:
            want = oracle.select_buffer(fh.read(npre * rb), **flags)
        with open(ref_out, "rb") as fh:
            prefix_ok = fh.read(len(want)) == want
        res["oracle_prefix"] = {"reads": npre, "identical": prefix_ok}
        ok = prefix_ok
        devs = list(range(ndev))
        # (.gz: the first passes allocate the page-locked output buffers of fresh lanes and the inflater's working sets)
        for tag, files, bb, rep in (("plain", plain, 0, 2), ("plain_1MiB_batches", plain, 1 << 20, 1), ("gz", gz, 0, 4)):
            outp, dt, dt_file, st = run(devs, files, bb, tag, rep)
            same = file_md5(outp) == ref_md5 and os.path.getsize(outp) == os.path.getsize(ref_out)
            ok = ok and same
            res[tag] = {"reads_per_s": round(total / dt, 1), "reads_per_s_to_file": round(total / dt_file, 1),
                        "identical_to_one_device": same, "batches": st["batches"],
                        "per_device_batches": st["device_batches"][:ndev], "fallback_batches": st["fallback_batches"]}
            os.unlink(outp)
        res["reads_per_s"] = res["plain"]["reads_per_s"]
        res["per_device_batches"] = res["plain"]["per_device_batches"]
        res["parity"] = "byte-identical" if ok else "MISMATCH"
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return res


def run_ours(args):
    D = Dist()
    main = measure_config(D, args.workload, args.reads, args.batch_reads, args.steps, args.warmup,
                          args.e2e_batch_reads, probe=not args.quick, sample_clocks=True)
    ctx, batches, rb = main["ctx"], main["batches"], main["rb"]

    # ---- CPU baseline (oracle port, 1 core) on a bounded sample + parity check -------------
    cpu = None
    if D.world == 1 and not args.no_cpu:
        from oracle import oracle
        sample = min(args.cpu_sample, args.batch_reads, main["nreads"])
        raw = batches[0].array[:sample * rb].tobytes()
        t0 = time.perf_counter()
        want = oracle.select_buffer(raw, **main["flags"])
        dt = time.perf_counter() - t0
        got, _ = ctx.scan_batch(raw)
        cpu = {"value": round(sample / dt, 1), "unit": "reads/s", "cores": 1, "kind": "port",
               "sample": "first %d reads of the workload, oracle/str_oracle.c single thread" % sample,
               "parity_vs_gpu": "byte-identical" if got == want else "MISMATCH"}
        if got != want:
            print("bench.py: GPU output differs from the oracle on the sample", file=sys.stderr)
            sys.exit(1)
    release(main)

    # ---- the other BASELINE.json configs ----------------------------------------------------
    others, c3 = {}, None
    if not args.no_other:
        o_steps, o_warm = max(2, min(args.steps, 5)), max(3, min(args.warmup, 3))
        if args.workload != "C3":
            m = measure_config(D, "C3", args.c3_reads, args.batch_reads, o_steps, o_warm, args.e2e_batch_reads)
            c3 = summary_block(m)
            c3["note"] = ("BASELINE.json configs[2] split over 8 GPUs = 12.5 M reads per GPU; weak scaling, every rank "
                          "scans its own block of the read index range")
            release(m)
            others["C3"] = c3
        if D.world == 1:
            for name, reads in (("C4", args.other_reads), ("C5", args.other_reads)):
                if name == args.workload:
                    continue
                m = measure_config(D, name, reads, min(args.batch_reads, 500_000 if name == "C4" else args.batch_reads),
                                   o_steps, o_warm, args.e2e_batch_reads, probe=False)
                others[name] = summary_block(m)
                release(m)

    # ---- the product's in-process multi-GPU path, with parity (rank 0 drives all GPUs) --------
    multi = None
    if D.world > 1 and not args.no_multi:
        D.barrier()
        if D.rank == 0:
            try:
                multi = multi_gpu_leg(D.world, args.shard_reads)
            except Exception as e:                                       # the line must say so, and the run must fail
                multi = {"devices": D.world, "parity": "ERROR", "error": "%s: %s" % (type(e).__name__, e)}
        D.host_barrier()

    if D.rank != 0:
        D.close()
        return

    parity_ref = None
    if D.world == 1 and not args.no_cpu:
        parity_ref = parity_vs_reference(D.local, args.parity_reads)

    # ---- file ingest (host side of bstr_run_files) on a bounded sample: plain and .gz ---------
    ingest = None
    if D.world == 1 and not args.no_ingest:
        ingest = measure_ingest(args.workload, args.ingest_reads, D.local)

    nxt = None
    if D.world == 1 and not args.no_other and not args.quick:
        nxt = next_stages(D.local)

    line = {
        "metric": METRIC, "value": round(main["value"], 1), "unit": "reads/s", "n_gpus": D.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(main["ms_step"], 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": main["workload"],
                   "reads_per_gpu": main["nreads"], "read_len": main["read_len"], "batch_reads": args.batch_reads,
                   "raw_fastq_bytes_per_gpu": main["in_bytes"], "selected_fraction": round(main["sel_frac"], 4),
                   "l2_policy": "inputs larger than L2 (%.2f GB read per step), no flush" % (main["in_bytes"] / 1e9),
                   "parallelism": "reads sharded across GPUs, no collective"},
        "gbases_per_s": round(main["value"] * main["read_len"] / 1e9, 3),
        "e2e": main["e2e"], "pcie_probe": main["pcie_probe"],
        "gpu_launches": main["launches_per_step"] * args.steps,
        "roofline": main["roofline"], "cpu_baseline": cpu, "clocks": main["clocks"],
        "c3": c3, "other_configs": others or None, "multi_gpu": multi, "parity_vs_reference": parity_ref,
        "ingest": ingest, "next_stages": nxt, "out_checksum_fnv1a": "%016x" % main["checksum"],
    }
    print(json.dumps(line))
    D.close()
    if multi is not None and multi.get("parity") != "byte-identical":
        print("bench.py: multi-GPU parity failed: %s" % json.dumps(multi), file=sys.stderr)
        sys.exit(1)


def next_stages(device):
    """SURVEY 8(f) ranks 2 and 3, the first slices of the stages after the path: the merge stage
    (bstr_merge_reads) and the k-mer counting pass of extend_STR_reads (bstr_kmers_*), each timed on a
    small synthetic input through the C ABI with host buffers, beside the reference's own code
    (oracle/_ref/downstream: its sources compiled in place) with a cmp of the results."""
    from baitstr_b200 import capi
    from tests import util
    down = os.path.join(ROOT, "oracle", "_ref", "downstream")
    tmp = tempfile.mkdtemp(prefix="bstr_next_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    out = {}
    try:
        with capi.Context(devices=[device]) as ctx:
            # merge: 2000 STR loci x 40 reads of 150 bp -> select (GPU) -> merge
            raw = b"".join(util.str_locus_reads(500, 40, 150, seed=200 + i).replace(b"@m.", b"@m%d." % i) for i in range(4))
            sel = ctx.scan_batch(raw)[0]
            nsel = sel.count(b"\n") // 4
            ctx.merge_reads(sel, 27)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter()
                got, st = ctx.merge_reads(sel, 27)
                best = min(best, time.perf_counter() - t0)
            m = {"reads": nsel, "klength": 27, "chains": st["chains"], "blocks_printed": st["printed"], "reads_per_s": round(nsel / best, 1),
                 "ms": round(best * 1e3, 2), "reference_reads_per_s": None, "identical_to_reference": None}
            exe = os.path.join(down, "merge_STR_reads")
            if os.path.exists(exe):
                p = os.path.join(tmp, "sel.fq")
                with open(p, "wb") as f:
                    f.write(sel)
                t0 = time.perf_counter()
                want = subprocess.run([exe, "27", p], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout
                m["reference_reads_per_s"] = round(nsel / (time.perf_counter() - t0), 1)
                m["identical_to_reference"] = want == got
            out["merge_STR_reads"] = m
            # k-mer counting: 100 k reads of 150 bp from one 1 Mb genome, k = 27
            raw = util.genome_reads(100000, 150, 1000000, seed=9)
            expected = 100000 * 124 // 2 + 1000000
            ctx.count_kmers([raw[: len(raw) // 20]], 27, expected_kmers=expected // 10)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter()
                kmers, counts, st = ctx.count_kmers([raw], 27, expected_kmers=expected)
                best = min(best, time.perf_counter() - t0)
            k = {"reads": 100000, "k": 27, "kmers": st["kmers"], "kept": int(len(kmers)), "reads_per_s": round(100000 / best, 1),
                 "kmers_per_s": round(st["kmers"] / best, 1), "ms": round(best * 1e3, 2), "reference_reads_per_s": None,
                 "identical_to_reference": None}
            exe = os.path.join(down, "kmer_count_ref")
            if os.path.exists(exe):
                p = os.path.join(tmp, "reads.fq")
                with open(p, "wb") as f:
                    f.write(raw)
                t0 = time.perf_counter()
                ref = subprocess.run([exe, "27", "2", "255", str(expected), p], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout
                k["reference_reads_per_s"] = round(100000 / (time.perf_counter() - t0), 1)
                want = [tuple(int(x) for x in ln.split()) for ln in ref.decode().splitlines()]
                k["identical_to_reference"] = want == list(zip(kmers.tolist(), counts.tolist()))
            out["extend_STR_reads_kmer_count"] = k
            # the whole extend stage: 50 k reads of a 0.25 Mb diploid genome -> select, merge (GPU) -> k-mer counting + walks
            raw = util.diploid_str_reads(250000, 250, 30, 150, seed=31)
            nreads = raw.count(b"\n") // 4
            recs = raw.split(b"\n@")
            r1, r2 = b"\n@".join(recs[:len(recs) // 2]) + b"\n", b"@" + b"\n@".join(recs[len(recs) // 2:])
            with capi.Context(devices=[device], n=3) as c3:
                sel = c3.scan_batch(raw)[0]
            merged, _ = ctx.merge_reads(sel, 27, 3, 10000, True)
            expected = nreads * 124 // 2 + 500000
            ctx.extend_reads(merged, [r1[: r1.index(b"\n@", len(r1) // 20) + 1]], 27, expected_kmers=expected // 10)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter()
                got, st = ctx.extend_reads(merged, [r1, r2], 27, expected_kmers=expected)
                best = min(best, time.perf_counter() - t0)
            e = {"reads": nreads, "k": 27, "blocks": st["blocks"], "contigs": st["contigs"], "reads_per_s": round(nreads / best, 1),
                 "ms": round(best * 1e3, 2), "reference_reads_per_s": None, "identical_to_reference": None}
            exe = os.path.join(down, "extend_STR_reads")
            if os.path.exists(exe):
                paths = []
                for fn, data in (("merged.fq", merged), ("r1.fq", r1), ("r2.fq", r2)):
                    with open(os.path.join(tmp, fn), "wb") as f:
                        f.write(data)
                    paths.append(os.path.join(tmp, fn))
                t0 = time.perf_counter()
                want = subprocess.run([exe, "250000", "30", "27"] + paths, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout
                e["reference_reads_per_s"] = round(nreads / (time.perf_counter() - t0), 1)
                e["identical_to_reference"] = want == got
            out["extend_STR_reads"] = e
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    for name, v in out.items():
        if v["identical_to_reference"] is False:
            print("bench.py: %s differs from the reference" % name, file=sys.stderr)
            sys.exit(1)
    return out


def measure_ingest(name, nreads, device):
    """FILES -> bstr_run_files -> /dev/null for a prefix of the workload, as plain FASTQ and as
    gzip -1 (the C4/C5 input format): what the reader / parallel inflater / stitcher sustain when
    the GPU is not the limit.  Not part of `value` or `e2e`."""
    from baitstr_b200 import capi, synth
    cfg, read_len, flags, _, _ = WORKLOADS[name]
    tmp = tempfile.mkdtemp(prefix="bstr_ingest_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    cores = len(os.sched_getaffinity(0))
    out = {"reads": nreads, "host_threads": cores, "sink": "/dev/null", "source": "tmpfs, two mate files"}
    try:
        plain = []
        for mate in (1, 2):
            p = os.path.join(tmp, "m%d.fq" % mate)
            synth.generate(cfg, read_len, 0, nreads // 2, mate).tofile(p)
            plain.append(p)
        procs = [subprocess.Popen("gzip -1 -c %s > %s.gz" % (p, p), shell=True) for p in plain]
        if any(pr.wait() for pr in procs):
            raise RuntimeError("gzip failed")
        ctx = capi.Context(devices=[device], host_threads=cores, **flags)
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
    cfg, read_len, _, workload, cli = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    runner = os.path.join(ROOT, "oracle", "ref_select.py")
    tmpdir = tempfile.mkdtemp(prefix="bstr_ref_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    outp = os.path.join(tmpdir, "ref.out")

    def one_pass(nreads, procs=cores):
        half = nreads // 2
        paths = []
        for mate in (1, 2):
            p = os.path.join(tmpdir, "syn_%d.fq" % mate)
            synth.generate(cfg, read_len, 0, half, mate).tofile(p)
            paths.append(p)
        t0 = time.perf_counter()
        with open(outp, "wb") as fo, open(os.devnull, "wb") as dn:      # tmpfs, so the output can be hashed
            subprocess.run([sys.executable, runner, "-p", str(procs)] + cli + paths, stdout=fo,
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
        out_md5, out_bytes = file_md5(outp), os.path.getsize(outp)
        dt1, n1 = one_pass(20000, procs=1)            # the -p 1 rate BASELINE.md section 3 asks for
    finally:
        shutil.rmtree(tmpdir, ignore_errors=True)
    value = reads / total
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 1), "unit": "reads/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(total / args.steps * 1e3, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload,
                   "read_len": read_len,
                   "note": "each step = the first %d reads of the workload (prefix), files in tmpfs, "
                           "stdout to a tmpfs file; interpreter CPython %d.%d + Cython shim build "
                           "(Python 2 unavailable); the FASTQ reader is the line-for-line Python restatement "
                           "of fastq.pyx (SURVEY 8c shim 4), whose serial parse in the parent bounds -p scaling "
                           "somewhat below an all-Cython reader" % (sample, sys.version_info[0], sys.version_info[1])},
        "gbases_per_s": round(value * read_len / 1e9, 6),
        "cpu_baseline": {"value": round(value, 1), "unit": "reads/s", "cores": cores, "kind": "reference",
                         "sample": "first %d reads per step, select_STR_reads -p %d" % (sample, cores)},
        "p1": {"value": round(n1 / dt1, 1), "unit": "reads/s", "cores": 1, "sample": "first %d reads, select_STR_reads -p 1" % n1},
        "output": {"reads": sample // 2 * 2, "md5": out_md5, "bytes": out_bytes,
                   "note": "md5 of the reference's stdout on the timed prefix; bench.py's own arm cmp's the GPU output "
                           "with the reference's in `parity_vs_reference`"},
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the oracle baseline and the reference parity legs")
    ap.add_argument("--no-ingest", action="store_true")
    ap.add_argument("--no-other", action="store_true", help="skip c3 / other_configs")
    ap.add_argument("--no-multi", action="store_true", help="skip the in-process multi-GPU leg (N > 1)")
    ap.add_argument("--ingest-reads", type=int, default=8_000_000)
    ap.add_argument("--c3-reads", type=int, default=12_500_000, help="C3 reads per GPU (100 M / 8)")
    ap.add_argument("--other-reads", type=int, default=4_000_000, help="C4 / C5 reads at N = 1")
    ap.add_argument("--shard-reads", type=int, default=2_000_000, help="reads per shard file in the multi-GPU leg")
    ap.add_argument("--parity-reads", type=int, default=60_000, help="prefix cmp'd with the reference per config")
    ap.add_argument("--e2e-batch-reads", type=int, default=1_000_000)
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--quick", action="store_true", help="kernel A/B runs: value, e2e and roofline of the workload only")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.quick:
        args.no_cpu = args.no_ingest = args.no_other = args.no_multi = True
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
