#!/usr/bin/env python
"""bench.py -- throughput of the MBG graph-build front end (BASELINE.json metric) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload chr20|ecoli|herro] [--scale F]

A "step" is one pass of the hot path (collapse -> microsatellite coding -> syncmer selection -> k-mer hashing ->
k-mer table: nodes + edges) over one batch of synthetic HiFi-shaped reads.  One process per GPU (torchrun for N > 1);
reads are sharded across ranks (weak scaling: genome and read count grow with N), selected k-mers are exchanged to
their owner rank with an NCCL all-to-all, and each rank counts its partition of the k-mer space.

`value`      raw bases/s with the 2-bit packed reads already resident in HBM (device time, CUDA events, max over ranks)
`e2e`        the same metric through the C ABI from pinned HOST buffers: ASCII reads -> H2D -> pack -> ... -> table
             -> D2H of the ordered node/edge table, every step
`roofline`   dominant kernel against the measured HBM copy peak (MEASURED_PEAKS.json)
`cpu_baseline` the reference's own hot path (oracle/_ref, all host cores) on a bounded sample of the same reads

--impl reference times ONLY the reference's CPU implementation (rank 0), same metric/unit/config.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: genome bp per GPU, coverage, read length range, error ppm, k, w
    "chr20": dict(genome=64_000_000, coverage=30, len_min=15000, len_max=20000, error_ppm=1000, k=1001, w=100,
                  desc="human chr20-shaped synthetic HiFi 30x (64 Mbp per GPU, 15-20 kb reads, 0.1% errors), MBG k=1001 w=100 collapse-msat"),
    "ecoli": dict(genome=5_000_000, coverage=24, len_min=11000, len_max=19000, error_ppm=1000, k=1001, w=100,
                  desc="E. coli-shaped synthetic HiFi 24x (5 Mbp, ~15 kb reads), MBG k=1001 w=100 collapse-msat"),
    "herro": dict(genome=64_000_000, coverage=30, len_min=30000, len_max=50000, error_ppm=7500, k=1001, w=100,
                  desc="HERRO/ONT-duplex-shaped synthetic reads 30x (64 Mbp per GPU, 30-50 kb reads, 0.75% errors), MBG k=1001 w=100 collapse-msat"),
}
HG002 = dict(genome=3_100_000_000, coverage=30, len_min=15000, len_max=20000, error_ppm=1000, k=1001, w=100, batch_reads=110_000,
             desc="HG002-shaped synthetic diploid HiFi (3.1 Gbp x 2 haplotypes, 0.1% SNPs, 15-20 kb reads, 0.1% errors), MBG k=1001 w=100 collapse-msat")
METRIC = "raw HiFi bases/s through the MBG graph-build front end (HPC+microsatellite compression, syncmer selection k=1001 w=100, 128-bit k-mer hashing, k-mer counting)"
SEED_GENOME, SEED_READS = 7, 11


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled from a thread every ~2 ms (the timed
    region of a default run is well under a second, so `nvidia-smi -lms` is too coarse); falls back to nvidia-smi."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, device):
        self.device, self.sm, self.bits, self.mx = device, [], 0, None
        self.stop_flag, self.thread, self.nvml = threading.Event(), None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.device
            if vis:
                ids = [x for x in vis.split(",") if x.strip() != ""]
                if self.device < len(ids) and ids[self.device].strip().isdigit():
                    idx = int(ids[self.device])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None
            return
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def _poll(self):
        n = self.nvml
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                try:
                    self.bits |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:
                    self.bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        if self.nvml is None:
            return self._smi_once()
        self.stop_flag.set()
        self.thread.join(timeout=2)
        reasons = [name for bit, name in self.REASONS if self.bits & bit]
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx, "reasons": reasons, "samples": len(self.sm), "source": "nvml, 2 ms polling"}

    def _smi_once(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={q}", "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=20).stdout
            r = [x.strip() for x in out.strip().split(",")]
            reasons = [n for n, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]) if v.lower().startswith("active")]
            return {"sm_mhz": float(r[0]), "sm_max_mhz": float(r[1]), "reasons": reasons, "samples": 1, "source": "nvidia-smi after the timed region (NVML unavailable)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_reads(wl, world, rank, scale, pinned_alloc=None, max_reads=None):
    """this rank's shard: the genome grows with the world size (weak scaling), reads are dealt in contiguous blocks"""
    from verkko_b200 import synth
    genome_len = int(wl["genome"] * scale) * world
    g = synth.genome(SEED_GENOME, genome_len)
    mean_len = (wl["len_min"] + wl["len_max"]) / 2
    per_rank = int(wl["genome"] * scale * wl["coverage"] / mean_len)
    if max_reads:
        per_rank = min(per_rank, max_reads)
    first = rank * per_rank
    offsets = synth.read_offsets(SEED_READS, g, per_rank, wl["len_min"], wl["len_max"], wl["error_ppm"], first_read=first)
    nbytes = int(offsets[-1])
    if pinned_alloc:
        buf = pinned_alloc(nbytes + 64)
        bases = buf.array[:nbytes]
    else:
        buf, bases = None, np.empty(nbytes, dtype=np.uint8)
    synth.fill_reads(SEED_READS, g, per_rank, wl["len_min"], wl["len_max"], offsets, bases, wl["error_ppm"], first_read=first)
    return bases, offsets, buf


def sample_size(wl):
    """reads of the bounded sample both arms push through the reference: ~25 s of CPU work at the producer-bound rate of
    the reference's reader thread, at most 20 000 reads (the first reads of rank 0's shard of the workload)"""
    cores = os.cpu_count() or 1
    est_rate = 25e6 * min(cores, 8)
    return max(64, int(min(est_rate * 25.0 / ((wl["len_min"] + wl["len_max"]) / 2), 20000)))


DETERMINISTIC = ("selected", "distinct", "distinct_edges", "node_xor", "node_sum", "edge_cov", "edge_sum", "edge_xor")


def write_sample_fasta(path, bases, offsets):
    with open(path, "wb") as f:
        for i in range(len(offsets) - 1):
            f.write(b">r%d\n" % i)
            f.write(bases[int(offsets[i]):int(offsets[i + 1])].tobytes())
            f.write(b"\n")


def quiet_reference(fn):
    """the reference prints progress on stderr; keep it out of the bench output"""
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)
    try:
        return fn()
    finally:
        os.dup2(saved, 2)
        os.close(devnull)
        os.close(saved)


def reference_arm(args, wl):
    """the reference's own CPU implementation of the path on all host cores, on a bounded sample of the workload"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C
    from oracle_lib import REF_SO, Oracle, have_reference
    cores = os.cpu_count() or 1
    bases, offsets, _ = make_reads(wl, 1, 0, args.scale, max_reads=sample_size(wl))
    n_reads = len(offsets) - 1
    n_bases = int(offsets[-1])
    times = []
    digest = None
    if have_reference():
        from oracle_lib import reference_file_digest
        kind = "reference"
        with tempfile.TemporaryDirectory() as d:
            fa = os.path.join(d, "sample.fa")
            write_sample_fasta(fa, bases, offsets)
            for it in range(args.warmup + args.steps):
                t, digest = quiet_reference(lambda: reference_file_digest("collapse-msat", wl["k"], wl["w"], fa, cores))
                if it >= args.warmup:
                    times.append(t)
            selected = digest["selected"]
    else:
        kind, cores = "port", 1
        orc = Oracle()
        for it in range(args.warmup + args.steps):
            t0 = time.time()
            r = orc.run("collapse-msat", wl["k"], wl["w"], (bases, offsets))
            if it >= args.warmup:
                times.append(time.time() - t0)
        selected = len(r.positions)
        from oracle_lib import table_digest
        digest = table_digest(r.node_hash, r.node_cov, r.node_tip, r.edges)
    sec = float(np.mean(times))
    value = n_bases / sec
    sample = f"{n_reads} reads / {n_bases} bases of the workload (FASTA file, page cache warm), {cores} worker threads + reader, loadReadsAsHashesMultithread"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "bases/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16/u64 integer", "data": "synthetic",
        "config": {"workload": wl["desc"] + f" [bounded sample: {n_reads} reads]", "k": wl["k"], "w": wl["w"], "error_masking": "collapse-msat"},
        "kmers_per_s": selected / sec,
        "cpu_baseline": {"value": value, "unit": "bases/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        # order-free digest of the table the reference built from the sample (oracle/ref_shim.cpp:hashlist_digest); the GPU arm
        # prints the digest of ITS table of the same reads under the same key
        "sample_digest": {f: digest[f] for f in DETERMINISTIC},
    }
    emit(json.dumps(line))      # stdout proper (fd 1 itself points at stderr during the run, see main)


def cpu_baseline(wl, scale, device=0):
    """the bounded sample (the one `--impl reference` times) through the reference's own code on the host cores, and --
    parity at bench size -- the same reads through the CUDA path: the two tables must have the same order-free digest
    (selected, distinct k-mers and edges, XOR / sum of canonical hashes x coverage; tests/oracle_lib.py:table_digest)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle, have_reference, reference_file_digest, table_digest
    from verkko_b200 import MbgFrontEnd
    cores = os.cpu_count() or 1
    bases, offsets, _ = make_reads(wl, 1, 0, scale, max_reads=sample_size(wl) if have_reference() else 300)
    n_reads, n_bases = len(offsets) - 1, int(offsets[-1])
    if have_reference():
        with tempfile.TemporaryDirectory() as d:
            fa = os.path.join(d, "sample.fa")
            write_sample_fasta(fa, bases, offsets)
            t, ref = quiet_reference(lambda: reference_file_digest("collapse-msat", wl["k"], wl["w"], fa, cores))
        out = {"value": n_bases / t, "unit": "bases/s", "cores": cores, "kind": "reference",
               "sample": f"{n_reads} reads / {n_bases} bases of the workload, MBG's loadReadsAsHashesMultithread with {cores} worker threads (oracle/_ref)"}
    else:
        orc = Oracle()
        t0 = time.time()
        r = orc.run("collapse-msat", wl["k"], wl["w"], (bases, offsets))
        t = time.time() - t0
        ref = table_digest(r.node_hash, r.node_cov, r.node_tip, r.edges)
        out = {"value": n_bases / t, "unit": "bases/s", "cores": 1, "kind": "port",
               "sample": f"{n_reads} reads / {n_bases} bases of the workload, oracle/mbg_oracle.c single thread"}
    fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat", device=device)
    try:
        fe.add_reads(bases, offsets)
        fe.run()
        mine = table_digest(fe.node_hashes(copy=False), fe.node_coverage(copy=False), fe.node_tip(copy=False), fe.edges(copy=False))
    finally:
        fe.close()
    bad = [f for f in DETERMINISTIC if mine[f] != ref[f]]
    return out, {f: mine[f] for f in DETERMINISTIC}, bad


def run_hg002(args, rank, world, local_rank, dist, torch, emit):
    """BASELINE.json configs[2]: the HG002-shaped diploid set sharded over the GPUs of one box.  TOTAL work is fixed (strong
    scaling): 2 x 3.1 Gbp x 30x x --scale bases, rank r takes every world-th block of reads, in batches of 110 000 reads (one
    chr20-sized resident batch each) through mbgfe_run_append; ONE exchange (or mbgfe_build_table at N = 1) counts the selected
    k-mers of all batches.  --scale < 1 lowers the coverage (host memory: every batch sits in pinned memory before the clock starts)."""
    from verkko_b200 import MbgFrontEnd, synth
    from verkko_b200.frontend import NO_TABLE
    wl = HG002
    hap_len = int(wl["genome"] * args.genome_scale)
    hap1 = synth.genome(SEED_GENOME, hap_len)
    rng = np.random.default_rng(5)
    hap2 = hap1.copy()
    idx = rng.integers(0, hap_len, hap_len // 1000)
    hap2[idx] = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, len(idx))]
    genome = np.concatenate([hap1, hap2])
    del hap1, hap2
    mean_len = (wl["len_min"] + wl["len_max"]) / 2
    total_reads = int(2 * hap_len * wl["coverage"] * args.scale / mean_len)
    per_rank = total_reads // world
    sharded = world > 1
    fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat", device=local_rank, rank=rank, world=world, flags=NO_TABLE if sharded else 0)
    if sharded:
        uid = [fe.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        fe.comm_init_nccl(uid[0])
    batches, keep = [], []
    first = rank * per_rank
    for b0 in range(0, per_rank, wl["batch_reads"]):
        n = min(wl["batch_reads"], per_rank - b0)
        offsets = synth.read_offsets(SEED_READS, genome, n, wl["len_min"], wl["len_max"], wl["error_ppm"], first_read=first + b0)
        buf = fe.pinned(int(offsets[-1]) + 64)
        synth.fill_reads(SEED_READS, genome, n, wl["len_min"], wl["len_max"], offsets, buf.array[:int(offsets[-1])], wl["error_ppm"], first_read=first + b0)
        batches.append((buf.array[:int(offsets[-1])], offsets))
        keep.append(buf)
    del genome
    n_bases = sum(int(o[-1]) for _, o in batches)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def finish():
        if sharded:
            fe.exchange()
        else:
            fe.build_table()
    # ---- resident per batch: upload untimed, front end timed on the device, then the table / exchange
    stage_sum = {}

    def resident_step():
        fe.reset()
        ms = 0.0
        for kname in list(stage_sum):
            stage_sum[kname] = 0.0
        for b, o in batches:
            fe.add_reads(b, o)
            fe.upload_append()
            fe.timer_start()
            fe.process_append()
            ms += fe.timer_stop()
            stb = fe.stats()
            for kname in ("ms_collapse", "ms_msat", "ms_select", "ms_hash", "ms_process"):      # per kernel stage, by events around the launches
                stage_sum[kname] = stage_sum.get(kname, 0.0) + stb[kname]
        stage_sum["ms_batches_timer"] = ms      # the same batches by one timer around each call (adds what the host leaves idle between launches)
        # every rank enters the collective together: the uploads between the timed regions are untimed, and without this a rank whose
        # uploads are quicker (the GPUs on the other socket of the box) waits inside its timed exchange for the ranks still uploading
        # -- measured at N = 8: 66 ms on four ranks, 150 ms on the other four for the same work
        barrier()
        fe.timer_start()
        finish()
        return ms + fe.timer_stop()
    for _ in range(args.warmup):
        resident_step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    dev_ms = [resident_step() for _ in range(args.steps)]
    barrier()
    clocks = sampler.stop()
    st = fe.stats()
    # ---- end to end: every batch from pinned host memory, the table (partition) read back
    e2e_ms, checksum = [], 0
    for it in range(1 + args.e2e_steps):
        barrier()
        t0 = time.time()
        fe.reset()
        for b, o in batches:
            fe.add_reads(b, o)
            fe.run_append()
        finish()
        cov = fe.node_coverage(copy=False)
        checksum = int(cov[::1024].sum(dtype=np.uint64)) + len(fe.node_hashes(copy=False)) + len(fe.edges(copy=False))
        barrier()
        if it > 0:
            e2e_ms.append((time.time() - t0) * 1e3)
    st_e = fe.stats()
    t = torch.tensor([float(np.mean(dev_ms)), float(np.mean(e2e_ms))], dtype=torch.float64)
    sm = torch.tensor([float(n_bases), float(st_e["selected"]), float(st_e["distinct"]), float(st_e["distinct_edges"])], dtype=torch.float64)
    per_rank_ms = [float(np.mean(dev_ms))]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        gathered = [None] * world
        dist.all_gather_object(gathered, float(np.mean(dev_ms)))
        per_rank_ms = gathered
    if rank == 0:
        peak, peak_src = measured_peak()
        line = {"metric": METRIC, "value": float(sm[0]) / (float(t[0]) * 1e-3), "unit": "bases/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": float(t[0]), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u16 codes / u64 hashes (integer)", "data": "synthetic",
                "config": {"workload": wl["desc"] + f", {wl['coverage'] * args.scale:g}x" + (f", genome x{args.genome_scale}" if args.genome_scale != 1 else ""), "k": wl["k"], "w": wl["w"],
                           "error_masking": "collapse-msat", "batches_per_gpu": len(batches), "bases_per_gpu": n_bases, "total_bases": float(sm[0]),
                           "parallelism": f"reads sharded over {world} GPU(s) in batches (mbgfe_run_append), k-mers owned by hash range, one mbgfe_exchange over NCCL",
                           "l2": "every batch (>= 0.48 GB packed) is larger than the 126 MB L2"},
                "kmers_per_s": float(sm[1]) / (float(t[0]) * 1e-3), "counts": {"selected_kmers": float(sm[1]), "distinct_kmers": float(sm[2]), "distinct_edges": float(sm[3])},
                "exchange": {"ms": st["ms_exchange"], "bytes_sent_rank0": st["exchange_bytes"], "launches": st["exchange_launches"],
                             "nvlink_GBps_rank0": (st["exchange_bytes"] / (st["ms_exchange"] * 1e-3) / 1e9) if st["ms_exchange"] > 0 else None, "nvlink_peak_GBps": 900.0},
                "clocks": clocks, "gpu_launches": int((st["launches"] * len(batches) + st["exchange_launches"]) * args.steps),
                "stage_ms_per_step_rank0": dict(stage_sum), "ms_per_step_by_rank": per_rank_ms,
                "e2e": {"value": float(sm[0]) / (float(t[1]) * 1e-3), "unit": "bases/s", "ms_per_step": float(t[1]), "h2d_bytes_per_step": int(st_e["h2d_bytes"] * len(batches)),
                        "d2h_bytes_per_step": int(st_e["d2h_bytes"]), "checksum": checksum}}
        emit(json.dumps(line))
    if world > 1:
        dist.barrier()
    fe.close()


def main():
    # stdout carries exactly ONE JSON line (rank 0).  Libraries chat on fd 1 (NCCL prints its version banner there when the
    # first communicator is made), so fd 1 is pointed at stderr for the whole run and the line goes to the saved descriptor.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    global emit
    def emit(text):
        os.write(json_fd, (text + "\n").encode())
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="chr20", choices=list(WORKLOADS) + ["hg002"])
    ap.add_argument("--genome-scale", type=float, default=1.0, help="hg002: shrink the genome (quick runs)")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the genome (and read count) for quick runs")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed end-to-end steps (0 = --steps)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-file", action="store_true", help="skip the from-a-file end-to-end leg (e2e_file)")
    ap.add_argument("--k", type=int, default=0, help="k-mer size override (BASELINE.json configs[4]: the k/w sweep)")
    ap.add_argument("--w", type=int, default=0, help="window size override")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload if args.workload != "hg002" else "chr20"])
    if args.k or args.w:
        wl["k"], wl["w"] = args.k or wl["k"], args.w or wl["w"]
        wl["desc"] = wl["desc"].replace("k=1001 w=100", f"k={wl['k']} w={wl['w']}")
    if args.e2e_steps <= 0:
        args.e2e_steps = args.steps
    if args.impl == "reference":
        reference_arm(args, wl)
        return

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: libmbgfe has no CPU path")
    torch.cuda.set_device(local_rank)
    # MBGFE_FORCE_EXCHANGE=1: run the sharded-table path (owner exchange, id replies, edge records) on ONE GPU, so that its
    # kernels can be profiled with ncu (which must not wrap a multi-rank command)
    force_exchange = world == 1 and os.environ.get("MBGFE_FORCE_EXCHANGE") == "1"
    if world > 1 or force_exchange:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"       # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
    if args.workload == "hg002":
        if world > 1:
            os.environ.setdefault("NCCL_DEBUG", "WARN")
            dist.init_process_group("gloo")
        run_hg002(args, rank, world, local_rank, dist, torch, emit)
        if world > 1:
            dist.destroy_process_group()
        return
    if world > 1:
        # torch.distributed is host plumbing only (gloo: the NCCL id hand-out, barriers, max / sum of a few scalars); the data path
        # collectives run inside libmbgfe on its own NCCL communicator (mbgfe_comm_init_nccl / mbgfe_exchange)
        dist.init_process_group("gloo")
    from verkko_b200 import MbgFrontEnd
    from verkko_b200.frontend import NO_TABLE

    sharded = world > 1 or force_exchange
    fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat", device=local_rank, rank=rank, world=world, flags=NO_TABLE if sharded else 0)
    bases, offsets, pinned = make_reads(wl, world, rank, args.scale, pinned_alloc=fe.pinned)
    n_bases = int(offsets[-1])

    class Exchange:
        """mbgfe_exchange: counts, node records, id replies, edge records -- all inside the library, over NCCL"""
        launches, last_exchange_ms, sent_bytes = 0, 0.0, 0

        def exchange_and_count(self):
            fe.exchange()
            st_x = fe.stats()
            self.launches, self.last_exchange_ms, self.sent_bytes = st_x["exchange_launches"], st_x["ms_exchange"], st_x["exchange_bytes"]

        def download_counts(self):
            cov = fe.node_coverage(copy=False)
            fe.node_hashes(copy=False)
            return int(cov.sum(dtype=np.uint64)) + len(fe.edges(copy=False))
    counter = None
    if sharded:
        uid = [fe.unique_id() if rank == 0 else None]
        if world > 1:
            dist.broadcast_object_list(uid, src=0)
        fe.comm_init_nccl(uid[0])
        counter = Exchange()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step():
        fe.process()
        if counter:
            counter.exchange_and_count()

    # ---- resident throughput
    fe.add_reads(bases, offsets)
    fe.upload()
    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    stage_ms = {}
    fe.timer_start()
    t0 = time.time()
    for _ in range(args.steps):
        step()
        st = fe.stats()
        for kname in ("ms_collapse", "ms_msat", "ms_select", "ms_hash", "ms_nodes", "ms_edges", "ms_process"):
            stage_ms[kname] = stage_ms.get(kname, 0.0) + st[kname]
        if counter:
            stage_ms["ms_exchange"] = stage_ms.get("ms_exchange", 0.0) + counter.last_exchange_ms
    device_ms = fe.timer_stop()
    barrier()
    wall_ms = (time.time() - t0) * 1e3
    clocks = sampler.stop()
    st = fe.stats()
    launches = st["launches"] * args.steps + (counter.launches * args.steps if counter else 0)
    t = torch.tensor([device_ms, wall_ms, float(n_bases), float(st["selected"])], dtype=torch.float64)
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        device_ms, wall_ms, total_bases, total_sel = float(mx[0]), float(mx[1]), float(sm[2]), float(sm[3])
    else:
        total_bases, total_sel = float(n_bases), float(st["selected"])
    ms_per_step = device_ms / args.steps
    value = total_bases / (ms_per_step * 1e-3)
    parts_stats = dict(st)
    fe.parts()
    st = fe.stats()

    # ---- end to end through the C ABI from pinned host buffers: ASCII reads in host memory -> ordered table in host memory
    e2e_ms, h2d, d2h, e2e_parts = [], 0, 0, {}
    for it in range(1 + args.e2e_steps):
        barrier()
        t0 = time.time()
        fe.add_reads(bases, offsets)
        fe.run()                      # pipelined: H2D pieces on a copy stream, K0-K3a per piece behind them, then hash + table
        t1 = t2 = time.time()
        st_up = fe.stats()
        if counter:
            counter.exchange_and_count()
            t2 = time.time()
        if world == 1:
            cov = fe.node_coverage(copy=False)       # views of the library's pinned result arrays (no extra host copy)
            hashes = fe.node_hashes(copy=False)
            edges = fe.edges(copy=False)
            checksum = int(cov[::1024].sum(dtype=np.uint64)) + len(hashes) + len(edges)   # touches the result arrays; a full sum would time numpy
        else:
            checksum = counter.download_counts()
        barrier()
        t3 = time.time()
        if it > 0:
            e2e_ms.append((t3 - t0) * 1e3)
            for name, dt in (("run_ms", t1 - t0), ("exchange_ms", t2 - t1), ("export_ms", t3 - t2)):
                e2e_parts[name] = e2e_parts.get(name, 0.0) + dt * 1e3 / args.e2e_steps
        st_e = fe.stats()
        h2d, d2h = st_up["h2d_bytes"], st_e["d2h_bytes"]
    # ---- end to end FROM A FILE (row a1 inside the timed region, as in the reference arm, which parses its FASTA while it
    # runs): mbgfe_add_file + mbgfe_run on the FASTA of this rank's reads (page cache warm, as for the reference) -> ordered table
    # in host memory.  Plain FASTA goes through the streaming loader (pread + scan + 2-bit packing on worker threads, parse ==
    # pack); a .gz is inflated by zlib on one thread, which is then the bound (measured on a bounded sample).
    e2e_file = None
    if world == 1 and not args.no_file:
        import gzip
        import shutil
        tmp_root = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 3 * n_bases else None
        with tempfile.TemporaryDirectory(dir=tmp_root) as d:
            fa = os.path.join(d, "reads.fasta")
            write_sample_fasta(fa, bases, offsets)
            file_bytes = os.path.getsize(fa)
            with open(fa, "rb") as f:              # warm the page cache the way a just-written corrected-read file is
                while f.read(1 << 26):
                    pass
            ms_file = []
            for it in range(1 + args.e2e_steps):
                barrier()
                t0 = time.time()
                fe.reset()
                fe.add_file(fa)
                fe.run()
                cov = fe.node_coverage(copy=False)
                hashes = fe.node_hashes(copy=False)
                edges = fe.edges(copy=False)
                checksum_file = int(cov[::1024].sum(dtype=np.uint64)) + len(hashes) + len(edges)
                barrier()
                if it > 0:
                    ms_file.append((time.time() - t0) * 1e3)
            st_f = fe.stats()
            e2e_file = {"value": n_bases / (float(np.mean(ms_file)) * 1e-3), "unit": "bases/s", "ms_per_step": float(np.mean(ms_file)), "file_bytes": file_bytes,
                        "h2d_bytes_per_step": int(st_f["h2d_bytes"]), "d2h_bytes_per_step": int(st_f["d2h_bytes"]), "checksum": checksum_file,
                        "same_table_as_e2e": bool(checksum_file == checksum), "input": "plain FASTA, one sequence line per read, page cache warm"}
            # .gz on a bounded sample (zlib inflate on one thread is the bound, as it is for the reference's zstr reader)
            n_gz = min(len(offsets) - 1, 3000)
            gz = os.path.join(d, "sample.fasta.gz")
            with gzip.open(gz, "wb", compresslevel=1) as f:
                for i in range(n_gz):
                    f.write(b">r%d\n" % i + bases[int(offsets[i]):int(offsets[i + 1])].tobytes() + b"\n")
            gz_bases = int(offsets[n_gz])
            ms_gz = []
            for it in range(2):
                t0 = time.time()
                fe.reset()
                fe.add_file(gz)
                fe.run()
                fe.node_coverage(copy=False)
                fe.edges(copy=False)
                ms_gz.append((time.time() - t0) * 1e3)
            e2e_file["gz"] = {"value": gz_bases / (min(ms_gz) * 1e-3), "unit": "bases/s", "sample": f"{n_gz} reads / {gz_bases} bases, gzip -1", "bound": "zlib inflate, one thread"}
            fe.reset()
            fe.add_reads(bases, offsets)          # back to the full in-memory batch for the stages below (untimed)
            fe.run()
    # ---- pass 3 of runMBG: the same reads with hpc variants registered (row a11); the re-hash runs per piece inside the
    # pipelined run, so the leg costs one more small kernel per piece and the run-length positions K1 has to write
    e2e_variants = None
    if world == 1 and not args.no_file:
        nh = fe.node_hashes(copy=True)
        pick = nh[:: max(1, len(nh) // 6600)]
        k_ = wl["k"]
        v_hash = np.repeat(pick, 3, axis=0)
        v_off = np.tile(np.array([k_ // 4, k_ // 2, (3 * k_) // 4], dtype=np.uint32), len(pick))
        v_len = np.tile(np.array([2, 4, 1000], dtype=np.uint32), len(v_off))
        v_begin = np.arange(len(v_off) + 1, dtype=np.uint64) * 3
        ms_var = []
        for it in range(1 + args.e2e_steps):
            barrier()
            t0 = time.time()
            fe.reset()
            fe.add_reads(bases, offsets)
            fe.set_variants(v_hash, v_off, v_begin, v_len)
            fe.run()
            cov = fe.node_coverage(copy=False)
            hashes = fe.node_hashes(copy=False)
            edges = fe.edges(copy=False)
            _ = int(cov[::1024].sum(dtype=np.uint64)) + len(hashes) + len(edges)
            barrier()
            if it > 0:
                ms_var.append((time.time() - t0) * 1e3)
        e2e_variants = {"value": n_bases / (float(np.mean(ms_var)) * 1e-3), "unit": "bases/s", "ms_per_step": float(np.mean(ms_var)), "variant_entries": int(len(v_off)),
                        "vs_pass1": float(np.mean(ms_var)) / float(np.mean(e2e_ms))}
        fe.reset()
        fe.add_reads(bases, offsets)          # back to the plain batch for the stages below (untimed)
        fe.run()
    # ---- the step after the table (rows a14 / f1), outside the timed regions: abundance filter + unitig chains on the device
    unitig_info = None
    if world == 1 and not sharded:
        fe.build_unitigs(1, 2.0)                      # Verkko's buildGraph flags: --kmer-abundance 1 --unitig-abundance 2
        u = fe.build_unitigs(1, 2.0)
        st_u = fe.stats()
        unitig_info = {"ms": st_u["ms_unitigs"], "kmer_abundance": 1, "unitig_abundance": 2.0, "nodes": int(len(u["kept"])), "kept_nodes": int(len(u["node_cov"])),
                       "kept_edges": int(len(u["edges"])), "unitigs": int(len(u["unitig_begin"]) - 1), "launches": st_u["launches"]}
        # the same with getUnitigGraph's oneCovHeuristic, as runMBG builds its final graph under Verkko's flags (MBG.cpp:1573)
        fe.build_unitigs(1, 2.0, one_cov_heuristic=True)
        u = fe.build_unitigs(1, 2.0, one_cov_heuristic=True)
        st_u = fe.stats()
        unitig_info["one_cov_heuristic"] = {"ms": st_u["ms_unitigs"], "kept_nodes": int(len(u["node_cov"])), "kept_edges": int(len(u["edges"])),
                                            "unitigs": int(len(u["unitig_begin"]) - 1), "unitig_edges": int(len(u["unitig_edges"])), "launches": st_u["launches"]}
        # the last step of row f1 under MBG's defaults: which k-mers stay inside the unitigs (filterKmersToUnitigKmers)
        fe.thin_unitigs()
        thin, _ = fe.thin_unitigs()
        st_n = fe.stats()
        unitig_info["thin_unitigs"] = {"ms": st_n["ms_unitigs"], "kmers_in": int(len(thin)), "kmers_kept": int(thin.sum()), "launches": st_n["launches"]}
        # row f2: every resident k-mer instance threaded through those unitigs (getReadPaths as a batched lookup)
        fe.thread_reads()
        th = fe.thread_reads()
        st_t = fe.stats()
        steps = np.bincount(th["step"], minlength=6)
        unitig_info["thread_reads"] = {"ms": st_t["ms_unitigs"], "instances": int(len(th["step"])), "launches": st_t["launches"],
                                       "located": int(steps[2:].sum()), "paths": int(steps[2] + steps[3]), "unitig_crossings": int(steps[5])}
    e2e_t = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = total_bases / (float(e2e_t[0]) * 1e-3)

    if rank == 0:
        peak, peak_src = measured_peak()
        R, C1, C2, S = st["raw_bases"], st["collapsed_symbols"], st["codes"], st["selected"]
        per_step = {k_: v / args.steps for k_, v in stage_ms.items()}
        # algorithmic bytes per launch, from this run's exact counts
        alg = {
            "k_collapse": (R / 4 + C1 / 4, per_step["ms_collapse"]),
            "k_msat": (C1 / 4 + 2 * C2, per_step["ms_msat"]),
            "k_select": (2 * C2 + C2 / 8, per_step["ms_select"]),
            "k_kmer_hash": (2 * C2 + 20 * S, per_step["ms_hash"]),
            "k_insert_nodes": (64 * S + 16 * S, per_step["ms_nodes"]),
            "k_insert_edges": (64 * S + 20 * S, per_step["ms_edges"]),
        }
        dominant = max(alg, key=lambda n: alg[n][1])
        byts, ms = alg[dominant]
        achieved = byts / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        line = {
            "metric": METRIC, "value": value, "unit": "bases/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16 codes / u64 hashes (integer)", "data": "synthetic",
            "config": {"workload": wl["desc"] + (f" x{args.scale}" if args.scale != 1.0 else ""), "k": wl["k"], "w": wl["w"], "error_masking": "collapse-msat",
                       "reads_per_gpu": len(offsets) - 1, "bases_per_gpu": n_bases, "parallelism": f"reads sharded over {world} GPU(s), k-mers owned by hash range",
                       "l2": "inputs and intermediates (>= 0.48 GB packed reads per pass) are larger than the 126 MB L2; no flush needed"},
            "kmers_per_s": total_sel / (ms_per_step * 1e-3),
            "counts": {"raw_bases": R, "collapsed_symbols": C1, "codes": C2, "selected_kmers": S, "distinct_kmers": parts_stats["distinct"], "distinct_edges": parts_stats["distinct_edges"]},
            "stage_ms_per_step": per_step, "wall_ms_per_step": wall_ms / args.steps,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "bases/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": float(e2e_t[0]), "checksum": checksum,
                    "wall_ms": e2e_parts, "device_ms": {"h2d_all_pieces": st_up["ms_upload"], "hash": st_up["ms_hash"], "nodes": st_up["ms_nodes"], "edges": st_up["ms_edges"], "export_sort_d2h": st_e["ms_export"]}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": peak_src,
                         # what ncu says limits each kernel (profiles/r02_ncu_full_front_scale1.txt); none of them is DRAM-bound
                         "limiter": {"k_select": "ALU pipe 83 % and LSU data pipe 83 %: 64-bit min / compare / xor / rotate are 2-4 32-bit ALU instructions each, table and ring reads; issue 64 %",
                                     "k_kmer_hash": "a chain of latencies: issue 50 %, LSU data pipe 79 % (shared-memory f loads), fmaheavy 53 % (Murmur64A multiplies), barrier stall 9.9 per issue",
                                     "k_msat": "instruction issue 80 %, ALU pipe 77 %, divergence in the per-lane candidate and stretch loops",
                                     "k_collapse": "LSU data pipe 91 % (byte-table lookups in shared memory), issue 59 %",
                                     "k_insert_nodes": "DRAM sectors + atomics (random 32-byte slots)", "k_insert_edges": "DRAM sectors + atomics (random 32-byte slots)"},
                         "all_kernels_GBps": {n: (b / (m * 1e-3) / 1e9 if m > 0 else None) for n, (b, m) in alg.items()}},
        }
        if e2e_file:
            line["e2e_file"] = e2e_file
        if e2e_variants:
            line["e2e_pass3_variants"] = e2e_variants
        if unitig_info:
            line["unitig_stage"] = unitig_info
        traffic_file = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(traffic_file) and args.scale == 1.0 and args.workload == "chr20" and not (args.k or args.w):
            # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed `ncu --set full` capture of this workload
            line["roofline"]["traffic"] = json.load(open(traffic_file)).get(dominant)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], line["sample_digest"], bad = cpu_baseline(wl, args.scale, local_rank)
            line["parity_sample"] = "ok" if not bad else "MISMATCH in " + ",".join(bad)
            if bad:
                emit(json.dumps(line))
                raise SystemExit("bench: the CUDA path and the reference disagree on the sample: " + ",".join(bad))
        emit(json.dumps(line))
    if world > 1:
        dist.barrier()
    fe.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
