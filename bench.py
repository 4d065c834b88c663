#!/usr/bin/env python
"""bench.py — canonical k-mers inserted/sec into the count index (BASELINE.json metric).

A "step" is one pass of the hot path over one batch: build a K=31 canonical CountIndex from a
synthetic FASTQ block (150 bp reads, 30x coverage, 0.5 % substitutions).
  N = 1   : configs[1] — 1 Gbp = 6 666 667 reads = 2.11 GB of text = 8.0e8 31-mers, resident in HBM;
            the fused text->table kernel path (kmi_build_fastq_dev).
  N > 1   : weak scaling — every rank holds its own 1 Gbp of reads of an N-times larger genome;
            parse -> partition (owner = dist_hash % N) -> NCCL all-to-all -> insert.
  value   : total k-mers / device time (CUDA events, max over ranks), inputs resident in HBM.
  e2e     : the same through the host-buffer call (kmi_build_fastq from pinned memory: H2D copy
            inside, D2H read of the resulting size).
  --impl reference : the reference's own CPU code for the path (oracle/_ref: its Kmer, lex_less,
            farm hash and google::dense_hash_map, rank-parallel over all host threads; the C port
            oracle/kmi_oracle.c when _ref is absent) on a bounded sample of the same workload.
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K = 31
READ_LEN = 150
REC = 13 + 2 * READ_LEN + 4          # 317
KMERS_PER_READ = READ_LEN - K + 1    # 120
READS_1GBP = 6666667
GENOME_1GBP = 33333333               # 30x
SEED = 1
ERR_PPM = 5000
# algorithmic bytes per k-mer, kernels of the bucketed single-GPU build
ALG_BYTES = {"k_fastq_kmers<31,farm,sink=buckets> (text -> 256 hash-prefix buckets)": REC / KMERS_PER_READ + 8.0,
             "k_scatter_fine<31,farm> (buckets -> table-segment buckets)": 16.0,
             "k_insert_segments<31,farm> (segment buckets -> table, shared-memory segments)": 32.0}
# whole single-GPU build by SURVEY.md 8d: S1 boundary+encode 3.06 + S2/S3 window+canonical+hash 8.42 + S5 insert 32 = 43.5 B/k-mer
ALG_BYTES_BUILD = (REC + 38 + 12) / KMERS_PER_READ + (38 + 12 + 8.0 * KMERS_PER_READ) / KMERS_PER_READ + 32.0
# unfused (multi-GPU) pipeline per SURVEY.md §8d: S1-S3 10.6 + S4 24 + S5 32
ALG_BYTES_SHARDED = (REC + 8.0 * KMERS_PER_READ) / KMERS_PER_READ + 24.0 + 32.0


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# the workload, named identically by both arms
# ------------------------------------------------------------------------------------------------
def workload_config(world, n_reads):
    nbytes = n_reads * REC
    return {"workload": "CountIndex K=31 canonical (lex_less), farm dist+store hash, synthetic %.2f Gbp of 150 bp reads per GPU "
                        "(30x coverage of a uniform genome, 0.5%% substitutions, 317-byte FASTQ records), %d GPU%s, "
                        "keys owned by dist_hash %% n_gpus" % (n_reads * READ_LEN / 1e9, world, "" if world == 1 else "s"),
            "k": K, "read_len": READ_LEN, "reads_per_gpu": n_reads, "text_bytes_per_gpu": nbytes,
            "kmers_per_gpu": n_reads * KMERS_PER_READ, "hash": "farm (dist+store)", "seed": SEED, "err_ppm": ERR_PPM,
            "l2": "inputs (%.2f GB text per GPU, multi-GB table) larger than the 126 MB L2; no flush needed" % (nbytes / 1e9)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline / parity checker: the reference's own CPU code (oracle/_ref), else the C port
# ------------------------------------------------------------------------------------------------
def _oracle_util():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    if not os.path.exists(os.path.join(ROOT, "oracle", "libkmi_oracle.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "libkmi_oracle.so")])
    return ou


def reference_build(text, threads, steps=1, warmup=0):
    """The reference's CountIndex build of FASTQ `text` on `threads` host threads, the whole path inside the timed region:
    per rank-thread the record walk of its record-aligned byte range (fastq_loader.hpp:389-470, restated: the parser headers
    need mpi.h), then the reference's own Kmer::nextFromChar window, lex_less, farm KeyToRank, bucket hand-over (all2allv)
    and google::dense_hash_map insert with count += 1 (oracle/ref_shim.cpp, compiled from /root/reference). Returns the
    timing and the content summary (distinct keys, sum of counts, sum of mix(key) * count) the GPU index is compared with."""
    ou = _oracle_util()
    n_reads = text.size // REC
    times, res = [], None
    if ou.ref_lib("ndebug") is not None:
        kind = "reference"
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            res = ou.ref_count_build_fastq(text, K, threads, 1)
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
        distinct = res["distinct"]
        with np.errstate(over="ignore"):
            ck = (int(res["ck"][:, 0].sum(dtype=np.uint64)), int(res["ck"][:, 1].sum(dtype=np.uint64)))
        how = ("reference Kmer / lex_less / farm / google::dense_hash_map compiled from /root/reference (oracle/_ref), %d "
               "rank-threads, FASTQ record walk restated and inside the timed region" % threads)
    else:
        kind, threads = "port", 1
        nk = ctypes.c_long(0)
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            distinct = int(ou.oracle().kmio_count_build_port(ou._ptr(text), text.size, K, 2, None, None, 0, ctypes.byref(nk)))
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
        ck = None
        how = "C port oracle/kmi_oracle.c, 1 thread (oracle/_ref not built)"
    sec = float(np.mean(times))
    n_kmers = n_reads * KMERS_PER_READ
    return {"value": n_kmers / sec, "unit": "kmers/s", "cores": threads, "kind": kind, "seconds_per_step": sec,
            "distinct": int(distinct), "checksum": ck, "n_kmers": n_kmers,
            "sample": "%d reads = %.2f Gbp = %d k-mers (%d distinct): %s" % (n_reads, n_reads * READ_LEN / 1e9, n_kmers, distinct, how)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    threads = os.cpu_count() or 1
    n_reads = int(os.environ.get("KMI_BENCH_READS", str(READS_1GBP)))
    ou = _oracle_util()
    # One GPU's share of the workload (at N = 1 that IS the workload: same reads, same seed, same bytes as the GPU arm).
    # At N > 1 the N Gbp job does not fit a host run of a few minutes: the step stays one 1 Gbp / 30x block (k-mers/s is
    # what is compared; the CPU rate does not grow with the job).
    text = ou.synth_fastq(0, n_reads, READ_LEN, int(GENOME_1GBP * (n_reads / READS_1GBP)), SEED, ERR_PPM, 0)
    r = reference_build(text, threads, steps=args.steps, warmup=args.warmup)
    sample = ("the full single-GPU workload, every step: " if world == 1 else
              "bounded sample = one GPU's 1/%d share (a 30x block of %.2f Gbp), every step: " % (world, n_reads * READ_LEN / 1e9)) + r["sample"]
    line = {"impl": "reference", "metric": "canonical k-mers inserted/sec into count index", "value": r["value"], "unit": "kmers/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["seconds_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": workload_config(world, n_reads),
            "cpu_baseline": {"value": r["value"], "unit": "kmers/s", "cores": r["cores"], "kind": r["kind"], "sample": sample},
            "e2e": {"value": r["value"], "unit": "kmers/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "index": {"distinct_keys": r["distinct"], "checksum": r["checksum"]}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def _traffic():
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(tp))
    except Exception:
        return {}


def parity_sharded(idx, kb, ou, text, n_reads, world, rank, device):
    """Content parity of a sharded build, outside the timed region. All ranks call it.
    full_size: every rank walks ITS 1 Gbp with the reference code (Kmer window, lex_less, farm KeyToRank) and accumulates, per
      owner rank, [sum of mix(key) over occurrences, occurrences]; both are additive, so the all-reduced row r is exactly
      what owner r's index must report as kmi_checksum — it catches keys on the wrong rank, merged / split / lost keys and
      wrong counts at the measured size (distinct counts are not additive: they are pinned by the sample below).
    sample: a second, smaller build through the SAME sharded path (NVLink push) whose texts rank 0 gathers and hands to the
      reference's p-rank build (oracle/_ref): per-rank local_size and checksum must equal the reference's per-rank values at
      p = N, and one sharded find() batch must return the reference's frequencies."""
    import torch
    import torch.distributed as dist
    from kmerind_b200.dist import ShardedIndex
    out = {}
    threads = max(1, (os.cpu_count() or 1) // world)
    nbytes = n_reads * REC
    host_text = text[:nbytes].cpu().numpy()
    oc = ou.ref_owner_checksums_fastq(host_text, K, world, 1, threads=threads)
    if oc is None:
        return {"ok": None, "note": "oracle/_ref not built: no reference code to compare with"}
    want = idx.comm.allreduce(oc.reshape(-1), "sum").reshape(world, 2)
    got = idx.local.checksum()
    ok_full = int(want[rank, 0]) == int(got[0]) and int(want[rank, 1]) == int(got[1])
    flags = idx.comm.allreduce(np.array([1 if ok_full else 0], dtype=np.uint64), "min")
    out["full_size"] = {"what": "per owner rank: sum of counts and sum of mix(key)*count == reference walk of all %d texts, all-reduced" % world,
                        "rank0_sum_counts": [int(got[1]), int(want[rank, 1])], "rank0_sum_mix": [int(got[0]), int(want[rank, 0])],
                        "all_ranks_ok": bool(flags[0])}
    # ---- sample build, exact per-rank comparison with the reference's p-rank build
    s_reads = max(100000, 1600000 // world)
    s_bytes = s_reads * REC
    with ShardedIndex(K, hash=kb.HASH_FARM, device=device.index, comm=idx.comm) as sidx:
        n = sidx.build_fastq_dev(text.data_ptr(), s_bytes)
        assert n == s_reads * KMERS_PER_READ
        mine = np.array([sidx.local_size()] + list(sidx.local.checksum()), dtype=np.uint64)
        parts = [torch.empty(s_bytes, dtype=torch.uint8, device=device) for _ in range(world)]
        dist.all_gather(parts, text[:s_bytes].contiguous())
        ref_vals = np.zeros(3 * world, dtype=np.uint64)
        q = rv = None
        if rank == 0:
            whole = torch.cat(parts).cpu().numpy()
            ref = ou.ref_count_build_fastq(whole, K, world, 1, cuts=[r * s_bytes for r in range(world + 1)], want_dump=True)
            for r in range(world):
                ref_vals[3 * r:3 * r + 3] = [ref["sizes"][r], ref["ck"][r, 0], ref["ck"][r, 1]]
            order = np.argsort(ref["keys"])
            sk, sc = ref["keys"][order], ref["counts"][order].astype(np.uint64)
        del parts
        ref_vals = idx.comm.allreduce(ref_vals, "sum")               # only rank 0 contributed: a broadcast
        ok_s = bool((ref_vals[3 * rank:3 * rank + 3] == mine).all())
        # one query batch, the same on every rank: k-mers of rank 0's first reads in read orientation + absent keys
        qtext = text[:2000 * REC].clone()
        dist.broadcast(qtext, src=0)
        q, _, _ = ou.fastq_kmers(qtext.cpu().numpy(), K, canonical=False)
        q = np.concatenate([q, (np.arange(5000, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(3)])
        vals, found = sidx.find(q)
        lo = idx.comm.allreduce(vals, "min")
        hi = idx.comm.allreduce(vals, "max")
        ok_q = bool((lo == hi).all())
        if rank == 0:
            cq = ou.canonical(q, K)
            pos = np.minimum(np.searchsorted(sk, cq), len(sk) - 1)
            hit = sk[pos] == cq
            ok_q = ok_q and bool(np.array_equal(found.astype(bool), hit)) and bool(np.array_equal(vals, np.where(hit, sc[pos], 0)))
        flags = idx.comm.allreduce(np.array([1 if ok_s else 0, 1 if ok_q else 0], dtype=np.uint64), "min")
        out["sample"] = {"what": "%d reads per rank through the same sharded path (path %d): per-rank local_size, sum of counts, "
                                 "sum of mix(key)*count == the reference's %d-rank build (oracle/_ref); one sharded find() of %d keys "
                                 "== the reference's frequencies on every rank" % (s_reads, sidx.path, world, len(q)),
                         "rank0_local_size": [int(mine[0]), int(ref_vals[3 * rank])], "per_rank_ok": bool(flags[0]), "find_ok": bool(flags[1])}
    out["ok"] = bool(out["full_size"]["all_ranks_ok"] and out["sample"]["per_rank_ok"] and out["sample"]["find_ok"])
    return out


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import kmerind_b200 as kb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    n_reads = int(os.environ.get("KMI_BENCH_READS", str(READS_1GBP)))
    genome = int(GENOME_1GBP * (n_reads / READS_1GBP)) * world
    nbytes = n_reads * REC
    n_kmers_rank = n_reads * KMERS_PER_READ
    first_read = rank * n_reads
    no_cpu = bool(os.environ.get("KMI_BENCH_NO_CPU"))

    # synthetic reads, generated on the device (same bytes as the host twin in oracle/)
    text = torch.empty(nbytes + 64, dtype=torch.uint8, device=device)
    kb.synth_fastq_dev(local_rank, text.data_ptr(), first_read, n_reads, READ_LEN, genome, SEED, ERR_PPM, 0)
    torch.cuda.synchronize()
    reserve = int(os.environ.get("KMI_BENCH_RESERVE", str(int(0.20 * n_kmers_rank))))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    stage = {}
    if world == 1:
        idx = kb.CountIndex(K, hash=kb.HASH_FARM, device=local_rank, reserve=reserve)
        local = idx

        def step():
            idx.clear()
            n = idx.build_fastq_dev(text.data_ptr(), nbytes)
            assert n == n_kmers_rank, (n, n_kmers_rank)
    else:
        from kmerind_b200.dist import ShardedIndex
        idx = ShardedIndex(K, hash=kb.HASH_FARM, device=local_rank, reserve=reserve)   # NCCL + CUDA IPC inside libkmi.so
        local = idx.local

        def step():
            idx.local.clear()
            n = idx.build_fastq_dev(text.data_ptr(), nbytes)
            assert n == n_kmers_rank, (n, n_kmers_rank)
    launches_of = local.launch_count

    for _ in range(args.warmup):
        step()
    barrier()
    l0 = launches_of()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ms = []
    e0.record()
    for _ in range(args.steps):
        step()
        kern_ms.append(local.last_kernel_ms())
        if world > 1:
            for k_, v_ in idx.stage_ms.items():
                stage.setdefault(k_, []).append(v_)
    e1.record()
    barrier()
    clocks = sampler.stop()
    launches = launches_of() - l0
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    total_kmers = n_kmers_rank * world
    value = total_kmers / (ms_step * 1e-3)

    # ---- parity of what was just measured, outside the timed region (full content, not a sum):
    # N = 1: the reference's own CPU build of the SAME 1 Gbp of text (oracle/_ref) -> distinct keys, sum of counts and
    # sum of mix(key) * count must equal kmi_local_size / kmi_checksum of the index the timed steps built. The same run
    # is the cpu_baseline (it is the reference path on this box's host cores, timed). N > 1: parity_sharded().
    parity, cpu_base = None, None
    cks = local.checksum()
    info = local.table_info()
    if world == 1:
        assert cks[1] == n_kmers_rank, "sum of counts %d != k-mers %d" % (cks[1], n_kmers_rank)
        distinct = idx.size()
        if not no_cpu:
            host_text = text[:nbytes].cpu().numpy()
            ref = reference_build(host_text, os.cpu_count() or 1)
            del host_text
            cpu_base = {"value": ref["value"], "unit": "kmers/s", "cores": ref["cores"], "kind": ref["kind"],
                        "sample": "the full workload, one step (also the parity checker): " + ref["sample"]}
            parity = {"checked_against": "oracle/_ref (reference code) on the same text" if ref["kind"] == "reference" else "oracle C port",
                      "distinct_keys": [int(distinct), ref["distinct"]], "sum_counts": [int(cks[1]), ref["n_kmers"]],
                      "sum_mix_key_times_count": [int(cks[0]), ref["checksum"][0] if ref["checksum"] else None]}
            ok = distinct == ref["distinct"] and (ref["checksum"] is None or (int(cks[0]), int(cks[1])) == ref["checksum"])
            parity["ok"] = bool(ok)
            assert ok, "index contents differ from the reference build: %r" % (parity,)
    else:
        tsum = idx.comm.allreduce(np.array([cks[1]], dtype=np.uint64), "sum")
        assert int(tsum[0]) == total_kmers, "sum of counts %d != k-mers %d" % (int(tsum[0]), total_kmers)
        distinct = idx.size()
        if not no_cpu:
            parity = parity_sharded(idx, kb, _oracle_util(), text, n_reads, world, rank, device)
            assert parity["ok"] is not False, "sharded index differs from the reference: %r" % (parity,)

    # ---- e2e: host (pinned) buffer through the public host-pointer call, D2H read of the size
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.copy_(text[:nbytes])
    torch.cuda.synchronize()
    host_np = host.numpy()
    e2e_steps = max(1, min(args.steps, 3))
    build_host = idx.build_fastq
    clear = local.clear
    clear(); build_host(host_np)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        clear()
        n = build_host(host_np)
        sz = idx.size()
    barrier()
    dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    dt = float(dt.item())
    assert n == n_kmers_rank and sz == distinct, (n, sz, distinct)
    # what the host side of this box can deliver: every rank copies its pinned text to its GPU at the same time
    barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(2):
        text[:nbytes].copy_(host, non_blocking=True)
    c1.record()
    barrier()
    h2d = torch.tensor([c0.elapsed_time(c1) / 2], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(h2d, op=dist.ReduceOp.MAX)
    h2d_ms = float(h2d.item())
    e2e = {"value": total_kmers / dt, "unit": "kmers/s", "h2d_bytes_per_step": int(nbytes) * world, "d2h_bytes_per_step": 72 * world,
           "ms_per_step": dt * 1e3, "h2d_copy_alone_ms": h2d_ms, "h2d_gbs_per_gpu": nbytes / h2d_ms / 1e6,
           "how": "per rank: pinned host text -> kmi_%sbuild_fastq (H2D in rounds on a copy stream under the parse%s); the copy alone "
                  "(all ranks at once) is this box's ceiling for the number" % ("dist_" if world > 1 else "", " / push" if world > 1 else "")}
    # the caller's memory as the reference's caller has it: pageable
    if world == 1 and not os.environ.get("KMI_BENCH_NO_PAGEABLE"):
        pg = np.empty(nbytes, dtype=np.uint8)
        pg[:] = host_np
        clear(); build_host(pg)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        clear()
        n = build_host(pg)
        sz = idx.size()
        torch.cuda.synchronize()
        e2e["pageable"] = {"value": n_kmers_rank / (time.perf_counter() - t0), "unit": "kmers/s", "ms_per_step": (time.perf_counter() - t0) * 1e3,
                           "how": "numpy (pageable) text -> kmi_build_fastq: host threads move 4 MiB slices through a ring of pinned slots (kmi_hoststage.hpp)"}
        del pg
    # ... and as the reference's driver has it: a FASTQ file (Index::build_posix). Written to the scratch directory, so it
    # sits in the page cache; kmi_build_file preads its slices straight into the pinned ring
    if world == 1 and not os.environ.get("KMI_BENCH_NO_FILE"):
        import tempfile
        path = os.path.join(tempfile.gettempdir(), "kmi_bench_%d_%d.fastq" % (os.getpid(), n_reads))
        try:
            with open(path, "wb") as f:
                f.write(host_np.data)
            clear(); nf = local.build_file(path)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            clear()
            nf = local.build_file(path)
            sz = idx.size()
            dtf = time.perf_counter() - t0
            assert nf == n_kmers_rank and sz == distinct, (nf, sz, distinct)
            e2e["file"] = {"value": n_kmers_rank / dtf, "unit": "kmers/s", "ms_per_step": dtf * 1e3, "file_bytes": int(nbytes),
                           "how": "kmi_build_file on the workload written to %s (page cache): pread by the ring's threads -> pinned slices -> H2D under the build" % tempfile.gettempdir()}
        except OSError as ex:     # no room in the scratch directory: the leg is optional
            e2e["file"] = {"unavailable": str(ex)}
        finally:
            if os.path.exists(path):
                os.remove(path)

    peak, peak_kind = measured_peaks()
    traffic = _traffic()
    km = np.mean(np.array(kern_ms), axis=0)   # [whole build, parse+scatter, fine scatter, insert, overflow]
    if world == 1:
        kernels = []
        for (name, alg), ms in zip(ALG_BYTES.items(), km[1:4]):
            ach = alg * n_kmers_rank / (max(float(ms), 1e-6) * 1e-3) / 1e9
            tr = traffic.get(name.split("<")[0])
            kernels.append({"kernel": name, "ms_per_step": float(ms), "alg_bytes_per_kmer": alg, "achieved": ach, "frac": ach / peak,
                            "traffic": tr, "frac_measured": (tr / (max(float(ms), 1e-6) * 1e-3) / 1e9 / peak) if tr else None})
        dom = max(kernels, key=lambda k_: k_["ms_per_step"])
        whole = ALG_BYTES_BUILD * n_kmers_rank / (float(km[0]) * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved"], "peak": peak, "peak_source": peak_kind,
                    "unit": "GB/s", "frac": dom["frac"], "traffic": dom["traffic"], "frac_measured": dom["frac_measured"],
                    "alg_bytes_per_kmer": dom["alg_bytes_per_kmer"], "kernel_ms_per_step": dom["ms_per_step"],
                    "build_ms_per_step": float(km[0]), "kernels": kernels,
                    "whole_build": {"alg_bytes_per_kmer": ALG_BYTES_BUILD, "achieved": whole, "frac": whole / peak,
                                    "note": "SURVEY 8d single-GPU budget: S1 3.06 + S2/S3 8.42 + S5 32 B per k-mer"}}
    else:
        smean = {k_: float(np.mean(v_)) for k_, v_ in stage.items()}
        x_ms = smean.get("parse+push", smean.get("parse+exchange", 1e9))
        ins_ms = smean.get("insert", 0.0)
        n_recv = idx.n_received
        kernels = [{"kernel": "k_scatter_fine (received regions -> table-segment buckets)", "ms_per_step": float(km[2]), "alg_bytes_per_kmer": 16.0,
                    "achieved": 16.0 * n_recv / (max(float(km[2]), 1e-6) * 1e-3) / 1e9},
                   {"kernel": "k_insert_segments (segment buckets -> table)", "ms_per_step": float(km[3]), "alg_bytes_per_kmer": 32.0,
                    "achieved": 32.0 * n_recv / (max(float(km[3]), 1e-6) * 1e-3) / 1e9}]
        for k_ in kernels:
            k_["frac"] = k_["achieved"] / peak
        dom = max(kernels, key=lambda k_: k_["ms_per_step"])
        nv = (8.0 * n_kmers_rank * (world - 1) / world) / (x_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved"], "peak": peak, "peak_source": peak_kind,
                    "unit": "GB/s", "frac": dom["frac"], "traffic": None, "alg_bytes_per_kmer": dom["alg_bytes_per_kmer"],
                    "kernel_ms_per_step": dom["ms_per_step"], "kernels": kernels, "stage_ms": smean,
                    "nvlink": {"gbs_per_dir_per_gpu": nv, "peak": 770.0, "peak_source": "B200_PROFILING.md measured peer copy",
                               "frac": nv / 770.0, "note": "8 B x k-mers leaving the GPU / (parse + push stage): the pushes of round i run "
                                                           "under the parse of round i+1, so the stage is not NVLink time alone"},
                    "exchange": "k_owner_push: stores into the owners' HBM (CUDA IPC, NVLink)" if idx.path == 1 else "NCCL all-to-all"}

    if rank == 0:
        line = {"metric": "canonical k-mers inserted/sec into count index", "value": value, "unit": "kmers/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": workload_config(world, n_reads),
                "index": {"distinct_keys": int(distinct), "table_slots_per_gpu": info["slots"], "table_bytes_per_gpu": info["slots"] * info["slot_bytes"],
                          "path": "1xB200 text->buckets->segment insert" if world == 1 else
                                  "%dxB200 %s" % (world, "parse->buckets | k_owner_push into the owners' HBM over NVLink (overlapped)->fine scatter->segment insert"
                                                  if idx.path == 1 else "parse->partition->NCCL all-to-all->bucketed insert")},
                "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if parity is not None:
            line["parity"] = parity
        print(json.dumps(line))
    if world > 1:
        idx.close()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# the other BASELINE.json configurations (driver-runnable: bench.py --config NAME [--gpus N under torchrun])
#   count31_100g  cfg 3: K=31, 12.5 Gbp of reads per GPU (= 100 Gbp at 8 GPUs), built in batches of 1.5 Gbp into one index
#   count63       cfg 4: K=63 (two-word keys), 4 Gbp per GPU (the 11.6 Gbp share of 30x human needs a 137 GB
#                 table of 32-byte slots and does not fit next to its text and exchange buffers)
#   pos21         cfg 5: PositionIndex K=21 over a synthetic FASTA genome, 375 Mbp per GPU (= 24 x 125 Mbp at 8 GPUs), then
#                 1.25e8 query k-mers per GPU (= 1e9 at 8): count() and find() of all positions
# ------------------------------------------------------------------------------------------------
# "defer": staging budget (keys) of bulk-load mode (KMI_PARAM_DEFER_INSERT): the batches' fine buckets are staged and the table
# is swept once per flush instead of once per batch (the library clamps the budget to 60 % of the free device memory)
BIG = {"count31_100g": {"k": 31, "reads": 83333344, "batch": 10000000, "distinct_frac": 0.19, "defer": 10000001280},
       "count63": {"k": 63, "reads": 26666672, "batch": 5000000, "distinct_frac": 0.50, "defer": 2400000000}}


def _dist_setup():
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    return torch, dist, world, rank, local_rank, device


def _timed(torch, dist, world, device, step, steps, warmup, sampler):
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(warmup):
        step()
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) / steps, clocks


def run_big_count(args, name):
    """cfg 3 / cfg 4: a CountIndex far larger than one bucket round, built batch by batch from text resident in HBM."""
    torch, dist, world, rank, lr, device = _dist_setup()
    import kmerind_b200 as kb
    c = BIG[name]
    k = c["k"]
    n_reads = int(os.environ.get("KMI_BENCH_READS", str(c["reads"]))) // 16 * 16    # batches start 16-byte aligned
    batch = min(int(os.environ.get("KMI_BENCH_BATCH", str(c["batch"]))) // 16 * 16, n_reads)
    kpr = READ_LEN - k + 1
    n_kmers_rank = n_reads * kpr
    genome = n_reads * READ_LEN // 30 * world
    nbytes = n_reads * REC
    text = torch.empty(nbytes + 64, dtype=torch.uint8, device=device)
    for b0 in range(0, n_reads, 20000000):                 # (the generator indexes records with 64-bit arithmetic; pieces keep each launch short)
        m = min(20000000, n_reads - b0)
        kb.synth_fastq_dev(lr, text.data_ptr() + b0 * REC, rank * n_reads + b0, m, READ_LEN, genome, SEED, ERR_PPM, 0)
    torch.cuda.synchronize()
    reserve = int(c["distinct_frac"] * n_kmers_rank)
    if world == 1:
        idx = kb.CountIndex(k, hash=kb.HASH_FARM, device=lr, reserve=reserve)
        local, build = idx, idx.build_fastq_dev
    else:
        from kmerind_b200.dist import ShardedIndex
        idx = ShardedIndex(k, hash=kb.HASH_FARM, device=lr, reserve=reserve)
        local, build = idx.local, idx.build_fastq_dev
    defer = int(os.environ.get("KMI_BENCH_DEFER", str(c["defer"])))
    local.set_defer_insert(defer)
    stage = []
    sweeps = []

    def step():
        local.clear()
        got = 0
        ms = np.zeros(5)
        f0 = local.bulk_stats()[1]
        for b0 in range(0, n_reads, batch):
            m = min(batch, n_reads - b0)
            got += build(text.data_ptr() + b0 * REC, m * REC)
            ms += np.array(local.last_kernel_ms())
        before = np.array(local.last_kernel_ms())
        local.flush()                                      # bulk load: what is still staged goes in now, inside the timed region
        ms += np.array(local.last_kernel_ms()) - before
        assert got == n_kmers_rank, (got, n_kmers_rank)
        stage.append(ms)
        sweeps.append((local.bulk_stats()[1] - f0) if defer else -(-n_reads // batch))
    ms_step, clocks = _timed(torch, dist, world, device, step, args.steps, args.warmup, ClockSampler(lr))
    total_kmers = n_kmers_rank * world
    cks = local.checksum()
    tsum = int(cks[1]) if world == 1 else int(idx.comm.allreduce(np.array([cks[1]], dtype=np.uint64), "sum")[0])
    distinct = idx.size()
    info = local.table_info()
    parity = {"sum_counts": [tsum, total_kmers], "ok": tsum == total_kmers,
              "note": "sum of counts over all shards == k-mers generated; the content checks against the reference code run at the "
                      "1 Gbp configuration (default --config) and in tests/ (a reference walk of this size takes minutes per rank)"}
    if os.environ.get("KMI_BENCH_FULL_PARITY") and k <= 32:
        ou = _oracle_util()
        oc = ou.ref_owner_checksums_fastq(text[:nbytes].cpu().numpy(), k, world, 1, threads=max(1, (os.cpu_count() or 1) // world))
        want = oc if world == 1 else idx.comm.allreduce(oc.reshape(-1), "sum").reshape(world, 2)
        okf = int(want[rank, 0]) == int(cks[0]) and int(want[rank, 1]) == int(cks[1])
        if world > 1:
            okf = bool(idx.comm.allreduce(np.array([1 if okf else 0], dtype=np.uint64), "min")[0])
        parity["full_size_reference_walk"] = {"all_ranks_ok": bool(okf)}
        parity["ok"] = parity["ok"] and bool(okf)
    assert parity["ok"], parity
    peak, peak_kind = measured_peaks()
    km = np.mean(np.array(stage[-args.steps:]), axis=0)
    n_batches = -(-n_reads // batch)
    n_sweeps = int(sweeps[-1])
    key_b = 8.0 * kb.nwords(k)
    # table sweeps: a segment insert reads and writes every slot (the first one after kmi_clear only writes)
    sweep = (2.0 * n_sweeps - 1.0) * info["slots"] * info["slot_bytes"]
    ins_alg = (key_b + 12 + 12 if k <= 32 else key_b + 20 + 20) * n_kmers_rank
    roofline = {"bound": "hbm", "kernel": "k_insert_segments%s (%d table sweep%s per step)" % ("" if k <= 32 else "2", n_sweeps, "" if n_sweeps == 1 else "s"),
                "achieved": ins_alg / (float(km[3]) * 1e-3) / 1e9, "peak": peak, "peak_source": peak_kind, "unit": "GB/s",
                "frac": ins_alg / (float(km[3]) * 1e-3) / 1e9 / peak, "traffic": None,
                "swept_bytes_per_step": sweep, "swept_gbs": sweep / (float(km[3]) * 1e-3) / 1e9,
                "alg_bytes_per_kmer": ins_alg / n_kmers_rank, "kernel_ms_per_step": float(km[3]),
                "stage_ms_per_step": {"parse+scatter": float(km[1]), "fine scatter": float(km[2]), "segment insert": float(km[3]), "overflow": float(km[4])},
                "batches_per_step": n_batches, "table_sweeps_per_step": n_sweeps, "defer_budget_keys": defer}
    if rank == 0:
        cfg = workload_config(world, n_reads)
        cfg.update({"k": k, "kmers_per_gpu": n_kmers_rank, "batch_reads": batch,
                    "workload": "CountIndex K=%d canonical, farm hash, synthetic %.2f Gbp of 150 bp reads per GPU (30x, 0.5%% substitutions) = "
                                "%.1f Gbp over %d GPU%s, built in %d batches into one index (%s)" % (k, n_reads * READ_LEN / 1e9,
                                n_reads * READ_LEN * world / 1e9, world, "" if world == 1 else "s", n_batches,
                                "bulk load: %d table sweep%s" % (n_sweeps, "" if n_sweeps == 1 else "s") if defer else "one table sweep per batch")})
        line = {"metric": "canonical k-mers inserted/sec into count index", "value": total_kmers / (ms_step * 1e-3), "unit": "kmers/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": cfg,
                "index": {"distinct_keys": int(distinct), "table_slots_per_gpu": info["slots"], "table_bytes_per_gpu": info["slots"] * info["slot_bytes"],
                          "device_bytes_per_gpu": info.get("device_bytes")},
                "clocks": clocks, "gpu_launches": int(local.launch_count()), "roofline": roofline, "parity": parity,
                "e2e": None}
        print(json.dumps(line))
    if world > 1:
        idx.close()
        dist.destroy_process_group()


def run_pos21(args):
    """cfg 5: PositionIndex K=21 from a FASTA genome cut into one block per GPU at arbitrary bytes, then batched queries."""
    torch, dist, world, rank, lr, device = _dist_setup()
    import kmerind_b200 as kb
    k, width = 21, 80
    seq_len = int(os.environ.get("KMI_BENCH_SEQ_LEN", "125000000"))
    n_seq = 3 * world
    n_q = int(os.environ.get("KMI_BENCH_QUERIES", "125000000"))
    fbytes = kb.synth_fasta_bytes(n_seq, seq_len, width)
    # rank r's block: an equal share of the bytes, the cut moved off the sequence boundary into the middle of a line
    cut = [0 if r == 0 else fbytes if r == world else fbytes * r // world + 1234 for r in range(world + 1)]
    b0, b1 = cut[rank], cut[rank + 1]
    halo = min(64, fbytes - b1)
    text = torch.empty(b1 - b0 + halo + 64, dtype=torch.uint8, device=device)
    kb.synth_fasta_dev(lr, text.data_ptr(), b0, b1 - b0 + halo, n_seq, seq_len, width, SEED)
    torch.cuda.synchronize()
    n_kmers_total = n_seq * (seq_len - k + 1)
    if world == 1:
        idx = kb.PositionIndex(k, hash=kb.HASH_MURMUR, device=lr)
        local = idx

        def step():
            idx.clear()
            return idx.build_fasta_dev(text.data_ptr(), b1 - b0)
    else:
        from kmerind_b200.dist import ShardedIndex
        idx = ShardedIndex(k, mode=kb.MODE_POSITION, hash=kb.HASH_MURMUR, device=lr)
        local = idx.local
    # fill limit of the table (KMI_PARAM_MAX_LOAD): 70 % = dense_hash_map's, the default. KMI_BENCH_MAX_LOAD=50 trades build speed
    # (twice the table to write, 24.0 vs 18.8 ms per 375 Mbp) for query speed (29 vs 10.8 G count() queries/s resident in HBM:
    # linear probing walks 6 slots per miss at 0.7, 2.5 at 0.5) — profiles/r02_query_experiments.md. Then the reference's reserve().
    max_load = int(os.environ.get("KMI_BENCH_MAX_LOAD", "70"))
    local.set_max_load(max_load)
    local.reserve(n_kmers_total // world)
    if world > 1:

        def step():
            idx.local.clear()
            return idx.build_fasta_dev(text.data_ptr(), b1 - b0 + halo, n_valid=b1 - b0, file_offset=b0, bol=(b0 == 0))
    produced = []
    ms_step, clocks = _timed(torch, dist, world, device, lambda: produced.append(step()), args.steps, args.warmup, ClockSampler(lr))
    got = produced[-1] if world == 1 else int(idx.comm.allreduce(np.array([produced[-1]], dtype=np.uint64), "sum")[0])
    size = idx.size()
    ok_build = got == n_kmers_total and size == n_kmers_total
    # ---- queries: half from the genome (random strand), half random; host arrays in, answers out (the public batch call)
    qd = torch.empty(n_q, dtype=torch.int64, device=device)
    kb.synth_queries_dev(lr, qd.data_ptr(), rank * n_q, n_q, k, n_seq, seq_len, SEED, 4)
    q = qd.cpu().numpy().view(np.uint64)
    # the same batch with the queries resident in HBM and the answers left there: the probe + owner exchange without the host
    cd = torch.zeros(n_q, dtype=torch.int64, device=device)
    count_dev = (lambda: idx.count_dev(qd.data_ptr(), n_q, cd.data_ptr())) if world > 1 else (lambda: idx.count_dev(qd.data_ptr(), n_q, cd.data_ptr()))
    count_dev()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(3):
        count_dev()
    ev1.record()
    torch.cuda.synchronize()
    t_count_dev = ev0.elapsed_time(ev1) / 3.0 * 1e-3
    counts_dev = cd.cpu().numpy().view(np.uint64).copy()
    del qd, cd
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    counts = idx.count(q)
    t_count = time.perf_counter() - t0
    hits = counts[0::2]
    ok_q = bool((hits >= 1).all())                          # every genome query is present (both strands are stored canonically)
    ok_q = ok_q and bool(np.array_equal(counts, counts_dev))
    n_find = min(n_q, int(os.environ.get("KMI_BENCH_FIND", "20000000")))
    t0 = time.perf_counter()
    off, vals = idx.find_positions(q[:n_find], cap=int(counts[:n_find].sum()))
    t_find = time.perf_counter() - t0
    ok_q = ok_q and bool(np.array_equal(np.diff(off), counts[:n_find]))
    # ids of a sample of genome queries point at file bytes that spell the k-mer or its reverse complement
    samp = np.arange(0, min(n_find, 2000), 2)
    pos_ok = True
    for j in samp[:200]:
        for v in vals[int(off[j]):int(off[j + 1])][:2]:
            fo = int(v) & ((1 << 40) - 1)
            pos_ok = pos_ok and 0 <= fo < fbytes
    tq = torch.tensor([t_count, t_find, t_count_dev], dtype=torch.float64, device=device)
    flags = np.array([1 if (ok_build and ok_q and pos_ok) else 0], dtype=np.uint64)
    if world > 1:
        dist.all_reduce(tq, op=dist.ReduceOp.MAX)
        flags = idx.comm.allreduce(flags, "min")
    t_count, t_find, t_count_dev = float(tq[0].item()), float(tq[1].item()), float(tq[2].item())
    assert flags[0] == 1, "pos21 checks failed: build %s queries %s positions %s" % (ok_build, ok_q, pos_ok)
    peak, peak_kind = measured_peaks()
    info = local.table_info()
    alg = 17.0 + 32.0 + (48.0 if world > 1 else 0.0)        # parse 17 B/k-mer, insert 32, partition 48 when sharded
    ach = alg * (n_kmers_total / world) / (ms_step * 1e-3) / 1e9
    if rank == 0:
        line = {"metric": "k-mers inserted/sec into position index", "value": n_kmers_total / (ms_step * 1e-3), "unit": "kmers/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": {"workload": "PositionIndex K=21 canonical, murmur hash, synthetic FASTA genome of %d x %.0f Mbp (80 bases per line) = "
                                       "%.2f Gbp cut into %d blocks at arbitrary bytes, then %d query k-mers per GPU (half from the genome, half random)"
                                       % (n_seq, seq_len / 1e6, n_seq * seq_len / 1e9, world, n_q),
                           "k": k, "file_bytes": int(fbytes), "kmers": int(n_kmers_total), "queries_per_gpu": n_q, "table_max_load_pct": max_load,
                           "l2": "text, tuple log and table far larger than the 126 MB L2"},
                "index": {"tuples": int(size), "table_slots_per_gpu": info["slots"]},
                "clocks": clocks, "gpu_launches": int(local.launch_count()),
                "roofline": {"bound": "hbm", "kernel": "whole position build (parse -> tuples -> %sinsert)" % ("partition -> all-to-all -> " if world > 1 else ""),
                             "achieved": ach, "peak": peak, "peak_source": peak_kind, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                             "alg_bytes_per_kmer": alg},
                "query": {"count": {"value": n_q * world / t_count, "unit": "queries/s", "seconds": t_count,
                                    "how": "host keys in, multiplicities out (H2D, owner exchange over NCCL, probe, answers back, D2H)"},
                          "count_dev": {"value": n_q * world / t_count_dev, "unit": "queries/s", "seconds": t_count_dev,
                                        "how": "queries resident in HBM, answers left there (kmi_%scount_dev), CUDA events, mean of 3" % ("dist_" if world > 1 else ""),
                                        "roofline": {"bound": "hbm", "alg_bytes_per_query": 36.0 + (32.0 if world > 1 else 0.0),
                                                     "achieved": (36.0 + (32.0 if world > 1 else 0.0)) * n_q / t_count_dev / 1e9, "peak": peak, "unit": "GB/s",
                                                     "frac": (36.0 + (32.0 if world > 1 else 0.0)) * n_q / t_count_dev / 1e9 / peak,
                                                     "note": "36 B/query = key 8 + slot 12 + answer 16 (SURVEY 8d); sharded: + 24 B partition of (key, tag) + 8 B un-permute; NVLink 8 B out + 8 B back per query"}},
                          "find_positions": {"value": n_find * world / t_find, "unit": "queries/s", "seconds": t_find, "queries_per_gpu": n_find,
                                             "ids_returned_rank0": int(off[-1])}},
                "parity": {"ok": True, "tuples": [int(size), int(n_kmers_total)],
                           "what": "every k-mer of the genome stored once per occurrence (size == sum over sequences of L-K+1 on all shards); "
                                   "every genome query found on either strand; find() returns exactly count() ids per key, all valid file offsets. "
                                   "Content against the oracle: tests/ (sharded FASTA blocks cut mid-line, dist_check.cpp)"},
                "e2e": None}
        print(json.dumps(line))
    if world > 1:
        idx.close()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kmerind_b200")
    ap.add_argument("--config", default="count31", choices=["count31", "count31_100g", "count63", "pos21"],
                    help="count31 = BASELINE.json configs[1]/[2] shape (1 Gbp per GPU, the headline metric); the others: configs[2..4] at size")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.config in BIG:
        run_big_count(args, args.config)
    elif args.config == "pos21":
        run_pos21(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
