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
ALG_BYTES_FUSED = REC / KMERS_PER_READ + 24.0
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

        def step():
            idx.clear()
            n = idx.build_fastq_dev(text.data_ptr(), nbytes)
            assert n == n_kmers_rank, (n, n_kmers_rank)
        launches_of = idx.launch_count
    else:
        from kmerind_b200.dist import ShardedIndex
        idx = ShardedIndex(K, hash=kb.HASH_FARM, device=local_rank, reserve=reserve)
        peer = False
        if not os.environ.get("KMI_BENCH_NCCL_A2A"):
            peer = idx.enable_peer_exchange(n_kmers_rank // world + n_kmers_rank // (4 * world) + 4096)

        def step():
            idx.local.clear()
            n = idx.build_fastq_dev(text, nbytes, max_kmers=n_kmers_rank)
            assert n == n_kmers_rank, (n, n_kmers_rank)
        launches_of = idx.local.launch_count

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
        if world == 1:
            kern_ms.append(idx.last_kernel_ms())
        else:
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
    # is the cpu_baseline (it is the reference path on this box's host cores, timed).
    parity, cpu_base = None, None
    if world == 1:
        cks = idx.checksum()
        assert cks[1] == n_kmers_rank, "sum of counts %d != k-mers %d" % (cks[1], n_kmers_rank)
        distinct = idx.size()
        info = idx.table_info()
        if not os.environ.get("KMI_BENCH_NO_CPU"):
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
        cks = idx.local.checksum()
        tsum = torch.tensor([cks[1]], dtype=torch.int64, device=device)
        dist.all_reduce(tsum)
        assert int(tsum.item()) == total_kmers, "sum of counts %d != k-mers %d" % (int(tsum.item()), total_kmers)
        distinct = idx.size()
        info = idx.local.table_info()

    # ---- e2e: host (pinned) buffer through the public host-pointer call, D2H read of the size
    e2e = None
    if world == 1:
        host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        host.copy_(text[:nbytes])
        torch.cuda.synchronize()
        host_np = host.numpy()
        e2e_steps = max(1, min(args.steps, 3))
        for _ in range(1):
            idx.clear(); idx.build_fastq(host_np)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            idx.clear()
            n = idx.build_fastq(host_np)
            sz = idx.size()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        e2e = {"value": n_kmers_rank / dt, "unit": "kmers/s", "h2d_bytes_per_step": int(nbytes), "d2h_bytes_per_step": 72,
               "ms_per_step": dt * 1e3}
        assert n == n_kmers_rank and sz == distinct
    else:
        # every rank's text starts in ITS pinned host buffer; 4 record-aligned chunks, copy of chunk i+1 under the
        # parse / exchange / insert of chunk i (ShardedIndex.build_fastq_host); D2H: the all-reduced index size
        host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        host.copy_(text[:nbytes])
        torch.cuda.synchronize()
        n_chunks = 4
        per = (n_reads + n_chunks - 1) // n_chunks
        cuts = [min(i * per, n_reads) * REC for i in range(n_chunks + 1)]
        cap_chunk = per * KMERS_PER_READ
        e2e_steps = max(1, min(args.steps, 3))
        idx.local.clear(); idx.build_fastq_host(host, cuts, cap_chunk)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            idx.local.clear()
            n = idx.build_fastq_host(host, cuts, cap_chunk)
            sz = idx.size()
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=device)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dt = float(dt.item())
        assert n == n_kmers_rank and sz == distinct, (n, sz, distinct)
        e2e = {"value": total_kmers / dt, "unit": "kmers/s", "h2d_bytes_per_step": int(nbytes) * world,
               "d2h_bytes_per_step": 72 * world, "ms_per_step": dt * 1e3,
               "how": "per rank: pinned host text -> 4 chunks, H2D of chunk i+1 under parse/exchange/insert of chunk i"}

    peak, peak_kind = measured_peaks()
    if world == 1:
        km = np.mean(np.array(kern_ms), axis=0)   # [whole build, parse+scatter, fine scatter, insert, overflow]
        kernels = []
        for (name, alg), ms in zip(ALG_BYTES.items(), km[1:4]):
            ach = alg * n_kmers_rank / (max(float(ms), 1e-6) * 1e-3) / 1e9
            kernels.append({"kernel": name, "ms_per_step": float(ms), "alg_bytes_per_kmer": alg, "achieved": ach, "frac": ach / peak})
        dom = max(kernels, key=lambda k_: k_["ms_per_step"])
        # traffic: dram bytes read+written per launch of the dominant kernel, from profiles/ (ncu --set full), if recorded
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get(dom["kernel"].split("<")[0])
            except Exception:
                traffic = None
        roofline = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved"], "peak": peak, "peak_source": peak_kind,
                    "unit": "GB/s", "frac": dom["frac"], "traffic": traffic, "alg_bytes_per_kmer": dom["alg_bytes_per_kmer"],
                    "kernel_ms_per_step": dom["ms_per_step"], "build_ms_per_step": float(km[0]), "kernels": kernels,
                    "whole_build": {"alg_bytes_per_kmer": ALG_BYTES_FUSED,
                                    "achieved": ALG_BYTES_FUSED * n_kmers_rank / (float(km[0]) * 1e-3) / 1e9,
                                    "frac": ALG_BYTES_FUSED * n_kmers_rank / (float(km[0]) * 1e-3) / 1e9 / peak}}
    else:
        smean = {k_: float(np.mean(v_)) for k_, v_ in stage.items()}
        k_ms = smean.get("insert", 0.0)
        # insert of received keys = coarse scatter (16) + fine scatter (16) + segment insert (32) algorithmic B/k-mer
        achieved = 64.0 * idx.n_received / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
        x_ms = smean.get("parse+scatter+exchange", smean.get("a2a", 1e9))
        roofline = {"bound": "hbm", "kernel": "insert of received keys (k_scatter_keys + k_scatter_fine + k_insert_segments)",
                    "achieved": achieved, "peak": peak, "peak_source": peak_kind, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": None, "alg_bytes_per_kmer": 64.0, "kernel_ms_per_step": k_ms, "stage_ms": smean,
                    "nvlink_gbs_per_dir": (8.0 * n_kmers_rank * (world - 1) / world) / (x_ms * 1e-3) / 1e9,
                    "exchange": "peer stores from the parse kernel (CUDA IPC)" if peer else "NCCL all_to_all"}

    if rank == 0:
        line = {"metric": "canonical k-mers inserted/sec into count index", "value": value, "unit": "kmers/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": workload_config(world, n_reads),
                "index": {"distinct_keys": int(distinct), "table_slots_per_gpu": info["slots"], "table_bytes_per_gpu": info["slots"] * info["slot_bytes"],
                          "path": "1xB200 text->buckets->segment insert" if world == 1 else
                                  "%dxB200 %s" % (world, "parse+owner scatter into peer HBM over NVLink (one kernel)->bucketed insert" if peer
                                                 else "parse+owner scatter->NCCL all-to-all->bucketed insert")},
                "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline}
        if e2e is not None:
            line["e2e"] = e2e
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if parity is not None:
            line["parity"] = parity
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kmerind_b200")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
