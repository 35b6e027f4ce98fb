#!/usr/bin/env python
"""bench.py -- the press hot path on B200: compress + decompress of a synthetic
NA12878-like read batch (BASELINE.json configs[1]: svb16-zd, 500 000 reads,
lognormal lengths), reported as GB/s of raw int16 signal.

    python bench.py --gpus N --steps K --warmup W            (N > 1: under torchrun)
    python bench.py --impl reference ...                     (the reference's CPU code)

A step = one pass of the hot path over the whole batch: every read compressed
by press_gpu_batch_press, then decompressed by press_gpu_batch_depress.  500 k
reads are 113 GB of raw signal, more than fits next to their output in 180 GB, so
the batch is streamed through the GPU in device-resident chunks of whole reads
(identical kernels, identical streams); the chunk's samples are generated on the
device (untimed) and only the press / depress calls sit between the CUDA events.

The ONE JSON line carries
  * the headline keys: svb16-zd, `--reads` reads PER GPU (weak scaling: rank k owns reads
    [k*reads, (k+1)*reads) of the batch), with `roofline`, `cpu_baseline`, `e2e`;
  * `methods`: the other north_star methods (ex_zd, shuff_vbe21_zd, svb_zd, vbz_svb16_zd) on
    `--method-reads` reads per GPU, each with value / compress / decompress / bytes per
    sample / roofline / cpu_baseline / e2e;
  * `scaling_strong`: ONE `--reads`-read batch cut over the N GPUs by sample-balanced read
    ranges (honours_b200.shard.partition_reads), all-gather padded to the largest shard.
The only cross-GPU step is the allgather of the per-read byte positions
(press_gpu_allgather_offsets: ncclAllGather + container index, inside the C ABI).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# data/reads.blow5.test -- 113 471 398 938 raw bytes, press + depress seconds, 1 thread Xeon Silver 4114
PUBLISHED_ROUNDTRIP_GBS = {
    "svb16_zd": 113.471398938 / (293.05 + 356.11),
    "svb_zd": 113.471398938 / (274.50 + 271.24),
    "vbe21_zd": 113.471398938 / (506.07 + 415.31),
    "shuff_vbe21_zd": 113.471398938 / (3337.78 + 3753.19),
}
CHUNK_RAW_BYTES = 30e9
DEFAULT_METHODS = "ex_zd,shuff_vbe21_zd,svb_zd,vbz_svb16_zd"
HUFFMAN_TABLE = os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--method", default="svb16_zd")
    ap.add_argument("--reads", type=int, default=500_000, help="reads per GPU per step (weak) / of the whole batch (strong)")
    ap.add_argument("--p-ex", type=float, default=None, help="exception rate (default NA12878's 4.28e-5)")
    ap.add_argument("--methods", default=DEFAULT_METHODS, help="further methods reported under `methods` ('' = none)")
    ap.add_argument("--method-reads", type=int, default=125_000, help="reads per GPU of each `methods` leg")
    ap.add_argument("--method-steps", type=int, default=3)
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (N > 1)")
    ap.add_argument("--strong-steps", type=int, default=3)
    ap.add_argument("--e2e-gb", type=float, default=2.0, help="raw GB of the host-buffer (e2e) sample")
    ap.add_argument("--cpu-reads", type=int, default=4000, help="reads of the CPU-baseline sample")
    ap.add_argument("--cpu-seconds", type=float, default=8.0)
    ap.add_argument("--len-cap", type=int, default=0, help="diagnostic: clip read lengths (0 = off; not the BASELINE workload)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-sidecar", action="store_true", help="Huffman methods: decode without the encoder's side-car")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# CPU legs (the only places bench.py touches oracle/; they load nothing of the product)

def cpu_leg(method: str, nreads: int, p_ex: float, seconds: float, steps: int | None, warmup: int):
    """Time the reference's own CPU code (oracle/_ref when it was compiled, else
    the oracle port) on the first `nreads` reads of the same synthetic batch,
    all host threads, per-read bound/press/depress/compare like press/test.c."""
    from benchdata import workload
    from oracle import cpu_impl
    kind = "ref" if cpu_impl.available("ref") else "oracle"
    if kind == "oracle" and not cpu_impl.available("oracle"):
        cpu_impl.build(ref=False)
    impl = cpu_impl.load(kind)
    if method.startswith("shuff_"):
        impl.set_huffman_table(open(HUFFMAN_TABLE, "rb").read())
    lengths, x0, shift = workload.read_table(nreads)
    offsets = workload.offsets_of(lengths)
    samples = workload.synth_host(offsets, x0, shift, p_ex=p_ex)
    cores = os.cpu_count() or 1
    raw = 2.0 * samples.size
    tp = td = 0.0
    nbytes = 0
    done = 0
    for _ in range(warmup):
        impl.run_batch(method, samples, offsets, threads=cores)
    t_start = time.time()
    while True:
        r = impl.run_batch(method, samples, offsets, threads=cores)
        if r["rc"] != 0:
            raise RuntimeError(f"CPU {kind} failed: {r}")
        tp += r["press_s"]
        td += r["depress_s"]
        nbytes = r["bytes"]
        done += 1
        if steps is not None:
            if done >= steps:
                break
        elif time.time() - t_start >= seconds:
            break
    return {
        "kind": "reference" if kind == "ref" else "port",
        "cores": cores,
        "passes": done,
        "raw_bytes": raw,
        "press_s": tp / done,
        "depress_s": td / done,
        "roundtrip_gbs": raw / ((tp + td) / done) / 1e9,
        "compress_gbs": raw / (tp / done) / 1e9,
        "decompress_gbs": raw / (td / done) / 1e9,
        "bytes_per_sample": nbytes / samples.size,
        "sample": f"first {nreads} reads of the batch ({samples.size} samples, {raw / 1e9:.2f} GB raw), "
                  f"{done} passes, {cores} threads over disjoint read ranges",
    }


def cpu_baseline_entry(r: dict) -> dict:
    return {"value": r["roundtrip_gbs"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"],
            "sample": r["sample"], "compress_gbs": r["compress_gbs"],
            "decompress_gbs": r["decompress_gbs"], "bytes_per_sample": r["bytes_per_sample"]}


def workload_config(args, p_ex):
    """Identical in both arms: it names the workload, not how an arm samples it."""
    from benchdata import workload
    lengths, _, _ = workload.read_table(args.reads)
    if args.len_cap:
        lengths = np.minimum(lengths, args.len_cap)
    total = int(lengths.sum())
    return {
        "workload": f"{args.method} compress+decompress, {args.reads} synthetic NA12878-like reads per GPU "
                    "(lognormal lengths, median 80304 / mean 113471 samples, zd histogram of "
                    "press/gen_huffman.c:11)",
        "method": args.method, "reads_per_gpu": args.reads, "p_ex": p_ex,
        "samples_per_gpu": total, "chunks_per_step": max(1, math.ceil(2.0 * total / CHUNK_RAW_BYTES)),
        "l2": "inputs larger than L2 (each chunk is tens of GB)",
        "sharding": "by read; weak scaling headline (every GPU its own --reads reads) and a strong-scaling leg "
                    "(one batch cut by sample-balanced read ranges); one ncclAllGather of per-read byte positions",
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from benchdata import workload
    p_ex = workload.P_EX if args.p_ex is None else args.p_ex
    r = cpu_leg(args.method, args.cpu_reads, p_ex, args.cpu_seconds, args.steps, args.warmup)
    ms = (r["press_s"] + r["depress_s"]) * 1e3
    line = {
        "impl": "reference",
        "metric": f"{args.method} compress+decompress GB/s of raw int16 signal",
        "value": r["roundtrip_gbs"], "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16", "data": "synthetic",
        "config": workload_config(args, p_ex),
        "compress_gbs": r["compress_gbs"], "decompress_gbs": r["decompress_gbs"],
        "bytes_per_sample": r["bytes_per_sample"],
        "cpu_baseline": {"value": r["roundtrip_gbs"], "unit": "GB/s", "cores": r["cores"],
                         "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": r["roundtrip_gbs"], "unit": "GB/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    # the other north_star methods, same bounded-sample procedure (short)
    methods = {}
    for m in [x for x in args.methods.split(",") if x and x != args.method]:
        rr = cpu_leg(m, max(200, args.cpu_reads // 4), p_ex, 3.0, None, 1)
        methods[m] = cpu_baseline_entry(rr)
    if methods:
        line["methods"] = methods
    _emit(line)


# ---------------------------------------------------------------------------
# clocks

class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_uuid: str | None):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        cmd = ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"]
        if gpu_uuid:
            cmd.insert(1, f"--id={gpu_uuid}")
        try:
            self.p = subprocess.Popen(cmd, stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.f:
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        hot = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm
        return {"sm_mhz": float(np.median(hot)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ---------------------------------------------------------------------------
# our arm

class Runner:
    """One rank's view of the job: device, communicator, and the legs it can run."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from honours_b200 import press
        from benchdata import workload
        self.torch, self.dist, self.press, self.workload = torch, dist, press, workload
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torchrun --nproc-per-node {args.gpus}")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        press.init(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
            # the product's own communicator (ncclAllGather inside the C ABI): rank 0 makes the id,
            # torch.distributed only carries its 128 bytes to the other ranks
            uid = [press.comm_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            press.comm_init(uid[0], self.rank, self.world)
        self.p_ex = workload.P_EX if args.p_ex is None else args.p_ex
        self.stream = torch.cuda.current_stream().cuda_stream
        self.huff_ready = False
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        self.peak = float(peaks.get("hbm_gbs", 6650.0))
        self.peak_src = ("MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks
                         else "fallback 6650 GB/s (B200_PROFILING.md)")

    # -- helpers ---------------------------------------------------------------
    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, vals, op):
        t = self.torch.tensor(vals, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return t.tolist()

    def need_table(self, method):
        if method.startswith("shuff_") and not self.huff_ready:
            self.press.set_huffman_table()
            self.huff_ready = True

    def make_chunks(self, lengths, x0, shift, first_read, pad_reads=None):
        """Cut this rank's reads [first_read, first_read + len(lengths)) into device-resident chunks of
        whole reads (about CHUNK_RAW_BYTES of raw signal each).  pad_reads: per-chunk read counts the
        all-gather is padded to (the same on every rank)."""
        torch = self.torch
        total = int(lengths.sum())
        nchunks = max(1, math.ceil(2.0 * total / CHUNK_RAW_BYTES))
        if pad_reads is not None:
            nchunks = len(pad_reads)
        cum = np.concatenate([[0], np.cumsum(lengths)])
        cuts = [0] + [int(np.searchsorted(cum, total * c / nchunks)) for c in range(1, nchunks)] + [lengths.size]
        chunks = []
        for c in range(nchunks):
            r0, r1 = cuts[c], cuts[c + 1]
            off = self.workload.offsets_of(lengths[r0:r1])
            chunks.append({
                "r0": r0, "nreads": r1 - r0, "ns": int(off[-1]), "first": first_read + r0,
                "d_off": torch.from_numpy(off.astype(np.int64)).to(self.dev),
                "d_x0": torch.from_numpy(np.ascontiguousarray(x0[r0:r1])).to(self.dev),
                "d_sh": torch.from_numpy(np.ascontiguousarray(shift[r0:r1])).to(self.dev),
            })
        return chunks

    def device_leg(self, method, chunks, K, W, pad_reads, clocks=False):
        """K timed passes over `chunks` (this rank's reads) -> dict of this JOB's numbers (max over
        ranks of the device time, sums of the work).  pad_reads[c] = read count of chunk c the
        all-gather uses on every rank (>= this rank's)."""
        torch, press, workload = self.torch, self.press, self.workload
        args, dev, stream, world = self.args, self.dev, self.stream, self.world
        self.need_table(method)
        p_ex = self.p_ex
        max_ns = max(c["ns"] for c in chunks)
        max_nr = max(pad_reads)
        cap = max(press.batch_bound(method, c["nreads"], c["ns"]) for c in chunks)
        if 4 <= press.METHODS[method] <= 10:
            # batch_bound is "every sample an exception" (6 B/sample); the container and the
            # exception pool are sized for this workload's exception rate instead (with slack) --
            # an overflow of either is reported by batch_status, never silent
            cap = min(cap, max(int((1.05 + 8.0 * p_ex) * c["ns"]) + 1100 * c["nreads"] + 4096 for c in chunks))
            press.set_exception_capacity(max(1.0 / 64, 1.3 * p_ex))
        sb = max(press.batch_scratch_bytes(method, c["nreads"], c["ns"]) for c in chunks)
        sidecar = press.sidecar_bytes(method, max(c["nreads"] for c in chunks), max_ns) if not args.no_sidecar else 0
        d_s = torch.empty(max_ns + 64, dtype=torch.int16, device=dev)
        d_back = torch.empty(max_ns + 64, dtype=torch.int16, device=dev)
        d_out = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
        d_oo = torch.empty(max_nr + 1, dtype=torch.int64, device=dev)
        d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
        d_side = torch.empty(sidecar + 64, dtype=torch.uint8, device=dev) if sidecar else None
        d_gath = torch.empty(world * (max_nr + 1), dtype=torch.int64, device=dev) if world > 1 else None
        d_sync = torch.zeros(1, dtype=torch.int32, device=dev) if world > 1 else None
        d_glob = torch.empty(world * max_nr + 1, dtype=torch.int64, device=dev) if world > 1 else None
        bad_log = []
        total_samples = sum(c["ns"] for c in chunks)

        def one_step():
            evs = []
            for ci, c in enumerate(chunks):
                workload.synth_device(d_s.data_ptr(), c["d_off"].data_ptr(), c["nreads"], c["d_x0"].data_ptr(),
                                      c["d_sh"].data_ptr(), p_ex=p_ex, first_read=c["first"], stream=stream)
                e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                if world > 1:
                    # the generator above is not part of the path and its duration differs per shard:
                    # line the ranks up on the stream before the timed region, so that the allgather
                    # below is charged with the exchange and the encoders' own spread, not with that
                    self.dist.all_reduce(d_sync)
                e0.record()
                press.batch_press(method, d_s.data_ptr(), c["d_off"].data_ptr(), c["nreads"], c["ns"],
                                  d_out.data_ptr(), cap, d_oo.data_ptr(), d_scr.data_ptr(), sb, stream,
                                  sidecar=d_side.data_ptr() if d_side is not None else 0, sidecar_bytes=sidecar)
                if world > 1:
                    # the path's one exchange, through the C ABI on the same stream right behind the
                    # encode: shards with fewer reads than the padded count repeat their total
                    npad = pad_reads[ci]
                    if npad > c["nreads"]:
                        d_oo[c["nreads"] + 1: npad + 1] = d_oo[c["nreads"]]
                    press.allgather_offsets(d_oo.data_ptr(), npad, d_gath.data_ptr(), d_glob.data_ptr(), stream)
                e1.record()
                # in_bytes: the container's true size is only known on the device; the
                # decode kernels bound their reads by it, so pass the capacity here
                press.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), cap, c["d_off"].data_ptr(),
                                    c["nreads"], c["ns"], d_back.data_ptr(), d_scr.data_ptr(), sb, stream,
                                    sidecar=d_side.data_ptr() if d_side is not None else 0, sidecar_bytes=sidecar)
                e2.record()
                evs.append((e0, e1, e2))
                if not args.no_verify:
                    # round trip on ALL samples (press/test.c:1804 only checks the first half)
                    bad = torch.zeros(1, dtype=torch.int64, device=dev)
                    step = 1 << 28
                    for i in range(0, c["ns"], step):
                        j = min(c["ns"], i + step)
                        bad.add_((d_s[i:j] != d_back[i:j]).sum())
                    bad_log.append(((len(bad_log), c["first"], c["ns"]), bad))
                c["comp_t"] = d_oo[c["nreads"]: c["nreads"] + 1].clone()
            torch.cuda.synchronize()
            tp = sum(a.elapsed_time(b) for a, b, _ in evs)
            td = sum(b.elapsed_time(c_) for _, b, c_ in evs)
            comp = sum(int(c["comp_t"].item()) for c in chunks)
            st = press.batch_status(d_scr.data_ptr(), stream)
            if st != 0:
                raise RuntimeError(f"{method}: device status 0x{st:x}: {press.last_error()}")
            return tp, td, comp

        for _ in range(W):
            one_step()
        press.profile_enable(True)
        press.profile_collect()
        sampler = None
        if clocks and self.rank == 0:
            uuid = None
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(self.local).uuid)
            except Exception:
                pass
            sampler = ClockSampler(uuid)
        self.barrier()
        l0 = press.launch_count()
        t_wall = time.time()
        tp = td = 0.0
        comp = 0
        for _ in range(K):
            a, b, comp = one_step()
            tp += a
            td += b
        self.barrier()
        t_wall = time.time() - t_wall
        launches = press.launch_count() - l0
        clk = sampler.stop() if sampler else None
        kms, kn = press.profile_collect()
        press.profile_enable(False)
        wrong = [(tag, int(b.item())) for tag, b in bad_log if int(b.item()) != 0]
        if wrong:
            raise RuntimeError(f"{method}: round trip mismatch (call index, first read, samples) -> bad samples: {wrong[:8]}")
        del d_s, d_back, d_out, d_scr, d_side
        torch.cuda.empty_cache()

        t_all, t_p, t_d = (x / K for x in self.reduce([tp + td, tp, td], "MAX"))      # ms per step, max over ranks
        job_samples, job_comp, job_launches = self.reduce([float(total_samples), float(comp), float(launches)], "SUM")
        raw = 2.0 * job_samples
        # roofline of the dominant kernel sequence (this rank's launches, CUDA events in the library)
        alg_bytes_step = 2.0 * total_samples + comp       # per pass over this rank's reads, press or depress
        kern = {}
        for i, nm in enumerate(("encode", "decode")):
            if kn[i]:
                ms_launch = kms[i] / kn[i]
                per_launch = alg_bytes_step * K / kn[i]
                kern[nm] = {"ms_per_launch": ms_launch, "launches": kn[i], "alg_bytes_per_launch": per_launch,
                            "achieved_gbs": per_launch / (ms_launch * 1e-3) / 1e9,
                            "frac": per_launch / (ms_launch * 1e-3) / 1e9 / self.peak}
        dom = min(kern, key=lambda k: kern[k]["achieved_gbs"]) if kern else None
        roofline = None
        if dom:
            roofline = {"bound": "hbm", "kernel": f"{method} {dom}", "achieved": kern[dom]["achieved_gbs"],
                        "peak": self.peak, "unit": "GB/s", "frac": kern[dom]["achieved_gbs"] / self.peak,
                        "traffic": None, "peak_source": self.peak_src,
                        "alg_bytes_per_launch": kern[dom]["alg_bytes_per_launch"],
                        "ms_per_launch": kern[dom]["ms_per_launch"]}
            # DRAM bytes of that kernel from the committed ncu --set full capture (profiles/traffic.json,
            # taken on a shorter run of this same command), scaled to this run's launch by the
            # algorithmic bytes of the two launches
            tr = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tr):
                try:
                    ent = json.load(open(tr)).get(f"{method}_{dom}")
                    if ent:
                        roofline["traffic"] = ent["dram_bytes"] * kern[dom]["alg_bytes_per_launch"] / ent["alg_bytes"]
                        roofline["traffic_source"] = ent.get("source")
                except Exception:
                    pass
        return {"value": raw / (t_all * 1e-3) / 1e9, "ms_per_step": t_all,
                "compress_gbs": raw / (t_p * 1e-3) / 1e9, "decompress_gbs": raw / (t_d * 1e-3) / 1e9,
                "bytes_per_sample": job_comp / job_samples, "raw_bytes_per_step": raw,
                "samples_per_step": job_samples, "wall_s_timed_region": t_wall, "roofline": roofline,
                "kernels": kern, "gpu_launches": int(job_launches), "clocks": clk, "steps": K, "warmup": W,
                "sidecar_bytes_per_sample": (sidecar / max_ns) if sidecar else 0.0}

    # -- e2e: host buffers through the C ABI, copies inside the timed region -------------------
    def e2e_leg(self, method, lengths, x0, shift, first_read, gb, K, W, link=None):
        """Same metric through press_gpu_batch_press_host / _depress_host: pinned host buffers in,
        pinned host buffers out, every copy inside the timed region.  Two figures: `serial` (a step =
        compress, then decompress, one after the other) and the headline `value` (duplex: one host
        thread compresses step k + 1's batch while another decompresses step k's container, so the
        H2D-heavy and the D2H-heavy half share the full-duplex link; K compressions + K
        decompressions inside one wall-clock region)."""
        torch, press, workload = self.torch, self.press, self.workload
        self.need_table(method)
        L = press.lib()
        m = press.METHODS[method]
        want = int(gb * 1e9 / 2)
        cum = np.cumsum(lengths)
        nreads = int(min(lengths.size, max(1, np.searchsorted(cum, want) + 1)))
        off = workload.offsets_of(lengths[:nreads])
        ns = int(off[-1])
        cap = press.batch_bound(method, nreads, ns)
        if 4 <= press.METHODS[method] <= 10:
            cap = min(cap, int((1.05 + 8.0 * self.p_ex) * ns) + 1100 * nreads + 4096)
        h_s = press.pinned_array(2 * ns + 64)
        h_back = press.pinned_array(2 * ns + 64)
        h_out = [press.pinned_array(cap + 64) for _ in range(3)]
        h_oo = [press.pinned_array(8 * (nreads + 1)) for _ in range(3)]
        h_off = press.pinned_array(8 * (nreads + 1))
        h_off.view(np.uint64)[:] = off
        # the sample is generated on the device (same bits as the host generator) and parked in pinned memory
        d_off = torch.from_numpy(off.astype(np.int64)).to(self.dev)
        d_x0 = torch.from_numpy(np.ascontiguousarray(x0[:nreads])).to(self.dev)
        d_sh = torch.from_numpy(np.ascontiguousarray(shift[:nreads])).to(self.dev)
        d_tmp = torch.empty(ns + 64, dtype=torch.int16, device=self.dev)
        workload.synth_device(d_tmp.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(),
                              p_ex=self.p_ex, first_read=first_read, stream=self.stream)
        torch.cuda.synchronize()
        h_s.view(np.int16)[:ns] = d_tmp[:ns].cpu().numpy()
        del d_tmp
        torch.cuda.empty_cache()

        def do_press(b):
            rc = L.press_gpu_batch_press_host(m, h_s.ctypes.data, h_off.ctypes.data, nreads, h_out[b].ctypes.data,
                                              cap, h_oo[b].ctypes.data)
            if rc != 0:
                raise RuntimeError("e2e press: " + press.last_error())

        def do_depress(b):
            rc = L.press_gpu_batch_depress_host(m, h_out[b].ctypes.data, h_oo[b].ctypes.data, h_off.ctypes.data,
                                                nreads, h_back.ctypes.data)
            if rc != 0:
                raise RuntimeError("e2e depress: " + press.last_error())

        def serial_step():
            t0 = time.perf_counter()
            do_press(0)
            t1 = time.perf_counter()
            do_depress(0)
            return t1 - t0, time.perf_counter() - t1

        for _ in range(max(1, min(W, 2))):
            serial_step()
        if not np.array_equal(h_back.view(np.int16)[:ns], h_s.view(np.int16)[:ns]):
            raise RuntimeError("e2e round trip mismatch")
        self.barrier()
        tp = td = 0.0
        for _ in range(K):
            a, b = serial_step()
            tp += a
            td += b
        comp = int(h_oo[0].view(np.uint64)[nreads])

        # duplex: two free-running host threads -- one compresses step 0, 1, 2 ... into a ring of NB
        # containers, the other decompresses container k as soon as it is complete; no barrier between
        # steps, so both link directions stay busy from the first container to the last
        NB = len(h_out)
        errs = []
        ready = [threading.Semaphore(0) for _ in range(NB)]     # container b holds a finished step
        free = [threading.Semaphore(1) for _ in range(NB)]      # container b may be overwritten

        def presser():
            try:
                for k in range(K):
                    free[k % NB].acquire()
                    do_press(k % NB)
                    ready[k % NB].release()
            except Exception as e:      # noqa: BLE001 -- re-raised below
                errs.append(e)
                for s in ready:
                    s.release()

        def depresser():
            try:
                for k in range(K):
                    ready[k % NB].acquire()
                    if errs:
                        return
                    do_depress(k % NB)
                    free[k % NB].release()
            except Exception as e:      # noqa: BLE001
                errs.append(e)
                for s in free:
                    s.release()

        h_back.view(np.int16)[:ns] = 0
        self.barrier()
        t0 = time.perf_counter()
        th = [threading.Thread(target=presser), threading.Thread(target=depresser)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        t_dup = time.perf_counter() - t0
        if errs:
            raise errs[0]
        if not np.array_equal(h_back.view(np.int16)[:ns], h_s.view(np.int16)[:ns]):
            raise RuntimeError("e2e (duplex) round trip mismatch")

        t_ser, t_p, t_d, t_dup = (x / K for x in self.reduce([tp + td, tp, td, t_dup], "MAX"))
        raw = self.reduce([2.0 * ns], "SUM")[0]
        idx = 8 * (nreads + 1)
        for a in [h_s, h_back, h_off] + h_out + h_oo:
            press.pinned_free(a)
        out = {
            "value": raw / t_dup / 1e9, "unit": "GB/s",
            "h2d_bytes_per_step": 2 * ns + comp + 3 * idx, "d2h_bytes_per_step": comp + idx + 2 * ns,
            "mode": "duplex: one host thread compresses step after step while another decompresses every container "
                    "as soon as it is complete (two pipelines in the library, both link directions busy); "
                    "K compressions + K decompressions in one wall-clock region",
            "serial_value": raw / t_ser / 1e9, "serial_compress_gbs": raw / t_p / 1e9,
            "serial_decompress_gbs": raw / t_d / 1e9,
            "sample": f"first {nreads} reads of each rank's shard ({ns} samples, {2 * ns / 1e9:.2f} GB raw) "
                      "from pinned host memory through press_gpu_batch_press_host + _depress_host, "
                      "host wall clock, max over ranks",
        }
        if link:
            # per direction a round trip moves 2 + C bytes per sample; both directions run at once
            bps = comp / ns
            dirbw = min(link["h2d_duplex_gbs"], link["d2h_duplex_gbs"])
            out["link_ceiling_gbs"] = self.world * 2.0 * dirbw / (2.0 + bps)
            out["frac_of_link_ceiling"] = out["value"] / out["link_ceiling_gbs"]
            out["link"] = link
        return out

    def link_leg(self):
        """pinned cudaMemcpyAsync bandwidth of every rank's GPU, all ranks at once (min over ranks)"""
        self.barrier()
        r = self.workload.link_probe()
        keys = ("h2d_gbs", "d2h_gbs", "h2d_duplex_gbs", "d2h_duplex_gbs")
        mins = self.reduce([r[k] for k in keys], "MIN")
        out = dict(zip(keys, mins))
        out["note"] = (f"plain pinned cudaMemcpyAsync, {r['chunk_bytes'] >> 20} MiB per copy, {r['bytes'] >> 20} MiB x "
                       f"{r['reps']}, all {self.world} rank(s) at once, min over ranks, GB/s per GPU per direction")
        return out


def run_ours(args):
    R = Runner(args)
    press, workload, rank, world = R.press, R.workload, R.rank, R.world
    method = args.method
    if method not in press.implemented_methods():
        raise SystemExit(f"method {method} is not implemented in this build of libpress_gpu (no CPU fallback)")
    K, W = args.steps, args.warmup

    # ---- headline: weak scaling, this rank's own `--reads` reads of the batch ----------------
    first_read = rank * args.reads
    lengths, x0, shift = workload.read_table(args.reads, first_read=first_read)
    if args.len_cap:
        lengths = np.minimum(lengths, args.len_cap)
    total_samples = int(lengths.sum())
    nchunks = max(1, math.ceil(2.0 * total_samples / CHUNK_RAW_BYTES))
    if world > 1:
        # every rank uses the same number of chunks; the all-gather is padded to the largest read count
        nchunks = int(R.reduce([float(nchunks)], "MAX")[0])
    chunks = R.make_chunks(lengths, x0, shift, first_read, pad_reads=[0] * nchunks)
    pad = [int(x) for x in R.reduce([float(c["nreads"]) for c in chunks], "MAX")]
    head = R.device_leg(method, chunks, K, W, pad, clocks=True)
    del chunks

    # ---- strong scaling: ONE batch of `--reads` reads cut into sample-balanced read ranges ---
    strong = None
    if world > 1 and not args.no_strong:
        from honours_b200 import shard
        all_len, all_x0, all_sh = workload.read_table(args.reads)
        if args.len_cap:
            all_len = np.minimum(all_len, args.len_cap)
        r0, r1 = shard.partition_reads(all_len, world)[rank]
        my_total = int(all_len[r0:r1].sum())
        nch = int(R.reduce([float(max(1, math.ceil(2.0 * my_total / CHUNK_RAW_BYTES)))], "MAX")[0])
        sch = R.make_chunks(all_len[r0:r1], all_x0[r0:r1], all_sh[r0:r1], r0, pad_reads=[0] * nch)
        spad = [int(x) for x in R.reduce([float(c["nreads"]) for c in sch], "MAX")]
        s = R.device_leg(method, sch, min(K, args.strong_steps), min(W, 3), spad)
        shard_samples = R.reduce([float(my_total)], "MAX")[0]
        strong = {"value": s["value"], "unit": "GB/s", "ms_per_step": s["ms_per_step"],
                  "compress_gbs": s["compress_gbs"], "decompress_gbs": s["decompress_gbs"],
                  "reads_total": args.reads, "samples_total": s["samples_per_step"],
                  "largest_shard_samples": shard_samples,
                  "imbalance": shard_samples * world / s["samples_per_step"],
                  "efficiency_vs_weak_per_gpu": s["value"] / head["value"],
                  "steps": s["steps"], "warmup": s["warmup"],
                  "sharding": "honours_b200.shard.partition_reads: contiguous read ranges cut where the prefix sum "
                              "of the lengths crosses k/N of the total; all-gather padded to the largest shard's "
                              "read count; efficiency_vs_weak_per_gpu = value / (this run's weak headline value), "
                              "i.e. against N GPUs each at its own full-size batch rate"}
        del sch

    # ---- the other north_star methods, one chunk per GPU --------------------------------------
    others = {}
    names = [m for m in args.methods.split(",") if m and m != method]
    for m in names:
        if m not in press.implemented_methods():
            others[m] = {"unavailable": "not implemented in this build"}
            continue
        mfirst = rank * args.method_reads
        ml, mx, msh = workload.read_table(args.method_reads, first_read=mfirst)
        if args.len_cap:
            ml = np.minimum(ml, args.len_cap)
        mch = R.make_chunks(ml, mx, msh, mfirst, pad_reads=[0])
        mpad = [int(x) for x in R.reduce([float(c["nreads"]) for c in mch], "MAX")]
        r = R.device_leg(m, mch, min(K, args.method_steps), min(W, 3), mpad)
        pub = PUBLISHED_ROUNDTRIP_GBS.get(m)
        others[m] = {"value": r["value"], "unit": "GB/s", "compress_gbs": r["compress_gbs"],
                     "decompress_gbs": r["decompress_gbs"], "bytes_per_sample": r["bytes_per_sample"],
                     "ms_per_step": r["ms_per_step"], "reads_per_gpu": args.method_reads,
                     "samples_per_step": r["samples_per_step"], "steps": r["steps"], "warmup": r["warmup"],
                     "roofline": r["roofline"], "kernels": r["kernels"], "gpu_launches": r["gpu_launches"],
                     "sidecar_bytes_per_sample": r["sidecar_bytes_per_sample"],
                     "vs_baseline": (r["value"] / pub) if pub else None}
        del mch

    # ---- e2e: host buffers through the C ABI -----------------------------------------------------
    e2e = None
    if not args.no_e2e:
        link = R.link_leg()
        e2e = R.e2e_leg(method, lengths, x0, shift, first_read, args.e2e_gb, K, W, link=link)
        for m in names:
            if "value" in others.get(m, {}):
                ml, mx, msh = workload.read_table(min(args.method_reads, 20000), first_read=rank * args.method_reads)
                others[m]["e2e"] = R.e2e_leg(m, ml, mx, msh, rank * args.method_reads, min(args.e2e_gb, 1.0),
                                             min(K, 3), 1, link=link)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_entry(cpu_leg(method, args.cpu_reads, R.p_ex, args.cpu_seconds, None, 1))
        for m in names:
            if "value" in others.get(m, {}):
                others[m]["cpu_baseline"] = cpu_baseline_entry(
                    cpu_leg(m, max(200, args.cpu_reads // 4), R.p_ex, 3.0, None, 1))
                if "e2e" in others[m]:
                    others[m]["e2e"]["vs_cpu_baseline"] = others[m]["e2e"]["value"] / others[m]["cpu_baseline"]["value"]

    if rank == 0:
        pub = PUBLISHED_ROUNDTRIP_GBS.get(method)
        cfg = workload_config(args, R.p_ex)
        launches = head["gpu_launches"] + sum(o.get("gpu_launches", 0) for o in others.values())
        line = {
            "metric": f"{method} compress+decompress GB/s of raw int16 signal",
            "value": head["value"], "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (head["value"] / pub) if pub else None, "dtype": "int16", "data": "synthetic",
            "config": cfg,
            "published_baseline": "data/reads.blow5.test (1 thread Xeon Silver 4114)" if pub else None,
            "compress_gbs": head["compress_gbs"], "decompress_gbs": head["decompress_gbs"],
            "bytes_per_sample": head["bytes_per_sample"],
            "raw_bytes_per_step": head["raw_bytes_per_step"], "wall_s_timed_region": head["wall_s_timed_region"],
            "roofline": head["roofline"], "kernels": head["kernels"],
            "roofline_vs_nominal_8TBs": (head["roofline"]["achieved"] / 8000.0) if head["roofline"] else None,
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": head["clocks"],
            "methods": others, "scaling_strong": strong,
        }
        _emit(line)
    if world > 1:
        R.dist.barrier()
        press.comm_destroy()
        R.dist.destroy_process_group()


def _emit(line: dict) -> None:
    """The contract's ONE line on the process's real stdout (see __main__)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1

if __name__ == "__main__":
    a = parse_args()
    # libraries print to fd 1 behind Python's back (NCCL's "NCCL version ..." banner at the first
    # communicator): keep the real stdout for the JSON line, send everything else to stderr
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
