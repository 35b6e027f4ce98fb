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
Multi-GPU is weak scaling: every rank processes its own `--reads` reads of the
batch (rank k owns reads [k*reads, (k+1)*reads)); the only cross-GPU step is the
NCCL allgather of the per-read byte positions that gives one container index.
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
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# data/reads.blow5.test:7 -- svb16-zd on 113 471 398 938 raw bytes: 293.05 s press + 356.11 s depress
PUBLISHED_ROUNDTRIP_GBS = {
    "svb16_zd": 113.471398938 / (293.05 + 356.11),
    "svb_zd": 113.471398938 / (274.50 + 271.24),
    "vbe21_zd": 113.471398938 / (506.07 + 415.31),
    "shuff_vbe21_zd": 113.471398938 / (3337.78 + 3753.19),
}
CHUNK_RAW_BYTES = 30e9


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--method", default="svb16_zd")
    ap.add_argument("--reads", type=int, default=500_000, help="reads per GPU per step")
    ap.add_argument("--p-ex", type=float, default=None, help="exception rate (default NA12878's 4.28e-5)")
    ap.add_argument("--e2e-gb", type=float, default=2.0, help="raw GB of the host-buffer (e2e) sample")
    ap.add_argument("--cpu-reads", type=int, default=4000, help="reads of the CPU-baseline sample")
    ap.add_argument("--cpu-seconds", type=float, default=8.0)
    ap.add_argument("--len-cap", type=int, default=0, help="diagnostic: clip read lengths (0 = off; not the BASELINE workload)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# CPU legs (the only places bench.py touches oracle/)

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
        impl.set_huffman_table(open(os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman"), "rb").read())
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


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    p_ex = args.p_ex
    from benchdata import workload
    if p_ex is None:
        p_ex = workload.P_EX
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
    _emit(line)


def workload_config(args, p_ex):
    return {
        "workload": f"{args.method} compress+decompress, {args.reads} synthetic NA12878-like reads per GPU "
                    "(lognormal lengths, median 80304 / mean 113471 samples, zd histogram of "
                    "press/gen_huffman.c:11)",
        "method": args.method, "reads_per_gpu": args.reads, "p_ex": p_ex,
        "l2": "inputs larger than L2 (each chunk is tens of GB)",
        "sharding": "by read, weak scaling, NCCL allgather of per-read byte positions (ranks aligned by a stream-level all_reduce before each timed chunk)",
    }


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

def run_ours(args):
    import torch
    import torch.distributed as dist
    from honours_b200 import press
    from benchdata import workload

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    press.init(local)
    method = args.method
    if method not in press.implemented_methods():
        raise SystemExit(f"method {method} is not implemented in this build of libpress_gpu (no CPU fallback)")
    if method.startswith("shuff_"):
        press.set_huffman_table()
    p_ex = workload.P_EX if args.p_ex is None else args.p_ex
    K, W = args.steps, args.warmup
    stream = torch.cuda.current_stream().cuda_stream

    # ---- this rank's reads, cut into device-resident chunks of whole reads ----
    first_read = rank * args.reads
    lengths, x0, shift = workload.read_table(args.reads, first_read=first_read)
    if args.len_cap:
        lengths = np.minimum(lengths, args.len_cap)
    total_samples = int(lengths.sum())
    nchunks = max(1, math.ceil(2.0 * total_samples / CHUNK_RAW_BYTES))
    cum = np.concatenate([[0], np.cumsum(lengths)])
    cuts = [0]
    for c in range(1, nchunks):
        cuts.append(int(np.searchsorted(cum, total_samples * c / nchunks)))
    cuts.append(args.reads)
    if world > 1:   # allgather wants the same read count per chunk on every rank
        per = math.ceil(args.reads / nchunks)
        cuts = [min(args.reads, c * per) for c in range(nchunks + 1)]
    chunks = []
    for c in range(nchunks):
        r0, r1 = cuts[c], cuts[c + 1]
        off = workload.offsets_of(lengths[r0:r1])
        chunks.append({
            "r0": r0, "nreads": r1 - r0, "ns": int(off[-1]),
            "d_off": torch.from_numpy(off.astype(np.int64)).to(dev),
            "d_x0": torch.from_numpy(x0[r0:r1]).to(dev),
            "d_sh": torch.from_numpy(shift[r0:r1]).to(dev),
        })
    max_ns = max(c["ns"] for c in chunks)
    max_nr = max(c["nreads"] for c in chunks)
    cap = max(press.batch_bound(method, c["nreads"], c["ns"]) for c in chunks)
    if press.METHODS[method] >= 4:
        # batch_bound is "every sample an exception" (6 B/sample); the container and the
        # exception pool are sized for this workload's exception rate instead (with slack) --
        # an overflow of either is reported by batch_status, never silent
        cap = min(cap, max(int((1.05 + 8.0 * p_ex) * c["ns"]) + 1100 * c["nreads"] + 4096 for c in chunks))
        press.set_exception_capacity(max(1.0 / 64, 1.3 * p_ex))
    sb = max(press.batch_scratch_bytes(method, c["nreads"], c["ns"]) for c in chunks)
    d_s = torch.empty(max_ns + 64, dtype=torch.int16, device=dev)
    d_back = torch.empty(max_ns + 64, dtype=torch.int16, device=dev)
    d_out = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(max_nr + 1, dtype=torch.int64, device=dev)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    d_gath = torch.empty(world * (max_nr + 1), dtype=torch.int64, device=dev) if world > 1 else None
    d_sync = torch.zeros(1, dtype=torch.int32, device=dev) if world > 1 else None
    d_glob = torch.empty(world * max_nr + 1, dtype=torch.int64, device=dev) if world > 1 else None
    bad_log = []

    def verify(ns, tag):
        # round trip on ALL samples (press/test.c:1804 only checks the first half)
        bad = torch.zeros(1, dtype=torch.int64, device=dev)
        step = 1 << 28
        for i in range(0, ns, step):
            j = min(ns, i + step)
            bad.add_((d_s[i:j] != d_back[i:j]).sum())
        bad_log.append((tag, bad))

    def one_step(timed: bool):
        """-> (press ms, depress ms, compressed bytes) of this rank for one pass over its reads"""
        evs = []
        comp = 0
        for c in chunks:
            workload.synth_device(d_s.data_ptr(), c["d_off"].data_ptr(), c["nreads"], c["d_x0"].data_ptr(),
                                  c["d_sh"].data_ptr(), p_ex=p_ex, first_read=first_read + c["r0"],
                                  stream=stream)
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            if world > 1:
                # the generator above is not part of the path and its duration differs per shard:
                # line the ranks up on the stream before the timed region, so that the allgather
                # below is charged with the exchange and the encoders' own spread, not with that
                dist.all_reduce(d_sync)
            e0.record()
            press.batch_press(method, d_s.data_ptr(), c["d_off"].data_ptr(), c["nreads"], c["ns"],
                              d_out.data_ptr(), cap, d_oo.data_ptr(), d_scr.data_ptr(), sb, stream)
            if world > 1:
                # the path's one exchange: allgather of the per-read positions + global_offsets ->
                # container index of the whole job, on the same stream right behind the encode
                # (tried beside the local decode on a side stream: NCCL's kernel then waits for
                # SMs until the decode drains and the step gets longer, 1918 vs 1980 GB/s)
                g = d_gath[: world * (c["nreads"] + 1)]
                dist.all_gather_into_tensor(g, d_oo[: c["nreads"] + 1])
                press.global_offsets(g.data_ptr(), world, c["nreads"], d_glob.data_ptr(), stream)
            e1.record()
            # in_bytes: the container's true size is only known on the device; the
            # decode kernels bound their reads by it, so pass the capacity here
            press.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), cap, c["d_off"].data_ptr(),
                                c["nreads"], c["ns"], d_back.data_ptr(), d_scr.data_ptr(), sb, stream)
            e2.record()
            evs.append((e0, e1, e2))
            if not args.no_verify:
                verify(c["ns"], (len(bad_log), c["r0"], c["ns"]))
            comp_t = d_oo[c["nreads"]: c["nreads"] + 1].clone()
            c["comp_t"] = comp_t
        torch.cuda.synchronize()
        tp = sum(a.elapsed_time(b) for a, b, _ in evs)
        td = sum(b.elapsed_time(c_) for _, b, c_ in evs)
        comp = sum(int(c["comp_t"].item()) for c in chunks)
        st = press.batch_status(d_scr.data_ptr(), stream)
        if st != 0:
            raise RuntimeError(f"device status 0x{st:x}: {press.last_error()}")
        return tp, td, comp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        one_step(False)
    press.profile_enable(True)
    press.profile_collect()
    uuid = None
    try:
        uuid = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        pass
    sampler = ClockSampler(uuid) if rank == 0 else None
    barrier()
    l0 = press.launch_count()
    t_wall = time.time()
    tp = td = 0.0
    comp = 0
    for _ in range(K):
        a, b, comp = one_step(True)
        tp += a
        td += b
    barrier()
    t_wall = time.time() - t_wall
    launches = press.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    kms, kn = press.profile_collect()
    press.profile_enable(False)
    wrong = [(tag, int(b.item())) for tag, b in bad_log if int(b.item()) != 0]
    if wrong:
        raise RuntimeError(f"round trip mismatch (call index, first read, samples) -> bad samples: {wrong[:8]}")

    t = torch.tensor([tp + td, tp, td], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(total_samples), float(comp), float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    t_all, t_p, t_d = (x / K for x in t.tolist())          # ms per step, max over ranks
    job_samples, job_comp, job_launches = tot.tolist()
    raw = 2.0 * job_samples
    value = raw / (t_all * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (this rank's launches, CUDA events in the library) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    alg_bytes_step = 2.0 * total_samples + comp       # per pass over this rank's reads, press or depress
    kern = {}
    for i, nm in enumerate(("encode", "decode")):
        if kn[i]:
            ms_launch = kms[i] / kn[i]
            per_launch = alg_bytes_step * K / kn[i]
            kern[nm] = {"ms_per_launch": ms_launch, "launches": kn[i],
                        "alg_bytes_per_launch": per_launch,
                        "achieved_gbs": per_launch / (ms_launch * 1e-3) / 1e9}
    dom = max(kern, key=lambda k: kern[k]["ms_per_launch"]) if kern else None
    roofline = None
    if dom:
        roofline = {"bound": "hbm", "kernel": f"{method} {dom}", "achieved": kern[dom]["achieved_gbs"],
                    "peak": peak, "unit": "GB/s", "frac": kern[dom]["achieved_gbs"] / peak,
                    "traffic": None, "peak_source": peak_src,
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

    # ---- e2e: host buffers through the C ABI, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        del d_s, d_back, d_out, d_scr
        torch.cuda.empty_cache()
        e2e = run_e2e(args, press, workload, dev, first_read, lengths, x0, shift, p_ex, world)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_leg(method, args.cpu_reads, p_ex, args.cpu_seconds, None, 1)
        cpu = {"value": r["roundtrip_gbs"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"],
               "sample": r["sample"], "compress_gbs": r["compress_gbs"],
               "decompress_gbs": r["decompress_gbs"], "bytes_per_sample": r["bytes_per_sample"]}

    if rank == 0:
        pub = PUBLISHED_ROUNDTRIP_GBS.get(method)
        cfg = workload_config(args, p_ex)
        cfg.update({"chunks_per_step": nchunks, "samples_per_gpu": total_samples,
                    "published_baseline": "data/reads.blow5.test (1 thread Xeon Silver 4114)" if pub else None})
        line = {
            "metric": f"{method} compress+decompress GB/s of raw int16 signal",
            "value": value, "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": t_all, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / pub) if pub else None, "dtype": "int16", "data": "synthetic",
            "config": cfg,
            "compress_gbs": raw / (t_p * 1e-3) / 1e9, "decompress_gbs": raw / (t_d * 1e-3) / 1e9,
            "bytes_per_sample": job_comp / job_samples,
            "raw_bytes_per_step": raw, "wall_s_timed_region": t_wall,
            "roofline": roofline, "kernels": kern,
            "roofline_vs_nominal_8TBs": (roofline["achieved"] / 8000.0) if roofline else None,
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(job_launches), "clocks": clocks,
        }
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_e2e(args, press, workload, dev, first_read, lengths, x0, shift, p_ex, world):
    """Same metric through press_gpu_batch_press_host / _depress_host: pinned host
    buffers in, pinned host buffers out, every copy inside the timed region."""
    import torch
    import torch.distributed as dist
    L = press.lib()
    m = press.METHODS[args.method]
    want = int(args.e2e_gb * 1e9 / 2)
    cum = np.cumsum(lengths)
    nreads = int(min(lengths.size, max(1, np.searchsorted(cum, want) + 1)))
    off = workload.offsets_of(lengths[:nreads])
    ns = int(off[-1])
    cap = press.batch_bound(args.method, nreads, ns)
    h_s = press.pinned_array(2 * ns + 64)
    h_back = press.pinned_array(2 * ns + 64)
    h_out = press.pinned_array(cap + 64)
    h_oo = press.pinned_array(8 * (nreads + 1))
    h_off = press.pinned_array(8 * (nreads + 1))
    h_off.view(np.uint64)[:] = off
    # the sample is generated on the device (same bits as the host generator) and parked in pinned memory
    d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
    d_x0 = torch.from_numpy(x0[:nreads]).to(dev)
    d_sh = torch.from_numpy(shift[:nreads]).to(dev)
    d_tmp = torch.empty(ns + 64, dtype=torch.int16, device=dev)
    workload.synth_device(d_tmp.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(),
                          p_ex=p_ex, first_read=first_read, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    h_s.view(np.int16)[:ns] = d_tmp[:ns].cpu().numpy()
    del d_tmp
    torch.cuda.empty_cache()

    def step():
        t0 = time.perf_counter()
        rc = L.press_gpu_batch_press_host(m, h_s.ctypes.data, h_off.ctypes.data, nreads, h_out.ctypes.data,
                                          cap, h_oo.ctypes.data)
        if rc != 0:
            raise RuntimeError("e2e press: " + press.last_error())
        t1 = time.perf_counter()
        rc = L.press_gpu_batch_depress_host(m, h_out.ctypes.data, h_oo.ctypes.data, h_off.ctypes.data,
                                            nreads, h_back.ctypes.data)
        if rc != 0:
            raise RuntimeError("e2e depress: " + press.last_error())
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    for _ in range(max(1, min(args.warmup, 2))):
        step()
    if not np.array_equal(h_back.view(np.int16)[:ns], h_s.view(np.int16)[:ns]):
        raise RuntimeError("e2e round trip mismatch")
    if world > 1:
        dist.barrier()
    tp = td = 0.0
    K = args.steps
    for _ in range(K):
        a, b = step()
        tp += a
        td += b
    comp = int(h_oo.view(np.uint64)[nreads])
    t = torch.tensor([tp + td, tp, td], dtype=torch.float64, device=dev)
    tot = torch.tensor([2.0 * ns], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    t_all, t_p, t_d = (x / K for x in t.tolist())
    raw = tot.item()
    idx = 16 * (nreads + 1)
    for a in (h_s, h_back, h_out, h_oo, h_off):
        press.pinned_free(a)
    return {
        "value": raw / t_all / 1e9, "unit": "GB/s",
        "h2d_bytes_per_step": 2 * ns + comp + idx + 8 * (nreads + 1),
        "d2h_bytes_per_step": comp + 8 * (nreads + 1) + 2 * ns,
        "compress_gbs": raw / t_p / 1e9, "decompress_gbs": raw / t_d / 1e9,
        "sample": f"first {nreads} reads of each rank's shard ({ns} samples, {2 * ns / 1e9:.2f} GB raw) "
                  "from pinned host memory through press_gpu_batch_press_host + _depress_host, "
                  "host wall clock, max over ranks",
    }


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
