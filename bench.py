#!/usr/bin/env python
"""bench.py -- Mpoints/s encode+decode of the learned point-cloud codec hot path on B200.

Step      one pass of the hot path over one synthetic frame per rank: encode (g_a, h_a, quantise, h_s,
          CDF indexes) + decode (coords down, dequantise, h_s, offsets, g_s with 3x generative up-sampling /
          occupancy head / top-k / prune, colour head).  The host range coder, G-PCC side channel and file I/O
          stay outside (north-star); symbols cross as int32 buffers.
Workload  BASELINE.json configs[1]: CVPR_inverse_scaling, full 10-bit longdress-shaped frame (~818 k voxels),
          random-init weights (torch.manual_seed(0)), q = (0.5, 0.5), one 1024-block.
value     whole-job Mpoints/s with the frame already resident in HBM; e2e: the same through host buffers
          (pinned H2D of the cloud, D2H+H2D of the symbol buffers, D2H of the reconstruction) per step.
N > 1     frames are independent: rank r codes frame seed r, no data-path collective ("weak" scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload frame10|block128]
"""
import argparse
import gc
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Mpoints/s encode+decode at 10-bit"
UNIT = "Mpoints/s"


def peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written per pool) or the profiling recipe's fallback.
    bf16_burst: cuBLAS timed alone at full clocks -- the denominator for kernels timed one by one with CUDA events (the roofline
    pass below) when the clock record shows no power cap; bf16_sustained: the same GEMM back to back under the 1 kW cap."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "fallback"}


def model_config():
    from tests.helpers import model_config as mc
    return mc(width=128)


def make_frame(workload, seed):
    from oracle import synth
    if workload == "block128":
        return synth.block128(seed=seed)
    if workload == "frame11":
        return synth.frame11(seed=seed)
    return synth.frame10(seed=seed, jitter=8 if seed else 0)


def cpu_sample(pts, rgb, workload):
    """The CPU arm codes the SAME unit as the GPU arm (one whole frame / block per step; ~7 s for the 10-bit frame on 16 cores)."""
    return pts, rgb, {"frame10": "the whole 10-bit frame", "frame11": "the whole 11-bit frame", "block128": "the whole 128^3 block"}[workload] + \
        f" ({len(pts)} voxels), encode + decode, one frame per step"


# ----------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, indexes):
        """Rank 0 samples the clocks of every GPU of the job.  nvidia-smi's start-up takes driver-wide locks (seen as a
        +20 ms first step on the other ranks when every rank started one after the barrier): wait for its first row here,
        before the barrier that opens the timed region."""
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", ",".join(str(i) for i in indexes), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.perf_counter()
            while not self.rows and time.perf_counter() - t0 < 3.0:
                time.sleep(0.01)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        """Timed region starts here: rows sampled before are dropped."""
        self.first = len(self.rows)

    def stop(self):
        self.rows = self.rows[getattr(self, "first", 0):] or self.rows
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------- CPU arm
def run_cpu(workload, steps, warmup, seed=0):
    """The reference's CPU path (oracle/pipeline_ref.py: ME-CPU algorithm in C++ + OpenBLAS, torch-CPU entropy ops)."""
    import numpy as np
    import torch
    from oracle import native, pipeline_ref
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    torch.manual_seed(0)
    cfg = model_config()
    net = model_mod.UnifiedModel(cfg)          # parameter container only (random init); arithmetic below is the oracle's
    net.update()
    codec = pipeline_ref.RefCodec(net.state_dict(), cfg)
    cores = native.set_threads(os.cpu_count() or 1)
    torch.set_num_threads(cores)
    pts, rgb = make_frame(workload, seed)
    pts, rgb, what = cpu_sample(pts, rgb, workload)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        enc = codec.encode_block(pts, rgb, [0.5, 0.5])
        codec.decode_block(enc["y_points"], enc["z_symbols"], enc["y_symbols"], enc["k"], [0.5, 0.5])
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return {"value": len(pts) / (ms * 1e3), "unit": UNIT, "cores": cores, "kind": "port", "sample": what,
            "ms_per_step": ms, "points": int(len(pts))}


# ----------------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--workload", default="frame10", choices=["frame10", "block128", "frame11"])
    ap.add_argument("--conv-mode", default="auto", choices=["auto", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--in-flight", type=int, default=2, help="frames in flight per GPU (host threads x CUDA streams); 1 = single stream")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything libraries print there (e.g. NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": f"CVPR_inverse_scaling {args.workload}: {max(1, args.in_flight)} synthetic surface-sampled frame(s) per rank per step, "
                          "compress+decompress, random-init weights, q=(0.5,0.5)",
              "sharding": f"{world} rank(s), independent frames on each, no collective"}

    if args.impl == "reference":
        if rank != 0:
            return
        w = max(1, min(args.warmup, 1))
        r = run_cpu(args.workload, max(1, min(args.steps, 5)), w)     # a 10-bit frame is ~7 s of CPU work: at most 5 timed steps
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": w, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        line["steps"] = max(1, min(args.steps, 5))
        line["config"]["workload"] = f"CVPR_inverse_scaling {args.workload}: 1 synthetic surface-sampled frame per step, compress+decompress, " \
                                     "random-init weights, q=(0.5,0.5) (CPU arm: one frame at a time on all host cores)"
        emit(line)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback for the graft)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as ge
    pkg = ge.build()
    so = pkg.sparse_ops
    so.set_conv_mode(args.conv_mode)
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    torch.manual_seed(0)
    net = model_mod.UnifiedModel(model_config()).to(dev).eval()
    net.update()
    q = torch.tensor([[0.5, 0.5]], device=dev)

    depth = max(1, args.in_flight)
    # a step codes `depth` frames (one per worker stream): frame seeds rank, rank + world, ...
    frames = [make_frame(args.workload, seed=rank + j * world) for j in range(depth)]
    n_points = sum(len(p) for p, _ in frames)
    host_clouds = [torch.from_numpy(np.concatenate([p.astype(np.float32), c], axis=1)).pin_memory() for p, c in frames]
    dev_clouds = [h.to(dev) for h in host_clouds]
    codec_mod = importlib.import_module(pkg.__name__ + ".codec")
    pipe = codec_mod.FramePipeline(dev, depth, model=net) if depth > 1 else None
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def step_resident(j=0):
        dev_cloud = dev_clouds[j]
        enc = net.encode_block(dev_cloud, q)
        ys = enc["y_symbols"]
        x_hat = net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], q)
        return net.finalize(x_hat.C, x_hat.F)

    io_frame = {}                                        # frame slot -> (h2d bytes, d2h bytes) of one step

    pinned = {}

    def to_host(name, t, slot=0):
        """D2H into a persistent pinned buffer (allocated on first use), asynchronous on the current stream."""
        t = t.contiguous()
        name = (slot, name)                              # one set of buffers per frame slot (a slot is coded by one worker at a time)
        buf = pinned.get(name)
        if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
            buf = pinned[name] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        buf.copy_(t, non_blocking=True)
        return buf

    def step_e2e(j=0):
        x = host_clouds[j].to(dev, non_blocking=True)
        enc = net.encode_block(x, q)
        host = {k: to_host(k, enc[k], j) for k in ("y_points", "z_symbols", "y_symbols", "y_indexes")}
        torch.cuda.current_stream().synchronize()
        # (host range coder runs here in a deployment; excluded by the metric)
        yp, zs, ys = (host[k].to(dev, non_blocking=True) for k in ("y_points", "z_symbols", "y_symbols"))
        x_hat = net.decode_block(yp, zs, lambda idx: ys, enc["k"], q)
        rec = to_host("rec", net.finalize(x_hat.C, x_hat.F), j)
        torch.cuda.current_stream().synchronize()
        io_frame[j] = (x.numel() * 4 + sum(host[k].numel() * host[k].element_size() for k in ("y_points", "z_symbols", "y_symbols")),
                       sum(v.numel() * v.element_size() for v in host.values()) + rec.numel() * rec.element_size())
        return rec

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, use_pipe=True):
        total = 0.0
        per_step = []
        host_ms = []
        for _ in range(steps):
            flush.fill_(1)                                   # evict L2 between timed iterations
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            t0 = time.perf_counter()
            if pipe is None or not use_pipe:
                fn()
            else:
                pipe.map(fn, range(depth), after=e0)             # `depth` frames, one per worker stream; this stream waits for all
            host_ms.append(1e3 * (time.perf_counter() - t0))     # host enqueue time (includes the step's own syncs)
            e1.record()
            flush[:1].fill_(2)                               # 1-block marker launch: end of the step in an ncu launch list
            torch.cuda.synchronize()
            per_step.append(e0.elapsed_time(e1))
            total += per_step[-1]
        print(f"[bench] {fn.__name__} per-step ms: " + " ".join(f"{t:.2f}" for t in per_step) +
              " | host enqueue ms: " + " ".join(f"{t:.2f}" for t in host_ms), file=sys.stderr)
        return total

    with torch.no_grad():
        for _ in range(max(args.warmup, 3)):
            for j in range(depth):
                step_resident(j)
        barrier()
        if pipe is not None:                                     # warm the worker streams too (scratch, allocator pools)
            for _ in range(2):
                pipe.map(step_resident, range(depth))
            barrier()
        # everything allocated so far (modules, torch, numpy) is long-lived: take it out of the collector's generations so that
        # a full collection cannot stall a timed step for ~100 ms (seen as a one-off 130 ms e2e step at N=2)
        gc.collect()
        gc.freeze()
        clocks = ClockSampler(range(local_rank, local_rank + world)) if rank == 0 else None
        barrier()
        if clocks is not None:
            clocks.mark()
        l0 = pkg._lib.launch_count()
        t_res = timed(step_resident, args.steps)
        launches = (pkg._lib.launch_count() - l0) // max(args.steps, 1)
        barrier()
        clk = clocks.stop() if clocks is not None else None
        for _ in range(3):
            for j in range(depth):
                step_e2e(j)
        if pipe is not None:
            pipe.map(step_e2e, range(depth))
        barrier()
        t_e2e = timed(step_e2e, args.steps)
        # the same frame (slot 0) alone on one stream, for reference next to the frames-in-flight numbers
        t_res1 = timed(step_resident, min(args.steps, 5), use_pipe=False) / min(args.steps, 5) if pipe is not None else None
        t_e2e1 = timed(step_e2e, min(args.steps, 5), use_pipe=False) / min(args.steps, 5) if pipe is not None else None
        barrier()
        # roofline pass: the same step with per-launch CUDA events around every convolution kernel
        so.profiler = so.ConvProfiler()
        step_resident()
        prof = so.profiler.summary()
        so.profiler = None

    def reduce_max(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(v):
        if world == 1:
            return v
        t = torch.tensor([float(v)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    ms_res, ms_e2e = reduce_max(t_res) / args.steps, reduce_max(t_e2e) / args.steps
    total_points = reduce_sum(n_points)
    if rank != 0:
        if pipe is not None:
            pipe.close()
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    # tensor-core kernels: k_conv_tc (output-stationary gather-GEMM) + k_gemm_first (dense half of the generative transposed conv)
    zero = {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "launches": 0, "by_shape": {}}
    tcs = [prof.get(k, zero) for k in ("conv_tc", "gemm_first")]
    tc_ms, tc_flops, tc_launches = (sum(t[k] for t in tcs) for k in ("ms", "flops", "launches"))
    all_conv_ms = sum(v["ms"] for k, v in prof.items() if k in ("conv_tc", "gemm_first", "gent_reduce", "conv_narrow", "conv_f32", "head_gather"))
    achieved = tc_flops / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    # the roofline pass times kernels one by one: burst peak unless the clock record of the timed region shows the power cap
    capped = bool(clk and "sw_power_cap" in (clk.get("reasons") or []))
    peak_kind = "bf16_tflops_sustained (sw_power_cap seen)" if capped else "bf16_tflops (burst: kernels timed one by one, no power cap seen)"
    bf16_peak = pk["bf16_sustained"] if capped else pk["bf16_burst"]

    def shapes(kind):
        d = prof.get(kind, zero)["by_shape"]
        return {"x".join(map(str, k)): {"launches": v[0], "ms": round(v[1], 3), "tflops": round(v[2] / (v[1] * 1e-3) / 1e12, 2) if v[1] > 0 else 0}
                for k, v in sorted(d.items(), key=lambda kv: -kv[1][1])}

    def hbm(kind):
        d = prof.get(kind, zero)
        gbs = d["bytes"] / (d["ms"] * 1e-3) / 1e9 if d["ms"] > 0 else 0.0
        return {"launches": d["launches"], "ms": round(d["ms"], 3), "algorithmic_gb": round(d["bytes"] / 1e9, 3), "achieved_gbs": round(gbs, 1),
                "frac_of_hbm_peak": round(gbs / pk["hbm_gbs"], 4),
                "ms_by_shape": {"x".join(map(str, k)): round(v[1], 3) for k, v in sorted(d["by_shape"].items(), key=lambda kv: -kv[1][1])}}

    traffic = None                      # DRAM bytes of the same launches from the committed ncu --set full capture
    tp = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(tp) and args.workload == "frame10":
        traffic = json.load(open(tp))["tcgen05_conv_dram_bytes_per_step"]
    roofline = {"bound": "tensor", "kernel": "k_conv_tc_mt / k_conv_tc + k_gemm_first (tcgen05 sparse conv kernels, all launches of one step)",
                "achieved": achieved, "peak": bf16_peak, "unit": "TFLOP/s", "frac": achieved / bf16_peak,
                "peak_kind": peak_kind, "frac_of_burst_peak": achieved / pk["bf16_burst"], "frac_of_sustained_peak": achieved / pk["bf16_sustained"],
                "traffic": traffic, "traffic_unit": "DRAM bytes per step over the same launches (ncu --set full; source named in profiles/r01_traffic.json)",
                "peak_source": pk["src"], "launches": tc_launches, "kernel_ms_per_step": tc_ms,
                "all_conv_ms_per_step": all_conv_ms, "share_of_step": tc_ms * depth / ms_res if ms_res else None,
                "per": "one frame (the roofline pass codes a single frame on one stream)",
                "algorithmic_gflop_per_step": tc_flops / 1e9,
                "by_shape": {"conv_tc": shapes("conv_tc"), "gemm_first": shapes("gemm_first")},
                "hbm_bound_kernels": {"gent_reduce (k_cells_reduce)": hbm("gent_reduce"), "conv_narrow (occupancy-head layers, k_conv_tc_mt<1, 32|64|128>)": hbm("conv_narrow"),
                                      "gemm_first (k_gemm_first, T write)": hbm("gemm_first"),
                                      "head_gather (k_gather_sum: second head layer, GEMM-first)": hbm("head_gather"),
                                      # north-star pieces 1 and 5: coordinate hash, kernel maps, compaction, entropy model (CUDA events per C-ABI call)
                                      "hash_build (k_hash_insert)": hbm("hash_build"),
                                      "stride_coords (k_insert_strided + compaction)": hbm("stride_coords"),
                                      "generate_coords (k_insert_generated + compaction)": hbm("generate_coords"),
                                      "kernel_map (k_kernel_map)": hbm("kernel_map"),
                                      "cells_kernel_map3 (k_cells_kernel_map3)": hbm("cells_kernel_map3"),
                                      "topk (k_topk_hist radix select + k_topk_above)": hbm("topk"),
                                      "gather_rows (prune: k_gather_rows)": hbm("gather_rows"),
                                      "features_at (k_features_at)": hbm("features_at"),
                                      "entropy (k_eb_* / k_gc_* / k_quantize / k_dequantize)": hbm("entropy")},
                "hbm_peak_gbs": pk["hbm_gbs"]}
    line = {"metric": METRIC, "value": total_points / (ms_res * 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_res, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.conv_mode == "auto" else "f32", "data": "synthetic",
            "config": dict(config, points_per_step=n_points, frames_per_step=depth,
                           in_flight=f"{depth} frame(s) in flight per GPU: {depth} host thread(s) x CUDA stream(s) (codec.FramePipeline)",
                           l2="512 MiB flush buffer written between timed steps", conv_mode=args.conv_mode),
            "clocks": clk,
            "e2e": {"value": total_points / (ms_e2e * 1e3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": sum(v[0] for v in io_frame.values()), "d2h_bytes_per_step": sum(v[1] for v in io_frame.values())},
            "gpu_launches": int(launches), "roofline": roofline}
    if t_res1 is not None:
        n1 = len(frames[0][0])
        line["single_stream"] = {"note": "one frame at a time on one stream, rank 0", "ms_per_frame": t_res1, "value": n1 / (t_res1 * 1e3),
                                 "e2e_ms_per_frame": t_e2e1, "e2e_value": n1 / (t_e2e1 * 1e3), "unit": UNIT}
    if world == 1 and not args.no_cpu_baseline:
        r = run_cpu(args.workload, 1, 0)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    emit(line)
    if pipe is not None:
        pipe.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
