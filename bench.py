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
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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
    ap.add_argument("--workload", default="frame10", choices=["frame10", "block128", "frame11", "stream11", "train256"],
                    help="frame10: BASELINE configs[1] (default, the metric's config); block128: configs[0]; stream11: configs[2], 64-frame 11-bit "
                         "stream sharded by frame (strong scaling); train256: configs[3], the training step on 256^3 blocks, batch 8 per GPU")
    ap.add_argument("--frames", type=int, default=64, help="stream11: frames in the sequence")
    ap.add_argument("--frames-per-step", type=int, default=0, help="frame10 / block128 / frame11: distinct frames a step codes through the "
                    "in-flight pipeline (default: 4 x --in-flight, or 1 for a single stream)")
    ap.add_argument("--serial-blocks", action="store_true", help="stream11: code the blocks of a frame one after the other (the reference's "
                    "loop, model.py:137-176) instead of as one batched sparse tensor")
    ap.add_argument("--topk-from", default="decoder", choices=["decoder", "encoder"],
                    help="encoder: prune every decoder level with the TRUE occupancy (what a trained model's top-k selects) instead of the "
                         "random-init logits, which keep a scattered set and make levels 2/3 ~1.5x heavier than any deployment sees")
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
    fps = args.frames_per_step or (4 * max(1, args.in_flight) if args.in_flight > 1 else 1)
    config = {"workload": f"CVPR_inverse_scaling {args.workload}: {fps} synthetic surface-sampled frame(s) per rank per step "
                          f"({max(1, args.in_flight)} in flight), compress+decompress, random-init weights, q=(0.5,0.5)",
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
    if world > 1 and os.environ.get("PCG_PIN", "1") == "1":
        # one slice of the host cores per rank: every rank runs a launching thread per frame in flight, and 8 ranks x 3 threads
        # migrating over 16 cores cost ~6 % of the 8-GPU throughput in round 1 (host enqueue is 80 % of a step's wall time)
        # (whole PHYSICAL cores: hyper-thread siblings stay together, so no two ranks share a core's execution units)
        cpus = sorted(os.sched_getaffinity(0))
        groups = {}
        for c in cpus:
            try:
                sib = open(f"/sys/devices/system/cpu/cpu{c}/topology/thread_siblings_list").read().strip()
            except OSError:
                sib = str(c)
            groups.setdefault(sib, []).append(c)
        phys = [g for _, g in sorted(groups.items(), key=lambda kv: min(kv[1]))]
        if len(phys) >= world:
            mine = [c for g in phys[local_rank::world] for c in g]
        else:
            per = max(1, len(cpus) // world)
            mine = cpus[local_rank * per:(local_rank + 1) * per] or cpus
        os.sched_setaffinity(0, set(mine))
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
    if args.workload in ("stream11", "train256"):
        ctx = argparse.Namespace(args=args, rank=rank, world=world, dev=dev, pkg=pkg, net=net, q=q, emit=emit, config=config)
        (run_stream if args.workload == "stream11" else run_train)(ctx)
        if world > 1:
            dist.destroy_process_group()
        return
    net.g_s.topk_from_encoder = args.topk_from == "encoder"

    depth = max(1, args.in_flight)
    # a step codes F distinct frames through the pipeline of `depth` worker streams: two generated frames (seeds rank, rank + world)
    # and integer translations of them (different coordinates, hence different hash tables / maps / symbols, at no generation cost).
    # With F = 4 x depth the frames of a step desynchronise like a real sequence does (one stream in its HBM-bound kernels while the
    # other is in its tensor-core kernels) instead of starting in lockstep every step, and the step's inputs (F x 20 MB) exceed L2.
    F = args.frames_per_step or (4 * depth if depth > 1 else 1)
    bases = [make_frame(args.workload, seed=rank + j * world) for j in range(min(2, F))]
    shift_rng = np.random.default_rng(1234 + rank)
    frames = []
    for j in range(F):
        p, c = bases[j % len(bases)]
        d = shift_rng.integers(0, 9, size=3) if j >= len(bases) else np.zeros(3, dtype=np.int64)
        frames.append(((p + d).astype(p.dtype), c))
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
        x_hat = net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], q, keep_sets=enc.get("keep_sets"))
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
        x_hat = net.decode_block(yp, zs, lambda idx: ys, enc["k"], q, keep_sets=enc.get("keep_sets"))
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
                pipe.map(fn, range(F), after=e0)                 # F frames through `depth` worker streams; this stream waits for all
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
        for w in range(max(args.warmup, 3)):
            for j in range(F if w == 0 else min(F, 2)):
                step_resident(j)
        barrier()
        if pipe is not None:                                     # warm the worker streams too (scratch, allocator pools)
            for _ in range(2):
                pipe.map(step_resident, range(F))
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
        for w in range(3):
            for j in range(F if w == 0 else min(F, 2)):
                step_e2e(j)
        if pipe is not None:
            pipe.map(step_e2e, range(F))
        barrier()
        t_e2e = timed(step_e2e, args.steps)
        # the same frame (slot 0) alone on one stream, for reference next to the frames-in-flight numbers
        t_res1 = timed(step_resident, min(args.steps, 5), use_pipe=False) / min(args.steps, 5) if pipe is not None else None
        t_e2e1 = timed(step_e2e, min(args.steps, 5), use_pipe=False) / min(args.steps, 5) if pipe is not None else None
        barrier()
        # compress() / decompress() as the reference defines them (model/model.py:94-250, component form): the host range coder and
        # its device<->host copies INSIDE the timed region, one frame at a time (reported beside the metric, which excludes them)
        hc_ms = None
        if rank == 0:
            reps = []
            for i in range(4):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                bits, shapes, ks, coords_, qv = net.compress(dev_clouds[0], q, block_size=1024)
                rec_ = net.decompress(coordinates=coords_, strings=bits, shape=shapes, k=ks, q_vals=qv)
                torch.cuda.synchronize()
                if i:
                    reps.append(1e3 * (time.perf_counter() - t0))
            hc_ms = sum(reps) / len(reps)
            hc_bytes = sum(len(s_[0]) for blk in bits for s_ in blk)
        # roofline pass: the same step with per-launch CUDA events around every convolution kernel
        so.profiler = so.ConvProfiler()
        step_resident()
        prof = so.profiler.summary()
        impl_bytes = dict(so.profiler.impl_bytes)
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
        """CUDA-event time of a kernel group and, where the group has a defined input and output, its ALGORITHMIC bytes
        (inputs read once + outputs written once; no intermediates, no neighbour tables) as GB/s."""
        d = prof.get(kind, zero)
        out = {"launches": d["launches"], "ms": round(d["ms"], 3),
               "ms_by_shape": {"x".join(map(str, k)): round(v[1], 3) for k, v in sorted(d["by_shape"].items(), key=lambda kv: -kv[1][1])}}
        if d["bytes"] > 0 and d["ms"] > 0:
            gbs = d["bytes"] / (d["ms"] * 1e-3) / 1e9
            out.update(algorithmic_gb=round(d["bytes"] / 1e9, 3), achieved_gbs=round(gbs, 1), frac_of_hbm_peak=round(gbs / pk["hbm_gbs"], 4))
        return out

    # DRAM traffic of the same launches: written by tools/traffic_from_ncu.py from an `ncu --set full` capture of THIS command
    # (profiles/r02_traffic.json names the capture); null when no capture of this workload is committed
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(tp) and args.workload == "frame10" and args.topk_from == "decoder":
        tj = json.load(open(tp))
        traffic, traffic_src = tj.get("tcgen05_conv_dram_bytes_per_step"), tj.get("source")
    roofline = {"bound": "tensor", "kernel": "k_conv_tc_mt / k_conv_tc + k_gemm_first (tcgen05 sparse conv kernels, all launches of one step)",
                "achieved": achieved, "peak": bf16_peak, "unit": "TFLOP/s", "frac": achieved / bf16_peak,
                "peak_kind": peak_kind, "frac_of_burst_peak": achieved / pk["bf16_burst"], "frac_of_sustained_peak": achieved / pk["bf16_sustained"],
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": pk["src"], "launches": tc_launches, "kernel_ms_per_step": tc_ms,
                "all_conv_ms_per_step": all_conv_ms, "share_of_step": tc_ms * F / ms_res if ms_res else None,
                "per": "one frame (the roofline pass codes a single frame on one stream)",
                "algorithmic_gflop_per_step": tc_flops / 1e9,
                "by_shape": {"conv_tc": shapes("conv_tc"), "gemm_first": shapes("gemm_first")},
                "intermediate_bytes_not_counted": {k: int(v) for k, v in impl_bytes.items()},
                "hbm_bound_kernels": {"gen_transpose (k_gemm_first + k_cells_reduce: parent rows in, child rows out)": hbm("gen_transpose"),
                                      "  of which k_gemm_first": hbm("gemm_first"), "  of which k_cells_reduce": hbm("gent_reduce"),
                                      "occupancy_head (first layer on tcgen05 + k_gather_sum: rows in, logits out)": hbm("occupancy_head"),
                                      "  of which k_gather_sum": hbm("head_gather"),
                                      "conv_narrow (unfused narrow head layers)": hbm("conv_narrow"),
                                      # north-star pieces 1 and 5: coordinate hash, kernel maps, compaction, entropy model (CUDA events per C-ABI call;
                                      # a kernel map's own product is its neighbour table, counted for that call only)
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
            "config": dict(config, points_per_step=n_points, frames_per_step=F,
                           in_flight=f"{depth} frame(s) in flight per GPU: {depth} host thread(s) x CUDA stream(s) (codec.FramePipeline)",
                           l2=f"512 MiB flush buffer written between timed steps; a step's inputs ({F} frames) are larger than L2 as well", conv_mode=args.conv_mode,
                           topk_from=args.topk_from + (" (measurement switch: every decoder level pruned with the true occupancy)" if args.topk_from == "encoder" else "")),
            "clocks": clk,
            "e2e": {"value": total_points / (ms_e2e * 1e3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": sum(v[0] for v in io_frame.values()), "d2h_bytes_per_step": sum(v[1] for v in io_frame.values())},
            "gpu_launches": int(launches), "roofline": roofline}
    if hc_ms is not None:
        n0 = len(frames[0][0])
        line["with_host_coder"] = {"note": "UnifiedModel.compress + decompress, component form, host rANS coder and its copies included, one frame at a time",
                                   "ms_per_frame": hc_ms, "value": n0 / (hc_ms * 1e3), "unit": UNIT, "payload_bytes": hc_bytes,
                                   "bits_per_point": 8.0 * hc_bytes / n0}
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


# ----------------------------------------------------------------------------------------- helpers shared by the extra workloads
def _reduce(ctx, v, op):
    import torch
    import torch.distributed as dist
    if ctx.world == 1:
        return float(v)
    t = torch.tensor([float(v)], device=ctx.dev, dtype=torch.float64)
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
    return float(t.item())


def _barrier(ctx):
    import torch
    import torch.distributed as dist
    if ctx.world > 1:
        dist.barrier()
    torch.cuda.synchronize()


# ----------------------------------------------------------------------------------------- configs[2]: the sharded 11-bit stream
def stream_frames(n_frames, n_base=4):
    """64-frame 11-bit Owlii-shaped sequence: `n_base` generated frames (oracle/synth.py::frame11, seeds 0..n_base-1,
    ~2.5 M voxels each; sampling 47 M surface points per frame on the host is ~15 s, hence not 64 of them) x integer translations
    of up to +-8 voxels.  Returns (loader(i) -> (N,6) float32 numpy, sizes)."""
    import numpy as np
    bases = [make_frame("frame11", seed=s) for s in range(min(n_base, n_frames))]
    rng = np.random.default_rng(11)
    shifts = rng.integers(-8, 9, size=(n_frames, 3))

    def load(i):
        p, c = bases[i % len(bases)]
        return np.concatenate([(p + shifts[i]).astype(np.float32), c], axis=1)
    return load, [len(bases[i % len(bases)][0]) for i in range(n_frames)]


def run_stream(ctx):
    """Strong scaling: the SAME n frames whatever the number of ranks; frames assigned largest-first (codec.lpt_schedule), each rank
    codes its frames through codec.StreamCodec at block_size 512 (evaluate.py:45: 6-8 blocks per 11-bit frame), two frames in
    flight, no collective on the data path; results (voxel counts + a checksum per frame) gathered on rank 0."""
    import numpy as np
    import torch
    args, dev, pkg, net, rank, world = ctx.args, ctx.dev, ctx.pkg, ctx.net, ctx.rank, ctx.world
    codec_mod = importlib.import_module(pkg.__name__ + ".codec")
    n_frames = args.frames
    load, sizes = stream_frames(n_frames)
    mine = codec_mod.my_units(sizes, world, rank)
    host = {i: torch.from_numpy(load(i)).pin_memory() for i in mine}
    resident = {i: h.to(dev) for i, h in host.items()}
    depth = max(1, args.in_flight)
    pipe = codec_mod.FramePipeline(dev, depth, model=net) if depth > 1 else None
    sc = codec_mod.StreamCodec(net, ctx.q, block_size=512, rank=rank, world=world, coder="bypass", batched=not args.serial_blocks)

    def checksum(rec):          # order-independent digest of a reconstruction: (voxels, sum of coordinates, sum of 8-bit colours)
        r = rec.double()
        return (int(rec.shape[0]), float(r[:, :3].sum()), float(torch.round(r[:, 3:] * 255).sum()))

    frame_ms = []

    def run(frames_of, keep):
        def timed_keep(rec):                      # host wall clock between consecutive frame completions of this rank (diagnostic)
            frame_ms.append(time.perf_counter())
            return keep(rec)
        return sc.run([(lambda i=i: frames_of(i)) for i in range(n_frames)], sizes=sizes, keep=timed_keep, pipeline=pipe)

    with torch.no_grad():
        # warm-up: packed weights, worker streams, and the caching allocator's pools -- one frame of every distinct size this rank
        # will see, largest first, on every worker stream (a first-seen larger frame costs cudaMalloc calls of GB-sized blocks:
        # 40-58 Mpoints/s from run to run before this)
        warm = sorted({sizes[i]: i for i in mine}.values(), key=lambda i: -sizes[i])
        for i in warm[:1]:
            sc.code_frame(resident[i])
        if pipe is not None:
            for _ in range(2):
                pipe.map(lambda i: sc.code_frame(resident[i]), [w for w in warm for _ in range(depth)])
        codec_mod.gather_results({-1 - rank: (0, 0.0, 0.0)}, world, rank)      # warm-up of the result gather (first object collective: ~1 s of set-up)
        _barrier(ctx)
        clocks = ClockSampler(range(world)) if rank == 0 else None
        _barrier(ctx)
        if clocks is not None:
            clocks.mark()
        l0 = pkg._lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = run(lambda i: resident[i], checksum)
        e1.record()
        torch.cuda.synchronize()
        ms_res = e0.elapsed_time(e1)
        launches = pkg._lib.launch_count() - l0
        stats = torch.cuda.memory_stats(dev)
        print(f"[bench] rank {rank}: cudaMalloc calls so far {stats.get('num_device_alloc', 0)}, frees {stats.get('num_device_free', 0)}, "
              f"reserved {stats.get('reserved_bytes.all.peak', 0) / 2**30:.1f} GiB", file=sys.stderr)
        _barrier(ctx)
        clk = clocks.stop() if clocks is not None else None
        # end to end: every frame from pinned host memory, reconstruction back to the host
        pinned_out = {}
        io = [0, 0]

        def e2e_frame(i):
            x = host[i].to(dev, non_blocking=True)
            rec = sc.code_frame(x)
            buf = pinned_out.get(i)
            if buf is None or buf.shape != rec.shape:
                buf = pinned_out[i] = torch.empty(rec.shape, dtype=rec.dtype, pin_memory=True)
            buf.copy_(rec, non_blocking=True)
            io[0] += x.numel() * 4
            io[1] += rec.numel() * rec.element_size()
            return rec
        _barrier(ctx)
        e0.record()
        if pipe is None:
            for i in mine:
                e2e_frame(i)
        else:
            pipe.map(e2e_frame, mine)
        e1.record()
        torch.cuda.synchronize()
        ms_e2e = e0.elapsed_time(e1)
    ms_res_max, ms_e2e_max = _reduce(ctx, ms_res, "MAX"), _reduce(ctx, ms_e2e, "MAX")
    ms_res_min = -_reduce(ctx, -ms_res, "MAX")
    h2d, d2h = _reduce(ctx, io[0], "SUM"), _reduce(ctx, io[1], "SUM")
    launches = _reduce(ctx, launches, "SUM")
    if pipe is not None:
        pipe.close()
    if rank != 0:
        return
    total = float(sum(sizes))
    loads = [sum(sizes[i] for i in codec_mod.my_units(sizes, world, r)) for r in range(world)]
    line = {"metric": "Mpoints/s encode+decode at 11-bit, 64-frame stream sharded by frame", "value": total / (ms_res_max * 1e3), "unit": UNIT,
            "n_gpus": world, "steps": 1, "warmup": 2, "ms_per_step": ms_res_max, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16" if args.conv_mode == "auto" else "f32", "data": "synthetic",
            "config": {"workload": f"CVPR_inverse_scaling stream11: {n_frames} synthetic 11-bit frames ({int(total)} voxels, 4 generated frames x integer "
                                   "translations), block_size 512 (" + ("blocks of a frame coded one after the other" if args.serial_blocks else "the 6-8 blocks of a frame coded as one batched sparse tensor") + "), compress+decompress through codec.StreamCodec, random-init weights, q=(0.5,0.5); "
                                   "a step = the whole sequence", "sharding": f"{world} rank(s), frames assigned largest-first (LPT), no collective",
                       "in_flight": depth, "l2": "every frame (40 MB of input, > 1 GB of intermediates) exceeds the 126 MB L2",
                       "load_imbalance": {"voxels_per_rank": loads, "max_over_mean": max(loads) / (sum(loads) / world),
                                          "rank_ms_max_over_min": ms_res_max / ms_res_min}},
            "clocks": clk, "e2e": {"value": total / (ms_e2e_max * 1e3), "unit": UNIT, "ms_per_step": ms_e2e_max, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(launches),
            "frames": {str(i): list(res[i]) for i in sorted(res)[:4]}, "frames_coded": len(res),
            "digest": float(sum(v[1] + v[2] for v in res.values())) if res else None}
    ctx.emit(line)



# ----------------------------------------------------------------------------------------- configs[3]: the training step
def training_batch(seed, n_blocks=8, block=256, min_points=300):
    """`n_blocks` 256^3 crops of the synthetic 10-bit cloud (configs/CVPR_inverse_scaling_256.yaml: block 256, min_points_train 300;
    BASELINE.json asks for batch 8 per GPU, the yaml has 4), coordinates local to the crop, one batch index per crop."""
    import numpy as np
    pts, rgb = make_frame("frame10", seed=seed)
    rng = np.random.default_rng(7 + seed)
    coords, feats = [], []
    while len(coords) < n_blocks:
        origin = pts[rng.integers(len(pts))] - rng.integers(32, block - 32, size=3)
        sel = np.all((pts >= origin) & (pts < origin + block), axis=1)
        if sel.sum() < min_points:
            continue
        b = len(coords)
        coords.append(np.concatenate([np.full((int(sel.sum()), 1), b, np.int32), (pts[sel] - origin).astype(np.int32)], axis=1))
        feats.append(rgb[sel].astype(np.float32))
    return np.concatenate(coords), np.concatenate(feats)


def run_train(ctx):
    """One optimiser step of /root/reference/train.py:195-234 per bench step: forward (train mode, tcgen05 convolutions), the
    configured losses, backward (input gradients on the inverse kernel maps, weight gradients on the tensor cores), gradient
    all-reduce over NCCL in buckets launched during backward, clip, both Adam steps.  Weak scaling: 8 blocks per rank."""
    import numpy as np
    import torch
    import torch.distributed as dist
    args, dev, pkg, net, rank, world = ctx.args, ctx.dev, ctx.pkg, ctx.net, ctx.rank, ctx.world
    training = importlib.import_module(pkg.__name__ + ".training")
    loss_mod = importlib.import_module(pkg.__name__ + ".model.loss")
    c_np, f_np = training_batch(rank)
    hc, hf = torch.from_numpy(c_np).pin_memory(), torch.from_numpy(f_np).pin_memory()
    dc, df = hc.to(dev), hf.to(dev)
    g = torch.Generator().manual_seed(100 + rank)
    q = torch.rand((8, 2), generator=g).to(dev)
    n_points = int(c_np.shape[0])
    n_params = sum(p.numel() for p in net.parameters())

    def time_steps(tr, steps, warmup, e2e=False):
        for _ in range(warmup):
            tr.step(dc, df, q)
        _barrier(ctx)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for _ in range(steps):
            if e2e:
                loss, aux, _ = tr.step(hc.to(dev, non_blocking=True), hf.to(dev, non_blocking=True), q)
                last = float(loss)                                   # device -> host read of the step's result
            else:
                loss, aux, _ = tr.step(dc, df, q)
        e1.record()
        torch.cuda.synchronize()
        return _reduce(ctx, e0.elapsed_time(e1), "MAX") / steps, (float(loss) if last is None else last)

    W = max(args.warmup, 3)
    tr = training.Trainer(net, loss_mod.default_loss_config(), world=world, overlap=True)
    clocks = ClockSampler(range(world)) if rank == 0 else None
    ms_warm, _ = time_steps(tr, 1, W)                                 # warm-up (allocator, packed weights, NCCL channels)
    if clocks is not None:
        clocks.mark()
    l0 = pkg._lib.launch_count()
    ms, loss = time_steps(tr, args.steps, 0)
    launches = (pkg._lib.launch_count() - l0) // max(args.steps, 1)
    clk = clocks.stop() if clocks is not None else None
    ms_e2e, _ = time_steps(tr, args.steps, 1, e2e=True)
    extra = {}
    if world > 1:
        # the same step with ONE flat all-reduce after backward (nothing overlapped), and the all-reduce alone
        tr_flat = training.Trainer(net, loss_mod.default_loss_config(), world=world, overlap=False)
        ms_flat, _ = time_steps(tr_flat, args.steps, 2)
        flat = torch.zeros(sum(p.numel() for p in tr.model_params), device=dev)
        for _ in range(3):
            dist.all_reduce(flat)
        _barrier(ctx)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dist.all_reduce(flat)
        e1.record()
        torch.cuda.synchronize()
        ar_ms = _reduce(ctx, e0.elapsed_time(e1), "MAX") / 10
        nbytes = flat.numel() * 4
        extra = {"allreduce": {"bytes": nbytes, "ms": ar_ms, "algbw_gbs": nbytes / ar_ms / 1e6, "busbw_gbs": nbytes / ar_ms / 1e6 * 2 * (world - 1) / world,
                               "buckets": len(tr.buckets.buckets)},
                 "ms_per_step_allreduce_after_backward": ms_flat, "exposed_comm_ms_removed_by_overlap": ms_flat - ms}
    total_points = _reduce(ctx, n_points, "SUM")
    if rank != 0:
        return
    line = {"metric": "Mpoints/s trained (forward + losses + backward + all-reduce + optimiser step), 256^3 blocks, batch 8 per GPU", "value": total_points / (ms * 1e3),
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": ms, "steps_per_s": 1e3 / ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.conv_mode == "auto" else "f32", "data": "synthetic",
            "config": {"workload": "CVPR_inverse_scaling_256 train256: 8 synthetic 256^3 crops of the 10-bit cloud per rank per step (the yaml's batch_size is 4), "
                                   f"{n_points} voxels on rank 0, widths 128 / 192, random-init weights, Adam 1e-4 / 1e-3, clip 1.0",
                       "sharding": f"{world} rank(s), data parallel, bucketed NCCL gradient all-reduce launched during backward", "parameters": n_params,
                       "l2": "activations of one step (> 2 GB) exceed the 126 MB L2", "conv_mode": args.conv_mode},
            "clocks": clk, "loss": loss,
            "e2e": {"value": total_points / (ms_e2e * 1e3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(hc.numel() * 4 + hf.numel() * 4) * world,
                    "d2h_bytes_per_step": 4 * world},
            "gpu_launches": int(launches)}
    line.update(extra)
    ctx.emit(line)



if __name__ == "__main__":
    main()
