"""bench.py -- motion-frames/s sampled by the B200-native DiffMimic sampler (100-step DDPM, horizon 64).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --steps K --warmup W    (CPU arm: the oracle port of the reference sampler)

A "step" is one complete p_sample_loop (100 denoising steps) over one batch of synthetic humanoid3d qpos motions:
BASELINE.json configs[1] (box-holding apply_conditioning, H=64, B=1024 per GPU).  Weak scaling: every rank
samples its own B motions, no data-path collective (samples are independent).  One JSON line on rank 0.
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

H, T, N_DDPM, DIM = 64, 35, 100, 32
NCU_DRAM_BYTES_PER_LAUNCH_B1024 = 39.75e9  # dram__bytes_read.sum + dram__bytes_write.sum, profiles/r1f_ncu_full_summary.md
MFLOP_PER_SAMPLE_STEP = 78.868  # algorithmic FLOPs of one U-Net forward at H=64 (hooks on the reference)
METRIC = "motion_frames_per_sec_sampled_100step_ddpm_h64"


def box_holding(x):  # sampling_config.py:23-32
    x[:, :, 13:16] = 0.0
    x[:, :, 16] = 1.57
    x[:, :, 17:20] = 0.0
    x[:, :, 20] = 1.57
    return x


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1424.7), d.get("hbm_gbs", 6539.5), "measured"
    return 1400.0, 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.stop, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                    str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([c.strip() for c in o.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_port_frames_per_sec(B, steps, warmup, threads):
    """The oracle port of the reference sampler (oracle/diffusion_oracle.py) on the host cores."""
    import torch
    from oracle import diffusion_oracle as O
    torch.set_num_threads(threads)
    sd = O.seeded_state_dict(H=H, seed=0)
    sched = O.make_schedule(N_DDPM)
    mask = torch.zeros(H, T, dtype=torch.bool)
    value = torch.zeros(H, T)
    mask[:, 13:21] = True
    value[:, 16] = value[:, 20] = 1.57
    g = torch.Generator().manual_seed(0)
    times = []
    for i in range(warmup + steps):
        noise = torch.randn(N_DDPM + 1, B, H, T, generator=g)
        t0 = time.perf_counter()
        O.p_sample_loop(sd, sched, noise, mask=mask, value=value)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return B * H / (ms / 1e3), ms


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    B = 16
    # bounded: at most 3 timed + 1 warm-up passes of a B=16 sample (a few seconds each on the box's host cores)
    steps, warm = min(args.steps, 3), min(args.warmup, 1)
    v, ms = cpu_port_frames_per_sec(B, steps, warm, threads)
    sample = f"B={B} of the B=1024 workload, {steps} timed full 100-step sampling passes, torch CPU fp32"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: walk + box-holding apply_conditioning, H=64, 100-step DDPM (CPU sample B=16)"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=1024, help="samples per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import deepmimic_diffusion_mujoco_b200 as dm
    from oracle import diffusion_oracle as O  # weights generator only (seeded synthetic state_dict)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.batch
    warm = max(args.warmup, 3)

    net = dm.TemporalUnet(H, T, T, dim=DIM)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, T, n_timesteps=N_DDPM, predict_epsilon=False, clip_denoised=False).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass(seed):
        return diff.p_sample_loop((B, H, T), {}, verbose=False, conditioning_fn=box_holding, seed=seed)

    for i in range(warm):
        one_pass(1000 * rank + i)
    barrier()

    # ---- device-resident timing: K passes, each bracketed by events on the launching stream; L2 flushed between ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        for k in range(args.steps):
            flush.zero_()
            ev[k][0].record()
            one_pass(2000 * rank + k)
            ev[k][1].record()
        barrier()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    launches = diff.last_launch_count * args.steps
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = t.item()
    ms_per_step = total_ms / args.steps
    value = world * B * H / (ms_per_step / 1e3)

    # ---- share of the step spent outside the conv kernel: K4 alone, same launch count, events on the stream ----
    import ctypes
    from deepmimic_diffusion_mujoco_b200 import _lib
    plan = net.plan_for(B, H)
    xk = torch.randn(B, H, T, device=dev)
    mo = torch.randn(B, H, T, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    nul = ctypes.POINTER(_lib.DmCond)()
    for rep in range(2):
        e0.record()
        for i in range(N_DDPM):
            _lib.check(_lib.lib().dm_ddpm_step(plan.handle, xk.data_ptr(), mo.data_ptr(), None, 1, i, 0, 0, nul,
                                               _lib.stream_ptr()))
        e1.record()
        torch.cuda.synchronize()
    k4_ms = e0.elapsed_time(e1)
    # the dominant kernel IS the step: one persistent sample_loop_kernel launch per pass (K1+K2+K4 fused)
    tf_peak, hbm_peak, which = peaks()
    flops_step = MFLOP_PER_SAMPLE_STEP * 1e6 * B * N_DDPM
    achieved = flops_step / (ms_per_step / 1e3) / 1e12
    roofline = {"kernel": "sample_loop_kernel", "bound": "tensor", "achieved": achieved, "peak": tf_peak,
                "unit": "TFLOP/s", "frac": achieved / tf_peak,
                "traffic": NCU_DRAM_BYTES_PER_LAUNCH_B1024 if B == 1024 else None, "traffic_unit": "bytes per launch (ncu --set full, r1f)", "peak_source": which + " bf16 sustained",
                "launches_per_step": diff.last_launch_count, "avg_launch_us": 1e3 * ms_per_step / diff.last_launch_count,
                "algorithmic_flops_per_launch": flops_step / diff.last_launch_count,
                "note": "fp16 hi/lo split: the tensor pipe executes 3x the algorithmic FLOPs; "
                        "k4_ms_per_step = the 100 DDPM updates timed alone as standalone kernels (informational)"}

    # ---- end to end through the public API with host buffers (pinned), copies inside the timed region ----
    host_value = torch.zeros(H, T).pin_memory()
    host_value[:, 16] = host_value[:, 20] = 1.57
    host_out = [torch.empty(B, H, T).pin_memory() for _ in range(2)]   # double-buffered result landing zone
    dev_value = torch.empty(H, T, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    done = [None, None]

    def cond_from_host(x):
        x[:, :, 13:21] = dev_value[:, 13:21]
        return x

    def e2e_pass(seed, k):
        # every pass: conditioning values host -> device, sample, trajectories device -> host.  The read-back runs on a
        # copy stream so that it overlaps the next pass's kernel, the way a streaming consumer of samples would do it;
        # a landing buffer is reused only after its previous copy has completed.
        if done[k % 2] is not None:
            done[k % 2].synchronize()
        dev_value.copy_(host_value, non_blocking=True)
        out = diff.p_sample_loop((B, H, T), {}, verbose=False, conditioning_fn=cond_from_host, seed=seed)
        copy_stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(copy_stream):
            host_out[k % 2].copy_(out.trajectories, non_blocking=True)
            out.trajectories.record_stream(copy_stream)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        done[k % 2] = ev

    e2e_pass(1, 0)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_pass(3000 * rank + k, k)
    torch.cuda.synchronize()      # every result has landed in host memory
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": world * B * H * args.steps / te.item(), "unit": "frames/s",
           "h2d_bytes_per_step": host_value.numel() * 4, "d2h_bytes_per_step": host_out[0].numel() * 4,
           "note": "public API (GaussianDiffusion.p_sample_loop), pinned host buffers, copies inside the timed region; the "
                   "read-back of pass k overlaps the kernel of pass k+1 (copy stream, double-buffered)"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, ms = cpu_port_frames_per_sec(16, 2, 1, threads)
        cpu = {"value": v, "unit": "frames/s", "cores": threads, "kind": "port",
               "sample": "B=16 of the B=1024 workload, 2 timed full 100-step passes, oracle port on torch CPU fp32"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16x2-split operands, f32 accumulate", "data": "synthetic",
            "config": {"workload": "configs[1]: walk + box-holding apply_conditioning (shoulder/elbow inpaint), H=64, "
                                   f"B={B} per GPU, 100-step DDPM, dim=32 TemporalUnet, Philox noise",
                       "l2": "256 MB flush between timed passes", "samples_per_gpu": B},
            "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
            "k4_ms_per_step": k4_ms, "cpu_baseline": cpu}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
