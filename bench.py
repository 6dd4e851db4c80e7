"""bench.py -- motion-frames/s sampled by the B200-native DiffMimic sampler (100-step DDPM, horizon 64).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --steps K --warmup W    (CPU arm: the UNMODIFIED reference sampler, oracle/_ref)

A "step" is one complete p_sample_loop (100 denoising steps) over one batch of synthetic humanoid3d qpos motions:
BASELINE.json configs[1] (box-holding apply_conditioning, H=64, B=1024 per GPU).  Weak scaling: every rank samples its
own B motions, no data-path collective (samples are independent).  After the timed sampling passes every rank also
runs the scoring half of the path on its samples (K5 FK + tracking reward, K6 FID moments, K8 SiFID window statistics),
the statistics all-reduce (the path's only collective, NCCL) and one configs[4]-sized sample-and-score job; those are
reported beside the headline, each kernel against its own roofline.  One JSON line on rank 0.
"""
import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, T, N_DDPM, DIM = 64, 35, 100, 32
MFLOP_PER_SAMPLE_STEP = 78.868  # algorithmic FLOPs of one U-Net forward at H=64 (hooks on the reference)
METRIC = "motion_frames_per_sec_sampled_100step_ddpm_h64"
WORKLOAD = "configs[1]: walk + box-holding apply_conditioning (shoulder/elbow inpaint), H=64, 100-step DDPM, dim=32 TemporalUnet"


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
        return {"tf_sustained": d.get("bf16_tflops_sustained", 1400.0), "tf_burst": d.get("bf16_tflops", 1650.0),
                "hbm": d.get("hbm_gbs", 6650.0), "source": "MEASURED_PEAKS.json"}
    return {"tf_sustained": 1400.0, "tf_burst": 1650.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def cited_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE sample_loop_kernel launch at the bench shape, from the committed
    ncu --set full capture (profiles/r2_roofline_traffic.json names the report it was read from)."""
    p = os.path.join(ROOT, "profiles", "r2_roofline_traffic.json")
    if os.path.exists(p):
        return json.load(open(p))
    return None


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


# ----------------------------------------------------------------------------------------------------------------------
# baselines: the unmodified reference (oracle/_ref, staged by oracle/make_ref.py) -- or, if it was never staged, the oracle
# port -- on the host cores, and the same reference code on the GPU through stock PyTorch kernels
# ----------------------------------------------------------------------------------------------------------------------
def reference_sampler(device="cpu"):
    """(callable B -> seconds for one 100-step box-holding p_sample_loop of B motions, kind)."""
    import torch
    from deepmimic_diffusion_mujoco_b200.synthetic import seeded_state_dict
    from oracle import make_ref
    sd = seeded_state_dict(H=H, seed=0)
    if make_ref.available():
        GD, TU = make_ref.import_reference()
        with contextlib.redirect_stdout(io.StringIO()):
            net = TU(H, T, T)
        net.load_state_dict(sd)
        diff = GD(net, H, T, n_timesteps=N_DDPM, predict_epsilon=False, clip_denoised=False).eval().to(device)

        def run(B):
            if device != "cpu":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = diff.p_sample_loop((B, H, T), {}, verbose=False, conditioning_fn=box_holding)
            if device != "cpu":
                torch.cuda.synchronize()
            assert tuple(out.trajectories.shape) == (B, H, T)
            return time.perf_counter() - t0
        return run, "reference"
    if device != "cpu":
        return None, "unavailable"
    from oracle import diffusion_oracle as O
    sched = O.make_schedule(N_DDPM)
    mask = torch.zeros(H, T, dtype=torch.bool)
    value = torch.zeros(H, T)
    mask[:, 13:21] = True
    value[:, 16] = value[:, 20] = 1.57
    g = torch.Generator().manual_seed(0)

    def run(B):
        noise = torch.randn(N_DDPM + 1, B, H, T, generator=g)
        t0 = time.perf_counter()
        O.p_sample_loop(sd, sched, noise, mask=mask, value=value)
        return time.perf_counter() - t0
    return run, "port"


def median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def cpu_reference_numbers(timed, warm, full):
    """C1 (B=1) and B=64 at H=64, all host threads and one thread, median of `timed` passes."""
    import torch
    run, kind = reference_sampler("cpu")
    cores = os.cpu_count() or 1
    out = {"kind": kind, "cores": cores}
    torch.set_num_threads(cores)
    for _ in range(warm):
        run(64)
    t64 = median([run(64) for _ in range(timed)])
    out["b64_all_threads"] = {"frames_per_s": 64 * H / t64, "s_per_pass": t64, "threads": torch.get_num_threads()}
    if full:
        t1 = median([run(1) for _ in range(timed)])
        out["c1_b1_all_threads"] = {"frames_per_s": H / t1, "s_per_pass": t1, "threads": torch.get_num_threads()}
        torch.set_num_threads(1)
        t1s = run(1)
        out["c1_b1_one_thread"] = {"frames_per_s": H / t1s, "s_per_pass": t1s, "threads": 1}
        t64s = run(64)
        out["b64_one_thread"] = {"frames_per_s": 64 * H / t64s, "s_per_pass": t64s, "threads": 1}
        torch.set_num_threads(cores)
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    timed, warm = max(1, min(args.steps, 3)), min(args.warmup, 1)
    r = cpu_reference_numbers(timed, warm, full=True)
    v = r["b64_all_threads"]["frames_per_s"]
    sample = (f"B=64 of the B=1024 workload: median of {timed} full 100-step box-holding p_sample_loop passes of the "
              f"{'unmodified reference (oracle/_ref)' if r['kind'] == 'reference' else 'oracle port'} on torch CPU fp32, "
              f"{r['b64_all_threads']['threads']} threads; C1 (B=1) and single-thread figures under cpu_baseline")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": timed,
        "warmup": warm, "ms_per_step": 1e3 * r["b64_all_threads"]["s_per_pass"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD + " (CPU sample B=64)"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": r["cores"], "kind": r["kind"], "sample": sample,
                         "c1_b1_all_threads": r.get("c1_b1_all_threads"), "c1_b1_one_thread": r.get("c1_b1_one_thread"),
                         "b64_one_thread": r.get("b64_one_thread")},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def gpu_eager_baseline(B):
    """The unmodified reference on the SAME B200 through stock PyTorch CUDA kernels (cuDNN conv, ATen GroupNorm / Mish):
    'the reference on the same box' of BASELINE.md section 4.  fp32 with TF32 off (the parity-grade setting) and on."""
    import torch
    run, kind = reference_sampler("cuda")
    if run is None:
        return {"unavailable": "oracle/_ref was not staged (python oracle/make_ref.py in the build container)"}
    out = {"kind": kind, "B": B, "note": "reference p_sample_loop on cuda, stock PyTorch kernels, median of 2 after 1 warm-up"}
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    try:
        for name, flag in (("fp32", False), ("tf32", True)):
            torch.backends.cudnn.allow_tf32 = flag
            torch.backends.cuda.matmul.allow_tf32 = flag
            run(B)
            t = median([run(B) for _ in range(2)])
            out[name] = {"frames_per_s": B * H / t, "ms_per_pass": 1e3 * t}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=1024, help="samples per GPU")
    ap.add_argument("--score-batch", type=int, default=8192, help="samples per GPU of the configs[4] sample-and-score job")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only: skip scoring kernels, collective, baselines")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    import deepmimic_diffusion_mujoco_b200 as dm
    from deepmimic_diffusion_mujoco_b200 import _lib, scoring, motion_dataset as MD
    from deepmimic_diffusion_mujoco_b200.synthetic import seeded_state_dict

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.batch
    warm = max(args.warmup, 3)
    pk = peaks()

    net = dm.TemporalUnet(H, T, T, dim=DIM)
    net.load_state_dict(seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, T, n_timesteps=N_DDPM, predict_epsilon=False, clip_denoised=False).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

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
            last = one_pass(2000 * rank + k)
            ev[k][1].record()
        barrier()
    total_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev))
    launches = diff.last_launch_count * args.steps
    ms_per_step = total_ms / args.steps
    value = world * B * H / (ms_per_step / 1e3)

    # the dominant kernel IS the step: one persistent sample_loop_kernel launch per pass (K1+K2+K4 fused)
    flops_step = MFLOP_PER_SAMPLE_STEP * 1e6 * B * N_DDPM
    achieved = flops_step / (ms_per_step / 1e3) / 1e12
    traffic = cited_traffic() if B == 1024 else None
    roofline = {"kernel": "sample_loop_kernel", "bound": "tensor", "achieved": achieved, "peak": pk["tf_sustained"],
                "unit": "TFLOP/s", "frac": achieved / pk["tf_sustained"],
                "traffic": traffic["bytes_per_launch"] if traffic else None,
                "traffic_source": traffic["source"] if traffic else "no ncu capture of this build committed yet",
                "peak_source": pk["source"] + " bf16 sustained (the kernel runs ~50 ms)",
                "launches_per_step": diff.last_launch_count, "avg_launch_us": 1e3 * ms_per_step / diff.last_launch_count,
                "algorithmic_flops_per_launch": flops_step / diff.last_launch_count,
                "note": "fp16 hi/lo split: the tensor pipe executes 3x the algorithmic FLOPs"}

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
            e = torch.cuda.Event()
            e.record(copy_stream)
        done[k % 2] = e

    e2e_pass(1, 0)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_pass(3000 * rank + k, k)
    torch.cuda.synchronize()      # every result has landed in host memory
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": world * B * H * args.steps / e2e_s, "unit": "frames/s",
           "h2d_bytes_per_step": host_value.numel() * 4, "d2h_bytes_per_step": host_out[0].numel() * 4,
           "note": "public API (GaussianDiffusion.p_sample_loop), pinned host buffers, copies inside the timed region; the "
                   "read-back of pass k overlaps the kernel of pass k+1 (copy stream, double-buffered)"}

    kernels, collective, score_job, cpu, eager = None, None, None, None, None
    if not args.no_extras:
        # ---- the scoring half on this rank's samples, each kernel alone against its roofline (events on the stream, L2
        #      flushed before every repetition, median of 5) ----
        def timed(fn, reps=5):
            fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(reps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            return max_over_ranks(median(ts))

        clip = np.load(os.path.join(ROOT, "tests", "golden", "mocap_clips.npz"))["spinkick_config"]
        ref = MD.tile_clip(torch.from_numpy(clip).float(), H).to(dev)
        # ---- configs[4] at this rank count: sample `score_batch` motions per GPU of ONE global batch (global Philox rows),
        #      FK + reward + FID moments, all-reduce; wall clock barrier to barrier ----
        Ns = args.score_batch

        def score_job_once():
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            evs[0].record()
            xs = diff.p_sample_loop((Ns, H, T), {}, verbose=False, seed=77, sample_offset=rank * Ns).trajectories
            evs[1].record()
            sc = scoring.fk_reward(xs, ref, 0.016666)
            s2, sxx2 = scoring.fid_moments(xs)
            evs[2].record()
            n = scoring.allreduce_statistics(Ns, s2, sxx2, sc["sums"])
            evs[3].record()
            torch.cuda.synchronize()
            return n, sc["sums"], [evs[i].elapsed_time(evs[i + 1]) for i in range(3)], xs

        score_job_once()
        barrier()
        t0 = time.perf_counter()
        n_tot, sums2, parts, xs = score_job_once()
        barrier()
        wall = max_over_ranks(time.perf_counter() - t0)
        score_job = {"workload": f"configs[4]: {world * Ns} samples sampled (100 steps), FK + tracking reward vs spinkick frames "
                                 "0-63, FID moments, statistics all-reduced", "samples": world * Ns, "wall_s": wall,
                     "frames_per_s": world * Ns * H / wall, "rank0_ms": {"sample": parts[0], "score": parts[1], "allreduce": parts[2]},
                     "count": n_tot, "mean_reward": float(sums2[0].item() / max(sums2[6].item(), 1.0))}

        # kernel-by-kernel numbers at the configs[4] per-GPU size (the bench batch, 9 MB, would be launch-latency sized)
        x = xs
        frames = Ns * H
        D = H * T
        s = torch.zeros(D, device=dev, dtype=torch.float64)
        sxx = torch.zeros(D, D, device=dev, dtype=torch.float64)
        plan = net.plan_for(B, H)
        xk, mo = torch.randn(B, H, T, device=dev), torch.randn(B, H, T, device=dev)
        nul = ctypes.POINTER(_lib.DmCond)()

        def hbm_entry(name, ms, nbytes, note):
            return {"name": name, "ms": ms, "bound": "hbm", "algorithmic_bytes": nbytes, "achieved_gbs": nbytes / ms / 1e6,
                    "frac": nbytes / ms / 1e6 / pk["hbm"], "peak_source": pk["source"] + " hbm_gbs", "note": note}

        kernels = []
        ms = timed(lambda: _lib.check(_lib.lib().dm_ddpm_step(plan.handle, xk.data_ptr(), mo.data_ptr(), None, 1, 50, 0, 0,
                                                              nul, _lib.stream_ptr())))
        kernels.append(hbm_entry("K4 ddpm_step_tiled_kernel (stand-alone launch at the bench batch, Philox noise)", ms, B * H * 12 * T,
                                 f"{B * H} frames x 420 B; at this size (9 MB) the launch is latency-sized -- inside the "
                                 "fused loop K4 is part of sample_loop_kernel"))
        ms = timed(lambda: scoring.fk_reward(x, ref, 0.016666))
        k5_note = ("; the kernel is FP32-issue bound, not HBM bound: 1 983 warp instructions per frame, issue slots 70 % busy, FMA pipe "
                   "56 %, DRAM traffic = algorithmic bytes (ncu, profiles/r1f_ncu_full_summary.md); `frac` is still quoted against HBM")
        kernels.append(hbm_entry("K5 fk_reward_kernel (reward only)", ms, frames * 144, f"{frames} frames x (140 B read + 4 B written)" + k5_note))
        kernels[-1]["limiter"] = "fp32-issue"
        ms = timed(lambda: scoring.fk_reward(x, ref, 0.016666, return_body_pos=True))
        kernels.append(hbm_entry("K5 fk_reward_kernel (+ body positions)", ms, frames * 324, f"{frames} frames x (140 + 4 + 180 B)"))
        ms = timed(lambda: scoring.fid_moments(x, s, sxx))
        fl = 2.0 * Ns * D * D
        kernels.append({"name": "K6 fid_transpose + fid_syrk (tcgen05) + fid_reduce", "ms": ms, "bound": "tensor",
                        "algorithmic_flops": fl, "achieved_tflops": fl / ms / 1e9, "frac": fl / ms / 1e9 / pk["tf_burst"],
                        "peak_source": pk["source"] + " bf16 burst", "note": f"N={Ns}, D={D}: 2 N D^2 (full matrix; the upper "
                        "triangle is computed, 3 MMAs per product for the fp16 hi/lo split)"})
        nw = min(Ns, 2048)
        ms = timed(lambda: scoring.window_moments(x[:nw], 10))
        kernels.append(hbm_entry("K8 window_moments_kernel (SiFID statistics)", ms, nw * (H * T * 4 + 350 * 4 + 350 * 350 * 4),
                                 f"{nw} samples x (input + 350 mean + 350x350 covariance, fp32)"))

        # ---- the one exchange step: all-reduce of {N, sum x, sum xx^T, reward sums} (fp64), NCCL ----
        sums = scoring.fk_reward(x, ref, 0.016666)["sums"]
        nbytes = (1 + s.numel() + sxx.numel() + sums.numel()) * 8
        if world > 1:
            scoring.allreduce_statistics(Ns, s, sxx, sums)
            barrier()
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                scoring.allreduce_statistics(Ns, s, sxx, sums)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = max_over_ranks(median(ts))
            collective = {"op": "all_reduce(SUM) fp64, NCCL", "bytes": nbytes, "ms": ms, "n_ranks": world,
                          "algbw_gbs": nbytes / ms / 1e6, "busbw_gbs": 2.0 * (world - 1) / world * nbytes / ms / 1e6,
                          "note": "includes packing / unpacking the statistics into one flat buffer (scoring.allreduce_statistics)"}
        else:
            collective = {"op": "all_reduce(SUM) fp64, NCCL", "bytes": nbytes, "ms": 0.0, "n_ranks": 1, "busbw_gbs": None,
                          "note": "single rank: nothing to exchange"}

        if rank == 0 and world == 1:
            eager = gpu_eager_baseline(B)
            if not args.no_cpu_baseline:
                r = cpu_reference_numbers(timed=2, warm=1, full=False)
                cpu = {"value": r["b64_all_threads"]["frames_per_s"], "unit": "frames/s", "cores": r["cores"], "kind": r["kind"],
                       "sample": "B=64 of the B=1024 workload: median of 2 full 100-step box-holding p_sample_loop passes (after 1 "
                                 "warm-up) of the " + ("unmodified reference (oracle/_ref)" if r["kind"] == "reference" else "oracle port") +
                                 f" on torch CPU fp32, {r['b64_all_threads']['threads']} threads"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16x2-split operands, f32 accumulate", "data": "synthetic",
            "config": {"workload": WORKLOAD + f", B={B} per GPU, Philox noise",
                       "l2": "256 MB flush between timed passes", "samples_per_gpu": B},
            "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "kernels": kernels,
            "collective": collective, "sample_and_score": score_job, "gpu_eager_baseline": eager, "cpu_baseline": cpu}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
