"""configs[4] end to end on N GPUs (torchrun, one rank per GPU): every rank samples its shard of ONE global batch
(same seed, sample_offset = first global row), scores it (FK + tracking reward against spinkick frames 0-63, FID
moments) and the packed statistics are all-reduced over NCCL -- the path's only collective.  Rank 0 then repeats the
job alone on the first 2 shards' rows and checks that the trajectories are bit-identical and the statistics agree
(1-GPU == N-GPU).  Prints one JSON line.
usage: torchrun --nproc-per-node N tools/run_config5_multi.py [samples_total] [ddpm_steps]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import scoring, motion_dataset as MD
from deepmimic_diffusion_mujoco_b200 import synthetic as O

total = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
H, T = 64, 35
lo, hi = scoring.shard_range(total, rank, world)
net = dm.TemporalUnet(H, T, T)
net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
diff = dm.GaussianDiffusion(net, H, T, n_timesteps=steps, predict_epsilon=False).to(dev)
clip = np.load(os.path.join(ROOT, "tests", "golden", "mocap_clips.npz"))["spinkick_config"]
ref = MD.tile_clip(torch.from_numpy(clip).float(), H).to(dev)


def job(n, offset, distributed):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record()
    x = diff.p_sample_loop((n, H, T), {}, verbose=False, seed=77, sample_offset=offset).trajectories
    ev[1].record()
    sc = scoring.fk_reward(x, ref, 0.016666)
    s, sxx = scoring.fid_moments(x)
    ev[2].record()
    cnt = n
    if distributed:
        cnt = scoring.allreduce_statistics(n, s, sxx, sc["sums"])
    ev[3].record()
    torch.cuda.synchronize()
    return x, s, sxx, sc["sums"], cnt, [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]


job(hi - lo, lo, world > 1)                                 # warm-up at the timed size (plan build, NCCL communicator)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
x, s, sxx, sums, cnt, ms = job(hi - lo, lo, world > 1)
if world > 1:
    dist.barrier()
wall = time.perf_counter() - t0
out = None
if rank == 0:
    mu, sigma = scoring.finalize_statistics(cnt, s, sxx)
    out = {"config": "configs[4]: sample-and-score", "n_gpus": world, "samples": total, "ddpm_steps": steps, "wall_s": wall,
           "frames_per_s": total * H / wall, "rank0_ms": {"sample": ms[0], "score": ms[1], "allreduce": ms[2]},
           "mean_reward": float(sums[0].item() / sums[6].item()), "frames_scored": int(sums[6].item()), "count": cnt,
           "trace_sigma": float(torch.trace(sigma).item())}
    # 1-GPU == N-GPU: rank 0 redoes the first two shards' rows as ONE batch
    if world > 1:
        n2 = scoring.shard_range(total, 1, world)[1]
        xa, sa, sxxa, sumsa, _, _ = job(n2, 0, False)
        out["shard0_rows_bit_identical"] = bool(torch.equal(xa[:hi - lo], x))
if world > 1:
    # every rank's trajectories against rank 0's single-batch rerun, through an order-sensitive 64-bit checksum of
    # the raw bits (one small all_gather; no point-to-point traffic)
    def checksum(t):
        bits = t.contiguous().view(torch.int32).to(torch.int64).reshape(-1)
        w = torch.arange(1, bits.numel() + 1, device=bits.device, dtype=torch.int64)
        return torch.stack([bits.sum(), (bits * (w % 65521)).sum()])
    mine = checksum(x)
    allsums = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allsums, mine)
    if rank == 0:
        l1, h1 = scoring.shard_range(total, 1, world)
        out["shard1_rows_bit_identical"] = bool(torch.equal(checksum(xa[l1:h1]), allsums[1]))
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
