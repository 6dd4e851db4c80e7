"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against the CPU
oracle and the committed reference-generated fixtures.  Tolerances: <= 1e-3 relative per denoising step
(north_star); conditioned channels bit-exact; FK <= 1e-5 m."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import diffusion_oracle as O  # noqa: E402
from oracle import scoring_oracle as SO  # noqa: E402

REL_STEP = 1e-3


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item()


@pytest.fixture(scope="module")
def dm():
    import deepmimic_diffusion_mujoco_b200 as m
    assert torch.cuda.is_available(), "GPU tests need a B200"
    assert os.path.exists(m._lib.LIB_PATH), "libdm_b200.so missing: the CUDA path must be the one that runs"
    return m


def box_cond(H):
    mask = torch.zeros(H, 35, dtype=torch.bool)
    value = torch.zeros(H, 35)
    mask[:, 13:21] = True
    value[:, 16] = 1.57
    value[:, 20] = 1.57
    return mask, value


def box_holding(x):
    x[:, :, 13:16] = 0.0
    x[:, :, 16] = 1.57
    x[:, :, 17:20] = 0.0
    x[:, :, 20] = 1.57
    return x


@pytest.mark.parametrize("B,H", [(3, 64), (17, 16), (1, 8), (2, 128), (16, 40)])
def test_unet_layers_match_oracle(dm, B, H):
    sd = O.seeded_state_dict(H=H, seed=B)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(B * 100 + H)
    x = torch.randn(B, H, 35, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, record=rec)
    y = net(x.cuda(), {}, t.cuda()).cpu()
    plan = net.plan_for(B, H)
    for i, (name, C, L) in enumerate(plan.layer_names()):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        assert rel(got, rec[name]) < 1e-4, (name, rel(got, rec[name]))
    assert rel(y, y_ref) < 1e-4


@pytest.mark.parametrize("B,H,T,dim", [(3, 32, 69, 128), (2, 32, 35, 64)])
def test_wide_unet_matches_oracle(dm, B, H, T, dim):
    """The later-generation configuration (temporal_v2.py dim=128, eval.py:189-197 horizon 32; T=69 features)."""
    sd = O.seeded_state_dict(H=H, T=T, dim=dim, seed=B)
    net = dm.TemporalUnet(H, T, T, dim=dim)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(dim + T)
    x = torch.randn(B, H, T, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, dim=dim, record=rec)
    y = net(x.cuda(), {}, t.cuda()).cpu()
    plan = net.plan_for(B, H)
    for i, (name, C, L) in enumerate(plan.layer_names()):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        assert rel(got, rec[name]) < 1e-4, (name, rel(got, rec[name]))
    assert rel(y, y_ref) < 1e-4


@pytest.mark.parametrize("name", ["unet_chain_plain", "unet_chain_box_eps_clip", "unet_chain_start_maxt"])
def test_forward_matches_reference_fixture(dm, golden_dir, name):
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    B, H, n_steps, pe, clip, cond, start, max_t, seed = [int(v) for v in z["meta"]]
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=seed))
    y = net(torch.from_numpy(z["fwd_x"]).cuda(), {}, torch.from_numpy(z["fwd_t"]).cuda()).cpu()
    assert rel(y, torch.from_numpy(z["fwd_y"])) < 1e-4


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", ["unet_chain_plain", "unet_chain_box_eps_clip", "unet_chain_start_maxt"])
def test_chain_matches_reference_fixture(dm, golden_dir, name, fused):
    """Full p_sample_loop with injected noise against the chain recorded from the UNMODIFIED reference."""
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    B, H, n_steps, pe, clip, cond, start, max_t, seed = [int(v) for v in z["meta"]]
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=seed))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n_steps, predict_epsilon=bool(pe), clip_denoised=bool(clip)).cuda()
    kw = {}
    if cond:
        kw["conditioning_fn"] = box_holding
    if start:
        kw["starting_motion"] = torch.from_numpy(z["start"])
    if max_t >= 0:
        kw["max_timesteps"] = max_t
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True,
                             sample_fn=dm.InjectedNoise(torch.from_numpy(z["noise"])), fused=fused, **kw)
    ref = torch.from_numpy(z["chain"])
    chain = out.chains.cpu()
    assert chain.shape == ref.shape
    for k in range(ref.shape[1]):
        assert rel(chain[:, k], ref[:, k]) < REL_STEP, (k, rel(chain[:, k], ref[:, k]))
    assert torch.equal(out.trajectories.cpu(), chain[:, -1])
    if cond:
        m = box_cond(H)[0].expand(B, H, 35)
        for k in range(ref.shape[1]):
            assert torch.equal(chain[:, k][m], ref[:, k][m]), "conditioned channels must be bit-exact"


@pytest.mark.parametrize("B,H", [(3, 64), (17, 16), (2, 24), (5, 128), (1, 512)])
def test_attention_unet_layers_match_oracle(dm, B, H):
    """attention=True (Residual(PreNorm(LinearAttention)), helpers.py:75-131): every layer, attention stages included."""
    sd = O.seeded_state_dict(H=H, seed=B + 40, attention=True)
    net = dm.TemporalUnet(H, 35, 35, attention=True)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(B * 100 + H + 7)
    x = torch.randn(B, H, 35, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, record=rec)
    y = net(x.cuda(), {}, t.cuda()).cpu()
    plan = net.plan_for(B, H)
    names = plan.layer_names()
    assert sum(1 for n, _, _ in names if n.endswith(".2") or n == "mid_attn") == 8
    for i, (name, C, L) in enumerate(names):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        assert rel(got, rec[name]) < 1e-4, (name, rel(got, rec[name]))
    assert rel(y, y_ref) < 1e-4


def test_wide_attention_unet_matches_oracle(dm):
    """dim=128 with attention=True: 128..1024-channel attention stages (weight chunks of 16 down to 2 rows)."""
    B, H, T, dim = 2, 32, 35, 128
    sd = O.seeded_state_dict(H=H, T=T, dim=dim, seed=9, attention=True)
    net = dm.TemporalUnet(H, T, T, dim=dim, attention=True)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(77)
    x = torch.randn(B, H, T, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, dim=dim, record=rec)
    y = net(x.cuda(), {}, t.cuda()).cpu()
    plan = net.plan_for(B, H)
    for i, (name, C, L) in enumerate(plan.layer_names()):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        assert rel(got, rec[name]) < 1e-4, (name, rel(got, rec[name]))
    assert rel(y, y_ref) < 1e-4


def test_attention_chain_matches_reference_fixture(dm, golden_dir):
    """p_sample_loop of the attention=True model against the chain recorded from the UNMODIFIED reference."""
    z = np.load(os.path.join(golden_dir, "unet_chain_attention.npz"))
    B, H, n_steps, pe, clip, cond, start, max_t, seed = [int(v) for v in z["meta"]]
    net = dm.TemporalUnet(H, 35, 35, attention=True)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=seed, attention=True))
    y = net(torch.from_numpy(z["fwd_x"]).cuda(), {}, torch.from_numpy(z["fwd_t"]).cuda()).cpu()
    assert rel(y, torch.from_numpy(z["fwd_y"])) < 1e-4
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n_steps, predict_epsilon=bool(pe), clip_denoised=bool(clip)).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, conditioning_fn=box_holding,
                             sample_fn=dm.InjectedNoise(torch.from_numpy(z["noise"])))
    ref = torch.from_numpy(z["chain"])
    chain = out.chains.cpu()
    for k in range(ref.shape[1]):
        assert rel(chain[:, k], ref[:, k]) < REL_STEP, (k, rel(chain[:, k], ref[:, k]))
    m = box_cond(H)[0].expand(B, H, 35)
    assert torch.equal(chain[:, -1][m], ref[:, -1][m])
    assert diff.last_launch_count == 1                    # the attention stages run inside the fused loop kernel
    out2 = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, conditioning_fn=box_holding,
                              sample_fn=dm.InjectedNoise(torch.from_numpy(z["noise"])), fused=False)
    assert diff.last_launch_count > n_steps * 40          # ... and as launches of their own, with the same arithmetic
    assert torch.equal(out2.chains, out.chains)


@pytest.mark.parametrize("B,H,n", [(3, 128, 4), (18, 64, 3), (1, 320, 2)])
def test_attention_chain_long_horizons(dm, B, H, n):
    """attention=True inside the fused loop at the horizons whose conv layers take the big-tile kernel variant (wave-staged
    epilogues from H = 128, TMEM rounds above 256) and with a second, partly filled 16-sample group: a short chain
    against the oracle, ONE launch, and bit-equal to the per-layer launches."""
    sd = O.seeded_state_dict(H=H, seed=21, attention=True)
    noise = torch.randn(n + 1, B, H, 35, generator=torch.Generator().manual_seed(H * 7 + B))
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35, attention=True)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                             conditioning_fn=box_holding)
    assert diff.last_launch_count == 1
    chain = out.chains.cpu()
    for k in range(n + 1):
        assert rel(chain[:, k], ref[:, k]) < REL_STEP, k
    m = mask.expand(B, H, 35)
    assert torch.equal(chain[:, -1][m], ref[:, -1][m])
    out2 = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                              conditioning_fn=box_holding, fused=False)
    assert torch.equal(out2.chains, out.chains)


@pytest.mark.parametrize("tag,sched", [("x0_cosine", "cosine"), ("eps_linear", "linear")])
@pytest.mark.parametrize("fused", [True, False])
def test_v4_sampler_matches_reference_fixture(dm, golden_dir, tag, sched, fused):
    """DiffusionV4.sample (diffusion_v4.py:82-149) on the CUDA loop vs the unmodified reference's output."""
    z = np.load(os.path.join(golden_dir, "v4_reference.npz"))
    B, H, T, px0, seed = [int(v) for v in z[tag + "_meta"]]
    sd = O.seeded_state_dict(H=H, seed=seed)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    d = dm.DiffusionV4(T, 1e-4, 0.02, 35, H, device="cuda", predict_x0=bool(px0), schedule_type=sched)
    noise = torch.from_numpy(z[tag + "_noise"])
    x, chain = d.sample(net.cuda(), B, noise=noise, return_chain=True, fused=fused)
    ref = torch.from_numpy(z[tag + "_x"])
    assert rel(x.cpu(), ref) < REL_STEP
    oc = O.v4_sample(sd, O.v4_betas(T, 1e-4, 0.02, sched), noise, bool(px0))
    for k in range(oc.shape[1]):
        assert rel(chain[:, k].cpu(), oc[:, k]) < REL_STEP, k
    assert d.last_launch_count == (1 if fused else (T - 1) * 41 + 1)
    # Philox path: finite, seeded, reproducible
    a = d.sample(net, 4, seed=9)
    b = d.sample(net, 4, seed=9)
    assert torch.isfinite(a).all() and torch.equal(a, b)


def test_hundred_step_chain_vs_oracle(dm):
    """100 steps, H=64, box-holding conditioning: free-running chain stays within 1e-3 of the fp32 oracle."""
    B, H, n = 2, 64, 100
    sd = O.seeded_state_dict(H=H, seed=5)
    g = torch.Generator().manual_seed(77)
    noise = torch.randn(n + 1, B, H, 35, generator=g)
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                             conditioning_fn=box_holding)
    chain = out.chains.cpu()
    worst = max(rel(chain[:, k], ref[:, k]) for k in range(n + 1))
    assert worst < REL_STEP, worst
    m = mask.expand(B, H, 35)
    assert torch.equal(chain[:, -1][m], ref[:, -1][m])


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("B,H", [(3, 128), (2, 256)])
def test_hundred_step_chain_long_horizons(dm, B, H, fused):
    """100 free-running steps at H = 128 / 256 with injected noise: these horizons take different code from H = 64
    (several l-tiles per unit, the one-issuer fallback at H = 256, non-staged epilogues), so their error accumulation
    over a whole chain is checked slot by slot against the fp32 oracle."""
    n = 100
    sd = O.seeded_state_dict(H=H, seed=6)
    noise = torch.randn(n + 1, B, H, 35, generator=torch.Generator().manual_seed(H + B))
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                             conditioning_fn=box_holding, fused=fused)
    chain = out.chains.cpu()
    worst = max(rel(chain[:, k], ref[:, k]) for k in range(n + 1))
    assert worst < REL_STEP, worst
    m = mask.expand(B, H, 35)
    assert torch.equal(chain[:, -1][m], ref[:, -1][m])


def test_full_batch_rows_match_oracle(dm):
    """configs[1] batch (B = 1024, 64 clusters): rows spread over the first, middle and last clusters -- and both CTAs
    of a cluster -- are compared VALUE by value with the oracle run on just those rows (samples never interact)."""
    B, H, n = 1024, 64, 10
    rows = [0, 7, 8, 15, 16, 500, 777, 1015, 1016, 1023]
    sd = O.seeded_state_dict(H=H, seed=0)
    noise = torch.randn(n + 1, B, H, 35, generator=torch.Generator().manual_seed(31))
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise[:, rows].contiguous(), mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                             conditioning_fn=box_holding)
    assert diff.last_launch_count == 1
    chain = out.chains[rows].cpu()
    for k in range(n + 1):
        assert rel(chain[:, k], ref[:, k]) < REL_STEP, k
    m = mask.expand(len(rows), H, 35)
    assert torch.equal(chain[:, -1][m], ref[:, -1][m])


@pytest.mark.parametrize("B,H,ctas,rows", [(512, 128, 4, [0, 3, 4, 9, 14, 300, 511]), (160, 256, 4, [1, 6, 11, 12, 159]),
                                             (608, 128, 2, [0, 9, 300, 607]), (608, 256, 2, [0, 15, 607]),
                                             (20, 160, 4, [0, 7, 13, 19]), (18, 480, 4, [2, 8, 17]), (608, 480, 2, [5, 607])])
def test_split_and_unsplit_layouts_at_long_horizons(dm, B, H, ctas, rows):
    """Small batches at H >= 128 run with FOUR CTAs per 16-sample group (two pairs, half of the positions each, GroupNorm
    sums exchanged between them); when the clusters of four would not all be resident at once (the device reports how many
    it holds: 33 on a 148-SM B200) the plan keeps one pair per group.  Both layouts, at a large split batch and the smallest
    batch that is unsplit on every device: rows from every CTA's share of the step kernel (samples 4r .. 4r+3
    of a split group) against the oracle run on just those rows, fused and per-layer launches.  H = 160 and 480 (the training
    and projection horizons of archive/3_5_long_projection.ipynb) split into halves that are NOT whole MMA tiles (padded last
    tile, direct stores); unsplit, H = 480 goes through TMEM in rounds."""
    n = 3
    sd = O.seeded_state_dict(H=H, seed=8)
    noise = torch.randn(n + 1, B, H, 35, generator=torch.Generator().manual_seed(B + H))
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise[:, rows].contiguous(), mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    for fused in (True, False):
        out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                                 conditioning_fn=box_holding, fused=fused)
        assert net.plan_for(B, H).ctas_per_group() == ctas
        chain = out.chains[rows].cpu()
        for k in range(n + 1):
            assert rel(chain[:, k], ref[:, k]) < REL_STEP, (fused, k)
        m = mask.expand(len(rows), H, 35)
        assert torch.equal(chain[:, -1][m], ref[:, -1][m])


@pytest.mark.parametrize("B,H,n", [(2, 480, 3), (1, 512, 2), (1, 1024, 1)])
def test_horizons_beyond_256_layers_and_chain(dm, B, H, n):
    """The reference samples at any horizon that is a multiple of 8 (diffusion.py:186-205 takes an arbitrary shape;
    archive/3_5_long_projection.ipynb cell 19 samples [1, 480, 35] from a model trained at 160).  Above H = 256 a layer's
    conv output no longer fits the 512 TMEM columns: the l-tiles go through TMEM in rounds and the GroupNorm epilogue
    parks the raw values in a scratch between its two passes.  Every layer, and a short chain, against the oracle."""
    sd = O.seeded_state_dict(H=H, seed=7)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(H)
    x = torch.randn(B, H, 35, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, record=rec)
    y = net(x.cuda(), {}, t.cuda()).cpu()
    plan = net.plan_for(B, H)
    for i, (name, C, L) in enumerate(plan.layer_names()):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        assert rel(got, rec[name]) < 1e-4, (name, rel(got, rec[name]))
    assert rel(y, y_ref) < 1e-4
    if n == 0:
        return
    noise = torch.randn(n + 1, B, H, 35, generator=g)
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    for fused in (True, False):
        out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                                 conditioning_fn=box_holding, fused=fused)
        chain = out.chains.cpu()
        for k in range(n + 1):
            assert rel(chain[:, k], ref[:, k]) < REL_STEP, (fused, k)
        m = mask.expand(B, H, 35)
        assert torch.equal(chain[:, -1][m], ref[:, -1][m])
        assert tuple(out.trajectories.shape) == (B, H, 35)


def test_plan_lifecycle_returns_device_memory(dm):
    """Plans of every layout (one pair per group, multicast clusters, L-split with its fallback build, attention blocks,
    TMEM rounds) are created, used for a pass and freed, twice over: the second round must not need more device memory
    than the first (dm_plan_free returns the packed weights, the per-layer activations, the schedule tables, the workspace
    the split layout's abandoned first build allocated, ...)."""
    import gc
    shapes = [(3, 64, False), (32, 64, False), (5, 128, False), (2, 480, False), (608, 480, False), (4, 64, True)]

    def one_round():
        for B, H, attn in shapes:
            net = dm.TemporalUnet(H, 35, 35, attention=attn)
            net.load_state_dict(O.seeded_state_dict(H=H, seed=1, attention=attn))
            diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=2, predict_epsilon=False).cuda()
            out = diff.p_sample_loop((B, H, 35), {}, verbose=False, seed=3)
            assert torch.isfinite(out.trajectories).all()
            del out, diff, net
        gc.collect()
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        return torch.cuda.mem_get_info()[0]

    free1 = one_round()
    free2 = one_round()
    assert free1 - free2 < 8 << 20, (free1, free2)     # (allocator granularity, not a plan's worth: the smallest plan is > 100 MB)


def test_ddpm_step_kernel_matches_oracle(dm):
    """K4 alone through dm_ddpm_step, both parameterisations, with clamp and a per-sample mask."""
    import ctypes
    from deepmimic_diffusion_mujoco_b200 import _lib
    B, H, n = 5, 16, 50
    sd = O.seeded_state_dict(H=H, seed=3)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n).cuda()
    plan = net.plan_for(B, H)
    plan.set_schedule(diff._schedule_bufs(), diff._schedule_id())
    sched = O.make_schedule(n)
    g = torch.Generator().manual_seed(9)
    for pe, clip, t in [(0, 0, 30), (1, 1, 7), (1, 0, 0), (0, 1, 49)]:
        x = torch.randn(B, H, 35, generator=g)
        out = torch.randn(B, H, 35, generator=g)
        z = torch.randn(B, H, 35, generator=g)
        mask = torch.rand(B, H, 35, generator=g) < 0.2
        value = torch.randn(B, H, 35, generator=g)
        ref = O.apply_cond(O.ddpm_step(sched, x, out, z, t, bool(pe), bool(clip)), mask, value)
        xd, od, zd = x.cuda(), out.cuda(), z.cuda()
        md, vd = mask.to(torch.uint8).cuda().contiguous(), value.cuda()
        c = _lib.DmCond(md.data_ptr(), vd.data_ptr(), 1, 1)
        _lib.check(_lib.lib().dm_ddpm_step(plan.handle, xd.data_ptr(), od.data_ptr(), zd.data_ptr(), 0, t, pe, clip,
                                           ctypes.byref(c), _lib.stream_ptr()))
        got = xd.cpu()
        assert rel(got, ref) < 1e-6
        assert torch.equal(got[mask], value[mask])


def test_philox_sampling_statistics_and_determinism(dm):
    """Built-in sampler (in-kernel Philox noise, persistent cluster kernel): deterministic per seed, finite,
    conditioned entries exact, different seeds differ, and bit-identical to the one-launch-per-layer path."""
    B, H = 32, 64
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=20, predict_epsilon=False).cuda()
    a = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=11).trajectories.clone()
    b = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=11).trajectories.clone()
    c = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=12).trajectories.clone()
    assert torch.isfinite(a).all()
    assert torch.equal(a, b)
    assert not torch.equal(a, c)
    assert torch.equal(a[:, :, 16], torch.full((B, H), 1.57, device="cuda"))
    assert torch.equal(a[:, :, 13:16], torch.zeros(B, H, 3, device="cuda"))
    assert diff.last_launch_count == 1
    d = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=11, fused=False).trajectories
    assert diff.last_launch_count > 20 * 40
    assert torch.equal(a, d)


def test_initial_noise_is_standard_normal(dm):
    B, H = 256, 64
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=10, predict_epsilon=False).cuda()
    x = diff.p_sample_loop((B, H, 35), {}, verbose=False, max_timesteps=0, seed=3).trajectories
    assert abs(x.mean().item()) < 0.01 and abs(x.std().item() - 1.0) < 0.01
    k = ((x ** 4).mean() / (x ** 2).mean() ** 2).item()
    assert abs(k - 3.0) < 0.1


def test_fk_positions_and_reward(dm):
    from deepmimic_diffusion_mujoco_b200 import scoring
    g = np.random.default_rng(0)
    N, H, dt = 6, 24, 1 / 60
    ref = np.zeros((H, 35))
    ref[:, 2] = 0.9
    ref[:, 3] = 1.0
    ref[:, 7:] = 0.3 * g.standard_normal((H, 28))
    ref[:, 3:7] += 0.1 * g.standard_normal((H, 4))
    qpos = ref[None] + 0.15 * g.standard_normal((N, H, 35))
    qpos[0] = ref  # a perfect tracker: reward 1
    out = scoring.fk_reward(torch.tensor(qpos, dtype=torch.float32).cuda(),
                            torch.tensor(ref, dtype=torch.float32), dt, return_body_pos=True)
    q32 = np.asarray(torch.tensor(qpos, dtype=torch.float32).numpy(), dtype=np.float64)
    r32 = np.asarray(torch.tensor(ref, dtype=torch.float32).numpy(), dtype=np.float64)
    bp = SO.body_positions(q32)
    assert np.abs(out["body_pos"].cpu().numpy() - bp).max() < 1e-5
    r, terms = SO.tracking_reward(q32, r32, dt)
    got = out["reward"].cpu().numpy()
    assert np.abs(got - r).max() < 2e-4
    assert abs(got[0].mean() - 1.0) < 1e-5
    s = out["sums"].cpu().numpy()
    assert abs(s[0] - r.sum()) < 1e-2 and s[6] == N * H
    assert np.allclose(s[2:6], terms.reshape(-1, 4).sum(0), atol=2e-2)


def test_fid_moments_match_oracle(dm):
    from deepmimic_diffusion_mujoco_b200 import scoring
    g = torch.Generator().manual_seed(4)
    N, H = 700, 8
    data = torch.randn(N, H, 35, generator=g) * 0.3 + torch.randn(1, H, 35, generator=g)
    fid = scoring.MotionFID()
    mu, sigma = fid.calculate_activation_statistics(data.cuda())
    mu_o, sig_o = SO.activation_statistics(data.numpy())
    assert np.abs(mu - mu_o).max() < 1e-5
    assert np.abs(sigma - sig_o).max() < 1e-4 * max(1.0, np.abs(sig_o).max())
    other = data[: N // 2] * 1.1 + 0.05
    mu2, sigma2 = fid.calculate_activation_statistics(other.cuda())
    f = fid.calculate_fid(mu, sigma, mu2, sigma2)
    f_o = SO.fid_from_statistics(*SO.activation_statistics(data.numpy()), *SO.activation_statistics(other.numpy()))
    assert abs(f - f_o) < 2e-2 * max(1.0, abs(f_o))
    # two halves of the same distribution: FID close to zero (test_motion_fid.py:106-124)
    h1, h2 = data[::2], data[1::2]
    f0 = fid.calculate_fid(*fid.calculate_activation_statistics(h1.cuda()), *fid.calculate_activation_statistics(h2.cuda()))
    f0_o = SO.fid_from_statistics(*SO.activation_statistics(h1.numpy()), *SO.activation_statistics(h2.numpy()))
    assert abs(f0 - f0_o) < 2e-2 * max(1.0, abs(f0_o))


def test_eval_metrics_match_reference_fixture(dm, golden_dir):
    """SiFID (batched window statistics kernel + batched SVD), inter / intra diversity vs eval.py:15-89."""
    from deepmimic_diffusion_mujoco_b200 import scoring
    z = np.load(os.path.join(golden_dir, "eval_metrics_reference.npz"))
    gen, gt = torch.from_numpy(z["gen"]).cuda(), torch.from_numpy(z["gt"]).cuda()
    assert np.array_equal(scoring.slice_motion_sample(gt, 10).cpu().numpy(), z["gt_slices"])
    mu, sig = scoring.window_moments(gen, 10)
    for i in range(gen.shape[0]):
        mo, so = SO.activation_statistics(SO.slice_windows(z["gen"][i], 10))
        assert np.abs(mu[i].cpu().numpy() - mo).max() < 1e-5
        assert np.abs(sig[i].cpu().numpy() - so).max() < 1e-5 * max(1.0, np.abs(so).max())
    assert abs(scoring.calc_sifid(gen, gt) - float(z["sifid"])) < 1e-2 * float(z["sifid"])
    assert abs(scoring.calc_inter_diversity(gen) - float(z["inter"])) < 1e-5 * float(z["inter"])
    assert abs(scoring.calc_intra_diversity(gen, offsets=z["offsets"]) - float(z["intra"])) < 1e-5 * float(z["intra"])
    with pytest.raises(dm.DmError):
        scoring.window_moments(gen[:, :15], 10)      # one window: covariance undefined
    with pytest.raises(dm.DmError):
        scoring.window_moments(gen.cpu(), 10)        # no CPU fallback


def test_errors_are_loud(dm):
    net = dm.TemporalUnet(64, 35, 35)
    with pytest.raises(dm.DmError):
        net(torch.zeros(1, 64, 35), {}, torch.zeros(1))          # CPU tensor: no fallback
    with pytest.raises(dm.DmError):
        net(torch.zeros(1, 60, 35).cuda(), {}, torch.zeros(1).cuda())   # H % 8 != 0
    diff = dm.GaussianDiffusion(net, 64, 35, n_timesteps=10).cuda()
    with pytest.raises(dm.DmError):
        diff.p_sample_loop((1, 64, 35), {}, verbose=False, conditioning_fn=lambda x: x * 0.5)
    with pytest.raises(dm.DmError):
        diff.p_sample_loop((1, 64, 35), {}, verbose=False, sample_fn=lambda *a: None)


def test_conditioning_structure_may_change_between_calls(dm):
    """The same function object goes from batch-uniform to per-sample values (and back) between calls: every call must
    inpaint each sample with ITS OWN constraints.  The kernels take the [H,T]-vs-[B,H,T] decision from device-resident
    flags, so no host-side cache can go stale; a function that stops being a pure overwrite poisons its samples with
    NaN and raises on a later call."""
    from deepmimic_diffusion_mujoco_b200 import sampling_config as SC
    B, H = 6, 16
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=1))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=3, predict_epsilon=False).cuda()
    keys = torch.zeros(B, H, 35, device="cuda")
    state = {"impure": False}

    def cond(x):
        if state["impure"]:
            return x * 0.5
        x[:, 3:9, 0:7] = keys[:, 3:9, 0:7]
        return x

    a = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=cond, seed=4).trajectories   # first sight: uniform zeros
    assert torch.equal(a[:, 3:9, 0:7], keys[:, 3:9, 0:7])
    keys.copy_(torch.randn(B, H, 35, generator=torch.Generator().manual_seed(2)))                        # now per-sample
    b = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=cond, seed=4).trajectories
    assert torch.equal(b[:, 3:9, 0:7], keys[:, 3:9, 0:7]) and not torch.equal(b[0, 3:9, 0:7], b[1, 3:9, 0:7])
    free = torch.ones(B, H, 35, dtype=torch.bool)
    free[:, 3:9, 0:7] = False
    assert torch.isfinite(b).all() and not torch.equal(a[free.cuda()], torch.zeros_like(a[free.cuda()]))
    keys[:] = keys[:1]                                                                                   # uniform again
    c = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=cond, seed=4).trajectories
    assert torch.equal(c[:, 3:9, 0:7], keys[:, 3:9, 0:7])
    state["impure"] = True
    bad = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=cond, seed=4).trajectories
    assert torch.isnan(bad).all()
    with pytest.raises(dm.DmError):
        SC.check_pending(block=True)
    state["impure"] = False
    ok = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=cond, seed=4).trajectories
    assert torch.equal(ok, c)


# ---------------------------------------------------------------------------------------------------
# BASELINE.json configs[2..4] at test sizes, and size-independent properties at the full configs[1] size
# ---------------------------------------------------------------------------------------------------
def _clip(golden_dir, name):
    z = np.load(os.path.join(golden_dir, "mocap_clips.npz"))
    return torch.from_numpy(z[name + "_config"]).float()


def test_config3_long_horizon_arm_limit(dm):
    """configs[2]: H=256 with the hand/arm joint limit inpaint (tiles that exceed the staged epilogue's budget)."""
    B, H, n = 3, 256, 4
    sd = O.seeded_state_dict(H=H, seed=2)
    g = torch.Generator().manual_seed(5)
    noise = torch.randn(n + 1, B, H, 35, generator=g)
    mask, value = box_cond(H)
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    for fused in (True, False):
        out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                                 conditioning_fn=box_holding, fused=fused)
        chain = out.chains.cpu()
        for k in range(n + 1):
            assert rel(chain[:, k], ref[:, k]) < REL_STEP, (fused, k)
        m = mask.expand(B, H, 35)
        assert torch.equal(chain[:, -1][m], ref[:, -1][m])


def test_config4_inbetweening_blend_of_two_clips(dm, golden_dir):
    """configs[3]: H=128, frames [0,32) from walk and [96,128) from spinkick kept fixed, the middle sampled."""
    B, H, n = 5, 128, 4
    walk, kick = _clip(golden_dir, "walk")[:32], _clip(golden_dir, "spinkick")[:32]
    start = torch.randn(B, H, 35, generator=torch.Generator().manual_seed(8))
    start[:, :32] = walk
    start[:, 96:] = kick

    def blend(x):   # 5_motion_blending.ipynb c20-21 pattern with two clips
        x[:, :32] = walk.to(x.device)
        x[:, 96:] = kick.to(x.device)
        return x

    sd = O.seeded_state_dict(H=H, seed=4)
    noise = torch.randn(n + 1, B, H, 35, generator=torch.Generator().manual_seed(6))
    mask = torch.zeros(H, 35, dtype=torch.bool)
    mask[:32] = True
    mask[96:] = True
    value = torch.zeros(H, 35)
    value[:32] = walk
    value[96:] = kick
    ref = O.p_sample_loop(sd, O.make_schedule(n), noise, mask=mask, value=value, starting_motion=start)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=dm.InjectedNoise(noise),
                             conditioning_fn=blend, starting_motion=start)
    chain = out.chains.cpu()
    for k in range(n + 1):
        assert rel(chain[:, k], ref[:, k]) < REL_STEP, k
    assert torch.equal(chain[:, -1, :32], walk.expand(B, 32, 35)) and torch.equal(chain[:, -1, 96:], kick.expand(B, 32, 35))


def test_config5_sample_and_score(dm, golden_dir):
    """configs[4]: sample, then FK + tracking reward against spinkick frames 0-63 and FID moments of the samples."""
    from deepmimic_diffusion_mujoco_b200 import scoring
    B, H = 64, 64
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=5, predict_epsilon=False).cuda()
    x = diff.p_sample_loop((B, H, 35), {}, verbose=False, seed=21).trajectories
    ref = _clip(golden_dir, "spinkick")[:64]
    out = scoring.fk_reward(x, ref, 0.016666)
    r, terms = SO.tracking_reward(x.double().cpu().numpy(), ref.double().numpy(), 0.016666)
    assert np.abs(out["reward"].cpu().numpy() - r).max() < 2e-4
    s, sxx = scoring.fid_moments(x)
    X = x.double().cpu().reshape(B, -1)
    assert torch.allclose(s.cpu(), X.sum(0), rtol=1e-6, atol=1e-4)
    assert torch.allclose(sxx.cpu(), X.T @ X, rtol=1e-5, atol=1e-3)
    # linearity of the moments: two half batches sum to the full batch (what the all-reduce relies on)
    s1, q1 = scoring.fid_moments(x[:20])
    s2, q2 = scoring.fid_moments(x[20:])
    assert torch.allclose(s1 + s2, s, rtol=1e-6, atol=1e-4) and torch.allclose(q1 + q2, sxx, rtol=1e-5, atol=1e-3)


def test_full_size_properties(dm):
    """configs[1] at full size (B=1024, H=64): finite, conditioned channels exact, deterministic, and every sample is
    independent of the batch it was drawn in (Philox keyed by global sample index) -- the sharding invariant."""
    H, n = 64, 6
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=n, predict_epsilon=False).cuda()
    a = diff.p_sample_loop((1024, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=5).trajectories.clone()
    b = diff.p_sample_loop((1024, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=5).trajectories.clone()
    small = diff.p_sample_loop((48, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=5).trajectories.clone()
    assert torch.isfinite(a).all() and torch.equal(a, b)
    assert torch.equal(a[:, :, 16], torch.full((1024, H), 1.57, device="cuda"))
    assert torch.equal(a[:, :, 17:20], torch.zeros(1024, H, 3, device="cuda"))
    assert torch.equal(a[:48], small)
    # a shard that starts at global sample 16 reproduces rows 16..63 of the global batch (same seed on every rank)
    shard = diff.p_sample_loop((48, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=5, sample_offset=16).trajectories
    assert torch.equal(a[16:64], shard)
    again = diff.p_sample_loop((48, H, 35), {}, verbose=False, conditioning_fn=box_holding, seed=5).trajectories
    assert torch.equal(again, small)          # the offset does not stick to the plan


def test_fk_invariants_at_full_size(dm):
    """configs[4] scoring size (65536 samples x 64 frames): properties of forward kinematics that hold whatever the
    pose -- FK has no reference implementation to pin against (MuJoCo is absent), so besides the float64 oracle the
    kernel is held to the model file's own numbers (assets/dp_env_v2.xml:20-107)."""
    from deepmimic_diffusion_mujoco_b200 import scoring
    N, H = 65536, 64
    g = torch.Generator(device="cuda").manual_seed(3)
    q = 0.5 * torch.randn(N, H, 35, device="cuda", generator=g)
    q[:, :, 3] += 1.0
    ref = q[0].clone()
    out = scoring.fk_reward(q, ref, 1 / 60, return_body_pos=True)
    bp = out["body_pos"]                                  # BODY_DEFS order
    ROOT, CHEST, NECK, RHIP, RKNEE, RANK, RSHO, RELB, RWR, LHIP, LKNEE, LANK, LSHO, LELB, LWR = range(15)
    assert torch.isfinite(bp).all() and torch.equal(bp[:, :, ROOT], q[:, :, 0:3])

    def dist(a, b):
        return (bp[:, :, a] - bp[:, :, b]).norm(dim=-1)
    # rigid links between body origins joined by hinges through the child's origin
    for a, b, length in [(CHEST, ROOT, 0.236151), (NECK, CHEST, 0.223894), (RWR, RELB, 0.258947), (LWR, LELB, 0.258947),
                         (RSHO, CHEST, (0.02405 ** 2 + 0.18311 ** 2 + 0.2435 ** 2) ** 0.5),
                         (LSHO, CHEST, (0.02405 ** 2 + 0.18311 ** 2 + 0.2435 ** 2) ** 0.5)]:
        assert (dist(a, b) - length).abs().max().item() < 5e-6, (a, b)
    # anchored hinges (elbow 0.028, knee 0.035, hip / ankle 0.02 off the body origin): the origin moves on a circle
    # about the anchor, so the link length stays within the anchor offset of the nominal one
    for a, b, length, slack in [(RELB, RSHO, 0.274788, 0.056), (RKNEE, RHIP, 0.421546, 0.07), (RANK, RKNEE, 0.40987, 0.04),
                                (LANK, LKNEE, 0.40987, 0.04)]:
        assert (dist(a, b) - length).abs().max().item() < slack + 1e-5
    # rigid motion of the root moves every body rigidly: rotate by a fixed quaternion r and translate
    r = torch.tensor([0.5, 0.5, -0.5, 0.5], device="cuda")           # unit quaternion (w, x, y, z)
    def qmul(a, b):
        aw, ax, ay, az = a.unbind(-1); bw, bx, by, bz = b.unbind(-1)
        return torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                            aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], -1)
    def rot(v):
        w, x, y, z = r.tolist()
        R = torch.tensor([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                          [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                          [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]], device="cuda")
        return v @ R.T
    sub = q[:2048].clone()
    qn = sub[:, :, 3:7] / sub[:, :, 3:7].norm(dim=-1, keepdim=True)
    moved = sub.clone()
    moved[:, :, 3:7] = qmul(r.expand_as(qn), qn)
    shift = torch.tensor([0.3, -1.2, 0.7], device="cuda")
    moved[:, :, 0:3] = rot(sub[:, :, 0:3]) + shift
    bp2 = scoring.fk_reward(moved, ref, 1 / 60, return_body_pos=True)["body_pos"]
    assert (bp2 - (rot(bp[:2048]) + shift)).abs().max().item() < 2e-5
    # reward: a perfect tracker scores 1 on every frame; rewards lie in (0, 1]
    assert (out["reward"][0] - 1.0).abs().max().item() < 1e-5
    assert out["reward"].min().item() > 0 and out["reward"].max().item() <= 1.0 + 1e-6
    assert out["sums"][6].item() == N * H


def test_on_device_loader_matches_oracle(dm, golden_dir, tmp_path):
    """dm_mocap_to_qpos / dm_motion_dataset vs the loader oracle on the three committed clips (fp32 cast of the
    float64 conversion: at most one fp32 ulp apart where the device's atan2 rounds the last double bit differently)."""
    import json
    from deepmimic_diffusion_mujoco_b200 import motion_dataset as MD
    from oracle import loader_oracle as LO
    z = np.load(os.path.join(golden_dir, "mocap_clips.npz"))
    for name in ["walk", "cartwheel", "spinkick"]:
        frames = z[name + "_frames"]
        got = MD.mocap_qpos_on_device(frames, torch.device("cuda")).cpu().numpy()
        want = z[name + "_config"].astype(np.float32)
        assert got.shape == want.shape
        assert np.abs(got - want).max() <= 2.4e-7 * max(1.0, np.abs(want).max())
        path = tmp_path / f"humanoid3d_{name}.txt"
        path.write_text(json.dumps({"Loop": "wrap", "Frames": frames.tolist()}))
        ds = MD.MotionDataset(str(path), device="cuda")
        ref = LO.motion_dataset(frames)
        assert ds.motion_data.is_cuda and tuple(ds.motion_data.shape) == ref.shape
        assert np.abs(ds.motion_data.cpu().numpy() - ref).max() <= 2.4e-7 * max(1.0, np.abs(ref).max())
        b = ds[3]
        assert torch.equal(b.conditions[0], b.trajectories[0]) and torch.equal(ds[0].trajectories[3], b.trajectories[0])
        # velocity rows (dm_mocap_to_qvel): one fp32 ulp of each value, relative to the row's scale
        gv = MD.mocap_qvel_on_device(frames, torch.device("cuda")).cpu().numpy()
        wv = z[name + "_vel"].astype(np.float32)
        assert gv.shape == wv.shape and not gv[0].any()
        assert (np.abs(gv - wv) <= 1.2e-7 * np.maximum(1.0, np.abs(wv)) + 1e-6 * 0).all() or \
            np.abs(gv - wv).max() <= 2.4e-7 * max(1.0, np.abs(wv).max())
        ds2 = MD.MotionDatasetV2(str(path), device="cuda")
        row = ds2[0].trajectories
        assert row.is_cuda and tuple(row.shape) == (frames.shape[0], 69) and int(ds2[0].motion_class) == MD.MOTION_CLASSES[f"humanoid3d_{name}"]
        want69 = LO.motion_dataset_v2_rows(frames)
        assert np.abs(row.cpu().numpy() - want69).max() <= 2.4e-7 * max(1.0, np.abs(want69).max())


def test_scoring_edge_sizes(dm):
    """Ragged / minimal sizes of the scoring kernels: one frame, one sample, a sample count that crosses the K6 chunk
    boundary (8192 + 1) and a feature width that is not a multiple of the 128-wide SYRK tile, the minimum of two
    SiFID windows, an unaligned qpos pointer."""
    from deepmimic_diffusion_mujoco_b200 import scoring
    g = np.random.default_rng(11)
    # FK + reward: N = 1, H = 1 (no previous frame), and H = 3 through an offset (16-byte misaligned) view
    for N, H in [(1, 1), (3, 3), (2, 129)]:
        q = 0.3 * g.standard_normal((N + 1, H, 35)); q[:, :, 3] += 1.0
        ref = q[0]
        qt = torch.tensor(q, dtype=torch.float32).cuda()[1:]          # storage offset H*35 floats
        out = scoring.fk_reward(qt, torch.tensor(ref, dtype=torch.float32), 1 / 30, return_body_pos=True)
        q32 = qt.cpu().double().numpy()
        r, _ = SO.tracking_reward(q32, torch.tensor(ref, dtype=torch.float32).double().numpy(), 1 / 30)
        assert np.abs(out["reward"].cpu().numpy() - r).max() < 2e-4
        assert np.abs(out["body_pos"].cpu().numpy() - SO.body_positions(q32)).max() < 1e-5
        assert out["sums"][6].item() == N * H
    # FID moments: N = 1; N = 8193 with D = 200 (two 128-wide tiles, the second one ragged)
    for N, D in [(1, 70), (8193, 200), (130, 129)]:
        X = torch.tensor(g.standard_normal((N, D)) * 0.5 + 0.2, dtype=torch.float32)
        s, sxx = scoring.fid_moments(X.cuda())
        Xd = X.double()
        assert torch.allclose(s.cpu(), Xd.sum(0), rtol=1e-6, atol=1e-4)
        assert torch.allclose(sxx.cpu(), Xd.T @ Xd, rtol=2e-5, atol=2e-3)
        assert torch.allclose(sxx, sxx.T, rtol=1e-6, atol=1e-4)
    # SiFID statistics: exactly two windows
    x = torch.tensor(g.standard_normal((3, 20, 35)), dtype=torch.float32)
    mu, sig = scoring.window_moments(x.cuda(), 10)
    for i in range(3):
        mo, so = SO.activation_statistics(SO.slice_windows(x[i].numpy(), 10))
        assert np.abs(mu[i].cpu().numpy() - mo).max() < 1e-5 and np.abs(sig[i].cpu().numpy() - so).max() < 1e-5
