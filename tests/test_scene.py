"""Full-scene tiling front / back end (BASELINE config 4) and the quality metrics of the north star's 0.5 % gate."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import metrics_ref
from oracle.inputs import make_inputs
from tmdiff_b200 import tiling
from util import rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def vec():
    return np.load(os.path.join(ROOT, "tests", "golden", "scene_vectors.npz"))


def test_patch_order_matches_reference(vec):
    tiles, cube = torch.from_numpy(vec["tiles16"]), torch.from_numpy(vec["cube16"])
    assert torch.equal(tiling.patch(tiles, 4, 4), cube)            # patch_16
    assert torch.equal(tiling.unpatch(cube, 8), tiles)             # unpatch_16
    big = torch.arange(2 * 24 * 40, dtype=torch.float32).reshape(2, 24, 40)
    assert torch.equal(tiling.patch(tiling.unpatch(big, 8), 3, 5), big)
    with pytest.raises(ValueError):
        tiling.unpatch(big, 16)
    with pytest.raises(ValueError):
        tiling.patch(tiles, 3, 4)


def test_upsample_matches_reference(vec):
    got = tiling.upsample_ms(torch.from_numpy(vec["lrms"]), 64)
    assert float((got - torch.from_numpy(vec["lms"])).abs().max()) < 1e-5      # cv2.resize INTER_LINEAR
    with pytest.raises(ValueError):
        tiling.upsample_ms(torch.zeros(4, 16, 16), 64)


def test_metrics_match_reference(vec):
    a, b = vec["metric_a"], vec["metric_b"]
    got = np.array([metrics_ref.sam(a, b), metrics_ref.ergas(a, b), metrics_ref.uiqc(a, b)])
    assert np.allclose(got, vec["metric_values"], rtol=1e-12)


class _FakeDiffusion:
    """Deterministic stand-in for GeneralDiffusion: 'fuses' a tile as MS + 0.25*PAN + 0.01*noise."""
    host_tiling_ok = True           # the sharding / stitching logic runs on CPU tensors under gloo

    def sample_by_dpmsolver(self, x_in, prompt, steps=20, x_T=None):
        return x_in["MS"] + 0.25 * x_in["PAN"] + 0.01 * x_T


def _scene(seed=0, C=4, size=128):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(1, size, size, generator=g), torch.rand(C, size // 4, size // 4, generator=g)


def test_sample_scene_stitches_tiles_in_order():
    pan, lrms = _scene()
    noise = torch.randn(4, 4, 64, 64, generator=torch.Generator().manual_seed(1))
    got = tiling.sample_scene(_FakeDiffusion(), pan, lrms, "QB", tile=64, batch=3, x_T=noise)
    want = tiling.upsample_ms(lrms[None], 128)[0] + 0.25 * pan + 0.01 * tiling.patch(noise, 2, 2)
    assert got.shape == (4, 128, 128) and torch.allclose(got, want, atol=1e-6)


def _scene_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pan, lrms = _scene(size=192)                                   # 9 tiles over 2 ranks: 5 + 4
    noise = torch.randn(9, 4, 64, 64, generator=torch.Generator().manual_seed(1))
    got = tiling.sample_scene(_FakeDiffusion(), pan, lrms, "QB", tile=64, batch=2, x_T=noise)
    if rank == 0:
        q.put(got)
    dist.barrier()
    dist.destroy_process_group()


def test_sample_scene_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 200
    procs = [ctx.Process(target=_scene_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    pan, lrms = _scene(size=192)
    noise = torch.randn(9, 4, 64, 64, generator=torch.Generator().manual_seed(1))
    want = tiling.sample_scene(_FakeDiffusion(), pan, lrms, "QB", tile=64, batch=16, x_T=noise)
    assert torch.equal(got, want)


@pytest.mark.gpu
def test_scene_on_gpu_equals_per_tile_sampling_and_metrics_gate():
    """A 128x128 QB scene (4 tiles) through sample_scene == the sampler on the same tiles; SAM / ERGAS / Q of the bf16
    and fp32 CUDA results within 0.5 % of the oracle's, against a seeded synthetic HR image (north star)."""
    from oracle import dpm_ref, wavbest_ref
    from tmdiff_b200 import GeneralDiffusion, WavBEST
    ch = (32, 64, 128, 256)
    sd, emb = wavbest_ref.seeded_weights(ch), wavbest_ref.seeded_text_embeddings()
    pan, lrms = _scene(seed=3)
    noise = torch.randn(4, 4, 64, 64, generator=torch.Generator().manual_seed(2))
    # the oracle on the same tiles
    lms = tiling.upsample_ms(lrms[None], 128)[0]
    pan_t, ms_t = tiling.unpatch(pan, 64), tiling.unpatch(lms, 64)
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    want_t = dpm_ref.Sampler(lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan_t, ms_t, emb["QB"], faithful=False), ns).sample(
        noise, steps=10, order=3) + ms_t
    want = tiling.patch(want_t, 2, 2)
    hr = (want + 0.02 * torch.randn(want.shape, generator=torch.Generator().manual_seed(9))).clamp(0.01, 1.0)
    to_hwc = lambda t: t.permute(1, 2, 0).double().numpy()
    ref_m = [f(to_hwc(hr), to_hwc(want.clamp(0.01, 1.0))) for f in (metrics_ref.sam, metrics_ref.ergas, metrics_ref.uiqc)]
    for dtype, tol in (("fp32", 1e-5), ("bf16", 2e-2)):
        m = WavBEST(channels=list(ch), dtype=dtype, text_embeddings=emb)
        m.load_state_dict(sd)
        m = m.cuda()
        G = GeneralDiffusion(m)
        G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, "cuda")
        got = tiling.sample_scene(G, pan.cuda(), lrms.cuda(), "QB", tile=64, steps=10, batch=3, x_T=noise).cpu()
        assert rel_l2(got - lms, want - lms) < tol          # the sampled residual itself: 1e-5 / 2e-2 (north star)
        got_m = [f(to_hwc(hr), to_hwc(got.clamp(0.01, 1.0))) for f in (metrics_ref.sam, metrics_ref.ergas, metrics_ref.uiqc)]
        for name, a, b in zip(("SAM", "ERGAS", "Q"), got_m, ref_m):
            assert abs(a - b) <= 5e-3 * abs(b), (dtype, name, a, b)
        m.release()


@pytest.mark.gpu
def test_scene_front_and_back_end_kernels(vec):
    """tmd_scene_to_tiles / tmd_tiles_to_scene against the reference's own helpers: Upsample (cv2.resize INTER_LINEAR) on
    committed vectors, DN scaling (/2047, /1023), the unpatch_16 / patch_16 order on a non-square grid, clamp + scale back."""
    lr = torch.from_numpy(vec["lrms"])                               # [2, 4, 16, 16] in [0, 1]; vec["lms"] = Upsample(lr, 64)
    for b in range(2):
        for scale in (1.0, 2047.0, 1023.0):
            lr_dn = (lr[b] * scale).cuda()
            pan_dn = (torch.rand(1, 64, 64, generator=torch.Generator().manual_seed(b)) * scale).cuda()
            pan_t, ms_t = tiling.scene_to_tiles(pan_dn, lr_dn, 32, 0, 4, img_scale=scale)
            want_ms = tiling.unpatch(torch.from_numpy(vec["lms"][b]), 32)
            assert float((ms_t.cpu() - want_ms).abs().max()) < 2e-6, (b, scale)
            assert torch.equal(pan_t.cpu(), tiling.unpatch(pan_dn.cpu() / scale, 32))
    # a 3 x 5 grid of 8-pixel tiles, a tile range in the middle, then back
    g = torch.Generator().manual_seed(3)
    pan, lr = torch.rand(1, 24, 40, generator=g).cuda(), torch.rand(3, 6, 10, generator=g).cuda()
    pan_t, ms_t = tiling.scene_to_tiles(pan, lr, 8, 4, 7)
    assert torch.equal(pan_t.cpu(), tiling.unpatch(pan.cpu(), 8)[4:11])
    lms = torch.nn.functional.interpolate(lr[None].cpu(), size=(24, 40), mode="bilinear", align_corners=False)[0]
    assert float((ms_t.cpu() - tiling.unpatch(lms, 8)[4:11]).abs().max()) < 2e-6
    tiles = torch.randn(15, 3, 8, 8, generator=g) * 0.7 + 0.5
    scene = torch.zeros(3, 24, 40).cuda()
    tiling.tiles_to_scene(tiles[:6].cuda(), scene, 0)
    tiling.tiles_to_scene(tiles[6:].cuda(), scene, 6)
    assert torch.equal(scene.cpu(), tiling.patch(tiles, 3, 5))
    tiling.tiles_to_scene(tiles.cuda(), scene, 0, img_scale=2047.0, clamp=True)
    assert torch.equal(scene.cpu(), tiling.patch(tiles, 3, 5).clamp(0, 1) * 2047.0)
    with pytest.raises(RuntimeError):
        tiling.scene_to_tiles(pan, lr, 8, 10, 7)                     # tile range outside the scene
    with pytest.raises(RuntimeError):
        tiling.scene_to_tiles(pan.cpu(), lr.cpu(), 8, 0, 1)          # no CPU path
