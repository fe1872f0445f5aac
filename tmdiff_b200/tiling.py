"""Full-scene front / back end of the sampler (BASELINE config 4): cut a scene into
non-overlapping patches in the reference's row-major `patch_16` order, up-sample the low-resolution MS to the PAN
grid, run the patches through the sampler (sharded over the ranks of the process group), stitch the result.

    unpatch / patch  data/LRHR_dataset.py:29-53 (patch_16 / unpatch_16, generalised from a 4x4 to any grid)
    upsample_ms      data/LRHR_dataset.py:58-84 (cv2.resize INTER_LINEAR == bilinear, half-pixel centres)

On CUDA tensors the front end (DN scaling, bilinear up-sampling, tile cut) and the back end (stitch, clamp, scale back) are one
kernel each (csrc/scene.cu); `unpatch` / `patch` / `upsample_ms` are the host-side mirrors of the reference helpers the tests
check them against.  Tiles have no halo: seams are part of the reference's tiling semantics.
"""
import torch
import torch.nn.functional as F

from . import _lib
from .parallel import gather_tiles, shard_range


def unpatch(scene, tile):
    """[C, gh*tile, gw*tile] -> [gh*gw, C, tile, tile], tile (i, j) at index i*gw + j (unpatch_16 order)."""
    Cc, Hs, Ws = scene.shape
    if Hs % tile or Ws % tile:
        raise ValueError("scene %dx%d is not a whole number of %d-pixel tiles" % (Hs, Ws, tile))
    gh, gw = Hs // tile, Ws // tile
    return scene.reshape(Cc, gh, tile, gw, tile).permute(1, 3, 0, 2, 4).reshape(gh * gw, Cc, tile, tile).contiguous()


def patch(tiles, grid_h, grid_w):
    """Inverse of `unpatch`: [gh*gw, C, h, w] -> [C, gh*h, gw*w] (patch_16 order)."""
    T, Cc, h, w = tiles.shape
    if T != grid_h * grid_w:
        raise ValueError("%d tiles do not fill a %dx%d grid" % (T, grid_h, grid_w))
    return tiles.reshape(grid_h, grid_w, Cc, h, w).permute(2, 0, 3, 1, 4).reshape(Cc, grid_h * h, grid_w * w).contiguous()


def upsample_ms(ms, resolution):
    """Upsample(img_MSs, resolution): [B, C, h, w] -> [B, C, resolution, resolution], bilinear with half-pixel centres
    and replicated borders (what cv2.INTER_LINEAR computes when enlarging)."""
    if ms.dim() != 4:
        raise ValueError("img_MSs must be a 4-dimensional array with shape [B, C, H, W].")
    return F.interpolate(ms.float(), size=(resolution, resolution), mode="bilinear", align_corners=False)


def scene_to_tiles(pan, lrms, tile, first, count, img_scale=1.0, ratio=4):
    """Tiles [first, first + count) of a scene as sampler inputs, in one kernel (tmd_scene_to_tiles): pan [1, Hs, Ws] and lrms
    [C, Hs/ratio, Ws/ratio] in sensor DN (img_scale 2047 / 1023, data/LRHR_dataset.py:134-139; 1.0: already in [0, 1]) ->
    PAN tiles [count, 1, tile, tile], bilinearly up-sampled MS tiles [count, C, tile, tile] (cv2 INTER_LINEAR, :97-122)."""
    if not (pan.is_cuda and lrms.is_cuda):
        raise RuntimeError("scene_to_tiles needs CUDA tensors (tmdiff_b200 has no CPU path)")
    pan, lrms = pan.contiguous().float(), lrms.contiguous().float()
    Cc = lrms.shape[0]
    Hs, Ws = pan.shape[-2:]
    pan_t = torch.empty(count, 1, tile, tile, device=pan.device)
    ms_t = torch.empty(count, Cc, tile, tile, device=pan.device)
    with torch.cuda.device(pan.device):
        _lib.check(_lib.load().tmd_scene_to_tiles(_lib.ptr(pan), _lib.ptr(lrms), _lib.ptr(pan_t), _lib.ptr(ms_t), Cc, Hs, Ws, ratio, tile, first,
                                                  count, float(img_scale), _lib.stream_ptr()))
    return pan_t, ms_t


def tiles_to_scene(tiles, scene, first, img_scale=1.0, clamp=False):
    """Fused tiles [count, C, tile, tile] into their places of `scene` [C, Hs, Ws] (patch_16 order), clamped to [0, 1] and scaled
    back to DN when asked (tmd_tiles_to_scene; general_sharpening_joint_random_batch_finetune.py:39-42, 136-142)."""
    if not (tiles.is_cuda and scene.is_cuda):
        raise RuntimeError("tiles_to_scene needs CUDA tensors (tmdiff_b200 has no CPU path)")
    tiles = tiles.contiguous().float()
    count, Cc, tile, _ = tiles.shape
    with torch.cuda.device(scene.device):
        _lib.check(_lib.load().tmd_tiles_to_scene(_lib.ptr(tiles), _lib.ptr(scene), Cc, scene.shape[1], scene.shape[2], tile, first, count,
                                                  float(img_scale), int(bool(clamp)), _lib.stream_ptr()))
    return scene


@torch.no_grad()
def sample_scene(diffusion, pan, lrms, prompt, tile=64, steps=20, batch=64, x_T=None, group=None, img_scale=1.0, ratio=4, clamp=False):
    """Pansharpen one scene: pan [1, H, W], lrms [C, H/ratio, W/ratio] -> [C, H, W].  img_scale = 1: inputs already in [0, 1]
    (sensor DN / 2047 or / 1023, data/LRHR_dataset.py:134-139); img_scale = 2047 / 1023: inputs and result in sensor DN (then
    clamp=True reproduces the reference's clamp to [0, 1] before scaling back).  Tiles are split contiguously over the ranks of
    `group` (no per-step communication) and gathered once at the end; every rank returns the whole scene.  On CUDA tensors the
    front and back end are one kernel each (scene_to_tiles / tiles_to_scene); CPU tensors are accepted only from test doubles
    that declare `host_tiling_ok` (the sharding / stitching logic under gloo)."""
    dev = pan.device
    Cc = lrms.shape[0]
    Hs, Ws = pan.shape[-2:]
    if Hs % tile or Ws % tile:
        raise ValueError("scene %dx%d is not a whole number of %d-pixel tiles" % (Hs, Ws, tile))
    T = (Hs // tile) * (Ws // tile)
    on_gpu = pan.is_cuda
    if not on_gpu:
        if not getattr(diffusion, "host_tiling_ok", False):
            raise RuntimeError("sample_scene needs CUDA tensors (tmdiff_b200 has no CPU path)")
        lms = F.interpolate((lrms[None].float() / img_scale), size=(Hs, Ws), mode="bilinear", align_corners=False)[0]
        pan_all, ms_all = unpatch(pan.float() / img_scale, tile), unpatch(lms, tile)
    dist_on = torch.distributed.is_available() and torch.distributed.is_initialized()
    rank = torch.distributed.get_rank(group) if dist_on else 0
    world = torch.distributed.get_world_size(group) if dist_on else 1
    lo, hi = shard_range(T, rank, world)
    outs = []
    for s in range(lo, hi, batch):
        e = min(s + batch, hi)
        if on_gpu:
            pan_t, ms_t = scene_to_tiles(pan, lrms, tile, s, e - s, img_scale, ratio)
        else:
            pan_t, ms_t = pan_all[s:e], ms_all[s:e]
        noise = torch.randn(e - s, Cc, tile, tile, device=dev) if x_T is None else x_T[s:e].to(dev)
        outs.append(diffusion.sample_by_dpmsolver({"PAN": pan_t, "MS": ms_t, "Res": noise}, prompt, steps=steps, x_T=noise))
    local = torch.cat(outs) if outs else pan.new_zeros((0, Cc, tile, tile), dtype=torch.float32)
    tiles = gather_tiles(local, T, group) if world > 1 else local
    if on_gpu:
        return tiles_to_scene(tiles, torch.empty(Cc, Hs, Ws, device=dev), 0, img_scale, clamp)
    out = patch(tiles, Hs // tile, Ws // tile)
    return (out.clamp(0.0, 1.0) if clamp else out) * img_scale
