"""Full-scene front / back end of the sampler (BASELINE config 4): cut a scene into
non-overlapping patches in the reference's row-major `patch_16` order, up-sample the low-resolution MS to the PAN
grid, run the patches through the sampler (sharded over the ranks of the process group), stitch the result.

    unpatch / patch  data/LRHR_dataset.py:29-53 (patch_16 / unpatch_16, generalised from a 4x4 to any grid)
    upsample_ms      data/LRHR_dataset.py:58-84 (cv2.resize INTER_LINEAR == bilinear, half-pixel centres)

Everything stays on the device as strided views / one interpolation; there is no host round trip.  Tiles have no halo:
seams are part of the reference's tiling semantics.
"""
import torch
import torch.nn.functional as F

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


@torch.no_grad()
def sample_scene(diffusion, pan, lrms, prompt, tile=64, steps=20, batch=64, x_T=None, group=None):
    """Pansharpen one scene: pan [1, H, W], lrms [C, H/4, W/4] in [0, 1] (sensor DN / 2047 or / 1023,
    data/LRHR_dataset.py:134-139) -> [C, H, W].  Patches are split contiguously over the ranks of `group` (no
    per-step communication) and gathered once at the end; every rank returns the whole scene."""
    dev = pan.device
    Cc = lrms.shape[0]
    Hs, Ws = pan.shape[-2:]
    lms = upsample_ms(lrms[None], Hs)[0] if Hs == Ws else F.interpolate(lrms[None].float(), size=(Hs, Ws), mode="bilinear",
                                                                        align_corners=False)[0]
    pan_t, ms_t = unpatch(pan, tile), unpatch(lms, tile)
    T = pan_t.shape[0]
    dist_on = torch.distributed.is_available() and torch.distributed.is_initialized()
    rank = torch.distributed.get_rank(group) if dist_on else 0
    world = torch.distributed.get_world_size(group) if dist_on else 1
    lo, hi = shard_range(T, rank, world)
    outs = []
    for s in range(lo, hi, batch):
        e = min(s + batch, hi)
        noise = torch.randn(e - s, Cc, tile, tile, device=dev) if x_T is None else x_T[s:e].to(dev)
        outs.append(diffusion.sample_by_dpmsolver({"PAN": pan_t[s:e], "MS": ms_t[s:e], "Res": noise}, prompt, steps=steps, x_T=noise))
    local = torch.cat(outs) if outs else pan_t.new_zeros((0, Cc, tile, tile))
    tiles = gather_tiles(local, T, group) if world > 1 else local
    return patch(tiles, Hs // tile, Ws // tile)
