"""Run under torchrun on >= 2 GPUs: the fused emit + exchange (CloudExchange, emit kernel writing
into every rank's peer-mapped cloud) must equal the NCCL all-gather of the per-rank clouds, which
in turn equals the single-GPU cloud of all views (bit-identical for G = 1, 2, 4, 8)."""
import importlib
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    L, mdi, syn, dmod = mod("_lib"), mod("monocular_depth_init"), mod("synthetic"), mod("distributed")
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = "adaptive"
    n_views = int(os.environ.get("MDI_CHECK_VIEWS", "11"))  # ragged blocks; small values leave ranks without views
    sc = syn.make_scene(n_views=n_views, H=240, W=320, n_sfm=3000, seed=5, device=str(dev), low_conf_every=4)
    mine = list(dmod.view_partition(n_views, world, rank))
    sel = lambda t: t[mine[0]:mine[-1] + 1] if mine else t[:0]
    names = [sc.names[i] for i in mine]
    args = (sel(sc.depth), sel(sc.mask), sel(sc.image), sel(sc.K), sel(sc.c2w), names, sc.points, sc.point_indices, dev,
            L.IMAGE_U8)
    exchanges = [dmod.CloudExchange(dev, capacity_rows=1024),                   # small on purpose: collective growth
                 dmod.CloudExchange(dev, capacity_rows=1024, byte_rgb=True)]    # rgb travels as 3 bytes per point
    for rep in range(4):  # the buffers are reused across scenes
        ex = exchanges[rep % 2]
        res = mdi.points_from_views(cfg, *args, exchange=ex)
        got_xyz, got_rgb = res.xyz.clone(), res.rgb.clone()
        loc = mdi.points_from_views(cfg, *args)
        none = torch.empty((0, 3), device=dev)
        ref_xyz, ref_rgb, counts = dmod.allgather_points(loc.xyz if loc.xyz is not None else none,
                                                         loc.rgb if loc.rgb is not None else none)
        assert got_xyz.shape == ref_xyz.shape, (got_xyz.shape, ref_xyz.shape)
        assert torch.equal(got_xyz, ref_xyz) and torch.equal(got_rgb, ref_rgb), f"rank {rank} rep {rep}: clouds differ"
        assert res.extras["rank_counts"] == [int(c) for c in counts]
    full = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points,
                                 sc.point_indices, dev, L.IMAGE_U8)
    assert torch.equal(full.xyz, got_xyz) and torch.equal(full.rgb, got_rgb), "differs from the single-GPU cloud"
    dist.barrier()
    if rank == 0:
        print(f"exchange ok: world {world}, {got_xyz.shape[0]} points, counts {res.extras['rank_counts']}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
