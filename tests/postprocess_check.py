"""Run under torchrun on >= 2 GPUs: the query-split post-processing ("stages that do not
shard by view") -- LOF with the kNN queries split over ranks, minimal extents split by point -- must be
bit-identical to the single-GPU result on every rank, for any number of ranks."""
import importlib
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "oracle"))   # test_native_subsample (cloud generators) imports the C oracle
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    knn, sub, cfgm = mod("knn"), mod("point_cloud_postprocess.subsampling"), mod("point_cloud_postprocess.config")
    post = mod("point_cloud_postprocess.postprocess")
    from test_native_subsample import _cams, _cloud
    for n, kind, k in [(200003, "surface", 40), (50001, "clusters", 10), (97, "gauss", 40), (7, "gauss", 3)]:
        pts_np, rgb_np = _cloud(kind, n, 41)
        pts, rgb = torch.from_numpy(pts_np).to(dev), torch.from_numpy(rgb_np).to(dev)
        out1, lof1 = knn.local_outlier_factor(pts, k)
        outg, lofg = knn.local_outlier_factor_sharded(pts, k, dist.group.WORLD)
        assert torch.equal(out1, outg) and torch.equal(lof1, lofg), f"rank {rank}: sharded LOF differs (n={n})"
        K, P, sizes = _cams(9, 42)
        params = cfgm.PointCloudSubsamplingParams()
        a = sub.subsample_pointcloud_device(pts, rgb, torch.from_numpy(K), torch.from_numpy(P), torch.from_numpy(sizes), params)
        b = sub.subsample_pointcloud_device(pts, rgb, torch.from_numpy(K), torch.from_numpy(P), torch.from_numpy(sizes), params,
                                            group=dist.group.WORLD)
        assert all(torch.equal(x, y) for x, y in zip(a, b)), f"rank {rank}: sharded extents / merge differ (n={n})"
    # through postprocess_point_cloud, both stages on
    pc = cfgm.PointCloudPostprocessConfig(outlier_removal=cfgm.OutlierRemovalMethod.lof, subsample=True)
    pts_np, rgb_np = _cloud("surface", 120000, 43)
    K, P, sizes = _cams(5, 44)
    args = (torch.from_numpy(pts_np).to(dev), torch.from_numpy(rgb_np), list(K), list(P), sizes, pc, dev)
    p1, c1 = post.postprocess_point_cloud(*args)
    p2, c2 = post.postprocess_point_cloud(*args, group=dist.group.WORLD)
    assert torch.equal(p1, p2) and torch.equal(c1.to(dev), c2.to(dev))
    dist.barrier()
    if rank == 0:
        print(f"postprocess ok: world {world}, {p1.shape[0]} of {pts_np.shape[0]} points kept")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
