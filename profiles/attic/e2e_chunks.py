# end-to-end (host buffers) step time against the number of copy/compute blocks
import sys, time, importlib
sys.path.insert(0, "/root/repo")
import torch, numpy as np
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=1, device="cpu", low_conf_every=50)
dh, mh, ih = sc.depth.pin_memory(), sc.mask.pin_memory(), sc.image.pin_memory()
ref = None
for chunks in (1, 2, 3, 4, 6, 8, 1):
    def step():
        return mdi.points_from_host_views(cfg, dh, mh, ih, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev,
                                          L.IMAGE_U8, chunks=chunks)
    for _ in range(3): r = step()
    ts = []
    for _ in range(8):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = step(); torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    if ref is None: ref = (r[0].xyz.clone(), r[1].clone())
    same = torch.equal(r[0].xyz, ref[0]) and torch.equal(r[1], ref[1])
    mdi._TRACE = []
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = step(); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
    tr = mdi._TRACE; mdi._TRACE = None
    print("   ", "  ".join(f"{l}: {tr[0][1].elapsed_time(e):.2f}" for l, e in tr[1:]), f" wall {wall:.2f}")
    print(f"chunks {chunks}: median {np.median(ts):.2f} ms  min {min(ts):.2f}  same_as_one_block={same}")
