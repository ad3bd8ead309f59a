#!/usr/bin/env python
"""Sweep of RunOptions.streams (view ranges of one step on separate CUDA streams) on a bench workload:
one scene, resident inputs, K timed steps per setting (CUDA events around the loop, like bench.py's
headline loop).  Prints one JSON line per setting and writes the best setting to --best-file.

  python profiles/attic/streams_sweep.py [--workload c2_...] [--steps 20] [--settings 1,2,3,4,6,8]
"""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default=bench.DEFAULT_WORKLOAD)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--settings", default="1,2,3,4,6,8,1,2,3,4")
    ap.add_argument("--best-file", default="")
    args = ap.parse_args()
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    L, mdi, emod = bench.mod("_lib"), bench.mod("monocular_depth_init"), bench.mod("engine")
    V, H, W, n_sfm, aligner, subsample = bench.WORKLOADS[args.workload]
    cfg = bench.make_config(aligner, subsample)
    sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
    sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
    torch.manual_seed(1234)

    def step(streams):
        return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points,
                                     sc.point_indices, dev, L.IMAGE_U8, sfm=sfm, streams=streams)

    best = {}
    for n in [int(x) for x in args.settings.split(",")]:
        for _ in range(4):
            step(n)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            r = step(n)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        best[n] = min(best.get(n, 1e9), ms)
        print(json.dumps({"workload": args.workload, "streams": n, "ms_per_step": round(ms, 4), "kept": int(r.total)}), flush=True)
    pick = min(best, key=best.get)
    # a setting has to beat one stream by more than run-to-run noise to be worth its extra launches
    if best[pick] > 0.97 * best.get(1, 1e9):
        pick = 1
    print(json.dumps({"best": pick, "ms": best}))
    if args.best_file:
        Path(args.best_file).write_text(str(pick))


if __name__ == "__main__":
    main()
