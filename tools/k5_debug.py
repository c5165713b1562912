import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np, torch
import oracle, ska_sdp_func_radler_b200 as rdl
from helpers import hogbom_case
oracle.build()
n, c, grid = 1024, 1, (8, 8)
res, psfs, weights, _ = hogbom_case(n, c, seed=41)
kw = dict(max_iterations=400, minor_loop_gain=0.1, major_loop_gain=0.7, clean_border_ratio=0.0)
ref = oracle.Reference("generic", n, n, c, weights=weights, use_sub_minor_optimization=True, grid=grid, threads=1, **kw)
r_res, r_model = res.copy(), np.zeros_like(res)
alg = rdl.GenericClean(use_sub_minor_optimization=True)
alg.apply(**kw)
d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, n, n), device="cuda")
d_psf = torch.from_numpy(psfs).cuda()
cube = rdl.Cube(n, n, c, 1, weights)
peak = np.abs(res).max()
for call in range(2):
    w = ref.execute(r_res, r_model, psfs)
    g = alg.parallel_execute_major_iteration(grid[0], grid[1], cube, d_res, d_model, d_psf)
    dr = np.abs(d_res.cpu().numpy() - r_res)[0]
    dm = np.abs(d_model.cpu().numpy() - r_model)[0]
    print("call", call, "res max rel", dr.max() / peak, "at", np.unravel_index(dr.argmax(), dr.shape), "model max", dm.max(), "n model diff", np.count_nonzero(dm > 0))
    facets = alg.facets()
    for i, f in enumerate(facets):
        sub = dr[f["y"]:f["y"] + f["height"], f["x"]:f["x"] + f["width"]]
        if sub.max() / peak > 1e-5:
            print("  facet", i, f["x"], f["y"], f["width"], f["height"], "max rel", sub.max() / peak, "conv size", rdl.convolution_size(0.0, f["width"], 1.1) if hasattr(rdl, "convolution_size") else "")
    if call == 1:
        y0, x0 = np.unravel_index(dr.argmax(), dr.shape)
        np.set_printoptions(precision=5, linewidth=200, suppress=True)
        print("diff neighbourhood (gpu - ref):")
        print((d_res.cpu().numpy()[0] - r_res[0])[y0 - 4:y0 + 5, x0 - 4:x0 + 5])
        print("ref model neighbourhood:")
        print(r_model[0][y0 - 4:y0 + 5, x0 - 4:x0 + 5])
        print("gpu model neighbourhood:")
        print(d_model.cpu().numpy()[0][y0 - 4:y0 + 5, x0 - 4:x0 + 5])
        big = np.argwhere(dr > 1e-4 * peak)
        print("pixels with diff > 1e-4 peak:", len(big), "bbox", big.min(0), big.max(0))
        for i, f in enumerate(alg.facets(with_masks=True)):
            if f["y"] <= y0 < f["y"] + f["height"] and f["x"] <= x0 < f["x"] + f["width"]:
                own = f["boundary_mask"][y0 - f["y"], x0 - f["x"]] if "boundary_mask" in f else None
                print("  in facet", i, "owned:", own, "iterations", f.get("iteration_number"), f.get("peak"))
