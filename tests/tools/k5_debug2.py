"""Which of the three is the odd one out at one facet of the K5 test: restated oracle, reference
sources (oracle/_ref), device. Standalone GenericClean path A on the facet box with its mask."""
import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np, torch
import oracle, ska_sdp_func_radler_b200 as rdl
from helpers import hogbom_case
oracle.build()
n, c, grid = 1024, 1, (8, 8)
res, psfs, weights, _ = hogbom_case(n, c, seed=41)
alg = rdl.GenericClean(use_sub_minor_optimization=True)
alg.apply(max_iterations=0, clean_border_ratio=0.0)
d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, n, n), device="cuda")
alg.parallel_execute_major_iteration(8, 8, rdl.Cube(n, n, c, 1, weights), d_res, d_model, torch.from_numpy(psfs).cuda())
facets = alg.facets(with_masks=True)
for fi in (3, 11, 0, 27):
    f = facets[fi]
    x, y, w, h = f["x"], f["y"], f["width"], f["height"]
    sub = np.ascontiguousarray(res[:, y:y + h, x:x + w])
    lo_y, lo_x = (n - h) // 2, (n - w) // 2
    sub_psf = np.ascontiguousarray(psfs[:, lo_y:lo_y + h, lo_x:lo_x + w])
    mask = np.ascontiguousarray(f["boundary_mask"]).astype(np.uint8)
    kw = dict(max_iterations=60, minor_loop_gain=0.1, major_loop_gain=0.7, clean_border_ratio=0.0)
    out = {}
    # reference sources
    ref = oracle.Reference("generic", w, h, c, weights=weights, use_sub_minor_optimization=True, **kw)
    ref.set_clean_mask(mask)
    r, m = sub.copy(), np.zeros_like(sub)
    for call in range(2):
        ref.execute(r, m, sub_psf)
    out["ref"] = (r, m)
    # device
    g = rdl.GenericClean(use_sub_minor_optimization=True)
    g.apply(**kw)
    g.set_clean_mask(mask)
    r, m = sub.copy(), np.zeros_like(sub)
    for call in range(2):
        g.execute_major_iteration(rdl.Cube(w, h, c, 1, weights), r, m, sub_psf)
    out["gpu"] = (r, m)
    # restatement
    o = oracle.GenericClean(True)
    r, m = sub.copy(), np.zeros_like(sub)
    try:
        it = 0
        for call in range(2):
            st = oracle.Settings(max_iterations=60, minor_loop_gain=0.1, major_loop_gain=0.7,
                                 clean_border_ratio=0.0, iteration_number=it)
            st.clean_mask = mask
            res_o = o.execute(oracle.Cube(w, h, c, 1, weights), st, r, m, sub_psf)
            it = res_o.iteration_number
        out["orc"] = (r, m)
    except Exception as e:  # noqa: BLE001
        print("oracle restatement failed:", e)
    peak = np.abs(sub).max()
    names = list(out)
    for i in range(len(names)):
        for j in range(i + 1, len(names)):
            a, b = out[names[i]], out[names[j]]
            dr = np.abs(a[0] - b[0])[0]
            print("facet", fi, (w, h), names[i], "vs", names[j], "res max rel", dr.max() / peak,
                  "at", np.unravel_index(dr.argmax(), dr.shape), "n>1e-4:", int((dr > 1e-4 * peak).sum()),
                  "model max", np.abs(a[1] - b[1]).max())
