"""world_size-2 `gloo` test of the row-sharded Hogbom protocol (include/radler_b200.h
"multi-GPU"): every rank searches and updates only its row band; the one exchange per minor
iteration is an all-gather of (64-bit peak key with the frame's row-major index, value at that
pixel). The band arithmetic is the CPU oracle's (bit-exact), the exchange and the merge are the
product's host logic (ska_sdp_func_radler_b200/distributed.py). The joined run must reproduce
the single-process oracle's peak sequence and band residuals bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, iters, gain, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist

    import oracle
    from helpers import hogbom_case
    from ska_sdp_func_radler_b200 import distributed as rd
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank,
                            world_size=world)
    oracle.set_threads(1)
    uid = rd.share_unique_id(lambda: bytes(range(128)), rank, world)
    assert uid == bytes(range(128))
    res, psfs, _, _ = hogbom_case(n, 1)
    img, psf = res[0].copy(), psfs[0]
    y0, y1 = rd.row_band(n, rank, world)
    trace = []
    for it in range(iters):
        v, x, y = oracle.find_peak(img, True, y0, y1, 0, 0)
        key = rd.make_peak_key(v, y * n + x) if v is not None else 0
        # all-gather of the candidates; the 64-bit key travels as two 32-bit halves
        mine = torch.tensor([key >> 32, key & 0xFFFFFFFF,
                             int(np.float32(v if v is not None else 0).view(np.int32))],
                            dtype=torch.int64)
        allc = [torch.zeros(3, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allc, mine)
        keys = [(int(c[0]) << 32) | int(c[1]) for c in allc]
        w = rd.merge_peak_keys(keys)
        if keys[w] == 0:
            break
        value = float(np.array([int(allc[w][2])], np.int64).astype(np.int32).view(np.float32)[0])
        index = (~(keys[w] >> 1)) & 0xFFFFFFFF
        px, py = index % n, index // n
        trace.append((px, py, value))
        oracle.partial_subtract_image(img, psf, px, py, np.float32(value) * np.float32(gain), y0, y1)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), trace=np.array(trace, np.float64),
             band=img[y0:y1], y0=y0, y1=y1)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_row_sharded_protocol_matches_single_process(tmp_path):
    import torch.multiprocessing as mp

    sys.path.insert(0, ROOT)
    import oracle
    from helpers import hogbom_case
    n, iters, gain, world = 96, 40, 0.1, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, iters, gain, str(tmp_path)), nprocs=world, join=True)
    res, psfs, _, _ = hogbom_case(n, 1)
    o_res, o_model = res.copy(), np.zeros_like(res)
    o = oracle.GenericClean(False).execute(
        oracle.Cube(n, n, 1, 1, None),
        oracle.Settings(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=gain,
                        record_trace=True), o_res, o_model, psfs)
    assert o.iteration_number == iters
    for r in range(world):
        g = np.load(tmp_path / f"rank{r}.npz")
        tr = g["trace"]
        assert len(tr) == iters
        assert np.array_equal(tr[:, 0], o.trace["x"]) and np.array_equal(tr[:, 1], o.trace["y"])
        assert np.array_equal(tr[:, 2].astype(np.float32), o.trace["value"])
        assert np.array_equal(g["band"], o_res[0, int(g["y0"]):int(g["y1"])])


def test_row_bands_partition_the_frame():
    from ska_sdp_func_radler_b200.distributed import facets_of_rank, row_band
    for h in (1, 7, 64, 8192, 16383):
        for world in (1, 2, 3, 8):
            bands = [row_band(h, r, world) for r in range(world)]
            assert bands[0][0] == 0 and bands[-1][1] == h
            assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
    assert sorted(sum((facets_of_rank(64, r, 8) for r in range(8)), [])) == list(range(64))
    from ska_sdp_func_radler_b200.distributed import cyclic_rows
    allrows = np.concatenate([cyclic_rows(256, r, 4, 16) for r in range(4)])
    assert sorted(allrows.tolist()) == list(range(256))
    assert cyclic_rows(256, 1, 4, 16)[:17].tolist() == list(range(16, 32)) + [80]


def test_peak_key_orders_like_the_reference_scan():
    from ska_sdp_func_radler_b200.distributed import make_peak_key, merge_peak_keys
    # larger |v| wins; equal |v| -> lower index wins (strict '>' scan, peak_finder.cc:39,117)
    assert make_peak_key(2.0, 50) > make_peak_key(-1.5, 3)
    assert make_peak_key(-2.0, 7) > make_peak_key(2.0, 9)
    assert merge_peak_keys([make_peak_key(1.0, 900), make_peak_key(1.0, 20), 0]) == 1
    assert make_peak_key(0.0, 1) == 0 and make_peak_key(float("nan"), 1) == 0
