"""GPU tests of the row-sharded GenericClean path (rdl_dist_*): single-GPU band runs against the
oracle, and — when the box has >= 2 GPUs — the NCCL joined arg-max under torchrun."""
import os
import subprocess
import sys

import numpy as np
import pytest

from helpers import hogbom_case

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_single_rank_whole_frame_band_is_bit_exact(rdl, oracle, torch_cuda):
    torch = torch_cuda
    script = os.path.join(HERE, "dist_hogbom_check.py")
    out = subprocess.run([sys.executable, script], capture_output=True, text=True, timeout=600)
    assert "DIST-OK 1" in out.stdout, out.stdout + out.stderr


def test_sub_band_matches_masked_oracle(rdl, oracle, torch_cuda):
    """A rank that owns rows [y0, y1) alone behaves like the reference with a clean mask of
    those rows: same peak sequence, same residual and model inside the band."""
    torch = torch_cuda
    n, c, iters, y0, y1 = 160, 2, 80, 37, 121
    res, psfs, weights, _ = hogbom_case(n, c)
    mask = np.zeros((n, n), np.uint8)
    mask[y0:y1] = 1
    o_res, o_model = res.copy(), np.zeros_like(res)
    o = oracle.GenericClean(False).execute(
        oracle.Cube(n, n, c, 1, weights),
        oracle.Settings(max_iterations=iters, clean_border_ratio=0.0, record_trace=True,
                        clean_mask=mask), o_res, o_model, psfs)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    alg.apply(max_iterations=iters, clean_border_ratio=0.0, trace_capacity=iters)
    alg.dist_init(bytes(128), 0, 1)
    alg.dist_set_row_band(y0, n)
    d_res = torch.from_numpy(np.ascontiguousarray(res[:, y0:y1])).cuda()
    d_model = torch.zeros_like(d_res)
    g = alg.execute_major_iteration(rdl.Cube(n, y1 - y0, c, 1, weights), d_res, d_model,
                                    torch.from_numpy(psfs).cuda())
    tr = alg.trace()
    assert g.iteration_number == o.iteration_number == iters
    assert np.array_equal(tr["x"], o.trace["x"]) and np.array_equal(tr["y"], o.trace["y"])
    assert np.array_equal(d_res.cpu().numpy(), o_res[:, y0:y1])
    assert np.array_equal(d_model.cpu().numpy(), o_model[:, y0:y1])
    # back to whole frames on the same handle
    alg.dist_set_row_band(0, 0)
    alg.set_iteration_number(0)
    f_res, f_model = torch.from_numpy(res).cuda(), torch.zeros((c, n, n), device="cuda")
    f = alg.execute_major_iteration(rdl.Cube(n, n, c, 1, weights), f_res, f_model,
                                    torch.from_numpy(psfs).cuda())
    assert f.iteration_number == iters


def test_sharded_run_rejects_unsupported_paths(rdl, torch_cuda):
    torch = torch_cuda
    alg = rdl.GenericClean(use_sub_minor_optimization=True)
    alg.dist_set_row_band(0, 64)
    z = torch.zeros((1, 32, 64), device="cuda")
    with pytest.raises(rdl.RadlerError):
        alg.execute_major_iteration(rdl.Cube(64, 32), z, z.clone(), torch.zeros((1, 64, 64), device="cuda"))


def test_two_ranks_nccl_joined_argmax(torch_cuda):
    torch = torch_cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run tests/dist_hogbom_check.py under torchrun with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                          "29517", os.path.join(HERE, "dist_hogbom_check.py")],
                         capture_output=True, text=True, timeout=900)
    assert "DIST-OK 2" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_facets_single_rank_dist_context_is_identical(torch_cuda):
    """world = 1: the facet-sharded code path (exchange + bookkeeping) must not change anything."""
    out = subprocess.run([sys.executable, os.path.join(HERE, "dist_facets_check.py")],
                         capture_output=True, text=True, timeout=600)
    assert "FACETS-OK 1" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_two_ranks_facets_sharded(torch_cuda):
    torch = torch_cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run tests/dist_facets_check.py under torchrun with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                          "29519", os.path.join(HERE, "dist_facets_check.py")],
                         capture_output=True, text=True, timeout=900)
    assert "FACETS-OK 2" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_channel_sharded_multiscale_single_rank(torch_cuda):
    """world = 1 with a distributed context: the channel-sharded code path must change nothing."""
    out = subprocess.run([sys.executable, os.path.join(HERE, "dist_multiscale_check.py")],
                         capture_output=True, text=True, timeout=900)
    assert "MS-DIST-OK 1" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_two_ranks_channel_sharded_multiscale(torch_cuda):
    torch = torch_cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run tests/dist_multiscale_check.py under torchrun with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                          "29521", os.path.join(HERE, "dist_multiscale_check.py")],
                         capture_output=True, text=True, timeout=900)
    assert "MS-DIST-OK 2" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
