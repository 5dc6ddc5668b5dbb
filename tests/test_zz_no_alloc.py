"""kd_build_gpu_noalloc: the build for callers that cannot allocate device memory.

Some users of this library run where cudaMalloc is off limits — inside another framework's fixed memory
plan, or in a service that gave the whole device to a pool at start-up — so they cannot hand kd_build_gpu a
workspace.  The no-alloc entry point keeps its scratch in a 16 MiB array inside the library.  This test
drives it through the C ABI (ctypes, raw pointers) and requires the same tree as the regular entry point,
bit for bit, on two calls in a row: the second call finds the first call's leftovers in the shared scratch
and must not be disturbed by them.  It also checks that a problem too large for the scratch is refused.

The file name sorts last on purpose: a `pytest -x` run reaches it only after every other test has run.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

SCRATCH_LIMIT = 16 << 20


def _regular(lib, pts):
    n, dim = pts.shape
    perm = torch.empty(n, dtype=torch.int32, device=pts.device)
    tp = torch.empty_like(pts)
    ws = torch.empty(lib.kd_build_workspace_bytes(n, dim), dtype=torch.uint8, device=pts.device)
    rc = lib.kd_build_gpu(pts.data_ptr(), n, dim, perm.data_ptr(), tp.data_ptr(), ws.data_ptr(), ws.numel(),
                          torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.kd_last_error()
    return perm, tp


def _noalloc(lib, pts):
    n, dim = pts.shape
    perm = torch.full((n,), -1, dtype=torch.int32, device=pts.device)   # outputs belong to the caller
    tp = torch.zeros_like(pts)
    rc = lib.kd_build_gpu_noalloc(pts.data_ptr(), n, dim, perm.data_ptr(), tp.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.kd_last_error()
    # the device-side error word lives in the built-in scratch; this is the only way for such a caller to read it
    rc = lib.kd_build_status_noalloc(n, dim, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.kd_last_error()
    return perm, tp


def test_noalloc_build_matches_regular_entry_point_twice_in_a_row(kd):
    lib = kd.lib()
    torch.cuda.set_device(0)
    # two different problems back to back: the second one runs on top of the first one's scratch contents
    problems = [kd.ops.generate_uniform(5, 3, 300000), kd.ops.generate_uniform(6, 2, 70001)]
    # quantised coordinates: many equal keys, so the deep digit passes and their histograms are exercised too
    problems.append(torch.randint(-2, 3, (150000, 3), generator=torch.Generator().manual_seed(1)).float())
    for host in problems:
        pts = host.cuda()
        assert lib.kd_build_workspace_bytes(*pts.shape) <= SCRATCH_LIMIT
        got_perm, got_tp = _noalloc(lib, pts)
        again_perm, again_tp = _noalloc(lib, pts)        # same input, second call in a row
        ref_perm, ref_tp = _regular(lib, pts)
        torch.cuda.synchronize()
        assert torch.equal(got_perm, ref_perm) and torch.equal(got_tp, ref_tp)
        assert torch.equal(again_perm, ref_perm) and torch.equal(again_tp, ref_tp)
        assert torch.equal(ref_perm.cpu(), kd.ops.build(host)[0])   # and both are the reference-rule tree


def test_noalloc_refuses_problems_beyond_its_scratch(kd):
    lib = kd.lib()
    n, dim = 1 << 22, 3
    assert lib.kd_build_workspace_bytes(n, dim) > SCRATCH_LIMIT
    pts = torch.zeros((n, dim), device="cuda")
    perm = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    rc = lib.kd_build_gpu_noalloc(pts.data_ptr(), n, dim, perm.data_ptr(), None,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.kd_last_error().decode()
    assert "16777216" in msg and "kd_build_gpu" in msg
    torch.cuda.synchronize()
    assert int(perm.min()) == -7 and int(perm.max()) == -7          # nothing was launched
