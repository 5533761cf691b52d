"""GPU test of m3r_iter_proj_noalloc (include/m3r_backends.h): the allocation-free variant of m3r_iter_proj, called through
the C ABI, must reproduce the existing entry point on two calls in a row (different shapes, so contents left in the shared
scratch array by the first call would show in the second), must reject inputs beyond its fixed scratch with a clear error,
and must not take device memory.  Collected last on purpose (file name).

No counterpart in the reference."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _call(fn, rays, pts, p0, mt):
    b, h, w, _ = rays.shape
    n = p0.shape[1]
    p_new = torch.full((b, n, 2), -7.0, dtype=torch.float32, device=rays.device)
    conv = torch.zeros((b, n), dtype=torch.bool, device=rays.device)
    st = C.c_void_p(torch.cuda.current_stream(rays.device).cuda_stream)
    rc = fn(_ptr(rays), _ptr(pts), _ptr(p0), _ptr(p_new), _ptr(conv), b, h, w, n, mt["max_iter"], mt["lambda_init"],
            mt["convergence_thresh"], st)
    return rc, p_new, conv


def test_iter_proj_noalloc_matches_iter_proj_on_two_calls_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_ysh_b200 import _lib, synthetic as syn
    lib = _lib.load()
    mt = syn.MATCHING
    # first the largest input the fixed scratch holds (one 512x384 image), then a smaller batch of two images
    cases = [syn.make_match_problem(1, 384, 512, seed=1235, device=DEV),
             syn.make_match_problem(2, 96, 128, seed=77, device=DEV, warm_start=True)]
    assert 384 * 512 == 196608  # == M3R_ITER_PROJ_NOALLOC_MAX_PIXELS
    with torch.cuda.device(DEV):
        want = []
        for m in cases:  # the existing entry point (allocates stream-ordered scratch)
            rc, p, c = _call(lib.m3r_iter_proj, m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt)
            assert rc == 0
            want.append((p, c))
        torch.cuda.synchronize()
        free_before = None
        for rep in range(2):  # the whole sequence twice: the second round starts from a scratch full of old contents
            for k, m in enumerate(cases):
                if rep == 1 and k == 0:
                    torch.cuda.synchronize()
                    free_before = torch.cuda.mem_get_info()[0]  # after the first round: modules are loaded by now
                rc, p, c = _call(lib.m3r_iter_proj_noalloc, m["rays_img_with_grad"], m["pts_3d_norm"], m["p_init"], mt)
                assert rc == 0, lib.m3r_last_error_string()
                torch.cuda.synchronize()
                err = float((p - want[k][0]).abs().max())
                assert err <= 1e-3, (rep, k, err)            # the tolerance of the iter_proj parity tests (1e-3 px)
                assert torch.equal(c, want[k][1]), (rep, k)  # converged flags equal
                assert torch.equal(p, want[k][0]), (rep, k)  # same kernels on the same inputs: in fact bit-identical
                assert float(c.float().mean()) > 0.05 and float((p - m["p_init"]).abs().max()) > 1.0  # real work was done
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info()[0]
        assert free_after >= free_before, (free_before, free_after)  # the calls took no device memory
        # beyond the fixed scratch: rejected with a clear message, outputs untouched
        big = syn.make_match_problem(2, 384, 512, seed=3, device=DEV)
        rc, p, c = _call(lib.m3r_iter_proj_noalloc, big["rays_img_with_grad"], big["pts_3d_norm"], big["p_init"], mt)
        torch.cuda.synchronize()
        assert rc == -1 and b"exceed the fixed scratch" in lib.m3r_last_error_string()
        assert bool((p == -7.0).all()) and not bool(c.any())
        # empty input is a no-op, as for m3r_iter_proj
        m = cases[1]
        rc, p, c = _call(lib.m3r_iter_proj_noalloc, m["rays_img_with_grad"], m["pts_3d_norm"][:, :0].contiguous(),
                         m["p_init"][:, :0].contiguous(), mt)
        assert rc == 0 and p.shape == (2, 0, 2)
