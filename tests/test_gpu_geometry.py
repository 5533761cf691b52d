"""Element-wise geometry kernels (csrc/geometry.cu) against the reference's torch ops (geometry.py:37-42, frame.py:41-105)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_constrain(img_size, Xs, K):  # geometry.py:37-42, :111-129 verbatim
    h, w = img_size
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")
    uv = torch.stack((u, v), dim=-1).unsqueeze(0).repeat(Xs.shape[0], 1, 1, 1).to(device=Xs.device, dtype=Xs.dtype)
    uv = uv.view(*Xs.shape[:-1], 2)
    z = Xs[..., 2:3]
    tmp1 = (uv[..., 0] - K[0, 2]) / K[0, 0]
    tmp2 = (uv[..., 1] - K[1, 2]) / K[1, 1]
    dP_dz = torch.empty(uv.shape[:-1] + (3, 1), device=z.device, dtype=K.dtype)
    dP_dz[..., 0, 0] = tmp1
    dP_dz[..., 1, 0] = tmp2
    dP_dz[..., 2, 0] = 1.0
    return torch.squeeze(z[..., None, :] * dP_dz, dim=-1)


@pytest.mark.parametrize("shape", [(1, 24, 32), (3, 48, 64), (2, 384, 512)])
def test_constrain_points_to_ray_bit_exact(shape):
    from mast3r_slam_ysh_b200 import geometry, synthetic as syn
    b, h, w = shape
    g = torch.Generator().manual_seed(b)
    Xs = (torch.randn(b, h * w, 3, generator=g) * 2.0).cuda()
    K = syn.intrinsics(h, w).float().cuda()
    out = geometry.constrain_points_to_ray((h, w), Xs, K)
    assert out.shape == Xs.shape and torch.equal(out, _ref_constrain((h, w), Xs, K))
    geometry.constrain_points_to_ray((h, w), Xs, K, out=Xs)  # in place
    assert torch.equal(Xs, out)


@pytest.mark.parametrize("mode", ["weighted_pointmap", "indep_conf"])
def test_update_pointmap_bit_exact(mode):
    from mast3r_slam_ysh_b200 import geometry
    g = torch.Generator().manual_seed(5)
    n = 384 * 512 + 7
    Xc, X = torch.randn(n, 3, generator=g).cuda(), torch.randn(n, 3, generator=g).cuda()
    Cc, C = (1.0 + 3.0 * torch.rand(n, 1, generator=g)).cuda(), (1.0 + 3.0 * torch.rand(n, 1, generator=g)).cuda()
    if mode == "weighted_pointmap":  # frame.py:73-76
        X_ref = ((Cc * Xc) + (C * X)) / (Cc + C)
        C_ref = Cc + C
    else:  # frame.py:68-72
        X_ref, C_ref = Xc.clone(), Cc.clone()
        new_mask = C > Cc
        X_ref[new_mask.repeat(1, 3)] = X[new_mask.repeat(1, 3)]
        C_ref[new_mask] = C[new_mask]
    geometry.update_pointmap(Xc, Cc, X, C, filtering_mode=mode)
    assert torch.equal(Xc, X_ref) and torch.equal(Cc, C_ref)


def test_update_pointmap_with_pose_matches_act_then_merge():
    """Fused Xkk = T_CkCf.act(Xkf) + weighted merge (tracker.py:96-97) vs the torch restatement of Sim3.act."""
    from mast3r_slam_ysh_b200 import geometry
    from oracle import tracker_torch as tt
    g = torch.Generator().manual_seed(6)
    n = 5000
    Xc, X = torch.randn(n, 3, generator=g).cuda(), torch.randn(n, 3, generator=g).cuda()
    Cc, C = (1.0 + torch.rand(n, 1, generator=g)).cuda(), (1.0 + torch.rand(n, 1, generator=g)).cuda()
    q = torch.nn.functional.normalize(torch.randn(4, generator=g), dim=0)
    T = torch.cat((torch.randn(3, generator=g) * 0.1, q, torch.tensor([1.05]))).view(1, 8).cuda()
    Xa = tt.Sim3(T).act(X)
    X_ref = ((Cc * Xc) + (C * Xa)) / (Cc + C)
    geometry.update_pointmap(Xc, Cc, X, C, T=T)
    assert torch.allclose(Xc, X_ref, rtol=0, atol=2e-6)
    with pytest.raises(NotImplementedError):
        geometry.update_pointmap(Xc, Cc, X, C, filtering_mode="best_score")  # a plain copy in the reference, no kernel


def test_update_pointmap_weighted_spherical_vs_reference_formulas():
    """filtering_mode 'weighted_spherical' (frame.py:76-101): the reference's own torch expressions on the same device."""
    from mast3r_slam_ysh_b200 import geometry
    g = torch.Generator().manual_seed(9)
    n = 384 * 512 + 3
    Xc = (torch.randn(n, 3, generator=g) + torch.tensor([0.0, 0.0, 3.0])).cuda()
    X = (Xc.cpu() + 0.02 * torch.randn(n, 3, generator=g)).cuda()
    Cc, C = (1.0 + 3.0 * torch.rand(n, 1, generator=g)).cuda(), (1.0 + 3.0 * torch.rand(n, 1, generator=g)).cuda()

    def cartesian_to_spherical(P):
        r = torch.linalg.norm(P, dim=-1, keepdim=True)
        x, y, z = torch.tensor_split(P, 3, dim=-1)
        return torch.cat((r, torch.atan2(y, x), torch.acos(z / r)), dim=-1)

    def spherical_to_cartesian(sph):
        r, phi, theta = torch.tensor_split(sph, 3, dim=-1)
        return torch.cat((r * torch.sin(theta) * torch.cos(phi), r * torch.sin(theta) * torch.sin(phi), r * torch.cos(theta)), dim=-1)

    sph = ((Cc * cartesian_to_spherical(Xc)) + (C * cartesian_to_spherical(X))) / (Cc + C)
    X_ref, C_ref = spherical_to_cartesian(sph), Cc + C
    geometry.update_pointmap(Xc, Cc, X, C, filtering_mode="weighted_spherical")
    assert torch.equal(Cc, C_ref)
    # acos(z / r) is ill-conditioned towards the poles (d acos = du / sqrt(1 - u^2)): one ulp of r (torch's norm reduction
    # vs the kernel's) moves theta by up to 1e-4 there, in the reference's own arithmetic as much as here
    err = (Xc - X_ref).abs().amax(-1)
    cos_t = (X_ref[:, 2] / X_ref.norm(dim=-1)).abs()
    scale = float(X_ref.abs().max())
    assert float(err[cos_t < 0.99].max()) <= 4e-6 * scale
    assert float(err.max()) <= 1e-3 * scale and float((err > 4e-6 * scale).float().mean()) < 1e-3


def test_keyframe_store_fuse_pointmap():
    from mast3r_slam_ysh_b200 import global_opt as go
    g = torch.Generator().manual_seed(8)
    s = go.KeyframeStore(100, capacity=2, device="cuda")
    X0, C0 = torch.randn(100, 3, generator=g).cuda(), (1 + torch.rand(100, 1, generator=g)).cuda()
    s.append(X0, C0, torch.tensor([0, 0, 0, 0, 0, 0, 1, 1.0]).cuda())
    X1, C1 = torch.randn(100, 3, generator=g).cuda(), (1 + torch.rand(100, 1, generator=g)).cuda()
    assert s.average_conf().data_ptr() == s.C.data_ptr()  # count 1 everywhere: a view, no copy
    s.fuse_pointmap(0, X1, C1)
    assert torch.equal(s.X[0], ((C0 * X0) + (C1 * X1)) / (C0 + C1)) and torch.equal(s.C[0], C0 + C1)
    # Frame.get_average_conf (frame.py:107-108): C / N with N = 2 after one weighted update
    assert s.update_count(0) == 2 and torch.equal(s.average_conf()[0], (C0 + C1) / 2)
    s.fuse_pointmap(0, X1, 10 * C1, filtering_mode="indep_conf")  # frame.py:68-72: N back to 1
    assert s.update_count(0) == 1 and torch.equal(s.average_conf()[0], s.C[0])
