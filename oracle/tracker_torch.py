"""TEST/BASELINE INFRASTRUCTURE ONLY -- the reference's CPU path for BASELINE.json configs[0].

Pure-torch restatement of FrameTracker.opt_pose_ray_dist_sim3 + solve (reference: mast3r_slam/tracker.py:173-214,
:156-171) with geometry.act_Sim3 / point_to_ray_dist (geometry.py:17-52), nonlinear_optimizer.huber /
check_convergence (nonlinear_optimizer.py:5-33).  The reference imports `lietorch` for Sim(3) objects; lietorch is an
un-pinned, un-vendored pip dependency (pyproject.toml:15) absent from this image, so `Sim3` below is a minimal
stand-in with the same `.data` layout [t(3), q_xyzw(4), s(1)] and the operations the tracker uses (act, inv, *, retr =
Exp(xi) o T with tangent order tau, omega, sigma -- consistent with gn_kernels.cu:392-413).  Results at that boundary
are therefore "parity unpinned" by the reference; this file is a REPORTED CPU BASELINE, not a parity oracle.
"""
import math

import torch

TRACKING = dict(max_iters=50, rel_error=1e-3, delta_norm=1e-3, huber=1.345, sigma_ray=0.003, sigma_dist=10.0,
                sigma_pixel=1.0, sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6)  # config/base.yaml:16-33


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack((aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz), -1)


def _qrot(q, X):
    qv = q[..., :3].expand_as(X)
    uv = 2.0 * torch.cross(qv, X, dim=-1)
    return X + q[..., 3:4] * uv + torch.cross(qv, uv, dim=-1)


class Sim3:
    def __init__(self, data):
        self.data = data  # (...,8)

    def act(self, X):
        return self.data[..., 7:8] * _qrot(self.data[..., 3:7], X) + self.data[..., 0:3]

    def inv(self):
        d = self.data
        qi = torch.cat((-d[..., 3:6], d[..., 6:7]), -1)
        si = 1.0 / d[..., 7:8]
        return Sim3(torch.cat((-si * _qrot(qi, d[..., 0:3]), qi, si), -1))

    def __mul__(self, o):
        a, b = self.data, o.data
        t = a[..., 7:8] * _qrot(a[..., 3:7], b[..., 0:3]) + a[..., 0:3]
        return Sim3(torch.cat((t, _qmul(a[..., 3:7], b[..., 3:7]), a[..., 7:8] * b[..., 7:8]), -1))

    @staticmethod
    def exp(xi):
        tau, phi, sigma = xi[..., 0:3], xi[..., 3:6], xi[..., 6]
        out = []
        for k in range(xi.shape[0]):  # tiny (one pose): scalar branches as in expSim3 (gn_kernels.cu:323-390)
            ph, sg, ta = phi[k], float(sigma[k]), tau[k]
            th2 = float((ph * ph).sum())
            th = math.sqrt(th2)
            sc = math.exp(sg)
            if th2 < 1e-6:
                imag, real = 0.5 - th2 / 48.0 + th2 * th2 / 3840.0, 1.0 - th2 / 8.0 + th2 * th2 / 384.0
            else:
                imag, real = math.sin(0.5 * th) / th, math.cos(0.5 * th)
            if abs(sg) < 1e-6:
                Cc = 1.0
                A, B = (0.5, 1.0 / 6.0) if th < 1e-6 else ((1 - math.cos(th)) / th2, (th - math.sin(th)) / (th2 * th))
            else:
                Cc = (sc - 1.0) / sg
                if th < 1e-6:
                    A = ((sg - 1) * sc + 1) / (sg * sg)
                    B = (sc * 0.5 * sg * sg + sc - 1 - sg * sc) / (sg ** 3)
                else:
                    a, b, c = sc * math.sin(th), sc * math.cos(th), th2 + sg * sg
                    A = (a * sg + (1 - b) * th) / (th * c)
                    B = (Cc - ((b - 1) * sg + a * th) / c) / th2
            cr = torch.linalg.cross(ph, ta)
            t = Cc * ta + A * cr + B * torch.linalg.cross(ph, cr)
            out.append(torch.cat((t, imag * ph, torch.tensor([real, sc], dtype=xi.dtype, device=xi.device))))
        return Sim3(torch.stack(out))

    def retr(self, xi):
        return Sim3.exp(xi) * self


def skew_sym(x):
    b = x.shape[:-1]
    x, y, z = x.unbind(dim=-1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).view(*b, 3, 3)


def point_to_ray_dist(X, jacobian=False):  # geometry.py:17-34
    b = X.shape[:-1]
    d = torch.linalg.norm(X, dim=-1, keepdim=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = torch.cat((r, d), dim=-1)
    if not jacobian:
        return rd
    d_inv_2 = d_inv ** 2
    I = torch.eye(3, device=X.device, dtype=X.dtype).repeat(*b, 1, 1)
    dr_dX = d_inv.unsqueeze(-1) * (I - d_inv_2.unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
    return rd, torch.cat((dr_dX, r.unsqueeze(-2)), dim=-2)


def act_Sim3(X, pC, jacobian=False):  # geometry.py:45-52
    pW = X.act(pC)
    if not jacobian:
        return pW
    dpC_dt = torch.eye(3, device=pW.device).repeat(*pW.shape[:-1], 1, 1)
    return pW, torch.cat([dpC_dt, -skew_sym(pW), pW.reshape(*pW.shape[:-1], -1, 1)], dim=-1)


def huber(r, k=1.345):  # nonlinear_optimizer.py:28-33
    unit = torch.ones((1), dtype=r.dtype, device=r.device)
    r_abs = torch.abs(r)
    return torch.where(r_abs < k, unit, k / r_abs)


def check_convergence(rel_error_threshold, delta_norm_threshold, old_cost, new_cost, delta):
    """nonlinear_optimizer.py:5-25 (without the print).  On the first iteration old_cost is inf (tracker.py:183):
    (inf - c) / inf is NaN and NaN < threshold is False, so only the step norm can stop the loop there."""
    cost_diff = old_cost - new_cost
    rel_dec = math.fabs(cost_diff / old_cost)
    delta_norm = torch.linalg.norm(delta)
    return bool(rel_dec < rel_error_threshold or delta_norm < delta_norm_threshold)


def solve(sqrt_info, r, J, cfg):  # tracker.py:156-171
    whitened_r = sqrt_info * r
    robust_sqrt_info = sqrt_info * torch.sqrt(huber(whitened_r, k=cfg["huber"]))
    mdim = J.shape[-1]
    A = (robust_sqrt_info[..., None] * J).view(-1, mdim)
    b = (robust_sqrt_info * r).view(-1, 1)
    H = A.T @ A
    g = -A.T @ b
    cost = 0.5 * (b.T @ b).item()
    L = torch.linalg.cholesky(H, upper=False)
    return torch.cholesky_solve(g, L, upper=False).view(1, -1), cost


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=TRACKING, max_iters=None):
    """tracker.py:173-214.  Returns (T_WCf, T_CkCf, iterations executed)."""
    sqrt_info_ray = 1 / cfg["sigma_ray"] * valid * torch.sqrt(Qk)
    sqrt_info_dist = 1 / cfg["sigma_dist"] * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_ray.repeat(1, 3), sqrt_info_dist), dim=1)
    T_CkCf = T_WCk.inv() * T_WCf
    rd_k = point_to_ray_dist(Xk, jacobian=False)
    old_cost = float("inf")
    iters = 0
    for step in range(max_iters or cfg["max_iters"]):
        Xf_Ck, dXf_Ck_dT = act_Sim3(T_CkCf, Xf, jacobian=True)
        rd_f_Ck, drd_dX = point_to_ray_dist(Xf_Ck, jacobian=True)
        r = rd_k - rd_f_Ck
        J = -drd_dX @ dXf_Ck_dT
        tau, new_cost = solve(sqrt_info, r, J, cfg)
        T_CkCf = T_CkCf.retr(tau)
        iters += 1
        if check_convergence(cfg["rel_error"], cfg["delta_norm"], old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return T_WCk * T_CkCf, T_CkCf, iters


def project_calib(P, K, img_size, jacobian=False, border=0, z_eps=0.0):  # geometry.py:63-108
    b = P.shape[:-1]
    K_rep = K.repeat(*b, 1, 1)
    p = (K_rep @ P[..., None]).squeeze(-1)
    p = p / p[..., 2:3]
    p = p[..., :2]
    u, v = p.split([1, 1], dim=-1)
    x, y, z = P.split([1, 1, 1], dim=-1)
    valid_u = (u > border) & (u < img_size[1] - 1 - border)
    valid_v = (v > border) & (v < img_size[0] - 1 - border)
    valid_z = z > z_eps
    valid = valid_u & valid_v & valid_z
    logz = torch.log(z)
    logz[torch.logical_not(valid_z)] = 0.0
    pz = torch.cat((p, logz), dim=-1)
    if not jacobian:
        return pz, valid
    fx, fy = K[0, 0], K[1, 1]
    z_inv = 1.0 / z[..., 0]
    dpz_dP = torch.zeros(*b + (3, 3), device=P.device, dtype=P.dtype)
    dpz_dP[..., 0, 0] = fx
    dpz_dP[..., 1, 1] = fy
    dpz_dP[..., 0, 2] = -fx * x[..., 0] * z_inv
    dpz_dP[..., 1, 2] = -fy * y[..., 0] * z_inv
    dpz_dP *= z_inv[..., None, None]
    dpz_dP[..., 2, 2] = z_inv
    return pz, dpz_dP, valid


def make_meas(Xk, img_size, depth_eps=1e-6):
    """meas_k / valid_meas_k as FrameTracker.get_points_poses builds them (tracker.py:141-150); Xk (h*w,3) is the
    keyframe pointmap on the identity pixel grid (already constrained to its rays)."""
    h, w = img_size
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")  # geometry.py:124-129
    uv_k = torch.stack((u, v), dim=-1).to(dtype=Xk.dtype).view(-1, 2)
    meas_k = torch.cat((uv_k, torch.log(Xk[..., 2:3])), dim=-1)
    valid_meas_k = Xk[..., 2:3] > depth_eps
    meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
    return meas_k, valid_meas_k


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, cfg=TRACKING,
                        max_iters=None):
    """tracker.py:216-266.  Returns (T_WCf, T_CkCf, iterations executed)."""
    sqrt_info_pixel = 1 / cfg["sigma_pixel"] * valid * torch.sqrt(Qk)
    sqrt_info_depth = 1 / cfg["sigma_depth"] * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_pixel.repeat(1, 2), sqrt_info_depth), dim=1)
    T_CkCf = T_WCk.inv() * T_WCf
    old_cost = float("inf")
    iters = 0
    for step in range(max_iters or cfg["max_iters"]):
        Xf_Ck, dXf_Ck_dT = act_Sim3(T_CkCf, Xf, jacobian=True)
        pzf_Ck, dpz_dX, valid_proj = project_calib(Xf_Ck, K, img_size, jacobian=True, border=cfg["pixel_border"],
                                                   z_eps=cfg["depth_eps"])
        valid2 = valid_proj & valid_meas_k
        sqrt_info2 = valid2 * sqrt_info
        r = meas_k - pzf_Ck
        J = -dpz_dX @ dXf_Ck_dT
        tau, new_cost = solve(sqrt_info2, r, J, cfg)
        T_CkCf = T_CkCf.retr(tau)
        iters += 1
        if check_convergence(cfg["rel_error"], cfg["delta_norm"], old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return T_WCk * T_CkCf, T_CkCf, iters
