"""Synthetic workloads for BASELINE.json's configs.

Closed-form geometry so ground-truth matches exist without the MASt3R network: cameras inside an axis-aligned
6 x 4 x 3 m room, pinhole K = TUM fr1 intrinsics scaled to 512x384 (reference: config/intrinsics.yaml,
dataloader.py:289-293), pointmaps by ray/box intersection + 5 mm noise, per-frame Sim(3) scale gauge,
confidences C = 1 + 3 Exp(1), match quality Q = 3 Exp(1), 24-dim unit descriptors from a world-space sinusoid
field.  Shapes/dtypes are exactly what the reference callers hand to `mast3r_slam_backends`
(global_opt.py:104-155, matching.py:25-85).

This is input generation only (plain torch); nothing here is on the measured path.
"""
import math

import torch

ROOM = (6.0, 4.0, 3.0)
H_FULL, W_FULL = 384, 512
K_FULL = ((413.84, 0.0, 254.9), (0.0, 413.2, 204.2), (0.0, 0.0, 1.0))

# local_opt / matching parameters of the reference (config/base.yaml:8-14,35-50)
MATCHING = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=0.1, radius=3, dilation_max=5)
LOCAL_OPT = dict(C_conf=0.0, Q_conf=1.5, max_iters=10, sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0,
                 sigma_depth=10.0, sigma_point=0.05, pixel_border=-10, depth_eps=1e-6, delta_norm=1e-8)


def intrinsics(h, w, dtype=torch.float64, device="cpu"):
    K = torch.tensor(K_FULL, dtype=dtype, device=device).clone()
    K[0] *= w / W_FULL
    K[1] *= h / H_FULL
    return K


# ----------------------------------------------------------------------------------------------- quaternions
def quat_mul(a, b):  # xyzw
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack((aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz), -1)


def quat_rotate(q, X):
    qv = q[..., :3]
    uv = 2.0 * torch.cross(qv.expand_as(X), X, dim=-1)
    return X + q[..., 3:4] * uv + torch.cross(qv.expand_as(X), uv, dim=-1)


def quat_inv(q):
    return torch.cat((-q[..., :3], q[..., 3:4]), -1)


def quat_from_yaw_pitch(yaw, pitch):
    """camera looks along +z; world up is +z_world; camera x right, y down."""
    # rotation taking camera axes to world: columns = (right, down, forward)
    cy, sy, cp, sp = math.cos(yaw), math.sin(yaw), math.cos(pitch), math.sin(pitch)
    fwd = torch.tensor([cy * cp, sy * cp, sp], dtype=torch.float64)
    up = torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64)
    right = torch.linalg.cross(fwd, up)
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    R = torch.stack((right, down, fwd), dim=1)
    return rotmat_to_quat(R)


def rotmat_to_quat(R):
    t = R[0, 0] + R[1, 1] + R[2, 2]
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = [(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s, (R[2, 1] - R[1, 2]) / s]
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s, (R[0, 2] - R[2, 0]) / s]
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s, (R[1, 0] - R[0, 1]) / s]
    return torch.tensor([float(x) for x in q], dtype=torch.float64)


def sim3_act(T, X):
    """T (...,8) = [t, q_xyzw, s]; X (...,3)."""
    return T[..., 7:8] * quat_rotate(T[..., 3:7], X) + T[..., 0:3]


def sim3_inv(T):
    qi = quat_inv(T[..., 3:7])
    si = 1.0 / T[..., 7:8]
    ti = -si * quat_rotate(qi, T[..., 0:3])
    return torch.cat((ti, qi, si), -1)


def sim3_mul(A, B):
    t = A[..., 7:8] * quat_rotate(A[..., 3:7], B[..., 0:3]) + A[..., 0:3]
    return torch.cat((t, quat_mul(A[..., 3:7], B[..., 3:7]), A[..., 7:8] * B[..., 7:8]), -1)


def sim3_exp_small(xi):
    """First-order-accurate Exp for perturbing initial poses (exactness irrelevant: it only makes inputs)."""
    tau, phi, sig = xi[..., 0:3], xi[..., 3:6], xi[..., 6:7]
    th = phi.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    q = torch.cat((torch.sin(0.5 * th) * phi / th, torch.cos(0.5 * th)), -1)
    return torch.cat((tau, q, torch.exp(sig)), -1)


# ----------------------------------------------------------------------------------------------- scene
def _ray_box_depth(origin, dirs_w):
    """distance along unit-less dirs from an origin strictly inside the room to the first wall."""
    half = torch.tensor([ROOM[0] / 2, ROOM[1] / 2, ROOM[2] / 2], dtype=dirs_w.dtype, device=dirs_w.device)
    o = origin.to(dirs_w.device)
    t_pos = (half - o) / dirs_w
    t_neg = (-half - o) / dirs_w
    t = torch.where(dirs_w > 0, t_pos, t_neg)
    t = torch.where(dirs_w.abs() < 1e-12, torch.full_like(t, float("inf")), t)
    return t.min(dim=-1).values


def camera_poses(num_kf, yaw_step_deg=None, seed=0):
    """GT poses (num_kf,8) float64: circle r=1 m, looking outward, scale gauge s ~ U(0.9,1.1)."""
    g = torch.Generator().manual_seed(seed)
    step = (360.0 / num_kf) if yaw_step_deg is None else yaw_step_deg
    s = 0.9 + 0.2 * torch.rand(num_kf, generator=g, dtype=torch.float64)
    T = torch.zeros(num_kf, 8, dtype=torch.float64)
    for c in range(num_kf):
        yaw = math.radians(step * c)
        T[c, 0:3] = torch.tensor([math.cos(yaw), math.sin(yaw), 0.1 * math.sin(3 * yaw)], dtype=torch.float64)
        T[c, 3:7] = quat_from_yaw_pitch(yaw, 0.05 * math.cos(2 * yaw))
        T[c, 7] = s[c]
    return T


def pointmap(T_gt, h, w, noise=0.005, seed=0, device="cpu"):
    """Canonical pointmap of one camera in ITS OWN frame divided by its scale gauge: (h*w,3) float64."""
    K = intrinsics(h, w, device=device)
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=device),
                          torch.arange(w, dtype=torch.float64, device=device), indexing="ij")
    d = torch.stack(((u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)), -1).view(-1, 3)
    T = T_gt.to(device)
    dw = quat_rotate(T[3:7], d)
    lam = _ray_box_depth(T[0:3], dw)
    X = lam[:, None] * d
    if noise > 0:
        g = torch.Generator(device=device).manual_seed(seed)
        X = X + noise * torch.randn(X.shape, generator=g, dtype=torch.float64, device=device)
    return X / T[7]


def descriptor_field(Xw, fdim=24, seed=7, noise=0.05):
    g = torch.Generator().manual_seed(seed)
    Wm = (2 * math.pi / 0.3) * torch.randn(fdim, 3, generator=g, dtype=torch.float64)
    ph = 2 * math.pi * torch.rand(fdim, generator=g, dtype=torch.float64)
    D = torch.sin(Xw @ Wm.to(Xw.device).T + ph.to(Xw.device))
    if noise > 0:
        g2 = torch.Generator(device=Xw.device).manual_seed(seed + 1)
        D = D + noise * torch.randn(D.shape, generator=g2, dtype=torch.float64, device=Xw.device)
    return torch.nn.functional.normalize(D, dim=-1)


# ----------------------------------------------------------------------------------------------- GN graphs
def ring_edges(num_kf, span, wrap=True):
    """undirected edges (i<j order not required): each kf linked to its next `span` kfs."""
    ii, jj = [], []
    for i in range(num_kf):
        for d in range(1, span + 1):
            j = i + d
            if j >= num_kf:
                if not wrap:
                    continue
                j -= num_kf
            if j == i:
                continue
            ii.append(i)
            jj.append(j)
    return ii, jj


def loop_closure_edges(num_kf, n_loop=3, seed=0):
    """consecutive + n_loop pseudo-random long-range edges per kf (configs[4])."""
    g = torch.Generator().manual_seed(seed)
    ii, jj = [], []
    for i in range(num_kf):
        ii.append(i)
        jj.append((i + 1) % num_kf)
        offs = torch.randint(2, max(3, num_kf // 8), (n_loop,), generator=g).tolist()
        for o in offs:
            ii.append(i)
            jj.append((i + o) % num_kf)
    return ii, jj


def make_gn_problem(num_kf, h, w, edges, seed=1236, device="cpu", calib=False, pose_noise=(0.02, 0.02, 0.01),
                    kf_ids=None, yaw_step_deg=None, point_noise=0.005, directed_subset=None):
    """Inputs of gauss_newton_rays/calib/points exactly as FactorGraph hands them over (global_opt.py:121-155).

    Returns dict with Twc (num_kf,8) f32 [perturbed], Twc_gt, Xs (num_kf,N,3), Cs (num_kf,N,1), ii, jj (E) int64
    [global kf ids], idx_ii2jj (E,N) int64, valid_match (E,N,1) bool, Q (E,N,1) f32, K (3,3) f32.
    E = 2 * len(edges) (both directions, prep_two_way_edges global_opt.py:104-110).
    directed_subset (1-D LongTensor of indices into the two-way edge list) restricts ii/jj/idx/valid/Q to one rank's
    shard; all_ii/all_jj always hold the whole graph.
    """
    N = h * w
    g = torch.Generator().manual_seed(seed)
    T_gt = camera_poses(num_kf, yaw_step_deg=yaw_step_deg, seed=seed)
    Kd = intrinsics(h, w, device=device)
    Xs = torch.stack([pointmap(T_gt[c], h, w, noise=point_noise, seed=seed * 1000 + c, device=device) for c in range(num_kf)])
    if calib:  # constrain_points_to_ray (geometry.py:37-42): X = z * K^-1 [u,v,1]
        v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=device),
                              torch.arange(w, dtype=torch.float64, device=device), indexing="ij")
        d = torch.stack(((u - Kd[0, 2]) / Kd[0, 0], (v - Kd[1, 2]) / Kd[1, 1], torch.ones_like(u)), -1).view(1, N, 3)
        Xs = Xs[..., 2:3] * d
    gd = torch.Generator(device=device).manual_seed(seed + 17)
    Cs = 1.0 + 3.0 * torch.empty(num_kf, N, 1, dtype=torch.float64, device=device).exponential_(1.0, generator=gd)
    e_i, e_j = edges
    ii = torch.tensor(list(e_i) + list(e_j), dtype=torch.int64)
    jj = torch.tensor(list(e_j) + list(e_i), dtype=torch.int64)
    all_ii, all_jj = ii, jj
    if directed_subset is not None:
        ii, jj = ii[directed_subset], jj[directed_subset]
    E = ii.numel()
    idx = torch.zeros(E, N, dtype=torch.int64, device=device)
    valid = torch.zeros(E, N, 1, dtype=torch.bool, device=device)
    Tg = T_gt.to(device)
    for e in range(E):
        i, j = int(ii[e]), int(jj[e])
        # pixel k of frame j -> world -> camera i (metric) -> pixel in i
        Pw = sim3_act(Tg[j], Xs[j])
        Pi = sim3_act(sim3_inv(Tg[i]), Pw)  # in i's canonical units
        z = Pi[:, 2]
        uu = Kd[0, 0] * Pi[:, 0] / z + Kd[0, 2]
        vv = Kd[1, 1] * Pi[:, 1] / z + Kd[1, 2]
        ur, vr = torch.round(uu).long(), torch.round(vv).long()
        inside = (z > 1e-3) & (ur >= 0) & (ur < w) & (vr >= 0) & (vr < h)
        lin = torch.where(inside, vr * w + ur, torch.zeros_like(ur))
        dist = (Xs[i][lin] - Pi).norm(dim=-1) * Tg[i, 7]
        ok = inside & (dist < 0.1)
        idx[e] = torch.where(ok, lin, torch.zeros_like(lin))
        valid[e, :, 0] = ok
    Q = 3.0 * torch.empty(E, N, 1, dtype=torch.float64, device=device).exponential_(1.0, generator=gd)
    xi = torch.randn(num_kf, 7, generator=g, dtype=torch.float64)
    xi[:, 0:3] *= pose_noise[0]
    xi[:, 3:6] *= pose_noise[1]
    xi[:, 6] *= pose_noise[2]
    xi[0] = 0  # the first (fixed) keyframe defines the gauge
    T_init = sim3_mul(sim3_exp_small(xi), T_gt)
    if kf_ids is not None:  # non-contiguous global keyframe ids exercise the unique/searchsorted prep
        ids = torch.tensor(kf_ids, dtype=torch.int64)
        ii, jj, all_ii, all_jj = ids[ii], ids[jj], ids[all_ii], ids[all_jj]
    return dict(Twc=T_init.float().to(device).contiguous(), Twc_gt=T_gt.float().to(device), Xs=Xs.float().contiguous(),
                Cs=Cs.float().contiguous(), ii=ii.to(device), jj=jj.to(device), idx_ii2jj=idx.contiguous(),
                valid_match=valid.contiguous(), Q=Q.float().contiguous(), K=Kd.float().contiguous(), h=h, w=w,
                all_ii=all_ii.to(device), all_jj=all_jj.to(device))


def make_track_problem(h, w, seed=1234, device="cpu", calib=False, yaw_step_deg=8.0, pose_noise=(0.02, 0.02, 0.01),
                       point_noise=0.005, C_conf=0.0, Q_conf=1.5, depth_eps=1e-6):
    """Inputs of FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 as FrameTracker.track prepares them
    (tracker.py:27-64, :128-154): keyframe k = camera 0 on the identity pixel grid, frame f = camera 1 gathered by the
    match index idx_f2k; valid = valid_match & C gates & Q gate.  Returns Xf, Xk (N,3), T_WCf, T_WCk (1,8) [f
    perturbed], T_WCf_gt, Qk (N,1), valid (N,1) bool, meas_k (N,3), valid_meas_k (N,1) bool, K, img_size."""
    g = make_gn_problem(2, h, w, ([0], [1]), seed=seed, device=device, calib=calib, pose_noise=pose_noise,
                        yaw_step_deg=yaw_step_deg, point_noise=point_noise)
    e = 1  # directed edge ii = 1 (frame f), jj = 0 (keyframe k): pixel n of k matches pixel idx[n] of f
    idx = g["idx_ii2jj"][e]
    Xf, Xk = g["Xs"][1][idx].contiguous(), g["Xs"][0].contiguous()
    Cf, Ck = g["Cs"][1][idx], g["Cs"][0]
    Qk = g["Q"][e]
    valid = g["valid_match"][e] & (Cf > C_conf) & (Ck > C_conf) & (Qk > Q_conf)
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float32, device=device),
                          torch.arange(w, dtype=torch.float32, device=device), indexing="ij")
    valid_meas_k = Xk[..., 2:3] > depth_eps
    meas_k = torch.cat((u.reshape(-1, 1), v.reshape(-1, 1), torch.log(Xk[..., 2:3])), dim=-1)
    meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
    return dict(Xf=Xf, Xk=Xk, T_WCf=g["Twc"][1:2].contiguous(), T_WCk=g["Twc"][0:1].contiguous(),
                T_WCf_gt=g["Twc_gt"][1:2], Qk=Qk.contiguous(), valid=valid.contiguous(), meas_k=meas_k.contiguous(),
                valid_meas_k=valid_meas_k.contiguous(), K=g["K"], img_size=(h, w))


# ----------------------------------------------------------------------------------------------- matching pairs
def img_gradient(img):
    """Scharr-like gradient with reflect padding (reference: image.py:5-38). img (b,c,h,w)."""
    c = img.shape[1]
    gx_k = (1.0 / 32.0) * torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]], device=img.device,
                                       dtype=img.dtype).repeat(c, 1, 1, 1)
    gy_k = (1.0 / 32.0) * torch.tensor([[-3.0, -10.0, -3.0], [0.0, 0.0, 0.0], [3.0, 10.0, 3.0]], device=img.device,
                                       dtype=img.dtype).repeat(c, 1, 1, 1)
    pad = torch.nn.functional.pad(img, (1, 1, 1, 1), mode="reflect")
    return (torch.nn.functional.conv2d(pad, gx_k, groups=c), torch.nn.functional.conv2d(pad, gy_k, groups=c))


def prep_for_iter_proj(X11, X21, idx_init=None):
    """Same tensors as matching.prep_for_iter_proj (matching.py:25-49)."""
    b, h, w, _ = X11.shape
    rays = torch.nn.functional.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    rays_with_grad = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts3d_norm = torch.nn.functional.normalize(X21.reshape(b, -1, 3), dim=-1).contiguous()
    if idx_init is None:
        idx_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = torch.stack((idx_init % w, idx_init // w), dim=-1).float().contiguous()
    return rays_with_grad, pts3d_norm, p_init


def make_match_problem(b, h, w, seed=1235, device="cpu", yaw_step_deg=12.0, fdim=24, warm_start=False,
                       point_noise=0.005):
    """b image pairs (frame 2c+1 seen from camera 2c): X11, X21 (b,h,w,3) f32, D11, D21 (b,h,w,fdim) f32 plus
    the three iter_proj inputs and the int64 p1 a caller would hand to refine_matches."""
    T = camera_poses(2 * b, yaw_step_deg=yaw_step_deg, seed=seed)
    X11, X21, D11, D21 = [], [], [], []
    for p in range(b):
        T1, T2 = T[2 * p].to(device), T[2 * p + 1].to(device)
        x1 = pointmap(T[2 * p], h, w, noise=point_noise, seed=seed * 100 + 2 * p, device=device) * T1[7]  # metric, camera 1
        x2 = pointmap(T[2 * p + 1], h, w, noise=point_noise, seed=seed * 100 + 2 * p + 1, device=device)
        x2_w = sim3_act(T2, x2)
        T1m = T1.clone()
        T1m[7] = 1.0
        x2_in1 = sim3_act(sim3_inv(T1m), x2_w)
        x1_w = sim3_act(T1m, x1)
        X11.append(x1.view(h, w, 3))
        X21.append(x2_in1.view(h, w, 3))
        D11.append(descriptor_field(x1_w, fdim, seed=seed).view(h, w, fdim))
        D21.append(descriptor_field(x2_w, fdim, seed=seed, noise=0.05).view(h, w, fdim))
    X11 = torch.stack(X11).float()
    X21 = torch.stack(X21).float()
    D11 = torch.stack(D11).float().contiguous()
    D21 = torch.stack(D21).float().contiguous()
    idx_init = None
    if warm_start:
        K = intrinsics(h, w, device=device)
        P = X21.double().view(b, -1, 3)
        uu = (K[0, 0] * P[..., 0] / P[..., 2] + K[0, 2]).round().clamp(0, w - 1).long()
        vv = (K[1, 1] * P[..., 1] / P[..., 2] + K[1, 2]).round().clamp(0, h - 1).long()
        g = torch.Generator().manual_seed(seed + 3)
        jit = torch.randint(-4, 5, (b, h * w, 2), generator=g).to(device)
        idx_init = (vv + jit[..., 1]).clamp(0, h - 1) * w + (uu + jit[..., 0]).clamp(0, w - 1)
    rays, pts, p_init = prep_for_iter_proj(X11, X21, idx_init)
    return dict(X11=X11, X21=X21, D11=D11, D21=D21, rays_img_with_grad=rays, pts_3d_norm=pts, p_init=p_init, h=h, w=w)
