"""Graph-resident mirror of mast3r_slam/global_opt.py `FactorGraph`.

The reference keeps one tensor per edge attribute and per direction, grows them with `torch.cat` on every
`add_factors` (global_opt.py:89-96), re-concatenates both directions on every solve (`prep_two_way_edges`, :104-110)
and re-stacks every keyframe's pointmap and confidence (`get_poses_points`, :112-119): about
(16 N N_kf + 2 * 13 N E/2) bytes of device copies before the first kernel of a solve (1.4 GB at 32 keyframes / 512
directed edges), comparable to a whole Gauss-Newton iteration.

Here the directed edges live in ONE set of capacity-doubling device buffers, both directions of a graph edge appended
next to each other ([e0 fwd, e0 bwd, e1 fwd, ...]); a solve hands the kernels views of those buffers -- no copy.  The
kernels do not care about edge order, so this is the same problem as the reference's [all fwd | all bwd] layout; only
the fixed summation order inside the assembly differs (results agree to rounding, tested).  Keyframe data comes from a
`KeyframeStore` with stacked buffers (the reference's `SharedKeyframes` holds the same stacked buffers, frame.py:
186-200, and copies out of them); when the keyframes that have edges are 0..K-1, the solve reads the store's buffers in
place, otherwise it gathers the used rows once.

Out of scope here: the MASt3R network call inside `add_factors` (`mast3r_match_symmetric`, global_opt.py:44-54).
`add_factors_from_matches` takes that call's eight results and applies the reference's bookkeeping (:56-99) verbatim.
"""
import torch

from . import backends as be
from . import geometry

# config/base.yaml:35-47 (`local_opt:`)
LOCAL_OPT = dict(pin=1, window_size=1e+6, C_conf=0.0, Q_conf=1.5, min_match_frac=0.1, pixel_border=-10, depth_eps=1e-6,
                 max_iters=10, sigma_ray=0.003, sigma_dist=1e+1, sigma_pixel=1.0, sigma_depth=1e+1, sigma_point=0.05,
                 delta_norm=1e-8, use_cuda=True)


class KeyframeStore:
    """Stacked keyframe buffers: X_canon (cap,N,3), ACCUMULATED confidence C (cap,N,1) with its update count N_upd (what
    Frame keeps as self.C and self.N, frame.py:41-105), T_WC (cap,8) [t, q_xyzw, s].  The solver gets the AVERAGE confidence
    C / N (Frame.get_average_conf, frame.py:107-108; global_opt.py:116) -- a view of C while every count is 1 (no fusion
    happened, or only the modes that reset the count), one element-wise pass otherwise."""

    def __init__(self, num_points, capacity=64, device="cuda"):
        self.device = torch.device(device)
        self.N = int(num_points)
        self.n = 0
        self._counts = []  # host copy of the update counts (the reference keeps Frame.N on the host as well)
        self._alloc(int(capacity))

    def _alloc(self, cap):
        X = torch.empty((cap, self.N, 3), dtype=torch.float32, device=self.device)
        Cc = torch.empty((cap, self.N, 1), dtype=torch.float32, device=self.device)
        T = torch.empty((cap, 8), dtype=torch.float32, device=self.device)
        if self.n:
            X[:self.n], Cc[:self.n], T[:self.n] = self.X[:self.n], self.C[:self.n], self.T_WC[:self.n]
        self.X, self.C, self.T_WC, self.cap = X, Cc, T, cap

    def __len__(self):
        return self.n

    def append(self, X_canon, C, T_WC, N_updates=1):
        """C: the keyframe's accumulated confidence, N_updates its count (Frame.C, Frame.N); 1 for a fresh keyframe."""
        if self.n == self.cap:
            self._alloc(2 * self.cap)
        k = self.n
        self.X[k] = X_canon.reshape(self.N, 3)
        self.C[k] = C.reshape(self.N, 1)
        self.T_WC[k] = T_WC.reshape(8)
        self._counts.append(int(N_updates))
        self.n += 1
        return k

    def update_pointmap(self, k, X_canon, C, N_updates=1):
        """Replacement of keyframe k's pointmap (filtering modes first / recent / best_score, frame.py:53-66: N = 1)."""
        self.X[k] = X_canon.reshape(self.N, 3)
        self.C[k] = C.reshape(self.N, 1)
        self._counts[k] = int(N_updates)

    def fuse_pointmap(self, k, X_new, C_new, T=None, filtering_mode="weighted_pointmap"):
        """Frame.update_pointmap (frame.py:41-105) on keyframe k's rows, in place, one kernel.  The weighted modes add the
        new confidence to C and count the update (frame.py:74-75, :99-100); indep_conf resets the count to 1 (:71)."""
        geometry.update_pointmap(self.X[k], self.C[k], X_new.reshape(self.N, 3).contiguous(),
                                 C_new.reshape(self.N, 1).contiguous(), T, filtering_mode)
        if filtering_mode in ("weighted_pointmap", "weighted_spherical"):
            self._counts[k] += 1
        else:
            self._counts[k] = 1

    def update_count(self, k):
        return self._counts[k]

    def average_conf(self, idx=None):
        """C / N of keyframes idx (LongTensor) or of all: Frame.get_average_conf (frame.py:107-108)."""
        ks = list(range(self.n)) if idx is None else [int(i) for i in idx.tolist()]
        Cs = self.C[:self.n] if idx is None else self.C.index_select(0, idx)
        if all(self._counts[k] == 1 for k in ks):
            return Cs
        cnt = torch.tensor([float(self._counts[k]) for k in ks], dtype=torch.float32, device=self.device)
        return Cs / cnt.view(-1, 1, 1)

    def update_T_WCs(self, T_WCs, idx):
        """SharedKeyframes.update_T_WCs (frame.py:258-262): scatter (m,8) poses to keyframes idx."""
        self.T_WC[idx] = T_WCs.reshape(-1, 8)


class _EdgeBuffers:
    """Directed edges in append order, capacity doubling: ii, jj (E), idx (E,N) int64, valid (E,N,1) bool, Q (E,N,1)."""

    def __init__(self, N, device, capacity=64):
        self.N, self.device, self.E, self.cap = N, device, 0, 0
        self._grow(capacity)

    def _grow(self, cap):
        new = dict(ii=torch.empty(cap, dtype=torch.int64, device=self.device),
                   jj=torch.empty(cap, dtype=torch.int64, device=self.device),
                   idx=torch.empty((cap, self.N), dtype=torch.int64, device=self.device),
                   valid=torch.empty((cap, self.N, 1), dtype=torch.bool, device=self.device),
                   Q=torch.empty((cap, self.N, 1), dtype=torch.float32, device=self.device))
        if self.E:
            for k, v in new.items():
                v[:self.E] = getattr(self, k)[:self.E]
        for k, v in new.items():
            setattr(self, k, v)
        self.cap = cap

    def append_pairs(self, ii, jj, idx_i2j, idx_j2i, valid_j, valid_i, Qj, Qi):
        m = ii.numel()
        if m == 0:
            return
        while self.E + 2 * m > self.cap:
            self._grow(2 * self.cap)
        s = slice(self.E, self.E + 2 * m)
        # interleave: directed edge 2k = (ii -> jj) with the forward match, 2k+1 = (jj -> ii) with the backward match
        self.ii[s] = torch.stack((ii, jj), 1).reshape(-1)
        self.jj[s] = torch.stack((jj, ii), 1).reshape(-1)
        self.idx[s] = torch.stack((idx_i2j, idx_j2i), 1).reshape(2 * m, self.N)
        self.valid[s] = torch.stack((valid_j, valid_i), 1).reshape(2 * m, self.N, 1)
        self.Q[s] = torch.stack((Qj, Qi), 1).reshape(2 * m, self.N, 1)
        self.E += 2 * m


class FactorGraph:
    """mast3r_slam.global_opt.FactorGraph without the network: same attributes' meaning, same solve methods."""

    def __init__(self, frames: KeyframeStore, K=None, device="cuda", cfg=None):
        self.frames = frames
        self.device = torch.device(device)
        self.cfg = dict(LOCAL_OPT)
        if cfg:
            self.cfg.update(cfg)
        self.K = K
        self.window_size = self.cfg["window_size"]
        self.edges = _EdgeBuffers(frames.N, self.device)

    # ---- views in the reference's one-direction layout (global_opt.py:17-26), for inspection and tests
    @property
    def ii(self):
        return self.edges.ii[0:self.edges.E:2]

    @property
    def jj(self):
        return self.edges.jj[0:self.edges.E:2]

    def add_factors_from_matches(self, ii, jj, idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij,
                                 min_match_frac, is_reloc=False):
        """global_opt.py:56-99 with the eight results of mast3r_match_symmetric (:44-54) as inputs."""
        batch_inds = torch.arange(idx_i2j.shape[0], device=idx_i2j.device)[:, None].repeat(1, idx_i2j.shape[1])
        Qj = torch.sqrt(Qii[batch_inds, idx_i2j] * Qji)
        Qi = torch.sqrt(Qjj[batch_inds, idx_j2i] * Qij)
        valid_Qj = Qj > self.cfg["Q_conf"]
        valid_Qi = Qi > self.cfg["Q_conf"]
        valid_j = valid_match_j & valid_Qj
        valid_i = valid_match_i & valid_Qi
        nj = valid_j.shape[1] * valid_j.shape[2]
        ni = valid_i.shape[1] * valid_i.shape[2]
        match_frac_j = valid_j.sum(dim=(1, 2)) / nj
        match_frac_i = valid_i.sum(dim=(1, 2)) / ni
        ii_tensor = torch.as_tensor(ii, device=self.device)
        jj_tensor = torch.as_tensor(jj, device=self.device)
        # both edge directions must be above the threshold to accept either (global_opt.py:71-74)
        invalid_edges = torch.minimum(match_frac_j, match_frac_i) < min_match_frac
        consecutive_edges = ii_tensor == (jj_tensor - 1)
        invalid_edges = (~consecutive_edges) & invalid_edges
        if invalid_edges.any() and is_reloc:
            return False
        valid_edges = ~invalid_edges
        self.edges.append_pairs(ii_tensor[valid_edges], jj_tensor[valid_edges], idx_i2j[valid_edges],
                                idx_j2i[valid_edges], valid_match_j[valid_edges], valid_match_i[valid_edges],
                                Qj[valid_edges], Qi[valid_edges])
        return valid_edges.sum() > 0

    def get_unique_kf_idx(self):
        E = self.edges.E
        return torch.unique(torch.cat([self.edges.ii[:E], self.edges.jj[:E]]), sorted=True)

    def prep_two_way_edges(self):
        """Views, not copies (the reference concatenates, global_opt.py:104-110)."""
        e, E = self.edges, self.edges.E
        return e.ii[:E], e.jj[:E], e.idx[:E], e.valid[:E], e.Q[:E]

    def get_poses_points(self, unique_kf_idx):
        """Xs, T_WCs (m,8), average confidences of the keyframes that have edges (global_opt.py:112-119); views of the
        store when those are 0..m-1 (and no keyframe has fused updates, see KeyframeStore.average_conf)."""
        m = unique_kf_idx.numel()
        if m == len(self.frames) and int(unique_kf_idx[-1]) == m - 1:
            return self.frames.X[:m], self.frames.T_WC[:m], self.frames.average_conf(), True
        return (self.frames.X.index_select(0, unique_kf_idx), self.frames.T_WC.index_select(0, unique_kf_idx),
                self.frames.average_conf(unique_kf_idx), False)

    def _finish(self, pose_data, unique_kf_idx, in_place):
        pin = self.cfg["pin"]
        if not in_place:  # gathered copy: write the optimised poses back (frames.update_T_WCs, global_opt.py:157)
            self.frames.update_T_WCs(pose_data[pin:], unique_kf_idx[pin:])

    def solve_GN_rays(self):
        """global_opt.py:121-157."""
        c = self.cfg
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= c["pin"]:
            return
        Xs, pose_data, Cs, in_place = self.get_poses_points(unique_kf_idx)
        ii, jj, idx_ii2jj, valid_match, Q_ii2jj = self.prep_two_way_edges()
        be.gauss_newton_rays(pose_data, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, c["sigma_ray"],
                             c["sigma_dist"], c["C_conf"], c["Q_conf"], c["max_iters"], c["delta_norm"])
        self._finish(pose_data, unique_kf_idx, in_place)

    def solve_GN_calib(self, img_size):
        """global_opt.py:159-215; `img_size` = (h, w) (the reference reads it off frames[0].img)."""
        c = self.cfg
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= c["pin"]:
            return
        Xs, pose_data, Cs, in_place = self.get_poses_points(unique_kf_idx)
        Xs = geometry.constrain_points_to_ray(img_size, Xs, self.K)
        ii, jj, idx_ii2jj, valid_match, Q_ii2jj = self.prep_two_way_edges()
        height, width = int(img_size[0]), int(img_size[1])
        be.gauss_newton_calib(pose_data, Xs, Cs, self.K, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, height, width,
                              c["pixel_border"], c["depth_eps"], c["sigma_pixel"], c["sigma_depth"], c["C_conf"],
                              c["Q_conf"], c["max_iters"], c["delta_norm"])
        self._finish(pose_data, unique_kf_idx, in_place)
