"""Graph-resident FactorGraph mirror (mast3r_slam_ysh_b200/global_opt.py) against a verbatim restatement of the
reference's bookkeeping (mast3r_slam/global_opt.py:56-110) and, on the GPU, against the drop-in solver fed the
reference's concatenated layout."""
import pytest
import torch


class RefGraph:
    """mast3r_slam/global_opt.py:17-26, :56-110 restated with torch.cat lists (test infrastructure)."""

    def __init__(self, Q_conf, device="cpu"):
        self.Q_conf, self.device = Q_conf, device
        mk = lambda dt: torch.as_tensor([], dtype=dt, device=device)
        self.ii, self.jj = mk(torch.long), mk(torch.long)
        self.idx_ii2jj, self.idx_jj2ii = mk(torch.long), mk(torch.long)
        self.valid_match_j, self.valid_match_i = mk(torch.bool), mk(torch.bool)
        self.Q_ii2jj, self.Q_jj2ii = mk(torch.float32), mk(torch.float32)

    def add(self, ii, jj, idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij, min_match_frac,
            is_reloc=False):
        batch_inds = torch.arange(idx_i2j.shape[0], device=idx_i2j.device)[:, None].repeat(1, idx_i2j.shape[1])
        Qj = torch.sqrt(Qii[batch_inds, idx_i2j] * Qji)
        Qi = torch.sqrt(Qjj[batch_inds, idx_j2i] * Qij)
        valid_j = valid_match_j & (Qj > self.Q_conf)
        valid_i = valid_match_i & (Qi > self.Q_conf)
        nj = valid_j.shape[1] * valid_j.shape[2]
        ni = valid_i.shape[1] * valid_i.shape[2]
        match_frac_j = valid_j.sum(dim=(1, 2)) / nj
        match_frac_i = valid_i.sum(dim=(1, 2)) / ni
        ii_tensor = torch.as_tensor(ii, device=self.device)
        jj_tensor = torch.as_tensor(jj, device=self.device)
        invalid_edges = torch.minimum(match_frac_j, match_frac_i) < min_match_frac
        consecutive_edges = ii_tensor == (jj_tensor - 1)
        invalid_edges = (~consecutive_edges) & invalid_edges
        if invalid_edges.any() and is_reloc:
            return False
        v = ~invalid_edges
        self.ii = torch.cat([self.ii, ii_tensor[v]])
        self.jj = torch.cat([self.jj, jj_tensor[v]])
        self.idx_ii2jj = torch.cat([self.idx_ii2jj, idx_i2j[v]])
        self.idx_jj2ii = torch.cat([self.idx_jj2ii, idx_j2i[v]])
        self.valid_match_j = torch.cat([self.valid_match_j, valid_match_j[v]])
        self.valid_match_i = torch.cat([self.valid_match_i, valid_match_i[v]])
        self.Q_ii2jj = torch.cat([self.Q_ii2jj, Qj[v]])
        self.Q_jj2ii = torch.cat([self.Q_jj2ii, Qi[v]])
        return v.sum() > 0

    def two_way(self):
        return (torch.cat((self.ii, self.jj)), torch.cat((self.jj, self.ii)),
                torch.cat((self.idx_ii2jj, self.idx_jj2ii)), torch.cat((self.valid_match_j, self.valid_match_i)),
                torch.cat((self.Q_ii2jj, self.Q_jj2ii)))


def _random_matches(m, N, g, frac_valid):
    idx_i2j = torch.randint(0, N, (m, N), generator=g)
    idx_j2i = torch.randint(0, N, (m, N), generator=g)
    vj = torch.rand(m, N, 1, generator=g) < frac_valid
    vi = torch.rand(m, N, 1, generator=g) < frac_valid
    Qs = [0.5 + 3.0 * torch.rand(m, N, 1, generator=g) for _ in range(4)]
    return (idx_i2j, idx_j2i, vj, vi, *Qs)


def test_bookkeeping_matches_reference_restatement():
    from mast3r_slam_ysh_b200 import global_opt as go
    g = torch.Generator().manual_seed(3)
    N = 96
    store = go.KeyframeStore(N, capacity=2, device="cpu")
    fg = go.FactorGraph(store, device="cpu")
    fg.edges = go._EdgeBuffers(N, torch.device("cpu"), capacity=2)  # force several capacity doublings
    ref = RefGraph(fg.cfg["Q_conf"])
    batches = [([0, 1, 0], [1, 2, 2], 0.9), ([2, 0, 1, 3], [3, 3, 3, 4], 0.02), ([4], [5], 0.02), ([0, 2], [5, 5], 0.9)]
    for ii, jj, frac in batches:
        args = _random_matches(len(ii), N, g, frac)
        a = fg.add_factors_from_matches(ii, jj, *args, min_match_frac=0.1)
        b = ref.add(ii, jj, *args, min_match_frac=0.1)
        assert bool(a) == bool(b)
    # a relocalisation batch with a weak non-consecutive edge is rejected as a whole (global_opt.py:76-77)
    args = _random_matches(2, N, g, 0.01)
    assert fg.add_factors_from_matches([0, 1], [4, 5], *args, min_match_frac=0.1, is_reloc=True) is False
    assert ref.add([0, 1], [4, 5], *args, min_match_frac=0.1, is_reloc=True) is False
    m = ref.ii.numel()
    assert m > 0 and fg.edges.E == 2 * m and torch.equal(fg.ii, ref.ii) and torch.equal(fg.jj, ref.jj)
    rii, rjj, ridx, rvalid, rQ = ref.two_way()
    ii, jj, idx, valid, Q = fg.prep_two_way_edges()
    perm = torch.cat((torch.arange(0, 2 * m, 2), torch.arange(1, 2 * m, 2)))  # [all fwd | all bwd] -> interleaved
    assert torch.equal(ii[perm], rii) and torch.equal(jj[perm], rjj) and torch.equal(idx[perm], ridx)
    assert torch.equal(valid[perm], rvalid) and torch.equal(Q[perm], rQ)
    assert torch.equal(fg.get_unique_kf_idx(), torch.unique(torch.cat([ref.ii, ref.jj]), sorted=True))
    # views, not copies
    assert idx.data_ptr() == fg.edges.idx.data_ptr() and Q.data_ptr() == fg.edges.Q.data_ptr()


def test_keyframe_store_growth_and_updates():
    from mast3r_slam_ysh_b200 import global_opt as go
    s = go.KeyframeStore(8, capacity=1, device="cpu")
    Xs = [torch.randn(8, 3) for _ in range(5)]
    for k, X in enumerate(Xs):
        assert s.append(X, torch.full((8, 1), float(k)), torch.arange(8.0) + k) == k
    assert len(s) == 5 and s.cap >= 5
    for k, X in enumerate(Xs):
        assert torch.equal(s.X[k], X) and torch.equal(s.C[k], torch.full((8, 1), float(k)))
    s.update_pointmap(2, Xs[0], torch.ones(8, 1))
    assert torch.equal(s.X[2], Xs[0])
    s.update_T_WCs(torch.zeros(2, 8), torch.tensor([1, 3]))
    assert s.T_WC[1].abs().sum() == 0 and s.T_WC[3].abs().sum() == 0 and s.T_WC[2, 0] == 2.0


def _graph_from_problem(g, dev, batches=2, cfg=None):
    """Feed a synthetic problem's graph edges through add_factors_from_matches in `batches` calls."""
    from mast3r_slam_ysh_b200 import global_opt as go
    nkf, N = g["Xs"].shape[0], g["Xs"].shape[1]
    store = go.KeyframeStore(N, capacity=4, device=dev)
    for k in range(nkf):
        store.append(g["Xs"][k], g["Cs"][k], g["Twc"][k])
    fg = go.FactorGraph(store, K=g["K"], device=dev, cfg=cfg)
    m = g["ii"].numel() // 2
    ones = torch.ones((m, N, 1), device=dev)
    Qf, Qb = g["Q"][:m], g["Q"][m:]
    bounds = torch.linspace(0, m, batches + 1).long().tolist()
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        s = slice(lo, hi)
        # Qj = sqrt(Qii[idx] * Qji): Qii = 1, Qji = Q^2 reproduces Q up to one rounding of the square root
        fg.add_factors_from_matches(g["ii"][s].tolist(), g["jj"][s].tolist(), g["idx_ii2jj"][:m][s], g["idx_ii2jj"][m:][s],
                                    g["valid_match"][:m][s], g["valid_match"][m:][s], ones[s], ones[s], Qf[s] ** 2,
                                    Qb[s] ** 2, min_match_frac=0.0)
    return fg, store


@pytest.mark.gpu
@pytest.mark.parametrize("calib", [False, True])
def test_resident_graph_solve_equals_dropin_on_reference_layout(be, calib):
    from mast3r_slam_ysh_b200 import synthetic as syn
    dev = torch.device("cuda:0")
    nkf, h, w = 6, 48, 64
    g = syn.make_gn_problem(nkf, h, w, syn.ring_edges(nkf, 2), seed=21, device=dev, calib=False, yaw_step_deg=10.0)
    fg, store = _graph_from_problem(g, dev)
    assert fg.edges.E == g["ii"].numel()
    lo = syn.LOCAL_OPT
    # reference path: concatenated [fwd | bwd] layout, sqrt-rounded Q as the graph stores it
    m = g["ii"].numel() // 2
    perm = torch.cat((torch.arange(0, 2 * m, 2), torch.arange(1, 2 * m, 2))).to(dev)
    ii, jj, idx, valid, Q = (t[perm].contiguous() for t in fg.prep_two_way_edges())
    assert torch.equal(ii, g["ii"]) and torch.equal(idx, g["idx_ii2jj"]) and torch.allclose(Q, g["Q"], rtol=1e-6)
    T_ref = g["Twc"].clone()
    if calib:
        vv, uu = torch.meshgrid(torch.arange(h, device=dev, dtype=torch.float32),
                                torch.arange(w, device=dev, dtype=torch.float32), indexing="ij")
        K = g["K"]
        t1 = ((uu - K[0, 2]) / K[0, 0]).reshape(1, -1)
        t2 = ((vv - K[1, 2]) / K[1, 1]).reshape(1, -1)
        z = g["Xs"][..., 2]
        Xc = torch.stack((z * t1, z * t2, z * 1.0), -1).contiguous()  # geometry.py:37-42, :111-120
        from mast3r_slam_ysh_b200 import geometry
        assert torch.equal(geometry.constrain_points_to_ray((h, w), g["Xs"], K), Xc)  # bit-exact
        be.gauss_newton_calib(T_ref, Xc, g["Cs"], K, ii, jj, idx, valid, Q, h, w, lo["pixel_border"], lo["depth_eps"],
                              lo["sigma_pixel"], lo["sigma_depth"], lo["C_conf"], lo["Q_conf"], lo["max_iters"],
                              lo["delta_norm"])
        fg.solve_GN_calib((h, w))
    else:
        be.gauss_newton_rays(T_ref, g["Xs"], g["Cs"], ii, jj, idx, valid, Q, lo["sigma_ray"], lo["sigma_dist"],
                             lo["C_conf"], lo["Q_conf"], lo["max_iters"], lo["delta_norm"])
        fg.solve_GN_rays()
    T_new = store.T_WC[:nkf]
    assert (T_ref - g["Twc"]).abs().max() > 1e-3  # the solve moved the poses
    assert torch.allclose(T_new, T_ref, rtol=0, atol=2e-5)
    assert torch.equal(T_new[0], g["Twc"][0])  # pinned keyframe


@pytest.mark.gpu
def test_resident_graph_is_zero_copy_and_handles_unused_keyframes(be):
    from mast3r_slam_ysh_b200 import synthetic as syn
    dev = torch.device("cuda:0")
    g = syn.make_gn_problem(5, 24, 32, syn.ring_edges(5, 1), seed=4, device=dev, yaw_step_deg=10.0)
    fg, store = _graph_from_problem(g, dev, batches=1)
    Xs, T, Cs, in_place = fg.get_poses_points(fg.get_unique_kf_idx())
    assert in_place and Xs.data_ptr() == store.X.data_ptr() and T.data_ptr() == store.T_WC.data_ptr()
    # a keyframe without any edge: the used rows are gathered and the poses scattered back
    store.append(g["Xs"][0], g["Cs"][0], g["Twc"][0])
    before = store.T_WC[:6].clone()
    Xs, T, Cs, in_place = fg.get_poses_points(fg.get_unique_kf_idx())
    assert not in_place and Xs.shape[0] == 5
    fg.solve_GN_rays()
    assert torch.equal(store.T_WC[5], before[5]) and torch.equal(store.T_WC[0], before[0])
    assert (store.T_WC[1:5] - before[1:5]).abs().max() > 1e-4


@pytest.mark.gpu
def test_solver_gets_average_confidence_after_fusion(be, orc):
    """C_conf > 0: after a weighted pointmap fusion the confidence gate must see C / N (frame.py:107-108, global_opt.py:116),
    not the accumulated sum -- checked against the oracle run on the averaged confidences."""
    import numpy as np
    from mast3r_slam_ysh_b200 import global_opt as go, synthetic as syn
    dev = torch.device("cuda:0")
    g = syn.make_gn_problem(5, 24, 32, syn.ring_edges(5, 2, wrap=False), seed=6, device=dev, yaw_step_deg=9.0)
    c_gate = float(g["Cs"].median())
    fg, store = _graph_from_problem(g, dev, batches=1, cfg=dict(C_conf=c_gate))
    # fuse a second observation into keyframe 2: accumulated C doubles (most points would now pass the gate on the sum),
    # its average stays where it was
    store.fuse_pointmap(2, g["Xs"][2], g["Cs"][2])
    assert store.update_count(2) == 2 and float((store.C[2] > c_gate).float().mean()) > 0.75  # vs ~0.5 for the average
    Cs_avg = store.average_conf()
    assert torch.allclose(Cs_avg[2], g["Cs"][2]) and Cs_avg.data_ptr() != store.C.data_ptr()
    T_o = g["Twc"].cpu().numpy().copy()
    lo = fg.cfg
    orc.gauss_newton_rays(T_o, store.X[:5].cpu().numpy(), Cs_avg.cpu().numpy(), *[t.cpu().numpy() for t in fg.prep_two_way_edges()],
                          lo["sigma_ray"], lo["sigma_dist"], lo["C_conf"], lo["Q_conf"], lo["max_iters"], lo["delta_norm"])
    fg.solve_GN_rays()
    assert np.linalg.norm(store.T_WC[:5].cpu().numpy() - T_o) / np.linalg.norm(T_o) <= 1e-4
