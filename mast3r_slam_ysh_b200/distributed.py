"""Multi-GPU drivers (one process per GPU, torch.distributed): edge-sharded Gauss-Newton and pair-sharded matching.

The reference has no multi-GPU path (no NCCL/distributed call anywhere); this is the sharding
BASELINE.json:north_star defines: "graph edges are split across GPUs and the block Hessian is all-reduced, and batched
keyframe-pair matching ... is split by pair".

GN: every rank holds a replica of Twc/Xs/Cs and a contiguous slice of the directed edges.  Per iteration each rank
assembles its partial dense [H | b] in fp64, the partial systems are summed over NVLink, and every rank runs the same
deterministic fp64 Cholesky + retraction, so pose replicas stay bit-identical without a broadcast.  The convergence
flag lives on the device (a pure function of the replicated dx), so no rank ever synchronises with its host inside the
loop.  Two exchange paths:
  * "peer" (default on CUDA when torch symmetric memory is available): the system buffers are peer-mapped and the fused
    solver kernel sums them itself between assembly and factorisation (m3r_gn_iterate_peer) -- one Hessian kernel and
    ONE cooperative launch per iteration, exactly as on a single GPU, and no collective library call in the loop;
  * "nccl": m3r_gn_accumulate, ONE dist.all_reduce of [H | b], m3r_gn_solve_retract.
"""
import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _lib


# ------------------------------------------------------------------------------------------------ partitioning (host)
def partition_edges(num_graph_edges, world_size):
    """Contiguous, balanced split of the UNDIRECTED graph edges; returns [(lo, hi)] per rank.  Both directions of a
    graph edge go to the same rank so the two keyframes' point tiles are shared in its L2."""
    base, rem = divmod(num_graph_edges, world_size)
    out, lo = [], 0
    for r in range(world_size):
        hi = lo + base + (1 if r < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def directed_slice(num_graph_edges, lo, hi):
    """Indices into the two-way edge arrays of prep_two_way_edges (global_opt.py:104-110: [forward..., backward...])
    that belong to undirected edges lo..hi-1."""
    fwd = torch.arange(lo, hi)
    return torch.cat((fwd, fwd + num_graph_edges))


def partition_pairs(num_pairs, world_size):
    return partition_edges(num_pairs, world_size)


def allreduce_system(sys_tensor, group=None):
    """The one collective of a GN iteration: sum of the partial [H | b] buffers (fp64)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(sys_tensor, op=dist.ReduceOp.SUM, group=group)
    return sys_tensor


# ------------------------------------------------------------------------------------------------ peer-mapped systems
class _PeerSystem:
    """One symmetric-memory allocation per (device, group): every rank's [H | b] buffer + arrival slots, mapped into
    every rank's address space (torch.distributed._symmetric_memory: CUDA IPC / fabric handles).  Allocated and
    rendezvoused ONCE and reused by every solve (the rendezvous is a collective and costs milliseconds); the epoch
    counter advances identically on all ranks because they make the same sequence of calls."""
    _cache = {}

    def __init__(self, count, dev, group):
        import torch.distributed._symmetric_memory as symm
        self.count = int(count)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.buf = symm.empty(self.count + 64, dtype=torch.float64, device=dev)  # tail: 64 x 8-byte arrival slots
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.sys_ptrs = (C.c_uint64 * self.world)(*ptrs)
        self.flag_ptrs = (C.c_uint64 * self.world)(*[p + 8 * self.count for p in ptrs])
        self.epoch = 0
        torch.cuda.synchronize(dev)
        dist.barrier(group)  # every rank's buffer is zero before anyone's kernel may write into it

    @classmethod
    def get(cls, count, dev, group):
        key = (dev.index, id(group))
        cur = cls._cache.get(key)
        if cur is None or cur.count < count:
            # grow with head-room so that a growing graph does not rendezvous at every solve
            cur = cls(max(int(count), 2 * cur.count if cur is not None else 0), dev, group)
            cls._cache[key] = cur
        return cur


def peer_exchange_available():
    try:
        import torch.distributed._symmetric_memory  # noqa: F401
    except Exception:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"


# ------------------------------------------------------------------------------------------------ sharded GN (device)
class ShardedGaussNewton:
    """Step-wise GN over the C ABI (m3r_gn_begin / iterate_peer | accumulate + all-reduce + solve_retract / end)."""

    def __init__(self, params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, group=None,
                 exchange="auto"):
        self.lib = _lib.load()
        self.params = params
        self.group = group
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        if exchange == "auto":
            exchange = os.environ.get("M3R_EXCHANGE", "auto")  # A/B switch for benchmarking: peer | nccl
        if exchange == "auto":
            exchange = "peer" if world > 1 and world <= 8 and peer_exchange_available() else "nccl"
        self.exchange = exchange if world > 1 else "local"
        self.Twc = Twc
        self.dev = Twc.device
        self.num_kf, self.num_pts = Xs.shape[0], Xs.shape[1]
        self.E = local_ii.shape[0]
        self.keep = (Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj)
        self.wsb = self.lib.m3r_gn_workspace_bytes(self.num_kf, self.num_pts, self.E)
        self.ws = torch.empty(self.wsb + 256, dtype=torch.uint8, device=self.dev)
        self.ws_ptr = self.ws.data_ptr() + ((-self.ws.data_ptr()) % 256)
        st = self._stream()
        p = lambda t: C.c_void_p(t.data_ptr()) if t.numel() > 0 else C.c_void_p(0)
        with torch.cuda.device(self.dev):
            rc = self.lib.m3r_gn_begin(C.byref(params), p(Twc), p(Xs), p(Cs), p(local_ii), p(local_jj), p(idx_ii2jj),
                                       p(valid_match), p(Q), self.num_kf, self.num_pts, self.E, p(all_ii), p(all_jj),
                                       all_ii.shape[0], C.c_void_p(self.ws_ptr), self.wsb, st)
        _lib.check(rc, "gn_begin")
        # torch view of the [H | b] region of the workspace for the collective
        cnt = self.lib.m3r_gn_system_count(self.num_kf)
        sys_ptr = self.lib.m3r_gn_system_ptr(C.c_void_p(self.ws_ptr), self.num_kf, self.num_pts, self.E)
        off = sys_ptr - self.ws.data_ptr()
        self.sys = self.ws[off:off + 8 * cnt].view(torch.float64)
        self.peer = None
        if self.exchange == "peer":
            try:
                self.peer = _PeerSystem.get(cnt, self.dev, group)
            except Exception as e:  # symmetric memory unavailable on this system: the NCCL path is equivalent
                import warnings
                warnings.warn(f"peer-mapped system buffers unavailable ({e!r}); using the NCCL all-reduce")
                self.exchange = "nccl"

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)

    def hessian_pass(self):
        with torch.cuda.device(self.dev):
            rc = self.lib.m3r_gn_hessian_pass(C.byref(self.params), C.c_void_p(self.Twc.data_ptr()), self.num_kf,
                                              self.num_pts, self.E, C.c_void_p(self.ws_ptr), self._stream())
        _lib.check(rc, "gn_hessian_pass")

    def iterate(self):
        if self.peer is not None:
            pr = self.peer
            pr.epoch += 1
            with torch.cuda.device(self.dev):
                rc = self.lib.m3r_gn_iterate_peer(C.byref(self.params), C.c_void_p(self.Twc.data_ptr()), self.num_kf,
                                                  self.num_pts, self.E, C.c_void_p(self.ws_ptr), pr.rank, pr.world,
                                                  pr.sys_ptrs, pr.flag_ptrs, pr.epoch, self._stream())
            _lib.check(rc, "gn_iterate_peer")
            return
        with torch.cuda.device(self.dev):
            rc = self.lib.m3r_gn_accumulate(C.byref(self.params), C.c_void_p(self.Twc.data_ptr()), self.num_kf,
                                            self.num_pts, self.E, C.c_void_p(self.ws_ptr), self._stream())
            _lib.check(rc, "gn_accumulate")
            allreduce_system(self.sys, self.group)
            rc = self.lib.m3r_gn_solve_retract(C.byref(self.params), C.c_void_p(self.Twc.data_ptr()), self.num_kf,
                                               self.num_pts, self.E, C.c_void_p(self.ws_ptr), self._stream())
            _lib.check(rc, "gn_solve_retract")

    def finish(self):
        dx = torch.zeros((max(self.num_kf - 1, 0), 7), dtype=torch.float32, device=self.dev)
        info = torch.zeros(4, dtype=torch.int32, device=self.dev)
        with torch.cuda.device(self.dev):
            rc = self.lib.m3r_gn_end(self.num_kf, self.num_pts, self.E, C.c_void_p(dx.data_ptr()) if dx.numel() else None,
                                     C.c_void_p(info.data_ptr()), C.c_void_p(self.ws_ptr), self._stream())
        _lib.check(rc, "gn_end")
        if self.peer is not None and (int(info[3].item()) & 0x40000000):
            # a peer rank never reached the in-kernel exchange (~4 s): the device stopped the solve with dx = 0 and the
            # poses of the last completed iteration; replicas may have diverged, so this must not pass silently
            raise RuntimeError("sharded Gauss-Newton: a peer rank never arrived at the in-kernel all-reduce "
                               "(timeout); poses were left at the last completed iteration. Re-run with "
                               "exchange='nccl' or check the other ranks.")
        return dx, info


def _gauss_newton_sharded(params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, max_iter,
                          group, exchange, keep=()):
    s = ShardedGaussNewton(params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, group,
                           exchange)
    s.keep = s.keep + tuple(keep)  # tensors the params struct points at must outlive the enqueued work
    for _ in range(int(max_iter)):
        s.iterate()
    dx, info = s.finish()
    return [dx], info


def gauss_newton_rays_sharded(Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, sigma_ray,
                              sigma_dist, C_thresh, Q_thresh, max_iter, delta_thresh, group=None, exchange="auto"):
    """gauss_newton_rays (gn.cpp:28-52) with the directed edges sharded over the ranks of `group`.
    local_* are this rank's edges, all_ii/all_jj the concatenation over ranks (keyframe numbering must agree)."""
    params = _lib.GnParams(_lib.M3R_GN_RAYS, _lib.M3R_ROBUST_HUBER, float(sigma_ray), float(sigma_dist), float(C_thresh),
                           float(Q_thresh), None, 0, 0, 0, 0.0, int(max_iter), float(delta_thresh))
    return _gauss_newton_sharded(params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj,
                                 max_iter, group, exchange)


def gauss_newton_points_sharded(Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, sigma_point,
                                C_thresh, Q_thresh, max_iter, delta_thresh, group=None, exchange="auto"):
    """gauss_newton_points (gn.cpp:3-26), edge-sharded like gauss_newton_rays_sharded."""
    params = _lib.GnParams(_lib.M3R_GN_POINTS, _lib.M3R_ROBUST_HUBER, float(sigma_point), 1.0, float(C_thresh),
                           float(Q_thresh), None, 0, 0, 0, 0.0, int(max_iter), float(delta_thresh))
    return _gauss_newton_sharded(params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj,
                                 max_iter, group, exchange)


def gauss_newton_calib_sharded(Twc, Xs, Cs, K, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj, height,
                               width, pixel_border, z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter,
                               delta_thresh, group=None, exchange="auto"):
    """gauss_newton_calib (gn.cpp:54-82), edge-sharded like gauss_newton_rays_sharded; K (3,3) f32 on the device."""
    params = _lib.GnParams(_lib.M3R_GN_CALIB, _lib.M3R_ROBUST_HUBER, float(sigma_pixel), float(sigma_depth),
                           float(C_thresh), float(Q_thresh), K.data_ptr(), int(height), int(width), int(pixel_border),
                           float(z_eps), int(max_iter), float(delta_thresh))
    return _gauss_newton_sharded(params, Twc, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj,
                                 max_iter, group, exchange, keep=(K,))


# ------------------------------------------------------------------------------------------------ sharded matching
def gather_ragged(t, group=None):
    """All-gather along dim 0 with per-rank lengths that may differ (pairs are split by partition_pairs, the last ranks
    may hold one pair less): pad to the longest, all_gather, trim, concatenate in rank order.  bool tensors travel as
    uint8.  Identity without an initialised process group or at world size 1."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) <= 1:
        return t
    ws = dist.get_world_size(group)
    is_bool = t.dtype == torch.bool
    src = t.to(torch.uint8) if is_bool else t
    counts = torch.zeros(ws, dtype=torch.int64, device=t.device)
    counts[dist.get_rank(group)] = t.shape[0]
    dist.all_reduce(counts, group=group)
    counts = counts.tolist()
    cmax = max(int(c) for c in counts)
    padded = torch.cat((src, src.new_zeros((cmax - src.shape[0],) + tuple(src.shape[1:]))), 0).contiguous()
    outs = [torch.empty_like(padded) for _ in range(ws)]
    dist.all_gather(outs, padded, group=group)
    out = torch.cat([o[:int(c)] for o, c in zip(outs, counts)], 0)
    return out.bool() if is_bool else out


def match_pairs_sharded(rays_img_with_grad, pts_3d_norm, p_init, D11, D21, matching_cfg, group=None, gather=False):
    """iter_proj + refine_matches on THIS rank's pairs (split by pair with partition_pairs, no data-path collective:
    pairs are independent units).  With gather=True the per-rank results are all-gathered in rank order
    (ragged pair counts allowed), so that one rank can assemble the graph.  Returns (p1 (b,n,2) int64, converged (b,n))."""
    from . import backends as be
    p, conv = be.iter_proj(rays_img_with_grad, pts_3d_norm, p_init, matching_cfg["max_iter"],
                           matching_cfg["lambda_init"], matching_cfg["convergence_thresh"])
    p1 = p.long()
    if matching_cfg["radius"] > 0:
        (p1,) = be.refine_matches(D11, D21, p1, matching_cfg["radius"], matching_cfg["dilation_max"])
    if gather:
        p1, conv = gather_ragged(p1, group), gather_ragged(conv, group)
    return p1, conv


def match_sharded(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, group=None, gather=False):
    """The fused device match (matching.match == mast3r_slam/matching.py:8-10) on THIS rank's pairs; gather as above."""
    from . import matching as mm
    idx, valid = mm.match(X11, X21, D11, D21, idx_1_to_2_init, cfg)
    if gather:
        idx, valid = gather_ragged(idx, group), gather_ragged(valid, group)
    return idx, valid
