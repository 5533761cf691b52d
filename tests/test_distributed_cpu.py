"""world_size-2 gloo test of the multi-GPU host logic: edge partitioning + the one all-reduce of [H | b] per GN
iteration reproduce the single-process result (the per-rank device work is stood in for by the CPU oracle)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mast3r_slam_ysh_b200 import distributed as md, synthetic as syn
    from oracle import oracle as orc
    lo = syn.LOCAL_OPT
    nkf, h, w = 6, 24, 32
    edges = syn.ring_edges(nkf, 2, wrap=False)
    g = syn.make_gn_problem(nkf, h, w, edges, seed=4, yaw_step_deg=9.0)
    n_graph = len(edges[0])
    lo_e, hi_e = md.partition_edges(n_graph, world)[rank]
    sel = md.directed_slice(n_graph, lo_e, hi_e)
    ii, jj = g["ii"].numpy(), g["jj"].numpy()
    uniq, ie, je = orc.index_prep(ii, jj)  # numbering from ALL edges, as m3r_gn_begin does with all_ii/all_jj
    Hs, gs = orc.gn_edge_blocks(orc.KIND_RAYS, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie[sel], je[sel],
                                g["idx_ii2jj"].numpy()[sel], g["valid_match"].numpy()[sel], g["Q"].numpy()[sel],
                                lo["C_conf"], lo["Q_conf"], sigma_ray=lo["sigma_ray"], sigma_dist=lo["sigma_dist"])
    _, _, H, b = orc.gn_solve(Hs, gs, ie[sel] - 1, je[sel] - 1, nkf - 1, return_system=True)
    sysbuf = torch.from_numpy(np.concatenate([H.ravel(), b]))
    md.allreduce_system(sysbuf)
    dim = 7 * (nkf - 1)
    Hr, br = sysbuf.numpy()[:dim * dim].reshape(dim, dim), sysbuf.numpy()[dim * dim:]
    dx = -np.linalg.solve(Hr, br)
    if rank == 0:
        Hs_a, gs_a = orc.gn_edge_blocks(orc.KIND_RAYS, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie, je,
                                        g["idx_ii2jj"].numpy(), g["valid_match"].numpy(), g["Q"].numpy(), lo["C_conf"],
                                        lo["Q_conf"], sigma_ray=lo["sigma_ray"], sigma_dist=lo["sigma_dist"])
        dx_ref, fail = orc.gn_solve(Hs_a, gs_a, ie - 1, je - 1, nkf - 1)
        np.save(out, np.array([np.linalg.norm(dx.reshape(-1, 7) - dx_ref) / np.linalg.norm(dx_ref), float(fail)]))
    dist.barrier()
    dist.destroy_process_group()


def test_partition_helpers():
    from mast3r_slam_ysh_b200 import distributed as md
    parts = md.partition_edges(10, 4)
    assert parts == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert md.partition_edges(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    sel = md.directed_slice(10, 3, 6)
    assert sel.tolist() == [3, 4, 5, 13, 14, 15]
    allsel = torch.cat([md.directed_slice(10, a, b) for a, b in parts]).sort().values
    assert allsel.tolist() == list(range(20))


def test_sharded_system_allreduce_gloo(tmp_path):
    out = str(tmp_path / "res.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    rel, fail = np.load(out)
    assert fail == 0 and rel < 1e-6  # dx of the oracle is rounded to fp32


def _gather_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mast3r_slam_ysh_b200 import distributed as md
    lo, hi = md.partition_pairs(5, world)[rank]           # 5 pairs over 2 ranks: 3 + 2 (ragged)
    idx = torch.arange(lo, hi)[:, None].repeat(1, 4) * 10 + torch.arange(4)[None, :]
    valid = (idx % 3 == 0)[..., None]
    gi, gv = md.gather_ragged(idx), md.gather_ragged(valid)
    if rank == 1:
        want = torch.arange(5)[:, None].repeat(1, 4) * 10 + torch.arange(4)[None, :]
        ok = torch.equal(gi, want) and torch.equal(gv, (want % 3 == 0)[..., None]) and gv.dtype == torch.bool
        np.save(out, np.array([float(ok), gi.shape[0]]))
    dist.barrier()
    dist.destroy_process_group()


def test_pair_sharded_gather_ragged_gloo(tmp_path):
    """Pair-sharded matching (configs[4]: 256 pair matches split by pair): the optional result gather keeps rank order
    and handles ragged pair counts; world size 2 over gloo."""
    out = str(tmp_path / "gather.npy")
    port = 31500 + os.getpid() % 2000
    mp.spawn(_gather_worker, args=(2, port, out), nprocs=2, join=True)
    ok, n = np.load(out)
    assert ok == 1.0 and n == 5
    from mast3r_slam_ysh_b200 import distributed as md
    assert md.partition_pairs(256, 8) == [(32 * r, 32 * r + 32) for r in range(8)]
    t = torch.arange(6)
    assert md.gather_ragged(t) is t  # no process group: identity
