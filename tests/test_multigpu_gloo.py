"""The N>1 path on CPU: two processes (gloo), each accumulates its shard of the plan into a partial dose grid and
the grids are summed with the package's reduce (raytracedicom_b200.sharding.reduce_dose).  The shards' doses come
from the CPU oracle here (the CUDA engine has no CPU fallback); what is under test is the partition, the
layer-shard semantics and the collective plumbing that bench.py uses unchanged with NCCL on the GPUs."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, rel_to_max


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    import sys
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "2"
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from raytracedicom_b200 import cases, sharding
    from oracle import pyoracle
    case = cases.make_case("mini")
    # two beams (the second one mirrored weights) so both beam- and layer-splitting occur
    b0 = case.beams[0]
    b1 = cases.Beam(b0.weights[:, ::-1, ::-1].copy(), b0.energies, b0.sigmas, b0.ray_spacing, b0.steps, b0.source_dist, b0.spot_delta,
                    b0.spot_offset, b0.gantry_to_im_m, b0.gantry_to_im_v)
    beams = [b0, b1]
    parts = sharding.partition(beams, case.luts, world, chunks_per_beam=3)
    mine = sharding.shard_beams(beams, parts[rank])
    nx, ny, nz = case.dose_dims
    dose = np.zeros((nz, ny, nx), np.float32)
    updates = 0
    for b in mine:
        r = pyoracle.run_beam(cases.Case("shard", case.ct, [b], case.luts), 0, dose=dose)
        updates += r["updates"]
    t = torch.from_numpy(dose)
    sharding.reduce_dose(t, dst=0)
    u = torch.tensor([updates], dtype=torch.int64)
    dist.reduce(u, dst=0)
    if rank == 0:
        whole = np.zeros_like(dose); total_u = 0
        for b in beams:
            r = pyoracle.run_beam(cases.Case("whole", case.ct, [b], case.luts), 0, dose=whole)
            total_u += r["updates"]
        np.savez(out_path, reduced=t.numpy(), whole=whole, u=int(u.item()), total_u=total_u,
                 n_items=np.array([len(p) for p in parts]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_plan_sums_to_whole_plan(tmp_path):
    out = str(tmp_path / "r.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    assert int(r["u"]) == int(r["total_u"])                    # the metric's unit count is shard-invariant
    assert all(n >= 1 for n in r["n_items"])
    assert rel_to_max(r["reduced"], r["whole"])[0] < 1e-6        # fp32 sum order differs between 1 and 2 ranks


def test_reduced_dose_slab_contains_every_beam(oracle):
    """sharding.dose_slab: the z-slab that is zeroed / reduced over NCCL is derived from geometry alone and must contain
    every voxel any beam can reach - checked with the oracle on two oblique diverging beams in a tall dose grid."""
    from raytracedicom_b200 import cases, sharding
    luts = cases.load_packaged_luts()
    nx, ny, nz = 96, 96, 200
    ct = np.full((nz, ny, nx), 1000.0, np.float32)
    beams = []
    for ang, zc in ((20.0, 70.0), (250.0, 120.0)):
        m, v = cases.gantry_to_im_idx(ang, (48.0, 48.0, zc), (0.0, 0.0, 0.0), (1.0, 1.0, 1.0))
        zstart, steps = cases.steps_and_start(m, v, (nx, ny, nz))
        w = np.full((2, 3, 3), 60.0, np.float32)
        beams.append(cases.Beam(w, np.array([75.0, 80.0], np.float32), np.full((2, 2), 3.0, np.float32), (1.0, 1.0), steps, (2000.0, 1800.0),
                                (4.0, 4.0, -1.0), (-4.0, -4.0, zstart), m, v))
    case = cases.Case("slab", ct, beams, luts)
    z0, z1 = sharding.dose_slab(case.beams, case.dose_dims)
    assert 0 < z0 < z1 < nz - 1                       # a real restriction for this plan
    dose = np.zeros((nz, ny, nx), np.float32)
    for b in range(2):
        oracle.run_beam(case, b, dose=dose)
    zs = np.nonzero(dose.reshape(nz, -1).any(axis=1))[0]
    assert zs.size > 0 and z0 <= zs.min() and zs.max() <= z1
    # each beam alone is inside its own slab too, and the plan's slab is the union
    s0, s1 = sharding.dose_slab(case.beams[:1], case.dose_dims), sharding.dose_slab(case.beams[1:], case.dose_dims)
    assert (min(s0[0], s1[0]), max(s0[1], s1[1])) == (z0, z1)
