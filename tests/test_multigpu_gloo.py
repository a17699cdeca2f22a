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
