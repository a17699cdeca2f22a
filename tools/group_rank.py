#!/usr/bin/env python
"""One rank of a multi-process plan through the library's group API (rtd_group_create / rtd_plan_dose_group), without
torch.distributed: the 64-byte IPC handle travels through a file. Used by tests/test_gpu_parity.py to run two ranks on ONE GPU
(two processes, CUDA IPC between them, the device-memory barriers time-sliced by the driver) and by hand on several GPUs.
usage: group_rank.py <rank> <world> <case> <device> <workdir> [host|device|relay]
(relay = host CT through the group's CT relay; the CT array is filled with NaN first, so a texel the relay missed shows in the dose)"""
import os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytracedicom_b200 import cases, engine


def main():
    rank, world, name, device, work = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5]
    mode = sys.argv[6] if len(sys.argv) > 6 else "device"
    case = cases.make_case(name)
    nx, ny, nz = case.dose_dims
    cnz, cny, cnx = case.ct.shape
    eng = engine.DoseEngine(device)
    eng.set_luts(case.luts); eng.set_dose_grid(nx, ny, nz)
    if mode == "device":
        eng.set_ct(case.ct)
    hpath = os.path.join(work, "handle.bin")
    if rank == 0:
        h = eng.group_create(0, world, nz * ny * nx)
        with open(hpath + ".tmp", "wb") as f:
            f.write(h)
        os.replace(hpath + ".tmp", hpath)
    else:
        t0 = time.time()
        while not os.path.exists(hpath):
            if time.time() - t0 > 60:
                raise SystemExit("rank 0 never published the handle")
            time.sleep(0.05)
        eng.group_create(rank, world, nz * ny * nx, open(hpath, "rb").read())
    if mode == "relay":
        eng.set_ct(np.full(case.ct.shape, np.nan, np.float32))
        mine = eng.group_ct_create(64 << 20)
        with open(os.path.join(work, f"ct_handle_{rank}.tmp"), "wb") as f:
            f.write(mine)
        os.replace(os.path.join(work, f"ct_handle_{rank}.tmp"), os.path.join(work, f"ct_handle_{rank}.bin"))
        t0 = time.time()
        while not all(os.path.exists(os.path.join(work, f"ct_handle_{r}.bin")) for r in range(world)):
            if time.time() - t0 > 60:
                raise SystemExit("not every rank published its CT relay handle")
            time.sleep(0.05)
        eng.group_ct_open([open(os.path.join(work, f"ct_handle_{r}.bin"), "rb").read() for r in range(world)])
        mode = "host"
    packed, keep = engine.pack_beams(case.beams)
    host = np.full((nz, ny, nx), 0.25, np.float32)
    ct = np.ascontiguousarray(case.ct, np.float32)
    upd = 0
    for rep in range(2):          # the second call checks that the barrier epochs keep in step
        if mode == "host":
            t = eng.plan_dose_group(ct.ctypes.data, cnx, cny, cnz, packed, len(case.beams), host.ctypes.data if rank == 0 else 0)
        else:
            t = eng.plan_dose_group(0, cnx, cny, cnz, packed, len(case.beams))
        upd = int(t.updates)
    if rank == 0:
        import torch

        class _Ext:                       # zero-copy torch view of the group's grid
            __cuda_array_interface__ = {"shape": (nz, ny, nx), "typestr": "<f4", "data": (eng.group_dose_ptr(), False), "version": 2}
        out = torch.as_tensor(_Ext(), device=f"cuda:{device}").cpu().numpy()
        np.save(os.path.join(work, "group_dose.npy"), out)
        if mode == "host":
            np.save(os.path.join(work, "host_dose.npy"), host)
    with open(os.path.join(work, f"done_{rank}.txt"), "w") as f:
        f.write(str(upd))
    # non-owners unmap before the owner frees
    if rank != 0:
        eng.group_destroy()
        open(os.path.join(work, f"closed_{rank}.txt"), "w").write("1")
    else:
        t0 = time.time()
        while not all(os.path.exists(os.path.join(work, f"closed_{r}.txt")) for r in range(1, world)) and time.time() - t0 < 30:
            time.sleep(0.05)
        eng.group_destroy()
    eng.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
