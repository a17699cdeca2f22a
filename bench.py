#!/usr/bin/env python
"""Benchmark of the proton dose path (BASELINE.json: plan dose-calc ms & spot-voxel updates/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL)

A "step" = one whole-plan dose calculation (all beams, all energy layers) of the named synthetic workload.
  value  : spot-voxel updates/s with CT and LUTs already resident in HBM (device dose buffer zeroed, all kernels,
           partial-dose reduce over NCCL when N > 1), timed with CUDA events on the engine's stream, max over ranks.
  e2e    : the same metric through the reference-facing host call (rtd_dose == cudaWrapperProtons' coverage):
           rtd_plan_dose with pinned HOST CT + HOST dose: H2D of the CT / dose-in sub-volumes the plan touches, all beams,
           D2H of the dose sub-volumes, every step.
  roofline / cpu_baseline: see DESIGN.md "Measurement".
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "spot-voxel updates/s (plan dose calculation)"
UNIT = "updates/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                r = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                   capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.rows.append([x.strip() for x in r.stdout.strip().splitlines()[0].split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")):
            if any(r[3 + i].lower().startswith("active") for r in self.rows if len(r) > 3 + i):
                reasons.append(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_baseline_port(case, budget_s=20.0):
    """CPU restatement (oracle/, test infrastructure) timed on the host cores on a bounded sample of the workload:
    whole beams of the plan, one after the other, until the time budget is used (at least one beam)."""
    from oracle import pyoracle
    from raytracedicom_b200 import cases
    cores = os.cpu_count() or 1
    pyoracle.lib()
    upd, dt, nb = 0, 0.0, 0
    dose = np.zeros((case.dose_dims[2], case.dose_dims[1], case.dose_dims[0]), np.float32)
    for b in case.beams:
        t0 = time.perf_counter()
        r = pyoracle.run_beam(cases.Case("sample", case.ct, [b], case.luts), 0, dose=dose)
        dt += time.perf_counter() - t0
        upd += r["updates"]; nb += 1
        if dt * (nb + 1) / nb > budget_s * 1.5:
            break
    return {"value": upd / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"beams [0,{nb}) of {len(case.beams)} of workload {case.name}, all energy layers: {upd} updates in {dt:.2f} s "
                      f"(oracle/rtd_oracle.c, OpenMP, {cores} threads; tracer + pre-convolution + IDD/sigma + superposition + transfer)"}


def cpu_stage_baseline(case, gpu_stage_ms=None):
    """"Baseline B" (BASELINE.md 2B): the reference's OWN CPU stage code, src/cpu_convolution_1d.cpp (xConvCpuSparse +
    yConvCpuSparse, compiled unmodified into oracle/_ref/libref_cpuconv.so), timed on stage K3a (spot weights -> ray weights)
    of every beam of the workload on this box's host cores: as shipped (1 thread) and with an OpenMP loop over the layers."""
    from oracle import ref_cpu_stage as rcs
    if not rcs.available():
        return {"unavailable": "oracle/_ref/libref_cpuconv.so not built (needs /root/reference at build time)"}
    try:
        probs = [rcs.k3a_problem(case, b) for b in range(len(case.beams))]
        t1 = sum(rcs.run(p, threads=1, repeats=3)[1] for p in probs)
        tn = sum(rcs.run(p, threads=0, repeats=3)[1] for p in probs)
        err = max(float(np.abs(rcs.run(p, 0)[0] - p["oracle_ray_weights"]).max() / p["oracle_ray_weights"].max()) for p in probs)
    except Exception as ex:          # noqa: BLE001 - e.g. non-integer spot pitch: the reference code cannot run the problem
        return {"unavailable": str(ex)[:200]}
    out = {"stage": "K3a spot weights -> ray weights, all beams and layers of the workload",
           "kind": "reference", "code": "src/cpu_convolution_1d.cpp xConvCpuSparse + yConvCpuSparse, g++ -O2",
           "ms_1_core": 1e3 * t1, "ms_all_cores": 1e3 * tn, "cores": rcs.max_threads(),
           "max_rel_diff_vs_oracle_stage": err}
    if gpu_stage_ms is not None:
        out["gpu_stage_ms"] = gpu_stage_ms
    return out


def run_reference(args, case, rank, world):
    """Reference arm: the UNMODIFIED reference (sources compiled in place from /root/reference into oracle/_ref by
    oracle/ref_harness/Makefile, sm_100) through its own entry point cudaWrapperProtons on this workload, one process
    per step (it ends with cudaDeviceReset).  Its own 'Total global execution time' covers H2D CT + LUTs, all beams,
    D2H dose and frees, i.e. an end-to-end figure.  It is single-GPU: rank 0 alone runs it."""
    if rank != 0:
        return None
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_dose")
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": case.name, "ct_dims": list(case.ct_dims), "beams": len(case.beams),
                       "layers": int(sum(b.n_layers for b in case.beams))}}
    from oracle import pyoracle
    cpu = cpu_baseline_port(case)
    line["cpu_baseline"] = cpu
    line["cpu_stage_baseline"] = cpu_stage_baseline(case)
    have_gpu = False
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        pass
    if os.path.exists(exe) and have_gpu:
        # U from the oracle-independent integer outputs is identical for both engines; take it from one run of ours
        from raytracedicom_b200.engine import DoseEngine
        nx, ny, nz = case.dose_dims
        with DoseEngine(0) as eng:
            eng.set_luts(case.luts); eng.set_ct(case.ct); eng.set_dose_grid(nx, ny, nz)
            U = eng.dose(case.beams, np.zeros((nz, ny, nx), np.float32)).updates
        tmp = tempfile.mkdtemp(prefix="rtdref_")
        cpath = os.path.join(tmp, "case.rtdcase"); case.write(cpath)
        times = []
        for i in range(args.warmup + args.steps):
            r = subprocess.run([exe, cpath, os.path.join(tmp, "dose.f32")], capture_output=True, text=True, timeout=600)
            ms = None
            for ln in r.stdout.splitlines():
                if ln.startswith("REF_TIMING"):
                    ms = float(ln.split("total_ms=")[1].split()[0])
            if r.returncode != 0 or ms is None or ms <= 0:
                line["unavailable_gpu_reference"] = (r.stderr or r.stdout)[-300:]
                break
            if i >= args.warmup:
                times.append(ms)
        if times:
            ms = sum(times) / len(times)
            line.update({"value": U / (ms * 1e-3), "ms_per_step": ms, "reference_kind": "reference rebuilt for sm_100 from /root/reference (oracle/_ref/ref_dose), "
                         "its own total timer: H2D CT+LUTs, all beams, D2H dose, frees; 1 GPU",
                         "e2e": {"value": U / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(case.ct.nbytes * 2), "d2h_bytes_per_step": int(case.ct.nbytes)},
                         "updates_per_step": int(U), "gpu_launches": None})
            return line
    # no rebuilt reference or no GPU: the CPU port is the reference arm
    line.update({"value": cpu["value"], "ms_per_step": None, "reference_kind": "CPU restatement (oracle port); rebuilt reference or GPU not available",
                 "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    return line


def main():
    import faulthandler
    faulthandler.enable()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", help="c3 (default: 512x512x200 CT, 3 fields), c1a, c2, c4, c5, mini")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--reduce", default="push", choices=["push", "peer", "nccl"],
                    help="N > 1, how the ranks' partial doses meet in ONE grid on rank 0 (mapped by every rank over CUDA IPC): 'push' = each "
                         "rank adds the float4 groups of its slab that hold dose with 16-byte vector reductions over NVLink; 'peer' = the "
                         "transfer kernels themselves add into the mapped grid (scalar reductions, NVLink packet-rate bound); "
                         "'nccl' = local partial grids + one dense NCCL reduce of the reachable slab")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))

    from raytracedicom_b200 import cases
    case = cases.make_case(args.workload)

    if args.impl == "reference":
        line = run_reference(args, case, rank, world)
        if rank == 0:
            print(json.dumps(line), flush=True)
        return 0

    import torch
    import torch.distributed as dist
    from raytracedicom_b200 import sharding
    from raytracedicom_b200.engine import DoseEngine, pack_beams
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the dose path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    eng = DoseEngine(local)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)
    nx, ny, nz = case.dose_dims
    cnx, cny, cnz = case.ct_dims
    eng.set_luts(case.luts)
    eng.set_dose_grid(nx, ny, nz)
    ct_host = torch.from_numpy(case.ct).pin_memory()
    ct_dev = ct_host.to(dev)
    torch.cuda.synchronize()
    eng.set_ct_ptr(ct_dev.data_ptr(), cnx, cny, cnz, device=True)
    parts = sharding.partition(case.beams, case.luts, world)
    mine = sharding.shard_beams(case.beams, parts[rank])
    packed, keep = pack_beams(mine)
    dose_dev = torch.zeros((nz, ny, nx), dtype=torch.float32, device=dev)
    dose_host = torch.zeros((nz, ny, nx), dtype=torch.float32).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    # N > 1: only the z-slab of the dose grid that the plan's beams can reach (geometry alone, same on every rank) is
    # zeroed and summed over NCCL; the rest of every partial grid stays zero (checked once after the warm-up)
    z0, z1 = (0, nz - 1) if world == 1 else sharding.dose_slab(case.beams, case.dose_dims)
    e2e_extra = {"mod": 0}
    # N > 1, --reduce peer: the transfer kernels of every rank accumulate straight into rank 0's grid over NVLink
    # (rtd_shared_dose_*, opts.atomicDose): the reduce is fused into the kernel that produces the data. Falls back to the
    # NCCL reduce when the IPC mapping is not available.
    peer = world > 1 and args.reduce in ("peer", "push")
    push = args.reduce == "push"
    shared_ptr, token = 0, torch.zeros(1, dtype=torch.int32, device=dev)
    if peer:
        handle = [None]
        try:
            if rank == 0:
                shared_ptr, handle[0] = eng.shared_dose_create(nz * ny * nx)
        except Exception as ex:          # noqa: BLE001 - reported in the JSON line
            handle[0] = f"error: {ex}"
        dist.broadcast_object_list(handle, src=0)
        ok = torch.ones(1, dtype=torch.int32, device=dev)
        if isinstance(handle[0], str):
            ok.zero_()
        elif rank != 0:
            try:
                shared_ptr = eng.shared_dose_open(handle[0])
            except Exception:            # noqa: BLE001
                ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        peer = bool(ok.item())
    if peer:
        class _Ext:                       # zero-copy torch view of the shared grid (rank 0 zeroes / reads it through torch)
            __cuda_array_interface__ = {"shape": (nz, ny, nx), "typestr": "<f4", "data": (shared_ptr, False), "version": 2}
        if rank == 0:
            dose_dev = torch.as_tensor(_Ext(), device=dev)
        # push: only rank 0 writes its grid while pushes of the others may land in it -> its own transfers are atomic too
        eng.opts.atomicDose = 1 if (not push or rank == 0) else 0
    push = push and peer
    slab = dose_dev[z0:z1 + 1]
    slab_off = z0 * ny * nx * 4

    def dose_target():
        return shared_ptr if (peer and (rank == 0 or not push)) else dose_dev.data_ptr()

    def before_dose():
        if not peer or push or rank == 0:
            slab.zero_()
        if peer:
            stream_barrier()                                      # rank 0's grid is zeroed before anything is added into it

    def after_dose():
        if push and rank != 0:
            eng.push_dose_to_peer(slab.data_ptr(), shared_ptr + slab_off, slab.numel())
        if peer:
            stream_barrier()                                      # every rank's contributions have landed
        else:
            sharding.reduce_dose(slab, dst=0)

    def stream_barrier():
        # orders the streams of all ranks without a host sync: a 4-byte all-reduce on the engine's stream
        dist.all_reduce(token)

    def step_device():
        with torch.cuda.stream(stream):
            before_dose()
            t = eng.dose_ptr(packed, len(mine), dose_target(), device=True)
            after_dose()
        return t

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    barrier()
    if world > 1 and (not peer or push or rank == 0) and (bool(dose_dev[:z0].any().item()) or bool(dose_dev[z1 + 1:].any().item())):
        raise SystemExit(f"bench.py: rank {rank} deposited dose outside the reduced slab [{z0}, {z1}]")
    sampler = ClockSampler(local); sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    stage = {}
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        with torch.cuda.stream(stream):
            flush.fill_(i & 0xFF)                     # L2 flush between timed iterations (outside the event pair)
            evs[i][0].record(stream)
        tm = step_device()
        with torch.cuda.stream(stream):
            evs[i][1].record(stream)
        for k, v in tm.as_dict().items():
            stage[k] = stage.get(k, 0.0) + v
    barrier()
    t_wall = time.perf_counter() - t_wall0
    ms_local = sum(a.elapsed_time(b) for a, b in evs) / args.steps
    ms_each_local = [round(a.elapsed_time(b), 4) for a, b in evs]      # rank 0's own per-step times (spread / outliers)
    # stage timings / roofline: a few extra steps with the beams serialised on one stream, so that the library's stage
    # events bracket each kernel alone (with concurrent beams the stages of different beams overlap in time)
    conc_stage = dict(stage)
    eng.opts.serialBeams = 1
    stage = {}
    n_serial = 3
    step_device()
    for _ in range(n_serial):
        with torch.cuda.stream(stream):
            flush.fill_(1)
        tm = step_device()
        for k, v in tm.as_dict().items():
            stage[k] = stage.get(k, 0.0) + v
    eng.opts.serialBeams = 0
    for k in stage:
        stage[k] = stage[k] * args.steps / n_serial      # same normalisation as the timed loop below
    torch.cuda.synchronize()
    upd_local = torch.tensor([int(stage["updates"] / args.steps), int(stage["ksTaps"] / args.steps)], dtype=torch.int64, device=dev)
    ms_t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(upd_local, op=dist.ReduceOp.SUM)
    ms_step = float(ms_t.item()); U = int(upd_local[0].item()); taps = int(upd_local[1].item())
    # balance of the sharding: every rank's own stage total of the serialised extra steps (no waiting on the others in it)
    own_ms = torch.tensor([stage["msTotal"] / args.steps], dtype=torch.float64, device=dev)
    own_all = [torch.zeros_like(own_ms) for _ in range(world)]
    if world > 1:
        dist.all_gather(own_all, own_ms)
    else:
        own_all = [own_ms]
    rank_stage_ms = [round(float(t.item()), 4) for t in own_all]
    checksum = float(dose_dev.double().sum().item()) if rank == 0 else 0.0
    dmax = float(dose_dev.max().item()) if rank == 0 else 0.0

    # ---- end-to-end through the host-facing call, host buffers, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        def step_e2e():
            # the call a user of the reference makes: cudaWrapperProtons(imVol, doseVol, beams, ...) == rtd_plan_dose with
            # HOST CT and HOST dose; the library copies the sub-volumes the plan touches (bytes reported by the library)
            if world == 1:
                return eng.plan_dose_ptr(ct_host.data_ptr(), cnx, cny, cnz, packed, len(mine), dose_host.data_ptr())
            # N > 1: every rank uploads the CT sub-volumes of its own work items and computes them into its partial grid;
            # one NCCL reduce of the reachable slab; rank 0 adds the summed dose onto the caller's pinned host grid in
            # place (rtd_add_device_to_host: only voxels that received dose cross PCIe)
            with torch.cuda.stream(stream):
                before_dose()
                t = eng.plan_dose_ptr(ct_host.data_ptr(), cnx, cny, cnz, packed, len(mine), dose_target(), device=True)
                after_dose()
            if rank == 0:
                e2e_extra["mod"] = eng.add_device_to_host(slab.data_ptr(), dose_host.data_ptr() + z0 * ny * nx * 4, slab.numel())
            else:
                stream.synchronize()
            return t
        n_e2e = max(3, min(args.steps, 10))
        for _ in range(2):
            dose_host.zero_(); step_e2e()
        tt = []
        for _ in range(n_e2e):
            dose_host.zero_()
            barrier(); t0 = time.perf_counter(); te = step_e2e(); barrier(); tt.append(time.perf_counter() - t0)
        e_ms = torch.tensor([1e3 * sum(tt) / len(tt)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
        io = torch.tensor([int(te.h2dBytes) + 4 * e2e_extra["mod"], int(te.d2hBytes) + 4 * e2e_extra["mod"]], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(io, op=dist.ReduceOp.SUM)
        e2e = {"value": U / (float(e_ms.item()) * 1e-3), "unit": UNIT, "ms_per_plan": float(e_ms.item()),
               "h2d_bytes_per_step": int(io[0].item()), "d2h_bytes_per_step": int(io[1].item()),
               "dense_bytes": {"h2d": int(ct_host.numel() * 4 + dose_host.numel() * 4), "d2h": int(dose_host.numel() * 4),
                               "note": "what the reference moves: whole CT + whole dose grid up, whole dose grid down"},
               "timing": "host wall clock between barriers around the host call rtd_plan_dose (HOST CT + HOST dose in, dose out; the "
                         "library copies only the CT / dose sub-volumes the plan touches), max over ranks",
               "library_events_ms": {"total": float(te.msTotal), "ct_upload_done_at": float(te.msUpload), "download": float(te.msDownload)},
               "host_syncs_per_plan": int(te.hostSyncs), "dose_voxels_modified": int(te.doseVoxelsModified),
               "dose_checksum_matches_device_run": None}
        if rank == 0:
            cs = float(dose_host.double().sum().item())
            e2e["dose_checksum_matches_device_run"] = bool(abs(cs - checksum) <= 1e-5 * abs(checksum))
        eng.set_ct_ptr(ct_dev.data_ptr(), cnx, cny, cnz, device=True)

    sampler.stop_flag = True; sampler.join(timeout=2)   # sampled over the device-resident and the end-to-end timed regions
    if rank == 0:
        peak, peak_src, sm_max = load_peaks()
        reps = [eng.beam_report(b) for b in range(len(mine))]
        live = [r for r in reps if r["beamFirstCalculatedPassive"] > r["beamFirstInside"]]
        # on C3 two ks2 launches per live beam do the work (radius classes <= 8 and 9..16); the launches of the 17..24 and
        # 25..32 classes find an empty item list there (a few microseconds inside msSuperp). Workloads with deep targets
        # (C2, C4, C5) also use those classes; their roofline line still divides by two launches per beam.
        ks_launches = max(1, 2 * len(live))
        traffic, traffic_src = None, None
        tp = os.path.join(ROOT, "profiles", "r01_ncu_ks2_traffic.json")
        if os.path.exists(tp) and case.name.startswith("c3") and world == 1:
            tj = json.load(open(tp)); traffic = tj["dram_bytes_per_launch"]; traffic_src = tj["source"]
        ks_ms = stage["msSuperp"] / args.steps
        u_rank0 = sum(r["updates"] for r in reps)
        bev_bytes = sum(4 * (r["rayDimX"] + 64) * (r["rayDimY"] + 64) * (r["beamFirstCalculatedPassive"] - r["beamFirstInside"]) for r in live)
        ks_alg = 8 * u_rank0 + bev_bytes          # read idd + rSigmaEff once per update, write the haloed BEV dose once
        taps0 = stage["ksTaps"] / args.steps
        achieved = ks_alg / (ks_ms * 1e-3) / 1e9 if ks_ms > 0 else 0.0
        fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        # whole-plan algorithmic bytes
        plan_bytes = 4 * nx * ny * nz
        for b, r in zip(mine, reps):
            R = r["rayDimX"] * r["rayDimY"]
            box = max(0, r["bboxMax"][0] - r["bboxMin"][0] + 1) * max(0, r["bboxMax"][1] - r["bboxMin"][1] + 1) * max(0, r["bboxMax"][2] - r["bboxMin"][2] + 1)
            plan_bytes += 4 * box + 8 * R * r["steps"] + 8 * r["updates"] + 8 * (r["rayDimX"] + 64) * (r["rayDimY"] + 64) * max(0, r["beamFirstCalculatedPassive"] - r["beamFirstInside"]) + 8 * box
        line = {"metric": METRIC, "value": U / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": case.name, "ct_dims": list(case.ct_dims), "beams": len(case.beams), "layers": int(sum(b.n_layers for b in case.beams)),
                           "spots": int(sum(b.weights.size for b in case.beams)), "updates_per_plan": U, "ks_taps_per_plan": taps,
                           "sharding": (f"(beam, layer-range) work items, LPT over {world} rank(s); " + ("every rank adds the dose-holding float4 groups of its slab into ONE grid on rank 0 over NVLink (CUDA IPC, red.global.add.v4.f32), no dense collective" if push else "every rank's transfer kernels add into ONE dose grid on rank 0 over NVLink (CUDA IPC, red.global.add.f32), no separate collective" if peer else "partial dose grids (reachable z-slab) summed by one NCCL reduce")) if world > 1 else "single GPU",
                           "l2": "256 MiB buffer written between timed iterations (L2 flush); CT 4*%d B and BEV arrays exceed L2" % (nx * ny * nz)},
                "plan_ms": ms_step, "updates_per_step": U,
                "ms_each_step_rank0": ms_each_local, "stage_total_ms_per_rank": rank_stage_ms, "work_items_per_rank": [len(p) for p in parts],
                "stage_ms_rank0": {k: stage[k] / args.steps for k in ("msTrace", "msSetup", "msConv", "msIddSigma", "msRadius", "msSuperp", "msTransfer", "msTotal")},
                "stage_note": "stage_ms_rank0 and roofline come from 3 extra steps with opts.serialBeams=1 (stages bracketed alone); the timed steps run one stream per beam",
                "serial_beams_plan_ms": stage["msTotal"] / args.steps,
                "gpu_launches": int(conc_stage["kernelLaunches"]), "host_syncs_per_plan": int(stage["hostSyncs"] / args.steps),
                "dose_max": dmax, "dose_checksum": checksum,
                "roofline": {"kernel": "ks2_kernel (layer-fused kernel superposition, 2 radius-class launches per beam)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                             "traffic_note": "DRAM bytes per launch from ncu; below the algorithmic 8 B/update because tiles without dose and layers past their range are skipped before their idd / rSigmaEff are read",
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": ks_alg / ks_launches,
                             "avg_launch_ms": ks_ms / ks_launches, "launches_per_plan_rank0": ks_launches,
                             "fp32": {"achieved_tflops": 2 * taps0 / (ks_ms * 1e-3) / 1e12 if ks_ms > 0 else 0.0, "peak_tflops": fp32_peak,
                                      "frac": (2 * taps0 / (ks_ms * 1e-3) / 1e12) / fp32_peak if ks_ms > 0 else 0.0,
                                      "note": "superposition is FMA-bound: 2 x taps / kernel time vs 148 SM x 128 FMA/clk x sm_max_mhz"}},
                "plan_roofline": {"algorithmic_bytes": plan_bytes, "achieved_gbs": plan_bytes / (ms_step * 1e-3) / 1e9, "frac_of_hbm_peak": plan_bytes / (ms_step * 1e-3) / 1e9 / peak,
                                  "note": "rank 0's share of the plan over the whole step" if world > 1 else "whole plan"},
                "clocks": sampler.summary(), "wall_s_timed_region": t_wall}
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_port(case)
            line["cpu_stage_baseline"] = cpu_stage_baseline(case, gpu_stage_ms=stage["msConv"] / args.steps)
        print(json.dumps(line), flush=True)
    # tear down in dependency order: torch tensors that were used on the engine's stream must die before the stream does
    del evs, ms_t, upd_local
    dose_dev = ct_dev = flush = None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    # torch's pinned-memory and device allocators record events on every stream a block was used on when the block is
    # released, including the engine's stream (ExternalStream): leave the context to process exit instead of destroying
    # that stream under them.
    sys.stdout.flush(); sys.stderr.flush()
    os._exit(0)


if __name__ == "__main__":
    sys.exit(main())
