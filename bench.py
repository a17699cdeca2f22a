#!/usr/bin/env python
"""Benchmark of the proton dose path (BASELINE.json: plan dose-calc ms & spot-voxel updates/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL)

A "step" = one whole-plan dose calculation (all beams, all energy layers) of the named synthetic workload.
  value  : spot-voxel updates/s with CT and LUTs already resident in HBM (device dose buffer zeroed, all kernels,
           partial doses of the ranks summed into one grid when N > 1), timed with CUDA events on the engine's stream,
           max over ranks.
  e2e    : the same metric through the reference-facing host call (cudaWrapperProtons' coverage):
           rtd_plan_dose with pinned HOST CT + HOST dose: H2D of the CT sub-volumes the plan touches, all beams,
           dose accumulated onto the host grid, every step, on a warm context.
  e2e_cold: the drop-in call as include/rtdose_shim.hpp makes it, in a fresh process: rtd_create -> rtd_set_luts ->
           rtd_plan_dose -> rtd_destroy, wall clock (context creation excluded, like the reference's own timer).
  c4     : the same measurement on BASELINE.json's "beams sharded across 8 GPUs" config, as an extra key.
  roofline / cpu_baseline: see DESIGN.md "Measurement".
Prints ONE JSON line on rank 0 and exits normally (the driver's exit hook records which .so files were loaded).
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
DOSE_TOL_VS_1GPU = 1e-5      # N > 1: max |D_N - D_1| / max D_1 above this fails the run


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def config_of(case):
    """The `config` object: identical in both arms (--impl ours / reference)."""
    return {"workload": case.name, "ct_dims": [int(v) for v in case.ct_dims], "beams": len(case.beams),
            "layers": int(sum(b.n_layers for b in case.beams)), "spots": int(sum(b.weights.size for b in case.beams))}


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

    def stop(self):
        self.stop_flag = True
        self.join(timeout=6)

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


def cpu_baseline_port(case, budget_s=20.0, all_beams=False):
    """CPU restatement (oracle/, test infrastructure) timed on the host cores on a bounded sample of the workload:
    whole beams of the plan, one after the other, until the time budget is used (at least one beam).
    all_beams: keep going (untimed) through the remaining beams so that the returned update count U covers the whole plan -
    the integer outputs of the oracle are pinned to the reference's (tests/test_oracle_golden.py), so U is the reference's U."""
    from oracle import pyoracle
    from raytracedicom_b200 import cases
    cores = os.cpu_count() or 1
    pyoracle.lib()
    upd, dt, nb, u_all = 0, 0.0, 0, 0
    dose = np.zeros((case.dose_dims[2], case.dose_dims[1], case.dose_dims[0]), np.float32)
    timing = True
    for b in case.beams:
        t0 = time.perf_counter()
        r = pyoracle.run_beam(cases.Case("sample", case.ct, [b], case.luts), 0, dose=dose)
        u_all += r["updates"]
        if timing:
            dt += time.perf_counter() - t0
            upd += r["updates"]; nb += 1
            if dt * (nb + 1) / nb > budget_s * 1.5:
                timing = False
                if not all_beams:
                    break
    out = {"value": upd / dt, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"beams [0,{nb}) of {len(case.beams)} of workload {case.name}, all energy layers: {upd} updates in {dt:.2f} s "
                     f"(oracle/rtd_oracle.c, OpenMP, {cores} threads; tracer + pre-convolution + IDD/sigma + superposition + transfer)"}
    return out, (u_all if all_beams else None)


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
    D2H dose and frees, i.e. an end-to-end figure.  It is single-GPU: rank 0 alone runs it.
    Nothing of this repo's engine runs or is loaded here: the update count U comes from the CPU oracle."""
    if rank != 0:
        return None
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_dose")
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(case)}
    cpu, U = cpu_baseline_port(case, all_beams=True)
    line["cpu_baseline"] = cpu
    line["cpu_stage_baseline"] = cpu_stage_baseline(case)
    line["updates_per_step"] = int(U)
    have_gpu = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).returncode == 0 if _which("nvidia-smi") else False
    if os.path.exists(exe) and have_gpu:
        tmp = tempfile.mkdtemp(prefix="rtdref_")
        cpath = os.path.join(tmp, "case.rtdcase"); case.write(cpath)

        def run_steps(n_warm, n_steps, env=None):
            times = []
            for i in range(n_warm + n_steps):
                r = subprocess.run([exe, cpath, os.path.join(tmp, "dose.f32")], capture_output=True, text=True, timeout=600, env=env)
                ms = None
                for ln in r.stdout.splitlines():
                    if ln.startswith("REF_TIMING"):
                        ms = float(ln.split("total_ms=")[1].split()[0])
                if r.returncode != 0 or ms is None or ms <= 0:
                    line["unavailable_gpu_reference"] = (r.stderr or r.stdout)[-300:]
                    return None
                if i >= n_warm:
                    times.append(ms)
            return times
        times = run_steps(args.warmup, args.steps)
        if times:
            ms = sum(times) / len(times)
            line.update({"value": U / (ms * 1e-3), "ms_per_step": ms, "ms_each_step": [round(t, 3) for t in times],
                         "reference_kind": "reference rebuilt for sm_100 from /root/reference (oracle/_ref/ref_dose), one fresh process per step "
                         "(cudaWrapperProtons ends with cudaDeviceReset), its own total timer: LUT + CT + dose H2D, ~35 cudaMalloc/cudaFree per "
                         "beam, lazy module load of its kernels, all beams, D2H dose, frees; 1 GPU",
                         "e2e": {"value": U / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(case.ct.nbytes * 2), "d2h_bytes_per_step": int(case.ct.nbytes)},
                         "gpu_launches": None})
            # the same binary with its kernels loaded eagerly: takes the lazy module load of 33 kernelSuperposition instantiations
            # out of its timer (context creation is outside it either way)
            eager = run_steps(0, max(1, min(args.steps, 3)), env=dict(os.environ, CUDA_MODULE_LOADING="EAGER"))
            if eager:
                em = sorted(eager)[len(eager) // 2]       # median: a side figure, and single runs of the binary scatter (50 .. 1500 ms seen)
                line["reference_eager_module_loading"] = {"ms_per_step": em, "ms_each_step": [round(t, 3) for t in eager], "value": U / (em * 1e-3),
                                                          "env": "CUDA_MODULE_LOADING=EAGER"}
            return line
    # no rebuilt reference or no GPU: the CPU port is the reference arm
    line.update({"value": cpu["value"], "ms_per_step": None, "reference_kind": "CPU restatement (oracle port); rebuilt reference or GPU not available",
                 "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    return line


def _which(name):
    import shutil
    return shutil.which(name)


def cold_child(args):
    """Fresh-process measurement of the drop-in call (include/rtdose_shim.hpp): rtd_create -> rtd_set_luts -> rtd_set_dose_grid ->
    rtd_plan_dose (pinned HOST CT + HOST dose) -> rtd_destroy, host wall clock. The CUDA context exists before the timer starts
    (the reference's timer starts after its cudaFree(0), kernel_wrapper.cu:388-415); loading the library's kernels, creating the
    LUT textures and the CT array and growing the workspaces are inside."""
    import torch
    from raytracedicom_b200 import cases
    from raytracedicom_b200.engine import DoseEngine, pack_beams
    case = cases.make_case(args.workload)
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda:0"); torch.cuda.synchronize()
    nx, ny, nz = case.dose_dims; cnx, cny, cnz = case.ct_dims
    ct_host = torch.from_numpy(case.ct).pin_memory()
    dose_host = torch.zeros((nz, ny, nx), dtype=torch.float32).pin_memory()
    packed, keep = pack_beams(case.beams)
    t0 = time.perf_counter()
    eng = DoseEngine(0)
    eng.set_luts(case.luts)
    eng.set_dose_grid(nx, ny, nz)
    t1 = time.perf_counter()
    tm = eng.plan_dose_ptr(ct_host.data_ptr(), cnx, cny, cnz, packed, len(case.beams), dose_host.data_ptr())
    t2 = time.perf_counter()
    upd = int(tm.updates)
    eng.close()
    t3 = time.perf_counter()
    print(json.dumps({"cold_ms": 1e3 * (t3 - t0), "create_and_luts_ms": 1e3 * (t1 - t0), "plan_dose_ms": 1e3 * (t2 - t1), "destroy_ms": 1e3 * (t3 - t2),
                      "updates": upd, "dose_checksum": float(dose_host.double().sum().item())}), flush=True)
    del ct_host, dose_host
    return 0


class Runner:
    """Per-process state shared by the workloads measured in one run: device, engine (on a torch-owned stream), process group."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from raytracedicom_b200.engine import DoseEngine
        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1")); self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device - the dose path has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", rank=self.rank, world_size=self.world, device_id=self.dev)
        # torch owns the stream the engine works on (rtd_create_on_stream): torch's allocators may record events on it for as long
        # as the process lives, and the process can exit normally
        self.stream = torch.cuda.Stream(device=self.dev)
        self.eng = DoseEngine(self.local, stream=self.stream.cuda_stream)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)   # > 126 MB L2

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def close(self):
        self.flush = None
        self.torch.cuda.synchronize()
        self.eng.close()
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def measure(R, case, steps, warmup, do_e2e, do_cpu, full):
    """One workload on the runner's GPUs. Returns the result dict on rank 0 (None elsewhere)."""
    torch, dist, eng, args = R.torch, R.dist, R.eng, R.args
    from raytracedicom_b200 import sharding
    from raytracedicom_b200.engine import pack_beams
    rank, world, dev, stream = R.rank, R.world, R.dev, R.stream
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

    # N > 1: only the z-slab of the dose grid that the plan's beams can reach (geometry alone, same on every rank) is
    # zeroed and summed; the rest of every partial grid stays zero (checked once after the warm-up)
    z0, z1 = (0, nz - 1) if world == 1 else sharding.dose_slab(case.beams, case.dose_dims, sigma_cutoff=float(eng.opts.convSigmaCutoff))
    e2e_extra = {"mod": 0}
    # N > 1, --reduce push (default): the library's own multi-GPU call. rtd_group_create puts ONE dose grid + a control block on rank 0
    # and maps it on every other rank over CUDA IPC; rtd_plan_dose_group partitions the plan, computes this rank's work items and
    # sums the partial doses in that grid (vector reductions over NVLink, two stream-ordered barriers in device memory - no NCCL
    # call and no torch op inside the step). --reduce nccl: the partition is applied here and the partial grids meet in one dense
    # NCCL reduce (the baseline the default is measured against).
    group = world > 1 and args.reduce == "push"
    packed_all, keep_all = pack_beams(case.beams)
    if group:
        handle = [None]
        if rank == 0:
            handle[0] = eng.group_create(0, world, nz * ny * nx)
        dist.broadcast_object_list(handle, src=0)
        if rank != 0:
            eng.group_create(rank, world, nz * ny * nx, handle[0])
        if args.ct_relay:
            # CT relay: the e2e call uploads every CT box of the plan once per node; the ranks fetch what they need over NVLink
            relay_bytes = int(case.ct.nbytes // world + (64 << 20))
            mine_h = eng.group_ct_create(relay_bytes)
            all_h = [None] * world
            dist.all_gather_object(all_h, mine_h)
            eng.group_ct_open(all_h)
        if rank == 0:
            class _Ext:                       # zero-copy torch view of the group's grid (rank 0 reads it through torch)
                __cuda_array_interface__ = {"shape": (nz, ny, nx), "typestr": "<f4", "data": (eng.group_dose_ptr(), False), "version": 2}
            dose_dev = torch.as_tensor(_Ext(), device=dev)
    slab = dose_dev[z0:z1 + 1]

    def step_device():
        with torch.cuda.stream(stream):
            if group:
                return eng.plan_dose_group(0, cnx, cny, cnz, packed_all, len(case.beams))
            slab.zero_()
            t = eng.dose_ptr(packed, len(mine), dose_dev.data_ptr(), device=True)
            sharding.reduce_dose(slab, dst=0)
        return t

    for _ in range(warmup):
        step_device()
    torch.cuda.synchronize()
    if world > 1 and (not group or rank == 0) and (bool(dose_dev[:z0].any().item()) or bool(dose_dev[z1 + 1:].any().item())):
        raise SystemExit(f"bench.py: rank {rank} deposited dose outside the reduced slab [{z0}, {z1}]")
    sampler = ClockSampler(R.local) if full else None
    if sampler:
        sampler.start()  # (before the barrier: starting it costs rank 0 ~0.5 ms, which the other ranks' first step would wait for)
    R.barrier()          # every rank enters the timed loop together (a late rank would be waited for inside the others' first step)
    if group:
        eng.group_barrier()      # and the ranks' streams line up on the device as well (stream-ordered, no host sync)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    stage = {}
    t_wall0 = time.perf_counter()
    for i in range(steps):
        with torch.cuda.stream(stream):
            R.flush.fill_(i & 0xFF)                   # L2 flush between timed iterations (outside the event pair)
            evs[i][0].record(stream)
        tm = step_device()
        with torch.cuda.stream(stream):
            evs[i][1].record(stream)
        for k, v in tm.as_dict().items():
            stage[k] = stage.get(k, 0.0) + v
    R.barrier()
    t_wall = time.perf_counter() - t_wall0
    ms_local = sum(a.elapsed_time(b) for a, b in evs) / steps
    ms_each_local = [round(a.elapsed_time(b), 4) for a, b in evs]      # rank 0's own per-step times (spread / outliers)
    conc_stage = {k: v / steps for k, v in stage.items()}
    # stage timings / roofline: a few extra steps with the beams serialised on one stream, so that the library's stage
    # events bracket each kernel alone (with concurrent beams the stages of different beams overlap in time)
    eng.opts.serialBeams = 1
    stage = {}
    n_serial = 3
    step_device()
    for _ in range(n_serial):
        with torch.cuda.stream(stream):
            R.flush.fill_(1)
        tm = step_device()
        for k, v in tm.as_dict().items():
            stage[k] = stage.get(k, 0.0) + v
    eng.opts.serialBeams = 0
    stage = {k: v / n_serial for k, v in stage.items()}
    torch.cuda.synchronize()
    upd_local = torch.tensor([int(stage["updates"]), int(stage["ksTaps"])], dtype=torch.int64, device=dev)
    ms_t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(upd_local, op=dist.ReduceOp.SUM)
    ms_step = float(ms_t.item()); U = int(upd_local[0].item()); taps = int(upd_local[1].item())
    # balance of the sharding: every rank's own stage total of the serialised extra steps (no waiting on the others in it)
    own_ms = torch.tensor([stage["msTotal"]], dtype=torch.float64, device=dev)
    own_all = [torch.zeros_like(own_ms) for _ in range(world)]
    if world > 1:
        dist.all_gather(own_all, own_ms)
    else:
        own_all = [own_ms]
    rank_stage_ms = [round(float(t.item()), 4) for t in own_all]
    each_all = [None] * world            # every rank's own per-step times and concurrent-mode stage sums (diagnostics)
    if world > 1:
        dist.all_gather_object(each_all, {"ms_each_step": ms_each_local, "stage_ms": {k: round(v, 4) for k, v in conc_stage.items() if k.startswith("ms")}})
    checksum = float(dose_dev.double().sum().item()) if rank == 0 else 0.0
    dmax = float(dose_dev.max().item()) if rank == 0 else 0.0
    reps = [eng.beam_report(b) for b in range(len(mine))]            # of the rank's own shard, before anything else runs

    # ---- N > 1: the summed grid against the whole plan recomputed by rank 0 alone, voxel by voxel ----
    err_vs_1gpu = None
    if world > 1:
        if rank == 0:
            ref_grid = torch.zeros((nz, ny, nx), dtype=torch.float32, device=dev)
            with torch.cuda.stream(stream):
                eng.dose_ptr(packed_all, len(case.beams), ref_grid.data_ptr(), device=True)
            torch.cuda.synchronize()
            err_vs_1gpu = float((dose_dev - ref_grid).abs().max().item() / max(float(ref_grid.max().item()), 1e-30))
            del ref_grid
        R.barrier()

    # ---- end-to-end through the host-facing call, host buffers, copies inside the timed region ----
    e2e = None
    if do_e2e:
        def step_e2e():
            # the call a user of the reference makes: cudaWrapperProtons(imVol, doseVol, beams, ...) == rtd_plan_dose with
            # HOST CT and HOST dose; the library copies the sub-volumes the plan touches (bytes reported by the library)
            if world == 1:
                return eng.plan_dose_ptr(ct_host.data_ptr(), cnx, cny, cnz, packed, len(mine), dose_host.data_ptr())
            # N > 1: rtd_plan_dose_group with the HOST CT and, on rank 0, the HOST dose grid: every rank uploads the CT sub-volumes of
            # its own work items, the partial doses meet in the group's grid, rank 0 adds the plan's dose onto the caller's pinned
            # host grid in place (only voxels that received dose cross PCIe)
            with torch.cuda.stream(stream):
                if group:
                    return eng.plan_dose_group(ct_host.data_ptr(), cnx, cny, cnz, packed_all, len(case.beams), dose_host.data_ptr() if rank == 0 else 0)
                slab.zero_()
                t = eng.plan_dose_ptr(ct_host.data_ptr(), cnx, cny, cnz, packed, len(mine), dose_dev.data_ptr(), device=True)
                sharding.reduce_dose(slab, dst=0)
            if rank == 0:
                e2e_extra["mod"] = eng.add_device_to_host(slab.data_ptr(), dose_host.data_ptr() + z0 * ny * nx * 4, slab.numel())
            else:
                stream.synchronize()
            return t
        n_e2e = max(3, min(steps, 10))
        for _ in range(2):
            dose_host.zero_(); step_e2e()
        tt = []
        for _ in range(n_e2e):
            dose_host.zero_()
            R.barrier(); t0 = time.perf_counter(); te = step_e2e(); R.barrier(); tt.append(time.perf_counter() - t0)
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
               "timing": "host wall clock between barriers around the host call rtd_plan_dose on a WARM context (HOST CT + HOST dose in, dose "
                         "out; the library copies only the CT / dose sub-volumes the plan touches), max over ranks; rtd_create, rtd_set_luts and "
                         "the first growth of the workspaces are outside - see e2e_cold for the call from a fresh process",
               "library_events_ms": {"total": float(te.msTotal), "ct_upload_done_at": float(te.msUpload), "download": float(te.msDownload)},
               "host_syncs_per_plan": int(te.hostSyncs), "dose_voxels_modified": int(te.doseVoxelsModified),
               "dose_checksum_matches_device_run": None}
        if rank == 0:
            cs = float(dose_host.double().sum().item())
            e2e["dose_checksum_matches_device_run"] = bool(abs(cs - checksum) <= 1e-5 * abs(checksum))
        eng.set_ct_ptr(ct_dev.data_ptr(), cnx, cny, cnz, device=True)

    if sampler:
        sampler.stop()   # sampled over the device-resident and the end-to-end timed regions
    res = None
    if rank == 0:
        peak, peak_src, sm_max = load_peaks()
        live = [r for r in reps if r["beamFirstCalculatedPassive"] > r["beamFirstInside"]]
        ks_launches = max(1, int(round(stage["ksLaunchesWithWork"])))      # (beam, radius class) launches that had items: counted on the device
        traffic, traffic_src = None, None
        tp = os.path.join(ROOT, "profiles", "r02_ncu_ks2_traffic.json")
        if os.path.exists(tp) and case.name.startswith("c3") and world == 1:
            tj = json.load(open(tp)); traffic = tj["dram_bytes_per_launch"]; traffic_src = tj["source"]
        smem = None
        sp = os.path.join(ROOT, "profiles", "r02_ncu_ks2_smem.json")
        if os.path.exists(sp) and case.name.startswith("c3") and world == 1:
            sj = json.load(open(sp))
            smem = {"frac": sj["frac_time_weighted"], "unit": "shared-memory wavefronts (128 B) per clock per SM, peak 1.0",
                    "per_launch": [{k: l[k] for k in ("kernel", "wavefronts_per_clk_per_sm")} for l in sj["launches"]],
                    "source": "profiles/r02_ncu_ks2_smem.json: the committed ncu --set full capture of this workload, not this run"}
        ks_ms = stage["msSuperp"]
        u_rank0 = sum(r["updates"] for r in reps)
        bev_bytes = sum(4 * (r["rayDimX"] + 64) * (r["rayDimY"] + 64) * (r["beamFirstCalculatedPassive"] - r["beamFirstInside"]) for r in live)
        ks_alg = 8 * u_rank0 + bev_bytes          # read idd + rSigmaEff once per update, write the haloed BEV dose once
        taps0 = stage["ksTaps"]
        achieved = ks_alg / (ks_ms * 1e-3) / 1e9 if ks_ms > 0 else 0.0
        fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        fp32_frac = (2 * taps0 / (ks_ms * 1e-3) / 1e12) / fp32_peak if ks_ms > 0 else 0.0
        # whole-plan algorithmic bytes
        plan_bytes = 4 * nx * ny * nz
        for b, r in zip(mine, reps):
            Rr = r["rayDimX"] * r["rayDimY"]
            box = max(0, r["bboxMax"][0] - r["bboxMin"][0] + 1) * max(0, r["bboxMax"][1] - r["bboxMin"][1] + 1) * max(0, r["bboxMax"][2] - r["bboxMin"][2] + 1)
            plan_bytes += 4 * box + 8 * Rr * r["steps"] + 8 * r["updates"] + 8 * (r["rayDimX"] + 64) * (r["rayDimY"] + 64) * max(0, r["beamFirstCalculatedPassive"] - r["beamFirstInside"]) + 8 * box
        how = ("rtd_plan_dose_group: every rank adds the dose-holding float4 groups of its slab into ONE grid on rank 0 over NVLink (CUDA IPC, "
               "red.global.add.v4.f32), ordered by two barriers in device memory; no NCCL call inside the step" if group
               else "partial dose grids (reachable z-slab) summed by one NCCL reduce")
        res = {"value": U / (ms_step * 1e-3), "ms_per_step": ms_step, "updates_per_step": U, "ks_taps_per_step": taps,
               "sharding": (f"(beam, layer-range) work items, contiguous equal-cost runs over {world} ranks; " + how) if world > 1 else "single GPU",
               "l2": "256 MiB buffer written between timed iterations (L2 flush); CT 4*%d B and BEV arrays exceed L2" % (nx * ny * nz),
               "ms_each_step_rank0": ms_each_local, "per_rank": each_all if world > 1 else None, "stage_total_ms_per_rank": rank_stage_ms, "work_items_per_rank": [len(p) for p in parts],
               "stage_ms_rank0": {k: stage[k] for k in ("msTrace", "msSetup", "msConv", "msIddSigma", "msRadius", "msSuperp", "msTransfer", "msTotal")},
               "stage_note": "stage_ms_rank0 and roofline come from 3 extra steps with opts.serialBeams=1 (stages bracketed alone); the timed steps run one stream per beam",
               "serial_beams_plan_ms": stage["msTotal"],
               "gpu_launches": int(round(conc_stage["kernelLaunches"] * steps)), "gpu_launches_per_step": int(round(conc_stage["kernelLaunches"])),
               "host_syncs_per_plan": int(round(stage["hostSyncs"])),
               "dose_max": dmax, "dose_checksum": checksum,
               "roofline": {"kernel": "ks2_kernel (layer-fused kernel superposition; one launch per beam and radius class that has items)",
                            "bound": "shared-memory bandwidth (fp32 SIMT): neither HBM nor tensor - the HBM fraction below is the mandated algorithmic form, "
                                     "smem.frac (where a committed capture exists) and fp32.frac are the fractions of the bounds that apply",
                            "smem": smem,
                            "achieved": achieved, "peak": peak, "unit": "GB/s",
                            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                            "traffic_note": "DRAM bytes per launch from a committed ncu capture of this workload (not from this run); below the algorithmic "
                                            "8 B/update because tiles without dose and layers past their range are skipped before their idd / rSigmaEff are read",
                            "peak_source": peak_src, "algorithmic_bytes_per_launch": ks_alg / ks_launches,
                            "avg_launch_ms": ks_ms / ks_launches, "launches_per_plan_rank0": ks_launches,
                            "fp32": {"achieved_tflops": 2 * taps0 / (ks_ms * 1e-3) / 1e12 if ks_ms > 0 else 0.0, "peak_tflops": fp32_peak, "frac": fp32_frac,
                                     "note": "2 x reference-equivalent taps / kernel time vs 148 SM x 128 FMA/clk x sm_max_mhz (the kernel issues about "
                                             "half of those taps thanks to the +-j symmetry)"}},
               "plan_roofline": {"algorithmic_bytes": plan_bytes, "achieved_gbs": plan_bytes / (ms_step * 1e-3) / 1e9, "frac_of_hbm_peak": plan_bytes / (ms_step * 1e-3) / 1e9 / peak,
                                 "note": "rank 0's share of the plan over the whole step" if world > 1 else "whole plan"},
               "wall_s_timed_region": t_wall}
        if sampler:
            res["clocks"] = sampler.summary()
        if err_vs_1gpu is not None:
            res["dose_max_rel_err_vs_1gpu"] = err_vs_1gpu
        if e2e:
            res["e2e"] = e2e
        if do_cpu and world == 1:
            res["cpu_baseline"] = cpu_baseline_port(case)[0]
            res["cpu_stage_baseline"] = cpu_stage_baseline(case, gpu_stage_ms=stage["msConv"])
    # ordered teardown of this workload: torch tensors first, then the IPC mapping (non-owners unmap before the owner frees)
    del evs, ms_t, upd_local, slab
    dose_dev = ct_dev = None
    torch.cuda.synchronize()
    if group:
        if rank != 0:
            eng.group_destroy()
        R.barrier()
        if rank == 0:
            eng.group_destroy()
    del ct_host, dose_host, keep_all
    return res


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
    ap.add_argument("--no-c4", action="store_true", help="skip the extra C4 measurement (key 'c4')")
    ap.add_argument("--no-cold", action="store_true", help="skip the fresh-process measurement (key 'e2e_cold')")
    ap.add_argument("--cold-child", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--ct-relay", action="store_true", help="N > 1: the e2e call uploads every CT box once per node and the ranks fetch theirs over NVLink (rtd_group_ct_*; measured slower up to N = 4, off by default)")
    ap.add_argument("--lib", default=None, help="measure another build of librtdose.so (A/B of kernel variants)")
    ap.add_argument("--reduce", default="push", choices=["push", "nccl"],
                    help="N > 1, how the ranks' partial doses meet: 'push' = rtd_plan_dose_group (ONE grid on rank 0 mapped by every rank over "
                         "CUDA IPC, 16-byte vector reductions over NVLink, barriers in device memory); 'nccl' = local partial grids + one dense "
                         "NCCL reduce of the reachable slab")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.lib:
        from raytracedicom_b200 import engine as _engine
        _engine.LIB_PATH = os.path.abspath(args.lib)

    if args.cold_child:
        return cold_child(args)

    from raytracedicom_b200 import cases
    case = cases.make_case(args.workload)

    if args.impl == "reference":
        line = run_reference(args, case, rank, world)
        if rank == 0:
            print(json.dumps(line), flush=True)
        return 0

    R = Runner(args)
    rc = 0
    main_res = measure(R, case, args.steps, args.warmup, do_e2e=not args.no_e2e, do_cpu=not args.no_cpu_baseline, full=True)
    c4_res = c4_cfg = None
    if not args.no_c4 and args.workload == "c3":
        c4_case = cases.make_case("c4")
        c4_cfg = config_of(c4_case)
        c4_res = measure(R, c4_case, max(3, min(args.steps, 5)), 3, do_e2e=not args.no_e2e, do_cpu=False, full=False)
        del c4_case
    R.close()
    if rank == 0:
        line = {"metric": METRIC, "value": main_res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": main_res["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config_of(case), "plan_ms": main_res["ms_per_step"]}
        line.update({k: v for k, v in main_res.items() if k not in ("value", "ms_per_step")})
        if c4_res is not None:
            keys = ("value", "ms_per_step", "updates_per_step", "stage_total_ms_per_rank", "stage_ms_rank0", "per_rank", "work_items_per_rank", "gpu_launches_per_step",
                    "host_syncs_per_plan", "dose_checksum", "dose_max_rel_err_vs_1gpu", "ms_each_step_rank0")
            line["c4"] = {"config": c4_cfg, **{k: c4_res[k] for k in keys if k in c4_res}}
            if "e2e" in c4_res:
                line["c4"]["e2e"] = {k: c4_res["e2e"][k] for k in ("value", "unit", "ms_per_plan", "h2d_bytes_per_step", "d2h_bytes_per_step")}
        # the drop-in call from a fresh process (rank 0, single GPU): what include/rtdose_shim.hpp costs a caller of cudaWrapperProtons
        if world == 1 and not args.no_cold and not args.no_e2e:
            try:
                cmd = [sys.executable, os.path.abspath(__file__), "--cold-child", "--workload", args.workload] + (["--lib", args.lib] if args.lib else [])
                runs = []
                for _ in range(2):
                    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
                    js = [l for l in r.stdout.splitlines() if l.startswith("{")]
                    if r.returncode == 0 and js:
                        runs.append(json.loads(js[-1]))
                if runs:
                    best = min(runs, key=lambda d: d["cold_ms"])
                    line["e2e_cold"] = {"value": main_res["updates_per_step"] / (best["cold_ms"] * 1e-3), "unit": UNIT, "ms_per_plan": best["cold_ms"],
                                        "breakdown_ms": {k: best[k] for k in ("create_and_luts_ms", "plan_dose_ms", "destroy_ms")},
                                        "all_runs_ms": [round(d["cold_ms"], 3) for d in runs],
                                        "what": "fresh process, CUDA context up, then rtd_create -> rtd_set_luts -> rtd_set_dose_grid -> rtd_plan_dose (pinned HOST "
                                                "CT + HOST dose) -> rtd_destroy, host wall clock: the coverage of the reference's own total timer "
                                                "(kernel_wrapper.cu:410-415, 1357) for one call of the drop-in shim"}
                else:
                    line["e2e_cold"] = {"unavailable": (r.stderr or r.stdout)[-300:]}
            except Exception as ex:      # noqa: BLE001
                line["e2e_cold"] = {"unavailable": str(ex)[:300]}
        if "dose_max_rel_err_vs_1gpu" in main_res and not (main_res["dose_max_rel_err_vs_1gpu"] <= DOSE_TOL_VS_1GPU):
            rc = 3
        if c4_res is not None and "dose_max_rel_err_vs_1gpu" in c4_res and not (c4_res["dose_max_rel_err_vs_1gpu"] <= DOSE_TOL_VS_1GPU):
            rc = 3
        if rc:
            line["error"] = f"multi-GPU dose differs from the single-GPU dose by more than {DOSE_TOL_VS_1GPU} of the maximum"
        print(json.dumps(line), flush=True)
    return rc


if __name__ == "__main__":
    sys.exit(main())
