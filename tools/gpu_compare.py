#!/usr/bin/env python
"""GPU-box parity driver: runs the rebuilt reference (oracle/_ref/ref_stages, ref_dose) and the new engine on
the same case and compares every stage. Index/integer outputs must be bit-exact; float arrays are reported as
bit-mismatch counts and max abs/rel error; dose must satisfy |D-Dref| <= 1e-3*max(Dref).

usage: python tools/gpu_compare.py --case mini [--golden out.npz] [--no-ref-dose] [--layers 0,1]
Writes gpurun_out/compare_<case>.json. Test infrastructure (uses oracle/_ref, which is allowed for checkers)."""
import argparse, json, os, subprocess, sys, tempfile, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytracedicom_b200 import cases
from raytracedicom_b200.engine import DoseEngine

REF = os.path.join(ROOT, "oracle", "_ref")


def parse_meta(path):
    m = {}
    for line in open(path):
        k, _, v = line.strip().partition("=")
        m[k] = v
    return m


def cmp_float(name, a, b, rep, valid=None):
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    if valid is not None:
        a = a[valid]; b = b[valid]
    bits = int(np.sum(a.view(np.uint32) != b.view(np.uint32)))
    both_nan = np.isnan(a) & np.isnan(b)
    fin = np.isfinite(a) & np.isfinite(b)
    d = np.abs(a[fin].astype(np.float64) - b[fin].astype(np.float64))
    mx = float(d.max()) if d.size else 0.0
    scale = float(np.abs(b[fin]).max()) if d.size else 0.0
    nonfin_mismatch = int(np.sum((~fin) & ~both_nan & (a != b)))
    rep[name] = dict(n=int(a.size), bit_mismatch=bits, max_abs=mx, ref_absmax=scale, rel_to_max=(mx / scale if scale else 0.0),
                     nonfinite_mismatch=nonfin_mismatch)
    return bits


def cmp_int(name, a, b, rep):
    a = np.asarray(a); b = np.asarray(b)
    mism = int(np.sum(a != b))
    rep[name] = dict(n=int(a.size), mismatch=mism)
    return mism


def gamma_pass_rate_cropped(dose, ref, spacing, dose_crit=0.01, dist_mm=1.0, cutoff=0.1):
    """Global gamma(1 %/1 mm) pass rate on the voxels with ref >= 10 % of max (BASELINE.json: >= 99.9 %). A voxel whose own dose
    difference is within the dose criterion passes at distance 0, so when max |D - Dref| <= 1 % of max the rate is exactly 1;
    otherwise the 3x3x3 (+ interpolated) neighbourhood search of tests/conftest.py runs on the bounding box of the evaluated voxels."""
    dmax = float(ref.max())
    mask = ref >= cutoff * dmax
    if not mask.any():
        return 1.0
    if float(np.abs(dose - ref)[mask].max()) <= dose_crit * dmax:
        return 1.0
    idx = np.argwhere(mask)
    lo = np.maximum(idx.min(0) - 1, 0); hi = np.minimum(idx.max(0) + 2, ref.shape)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import gamma_pass_rate
    sl = tuple(slice(int(a), int(b)) for a, b in zip(lo, hi))
    return gamma_pass_rate(dose[sl], ref[sl], spacing, dose_crit, dist_mm, cutoff)     # the box holds the maximum of ref


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="mini")
    ap.add_argument("--golden", default=None)
    ap.add_argument("--no-ref-dose", action="store_true")
    ap.add_argument("--no-ref", action="store_true", help="run only the new engine (e.g. under compute-sanitizer)")
    ap.add_argument("--layers", default="all")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out"))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    case = cases.make_case(args.case)
    rep = {"case": case.name, "ct_dims": case.ct_dims, "n_beams": len(case.beams)}
    tmp = tempfile.mkdtemp(prefix="rtd_")
    cpath = os.path.join(tmp, "case.rtdcase")
    case.write(cpath)

    eng = DoseEngine(0)
    eng.set_luts(case.luts)
    eng.set_ct(case.ct)
    nx, ny, nz = case.dose_dims
    eng.set_dose_grid(nx, ny, nz)
    dose = np.zeros((nz, ny, nx), np.float32)
    t0 = time.time(); tm = eng.dose(case.beams, dose); t1 = time.time()
    rep["engine_timing_1st_call"] = tm.as_dict()
    # second run with opts.fullIntermediates: the depth scan writes every (ray, step) of idd / rSigmaEff (the default leaves row
    # segments without dose unwritten), so that the arrays can be compared with the reference's dumps below; the two doses must agree
    dose2 = np.zeros_like(dose)
    eng.opts.fullIntermediates = 1
    tm = eng.dose(case.beams, dose2)
    rep["engine_timing_2nd_call"] = tm.as_dict()
    rep["engine_wall_first_call_s"] = t1 - t0
    rep["engine_dose_max"] = float(dose.max())
    rep["engine_rerun_max_abs_diff"] = float(np.abs(dose - dose2).max())
    rep["engine_sparse_vs_full_intermediates_max_rel"] = float(np.abs(dose - dose2).max() / max(float(dose2.max()), 1e-30))
    rep["beams"] = []
    if args.no_ref:
        for b in range(len(case.beams)):
            rep["beams"].append(eng.beam_report(b))
        json.dump(rep, open(os.path.join(args.out, f"compare_{args.case}_noref.json"), "w"), indent=1)
        print(json.dumps(rep)[:2000])
        return 0

    golden = {}
    ok = True
    total_ref = np.zeros_like(dose)
    for b in range(len(case.beams)):
        br = eng.beam_report(b)
        d = os.path.join(tmp, f"b{b}"); os.makedirs(d)
        r = subprocess.run([os.path.join(REF, "ref_stages"), cpath, d, str(b), args.layers], capture_output=True, text=True, timeout=1200)
        print(r.stdout[-500:], r.stderr[-2000:])
        if r.returncode != 0:
            rep["beams"].append({"ref_stages_failed": r.stderr[-500:]}); ok = False; continue
        m = parse_meta(os.path.join(d, "meta.txt"))
        brep = {"engine": br, "ref_meta": {k: v for k, v in m.items() if not k.startswith("hist") and not k.startswith("layer")}}
        Rx, Ry, S, L = int(m["Rx"]), int(m["Ry"]), int(m["S"]), int(m["L"])
        R = Rx * Ry
        ints = dict(lSteps="lSteps", bSteps="bSteps", rSteps="rSteps", tSteps="tSteps", rayDimX="Rx", rayDimY="Ry",
                    beamFirstInside="beamFirstInside", beamFirstOutside="beamFirstOutside",
                    beamFirstGuaranteedPassive="beamFirstGuaranteedPassive", beamFirstCalculatedPassive="beamFirstCalculatedPassive",
                    updates="updates")
        sc = {k: (br[k], int(m[v])) for k, v in ints.items()}
        sc["bboxMin"] = (br["bboxMin"], [int(x) for x in m["bboxMin"].split(",")])
        sc["bboxMax"] = (br["bboxMax"], [int(x) for x in m["bboxMax"].split(",")])
        brep["scalar_mismatches"] = {k: v for k, v in sc.items() if v[0] != v[1]}
        ok &= not brep["scalar_mismatches"]
        if (br["rayDimX"], br["rayDimY"]) != (Rx, Ry):
            rep["beams"].append(brep); continue
        bfi, Zg = int(m["beamFirstInside"]), int(m["beamFirstGuaranteedPassive"])
        ld = lambda n, dt, shape: np.fromfile(os.path.join(d, n), dt).reshape(shape)
        ref_density = ld("density.f32", np.float32, (S, Ry, Rx)); ref_wepl = ld("wepl.f32", np.float32, (S, Ry, Rx))
        ok &= cmp_float("density", eng.fetch(b, "density"), ref_density, brep) == 0
        ok &= cmp_float("wepl", eng.fetch(b, "wepl"), ref_wepl, brep) == 0
        ref_fi = ld("first_inside.i32", np.int32, (Ry, Rx)); ref_fo = ld("first_outside.i32", np.int32, (Ry, Rx))
        ok &= cmp_int("first_inside", eng.fetch(b, "first_inside"), ref_fi, brep) == 0
        ok &= cmp_int("first_outside", eng.fetch(b, "first_outside"), ref_fo, brep) == 0
        ok &= cmp_float("wepl_min", eng.fetch(b, "wepl_min"), ld("wepl_min.f32", np.float32, (S,)), brep) == 0
        ref_rw = ld("ray_weights.f32", np.float32, (L, Ry, Rx))
        cmp_float("ray_weights", eng.fetch(b, "ray_weights"), ref_rw, brep)
        if b == 0:
            golden.update(density=ref_density, wepl=ref_wepl, first_inside=ref_fi, first_outside=ref_fo, ray_weights=ref_rw)
        layers = range(L) if args.layers == "all" else [int(x) for x in args.layers.split(",")]
        lay = {}
        for l in range(L):
            lr = eng.layer_report(b, l)
            kv = dict(x.split(":") for x in m[f"layer{l}"].split(","))
            hist = [int(x) for x in m[f"hist{l}"].split(",")]
            e = {}
            if lr["afterLastStep"] != int(kv["afterLast"]): e["afterLast"] = (lr["afterLastStep"], int(kv["afterLast"]))
            if lr["layerFirstPassive"] != int(kv["layerFirstPassive"]): e["layerFirstPassive"] = (lr["layerFirstPassive"], int(kv["layerFirstPassive"]))
            if lr["radiusHist"] != hist: e["hist"] = (lr["radiusHist"], hist)
            for k1, k2 in (("energyIdx", "energyIdx"), ("energyScaleFact", "scale"), ("peakDepth", "peak"), ("entrySigmaX", "entrySigX"), ("entrySigmaY", "entrySigY")):
                if abs(lr[k1] - float(kv[k2])) > 2e-6 * max(1.0, abs(lr[k1])): e[k1] = (lr[k1], float(kv[k2]))
            if l in layers:
                al = int(kv["afterLast"])
                if al > bfi:
                    ref_idd = ld(f"idd_l{l}.f32", np.float32, (S, Ry, Rx))[bfi:al]
                    ref_rs = ld(f"rsig_l{l}.f32", np.float32, (S, Ry, Rx))[bfi:al]
                    mine_idd = eng.fetch(b, "idd", l)[:al - bfi]; mine_rs = eng.fetch(b, "rsigma", l)[:al - bfi]
                    cmp_float("idd", mine_idd, ref_idd, e); cmp_float("rsig", mine_rs, ref_rs, e)
                    if b == 0 and l == layers[0]:
                        golden.update(idd_l=ref_idd, rsig_l=ref_rs, golden_layer=np.int32(l))
                ref_fp = ld(f"first_passive_l{l}.i32", np.int32, (Ry, Rx))
                cmp_int("first_passive", eng.fetch(b, "first_passive", l), ref_fp, e)
                if b == 0 and l == layers[0]:
                    golden.update(first_passive_l=ref_fp)
            if b == 0:
                golden[f"hist_{l}"] = np.array(hist, np.int32)
                golden[f"layer_ints_{l}"] = np.array([int(kv["afterLast"]), int(kv["layerFirstPassive"])], np.int32)
            lay[l] = e
        brep["layers"] = lay
        bev_ref = ld("bev_dose.f32", np.float32, (Zg, Ry + 64, Rx + 64))
        cmp_float("bev_dose", eng.fetch(b, "bev_dose"), bev_ref, brep)
        dref = ld("dose.f32", np.float32, (nz, ny, nx))
        total_ref += dref
        if b == 0:
            golden.update(bev_dose=bev_ref, dose_beam0=dref,
                          meta_ints=np.array([int(m[k]) for k in ("lSteps", "bSteps", "rSteps", "tSteps", "Rx", "Ry", "beamFirstInside", "beamFirstOutside",
                                                                    "beamFirstGuaranteedPassive", "beamFirstCalculatedPassive", "updates")], np.int64),
                          bbox=np.array([int(x) for x in (m["bboxMin"] + "," + m["bboxMax"]).split(",")], np.int32))
        rep["beams"].append(brep)
    cmp_float("dose_vs_ref_stages", dose, total_ref, rep)
    dmax = float(total_ref.max())
    rep["dose_max_rel_err"] = float(np.abs(dose - total_ref).max() / dmax) if dmax > 0 else None
    rep["gamma_1pct_1mm_pass_rate"] = gamma_pass_rate_cropped(dose, total_ref, case.spacing) if dmax > 0 else None
    if not args.no_ref_dose:
        dpath = os.path.join(tmp, "ref_dose.f32")
        r = subprocess.run([os.path.join(REF, "ref_dose"), cpath, dpath], capture_output=True, text=True, timeout=1200)
        rep["ref_dose_stdout_tail"] = r.stdout[-400:]; rep["ref_dose_rc"] = r.returncode
        if r.returncode == 0:
            dfull = np.fromfile(dpath, np.float32).reshape(nz, ny, nx)
            cmp_float("dose_vs_ref_wrapper", dose, dfull, rep)
            cmp_float("ref_stages_vs_ref_wrapper", total_ref, dfull, rep)
            rep["dose_vs_wrapper_max_rel_err"] = float(np.abs(dose - dfull).max() / dfull.max())
        r = subprocess.run([os.path.join(REF, "ref_dose_timed"), cpath, dpath], capture_output=True, text=True, timeout=1200)
        rep["ref_dose_timed_tail"] = r.stdout[-3000:]
    if args.golden:
        np.savez_compressed(args.golden, **golden)
        rep["golden_bytes"] = os.path.getsize(args.golden)
    rep["index_and_trace_bit_exact"] = bool(ok)
    json.dump(rep, open(os.path.join(args.out, f"compare_{args.case}.json"), "w"), indent=1, default=str)
    # compact print
    def brief(o, depth=0):
        if isinstance(o, dict):
            return {k: brief(v, depth + 1) for k, v in o.items() if k not in ("ref_dose_timed_tail",)}
        return o
    print(json.dumps(brief(rep), default=str)[:6000])
    return 0


if __name__ == "__main__":
    sys.exit(main())
