#!/usr/bin/env python
"""Benchmark of the Groth16 proving path (BASELINE.json metric), one JSON line on stdout.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME]

A step = one partition proof (witness -> 192-byte proof) of a PoRep-C2-shaped synthetic R1CS
through libb2g16 (hand-written sm_100a kernels behind the C ABI of include/b2g16.h).
`value` times K proofs with the witness already in HBM; `e2e` times the same through the
host-buffer entry point b2g16_prove (pinned witness H2D + proof D2H inside the timed region).
The line also carries the two standalone figures the metric names (G1 MSM Mpts/s, Fr NTT ms)
with their rooflines, and a CPU baseline (the oracle's OpenMP prover on the box's host cores,
bounded sample).  `--impl reference` times only that CPU prover (the reference's own
implementation of this path is absent from its tree: DESIGN.md section 3).
Under torchrun (N > 1) every rank proves its own partitions (same key, different witness):
weak scaling, no data-path collective; times are max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np

# workload name -> (log_m, challenges per partition, description); n_in = 3 + 18 * challenges
WORKLOADS = {
    "c1-2^16": (16, 0, "Groth16 prove, synthetic 2^16-domain R1CS (BASELINE configs[0])"),
    "porep-c2-2KiB": (22, 2, "PoRep C2 partition, 2 KiB-sector-shaped circuit: domain 2^22, 39 public inputs (BASELINE configs[1])"),
    "porep-c2-2^24": (24, 18, "PoRep C2 partition shape scaled to a 2^24 domain, 327 public inputs"),
    "porep-c2-32GiB": (27, 18, "PoRep C2 partition, 32 GiB-sector-shaped circuit: domain 2^27, 327 public inputs (BASELINE configs[4], one partition)"),
    # WindowPoSt, 32 GiB sectors: 2349 sectors x (comm_r + 10 challenge paths) public inputs ([UPSTREAM-RECALL]); 125,279,217
    # constraints (F/constants.hpp:88) round up to the same 2^27 domain
    "window-post-32GiB": (27, -25839, "WindowPoSt-shaped proof, 32 GiB sector parameters: domain 2^27, 25,839 public inputs (BASELINE configs[3])"),
}
MAD32_PER_G1_POINT = 48000.0      # 16 windows x 10 Fq products x 300 MAD32
MAD32_PER_G2_POINT = 144000.0
METRIC = "groth16_c2_partition_prove_s"


def n_in_of(workload):
    from b2g16 import shapes
    log_m, ch, _ = WORKLOADS[workload]
    if ch < 0:
        return -ch                       # public input count given directly
    return 1 if ch == 0 else shapes.porep_public_inputs(ch)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------- clocks
CLOCK_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
               "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, device):
        self.path = "/tmp/b2g16_clocks_%d_%d.csv" % (os.getpid(), device)
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + CLOCK_QUERY, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.f.close()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as f:
            for line in f:
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1])); smax.append(float(c[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, c[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(smax), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- instance
def build_instance(workload, seed=2026):
    """Host-side synthetic R1CS (numpy).  Shape: SURVEY 8(d) -- 3 nnz/row in A and B, 1 in C,
    n_cons = 2^log_m - n_in - 1 (domain exactly 2^log_m)."""
    from b2g16 import synth
    log_m = WORKLOADS[workload][0]
    return synth.SynthR1cs(log_m, n_in_of(workload), seed, n_free=8)


def toxic_from_seed(seed):
    from b2g16 import codec
    rng = np.random.Generator(np.random.PCG64(seed ^ 0x70C51C))
    t = codec.random_fr(7, rng)
    return t[:5], t[5], t[6]


class _Cs:
    """what lib._r1cs_view reads"""

    def __init__(self, s):
        self.log_m, self.n_in, self.n_aux, self.n_cons = s.log_m, s.n_in, s.n_aux, s.n_cons
        self.n_vars = s.n_vars
        self.mats = [s.A, s.B, s.C]


def dev_random_fr(torch, n, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    x = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device=device, generator=g)
    x[:, 31] &= 0x3F                      # < 2^254 < r: canonical, near-uniform (all 16 signed windows populated)
    return x


def dev_filecoin_like_fr(torch, n, seed, device):
    """The skewed scalar set of SURVEY 8(d): 60 % zero, 20 % one, 10 % below 2^16, 10 % uniform -- what a real
    Filecoin witness (boolean wires of the SHA-256 / bit-decomposition gadgets) looks like to the multiexp."""
    x = dev_random_fr(torch, n, seed, device)
    g = torch.Generator(device=device)
    g.manual_seed(seed + 1)
    u = torch.rand(n, device=device, generator=g)
    x[u < 0.9, 2:] = 0                     # below 2^16
    x[u < 0.8] = 0
    x[(u >= 0.6) & (u < 0.8), 0] = 1       # one
    return x


# ------------------------------------------------------------------------ CPU baseline
def cpu_prove_sample(workload, sample_log_m, steps=1, warmup=0, seed=2026):
    """The oracle's OpenMP Groth16 prover (oracle/c/oracle.cpp) on a 2^sample_log_m instance of the
    same shape, all host threads; returns (seconds per proof scaled linearly to the workload's
    domain, threads, sample description).  Setup (CPU) is outside the timed region."""
    from oracle import coracle
    from b2g16 import synth
    coracle.build()
    # all host cores, whatever the launcher exported (torchrun sets OMP_NUM_THREADS=1 for its workers)
    try:
        coracle.set_threads(len(os.sched_getaffinity(0)))
    except AttributeError:
        coracle.set_threads(os.cpu_count() or 1)
    log_m = WORKLOADS[workload][0]
    n_in = n_in_of(workload)
    sample_log_m = min(sample_log_m, log_m)
    s = synth.SynthR1cs(sample_log_m, n_in, seed, n_free=8)
    cs = coracle.R1csArrays(s.log_m, s.n_in, s.n_aux, s.A, s.B, s.C)
    z = coracle.synth_witness(cs, s.z_head, s.n_free)
    toxic, r, sr = toxic_from_seed(seed)
    key = coracle.groth16_setup(cs, toxic)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        proof, _ = coracle.groth16_prove(key, z, r, sr, want_trace=False)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    assert len(proof) == 192
    scale = float(1 << (log_m - sample_log_m))
    per = sum(times) / len(times)
    sample = "oracle C++/OpenMP Groth16 prove at domain 2^%d (%.2f s/proof measured), scaled x%d linearly to 2^%d" % (sample_log_m, per, int(scale), log_m)
    return per * scale, coracle.threads(), sample, sum(times)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_log = args.cpu_sample_log_m or (18 if WORKLOADS[args.workload][0] >= 18 else WORKLOADS[args.workload][0])
    t0 = time.perf_counter()
    v, threads, sample, timed_s = cpu_prove_sample(args.workload, sample_log, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "s", "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * timed_s / args.steps, "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "u32-limb Montgomery (integer)",
        "data": "synthetic", "config": {"workload": args.workload, "describe": WORKLOADS[args.workload][2], "log_m": WORKLOADS[args.workload][0], "n_in": n_in_of(args.workload)},
        "cpu_baseline": {"value": v, "unit": "s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
        "note": "the reference's own prover sources are absent from its tree (un-vendored crypto3 submodules), so kind=port: the oracle restatement",
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------- B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from b2g16 import codec, lib

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the CPU prover)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)             # torch work, its events and the library share one stream
    ctx = lib.Context(local, ts.cuda_stream)
    hbm_peak, peak_src = peaks()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed_region(fn, reps):
        """barrier + sync, events on the library's stream around `reps` calls, sync + barrier; max over ranks"""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    # ---- instance + key generated on the device (hashed synthetic R1CS, b2g16_pk_generate_synth), witness
    #      evaluated on the device: all outside the timed region
    t_setup = time.perf_counter()
    log_m, n_in, n_free, seed = WORKLOADS[args.workload][0], n_in_of(args.workload), 8, 2026
    toxic, r, sr = toxic_from_seed(seed)
    pk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic)
    shape = pk.shape()
    nv = shape["n_vars"]
    rng = np.random.Generator(np.random.PCG64(9000 + rank))     # a different witness per rank: own inputs and free variables
    head = codec.random_fr(1 + n_in + n_free, rng)
    head[0] = 0
    head[0, 0] = 1
    dz = torch.zeros((nv, 32), dtype=torch.uint8, device=dev)
    dz[: head.shape[0]] = torch.from_numpy(head).to(dev)
    ctx.synth_witness(pk, dz.data_ptr(), n_free)
    hz = torch.empty((nv, 32), dtype=torch.uint8, pin_memory=True)
    hz.copy_(dz)
    torch.cuda.synchronize()
    hz_np = hz.numpy()
    t_setup = time.perf_counter() - t_setup

    proofs = []
    step_dev = lambda: proofs.append(ctx.prove_device(pk, dz.data_ptr(), r, sr))
    step_e2e = lambda: proofs.append(ctx.prove(pk, hz_np, r, sr))
    for _ in range(args.warmup):
        step_dev()
    for _ in range(min(args.warmup, 2)):
        step_e2e()
    assert len(set(proofs)) == 1 and len(proofs[0]) == 192, "device-resident and host-buffer entry points disagree"

    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ctx.launches
    ms = timed_region(step_dev, args.steps)
    launches = ctx.launches - l0
    ms_e2e = timed_region(step_e2e, args.steps)
    clocks = sampler.stop() if sampler else None
    assert len(set(proofs)) == 1

    # ---- dominant kernel: bucket accumulation of the G1 MSM; roofline on the int32 multiply pipe
    mad_peak, _ = ctx.mad32_peak()
    n_h = (1 << log_m) - 1
    kk = dev_random_fr(torch, n_h, 99, dev)
    bases = ctx.bases_generate(1, kk.data_ptr(), n_h)
    del kk
    sc = dev_random_fr(torch, n_h, 7, dev)
    out = torch.zeros(192, dtype=torch.uint8, device=dev)
    ctx.msm_resident_async(bases, sc.data_ptr(), n_h, out.data_ptr())
    ctx.set_msm_timing(True)
    acc_ms, tot_ms = [], []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.msm_resident_async(bases, sc.data_ptr(), n_h, out.data_ptr()); e1.record()
        torch.cuda.synchronize()
        acc_ms.append(ctx.last_msm_timing()["accumulate"]); tot_ms.append(e0.elapsed_time(e1))
    ctx.set_msm_timing(False)
    bases.free()
    del sc
    acc = sum(acc_ms) / len(acc_ms)
    achieved = n_h * MAD32_PER_G1_POINT / (acc * 1e-3) / 1e12
    roofline = {"kernel": "msm_accumulate_kernel<Fq> (bucket accumulation, G1 h-query size)", "bound": "int32-imad", "achieved": achieved,
                "peak": mad_peak / 1e12, "unit": "TMAD32/s", "frac": achieved / (mad_peak / 1e12), "traffic": ncu_traffic(log_m),
                "traffic_note": "DRAM bytes per accumulate launch, ncu --set full of the same kernel at 2^22 (profiles/r01_ncu_full_s13.json); 16 windows x 2^22 gathers of 96-byte points = 6.4 GB is the kernel's own minimum, SURVEY's N*128 B counts one read of the inputs; 384 GB/s = 6 % of HBM, not the bound",
                "launch_ms": acc, "units_per_launch": n_h, "alg_mad32_per_point": MAD32_PER_G1_POINT,
                "peak_source": "b2g16_measure_mad32_peak (dependency-free mad.wide.u32 loop, this run)",
                "hbm_view": {"alg_bytes": n_h * 128, "GBps": n_h * 128 / (acc * 1e-3) / 1e9, "peak_GBps": hbm_peak, "peak_source": peak_src},
                "msm_total_ms": sum(tot_ms) / len(tot_ms)}

    line = {
        "metric": METRIC, "value": ms / 1e3 / (args.steps * world), "unit": "s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32-limb Montgomery (integer; Fr 8 limbs, Fq 12 limbs)", "data": "synthetic",
        "config": {"workload": args.workload, "describe": WORKLOADS[args.workload][2], "log_m": log_m, "n_in": n_in, "n_cons": shape["n_cons"], "n_vars": nv,
                   "partitions_per_step_per_gpu": 1,
                   "query_density": "about 0.64 of the a / b_g1 / b_g2 bases are non-identity in this synthetic instance (profiles/r01_key_cache_time_2p22.json); identity bases take no multiexp slot, as with bellman's density-filtered keys; l and h are dense",
                   "l2": "inputs larger than L2 (key queries %.0f MB, witness %.0f MB per proof)" % (5 * nv * 96 / 1e6, nv * 32 / 1e6),
                   "parallelism": "one partition per GPU, key replicated, no data-path collective"},
        "e2e": {"value": ms_e2e / 1e3 / (args.steps * world), "unit": "s", "h2d_bytes_per_step": nv * 32 + 64, "d2h_bytes_per_step": 192 + 672,
                "ms_per_step": ms_e2e / args.steps, "api": "b2g16_prove (host witness buffer, pinned)"},
        "gpu_launches": int(launches), "roofline": roofline, "clocks": clocks, "setup_s": t_setup,
    }

    if world > 1:
        # latency mode: the same proof with its five multiexps range-sharded over all ranks;
        # rank 0's witness everywhere (the other ranks' own witnesses belong to the weak-scaling figure above)
        from b2g16 import dist as bd
        dist.broadcast(dz, src=0)
        strong = []
        p_ref = bd.prove_sharded(ctx, pk, dz.data_ptr(), r, sr)
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); e0.record(); p_s = bd.prove_sharded(ctx, pk, dz.data_ptr(), r, sr); e1.record(); barrier()
            strong.append(max_over_ranks(e0.elapsed_time(e1)))
            assert p_s == p_ref
        full, peer_err = [], None
        peer = None
        if world in (2, 4, 8):
            peer, peer_err = try_peer_ntt(ctx, torch, dist, dev, log_m)
        if peer is not None:
            assert bd.prove_sharded_full(ctx, pk, peer, dz.data_ptr(), r, sr) == p_ref
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                barrier(); e0.record(); p_f = bd.prove_sharded_full(ctx, pk, peer, dz.data_ptr(), r, sr); e1.record(); barrier()
                full.append(max_over_ranks(e0.elapsed_time(e1)))
                assert p_f == p_ref
            peer.close()
        if rank == 0:
            assert p_ref == ctx.prove_device(pk, dz.data_ptr(), r, sr), "sharded proof differs from the single-GPU proof"
            line["one_proof_sharded"] = {"n_gpus": world, "prove_s": min(strong) / 1e3, "all_s": [t / 1e3 for t in strong], "scaling": "strong",
                                         "how": "every rank runs the witness map and reduces 1/N of each of the five queries (b2g16_prove_partial); all_gather of N x 576 B over NCCL; b2g16_prove_assemble",
                                         "speedup_vs_1gpu_step": (ms / args.steps) / min(strong),
                                         "witness_map_sharded_too": ({"unavailable": peer_err} if peer_err else None) if not full else {
                                             "prove_s": min(full) / 1e3, "all_s": [t / 1e3 for t in full], "speedup_vs_1gpu_step": (ms / args.steps) / min(full),
                                             "how": "rows k1 + N k2 of a, b, c per rank, seven transforms on the fused peer-memory path (b2g16.dist.PeerNtt), pointwise quotient, each rank's H block feeds its share of the H query (b2g16_prove_partial_h)"}}
    pk.free()
    del dz, hz
    if rank == 0 and not args.skip_standalone:
        line["standalone"] = standalone(ctx, torch, dev, args, hbm_peak, peak_src, mad_peak)
    if world > 1 and not args.skip_standalone:
        sh = msm_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, mad_peak)
        if rank == 0:
            line["msm_g1_sharded"] = sh
        if world in (2, 4, 8):
            shn = ntt_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, hbm_peak)
            if rank == 0:
                line["ntt_fr_sharded"] = shn
    if rank == 0 and world == 1 and not args.skip_32gib and args.workload != "porep-c2-32GiB":
        line["porep_c2_32GiB_partition"] = big_partition(ctx, torch, dev)
    if rank == 0 and world == 1 and not args.skip_cpu:
        sample_log = args.cpu_sample_log_m or min(18, log_m)
        v, threads, sample, _ = cpu_prove_sample(args.workload, sample_log)
        line["cpu_baseline"] = {"value": v, "unit": "s", "cores": threads, "kind": "port", "sample": sample}
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


def try_peer_ntt(ctx, torch, dist, dev, log_n):
    """PeerNtt needs CUDA IPC between the ranks' processes; if the environment forbids it, the add-on figures that
    use it are reported as unavailable instead of taking the whole bench line down.  Collective: every rank calls it."""
    from b2g16 import dist as bd
    peer, err = None, None
    try:
        peer = bd.PeerNtt(ctx, log_n)
    except Exception as e:               # noqa: BLE001 -- any failure here only disables an optional figure
        err = "%s: %s" % (type(e).__name__, e)
    ok = torch.tensor([0 if peer is None else 1], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) == 0:
        return None, err or "peer memory unavailable on another rank"
    return peer, None


def msm_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, mad_peak):
    """One G1 multiexp of 2^log_msm points sharded by contiguous point range over the ranks (strong
    scaling): every rank reduces its resident slice to an affine point, the partials are all-gathered
    over NCCL (96 B per rank) and summed on every rank.  Max over ranks, best of 3."""
    from b2g16 import dist as bd
    n = 1 << args.log_msm
    first, last = bd.shard_range(n, rank, world)
    cnt = last - first
    kk = dev_random_fr(torch, cnt, 1234 + rank, dev)
    bases = ctx.bases_generate(1, kk.data_ptr(), cnt)
    del kk
    sc = dev_random_fr(torch, cnt, 4321 + rank, dev)
    part = torch.zeros(96, dtype=torch.uint8, device=dev)
    allp = torch.zeros((world, 96), dtype=torch.uint8, device=dev)

    def step():
        ctx.msm_resident_async(bases, sc.data_ptr(), cnt, part.data_ptr())
        dist.all_gather_into_tensor(allp, part)
        return ctx.g1_sum(allp.cpu().numpy())
    res = step()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record(); res = step(); e1.record(); barrier()
        best = min(best, max_over_ranks(e0.elapsed_time(e1)))
    bases.free()
    ach = n * MAD32_PER_G1_POINT / (best * 1e-3)
    return {"log_n": args.log_msm, "n_gpus": world, "points_per_gpu": cnt, "ms": best, "Mpts_per_s": n / best / 1e3, "scaling": "strong",
            "collective": "all_gather of %d affine partials (96 B each) over NCCL, then b2g16_g1_sum" % world,
            "roofline": {"bound": "int32-imad", "achieved": ach / 1e12, "peak": world * mad_peak / 1e12, "unit": "TMAD32/s", "frac": ach / (world * mad_peak)},
            "result_head": bytes(res[:8]).hex()}


def ntt_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, hbm_peak):
    """One Fr transform of 2^log_ntt elements sharded over the ranks (strong scaling): rank r holds the
    block [r M, (r+1) M); two all-to-all exchanges around the cross-rank step + the local size-M transform
    give the spectrum in cyclic order, a third exchange + transpose the natural block order (SURVEY 8e row 3,
    b2g16.dist.ntt_sharded).  Max over ranks, best of 3; forward followed by inverse must return the input."""
    from b2g16 import dist as bd
    log_n = args.log_ntt
    M = (1 << log_n) // world
    ops = bd.DeviceNttOps(ctx)
    x = dev_random_fr(torch, M, 777 + rank, dev)
    out = {"log_n": log_n, "n_gpus": world, "elements_per_gpu": M, "scaling": "strong",
           "collective": "all_to_all_single over NCCL: %d exchanges of %.0f MB per rank (natural order), 2 (cyclic order)" % (3, M * 32 * (world - 1) / world / 1e6)}
    for order in ("natural", "cyclic"):
        spec = bd.ntt_sharded(ops, x, log_n, order=order)
        back = bd.ntt_sharded(ops, spec, log_n, inverse=True, order=order)
        assert bool(torch.equal(back, x)), "sharded forward + inverse transform is not the identity"
        best = 1e30
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); e0.record(); spec = bd.ntt_sharded(ops, x, log_n, order=order); e1.record(); barrier()
            best = min(best, max_over_ranks(e0.elapsed_time(e1)))
        gbps = 2.0 * (1 << log_n) * 32 / (best * 1e-3) / 1e9
        out[order] = {"ms": best, "roofline": {"bound": "hbm", "achieved": gbps, "peak": world * hbm_peak, "unit": "GB/s", "frac": gbps / (world * hbm_peak)}}
        del spec, back
    # cross step fused with its two exchanges over NVLink peer memory (b2g16.dist.PeerNtt), cyclic spectrum order
    pn, peer_err = try_peer_ntt(ctx, torch, dist, dev, log_n)
    if pn is None:
        out["cyclic_p2p_fused"] = {"unavailable": peer_err}
        return out
    pn.x.copy_(x)
    pn.forward()
    assert bool(torch.equal(pn.inverse(), x)), "peer-memory forward + inverse transform is not the identity"
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record(); pn.forward(); e1.record(); barrier()
        best = min(best, max_over_ranks(e0.elapsed_time(e1)))
    gbps = 2.0 * (1 << log_n) * 32 / (best * 1e-3) / 1e9
    out["cyclic_p2p_fused"] = {"ms": best, "how": "one kernel loads its column chunk from every rank's block and stores into every rank's spectrum buffer over NVLink (CUDA IPC peer memory); two barriers instead of two all-to-alls",
                               "roofline": {"bound": "hbm", "achieved": gbps, "peak": world * hbm_peak, "unit": "GB/s", "frac": gbps / (world * hbm_peak)}}
    pn.close()
    return out


def big_partition(ctx, torch, dev, log_m=27, n_in=327, reps=2):
    """The target of BASELINE.json's north_star: one 32 GiB-sector-shaped PoRep C2 partition (domain 2^27,
    134,217,400 constraints, 327 public inputs) proved on one GPU with the key resident (~150 GB of HBM).
    Device-generated instance; best of `reps` device-resident proofs."""
    from b2g16 import codec
    seed, n_free = 2027, 8
    toxic, r, sr = toxic_from_seed(seed)
    t0 = time.perf_counter()
    pk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic)
    shape = pk.shape()
    rng = np.random.Generator(np.random.PCG64(seed))
    head = codec.random_fr(1 + n_in + n_free, rng)
    head[0] = 0
    head[0, 0] = 1
    dz = torch.zeros((shape["n_vars"], 32), dtype=torch.uint8, device=dev)
    dz[: head.shape[0]] = torch.from_numpy(head).to(dev)
    ctx.synth_witness(pk, dz.data_ptr(), n_free)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    times = []
    for _ in range(reps + 1):                               # first one is the warm-up (domain tables, pool growth)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); proof = ctx.prove_device(pk, dz.data_ptr(), r, sr); e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) / 1e3)
    free_b, total_b = torch.cuda.mem_get_info()
    pk.free()
    del dz
    return {"log_m": log_m, "n_cons": shape["n_cons"], "n_in": n_in, "prove_s": min(times[1:]), "all_s": times, "setup_s": setup_s,
            "hbm_used_GB": (total_b - free_b) / 1e9, "proof_len": len(proof),
            "parity": "size-independent property: proofs of this shape pass the pairing check (tools/large_proof.py; tests/test_gpu_parity.py::test_large_proof_verifies at 2^20)"}


def standalone(ctx, torch, dev, args, hbm_peak, peak_src, mad_peak):
    """The two standalone figures BASELINE.json's metric names: G1 MSM Mpts/s at 2^log_msm and
    Fr NTT ms at 2^log_ntt, device-resident inputs, best of 3 after a warm-up."""
    out = {}
    n = 1 << args.log_msm
    kk = dev_random_fr(torch, n, 1234, dev)
    bases = ctx.bases_generate(1, kk.data_ptr(), n)
    del kk
    sc = dev_random_fr(torch, n, 4321, dev)
    res = torch.zeros(192, dtype=torch.uint8, device=dev)
    ctx.msm_resident_async(bases, sc.data_ptr(), n, res.data_ptr())
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); ctx.msm_resident_async(bases, sc.data_ptr(), n, res.data_ptr()); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del sc
    sk = dev_filecoin_like_fr(torch, n, 2468, dev)
    ctx.msm_resident_async(bases, sk.data_ptr(), n, res.data_ptr())
    best_sk = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); ctx.msm_resident_async(bases, sk.data_ptr(), n, res.data_ptr()); e1.record(); torch.cuda.synchronize()
        best_sk = min(best_sk, e0.elapsed_time(e1))
    bases.free()
    del sk
    ach = n * MAD32_PER_G1_POINT / (best * 1e-3)
    out["msm_g1"] = {"log_n": args.log_msm, "ms": best, "Mpts_per_s": n / best / 1e3,
                     "roofline": {"bound": "int32-imad", "achieved": ach / 1e12, "peak": mad_peak / 1e12, "unit": "TMAD32/s", "frac": ach / mad_peak},
                     "filecoin_like_scalars": {"ms": best_sk, "Mpts_per_s": n / best_sk / 1e3,
                                               "mix": "60 % zero, 20 % one, 10 % < 2^16, 10 % uniform (SURVEY 8d); reported beside the uniform worst case, not instead of it"}}
    n = 1 << args.log_ntt
    x = dev_random_fr(torch, n, 77, dev)
    ctx.ntt_device(x.data_ptr(), args.log_ntt)
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); ctx.ntt_device(x.data_ptr(), args.log_ntt); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del x
    gbs = 2.0 * n * 32 / (best * 1e-3) / 1e9
    out["ntt_fr"] = {"log_n": args.log_ntt, "ms": best, "order": "natural in, natural out, canonical bytes both sides",
                     "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "peak_source": peak_src,
                                  "alg_bytes": 2.0 * n * 32,
                                  # SURVEY 8(d) asks for both views: (N/2) log2 N butterfly products x 136 MAD32 (8-limb CIOS)
                                  "int32_view": {"alg_mad32": 0.5 * n * args.log_ntt * 136.0, "achieved": 0.5 * n * args.log_ntt * 136.0 / (best * 1e-3) / 1e12,
                                                 "peak": mad_peak / 1e12, "unit": "TMAD32/s", "frac": 0.5 * n * args.log_ntt * 136.0 / (best * 1e-3) / mad_peak,
                                                 "note": "the transform is bound by the integer multiply pipe on this hardware (ncu: math-pipe throttle, DRAM at 7 % of peak), not by HBM"}}}
    return out


def ncu_traffic(log_m):
    """dram__bytes_read.sum + dram__bytes_write.sum of one accumulate launch from the committed
    `ncu --set full` capture (profiles/r01_ncu_full_s13.json: G1 MSM at 2^22); null at other sizes."""
    path = os.path.join(ROOT, "profiles", "r01_ncu_full_s13.json")
    if log_m != 22 or not os.path.exists(path):
        return None
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    for l in json.load(open(path))["launches"]:
        if "msm_accumulate" in l.get("Kernel Name", ""):
            tot = 0.0
            for k, v in l.items():
                if k.startswith("dram__bytes_read.sum") or k.startswith("dram__bytes_write.sum"):
                    tot += float(v) * scale[k.split("[")[1].rstrip("]")]
            return tot
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="porep-c2-2KiB", choices=sorted(WORKLOADS))
    ap.add_argument("--log-msm", type=int, default=26)
    ap.add_argument("--log-ntt", type=int, default=27)
    ap.add_argument("--cpu-sample-log-m", type=int, default=0)
    ap.add_argument("--skip-standalone", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-32gib", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
