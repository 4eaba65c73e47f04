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
BENCH_SEED, BENCH_N_FREE = 2026, 8


def witness_head(n_in, rank=0):
    """Inputs and free variables of the benchmark witness (the rest follows by evaluating the rows forward)."""
    from b2g16 import codec
    rng = np.random.Generator(np.random.PCG64(9000 + rank))
    head = codec.random_fr(1 + n_in + BENCH_N_FREE, rng)
    head[0] = 0
    head[0, 0] = 1
    return head


def cpu_threads(coracle):
    # all host cores, whatever the launcher exported (torchrun sets OMP_NUM_THREADS=1 for its workers)
    try:
        coracle.set_threads(len(os.sched_getaffinity(0)))
    except AttributeError:
        coracle.set_threads(os.cpu_count() or 1)
    return coracle.threads()


def cpu_instance(log_m, n_in):
    """The benchmark instance on the host: synth.HashedR1cs is the numpy mirror of the constraint system
    b2g16_pk_generate_synth builds on the device (same seed -> same matrices, tests/test_gpu_parity.py
    ::test_hashed_instance_device_equals_host), the witness is evaluated forward by the oracle."""
    from oracle import coracle
    from b2g16 import synth
    coracle.build()
    s = synth.HashedR1cs(log_m, n_in, BENCH_SEED, n_free=BENCH_N_FREE)
    cs = coracle.R1csArrays(s.log_m, s.n_in, s.n_aux, s.A, s.B, s.C)
    z = coracle.synth_witness(cs, witness_head(n_in), BENCH_N_FREE)
    return coracle, s, cs, z


def cpu_prove_timed(coracle, key, z, r, sr, steps, warmup, budget_s):
    """Times the oracle's OpenMP Groth16 prover (oracle/c/oracle.cpp) on all host threads: `warmup` untimed proofs,
    then up to `steps` timed ones, stopping early once `budget_s` seconds of timed work are spent (the count that was
    actually run is returned).  -> (seconds per proof, [per-step seconds], proof bytes)"""
    times, proof = [], None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        proof, _ = coracle.groth16_prove(key, z, r, sr, want_trace=False)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            if sum(times) + dt > budget_s and len(times) < steps:
                break
    assert len(proof) == 192
    return sum(times) / len(times), times, proof


def run_reference(args):
    """`--impl reference`: the CPU prover on the SAME instance the B200 arm proves (same constraint system, same
    witness, same r and s, hence the same 192 bytes), all host cores, no GPU anywhere on the path -- the key comes from
    the oracle's own CPU setup (untimed).  The reference's own prover sources are absent from its tree,
    so this is the oracle restatement: kind = "port"."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t_wall = time.perf_counter()
    log_m, n_in = WORKLOADS[args.workload][0], n_in_of(args.workload)
    same_config = True
    if args.cpu_sample_log_m and args.cpu_sample_log_m < log_m:
        log_m, same_config = args.cpu_sample_log_m, False            # explicit request for a reduced sample
    elif log_m > 23:
        log_m, same_config = 22, False                               # a 2^27 CPU proof takes ~25 min: bounded sample, says so below
    t0 = time.perf_counter()
    coracle, s, cs, z = cpu_instance(log_m, n_in)
    threads = cpu_threads(coracle)
    toxic, r, sr = toxic_from_seed(BENCH_SEED)
    key = coracle.groth16_setup(cs, toxic)
    setup_s = time.perf_counter() - t0
    budget = float(os.environ.get("B2G16_REF_BUDGET_S", "600"))
    per, times, proof = cpu_prove_timed(coracle, key, z, r, sr, args.steps, min(args.warmup, 1), budget)
    full_log_m = WORKLOADS[args.workload][0]
    scale = float(1 << (full_log_m - log_m))
    value = per * scale
    sample = ("oracle C++/OpenMP Groth16 prover, %d proofs of the benchmark's own instance (domain 2^%d, %d constraints, %d variables), %.2f s/proof"
              % (len(times), log_m, s.n_cons, s.n_vars, per))
    if not same_config:
        sample += "; scaled x%d linearly to the 2^%d workload (an extrapolation, not a measurement)" % (int(scale), full_log_m)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "s", "n_gpus": args.gpus, "steps": len(times), "steps_requested": args.steps,
        "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * per, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64-limb Montgomery (integer)", "data": "synthetic",
        "config": {"workload": args.workload, "describe": WORKLOADS[args.workload][2], "log_m": log_m, "n_in": n_in, "n_cons": s.n_cons, "n_vars": s.n_vars,
                   "partitions_per_step_per_gpu": 1, "same_instance_as_b200_arm": same_config},
        "cpu_baseline": {"value": value, "unit": "s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "setup_s": setup_s, "wall_s": time.perf_counter() - t_wall, "all_s": times, "proof_head": proof[:8].hex(),
        "note": "the reference's own prover sources are absent from its tree (un-vendored crypto3 submodules), so kind=port: the oracle restatement; "
                "steps are capped by B2G16_REF_BUDGET_S (%.0f s of timed work), the count run is `steps`" % budget,
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
    log_m, n_in, n_free, seed = WORKLOADS[args.workload][0], n_in_of(args.workload), BENCH_N_FREE, BENCH_SEED
    toxic, r, sr = toxic_from_seed(seed)
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
    shape = pk.shape()
    nv = shape["n_vars"]
    head = witness_head(n_in, rank)                              # a different witness per rank: own inputs and free variables
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
    proof = proofs[0]

    # ---- parity of the proof that was timed (outside the timed region): the native verifier (GPU multiexp over the IC
    #      bases + host pairing, csrc/verify.cu) accepts it and rejects it for a perturbed public input
    pvk = ctx.vk_prepare(vk)
    pub = head[1:1 + n_in]
    bad_pub = pub.copy(); bad_pub[0, 0] ^= 1
    parity = {"native_verify": bool(ctx.verify(pvk, pub, proof)), "perturbed_input_rejected": not ctx.verify(pvk, bad_pub, proof)}
    pvk.free()
    assert parity["native_verify"] and parity["perturbed_input_rejected"], "the timed proof does not verify"

    # ---- dominant kernel, timed on the proof's own launches: one more proof of the same instance with its five
    #      multiexps run one after the other (so that no two accumulate kernels share the SMs) and CUDA events around
    #      each bucket-accumulation launch; roofline on the int32 multiply pipe
    mad_peak, _ = ctx.mad32_peak()
    roofline, acc_log = proof_roofline(ctx, lambda: ctx.prove_device(pk, dz.data_ptr(), r, sr), proof, log_m, mad_peak, hbm_peak, peak_src)

    line = {
        "metric": METRIC, "value": ms / 1e3 / (args.steps * world), "unit": "s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32-limb Montgomery (integer; Fr 8 limbs, Fq 12 limbs)", "data": "synthetic",
        "config": {"workload": args.workload, "describe": WORKLOADS[args.workload][2], "log_m": log_m, "n_in": n_in, "n_cons": shape["n_cons"], "n_vars": nv,
                   "partitions_per_step_per_gpu": 1,
                   "query_density": "about 0.64 of the a / b_g1 / b_g2 bases are non-identity in this synthetic instance; identity bases take no multiexp slot, as with bellman's density-filtered keys; l and h are dense",
                   "l2": "inputs larger than L2 (key queries %.0f MB, witness %.0f MB per proof)" % (5 * nv * 96 / 1e6, nv * 32 / 1e6),
                   "parallelism": "one partition per GPU, key replicated, no data-path collective"},
        "e2e": {"value": ms_e2e / 1e3 / (args.steps * world), "unit": "s", "h2d_bytes_per_step": nv * 32 + 64, "d2h_bytes_per_step": 192 + 672,
                "ms_per_step": ms_e2e / args.steps, "api": "b2g16_prove (host witness buffer, pinned)"},
        "gpu_launches": int(launches), "roofline": roofline, "accumulate_launches_in_proof": acc_log, "clocks": clocks, "setup_s": t_setup,
        "parity_checked": True, "parity": parity,
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
                                         "speedup_vs_1gpu_step": (ms / args.steps) / min(strong), "equals_single_gpu_proof": True,
                                         "witness_map_sharded_too": ({"unavailable": peer_err} if peer_err else None) if not full else {
                                             "prove_s": min(full) / 1e3, "all_s": [t / 1e3 for t in full], "speedup_vs_1gpu_step": (ms / args.steps) / min(full),
                                             "how": "rows k1 + N k2 of a, b, c per rank, seven transforms on the fused peer-memory path (b2g16.dist.PeerNtt), pointwise quotient, each rank's H block feeds its share of the H query (b2g16_prove_partial_h)"}}
    # ---- CPU baseline leg (rank 0, N = 1): the oracle's prover on the SAME instance with the SAME (device-generated,
    #      exported) key -- one full proof on all host cores, and its 192 bytes must equal the GPU's
    cpu_fit = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        line["cpu_baseline"], cpu_fit, par = cpu_baseline_leg(ctx, torch, dev, args, pk, vk, hz_np, head, proof, log_m, n_in, r, sr)
        line["parity"].update(par)
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
    if not args.skip_32gib and args.workload not in ("porep-c2-32GiB", "window-post-32GiB"):
        # BASELINE configs[3] / configs[4] at this GPU count: one 32 GiB-shaped PoRep partition and one WindowPoSt-shaped
        # proof (2^27 domain) on one GPU and spread over all ranks; at 8 GPUs also all 10 partitions of a 32 GiB PoRep
        big = {}
        for wl in ("porep-c2-32GiB", "window-post-32GiB"):
            big[wl] = big_proofs(ctx, torch, dist if world > 1 else None, dev, wl, rank, world, barrier, max_over_ranks, mad_peak, hbm_peak, peak_src, cpu_fit)
        if rank == 0:
            line["porep_c2_32GiB_partition"] = big["porep-c2-32GiB"]
            line["window_post_32GiB"] = big["window-post-32GiB"]
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


def proof_roofline(ctx, prove_fn, want_proof, log_m, mad_peak, hbm_peak, peak_src):
    """One proof with its multiexps serialised and the accumulate launches bracketed by CUDA events (b2g16_kernel_log).
    The dominant launch is the H query's (m - 1 dense bases, uniform 255-bit scalars): `achieved` = its algorithmic
    MAD32 (48,000 per point) / its duration."""
    ctx.set_tuning("serial_above_log", 0)
    ctx.set_tuning("kernel_log", 1)
    try:
        p = prove_fn()
        log = ctx.kernel_log()
    finally:
        ctx.set_tuning("kernel_log", 0)
        ctx.set_tuning("serial_above_log", 24)
    assert want_proof is None or p == want_proof, "serialised proof differs"
    n_h = (1 << log_m) - 1
    h = [e for e in log if e["group"] == 1 and e["n"] == n_h]
    assert h, "H-query launch missing from the kernel log"
    e = h[-1]
    g2 = [x for x in log if x["group"] == 2]
    achieved = n_h * MAD32_PER_G1_POINT / (e["ms"] * 1e-3) / 1e12
    kname = {1: "tree_down0_kernel<Fq> + tree_up0_kernel<Fq> (batched-affine tree levels, ~85 % of the stage) + msm_accumulate_chunked_kernel<Fq,128,2,direct> (XYZZ tail)",
             2: "tree_down0_kernel<Fq2> + tree_up0_kernel<Fq2> + msm_accumulate_chunked_kernel<Fq2,64,4,direct>"}
    # what the stage really executes: an affine addition with a shared inversion is 6 field products (G1: 6 x 300 MAD32;
    # G2: 5 lazy-reduction Fq2 products of 744 + one Fq2 squaring of 600 = 4320), not the 10 / 30 of the mixed XYZZ addition
    # that SURVEY's per-point figure assumes -- so `frac`, which keeps SURVEY's algorithmic work as the numerator, can
    # exceed the share of the multiplier the stage keeps busy, reported beside it as pipe_utilisation_actual.
    roofline = {"kernel": kname[1] + " -- bucket accumulation stage of the H query, timed inside a proof (multiexps serialised for this one proof)",
                "bound": "int32-imad", "achieved": achieved, "peak": mad_peak / 1e12, "unit": "TMAD32/s", "frac": achieved / (mad_peak / 1e12),
                "traffic": ncu_traffic(log_m),
                "traffic_note": "DRAM bytes of the whole accumulation stage of one 2^22 G1 multiexp (all its launches) from the committed ncu --set full capture (profiles/r02_ncu_full_msm_g1_2p22.json); every tree level re-reads its entries twice (prefix pass + addition pass), SURVEY's N*128 B counts one read of the inputs",
                "launch_ms": e["ms"], "units_per_launch": n_h, "window_bits": e["c"], "additions_in_launch": e["entries"],
                "alg_mad32_per_point": MAD32_PER_G1_POINT,
                "actual_mad32_per_addition": 1800, "pipe_utilisation_actual": e["entries"] * 1800.0 / (e["ms"] * 1e-3) / mad_peak,
                "peak_source": "b2g16_measure_mad32_peak (dependency-free mad.wide.u32 loop, this run)",
                "hbm_view": {"alg_bytes": n_h * 128, "GBps": n_h * 128 / (e["ms"] * 1e-3) / 1e9, "peak_GBps": hbm_peak, "peak_source": peak_src}}
    if g2:
        x = g2[-1]
        roofline["g2_launch"] = {"kernel": kname[2], "launch_ms": x["ms"], "window_bits": x["c"], "additions_in_launch": x["entries"], "bases_in_slice": x["n"],
                                 "actual_mad32_per_addition": 4320, "pipe_utilisation_actual": x["entries"] * 4320.0 / (x["ms"] * 1e-3) / mad_peak,
                                 "note": "b_g2 query: identity bases and zero scalars take no slot, so the additions, not the slice length, are the work"}
    return roofline, log


def cpu_baseline_leg(ctx, torch, dev, args, pk, vk, hz_np, head, proof, log_m, n_in, r, sr):
    """-> (cpu_baseline dict, fit points {log_m: seconds}, parity dict).  The CPU prover runs the benchmark's own
    instance when its domain is <= 2^23 (one full proof, bit-exact against the GPU's); smaller samples of the same
    shape (2^18, 2^20, device-generated keys exported to the oracle, each checked bit-exact too) give the fit that
    the 2^27 figures are extrapolated with."""
    from b2g16 import synth
    toxic, _, _ = toxic_from_seed(BENCH_SEED)
    budget = float(os.environ.get("B2G16_CPU_BUDGET_S", "120"))
    fit, par = {}, {}

    def one(lm, pk_l, z_l, want):
        coracle, s, cs, z = cpu_instance(lm, n_in)
        threads = cpu_threads(coracle)
        if z_l is not None:
            assert np.array_equal(z, z_l), "device witness differs from the oracle's"
        key = coracle.Key(cs)
        m, nvv = 1 << lm, cs.n_vars
        for which, n, cnt in ((0, "a", nvv), (1, "b_g1", nvv), (2, "b_g2", nvv), (3, "l", cs.n_aux), (4, "h", m - 1)):
            setattr(key, n, pk_l.export(which, cnt))
        for which, n in ((5, "alpha_g1"), (6, "beta_g1"), (7, "delta_g1"), (8, "beta_g2"), (9, "delta_g2")):
            setattr(key, n, pk_l.export(which, 1)[0])
        per, times, cpu_proof = cpu_prove_timed(coracle, key, z, r, sr, 1, 0, budget)
        if want is None:
            dzl = torch.from_numpy(z).to(dev)
            want = ctx.prove_device(pk_l, dzl.data_ptr(), r, sr)
            del dzl
        assert cpu_proof == want, "GPU proof differs from the CPU oracle's at domain 2^%d" % lm
        return per, threads, s, coracle, z

    full = log_m <= 23
    top = log_m if full else 22
    if full:
        per, threads, s, coracle, z = one(log_m, pk, hz_np, proof)
        import util
        par["oracle_pairing_verify"] = bool(util.pairing_verify(coracle, vk, head[1:1 + n_in], proof))
        assert par["oracle_pairing_verify"]
    else:
        pk_t = ctx.pk_generate_synth(top, n_in, BENCH_N_FREE, BENCH_SEED, toxic)
        per, threads, s, coracle, z = one(top, pk_t, None, None)
        pk_t.free()
    fit[top] = per
    par["bit_exact_vs_cpu_oracle"] = {"domain_log2": top, "equal": True}
    for lm in (18, 20):
        if lm < top:
            pk_s = ctx.pk_generate_synth(lm, n_in, BENCH_N_FREE, BENCH_SEED, toxic)
            fit[lm] = one(lm, pk_s, None, None)[0]
            pk_s.free()
    scale = float(1 << (log_m - top))
    sample = ("oracle C++/OpenMP Groth16 prover, 1 proof of the benchmark's own instance (domain 2^%d, %d constraints, %d variables, device-generated key exported to the host): %.2f s; its 192 bytes equal the GPU proof's"
              % (top, s.n_cons, s.n_vars, per))
    if not full:
        sample += "; scaled x%d linearly to 2^%d (an extrapolation)" % (int(scale), log_m)
    return {"value": per * scale, "unit": "s", "cores": threads, "kind": "port", "sample": sample, "same_config": full,
            "other_sizes_s": {"2^%d" % k: v for k, v in sorted(fit.items())}}, fit, par


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
    from b2g16 import codec
    kk = dev_random_fr(torch, cnt, 1234 + rank, dev)
    bases = ctx.bases_generate(1, kk.data_ptr(), cnt)
    sc = dev_random_fr(torch, cnt, 4321 + rank, dev)
    # scalar side of the linearity identity: every rank's sum_i s_i k_i, gathered and added mod r
    dots = bd.all_gather_bytes(ctx.fr_dot_device(kk.data_ptr(), sc.data_ptr(), cnt))
    tot = sum(codec.fr_to_ints(dots.reshape(world, 32))) % codec.R_MOD
    del kk
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
    one = torch.from_numpy(codec.fr_from_ints([tot])).to(dev)
    b1 = ctx.bases_generate(1, one.data_ptr(), 1)
    want = b1.download()[0]
    b1.free()
    assert np.array_equal(np.asarray(res).reshape(-1)[:96], want), "sharded multiexp violates sum s_i (k_i G) = (sum s_i k_i) G"
    ach = n * MAD32_PER_G1_POINT / (best * 1e-3)
    return {"log_n": args.log_msm, "n_gpus": world, "points_per_gpu": cnt, "parity_checked": True,
            "parity": "sum over all ranks of s_i (k_i G) equals (sum_i s_i k_i) G -- the value a single GPU would produce", "ms": best, "Mpts_per_s": n / best / 1e3, "scaling": "strong",
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


def big_proofs(ctx, torch, dist, dev, workload, rank, world, barrier, max_over_ranks, mad_peak, hbm_peak, peak_src, cpu_fit, reps=2):
    """BASELINE configs[3] / configs[4]: a 2^27-domain proof (32 GiB-sector-shaped PoRep C2 partition, 327 public inputs;
    WindowPoSt-shaped, 25,839 public inputs) with the key resident (~150 GB of HBM).
      * on one GPU (every rank proves the same instance at the same time; max over ranks), with the accumulate kernel's
        roofline from that proof's own launches (multiexps already run one after the other above 2^24);
      * world > 1: the same proof spread over all ranks (five multiexps range-sharded; then the witness map sharded too);
      * world == 8 and PoRep: all 10 partitions of a 32 GiB sector under one key (8 x 1 GPU + 2 x 4 GPUs).
    Every proof is checked by the native verifier; sharded proofs must equal the single-GPU bytes."""
    from b2g16 import codec, shapes
    log_m, n_in, n_free, seed = WORKLOADS[workload][0], n_in_of(workload), BENCH_N_FREE, 2027
    torch.cuda.synchronize()
    torch.cuda.empty_cache()          # blocks torch's allocator still caches from the earlier legs are invisible to the library's pool
    toxic, r, sr = toxic_from_seed(seed)
    t0 = time.perf_counter()
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
    shape = pk.shape()
    nv = shape["n_vars"]
    head = codec.random_fr(1 + n_in + n_free, np.random.Generator(np.random.PCG64(seed)))
    head[0] = 0
    head[0, 0] = 1
    dz = torch.zeros((nv, 32), dtype=torch.uint8, device=dev)
    dz[: head.shape[0]] = torch.from_numpy(head).to(dev)
    ctx.synth_witness(pk, dz.data_ptr(), n_free)           # same seed and head on every rank: identical witnesses
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0

    def timed(fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record(); p = fn(); e1.record(); barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / 1e3, p

    one, proof = [], None
    for _ in range(reps + 1):                               # first one is the warm-up (domain tables, pool growth)
        t, p = timed(lambda: ctx.prove_device(pk, dz.data_ptr(), r, sr))
        assert proof is None or p == proof
        one.append(t); proof = p
    roofline, acc_log = proof_roofline(ctx, lambda: ctx.prove_device(pk, dz.data_ptr(), r, sr), proof, log_m, mad_peak, hbm_peak, peak_src)
    free_b, total_b = torch.cuda.mem_get_info()
    out = {"workload": workload, "describe": WORKLOADS[workload][2], "log_m": log_m, "n_cons": shape["n_cons"], "n_in": n_in, "n_vars": nv,
           "prove_s": min(one[1:]), "all_s": one, "setup_s": setup_s, "hbm_used_GB": (total_b - free_b) / 1e9, "proof_len": len(proof),
           "roofline": roofline, "accumulate_launches_in_proof": acc_log}
    if rank == 0:
        pvk = ctx.vk_prepare(vk)
        pub = head[1:1 + n_in]
        bad = pub.copy(); bad[n_in - 1, 0] ^= 1
        out["parity"] = {"native_verify": bool(ctx.verify(pvk, pub, proof)), "perturbed_input_rejected": not ctx.verify(pvk, bad, proof),
                         "note": "size-independent property (prove -> verify round trip); bit-exact parity against the CPU oracle is run at 2^20 and 2^22 (tests/test_gpu_parity.py, bench parity)"}
        assert out["parity"]["native_verify"] and out["parity"]["perturbed_input_rejected"]
        pvk.free()
        if cpu_fit and len(cpu_fit) >= 2:
            # t = a * n^b through the measured CPU proofs (least squares in log-log), evaluated at 2^27: an extrapolation
            xs = np.array(sorted(cpu_fit)); ys = np.log2([cpu_fit[k] for k in xs])
            bcoef, acoef = np.polyfit(xs, ys, 1)
            est = float(2 ** (acoef + bcoef * log_m))
            out["cpu_baseline"] = {"value": est, "unit": "s", "kind": "port", "extrapolated": True,
                                   "sample": "fit t = 2^(%.3f + %.3f log2 n) through CPU proofs measured in this run at %s" % (acoef, bcoef, ", ".join("2^%d: %.2f s" % (k, cpu_fit[k]) for k in xs))}
    if world > 1:
        from b2g16 import dist as bd
        sh = []
        for _ in range(reps + 1):
            t, p = timed(lambda: bd.prove_sharded(ctx, pk, dz.data_ptr(), r, sr))
            assert p == proof, "sharded proof differs from the single-GPU proof"
            sh.append(t)
        out["sharded_over_all_gpus"] = {"n_gpus": world, "prove_s": min(sh[1:]), "all_s": sh, "speedup_vs_1gpu": min(one[1:]) / min(sh[1:]), "equals_single_gpu_proof": True,
                                        "how": "five multiexps range-sharded over the ranks (b2g16_prove_partial), witness map on every rank"}
        peer, peer_err = (None, "needs 2, 4 or 8 ranks")
        if world in (2, 4, 8):
            peer, peer_err = try_peer_ntt(ctx, torch, dist, dev, log_m)
        if peer is not None:
            fu = []
            for _ in range(reps + 1):
                t, p = timed(lambda: bd.prove_sharded_full(ctx, pk, peer, dz.data_ptr(), r, sr))
                assert p == proof, "fully sharded proof differs from the single-GPU proof"
                fu.append(t)
            peer.close()
            out["sharded_over_all_gpus"]["witness_map_sharded_too"] = {"prove_s": min(fu[1:]), "all_s": fu, "speedup_vs_1gpu": min(one[1:]) / min(fu[1:]), "equals_single_gpu_proof": True,
                                                                        "how": "rows k1 + N k2 of a, b, c per rank, seven transforms on the fused peer-memory path, each rank's H block feeds its share of the H query"}
        else:
            out["sharded_over_all_gpus"]["witness_map_sharded_too"] = {"unavailable": peer_err}
        if world == 8 and workload == "porep-c2-32GiB":
            n_part = shapes.POREP_PARTITIONS[32 << 30]
            sched = bd.PartitionScheduler(n_part, ctx=ctx, log_m=log_m)
            heads, dzs = {}, {}
            for k in range(n_part):
                heads[k] = codec.random_fr(1 + n_in + n_free, np.random.Generator(np.random.PCG64(5000 + k)))
                heads[k][0] = 0; heads[k][0, 0] = 1
            dz = None
            for k, ranks in sched.plan:
                if rank in ranks:
                    d = torch.zeros((nv, 32), dtype=torch.uint8, device=dev)
                    d[: heads[k].shape[0]] = torch.from_numpy(heads[k]).to(dev)
                    ctx.synth_witness(pk, d.data_ptr(), n_free)
                    dzs[k] = d
            rs = codec.random_fr(n_part, np.random.Generator(np.random.PCG64(77)))
            ss = codec.random_fr(n_part, np.random.Generator(np.random.PCG64(78)))
            times, multi = [], None
            for _ in range(2):
                t, m = timed(lambda: sched.prove_all(ctx, pk, lambda k: dzs[k].data_ptr(), rs, ss))
                assert multi is None or m == multi
                times.append(t); multi = m
            res = {"partitions": n_part, "n_gpus": world, "multi_proof_bytes": len(multi), "prove_all_s": min(times), "all_s": times,
                   "schedule": [[k, ranks] for k, ranks in sched.plan], "one_gpu_all_partitions_s": n_part * min(one[1:]),
                   "speedup_vs_1gpu": n_part * min(one[1:]) / min(times)}
            if rank == 0:
                pvk = ctx.vk_prepare(vk)
                pubs = np.stack([heads[k][1:1 + n_in] for k in range(n_part)])
                res["batch_verified_natively"] = bool(ctx.verify_batch(pvk, pubs, multi))
                badp = pubs.copy(); badp[n_part - 1, 0, 0] ^= 1
                res["perturbed_input_rejected"] = not ctx.verify_batch(pvk, badp, multi)
                assert res["batch_verified_natively"] and res["perturbed_input_rejected"]
                pvk.free()
            sched.close()
            dzs.clear()
            out["porep_c2_32GiB_all_partitions"] = res
    pk.free()
    return out


def standalone(ctx, torch, dev, args, hbm_peak, peak_src, mad_peak):
    """The two standalone figures BASELINE.json's metric names: G1 MSM Mpts/s at 2^log_msm and
    Fr NTT ms at 2^log_ntt, device-resident inputs, best of 3 after a warm-up."""
    out = {}
    n = 1 << args.log_msm
    kk = dev_random_fr(torch, n, 1234, dev)
    bases = ctx.bases_generate(1, kk.data_ptr(), n)
    sc = dev_random_fr(torch, n, 4321, dev)
    res = torch.zeros(192, dtype=torch.uint8, device=dev)
    ctx.msm_resident_async(bases, sc.data_ptr(), n, res.data_ptr())
    # parity at full size through linearity: the bases are k_i G, so sum_i s_i (k_i G) must equal (sum_i s_i k_i) G
    linear_ok = msm_linearity_check(ctx, torch, dev, kk, sc, n, res)
    assert linear_ok, "G1 multiexp at 2^%d violates sum s_i (k_i G) = (sum s_i k_i) G" % args.log_msm
    del kk
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
    out["msm_g1"] = {"log_n": args.log_msm, "ms": best, "Mpts_per_s": n / best / 1e3, "parity_checked": True,
                     "parity": "linearity identity sum_i s_i (k_i G) = (sum_i s_i k_i) G holds for the timed inputs (scalar side by b2g16_fr_dot_device)",
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


def msm_linearity_check(ctx, torch, dev, kk, sc, n, res, extra_dot=None):
    """res (device, ABI affine) == (sum_i sc_i kk_i [+ extra_dot]) G ?  The dot product runs on the device
    (b2g16_fr_dot_device), the single fixed-base multiplication through b2g16_bases_generate."""
    from b2g16 import codec
    tot = codec.fr_to_ints(ctx.fr_dot_device(kk.data_ptr(), sc.data_ptr(), n).reshape(1, 32))[0]
    if extra_dot is not None:
        tot = (tot + extra_dot) % codec.R_MOD
    one = torch.from_numpy(codec.fr_from_ints([tot])).to(dev)
    b = ctx.bases_generate(1, one.data_ptr(), 1)
    want = b.download()[0]
    b.free()
    torch.cuda.synchronize()
    return bool(np.array_equal(res[:96].cpu().numpy(), want))


def ncu_traffic(log_m):
    """dram__bytes_read.sum + dram__bytes_write.sum over the accumulation-stage launches of one 2^22 G1 multiexp, from the
    committed `ncu --set full` capture (profiles/r02_ncu_full_msm_g1_2p22.json, written by tools/ncu_stage_summary.py);
    null at other sizes."""
    path = os.path.join(ROOT, "profiles", "r02_ncu_full_msm_g1_2p22.json")
    if log_m != 22 or not os.path.exists(path):
        return None
    return float(json.load(open(path))["total_dram_bytes"])


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
