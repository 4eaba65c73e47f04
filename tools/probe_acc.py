"""Accumulate-kernel variant sweep (B2G16_ACC_VARIANT) at one MSM size."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
    import torch
    from b2g16 import lib
    from probe import dev_random_fr, timed
    log_n = int(sys.argv[2])
    ts = torch.cuda.Stream(); torch.cuda.set_stream(ts)
    ctx = lib.Context(0, ts.cuda_stream)
    k = dev_random_fr(1 << log_n, 99); s = dev_random_fr(1 << log_n, 7)
    bases = ctx.bases_generate(1, k.data_ptr(), 1 << log_n)
    out = torch.zeros(192, dtype=torch.uint8, device="cuda")
    ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
    t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr()), reps=3)
    ctx.set_msm_timing(True); ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
    print(json.dumps({"variant": os.environ.get("B2G16_ACC_VARIANT"), "log_n": log_n, "ms": round(t, 2), "stages": {k: round(v, 2) for k, v in ctx.last_msm_timing().items()},
                      "sum": int(out.sum().item())}), flush=True)
else:
    for v in sys.argv[2:]:
        env = dict(os.environ, B2G16_ACC_VARIANT=v)
        subprocess.call([sys.executable, os.path.abspath(__file__), "child", sys.argv[1]], env=env)
