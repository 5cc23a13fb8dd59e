"""configs[2] decode forward (GPT-2 small bf16, B=64, cache 512+) under environment variants, CUDA events over graph replays.
Usage: python tools/bench_decode_step.py [VAR=VALUE[,VAR=VALUE] ...]   (the empty variant "-" is the default build)"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] != "--run":
    for variant in sys.argv[1:]:
        env = dict(os.environ)
        if variant != "-":
            for kv in variant.split(","):
                k, v = kv.split("=")
                env[k] = v
        out = subprocess.run([sys.executable, __file__, "--run"], env=env, capture_output=True, text=True)
        print(f"{variant:30s} {out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:]}", flush=True)
    sys.exit(0)
import torch
import lms_b200
from oracle import reference_path as rp
B = int(os.environ.get("BENCH_B", "64"))
hf = rp.build_gpt2(os.environ.get("BENCH_ARCH", "small"), seed=0)
eng = lms_b200.B200GPT2LMHeadModel.from_hf(hf, dtype="bf16")
ids = rp.gpt2_prompt_ids(B, 512, seed=1)
stream = torch.cuda.ExternalStream(eng.stream)
eng.prefill(ids, max_new_tokens=256)
eng.step(10)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
eng.step(240)
e1.record(stream)
torch.cuda.synchronize()
out = eng.read_ids()
print(f"{e0.elapsed_time(e1) / 240 * 1e3:.1f} us per forward; checksum {int(out.sum())}")
