"""Same-box A/B of the dense channels-last MAC kernels at the in-step shape (two 512 x 16 x 16 f32 maps per image, B = 256).
Two timings per variant: `cold` = one launch after an L2 flush, median of N (what the step sees: launches 17 ms apart);
`stream` = 30 back-to-back launches rotating over 3 input sets (807 MB > L2), per launch (steady-state streaming rate).
CVPCE_B200_SO selects the library build; --tune name=value sets library knobs."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import ops, _lib

ap = argparse.ArgumentParser()
ap.add_argument("--tune", nargs="*", default=[])
ap.add_argument("--B", type=int, default=256)
ap.add_argument("--hw", type=int, default=16)
args = ap.parse_args()
for t in args.tune:
    k, v = t.split("=")
    _lib.tuning_set(k, int(v))
dev = torch.device("cuda", 0)
B, hw = args.B, args.hw
sets = [[torch.rand((B, 512, hw, hw), device=dev).contiguous(memory_format=torch.channels_last) for _ in range(2)] for _ in range(3)]
out = torch.empty((B, 1024), device=dev)
nbytes = 4.0 * 2 * B * 512 * hw * hw + 4.0 * 1024 * B
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for s in sets:
    ops.region_pool_embed(s, out=out)
torch.cuda.synchronize()
cold = []
for it in range(15):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.region_pool_embed(sets[it % 3], out=out); e1.record(); torch.cuda.synchronize()
    cold.append(e0.elapsed_time(e1))
cold = sorted(cold[3:])[len(cold[3:]) // 2]
stream = []
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(30):
        ops.region_pool_embed(sets[it % 3], out=out)
    e1.record(); torch.cuda.synchronize()
    stream.append(e0.elapsed_time(e1) / 30)
stream = sorted(stream)[2]
print(json.dumps({"so": os.path.basename(os.environ.get("CVPCE_B200_SO", "current")), "tune": args.tune, "B": B, "hw": hw,
                  "cold_us": round(cold * 1e3, 1), "cold_GBs": round(nbytes / cold / 1e6, 1),
                  "stream_us": round(stream * 1e3, 1), "stream_GBs": round(nbytes / stream / 1e6, 1)}))
