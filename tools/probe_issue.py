"""Issue-slot probe of packed FP32 (rm_probe_issue): prints ms and G-instr/s for the four mixes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from ray_marching_b200 import _lib

lib = _lib.lib()
sink = torch.zeros(4, device="cuda")
st = torch.cuda.current_stream().cuda_stream
blocks, iters = 148 * 8, 4096
names = {0: "128 FFMA", 1: "64 FFMA2", 2: "64 FFMA + 32 FMNMX + 16 MUFU", 3: "32 FFMA2 + 32 FMNMX + 16 MUFU"}
for kind in range(4):
    for _ in range(2):
        _lib.check(lib.rm_probe_issue(kind, blocks, iters, sink.data_ptr(), st), "probe")
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.rm_probe_issue(kind, blocks, iters, sink.data_ptr(), st), "probe")
        e1.record(); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    fma = blocks * 256 * iters * (128 if kind < 2 else 64)
    print(f"kind {kind} ({names[kind]} per thread-iteration): {best:.3f} ms, {fma * 2 / best / 1e9:.1f} TFLOP/s of FMA")
