"""C2 latency (batch 1, 50 phonemes, host ids -> int16 audio on the host): p50 / p10 / p90 over 200 calls after 20 warm-ups.
MASG_PROBE_ONE=1: warm up, then ONE call inside cudaProfilerStart/Stop (for an ncu launch list)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from masg_b200 import synthetic as syn
B = int(os.environ.get("PROBE_B", "1")); L = int(os.environ.get("PROBE_L", "50"))
dev = torch.device("cuda", 0)
fs2, voc, synth = bench.build_models("bf16", dev)
b1 = syn.make_batch(B, L, seed=1235, accents_const=4)
emb1 = (torch.randn(B, 256, generator=torch.Generator().manual_seed(1)) * 0.5).pin_memory()
packed = os.environ.get("PROBE_PACKED", "0") == "1"
def call():
    if packed:
        return synth.synthesize_host_packed(b1["texts"], b1["accents"], b1["src_lens"], emb1, B * (3 * L + 8) * 256)
    return synth.synthesize_host(b1["texts"], b1["accents"], b1["src_lens"], emb1, 3 * L + 8)
for _ in range(20):
    call()
if os.environ.get("MASG_PROBE_ONE"):
    torch.cuda.synchronize(); torch.cuda.profiler.start(); call(); torch.cuda.synchronize(); torch.cuda.profiler.stop(); sys.exit(0)
lat = []
n0 = synth.ctx.launch_count()
for i in range(200):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); call(); lat.append((time.perf_counter() - t0) * 1e3)
lat.sort()
print("B=%d L=%d packed=%d PDL=%s GRAPH=%s: p10 %.3f p50 %.3f p90 %.3f ms, %d launches per call" % (
    B, L, packed, os.environ.get("MASG_PDL", "-"), os.environ.get("MASG_GRAPH", "-"), lat[20], lat[100], lat[180], (synth.ctx.launch_count() - n0) // 200))
