"""Where does the bf16-mode mel error at the C3 shape (T = 600) come from?  Taps against the oracle, per row / per frame."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import masg_b200
from masg_b200 import synthetic as syn
from oracle import restate
from tests import helpers
L = int(os.environ.get("DBG_L", "200")); B = int(os.environ.get("DBG_B", "4"))
prec = os.environ.get("DBG_PREC", "bf16")
fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", prec)
b = syn.make_batch(64, L, seed=1236)
b = {k: (v[:B] if torch.is_tensor(v) else v) for k, v in b.items()}
g = torch.Generator().manual_seed(3)
emb = (torch.randn(64, 256, generator=g) * 0.5)[:B]
gb = helpers.to_gpu(b)
fs2._context().enable_taps(True)
out = fs2.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], L, accents=gb["accents"], speaker_emb=emb.cuda())
torch.cuda.synchronize()
T = out[0].shape[1]
ctx = fs2._context()
taps_g = {n: ctx.tap(n, s).cpu() for n, s in (("encoder_out", (B, L, 256)), ("lr_out", (B, T, 256)), ("decoder_out", (B, T, 256)))}
taps = {}
ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], L, emb, taps=taps)
for n in ("encoder_out", "lr_out", "decoder_out"):
    e = (taps_g[n] - taps[n]).abs()
    print(n, "max-abs %.3e" % e.max().item(), "mean %.3e" % e.mean().item(), "|ref| max %.2f" % taps[n].abs().max().item(),
          "per-row max", [round(x, 4) for x in e.flatten(1).max(1).values.tolist()])
for i, n in ((0, "mel"), (1, "postnet")):
    e = (out[i].cpu() - ref[i]).abs()
    print(n, "max-abs %.3e" % e.max().item(), "mean %.3e" % e.mean().item(), "|ref| max %.2f" % ref[i].abs().max().item(),
          "per-row max", [round(x, 4) for x in e.flatten(1).max(1).values.tolist()])
    fr = e.max(2).values
    for r in range(B):
        worst = torch.topk(fr[r], 5)
        print("   row", r, "worst frames", worst.indices.tolist(), [round(x, 4) for x in worst.values.tolist()])
e = (taps_g["decoder_out"] - taps["decoder_out"]).abs().max(2).values
print("decoder_out err by frame decile (row 0):", [round(e[0, i * T // 10:(i + 1) * T // 10].max().item(), 4) for i in range(10)])
