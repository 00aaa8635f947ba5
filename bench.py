#!/usr/bin/env python
"""Benchmark of the batched synthesis hot path (BASELINE.json: audio-sec/sec, FS2 + HiFi-GAN, bs64).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16|fp32]

One "step" = one pass of the hot path over one synthetic batch of config 3 (BASELINE.json configs[2]):
64 speakers (32 sampled from the TacoSpawn GMM prior + 32 from InterpolateGMM mid-attribute priors) x 200
phonemes -> FastSpeech2 (3 frames/phoneme, T = 600) -> HiFi-GAN -> int16 waveforms (445.8 s of audio).
N > 1 (launched by torch.distributed.run, one rank per GPU): every rank runs its own 64-utterance batch
(weak scaling, no data-path collective); the time is the max over ranks.

Printed JSON (rank 0, one line): `value` = audio-s/s with inputs resident in HBM, device-timed with CUDA
events; `e2e` = the same metric through the host-buffer C-ABI call (masg_synthesize_host: H2D of ids/lengths/
embeddings, D2H of int16 audio inside the timed region); `roofline` = the tcgen05 implicit-GEMM kernel's
achieved algorithmic TFLOP/s, measured with per-launch CUDA events on one extra identical step, against
MEASURED_PEAKS.json; `cpu_baseline` = the CPU oracle (oracle/restate.py, the reference's fp32 PyTorch
algorithm) on a bounded sample on this box's host cores.

--impl reference: the reference's CPU path (the oracle port; /root/reference itself cannot travel to the
GPU box) timed on all host threads, each step a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "audio-sec/sec (FS2+HiFi-GAN, bs64)"
UNIT = "audio-s/s"
B_C3, L_C3 = 64, 200
WORKLOAD = "C3: 64 sampled/interpolated speakers x 200 phonemes, FastSpeech2 + HiFi-GAN, 3 frames/phoneme (T=600), int16 audio out"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops": d.get("bf16_tflops_sustained", 1400.0), "hbm": d.get("hbm_gbs", 6650.0), "src": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback (B200_PROFILING.md)"}


def _ncu_traffic():
    """DRAM bytes (read + written) per k_conv_tc / k_respair_tc launch, from the newest committed `ncu --set full` capture of one C3 step
    (profiles/r*_ncu_full_conv_tc_*.csv, written by tools/ncu_summary.py).  A profiler number, reported beside the live ones."""
    import csv, glob
    # r<round>_ncu_<full|dram>_conv_tc_<tag>.csv: `--set full` captures and the cheaper DRAM / tensor-pipe metric captures; newest tag wins
    import re

    def order(f):
        m = re.match(r"r(\d+)_ncu_.*_conv_tc_(.*)\.csv", os.path.basename(f))
        return (int(m.group(1)), m.group(2)) if m else (0, "")
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_*_conv_tc_*.csv")), key=order)
    if not files:
        return None, None
    rows = list(csv.DictReader(open(files[-1])))
    if not rows:
        return None, None
    tot = sum((float(r["dram_rd_MB"]) + float(r["dram_wr_MB"])) * 1e6 for r in rows)
    return tot / len(rows), "%s: %d launches of one C3 step, %.1f GB in total" % (os.path.basename(files[-1]), len(rows), tot / 1e9)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons, mx = [], set(), None
        try:
            for line in open(self.path):
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1])); mx = float(c[2])
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))
        return out


def _dist_setup(n_gpus):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if n_gpus > 1 and world == 1:
        # plain `python bench.py --gpus N`: re-launch under torch.distributed.run, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n_gpus), "--master-addr", "127.0.0.1",
               "--master-port", os.environ.get("MASTER_PORT", "29531"), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return world, rank, local


# ----------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline / eager-GPU baseline: the UNMODIFIED reference (oracle/_ref staged by oracle/make_ref.py, or
# /root/reference in the build container) through its own public API, else the oracle port.  The only place bench.py
# executes oracle/.
# ----------------------------------------------------------------------------------------------------------
N_REF_SAMPLE = 8            # utterances of the C3 batch per reference step (55.7 s of audio; ~4 s of CPU work on 16 threads)


def c3_config(world):
    """The `config` object of the bench line — identical in both arms (the reference arm runs on OUR arm's config)."""
    return {"workload": WORKLOAD, "global_batch": B_C3 * world, "phonemes": L_C3, "frames_per_utt": 3 * L_C3, "parallelism": "dp%d" % world,
            "weights": "seeded random init (no checkpoint in the reference tree)",
            "l2": "256 MiB flush buffer written before every timed step; activations (GBs per step) far exceed the 126 MB L2"}


class ReferencePath:
    """FastSpeech2 + speaker prior + HiFi-GAN of the reference on `device` ("cpu" or "cuda"): the unmodified modules driven the
    way examples_gen.py:114-140 + utils/tools.py:264-270 drive them (synthesize_from_speaker_emb -> vocoder_infer), or the oracle
    port (oracle/restate.py) when no reference tree is present."""

    def __init__(self, device="cpu"):
        import torch
        import masg_b200  # noqa: F401
        from masg_b200 import synthetic as syn
        from oracle import ref_import, restate
        self.torch, self.syn, self.restate, self.device = torch, syn, restate, device
        self.sd = syn.make_fs2_state_dict(seed=1234, duration_mode="fixed3")
        hsd = syn.make_hifigan_state_dict(seed=4321)
        self.kind = "reference" if ref_import.available() else "port"
        self.source = ref_import.kind()
        if self.kind == "reference":
            self.ns = ref_import.load()
            self.pre, self.mod = ref_import.load_configs()
            fs2 = ref_import.build_fs2(self.sd, syn.speakers_json(), syn.STATS_JSON)
            # the mid-attribute prior is built on the CPU as the reference does (model/distributions.py mixes numpy / CPU tensors
            # into the components, so an InterpolateGMM of CUDA distributions cannot be sampled); it is sampled on the CPU too
            metas = torch.tensor(syn.META_COMBOS, dtype=torch.float32)
            with torch.no_grad():
                d = [fs2.speaker_distribution(metas[i:i + 1]) for i in (0, 3)]
            t0 = time.perf_counter()
            self.interp = self.ns.InterpolateGMM(d[0], d[1])          # scipy sqrtm/inv on 256x256 matrices
            self.interp_build_s = time.perf_counter() - t0
            self.fs2 = fs2.to(device)
            self.voc = ref_import.build_hifigan(hsd).to(device)
        else:
            self.hsd = restate.fold_weight_norm(hsd)

    def embeddings(self, n, seed):
        """n speaker embeddings the way config 3 draws them: half from speaker_gen, half from an InterpolateGMM prior."""
        torch = self.torch
        torch.manual_seed(seed)
        if self.kind != "reference":
            return torch.randn(n, 256) * 0.5
        meta = torch.tensor([self.syn.META_COMBOS[i % 4] for i in range(n - n // 2)], dtype=torch.float32, device=self.device)
        with torch.no_grad():
            embs = [self.fs2.speaker_gen(meta)] + [self.interp.sample().to(self.device) for _ in range(n // 2)]
        return torch.cat(embs, 0).float()

    def run(self, b, emb, vocoder=True):
        """One pass: ids -> mel (-> int16 waveforms cut to length).  Returns audio seconds produced."""
        torch, syn = self.torch, self.syn
        if self.kind == "reference":
            dev = self.device
            with torch.no_grad():
                out = self.fs2.synthesize_from_speaker_emb(None, b["texts"].to(dev), b["src_lens"].to(dev), b["max_src_len"],
                                                           accents=b["accents"].to(dev), speaker_emb=emb.to(dev))
                if vocoder:
                    lengths = out[9] * self.pre["stft"]["hop_length"]
                    self.ns.vocoder_infer(out[1].transpose(1, 2), self.voc, self.mod, self.pre, lengths=lengths)
                elif dev != "cpu":
                    out[1].cpu()
        else:
            out = self.restate.fs2_forward(self.sd, b["texts"], b["accents"], b["src_lens"], b["max_src_len"], emb)
            if vocoder:
                self.restate.wav_to_int16(self.restate.hifigan_forward(self.hsd, out[1].transpose(1, 2)).squeeze(1))
        return float(out[9].sum()) * syn.HOP / syn.SAMPLING_RATE


def _p50(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores (all threads), each step
    a bounded sample (N_REF_SAMPLE of the 64 utterances) of our arm's workload.  CUDA is hidden from this process so the
    reference's module-level `device` globals (model/modules.py:14, utils/tools.py:15) resolve to the CPU."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
    cores = torch.get_num_threads()
    ref = ReferencePath("cpu")
    syn = ref.syn
    n_utts = max(1, min(B_C3, args.ref_sample))
    ref.run(syn.make_batch(1, 16, seed=3), ref.embeddings(1, 0))               # tiny warm-up (thread pools, oneDNN primitives)
    b = syn.make_batch(B_C3, L_C3, seed=1234 + 2)
    b = {k: (v[:n_utts] if torch.is_tensor(v) else v) for k, v in b.items()}
    times, audio = [], 0.0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        emb = ref.embeddings(n_utts, 77 + i)
        audio = ref.run(b, emb)
        times.append(time.perf_counter() - t0)
    timed = times[args.warmup:]
    sec = sum(timed) / len(timed)
    val = audio / sec
    # BASELINE.json config 1 (B = 1, L = 50, mel only, CPU) and config 2 (+ HiFi-GAN) latencies on the same cores
    b1 = syn.make_batch(1, 50, seed=1235, accents_const=4)
    e1 = ref.embeddings(1, 5)
    lat1, lat2 = [], []
    for i in range(1 + args.latency_reps):
        t0 = time.perf_counter(); ref.run(b1, e1, vocoder=False); lat1.append((time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter(); ref.run(b1, e1, vocoder=True); lat2.append((time.perf_counter() - t0) * 1e3)
    what = ("the unmodified reference modules (%s copy), synthesize_from_speaker_emb + vocoder_infer" % ref.source) if ref.kind == "reference" \
        else "oracle/restate.py (CPU port of the reference's algorithm; no reference tree on this box)"
    sample = "%d of the 64 utterances of the C3 batch per step (200 phonemes, 600 frames each; %.1f s of audio), %s, fp32 PyTorch CPU ops, %d threads" % (
        n_utts, audio, what, cores)
    cpu = {"value": val, "unit": UNIT, "cores": cores, "kind": ref.kind, "sample": sample,
           "c1_latency_p50_ms": {"value": _p50(lat1[1:]), "config": "C1: batch 1, 50 phonemes, mel only, CPU, p50 of %d runs after 1 warm-up" % args.latency_reps},
           "c2_latency_p50_ms": {"value": _p50(lat2[1:]), "config": "C2 on the CPU: batch 1, 50 phonemes, FastSpeech2 + HiFi-GAN -> int16, p50 of %d runs" % args.latency_reps}}
    if ref.kind == "reference":
        cpu["interpolate_gmm_build_s"] = ref.interp_build_s
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": c3_config(args.gpus),
            "cpu_baseline": cpu,
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def cpu_baseline_subprocess():
    """cpu_baseline leg of our arm: the reference arm above in a child process with CUDA hidden (one warm-up + one timed step)."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2", "--warmup", "1", "--latency-reps", "5"],
                           capture_output=True, text=True, env=env, timeout=900)
        for ln in reversed(r.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
        return {"error": (r.stderr or r.stdout)[-400:]}
    except Exception as e:                                                   # noqa: BLE001 - a missing baseline must not lose the bench line
        return {"error": repr(e)[:400]}


def gpu_eager_baseline(dev, steps=3):
    """The unmodified reference modules on the B200 through PyTorch eager (fp32; TF32 off and on), full C3 batch — SURVEY §8(d)
    'reference GPU path': not a published number, the honest thing to beat.  Same call sequence as the CPU arm."""
    import torch
    from oracle import ref_import
    if not ref_import.available():
        return {"unavailable": "no reference tree on this box (oracle/_ref not staged)"}
    out = {}
    try:
        ref = ReferencePath(str(dev))
        b = ref.syn.make_batch(B_C3, L_C3, seed=1234 + 2)
        audio = 0.0
        old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        for name, tf32 in (("fp32_tf32_off", False), ("fp32_tf32_on", True)):
            torch.backends.cudnn.allow_tf32 = tf32
            torch.backends.cuda.matmul.allow_tf32 = tf32
            ts = []
            for i in range(1 + steps):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                audio = ref.run(b, ref.embeddings(B_C3, 77 + i))              # ends with the int16 audio on the host (vocoder_infer .cpu())
                torch.cuda.synchronize()
                ts.append(time.perf_counter() - t0)
            sec = sum(ts[1:]) / steps
            out[name] = {"value": audio / sec, "unit": UNIT, "ms_per_step": sec * 1e3}
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
        out["config"] = "the unmodified reference modules on cuda:0 (PyTorch eager), full C3 batch of 64, host ids -> int16 audio on the host, %d timed steps after 1 warm-up" % steps
        del ref
        torch.cuda.empty_cache()
    except Exception as e:                                                   # noqa: BLE001
        out["error"] = repr(e)[:400]
    return out


# ----------------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------------
def build_models(precision, device):
    import torch
    import masg_b200  # noqa: F401
    from masg_b200 import synthetic as syn
    from masg_b200.fastspeech2 import FastSpeech2
    from masg_b200.hifigan import Generator, AttrDict
    from masg_b200.synth import Synthesizer
    d = tempfile.mkdtemp(prefix="masg_cfg_")
    json.dump(syn.speakers_json(), open(os.path.join(d, "speakers.json"), "w"))
    json.dump(syn.STATS_JSON, open(os.path.join(d, "stats.json"), "w"))
    fs2 = FastSpeech2(syn.PREPROCESS_CONFIG, syn.MODEL_CONFIG, d, precision=precision)
    fs2.load_state_dict(syn.make_fs2_state_dict(seed=1234, duration_mode="fixed3"))
    fs2.eval().to(device)
    voc = Generator(AttrDict(syn.HIFIGAN_CONFIG), precision=precision)
    voc.load_state_dict(syn.make_hifigan_state_dict(seed=4321))
    voc.eval()
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        voc.remove_weight_norm()
    voc.to(device)
    return fs2, voc, Synthesizer(fs2, voc, precision)


class SpeakerSampler:
    """Config-3 speaker set, one batch per call: 32 embeddings from speaker_gen (8 per metadata combo) + 32 from the
    InterpolateGMM mid-attribute priors over the 6 unordered combo pairs x t in {.25,.5,.75}.
    The priors are built once (as the reference loads a pickled distribution once); sampling runs every step."""

    def __init__(self, fs2, n_sampled, n_interp, seed):
        import torch
        from masg_b200 import synthetic as syn
        from masg_b200.distributions import InterpolateGMM
        self.fs2, self.ctx, self.seed, self.calls = fs2, fs2._context(), seed, 0
        dev = self.ctx.device
        self.meta = torch.tensor([syn.META_COMBOS[i % 4] for i in range(n_sampled)], dtype=torch.float32, device=dev)
        combos = torch.tensor(syn.META_COMBOS, dtype=torch.float32, device=dev)
        dists = [fs2.speaker_distribution(combos[i:i + 1]) for i in range(4)]
        pairs = [(a, b) for a in range(4) for b in range(a + 1, 4)]
        rates = (0.25, 0.5, 0.75)
        cache, pis, mus, scs = {}, [], [], []
        for i in range(n_interp):
            a, b = pairs[i % len(pairs)]
            t = rates[(i // len(pairs)) % len(rates)]
            if (a, b) not in cache:
                cache[(a, b)] = InterpolateGMM(dists[a], dists[b])
            ig = cache[(a, b)]
            ig.interpolate_rate(t)
            comp = ig.component_distribution.base_dist
            pis.append(ig.mixture_distribution.probs.reshape(1, -1).clone())
            mus.append(comp.loc.reshape(1, -1, 256).clone())
            scs.append(comp.scale.reshape(1, -1, 256).clone())
        self.n_interp = n_interp
        if n_interp:
            self.pi, self.mu, self.sc = torch.cat(pis).contiguous(), torch.cat(mus).contiguous(), torch.cat(scs).contiguous()

    def sample(self):
        import torch
        from masg_b200.fastspeech2 import gmm_sample
        self.calls += 1
        embs = [self.fs2.speaker_gen(self.meta, seed=self.seed + 2 * self.calls)]
        if self.n_interp:
            embs.append(gmm_sample(self.ctx, self.pi, self.mu, self.sc, self.n_interp, seed=self.seed + 2 * self.calls + 1))
        return torch.cat(embs, 0)


def run_ours(args):
    import torch
    world, rank, local = _dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); the synthesis path has no CPU fallback")
    import masg_b200  # noqa: F401
    from masg_b200 import synthetic as syn
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    fs2, voc, synth = build_models(args.precision, dev)
    ctx = synth.ctx
    B, L = B_C3, L_C3
    batch = syn.make_batch(B, L, seed=1234 + 2 + 1000 * rank)
    texts, accents, lens = batch["texts"].to(dev), batch["accents"].to(dev), batch["src_lens"].to(dev)
    sampler = SpeakerSampler(fs2, B // 2, B - B // 2, seed=77 + 100000 * rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)            # > 126 MB L2
    audio_per_step = B * 3 * L * syn.HOP / syn.SAMPLING_RATE

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    # --gather (N > 1): the optional NCCL all-gather of every rank's int16 audio, issued on a side stream after each
    # step and waited for one step later, so the transfer of batch n overlaps the kernels of batch n + 1
    gather = None
    if args.gather and world > 1:
        from masg_b200.synth import WaveformGather
        gather = WaveformGather()
        g_off = torch.arange(B + 1, dtype=torch.int64, device=dev) * (3 * L * syn.HOP)
    pending = [None]

    def step_device():
        emb = sampler.sample()
        w16, mel_lens, _ = synth.synthesize(texts, accents, lens, emb)
        if gather is not None:
            if pending[0] is not None:
                pending[0].wait()
            pending[0] = gather.start(w16.view(-1), g_off, n_max=w16.numel(), b_max=B)
        return w16, mel_lens

    # ---- device-resident throughput -----------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    if args.profile_range:
        # ncu --profile-from-start off: exactly one step between cudaProfilerStart/Stop, nothing else measured
        torch.cuda.profiler.start()
        step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        if rank == 0:
            print(json.dumps({"profile_range": "one C3 step between cudaProfilerStart/Stop", "precision": args.precision}))
        return 0
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ctx.launch_count(reset=True)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for i in range(args.steps):
        flush.fill_(i & 0xFF)                                                # L2 flush between timed iterations (untimed)
        ev[i][0].record()
        w16, mel_lens = step_device()
        ev[i][1].record()
    if pending[0] is not None:
        got = pending[0].wait()                                              # the last step's gather is inside the barrier-to-barrier span
        pending[0] = None
        assert len(got) == world and all(int(o[-1]) == B * 3 * L * syn.HOP for _, o in got)
    barrier()
    launches = ctx.launch_count()
    ms_dev = sum(a.elapsed_time(b) for a, b in ev)
    ms_dev = max_over_ranks(ms_dev)
    gather_was_on, gather = gather is not None, None                         # the e2e / roofline legs below run without it
    assert int(mel_lens.sum().item()) == B * 3 * L, "unexpected mel lengths (benchmark shape is 3 frames/phoneme)"
    value = world * audio_per_step * args.steps / (ms_dev / 1e3)

    # ---- end to end through the host-buffer C-ABI call ------------------------------------------------
    t_cap = 3 * L + 8
    pin_emb = torch.empty(B, 256, dtype=torch.float32).pin_memory()
    h_texts, h_acc, h_lens = batch["texts"], batch["accents"], batch["src_lens"]

    def step_host():
        emb = sampler.sample()
        pin_emb.copy_(emb, non_blocking=True)
        torch.cuda.current_stream().synchronize()                             # embeddings reach the host (examples_gen.py:268 does .cpu())
        return synth.synthesize_host(h_texts, h_acc, h_lens, pin_emb, t_cap)

    for _ in range(3):
        step_host()
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        h_w, h_m, T, h2d, d2h = step_host()
    e1.record()
    barrier()
    # the call blocks the host until the audio is back, so host wall-clock between the two barriers is the end-to-end time
    # (the CUDA-event span over the same region is reported beside it)
    ms_e2e_sync = max_over_ranks(max((time.perf_counter() - t0) * 1e3, e0.elapsed_time(e1)))
    d2h += B * 256 * 4                # sampled embeddings copied to the pinned host buffer (their H2D is counted by synthesize_host)

    # the same through the deferred-copy call (masg_synthesize_host_packed_async): every step still copies its inputs H2D and its
    # audio D2H inside the timed region and every result is waited for, but the D2H of step n overlaps phase A of step n + 1 —
    # what a generation job / service does with a stream of batches.  This is the end-to-end number; the blocking form is beside it.
    cap = B * t_cap * syn.HOP

    def step_host_async():
        emb = sampler.sample()
        pin_emb.copy_(emb, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return synth.synthesize_host_packed_async(h_texts, h_acc, h_lens, pin_emb, cap)

    prev = None
    for _ in range(3):
        cur = step_host_async()
        if prev is not None:
            prev.wait()
        prev = cur
    prev.wait()
    barrier()
    t0 = time.perf_counter()
    prev = None
    for i in range(args.steps):
        cur = step_host_async()
        if prev is not None:
            h_wp, h_o, h_m2, T2, h2d2, d2h2 = prev.wait()
        prev = cur
    h_wp, h_o, h_m2, T2, h2d2, d2h2 = prev.wait()
    torch.cuda.synchronize()
    ms_e2e = max_over_ranks((time.perf_counter() - t0) * 1e3)
    barrier()
    # (every step draws new speakers, so the audio differs from step to step; bit-equality of the two calls on the same inputs is
    # tests/test_gpu_generation.py::test_deferred_copy_calls_match_blocking_calls)
    assert int(h_o[B]) == B * 3 * L * syn.HOP and int(h_wp[: int(h_o[B])].abs().max()) > 0, "deferred-copy call returned no audio"
    e2e_value = world * audio_per_step * args.steps / (ms_e2e / 1e3)
    d2h = d2h2 + B * 256 * 4
    clk = clocks.stop() if rank == 0 else None

    # ---- roofline of the dominant kernel: one extra identical step with per-launch CUDA events ---------
    barrier()
    ctx.profile(True)
    step_device()
    k_ms, k_flops, k_n = ctx.profile_read()
    ctx.profile(False)
    e0.record(); step_device(); e1.record(); torch.cuda.synchronize()
    step_ms_plain = e0.elapsed_time(e1)
    pk = _peaks()
    traffic, traffic_src = _ncu_traffic()

    # ---- p50 per-utterance latency (config 2: batch 1, 50 phonemes, host ids in -> int16 audio on the host) ----
    b1 = syn.make_batch(1, 50, seed=1235, accents_const=4)
    emb1 = sampler.sample()[:1].cpu().pin_memory()
    lat = []
    for i in range(220):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        synth.synthesize_host(b1["texts"], b1["accents"], b1["src_lens"], emb1, 160)
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = sorted(lat[20:])

    if rank != 0:
        return 0
    # ---- CPU baseline (oracle port) on this box's host cores, bounded sample, N = 1 only ----------------
    cpu = eager = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_subprocess()
        eager = gpu_eager_baseline(dev)
    ach = k_flops / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if args.precision.startswith("bf16") else "f32", "data": "synthetic",
        "config": c3_config(world),
        "precision": args.precision + (" (HiFi-GAN bf16; FastSpeech2 encoder + variance predictors split-bf16 hi/lo, 3 MMAs per K step; decoder, mel_linear, PostNet fp16 operands; fp32 accumulation everywhere)" if args.precision == "bf16" else ""),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / args.steps,
                "call": "masg_synthesize_host_packed_async + masg_synthesize_host_wait (host buffers; the D2H of a step overlaps the next step's phase A; every result waited for)",
                "blocking_call": {"value": world * audio_per_step * args.steps / (ms_e2e_sync / 1e3), "ms_per_step": ms_e2e_sync / args.steps,
                                  "call": "masg_synthesize_host (returns with the audio on the host)"}},
        "gpu_launches": int(launches),
        "gather": ({"collective": "NCCL all_gather of every rank's int16 audio (padded-to-max exchange + counts), side stream, overlapped with the next step",
                    "bytes_per_rank_per_step": B * 3 * L * syn.HOP * 2} if gather_was_on else None),
        "clocks": clk,
        "roofline": {"bound": "tensor", "kernel": "k_conv_tc + k_respair_tc + k_respair2_tc (tcgen05 implicit-GEMM conv1d / linear / convT and the fused ResBlock pairs, one-SM and CTA-pair)", "achieved": ach, "peak": pk["tflops"],
                     "unit": "TFLOP/s", "frac": ach / pk["tflops"], "traffic": traffic, "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
                     "traffic_source": traffic_src, "peak_source": pk["src"],
                     "launches_per_step": k_n, "kernel_ms_per_step": k_ms, "step_ms": step_ms_plain,
                     "share_of_step": k_ms / step_ms_plain if step_ms_plain else None,
                     "how": "per-launch CUDA events on the launching stream over one extra identical step; algorithmic FLOPs = 2*rows*K*N per conv (both convs of a fused pair)"},
        "cpu_baseline": cpu,
        "gpu_eager_baseline": eager,
        "latency_p50_ms": {"value": lat[len(lat) // 2], "config": "C2: batch 1, 50 phonemes, host ids -> int16 audio on host (masg_synthesize_host), 200 timed runs after 20 warm-ups"},
    }
    print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0

# ----------------------------------------------------------------------------------------------------------
# the other shard-able configurations of BASELINE.json: C4 and C5, strong scaling
# ----------------------------------------------------------------------------------------------------------
def run_sharded(args):
    """--workload c4: 4096 utterances, L = default_rng(0).integers(20, 301, 4096), sorted into length-bucketed batches under a
    padded-frame budget, batches assigned to the ranks longest-processing-time-first (synth.plan_batches / assign_batches);
    --workload c5: 16 utterances x 1000 phonemes (T = 3000), contiguous split of the batch over the ranks.
    One step = one pass over the whole job (total work fixed -> "strong"); value = job audio-seconds / max-over-ranks time."""
    import numpy as np
    import torch
    world, rank, local = _dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); the synthesis path has no CPU fallback")
    import masg_b200  # noqa: F401
    from masg_b200 import synthetic as syn
    from masg_b200.synth import plan_batches, assign_batches, shard_batch
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    fs2, voc, synth = build_models(args.precision, dev)
    ctx = synth.ctx
    if args.workload == "c4":
        lengths = syn.c4_lengths(4096)
        batches = plan_batches(lengths, frame_budget=args.frame_budget, max_batch=args.max_batch)
        mine = [batches[i] for i in assign_batches(batches, lengths, world)[rank]]
        name = ("C4: 4096 utterances, 20-300 phonemes (numpy default_rng(0)), length-bucketed batches of <= %d utterances / <= %d padded "
                "frames, LPT assignment to ranks, 3 frames/phoneme, int16 audio out" % (args.max_batch, args.frame_budget))
        seed0 = 1234 + 3
    else:
        lengths = np.full(16, 1000)
        lo, hi = shard_batch(16, rank, world)
        mine = [np.arange(lo, hi)] if hi > lo else []
        batches = [np.arange(16)]
        name = "C5: 16 utterances x 1000 phonemes (T = 3000), contiguous batch split over the ranks, int16 audio out"
        seed0 = 1234 + 4
    job_audio = float(lengths.sum()) * 3 * syn.HOP / syn.SAMPLING_RATE
    nb = max(len(b) for b in batches)
    sampler = SpeakerSampler(fs2, nb // 2, nb - nb // 2, seed=77 + 100000 * rank)
    work = []
    for bi, idx in enumerate(mine):
        b = syn.make_batch(len(idx), 0, seed=seed0 + 7919 * int(idx[0]), lens=lengths[idx])
        work.append({"n": len(idx), "host": b, "dev": tuple(b[k].to(dev) for k in ("texts", "accents", "src_lens")),
                     "frames": int(lengths[idx].sum()) * 3, "t_cap": 3 * int(lengths[idx].max()) + 8})
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def pass_device():
        frames = 0
        for w in work:
            emb = sampler.sample()[:w["n"]]
            w16, mel_lens, _ = synth.synthesize(*w["dev"], emb)
            frames += int(mel_lens.sum().item()) if args.check_lens else w["frames"]
        return frames

    pin_emb = torch.empty(nb, 256, dtype=torch.float32).pin_memory()

    def pass_host():
        h2d = d2h = 0
        for w in work:
            emb = sampler.sample()[:w["n"]]
            pin_emb[:w["n"]].copy_(emb, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            b = w["host"]
            _, _, T, a, c = synth.synthesize_host(b["texts"], b["accents"], b["src_lens"], pin_emb[:w["n"]], w["t_cap"])
            h2d += a; d2h += c + w["n"] * 256 * 4
        return h2d, d2h

    args.check_lens = True
    my_frames = pass_device()                                               # warm-up pass 1 also checks the 3-frames-per-phoneme lengths
    assert my_frames == sum(w["frames"] for w in work), "unexpected mel lengths"
    args.check_lens = False
    for _ in range(max(args.warmup, 3) - 1):
        pass_device()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ctx.launch_count(reset=True)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for i in range(args.steps):
        flush.fill_(i & 0xFF)
        ev[i][0].record()
        pass_device()
        ev[i][1].record()
    barrier()
    launches = ctx.launch_count()
    ms_dev = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev))
    value = job_audio * args.steps / (ms_dev / 1e3)
    pass_host()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        h2d, d2h = pass_host()
    torch.cuda.synchronize()
    ms_e2e = max_over_ranks((time.perf_counter() - t0) * 1e3)
    barrier()
    clk = clocks.stop() if rank == 0 else None
    ctx.profile(True)
    pass_device()
    k_ms, k_flops, k_n = ctx.profile_read()
    ctx.profile(False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); pass_device(); e1.record(); torch.cuda.synchronize()
    pass_ms = e0.elapsed_time(e1)
    if rank != 0:
        return 0
    pk = _peaks()
    ach = k_flops / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": name, "utterances": int(len(lengths)), "batches": len(batches), "batches_rank0": len(work), "job_audio_s": job_audio,
                   "parallelism": "dp%d" % world, "precision": args.precision, "weights": "seeded random init",
                   "l2": "256 MiB flush buffer written before every timed pass"},
        "e2e": {"value": job_audio * args.steps / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": ms_e2e / args.steps, "note": "rank 0's bytes; one masg_synthesize_host call per batch"},
        "gpu_launches": int(launches), "clocks": clk,
        "roofline": {"bound": "tensor", "kernel": "k_conv_tc + k_respair_tc + k_respair2_tc", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": ach / pk["tflops"],
                     "traffic": None, "peak_source": pk["src"], "launches_per_step": k_n, "kernel_ms_per_step": k_ms, "step_ms": pass_ms,
                     "share_of_step": k_ms / pass_ms if pass_ms else None,
                     "how": "per-launch CUDA events over one extra pass of rank 0's share; FLOPs counted over the padded rows each launch processes"},
        "cpu_baseline": None,
    }
    print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32", "bf16x3"],
                    help="bf16 = the benchmarked mode; fp32 = CUDA-core strict parity; bf16x3 = 1e-4 mel parity on the tensor cores (split-bf16 through all of FastSpeech2)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline and gpu_eager_baseline legs")
    ap.add_argument("--ref-sample", type=int, default=N_REF_SAMPLE, help="reference arm: utterances of the C3 batch per step")
    ap.add_argument("--latency-reps", type=int, default=5, help="reference arm: timed runs of the C1 / C2 CPU latency")
    ap.add_argument("--workload", default="c3", choices=["c3", "c4", "c5"], help="c3 = the headline configuration (default); c4 / c5 = the sharded configurations of BASELINE.json")
    ap.add_argument("--frame-budget", type=int, default=40000, help="c4: padded frames per batch")
    ap.add_argument("--max-batch", type=int, default=64, help="c4: utterances per batch")
    ap.add_argument("--gather", action="store_true", help="N > 1: all-gather every rank's int16 audio over NCCL each step (optional collective of SURVEY 8(e))")
    ap.add_argument("--profile-range", action="store_true", help="warm up, then run ONE step inside cudaProfilerStart/Stop and exit (for ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        os.environ["CUDA_VISIBLE_DEVICES"] = ""        # the CPU arm: the reference's `device` globals must resolve to the CPU
        if int(os.environ.get("RANK", "0")) != 0:
            return 0                                   # under torchrun rank 0 alone runs it
    import __graft_entry__
    __graft_entry__.build_if_needed()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload != "c3":
        return run_sharded(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
