"""Benchmark of the Tacotron2 hot path (BASELINE.json metric: mel frames/sec, AR decode).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): vanilla-ljspeech-stop autoregressive inference,
batch 64 x 200 characters, 1000 mel frames per utterance (the random-init stop gate is biased so it
never fires: the full 1000 frames are decoded, nothing is skipped), synthetic seeded inputs and weights.
A "step" = one whole Tacotron2.forward: encoder -> memory projection -> 1000-frame persistent decoder
kernel -> postnet -> residual/masking.  N > 1: one process per GPU, each with its own utterance shard
(weak scaling, no data-path collective); timing = max over ranks of device time.
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tacotron2_b200 import synth  # noqa: E402

B, L, T = 64, 200, 1000
METRIC, UNIT = "mel_frames_per_sec_ar_decode", "frames/s"
# dram__bytes_read.sum + dram__bytes_write.sum of the persistent kernels per frame, from the committed `ncu --set full` captures of the
# CURRENT kernels: profiles/ncu_traffic.json is rewritten by tools/evidence.sh together with the capture it names (the captures are
# 200-frame launches of the same batches; traffic is linear in the frame count).  null when no capture of the current kernel exists.
def ncu_traffic(kernel):
    try:
        e = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kernel]
        return float(e["dram_bytes_per_frame"]), e["source"]
    except Exception:  # noqa: BLE001
        return None, "no ncu capture of the current kernel committed"


WORKLOAD = "C2: vanilla-ljspeech-stop AR inference, batch 64 x 200 chars x 1000 frames, fp32, stop gate evaluated (never fires)"
# kernels launched by one Tacotron2.forward (AR): embed 1, encoder conv 3, BiLSTM input proj 1, BiLSTM 1, memory split 1,
# att_encoder 1, persistent decoder (decoder_ar_kernel) 1, abort-poison guard 1, mel split 1, postnet conv 5, finalize 1  (memsets / memcpys not counted)
LAUNCHES_PER_STEP = 17


def algorithmic_bytes_per_frame_step(a=synth.VANILLA, b=B, l=L):
    """bytes one decoder step must move for the whole batch, fp32, LSTM weights
    not on-chip: memory + processed memory read once, attention-weight traffic, recurrent state, weights."""
    ef, h, s = a.encoded_full_dim, a.att_rnn_dim, 4
    act = b * l * (ef + a.att_dim) * s + 5 * b * l * s + b * (8 * h + ef + a.num_mels + a.prenet_dim) * s
    lstm = 4 * h * (a.prenet_dim + ef + h) + 4 * h * (h + ef + a.controls_dim + h)
    small = a.att_dim * h + (a.num_mels + 1) * (h + ef) + a.prenet_dim * (a.num_mels + a.prenet_dim) + 32 * 62 + a.att_dim * 33
    return act + (lstm + small) * s


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index, self.begin = [], None, index, 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def mark(self):
        """Start of the timed region: only later samples count (nvidia-smi needs a second or more to deliver its first row on an
        8-GPU box, so the process is started before the warm-up and bench.py keeps the GPU loaded until it reports)."""
        self.begin = len(self.rows)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        if len(self.rows) > self.begin:
            self.rows = self.rows[self.begin:]
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def _reference_available():
    from oracle import ref_model as R   # checker / baseline infrastructure: never on the product path
    return R.available()


def reference_step_cpu(frames, threads):
    """The reference's own CPU path on the workload: the UNMODIFIED model/tacotron2.py staged under oracle/_ref (by
    oracle/stage_ref.py, where /root/reference is readable) -- Tacotron2.forward(teacher_forcing=False, max_len_override=frames) in
    eval mode on all host threads, its own AlwaysDropout draws, its own per-frame stop check.  When oracle/_ref is absent the oracle
    port (oracle/tacotron2_oracle.py) stands in.  Returns (frames/s, seconds, kind)."""
    torch.set_num_threads(threads)
    sd = synth.synth_state_dict(synth.VANILLA, seed=0, gate_shift=5.0)
    b = synth.synth_batch(synth.VANILLA, B, L, 8, seed=1234)
    if _reference_available():
        from oracle import ref_model as R
        m = R.reference_tacotron2(synth.VANILLA, sd).eval()
        t0 = time.perf_counter()
        with torch.no_grad():
            out = m(b["chars_idx"], b["chars_idx_len"], False, max_len_override=frames)
        dt = time.perf_counter() - t0
        kind = "reference"
    else:
        from oracle import tacotron2_oracle as O
        masks = synth.synth_masks(synth.VANILLA, B, L, frames, seed=7)
        t0 = time.perf_counter()
        with torch.no_grad():
            out = O.tacotron2_forward(sd, synth.VANILLA, b["chars_idx"], b["chars_idx_len"], False, max_len_override=frames, masks=masks)
        dt = time.perf_counter() - t0
        kind = "port"
    assert out[0].shape[1] == frames, "frames were skipped"
    return B * frames / dt, dt, kind


def reference_cuda_eager(dev):
    """The comparator BASELINE.json's north_star names: the reference's own PyTorch-CUDA decoder step at batch 64 -- the UNMODIFIED
    model/tacotron2.py (oracle/_ref) moved to this GPU, eval mode, Tacotron2.forward(teacher_forcing=False) with its per-frame
    `done.all()` host sync, under the reference's own settings: float32_matmul_precision 'high' (run/train.py:170) in fp32, and
    under 16-mixed autocast (run/train.py:239).  us per decoder frame = (t[400 frames] - t[100 frames]) / 300, so the encoder and
    postnet of the call cancel.  A reported baseline: none of this repo's kernels runs here."""
    if not _reference_available():
        return {"unavailable": "oracle/_ref not staged (build() stages it where /root/reference is readable)"}
    from oracle import ref_model as R
    prev = torch.get_float32_matmul_precision()
    torch.set_float32_matmul_precision("high")
    out = {"what": f"UNMODIFIED reference Tacotron2.forward (oracle/_ref/model/tacotron2.py:276-325) in eval mode on this GPU, B={B} L={L}, "
                   "teacher_forcing=False, per-frame stop check with host sync; us/frame from the 400- minus the 100-frame call"}
    try:
        sd = synth.synth_state_dict(synth.VANILLA, seed=0, gate_shift=5.0)
        m = R.reference_tacotron2(synth.VANILLA, sd, device=dev).eval()
        b = synth.synth_batch(synth.VANILLA, B, L, 8, seed=1234)
        c, n = b["chars_idx"].to(dev), b["chars_idx_len"].to(dev)

        def run(frames):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            o = m(c, n, False, max_len_override=frames)
            e1.record()
            e1.synchronize()
            assert o[0].shape[1] == frames
            return e0.elapsed_time(e1)

        for key, ctx in (("fp32_high", None), ("amp_16_mixed", torch.autocast("cuda", dtype=torch.float16))):
            with torch.no_grad():
                if ctx is not None:
                    ctx.__enter__()
                try:
                    run(20)
                    t_short, t_long = run(100), run(400)
                finally:
                    if ctx is not None:
                        ctx.__exit__(None, None, None)
            us = 1e3 * (t_long - t_short) / 300
            out[key] = {"us_per_decoder_step": us, "frames_per_s": B * 1e6 / us, "ms_400_frames": t_long, "ms_100_frames": t_short}
        out["us_per_decoder_step"] = out["fp32_high"]["us_per_decoder_step"]
        del m
    finally:
        torch.set_float32_matmul_precision(prev)
        torch.cuda.empty_cache()
    return out


def reference_cuda_train(dev, steps=2):
    """Training half of the comparator: the UNMODIFIED reference's training_step arithmetic (Tacotron2.forward in train() mode with
    teacher forcing, the three losses of model/tts_model.py:197-201, loss.backward()) at C3 on this GPU, eager PyTorch, in fp32 with
    matmul precision 'high' and under 16-mixed autocast + GradScaler-free backward (run/train.py:170,239).  No optimizer step."""
    if not _reference_available():
        return {"unavailable": "oracle/_ref not staged"}
    from oracle import ref_model as R
    prev = torch.get_float32_matmul_precision()
    torch.set_float32_matmul_precision("high")
    out = {"what": f"UNMODIFIED reference Tacotron2 (oracle/_ref) train() forward + loss + backward on this GPU, C3: B={TRAIN_B} L={TRAIN_L} "
                   f"T={TRAIN_T}, speaker tokens; eager PyTorch"}
    try:
        a = synth.HIFI
        m = R.reference_tacotron2(a, synth.synth_state_dict(a, seed=0), device=dev).train()
        b = {k: v.to(dev) for k, v in synth.synth_batch(a, TRAIN_B, TRAIN_L, TRAIN_T, seed=1234, ragged=False).items()}

        def step(amp):
            m.zero_grad(set_to_none=True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            with torch.autocast("cuda", dtype=torch.float16, enabled=amp):
                mels, post, gates, _ = m(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
                                         speaker_id=b["speaker_id"])
                loss = R.reference_loss(mels.float(), post.float(), gates.float(), b["mel_spectrogram"], b["gate"])
            (loss * (1024.0 if amp else 1.0)).backward()   # static loss scale stands in for the GradScaler
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1)

        for key, amp in (("fp32_high", False), ("amp_16_mixed", True)):
            step(amp)
            ms = min(step(amp) for _ in range(steps))
            out[key] = {"ms_per_step": ms, "frames_per_s": TRAIN_B * TRAIN_T / (ms * 1e-3)}
        del m
    finally:
        torch.set_float32_matmul_precision(prev)
        torch.cuda.empty_cache()
    return out


def reference_c1_cpu(threads, iters=3):
    """SURVEY 8d / BASELINE.md section 3: config C1 exactly on the host cores -- vanilla-ljspeech-stop, batch 8 x 150 characters x 600 frames,
    train() mode, teacher-forced forward + loss + backward, fp32; best and mean of `iters` iterations.  The reference itself
    (oracle/_ref) when staged, else the oracle port."""
    torch.set_num_threads(threads)
    a, Bc, Lc, Tc = synth.VANILLA, 8, 150, 600
    sd = synth.synth_state_dict(a, seed=0)
    b = synth.synth_batch(a, Bc, Lc, Tc, seed=1234)
    times = []
    if _reference_available():
        from oracle import ref_model as R
        m = R.reference_tacotron2(a, sd).train()
        kind = "reference"
        for _ in range(iters):
            m.zero_grad(set_to_none=True)
            t0 = time.perf_counter()
            mels, post, gates, _ = m(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"])
            R.reference_loss(mels, post, gates, b["mel_spectrogram"], b["gate"]).backward()
            times.append(time.perf_counter() - t0)
    else:
        from oracle import tacotron2_oracle as O
        kind = "port"
        masks = synth.synth_masks(a, Bc, Lc, Tc, seed=7, train=True)
        for _ in range(iters):
            sdg = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v) for k, v in sd.items()}
            t0 = time.perf_counter()
            mels, post, gates, _ = O.tacotron2_forward(sdg, a, b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"],
                                                       b["mel_spectrogram_len"], masks=masks, train=True)
            O.loss_fn(mels, post, gates, b["mel_spectrogram"], b["gate"]).backward()
            times.append(time.perf_counter() - t0)
    frames = int(b["mel_spectrogram_len"].sum())
    return {"value": frames / min(times), "unit": "frames/s", "cores": threads, "kind": kind,
            "best_s": min(times), "mean_s": sum(times) / len(times), "iters": iters, "valid_frames": frames,
            "sample": f"C1 exactly: vanilla B={Bc} L={Lc} T={Tc} (ragged lengths), train() teacher-forced forward + loss + backward, torch CPU fp32, "
                      f"whole step, best of {iters}"}


def bf16_mode_us_per_step(dev, steps=3):
    """AR decode of the same workload in the library's bf16 mode (one tensor-core pass instead of three (hi,lo) passes; tolerance in
    tests/test_gpu_parity.py): us per decoder frame."""
    from tacotron2_b200.model import Tacotron2
    old = os.environ.get("T2B200_PRECISION")
    os.environ["T2B200_PRECISION"] = "bf16"
    try:
        a = synth.VANILLA
        m = Tacotron2(**a.kwargs()).to(dev).eval()
        m.load_state_dict(synth.synth_state_dict(a, seed=0, gate_shift=5.0))
        b = synth.synth_batch(a, B, L, 8, seed=1234)
        c, n = b["chars_idx"].to(dev), b["chars_idx_len"].to(dev)
        with torch.no_grad():
            for _ in range(2):
                m(c, n, False, max_len_override=T)
            import ctypes
            from tacotron2_b200 import _lib as L_
            L_.check(L_.load().t2_debug_set_timing(m._ctx, 1))
            tot = 0.0
            for _ in range(steps):
                out = m(c, n, False, max_len_override=T)[1]
                torch.cuda.synchronize()
                assert out.shape[1] == T
                ms_ = ctypes.c_float(0.0)
                L_.check(L_.load().t2_debug_kernel_ms(m._ctx, 0, ctypes.byref(ms_)))
                tot += ms_.value
        del m
        return 1e3 * tot / steps / T
    finally:
        if old is None:
            os.environ.pop("T2B200_PRECISION", None)
        else:
            os.environ["T2B200_PRECISION"] = old


# profiling slots of decoder_ar_kernel (CTA 0, thread 0; csrc/decoder_ar.cu): the dependent chain of one frame, in slot order
PHASE_NAMES = ["att_lstm_tail(exchange+pointwise+h stores)", "query_partials+Q_barrier", "C_barrier(+toeplitz build beside it)", "dec_lstm_tail+R_barrier",
               "projection", "prenet1", "prenet2", "X_barrier+stop_check", "wait_att_accumulator(prenet K-block)", "wait_dec_accumulator(ctx K-blocks)",
               "attn_query_sum+location_term_wait", "attn_energies(tmem)", "attn_exchange", "attn_softmax", "attn_context+stores", "projection_ctx_half"]

TRAIN_B, TRAIN_L, TRAIN_T = 64, 150, 800
TRAIN_WORKLOAD = ("C3: vanilla-lj-hifi-stop (speaker tokens, 4 speakers) teacher-forced training step, batch 64/GPU x 150 chars x 800 frames, "
                  "fp32-parity arithmetic: forward + loss + backward (BPTT) + gradient all-reduce (N>1) + grad-clip 1.0 + Adam(1e-3, wd 1e-6)")


def melspec_bench(dev, steps=5):
    """Log-mel extraction for one C3-sized batch: 64 waveforms of 800 frames (204,800 samples each), resident in HBM.
    Reported as frames/s and as achieved HBM traffic (waveform read once + log-mel written once) against the measured copy peak --
    the kernel is FFT / shared-memory bound, the fraction says how far from the HBM floor it sits."""
    from tacotron2_b200.melspec import TacotronMelSpectrogram
    B, Tm = 64, 800
    S = (Tm - 1) * 256
    wav = torch.from_numpy(np.stack([synth.synth_wav(S, seed=i % 4) for i in range(4)])).repeat(B // 4, 1).to(dev)
    lens = torch.full((B,), S, dtype=torch.int64, device=dev)
    tr = TacotronMelSpectrogram().to(dev)
    for _ in range(3):
        out, n = tr.batch(wav, lens)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out, n = tr.batch(wav, lens)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    frames = int(n.sum())
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
    except Exception:  # noqa: BLE001
        peak = 6650.0
    gbs = (wav.numel() * 4 + frames * 80 * 4) / (ms * 1e-3) / 1e9
    return {"workload": "64 waveforms x 204,800 samples -> 64 x 800 x 80 log-mel, inputs resident", "ms": ms, "frames_per_s": frames / (ms * 1e-3),
            "achieved_gbs": gbs, "hbm_peak_gbs": peak, "frac_of_hbm": gbs / peak}


def postnet_bench(dev, steps=10):
    """The tensor-pipe piece of the path with the simplest roofline: Postnet.forward (model/postnet.py:51-52; five k=5 convs as
    implicit GEMMs on tcgen05) on one C2-sized batch, 64 x 1000 frames, inputs resident.  FLOPs per frame 8.68 M;
    fp32-parity arithmetic issues three bf16 passes, so the tensor pipe executes 3x that -- both rates are reported against the
    measured dense bf16 peak."""
    from tacotron2_b200.model import Tacotron2
    m = Tacotron2(**synth.VANILLA.kwargs()).to(dev).eval()
    m.load_state_dict(synth.synth_state_dict(synth.VANILLA, seed=0))
    x = torch.randn(B, 80, T, device=dev)
    with torch.no_grad():
        for _ in range(3):
            y = m.postnet(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            y = m.postnet(x)
        e1.record()
        e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    assert y.shape == x.shape
    flops = 2 * 5 * (80 * 512 + 3 * 512 * 512 + 512 * 80) * B * T
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak = float(pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 1342.0)))
    except Exception:  # noqa: BLE001
        peak = 1342.0
    tf = flops / (ms * 1e-3) / 1e12
    return {"workload": f"Postnet.forward eval, {B} x {T} frames (5 convs k=5, BN folded, tanh), fp32-parity = 3 bf16 passes; includes the two layout transposes",
            "ms": ms, "frames_per_s": B * T / (ms * 1e-3), "algorithmic_tflops": tf, "issued_bf16_tflops": 3 * tf, "bound": "tensor",
            "peak_tflops": peak, "frac_algorithmic": tf / peak, "frac_issued": 3 * tf / peak}


def hifigan_bench(dev):
    from tacotron2_b200 import hifigan
    return hifigan.bench(dev)


def collate_bench(dev, steps=20):
    """Batch assembly of one C3-sized batch: 64 ragged items (<= 800 frames, <= 150 characters), items
    resident in HBM; one t2_collate launch pads characters and mels and writes the gate targets.  HBM-bound byte movement: achieved
    = (valid bytes read + padded bytes written) / kernel time.  Beside it: the reference's own pad_sequence collate on the host."""
    from tacotron2_b200.collate import CollatePlan
    g = torch.Generator().manual_seed(3)
    Ts = torch.randint(400, 801, (64,), generator=g).tolist()
    Ls = torch.randint(75, 151, (64,), generator=g).tolist()
    Ts[0], Ls[0] = 800, 150
    mel_h = [torch.randn(t, 80, generator=g) for t in Ts]
    chr_h = [torch.randint(1, 40, (l,), generator=g, dtype=torch.int64) for l in Ls]
    plan = CollatePlan({"chars_idx": [c.to(dev) for c in chr_h], "mel_spectrogram": [m.to(dev) for m in mel_h]}, dev, gate_from="mel_spectrogram")
    for _ in range(3):
        plan.launch()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        plan.launch()
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    t0 = time.perf_counter()
    for _ in range(3):   # datasets/tts_dataloader.py:27-28 on the host cores
        torch.nn.utils.rnn.pad_sequence(mel_h, batch_first=True), torch.nn.utils.rnn.pad_sequence(chr_h, batch_first=True)
    cpu_ms = (time.perf_counter() - t0) / 3 * 1e3
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
    except Exception:  # noqa: BLE001
        peak = 6650.0
    gbs = (plan.bytes_read + plan.bytes_written) / (ms * 1e-3) / 1e9
    return {"workload": "64 ragged items (400..800 frames x 80 mels, 75..150 chars) -> padded chars, mels, gate targets; items resident",
            "ms": ms, "bytes_read": plan.bytes_read, "bytes_written": plan.bytes_written, "achieved_gbs": gbs, "hbm_peak_gbs": peak,
            "frac_of_hbm": gbs / peak, "cpu_pad_sequence_ms": cpu_ms, "note": "one launch moves 29 MB; repeated launches find the items in L2, and launch latency rather than bandwidth sets the time"}


def train_bench(dev, rank, world, steps, warmup=3):
    """BASELINE.json metric, training half: mel frames/s of the whole training step (model/tts_model.py:165-202 +
    run/train.py:239-240 settings) on synthetic C3 batches; weak scaling (64 utterances per GPU), gradients averaged over NCCL."""
    import torch.distributed as dist
    from tacotron2_b200.model import Tacotron2
    from tacotron2_b200.parallel import GradSync

    a = synth.HIFI
    model = Tacotron2(**a.kwargs()).to(dev).train()
    model.load_state_dict(synth.synth_state_dict(a, seed=0))
    lib = model._ready(dev)
    from tacotron2_b200 import _lib as L_
    L_.check(lib.t2_debug_set_timing(model._ctx, 1))   # CUDA events around the two persistent kernels, on the launching stream
    kern_ms = [0.0, 0.0]
    sync = GradSync(model) if world > 1 else None
    batch = synth.synth_batch(a, TRAIN_B, TRAIN_L, TRAIN_T, seed=1234 + rank, ragged=False)
    host = {k: v.pin_memory() for k, v in batch.items()}
    from tacotron2_b200.optim import ClipAdam
    params = [p for p in model.parameters()]
    opt = ClipAdam(params, lr=1e-3, weight_decay=1e-6, max_norm=1.0)   # model/tts_model.py:78-91 + run/train.py:239-240, 3 launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    phase = [0.0, 0.0, 0.0, 0.0]
    per_step = []
    loss_h = torch.empty((), pin_memory=True)

    def step(timed):
        b = {k: v.to(dev, non_blocking=True) for k, v in host.items()}  # host -> device copy of the step's inputs (e2e)
        opt.zero_grad(set_to_none=True)
        ev[0].record()
        mels, post, gates, _ = model(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
                                     speaker_id=b["speaker_id"])
        loss = model.loss(mels, post, gates, b["mel_spectrogram"], b["gate"])
        ev[1].record()
        loss.backward()
        ev[2].record()
        if sync is not None:
            sync.finish()
        ev[3].record()
        opt.step()
        ev[4].record()
        loss_h.copy_(loss.detach(), non_blocking=True)  # device -> host read of the step's result
        torch.cuda.current_stream().synchronize()
        if timed:
            import ctypes
            for i in range(4):
                phase[i] += ev[i].elapsed_time(ev[i + 1])
            per_step.append([round(ev[i].elapsed_time(ev[i + 1]), 2) for i in range(2)])
            for which in (0, 1):
                ms_ = ctypes.c_float(0.0)
                L_.check(lib.t2_debug_kernel_ms(model._ctx, which, ctypes.byref(ms_)))
                kern_ms[which] += ms_.value
        return float(loss_h)

    # two untimed steps on top of W: this leg runs after the AR leg has grown and released its buffers, and the first steps re-shape
    # the caching allocator's pools (an 8 GB tape per step) -- a 2-GPU run showed one 97 ms forward among 24.5 ms ones with W = 3
    train_warmup = warmup + 2
    for _ in range(train_warmup):
        step(False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # no Python garbage-collection pass inside the timed region (a full pass over a process that has built two models takes
    # milliseconds of host time in front of a step's first launch; one r02_v11 run showed a 42.6 ms forward among 23.8 ms ones)
    gc.collect()
    gc.disable()
    t0.record()
    losses = [step(True) for _ in range(steps)]
    t1.record()
    t1.synchronize()
    gc.enable()
    total = t0.elapsed_time(t1)
    if world > 1:
        tt = torch.tensor([total], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total = float(tt[0])
    out = {"metric": "mel_frames_per_sec_train_step", "unit": "frames/s", "value": world * TRAIN_B * TRAIN_T * steps / (total * 1e-3),
           "ms_per_step": total / steps, "steps": steps, "warmup": train_warmup, "config": {"workload": TRAIN_WORKLOAD, "utterances_per_gpu": TRAIN_B,
                                                                    "parallelism": f"dp{world}" + (", NCCL bucketed all-reduce overlapped with backward" if world > 1 else "")},
           "phase_ms": {"forward_loss": phase[0] / steps, "backward": phase[1] / steps, "allreduce_exposed": phase[2] / steps,
                        "clip_adam": phase[3] / steps},
           "forward_backward_ms_each_step": per_step,
           "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in host.values()), "d2h_bytes_per_step": 4,
           "loss_first_last": [losses[0], losses[-1]], "grad_bytes_allreduced_per_step": 4 * sum(p.numel() for p in params) if world > 1 else 0}
    try:
        import ctypes
        from tacotron2_b200 import _lib
        cyc = (ctypes.c_ulonglong * 16)()
        _lib.check(_lib.load().t2_debug_phase_cycles(model._ctx, cyc))
        names = ["input_grad_epilogue_barrier", "attention_bwd_tail", "pointwise_barrier", "att_toeplitz_build", "att_ctx_dot", "att_softmax_dense_ds",
                 "att_df_dWd", "att_dwcat", "pointwise_loads_query_partials", "pointwise_query_reduce", "pointwise_cells", "att_csync_df", "att_stage_loads", "input_grad_gemm_stream"]
        out["decoder_bwd_phase_us_per_frame"] = {n: round(cyc[i] / 1965.0 / TRAIN_T, 2) for i, n in enumerate(names)}
        # roofline of the step's dominant kernel (decoder_bwd_kernel): algorithmic bytes per frame for 64 utterances at C3 =
        # transposed LSTM weights 71.3 MB + memory / processed memory 24.6 MB + gate tapes 5.2 MB + operand rows 2.4 MB +
        # taped location term and conv features 6.1 MB; kernel time = CTA 0's cycle counters over all phases
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
        except Exception:  # noqa: BLE001
            peak = 6650.0
        kernel_ms = kern_ms[1] / steps   # decoder_bwd_kernel, CUDA events on the launching stream, mean over the timed steps
        out["decoder_fwd_kernel_ms"] = kern_ms[0] / steps
        traffic_pf, traffic_src = ncu_traffic("decoder_bwd_kernel")
        alg = 109.6e6 * TRAIN_T
        out["roofline"] = {"kernel": "decoder_bwd_kernel", "bound": "hbm", "achieved": alg / (kernel_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                           "frac": alg / (kernel_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": alg, "kernel_ms": kernel_ms,
                           "share_of_step": kernel_ms / (total / steps), "traffic": None if traffic_pf is None else traffic_pf * TRAIN_T,
                           "traffic_source": traffic_src, "kernel_ms_source": "CUDA events around the launch (t2_debug_kernel_ms)"}
    except Exception as e:  # noqa: BLE001
        out["decoder_bwd_phase_us_per_frame"] = str(e)
    model.check_status()
    if sync is not None:
        sync.detach()
    del model, opt
    torch.cuda.empty_cache()
    return out


def bench_config(world):
    """`config` of the JSON line: identical for both arms (the driver compares them)."""
    return {"workload": WORKLOAD, "utterances_per_gpu": B, "chars": L, "frames": T, "parallelism": f"replicas x{world}, no collective"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref: the unmodified model/tacotron2.py) on the
    host cores, all threads, on the SAME workload as the GPU arm: every step is one whole C2 step -- batch 64 x 200 characters, all
    1000 frames, encoder + frame loop + postnet (no sampling: a step takes a few seconds).  Rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        reference_step_cpu(40, threads)   # untimed warm-up (thread pool, allocator): a short call is enough on the CPU
    t, kind = [], "reference"
    for _ in range(args.steps):
        v, dt, kind = reference_step_cpu(T, threads)
        t.append(dt)
    ms = 1e3 * sum(t) / len(t)
    value = B * T / (ms / 1e3)
    sample = f"whole step: B={B} L={L}, all {T} AR frames + encoder + postnet, torch CPU fp32, {threads} threads"
    cfg = bench_config(1)
    cfg["implementation"] = ("oracle/_ref: unmodified reference model/tacotron2.py" if kind == "reference" else "oracle port (oracle/_ref not staged)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step measurement (extra `train` object of the JSON line)")
    ap.add_argument("--no-extras", action="store_true", help="skip the log-mel, bf16-mode and torch-eager side measurements (profiling runs)")
    ap.add_argument("--train-only", action="store_true", help="print only the training-step measurement (profiling aid)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    from tacotron2_b200.model import Tacotron2

    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.train_only:
        out = train_bench(dev, rank, world, max(args.steps, 1))
        if rank == 0:
            print(json.dumps(out))
        if world > 1:
            dist.destroy_process_group()
        return

    a = synth.VANILLA
    model = Tacotron2(**a.kwargs()).to(dev).eval()
    model.load_state_dict(synth.synth_state_dict(a, seed=0, gate_shift=5.0))
    batch = synth.synth_batch(a, B, L, 8, seed=1234 + rank)
    chars_h, lens_h = batch["chars_idx"].pin_memory(), batch["chars_idx_len"].pin_memory()
    chars_d, lens_d = chars_h.to(dev), lens_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    out_h = torch.empty(B, T, a.num_mels).pin_memory()
    gate_h = torch.empty(B, T, 1).pin_memory()   # the masked gates carry the utterance lengths a caller cuts by (run/test.py:167)
    import ctypes
    from tacotron2_b200 import _lib as L_
    lib = model._ready(dev)
    L_.check(lib.t2_debug_set_timing(model._ctx, 1))   # CUDA events around decoder_persistent_kernel, on the launching stream

    def step_resident():
        with torch.no_grad():
            return model(chars_d, lens_d, False, max_len_override=T)[1]

    def step_e2e():
        with torch.no_grad():
            c, n = chars_h.to(dev, non_blocking=True), lens_h.to(dev, non_blocking=True)
            mels, post, gates, al = model(c, n, False, max_len_override=T)
            out_h[:, : post.shape[1]].copy_(post, non_blocking=True)
            gate_h[:, : gates.shape[1]].copy_(gates, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return post

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        step_resident()
    for _ in range(2):
        step_e2e()
    t_wait = time.perf_counter()
    while rank == 0 and not sampler.rows and sampler.proc is not None and time.perf_counter() - t_wait < 20.0:
        step_resident()   # keep the GPU under load until the clock sampler delivers (untimed extra warm-up)

    def timed(fn, k):
        total, dec = 0.0, 0.0
        for _ in range(k):
            flush.zero_()  # flush L2 between timed iterations (untimed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            e1.synchronize()
            assert out.shape[1] == T, "frames were skipped"
            total += e0.elapsed_time(e1)
            ms_ = ctypes.c_float(0.0)
            L_.check(lib.t2_debug_kernel_ms(model._ctx, 0, ctypes.byref(ms_)))   # the persistent kernel alone
            dec += ms_.value
        return total, dec

    barrier()
    sampler.mark()
    gc.collect()
    gc.disable()   # (see train_bench)
    t_res, t_dec = timed(step_resident, args.steps)
    barrier()
    t_e2e, _ = timed(step_e2e, args.steps)
    gc.enable()
    barrier()
    clocks = sampler.stop()
    if world > 1:
        tt = torch.tensor([t_res, t_e2e, t_dec], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_res, t_e2e, t_dec = (float(x) for x in tt)
    model.check_status()
    train = None
    if not args.no_train:
        del flush
        torch.cuda.empty_cache()
        try:
            train = train_bench(dev, rank, world, max(2, min(args.steps, 5)))
        except Exception as e:  # noqa: BLE001  (the AR line must still be printed)
            train = {"error": f"{type(e).__name__}: {e}"}
    if rank != 0:
        dist.destroy_process_group()
        return
    ms = t_res / args.steps
    frames = world * B * T
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    dec_ms = t_dec / args.steps
    bytes_launch = algorithmic_bytes_per_frame_step() * T
    achieved = bytes_launch / (dec_ms * 1e-3) / 1e9
    traffic_pf, traffic_src = ncu_traffic("decoder_ar_kernel")
    cfg = bench_config(world)
    cfg.update({"l2": "flushed between timed iterations (256 MiB write)", "us_per_decoder_step": 1e3 * dec_ms / T})
    line = {
        "metric": METRIC, "value": frames / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "e2e": {"value": frames / (t_e2e / args.steps * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": chars_h.numel() * 8 + lens_h.numel() * 8, "d2h_bytes_per_step": out_h.numel() * 4 + gate_h.numel() * 4 + 4},
        "gpu_launches": LAUNCHES_PER_STEP * args.steps,
        "clocks": clocks,
        "roofline": {"kernel": "decoder_ar_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": None if traffic_pf is None else traffic_pf * T, "traffic_source": traffic_src,
                     "peak_source": "measured" if peaks else "fallback",
                     "algorithmic_bytes_per_launch": bytes_launch, "kernel_ms": dec_ms, "share_of_step": dec_ms / ms,
                     "kernel_ms_source": "CUDA events around the launch (t2_debug_kernel_ms)"},
    }
    try:  # per-phase breakdown of the last decode launch (CTA 0's SM-clock cycles -> us per frame)
        cyc = (ctypes.c_ulonglong * 16)()
        L_.check(lib.t2_debug_phase_cycles(model._ctx, cyc))
        mhz = clocks.get("sm_mhz") or 1965
        line["config"]["phase_us_per_frame"] = {n: round(cyc[i] / mhz / T, 2) for i, n in enumerate(PHASE_NAMES)}
    except Exception as e:  # noqa: BLE001
        line["config"]["phase_us_per_frame"] = str(e)
    if train is not None:
        line["train"] = train
        if "error" not in train:
            # the partitioned path's numbers where the driver keeps them (SCALE_rNN.json retains `config`, not extra keys)
            line["config"].update({"train_workload": TRAIN_WORKLOAD, "train_frames_per_s": train["value"], "train_ms_per_step": train["ms_per_step"],
                                   "train_parallelism": train["config"]["parallelism"],
                                   "allreduce_exposed_ms": train["phase_ms"]["allreduce_exposed"],
                                   "grad_bytes_allreduced": train["grad_bytes_allreduced_per_step"]})
    extras = () if args.no_extras else (("melspec", melspec_bench), ("collate", collate_bench), ("postnet", postnet_bench),
                                        ("hifigan", hifigan_bench), ("bf16_mode_us_per_decoder_step", bf16_mode_us_per_step),
                                        ("torch_cuda_eager", reference_cuda_eager), ("torch_cuda_eager_train", reference_cuda_train))
    for key, fn in extras:
        try:
            line[key] = fn(dev)
        except Exception as e:  # noqa: BLE001
            line[key] = {"error": f"{type(e).__name__}: {e}"}
    eager = line.get("torch_cuda_eager", {})
    if isinstance(eager, dict) and "us_per_decoder_step" in eager:
        line["config"]["reference_cuda_eager_us_per_decoder_step"] = eager["us_per_decoder_step"]
        line["config"]["speedup_vs_reference_cuda_eager_per_frame"] = eager["us_per_decoder_step"] / (1e3 * dec_ms / T)
    if not args.no_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        if train is not None and "error" not in train:
            train["cpu_baseline"] = reference_c1_cpu(threads)
        reference_step_cpu(40, threads)  # warm-up
        v, dt, kind = reference_step_cpu(T, threads)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
                                "sample": f"one whole step of the workload: B={B} L={L}, all {T} AR frames + encoder + postnet, torch CPU fp32, {dt:.1f} s"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    # Anything a library writes to fd 1 (NCCL's version banner, ...) goes to stderr; the JSON line alone reaches stdout.
    sys.stdout.flush()
    _out = os.dup(1)
    os.dup2(2, 1)
    _real, sys.stdout = sys.stdout, open(_out, "w", closefd=False)
    try:
        main()
    finally:
        sys.stdout.flush()
