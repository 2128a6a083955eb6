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
# dram__bytes_read.sum + dram__bytes_write.sum of decoder_persistent_kernel from the committed ncu --set full capture, per decoded
# frame (the capture is a 200-frame launch of the same C2 batch; traffic is linear in the frame count)
NCU_DRAM_BYTES_PER_FRAME = (13.756307e9 + 81.929216e6) / 200
NCU_DRAM_SOURCE = "profiles/r01_v13_decoder_fwd_full_raw.csv (ncu --set full, 200-frame launch; per frame x frames)"
WORKLOAD = "C2: vanilla-ljspeech-stop AR inference, batch 64 x 200 chars x 1000 frames, fp32, stop gate evaluated (never fires)"
# kernels launched by one Tacotron2.forward (AR): embed 1, encoder conv 3, BiLSTM input proj 1, BiLSTM 1, memory split 1,
# att_encoder 1, persistent decoder 1, mel split 1, postnet conv 5, finalize 1  (memsets / memcpys not counted)
LAUNCHES_PER_STEP = 16


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


def oracle_sample(frames, threads):
    """Times the CPU oracle (port of the reference's torch-CPU path) on the same workload, bounded to
    `frames` decoder frames (+ encoder + postnet on those frames).  Returns frames/s."""
    from oracle import tacotron2_oracle as O  # the one place bench.py runs the oracle: the CPU baseline

    torch.set_num_threads(threads)
    sd = synth.synth_state_dict(synth.VANILLA, seed=0, gate_shift=5.0)
    b = synth.synth_batch(synth.VANILLA, B, L, 8, seed=1234)
    masks = synth.synth_masks(synth.VANILLA, B, L, frames, seed=7)
    t0 = time.perf_counter()
    with torch.no_grad():
        out = O.tacotron2_forward(sd, synth.VANILLA, b["chars_idx"], b["chars_idx_len"], False, max_len_override=frames, masks=masks)
    dt = time.perf_counter() - t0
    assert out[0].shape[1] == frames
    return B * frames / dt, dt


def torch_eager_cuda_step(dev, frames=100):
    """The comparator BASELINE.json's north_star names: the reference's own PyTorch-CUDA decoder step at batch 64.  The reference
    cannot travel to the GPU box, so its per-frame op sequence (model/tacotron2.py:276-325 -> model/decoder.py:53-119 ->
    model/attention.py:34-69) is restated here with the same stock torch.nn modules it is built from (LSTMCell, Linear, Conv1d,
    softmax, bmm, F.dropout, the per-frame ``done.all()`` host sync), in eval mode, fp32 with float32_matmul_precision 'high'
    (run/train.py:170).  A reported baseline only: nothing of this repo's kernels runs here and nothing else uses it."""
    import torch.nn as nn
    import torch.nn.functional as F
    a = synth.VANILLA
    E, H, A, M, P = a.encoded_dim, a.att_rnn_dim, a.att_dim, a.num_mels, a.prenet_dim
    prev = torch.get_float32_matmul_precision()
    torch.set_float32_matmul_precision("high")
    g = torch.Generator(device="cpu").manual_seed(0)
    with torch.no_grad():
        pre1, pre2 = nn.Linear(M, P, bias=False).to(dev), nn.Linear(P, P, bias=False).to(dev)
        att_rnn, dec_rnn = nn.LSTMCell(P + E, H).to(dev), nn.LSTMCell(H + E, H).to(dev)
        q_l, v_l = nn.Linear(H, A, bias=False).to(dev), nn.Linear(A, 1, bias=False).to(dev)
        loc_c, loc_d = nn.Conv1d(2, 32, 31, padding=15, bias=False).to(dev), nn.Linear(32, A, bias=False).to(dev)
        mel_l, gate_l = nn.Linear(H + E, M).to(dev), nn.Linear(H + E, 1).to(dev)
        memory = torch.randn(B, L, E, generator=g).to(dev)
        pmem = torch.randn(B, L, A, generator=g).to(dev)
        mask = (torch.arange(L)[None, :] >= torch.randint(120, L + 1, (B,), generator=g)[:, None]).to(dev)

        def run(n):
            z = lambda *sh: torch.zeros(*sh, device=dev)  # noqa: E731
            ah, ac, ctx, w, wc, rh, rc = z(B, H), z(B, H), z(B, E), z(B, L), z(B, L), z(B, H), z(B, H)
            x = F.dropout(F.relu(pre2(F.dropout(F.relu(pre1(z(B, M))), 0.5, True))), 0.5, True)
            done, lengths = torch.zeros(B, dtype=torch.bool, device=dev), torch.zeros(B, dtype=torch.long, device=dev)
            mels, gates, als = [], [], []
            for _ in range(n):
                ah, ac = att_rnn(torch.cat([x, ctx], -1), (ah, ac))
                wcat = torch.cat([w.unsqueeze(1), wc.unsqueeze(1)], 1)
                e = v_l(torch.tanh(q_l(ah.unsqueeze(1)) + loc_d(loc_c(wcat).transpose(1, 2)) + pmem)).squeeze(-1)
                e.masked_fill_(mask, -float("inf"))
                w = F.softmax(e, dim=1)
                ctx = torch.bmm(w.unsqueeze(1), memory).squeeze(1)
                wc = wc + w
                rh, rc = dec_rnn(torch.cat([ah, ctx], -1), (rh, rc))
                hc = torch.cat([rh, ctx], -1)
                mel, gate = mel_l(hc), gate_l(hc)
                mels.append(mel); gates.append(gate); als.append(w)  # noqa: E702
                done[gate.squeeze(-1) < 0] = True
                lengths[gate.squeeze(-1) >= 0] += 1
                if bool(done.all()):
                    break
                x = F.dropout(F.relu(pre2(F.dropout(F.relu(pre1(mel)), 0.5, True))), 0.5, True)
            return torch.stack(mels, 1), len(mels)

        gate_l.bias.fill_(5.0)   # the random-init gate never fires: same as the measured workload
        run(10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, n = run(frames)
        e1.record()
        e1.synchronize()
    torch.set_float32_matmul_precision(prev)
    us = 1e3 * e0.elapsed_time(e1) / n
    return {"us_per_decoder_step": us, "frames_per_s": B * 1e6 / us, "frames": n,
            "what": f"stock torch.nn eager restatement of the reference decoder step on this GPU, B={B} L={L}, fp32 matmul precision 'high', "
                    "eval mode, per-frame stop check with host sync as in model/tacotron2.py:319-322; decoder loop only (no encoder/postnet)"}


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
            tot = 0.0
            for _ in range(steps):
                m._prof = []
                out = m(c, n, False, max_len_override=T)[1]
                torch.cuda.synchronize()
                assert out.shape[1] == T
                tot += sum(x.elapsed_time(y) for x, y in m._prof)
            m._prof = None
        del m
        return 1e3 * tot / steps / T
    finally:
        if old is None:
            os.environ.pop("T2B200_PRECISION", None)
        else:
            os.environ["T2B200_PRECISION"] = old


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
            for i in range(4):
                phase[i] += ev[i].elapsed_time(ev[i + 1])
            per_step.append([round(ev[i].elapsed_time(ev[i + 1]), 2) for i in range(2)])
        return float(loss_h)

    for _ in range(warmup):
        step(False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    losses = [step(True) for _ in range(steps)]
    t1.record()
    t1.synchronize()
    total = t0.elapsed_time(t1)
    if world > 1:
        tt = torch.tensor([total], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total = float(tt[0])
    out = {"metric": "mel_frames_per_sec_train_step", "unit": "frames/s", "value": world * TRAIN_B * TRAIN_T * steps / (total * 1e-3),
           "ms_per_step": total / steps, "steps": steps, "config": {"workload": TRAIN_WORKLOAD, "utterances_per_gpu": TRAIN_B,
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
        names = ["input_grad_gemms", "attention_bwd_tail", "pointwise_dec_and_barrier", "att_stage_conv", "att_ctx_dot", "att_softmax_dense_ds",
                 "att_df_dWd", "att_dwcat", "pointwise_dq_stage", "pointwise_query_bwd", "pointwise_att", "att_csync_df", "att_halo", "att_dWc"]
        out["decoder_bwd_phase_us_per_frame"] = {n: round(cyc[i] / 1965.0 / TRAIN_T, 2) for i, n in enumerate(names)}
        # roofline of the step's dominant kernel (decoder_bwd_kernel): algorithmic bytes per frame for 64 utterances at C3 =
        # transposed LSTM weights 71.3 MB + memory / processed memory 24.6 MB + gate tapes 5.2 MB + operand rows 2.4 MB +
        # taped location term and conv features 6.1 MB; kernel time = CTA 0's cycle counters over all phases
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
        except Exception:  # noqa: BLE001
            peak = 6650.0
        kernel_ms = sum(cyc[i] for i in range(len(names))) / 1965.0 / 1e3
        alg = 109.6e6 * TRAIN_T
        out["roofline"] = {"kernel": "decoder_bwd_kernel", "bound": "hbm", "achieved": alg / (kernel_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                           "frac": alg / (kernel_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": alg, "kernel_ms": kernel_ms,
                           "share_of_step": kernel_ms / (total / steps), "traffic": 60.716650e9 + 2.027008e9,
                           "traffic_source": "profiles/r01_v14_decoder_bwd_full_raw.csv (ncu --set full, same launch shape)"}
    except Exception as e:  # noqa: BLE001
        out["decoder_bwd_phase_us_per_frame"] = str(e)
    if sync is not None:
        sync.detach()
    del model, opt
    torch.cuda.empty_cache()
    return out


def oracle_train_sample(threads, B=8, L=150, T=24):
    """CPU baseline of the training half: the oracle's teacher-forced forward + loss + backward (torch CPU autograd), bounded sample."""
    from oracle import tacotron2_oracle as O

    torch.set_num_threads(threads)
    a = synth.VANILLA
    sd = synth.synth_state_dict(a, seed=0)
    sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v) for k, v in sd.items()}
    b = synth.synth_batch(a, B, L, T, seed=1234, ragged=False)
    masks = synth.synth_masks(a, B, L, T, seed=7, train=True)
    t0 = time.perf_counter()
    mels, post, gates, _ = O.tacotron2_forward(sd, a, b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
                                               masks=masks, train=True)
    O.loss_fn(mels, post, gates, b["mel_spectrogram"], b["gate"]).backward()
    dt = time.perf_counter() - t0
    return B * T / dt, dt, f"vanilla B={B} L={L} T={T} teacher-forced fwd + loss + bwd, torch CPU fp32 autograd over the oracle, {dt:.1f} s"


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    frames = 24  # bounded sample per step: ~2-4 s of host time
    for _ in range(args.warmup):
        oracle_sample(frames, threads)
    t = []
    for _ in range(args.steps):
        v, dt = oracle_sample(frames, threads)
        t.append(dt)
    ms = 1e3 * sum(t) / len(t)
    value = B * frames / (ms / 1e3)
    sample = f"B={B} L={L}, first {frames} of {T} AR frames + encoder + postnet per step, torch CPU fp32"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
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
    len_h = torch.empty(B, dtype=torch.int64).pin_memory()

    def step_resident():
        with torch.no_grad():
            return model(chars_d, lens_d, False, max_len_override=T)[1]

    def step_e2e():
        with torch.no_grad():
            c, n = chars_h.to(dev, non_blocking=True), lens_h.to(dev, non_blocking=True)
            mels, post, gates, al = model(c, n, False, max_len_override=T)
            out_h[:, : post.shape[1]].copy_(post, non_blocking=True)
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
            model._prof = []
            e0.record()
            out = fn()
            e1.record()
            e1.synchronize()
            assert out.shape[1] == T, "frames were skipped"
            total += e0.elapsed_time(e1)
            dec += sum(x.elapsed_time(y) for x, y in model._prof)
        model._prof = None
        return total, dec

    barrier()
    sampler.mark()
    t_res, t_dec = timed(step_resident, args.steps)
    barrier()
    t_e2e, _ = timed(step_e2e, args.steps)
    barrier()
    clocks = sampler.stop()
    if world > 1:
        tt = torch.tensor([t_res, t_e2e, t_dec], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_res, t_e2e, t_dec = (float(x) for x in tt)
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
    line = {
        "metric": METRIC, "value": frames / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "l2": "flushed between timed iterations (256 MiB write)", "utterances_per_gpu": B,
                   "us_per_decoder_step": 1e3 * dec_ms / T, "parallelism": f"replicas x{world}, no collective"},
        "e2e": {"value": frames / (t_e2e / args.steps * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": chars_h.numel() * 8 + lens_h.numel() * 8, "d2h_bytes_per_step": out_h.numel() * 4 + 4},
        "gpu_launches": LAUNCHES_PER_STEP * args.steps,
        "clocks": clocks,
        "roofline": {"kernel": "decoder_persistent_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": NCU_DRAM_BYTES_PER_FRAME * T, "traffic_source": NCU_DRAM_SOURCE,
                     "peak_source": "measured" if peaks else "fallback",
                     "algorithmic_bytes_per_launch": bytes_launch, "kernel_ms": dec_ms, "share_of_step": dec_ms / ms},
    }
    try:  # per-phase breakdown of the last decode launch (CTA 0's SM-clock cycles -> us per frame)
        import ctypes
        from tacotron2_b200 import _lib
        cyc = (ctypes.c_ulonglong * 16)()
        _lib.check(_lib.load().t2_debug_phase_cycles(model._ctx, cyc))
        mhz = clocks.get("sm_mhz") or 1965
        names = ["att_lstm_tail", "query_partials_barrier", "attention_ctx_store_barrier", "dec_lstm_tail_barrier", "proj(beside att gemm)",
                 "prenet1(beside att gemm)", "prenet2(beside att gemm)", "x_barrier_and_prenet_kblock_gemm", "att_lstm_gemm_with_output_stage_beside",
                 "dec_lstm_gemm", "attn_query_sum", "attn_energies", "attn_csync", "attn_softmax", "attn_context",
                 "location_precompute(beside dec gemm)"]
        line["config"]["phase_us_per_frame"] = {n: round(cyc[i] / mhz / T, 2) for i, n in enumerate(names)}
    except Exception as e:  # noqa: BLE001
        line["config"]["phase_us_per_frame"] = str(e)
    if train is not None:
        line["train"] = train
    extras = () if args.no_extras else (("melspec", melspec_bench), ("collate", collate_bench), ("postnet", postnet_bench), ("bf16_mode_us_per_decoder_step", bf16_mode_us_per_step),
                                        ("torch_cuda_eager", torch_eager_cuda_step))
    for key, fn in extras:
        try:
            line[key] = fn(dev)
        except Exception as e:  # noqa: BLE001
            line[key] = {"error": f"{type(e).__name__}: {e}"}
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        if train is not None and "error" not in train:
            oracle_train_sample(threads, T=4)  # warm-up
            v, dt, sample = oracle_train_sample(threads)
            train["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample}
        oracle_sample(4, threads)  # warm-up
        v, dt = oracle_sample(T, threads)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
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
