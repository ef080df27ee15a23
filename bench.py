#!/usr/bin/env python
"""bench.py — headline benchmark of the transducer hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

Metric: RNN-T loss fwd+bwd utterances/second at B=32, T=500, U=100, V=5000 (cfg3 of
BASELINE.json: bf16 logits).  One "step" = one forward+backward pass of the hot path
over one synthetic batch.  Prints ONE JSON line on rank 0.

Workloads (``--workload``):
  rnnt_logits   drop-in rnnt_loss(logits, ...) over materialised bf16 logits resident in HBM
                (HBM-bound kernels: log-sum-exp pass, lattice, gradient pass)

``--impl reference`` times the reference's own CPU implementation of the path (the
oracle port: torch joint + torchaudio.functional.rnnt_loss on the host cores) on a
bounded sample of the same workload.  It is the only mode that executes oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

CFG3 = dict(B=32, T=500, U=100, V=5000, E=256, P=256, J=512)
METRIC = "rnnt_loss_fwd_bwd_utts_per_sec"
UNIT = "utt/s"


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]),
                    bf16_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.nvml = None  # (module, handle): NVML polled every 5 ms from a thread - the timed region is a
        self.stop_flag = False  # few hundred ms, nvidia-smi -lms 100 gave 1-4 samples (often one idle sample)
        self.samples = []

    def _nvml_open(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:  # noqa: BLE001
                h = nv.nvmlDeviceGetHandleByIndex(self.index)
            return nv, h
        except Exception:  # noqa: BLE001
            return None

    def _poll(self):
        nv, h = self.nvml
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.stop_flag:
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM), int(reasons_fn(h))))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def start(self):
        self.nvml = self._nvml_open()
        if self.nvml is not None:
            self.thr = threading.Thread(target=self._poll, daemon=True)
            self.thr.start()
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thr.join(timeout=1.0)
            bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
            reasons = sorted({n for _, _, r in self.samples for bit, n in bits.items() if r & bit})
            sm = [c for c, _, _ in self.samples]
            return {"sm_mhz": statistics.median(sm) if sm else None,
                    "sm_max_mhz": max((m for _, m, _ in self.samples), default=None), "samples": len(sm),
                    "reasons": reasons, "source": "nvml, 5 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- synthetic data
def synth_lengths(B, T, U, gen, ragged):
    if not ragged:
        return torch.full((B,), T, dtype=torch.int32), torch.full((B,), U, dtype=torch.int32)
    tl = torch.randint(int(0.6 * T), T + 1, (B,), generator=gen, dtype=torch.int32)
    ul = torch.randint(int(0.5 * U), U + 1, (B,), generator=gen, dtype=torch.int32)
    tl[0], ul[-1] = T, U
    return tl, ul


class RnntLogitsWorkload:
    """rnnt_loss(logits, ...) fwd+bwd over materialised bf16 logits (HBM-resident)."""
    name = "rnnt_logits"

    def __init__(self, cfg, rank, ragged=False, world=1):
        import wenet_celoss_b200 as pkg
        self.pkg = pkg
        self.cfg = cfg
        B, T, U, V = cfg["B"], cfg["T"], cfg["U"], cfg["V"]
        self.dev = torch.device("cuda", torch.cuda.current_device())
        g = torch.Generator().manual_seed(1234 + rank)
        self.tl_h, self.ul_h = synth_lengths(B, T, U, g, ragged)
        self.targets_h = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32)
        gd = torch.Generator(device=self.dev).manual_seed(1234 + rank)
        self.logits = torch.empty(B, T, U + 1, V, dtype=torch.bfloat16, device=self.dev)
        for b in range(B):  # fill in slices: no fp32 temporary of the whole tensor
            self.logits[b] = torch.randn(T, U + 1, V, generator=gd, device=self.dev, dtype=torch.float32)
        self.logits.requires_grad_(True)
        self.targets = self.targets_h.to(self.dev)
        self.tl = self.tl_h.to(self.dev)
        self.ul = self.ul_h.to(self.dev)
        self.N = B * T * (U + 1) * V
        self.elem = 2

    def step(self):
        self.logits.grad = None
        loss = self.pkg.rnnt_loss(self.logits, self.targets, self.tl, self.ul, blank=0,
                                  reduction="mean")
        loss.backward()
        return loss

    def algorithmic_bytes(self):
        return 3 * self.N * self.elem

    def kernel_breakdown(self, iters=5):
        """CUDA-event time of each C-ABI pass on the launching stream."""
        from wenet_celoss_b200 import _lib
        from wenet_celoss_b200._lib import f32, i32, ptr
        from wenet_celoss_b200.rnnt import rnnt_grad_coef
        B, T, U, V = (self.cfg[k] for k in "BTUV")
        U1 = U + 1
        x = self.logits.detach()
        lat = torch.empty(5, B, T, U1, dtype=torch.float32, device=self.dev)
        lse, lpb, lpe, alpha, beta = lat.unbind(0)
        costs = torch.empty(B, dtype=torch.float32, device=self.dev)
        dy = torch.full((B,), 1.0 / B, dtype=torch.float32, device=self.dev)
        doff = torch.empty(B, 2, T + U1, dtype=torch.float64, device=self.dev)
        grad = torch.empty_like(x)
        st = _lib.stream_handle(self.dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        coef_box = {}

        def k_logprobs():
            _lib.call("b200_rnnt_logprobs", ptr(x), i32(_lib.BF16), ptr(self.targets), ptr(self.tl),
                      ptr(self.ul), i32(B), i32(T), i32(U1), i32(V), i32(0), ptr(lse), ptr(lpb),
                      ptr(lpe), st)

        def k_lattice():
            _lib.call("b200_rnnt_lattice", ptr(lpb), ptr(lpe), ptr(self.tl), ptr(self.ul), i32(B),
                      i32(T), i32(U1), ptr(alpha), ptr(beta), ptr(costs), ptr(doff), st)

        def k_coef():
            coef_box["c"] = rnnt_grad_coef(lse, alpha, beta, doff, dy, self.targets, self.tl,
                                           self.ul, 0)

        def k_grad():
            _lib.call("b200_rnnt_grad", ptr(x), i32(_lib.BF16), ptr(self.targets), ptr(coef_box["c"]),
                      i32(B), i32(T), i32(U1), i32(V), i32(0), f32(-1.0), ptr(grad), st)

        out = {}
        for name, fn in (("rnnt_logprobs_kernel", k_logprobs), ("rnnt_lattice_kernel", k_lattice),
                         ("rnnt_grad_coef_kernel", k_coef), ("rnnt_grad_kernel", k_grad)):
            ts = []
            for _ in range(iters + 1):
                flush.zero_()  # evict L2 between timed launches (small-tensor kernels)
                e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
                e0.record()
                fn()
                e1.record()
                e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            out[name] = statistics.mean(ts[1:])
        return out



    # ---- contract hooks -------------------------------------------------------------
    dtype = "bf16"

    def e2e_setup(self):
        self.tgt_pin = self.targets_h.pin_memory()
        self.tl_pin, self.ul_pin = self.tl_h.pin_memory(), self.ul_h.pin_memory()
        self.loss_pin = torch.empty((), dtype=torch.float32).pin_memory()

    def e2e_step(self):
        self.targets.copy_(self.tgt_pin, non_blocking=True)
        self.tl.copy_(self.tl_pin, non_blocking=True)
        self.ul.copy_(self.ul_pin, non_blocking=True)
        l = self.step()
        self.loss_pin.copy_(l.detach().float(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return l

    def e2e_bytes(self):
        return int(4 * (self.tgt_pin.numel() + self.tl_pin.numel() + self.ul_pin.numel())), 4

    e2e_note = "rnnt_loss(logits,...) API: logits are device-produced; host inputs are targets+lengths"

    def roofline(self, pk, ms_step):
        kb = self.kernel_breakdown()
        dom = "rnnt_grad_kernel"
        dom_bytes = 2 * self.N * self.elem
        achieved = dom_bytes / (kb[dom] * 1e-3) / 1e9
        lp = self.N * self.elem / (kb["rnnt_logprobs_kernel"] * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s",
                "frac": achieved / pk["hbm"], "traffic": None, "peak_source": pk["source"],
                "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms": kb,
                "other": {"rnnt_logprobs_kernel": {"achieved": lp, "frac": lp / pk["hbm"]}},
                "step_frac_of_hbm_roofline":
                    self.algorithmic_bytes() / (ms_step * 1e-3) / 1e9 / pk["hbm"]}

    def describe(self):
        return ("rnnt_logits: rnnt_loss fwd+bwd over materialised bf16 logits, B=32 T=500 U=100 "
                "V=5000 per GPU (cfg3)", "inputs (16.2 GB logits) larger than L2")

    def allreduce(self, dist, loss):
        return None  # independent utterances, no data-path collective


# --------------------------------------------------------------------------- reference arm
FUSED_DESC = ("rnnt_fused: joint (enc_ffn, pred_ffn, tanh, ffn_out) + RNN-T loss fwd+bwd, logits "
              "never materialised, B=32 T=500 U=100 V=5000 J=512 per GPU (cfg3), bf16 MMA operands")


def reference_step(cfg, T_s, threads, seed=0):
    """One fwd+bwd of the reference path (oracle port: torch joint + torchaudio CPU
    rnnt_loss, wenet/transducer/transducer.py:132-147) on one utterance truncated to
    T_s frames.  Returns seconds."""
    import torchaudio
    from oracle import reference as orc
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(seed)
    U, V, E, P, J = cfg["U"], cfg["V"], cfg["E"], cfg["P"], cfg["J"]
    joint = orc.OracleJoint(V, E, P, J)
    enc = torch.randn(1, T_s, E, generator=g, requires_grad=True)
    pred = torch.randn(1, U + 1, P, generator=g, requires_grad=True)
    tgt = torch.randint(1, V, (1, U), generator=g, dtype=torch.int32)
    tl = torch.tensor([T_s], dtype=torch.int32)
    ul = torch.tensor([U], dtype=torch.int32)
    t0 = time.perf_counter()
    logits = joint(enc, pred)
    loss = torchaudio.functional.rnnt_loss(logits, tgt, tl, ul, blank=0, reduction="mean")
    loss.backward()
    return time.perf_counter() - t0


def run_reference(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    budget_s = 120.0
    T_s = 50
    reference_step(cfg, T_s, threads)  # warms the thread pool / allocator
    t_probe = reference_step(cfg, T_s, threads)
    per_frame = t_probe / T_s
    n_total = args.steps + args.warmup
    T_s = int(max(10, min(cfg["T"], budget_s / max(n_total, 1) / per_frame)))
    for _ in range(args.warmup):
        reference_step(cfg, T_s, threads)
    ts = [reference_step(cfg, T_s, threads, seed=i + 1) for i in range(args.steps)]
    sec = sum(ts)
    frac = T_s / cfg["T"]  # fraction of one utterance's lattice processed per step
    value = args.steps * frac / sec
    sample = (f"1 utterance x T={T_s} of {cfg['T']} frames (U={cfg['U']}, V={cfg['V']}), fp32, "
              f"joint + torchaudio.functional.rnnt_loss fwd+bwd per step; utt/s scaled by T_s/T")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": FUSED_DESC, "global_batch": cfg["B"], **cfg, "ragged": False,
                   "parallelism": "cpu (rank 0 only; reference = torch joint + torchaudio rnnt_loss, fp32)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- fused workload
class _Joint(torch.nn.Module):
    """Parameter container with the reference TransducerJoint's names (joint.py:34-43)."""

    def __init__(self, V, E, P, J):
        super().__init__()
        self.enc_ffn = torch.nn.Linear(E, J)
        self.pred_ffn = torch.nn.Linear(P, J)
        self.ffn_out = torch.nn.Linear(J, V)


class RnntFusedWorkload:
    """Headline: joint + RNN-T loss fwd+bwd from encoder / predictor outputs (cfg3), logits
    never materialised; gradients w.r.t. enc_out, pred_out and all joint parameters."""
    name = "rnnt_fused"
    dtype = "bf16"

    def __init__(self, cfg, rank, ragged=False, world=1):
        import wenet_celoss_b200 as pkg
        from wenet_celoss_b200.dist import shard_batch
        self.pkg = pkg
        self.cfg = cfg
        B, T, U, V, E, P, J = (cfg[k] for k in ("B", "T", "U", "V", "E", "P", "J"))
        self.dev = torch.device("cuda", torch.cuda.current_device())
        torch.manual_seed(1234)  # identical replicated weights on every rank
        self.joint = _Joint(V, E, P, J).to(self.dev)
        if ragged and world > 1:
            # ONE global ragged batch of world*B utterances (same seed on every rank), dealt by work
            # T_b*U_b as dist.shard_batch does (sort by work, round-robin)
            gg = torch.Generator().manual_seed(1234)
            tl_g, ul_g = synth_lengths(world * B, T, U, gg, True)
            idx = shard_batch(world * B, rank, world, weights=(tl_g.double() * (ul_g.double() + 1)).tolist())
            self.tl_h, self.ul_h = tl_g[idx].contiguous(), ul_g[idx].contiguous()
            self.tl_h[0], self.ul_h[-1] = T, U  # torchaudio's contract: the maxima equal the tensor sizes
            g = torch.Generator().manual_seed(1234 + rank)
        else:
            g = torch.Generator().manual_seed(1234 + rank)
            self.tl_h, self.ul_h = synth_lengths(B, T, U, g, ragged)
        self.targets_h = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32)
        self.enc_h = torch.randn(B, T, E, generator=g)
        self.pred_h = torch.randn(B, U + 1, P, generator=g)
        self.enc = self.enc_h.to(self.dev).requires_grad_(True)
        self.pred = self.pred_h.to(self.dev).requires_grad_(True)
        self.targets = self.targets_h.to(self.dev)
        self.tl = self.tl_h.to(self.dev)
        self.ul = self.ul_h.to(self.dev)
        self.cells = B * T * (U + 1)
        self.live_cells = int((self.tl_h.long() * (self.ul_h.long() + 1)).sum())
        self.params = list(self.joint.parameters())
        self.reduction = "sum" if world > 1 else "mean"

    def step(self):
        self.enc.grad = None
        self.pred.grad = None
        for p in self.params:
            p.grad = None
        loss = self.pkg.fused_joint_rnnt_loss(self.joint, self.enc, self.pred, self.targets, self.tl,
                                              self.ul, blank=0, reduction=self.reduction,
                                              precision=self.cfg.get("precision", "bf16"))
        loss.backward()
        return loss

    def allreduce(self, dist, loss):
        """The package's own exchange (wenet_celoss_b200/dist.py: what DDP does for these parameters at
        wenet/bin/train.py:231): gradients of the local SUM of costs + {sum, batch} in ONE flat NCCL
        all-reduce, then the global mean.  Returns the global mean loss."""
        from wenet_celoss_b200.dist import allreduce_step
        mean, _ = allreduce_step(self.params, loss.detach(), self.cfg["B"])
        return mean

    def e2e_setup(self):
        self.pins = [t.pin_memory() for t in (self.enc_h, self.pred_h, self.targets_h, self.tl_h,
                                              self.ul_h)]
        self.loss_pin = torch.empty((), dtype=torch.float32).pin_memory()

    def e2e_step(self):
        with torch.no_grad():
            for dst, src in zip((self.enc, self.pred, self.targets, self.tl, self.ul), self.pins):
                dst.copy_(src, non_blocking=True)
        l = self.step()
        self.loss_pin.copy_(l.detach().float(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return l

    def e2e_bytes(self):
        return int(sum(t.numel() * t.element_size() for t in self.pins)), 4

    e2e_note = ("fused_joint_rnnt_loss API: encoder/predictor outputs, targets and lengths copied "
                "from pinned host memory every step; loss read back")

    def kernel_breakdown(self, iters=3):
        from wenet_celoss_b200 import _lib
        from wenet_celoss_b200._lib import i32, ptr
        from wenet_celoss_b200.joint import _i64, _pad_bias, joint_logprobs
        from wenet_celoss_b200.rnnt import rnnt_grad_coef, rnnt_lattice
        cfg = self.cfg
        B, T, U, V, J = (cfg[k] for k in ("B", "T", "U", "V", "J"))
        U1 = U + 1
        with torch.no_grad():
            e = self.joint.enc_ffn(self.enc).float().contiguous()
            p = self.joint.pred_ffn(self.pred).float().contiguous()
            w = self.joint.ffn_out.weight.to(torch.bfloat16).contiguous()
            bias = _pad_bias(self.joint.ffn_out.bias, V)
        st = _lib.stream_handle(self.dev)
        box = {}

        per_utt = T * U1
        H_all = torch.empty(B * per_utt, J, dtype=torch.bfloat16, device=self.dev)

        def k_fwd():  # the training forward: also stores H = tanh(e + p) for the backward kernels
            box["lat"] = joint_logprobs(e, p, w, bias, self.targets, self.tl, self.ul, V, 0, H=H_all)

        lat_out = [torch.empty(B, T, U1, device=self.dev), torch.empty(B, T, U1, device=self.dev),
                   torch.empty(B, device=self.dev),
                   torch.empty(B, 2, T + U1, dtype=torch.float64, device=self.dev)]

        def k_lat():
            _lib.call("b200_rnnt_lattice", ptr(box["lat"][1]), ptr(box["lat"][2]), ptr(self.tl),
                      ptr(self.ul), i32(B), i32(T), i32(U1), ptr(lat_out[0]), ptr(lat_out[1]),
                      ptr(lat_out[2]), ptr(lat_out[3]), st)
            box["ab"] = lat_out

        ldg = (V + 15) // 16 * 16
        G = torch.empty(B * per_utt, ldg, dtype=torch.bfloat16, device=self.dev)
        dbp = torch.zeros(16, (V + 127) // 128 * 128, device=self.dev)
        dy = torch.full((B,), 1.0 / B, device=self.dev)

        def k_grad():  # the launch the step makes: all B utterances in one chunk, H from the forward
            _lib.call("b200_joint_grad_chunk_h", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(self.targets),
                      ptr(box["coef"]), i32(B), i32(T), i32(U1), i32(V), i32(J), i32(0), _i64(0),
                      _i64(B * per_utt), ptr(G), i32(ldg), ptr(H_all), i32(J), ptr(dbp), i32(0), st)

        from wenet_celoss_b200.joint import joint_dw
        dW8 = torch.zeros(V, J, device=self.dev)

        def k_dw():
            joint_dw(G, H_all, B * per_utt, V, J, dW8)

        de = torch.zeros(B, T, J, device=self.dev)
        dp = torch.zeros(B, U1, J, device=self.dev)

        def k_dh():  # dH = G.W with tanh' and the sums over u / t in the epilogue: d e, d p
            de.zero_(); dp.zero_()
            _lib.call("b200_joint_dh", ptr(G), i32(ldg), ptr(w), ptr(H_all), i32(J), i32(B), i32(T), i32(U1),
                      i32(V), i32(J), ptr(self.tl), ptr(self.ul), ptr(de), ptr(dp), i32(0), st)

        # the no-recompute backward (what the step runs when the logit bound allows: joint.py)
        from wenet_celoss_b200 import joint as jmod
        from wenet_celoss_b200.joint import joint_logit_bound
        nr = jmod.LAST_BACKWARD_PATH == "no_recompute"
        bound = joint_logit_bound(w, bias, V)
        rowscale = torch.empty(B * per_utt, device=self.dev)
        Hs = torch.empty_like(H_all) if nr else None
        db = torch.zeros(V, device=self.dev)

        def k_h():  # H = tanh(e + p) (bf16): the A operand of the forward kernel and of the backward GEMMs
            _lib.call("b200_joint_h", ptr(e), ptr(p), i32(B), i32(T), i32(U1), i32(J), ptr(H_all), i32(J), st)

        def k_fwdp():  # joint_kernel<FWDP> on the ready H: statistics and Pt = exp(logit) (bf16) into G
            lat = torch.empty(3, B, T, U1, device=self.dev)
            _lib.call("b200_joint_fwd_p_h", ptr(e), ptr(p), ptr(w), ptr(bias), ptr(self.targets), ptr(self.tl),
                      ptr(self.ul), i32(B), i32(T), i32(U1), i32(V), i32(J), i32(0), ptr(lat[0]), ptr(lat[1]),
                      ptr(lat[2]), ptr(bound), ptr(H_all), i32(J), ptr(G), i32(ldg), i32(0), st)
            box["lat"] = lat.unbind(0)

        def k_rowscale():
            _lib.call("b200_joint_rowscale", ptr(box["coef"]), ptr(box["lat"][0]), ptr(box["lat"][1]),
                      ptr(box["lat"][2]), ptr(self.targets), i32(B), i32(T), i32(U1), i32(V), i32(J), i32(0),
                      ptr(G), i32(ldg), ptr(H_all), i32(J), ptr(Hs), ptr(rowscale), st)

        def k_dw_db():
            _lib.call("b200_joint_dw_db", ptr(G), i32(ldg), ptr(Hs), i32(J), _i64(B * per_utt), i32(V), i32(J),
                      ptr(dW8), ptr(rowscale), ptr(db), i32(0), st)

        def k_dh_scaled():
            de.zero_(); dp.zero_()
            _lib.call("b200_joint_dh_scaled", ptr(G), i32(ldg), ptr(w), ptr(H_all), i32(J), ptr(rowscale), i32(B),
                      i32(T), i32(U1), i32(V), i32(J), ptr(self.tl), ptr(self.ul), ptr(de), ptr(dp), i32(0), st)

        out = {}
        if nr:
            k_h()
        (k_fwdp if nr else k_fwd)(); k_lat()
        alpha, beta, costs, doff = box["ab"]
        box["coef"] = rnnt_grad_coef(box["lat"][0], alpha, beta, doff, dy, self.targets, self.tl, self.ul, 0)
        if nr:
            plan = (("joint_h_kernel", k_h), ("joint_kernel<FWDP> (all cells)", k_fwdp), ("rnnt_lattice_kernel", k_lat),
                    ("joint_rowscale_kernel", k_rowscale), ("joint_dw_kernel (all cells, with d b)", k_dw_db),
                    ("joint_dh_kernel (all cells)", k_dh_scaled))
        else:
            plan = (("joint_kernel<FWD> (all cells)", k_fwd), ("rnnt_lattice_kernel", k_lat),
                    ("joint_kernel<GRAD> (all cells, one chunk)", k_grad),
                    ("joint_dw_kernel (all cells)", k_dw), ("joint_dh_kernel (all cells)", k_dh))
        out["backward_path"] = jmod.LAST_BACKWARD_PATH
        for name, fn in plan:
            ts = []
            for _ in range(iters + 1):
                e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
                e0.record()
                fn()
                e1.record()
                e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            out[name] = statistics.mean(ts[1:])
        return out

    # entry point (include/b200_transducer.h) -> the kernel it launches, for the in-step trace
    ENTRY_KERNEL = {
        "b200_joint_h": "joint_h_kernel", "b200_joint_fwd_p_h": "joint_kernel<FWDP> (all cells)",
        "b200_joint_fwd_p": "joint_h_kernel + joint_kernel<FWDP> (all cells)",
        "b200_joint_fwd": "joint_kernel<FWD> (all cells)",
        "b200_joint_fwd_h": "joint_kernel<FWD> (all cells)", "b200_rnnt_lattice": "rnnt_lattice_kernel",
        "b200_rnnt_grad_coef": "rnnt_grad_coef_kernel", "b200_joint_rowscale": "joint_rowscale_kernel",
        "b200_joint_dw_db": "joint_dw_kernel (all cells, with d b)",
        "b200_joint_dh_scaled": "joint_dh_kernel (all cells)",
        "b200_joint_grad_chunk_h": "joint_kernel<GRAD> (all cells, one chunk)",
        "b200_joint_dw": "joint_dw_kernel (all cells)", "b200_joint_dh": "joint_dh_kernel (all cells)",
    }

    def in_step_kernel_ms(self, steps=5):
        """Per-kernel device time INSIDE real steps: the same ``self.step()`` the timed region runs, with
        every C-ABI call bracketed by CUDA events on the launching stream (``_lib.TRACE``).  Average per
        step; one warm step first.  These are the launch durations the roofline block is computed from."""
        from wenet_celoss_b200 import _lib
        for _ in range(3):
            self.step()
        torch.cuda.synchronize()
        _lib.TRACE = trace = []
        try:
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            for _ in range(steps):
                self.step()
            e1.record()
            e1.synchronize()
        finally:
            _lib.TRACE = None
        out = {}
        for name, a, b in trace:
            k = self.ENTRY_KERNEL.get(name, name)
            out[k] = out.get(k, 0.0) + a.elapsed_time(b) / steps
        out["_traced_step_ms"] = e0.elapsed_time(e1) / steps
        return out

    def roofline(self, pk, ms_step):
        """Roofline of the kernel with the LARGEST time per step (tensor-pipe bound); the other
        tensor-core kernels of the step under ``other``.  Launch durations are measured inside real
        steps (``in_step_kernel_ms``) and set against the SUSTAINED bf16 rate of MEASURED_PEAKS.json (the
        figure for a kernel timed inside a long step); the same kernels timed alone, one launch between
        two synchronisations, and their fraction of the BURST rate are kept under ``alone``.
        ``traffic`` = DRAM bytes per launch of that kernel from the committed ncu --set full capture of
        this round (profiles/r02_traffic.json, keyed by kernel; null when no capture of the current
        kernel exists)."""
        kb = self.in_step_kernel_ms()  # first: the device is still in the timed region's clock / power state
        traced_step_ms = kb.pop("_traced_step_ms")
        alone = self.kernel_breakdown()
        alone_path = alone.pop("backward_path", None)
        kb["backward_path"] = alone_path
        cfg = self.cfg
        flops = 2.0 * self.cells * cfg["J"] * cfg["V"]  # per GEMM-shaped launch (2*cells*J*V)
        path = kb.pop("backward_path", None)
        tensor_kernels = {k: v for k, v in kb.items()
                          if k.startswith(("joint_kernel", "joint_dw_kernel", "joint_dh_kernel"))}
        dom = max(tensor_kernels, key=tensor_kernels.get)
        ach = {k: flops / (v * 1e-3) / 1e12 for k, v in tensor_kernels.items()}
        counted = 6.0 * self.cells * cfg["J"] * cfg["V"]
        traffic, traffic_src = None, None
        try:
            with open(ROOT / "profiles" / "r02_traffic.json") as fh:
                tj = json.load(fh)
            key = dom.split(" ")[0]
            traffic, traffic_src = tj[key]["bytes"], tj[key].get("source")
        except (OSError, KeyError, ValueError):
            pass
        alone_tc = {k: v for k, v in alone.items()
                    if k.startswith(("joint_kernel", "joint_dw_kernel", "joint_dh_kernel"))}
        peak = pk["bf16_sustained"]
        return {"bound": "tensor", "kernel": dom, "achieved": ach[dom], "peak": peak,
                "unit": "TFLOP/s", "frac": ach[dom] / peak, "traffic": traffic,
                "traffic_source": traffic_src,
                "peak_source": pk["source"] + " (sustained bf16: launch durations measured inside the step, "
                               "CUDA events around every C-ABI call on the launching stream)",
                "algorithmic_flops_per_launch": flops, "kernel_ms": kb, "backward_path": path,
                "kernel_ms_sum": sum(v for v in kb.values()), "traced_step_ms": traced_step_ms,
                "other": {k: {"achieved": a, "frac": a / peak} for k, a in ach.items() if k != dom},
                "alone": {"kernel_ms": alone, "peak": pk["bf16"],
                          "frac_of_burst": {k: flops / (v * 1e-3) / 1e12 / pk["bf16"] for k, v in alone_tc.items()},
                          "note": "each kernel launched alone between two synchronisations (cold clocks / "
                                  "power state differ from the step), against the burst bf16 rate"},
                "step_frac_of_tensor_roofline":
                    counted / (ms_step * 1e-3) / 1e12 / pk["bf16_sustained"],
                "step_counted_flops": counted}

    def self_check(self):
        """Before timing: the fused loss of utterance 0 against (a) this package's rnnt_loss and
        (b) torchaudio's CUDA rnnt_loss on the SAME bf16-rounded operands, materialised for that one
        utterance (1 GB of fp32 logits).  Raises when they disagree (rtol 2e-3)."""
        import torchaudio
        with torch.no_grad():
            j = self.joint
            fused = self.pkg.fused_joint_rnnt_loss(j, self.enc, self.pred, self.targets, self.tl, self.ul,
                                                   blank=0, reduction="none", precision="bf16")
            e = j.enc_ffn(self.enc[:1]).float()
            p = j.pred_ffn(self.pred[:1]).float()
            h = torch.tanh(e[:, :, None] + p[:, None]).to(torch.bfloat16)
            w = j.ffn_out.weight.to(torch.bfloat16)
            logits = (h.reshape(-1, h.shape[-1]) @ w.t()).float().add_(j.ffn_out.bias.float())
            logits = logits.view(1, e.shape[1], p.shape[1], -1).contiguous()
            tl0, ul0 = self.tl[:1].clone(), self.ul[:1].clone()
            Tm, Um = int(tl0[0]), int(ul0[0])
            logits = logits[:, :Tm, :Um + 1].contiguous()
            tg = self.targets[:1, :Um].contiguous()
            mine = self.pkg.rnnt_loss(logits, tg, tl0, ul0, blank=0, reduction="none")
            ta = torchaudio.functional.rnnt_loss(logits, tg, tl0, ul0, blank=0, reduction="none")
        f0, m0, t0 = float(fused[0]), float(mine[0]), float(ta[0])
        ok = abs(f0 - t0) <= 2e-3 * abs(t0) and abs(m0 - t0) <= 1e-4 * abs(t0)
        res = {"fused_cost_utt0": f0, "materialised_b200_cost_utt0": m0,
               "torchaudio_cuda_cost_utt0": t0, "rtol_fused": 2e-3, "ok": bool(ok)}
        if not ok:
            raise AssertionError(f"bench self-check failed: {res}")
        return res

    def describe(self):
        return (FUSED_DESC,
                "L2 flushed by the step itself: each step streams >16 GB of gradient scratch")


# --------------------------------------------------------------------------- other rows of the path
class _Predictor(torch.nn.Module):
    """RNNPredictor parameter layout (wenet/transducer/predictor.py:58-86)."""

    def __init__(self, V, emb, out, hidden, layers):
        super().__init__()
        self.embed = torch.nn.Embedding(V, emb)
        self.rnn = torch.nn.LSTM(emb, hidden, layers, batch_first=True)
        self.projection = torch.nn.Linear(hidden, out)


class _SearchModel(torch.nn.Module):
    def __init__(self, cfg):
        super().__init__()
        import wenet_celoss_b200 as pkg
        V, E, P, J = cfg["V"], cfg["E"], cfg["P"], cfg["J"]
        self.blank = 0
        self.predictor = _Predictor(V, 256, P, 256, 2)
        self.joint = _Joint(V, E, P, J)
        self.ctc = pkg.CTC(V, E)
        with torch.no_grad():  # random-init weights emit n_steps symbols per frame:
            self.joint.ffn_out.weight.mul_(3.0)   # rescale so blank / non-blank steps interleave
            self.joint.ffn_out.bias[0] += 7.5
            self.predictor.projection.weight.mul_(16.0)
            self.predictor.embed.weight.mul_(4.0)


class _Mha(torch.nn.Module):
    """MultiHeadedAttention parameter layout (wenet/transformer/attention.py:35-46)."""

    def __init__(self, heads, dim):
        super().__init__()
        self.h = heads
        self.linear_q, self.linear_k = torch.nn.Linear(dim, dim), torch.nn.Linear(dim, dim)
        self.linear_v, self.linear_out = torch.nn.Linear(dim, dim), torch.nn.Linear(dim, dim)

    def forward(self, q, k, v):
        B, Tq, D = q.shape
        dk = D // self.h
        sp = lambda x, lin: lin(x).view(x.size(0), -1, self.h, dk).transpose(1, 2)
        att = torch.softmax(sp(q, self.linear_q) @ sp(k, self.linear_k).transpose(-2, -1) / dk ** 0.5, -1)
        return self.linear_out((att @ sp(v, self.linear_v)).transpose(1, 2).reshape(B, Tq, D)), None


class _ContextBias(torch.nn.Module):
    """The slice of ContextBias (wenet/transformer/context_bias.py:159-399) the greedy search reads,
    random-init, dims of conf/encoder_bias_conformer_rnnt_4_head_bi_1_layer_pred.yaml (256, 4 heads,
    unified_hw_odim 256); the phrase vectors stand in for the BLSTM extractor's output."""

    def __init__(self, D, n_phrases, heads=4):
        super().__init__()
        self.encoder_bias, self.predictor_bias, self.hw_bias = _Mha(heads, D), _Mha(heads, D), _Mha(heads, D)
        LN, Lin = torch.nn.LayerNorm, torch.nn.Linear
        self.encoder_bias_combine, self.predictor_bias_combine = Lin(2 * D, D), Lin(2 * D, D)
        self.encdoer_bias_bias_norm, self.encdoer_bias_out_norm = LN(D), LN(D)
        self.predictor_bias_bias_norm, self.predictor_bias_out_norm = LN(D), LN(D)
        self.hw_bias_norm, self.hw_output_layer = LN(D), Lin(D, 2)
        self.register_buffer("phrases", torch.randn(1, n_phrases, D))
        with torch.no_grad():   # LayerNorm outputs are unit scale: restore the stream scales of _SearchModel
            self.hw_output_layer.weight.mul_(20.0)
            self.predictor_bias_out_norm.weight.mul_(24.0)
            self.encdoer_bias_out_norm.weight.mul_(2.0)

    def forward_bias_hidden(self, context_list, context_lengths):
        return self.phrases[:, :1] if context_list.numel() == 1 else self.phrases

    def forward_encoder_bias(self, bias_hidden, encoder_out):
        bh = bias_hidden.expand(encoder_out.shape[0], -1, -1)
        a = self.encdoer_bias_bias_norm(self.encoder_bias(encoder_out, bh, bh)[0])
        return self.encdoer_bias_out_norm(self.encoder_bias_combine(torch.cat([encoder_out, a], -1))), a


def extra_rows(cfg, dev, pk):
    """Secondary numbers of the same hot path (BASELINE configs 4 and 5), timed with CUDA events."""
    import wenet_celoss_b200 as pkg
    out = {}
    torch.manual_seed(4321)
    g = torch.Generator().manual_seed(4321)

    def timed_run(fn, steps=10, warm=3):
        """ms per step of `steps` back-to-back steps (the bracket bench.py's headline uses: one
        synchronisation on each side of the K steps, none between them)."""
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1) / steps, r

    def timed(fn, reps=3):
        fn()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            r = fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        return statistics.median(ts), r

    # cfg4: greedy + beam(10) over synthetic encoder outputs B=64 T=500 V=5000
    B4, T4 = 64, 500
    model = _SearchModel(cfg).to(dev).eval()
    enc = (torch.randn(B4, T4, cfg["E"], generator=g) * 2.0).to(dev)
    lens = torch.full((B4,), T4, dtype=torch.int32, device=dev)
    ms, (hyps, rounds) = timed(lambda: pkg.greedy_search_batch(model, enc, lens, n_steps=64,
                                                               return_rounds=True))
    ntok = sum(len(h) for h in hyps)
    out["greedy"] = {"value": B4 / (ms * 1e-3), "unit": "utt/s", "ms": ms, "rounds": rounds,
                     "symbols_per_frame": ntok / (B4 * T4),
                     "workload": "cfg4: B=64 T=500 V=5000 greedy, random-init predictor/joint"}
    # the fork's default decode: hot-word gate + go-back (context_filter_state='on'), 100 phrases
    model.context_bias = _ContextBias(cfg["P"], 100).to(dev).eval()
    with torch.no_grad():   # ~0.3 symbols per frame on the biased streams (tuned with the CPU oracle)
        model.joint.ffn_out.bias[0] -= 2.0
    ctx_list = torch.zeros(100, 4, dtype=torch.int32, device=dev)
    ctx_len = torch.full((100,), 4, dtype=torch.int32, device=dev)
    ms, (hyps_g, dec_g, rounds_g) = timed(lambda: pkg.greedy_search_batch(
        model, enc, lens, n_steps=64, context_list=ctx_list, context_lengths=ctx_len,
        context_filter_state="on", return_decisions=True, return_rounds=True), reps=2)
    n0 = sum(d.count(0) for d in dec_g)
    out["greedy_hotword_gated"] = {
        "value": B4 / (ms * 1e-3), "unit": "utt/s", "ms": ms, "rounds": rounds_g,
        "symbols_per_frame": sum(len(h) for h in hyps_g) / (B4 * T4),
        "decisions": sum(len(d) for d in dec_g), "decision0_share": n0 / max(1, sum(len(d) for d in dec_g)),
        "workload": "cfg4 shape, context_filter_state='on' (the fork's default greedy): 100 phrase vectors, "
                    "gate head + predictor bias per predictor call, go-back re-decodes; one cooperative launch "
                    "(greedy_persistent_kernel<GATED>, gate chain per flagged utterance on a cluster of 8 CTAs)"}
    del model.context_bias
    with torch.no_grad():
        model.joint.ffn_out.bias[0] += 2.0
    ms, beams = timed(lambda: pkg.prefix_beam_search_batch(model, enc, lens, beam_size=10), reps=2)
    out["beam10"] = {"value": B4 / (ms * 1e-3), "unit": "utt/s", "ms": ms,
                     "workload": "cfg4: B=64 T=500 V=5000 prefix beam search, beam=10, ctc 0.3 / rnnt 0.7"}
    try:  # the reference's per-utterance Python loops with the model on the GPU (baseline leg)
        out["search_gpu_reference"] = gpu_reference_search(cfg, dev, model, enc, lens)
        out["greedy"]["vs_python_cuda"] = (out["greedy"]["value"] /
                                           out["search_gpu_reference"]["greedy_python_cuda_utts_per_sec"])
        out["beam10"]["vs_python_cuda"] = (out["beam10"]["value"] /
                                           out["search_gpu_reference"]["beam10_python_cuda_utts_per_sec"])
    except Exception as exc:  # noqa: BLE001
        out["search_gpu_reference"] = {"error": repr(exc)[:300]}
    # cfg5: CTC module fwd+bwd B=64 T=1000 V=5000
    B5, T5, L5, V = 64, 1000, 100, cfg["V"]
    ctc = pkg.CTC(V, cfg["E"]).to(dev)
    hs = torch.randn(B5, T5, cfg["E"], generator=g).to(dev).requires_grad_(True)
    yl = torch.randint(50, L5 + 1, (B5,), generator=g)
    ys = torch.randint(1, V, (B5, L5), generator=g)
    ys = torch.where(torch.arange(L5)[None] < yl[:, None], ys, torch.full_like(ys, -1)).to(dev)
    hl = torch.full((B5,), T5, dtype=torch.int32, device=dev)
    yl = yl.to(dev)

    def ctc_step():
        hs.grad = None
        for p_ in ctc.parameters():
            p_.grad = None
        l = ctc(hs, hl, ys, yl)
        l.backward()
        return l
    ms, _ = timed_run(ctc_step)
    logits = torch.randn(B5, T5, V, generator=g).to(dev).requires_grad_(True)

    def ctc_loss_step():
        logits.grad = None
        l = pkg.ctc_loss(logits, ys, hl, yl, reduction="sum") / B5
        l.backward()
        return l
    ms_k, _ = timed_run(ctc_loss_step)
    out["ctc"] = {"value": B5 / (ms * 1e-3), "unit": "utt/s", "ms": ms,
                  "loss_only_ms": ms_k, "loss_only_utts_per_sec": B5 / (ms_k * 1e-3),
                  "loss_only_frac_of_hbm_roofline": 3.0 * B5 * T5 * V * 4 / (ms_k * 1e-3) / 1e9 / pk["hbm"],
                  "timing": "10 back-to-back steps between two synchronisations (as the headline)",
                  "workload": "cfg5: CTC.forward+backward B=64 T=1000 V=5000 L<=100 fp32; 'ms' = the fused "
                              "module (ctc_lo + log_softmax + loss in split_gemm, no logits in HBM), 'loss_only' "
                              "is ctc_loss(logits) fwd+bwd on given logits (3*B*T*V*4 algorithmic bytes)"}
    # SURVEY 8(f) rows 2 and 4 on the cfg5 shape: CTC prefix beam search (beam 10) and forced alignment
    from wenet_celoss_b200.ctc_search import ctc_prefix_beam_search_batch, forced_align_batch
    with torch.no_grad():
        lp5 = torch.log_softmax(logits.detach(), dim=2)
    ms_b, hyps5 = timed(lambda: ctc_prefix_beam_search_batch(lp5, hl, 10), reps=2)
    out["ctc_prefix_beam10"] = {
        "value": B5 / (ms_b * 1e-3), "unit": "utt/s", "ms": ms_b, "frames_per_s": B5 * T5 / (ms_b * 1e-3),
        "workload": "cfg5 shape: B=64 T=1000 V=5000 CTC prefix beam search, beam=10, log-probs resident; "
                    "one CTA per utterance, T dependent frames (bound: per-frame latency chain)"}
    ys_a = ys.clamp(min=1)
    ms_a, _ = timed(lambda: forced_align_batch(lp5, ys_a, hl, yl), reps=3)
    out["ctc_forced_align"] = {
        "value": B5 / (ms_a * 1e-3), "unit": "utt/s", "ms": ms_a,
        "workload": "cfg5 shape: B=64 T=1000 V=5000 L in [50,100] Viterbi alignment, log-probs resident"}
    try:  # reference algorithm on the host for one short utterance (oracle = checker, bounded sample)
        import time as _t
        from oracle import reference as orc
        lp_h = lp5[0, :100].cpu()
        t0 = _t.perf_counter(); orc.ctc_prefix_beam_search(lp_h, 10); dt = _t.perf_counter() - t0
        out["ctc_prefix_beam10"]["cpu_port_utts_per_sec"] = 1.0 / (dt * T5 / 100)
        out["ctc_prefix_beam10"]["cpu_port_sample"] = "1 utterance x 100 of 1000 frames, python port, 1 core, scaled"
        t0 = _t.perf_counter(); orc.forced_align(lp_h, ys_a[0, :10].cpu(), 0); dt = _t.perf_counter() - t0
        out["ctc_forced_align"]["cpu_port_utts_per_sec"] = 1.0 / (dt * (T5 / 100) * (75 / 10))
        out["ctc_forced_align"]["cpu_port_sample"] = "1 utterance x 100 frames x 10 labels, python port, 1 core, scaled by T and L"
    except Exception as exc:  # noqa: BLE001
        out["ctc_prefix_beam10"]["cpu_port_error"] = repr(exc)
    return out



# --------------------------------------------------------------------------- GPU reference
CFG2 = dict(B=32, T=250, U=40, V=4233, E=256, P=256, J=512)


class RnntFusedFp32Workload(RnntFusedWorkload):
    """BASELINE configs[1]: the same fused step in the reference's fp32 configuration (B=32 T=250 U=40
    V=4233 J=512), vocabulary GEMMs as split-bf16 tcgen05 products (rtol 1e-4: tests/test_fp32_path_gpu.py)."""
    name, dtype, base_cfg = "rnnt_fused_fp32", "f32", dict(CFG2, precision="fp32")
    self_check = None

    def roofline(self, pk, ms_step):
        # the three GEMMs of joint.py:68-69 and its autograd over all cells: logits, d W_out, d H
        c = self.cfg
        flops = 3 * 2.0 * self.cells * c["J"] * c["V"]
        tf = flops / (ms_step * 1e-3) / 1e12
        return {"bound": "tensor", "kernel": "whole step (split_gemm LSE / EXP, joint_dw, joint_dh, lattice, enc/pred ffn)",
                "achieved": tf, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": tf / pk["bf16"], "traffic": None,
                "peak_source": pk["source"], "algorithmic_flops_per_step": flops, "issued_flops_per_step": 4.0 * flops,
                "note": "fp32-accurate mode: 3 bf16 products per GEMM (hi.hi + lo.hi + hi.lo) and the logits are "
                        "recomputed in the backward pass, i.e. 4x the algorithmic flops are issued"}

    def describe(self):
        c = self.cfg
        return (f"rnnt_fused_fp32: joint + RNN-T loss fwd+bwd, logits never materialised, B={c['B']} T={c['T']} "
                f"U={c['U']} V={c['V']} J={c['J']} per GPU (cfg2), fp32-accurate split-bf16 MMA operands",
                "L2 flushed by the step itself: each step streams 5.6 GB of split gradient scratch")


def _eager_joint(joint, enc, pred):
    """TransducerJoint.forward as the reference runs it (wenet/transducer/joint.py:55-69): two Linear
    layers, broadcast add, tanh, ffn_out -> materialised [B,T,U1,V] logits (library kernels)."""
    e = joint.enc_ffn(enc)
    p = joint.pred_ffn(pred)
    return joint.ffn_out(torch.tanh(e.unsqueeze(2) + p.unsqueeze(1)))


def _timed_cuda(fn, reps=2, warm=1):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def gpu_reference(cfg, dev, pk):
    """The reference's own GPU path on THIS B200 (SURVEY 8d last row): eager joint +
    torchaudio CUDA rnnt_loss (wenet/transducer/transducer.py:132-147), torchaudio CUDA rnnt_loss alone
    on materialised logits, log_softmax + torch CTC loss (wenet/transformer/ctc.py:57-63), next to the
    same step through this package.  Library kernels only on the reference side; CUDA events."""
    import torchaudio
    import wenet_celoss_b200 as pkg
    out = {"note": "reference call sites reproduced on cuda:0 with the installed torchaudio/ATen sm_100 "
                   "kernels (fp32, allow_tf32=False as torch defaults); 'ours' = this package on the same inputs"}
    g = torch.Generator().manual_seed(777)

    def fused_pair(c, tag, dtype_note, precisions=("bf16",)):
        B, T, U, V, E, P, J = (c[k] for k in ("B", "T", "U", "V", "E", "P", "J"))
        torch.manual_seed(1234)
        joint = _Joint(V, E, P, J).to(dev)
        params = list(joint.parameters())
        tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
        row = {"shape": dict(c), "dtype": dtype_note}
        ours_ms = None
        # torchaudio's CUDA kernels index the logits with 32-bit ints: B*T*U1*V must stay below 2^31 (cfg3 at
        # B=32 is 8.08e9 elements and faults with an illegal memory access - observed on this box)
        b_ref = B
        while b_ref > 1 and b_ref * T * (U + 1) * V >= 2 ** 31:
            b_ref //= 2
        row["reference_batch_limit"] = ("largest power-of-two fraction of B with B*T*U1*V < 2^31 "
                                        "(32-bit indexing in torchaudio's CUDA rnnt_loss)")
        while b_ref >= 1:
            try:
                torch.cuda.empty_cache()
                torch.cuda.reset_peak_memory_stats()
                enc = torch.randn(b_ref, T, E, generator=g).to(dev).requires_grad_(True)
                pred = torch.randn(b_ref, U + 1, P, generator=g).to(dev).requires_grad_(True)
                tl = torch.full((b_ref,), T, dtype=torch.int32, device=dev)
                ul = torch.full((b_ref,), U, dtype=torch.int32, device=dev)

                def ref_step():
                    enc.grad = pred.grad = None
                    for p_ in params:
                        p_.grad = None
                    loss = torchaudio.functional.rnnt_loss(_eager_joint(joint, enc, pred), tgt[:b_ref], tl, ul,
                                                           blank=0, reduction="mean")
                    loss.backward()
                    return loss

                ms = _timed_cuda(ref_step)
                row.update(reference_B=b_ref, reference_ms=ms, reference_utts_per_sec=b_ref / (ms * 1e-3),
                           reference_peak_mem_gb=torch.cuda.max_memory_allocated() / 1e9,
                           reference_loss=float(ref_step().item()))

                for k, prec in enumerate(precisions):
                    def our_step():
                        enc.grad = pred.grad = None
                        for p_ in params:
                            p_.grad = None
                        loss = pkg.fused_joint_rnnt_loss(joint, enc, pred, tgt[:b_ref], tl, ul, blank=0,
                                                         reduction="mean", precision=prec)
                        loss.backward()
                        return loss

                    ours_ms = _timed_cuda(our_step, reps=3)
                    sfx = "" if k == 0 else "_" + prec
                    row.update({"ours_precision" + sfx: prec, "ours_ms" + sfx: ours_ms,
                                "ours_utts_per_sec" + sfx: b_ref / (ours_ms * 1e-3),
                                "ours_loss" + sfx: float(our_step().item()), "speedup" + sfx: ms / ours_ms})
                break
            except torch.OutOfMemoryError:
                b_ref //= 2
        out[tag] = row
        del joint
        torch.cuda.empty_cache()

    fused_pair(CFG2, "cfg2_joint_rnnt", "reference fp32 (SIMT GEMMs); ours precision='fp32' = split-bf16 tcgen05 "
               "(3 products per GEMM, rtol 1e-4: tests/test_fp32_path_gpu.py); '_bf16' keys = bf16 MMA operands",
               precisions=("fp32", "bf16"))
    fused_pair(cfg, "cfg3_joint_rnnt", "reference fp32 (SIMT GEMMs); ours bf16 MMA operands")

    # (b) rnnt_loss alone on materialised logits: torchaudio CUDA vs this package, same tensor
    for tag, c, dt in (("cfg2_rnnt_loss_fp32", CFG2, torch.float32), ("cfg3_rnnt_loss_fp16", cfg, torch.float16)):
        B, T, U, V = c["B"], c["T"], c["U"], c["V"]
        while B > 1 and B * T * (U + 1) * V >= 2 ** 31:  # 32-bit indexing in the reference kernel
            B //= 2
        try:
            torch.cuda.empty_cache()
            logits = torch.empty(B, T, U + 1, V, dtype=dt, device=dev)
            gd = torch.Generator(device=dev).manual_seed(5)
            for b in range(B):
                logits[b] = torch.randn(T, U + 1, V, generator=gd, device=dev)
            logits.requires_grad_(True)
            tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
            tl = torch.full((B,), T, dtype=torch.int32, device=dev)
            ul = torch.full((B,), U, dtype=torch.int32, device=dev)

            def step(fn):
                def run():
                    logits.grad = None
                    l = fn(logits, tgt, tl, ul, blank=0, reduction="mean")
                    l.backward()
                    return l
                return run
            ms_ref = _timed_cuda(step(torchaudio.functional.rnnt_loss))
            ms_our = _timed_cuda(step(pkg.rnnt_loss), reps=3)
            nbytes = 3.0 * logits.numel() * logits.element_size()
            out[tag] = {"B": B, "reference_ms": ms_ref, "ours_ms": ms_our, "speedup": ms_ref / ms_our,
                        "reference_utts_per_sec": B / (ms_ref * 1e-3), "ours_utts_per_sec": B / (ms_our * 1e-3),
                        "ours_frac_of_hbm_roofline": nbytes / (ms_our * 1e-3) / 1e9 / pk["hbm"],
                        "reference_frac_of_hbm_roofline": nbytes / (ms_ref * 1e-3) / 1e9 / pk["hbm"],
                        "reference_loss": float(step(torchaudio.functional.rnnt_loss)().item()),
                        "ours_loss": float(step(pkg.rnnt_loss)().item())}
            del logits
        except torch.OutOfMemoryError as exc:
            out[tag] = {"error": repr(exc)[:200]}
        torch.cuda.empty_cache()

    # (c) cfg5 CTC: CTC.forward of the reference (ctc_lo, log_softmax on [T,B,V], torch CTC loss) vs ours
    B5, T5, L5, V = 64, 1000, 100, cfg["V"]
    ctc = pkg.CTC(V, cfg["E"]).to(dev)
    hs = torch.randn(B5, T5, cfg["E"], generator=g).to(dev).requires_grad_(True)
    yl = torch.randint(50, L5 + 1, (B5,), generator=g)
    ys = torch.randint(1, V, (B5, L5), generator=g)
    ys = torch.where(torch.arange(L5)[None] < yl[:, None], ys, torch.full_like(ys, -1)).to(dev)
    hl = torch.full((B5,), T5, dtype=torch.int32, device=dev)
    yl = yl.to(dev)

    def ctc_ref_step():  # wenet/transformer/ctc.py:57-63
        hs.grad = None
        for p_ in ctc.parameters():
            p_.grad = None
        ys_hat = ctc.ctc_lo(hs).transpose(0, 1).log_softmax(2)
        l = torch.nn.functional.ctc_loss(ys_hat, ys, hl, yl, reduction="sum") / B5
        l.backward()
        return l

    def ctc_our_step():
        hs.grad = None
        for p_ in ctc.parameters():
            p_.grad = None
        l = ctc(hs, hl, ys, yl)
        l.backward()
        return l
    ms_ref = _timed_cuda(ctc_ref_step, reps=3)
    ms_our = _timed_cuda(ctc_our_step, reps=5)
    logits5 = torch.randn(B5, T5, V, generator=g).to(dev).requires_grad_(True)

    def ctc_loss_ref():
        logits5.grad = None
        l = torch.nn.functional.ctc_loss(logits5.transpose(0, 1).log_softmax(2), ys, hl, yl,
                                         reduction="sum") / B5
        l.backward()
        return l

    def ctc_loss_our():
        logits5.grad = None
        l = pkg.ctc_loss(logits5, ys, hl, yl, reduction="sum") / B5
        l.backward()
        return l
    ms_ref_k = _timed_cuda(ctc_loss_ref, reps=3)
    ms_our_k = _timed_cuda(ctc_loss_our, reps=5)
    out["cfg5_ctc"] = {"module_reference_ms": ms_ref, "module_ours_ms": ms_our, "module_speedup": ms_ref / ms_our,
                       "loss_only_reference_ms": ms_ref_k, "loss_only_ours_ms": ms_our_k,
                       "loss_only_speedup": ms_ref_k / ms_our_k,
                       "reference_loss": float(ctc_ref_step().item()), "ours_loss": float(ctc_our_step().item())}
    del logits5
    torch.cuda.empty_cache()
    return out


def logits_row(cfg, pk, steps=5):
    """The metric's "% HBM peak" half: rnnt_loss(logits) fwd+bwd over materialised bf16 logits at cfg3
    (algorithmic bytes 3*N*2), whole step and per kernel, CUDA events."""
    wl = RnntLogitsWorkload(cfg, 0)
    for _ in range(3):
        wl.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(steps):
        wl.step()
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    roof = wl.roofline(pk, ms)
    desc, l2 = wl.describe()
    return {"value": cfg["B"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "workload": desc, "l2": l2,
            "algorithmic_bytes_per_step": wl.algorithmic_bytes(),
            "step_frac_of_hbm_roofline": roof["step_frac_of_hbm_roofline"],
            "kernel_ms": roof["kernel_ms"],
            "kernel_frac_of_hbm_peak": {
                "rnnt_grad_kernel": roof["frac"],
                "rnnt_logprobs_kernel": roof["other"]["rnnt_logprobs_kernel"]["frac"]},
            "hbm_peak_gbs": pk["hbm"], "peak_source": pk["source"]}


def gpu_reference_search(cfg, dev, model, enc, lens):
    """Per-utterance Python greedy / prefix beam search of the reference (search/greedy_search.py:34-176,
    prefix_beam_search.py:42-148; restated in oracle/reference.py) with the model ON THE GPU: ~10 kernel
    launches and a device->host sync per lattice step, i.e. launch-bound.  Baseline leg only."""
    from oracle import reference as orc
    om = orc.OracleModel(cfg["V"], cfg["E"], cfg["P"], cfg["J"], 256, 256, 2, blank=0).to(dev).eval()
    om.load_state_dict(model.state_dict(), strict=False)
    res = {}
    T = int(lens[0])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 2
    for b in range(n):
        orc.greedy_search(om, enc[b:b + 1], T, n_steps=64)
    torch.cuda.synchronize()
    res["greedy_python_cuda_utts_per_sec"] = n / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    orc.prefix_beam_search(om, enc[0:1], beam_size=10)
    torch.cuda.synchronize()
    res["beam10_python_cuda_utts_per_sec"] = 1.0 / (time.perf_counter() - t0)
    res["sample"] = f"greedy: {n} utterances x T={T}; beam 10: 1 utterance x T={T}; model on cuda:0, launch-bound"
    return res


# --------------------------------------------------------------------------- sharded secondary workloads
CFG5 = dict(B=64, T=1000, L=100, V=5000, E=256)
CFG4 = dict(B=64, T=500, V=5000, E=256, P=256, J=512)


class CtcWorkload:
    """BASELINE configs[4]: CTC.forward + backward (wenet/transformer/ctc.py:46-64), B=64 T=1000 V=5000,
    utterances sharded over the ranks; the ctc_lo gradients are the only thing that crosses ranks."""
    name, dtype, base_cfg, metric = "ctc", "f32", CFG5, "ctc_loss_fwd_bwd_utts_per_sec"

    def __init__(self, cfg, rank, ragged=False, world=1):
        import wenet_celoss_b200 as pkg
        self.pkg, self.cfg = pkg, cfg
        B, T, L, V, E = (cfg[k] for k in ("B", "T", "L", "V", "E"))
        self.dev = torch.device("cuda", torch.cuda.current_device())
        torch.manual_seed(1234)
        self.ctc = pkg.CTC(V, E).to(self.dev)
        self.params = list(self.ctc.parameters())
        g = torch.Generator().manual_seed(4321 + rank)
        self.hs_h = torch.randn(B, T, E, generator=g)
        yl = torch.randint(L // 2, L + 1, (B,), generator=g)
        ys = torch.randint(1, V, (B, L), generator=g)
        self.ys_h = torch.where(torch.arange(L)[None] < yl[:, None], ys, torch.full_like(ys, -1))
        hl = torch.full((B,), T, dtype=torch.int32)
        if ragged:
            hl = torch.randint(int(0.6 * T), T + 1, (B,), generator=g, dtype=torch.int32)
            hl[0] = T
        self.hl_h, self.yl_h = hl, yl.to(torch.int32)
        self.hs = self.hs_h.to(self.dev).requires_grad_(True)
        self.ys, self.hl, self.yl = self.ys_h.to(self.dev), self.hl_h.to(self.dev), self.yl_h.to(self.dev)

    def step(self):
        self.hs.grad = None
        for p in self.params:
            p.grad = None
        loss = self.ctc(self.hs, self.hl, self.ys, self.yl) * self.cfg["B"]  # local SUM of nll
        loss.backward()
        return loss

    def allreduce(self, dist, loss):
        from wenet_celoss_b200.dist import allreduce_step
        return allreduce_step(self.params, loss.detach(), self.cfg["B"])[0]

    def e2e_setup(self):
        self.pins = [t.pin_memory() for t in (self.hs_h, self.ys_h, self.hl_h, self.yl_h)]
        self.loss_pin = torch.empty((), dtype=torch.float32).pin_memory()

    def e2e_step(self):
        with torch.no_grad():
            for dst, src in zip((self.hs, self.ys, self.hl, self.yl), self.pins):
                dst.copy_(src, non_blocking=True)
        l = self.step()
        self.loss_pin.copy_(l.detach().float(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return l

    def e2e_bytes(self):
        return int(sum(t.numel() * t.element_size() for t in self.pins)), 4

    e2e_note = "CTC.forward API: encoder outputs, padded labels and lengths copied from pinned host memory every step"

    def roofline(self, pk, ms_step):
        # fused module: no [B,T,V] logits in HBM; the step is the three GEMMs of ctc.py:57 and its autograd
        # (logits, d ctc_lo.weight, d hs: 6*B*T*E*V flops).  Issued: 4 GEMMs (the logits are recomputed in the
        # backward pass) x 3 bf16 products each (fp32-accurate split operands)
        c = self.cfg
        flops = 6.0 * c["B"] * c["T"] * c["E"] * c["V"]
        tf = flops / (ms_step * 1e-3) / 1e12
        return {"bound": "tensor", "kernel": "CTC.forward+backward (whole step: split_gemm LSE / EXP, dW, d hs, trellis)",
                "achieved": tf, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": tf / pk["bf16"], "traffic": None,
                "peak_source": pk["source"], "algorithmic_flops_per_step": flops,
                "issued_flops_per_step": 4.0 * flops,
                "note": "fp32-accurate mode issues 3 bf16 products per GEMM and recomputes the logits in the "
                        "backward pass; the serial trellis (T dependent frames) is inside the step"}

    def describe(self):
        c = self.cfg
        return (f"ctc: CTC.forward+backward B={c['B']} T={c['T']} V={c['V']} L<={c['L']} per GPU (cfg5), fp32",
                "split G scratch (2 x 0.64 GB at B=64) larger than L2")


class SearchWorkload:
    """BASELINE configs[3]: batched greedy / prefix beam search (beam 10) over synthetic encoder outputs,
    utterances sharded over the ranks, no collective (results are gathered on the host by the caller)."""
    dtype, base_cfg = "f32", CFG4
    mode = "greedy"

    def __init__(self, cfg, rank, ragged=False, world=1):
        import wenet_celoss_b200 as pkg
        self.pkg, self.cfg = pkg, cfg
        self.dev = torch.device("cuda", torch.cuda.current_device())
        torch.manual_seed(4321)
        self.model = _SearchModel(dict(CFG3, **cfg)).to(self.dev).eval()
        g = torch.Generator().manual_seed(4321 + rank)
        B, T = cfg["B"], cfg["T"]
        self.enc_h = torch.randn(B, T, cfg["E"], generator=g) * 2.0
        self.lens_h = torch.full((B,), T, dtype=torch.int32)
        if ragged:
            self.lens_h = torch.randint(int(0.6 * T), T + 1, (B,), generator=g, dtype=torch.int32)
        self.enc, self.lens = self.enc_h.to(self.dev), self.lens_h.to(self.dev)
        self.metric = f"rnnt_{self.mode}_search_utts_per_sec"
        self.name = self.mode

    def step(self):
        if self.mode == "greedy":
            self.out = self.pkg.greedy_search_batch(self.model, self.enc, self.lens, n_steps=64)
        else:
            self.out = self.pkg.prefix_beam_search_batch(self.model, self.enc, self.lens, beam_size=10)
        return None

    def allreduce(self, dist, loss):
        return None

    def e2e_setup(self):
        self.pins = [self.enc_h.pin_memory(), self.lens_h.pin_memory()]

    def e2e_step(self):
        self.enc.copy_(self.pins[0], non_blocking=True)
        self.lens.copy_(self.pins[1], non_blocking=True)
        self.step()  # the token lists come back to the host inside the call
        return None

    def e2e_bytes(self):
        n_out = sum(len(h) if self.mode == "greedy" else sum(len(s.hyp) for s in h) for h in self.out)
        return int(sum(t.numel() * t.element_size() for t in self.pins)), int(4 * n_out)

    e2e_note = "encoder outputs copied from pinned host memory every step; token lists returned as Python lists"

    def roofline(self, pk, ms_step):
        return {"bound": "hbm", "kernel": f"{self.mode} decode (dependency chain of T + emitted lattice steps; weights "
                "L2-resident: no bandwidth roofline applies, SURVEY 8d-v)", "achieved": None, "peak": pk["hbm"],
                "unit": "GB/s", "frac": None, "traffic": None, "peak_source": pk["source"]}

    def describe(self):
        c = self.cfg
        return (f"{self.mode}: B={c['B']} T={c['T']} V={c['V']} per GPU (cfg4), random-init predictor/joint"
                + (", beam=10, ctc 0.3 / rnnt 0.7" if self.mode == "beam" else ""), "weights L2-resident")


class BeamWorkload(SearchWorkload):
    mode = "beam"


WORKLOADS = {"rnnt_logits": RnntLogitsWorkload, "rnnt_fused": RnntFusedWorkload,
             "rnnt_fused_fp32": RnntFusedFp32Workload, "ctc": CtcWorkload,
             "greedy": SearchWorkload, "beam": BeamWorkload}


# --------------------------------------------------------------------------- main arm
def cpu_baseline(cfg):
    threads = os.cpu_count() or 1
    T_s = 50
    reference_step(cfg, T_s, threads)
    t = reference_step(cfg, T_s, threads)
    T_s = int(max(10, min(cfg["T"], 10.0 / (t / T_s))))
    t = min(reference_step(cfg, T_s, threads, seed=1), reference_step(cfg, T_s, threads, seed=2))
    return {"value": (T_s / cfg["T"]) / t, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"1 utterance x T={T_s}/{cfg['T']} frames, U={cfg['U']}, V={cfg['V']}, fp32 "
                      f"torch joint + torchaudio rnnt_loss fwd+bwd on the host, best of 2, scaled by T_s/T"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="rnnt_fused", choices=sorted(WORKLOADS))
    ap.add_argument("--ragged", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the config's batch PER GPU; strong: the config's batch split over the GPUs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    cfg = dict(CFG3)

    if args.impl == "reference":
        run_reference(args, cfg)
        return

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import wenet_celoss_b200 as pkg
    wcls = WORKLOADS[args.workload]
    cfg = dict(getattr(wcls, "base_cfg", CFG3))
    global_B = cfg["B"] * world if args.scaling == "weak" else cfg["B"]
    if args.scaling == "strong":
        assert cfg["B"] % world == 0, "strong scaling splits the config's batch evenly"
        cfg["B"] //= world
    wl = wcls(cfg, rank, ragged=args.ragged, world=world)
    pk = peaks()

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms, [ms]
        t = torch.tensor([ms], device="cuda")
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per = [float(x.item()) for x in allt]
        return max(per), per

    def full_step():
        loss = wl.step()
        if dist is not None:
            red = wl.allreduce(dist, loss)
            loss = red if red is not None else loss
        return loss

    check = wl.self_check() if getattr(wl, "self_check", None) else None
    for _ in range(args.warmup):
        full_step()
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(args.steps):
        loss = full_step()
    e1.record()
    sync_all()
    ms, per_rank_ms = max_over_ranks(e0.elapsed_time(e1))
    launches = pkg.launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    B = cfg["B"]
    value = world * B * args.steps / (ms * 1e-3)

    # e2e: same step through the public API with host buffers (pinned), H2D + D2H inside
    wl.e2e_setup()
    wl.e2e_step()
    sync_all()
    f0, f1 = torch.cuda.Event(True), torch.cuda.Event(True)
    f0.record()
    for _ in range(args.steps):
        l_e2e = wl.e2e_step()
        if dist is not None:
            wl.allreduce(dist, l_e2e)
    f1.record()
    sync_all()
    ms_e2e, _ = max_over_ranks(f0.elapsed_time(f1))
    h2d, d2h = wl.e2e_bytes()
    e2e = {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "note": wl.e2e_note}

    exchange = None
    if dist is not None and hasattr(wl, "params"):
        # the collective alone (same flat payload, back to back) and the step WITHOUT it: names what the
        # weak-scaling loss is made of (max over ranks of power-capped GPUs vs the exchange itself)
        from wenet_celoss_b200.dist import flatten_grads
        flat, _ = flatten_grads(wl.params, [torch.zeros(2, device="cuda")])
        for _ in range(3):
            dist.all_reduce(flat)
        sync_all()
        a0, a1 = torch.cuda.Event(True), torch.cuda.Event(True)
        a0.record()
        for _ in range(20):
            dist.all_reduce(flat)
        a1.record()
        sync_all()
        ar_ms, _ = max_over_ranks(a0.elapsed_time(a1) / 20)
        c0, c1 = torch.cuda.Event(True), torch.cuda.Event(True)
        c0.record()
        for _ in range(args.steps):
            wl.step()
        c1.record()
        sync_all()
        _, local_ms = max_over_ranks(c0.elapsed_time(c1) / args.steps)
        exchange = {"payload_bytes": int(flat.numel() * 4), "allreduce_alone_ms": ar_ms,
                    "step_without_exchange_ms_per_rank": local_ms,
                    "step_with_exchange_ms_per_rank": [t / args.steps for t in per_rank_ms],
                    "api": "wenet_celoss_b200.dist.allreduce_step (one flat NCCL all-reduce per step)"}

    if rank == 0:
        roof = wl.roofline(pk, ms / args.steps)
        # the CPU port of the SAME workload: the joint + rnnt_loss chain at this workload's shape (cfg3, or cfg2
        # for the fp32 row); the secondary workloads (ctc, searches) carry their CPU ports inside `extra`
        cb_cfg = dict(getattr(wcls, "base_cfg", CFG3)) if args.workload.startswith("rnnt") else None
        cb = None if (args.no_cpu_baseline or cb_cfg is None) else cpu_baseline(cb_cfg)
        desc, l2 = wl.describe()
        line = {
            "metric": getattr(wl, "metric", METRIC), "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": wl.dtype,
            "data": "synthetic",
            "config": {"workload": desc, "global_batch": global_B, **cfg,
                       "ragged": bool(args.ragged), "l2": l2, "parallelism": f"dp{world}"},
            "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cb,
            "clocks": clocks, "loss": float(loss.item()) if torch.is_tensor(loss) else None,
            "self_check": check, "exchange": exchange,
        }
        if args.ragged and hasattr(wl, "live_cells"):
            line["config"]["live_cell_fraction_rank0"] = wl.live_cells / wl.cells
        if not args.no_extras and world == 1 and args.workload == "rnnt_fused":
            dev = wl.dev
            del wl, loss
            torch.cuda.empty_cache()
            extra = {}
            for name, fn in (("rows", lambda: extra_rows(dict(CFG3), dev, pk)),
                             ("rnnt_logits", lambda: logits_row(dict(CFG3), pk)),
                             ("gpu_reference", lambda: gpu_reference(dict(CFG3), dev, pk))):
                try:  # secondary rows never hide the headline
                    r = fn()
                    if name == "rows":
                        extra.update(r)
                    else:
                        extra[name] = r
                except Exception as exc:  # noqa: BLE001
                    extra[name] = {"error": repr(exc)[:400]}
                torch.cuda.empty_cache()
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
