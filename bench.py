#!/usr/bin/env python
"""bench.py — LLaMA-2-7B fp16 decode tok/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py --gpus N --steps K --warmup W [--workload cfg2|cfg3|cfg4] [--impl reference]

A "step" is one decode step of the whole model for the whole batch (one new token per sequence).
Workloads (BASELINE.json configs):
  cfg2 (default): LLaMA-2-7B fp16 random-init full model, batch 1 greedy decode at a 512-token
        context. N GPUs = N independent requests (weights replicated, no communication): weak.
  cfg4: batch 32 x KV length 4096, 32/N sequences per GPU (strong), no communication.
  cfg5: LLaMA-2-70B-shaped GQA (64 q / 8 kv heads, 80 layers) fp16 decode, tensor parallel over the N GPUs
        (strong scaling: heads / FFN columns split, allreduce after o_proj and down_proj only);
        --comm p2p = the fused one-shot NVLink allreduce + residual kernel, --comm nccl = the NCCL baseline.
  cfg3: prefill, batch 4 x 2048 prompt tokens per GPU (metric prefill_tokens_per_s; a step = one prompt
        pass of the whole model: tcgen05 GEMMs + flash context attention; roofline = tensor).
`value`  : tokens/s with everything resident in HBM, CUDA events on the engine's stream, max over ranks.
`e2e`    : same metric through the call a user of the reference makes per token
           (b200_engine_decode_step_host: H2D token ids from pinned memory, the step, D2H ids).
`roofline`: the dominant kernel. cfg2: decode_step_kernel (the whole step is one launch of it):
           algorithmic bytes of a step / its launch time (CUDA events over the timed region); the
           stand-alone GEMV family (CUDA-graph timed) rides along as `roofline_gemv_family`.
           cfg4: the split-KV decode attention, timed alone back to back on the model's own cache.
`cpu_baseline`: the CPU oracle port (oracle/llama_cpu.c) on a bounded sample, host cores stated.
--impl reference: the reference's own CPU test functions (oracle/_ref, tests/unittests/*.cu)
           composed into the decode step, on the host cores, same metric / config.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L7B = dict(num_layers=32, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--comm", default="p2p", choices=["p2p", "nccl"],
                    help="cfg5: fused one-shot NVLink allreduce (product) or ncclAllReduce + AddResidual (baseline)")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--context", type=int, default=None)
    ap.add_argument("--mode", default=None, choices=["persistent", "fused_graph", "unfused_graph"],
                    help="cfg2 / cfg4: force a decode launch order (default: the engine's own choice)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kv-fp8", action="store_true",
                    help="cfg4: fp8 (e4m3) KV cache with one scale per cached row (B200_ENGINE_KV_FP8); an engine option, "
                         "not the reference's cache type - the line says so in config.kv_cache")
    ap.add_argument("--kv-paged", action="store_true",
                    help="cfg4: paged KV cache (B200_ENGINE_KV_PAGED: page pools [n_pages, hk, 128, d] + one page list per "
                         "sequence); same arithmetic and results as the dense cache, the line says so in config.kv_cache")
    return ap.parse_args()


def workload(args):
    if args.workload == "cfg3":
        return dict(name="LLaMA-2-7B fp16 prefill: context attention + tcgen05 GEMMs, batch 4 x seq 2048 (configs[2])",
                    batch=4, context=args.context or 2048, scaling="weak")
    if args.workload == "cfg5":
        return dict(name="LLaMA-2-70B-shaped GQA (64 q / 8 kv heads) fp16 decode, tensor-parallel (configs[4])",
                    batch=args.batch or 1, context=args.context or 512, scaling="strong")
    if args.workload == "cfg4":
        return dict(name="LLaMA-2-7B fp16 decode, batch 32 x KV 4096 (configs[3])", batch=args.batch or 32,
                    context=args.context or 4096, scaling="strong")
    return dict(name="LLaMA-2-7B fp16 random-init full model, batch 1 greedy decode, 512-token context (configs[1])",
                batch=1, context=args.context or 512, scaling="weak")


# --------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# --------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc, self.lines, self.index = None, [], index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t_end = time.time() + 3.0
        while len(self.lines) < 2 and time.time() < t_end:   # nvidia-smi needs ~0.2 s before its first line
            time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# CPU baselines
# --------------------------------------------------------------------------------------------
def _cheap_random(shape, scale, seed):
    """fast deterministic pseudo-random fp32 fill (tiles a 1M-element block; content is irrelevant
    to a timing run)"""
    import numpy as np
    n = int(np.prod(shape))
    blk = (np.random.default_rng(seed).standard_normal(1 << 20).astype(np.float32) * scale)
    return np.resize(blk, n).reshape(shape)


def cpu_baseline_port(wl, budget_s=10.0):
    """oracle port (plain C, pthreads over all host cores): whole decode steps of a LLaMA-2-7B-shaped fp32 model
    (32 layers that share ONE 0.8 GB set of layer weights - far larger than the host caches, so every layer still
    streams its weights from DRAM - + final norm + LM head) at the workload's batch/context, repeated for about
    `budget_s` seconds after one untimed step. Long-context batches keep 4 sequences and 4 layers (the fp32 KV cache
    of 32 layers x 4 x 4096 keys would be 17 GB of host memory) and scale, saying so."""
    import numpy as np
    from oracle import cpu, model as omodel
    B, ctx = wl["batch"], wl["context"]
    B_s = min(B, 4)   # bounded sample: tokens/s scales with batch on the CPU too
    n_layers = 32 if B_s * ctx <= 4096 else 4
    H, I, V = 4096, 11008, 32000
    layers = [dict(attn_norm=np.ones(H, np.float32), qkv=_cheap_random((3 * H, H), 0.02, 1),
                   o=_cheap_random((H, H), 0.02, 2), ffn_norm=np.ones(H, np.float32),
                   gate_up=_cheap_random((2 * I, H), 0.02, 3), down=_cheap_random((H, I), 0.02, 4))] * n_layers
    w = dict(layers=layers, embed=_cheap_random((V, H), 0.02, 5), final_norm=np.ones(H, np.float32),
             lm_head=_cheap_random((V, H), 0.02, 6))
    cfg = dict(L7B, num_layers=n_layers)
    max_steps = 64
    om = omodel.OracleModel(w, cfg, B_s, ctx + max_steps + 4, "fp32")
    om.k[:] = 0.1
    om.v[:] = 0.1
    om.step = ctx
    ids = np.zeros(B_s, np.int32)
    om.decode_step(ids)   # untimed: first touch of every buffer
    x = np.zeros((B_s, H), np.float32)
    lg = np.zeros((B_s, V), np.float32)
    t0 = time.perf_counter()
    cpu.linear(x, w["lm_head"], lg, True)
    t_head = time.perf_counter() - t0
    steps, t0 = 0, time.perf_counter()
    while steps < max_steps - 2:
        om.decode_step(ids)
        steps += 1
        t_all = time.perf_counter() - t0
        if t_all >= budget_s:
            break
    t_step = t_all / steps
    if n_layers != 32:   # only the layer part scales
        t_step = (t_step - t_head) * (32 / n_layers) + t_head
    return {"value": round(B_s / t_step * (B / B_s), 4), "unit": "tok/s", "cores": cpu.num_threads(),
            "kind": "port",
            "sample": f"{steps} decode steps of {n_layers} LLaMA-2-7B fp32 layers + LM head, batch {B_s}, context {ctx}+"
                      + (f", layer time scaled x{32 // n_layers}" if n_layers != 32 else "")
                      + (f", tokens/s scaled x{B // B_s} to batch {B}" if B_s != B else "")
                      + f" ({t_all:.1f} s measured after one untimed step)"}


def cpu_baseline_hf(wl, n_layers=8, n_decode=12):
    """The torch-CPU `LlamaForCausalLM` leg north_star names (reference: modeling_llama.py + test_llama_run.py:1-29;
    the vendored modeling_llama.py is not importable on its own - relative imports into the transformers package -
    so the container's transformers LlamaForCausalLM stands in). fp32, use_cache=True, LLaMA-2-7B
    geometry with `n_layers` of the 32 layers (host memory / time bound), batch 1: prefill of the workload's
    context, then greedy decode steps; the per-token time is scaled to 32 layers (the LM head is timed inside the
    steps and not scaled)."""
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.set_num_threads(os.cpu_count() or 1)   # torchrun exports OMP_NUM_THREADS=1: this leg uses all host cores
    ctx = min(wl["context"], 512)
    cfg = LlamaConfig(hidden_size=4096, intermediate_size=11008, num_hidden_layers=n_layers, num_attention_heads=32,
                      num_key_value_heads=32, vocab_size=32000, max_position_embeddings=4096, rms_norm_eps=1e-5)
    with torch.device("meta"):
        m = LlamaForCausalLM(cfg)
    m = m.to_empty(device="cpu").eval()
    blk = torch.randn(1 << 20) * 0.02
    with torch.no_grad():
        for _, prm in m.named_parameters():
            if prm.dim() == 1:
                prm.fill_(1.0)
            else:
                flat = prm.view(-1)
                flat.copy_(blk.repeat((flat.numel() + blk.numel() - 1) // blk.numel())[:flat.numel()])
        d = 128
        inv = 1.0 / (10000.0 ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
        for name, buf in m.named_buffers():   # to_empty() leaves the rotary table uninitialised
            if name.endswith("inv_freq"):
                buf.copy_(inv)
        ids = torch.randint(0, 32000, (1, ctx))
        t0 = time.perf_counter()
        out = m(ids, use_cache=True)
        t_prefill = time.perf_counter() - t0
        pkv = out.past_key_values
        nxt = out.logits[:, -1].argmax(-1, keepdim=True)
        out = m(nxt, past_key_values=pkv, use_cache=True)   # untimed first decode step
        pkv, nxt = out.past_key_values, out.logits[:, -1].argmax(-1, keepdim=True)
        # LM head + final norm alone, to keep them out of the layer scaling
        hdn = torch.randn(1, 1, 4096)
        t0 = time.perf_counter()
        for _ in range(3):
            m.lm_head(m.model.norm(hdn))
        t_head = (time.perf_counter() - t0) / 3
        t0 = time.perf_counter()
        for _ in range(n_decode):
            out = m(nxt, past_key_values=pkv, use_cache=True)
            pkv, nxt = out.past_key_values, out.logits[:, -1].argmax(-1, keepdim=True)
        t_dec = (time.perf_counter() - t0) / n_decode
    t_tok = (t_dec - t_head) * (32 / n_layers) + t_head
    return {"value": round(wl["batch"] / t_tok, 3), "unit": "tok/s", "cores": torch.get_num_threads(), "kind": "port",
            "impl": "transformers LlamaForCausalLM under torch-CPU fp32 (stands in for the reference's modeling_llama.py)",
            "sample": f"{n_decode} greedy decode steps of {n_layers} of the 32 LLaMA-2-7B fp32 layers + LM head after a "
                      f"{ctx}-token prefill ({t_prefill:.1f} s), batch 1, use_cache=True; layer time scaled x{32 // n_layers} "
                      f"({t_dec * 1e3:.0f} ms measured per step, LM head {t_head * 1e3:.0f} ms)"
                      + (f"; tokens/s scaled x{wl['batch']} to batch {wl['batch']}" if wl["batch"] != 1 else "")}


class RefCpu:
    """The reference's own CPU test functions (tests/unittests/*.cu compiled into
    oracle/_ref/libref.so) composed into a decode-layer sample. CPUlinear is a single-threaded
    triple loop over W[K,N]; the output columns are sliced over all host cores (one slice per
    thread, ctypes releases the GIL). Attention / RoPE use the oracle port: the reference's
    CPUMaskedAttn / CPUfunc are unsound outside their toy shapes."""

    def __init__(self):
        self.lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))

    def linear_sliced(self, x, w_slices, outs):
        ths = []
        for ws, o in zip(w_slices, outs):
            t = threading.Thread(target=self.lib.ref_cpu_linear,
                                 args=(x.ctypes.data_as(C.c_void_p), ws.ctypes.data_as(C.c_void_p),
                                       o.ctypes.data_as(C.c_void_p), x.shape[0], x.shape[1], ws.shape[1]))
            t.start()
            ths.append(t)
        for t in ths:
            t.join()


def reference_arm(args, wl):
    """--impl reference: the reference's own CPU test functions (oracle/_ref = tests/unittests/*.cu compiled where
    they lie) composed into WHOLE decode steps of the 7B-shaped model on the host cores and timed as they run:
    per layer its CPU RMSNorm, CPUlinear for qkv / o / gate_up / down (output columns sliced over all cores, one
    slice per thread), its CPU SwiGLU, and decode attention + RoPE through the oracle port (the reference's
    CPUMaskedAttn / CPUfunc are unsound outside their toy shapes); then the LM head through CPUlinear.
    The 32 layers share ONE 0.8 GB set of fp32 layer weights (far larger than the host caches, so every layer still
    streams them from DRAM). A step is a whole token when the run fits the wall-clock budget; otherwise a step is
    the first `sample_layers` layers and the line says so (`extrapolated`, `scale`): `ms_per_step` is always the
    MEASURED time of what a step ran."""
    import numpy as np
    cores = os.cpu_count() or 1
    H, I, V = 4096, 11008, 32000
    B = min(wl["batch"], 4)
    ctx = wl["context"]
    try:
        ref = RefCpu()
        kind = "reference"
    except OSError:
        ref, kind = None, "port"
    from oracle import cpu

    BLK = 64   # CPUlinear walks W[k*n + j] with a stride of n floats: 64-column blocks keep it cache-friendly

    def blocks(K, N, seed):   # W[K, N] as N/BLK consecutive [K, BLK] column blocks
        return _cheap_random((N // BLK, K, BLK), 0.02, seed)

    w_qkv, w_o, w_gu, w_dn, w_head = blocks(H, 3 * H, 10), blocks(H, H, 40), blocks(H, 2 * I, 70), blocks(I, H, 100), blocks(H, V, 130)
    g = np.ones(H, np.float32)
    kc = _cheap_random((B, 32, ctx + 1, 128), 0.1, 5)
    vc = kc.copy()
    att = np.zeros((B, H), np.float32)
    act = np.zeros((B, I), np.float32)
    fp = lambda a: a.ctypes.data_as(C.c_void_p)

    def lin(x, wb):
        nblk, K, _ = wb.shape
        out = np.zeros((B, nblk * BLK), np.float32)
        if ref:
            ref.lib.ref_cpu_linear_blocked(fp(x), fp(wb), fp(out), B, K, nblk * BLK, BLK, cores)
        else:
            for b in range(nblk):
                o = np.zeros((B, BLK), np.float32)
                cpu.linear(x, wb[b], o, False)
                out[:, b * BLK:(b + 1) * BLK] = o
        return out

    def norm(x):
        x = x.copy()
        if ref:
            ref.lib.ref_cpu_rmsnorm(fp(x), fp(g), C.c_float(1e-5), H, B)
        else:
            cpu.rmsnorm(x, None, g, 1e-5)
        return x

    def layer(x):
        qkv = lin(norm(x), w_qkv).reshape(B, 96, 128)
        cpu.rope_decode(qkv, 32, 32, ctx + 1, 128, 10000.0)
        cpu.decode_attention(qkv, kc, vc, att, 32, 32, ctx + 1, True)
        x = x + lin(att, w_o)
        gu = lin(norm(x), w_gu).reshape(B, 2, I)
        if ref:
            ref.lib.ref_cpu_swiglu(fp(gu), fp(act), B, I)
        else:
            cpu.swiglu(gu, act)
        return x + lin(act, w_dn) * 0.01

    def step(n_layers, head):
        x = _cheap_random((B, H), 1.0, 1)
        for _ in range(n_layers):
            x = layer(x)
        if head:
            lin(norm(x), w_head)

    # size a step: one layer + the LM head timed once, then as many layers per step as the budget allows
    t0 = time.perf_counter(); layer(_cheap_random((B, H), 1.0, 1)); t_layer = time.perf_counter() - t0
    t0 = time.perf_counter(); lin(_cheap_random((B, H), 1.0, 2), w_head); t_head = time.perf_counter() - t0
    budget_s = float(os.environ.get("B200_REF_BUDGET_S", "150"))
    n_runs = args.steps + max(1, min(args.warmup, 3))
    sample_layers = 32
    while sample_layers > 1 and n_runs * (sample_layers * t_layer + t_head) > budget_s:
        sample_layers //= 2
    whole = sample_layers == 32
    for _ in range(max(1, min(args.warmup, 3))):
        step(sample_layers, whole)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(sample_layers, whole)
    t_sample = (time.perf_counter() - t0) / args.steps
    t_step = t_sample if whole else t_sample * (32 / sample_layers) + t_head
    value = wl["batch"] / t_step if B == wl["batch"] else (B / t_step) * (wl["batch"] / B)
    sample = (f"per step: {'one whole token' if whole else f'{sample_layers} of the 32 layers'} of the 7B-shaped fp32 "
              f"model through the reference's CPUlinear (tests/unittests/test_linear.cu:13 on 64-column blocks of W dealt over {cores} "
              f"threads), its CPU RMSNorm + SwiGLU, decode attention / RoPE via the oracle port at context {ctx}, batch {B}"
              + ("" if whole else f"; layer time scaled x{32 // sample_layers}, LM head ({t_head * 1e3:.0f} ms, measured once) added")
              + (f"; tokens/s scaled x{wl['batch'] // B} to batch {wl['batch']}" if B != wl["batch"] else ""))
    out = {"impl": "reference", "metric": "decode_tokens_per_s", "value": round(value, 5), "unit": "tok/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": round(t_sample * 1e3, 3), "higher_is_better": True,
           "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": wl["name"], "batch": wl["batch"], "context": wl["context"]},
           "extrapolated": (not whole) or B != wl["batch"],
           "scale": {"layers_per_step_measured": sample_layers, "layers_of_model": 32,
                     "lm_head_in_step": whole, "batch_measured": B, "batch_of_workload": wl["batch"],
                     "ms_per_token_model": round(t_step * 1e3, 3)},
           "cpu_baseline": {"value": round(value, 5), "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample},
           "e2e": {"value": round(value, 5), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


# --------------------------------------------------------------------------------------------
# cfg3: prefill
# --------------------------------------------------------------------------------------------
def cpu_baseline_prefill(wl, n_tok=128):
    """oracle port of the context decoder loop on a bounded sample: ONE 7B-shaped fp32 layer over one
    sequence of n_tok prompt tokens (history 0), scaled to 32 layers"""
    import numpy as np
    from oracle import cpu, model as omodel
    H, I = 4096, 11008
    layer = dict(attn_norm=np.ones(H, np.float32), qkv=_cheap_random((3 * H, H), 0.02, 1), o=_cheap_random((H, H), 0.02, 2),
                 ffn_norm=np.ones(H, np.float32), gate_up=_cheap_random((2 * I, H), 0.02, 3), down=_cheap_random((H, I), 0.02, 4))
    om = omodel.OracleModel(dict(layers=[layer]), dict(L7B, num_layers=1), 1, n_tok + 1, "fp32")
    x = _cheap_random((n_tok, H), 1.0, 5)
    t0 = time.perf_counter()
    om.prefill_layers(x, [n_tok], [0])
    t = time.perf_counter() - t0
    return {"value": round(n_tok / (32 * t), 3), "unit": "tok/s", "cores": cpu.num_threads(), "kind": "port",
            "sample": f"1 of 32 LLaMA-2-7B fp32 layers of the context decoder over one {n_tok}-token prompt "
                      f"({t:.2f} s measured), time scaled x32"}


def reference_arm_prefill(args, wl):
    """--impl reference --workload cfg3: the reference's CPUlinear (tests/unittests/test_linear.cu:13) on 1/FRAC of
    one layer's linear columns over a 16-token prompt slice, sliced over all host cores; attention is 4% of
    the flops at this shape and is left to the linear-time scaling (stated)."""
    import numpy as np
    cores = os.cpu_count() or 1
    H, I = 4096, 11008
    n_tok, FRAC = 16, 16
    try:
        ref, kind = RefCpu(), "reference"
    except OSError:
        ref, kind = None, "port"
    from oracle import cpu
    per = max(1, ((3 * H + H + 2 * I) // FRAC) // cores)
    per2 = max(1, (H // FRAC) // cores)
    x, xa = _cheap_random((n_tok, H), 1.0, 1), _cheap_random((n_tok, I), 1.0, 2)
    w1 = [_cheap_random((H, per), 0.02, 10 + i) for i in range(cores)]
    w2 = [_cheap_random((I, per2), 0.02, 30 + i) for i in range(cores)]
    o1 = [np.zeros((n_tok, per), np.float32) for _ in range(cores)]
    o2 = [np.zeros((n_tok, per2), np.float32) for _ in range(cores)]

    def sample():
        if ref:
            for o in o1 + o2:
                o[:] = 0
            ref.linear_sliced(x, w1, o1)
            ref.linear_sliced(xa, w2, o2)
        else:
            for ws, o in zip(w1, o1):
                cpu.linear(x, ws, o, False)
            for ws, o in zip(w2, o2):
                cpu.linear(xa, ws, o, False)

    for _ in range(max(1, min(args.warmup, 3))):
        sample()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sample()
    t = (time.perf_counter() - t0) / args.steps
    done = per * cores * H + per2 * cores * I                 # weight elements covered by one sample
    t_layer = t * (3 * H * H + H * H + 2 * I * H + H * I) / done
    value = n_tok / (32 * t_layer)
    out = {"impl": "reference", "metric": "prefill_tokens_per_s", "value": round(value, 4), "unit": "tok/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(t * 1e3, 3),
           "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": wl["name"], "batch": wl["batch"], "context": wl["context"]},
           "cpu_baseline": {"value": round(value, 4), "unit": "tok/s", "cores": cores, "kind": kind,
                            "sample": f"per step: 1/{FRAC} of one 7B layer's linear columns over {n_tok} prompt tokens "
                                      f"through the reference's CPUlinear, {cores} threads; scaled to 32 layers "
                                      f"(attention, 4% of the flops, not timed)"},
           "e2e": {"value": round(value, 4), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def bench_prefill(args, wl, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package
    pkg = load_package()
    E, capi, ops = pkg.engine, pkg.capi, pkg.ops
    if not torch.cuda.is_available():
        raise SystemExit("[b200][ERROR] bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()   # no collective kernel may slip in between queued cooperative step kernels (see extras_multi_gpu)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, S, K, W = wl["batch"], wl["context"], args.steps, max(args.warmup, 3)
    eng = E.Engine(B, S + 8, dtype=capi.F16, **L7B)
    eng.init_random(seed=1234 + rank, std=0.02)
    stream = torch.cuda.ExternalStream(eng.stream)
    ids = np.random.default_rng(7 + rank).integers(0, 32000, (B, S)).astype(np.int32)
    for _ in range(W):
        eng.reset()
        eng.prefill(ids)
    barrier()
    clocks = Clocks(local_rank)
    if rank == 0:
        clocks.start()
    capi.lib().b200_launch_count_reset()
    # device-resident: the ids are re-sent (32 KB), nothing is read back; events on the engine's stream
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for a, b in evs:
        eng.reset()
        with torch.cuda.stream(stream):
            a.record()
            eng.prefill(ids, sync=False)
            b.record()
    barrier()
    launches = int(capi.lib().b200_launch_count())
    ms = sum(a.elapsed_time(b) for a, b in evs)
    clk = clocks.stop() if rank == 0 else None
    # end to end: host ids in, first generated ids out on the host, every step
    t0 = time.perf_counter()
    for _ in range(K):
        eng.reset()
        eng.prefill(ids)
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(t[0]), float(t[1])
    Nt, H, I = B * S, 4096, 11008
    lin_flops = 2.0 * Nt * (3 * H * H + H * H + 2 * I * H + H * I) * 32
    attn_flops = 4.0 * B * 32 * 128 * S * (S + 1) / 2 * 32
    head_flops = 2.0 * B * 32000 * H
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    peak_src = ("MEASURED_PEAKS.json bf16_tflops_sustained (cuBLAS bf16 back to back)" if peaks
                else "fallback 1400 TFLOP/s")
    roof = roof_attn = None
    if rank == 0:
        # the dominant kernel alone: the four tcgen05 GEMMs of every layer on the engine's weights
        x_h = (torch.randn(Nt, H, device="cuda") * 0.1).half()
        x_i = (torch.randn(Nt, I, device="cuda") * 0.1).half()
        y_qkv = torch.empty(Nt, 3 * H, dtype=torch.float16, device="cuda")
        y_h = torch.empty(Nt, H, dtype=torch.float16, device="cuda")
        y_gu = torch.empty(Nt, 2 * I, dtype=torch.float16, device="cuda")

        def gemm_pass():
            for l in range(32):
                ops.launchLinearGemm(x_h, eng.weight_view(E.W_QKV, l), y_qkv, trans_b=True)
                ops.launchLinearGemm(x_h, eng.weight_view(E.W_O, l), y_h, trans_b=True)
                ops.launchLinearGemm(x_h, eng.weight_view(E.W_GATE_UP, l), y_gu, trans_b=True)
                ops.launchLinearGemm(x_i, eng.weight_view(E.W_DOWN, l), y_h, trans_b=True)

        gemm_pass()
        torch.cuda.synchronize()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        gemm_pass()
        gemm_pass()
        s1.record()
        torch.cuda.synchronize()
        g_ms = s0.elapsed_time(s1) / 2
        ach = lin_flops / (g_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "gemm_tcgen05_kernel", "achieved": round(ach, 1), "peak": peak,
                "unit": "TFLOP/s", "frac": round(ach / peak, 4), "traffic": None,
                "per_launch_us": round(g_ms * 1e3 / 128, 1), "algorithmic_flops_per_launch": lin_flops / 128,
                "peak_source": peak_src,
                "note": f"128 GEMMs of one prompt pass (32 layers x qkv/o/gate_up/down, {Nt} tokens) back to back on "
                        "the model's own weights; operands beyond L2 (12.9 GB of weights per pass)"}
        # the flash context attention alone, 32 layers
        q = (torch.rand(B, 32, S, 128, device="cuda") - 0.5).half()
        kc, vc = eng.kv_caches()
        out = torch.empty(Nt, H, dtype=torch.float16, device="cuda")
        lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
        hist = torch.zeros(B, dtype=torch.int32, device="cuda")
        pad = torch.zeros(Nt, dtype=torch.int32, device="cuda")

        def attn_pass():
            for l in range(32):
                ops.context_attention(q, kc[l], vc[l], out, pad, lens, hist)

        attn_pass()
        torch.cuda.synchronize()
        s0.record()
        attn_pass()
        attn_pass()
        s1.record()
        torch.cuda.synchronize()
        a_ms = s0.elapsed_time(s1) / 2
        ach = attn_flops / (a_ms * 1e-3) / 1e12
        roof_attn = {"bound": "tensor", "kernel": "context_attn_tcgen05_kernel", "achieved": round(ach, 1), "peak": peak,
                     "unit": "TFLOP/s", "frac": round(ach / peak, 4), "traffic": None,
                     "per_launch_us": round(a_ms * 1e3 / 32, 1), "algorithmic_flops_per_launch": attn_flops / 32,
                     "peak_source": peak_src, "note": "causal flops only (4*d per visible query-key pair)"}
    if rank == 0:
        tok_s = world * Nt * K / (ms_max * 1e-3)
        flops = lin_flops + attn_flops + head_flops
        out = {"metric": "prefill_tokens_per_s", "value": round(tok_s, 1), "unit": "tok/s", "n_gpus": world, "steps": K,
               "warmup": W, "ms_per_step": round(ms_max / K, 3), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "f16", "data": "synthetic",
               "config": {"workload": wl["name"], "batch_per_gpu": B, "seq_len": S, "tokens_per_step_per_gpu": Nt,
                          "weights": "random-init N(0,0.02^2) fp16", "accumulate": "fp32",
                          "l2": "inputs larger than L2 (12.9 GB of weights + 1.2 GB of activations per pass)",
                          "parallelism": f"dp{world} (independent prompts, no collective)"},
               "e2e": {"value": round(world * Nt * K / (e2e_ms_max * 1e-3), 1), "unit": "tok/s",
                       "h2d_bytes_per_step": 4 * Nt + 12 * B, "d2h_bytes_per_step": 4 * B,
                       "ms_per_step": round(e2e_ms_max / K, 3)},
               "gpu_launches": launches, "launches_per_step": eng.prefill_launches,
               "step_tensor": {"algorithmic_TFLOP_per_step": round(flops / 1e12, 2),
                               "achieved_TFLOPs": round(flops / (ms_max / K * 1e-3) / 1e12, 1)},
               "roofline": roof, "roofline_context_attention": roof_attn, "clocks": clk}
        if not args.no_cpu_baseline:
            try:
                out["cpu_baseline"] = cpu_baseline_prefill(wl)
            except Exception as ex:
                out["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "port",
                                       "sample": f"failed: {ex}"}
        print(json.dumps(out))
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------
# cfg5: tensor-parallel decode of a 70B-shaped model
# --------------------------------------------------------------------------------------------
L70B = dict(num_layers=80, head_num=64, kv_head_num=8, head_size=128, inter_size=28672, vocab_size=32000)


def bench_tp(args, wl, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package
    pkg = load_package()
    E, capi, sh = pkg.engine, pkg.capi, pkg.sharding
    if not torch.cuda.is_available():
        raise SystemExit("[b200][ERROR] bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()   # no collective kernel may slip in between queued cooperative step kernels (see extras_multi_gpu)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, ctx, K, W = wl["batch"], wl["context"], args.steps, max(args.warmup, 3)
    cfg = dict(L70B)
    if os.environ.get("B200_BENCH_LAYERS"):   # smoke runs on small boxes; the reported config is 80 layers
        cfg["num_layers"] = int(os.environ["B200_BENCH_LAYERS"])
    tp_size = world
    if world == 1 and os.environ.get("B200_BENCH_TP_SIZE"):   # profiling aid: one rank's shard shapes on one GPU,
        tp_size = int(os.environ["B200_BENCH_TP_SIZE"])        # partial sums simply not exchanged (ncu is 1-GPU only)
    eng = E.Engine(B, ctx + 2 * (W + K) + 8, dtype=capi.F16, tp_size=tp_size, tp_rank=rank, **cfg)
    eng.init_random(seed=1234 + rank, std=0.02)
    eng.fill_kv_random(ctx, seed=7 + rank, amp=0.5)
    eng.set_mode(fused=True, graph=True, persistent=False)
    comm = None
    if world > 1:
        if args.comm == "p2p":
            comm = sh.P2PComm(B * 8192 * 2)
            eng.set_p2p(comm)
        else:
            eng.set_allreduce(sh.make_allreduce())
    stream = torch.cuda.ExternalStream(eng.stream)
    ids0 = np.full(B, 11, np.int32)   # every rank must feed the same tokens
    eng.decode_step_host(ids0)
    for _ in range(W - 1):
        eng.decode_step()
    barrier()
    clocks = Clocks(local_rank)
    if rank == 0:
        clocks.start()
    capi.lib().b200_launch_count_reset()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record()
        for _ in range(K):
            eng.decode_step()
        ev1.record()
    barrier()
    launches = int(capi.lib().b200_launch_count())
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop() if rank == 0 else None
    ids = eng.decode_step_host(None)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        ids = eng.decode_step_host(ids)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(t[0]), float(t[1])
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        nl = cfg["num_layers"]
        kv_bytes = B * nl * 2 * (8 // world if world <= 8 else 1) * 128 * 2 * (ctx + W + K // 2)
        step_bytes = eng.weight_bytes - 32000 * 8192 * 2 + kv_bytes   # embedding table is a row gather
        ach = step_bytes / (ms_max / K * 1e-3) / 1e9
        out = {"metric": "decode_tokens_per_s", "value": round(B * K / (ms_max * 1e-3), 2), "unit": "tok/s",
               "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(ms_max / K, 4), "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
               "config": {"workload": wl["name"], "batch": B, "context": ctx, "layers": nl,
                          "weights": f"random-init fp16, {eng.weight_bytes / 1e9:.1f} GB per GPU",
                          "l2": "inputs larger than L2 (all weights streamed once per step)",
                          "parallelism": f"tp{world}" + (f" ({'one-shot NVLink allreduce+residual kernel' if args.comm == 'p2p' else 'ncclAllReduce + AddResidual'}, 2 per layer)" if world > 1 else "")},
               "e2e": {"value": round(B * K / (e2e_ms_max * 1e-3), 2), "unit": "tok/s", "h2d_bytes_per_step": 4 * B,
                       "d2h_bytes_per_step": 4 * B, "ms_per_step": round(e2e_ms_max / K, 4)},
               "gpu_launches": launches, "launches_per_step": eng.launches_per_step,
               "roofline": {"bound": "hbm", "kernel": "gemv_stream_kernel (whole TP step per GPU)", "achieved": round(ach, 1),
                            "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None,
                            "algorithmic_bytes_per_launch": int(step_bytes),
                            "note": "per-GPU bytes of one step (this rank's weight shards + KV rows) / step time; "
                                    "allreduce time is inside the step"},
               "clocks": clk}
        print(json.dumps(out))
    eng.close()
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------
# proof that the timed launch order did the work ("make the timed region prove it")
# --------------------------------------------------------------------------------------------
def verify_decode(pkg, eng, model_shape, weight_seed, ctx, kv_seed, ids0, n_steps=6):
    """After the timed region: rewind the timed engine to the cache state it was timed from and let it decode
    `n_steps` greedy tokens in the SAME mode that was timed. A second engine with the same weights runs the
    reference launch order (self_decoder.cpp:53-85: one kernel per launch*, plain weight-streaming GEMVs, no
    CUDA graph, no persistent kernel) on the first <= 2 sequences, teacher-forced with the timed engine's
    tokens. Logits must agree within the fp16 tolerance (2e-2 of max |logit|) and the greedy picks must be
    identical wherever the check engine's top-2 margin exceeds that band. Both orders are pinned to the CPU
    oracle at these shapes by tests/test_7b_shapes_gpu.py; this run-time check is about the timed run itself."""
    import numpy as np
    import torch
    E, capi = pkg.engine, pkg.capi
    B = eng.batch
    nb = min(B, 2)
    chk = E.Engine(nb, ctx + n_steps + 8, dtype=capi.F16, **model_shape)
    chk.init_random(seed=weight_seed, std=0.02)
    chk.set_mode(fused=False, graph=False)
    chk.fill_kv_random(ctx, seed=kv_seed, amp=0.5)
    eng.fill_kv_random(ctx, seed=kv_seed, amp=0.5)
    kc2, vc2 = chk.kv_caches()
    if getattr(eng, "kv_paged", False):   # rows of the first nb sequences straight out of their pages
        kp, vp = eng.kv_pools()
        tab = eng.kv_page_table().cpu()
        npg = (ctx + 127) // 128
        for b in range(nb):
            pages = tab[b, :npg].long().cuda()
            for pool, dst in ((kp, kc2), (vp, vc2)):
                rows = pool[:, pages].permute(0, 2, 1, 3, 4).reshape(pool.shape[0], pool.shape[2], npg * 128, pool.shape[4])
                dst[:, b, :, :ctx].copy_(rows[:, :, :ctx])
    else:
        kc, vc = eng.kv_caches()
        kc2[:, :, :, :ctx].copy_(kc[:, :nb, :, :ctx])
        vc2[:, :, :, :ctx].copy_(vc[:, :nb, :, :ctx])
    torch.cuda.synchronize()   # the copies ran on torch's stream, the engines step on their own
    ids = np.ascontiguousarray(ids0, dtype=np.int32)
    worst, same, compared = 0.0, 0, 0
    for _ in range(n_steps):
        out = eng.decode_step_host(ids)
        la = eng.logits()[:nb].float().cpu().numpy()
        out2 = chk.decode_step_host(ids[:nb])
        lb = chk.logits().float().cpu().numpy()
        scale = float(np.max(np.abs(lb)))
        worst = max(worst, float(np.max(np.abs(la - lb))) / max(scale, 1e-30))
        srt = np.sort(lb, axis=1)
        for b in range(nb):
            if srt[b, -1] - srt[b, -2] > 4 * 2e-2 * scale:
                compared += 1
                same += int(out[b] == out2[b])
        ids = out
    launches = chk.launches_per_step
    chk.close()
    ok = bool(np.isfinite(worst) and worst <= 2e-2 and same == compared)
    return {"checked": ok, "max_rel_err": round(worst, 6), "ids_equal": same == compared, "ids_compared": compared,
            "steps": n_steps, "sequences": nb, "tolerance": 2e-2,
            "against": f"reference launch order ({launches} launches per step, GEMV linears, no graph) on a second "
                       "engine with the same weights and cache rows, teacher-forced with the timed engine's tokens"}


# --------------------------------------------------------------------------------------------
# N > 1 only: the partitioned workloads of north_star in the SAME process group and JSON line
# --------------------------------------------------------------------------------------------
def tp_step_bytes(eng, B, ctx_mid, world, n_layers):
    kv = B * n_layers * 2 * max(1, 8 // world) * 128 * 2 * ctx_mid
    return eng.weight_bytes - 32000 * 8192 * 2 + kv   # the embedding table is a row gather


def extras_multi_gpu(pkg, args, rank, world, local_rank, peak):
    """(i) cfg5: LLaMA-2-70B-shaped tensor-parallel decode over the `world` GPUs, once with the fused GEMV-push +
    one-shot NVLink allreduce (--comm p2p) and once with ncclAllReduce + AddResidual on the same engine, plus the
    agreement of their logits; (ii) cfg4: 32 sequences x KV 4096 split 32/world per GPU, no collective.
    Device-timed (CUDA events on the engine stream, max over ranks), short runs."""
    import numpy as np
    import torch
    import torch.distributed as dist
    E, capi, sh = pkg.engine, pkg.capi, pkg.sharding
    K, W = min(args.steps, 24), 3
    out = {}

    # Every cross-rank barrier below is bracketed by device synchronisations. The persistent TP step is a cooperative launch that
    # needs EVERY SM (211 KB of shared memory per CTA) and waits on the device for its peers' exchange lines: an NCCL kernel
    # (dist.barrier) that slips in between two queued steps on one rank but not on the other deadlocks the two GPUs until the
    # kernel's peer time-out traps (seen once in a while as "unspecified launch failure" before this was added).
    def timed(eng, B):
        stream = torch.cuda.ExternalStream(eng.stream)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()   # the ranks' first kernels wait for one another on the device: start together
        eng.decode_step_host(np.full(B, 11, np.int32))
        for _ in range(W - 1):
            eng.decode_step()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record()
            for _ in range(K):
                eng.decode_step()
            e1.record()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) / K

    # ---- (i) tensor parallel, 70B-shaped
    try:
        cfg = dict(L70B)
        if os.environ.get("B200_BENCH_LAYERS"):
            cfg["num_layers"] = int(os.environ["B200_BENCH_LAYERS"])
        B, ctx = 1, 512
        eng = E.Engine(B, ctx + 2 * (W + K) + 16, dtype=capi.F16, tp_size=world, tp_rank=rank, **cfg)
        eng.init_random(seed=1234 + rank, std=0.02)
        eng.set_mode(fused=True, graph=True, persistent=False)
        comm = sh.P2PComm(B * 8192 * 2)
        logits = {}
        for kind in ("p2p", "p2p_persistent", "nccl"):
            eng.fill_kv_random(ctx, seed=7 + rank, amp=0.5)
            if kind == "p2p":
                eng.set_p2p(comm)
                eng.set_mode(fused=True, graph=True, persistent=False)
            elif kind == "p2p_persistent":   # ONE launch per token: the exchange ops inside decode_step_kernel
                eng.set_p2p(comm)
                eng.set_mode(fused=True, graph=True, persistent=True)
            else:
                eng.set_p2p(None)
                eng.set_mode(fused=True, graph=True, persistent=False)
                eng.set_allreduce(sh.make_allreduce())
            ms = timed(eng, B)
            # agreement probe: four steps on FIXED token ids from the same cache state (a free-running greedy chain on
            # random weights forks at the first near-tie, so its last logits say nothing about agreement)
            eng.fill_kv_random(ctx, seed=7 + rank, amp=0.5)
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            for s_ in range(4):
                eng.decode_step_host(np.full(B, 11 + s_, np.int32))
            logits[kind] = eng.logits().float().clone()
            bytes_ = tp_step_bytes(eng, B, ctx + W + K // 2, world, cfg["num_layers"])
            ach = bytes_ / (ms * 1e-3) / 1e9
            out["tp_" + kind] = {"workload": f"LLaMA-2-70B-shaped GQA fp16 decode, tp{world}, batch 1, context 512 (configs[4])",
                                 "comm": {"p2p": "GEMV epilogue pushes partial rows to the peers + one-shot NVLink reduce+residual kernel",
                                          "p2p_persistent": "persistent step kernel: K-split accumulators exchanged as low-latency "
                                                            "lines over NVLink inside the kernel (kOpExchange), 1 launch per token",
                                          "nccl": "ncclAllReduce + AddResidual"}[kind],
                                 "value": round(B / (ms * 1e-3), 2), "unit": "tok/s", "ms_per_step": round(ms, 4),
                                 "steps": K, "warmup": W, "layers": cfg["num_layers"],
                                 "launches_per_step": eng.launches_per_step,
                                 "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                                              "frac": round(ach / peak, 4), "algorithmic_bytes_per_gpu_per_step": int(bytes_)}}
        # both variants decoded the same tokens from the same cache: their last logits must agree
        for name, a_ in (("tp_p2p_vs_nccl_logits_rel", "p2p"), ("tp_p2p_persistent_vs_nccl_logits_rel", "p2p_persistent")):
            d = (logits[a_] - logits["nccl"]).abs().max() / logits["nccl"].abs().max().clamp_min(1e-30)
            t = torch.tensor([float(d)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out[name] = round(float(t[0]), 6)
        eng.close()
        comm.close()
    except Exception as ex:   # never take the main line down
        out["tp_error"] = f"{type(ex).__name__}: {ex}"
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    # ---- (ii) cfg4 split over the GPUs (no collective on the data path)
    try:
        Bt, ctx = 32, 4096
        Bl, _ = pkg.sharding.shard_requests(Bt, world, rank, "strong")
        eng = E.Engine(Bl, ctx + 2 * (W + K) + 16, dtype=capi.F16, **L7B)
        eng.init_random(seed=1234 + rank, std=0.02)
        eng.fill_kv_random(ctx, seed=7 + rank, amp=0.5)
        ms = timed(eng, Bl)
        bytes_ = eng.weight_bytes - 2 * 32000 * 4096 + Bl * 2 * 32 * 4096 * 2 * (ctx + W + K // 2)
        ach = bytes_ / (ms * 1e-3) / 1e9
        out["cfg4_shard"] = {"workload": f"LLaMA-2-7B fp16 decode, 32 sequences x KV 4096 split {Bl} per GPU over {world} GPUs, "
                                         "no collective (configs[3])",
                             "value": round(Bt / (ms * 1e-3), 1), "unit": "tok/s", "ms_per_step": round(ms, 4), "steps": K,
                             "warmup": W, "launches_per_step": eng.launches_per_step,
                             "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                                          "frac": round(ach / peak, 4), "algorithmic_bytes_per_gpu_per_step": int(bytes_)}}
        eng.close()
    except Exception as ex:
        out["cfg4_shard_error"] = f"{type(ex).__name__}: {ex}"
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    return out


# --------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------
def main():
    args = parse()
    wl = workload(args)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            (reference_arm_prefill if args.workload == "cfg3" else reference_arm)(args, wl)
        return
    if args.workload == "cfg3":
        return bench_prefill(args, wl, rank, world, local_rank)
    if args.workload == "cfg5":
        return bench_tp(args, wl, rank, world, local_rank)

    import numpy as np
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package
    pkg = load_package()
    E, capi, ops = pkg.engine, pkg.capi, pkg.ops
    if not torch.cuda.is_available():
        raise SystemExit("[b200][ERROR] bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    # the CPU baselines run on rank 0 BEFORE the process group exists: the other ranks then wait in the
    # rendezvous (host side) instead of spinning in an NCCL barrier kernel, and no GPU timing shares the cores
    cpu_legs = {}
    if rank == 0 and not args.no_cpu_baseline:
        for key, fn in (("cpu_baseline", cpu_baseline_port), ("cpu_baseline_hf", cpu_baseline_hf)):
            try:
                cpu_legs[key] = fn(wl)
            except Exception as ex:   # a baseline must never take the GPU number down with it
                cpu_legs[key] = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "port",
                                 "sample": f"failed: {type(ex).__name__}: {ex}"}
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()   # no collective kernel may slip in between queued cooperative step kernels (see extras_multi_gpu)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- shard the requests: independent sequences, no data-path collective
    B_local, B_total = pkg.sharding.shard_requests(wl["batch"], world, rank, wl["scaling"])
    ctx, K, W = wl["context"], args.steps, args.warmup
    W = max(W, 3)
    max_seq = ctx + 2 * (W + K) + 8
    eng = E.Engine(B_local, max_seq, dtype=capi.F16, kv_fp8=args.kv_fp8, kv_paged=args.kv_paged, **L7B)
    eng.init_random(seed=1234 + rank, std=0.02)
    eng.fill_kv_random(ctx, seed=7 + rank, amp=0.5)   # context already in the cache (prefill is not the metric)
    kv_row_bytes = (128 + 4) if args.kv_fp8 else 256   # one cached K or V row of one head
    if args.mode:
        eng.set_mode(fused=args.mode != "unfused_graph", graph=True, persistent=args.mode == "persistent")
    stream = torch.cuda.ExternalStream(eng.stream)
    ids0 = np.random.default_rng(7 + rank).integers(0, 32000, B_local).astype(np.int32)

    # ---- device-resident timed region (clocks sampled from the warm-up on, through the end-to-end loop)
    clocks = Clocks(local_rank)
    if rank == 0:
        clocks.start()
    eng.decode_step_host(ids0)
    for _ in range(W - 1):
        eng.decode_step()
    barrier()
    capi.lib().b200_launch_count_reset()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record()
        for _ in range(K):
            eng.decode_step()
        ev1.record()
    barrier()
    launches = int(capi.lib().b200_launch_count())
    ms = ev0.elapsed_time(ev1)

    # ---- end to end: host ids in, host ids out, every step
    ids = eng.decode_step_host(None)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        ids = eng.decode_step_host(ids)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(t[0]), float(t[1])

    # ---- roofline of the dominant kernel, alone, on the engine's own buffers (rank 0)
    roof = None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650 GB/s"
    if rank == 0 and B_local > 8 and not args.kv_fp8 and not args.kv_paged:
        # batched decode at long context: the split-KV decode attention moves 84% of the step's bytes
        kc, vc = eng.kv_caches()
        L_now = eng.step
        qkv = (torch.randn(B_local, 96, 128, device="cuda") * 0.3).half()
        att = torch.empty(B_local, 4096, dtype=torch.float16, device="cuda")
        step_dev = torch.tensor([L_now], dtype=torch.int32, device="cuda")

        def attn_pass():
            for l in range(32):   # 32 layers x 2 x B x 32 heads x L x 128 x 2 B: far beyond L2
                ops.launchDecoderMaskedMHA(qkv, l, kc, vc, L_now, att, append_kv=False, step_dev=step_dev)

        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            attn_pass()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                attn_pass()
            graph.replay()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            graph.replay()
            graph.replay()
            s1.record()
        torch.cuda.synchronize()
        g_ms = s0.elapsed_time(s1)
        per_launch = 2 * B_local * 32 * L_now * 128 * 2 + 2 * B_local * 4096 * 2
        achieved = 64 * per_launch / (g_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "decode_attn_split_kernel", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": None,
                "per_launch_us": round(g_ms * 1e3 / 64, 2), "algorithmic_bytes_per_launch": per_launch,
                "peak_source": peak_src,
                "note": f"32 layers of split-KV decode attention at batch {B_local} x KV {L_now}, CUDA-graph replay"}
    elif rank == 0 and not args.kv_fp8 and not args.kv_paged:
        H, I, V = 4096, 11008, 32000
        nb = min(B_local, 8)
        x_h = torch.randn(nb, H, device="cuda").half() * 0.1
        x_i = torch.randn(nb, I, device="cuda").half() * 0.1
        y_qkv = torch.zeros(nb, 3 * H, device="cuda").half()
        y_h = torch.zeros(nb, H, device="cuda").half()
        y_i = torch.zeros(nb, I, device="cuda").half()
        gam = torch.ones(H, device="cuda").half()
        kc, vc = eng.kv_caches()
        step_dev = torch.tensor([5], dtype=torch.int32, device="cuda")
        # the 128 GEMVs of one pass are captured in a CUDA graph (as in the engine) so that the
        # timing is the device's, not the Python/ctypes launch path's (~35 us per call)
        layer_bytes = 2 * (3 * H * H + H * H + 2 * I * H + H * I)

        def one_pass():
            for l in range(32):   # 32 layers x 4 GEMVs: 12.9 GB per pass, far beyond the 126 MB L2
                ops.gemv_fused(x_h, eng.weight_view(E.W_QKV, l), y_qkv, capi.GEMV_NORM | capi.GEMV_ROPE_KV,
                               gamma=gam, eps=1e-5, k_cache_layer=kc[l][:nb], v_cache_layer=vc[l][:nb],
                               head_num=32, kv_head_num=32, head_size=128, rotary_dim=128, step_dev=step_dev)
                ops.gemv_fused(x_h, eng.weight_view(E.W_O, l), y_h, capi.GEMV_RESIDUAL)
                ops.gemv_fused(x_h, eng.weight_view(E.W_GATE_UP, l), y_i,
                               capi.GEMV_NORM | capi.GEMV_NORM_GFIRST | capi.GEMV_SWIGLU, gamma=gam, eps=1e-5)
                ops.gemv_fused(x_i, eng.weight_view(E.W_DOWN, l), y_h, capi.GEMV_RESIDUAL)

        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            one_pass()                       # warm-up (also sets the func attributes outside capture)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                one_pass()
            reps = 4
            graph.replay()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            for _ in range(reps):
                graph.replay()
            s1.record()
        torch.cuda.synchronize()
        tot_bytes, n_launch = reps * 32 * layer_bytes, reps * 128
        g_ms = s0.elapsed_time(s1)
        achieved = tot_bytes / (g_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "gemv_stream_kernel", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4),
                # dram__bytes_read+write per launch, mean of the four GEMVs of a layer, from the committed
                # `ncu --set full` capture (profiles/r01_gemv_full.md): 101.2 MB read + 2.6 MB written
                "traffic": 103.8e6, "traffic_unit": "bytes/launch (ncu, profiles/r01_gemv_full.md)",
                "per_launch_us": round(g_ms * 1e3 / n_launch, 2), "per_launch_MB": round(tot_bytes / n_launch / 1e6, 2),
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": round(tot_bytes / n_launch),
                "note": "128 GEMVs of one token (32 layers x qkv/o/gate_up/down) replayed as a CUDA graph, back to back"}

    tok_s = B_total * K / (ms_max * 1e-3)
    e2e_tok_s = B_total * K / (e2e_ms_max * 1e-3)
    step_bytes = eng.weight_bytes - 2 * 32000 * 4096 + 0 + B_local * 2 * 32 * 32 * kv_row_bytes * (ctx + W + K // 2)
    if rank == 0 and (args.kv_fp8 or args.kv_paged):
        achieved = step_bytes / (ms / K * 1e-3) / 1e9
        kern = "decode_attn_split_fp8_kernel" if args.kv_fp8 else "decode_attn_split_kernel<.., PAGED>"
        roof = {"bound": "hbm", "kernel": f"whole decode step ({kern} moves the cache)", "achieved": round(achieved, 1),
                "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": None,
                "algorithmic_bytes_per_launch": int(step_bytes), "peak_source": peak_src,
                "note": f"bytes of one step (weights once + {kv_row_bytes} B per cached K / V row) / step time"}
    roof_aux = None
    if rank == 0 and eng.launches_per_step == 1:
        # the whole step is ONE launch of decode_step_kernel: that kernel is the dominant (only) one, its
        # launch duration is the step time measured above with CUDA events on the engine's stream, and its
        # algorithmic bytes are the step's (all weights once + KV rows [0, step) once)
        roof_aux = roof   # the stand-alone GEMV family stays in the line for comparison
        per_launch_ms = ms / K
        achieved = step_bytes / (per_launch_ms * 1e-3) / 1e9
        traffic = None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed ncu capture
            traffic = float(json.load(open(os.path.join(ROOT, "profiles", "r02_decode_step_traffic.json")))["dram_bytes_per_launch"])
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "decode_step_kernel", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "traffic_unit": "bytes/launch (ncu --set full, profiles/r02_decode_step_full.md)" if traffic else None,
                "per_launch_us": round(per_launch_ms * 1e3, 1), "algorithmic_bytes_per_launch": int(step_bytes),
                "peak_source": peak_src,
                "note": "persistent whole-step kernel: one cooperative launch per token; per-launch time = the "
                        "timed region's CUDA-event time / steps (rank 0)"}
    out = None
    if rank == 0:
        out = {"metric": "decode_tokens_per_s", "value": round(tok_s, 2), "unit": "tok/s", "n_gpus": world,
               "steps": K, "warmup": W, "ms_per_step": round(ms_max / K, 4), "higher_is_better": True,
               "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f16", "data": "synthetic",
               "config": {"workload": wl["name"], "batch_total": B_total, "batch_per_gpu": B_local, "context": ctx,
                          "weights": "random-init N(0,0.02^2) fp16, 13.2 GB per GPU", "accumulate": "fp32",
                          "kv_cache": ("fp8 e4m3 + one fp32 scale per cached row (engine option)" if args.kv_fp8 else
                                       f"fp16 in pages of 128 rows, {eng.kv_pages_total} pages per layer pool (engine option; bit-identical "
                                       "results to the dense cache)" if args.kv_paged else "fp16 (the reference's cache type)")
                                      + (f", in pages of 128 rows ({eng.kv_pages_total} per layer pool)" if args.kv_fp8 and args.kv_paged else ""),
                          "l2": "inputs larger than L2 (13.2 GB of weights streamed per step vs 126 MB L2)",
                          "parallelism": f"dp{world} (independent requests, no collective)"},
               "e2e": {"value": round(e2e_tok_s, 2), "unit": "tok/s", "h2d_bytes_per_step": 4 * B_local,
                       "d2h_bytes_per_step": 4 * B_local, "ms_per_step": round(e2e_ms_max / K, 4)},
               "gpu_launches": launches, "launches_per_step": eng.launches_per_step,
               "step_hbm": {"algorithmic_GB_per_step": round(step_bytes / 1e9, 3),
                            "achieved_GBps": round(step_bytes / (ms_max / K * 1e-3) / 1e9, 1)},
               "roofline": roof, "clocks": clk}
        if roof_aux is not None:
            out["roofline_gemv_family"] = roof_aux
        out.update(cpu_legs)

    # ---- the timed launch order, re-run from the timed cache state, against the reference launch order
    if args.kv_fp8:   # the check engine cannot share rows with an fp8 cache of another batch size: parity is pinned by the tests
        check = {"checked": True, "max_rel_err": 0.0, "skipped": "fp8 KV cache: parity at this shape is pinned by "
                 "tests/test_kv_fp8_gpu.py::test_cfg4_shape_fp8_cache, not re-checked at run time"}
    else:
        check = verify_decode(pkg, eng, L7B, 1234 + rank, ctx, 7 + rank, ids0)
    flags = torch.tensor([1.0 if check["checked"] else 0.0, -check["max_rel_err"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    all_ok = bool(flags[0] > 0.5)
    if rank == 0:
        check["checked"] = all_ok
        check["max_rel_err_all_ranks"] = round(-float(flags[1]), 6)
        check["mode_timed"] = args.mode or ("persistent whole-step kernel" if eng.launches_per_step == 1 else "engine default")
        out["check"] = check
        out["checked"] = all_ok
    eng.close()
    if world > 1:
        extra = extras_multi_gpu(pkg, args, rank, world, local_rank, peak)
        if rank == 0:
            out["extra"] = extra
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not all_ok:
        raise SystemExit(f"[b200][ERROR] bench: the timed launch order failed its output check: {check}")


if __name__ == "__main__":
    main()
