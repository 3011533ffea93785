"""ctypes front end of include/b200_engine.h (the decode-step driver). Plumbing only."""
import ctypes as C

import numpy as np

from . import capi

W_ATTN_NORM, W_QKV, W_O, W_FFN_NORM, W_GATE_UP, W_DOWN, W_EMBED, W_FINAL_NORM, W_LM_HEAD = range(9)
LAYER_SLOTS = (W_ATTN_NORM, W_QKV, W_O, W_FFN_NORM, W_GATE_UP, W_DOWN)
MODE_FUSED, MODE_GRAPH, MODE_PERSISTENT = 1, 2, 4
KV_FP8 = 1   # creation flag B200_ENGINE_KV_FP8
KV_PAGED = 2   # creation flag B200_ENGINE_KV_PAGED


class ModelConfig(C.Structure):
    _fields_ = [("num_layers", C.c_int), ("head_num", C.c_int), ("kv_head_num", C.c_int),
                ("head_size", C.c_int), ("inter_size", C.c_int), ("vocab_size", C.c_int),
                ("max_seq_len", C.c_int), ("max_batch", C.c_int), ("rms_eps", C.c_float),
                ("rotary_dim", C.c_int), ("rotary_base", C.c_float), ("dtype", C.c_int),
                ("tp_size", C.c_int), ("tp_rank", C.c_int)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p)

_vp, _i = C.c_void_p, C.c_int
ENGINE_PROTOTYPES = {
    "b200_engine_create": (_vp, [C.POINTER(ModelConfig), _i]),
    "b200_engine_create_ex": (_vp, [C.POINTER(ModelConfig), _i, C.c_uint]),
    "b200_engine_create_paged": (_vp, [C.POINTER(ModelConfig), _i, C.c_uint, _i]),
    "b200_engine_reset_slot": (_i, [_vp, _i]),
    "b200_engine_kv_is_paged": (_i, [_vp]),
    "b200_engine_kv_pages_total": (_i, [_vp]),
    "b200_engine_kv_pages_free": (_i, [_vp]),
    "b200_engine_kv_page_table": (_vp, [_vp]),
    "b200_engine_release_finished": (_i, [_vp]),
    "b200_engine_kv_gather": (_i, [_vp, _vp, _vp]),
    "b200_engine_kv_is_fp8": (_i, [_vp]),
    "b200_engine_set_eos": (_i, [_vp, _i]),
    "b200_engine_get_finished": (_i, [_vp, _vp]),
    "b200_engine_kv_dequant": (_i, [_vp, _vp, _vp]),
    "b200_engine_destroy": (None, [_vp]),
    "b200_engine_last_error": (C.c_char_p, []),
    "b200_engine_weight_bytes": (C.c_size_t, [_vp]),
    "b200_engine_kv_bytes": (C.c_size_t, [_vp]),
    "b200_engine_weight_elems": (C.c_size_t, [_vp, _i]),
    "b200_engine_set_mode": (_i, [_vp, C.c_uint]),
    "b200_engine_set_allreduce": (_i, [_vp, ALLREDUCE_FN, _vp]),
    "b200_engine_stream": (_vp, [_vp]),
    "b200_engine_init_random": (_i, [_vp, C.c_uint64, C.c_float]),
    "b200_engine_set_weight": (_i, [_vp, _i, _i, _vp, C.c_size_t]),
    "b200_engine_get_weight": (_i, [_vp, _i, _i, _vp, C.c_size_t]),
    "b200_engine_fill_kv_random": (_i, [_vp, _i, C.c_uint64, C.c_float]),
    "b200_engine_reset": (_i, [_vp]),
    "b200_engine_get_step": (_i, [_vp]),
    "b200_engine_decode_step": (_i, [_vp, _vp]),
    "b200_engine_decode_step_host": (_i, [_vp, _vp, _vp]),
    "b200_engine_logits": (_vp, [_vp]),
    "b200_engine_ids_out": (_vp, [_vp]),
    "b200_engine_hidden": (_vp, [_vp]),
    "b200_engine_k_cache": (_vp, [_vp]),
    "b200_engine_v_cache": (_vp, [_vp]),
    "b200_engine_launches_per_step": (_i, [_vp]),
    "b200_engine_step_trace": (_i, [_vp, _vp, _i]),
    "b200_engine_prefill": (_i, [_vp, _vp, _i, _vp]),
    "b200_engine_prefill_ragged": (_i, [_vp, _vp, _vp, _vp]),
    "b200_engine_get_seq_steps": (_i, [_vp, _vp]),
    "b200_engine_set_p2p": (_i, [_vp, _vp]),
    "b200_engine_generate": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "b200_engine_generate_ragged": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp]),
    "b200_engine_load_weights": (_i, [_vp, C.c_char_p]),
    "b200_p2p_handle_bytes": (C.c_size_t, []),
    "b200_p2p_create": (_vp, [_i, _i, C.c_size_t]),
    "b200_p2p_local_handle": (_i, [_vp, _vp]),
    "b200_p2p_connect": (_i, [_vp, _vp]),
    "b200_p2p_allreduce_add": (_i, [_vp, _vp, _vp, C.c_size_t, _i, _vp]),
    "b200_p2p_allreduce": (_i, [_vp, _vp, C.c_size_t, _i, _vp]),
    "b200_p2p_reduce_add": (_i, [_vp, _vp, C.c_size_t, _i, _vp]),
    "b200_p2p_max_message_bytes": (C.c_size_t, [_vp]),
    "b200_p2p_destroy": (None, [_vp]),
    "b200_engine_prefill_launches": (_i, [_vp]),
    "b200_engine_weight_ptr": (_vp, [_vp, _i, _i]),
}

_bound = False


def _lib():
    global _bound
    lib = capi.lib()
    if not _bound:
        for name, (res, args) in ENGINE_PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _bound = True
    return lib


LLAMA2_7B = dict(num_layers=32, head_num=32, kv_head_num=32, head_size=128, inter_size=11008,
                 vocab_size=32000)
LLAMA2_70B = dict(num_layers=80, head_num=64, kv_head_num=8, head_size=128, inter_size=28672,
                  vocab_size=32000)


class Engine:
    """Owns one b200_engine on the current CUDA device."""

    def __init__(self, batch, max_seq_len, dtype=capi.F16, rms_eps=1e-5, rotary_base=10000.0,
                 rotary_dim=None, tp_size=1, tp_rank=0, kv_fp8=False, kv_paged=False, kv_pages=0, **shape):
        cfg = ModelConfig()
        for k, v in shape.items():
            setattr(cfg, k, v)
        cfg.max_seq_len, cfg.max_batch = max_seq_len, batch
        cfg.rms_eps, cfg.rotary_base = rms_eps, rotary_base
        cfg.rotary_dim = rotary_dim if rotary_dim is not None else cfg.head_size
        cfg.dtype, cfg.tp_size, cfg.tp_rank = dtype, tp_size, tp_rank
        self.cfg, self.batch, self.dtype = cfg, batch, dtype
        self.np_dtype = np.float16 if dtype == capi.F16 else np.float32
        self.lib = _lib()
        self.kv_fp8 = bool(kv_fp8)
        self.kv_paged = bool(kv_paged)
        flags = (KV_FP8 if kv_fp8 else 0) | (KV_PAGED if kv_paged else 0)
        self.h = self.lib.b200_engine_create_paged(C.byref(cfg), batch, flags, int(kv_pages))
        if not self.h:
            raise RuntimeError("[b200][ERROR] engine_create: " + self.lib.b200_engine_last_error().decode())
        self._cb = None

    def close(self):
        if self.h:
            self.lib.b200_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"[b200][ERROR] {what}: rc={rc} {self.lib.b200_engine_last_error().decode()}")

    # ---- properties
    @property
    def weight_bytes(self):
        return self.lib.b200_engine_weight_bytes(self.h)

    @property
    def kv_bytes(self):
        return self.lib.b200_engine_kv_bytes(self.h)

    @property
    def step(self):
        return self.lib.b200_engine_get_step(self.h)

    @property
    def launches_per_step(self):
        return self.lib.b200_engine_launches_per_step(self.h)

    @property
    def stream(self):
        return self.lib.b200_engine_stream(self.h)

    # ---- control
    def step_trace(self):
        """numpy uint64 [n_ops, 4]: %globaltimer stamps (op done, barrier passed, x staged) + ring stages landed at GEMV
        start, or None (B200_STEP_TRACE=1). Also sets self.trace_starved [2, n_ops] (cycles warp 0 of the first / last CTA
        waited for weights) and self.trace_done [n_ops, grid] (op-done stamp of every CTA)."""
        import torch
        grid = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
        kmax = 5 * self.cfg.num_layers + 2
        buf = np.zeros(6 * kmax + 1 + kmax * grid, np.uint64)
        n = self.lib.b200_engine_step_trace(self.h, buf.ctypes.data_as(C.c_void_p), buf.size)
        if not n:
            return None
        k = (n - 1) // (6 + grid)   # ops of the program: 4 per layer + 2 with the fused FFN op, else 5 per layer + 2
        self.prefetch_stages = int(buf[4 * k])
        self.trace_starved = buf[4 * k + 1:6 * k + 1].reshape(2, k).astype(np.int64)
        self.trace_done = buf[6 * k + 1:6 * k + 1 + k * grid].reshape(k, grid).astype(np.int64)
        return np.concatenate([buf[:2 * k].reshape(k, 2), buf[2 * k:3 * k].reshape(k, 1),
                               buf[3 * k:4 * k].reshape(k, 1)], axis=1)

    def set_mode(self, fused=True, graph=True, persistent=False):
        """A fresh engine runs the persistent whole-step kernel for fp16 batches 1-2 and the CUDA-graphed
        fused-GEMV sequence for fp16 batches >= 3 (the measured best, engine.cu b200_engine_create);
        an explicit set_mode() selects the multi-kernel paths unless persistent=True (which applies up to batch 4)."""
        flags = (MODE_FUSED if fused else 0) | (MODE_GRAPH if graph else 0) | (MODE_PERSISTENT if persistent else 0)
        self._ck(self.lib.b200_engine_set_mode(self.h, flags), "set_mode")

    def set_p2p(self, comm):
        """comm: sharding.P2PComm (or None): the fused one-shot NVLink allreduce + residual add"""
        self._p2p = comm
        self._ck(self.lib.b200_engine_set_p2p(self.h, comm.handle if comm is not None else None), "set_p2p")

    def set_allreduce(self, pyfunc):
        """pyfunc(buf_ptr:int, count:int, dtype:int, stream_ptr:int) -> int"""
        self._cb = ALLREDUCE_FN(lambda buf, n, dt, st, user: int(pyfunc(buf, n, dt, st) or 0))
        self._ck(self.lib.b200_engine_set_allreduce(self.h, self._cb, None), "set_allreduce")

    def set_eos(self, eos_id):
        """sequence retirement (the reference's `finished` flags): -1 switches it off"""
        self._ck(self.lib.b200_engine_set_eos(self.h, int(eos_id)), "set_eos")

    def finished(self):
        out = np.zeros(self.batch, np.uint8)
        self._ck(self.lib.b200_engine_get_finished(self.h, out.ctypes.data_as(C.c_void_p)), "get_finished")
        return out.astype(bool)

    def init_random(self, seed=1234, std=0.02):
        self._ck(self.lib.b200_engine_init_random(self.h, seed, std), "init_random")

    def weight_shape(self, slot):
        c = self.cfg
        tp = c.tp_size
        H, h, hk, d, I, V = c.head_num * c.head_size, c.head_num // tp, c.kv_head_num // tp, c.head_size, c.inter_size // tp, c.vocab_size
        return {W_ATTN_NORM: (H,), W_FFN_NORM: (H,), W_FINAL_NORM: (H,), W_QKV: ((h + 2 * hk) * d, H),
                W_O: (H, h * d), W_GATE_UP: (2 * I, H), W_DOWN: (H, I), W_EMBED: (V, H), W_LM_HEAD: (V, H)}[slot]

    def set_weight(self, slot, layer, array):
        a = np.ascontiguousarray(array, dtype=self.np_dtype)
        assert a.shape == self.weight_shape(slot), (a.shape, self.weight_shape(slot))
        self._ck(self.lib.b200_engine_set_weight(self.h, slot, layer, a.ctypes.data_as(C.c_void_p), a.nbytes), "set_weight")

    def get_weight(self, slot, layer=0):
        a = np.empty(self.weight_shape(slot), self.np_dtype)
        self._ck(self.lib.b200_engine_get_weight(self.h, slot, layer, a.ctypes.data_as(C.c_void_p), a.nbytes), "get_weight")
        return a

    def fill_kv_random(self, length, seed=7, amp=0.5):
        self._ck(self.lib.b200_engine_fill_kv_random(self.h, length, seed, amp), "fill_kv_random")

    def reset(self):
        self._ck(self.lib.b200_engine_reset(self.h), "reset")

    # ---- stepping
    def decode_step(self, ids_dev_ptr=None):
        self._ck(self.lib.b200_engine_decode_step(self.h, ids_dev_ptr), "decode_step")

    def decode_step_host(self, ids_in):
        """numpy int32 [batch] (or None to continue from the previous output) -> numpy int32 [batch]"""
        out = np.empty(self.batch, np.int32)
        inp = None
        if ids_in is not None:
            ids_in = np.ascontiguousarray(ids_in, dtype=np.int32)
            inp = ids_in.ctypes.data_as(C.c_void_p)
        self._ck(self.lib.b200_engine_decode_step_host(self.h, inp, out.ctypes.data_as(C.c_void_p)), "decode_step_host")
        return out

    def prefill(self, ids, sync=True):
        """numpy int32 [batch, len] prompt tokens -> numpy int32 [batch] first generated tokens
        (b200_engine_prefill; sync=False skips the device->host read and returns None)"""
        ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(self.batch, -1)
        out = np.empty(self.batch, np.int32) if sync else None
        self._ck(self.lib.b200_engine_prefill(self.h, ids.ctypes.data_as(C.c_void_p), ids.shape[1],
                                              out.ctypes.data_as(C.c_void_p) if sync else None), "prefill")
        return out

    def prefill_ragged(self, prompts, sync=True):
        """list of `batch` int sequences of different lengths -> numpy int32 [batch] first generated tokens
        (b200_engine_prefill_ragged: ids packed back to back, per-sequence steps from here on)"""
        assert len(prompts) == self.batch
        lens = np.array([len(p) for p in prompts], np.int32)
        ids = np.ascontiguousarray(np.concatenate([np.asarray(p, np.int32).reshape(-1) for p in prompts]))
        out = np.empty(self.batch, np.int32) if sync else None
        self._ck(self.lib.b200_engine_prefill_ragged(self.h, ids.ctypes.data_as(C.c_void_p), lens.ctypes.data_as(C.c_void_p),
                                                     out.ctypes.data_as(C.c_void_p) if sync else None), "prefill_ragged")
        return out

    @property
    def seq_steps(self):
        """tokens cached per sequence (numpy int32 [batch])"""
        out = np.empty(self.batch, np.int32)
        self._ck(self.lib.b200_engine_get_seq_steps(self.h, out.ctypes.data_as(C.c_void_p)), "get_seq_steps")
        return out

    def generate(self, prompt_ids, max_new_tokens, eos_id=2):
        """b200_engine_generate: prompt [batch, len] -> (ids [batch, max_new_tokens] padded with eos_id, counts [batch])"""
        ids = np.ascontiguousarray(prompt_ids, dtype=np.int32).reshape(self.batch, -1)
        out = np.empty((self.batch, max_new_tokens), np.int32)
        cnt = np.empty(self.batch, np.int32)
        rc = self.lib.b200_engine_generate(self.h, ids.ctypes.data_as(C.c_void_p), ids.shape[1], max_new_tokens, eos_id,
                                           out.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p))
        if rc < 0:
            raise RuntimeError(f"[b200][ERROR] generate: {rc} " + self.lib.b200_engine_last_error().decode())
        return out, cnt

    def generate_ragged(self, prompts, max_new_tokens, eos_id=2):
        """b200_engine_generate_ragged: `batch` prompts of different lengths -> (ids [batch, max_new_tokens], counts)"""
        assert len(prompts) == self.batch
        lens = np.array([len(p) for p in prompts], np.int32)
        ids = np.ascontiguousarray(np.concatenate([np.asarray(p, np.int32).reshape(-1) for p in prompts]))
        out = np.empty((self.batch, max_new_tokens), np.int32)
        cnt = np.empty(self.batch, np.int32)
        rc = self.lib.b200_engine_generate_ragged(self.h, ids.ctypes.data_as(C.c_void_p), lens.ctypes.data_as(C.c_void_p),
                                                  max_new_tokens, eos_id, out.ctypes.data_as(C.c_void_p),
                                                  cnt.ctypes.data_as(C.c_void_p))
        if rc < 0:
            raise RuntimeError(f"[b200][ERROR] generate_ragged: {rc} " + self.lib.b200_engine_last_error().decode())
        return out, cnt

    def load_weights(self, directory):
        """the reference's weight directory (fp32 .bin files, llama_weights.cc:41-53 naming)"""
        self._ck(self.lib.b200_engine_load_weights(self.h, str(directory).encode()), "load_weights")

    @property
    def prefill_launches(self):
        return int(self.lib.b200_engine_prefill_launches(self.h))

    def _view(self, ptr, shape, torch_dtype):
        """zero-copy torch view of engine-owned device memory (tests / bench only)"""
        import torch
        n = int(np.prod(shape))
        itemsize = torch.tensor([], dtype=torch_dtype).element_size()

        class _Mem:
            pass
        m = _Mem()
        m.__cuda_array_interface__ = {"shape": (n * itemsize,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
        return torch.as_tensor(m, device="cuda").view(torch_dtype).view(*shape)

    def logits(self):
        import torch
        td = torch.float16 if self.dtype == capi.F16 else torch.float32
        return self._view(self.lib.b200_engine_logits(self.h), (self.batch, self.cfg.vocab_size), td)

    def hidden(self):
        import torch
        td = torch.float16 if self.dtype == capi.F16 else torch.float32
        return self._view(self.lib.b200_engine_hidden(self.h), (self.batch, self.cfg.head_num * self.cfg.head_size), td)

    def weight_view(self, slot, layer=0):
        """torch view of one weight tensor in the engine's arena"""
        import torch
        td = torch.float16 if self.dtype == capi.F16 else torch.float32
        return self._view(self.lib.b200_engine_weight_ptr(self.h, slot, layer), self.weight_shape(slot), td)

    def kv_caches(self):
        """torch views [layers, batch, kv_heads(local), max_seq, d] of K and V (fp8 cache: dequantised fp16 COPIES)"""
        import torch
        td = torch.float16 if self.dtype == capi.F16 else torch.float32
        c = self.cfg
        shape = (c.num_layers, self.batch, c.kv_head_num // c.tp_size, c.max_seq_len, c.head_size)
        if self.kv_paged:   # gathered COPIES laid out as the dense cache; rows beyond a sequence's length are zero
            k = torch.zeros(shape, dtype=td, device="cuda")
            v = torch.zeros(shape, dtype=td, device="cuda")
            # the zero fill runs on torch's stream, the gather on the engine's (non-blocking) stream: without this the fill can
            # land on top of rows the gather has already written (seen in the full test suite, where the block comes out of
            # torch's cache and the fill is still queued when the gather starts)
            torch.cuda.synchronize()
            self._ck(self.lib.b200_engine_kv_gather(self.h, C.c_void_p(k.data_ptr()), C.c_void_p(v.data_ptr())), "kv_gather")
            return k, v
        if self.kv_fp8:
            k = torch.empty(shape, dtype=torch.float16, device="cuda")
            v = torch.empty(shape, dtype=torch.float16, device="cuda")
            self._ck(self.lib.b200_engine_kv_dequant(self.h, C.c_void_p(k.data_ptr()), C.c_void_p(v.data_ptr())), "kv_dequant")
            return k, v
        return (self._view(self.lib.b200_engine_k_cache(self.h), shape, td),
                self._view(self.lib.b200_engine_v_cache(self.h), shape, td))

    # ---- paged cache
    @property
    def kv_pages_total(self):
        return self.lib.b200_engine_kv_pages_total(self.h)

    @property
    def kv_pages_free(self):
        return self.lib.b200_engine_kv_pages_free(self.h)

    def kv_page_table(self):
        """torch view [batch, ceil(max_seq_len / 128)] int32 of the DEVICE page lists"""
        import torch
        n = (self.cfg.max_seq_len + 127) // 128
        return self._view(self.lib.b200_engine_kv_page_table(self.h), (self.batch, n), torch.int32)

    def kv_pools(self):
        """torch views [layers, n_pages, kv_heads(local), 128, d] of the K and V page pools"""
        import torch
        c = self.cfg
        shape = (c.num_layers, self.kv_pages_total, c.kv_head_num // c.tp_size, 128, c.head_size)
        dt = torch.uint8 if self.kv_fp8 else torch.float16   # fp8 pages: raw e4m3 bytes (scales are not exposed here)
        return (self._view(self.lib.b200_engine_k_cache(self.h), shape, dt),
                self._view(self.lib.b200_engine_v_cache(self.h), shape, dt))

    def reset_slot(self, slot):
        """one sequence slot back to 'nothing cached' while the others keep theirs (then prefill_ragged with [] for the others)"""
        self._ck(self.lib.b200_engine_reset_slot(self.h, int(slot)), "reset_slot")

    def release_finished(self):
        """pages of the sequences the device has retired go back to the pool; returns the number released"""
        n = self.lib.b200_engine_release_finished(self.h)
        if n < 0:
            self._ck(n, "release_finished")
        return n

    def ids_out(self):
        import torch
        return self._view(self.lib.b200_engine_ids_out(self.h), (self.batch,), torch.int32)

    def sync(self):
        import torch
        torch.cuda.synchronize()
