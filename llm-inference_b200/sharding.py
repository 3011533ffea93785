"""Host-side sharding rules of the decode path across GPUs (the reference itself is
single-GPU, src/models/llama/llama.cpp has no communicator).

Two ways, one process per GPU in both:
  * independent requests: weights replicated, sequences partitioned, NO data-path collective
    (`shard_requests`); timing is the max over ranks (`max_over_ranks`).
  * tensor parallel (Megatron split of one decoder layer, the weight layout of
    src/weights/llama/layer_weights.cc:21-47): qkv and gate_up are split by output rows (each rank
    owns head_num/tp query heads, kv_head_num/tp kv heads, inter_size/tp FFN columns), o_proj and
    down_proj by input columns, so each rank produces a partial [B,H] sum that is all-reduced after
    o_proj and after down_proj only (`tp_shard_layer`, `make_allreduce`). Norm weights, embedding
    and LM head are replicated.
No torch import at module load: the CPU tests import this without a GPU."""
import numpy as np


def shard_requests(batch, world, rank, scaling):
    """-> (sequences this rank decodes, sequences of the whole job).
    weak: every rank decodes `batch` sequences of its own (N independent requests per GPU);
    strong: the `batch` sequences of the job are split evenly, first ranks take the remainder."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    if scaling == "weak":
        return batch, batch * world
    if scaling != "strong":
        raise ValueError(f"scaling must be weak or strong, not {scaling!r}")
    if batch < world:
        raise ValueError(f"{batch} sequences cannot be split over {world} ranks")
    base, rem = divmod(batch, world)
    return base + (1 if rank < rem else 0), batch


def request_slice(batch, world, rank):
    """half-open range of the job's sequences that `rank` owns under strong scaling"""
    base, rem = divmod(batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def tp_local_cfg(cfg, tp):
    """per-rank shape of a layer: heads and FFN columns divided, hidden size kept"""
    for key in ("head_num", "kv_head_num", "inter_size"):
        if cfg[key] % tp:
            raise ValueError(f"{key}={cfg[key]} is not divisible by tp={tp}")
    out = dict(cfg)
    out["hidden_size"] = cfg["head_num"] * cfg["head_size"]
    out["head_num"] //= tp
    out["kv_head_num"] //= tp
    out["inter_size"] //= tp
    return out


def tp_shard_layer(w, cfg, tp, rank):
    """w: dict attn_norm [H], qkv [(h+2hk)d, H], o [H, h d], ffn_norm [H], gate_up [2I, H],
    down [H, I] of the unsharded layer (cfg holds the GLOBAL head counts) -> this rank's dict."""
    h, hk, d, I = cfg["head_num"], cfg["kv_head_num"], cfg["head_size"], cfg["inter_size"]
    tp_local_cfg(cfg, tp)
    hl, hkl, Il = h // tp, hk // tp, I // tp
    H = w["qkv"].shape[1]
    qkv = np.asarray(w["qkv"]).reshape(h + 2 * hk, d, H)
    q = qkv[rank * hl:(rank + 1) * hl]
    k = qkv[h + rank * hkl:h + (rank + 1) * hkl]
    v = qkv[h + hk + rank * hkl:h + hk + (rank + 1) * hkl]
    gu = np.asarray(w["gate_up"]).reshape(2, I, H)
    return dict(
        attn_norm=np.ascontiguousarray(w["attn_norm"]),
        ffn_norm=np.ascontiguousarray(w["ffn_norm"]),
        qkv=np.ascontiguousarray(np.concatenate([q, k, v]).reshape((hl + 2 * hkl) * d, H)),
        o=np.ascontiguousarray(np.asarray(w["o"])[:, rank * hl * d:(rank + 1) * hl * d]),
        gate_up=np.ascontiguousarray(gu[:, rank * Il:(rank + 1) * Il].reshape(2 * Il, H)),
        down=np.ascontiguousarray(np.asarray(w["down"])[:, rank * Il:(rank + 1) * Il]),
    )


def max_over_ranks(ms, device=None):
    """the job's step time is the slowest rank's (bench.py contract); no-op without a process group"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(ms)
    t = torch.tensor([float(ms)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def make_allreduce(group=None):
    """-> callable(buf_ptr, count, dtype, stream_ptr) for Engine.set_allreduce: in-place sum of the
    engine's partial [B,H] buffer over the ranks of `group`.
    nccl backend: the collective is enqueued on the engine's own stream (capturable in the step's
    CUDA graph). gloo backend (CPU tests, or two ranks sharing one GPU): the buffer is staged
    through host memory after a stream sync — only valid with graph capture off."""
    import torch
    import torch.distributed as dist
    from . import capi
    backend = dist.get_backend(group)

    def view(ptr, count, dtype):
        td = torch.float16 if dtype == capi.F16 else torch.float32

        class _Mem:
            pass
        m = _Mem()
        m.__cuda_array_interface__ = {"shape": (count * (2 if dtype == capi.F16 else 4),), "typestr": "|u1",
                                      "data": (int(ptr), False), "version": 2}
        return torch.as_tensor(m, device="cuda").view(td)

    def allreduce(ptr, count, dtype, stream_ptr):
        t = view(ptr, count, dtype)
        ext = torch.cuda.ExternalStream(int(stream_ptr)) if stream_ptr else torch.cuda.current_stream()
        with torch.cuda.stream(ext):
            if backend == "nccl":
                dist.all_reduce(t, group=group)
            else:
                ext.synchronize()
                # the sum is formed in fp32 on the host and stored back in the buffer's type, the
                # same result a two-rank fp16 ring produces (one rounding of a two-term sum)
                host = t.float().cpu()
                dist.all_reduce(host, group=group)
                t.copy_(host.to(t.dtype))
                ext.synchronize()
        return 0

    return allreduce


class P2PComm:
    """One-shot NVLink allreduce communicator (include/b200_engine.h b200_p2p_*, csrc/p2p_allreduce.cu):
    one per process / GPU. The 64-byte cudaIpc handles travel through torch.distributed
    (all_gather_object); after that no library is involved in the collective — the kernel pushes
    partials into the peers' mapped buffers over NVLink and sums locally."""

    def __init__(self, max_message_bytes, group=None):
        import ctypes as C
        import torch.distributed as dist
        from . import engine as E
        self.lib = E._lib()
        self.world, self.rank, self.group = dist.get_world_size(group), dist.get_rank(group), group
        self.handle = self.lib.b200_p2p_create(self.world, self.rank, int(max_message_bytes))
        if not self.handle:
            raise RuntimeError("[b200][ERROR] b200_p2p_create failed")
        n = int(self.lib.b200_p2p_handle_bytes())
        mine = C.create_string_buffer(n)
        rc = self.lib.b200_p2p_local_handle(self.handle, mine)
        if rc != 0:
            raise RuntimeError(f"[b200][ERROR] b200_p2p_local_handle: {rc}")
        gathered = [None] * self.world
        dist.all_gather_object(gathered, bytes(mine.raw), group=group)
        rc = self.lib.b200_p2p_connect(self.handle, b"".join(gathered))
        if rc != 0:
            raise RuntimeError(f"[b200][ERROR] b200_p2p_connect: CUDA error {rc} (peer access over cudaIpc unavailable?)")
        dist.barrier(group=group)

    def allreduce_add(self, partial, hidden):
        """hidden += sum over ranks of partial (torch CUDA tensors, same dtype/shape), on torch's current stream"""
        import ctypes as C
        import torch
        from . import capi
        dt = capi.F16 if partial.dtype == torch.float16 else capi.F32
        rc = self.lib.b200_p2p_allreduce_add(self.handle, C.c_void_p(partial.data_ptr()), C.c_void_p(hidden.data_ptr()),
                                             partial.numel(), dt, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise RuntimeError(f"[b200][ERROR] b200_p2p_allreduce_add: {rc}")

    def allreduce(self, buf):
        import ctypes as C
        import torch
        from . import capi
        dt = capi.F16 if buf.dtype == torch.float16 else capi.F32
        rc = self.lib.b200_p2p_allreduce(self.handle, C.c_void_p(buf.data_ptr()), buf.numel(), dt,
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise RuntimeError(f"[b200][ERROR] b200_p2p_allreduce: {rc}")

    def close(self):
        """quiesce, then free: every rank synchronises its device and meets the others at a barrier before any of them
        unmaps / frees the receive buffers (b200_p2p_destroy's contract: a lagging peer must not push into freed memory)"""
        if self.handle:
            import torch
            import torch.distributed as dist
            torch.cuda.synchronize()
            if dist.is_available() and dist.is_initialized() and self.world > 1:
                dist.barrier(group=self.group)
            self.lib.b200_p2p_destroy(self.handle)
            self.handle = None
