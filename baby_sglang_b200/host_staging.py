"""Host-buffer entry to the attention path: q/k/v arrive in pinned host memory, the result is read
back on the host.

The reference keeps every activation on one device and has no such call; this is the plugin-style
entry a caller with HOST buffers uses (bench.py's `e2e` leg).  It drives the same drop-in
`SimpleAttentionBackend.forward` (attn_backend.py:8-32) and only adds the plumbing around it:

  copy-in stream   qkv of layer l+1 is uploaded while layer l computes (two staging buffers)
  compute stream   the caller's current stream: K1 kv_store + K2/K2c (or K3) per layer
  copy-out stream  layer l's output is downloaded while layer l+1 computes

so the PCIe transfers in both directions overlap the HBM-bound kernels instead of serialising
with them.  No CPU fallback: CUDA tensors and pinned host buffers are required.
"""
from __future__ import annotations

import os
import sys
from typing import List, Sequence

import torch

from . import DROPIN_DIR


class HostStagedAttention:
    """Runs `n_layers` attention layers of one forward batch from pinned host q/k/v.

    `qkv_host[l]` is the fused projection output `[T, (Hq + 2*Hkv) * D]` (qwen2.py:189) of layer l,
    `out_host[l]` receives `[T, Hq * D]`.  q and k are made contiguous on the device (RoPE does that
    in the real model, qwen2.py:268); v stays the strided view the reference hands over
    (qwen2.py:288-289).
    """

    def __init__(self, n_q_heads: int, n_kv_heads: int, head_dim: int, device: torch.device):
        if DROPIN_DIR not in sys.path:
            sys.path.insert(0, DROPIN_DIR)
        import attn_backend  # noqa: E402  (the drop-in module)
        self._backend = attn_backend.SimpleAttentionBackend
        self.Hq, self.Hkv, self.D = int(n_q_heads), int(n_kv_heads), int(head_dim)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostStagedAttention needs a CUDA device (there is no CPU path)")
        self.h2d = torch.cuda.Stream(device=self.device)
        self.d2h = torch.cuda.Stream(device=self.device)
        self._stage: List[torch.Tensor] = []

    def _staging(self, like: torch.Tensor) -> List[torch.Tensor]:
        if not self._stage or self._stage[0].shape != like.shape or self._stage[0].dtype != like.dtype:
            self._stage = [torch.empty(like.shape, dtype=like.dtype, device=self.device) for _ in range(2)]
        return self._stage

    def _split(self, qkv: torch.Tensor):
        T = qkv.shape[0]
        q, k, v = qkv.split([self.Hq * self.D, self.Hkv * self.D, self.Hkv * self.D], dim=-1)
        return (q.contiguous().view(T, self.Hq, self.D), k.contiguous().view(T, self.Hkv, self.D),
                v.view(T, self.Hkv, self.D))

    @torch.no_grad()
    def forward_layers(self, qkv_host: Sequence[torch.Tensor], out_host: Sequence[torch.Tensor], layers,
                       forward_batch, after=None, out_rows=None) -> None:
        """One step: every layer's attention, uploads and downloads overlapped with compute.
        `after(out)` (optional) post-processes a layer's device output before it is downloaded.
        `out_rows = (lo, hi)` downloads only rows lo..hi-1 of each layer's result into `out_host[l]` (under tensor
        parallelism the all-reduced result is replicated: each rank then brings home its 1/N of the rows instead of
        every rank downloading all of them).  Returns once `out_host` is complete."""
        L = len(layers)
        if len(qkv_host) != L or len(out_host) != L:
            raise RuntimeError("qkv_host / out_host / layers must have one entry per layer")
        for t in list(qkv_host) + list(out_host):
            if t.is_cuda or not t.is_pinned():
                raise RuntimeError("host buffers must be pinned CPU tensors")
        cur = torch.cuda.current_stream(self.device)
        stage = self._staging(qkv_host[0])
        up = [torch.cuda.Event() for _ in range(L)]
        done = [torch.cuda.Event() for _ in range(L)]
        copied = [torch.cuda.Event() for _ in range(L)]
        self.h2d.wait_stream(cur)          # staging buffers may still be read by earlier work
        with torch.cuda.stream(self.h2d):
            stage[0].copy_(qkv_host[0], non_blocking=True)
            up[0].record(self.h2d)
        keep = []
        for l in range(L):
            if l + 1 < L:
                with torch.cuda.stream(self.h2d):
                    if l >= 1:
                        self.h2d.wait_event(done[l - 1])      # layer l-1 has consumed stage[(l+1) % 2]
                    stage[(l + 1) % 2].copy_(qkv_host[l + 1], non_blocking=True)
                    up[l + 1].record(self.h2d)
            cur.wait_event(up[l])
            q, k, v = self._split(stage[l % 2])
            out = self._backend.forward(q, k, v, layers[l], forward_batch, True)
            if after is not None:
                # `after` may hand back a view of a buffer that its NEXT call overwrites (K8's communicator has a
                # single output buffer, valid until the following call): layer l-1's download must have left it
                # before layer l's call starts.  Waiting on this rank's compute stream also holds back the peers'
                # stores into this rank's buffer, which they issue only after seeing this rank enter call l.
                if l >= 1:
                    cur.wait_event(copied[l - 1])
                out = after(out)
            done[l].record(cur)
            out.record_stream(self.d2h)
            keep.append(out)
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(done[l])
                out_host[l].copy_(out if out_rows is None else out[out_rows[0]:out_rows[1]], non_blocking=True)
                copied[l].record(self.d2h)
        cur.wait_stream(self.d2h)
        cur.synchronize()                  # the caller reads out_host next
