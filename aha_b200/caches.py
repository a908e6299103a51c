"""The reference's three KV-cache policies on the native device ring buffer, behind the HF `Cache` API.

    SinkCache(window_length, num_sink_tokens)      test/sink_cache.py:7-164
    SlidingWindowCache(window_length)              test/sliding_window_cache.py:5-53
    TrulyStaticCache(window_size)                  test/static_cache.py:5-47

Same constructors, `update / get_seq_length / get_max_length / get_max_cache_shape`, `key_cache / value_cache`,
`window_length / window_size`, `num_sink_tokens`, `_seen_tokens`, `is_sliding`.  Storage and every byte moved live in
libaha_b200.so (aha_cache_* in include/aha_b200.h); `update()` returns a logical-order materialisation for callers that
run their own attention (HF interop, bit-exactness tests) - the fused engine (`AhaModel.forward`) never materialises it.
"""
from __future__ import annotations

import ctypes
from typing import Any, Dict, Optional, Tuple

import torch
from transformers import Cache

from . import _lib


class _LayerView:
    """`cache.key_cache[layer]` / `cache.value_cache[layer]` -> logical-order tensor [1, kvH, len, d]."""

    def __init__(self, cache: "_NativeCache", which: int):
        self._c, self._w = cache, which

    def __len__(self):
        return self._c._populated_layers()

    def __getitem__(self, layer):
        if isinstance(layer, slice):                 # the reference's key_cache is a plain list
            return [self[i] for i in range(*layer.indices(len(self)))]
        if layer < 0:
            layer += len(self)
        if not 0 <= layer < len(self):
            raise IndexError(layer)
        return self._c._export(layer)[self._w]

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class _NativeCache(Cache):
    _policy = _lib.POLICY_SINK

    def __init__(self, window: int, sinks: int = 0, num_layers: int = 0, max_chunk: Optional[int] = None):
        try:
            super().__init__()                       # transformers 4.49 (the reference's pin)
        except (TypeError, ValueError):
            super().__init__(layers=[])              # transformers >= 4.54 wants a layer list
        if window is None or int(window) <= 0:
            raise ValueError("window_length must be positive")
        self._window, self._sinks = int(window), int(sinks)
        self._num_layers_hint, self._max_chunk = int(num_layers), max_chunk
        self._h = None
        self._kvh = self._d = None
        self._device = None
        self.key_cache = _LayerView(self, 0)
        self.value_cache = _LayerView(self, 1)

    # ---- native handle ----
    def _ensure(self, kvh: int, d: int, q: int, device: torch.device):
        if self._h is not None:
            if (kvh, d) != (self._kvh, self._d):
                raise ValueError(f"key/value shape changed: got heads={kvh}, head_dim={d}, cache holds {self._kvh}, {self._d}")
            return
        lib = _lib.load()
        max_chunk = max(int(self._max_chunk or 0), q, 1)
        h = ctypes.c_void_p()
        dev = device.index if device.index is not None else torch.cuda.current_device()
        _lib.check(lib.aha_cache_create(self._num_layers_hint, kvh, d, self._policy, self._window, self._sinks, max_chunk,
                                        dev, ctypes.byref(h)))
        self._h, self._kvh, self._d, self._device = h, kvh, d, torch.device("cuda", dev)

    def allocate(self, kv_heads: int, head_dim: int, num_layers: int, max_chunk: int, device=None):
        """Eager creation (the engine path): all layer slabs at once."""
        self._num_layers_hint, self._max_chunk = num_layers, max_chunk
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self._ensure(kv_heads, head_dim, 1, dev)
        return self

    @property
    def handle(self):
        return self._h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                _lib.load().aha_cache_destroy(h)
            except Exception:
                pass

    def _populated_layers(self) -> int:
        n = 0
        while self.get_seq_length(n) > 0:
            n += 1
        return n

    def _export(self, layer: int, provenance: bool = False):
        n = self.get_seq_length(layer)
        K = torch.empty(1, self._kvh, n, self._d, dtype=torch.bfloat16, device=self._device)
        V = torch.empty_like(K)
        ids = torch.empty(n, dtype=torch.int32, device=self._device) if provenance else None
        if n > 0:
            _lib.check(_lib.load().aha_cache_export(self._h, layer, _lib.ptr(K), _lib.ptr(V), _lib.ptr(ids), _lib.stream_ptr()))
        return (K, V, ids) if provenance else (K, V)

    def provenance(self, layer: int = 0) -> torch.Tensor:
        """Global token id held by every logical slot (test hook for the eviction order)."""
        return self._export(layer, provenance=True)[2]

    # ---- HF Cache API ----
    def get_seq_length(self, layer_idx: Optional[int] = 0) -> int:
        if self._h is None:
            return 0
        out = ctypes.c_int(0)
        _lib.check(_lib.load().aha_cache_seq_length(self._h, int(layer_idx or 0), ctypes.byref(out)))
        return out.value

    def get_max_length(self) -> Optional[int]:
        return self._window

    def get_max_cache_shape(self) -> Optional[int]:
        return self._window

    def get_mask_sizes(self, cache_position, layer_idx: int = 0):
        # transformers >= 4.54 sizes masks through this; kv_length = what update() will return, no offset
        q = int(cache_position.shape[0]) if hasattr(cache_position, "shape") else int(cache_position)
        L = self.get_seq_length(layer_idx)
        return self._klen_after(L, q), 0

    def _klen_after(self, L: int, q: int) -> int:
        raise NotImplementedError

    def reset(self):
        if self._h is not None:
            _lib.check(_lib.load().aha_cache_reset(self._h))

    def update(self, key_states: torch.Tensor, value_states: torch.Tensor, layer_idx: int,
               cache_kwargs: Optional[Dict[str, Any]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if key_states.dim() != 4 or key_states.shape[0] != 1:
            raise ValueError("aha_b200 caches hold one stream: key_states must be [1, kv_heads, q, head_dim]")
        if key_states.dtype != torch.bfloat16 or value_states.dtype != torch.bfloat16 or not key_states.is_cuda:
            raise TypeError("aha_b200 caches store bf16 CUDA tensors only (no CPU / fp32 fallback)")
        if key_states.shape != value_states.shape:
            raise ValueError("key_states and value_states must have the same shape")
        _, kvh, q, d = key_states.shape
        self._ensure(kvh, d, q, key_states.device)
        if key_states.stride(3) != 1:
            key_states = key_states.contiguous()
        if value_states.stride(3) != 1:
            value_states = value_states.contiguous()
        cos = sin = None
        if cache_kwargs:
            cos, sin = cache_kwargs.get("cos"), cache_kwargs.get("sin")
        if cos is not None and sin is not None:
            cos = cos.reshape(-1, cos.shape[-1]).to(torch.bfloat16).contiguous()
            sin = sin.reshape(-1, sin.shape[-1]).to(torch.bfloat16).contiguous()
            if cos.shape != (q, d) or sin.shape != (q, d):
                raise ValueError(f"cos/sin must be [1, q, head_dim] = [1, {q}, {d}], got {tuple(cos.shape)}")
        else:
            cos = sin = None
        _lib.check(_lib.load().aha_cache_update(
            self._h, int(layer_idx), _lib.ptr(key_states), _lib.ptr(value_states), key_states.stride(1),
            key_states.stride(2), value_states.stride(1), value_states.stride(2), _lib.ptr(cos), _lib.ptr(sin), q,
            _lib.stream_ptr()))
        return self._export(int(layer_idx))


class _SlidingFlag(list):
    """`is_sliding = True` of the reference's SinkCache (test/sink_cache.py:8) in a form transformers >= 4.54 can also
    consume: there `Cache.is_sliding` is a per-layer list that the mask builder searches (`False in cache.is_sliding`)."""

    def __init__(self):
        super().__init__([True])

    def __eq__(self, other):
        return other is True or list.__eq__(self, other)

    __hash__ = None


class SinkCache(_NativeCache):
    """test/sink_cache.py:7-164."""
    is_sliding = _SlidingFlag()
    _policy = _lib.POLICY_SINK

    def __init__(self, window_length: int, num_sink_tokens: int, num_layers: int = 0, max_chunk: Optional[int] = None) -> None:
        if num_sink_tokens < 0 or num_sink_tokens >= window_length:
            raise ValueError("num_sink_tokens must be in [0, window_length)")
        super().__init__(window_length, num_sink_tokens, num_layers, max_chunk)
        self.window_length = window_length
        self.num_sink_tokens = num_sink_tokens

    @property
    def _seen_tokens(self) -> int:
        if self._h is None:
            return 0
        out = ctypes.c_longlong(0)
        _lib.check(_lib.load().aha_cache_seen_tokens(self._h, ctypes.byref(out)))
        return out.value

    def _klen_after(self, L, q):
        return L + q if (L == 0 or L + q < self.window_length) else self.window_length


class SlidingWindowCache(_NativeCache):
    """test/sliding_window_cache.py:5-53 (no sinks, no re-rotation)."""
    _policy = _lib.POLICY_SLIDING

    def __init__(self, window_length: int, num_layers: int = 0, max_chunk: Optional[int] = None):
        super().__init__(window_length, 0, num_layers, max_chunk)
        self.window_length = window_length
        self._seen_tokens = 0          # never updated by the reference either (sliding_window_cache.py:15)

    def _klen_after(self, L, q):
        return q if L == 0 else min(L + q, self.window_length)


class TrulyStaticCache(_NativeCache):
    """test/static_cache.py:5-47: the first update() is frozen (cropped to window_size), later ones are ignored."""
    _policy = _lib.POLICY_STATIC

    def __init__(self, window_size: int, num_layers: int = 0, max_chunk: Optional[int] = None):
        super().__init__(window_size, 0, num_layers, max_chunk)
        self.window_size = window_size
        self._seen_tokens = 0

    def _klen_after(self, L, q):
        return min(q, self.window_size) if L == 0 else L


def make_cache(alt_cache: str = "default_sink", window_length: int = 2048, num_sink_tokens: int = 32,
               sink_cache: bool = False, n_sys: int = 0, **kw) -> _NativeCache:
    """LiveInferForBenchmark._init_cache (test/inference.py:133-155)."""
    if sink_cache:
        return SinkCache(window_length=window_length + num_sink_tokens - n_sys, num_sink_tokens=n_sys, **kw)
    if alt_cache == "default_sink":
        return SinkCache(window_length=window_length, num_sink_tokens=num_sink_tokens, **kw)
    if alt_cache == "sliding_window":
        return SlidingWindowCache(window_length=window_length, **kw)
    if alt_cache == "static":
        return TrulyStaticCache(window_size=window_length, **kw)
    raise ValueError(f"unknown alt_cache {alt_cache!r}")
