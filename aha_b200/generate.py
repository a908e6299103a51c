"""Greedy response decode against the same KV cache: fast_greedy_generate with the reference's
signature and return value (models/modeling_live.py:64-90).  Every step is one q=1 prefill of the native engine with
last-token logits, one `aha_greedy_pick` launch (repetition penalty + argmax on the device) and one embedding lookup; the
chosen id is read back once per token, exactly where the reference tests it against eos."""
from __future__ import annotations

from typing import List, Optional

import torch

from . import _lib


def greedy_pick(logits: torch.Tensor, generated_token_ids: Optional[List[int]] = None,
                repetition_penalty: Optional[float] = None) -> torch.Tensor:
    """logits [..., V] float32 CUDA (last position is used) -> int64 tensor [1, 1] on the device."""
    row = logits.reshape(-1, logits.shape[-1])[-1].contiguous()
    if row.dtype != torch.float32 or not row.is_cuda:
        raise TypeError("logits must be a float32 CUDA tensor")
    out = torch.empty(1, 1, dtype=torch.int64, device=row.device)
    gen = scratch = None
    n = 0
    pen = 1.0
    if repetition_penalty is not None and generated_token_ids:
        gen = torch.tensor(generated_token_ids, dtype=torch.int64, device=row.device)
        scratch = torch.empty_like(row)
        n, pen = gen.numel(), float(repetition_penalty)
    _lib.check(_lib.load().aha_greedy_pick(_lib.ptr(row), row.numel(), _lib.ptr(gen), n, pen, _lib.ptr(scratch), _lib.ptr(out),
                                           _lib.stream_ptr()))
    return out


def fast_greedy_generate(*, model, inputs_embeds: torch.Tensor, past_key_values, eos_token_id: int,
                         inplace_output_ids: torch.Tensor, repetition_penalty=None, generated_token_ids=None):
    if generated_token_ids is None:
        generated_token_ids = []
    if repetition_penalty is not None:
        assert isinstance(repetition_penalty, float)
    i = 0
    for i in range(inplace_output_ids.size(1)):
        outputs = model(inputs_embeds=inputs_embeds, past_key_values=past_key_values, use_cache=True, return_dict=True,
                        logits_mode="last")
        past_key_values = outputs.past_key_values
        new_token_id = greedy_pick(outputs.logits, generated_token_ids, repetition_penalty)
        tok = int(new_token_id.item())                      # the reference's `new_token_id == eos_token_id` sync
        if repetition_penalty is not None and tok != eos_token_id:   # special tokens should not be penalized
            generated_token_ids.append(tok)
        inplace_output_ids[:, i] = new_token_id[0]
        if tok == eos_token_id:
            break
        inputs_embeds = model.get_input_embeddings()(new_token_id)
    return inplace_output_ids[:, :i + 1], past_key_values, generated_token_ids
