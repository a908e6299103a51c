"""N2: folding peft LoRA adapters into the base weights at load time (aha_b200/weights.py).  CPU: the merged Linear equals
the unmerged y = W x + (alpha / r) B (A x) of peft within bf16 rounding, key handling, error cases."""
import pytest
import torch

from aha_b200.weights import merge_lora_state_dict


def _mk(seed=0, out_f=96, in_f=64, r=16):
    g = torch.Generator().manual_seed(seed)
    W = (torch.randn(out_f, in_f, generator=g) * 0.05).bfloat16()
    A = (torch.randn(r, in_f, generator=g) * 0.05).bfloat16()
    B = (torch.randn(out_f, r, generator=g) * 0.05).bfloat16()
    return W, A, B


@pytest.mark.parametrize("infix", [".default", ""])
def test_merged_equals_unmerged(infix):
    W, A, B = _mk()
    name = "model.layers.3.self_attn.q_proj"
    base = {name + ".weight": W, "informative_head.weight": torch.zeros(2, 64).bfloat16(), "other": torch.ones(3)}
    head = torch.randn(2, 64).bfloat16()
    adapter = {f"base_model.model.{name}.lora_A{infix}.weight": A, f"base_model.model.{name}.lora_B{infix}.weight": B,
               f"base_model.model.informative_head.modules_to_save{infix or '.default'}.weight": head}
    merged = merge_lora_state_dict(base, adapter, lora_alpha=32.0, r=16)
    x = torch.randn(7, 64).bfloat16()
    ref = x.float() @ W.float().t() + (32.0 / 16) * ((x.float() @ A.float().t()) @ B.float().t())
    got = x.float() @ merged[name + ".weight"].float().t()
    assert merged[name + ".weight"].dtype == torch.bfloat16
    assert (got - ref).abs().max() <= 2e-2 * ref.abs().max()
    assert torch.equal(merged["informative_head.weight"], head) and torch.equal(merged["other"], base["other"])
    assert torch.equal(base[name + ".weight"], W), "the base dict must not be modified"


def test_merge_errors():
    W, A, B = _mk()
    with pytest.raises(ValueError, match="unpaired"):
        merge_lora_state_dict({"a.weight": W}, {"base_model.model.a.lora_A.weight": A})
    with pytest.raises(KeyError):
        merge_lora_state_dict({"a.weight": W}, {"b.lora_A.weight": A, "b.lora_B.weight": B})
    with pytest.raises(ValueError, match="rank"):
        merge_lora_state_dict({"a.weight": W}, {"a.lora_A.weight": A, "a.lora_B.weight": B}, r=8)
