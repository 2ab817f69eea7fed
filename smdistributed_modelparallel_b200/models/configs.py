"""Named model configurations expressed as ``DistributedTransformerLMHead`` keyword arguments.

Shapes are the public architectures (GPT-2, GPT-J 6B, GPT-NeoX 20B, BERT-large); the keyword mapping is
the one the reference's HuggingFace init hooks produce (reference smdistributed/modelparallel/torch/nn/
huggingface/gpt2.py, gptj.py:34-83, gptneox.py:35-90, bert.py:171-197).
"""

from __future__ import annotations


def _gpt2(layers, hidden, heads, seq=1024, vocab=50257):
    return dict(num_layers=layers, num_attention_heads=heads, attention_head_size=hidden // heads, hidden_size=hidden,
                intermediate_size=4 * hidden, vocab_size=vocab, num_positions=seq, attention_dropout_prob=0.0,
                hidden_dropout_prob=0.0, embedding_dropout_prob=0.0, activation="gelu_new", layernorm_epsilon=1e-5,
                causal_mask_size=seq, add_lm_head=True, pre_layernorm=True, post_layernorm=False, final_layernorm=True,
                use_normal_initialization=True, initializer_range=0.02)


def _gptj(layers, hidden, heads, seq=2048, vocab=50400, rotary=64):
    return dict(num_layers=layers, num_attention_heads=heads, attention_head_size=hidden // heads, hidden_size=hidden,
                intermediate_size=4 * hidden, vocab_size=vocab, num_positions=seq, attention_dropout_prob=0.0,
                hidden_dropout_prob=0.0, embedding_dropout_prob=0.0, activation="gelu_new", layernorm_epsilon=1e-5,
                causal_mask_size=seq, add_lm_head=True, pre_layernorm=True, post_layernorm=False, final_layernorm=True,
                use_normal_initialization=True, initializer_range=0.02, rotary_dim=rotary, rotary_emb_base=10000,
                use_positional_embedding=False, parallel_attn_output=True, single_pre_layernorm=True,
                use_qkv_bias=False, use_attn_dense_bias=False, tie_input_output_embedding=False, use_lm_head_bias=True)


def _neox(layers, hidden, heads, seq=2048, vocab=50432, rotary_pct=0.25):
    d = hidden // heads
    return dict(num_layers=layers, num_attention_heads=heads, attention_head_size=d, hidden_size=hidden,
                intermediate_size=4 * hidden, vocab_size=vocab, num_positions=seq, attention_dropout_prob=0.0,
                hidden_dropout_prob=0.0, embedding_dropout_prob=0.0, activation="gelu_fast", layernorm_epsilon=1e-5,
                causal_mask_size=seq, add_lm_head=True, pre_layernorm=True, post_layernorm=False, final_layernorm=True,
                use_normal_initialization=True, initializer_range=0.02, rotary_dim=int(d * rotary_pct),
                rotary_emb_base=10000, gpt_neox_type_rotary=True, use_positional_embedding=False,
                parallel_attn_output=True, tie_input_output_embedding=False, use_lm_head_bias=False)


def _bert(layers, hidden, heads, seq=512, vocab=30522):
    return dict(num_layers=layers, num_attention_heads=heads, attention_head_size=hidden // heads, hidden_size=hidden,
                intermediate_size=4 * hidden, vocab_size=vocab, num_positions=seq, attention_dropout_prob=0.0,
                hidden_dropout_prob=0.0, embedding_dropout_prob=0.0, activation="gelu", layernorm_epsilon=1e-12,
                causal_mask_size=None, add_lm_head=True, pre_layernorm=False, post_layernorm=True, num_token_types=2,
                use_normal_initialization=True, initializer_range=0.02)


MODEL_CONFIGS = {
    "gpt2-small": _gpt2(12, 768, 12),
    "gpt2-xl": _gpt2(48, 1600, 25),
    "gpt-j-6b": _gptj(28, 4096, 16),
    "gpt-neox-20b": _neox(44, 6144, 64),
    "bert-large": _bert(24, 1024, 16),
    "roberta-large": _bert(24, 1024, 16, seq=514, vocab=50265),
    # test-suite shape of the reference (6 layers, H=768, 12 heads, S=512): test/torch/model_zoo/gpt_models.py:34-46
    "gpt2-test": _gpt2(6, 768, 12, seq=512),
    "gptj-test": _gptj(6, 768, 12, seq=512, rotary=16),
    "tiny": _gptj(2, 64, 4, seq=32, vocab=128, rotary=8),
}


def param_count(cfg: dict) -> int:
    H, L, V, I = cfg["hidden_size"], cfg["num_layers"], cfg["vocab_size"], cfg["intermediate_size"]
    per_layer = 4 * H * H + 2 * H * I + I + H
    emb = V * H + (cfg["num_positions"] * H if cfg.get("use_positional_embedding", True) else 0)
    head = 0 if cfg.get("tie_input_output_embedding", True) else V * H
    return L * per_layer + emb + head


def train_flops_per_token(cfg: dict, seq_len: int) -> float:
    """6 * params(matmul) + attention score/value FLOPs (forward 2*2*S*H per layer, x3 for training)."""
    H, L, V, I = cfg["hidden_size"], cfg["num_layers"], cfg["vocab_size"], cfg["intermediate_size"]
    dense = L * (4 * H * H + 2 * H * I) + V * H
    return 6.0 * dense + 12.0 * L * seq_len * H
