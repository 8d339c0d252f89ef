"""Model dimensions for the three Llama configurations BASELINE.json names.

The reference (Berkay2002/fasttalk-llm-microservice) carries no model code: it
names models only by tag (`app/utils/config.py:86` "llama3.2:1b",
`docker-compose.vllm.yml:39`).  The dimensions below are the public HF configs
of those models.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict
from typing import Dict


@dataclass(frozen=True)
class ModelSpec:
    name: str
    hidden: int
    ffn: int
    layers: int
    n_q_heads: int
    n_kv_heads: int
    head_dim: int
    vocab: int = 128256
    tied_embeddings: bool = True
    rope_theta: float = 500000.0
    rope_factor: float = 32.0
    rope_low_freq_factor: float = 1.0
    rope_high_freq_factor: float = 4.0
    rope_original_max_pos: int = 8192
    rms_eps: float = 1e-5
    max_position: int = 131072

    # ---- derived sizes -------------------------------------------------
    @property
    def q_dim(self) -> int:
        return self.n_q_heads * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_kv_heads * self.head_dim

    @property
    def qkv_dim(self) -> int:
        return self.q_dim + 2 * self.kv_dim

    @property
    def group(self) -> int:
        return self.n_q_heads // self.n_kv_heads

    def param_count(self) -> int:
        per_layer = (
            self.qkv_dim * self.hidden          # fused q,k,v projection
            + self.hidden * self.q_dim          # o projection
            + 2 * self.ffn * self.hidden        # gate + up
            + self.hidden * self.ffn            # down
            + 2 * self.hidden                   # two RMSNorm weights
        )
        emb = self.vocab * self.hidden
        head = 0 if self.tied_embeddings else self.vocab * self.hidden
        return self.layers * per_layer + emb + head + self.hidden

    def weight_bytes(self) -> int:
        """bf16 bytes streamed once per decode step (SURVEY.md §8d P_bytes)."""
        return 2 * self.param_count()

    def kv_bytes_per_token(self) -> int:
        """16-bit (fp16 in the cache, bf16 in SURVEY.md §8d: same size) K+V bytes one cached token occupies over all layers."""
        return 2 * self.layers * self.n_kv_heads * self.head_dim * 2

    def decode_step_bytes(self, ctx_lens) -> int:
        """Algorithmic HBM bytes of one decode step."""
        return self.weight_bytes() + int(sum(ctx_lens)) * self.kv_bytes_per_token()

    def to_dict(self) -> Dict:
        return asdict(self)


LLAMA_3_2_1B = ModelSpec(
    name="llama-3.2-1b", hidden=2048, ffn=8192, layers=16,
    n_q_heads=32, n_kv_heads=8, head_dim=64, tied_embeddings=True, rope_factor=32.0,
)
LLAMA_3_2_3B = ModelSpec(
    name="llama-3.2-3b", hidden=3072, ffn=8192, layers=28,
    n_q_heads=24, n_kv_heads=8, head_dim=128, tied_embeddings=True, rope_factor=32.0,
)
LLAMA_3_1_8B = ModelSpec(
    name="llama-3.1-8b", hidden=4096, ffn=14336, layers=32,
    n_q_heads=32, n_kv_heads=8, head_dim=128, tied_embeddings=False, rope_factor=8.0,
)

# Small shapes of the same architecture for tests (all GEMM dims stay multiples
# of the 128-row / 64-column tiles the tcgen05 kernel needs).
TINY_D64 = ModelSpec(
    name="tiny-d64", hidden=256, ffn=512, layers=2,
    n_q_heads=8, n_kv_heads=2, head_dim=64, vocab=1024, tied_embeddings=True,
)
TINY_D128 = ModelSpec(
    name="tiny-d128", hidden=384, ffn=512, layers=2,
    n_q_heads=3, n_kv_heads=1, head_dim=128, vocab=1280, tied_embeddings=False,
    rope_factor=8.0,
)

# splits over two tensor-parallel ranks in whole tiles (4 q heads, 2 kv heads, ffn 256, vocab 512 each)
TINY_TP = ModelSpec(
    name="tiny-tp", hidden=256, ffn=512, layers=2,
    n_q_heads=8, n_kv_heads=4, head_dim=64, vocab=1024, tied_embeddings=False,
)

SPECS: Dict[str, ModelSpec] = {
    s.name: s for s in (LLAMA_3_2_1B, LLAMA_3_2_3B, LLAMA_3_1_8B, TINY_D64, TINY_D128, TINY_TP)
}
# the tags the reference's config uses (app/utils/config.py:86, README.md:474)
ALIASES = {
    "llama3.2:1b": "llama-3.2-1b",
    "llama3.2:3b": "llama-3.2-3b",
    "llama3.1:8b": "llama-3.1-8b",
    "llama3:8b": "llama-3.1-8b",
    "meta-llama/Llama-3.2-1B-Instruct": "llama-3.2-1b",
    "meta-llama/Llama-3.2-3B-Instruct": "llama-3.2-3b",
    "meta-llama/Llama-3.1-8B-Instruct": "llama-3.1-8b",
}


def get_spec(name: str) -> ModelSpec:
    key = ALIASES.get(name, name)
    if key not in SPECS:
        raise KeyError(f"unknown model '{name}'; known: {sorted(SPECS)} + {sorted(ALIASES)}")
    return SPECS[key]
