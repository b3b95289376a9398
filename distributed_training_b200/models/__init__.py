"""Model zoo: toy / wide MLP and GPT-2."""
from .mlp import MLP, wide_mlp  # noqa: F401
from .gpt2 import GPT2, GPT2Config  # noqa: F401
