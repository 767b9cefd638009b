"""B200-native (sm_100a) quantized Llama decode path of proxy-inference-engine: Python host over libpie_b200.so."""
from . import _lib  # noqa: F401

__all__ = ["_lib", "ops", "nn", "cache", "llama", "decode", "checkpoint"]
