"""B200-native GGUF decode engine: host-side mirror above the C-ABI library.

Replaces the one data-parallel hot path behind LittleSound/ollama-operator -- the
decode loop of the server container every `Model` CR deploys (reference
pkg/model/pod.go:14-66).  The compute lives in csrc/ (hand-written sm_100a CUDA behind
`include/b2gguf.h`); this package only binds it.
"""
from . import engine, synth  # noqa: F401

__all__ = ["engine", "synth"]
