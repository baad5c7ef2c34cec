"""f5e_tts_b200 - B200-native (sm_100a) inference hot path of F5E-TTS behind the reference's Python API."""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
