"""piper_b200: host-side Python mirror of the reference's PiperMetalRuntime over the
B200-native C-ABI library (include/piper_b200.h).  See runtime.py."""
from .spec import VoiceSpec, SPECS, MEDIUM, HIGH, MEDIUM_MULTISPEAKER, TINY, FIXTURE_IDS, FIXTURE_SCALES, DEFAULT_SEED  # noqa: F401
