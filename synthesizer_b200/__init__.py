"""
synthesizer_b200 — a B200-native (sm_100a) render backend with the API of irmen/synthesizer's
``synthplayer`` render path: oscillators with FM, the filter chain, and Sample mixing,
resampling and amplitude operations.  Host code is Python over a C ABI
(include/synth_b200.h -> synthesizer_b200/libsynth_b200.so); the kernels are hand-written
CUDA.  There is no CPU path: importing needs the built library, rendering needs a GPU.
"""
from . import params  # noqa: F401
from . import _cabi  # noqa: F401  (raises ImportError when libsynth_b200.so is not built)
from .oscillators import *  # noqa: F401,F403
from .sample import Sample  # noqa: F401
from .synth import WaveSynth  # noqa: F401
from .voicebank import VoiceBank  # noqa: F401

__version__ = "0.1.0"
