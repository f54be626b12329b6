"""B200-native DDSP synthesis hot path (oscillator bank, filtered noise, reverb, multi-scale spectral loss)
behind the module API of kureta/ddsp-pytorch.  Arithmetic lives in lib/libddsp_b200.so (csrc/*.cu, sm_100a)."""
from .config import CFG1, CFG4, SHIPPED, Config  # noqa: F401
from .modules import (FilteredNoise, MSSLoss, OscillatorBank, Reverb, SSSLoss,  # noqa: F401
                      amp_to_impulse_response, fft_convolve)
from . import ops  # noqa: F401
