from ddsp_pytorch_b200.modules import OscillatorBank  # noqa: F401  (replaces model/ddsp/harmonic_oscillator.py)
