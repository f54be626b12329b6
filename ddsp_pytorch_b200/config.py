"""Hyper-parameters of the path.  Field names and shipped values follow the reference's ``Config``
(config/default.py:8-24) so objects of either class can be handed to the modules; the named overrides are
the shapes BASELINE.json quotes."""
from __future__ import annotations

from dataclasses import dataclass, replace


@dataclass
class Config:
    data_dir: str = "/home/kureta/Music/flute"
    example_duration: int = 2          # seconds
    example_overlap: float = 0.5
    sample_rate: int = 44100
    n_fft: int = 2048
    hop_length: int = 512
    crepe_capacity: str = "full"
    n_harmonics: int = 180
    n_noise_filters: int = 195
    decoder_mlp_units: int = 512
    decoder_mlp_layers: int = 3
    decoder_gru_units: int = 512
    decoder_gru_layers: int = 1
    batch_size: int = 16

    def with_(self, **kw) -> "Config":
        return replace(self, **kw)


SHIPPED = Config()
# cfg 1 of BASELINE.json: 16 x 4 s @ 16 kHz, 100 harmonics, 65 noise bands
CFG1 = Config(sample_rate=16000, example_duration=4, n_harmonics=100, n_noise_filters=65, batch_size=16)
# cfg 4: 1024 voices x 4 s @ 48 kHz, K in {100,128,180,256}, 65 noise bands
CFG4 = Config(sample_rate=48000, example_duration=4, n_harmonics=256, n_noise_filters=65, batch_size=1024)
