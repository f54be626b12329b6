"""The drop-in shims resolve the reference's import paths to the B200 modules; when the reference checkout is
present (build container only) its untouched Decoder is built on top of them and the state-dict contract of
SURVEY.md §5/§8b is compared key by key."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "ddsp_pytorch_b200", "dropin")
REF = "/root/reference"


def test_shim_import_paths():
    code = ("from model.ddsp.harmonic_oscillator import OscillatorBank; from model.ddsp.filtered_noise import FilteredNoise, "
            "fft_convolve, amp_to_impulse_response; from model.ddsp.reverb import Reverb; from loss.mss_loss import MSSLoss, SSSLoss; "
            "print(OscillatorBank.__module__, Reverb.__module__, MSSLoss.__module__)")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([DROPIN, ROOT]))
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert out.stdout.split() == ["ddsp_pytorch_b200.modules"] * 3


def test_module_state_dict_contract():
    import ddsp_pytorch_b200 as d
    conf = d.Config()
    bank, rv, loss = d.OscillatorBank(conf), d.Reverb(conf), d.MSSLoss([2048, 64])
    sd = bank.state_dict()
    assert sd["harmonics"].dtype == torch.int64 and sd["last_phases"].dtype == torch.int64
    assert sd["harmonics"].tolist() == list(range(1, conf.n_harmonics + 1))
    assert {k: tuple(v.shape) for k, v in rv.state_dict().items()} == {
        "noise": (44100,), "decay": (), "wet": (), "t": (1, 44100), "buffer": (1, 44100)}
    assert [n for n, p in rv.named_parameters() if p.requires_grad] == ["noise", "decay", "wet"]
    assert set(loss.state_dict()) == {"losses.0.stft.window", "losses.1.stft.window"}
    assert d.FilteredNoise(conf).block_size == conf.hop_length and rv.length == rv.sampling_rate == 44100


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout only exists in the build container")
def test_reference_decoder_builds_on_the_shims_and_state_dicts_match():
    code = r'''
import sys, torch
from model.autoencoder.decoder import Decoder      # the reference's own, untouched file
from loss.mss_loss import MSSLoss
from config.default import Config
dec = Decoder(Config())
assert type(dec.harmonics).__module__ == type(dec.reverb).__module__ == type(dec.noise).__module__ == "ddsp_pytorch_b200.modules"
mine = {k: (tuple(v.shape), str(v.dtype)) for k, v in dec.state_dict().items()}
lossk = {k: (tuple(v.shape), str(v.dtype)) for k, v in MSSLoss([2048, 1024, 512, 256, 128, 64]).state_dict().items()}
for m in [m for m in list(sys.modules) if m.split(".")[0] in ("model", "loss")]: del sys.modules[m]
sys.path[:] = [p for p in sys.path if "dropin" not in p]
from model.autoencoder.decoder import Decoder as RefDecoder
from loss.mss_loss import MSSLoss as RefLoss
ref = RefDecoder(Config())
assert type(ref.harmonics).__module__ == "model.ddsp.harmonic_oscillator"
theirs = {k: (tuple(v.shape), str(v.dtype)) for k, v in ref.state_dict().items()}
assert mine == theirs, set(mine.items()) ^ set(theirs.items())
assert lossk == {k: (tuple(v.shape), str(v.dtype)) for k, v in RefLoss([2048, 1024, 512, 256, 128, 64]).state_dict().items()}
dec.load_state_dict(ref.state_dict(), strict=True)   # rt/synth.py:17 loads strictly
train = lambda m: sorted(n for n, p in m.named_parameters() if p.requires_grad)
assert train(dec) == train(ref)
print("ok")
'''
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([DROPIN, ROOT, REF]))
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]
