"""Streaming blocks through RealtimeSynth == the same decoder run block by block against the fp64 oracle's live path."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_blocks_match_oracle_live_path_and_keep_state():
    import test_train_gloo as H
    from ddsp_pytorch_b200.realtime import RealtimeSynth
    from ddsp_pytorch_b200.train import synthetic_batch
    from oracle import ddsp_oracle as O
    from oracle.philox import noise_uniform
    conf = H.small_conf()
    conf.sample_rate, conf.hop_length, conf.n_noise_filters = 4000, 64, 9
    torch.manual_seed(1)
    rt = RealtimeSynth(conf, "cuda:0", frames_per_block=4, noise_seed=3)
    rt.decoder.reverb.wet.data.fill_(1.0)                       # make the reverb tail audible
    dec = rt.decoder
    # oracle-side state
    last = torch.zeros(conf.n_harmonics, dtype=torch.float64)
    buf = torch.zeros(1, conf.sample_rate, dtype=torch.float64)
    ir = O.reverb_ir(dec.reverb.noise.detach().cpu().double(), dec.reverb.decay.detach().cpu().double(),
                     dec.reverb.wet.detach().cpu().double(), conf.sample_rate)
    for blk in range(5):
        feats = synthetic_batch(conf, 1, "cpu", seed=10 + blk)
        z = {k: v[:, :4].contiguous() for k, v in feats.items() if k != "audio"}
        got = torch.from_numpy(rt.process(z).copy()).double()
        with torch.no_grad():
            ctrl, _ = dec.controller({k: v.cuda() for k, v in z.items()}, rt.hidden)
        c, a, Hh, f0 = (ctrl[k].detach().cpu().double() for k in ("c", "a", "H", "f0"))
        harm, last = O.oscillator_bank_live(f0, c, a, last, conf.n_harmonics, conf.sample_rate, conf.hop_length)
        u = torch.from_numpy(noise_uniform(1, 4, conf.hop_length, 3, blk + 1, 0)).double()   # device call counter: 1, 2, ..
        dry = harm + O.filtered_noise(Hh, conf.hop_length, u)
        want, buf = O.reverb_live(buf, dry, ir)
        err = float((got - want.view(-1)).abs().max())
        print(f"block {blk}: |cuda - oracle64| = {err:.2e}, max|y| = {float(want.abs().max()):.3f}")
        assert err <= 1e-4 * max(1.0, float(want.abs().max()))
    st = rt.stats()
    assert st["blocks"] == 5 and st["p50_ms"] > 0


def test_graph_replay_continues_the_eager_stream():
    """graph=True (blocks 0-1 eager, then CUDA-graph replays) produces the same audio as the all-eager pipeline."""
    import test_train_gloo as H
    from ddsp_pytorch_b200.realtime import RealtimeSynth
    from ddsp_pytorch_b200.train import synthetic_batch
    conf = H.small_conf()
    conf.sample_rate, conf.hop_length, conf.n_noise_filters = 4000, 64, 9
    outs = []
    for graph in (False, True):
        torch.manual_seed(1)
        rt = RealtimeSynth(conf, "cuda:0", frames_per_block=4, noise_seed=3, graph=graph)
        rt.decoder.reverb.wet.data.fill_(1.0)
        blocks = []
        for blk in range(7):
            feats = synthetic_batch(conf, 1, "cpu", seed=10 + blk)
            z = {k: v[:, :4].contiguous() for k, v in feats.items() if k != "audio"}
            blocks.append(torch.from_numpy(rt.process(z).copy()))
        outs.append(torch.stack(blocks))
        assert (rt._g is not None) == graph
    eager, graphed = outs
    assert float(eager.abs().max()) > 1e-3
    assert float((eager - graphed).abs().max()) <= 1e-6 * float(eager.abs().max())
    assert not torch.equal(graphed[3], graphed[4])            # state and noise advance between replays


def test_live_path_from_a_callback_thread():
    """rt/synth.py:40-55 calls the decoder from the JACK process-callback thread while the main thread waits (SURVEY 8b,
    threading row): the live ops keep no thread-affine state -- blocks alternated between a worker thread and the main
    thread continue one stream, bit for bit, with a twin that stays on the main thread."""
    import threading
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    conf = d.Config(sample_rate=4000, hop_length=64, n_harmonics=20, n_noise_filters=9)
    torch.manual_seed(5)
    mods = []
    for _ in range(2):
        torch.manual_seed(5)
        mods.append((d.OscillatorBank(conf).cuda(), d.FilteredNoise(conf, seed=11), d.Reverb(conf, initial_wet=1.0).cuda()))
    mods[1][2].load_state_dict(mods[0][2].state_dict())

    def block(m, blk):
        f0, c, a, H = (t.cuda() for t in synthetic_controls(1, 4, conf.n_harmonics, conf.n_noise_filters, conf.sample_rate,
                                                            conf.hop_length, seed=40 + blk))
        bank, fn, rv = m
        with torch.no_grad():
            y = rv.live_forward(bank.live({"f0": f0, "c": c, "a": a}) + fn({"H": H}))
        return y.cpu()

    for blk in range(4):
        want = block(mods[0], blk)
        if blk % 2 == 0:
            box = {}
            th = threading.Thread(target=lambda: box.update(y=block(mods[1], blk)))
            th.start()
            th.join()
            got = box["y"]
        else:
            got = block(mods[1], blk)
        assert got.shape == (1, 4 * conf.hop_length)
        assert torch.equal(got, want), f"block {blk} differs between threads"
