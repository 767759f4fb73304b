"""-m gpu: the path bench.py TIMES, at the sizes it times it, against the CPU oracle.

``AudioPipeline.embed_device`` (device-resident `value`) and ``AudioPipeline.run_pinned(overlap=True)`` (the `e2e` leg: two
half-batches, twin plans, two streams) with the full 50 + 20 block stack, on

  * cfg3 (BASELINE.json config 3): 64 utterances of 1-30 s (seed 0), packed offsets - EVERY utterance compared;
  * a cfg5-sized batch (360 x 10 s = 60 120 rows: 256 x 256 CTA-pair tiles, folded-LN producer/consumer chains 70 deep,
    32-row FSMN CTAs, dual-stream halves) - 16 sampled utterances compared;
  * cfg2 (16 x 10 s: the CUDA-graph replay path) - every utterance compared.

Oracle = numpy fbank/LFR/CMVN (torchaudio's own fp32 tables) -> fp32 PyTorch SAN-M encoder -> adaptor, per utterance, on the
same bf16-rounded weights.  Tolerance (BASELINE.json: "a stated bf16 relative tolerance"): relative-to-RMS error <= 2e-2 per
utterance; the measured maxima are printed."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 2e-2


def _rel(got, want):
    got, want = got.double().cpu(), want.double()
    return float((got - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt())


@pytest.fixture(scope="module")
def stack(frontend_golden):
    import funasr_b200 as fb
    from oracle import encoder as oe

    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    enc, ad = fb.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF), fb.Transformer(**oe.NANO_ADAPTOR_CONF)
    enc.load_state_dict(o_enc.state_dict(), strict=True)
    ad.load_state_dict(o_ad.state_dict(), strict=True)
    rng = np.random.default_rng(99)
    cmvn = np.stack([-(8 + 4 * rng.random(560)), 0.1 + 0.2 * rng.random(560)]).astype(np.float32)
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=torch.from_numpy(cmvn), device="cuda")
    pipe = fb.AudioPipeline(fe, fb.AudioTower(enc.cuda(), ad.cuda()), "cuda")
    tables = {"window": frontend_golden["window_povey"], "mel": frontend_golden["mel_banks"]}
    return pipe, o_enc, o_ad, cmvn, tables


def _pack(lengths, seed):
    g = torch.Generator().manual_seed(seed)
    base = [0.5 * (2 * torch.rand(int(lengths.max()), generator=g) - 1) for _ in range(min(8, len(lengths)))]
    wavs = [base[i % len(base)][: int(n)] for i, n in enumerate(lengths)]
    starts = np.zeros(len(wavs), np.int64)
    starts[1:] = np.cumsum((lengths[:-1] + 3) // 4 * 4)
    total = int(starts[-1] + (lengths[-1] + 3) // 4 * 4)
    host = torch.zeros(total, dtype=torch.float32).pin_memory()
    for w, s0, n in zip(wavs, starts, lengths):
        host[s0:s0 + n] = w
    return wavs, starts, host


def _oracle_rows(stack, wav):
    from oracle import frontend as of

    _, o_enc, o_ad, cmvn, tables = stack
    feats, lens = of.wav_frontend([wav.numpy()], "povey", cmvn, tables=tables)
    x, il = torch.from_numpy(feats), torch.from_numpy(lens)
    y, ol = o_enc(x, il)
    z, _ = o_ad(y, ol)
    return z[0, : int(lens[0])]


def _check(stack, lengths, which, seed, label):
    pipe = stack[0]
    wavs, starts, host = _pack(lengths, seed)
    _, T = pipe.frontend.lengths(lengths)
    rows = int(T.sum())
    # (1) device-resident path = bench `value`
    out_dev, cu = pipe.embed_device(host.cuda(), starts, lengths)
    # (2) pinned host in / pinned host out with the overlapped two-half schedule = bench `e2e`
    out_host = torch.zeros((rows, 1024), dtype=torch.bfloat16).pin_memory()
    cu2 = pipe.run_pinned(host, starts, lengths, out_host, overlap=True)
    torch.cuda.synchronize()
    assert np.array_equal(cu, cu2) and cu[-1] == rows
    assert torch.equal(out_dev.cpu(), out_host), "the overlapped e2e schedule and the device-resident pass disagree"
    worst = 0.0
    for b in which:
        want = _oracle_rows(stack, wavs[b])
        e = _rel(out_host[cu[b]:cu[b + 1]].float(), want)
        worst = max(worst, e)
        assert e < REL_TOL, f"{label}: utterance {b} ({lengths[b] / 16000:.1f} s, {int(T[b])} rows): rel-rms {e:.3e}"
    print(f"{label}: {rows} rows, {len(which)} utterances checked against the fp32 oracle, worst rel-rms {worst:.3e}")


def test_cfg3_every_utterance_full_depth(stack):
    lengths = (np.random.default_rng(0).uniform(1.0, 30.0, 64) * 16000).astype(np.int64)  # bench.py workload_lengths("cfg3")
    _check(stack, lengths, range(64), seed=3, label="cfg3")


def test_cfg5_size_sampled_utterances_full_depth(stack):
    lengths = np.full(360, 160000, np.int64)
    which = np.linspace(0, 359, 16).astype(int)  # spans both halves of every two-way split
    _check(stack, lengths, which, seed=5, label="cfg5")


def test_cfg2_graph_replay_full_depth(stack):
    lengths = np.full(16, 160000, np.int64)
    _check(stack, lengths, range(16), seed=2, label="cfg2")
