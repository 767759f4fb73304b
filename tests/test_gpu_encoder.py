"""-m gpu: SAN-M encoder + adaptor through the drop-in modules (C ABI underneath) vs the fp32 CPU oracle on the same
bf16-rounded weights.  Tolerance (BASELINE.json: "a stated bf16 relative tolerance"): relative-to-RMS error
<= 2e-2 after the full stack; bf16 activations feed every GEMM, accumulation / residual stream / LN / softmax are fp32."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 2e-2


def _pair(n_blocks, tp_blocks, seed=4321):
    import funasr_b200 as fb
    from oracle import encoder as oe

    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=n_blocks, tp_blocks=tp_blocks)
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=seed)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=seed + 1)
    enc = fb.SenseVoiceEncoderSmall(560, **conf)
    ad = fb.Transformer(**oe.NANO_ADAPTOR_CONF)
    enc.load_state_dict(o_enc.state_dict(), strict=True)
    ad.load_state_dict(o_ad.state_dict(), strict=True)
    return o_enc, o_ad, enc.cuda(), ad.cuda()


def _rel(got, want):
    got, want = got.double().cpu(), want.double()
    return float((got - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt())


def _batch(lens, seed=55):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(len(lens), max(lens), 560, generator=g)
    for b, t in enumerate(lens):
        x[b, t:] = 0
    return x, torch.tensor(lens, dtype=torch.int32)


def test_golden_nano_width_vs_vllm_classes(encoder_golden):
    """Committed outputs of vLLM's own SenseVoiceEncoderSmall / Transformer (2+1 blocks, Nano widths)."""
    g = encoder_golden
    _, _, enc, ad = _pair(2, 1)
    x, lens = torch.from_numpy(g["B_x"]), torch.from_numpy(g["B_lens"])
    y, ol = enc(x.cuda(), lens)
    assert ol.dtype == torch.int32 and tuple(y.shape) == (2, 9, 512)
    for b in range(2):
        t = int(lens[b])
        assert _rel(y[b, :t], torch.from_numpy(g["B_y_pad"][b, :t])) < REL_TOL
        assert torch.all(y[b, t:] == 0)
        z, _ = ad(torch.from_numpy(g["B_y_pad"][b:b + 1, :t]).cuda(), lens[b:b + 1])
        assert _rel(z[0], torch.from_numpy(g[f"B_z_single_{b}"])) < REL_TOL


@pytest.mark.parametrize("lens", [[167, 40, 1, 98], [500], [17, 17]])
def test_encoder_adaptor_vs_oracle_ragged(lens):
    o_enc, o_ad, enc, ad = _pair(3, 2)
    x, il = _batch(lens)
    x0 = x.clone()
    want_y, _ = o_enc(x, il)
    want_z, _ = o_ad(want_y, il)
    xg = x0.cuda()
    y, ol = enc(xg, il.cuda())
    assert torch.equal(xg.cpu(), x0)  # input not modified (the reference scales it in place)
    z, zl = ad(y, ol)
    for b, t in enumerate(lens):
        e_y, e_z = _rel(y[b, :t], want_y[b, :t]), _rel(z[b, :t], want_z[b, :t])
        print(f"len {t}: encoder rel-rms {e_y:.3e}, adaptor rel-rms {e_z:.3e}")
        assert e_y < REL_TOL and e_z < REL_TOL
    assert zl.tolist() == lens


def test_fused_tower_equals_separate_calls_and_is_batch_invariant():
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    x, il = _batch([60, 33, 7])
    y, ol = enc(x.cuda(), il)
    z, _ = ad(y, ol)
    tower = fb.AudioTower(enc, ad)
    zf, _ = tower.forward_export(x.cuda(), il)
    for b, t in enumerate(il.tolist()):
        assert _rel(zf[b, :t], z[b, :t].cpu()) < 5e-3  # the fused path keeps the encoder output in bf16+fp32 workspace
    # packed == padded, and an utterance's result does not depend on its batch neighbours
    feats = torch.cat([x[b, :t] for b, t in enumerate(il.tolist())]).cuda()
    zp, cu = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
    for b, t in enumerate(il.tolist()):
        assert torch.equal(zp[cu[b]:cu[b + 1]], zf[b, :t])
    z1, _ = tower.forward_export(x[1:2, :33].cuda(), il[1:2])
    assert torch.equal(z1[0], zf[1, :33])


def test_graph_replay_is_invisible():
    """fa_plan_set_graph_rows: a batch that replays the captured CUDA graph of its (rows, utterances, longest utterance)
    bucket must produce the same bits as the launch-by-launch schedule - also when the graph was captured for ANOTHER batch
    of the bucket (row count, utterance count and boundaries all differ: they are read from cu_seqlens on the device), and
    for the separate encoder / adaptor calls with their padded fp32 hand-over."""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    tower = fb.AudioTower(enc, ad)
    batches = [[167, 40, 1, 98, 130, 5, 64, 129, 33, 255, 17, 80],   # 1019 rows, 12 utterances -> bucket (1024, 16, 256)
               [255, 200, 167, 98, 130, 64, 60, 31, 1],              # 1006 rows,  9 utterances -> same bucket
               [500, 333, 167],                                      # another max_len bucket
               [3]]
    for lens in batches:
        x, il = _batch(lens, seed=55 + len(lens))
        feats = torch.cat([x[b, :t] for b, t in enumerate(lens)]).cuda()
        tower.plan().set_graph_rows(0)
        want, cu = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
        want = want.clone()
        tower.plan().set_graph_rows(8192)
        for _ in range(2):  # capture, then replay
            got, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
            assert torch.equal(got, want), lens
        zp, _ = tower.forward_export(x.cuda(), il)  # padded in / padded out (/root/reference/model.py:251-254)
        for b, t in enumerate(lens):
            assert torch.equal(zp[b, :t], want[cu[b]:cu[b + 1]])
            assert torch.all(zp[b, t:] == 0)
    # separate encoder / adaptor calls with their own plans
    enc._shared_plan = ad._shared_plan = None
    x, il = _batch(batches[0])
    enc.plan().set_graph_rows(0)
    ad.plan().set_graph_rows(0)
    y0, ol = enc(x.cuda(), il)
    z0, _ = ad(y0, ol)
    enc.plan().set_graph_rows(8192)
    ad.plan().set_graph_rows(8192)
    for _ in range(2):
        y, _ = enc(x.cuda(), il)
        z, _ = ad(y, ol)
        assert torch.equal(y, y0) and torch.equal(z, z0)


def test_dual_stream_schedule_is_invisible():
    """fa_plan_set_streams(2): the two half-batches run on two streams (forked / joined inside the call); results are
    bit-identical to the single-stream schedule, for the fused tower and for separate encoder / adaptor calls.  (Graph
    replay off: it takes precedence for batches this small.)"""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    lens = [167] * 9 + [40, 500, 98, 130, 5, 64, 129, 333, 255, 17, 80]  # 2854 rows >= the 2048-row threshold
    x, il = _batch(lens)
    feats = torch.cat([x[b, :t] for b, t in enumerate(lens)]).cuda()
    tower = fb.AudioTower(enc, ad)
    tower.plan().set_graph_rows(0)
    tower.plan().set_streams(1)
    want, cu = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
    want = want.clone()
    tower.plan().set_streams(2)
    for _ in range(3):  # repeated calls re-use the plan's auxiliary stream / events
        got, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
        assert torch.equal(got, want)
    zp, _ = tower.forward_export(x.cuda(), il)
    for b, t in enumerate(lens):
        assert torch.equal(zp[b, :t], want[cu[b]:cu[b + 1]])
        assert torch.all(zp[b, t:] == 0)
    enc._shared_plan = ad._shared_plan = None
    enc.plan().set_graph_rows(0)
    ad.plan().set_graph_rows(0)
    enc.plan().set_streams(1)
    ad.plan().set_streams(1)
    y0, ol = enc(x.cuda(), il)
    z0, _ = ad(y0, ol)
    enc.plan().set_streams(2)
    ad.plan().set_streams(2)
    y, _ = enc(x.cuda(), il)
    z, _ = ad(y, ol)
    assert torch.equal(y, y0) and torch.equal(z, z0)


def test_chain_schedule_is_invisible():
    """fa_plan_set_chain: out-projection -> ffn1 -> ffn2 -> next qkv of a block as ONE launch with per-row-block dependency
    counters (csrc/gemm_chain.cu) against one launch per GEMM: same tile arithmetic, so bit-identical - for one partial row
    block, a ragged multi-row-block batch, the dual-stream schedule and graph replay; repeated calls re-use the counters.
    (4 + 2 blocks: layers 1, 2 and 4 are chained; 2 + 1 has no chainable layer.)"""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(4, 2)
    tower = fb.AudioTower(enc, ad)
    plan = tower.plan()
    for lens, streams, graph_rows in [([50], 1, 0), ([167] * 9 + [40, 500, 98, 130, 5, 64, 129, 333, 255, 17, 80], 1, 0),
                                      ([167] * 9 + [40, 500, 98, 130, 5, 64, 129, 333, 255, 17, 80], 2, 0),
                                      ([167] * 16, 1, 8192), ([257, 1, 255], 1, 8192)]:
        x, il = _batch(lens)
        feats = torch.cat([x[b, :t] for b, t in enumerate(lens)]).cuda()
        plan.set_graph_rows(graph_rows)
        plan.set_streams(streams)
        plan.set_chain(False)
        want, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
        want = want.clone()
        plan.set_chain(True)
        for _ in range(3):
            got, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
            assert torch.equal(got, want), (lens[:3], streams, graph_rows)


def test_full_depth_50_20_one_utterance():
    """The Nano stack (50 + 20 blocks + adaptor) on one 3 s utterance: error growth through 70 layers stays in budget."""
    o_enc, o_ad, enc, ad = _pair(50, 20)
    x, il = _batch([50], seed=9)
    want_y, _ = o_enc(x, il)
    want_z, _ = o_ad(want_y, il)
    y, ol = enc(x.cuda(), il)
    z, _ = ad(y, ol)
    e_y, e_z = _rel(y[0], want_y[0]), _rel(z[0], want_z[0])
    print(f"70 layers: encoder rel-rms {e_y:.3e}, adaptor rel-rms {e_z:.3e}")
    assert e_y < REL_TOL and e_z < REL_TOL


def test_state_dict_reload_invalidates_plan():
    o_enc, _, enc, _ = _pair(2, 0)
    x, il = _batch([20])
    y0, _ = enc(x.cuda(), il)
    sd = {k: v.clone() for k, v in enc.state_dict().items()}
    sd["after_norm.bias"] += 1.0
    enc.load_state_dict(sd)
    y1, _ = enc(x.cuda(), il)
    assert not torch.equal(y0, y1)


def test_overlapped_pipeline_is_invisible():
    """AudioPipeline.run_pinned(overlap=True): two half-batches, each with its own stream and workspace (Plan.twin shares the
    weights), H2D of the second half and D2H of the first half under the other half's compute.  Rows must be bit-identical
    to the single-pass schedule, call after call."""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=None, device="cuda")
    g = torch.Generator().manual_seed(8)
    n = np.array([16000, 48000, 160000, 8000, 24007, 33333, 100000, 400, 64000, 16001], np.int64)
    starts = np.zeros(len(n), np.int64)
    starts[1:] = np.cumsum((n[:-1] + 3) // 4 * 4)
    total = int(starts[-1] + (n[-1] + 3) // 4 * 4)
    host = torch.zeros(total).pin_memory()
    for s0, k in zip(starts, n):
        host[s0:s0 + k] = 0.5 * (2 * torch.rand(int(k), generator=g) - 1)
    pipe = fb.AudioPipeline(fe, fb.AudioTower(enc, ad), "cuda")
    _, T = fe.lengths(n)
    rows = int(T.sum())
    want = torch.empty((rows, 1024), dtype=torch.bfloat16).pin_memory()
    cu0 = pipe.run_pinned(host, starts, n, want, overlap=False)
    torch.cuda.synchronize()
    assert 0 < pipe._cut(T) < len(n)
    for _ in range(3):
        got = torch.zeros((rows, 1024), dtype=torch.bfloat16).pin_memory()
        cu1 = pipe.run_pinned(host, starts, n, got, overlap=True)
        torch.cuda.synchronize()
        assert np.array_equal(cu0, cu1)
        assert torch.equal(got, want)
    for pieces in (3, 4, 64):  # more pieces than two: alternating streams, copies back on a stream of their own
        got = torch.zeros((rows, 1024), dtype=torch.bfloat16).pin_memory()
        cu1 = pipe.run_pinned(host, starts, n, got, overlap=True, pieces=pieces)
        torch.cuda.synchronize()
        assert np.array_equal(cu0, cu1)
        assert torch.equal(got, want), f"pieces={pieces}"


def test_pipeline_int16_host_pcm_matches_float_host_pcm():
    """AudioPipeline.run_pinned on int16 host PCM (what a 16 kHz wav file holds): the GPU-side x / 32768 gives the very f32 samples
    ``torchaudio.load`` hands the reference, so the embeddings are bit-identical to the run on the f32 host buffer, in the
    single-pass and in the overlapped schedule, call after call (the starts / lengths buffer is re-used)."""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=None, device="cuda")
    g = torch.Generator().manual_seed(11)
    n = np.array([16000, 48000, 8000, 24007, 33333, 400, 64000, 16001], np.int64)
    starts = np.zeros(len(n), np.int64)
    starts[1:] = np.cumsum((n[:-1] + 7) // 8 * 8)  # 16-byte aligned for both sample types
    total = int(starts[-1] + (n[-1] + 7) // 8 * 8)
    raw = torch.zeros(total, dtype=torch.int16).pin_memory()
    for s0, k in zip(starts, n):
        raw[s0:s0 + k] = torch.randint(-32768, 32768, (int(k),), generator=g).to(torch.int16)
    raw[int(starts[1])] = -32768
    raw[int(starts[1]) + 1] = 32767
    host = (raw.float() / 32768.0).pin_memory()  # the reference's conversion (torchaudio.load normalisation), exact in f32
    pipe = fb.AudioPipeline(fe, fb.AudioTower(enc, ad), "cuda")
    _, T = fe.lengths(n)
    rows = int(T.sum())
    want = torch.empty((rows, 1024), dtype=torch.bfloat16).pin_memory()
    cu0 = pipe.run_pinned(host, starts, n, want, overlap=False)
    torch.cuda.synchronize()
    for overlap in (False, True, True, False):
        got = torch.zeros((rows, 1024), dtype=torch.bfloat16).pin_memory()
        cu1 = pipe.run_pinned(raw, starts, n, got, overlap=overlap)
        torch.cuda.synchronize()
        assert np.array_equal(cu0, cu1)
        assert torch.equal(got, want), f"overlap={overlap}"
    with pytest.raises(TypeError):
        pipe.run_pinned(host.double(), starts, n, want)


def test_streaming_pipeline_matches_single_batches():
    """StreamingPipeline: batches submitted back to back on two lanes (copies of one batch under the compute of another) give the
    rows of AudioPipeline.run_pinned on each batch alone, f32 and int16 host PCM, whatever the lane."""
    import funasr_b200 as fb

    _, _, enc, ad = _pair(2, 1)
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=None, device="cuda")
    tower = fb.AudioTower(enc, ad)
    g = torch.Generator().manual_seed(21)
    batches = []
    for n in ([16000, 48000, 8000, 24007, 33333, 400, 64000, 16001], [160000, 9000, 31999], [4000] * 9):
        n = np.array(n, np.int64)
        starts = np.zeros(len(n), np.int64)
        starts[1:] = np.cumsum((n[:-1] + 7) // 8 * 8)
        total = int(starts[-1] + (n[-1] + 7) // 8 * 8)
        raw = torch.zeros(total, dtype=torch.int16).pin_memory()
        for s0, k in zip(starts, n):
            raw[s0:s0 + k] = torch.randint(-20000, 20000, (int(k),), generator=g).to(torch.int16)
        batches.append((raw, (raw.float() / 32768.0).pin_memory(), starts, n))
    single = fb.AudioPipeline(fe, tower, "cuda")
    want = []
    for raw, host, starts, n in batches:
        rows = int(fe.lengths(n)[1].sum())
        w = torch.empty((rows, 1024), dtype=torch.bfloat16).pin_memory()
        single.run_pinned(host, starts, n, w, overlap=False)
        torch.cuda.synchronize()
        want.append(w.clone())
    sp = fb.StreamingPipeline(fe, tower, "cuda")
    own = [p._own_plan() for p in sp.pipes]  # no workspace is shared between lanes, nor with the tower's own plan
    assert own[0] is not own[1] and tower.plan() not in own
    jobs = []
    for rep in range(3):
        for b, (raw, host, starts, n) in enumerate(batches):
            got = torch.zeros_like(want[b]).pin_memory()
            src = raw if (rep + b) % 2 else host
            ev = sp.submit(src, starts, n, got, overlap=bool(rep % 2) if len(n) >= 2 else False)
            jobs.append((b, got, ev))
    for b, got, ev in jobs:
        ev.synchronize()
        assert torch.equal(got, want[b]), f"batch {b}"
    sp.drain()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process():
    """kernel attributes (dynamic shared memory opt-in) and the SM count are per DEVICE; one process that drives two
    GPUs must get the same results on the second one (the frontend, the plan and every kernel class run on cuda:1 after
    cuda:0 has used them)."""
    import funasr_b200 as fb
    from oracle import encoder as oe

    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    g = torch.Generator().manual_seed(3)
    wavs = [0.5 * (2 * torch.rand(n, generator=g) - 1) for n in (16000, 32000, 8000)]
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        enc, ad = fb.SenseVoiceEncoderSmall(560, **conf), fb.Transformer(**oe.NANO_ADAPTOR_CONF)
        enc.load_state_dict(o_enc.state_dict())
        ad.load_state_dict(o_ad.state_dict())
        enc.to(dev), ad.to(dev)
        fe = fb.WavFrontend(window="povey", lfr_m=7, lfr_n=6, dither=0.0, device=dev)
        with torch.cuda.device(dev):
            feats, lens = fb.extract_fbank(wavs, frontend=fe, is_final=True)
            y, ol = enc(feats, lens)
            z, _ = ad(y, ol)
            torch.cuda.synchronize()
        assert z.device == torch.device(dev)
        outs.append(z.cpu())
    assert torch.equal(outs[0], outs[1])
