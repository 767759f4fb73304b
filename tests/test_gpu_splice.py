"""-m gpu: fa_splice_embeds against the reference's own formulation, BIT-EXACT.

The reference builds one prompt_embeds tensor per utterance with
``torch.cat((prefix_embedding, encoder_out[i, :lens[i], :], suffix_embedding), dim=0)`` (/root/reference/infer.py:462-466) and,
in the single-utterance path, takes ``fake_token_len`` rows of the adaptor output (model.py:391-396, 513-541).  This is index
and byte work: rows must come out with exactly the source bits (an f32 -> bf16 narrowing, where asked for, is the same
round-to-nearest-even ``Tensor.to(torch.bfloat16)`` performs)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _want(speech, cu, prefix, suffix, take, out_dtype):
    parts = []
    for b in range(len(cu) - 1):
        n = cu[b + 1] - cu[b] if take is None else min(max(int(take[b]), 0), cu[b + 1] - cu[b])
        rows = speech[cu[b]:cu[b] + n].to(out_dtype)  # infer.py:464 encoder_out[i, :lens[i], :]
        cat = [t for t in (prefix, rows, suffix) if t is not None]
        parts.append(torch.cat(cat, dim=0))  # infer.py:465
    return parts


@pytest.mark.parametrize("speech_dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("P,S", [(7, 5), (0, 3), (4, 0), (0, 0)])
@pytest.mark.parametrize("use_take", [False, True])
def test_splice_bit_exact_vs_torch_cat(speech_dtype, out_dtype, P, S, use_take):
    import funasr_b200 as fb
    from funasr_b200.splice import splice_embeds

    lens = [167, 1, 40, 500, 2, 98]  # includes 1-row and 2-row utterances (fake_token_len(1) = 1, (2) = 1)
    cu = np.zeros(len(lens) + 1, np.int32)
    cu[1:] = np.cumsum(lens)
    g = torch.Generator().manual_seed(P * 10 + S)
    D = 1024
    speech = (torch.randn(int(cu[-1]), D, generator=g) * 3).to(speech_dtype).cuda()
    prefix = torch.randn(P, D, generator=g).to(out_dtype).cuda() if P else None
    suffix = torch.randn(S, D, generator=g).to(out_dtype).cuda() if S else None
    take = [fb.fake_token_len(t, True) for t in lens] if use_take else None  # model.py:391-396
    if use_take:
        assert take == [21, 1, 5, 63, 1, 13]
    flat, off = splice_embeds(speech, cu, prefix, suffix, take=take, out_dtype=out_dtype)
    torch.cuda.synchronize()
    want = _want(speech, cu, prefix, suffix, take, out_dtype)
    assert off[-1] == flat.shape[0] == sum(w.shape[0] for w in want)
    for b, w in enumerate(want):
        got = flat[off[b]:off[b + 1]]
        assert got.shape == w.shape and got.dtype == out_dtype
        assert torch.equal(got, w), f"utterance {b} differs from torch.cat"


def test_splice_take_is_clamped_to_the_utterance():
    """model.py:527-539: a speech_token_len larger than the rows the adaptor produced falls back to the rows that exist."""
    from funasr_b200.splice import splice_embeds

    cu = np.array([0, 3, 10], np.int32)
    speech = torch.arange(10 * 8, dtype=torch.float32).reshape(10, 8).cuda()
    flat, off = splice_embeds(speech, cu, None, None, take=[99, -4], out_dtype=torch.float32)
    assert off.tolist() == [0, 3, 3]
    assert torch.equal(flat, speech[:3])
