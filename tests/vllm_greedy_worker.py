"""cfg4 of BASELINE.json, run as a separate process by tests/test_gpu_zz_vllm_greedy.py:

    fbank -> SAN-M encoder -> adaptor -> splice(prefix, speech, suffix) -> vLLM Qwen3-0.6B `prompt_embeds` greedy decode

exactly as the reference's batched loop hands over (infer.py:381-388 engine, :462-475 hand-off), once with the prompt
embeddings of the CUDA path and once with those of the fp32 CPU oracle, same random-init weights on both sides.  There is
no checkpoint and no tokenizer on the box, so the LLM is a random-init Qwen3-0.6B (HF initialiser, seed 7) written to a
temporary directory, the prefix / suffix rows are embedding-table rows of fixed random token ids, and the engine runs with
skip_tokenizer_init.  Prints ONE JSON line with the greedy-token agreement:
    first_token   fraction of utterances whose first generated token agrees
    prefix_mean   mean length of the common token prefix (out of max_tokens)
    control_*     the same for the oracle embeddings submitted twice (different batch slots) = the engine's own
                  batch-position noise floor
and, so that the claim is well conditioned instead of a loose percentage, the MARGIN analysis: the same random-init
Qwen3-0.6B is run by HF transformers in fp32 on both embedding sets, teacher-forced on the tokens vLLM generated from the
oracle embeddings.  Per utterance and generated position it records the oracle's top-1 / top-2 logit margin and the shift
of that very gap caused by swapping in the CUDA path's embeddings (the measured logit perturbation).  A greedy decision can
only flip where the margin is below the perturbation, so:
    theta                = max gap shift over all positions (fp32) + 3 x the engine's own bf16-vs-fp32 gap noise
    decidable_fraction   positions whose oracle margin exceeds theta
    mismatches           first diverging position of every utterance whose vLLM tokens differ: (utterance, position,
                         oracle margin, gap shift) - each must sit at a margin <= theta (a near-tie of the random LLM)
    violations           mismatches at a decidable position: must be empty (asserted by the test)
It also times the prompt_embeds hop of the hand-off: prefill of the same 16 prompts submitted as
prompt_embeds (D2H + msgpack + IPC + H2D, VLLM/renderers/base.py:755-758) against token-id prompts of the same lengths.
"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fun-asr-vllm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_UTT = int(os.environ.get("FA_CFG4_UTTS", "16"))
SECONDS = float(os.environ.get("FA_CFG4_SECONDS", "10"))
MAX_TOKENS = int(os.environ.get("FA_CFG4_MAX_TOKENS", "16"))
FULL = os.environ.get("FA_CFG4_SMALL", "0") != "1"


def build_llm_dir(path: str, seed: int = 7):
    """Random-init Qwen3-0.6B (the consumer named by BASELINE.json cfg4) in HF format."""
    import torch
    from transformers import Qwen3Config, Qwen3ForCausalLM

    cfg = Qwen3Config(vocab_size=151936, hidden_size=1024, intermediate_size=3072, num_hidden_layers=28,
                      num_attention_heads=16, num_key_value_heads=8, head_dim=128, max_position_embeddings=40960,
                      rms_norm_eps=1e-6, tie_word_embeddings=True)
    torch.manual_seed(seed)
    m = Qwen3ForCausalLM(cfg).to(torch.bfloat16)
    table = m.get_input_embeddings().weight.detach().clone()
    m.save_pretrained(path, safe_serialization=True)
    return table


def main():
    import numpy as np
    import torch

    import funasr_b200 as fb
    from oracle import encoder as oe
    from oracle import frontend as of

    t00 = time.time()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    enc_conf = dict(oe.NANO_ENCODER_CONF) if FULL else dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **enc_conf), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    enc, ad = fb.SenseVoiceEncoderSmall(560, **enc_conf), fb.Transformer(**oe.NANO_ADAPTOR_CONF)
    enc.load_state_dict(o_enc.state_dict())
    ad.load_state_dict(o_ad.state_dict())
    enc.to(dev), ad.to(dev)
    rng = np.random.default_rng(99)
    cmvn = np.stack([-(8 + 4 * rng.random(560)), 0.1 + 0.2 * rng.random(560)]).astype(np.float32)
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=torch.from_numpy(cmvn), device=dev)
    g = torch.Generator().manual_seed(1234)
    n = int(SECONDS * 16000)
    wavs = [0.5 * (2 * torch.rand(n, generator=g) - 1) for _ in range(N_UTT)]

    # ---- the CUDA path, through the reference-facing calls (infer.py:446-458)
    feats, lens = fb.extract_fbank(wavs, frontend=fe, is_final=True)
    y, ol = enc(feats, lens)
    z, zl = ad(y, ol)
    torch.cuda.synchronize()
    # ---- the fp32 CPU oracle on the same inputs and weights
    want_f, want_l = of.wav_frontend([w.numpy() for w in wavs], "povey", cmvn)
    with torch.inference_mode():
        want_y, _ = o_enc(torch.from_numpy(want_f), torch.from_numpy(want_l))
        want_z, _ = o_ad(want_y, torch.from_numpy(want_l))
    t_audio = time.time() - t00

    d = tempfile.mkdtemp(prefix="fa_cfg4_llm_")
    table = build_llm_dir(d)
    ids = np.random.default_rng(5).integers(0, 151936, size=24)
    prefix, suffix = table[ids[:19]].to(dev), table[ids[19:]].to(dev)  # 19 + 5 rows, like the reference's chat template
    cu = np.zeros(N_UTT + 1, np.int32)
    cu[1:] = np.cumsum(want_l)
    ours_packed = torch.cat([z[b, :t].to(torch.bfloat16) for b, t in enumerate(want_l.tolist())])
    flat, off = fb.splice_embeds(ours_packed, cu, prefix, suffix)  # fa_splice_embeds (infer.py:462-466)
    flat = flat.cpu()
    ours = [flat[int(off[b]):int(off[b + 1])].clone() for b in range(N_UTT)]
    ref = [torch.cat([prefix.cpu(), want_z[b, :t].to(torch.bfloat16), suffix.cpu()]) for b, t in enumerate(want_l.tolist())]
    rel = [float((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt()) for a, b in zip(ours, ref)]
    del enc, ad, fe, z, y, feats
    torch.cuda.empty_cache()

    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    os.environ.setdefault("VLLM_WORKER_MULTIPROC_METHOD", "spawn")
    from vllm import LLM, SamplingParams

    t0 = time.time()
    llm = LLM(model=d, enable_prompt_embeds=True, gpu_memory_utilization=0.4, dtype="bfloat16", skip_tokenizer_init=True,
              enforce_eager=True, max_model_len=2048, seed=0)
    sp = SamplingParams(temperature=0.0, max_tokens=MAX_TOKENS, detokenize=False, ignore_eos=True, logprobs=2)
    t_engine = time.time() - t0
    print("FA_CFG4_ENGINE_STARTED", flush=True)  # from here on a failure is a failure of the hand-off, not of the box
    prompts = [{"prompt_embeds": e} for e in ours] + [{"prompt_embeds": e} for e in ref] + [{"prompt_embeds": e} for e in ref]
    outs = llm.generate(prompts, sp, use_tqdm=False)
    toks = [list(o.outputs[0].token_ids) for o in outs]
    a, b, c = toks[:N_UTT], toks[N_UTT:2 * N_UTT], toks[2 * N_UTT:]

    def vllm_gaps(o):
        """top-1 minus top-2 logprob (= logit gap) per generated position, in the engine's own bf16 arithmetic"""
        gaps = []
        for pos in o.outputs[0].logprobs:
            v = sorted((lp.logprob for lp in pos.values()), reverse=True)
            gaps.append(v[0] - v[1] if len(v) >= 2 else float("nan"))
        return gaps

    gaps_vllm_ref = [vllm_gaps(o) for o in outs[N_UTT:2 * N_UTT]]

    # ---- the prompt_embeds hop: prefill-only generate, embeds vs token ids of the same lengths
    sp1 = SamplingParams(temperature=0.0, max_tokens=1, detokenize=False, ignore_eos=True)
    rng_ids = np.random.default_rng(11)
    id_prompts = [{"prompt_token_ids": rng_ids.integers(0, 151936, size=int(e.shape[0])).tolist()} for e in ours]
    flat_dev = flat.to(dev)

    def hop_embeds():
        t = time.perf_counter()
        host = flat_dev.cpu()  # the D2H the reference's hand-off pays (vLLM forces CPU tensors)
        llm.generate([{"prompt_embeds": host[int(off[i]):int(off[i + 1])]} for i in range(N_UTT)], sp1, use_tqdm=False)
        return time.perf_counter() - t

    def hop_ids():
        t = time.perf_counter()
        llm.generate(id_prompts, sp1, use_tqdm=False)
        return time.perf_counter() - t

    for f in (hop_embeds, hop_ids):
        f()
    t_emb = min(hop_embeds() for _ in range(5))
    t_ids = min(hop_ids() for _ in range(5))

    # ---- margin analysis in fp32 (HF transformers), teacher-forced on the tokens vLLM produced from the ORACLE embeddings
    from transformers import Qwen3ForCausalLM

    hf = Qwen3ForCausalLM.from_pretrained(d, torch_dtype=torch.float32).to(dev).eval()
    tab32 = hf.get_input_embeddings().weight.detach()
    margins, shifts, hf_top1_is_vllm = [], [], []
    with torch.inference_mode():
        for u in range(N_UTT):
            gen = torch.tensor(b[u][:-1], dtype=torch.long, device=dev)
            tail = tab32[gen]
            lr = hf(inputs_embeds=torch.cat([ref[u].to(dev).float(), tail])[None]).logits[0, -MAX_TOKENS:].double()
            lo = hf(inputs_embeds=torch.cat([ours[u].to(dev).float(), tail])[None]).logits[0, -MAX_TOKENS:].double()
            top = lr.topk(2, dim=-1)
            m = (top.values[:, 0] - top.values[:, 1])
            g_ours = lo.gather(1, top.indices[:, :1])[:, 0] - lo.gather(1, top.indices[:, 1:2])[:, 0]
            margins.append(m.cpu().numpy())
            shifts.append((g_ours - m).abs().cpu().numpy())
            hf_top1_is_vllm.append((top.indices[:, 0].cpu().numpy() == np.asarray(b[u])))
    margins, shifts = np.stack(margins), np.stack(shifts)
    gv = np.asarray(gaps_vllm_ref, dtype=np.float64)
    same_top1 = np.stack(hf_top1_is_vllm)
    noise = float(np.nanmax(np.where(same_top1, np.abs(gv - margins), np.nan))) if same_top1.any() else float("nan")
    tau = float(shifts.max())
    theta = tau + 3.0 * (noise if np.isfinite(noise) else 0.0)
    mismatches, violations = [], []
    for u in range(N_UTT):
        k = next((t for t in range(MAX_TOKENS) if a[u][t] != b[u][t]), None)
        if k is None:
            continue
        rec = {"utterance": u, "position": k, "oracle_margin": float(margins[u, k]), "gap_shift": float(shifts[u, k]),
               "engine_gap": float(gv[u, k])}
        mismatches.append(rec)
        if margins[u, k] > theta:
            violations.append(rec)
    del hf

    def agree(x, y):
        first = float(np.mean([p[0] == q[0] for p, q in zip(x, y)]))
        pre = []
        for p, q in zip(x, y):
            k = 0
            while k < min(len(p), len(q)) and p[k] == q[k]:
                k += 1
            pre.append(k)
        return first, float(np.mean(pre)), float(np.mean([k == MAX_TOKENS for k in pre]))

    f1, p1, e1 = agree(a, b)
    f0, p0, e0 = agree(b, c)
    print(json.dumps({"cfg4": "fbank->encoder->adaptor->splice->vLLM Qwen3-0.6B prompt_embeds greedy", "utterances": N_UTT,
                      "seconds_each": SECONDS, "max_tokens": MAX_TOKENS, "encoder_blocks": [enc_conf["num_blocks"], enc_conf["tp_blocks"]],
                      "first_token": f1, "prefix_mean": p1, "exact": e1, "control_first_token": f0, "control_prefix_mean": p0,
                      "control_exact": e0, "embed_rel_rms_max": max(rel), "prompt_rows": [int(e.shape[0]) for e in ours[:4]],
                      "margin_analysis": {"llm": "HF Qwen3ForCausalLM fp32, teacher-forced on the engine's oracle-side tokens",
                                          "positions": int(margins.size), "oracle_margin_median": float(np.median(margins)),
                                          "oracle_margin_p10": float(np.percentile(margins, 10)),
                                          "gap_shift_median": float(np.median(shifts)), "gap_shift_max": tau,
                                          "engine_bf16_gap_noise_max": noise, "theta": theta,
                                          "decidable_fraction": float((margins > theta).mean()),
                                          "hf_top1_equals_engine_token": float(same_top1.mean()),
                                          "mismatches": mismatches, "violations": violations},
                      "handoff_hop": {"prompts": N_UTT, "rows_total": int(off[-1]), "bytes": int(flat.numel() * 2),
                                      "prefill_via_prompt_embeds_ms": round(1e3 * t_emb, 2), "prefill_via_token_ids_ms": round(1e3 * t_ids, 2),
                                      "hop_ms": round(1e3 * (t_emb - t_ids), 2),
                                      "note": "best of 5; embeds leg = D2H of the spliced rows + llm.generate(prompt_embeds), max_tokens 1"},
                      "audio_path_s": round(t_audio, 1), "engine_start_s": round(t_engine, 1)}), flush=True)


if __name__ == "__main__":
    main()
