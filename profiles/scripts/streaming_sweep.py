"""Streaming loop (funasr_b200.StreamingPipeline) at cfg5 / cfg3 / cfg2: lanes x schedule inside a lane; ms per batch, end to end.
Usage (GPU box): python profiles/scripts/streaming_sweep.py  ->  gpurun_out/streaming_sweep.json"""
import sys, os, json
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "fun-asr-vllm_b200")]
import numpy as np, torch
import bench as B
import funasr_b200 as fb
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
enc = B.random_init_(fb.SenseVoiceEncoderSmall(560, **B.ENC_CONF), 4321).to(dev)
ad = B.random_init_(fb.Transformer(**B.AD_CONF), 4322).to(dev)
fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                    cmvn=torch.from_numpy(B.make_cmvn()), device=dev)
tower = fb.AudioTower(enc, ad, max_tokens=60160)
res = {}
for name, K in (("cfg5", 8), ("cfg3", 24), ("cfg2", 48)):
    inp = B.Inputs(B.workload_lengths(name), 2000, dev, 1024)
    for lanes in (2, 3):
        for ov in (None, False):
            sp = fb.StreamingPipeline(fe, tower, dev, lanes=lanes, overlap=ov)
            outs = [torch.empty_like(inp.out_host).pin_memory() for _ in range(lanes)]
            def run(k):
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for i in range(k):
                    sp.submit(inp.host_pcm, inp.starts, inp.lengths, outs[i % lanes])
                for st in sp.streams:
                    torch.cuda.current_stream().wait_stream(st)
                b.record(); torch.cuda.synchronize()
                return a.elapsed_time(b) / k
            run(2 * lanes)
            res[f"{name} lanes={lanes} overlap={ov}"] = round(min(run(K), run(K)), 3)
            del sp, outs
    print(name, {k: v for k, v in res.items() if k.startswith(name)}, flush=True)
json.dump(res, open("gpurun_out/streaming_sweep.json", "w"), indent=1)
