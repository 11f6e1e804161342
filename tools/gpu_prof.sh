#!/bin/bash
# final profile set: ncu --set full of the streaming kernel (C2), yin_obs_kernel, viterbi_kernel (full rounds), the tcgen05
# prototype (tensor pipe), and the throughput of the run-time-hop variant beside the stock one
set -u
mkdir -p gpurun_out
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
python - <<'PY' > gpurun_out/runtime_hop.jsonl 2>gpurun_out/runtime_hop.err
import json, numpy as np, torch
from audio_feature_extraction_b200 import AudioFeatureExtractor
from audio_feature_extraction_b200.synth import make_batch_torch
dev = torch.device("cuda", 0)
for sr, fl, hops in ((22050, 1024, (256, 200, 160)), (16000, 512, (128, 160)), (44100, 2048, (512, 441))):
    audio, off, ln = make_batch_torch(4000, sr=sr, seconds=10.0, seed=3, device=dev, align=8)
    off, ln = off.numpy(), ln.numpy()
    for hop in hops:
        ex = AudioFeatureExtractor(sr=sr, frame_length=fl, hop_length=hop, n_mfcc=13)
        for _ in range(2): ex.extract_batch_device(audio, off, ln, ("mfcc", "energy"))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): s, m, _ = ex.extract_batch_device(audio, off, ln, ("mfcc", "energy"))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        frames = int(m[:, 3].sum().item())
        print(json.dumps({"sr": sr, "frame_length": fl, "hop_length": hop, "clips": 4000, "ms": ms, "audio_hours_per_s": 4000 * 10 / 3600 / (ms * 1e-3),
                          "ns_per_frame": ms * 1e6 / frames, "clips_ok": int((m[:, 0] == 0).sum().item())}))
    del audio
PY
echo "runtime hop exit $?"; cat gpurun_out/runtime_hop.jsonl
SMALL="python bench.py --clips 2368 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 1 -c 1 -f -o gpurun_out/prof_stream_kernel $SMALL > gpurun_out/ncu_full.log 2>&1; echo "ncu stream exit $?"
F0="python bench.py --clips 1480 --steps 1 --warmup 1 --features f0,mfcc,energy --no-e2e --no-cpu-baseline --no-legs"
$F0 > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none -k regex:"yin_obs_kernel|viterbi_kernel" -s 0 -c 2 -f -o gpurun_out/prof_pyin $F0 > gpurun_out/ncu_f0.log 2>&1; echo "ncu pyin exit $?"
timeout 120 tools/tc_dft/tc_dft 18944 8 > gpurun_out/tc_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:tc_dft_kernel -s 1 -c 1 -f -o gpurun_out/prof_tc_dft tools/tc_dft/tc_dft 18944 8 > gpurun_out/ncu_tc.log 2>&1; echo "ncu tc exit $?"
