"""First-light diagnostic: CUDA path vs oracle, stage by stage, on a few seeded clips."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
import __graft_entry__ as g
g.build()
from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
from audio_feature_extraction_b200.synth import make_clip
from oracle import OracleFeatureExtractor
import oracle as O

def main():
    secs = [2.0, 5.0, 3.3, 10.0]
    clips = [make_clip(s, seconds=sec) for s, sec in enumerate(secs)]
    ex = AudioFeatureExtractor()
    ora = OracleFeatureExtractor()
    buf, offsets, lengths = ex.pack_clips(clips)
    audio = buf.cuda()
    groups = ("mfcc", "energy")
    summ, meta, frames = ex.extract_batch_device(audio, offsets, lengths, groups, return_frames=True)
    torch.cuda.synchronize()
    summ = summ.cpu().numpy(); meta = meta.cpu().numpy(); frames = frames.cpu().numpy()
    ws = ex._ws
    hb = N.lib().afx_debug_header_bytes(len(clips))
    tot = sum(ex.max_frames(l) for l in lengths)
    logmel = ws[hb:hb + tot * 128 * 4].view(torch.float32).cpu().numpy().reshape(tot, 128)
    n = ex.n_mfcc
    row = 0
    for i, y in enumerate(clips):
        s, e = ora.trim_indices(y)
        yt = ora.preprocess_audio(y)
        print(f"clip {i}: len {len(y)} oracle trim ({s},{e}) gpu meta {meta[i].tolist()}")
        m, d1, d2 = ora.mfcc_frames(yt)
        _, D = O.mfcc(yt, sr=ex.sr, n_mfcc=n, n_fft=ex.frame_length, hop_length=ex.hop_length, return_logmel=True)
        M = O.melspectrogram(yt, ex.sr, ex.frame_length, ex.hop_length)
        Draw = 10 * np.log10(np.maximum(1e-10, M))
        r = ora.rms_frames(yt)[0]
        T = m.shape[1]
        Tg = int(meta[i, 3])
        if Tg != T:
            print("  !! frame count mismatch", Tg, T)
        Tm = min(T, Tg)
        fr = frames[row:row + Tm]
        lm = logmel[row:row + Tm]
        dl = np.abs(lm.T - Draw[:, :Tm])
        print(f"  logmel(raw) max err {dl.max():.3e} at {np.unravel_index(dl.argmax(), dl.shape)} mean {dl.mean():.3e}; oracle max {Draw.max():.3f} min {Draw.min():.3f}")
        print(f"  mfcc frames max err {np.abs(fr[:, :n].T - m[:, :Tm]).max():.3e}")
        print(f"  delta  max err {np.abs(fr[:, n:2*n].T - d1[:, :Tm]).max():.3e}  delta2 {np.abs(fr[:, 2*n:3*n].T - d2[:, :Tm]).max():.3e}")
        rel = np.abs(fr[:, 3 * n] - r[:Tm]) / np.maximum(np.abs(r[:Tm]), 1e-12)
        print(f"  rms max rel err {rel.max():.3e}")
        ref = ora.extract_features_array(y, features_to_extract=groups)
        got = ex.rows_to_dicts(summ[i:i+1], meta[i:i+1], ["x"], groups)[0]
        for k in ("mfcc_mean", "mfcc_std", "mfcc_delta_mean", "mfcc_delta2_mean"):
            print(f"  {k} max abs err {np.max(np.abs(np.asarray(got[k]) - np.asarray(ref[k]))):.3e}")
        for k in ("energy_mean", "energy_std", "energy_range"):
            print(f"  {k} rel err {abs(got[k]-ref[k])/abs(ref[k]):.3e}")
        row += ex.max_frames(lengths[i])
    # quick timing
    from audio_feature_extraction_b200.synth import make_batch_torch
    nclip = 2000
    a, off, ln = make_batch_torch(nclip, device="cuda")
    torch.cuda.synchronize()
    for it in range(3):
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        sm, mt, _ = ex.extract_batch_device(a, off.numpy(), ln.numpy(), groups)
        t1.record(); torch.cuda.synchronize()
        ms = t0.elapsed_time(t1)
        hours = nclip * 10.0 / 3600
        print(f"timing iter {it}: {ms:.2f} ms for {nclip} clips -> {hours / (ms / 1e3):.1f} audio-hours/s; status ok {(mt[:,0]==0).sum().item()}")
    print("launches", N.lib().afx_launch_count())

main()
