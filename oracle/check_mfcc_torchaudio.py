"""ORACLE cross-check (test infrastructure): oracle/mfcc.py (restatement of python_speech_features 0.6, which is absent from this image)
against torchaudio 2.11 — the installed second source SURVEY.md §8f N3 names — stage by stage:
  1. framing + |rfft|^2 / nfft   vs  torch.stft(window = ones, center = False)                       (same definition: must agree to rounding)
  2. DCT-II ortho, 13 x 26        vs  torchaudio.functional.create_dct(13, 26, norm="ortho")          (same definition: must agree to rounding)
  3. 26 mel filters (HTK formula) vs  torchaudio.functional.melscale_fbanks(norm=None, mel_scale="htk")
     python_speech_features places the triangle corners on FLOORED FFT bins, torchaudio on the exact frequencies: a documented difference
     of the two libraries, recorded as a residual, not an error of the restatement
  4. cepstra and final image with torchaudio's filterbank substituted                                  (residual caused by 3. alone)
Run here (needs /root/reference for the three assets): python oracle/check_mfcc_torchaudio.py  -> tests/golden/mfcc_torchaudio_residual.json
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mfcc as om  # noqa: E402


def stages(signal, samplerate):
    """Returns the residuals of stages 1-4 for one signal (float64 numpy [n] or [n, 2])."""
    import torchaudio.functional as AF

    nfft = om.calculate_nfft(samplerate, 0.025)
    sig = np.asarray(signal)
    pre = np.append(sig[0], sig[1:] - 0.97 * sig[:-1]).astype(np.float64)
    frame_len, frame_step, numframes = om.frame_geometry(len(pre), samplerate)
    padlen = int((numframes - 1) * frame_step + frame_len)
    pad = np.concatenate((pre, np.zeros(padlen - len(pre))))
    # 1. power spectrum
    idx = np.arange(frame_len)[None, :] + (np.arange(numframes) * frame_step)[:, None]
    pspec_o = 1.0 / nfft * np.square(np.absolute(np.fft.rfft(pad[idx], nfft)))
    win = torch.zeros(nfft, dtype=torch.float64)
    win[:frame_len] = 1.0  # rectangular window of frame_len samples, zero-padded to nfft exactly as rfft(frames, nfft) does
    x = torch.from_numpy(np.concatenate((pad, np.zeros(nfft))))  # room for the last frame's zero padding
    st = torch.stft(x, n_fft=nfft, hop_length=frame_step, win_length=nfft, window=win, center=False, return_complex=True)[:, :numframes]
    pspec_t = (st.abs() ** 2 / nfft).T.numpy()
    r1 = float(np.abs(pspec_o - pspec_t).max() / max(np.abs(pspec_o).max(), 1e-300))
    # 2. DCT
    from scipy.fftpack import dct

    z = np.random.default_rng(0).standard_normal((7, 26))
    d_o = dct(z, type=2, axis=1, norm="ortho")[:, :13]
    d_t = z @ AF.create_dct(13, 26, norm="ortho").double().numpy()
    r2 = float(np.abs(d_o - d_t).max() / np.abs(d_o).max())
    # 3. filterbank
    fb_o = om.get_filterbanks(26, nfft, samplerate)
    fb_t = AF.melscale_fbanks(nfft // 2 + 1, 0.0, samplerate / 2.0, 26, samplerate, norm=None, mel_scale="htk").double().numpy().T
    r3 = float(np.linalg.norm(fb_o - fb_t) / np.linalg.norm(fb_o))
    # 4. cepstra / image with torchaudio's filterbank substituted
    cep_o = om.psf_mfcc(signal, samplerate)
    energy = np.where(pspec_o.sum(1) == 0, np.finfo(float).eps, pspec_o.sum(1))
    feat = pspec_o @ fb_t.T
    feat = np.log(np.where(feat == 0, np.finfo(float).eps, feat))
    feat = dct(feat, type=2, axis=1, norm="ortho")[:, :13]
    feat = (1 + (22 / 2.0) * np.sin(np.pi * np.arange(13) / 22)) * feat
    feat[:, 0] = np.log(energy)
    r4 = float(np.linalg.norm(cep_o - feat) / np.linalg.norm(cep_o))
    img_o, img_t = om.resize_normalize(cep_o), om.resize_normalize(feat)
    r5 = float(np.abs(img_o - img_t).max())
    return {"frames": int(numframes), "nfft": int(nfft), "powspec_vs_torch_stft_max_rel": r1, "dct_vs_torchaudio_create_dct_max_rel": r2,
            "filterbank_vs_torchaudio_melscale_fbanks_rel_l2": r3, "cepstra_with_torchaudio_filterbank_rel_l2": r4,
            "image_with_torchaudio_filterbank_max_abs": r5}


def main():
    from scipy.io import wavfile

    out = {"torchaudio": __import__("torchaudio").__version__,
           "what": "oracle/mfcc.py vs torchaudio building blocks; stages 1-2 share a definition (agree to rounding), stage 3 differs by design "
                   "(python_speech_features floors the triangle corners to FFT bins), stages 4-5 = the effect of that difference alone"}
    assets = "/root/reference/environments/Minecraft/assets"
    for name in ("noise", "target_good", "target_bad"):
        sr, wav = wavfile.read(os.path.join(assets, name + ".wav"))
        out[name] = dict(samplerate=int(sr), shape=list(wav.shape), **stages(wav, sr))
        print(name, out[name])
    with open(os.path.join(ROOT, "tests", "golden", "mfcc_torchaudio_residual.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
