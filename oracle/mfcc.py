"""ORACLE (test infrastructure): the reference's audio front-end wav -> MFCC image, restated in numpy (float64).

Follows /root/reference/environments/Minecraft/Minecraft.py:226-245 (`_get_mfcc`, `_wav_to_mfcc`; the same code is in
environments/Doorpuzzle/Doorpuzzle.py:264-282):
    mfcc_data = python_speech_features.mfcc(signal=wav, samplerate=samplerate)      # all defaults
    mfcc_data = np.swapaxes(mfcc_data, 0, 1)                                         # time on the x axis -> [13, frames]
    mfcc_data = np.array(Image.fromarray(mfcc_data).resize(size=(W, H), resample=Image.BICUBIC))
    mfcc_data = 1. * (mfcc_data - min) / (max - min);  mfcc_data = mfcc_data * 2. - 1.

PARITY UNPINNED for the cepstra: `python_speech_features` (pinned 0.6 in the reference's poetry.lock) is a third-party dependency that is
neither vendored under /root/reference nor installed in this image.  `psf_mfcc` below restates its published algorithm (base.py /
sigproc.py of release 0.6) with the same numpy / scipy calls it makes itself (numpy.fft.rfft, scipy.fftpack.dct); the glue between
those calls is the restated part.  The image step is NOT restated: it calls Pillow (installed here and on the GPU box) exactly as
the reference does, and the min-max step is the reference's own two lines.
"""
from __future__ import annotations

import decimal
import math

import numpy as np


def round_half_up(number):  # sigproc.round_half_up
    return int(decimal.Decimal(number).quantize(decimal.Decimal("1"), rounding=decimal.ROUND_HALF_UP))


def calculate_nfft(samplerate, winlen):  # base.calculate_nfft (0.6): smallest power of two >= the window length in samples
    window_length_samples = winlen * samplerate
    nfft = 1
    while nfft < window_length_samples:
        nfft *= 2
    return nfft


def hz2mel(hz):
    return 2595 * np.log10(1 + hz / 700.0)


def mel2hz(mel):
    return 700 * (10 ** (mel / 2595.0) - 1)


def filterbank_bins(nfilt, nfft, samplerate, lowfreq=0, highfreq=None):
    highfreq = highfreq or samplerate / 2
    melpoints = np.linspace(hz2mel(lowfreq), hz2mel(highfreq), nfilt + 2)
    return np.floor((nfft + 1) * mel2hz(melpoints) / samplerate)


def get_filterbanks(nfilt=26, nfft=512, samplerate=16000, lowfreq=0, highfreq=None):  # base.get_filterbanks
    b = filterbank_bins(nfilt, nfft, samplerate, lowfreq, highfreq)
    fbank = np.zeros([nfilt, nfft // 2 + 1])
    for j in range(nfilt):
        for i in range(int(b[j]), int(b[j + 1])):
            fbank[j, i] = (i - b[j]) / (b[j + 1] - b[j])
        for i in range(int(b[j + 1]), int(b[j + 2])):
            fbank[j, i] = (b[j + 2] - i) / (b[j + 2] - b[j + 1])
    return fbank


def frame_geometry(n_samples, samplerate, winlen=0.025, winstep=0.01):
    frame_len, frame_step = round_half_up(winlen * samplerate), round_half_up(winstep * samplerate)
    if n_samples <= frame_len:
        numframes = 1
    else:
        numframes = 1 + int(math.ceil((1.0 * n_samples - frame_len) / frame_step))
    return frame_len, frame_step, numframes


def psf_mfcc(signal, samplerate=16000, winlen=0.025, winstep=0.01, numcep=13, nfilt=26, nfft=None, lowfreq=0, highfreq=None,
             preemph=0.97, ceplifter=22, appendEnergy=True):
    """python_speech_features.mfcc (0.6) with its default rectangular window; returns float64 [frames, numcep]."""
    from scipy.fftpack import dct

    nfft = nfft or calculate_nfft(samplerate, winlen)
    signal = np.asarray(signal)
    # sigproc.preemphasis: on a stereo [n, 2] array (the reference passes scipy.io.wavfile's array straight in) the difference acts
    # per channel row, then numpy.append flattens everything to interleaved samples, which the framing below walks
    signal = np.append(signal[0], signal[1:] - preemph * signal[:-1])
    slen = len(signal)
    frame_len, frame_step, numframes = frame_geometry(slen, samplerate, winlen, winstep)
    padlen = int((numframes - 1) * frame_step + frame_len)
    padsignal = np.concatenate((signal, np.zeros((padlen - slen,))))
    idx = np.tile(np.arange(0, frame_len), (numframes, 1)) + np.tile(np.arange(0, numframes * frame_step, frame_step), (frame_len, 1)).T
    frames = padsignal[idx.astype(np.int32)]  # rectangular window: x ones
    pspec = 1.0 / nfft * np.square(np.absolute(np.fft.rfft(frames, nfft)))  # sigproc.powspec
    energy = np.sum(pspec, 1)
    energy = np.where(energy == 0, np.finfo(float).eps, energy)
    fb = get_filterbanks(nfilt, nfft, samplerate, lowfreq, highfreq)
    feat = np.dot(pspec, fb.T)
    feat = np.where(feat == 0, np.finfo(float).eps, feat)
    feat = np.log(feat)
    feat = dct(feat, type=2, axis=1, norm="ortho")[:, :numcep]
    if ceplifter > 0:  # base.lifter
        n = np.arange(feat.shape[1])
        feat = (1 + (ceplifter / 2.0) * np.sin(np.pi * n / ceplifter)) * feat
    if appendEnergy:
        feat[:, 0] = np.log(energy)
    return feat


def resize_normalize(mfcc_data, size=(84, 84)):
    """Minecraft.py:234-243: [frames, 13] float64 -> swapaxes -> Pillow bicubic resize to size=(W, H) -> min-max to [-1, 1] (float32)."""
    from PIL import Image

    mfcc_data = np.swapaxes(mfcc_data, 0, 1)
    mfcc_data = np.array(Image.fromarray(mfcc_data).resize(size=size, resample=Image.BICUBIC))
    min_data = np.min(mfcc_data.flatten())
    max_data = np.max(mfcc_data.flatten())
    mfcc_data = 1.0 * (mfcc_data - min_data) / (max_data - min_data)
    mfcc_data = mfcc_data * 2.0 - 1.0
    return mfcc_data


def wav_to_mfcc(samplerate, wav, size=(84, 84)):
    """Minecraft.py:231-245 `_wav_to_mfcc`."""
    return resize_normalize(psf_mfcc(wav, samplerate), size)
