"""Audio front-end wav -> MFCC image (REF environments/Minecraft/Minecraft.py:226-245) through the C ABI vs oracle/mfcc.py.

* image step (Pillow bicubic resample of a float32 image + the reference's float32 min-max): BIT-EXACT (`==`) against Pillow itself on
  the same cepstra — Pillow is the reference's own call and is installed on the GPU box;
* cepstra (python_speech_features 0.6 restated, float64): the device FFT / mel / DCT run in double with a different summation order
  than numpy's pocketfft, tolerance 1e-9 relative to the cepstra's range (measured ~1e-13);
* end to end: image max-abs error <= 2e-6 (one float32 ulp at 1.0 is 1.2e-7; the float32 cast of cepstra that differ in their last
  double bits can flip one float32 ulp before the normalisation);
* integer rules (`==`): frame count, FFT size and the 28 filter-bank edge bins for the sample rates of the reference's assets.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _signal(n, sr, seed, stereo=False, silence=False):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / sr
    x = 6000 * np.sin(2 * np.pi * 440 * t) + 3000 * np.sin(2 * np.pi * 3100 * t * (1 + 0.2 * t)) + 800 * rng.standard_normal(n)
    x *= 0.5 + 0.5 * np.sin(2 * np.pi * 1.5 * t) ** 2
    if silence:
        x[n // 3: n // 3 + 3 * int(0.025 * sr)] = 0  # whole frames of digital silence: the eps branches of fbank()
    x = np.clip(np.round(x), -32768, 32767).astype(np.int16)
    if stereo:
        y = np.clip(np.round(0.7 * x + 500 * rng.standard_normal(n)), -32768, 32767).astype(np.int16)
        x = np.stack([x, y], axis=1)
    return x


@pytest.mark.parametrize("sr", [44100, 24000, 16000, 8000, 48000])
def test_geometry_and_filterbank_bins(sr):
    from mtrl_b200 import _lib
    from oracle import mfcc as O

    lib = _lib.load()
    nfft, bins = C.c_int(), (C.c_int * 28)()
    _lib.check(lib.dtb_mfcc_filterbank_bins(sr, C.byref(nfft), bins))
    assert nfft.value == O.calculate_nfft(sr, 0.025)
    assert list(bins) == [int(b) for b in O.filterbank_bins(26, nfft.value, sr)]
    for n in [1, 10, int(0.025 * sr), int(0.025 * sr) + 1, sr, 3 * sr + 17]:
        assert lib.dtb_mfcc_num_frames(n, sr) == O.frame_geometry(n, sr)[2], n


@pytest.mark.parametrize("sr,n,stereo,silence", [(44100, 44100 * 2 + 123, False, False), (24000, 30000, True, False),
                                                 (16000, 16000, False, True), (44100, 700, False, False), (8000, 12345, False, False)])
def test_cepstra_vs_oracle(cuda_device, sr, n, stereo, silence):
    from mtrl_b200 import audio
    from oracle import mfcc as O

    x = _signal(n, sr, seed=n, stereo=stereo, silence=silence)
    ref = O.psf_mfcc(x, sr)
    got = audio.mfcc(torch.from_numpy(x), sr, device=cuda_device).cpu().numpy()
    assert got.shape == ref.shape and got.dtype == np.float64
    scale = np.abs(ref).max()
    assert np.abs(got - ref).max() <= 1e-9 * scale, np.abs(got - ref).max() / scale
    # float32 samples take the same path
    got32 = audio.mfcc(torch.from_numpy(x.astype(np.float32)), sr, device=cuda_device).cpu().numpy()
    assert np.array_equal(got32, got)


@pytest.mark.parametrize("frames,size", [(1742, (84, 84)), (984, (84, 84)), (84, (84, 84)), (40, (64, 64)), (1, (84, 84)), (300, (100, 13)),
                                         (5000, (32, 48))])
def test_resize_normalize_bit_exact_vs_pillow(cuda_device, frames, size):
    from mtrl_b200 import audio
    from oracle import mfcc as O

    rng = np.random.default_rng(frames)
    cep = rng.standard_normal((frames, 13)) * np.array([8.0] + [20.0] * 12) + np.array([18.0] + [0.0] * 12)
    if frames == 1:
        cep[0] = np.linspace(-30, 30, 13)
    ref = O.resize_normalize(cep, size)
    got = audio.resize_normalize(torch.from_numpy(cep).to(cuda_device), size).cpu().numpy()
    assert got.shape == ref.shape == (size[1], size[0]) and got.dtype == np.float32
    assert np.array_equal(got, ref), np.abs(got - ref).max()


@pytest.mark.parametrize("sr,n,stereo", [(44100, 44100 * 3, False), (24000, 50000, True), (16000, 8000, False)])
def test_wav_to_mfcc_image_vs_oracle(cuda_device, sr, n, stereo):
    from mtrl_b200 import audio
    from oracle import mfcc as O

    x = _signal(n, sr, seed=7 + n, stereo=stereo)
    ref = O.wav_to_mfcc(sr, x)
    got = audio.wav_to_mfcc(sr, torch.from_numpy(x), device=cuda_device).cpu().numpy()
    assert got.shape == (84, 84) and got.min() == -1.0 and got.max() == 1.0
    assert np.abs(got - ref).max() <= 2e-6, np.abs(got - ref).max()


def test_batch_of_clips_matches_single_clips(cuda_device):
    from mtrl_b200 import audio

    sr, n = 24000, 12000
    clips = np.stack([_signal(n, sr, seed=s) for s in range(5)])
    batch = audio.wav_batch_to_mfcc(sr, torch.from_numpy(clips), device=cuda_device)
    assert batch.shape == (5, 84, 84)
    for i in range(5):
        one = audio.wav_to_mfcc(sr, torch.from_numpy(clips[i]), device=cuda_device)
        assert torch.equal(batch[i], one)


def test_golden_asset_excerpt(cuda_device):
    """1 s excerpt of the reference's own stereo asset (assets/target_good.wav, 24 kHz) and the oracle's image of it
    (tests/golden/mfcc_ref.npz, made by make_golden.py where /root/reference exists)."""
    from mtrl_b200 import audio

    d = np.load(os.path.join(GOLD, "mfcc_ref.npz"))
    got = audio.wav_to_mfcc(int(d["samplerate"]), torch.from_numpy(d["wav"]), device=cuda_device).cpu().numpy()
    assert np.abs(got - d["image"]).max() <= 2e-6
    cep = audio.mfcc(torch.from_numpy(d["wav"]), int(d["samplerate"]), device=cuda_device).cpu().numpy()
    assert np.abs(cep - d["cepstra"]).max() <= 1e-9 * np.abs(d["cepstra"]).max()


def test_feeds_the_audio_tower(cuda_device):
    """MFCC images are what the audio tower consumes (obs channel 1, Minecraft.py:128-132): 84x84 in [-1, 1]."""
    from mtrl_b200 import audio
    from mtrl_b200.encoder import MultimodalStateEncoder

    sr = 24000
    clips = np.stack([_signal(9000, sr, seed=s) for s in range(6)])
    img = audio.wav_batch_to_mfcc(sr, torch.from_numpy(clips), device=cuda_device)  # [6, 84, 84]
    enc = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), "small", "concat", device=cuda_device)
    frames = torch.zeros(2, 3, 1, 84, 84, device=cuda_device)
    states = enc.forward_raw(frames, img.view(2, 3, 1, 84, 84))
    torch.cuda.synchronize()
    assert states.shape == (2, 3, 512) and bool(torch.isfinite(states.float()).all())
