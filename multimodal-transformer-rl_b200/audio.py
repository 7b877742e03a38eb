"""Host-side mirror of the reference's audio front-end (REF environments/Minecraft/Minecraft.py:226-245 `_get_mfcc` /
`_wav_to_mfcc`; Doorpuzzle.py:264-282): wav samples -> MFCC image in [-1, 1], the input of the audio tower.

``wav_to_mfcc(samplerate, wav, size=(84, 84))`` keeps the reference's argument order and meaning (``wav`` = what
``scipy.io.wavfile.read`` returned: int16 [n] or stereo [n, 2], which python_speech_features flattens to interleaved samples) and
returns the float32 image on the GPU.  Compute runs only in ``libdtb200.so`` (mfcc.cu: shared-memory FFT in double per frame, Pillow's
bicubic resample + min-max in one CTA per clip); there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


_TABLES = {}


def _tables(samplerate, dev):
    """Per (device, sample rate) tables of the frame kernel (twiddles, DCT x lifter, mel weights): built once, reused by every call."""
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), int(samplerate))
    t = _TABLES.get(key)
    if t is None:
        lib = _lib.load()
        n = int(lib.dtb_mfcc_table_doubles(int(samplerate)))
        if n == 0:
            raise _lib.DtbError(f"mfcc: unsupported sample rate {samplerate}")
        t = torch.empty(n, dtype=torch.float64, device=dev)
        _lib.check(lib.dtb_mfcc_build_tables(_lib.stream_ptr(), int(samplerate), _lib.ptr(t)))
        _TABLES[key] = t
    return t


def _as_device_signal(wav, device):
    t = torch.as_tensor(wav)
    if t.dtype not in (torch.int16, torch.float32):
        t = t.to(torch.float32)
    return t.to(device).contiguous()


def mfcc_num_frames(n_samples: int, samplerate: int) -> int:
    return int(_lib.check(_lib.load().dtb_mfcc_num_frames(int(n_samples), int(samplerate))))


def mfcc(signal, samplerate, device="cuda"):
    """python_speech_features.mfcc(signal, samplerate) (release 0.6 defaults) -> float64 [frames, 13] on the device.
    signal: [n] / [n, channels] (flattened to interleaved samples, as numpy.append does inside the reference's call)."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.DtbError("mfcc runs only on an sm_100a CUDA device through libdtb200.so (no CPU fallback)")
    sig = _as_device_signal(signal, dev)
    channels = 1 if sig.dim() == 1 else int(sig.shape[1])  # pre-emphasis runs per channel before numpy.append flattens
    sig = sig.reshape(-1)
    n = sig.numel()
    frames = mfcc_num_frames(n, samplerate)
    out = torch.empty((frames, 13), dtype=torch.float64, device=dev)
    lib = _lib.load()
    tables = _tables(samplerate, sig.device)
    _lib.check(lib.dtb_mfcc_cepstra(_lib.stream_ptr(), _lib.ptr(sig), int(sig.dtype == torch.int16), channels, 1, n, n, int(samplerate),
                                    _lib.ptr(tables), _lib.ptr(out)))
    return out


def resize_normalize(cepstra, size=(84, 84)):
    """[frames, 13] (or [clips, frames, 13]) float64 -> Pillow-bicubic image(s) of size=(W, H), min-max normalised to [-1, 1]."""
    c = cepstra.to(torch.float64).contiguous()
    if not c.is_cuda:
        raise _lib.DtbError("resize_normalize runs only on an sm_100a CUDA device through libdtb200.so (no CPU fallback)")
    batched = c.dim() == 3
    c3 = c if batched else c.unsqueeze(0)
    clips, frames = c3.shape[0], c3.shape[1]
    W, H = int(size[0]), int(size[1])
    scratch = torch.empty((clips, 13, W), dtype=torch.float32, device=c.device)
    img = torch.empty((clips, H, W), dtype=torch.float32, device=c.device)
    _lib.check(_lib.load().dtb_mfcc_resize_normalize(_lib.stream_ptr(), _lib.ptr(c3), clips, frames, W, H, _lib.ptr(scratch), _lib.ptr(img)))
    return img if batched else img[0]


def wav_to_mfcc(samplerate, wav, size=(84, 84), device="cuda"):
    """`_wav_to_mfcc(samplerate, wav)` of the reference (Minecraft.py:231-245) -> float32 [H, W] on the device.
    Batches of equally long clips: wav_batch_to_mfcc."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.DtbError("wav_to_mfcc runs only on an sm_100a CUDA device through libdtb200.so (no CPU fallback)")
    t = _as_device_signal(wav, dev)
    channels = 1 if t.dim() == 1 else int(t.shape[1])
    return _wav_batch(samplerate, t.reshape(1, -1), size, channels)[0]


def wav_batch_to_mfcc(samplerate, clips, size=(84, 84), device="cuda"):
    """clips [n_clips, n_samples] (int16 or float32, mono / already interleaved) -> float32 [n_clips, H, W]: per-step audio of a batch
    of environments in one launch pair."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.DtbError("wav_batch_to_mfcc runs only on an sm_100a CUDA device through libdtb200.so (no CPU fallback)")
    t = _as_device_signal(clips, dev)
    if t.dim() != 2:
        raise ValueError("clips must be [n_clips, n_samples]")
    return _wav_batch(samplerate, t, size)


def _wav_batch(samplerate, sig, size, channels=1):
    lib = _lib.load()
    clips, n = sig.shape
    W, H = int(size[0]), int(size[1])
    ws = torch.empty(int(lib.dtb_mfcc_workspace_bytes(clips, n, int(samplerate), W)), dtype=torch.uint8, device=sig.device)
    img = torch.empty((clips, H, W), dtype=torch.float32, device=sig.device)
    _lib.check(lib.dtb_wav_to_mfcc_image(_lib.stream_ptr(), _lib.ptr(sig), int(sig.dtype == torch.int16), channels, clips, n, sig.stride(0),
                                         int(samplerate), W, H, _lib.ptr(ws), ws.numel(), _lib.ptr(_tables(samplerate, sig.device)),
                                         _lib.ptr(img)))
    return img
