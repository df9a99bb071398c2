"""paddle.audio.backends (reference: python/paddle/audio/backends/{backend,wave_backend,init_backend}.py): PCM16 WAV
I/O through the standard library's ``wave`` module.  (The reference can switch to paddleaudio's soundfile back end
when that package is installed; it is not here, so ``wave_backend`` is the only entry.)"""
from __future__ import annotations

import wave

import numpy as np
import torch

__all__ = ["get_current_backend", "list_available_backends", "set_backend"]


class AudioInfo:
    """Meta data of an audio file: sample_rate, num_samples (frames), num_channels, bits_per_sample, encoding."""

    def __init__(self, sample_rate, num_samples, num_channels, bits_per_sample, encoding):
        self.sample_rate, self.num_samples, self.num_channels = sample_rate, num_samples, num_channels
        self.bits_per_sample, self.encoding = bits_per_sample, encoding

    def __repr__(self):
        return (f"AudioInfo(sample_rate={self.sample_rate}, num_samples={self.num_samples}, num_channels={self.num_channels}, "
                f"bits_per_sample={self.bits_per_sample}, encoding={self.encoding})")


def _open(filepath):
    file_obj = filepath if hasattr(filepath, "read") else open(filepath, "rb")
    try:
        return file_obj, wave.open(file_obj)
    except (wave.Error, EOFError):
        if file_obj is not filepath:
            file_obj.close()
        raise NotImplementedError("Currently only 16-bit PCM WAV files are supported by the wave back end")


def info(filepath):
    file_obj, f = _open(filepath)
    try:
        return AudioInfo(f.getframerate(), f.getnframes(), f.getnchannels(), f.getsampwidth() * 8, "PCM_S")
    finally:
        f.close()
        if file_obj is not filepath:
            file_obj.close()


def load(filepath, frame_offset=0, num_frames=-1, normalize=True, channels_first=True):
    """-> (waveform, sample_rate).  ``normalize``: float32 in [-1, 1) (int16 / 2^15), else the raw int16 values as
    float32; ``channels_first``: ``[channels, frames]`` (default) or ``[frames, channels]``."""
    file_obj, f = _open(filepath)
    try:
        channels, sample_rate, width, frames = f.getnchannels(), f.getframerate(), f.getsampwidth(), f.getnframes()
        if width != 2:
            raise NotImplementedError(f"only 16-bit PCM is supported, the file has {8 * width}-bit samples")
        raw = f.readframes(frames)
    finally:
        f.close()
        if file_obj is not filepath:
            file_obj.close()
    audio = np.frombuffer(raw, dtype=np.int16).astype(np.float32)
    if normalize:
        audio = audio / 2 ** 15
    wav = torch.from_numpy(audio.reshape(frames, channels))
    if num_frames != -1:
        wav = wav[frame_offset:frame_offset + num_frames]
    elif frame_offset:
        wav = wav[frame_offset:]
    if channels_first:
        wav = wav.t().contiguous()
    return wav, sample_rate


def save(filepath, src, sample_rate, channels_first=True, encoding=None, bits_per_sample=16):
    """Write a 2-D float waveform in [-1, 1] as 16-bit PCM."""
    assert src.dim() == 2, "Expected 2D tensor"
    if bits_per_sample not in (None, 16):
        raise ValueError("Invalid bits_per_sample, only support 16 bit")
    audio = src.detach().cpu().numpy()
    if channels_first:
        audio = audio.T
    if audio.dtype.kind == "f":
        audio = np.clip(audio * 2 ** 15, -2 ** 15, 2 ** 15 - 1)
    audio = np.ascontiguousarray(audio.astype("<i2"))
    with wave.open(filepath, "wb") as f:
        f.setnchannels(audio.shape[1])
        f.setsampwidth(2)
        f.setframerate(int(sample_rate))
        f.writeframes(audio.tobytes())


_current = ["wave_backend"]


def list_available_backends():
    return ["wave_backend"]


def get_current_backend():
    return _current[0]


def set_backend(backend_name):
    if backend_name not in list_available_backends():
        raise NotImplementedError(f"backend {backend_name!r} is not available; installed: {list_available_backends()}")
    _current[0] = backend_name
